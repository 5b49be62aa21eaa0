// Pipelined cloud uploads: ndtb_stage_target / ndtb_stage_source copy the NEXT target / source to the device on a copy stream
// while the current map build and align run on the handle's stream; ndtb_commit_target / ndtb_commit_source then make the oldest
// staged cloud current by swapping buffers (setInputTarget / setInputSource of the reference, ndt_omp.h:126-130, without the
// copy on the critical path).  A FIFO of two staged clouds per kind, so that a caller can stage cloud i + 1 before committing
// cloud i.  From pinned host memory the copy is asynchronous; from pageable memory cudaMemcpyAsync degrades to a staged copy.
#include "handle.cuh"
#include <cstring>
#include <utility>

namespace ndtb {

void stage_free(ndtb_handle_s* h) {
  for (int k = 0; k < 2; ++k)
    for (int i = 0; i < 2; ++i) {
      ndtb_handle_s::StagedCloud& s = h->staged[k][i];
      cudaFree(s.d); cudaFree(s.d_raw);
      if (s.ready) cudaEventDestroy(s.ready);
      if (s.reusable) cudaEventDestroy(s.reusable);
      s = ndtb_handle_s::StagedCloud();
    }
  if (h->stage_stream) { cudaStreamDestroy(h->stage_stream); h->stage_stream = nullptr; }
  h->staged_count[0] = h->staged_count[1] = 0;
}

#define CKS(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { h->last_error = std::string(#x) + ": " + cudaGetErrorString(e_); return NDTB_ERR_CUDA; } } while (0)

static int stage_cloud(ndtb_handle_s* h, int kind, const float* xyz, size_t n, size_t stride_bytes) {
  if (!h) return NDTB_ERR_INVALID;
  CKS(cudaSetDevice(h->device));
  if (!xyz || n == 0) return set_error(h, NDTB_ERR_EMPTY, kind == 0 ? "empty target cloud" : "empty source cloud");
  if ((stride_bytes != 12 && stride_bytes < sizeof(float4)) || (stride_bytes & 3u)) return set_error(h, NDTB_ERR_INVALID, "stride_bytes must be 12 or >= 16, and a multiple of 4");
  if (h->staged_count[kind] == 2) return set_error(h, NDTB_ERR_NOT_READY, "two clouds are already staged: commit one first");
  if (!h->stage_stream) CKS(cudaStreamCreateWithFlags(&h->stage_stream, cudaStreamNonBlocking));
  ndtb_handle_s::StagedCloud& s = h->staged[kind][(h->staged_head[kind] + h->staged_count[kind]) & 1];
  if (!s.ready) {
    CKS(cudaEventCreateWithFlags(&s.ready, cudaEventDisableTiming));
    CKS(cudaEventCreateWithFlags(&s.reusable, cudaEventDisableTiming));
    CKS(cudaEventRecord(s.reusable, h->stream));
  }
  if (n > s.alloc) {
    CKS(cudaStreamSynchronize(h->stage_stream));
    cudaFree(s.d);
    s.d = nullptr; s.alloc = 0;
    const size_t cap = n + n / 8 + 256;
    CKS(cudaMalloc(&s.d, cap * sizeof(float4)));
    s.alloc = cap;
  }
  // the buffer may be a former d_target / d_source: whatever the handle's stream still had to read from it comes first
  CKS(cudaStreamWaitEvent(h->stage_stream, s.reusable, 0));
  {
    const int rc = copy_cloud_to_device(h, xyz, n, stride_bytes, s.d, h->stage_stream, &s.d_raw, &s.raw_alloc);
    if (rc != NDTB_OK) return rc;
  }
  CKS(cudaEventRecord(s.ready, h->stage_stream));
  s.n = n;
  h->staged_count[kind] += 1;
  return NDTB_OK;
}

// Oldest staged cloud -> d_target / d_source (buffers swapped; the previous current buffer becomes the free staging slot).
static int commit_cloud(ndtb_handle_s* h, int kind) {
  if (!h) return NDTB_ERR_INVALID;
  CKS(cudaSetDevice(h->device));
  if (h->staged_count[kind] == 0) return set_error(h, NDTB_ERR_NOT_READY, kind == 0 ? "no staged target" : "no staged source");
  ndtb_handle_s::StagedCloud& s = h->staged[kind][h->staged_head[kind]];
  CKS(cudaStreamWaitEvent(h->stream, s.ready, 0));
  float4** cur = kind == 0 ? &h->d_target : &h->d_source;
  size_t* cur_alloc = kind == 0 ? &h->target_alloc : &h->source_alloc;
  bool* owned = kind == 0 ? &h->target_owned : &h->source_owned;
  if (!*owned) { *cur = nullptr; *cur_alloc = 0; *owned = true; }   // a caller's device buffer (ndtb_set_*_device) is not ours to recycle
  std::swap(*cur, s.d);
  std::swap(*cur_alloc, s.alloc);
  (kind == 0 ? h->n_target : h->n_source) = s.n;
  s.n = 0;
  CKS(cudaEventRecord(s.reusable, h->stream));   // everything queued so far on the handle's stream may still read the old buffer
  h->staged_head[kind] ^= 1;
  h->staged_count[kind] -= 1;
  return NDTB_OK;
}

}  // namespace ndtb

extern "C" {

int ndtb_stage_target(ndtb_handle h, const float* xyz, size_t n, size_t stride_bytes) { return ndtb::stage_cloud(h, 0, xyz, n, stride_bytes); }
int ndtb_stage_source(ndtb_handle h, const float* xyz, size_t n, size_t stride_bytes) { return ndtb::stage_cloud(h, 1, xyz, n, stride_bytes); }

int ndtb_commit_target(ndtb_handle h) {
  const int rc = ndtb::commit_cloud(h, 0);
  if (rc != NDTB_OK) return rc;
  return ndtb::rebuild_grid(h);
}

int ndtb_commit_source(ndtb_handle h) { return ndtb::commit_cloud(h, 1); }

int ndtb_stage_sync(ndtb_handle h) {
  if (!h) return NDTB_ERR_INVALID;
  if (!h->stage_stream) return NDTB_OK;
  if (cudaSetDevice(h->device) != cudaSuccess || cudaStreamSynchronize(h->stage_stream) != cudaSuccess) return ndtb::set_error(h, NDTB_ERR_CUDA, "staged copy failed");
  return NDTB_OK;
}

int ndtb_staged_count(ndtb_handle h, int* targets, int* sources) {
  if (!h) return NDTB_ERR_INVALID;
  if (targets) *targets = h->staged_count[0];
  if (sources) *sources = h->staged_count[1];
  return NDTB_OK;
}

}  // extern "C"
