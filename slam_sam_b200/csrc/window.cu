// The step before setInputTarget in the reference's pipelines: the sliding-window target is assembled from the
// last few archived keyframe clouds, each moved to the map frame with its current pose estimate and concatenated
//   run/pipeline_ligo_tc.cpp:519-526   for currID in targetID: pcl::transformPointCloud(*archive.points, *cur, Tb2m.matrix()); *target += *cur;
//   run/pipeline_lo_svn.cpp            (same loop)
// Here the archive lives in HBM (one upload per keyframe, when it is created) and the window is assembled by one
// kernel straight into the target buffer the voxel build reads: the per-frame host -> device copy of the whole
// map window (16 MB for 1 M points) disappears; only the new scan crosses PCIe.
//
// Arithmetic: pcl::transformPointCloud<PointT, double> (PCL >= 1.14, common/impl/transforms.hpp, generic
// Transformer<Scalar>::se3 -- the reference builds without -mavx, so the AVX specialisation is not used):
//   out[r] = float( T(r,0)*double(x) + T(r,1)*double(y) + T(r,2)*double(z) + T(r,3) ),  left to right;  out[3] = 1.
#include "handle.cuh"
#include <algorithm>
#include <cstring>
#include <string>
#include <vector>

using namespace ndtb;

#define CKH(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { h->last_error = std::string(#x) + ": " + cudaGetErrorString(e_); return NDTB_ERR_CUDA; } } while (0)

namespace ndtb {

constexpr int WINDOW_MAX_SEGMENTS = 16;  // keyframes per launch (the reference uses <= 5); longer windows take several launches

struct WindowSegments {
  const float4* src[WINDOW_MAX_SEGMENTS];
  unsigned long long begin[WINDOW_MAX_SEGMENTS + 1];  // output offset of every segment, begin[count] = total
  double T[WINDOW_MAX_SEGMENTS][12];                  // rows 0..2 of the keyframe pose
  int count;
};

__global__ void __launch_bounds__(256) window_assemble_kernel(const __grid_constant__ WindowSegments seg, float4* __restrict__ out) {
  const unsigned long long total = seg.begin[seg.count];
  for (unsigned long long i = blockIdx.x * static_cast<unsigned long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<unsigned long long>(gridDim.x) * blockDim.x) {
    int s = 0;
#pragma unroll
    for (int k = 1; k < WINDOW_MAX_SEGMENTS; ++k) s += (k < seg.count && i >= seg.begin[k]) ? 1 : 0;
    const float4 p = __ldg(seg.src[s] + (i - seg.begin[s]));
    const double x = static_cast<double>(p.x), y = static_cast<double>(p.y), z = static_cast<double>(p.z);
    const double* T = seg.T[s];
    float4 o;
    o.x = static_cast<float>(((T[0] * x + T[1] * y) + T[2] * z) + T[3]);
    o.y = static_cast<float>(((T[4] * x + T[5] * y) + T[6] * z) + T[7]);
    o.z = static_cast<float>(((T[8] * x + T[9] * y) + T[10] * z) + T[11]);
    o.w = 1.0f;
    out[i] = o;
  }
}

}  // namespace ndtb

extern "C" {

int ndtb_keyframe_put(ndtb_handle h, int64_t id, const float* xyz, size_t n, size_t stride_bytes) {
  if (!h) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  if (!xyz || n == 0) return set_error(h, NDTB_ERR_EMPTY, "empty keyframe cloud");
  ndtb_handle_s::Keyframe& k = h->keyframes[id];
  size_t alloc = k.d_points ? k.n : 0;
  bool owned = true;
  if (k.d_points && k.n < n) { cudaFree(k.d_points); k.d_points = nullptr; alloc = 0; }
  const int rc = upload_cloud(h, xyz, n, stride_bytes, &k.d_points, &alloc, &owned);
  if (rc != NDTB_OK) { if (!k.d_points) h->keyframes.erase(id); return rc; }
  k.n = n;
  return NDTB_OK;
}

int ndtb_keyframe_drop(ndtb_handle h, int64_t id) {
  if (!h) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  auto it = h->keyframes.find(id);
  if (it == h->keyframes.end()) return set_error(h, NDTB_ERR_INVALID, "unknown keyframe id");
  cudaStreamSynchronize(h->stream);
  cudaFree(it->second.d_points);
  h->keyframes.erase(it);
  return NDTB_OK;
}

int ndtb_keyframe_count(ndtb_handle h, size_t* count, size_t* total_points) {
  if (!h) return NDTB_ERR_INVALID;
  size_t pts = 0;
  for (const auto& kv : h->keyframes) pts += kv.second.n;
  if (count) *count = h->keyframes.size();
  if (total_points) *total_points = pts;
  return NDTB_OK;
}

int ndtb_set_target_window(ndtb_handle h, int n_keyframes, const int64_t* ids, const double* poses16) {
  if (!h || n_keyframes < 0 || (n_keyframes && (!ids || !poses16))) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  size_t total = 0;
  for (int i = 0; i < n_keyframes; ++i) {
    auto it = h->keyframes.find(ids[i]);
    if (it == h->keyframes.end()) return set_error(h, NDTB_ERR_INVALID, "unknown keyframe id in the window");
    total += it->second.n;
  }
  if (total == 0) { h->n_target = 0; h->grid.built = false; return set_error(h, NDTB_ERR_EMPTY, "empty target cloud"); }
  if (!h->target_owned) { h->d_target = nullptr; h->target_alloc = 0; h->target_owned = true; }
  if (total > h->target_alloc) {
    cudaFree(h->d_target);
    h->d_target = nullptr;
    const size_t cap = total + total / 8 + 256;
    CKH(cudaMalloc(&h->d_target, cap * sizeof(float4)));
    h->target_alloc = cap;
  }
  size_t done = 0;
  for (int first = 0; first < n_keyframes; first += WINDOW_MAX_SEGMENTS) {
    WindowSegments seg;
    std::memset(&seg, 0, sizeof(seg));
    seg.count = std::min(WINDOW_MAX_SEGMENTS, n_keyframes - first);
    size_t off = 0;
    for (int k = 0; k < seg.count; ++k) {
      const ndtb_handle_s::Keyframe& kf = h->keyframes[ids[first + k]];
      seg.src[k] = kf.d_points;
      seg.begin[k] = off;
      std::memcpy(seg.T[k], poses16 + 16 * static_cast<size_t>(first + k), 12 * sizeof(double));
      off += kf.n;
    }
    seg.begin[seg.count] = off;
    const unsigned long long blocks_needed = (off + 255) / 256;
    const int blocks = static_cast<int>(std::min<unsigned long long>(blocks_needed, 148ull * 16));
    if (blocks > 0) {
      window_assemble_kernel<<<blocks, 256, 0, h->stream>>>(seg, h->d_target + done);
      h->launches += 1;
    }
    done += off;
  }
  CKH(cudaGetLastError());
  h->n_target = total;
  return rebuild_grid(h);
}

// pcl::VoxelGrid<PointT>::filter (leaf x leaf x leaf) of a host cloud on the device.  dest: 0 = copy the centroids back to
// `out` (packed x,y,z,1 floats, at most `capacity` points), 1 = they become the SOURCE cloud, 2 = they become the TARGET
// cloud (and the voxel map is rebuilt); in every case *n_out = number of occupied leaves = points produced.
int ndtb_voxelgrid_filter(ndtb_handle h, const float* xyz, size_t n, size_t stride_bytes, float leaf, int dest, float* out, size_t capacity,
                          size_t* n_out) {
  if (!h || dest < 0 || dest > 2) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  if (!xyz || n == 0) return set_error(h, NDTB_ERR_EMPTY, "empty cloud");
  int rc = upload_cloud(h, xyz, n, stride_bytes, &h->d_ds_in, &h->ds_in_alloc, &h->ds_in_owned);
  if (rc != NDTB_OK) return rc;
  float4** d_out = &h->d_ds_out;
  size_t* alloc = &h->ds_out_alloc;
  if (dest == 1) { if (!h->source_owned) { h->d_source = nullptr; h->source_alloc = 0; h->source_owned = true; } d_out = &h->d_source; alloc = &h->source_alloc; }
  if (dest == 2) { if (!h->target_owned) { h->d_target = nullptr; h->target_alloc = 0; h->target_owned = true; } d_out = &h->d_target; alloc = &h->target_alloc; }
  if (n > *alloc) {
    CKH(cudaStreamSynchronize(h->stream));
    cudaFree(*d_out);
    *d_out = nullptr;
    const size_t cap = n + n / 8 + 256;
    CKH(cudaMalloc(d_out, cap * sizeof(float4)));
    *alloc = cap;
  }
  size_t produced = 0;
  rc = voxelgrid_downsample(h->ds_grid, h->d_ds_in, n, leaf, h->params.is_dense, *d_out, &produced, h->stream);
  if (rc == NDTB_ERR_OVERFLOW) return set_error(h, rc, "leaf size too small for the input dataset: integer indices would overflow");
  if (rc != NDTB_OK) return set_error(h, rc, "voxel-grid filter failed");
  if (n_out) *n_out = produced;
  if (dest == 0) {
    const size_t m = std::min(produced, capacity);
    if (out && m) {
      CKH(cudaMemcpyAsync(out, *d_out, m * sizeof(float4), cudaMemcpyDeviceToHost, h->stream));
      CKH(cudaStreamSynchronize(h->stream));
    }
    return NDTB_OK;
  }
  if (dest == 1) { h->n_source = produced; return produced ? NDTB_OK : set_error(h, NDTB_ERR_EMPTY, "empty source cloud"); }
  h->n_target = produced;
  return rebuild_grid(h);
}

int ndtb_get_target_points(ndtb_handle h, float* xyzw, size_t capacity_points, size_t* n_points) {
  if (!h) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  if (n_points) *n_points = h->n_target;
  if (!xyzw || capacity_points == 0) return NDTB_OK;
  const size_t n = std::min(capacity_points, h->n_target);
  if (n == 0) return NDTB_OK;
  CKH(cudaMemcpyAsync(xyzw, h->d_target, n * sizeof(float4), cudaMemcpyDeviceToHost, h->stream));
  CKH(cudaStreamSynchronize(h->stream));
  return NDTB_OK;
}

}  // extern "C"
