// Scan producer on the device: the reference's LidarCallback (the step before setInputSource).
//   LidarCallback::Initialize         src/lidarcallback.cpp:191-341   lookup tables (host, once)
//   LidarCallback::DecodePacketRng19  src/lidarcallback.cpp:631-878   packets -> LidarFrame (include/dataframe.hpp)
// The reference decodes packet by packet on one core and push_back()s every surviving return.  Here the packets of
// the running frame are DMA-ed into HBM as they arrive and, when a packet with a new frame id closes the frame, ONE
// kernel decodes the whole revolution: one CTA per packet (or per four columns of it), one warp per column, one lane per (strided) channel.
// The reference's output order (packet, column, channel) is kept by an order-preserving compaction: warp ballots
// inside a column, a 32-entry scan over the columns of a packet, and a decoupled look-back over the packets (every
// CTA publishes {count, time stamp of its first point} under a per-launch epoch; CTAs take their packet through a
// ticket, so a CTA only ever waits for CTAs that are already running).
// Arithmetic: the scalar branch of the reference (-mavx2 -mfma are commented out in CMakeLists.txt:85-89):
//   range_m = float(range_mm) * 0.001f;  pt = range_m * x_1 + x_2   (two roundings; this library is built -fmad=false)
// HBM-bound byte work: per frame the packets are read once (12 B per decoded pixel + headers), 12 B of lookup table
// per pixel, 42 B written per surviving point.
#include "../../include/ndtb.h"
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <limits>
#include <string>
#include <vector>

namespace {

constexpr int LIDAR_MAX_ROUNDS = 4;  // 32 lanes x 4 = 128 decoded channels per column
constexpr int LIDAR_MAX_SLOTS = 64;

struct LidarKParams {
  int cpf, ppc, cpp, stride, subset, rounds;
  int cpc, upp;  // columns per CTA ("unit"), units per packet
  int block_bytes, pitch;
  int legacy;    // LEGACY profile (DecodePacketLegacy): no packet header, 16-byte column header with the frame id, block status word
  float r_min, r_max, z_min, z_max;
  int refl_thr;
  float bmin[3], bmax[3];
};
// LEGACY frames are cut at COLUMN granularity (the frame id travels in every column header): unit t of a frame is a packet with
// the column range [cb, ce) that belongs to the frame, and the host-tracked latest_timestamp_s at the start of that packet.
constexpr int LEGACY_MAX_UNITS = 288;
struct LegacyUnits {
  unsigned char cb[LEGACY_MAX_UNITS], ce[LEGACY_MAX_UNITS];
  double latest_at_start[LEGACY_MAX_UNITS];
  double prev_completion_ts;   // latest_timestamp_s when the previous frame was completed by this frame's first packet, < 0: none
  double end_ts;               // latest_timestamp_s when this frame was completed (timestamp_end)
};
struct PacketRec { double first_ts; unsigned count; unsigned flag; };
struct Ctl { unsigned ticket, done; };
struct FrameArrays {
  float4* xyzw; unsigned short *c_id, *m_id, *refl, *sig, *nir; double* ts; float* rel; unsigned capacity;
};

__device__ __forceinline__ unsigned ld_volatile_u32(const unsigned* p) { return *reinterpret_cast<const volatile unsigned*>(p); }

__global__ void __launch_bounds__(1024) lidar_decode_kernel(const __grid_constant__ LidarKParams P, const unsigned char* __restrict__ packets, int n_units,
                                                            const float* __restrict__ x1, const float* __restrict__ y1, const float* __restrict__ z1,
                                                            const float* __restrict__ x2, const float* __restrict__ y2, const float* __restrict__ z2,
                                                            Ctl* ctl, PacketRec* recs, unsigned epoch, double* chain, const __grid_constant__ FrameArrays out,
                                                            ndtb_lidar_frame_info* info, const __grid_constant__ LegacyUnits U) {
  __shared__ unsigned s_cnt[32], s_off[32];
  __shared__ double s_ts[32];
  __shared__ unsigned s_ticket, s_base;
  __shared__ double s_frame_ts;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;  // w = column inside the unit
  if (threadIdx.x == 0) s_ticket = atomicAdd(&ctl->ticket, 1u);
  __syncthreads();
  const unsigned t = s_ticket;  // unit index inside the frame: packet t / upp, columns (t % upp) * cpc ...
  const unsigned pk = t / static_cast<unsigned>(P.upp);
  const int col = static_cast<int>(t - pk * P.upp) * P.cpc + w;
  const unsigned* blk = reinterpret_cast<const unsigned*>(packets + static_cast<size_t>(pk) * P.pitch + (P.legacy ? 0 : 32) + static_cast<size_t>(col) * P.block_bytes);
  // column header (:676-700): 64-bit time stamp, measurement id, status; LEGACY (:391-441): time stamp, measurement id, frame id,
  // encoder count, and the block status word behind the pixels
  const unsigned ts_lo = __ldg(blk), ts_hi = __ldg(blk + 1), hdr2 = __ldg(blk + 2);
  const unsigned long long ts_ns = (static_cast<unsigned long long>(ts_hi) << 32) | ts_lo;
  const double col_ts = fmod(static_cast<double>(ts_ns) * 1e-9, 86400.0);
  const unsigned m_id = hdr2 & 0xffffu;
  bool col_ok = m_id < static_cast<unsigned>(P.cpf);
  if (P.legacy) col_ok = col_ok && col >= static_cast<int>(U.cb[t]) && col < static_cast<int>(U.ce[t]) && __ldg(blk + 4 + 3 * P.ppc) == 0xFFFFFFFFu;
  else col_ok = col_ok && ((hdr2 >> 16) & 1u);
  const int px0 = P.legacy ? 4 : 3;                                    // first pixel word behind the column header
  const unsigned range_mask = P.legacy ? 0x000FFFFFu : 0x0007FFFFu;    // :480 (20 bits) / :807-810 (19 bits)
  const float ox = col_ok ? __ldg(x2 + m_id) : 0.f, oy = col_ok ? __ldg(y2 + m_id) : 0.f, oz = col_ok ? __ldg(z2 + m_id) : 0.f;

  float px[LIDAR_MAX_ROUNDS], py[LIDAR_MAX_ROUNDS], pz[LIDAR_MAX_ROUNDS];
  unsigned w0[LIDAR_MAX_ROUNDS], w1[LIDAR_MAX_ROUNDS], w2[LIDAR_MAX_ROUNDS], ballot[LIDAR_MAX_ROUNDS];
  // the pixel words do not depend on the column header: all loads of the column are in flight together
#pragma unroll
  for (int r = 0; r < LIDAR_MAX_ROUNDS; ++r) {
    const int s = r * 32 + lane;
    w0[r] = w1[r] = w2[r] = 0u;
    if (r < P.rounds && s < P.subset) {
      const unsigned* px_w = blk + px0 + 3 * (s * P.stride);
      w0[r] = __ldg(px_w); w1[r] = __ldg(px_w + 1); w2[r] = __ldg(px_w + 2);
    }
  }
  unsigned cnt = 0;
#pragma unroll
  for (int r = 0; r < LIDAR_MAX_ROUNDS; ++r) {
    bool keep = false;
    px[r] = py[r] = pz[r] = 0.f;
    const int s = r * 32 + lane;
    if (r < P.rounds && col_ok && s < P.subset) {
      const int c_id = s * P.stride;
      const unsigned range_mm = w0[r] & range_mask;
      const float range_m = static_cast<float>(range_mm) * 0.001f;
      if (!(range_m < P.r_min || range_m > P.r_max || range_m == 0.f)) {  // :813
        const size_t k = static_cast<size_t>(m_id) * P.ppc + c_id;
        px[r] = range_m * __ldg(x1 + k) + ox;                     // :830-832
        py[r] = range_m * __ldg(y1 + k) + oy;
        pz[r] = range_m * __ldg(z1 + k) + oz;
        const bool in_box = px[r] >= P.bmin[0] && px[r] <= P.bmax[0] && py[r] >= P.bmin[1] && py[r] <= P.bmax[1] && pz[r] >= P.bmin[2] && pz[r] <= P.bmax[2];
        const int refl = static_cast<int>(w1[r] & 0xffu);
        keep = !in_box && ((pz[r] >= P.z_min && pz[r] <= P.z_max) || refl >= P.refl_thr);  // :839
      }
    }
    ballot[r] = __ballot_sync(0xffffffffu, keep);
    cnt += __popc(ballot[r]);
  }
  if (lane == 0) { s_cnt[w] = cnt; s_ts[w] = col_ts; }
  __syncthreads();

  if (w == 0) {
    // columns of this packet: exclusive scan of the counts, time stamp of the first point
    const unsigned c = lane < P.cpc ? s_cnt[lane] : 0u;
    unsigned incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    s_off[lane] = incl - c;
    const unsigned total = __shfl_sync(0xffffffffu, incl, 31);
    const unsigned nonempty = __ballot_sync(0xffffffffu, c > 0u);
    const double own_first_ts = nonempty ? s_ts[__ffs(nonempty) - 1] : 0.0;
    if (lane == 0) {  // publish, then look back
      recs[t].first_ts = own_first_ts;
      recs[t].count = total;
      __threadfence();
      *reinterpret_cast<volatile unsigned*>(&recs[t].flag) = epoch;
    }
    unsigned base = 0, first_idx = 0xffffffffu;
    for (unsigned i = lane; i < t; i += 32) {
      while (ld_volatile_u32(&recs[i].flag) != epoch) { }
      __threadfence();
      const unsigned ci = ld_volatile_u32(&recs[i].count);
      base += ci;
      if (ci > 0u && i < first_idx) first_idx = i;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      base += __shfl_xor_sync(0xffffffffu, base, o);
      first_idx = min(first_idx, __shfl_xor_sync(0xffffffffu, first_idx, o));
    }
    if (lane == 0) {
      double frame_ts = 0.0;
      bool have_first = false;
      if (first_idx != 0xffffffffu) { frame_ts = *reinterpret_cast<const volatile double*>(&recs[first_idx].first_ts); have_first = true; }
      else if (total > 0u) { frame_ts = own_first_ts; first_idx = t; have_first = true; }
      s_base = base;
      s_frame_ts = frame_ts;
      if (t == static_cast<unsigned>(n_units) - 1u) {  // the frame's scalars (:659-663, 856-862) and the hand-over to the next frame
        const unsigned n = base + total;
        const double prev = *chain;
        info->numberpoints = n;
        info->timestamp = have_first ? frame_ts : -1.7976931348623157e308;
        if (P.legacy) {
          // prev_frame_completed_latest_ts (:385-388, :428): latest_timestamp_s at the start of the packet that holds the first
          // point -- unless that packet also completed the previous frame (with points), then the value at that completion
          double pv = U.latest_at_start[first_idx < static_cast<unsigned>(LEGACY_MAX_UNITS) ? first_idx : 0];
          if (first_idx == 0u && prev > 0.0 && U.prev_completion_ts >= 0.0) pv = U.prev_completion_ts;
          info->timestamp_end = U.end_ts;
          info->interframe_timedelta = have_first ? (pv > 0.0 ? fmax(0.0, frame_ts - pv) : 0.0) : -1.7976931348623157e308;
          *chain = n > 0u ? 1.0 : 0.0;   // "the frame before had points"
        } else {
          const double latest = s_ts[P.cpc - 1];  // latest_timestamp_s: every column updates it, valid or not (:685)
          info->timestamp_end = latest;
          // interframe_timedelta only sees the previous frame's end when the first point comes from the packet that closed it
          info->interframe_timedelta = have_first ? ((first_idx < static_cast<unsigned>(P.upp) && prev > 0.0) ? fmax(0.0, frame_ts - prev) : 0.0) : -1.7976931348623157e308;
          *chain = n > 0u ? latest : 0.0;
        }
      }
    }
  }
  __syncthreads();

  const double frame_ts = s_frame_ts;
  const float rel = frame_ts > 0.0 ? static_cast<float>(fmax(0.0, col_ts - frame_ts)) : 0.0f;  // :842-844
  unsigned pos = s_base + s_off[w];
  const unsigned lt = (1u << lane) - 1u;
#pragma unroll
  for (int r = 0; r < LIDAR_MAX_ROUNDS; ++r) {
    if ((ballot[r] >> lane) & 1u) {
      const unsigned k = pos + __popc(ballot[r] & lt);
      if (k < out.capacity) {
        out.xyzw[k] = make_float4(px[r], py[r], pz[r], 1.0f);
        out.c_id[k] = static_cast<unsigned short>((r * 32 + lane) * P.stride);
        out.m_id[k] = static_cast<unsigned short>(m_id);
        out.ts[k] = col_ts;
        out.rel[k] = rel;
        out.refl[k] = static_cast<unsigned short>(w1[r] & 0xffu);
        out.sig[k] = static_cast<unsigned short>(w1[r] >> 16);
        out.nir[k] = static_cast<unsigned short>(w2[r] & 0xffffu);
      }
    }
    pos += __popc(ballot[r]);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned d = atomicAdd(&ctl->done, 1u);
    if (d == static_cast<unsigned>(n_units) - 1u) { ctl->ticket = 0u; ctl->done = 0u; }  // ready for the next frame's launch
  }
}

// adjugate / determinant, the arithmetic assumed for Eigen::Matrix4d::inverse (lidarcallback.cpp:280)
bool invert4(const double* m, double* inv) {
  double a[16];
  a[0] = m[5] * m[10] * m[15] - m[5] * m[11] * m[14] - m[9] * m[6] * m[15] + m[9] * m[7] * m[14] + m[13] * m[6] * m[11] - m[13] * m[7] * m[10];
  a[4] = -m[4] * m[10] * m[15] + m[4] * m[11] * m[14] + m[8] * m[6] * m[15] - m[8] * m[7] * m[14] - m[12] * m[6] * m[11] + m[12] * m[7] * m[10];
  a[8] = m[4] * m[9] * m[15] - m[4] * m[11] * m[13] - m[8] * m[5] * m[15] + m[8] * m[7] * m[13] + m[12] * m[5] * m[11] - m[12] * m[7] * m[9];
  a[12] = -m[4] * m[9] * m[14] + m[4] * m[10] * m[13] + m[8] * m[5] * m[14] - m[8] * m[6] * m[13] - m[12] * m[5] * m[10] + m[12] * m[6] * m[9];
  a[1] = -m[1] * m[10] * m[15] + m[1] * m[11] * m[14] + m[9] * m[2] * m[15] - m[9] * m[3] * m[14] - m[13] * m[2] * m[11] + m[13] * m[3] * m[10];
  a[5] = m[0] * m[10] * m[15] - m[0] * m[11] * m[14] - m[8] * m[2] * m[15] + m[8] * m[3] * m[14] + m[12] * m[2] * m[11] - m[12] * m[3] * m[10];
  a[9] = -m[0] * m[9] * m[15] + m[0] * m[11] * m[13] + m[8] * m[1] * m[15] - m[8] * m[3] * m[13] - m[12] * m[1] * m[11] + m[12] * m[3] * m[9];
  a[13] = m[0] * m[9] * m[14] - m[0] * m[10] * m[13] - m[8] * m[1] * m[14] + m[8] * m[2] * m[13] + m[12] * m[1] * m[10] - m[12] * m[2] * m[9];
  a[2] = m[1] * m[6] * m[15] - m[1] * m[7] * m[14] - m[5] * m[2] * m[15] + m[5] * m[3] * m[14] + m[13] * m[2] * m[7] - m[13] * m[3] * m[6];
  a[6] = -m[0] * m[6] * m[15] + m[0] * m[7] * m[14] + m[4] * m[2] * m[15] - m[4] * m[3] * m[14] - m[12] * m[2] * m[7] + m[12] * m[3] * m[6];
  a[10] = m[0] * m[5] * m[15] - m[0] * m[7] * m[13] - m[4] * m[1] * m[15] + m[4] * m[3] * m[13] + m[12] * m[1] * m[7] - m[12] * m[3] * m[5];
  a[14] = -m[0] * m[5] * m[14] + m[0] * m[6] * m[13] + m[4] * m[1] * m[14] - m[4] * m[2] * m[13] - m[12] * m[1] * m[6] + m[12] * m[2] * m[5];
  a[3] = -m[1] * m[6] * m[11] + m[1] * m[7] * m[10] + m[5] * m[2] * m[11] - m[5] * m[3] * m[10] - m[9] * m[2] * m[7] + m[9] * m[3] * m[6];
  a[7] = m[0] * m[6] * m[11] - m[0] * m[7] * m[10] - m[4] * m[2] * m[11] + m[4] * m[3] * m[10] + m[8] * m[2] * m[7] - m[8] * m[3] * m[6];
  a[11] = -m[0] * m[5] * m[11] + m[0] * m[7] * m[9] + m[4] * m[1] * m[11] - m[4] * m[3] * m[9] - m[8] * m[1] * m[7] + m[8] * m[3] * m[5];
  a[15] = m[0] * m[5] * m[10] - m[0] * m[6] * m[9] - m[4] * m[1] * m[10] + m[4] * m[2] * m[9] + m[8] * m[1] * m[6] - m[8] * m[2] * m[5];
  const double det = m[0] * a[0] + m[1] * a[4] + m[2] * a[8] + m[3] * a[12];
  if (det == 0.0) return false;
  for (int i = 0; i < 16; ++i) inv[i] = a[i] / det;
  return true;
}

struct Slot {
  unsigned char* d_packets = nullptr;
  FrameArrays arr{};
  ndtb_lidar_frame_info* h_info = nullptr;  // pinned; the kernel writes the frame's scalars here
  cudaEvent_t done = nullptr, k0 = nullptr, k1 = nullptr;
  int n_packets = 0;
  unsigned short frame_id = 0;
  LegacyUnits lu{};                 // LEGACY: column ranges of the units, time stamps the host tracked
  unsigned long long closing_seq = 0;  // LEGACY: number of the packet whose column closed the frame
};

}  // namespace

struct ndtb_lidar_s {
  int device = 0;
  cudaStream_t stream = nullptr;
  ndtb_lidar_config cfg{};
  LidarKParams kp{};
  size_t expected_size = 0;
  int cap_packets = 0;
  unsigned cap_points = 0;
  std::vector<float> x1, y1, z1, x2, y2, z2;
  float *d_x1 = nullptr, *d_y1 = nullptr, *d_z1 = nullptr, *d_x2 = nullptr, *d_y2 = nullptr, *d_z2 = nullptr;
  Ctl* d_ctl = nullptr;
  PacketRec* d_recs = nullptr;
  double* d_chain = nullptr;
  unsigned epoch = 0;
  std::vector<Slot*> slots;
  std::deque<Slot*> free_slots, closed;
  Slot* active = nullptr;
  Slot* current = nullptr;  // the popped frame
  unsigned short frame_id_ = 0;  // LidarCallback::frame_id_
  double latest_ts = 0.0;        // LEGACY: LidarCallback::latest_timestamp_s, tracked on the host from the column headers
  unsigned long long packet_seq = 0;
  uint64_t launches = 0;
  float last_decode_us = 0.f;
  std::string last_error;
};

namespace {

#define CKL(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { l->last_error = std::string(#x) + ": " + cudaGetErrorString(e_); return NDTB_ERR_CUDA; } } while (0)

int lidar_fail(ndtb_lidar_s* l, int code, const char* what) { l->last_error = what; return code; }

int new_slot(ndtb_lidar_s* l, Slot** out) {
  if (l->slots.size() >= static_cast<size_t>(LIDAR_MAX_SLOTS)) return lidar_fail(l, NDTB_ERR_OVERFLOW, "too many closed frames waiting: call ndtb_lidar_pop_frame");
  Slot* s = new Slot();
  l->slots.push_back(s);
  const size_t n = l->cap_points;
  CKL(cudaMalloc(&s->d_packets, static_cast<size_t>(l->cap_packets) * l->kp.pitch));
  CKL(cudaMalloc(&s->arr.xyzw, n * sizeof(float4)));
  CKL(cudaMalloc(&s->arr.ts, n * sizeof(double)));
  CKL(cudaMalloc(&s->arr.rel, n * sizeof(float)));
  CKL(cudaMalloc(&s->arr.c_id, n * 2)); CKL(cudaMalloc(&s->arr.m_id, n * 2)); CKL(cudaMalloc(&s->arr.refl, n * 2));
  CKL(cudaMalloc(&s->arr.sig, n * 2)); CKL(cudaMalloc(&s->arr.nir, n * 2));
  s->arr.capacity = l->cap_points;
  CKL(cudaHostAlloc(&s->h_info, sizeof(ndtb_lidar_frame_info), cudaHostAllocMapped));
  std::memset(s->h_info, 0, sizeof(ndtb_lidar_frame_info));
  CKL(cudaEventCreateWithFlags(&s->done, cudaEventDisableTiming));
  CKL(cudaEventCreate(&s->k0)); CKL(cudaEventCreate(&s->k1));
  *out = s;
  return NDTB_OK;
}

int take_slot(ndtb_lidar_s* l, Slot** out) {
  if (!l->free_slots.empty()) { *out = l->free_slots.front(); l->free_slots.pop_front(); }
  else { const int rc = new_slot(l, out); if (rc != NDTB_OK) return rc; }
  (*out)->n_packets = 0;
  (*out)->lu.prev_completion_ts = -1.0; (*out)->lu.end_ts = 0.0; (*out)->closing_seq = 0;
  return NDTB_OK;
}

// The running frame is complete (:658-669): decode its packets with one kernel.
int close_active(ndtb_lidar_s* l) {
  Slot* s = l->active;
  if (!s || s->n_packets == 0) return NDTB_OK;
  s->frame_id = l->frame_id_;
  ++l->epoch;
  if (l->epoch == 0) ++l->epoch;
  CKL(cudaEventRecord(s->k0, l->stream));
  lidar_decode_kernel<<<s->n_packets * l->kp.upp, 32 * l->kp.cpc, 0, l->stream>>>(l->kp, s->d_packets, s->n_packets * l->kp.upp, l->d_x1, l->d_y1, l->d_z1, l->d_x2, l->d_y2, l->d_z2,
                                                                     l->d_ctl, l->d_recs, l->epoch, l->d_chain, s->arr, s->h_info, s->lu);
  CKL(cudaGetLastError());
  CKL(cudaEventRecord(s->k1, l->stream));
  CKL(cudaEventRecord(s->done, l->stream));
  ++l->launches;
  l->closed.push_back(s);
  l->active = nullptr;
  return take_slot(l, &l->active);
}

void free_slot(Slot* s) {
  cudaFree(s->d_packets); cudaFree(s->arr.xyzw); cudaFree(s->arr.ts); cudaFree(s->arr.rel); cudaFree(s->arr.c_id); cudaFree(s->arr.m_id);
  cudaFree(s->arr.refl); cudaFree(s->arr.sig); cudaFree(s->arr.nir);
  if (s->h_info) cudaFreeHost(s->h_info);
  if (s->done) cudaEventDestroy(s->done);
  if (s->k0) cudaEventDestroy(s->k0);
  if (s->k1) cudaEventDestroy(s->k1);
  delete s;
}

// Initialize(), :191-341: the spherical -> cartesian lookup tables in the body frame.
bool build_tables(ndtb_lidar_s* l, const float* az_deg, const float* alt_deg) {
  const ndtb_lidar_config& c = l->cfg;
  const int W = c.columns_per_frame, Hh = c.pixels_per_column;
  l->x1.assign(static_cast<size_t>(W) * Hh, 0.f); l->y1 = l->x1; l->z1 = l->x1;
  l->x2.assign(W, 0.f); l->y2 = l->x2; l->z2 = l->x2;
  const double* R = c.body_to_lidar_rotation;
  const double* tr = c.body_to_lidar_translation;
  const double b2l[16] = {R[0], R[1], R[2], tr[0], R[3], R[4], R[5], tr[1], R[6], R[7], R[8], tr[2], 0, 0, 0, 1};
  double l2b[16];
  if (!invert4(b2l, l2b)) return false;
  std::vector<float> ca(Hh), sa(Hh), beam_az(Hh);
  const float pi_f = static_cast<float>(M_PI);
  for (int i = 0; i < Hh; ++i) {
    const float alt = alt_deg[i] * pi_f / 180.0f;
    sa[i] = sinf(alt);
    ca[i] = cosf(alt);
    beam_az[i] = az_deg[i] * pi_f / 180.0f;
  }
  const float beam_off = c.lidar_origin_to_beam_origin_mm * 0.001f;
  for (int m = 0; m < W; ++m) {
    const float az = 2.0f * pi_f * (1.0f - (static_cast<float>(m) / static_cast<float>(W)));
    const double o0 = static_cast<double>(beam_off * cosf(az)), o1 = static_cast<double>(beam_off * sinf(az));
    l->x2[m] = static_cast<float>(((l2b[0] * o0 + l2b[1] * o1) + l2b[2] * 0.0) + l2b[3] * 1.0);
    l->y2[m] = static_cast<float>(((l2b[4] * o0 + l2b[5] * o1) + l2b[6] * 0.0) + l2b[7] * 1.0);
    l->z2[m] = static_cast<float>(((l2b[8] * o0 + l2b[9] * o1) + l2b[10] * 0.0) + l2b[11] * 1.0);
    for (int ch = 0; ch < Hh; ++ch) {
      const float tot = az + beam_az[ch];
      const double d0 = static_cast<double>(ca[ch] * cosf(tot)), d1 = static_cast<double>(ca[ch] * sinf(tot)), d2 = static_cast<double>(sa[ch]);
      const size_t k = static_cast<size_t>(m) * Hh + ch;
      l->x1[k] = static_cast<float>((l2b[0] * d0 + l2b[1] * d1) + l2b[2] * d2);
      l->y1[k] = static_cast<float>((l2b[4] * d0 + l2b[5] * d1) + l2b[6] * d2);
      l->z1[k] = static_cast<float>((l2b[8] * d0 + l2b[9] * d1) + l2b[10] * d2);
    }
  }
  return true;
}

}  // namespace

extern "C" {

int ndtb_lidar_default_config(ndtb_lidar_config* c) {
  if (!c) return NDTB_ERR_INVALID;
  std::memset(c, 0, sizeof(*c));
  c->columns_per_frame = 2048; c->pixels_per_column = 128; c->columns_per_packet = 16; c->channel_stride = 1;
  c->reflectivity_threshold = 0;
  c->range_min = 1.0f; c->range_max = 200.0f; c->z_min = -300.0f; c->z_max = 0.0f;
  return NDTB_OK;
}

int ndtb_lidar_create(int cuda_device, const ndtb_lidar_config* cfg, const float* az_deg, const float* alt_deg, ndtb_lidar* out) {
  if (!cfg || !az_deg || !alt_deg || !out) return NDTB_ERR_INVALID;
  *out = nullptr;
  const int N = cfg->channel_stride;
  if (N != 1 && N != 2 && N != 4 && N != 8 && N != 16) return NDTB_ERR_INVALID;  // :41-46
  if (cfg->pixels_per_column <= 0 || N > cfg->pixels_per_column) return NDTB_ERR_INVALID;  // :47-49
  if (cfg->columns_per_frame <= 0 || cfg->columns_per_frame > 65535 || cfg->columns_per_packet <= 0 || cfg->columns_per_packet > 32) return NDTB_ERR_INVALID;
  const int subset = (cfg->pixels_per_column + N - 1) / N;
  if (subset > 32 * LIDAR_MAX_ROUNDS) return NDTB_ERR_INVALID;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || cuda_device < 0 || cuda_device >= count) return NDTB_ERR_CUDA;  // no CPU fallback
  if (cudaSetDevice(cuda_device) != cudaSuccess) return NDTB_ERR_CUDA;
  ndtb_lidar_s* l = new ndtb_lidar_s();
  l->device = cuda_device;
  l->cfg = *cfg;
  if (!build_tables(l, az_deg, alt_deg)) { delete l; return NDTB_ERR_INVALID; }
  LidarKParams& k = l->kp;
  k.cpf = cfg->columns_per_frame; k.ppc = cfg->pixels_per_column; k.cpp = cfg->columns_per_packet; k.stride = N; k.subset = subset;
  k.rounds = (subset + 31) / 32;
  // columns per CTA: one CTA per packet measured faster (21.7 us) than four columns per CTA (26.0 us, 256 CTAs: longer look-back)
  k.cpc = (std::getenv("NDTB_LIDAR_CPC") && k.cpp % 4 == 0) ? 4 : k.cpp;
  k.upp = k.cpp / k.cpc;
  k.legacy = cfg->udp_profile_legacy ? 1 : 0;
  if (k.legacy) {  // :198-203: no packet header / footer, 16-byte column header, 4-byte block status; a unit is a whole packet
    k.cpc = k.cpp; k.upp = 1;
    k.block_bytes = 16 + cfg->pixels_per_column * 12 + 4;
    l->expected_size = static_cast<size_t>(k.cpp) * k.block_bytes;
  } else {
    k.block_bytes = 12 + cfg->pixels_per_column * 12;
    l->expected_size = 32 + static_cast<size_t>(k.cpp) * k.block_bytes + 32;
  }
  k.pitch = static_cast<int>((l->expected_size + 15) / 16 * 16);
  k.r_min = cfg->range_min; k.r_max = cfg->range_max; k.z_min = cfg->z_min; k.z_max = cfg->z_max;
  k.refl_thr = static_cast<int>(static_cast<unsigned char>(static_cast<float>(cfg->reflectivity_threshold)));
  for (int a = 0; a < 3; ++a) {  // :325-327
    const float half = cfg->box_dimensions[a] / 2.0f;
    k.bmin[a] = cfg->box_center[a] - half;
    k.bmax[a] = cfg->box_center[a] + half;
  }
  l->cap_packets = 2 * ((k.cpf + k.cpp - 1) / k.cpp) + 2;
  if (k.legacy && l->cap_packets > LEGACY_MAX_UNITS) { delete l; return NDTB_ERR_INVALID; }  // unit table travels as a kernel parameter
  l->cap_points = static_cast<unsigned>(l->cap_packets) * k.cpp * subset;
  auto fail = [&](int code) { ndtb_lidar_destroy(l); return code; };
  if (cudaStreamCreateWithFlags(&l->stream, cudaStreamNonBlocking) != cudaSuccess) return fail(NDTB_ERR_CUDA);
  const size_t nt = l->x1.size() * 4, nc = l->x2.size() * 4;
  if (cudaMalloc(&l->d_x1, nt) || cudaMalloc(&l->d_y1, nt) || cudaMalloc(&l->d_z1, nt) || cudaMalloc(&l->d_x2, nc) || cudaMalloc(&l->d_y2, nc) ||
      cudaMalloc(&l->d_z2, nc) || cudaMalloc(&l->d_ctl, sizeof(Ctl)) || cudaMalloc(&l->d_recs, sizeof(PacketRec) * l->cap_packets * l->kp.upp) ||
      cudaMalloc(&l->d_chain, sizeof(double)))
    return fail(NDTB_ERR_CUDA);
  cudaMemcpyAsync(l->d_x1, l->x1.data(), nt, cudaMemcpyHostToDevice, l->stream);
  cudaMemcpyAsync(l->d_y1, l->y1.data(), nt, cudaMemcpyHostToDevice, l->stream);
  cudaMemcpyAsync(l->d_z1, l->z1.data(), nt, cudaMemcpyHostToDevice, l->stream);
  cudaMemcpyAsync(l->d_x2, l->x2.data(), nc, cudaMemcpyHostToDevice, l->stream);
  cudaMemcpyAsync(l->d_y2, l->y2.data(), nc, cudaMemcpyHostToDevice, l->stream);
  cudaMemcpyAsync(l->d_z2, l->z2.data(), nc, cudaMemcpyHostToDevice, l->stream);
  cudaMemsetAsync(l->d_ctl, 0, sizeof(Ctl), l->stream);
  cudaMemsetAsync(l->d_recs, 0, sizeof(PacketRec) * l->cap_packets * l->kp.upp, l->stream);
  cudaMemsetAsync(l->d_chain, 0, sizeof(double), l->stream);
  if (cudaStreamSynchronize(l->stream) != cudaSuccess) return fail(NDTB_ERR_CUDA);
  if (take_slot(l, &l->active) != NDTB_OK) return fail(NDTB_ERR_CUDA);
  *out = l;
  return NDTB_OK;
}

int ndtb_lidar_destroy(ndtb_lidar l) {
  if (!l) return NDTB_ERR_INVALID;
  cudaSetDevice(l->device);
  if (l->stream) cudaStreamSynchronize(l->stream);
  for (Slot* s : l->slots) free_slot(s);
  cudaFree(l->d_x1); cudaFree(l->d_y1); cudaFree(l->d_z1); cudaFree(l->d_x2); cudaFree(l->d_y2); cudaFree(l->d_z2);
  cudaFree(l->d_ctl); cudaFree(l->d_recs); cudaFree(l->d_chain);
  if (l->stream) cudaStreamDestroy(l->stream);
  delete l;
  return NDTB_OK;
}

const char* ndtb_lidar_last_error(ndtb_lidar l) { return l ? l->last_error.c_str() : "null handle"; }

int ndtb_lidar_packet_bytes(ndtb_lidar l, size_t* expected_size) {
  if (!l || !expected_size) return NDTB_ERR_INVALID;
  *expected_size = l->expected_size;
  return NDTB_OK;
}

int ndtb_lidar_get_tables(ndtb_lidar l, float* x1, float* y1, float* z1, float* x2, float* y2, float* z2) {
  if (!l) return NDTB_ERR_INVALID;
  if (x1) std::memcpy(x1, l->x1.data(), l->x1.size() * 4);
  if (y1) std::memcpy(y1, l->y1.data(), l->y1.size() * 4);
  if (z1) std::memcpy(z1, l->z1.data(), l->z1.size() * 4);
  if (x2) std::memcpy(x2, l->x2.data(), l->x2.size() * 4);
  if (y2) std::memcpy(y2, l->y2.data(), l->y2.size() * 4);
  if (z2) std::memcpy(z2, l->z2.data(), l->z2.size() * 4);
  return NDTB_OK;
}

int ndtb_lidar_push_packets(ndtb_lidar l, const uint8_t* packets, size_t n_packets, size_t packet_bytes, size_t pitch_bytes, int* frames_closed) {
  if (!l) return NDTB_ERR_INVALID;
  if (frames_closed) *frames_closed = 0;
  if (n_packets == 0) return NDTB_OK;
  if (!packets || pitch_bytes < packet_bytes) return lidar_fail(l, NDTB_ERR_INVALID, "bad packet buffer");
  if (packet_bytes != l->expected_size) return NDTB_OK;  // "Invalid packet size": every packet is ignored (:633-636)
  cudaSetDevice(l->device);
  int closed = 0, rc = NDTB_OK;
  size_t run_begin = 0, run_len = 0;  // consecutive accepted packets of the running frame, copied with one DMA
  auto flush_run = [&]() -> int {
    if (run_len == 0) return NDTB_OK;
    Slot* s = l->active;
    size_t take = run_len;
    int over = NDTB_OK;
    if (s->n_packets + take > static_cast<size_t>(l->cap_packets)) {
      take = static_cast<size_t>(l->cap_packets - s->n_packets);
      over = lidar_fail(l, NDTB_ERR_OVERFLOW, "more packets in one frame than two revolutions hold: packets dropped");
    }
    if (take > 0) {
      unsigned char* dst = s->d_packets + static_cast<size_t>(s->n_packets) * l->kp.pitch;
      const uint8_t* src = packets + run_begin * pitch_bytes;
      if (pitch_bytes == packet_bytes && static_cast<size_t>(l->kp.pitch) == packet_bytes) {
        CKL(cudaMemcpyAsync(dst, src, take * packet_bytes, cudaMemcpyHostToDevice, l->stream));
      } else {
        CKL(cudaMemcpy2DAsync(dst, l->kp.pitch, src, pitch_bytes, packet_bytes, take, cudaMemcpyHostToDevice, l->stream));
      }
      s->n_packets += static_cast<int>(take);
    }
    run_len = 0;
    return over;
  };
  if (l->kp.legacy) {
    // DecodePacketLegacy (:376-629): the frame id travels in every column header.  The host reads the 12 header bytes of every
    // column, cuts the stream at the column whose id differs, and hands every frame its packets with the column range that
    // belongs to it (a packet that straddles two frames is copied into both).
    const int cpp = l->kp.cpp, cpf = l->kp.cpf, bb = l->kp.block_bytes;
    auto add_unit = [&](const uint8_t* p, int cb, int ce, double las) -> int {
      Slot* s = l->active;
      if (s->n_packets >= l->cap_packets || s->n_packets >= LEGACY_MAX_UNITS)
        return lidar_fail(l, NDTB_ERR_OVERFLOW, "more packets in one frame than two revolutions hold: packets dropped");
      CKL(cudaMemcpyAsync(s->d_packets + static_cast<size_t>(s->n_packets) * l->kp.pitch, p, packet_bytes, cudaMemcpyHostToDevice, l->stream));
      s->lu.cb[s->n_packets] = static_cast<unsigned char>(cb);
      s->lu.ce[s->n_packets] = static_cast<unsigned char>(ce);
      s->lu.latest_at_start[s->n_packets] = las;
      s->n_packets += 1;
      return NDTB_OK;
    };
    for (size_t i = 0; i < n_packets; ++i) {
      const uint8_t* p = packets + i * pitch_bytes;
      const double las = l->latest_ts > 0.0 ? l->latest_ts : 0.0;   // :385-388
      int seg_begin = 0;
      for (int col = 0; col < cpp; ++col) {
        const uint8_t* hp = p + static_cast<size_t>(col) * bb;
        unsigned long long ts_ns = 0;
        for (int b = 7; b >= 0; --b) ts_ns = (ts_ns << 8) | hp[b];
        const unsigned m_id = hp[8] | (hp[9] << 8);
        if (m_id >= static_cast<unsigned>(cpf)) continue;            // :407-410: the column leaves no trace
        const unsigned short fid = static_cast<unsigned short>(hp[10] | (hp[11] << 8));
        if (fid != l->frame_id_) {                                    // :420-431
          if (col > seg_begin) { const int r = add_unit(p, seg_begin, col, las); if (r != NDTB_OK) rc = r; }
          double completed_at = -1.0;
          if (l->active->n_packets > 0) {
            l->active->lu.end_ts = l->latest_ts;
            l->active->closing_seq = l->packet_seq;
            completed_at = l->latest_ts;
            const int c = close_active(l);
            if (c != NDTB_OK) return c;
            ++closed;
          }
          l->active->lu.prev_completion_ts = completed_at;
          l->frame_id_ = fid;
          seg_begin = col;
        }
        l->latest_ts = fmod(static_cast<double>(ts_ns) * 1e-9, 86400.0);   // :396, :433
      }
      if (cpp > seg_begin) { const int r = add_unit(p, seg_begin, cpp, las); if (r != NDTB_OK) rc = r; }
      ++l->packet_seq;
    }
    if (frames_closed) *frames_closed = closed;
    return rc;
  }
  for (size_t i = 0; i < n_packets; ++i) {
    const uint8_t* p = packets + i * pitch_bytes;
    const unsigned packet_type = p[0] | (p[1] << 8);
    if (packet_type != 0x0001u) {  // :640-646
      const int r = flush_run();
      if (r != NDTB_OK) rc = r;
      continue;
    }
    const unsigned short fid = static_cast<unsigned short>(p[2] | (p[3] << 8));
    if (fid != l->frame_id_) {  // :658-669
      const int r = flush_run();
      if (r != NDTB_OK) rc = r;
      if (l->active->n_packets > 0) {
        const int c = close_active(l);
        if (c != NDTB_OK) return c;
        ++closed;
      }
      l->frame_id_ = fid;
    }
    if (run_len == 0) run_begin = i;
    ++run_len;
  }
  const int r = flush_run();
  if (r != NDTB_OK) rc = r;
  if (frames_closed) *frames_closed = closed;
  return rc;
}

int ndtb_lidar_pop_frame(ndtb_lidar l, ndtb_lidar_frame_info* info, int* available) {
  if (!l || !available) return NDTB_ERR_INVALID;
  *available = 0;
  cudaSetDevice(l->device);
  if (l->current) { l->free_slots.push_back(l->current); l->current = nullptr; }  // ReturnFrameToPool
  while (!l->closed.empty()) {
    Slot* s = l->closed.front();
    l->closed.pop_front();
    CKL(cudaEventSynchronize(s->done));
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, s->k0, s->k1) == cudaSuccess) l->last_decode_us = ms * 1000.f;
    if (s->h_info->numberpoints == 0) { l->free_slots.push_back(s); continue; }  // never emitted by the reference (:659)
    if (l->kp.legacy) {
      // one DecodePacketLegacy call returns ONE frame: when the same packet completed a later frame with points as well, that
      // one overwrote `completed_frame` (:426) and this one was never seen by the caller
      bool overwritten = false;
      for (Slot* later : l->closed) {
        if (later->closing_seq != s->closing_seq) break;
        CKL(cudaEventSynchronize(later->done));
        if (later->h_info->numberpoints > 0) { overwritten = true; break; }
      }
      if (overwritten) { l->free_slots.push_back(s); continue; }
    }
    s->h_info->frame_id = s->frame_id;
    s->h_info->packets = static_cast<uint16_t>(s->n_packets);
    l->current = s;
    if (info) *info = *s->h_info;
    *available = 1;
    return NDTB_OK;
  }
  return NDTB_OK;
}

int ndtb_lidar_read_frame(ndtb_lidar l, size_t capacity, float* x, float* y, float* z, uint16_t* c_id, uint16_t* m_id, double* ts, float* rel,
                          uint16_t* refl, uint16_t* sig, uint16_t* nir) {
  if (!l) return NDTB_ERR_INVALID;
  if (!l->current) return lidar_fail(l, NDTB_ERR_NOT_READY, "no popped frame");
  cudaSetDevice(l->device);
  Slot* s = l->current;
  const size_t n = s->h_info->numberpoints;
  if (capacity < n) return lidar_fail(l, NDTB_ERR_INVALID, "capacity smaller than numberpoints");
  if (x || y || z) {
    std::vector<float4> tmp(n);
    CKL(cudaMemcpy(tmp.data(), s->arr.xyzw, n * sizeof(float4), cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < n; ++i) {
      if (x) x[i] = tmp[i].x;
      if (y) y[i] = tmp[i].y;
      if (z) z[i] = tmp[i].z;
    }
  }
  if (c_id) CKL(cudaMemcpy(c_id, s->arr.c_id, n * 2, cudaMemcpyDeviceToHost));
  if (m_id) CKL(cudaMemcpy(m_id, s->arr.m_id, n * 2, cudaMemcpyDeviceToHost));
  if (ts) CKL(cudaMemcpy(ts, s->arr.ts, n * 8, cudaMemcpyDeviceToHost));
  if (rel) CKL(cudaMemcpy(rel, s->arr.rel, n * 4, cudaMemcpyDeviceToHost));
  if (refl) CKL(cudaMemcpy(refl, s->arr.refl, n * 2, cudaMemcpyDeviceToHost));
  if (sig) CKL(cudaMemcpy(sig, s->arr.sig, n * 2, cudaMemcpyDeviceToHost));
  if (nir) CKL(cudaMemcpy(nir, s->arr.nir, n * 2, cudaMemcpyDeviceToHost));
  return NDTB_OK;
}

int ndtb_lidar_frame_points(ndtb_lidar l, const void** d_xyzw, size_t* n_points) {
  if (!l || !d_xyzw || !n_points) return NDTB_ERR_INVALID;
  if (!l->current) return lidar_fail(l, NDTB_ERR_NOT_READY, "no popped frame");
  *d_xyzw = l->current->arr.xyzw;
  *n_points = l->current->h_info->numberpoints;
  return NDTB_OK;
}

int ndtb_lidar_sync(ndtb_lidar l) {
  if (!l) return NDTB_ERR_INVALID;
  cudaSetDevice(l->device);
  CKL(cudaStreamSynchronize(l->stream));
  return NDTB_OK;
}

int ndtb_lidar_get_stats(ndtb_lidar l, uint64_t* launches, float* last_decode_us) {
  if (!l) return NDTB_ERR_INVALID;
  if (launches) *launches = l->launches;
  if (last_decode_us) *last_decode_us = l->last_decode_us;
  return NDTB_OK;
}

}  // extern "C"
