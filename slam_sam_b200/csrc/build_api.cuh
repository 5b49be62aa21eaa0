// Host-side interface of the voxel-map build (implemented in voxel_build.cu / voxel_finalize.cu).
#pragma once
#include "ndtb_internal.cuh"
#include <stddef.h>

namespace ndtb {

struct DeviceGrid {
  GridMeta meta{};
  bool built = false;
  uint32_t V = 0;          // leaves (pclomp keeps under-populated / rejected leaves, flagged)
  uint32_t n_valid = 0;    // leaves that can be returned by a neighbour search
  uint64_t launches = 0;   // kernels launched by builds on this grid
  // hot
  VoxRec* recs = nullptr;
  HashSlot* table = nullptr;
  size_t table_cap = 0, table_alloc = 0;
  // cold (FP64 export path)
  uint32_t* c_key = nullptr;
  int32_t* c_n = nullptr;
  double *c_mean = nullptr, *c_cov = nullptr, *c_icov = nullptr, *c_evecs = nullptr, *c_evals = nullptr;
  size_t vox_alloc = 0;
  // scratch sized by the target cloud
  uint32_t *keys[2] = {nullptr, nullptr}, *vals[2] = {nullptr, nullptr};
  uint32_t *hist = nullptr, *seg_start = nullptr, *block_counts = nullptr;
  size_t pts_alloc = 0, hist_alloc = 0;
  int* d_bbox = nullptr;            // 6 ordered ints
  unsigned int* d_counters = nullptr;  // [0] skipped points, [1] V, [2] valid leaves
  int* h_pinned = nullptr;          // 16 ints of pinned host memory for the small read-backs
  double* sorted_xyz = nullptr;     // target points in voxel order (input order inside a voxel), widened to FP64, 3 per point
  double* vox_sums = nullptr;       // [V][9] Sum x, Sum x x^T per voxel (k4a -> k4b)
  uint32_t* digit_totals = nullptr; // 256 per-digit totals of the current radix pass
  unsigned int* d_ticket = nullptr; // last-block-done tickets of the bbox / finalize kernels
  volatile unsigned long long* h_slots = nullptr;  // pinned self-validating slots the kernels publish to (bbox 0..5, counters 8..10)
  unsigned long long slot_seq = 0;
};

// Returns NDTB_OK / NDTB_ERR_* ; synchronises the stream twice (bbox read-back, final counters).
int build_voxel_map(DeviceGrid& g, const float4* d_pts, size_t n, float resolution, int min_pts, double eig_mult, int flavour,
                    int is_dense, cudaStream_t stream);
void free_device_grid(DeviceGrid& g);
int voxelgrid_downsample(DeviceGrid& scratch, const float4* d_pts, size_t n, float leaf, int is_dense, float4* d_out, size_t* n_out,
                         cudaStream_t stream);

// voxel_finalize.cu
void launch_voxel_finalize(const DeviceGrid& g, uint32_t n_points_upper, unsigned long long publish_key, cudaStream_t stream);

}  // namespace ndtb
