// Host-side interface of the voxel-map build (implemented in voxel_build.cu / voxel_finalize.cu).
#pragma once
#include "ndtb_internal.cuh"
#include <stddef.h>

namespace ndtb {

// Everything the build decides ON THE DEVICE (no host round trip inside a build): the grid geometry from the bounding box, the
// radix-sort plan from the key range, the leaf count, the hash capacity.  Lives in device memory (DeviceGrid::d_meta), read by
// every build kernel, published to the host once, by the last kernel of the build.
constexpr int HEAVY_LEAF_POINTS = 48;  // leaves with more points are summed by a whole warp, the others by one lane (voxel_stats_kernel)
constexpr int RADIX_MAX_BITS = 10;      // widest digit: 1024 bins (20 key bits -> 2 passes of 10 bits)
struct BuildMeta {
  GridMeta m;
  int error;                  // 0 | NDTB_ERR_OVERFLOW | NDTB_ERR_CUDA (a grid barrier of the sort kernel timed out)
  int bits, digit_bits, passes;
  int final_buf;              // pairs[final_buf] holds the sorted (key << 32 | point index) list: (passes - 1) & 1
  unsigned int n_skipped;     // non-finite points (never assigned to a leaf)
  unsigned int V;             // leaves (pclomp keeps under-populated / rejected leaves, flagged)
  unsigned int n_valid;       // leaves a neighbour search can return
  unsigned int chunk_counter; // dynamic work hand-out of the statistics kernel
  unsigned int n_heavy;       // leaves with more than HEAVY_LEAF_POINTS points (summed warp-wide at the end of the sort kernel)
  unsigned int tickets[3];    // last-block-done tickets: -, statistics, centroid
  unsigned int n_flagged;     // pclomp leaves flagged nr_points_ = -1 that stay reachable by radius searches
};
static_assert(sizeof(BuildMeta) % 8 == 0, "BuildMeta is published to the host in 8-byte words");
constexpr int BUILD_META_WORDS = static_cast<int>(sizeof(BuildMeta) / 8);

struct BuildParams {  // what the host knows before the build starts
  float resolution, inv_leaf;
  int min_pts, flavour, is_dense;
  double eig_mult;
  int n;
  int reserved;
};

struct DeviceGrid {
  GridMeta meta{};         // host copy, valid after a successful build
  bool built = false;
  uint32_t V = 0;          // leaves
  uint32_t n_valid = 0;    // leaves that can be returned by a neighbour search
  uint32_t n_skipped = 0;  // non-finite points of the last build
  uint32_t n_flagged = 0;  // leaves flagged -1 (reachable by the radius searches only)
  int final_buf = 0;       // pairs[final_buf]: sorted (key << 32 | point index) of the last build
  uint64_t launches = 0;   // kernels launched by builds on this grid
  int last_build_launches = 0;
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
  unsigned long long* pairs[2] = {nullptr, nullptr};   // (voxel key << 32 | point index), ping-pong of the radix sort
  uint32_t *hist = nullptr, *seg_start = nullptr;
  double* vox_sums = nullptr;               // [V][9] Sum x, Sum x x^T of every leaf (sums kernel -> finalize kernel)
  float4* sorted_pts = nullptr;             // the target's points in sorted order (w = point index), written by the final radix pass
  uint32_t* tile_heads = nullptr;           // segment heads per tile (sort kernel)
  size_t pts_alloc = 0;
  int* d_bbox = nullptr;            // [0..5] ordered ints of the box, [8..9] barrier / exit counters of the sort kernel
  uint32_t* digit_base = nullptr;   // digit totals of the current radix pass (1024)
  BuildMeta* d_meta = nullptr;
  volatile unsigned long long* h_slots = nullptr;  // pinned self-validating slots: BUILD_META_WORDS words of BuildMeta
  unsigned long long slot_seq = 0;
};

// Returns NDTB_OK / NDTB_ERR_*.  The three kernels of a build run back to back; the host waits once, for the published BuildMeta.
int build_voxel_map(DeviceGrid& g, const float4* d_pts, size_t n, float resolution, int min_pts, double eig_mult, int flavour,
                    int is_dense, cudaStream_t stream);
void free_device_grid(DeviceGrid& g);
int voxelgrid_downsample(DeviceGrid& scratch, const float4* d_pts, size_t n, float leaf, int is_dense, float4* d_out, size_t* n_out,
                         cudaStream_t stream);

// voxel_finalize.cu: per-leaf statistics: sums in input order over DeviceGrid::sorted_pts (heavy leaves arrive summed by the sort
// kernel), mean / covariance / eigen-inflation / inverse / hot record / hash insert; the last block publishes BuildMeta to the host slots.
void launch_voxel_finalize(const DeviceGrid& g, uint32_t n_points, unsigned long long publish_key, cudaStream_t stream);

}  // namespace ndtb
