// ORACLE — TEST INFRASTRUCTURE ONLY (see small_linalg.hpp header).
//
// CPU restatement of the reference's two NDT target voxel maps:
//   flavour PCLOMP: extern/ndt_omp/include/pclomp/voxel_grid_covariance_omp_impl.hpp:55-439
//                   (container std::map<size_t,Leaf>, voxel_grid_covariance_omp.h:187)
//   flavour SVN   : extern/svn_ndt/include/voxel_grid_covariance_impl.hpp:46-71,77-379,560-615
//                   and voxel_grid_covariance.h:262-304 (container tsl::robin_map, .h:144-145)
// PCL pieces restated from their published behaviour (PCL >= 1.14, not in tree):
//   pcl::getMinMax3D (float min/max over finite points), pcl::VoxelGrid::setLeafSize
//   (inverse_leaf_size_ = 1.0f / leaf_size_ in float).
#pragma once
#include <cstdint>
#include <cstddef>
#include <algorithm>
#include <map>
#include <unordered_map>
#include <utility>
#include <vector>
#include <cmath>
#include <limits>
#include "small_linalg.hpp"

namespace oracle {

enum Flavour { FLAVOUR_PCLOMP = 0, FLAVOUR_SVN = 1 };

struct Leaf {
  int nr_points = 0;
  double mean[3] = {0, 0, 0};
  // Both reference Leaf constructors start cov_ at IDENTITY (voxel_grid_covariance_omp.h:113,
  // svn voxel_grid_covariance.h:109) and applyFilter accumulates x x^T on top of it
  // (voxel_grid_covariance_omp_impl.hpp:182,239; voxel_grid_covariance_impl.hpp:206,240): the
  // single-pass second moment is I + Sum x x^T, i.e. every covariance carries an extra I/n term.
  double cov[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  double icov[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  double evecs[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  double evals[3] = {0, 0, 0};
  // Float running sum -> centroid (Leaf::centroid_, voxel_grid_covariance_omp_impl.hpp:244-245,285;
  // multi_voxel_grid_covariance_omp_impl.hpp:404-405,358): what the multigrid kd-tree is built over.
  float centroid[3] = {0, 0, 0};
};

// Open-addressing Robin-Hood map restating the keying the svn grid inherits from tsl::robin_map
// (extern/robin-map/include/tsl/robin_map.h:86-90, robin_growth_policy.h:124-126,
// robin_hash.h:1515): std::hash<size_t> = identity, bucket = hash & (2^m - 1), max load 0.5,
// linear probing with Robin-Hood displacement, backward-shift erase.
class RobinMap {
 public:
  RobinMap() { rehash(16); }
  void clear() { slots_.clear(); count_ = 0; rehash(16); }
  size_t size() const { return count_; }
  Leaf& operator[](size_t key) {
    for (size_t i = key & mask_, d = 0;; i = (i + 1) & mask_, ++d) {
      if (slots_[i].dist < 0 || (size_t)slots_[i].dist < d) break;
      if (slots_[i].key == key) return slots_[i].leaf;
    }
    if ((count_ + 1) * 2 > slots_.size()) rehash(slots_.size() * 2);
    Slot cur; cur.key = key; cur.dist = 0; cur.leaf = Leaf();
    Leaf* ret = nullptr;
    for (size_t i = key & mask_;; i = (i + 1) & mask_) {
      if (slots_[i].dist < 0) { slots_[i] = cur; if (!ret) ret = &slots_[i].leaf; break; }
      if (slots_[i].dist < cur.dist) { std::swap(slots_[i], cur); if (!ret) ret = &slots_[i].leaf; }
      ++cur.dist;
    }
    ++count_;
    return *ret;
  }
  const Leaf* find(size_t key) const {
    for (size_t i = key & mask_, d = 0;; i = (i + 1) & mask_, ++d) {
      if (slots_[i].dist < 0 || (size_t)slots_[i].dist < d) return nullptr;
      if (slots_[i].key == key) return &slots_[i].leaf;
    }
  }
  void erase(size_t key) {
    size_t i = key & mask_;
    for (size_t d = 0;; i = (i + 1) & mask_, ++d) {
      if (slots_[i].dist < 0 || (size_t)slots_[i].dist < d) return;
      if (slots_[i].key == key) break;
    }
    for (;;) {
      size_t nx = (i + 1) & mask_;
      if (slots_[nx].dist <= 0) { slots_[i].dist = -1; break; }
      slots_[i] = slots_[nx]; --slots_[i].dist; i = nx;
    }
    --count_;
  }
  template <class F> void for_each(F f) const { for (auto& s : slots_) if (s.dist >= 0) f(s.key, s.leaf); }
  std::vector<size_t> keys() const { std::vector<size_t> k; k.reserve(count_); for (auto& s : slots_) if (s.dist >= 0) k.push_back(s.key); return k; }
 private:
  struct Slot { size_t key = 0; int dist = -1; Leaf leaf; };
  void rehash(size_t n) {
    std::vector<Slot> old; old.swap(slots_);
    slots_.assign(n, Slot()); mask_ = n - 1; size_t c = 0;
    for (auto& s : old) if (s.dist >= 0) {
      Slot cur = s; cur.dist = 0;
      for (size_t i = cur.key & mask_;; i = (i + 1) & mask_) {
        if (slots_[i].dist < 0) { slots_[i] = cur; break; }
        if (slots_[i].dist < cur.dist) std::swap(slots_[i], cur);
        ++cur.dist;
      }
      ++c;
    }
    count_ = c;
  }
  std::vector<Slot> slots_;
  size_t mask_ = 0, count_ = 0;
};

struct VoxelGrid {
  int flavour = FLAVOUR_PCLOMP;
  float leaf_size = 1.0f;           // cubic leaves (ndt_omp.h:324-326, svn_ndt_impl.hpp:148)
  float inverse_leaf_size = 1.0f;   // pcl::VoxelGrid::setLeafSize: 1.0f / leaf_size
  int min_points_per_voxel = 6;     // voxel_grid_covariance_omp.h:195 / voxel_grid_covariance.h:153
  double min_covar_eigvalue_mult = 0.01;
  int min_b[3] = {0, 0, 0}, max_b[3] = {0, 0, 0}, div_b[3] = {0, 0, 0}, divb_mul[3] = {0, 0, 0};
  bool built = false;
  std::map<size_t, Leaf> leaves_map;   // PCLOMP
  RobinMap leaves_robin;               // SVN

  void set_leaf_size(float r) { leaf_size = r; inverse_leaf_size = 1.0f / r; }

  // Returns 0 ok, -1 empty input, -2 index overflow (reference warns and leaves the grid empty).
  int build(const float* xyz, size_t n, size_t stride_floats, bool is_dense);

  const Leaf* find_index(size_t idx) const {
    if (flavour == FLAVOUR_PCLOMP) {
      auto it = leaves_map.find(idx);
      if (it != leaves_map.end() && it->second.nr_points >= min_points_per_voxel) return &it->second;
      return nullptr;
    }
    const Leaf* l = leaves_robin.find(idx);
    if (l && l->nr_points >= min_points_per_voxel) return l;
    return nullptr;
  }

  // svn: voxel_grid_covariance_impl.hpp:46-71 + voxel_grid_covariance.h:280-304
  const Leaf* svn_get_leaf(float px, float py, float pz) const {
    if (!built) return nullptr;
    const float p[3] = {px, py, pz};
    for (int a = 0; a < 3; ++a) {
      const float lo = static_cast<float>(min_b[a]) * leaf_size;
      const float hi = static_cast<float>(max_b[a] + 1) * leaf_size;
      if (!(p[a] >= lo && p[a] < hi)) return nullptr;
    }
    const int i0 = static_cast<int>(std::floor(p[0] * inverse_leaf_size) - min_b[0]);
    const int i1 = static_cast<int>(std::floor(p[1] * inverse_leaf_size) - min_b[1]);
    const int i2 = static_cast<int>(std::floor(p[2] * inverse_leaf_size) - min_b[2]);
    const size_t idx = static_cast<size_t>(i0 * divb_mul[0] + i1 * divb_mul[1] + i2 * divb_mul[2]);
    return find_index(idx);
  }

  // Neighbour search, n_probe = 7 (DIRECT7) or 1 (DIRECT1).  Returns number of leaves found,
  // in the reference's probe order (centre,+x,-x,+y,-y,+z,-z).
  //   pclomp: voxel_grid_covariance_omp_impl.hpp:368-399, 415-439
  //   svn   : voxel_grid_covariance_impl.hpp:560-615
  int neighbours(float px, float py, float pz, int n_probe, const Leaf** out) const {
    static const int off[7][3] = {{0, 0, 0}, {1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}};
    int found = 0;
    if (flavour == FLAVOUR_PCLOMP) {
      if (!built) return 0;
      const int ijk[3] = {static_cast<int>(std::floor(px / leaf_size)),
                          static_cast<int>(std::floor(py / leaf_size)),
                          static_cast<int>(std::floor(pz / leaf_size))};
      for (int ni = 0; ni < n_probe; ++ni) {
        bool in = true;
        for (int a = 0; a < 3; ++a) {
          if (!(min_b[a] - ijk[a] <= off[ni][a])) in = false;
          if (!(max_b[a] - ijk[a] >= off[ni][a])) in = false;
        }
        if (!in) continue;
        const int lin = (ijk[0] + off[ni][0] - min_b[0]) * divb_mul[0] +
                        (ijk[1] + off[ni][1] - min_b[1]) * divb_mul[1] +
                        (ijk[2] + off[ni][2] - min_b[2]) * divb_mul[2];
        const Leaf* l = find_index(static_cast<size_t>(lin));
        if (l) out[found++] = l;
      }
      return found;
    }
    const Leaf* c = svn_get_leaf(px, py, pz);
    if (c) out[found++] = c;
    if (n_probe == 1) return found;
    const float w = leaf_size;
    if (w <= 0) return found;
    const Leaf* l;
    l = svn_get_leaf(px + w, py, pz); if (l) out[found++] = l;
    l = svn_get_leaf(px - w, py, pz); if (l) out[found++] = l;
    l = svn_get_leaf(px, py + w, pz); if (l) out[found++] = l;
    l = svn_get_leaf(px, py - w, pz); if (l) out[found++] = l;
    l = svn_get_leaf(px, py, pz + w); if (l) out[found++] = l;
    l = svn_get_leaf(px, py, pz - w); if (l) out[found++] = l;
    return found;
  }

  // DIRECT26 (pclomp): voxel_grid_covariance_omp_impl.hpp:366-411 with pcl::getAllNeighborCellIndices(), which is
  // the 13 "half" displacements followed by their negatives -- 26 cells, the CENTRE CELL IS NOT AMONG THEM
  // (PCL voxel_grid.h getHalfNeighborCellIndices / getAllNeighborCellIndices, restated from the published source).
  int neighbours26(float px, float py, float pz, const Leaf** out) const {
    if (!built) return 0;
    int half[13][3]; int n = 0;
    for (int i = -1; i < 2; ++i) for (int j = -1; j < 2; ++j) { half[n][0] = i; half[n][1] = j; half[n][2] = -1; ++n; }
    for (int i = -1; i < 2; ++i) { half[n][0] = i; half[n][1] = -1; half[n][2] = 0; ++n; }
    half[n][0] = -1; half[n][1] = 0; half[n][2] = 0; ++n;
    const int ijk[3] = {static_cast<int>(std::floor(px / leaf_size)), static_cast<int>(std::floor(py / leaf_size)),
                        static_cast<int>(std::floor(pz / leaf_size))};
    int found = 0;
    for (int ni = 0; ni < 26; ++ni) {
      int off[3];
      for (int a = 0; a < 3; ++a) off[a] = ni < 13 ? half[ni][a] : -half[ni - 13][a];
      bool in = true;
      for (int a = 0; a < 3; ++a) if (!(min_b[a] - ijk[a] <= off[a]) || !(max_b[a] - ijk[a] >= off[a])) in = false;
      if (!in) continue;
      const int lin = (ijk[0] + off[0] - min_b[0]) * divb_mul[0] + (ijk[1] + off[1] - min_b[1]) * divb_mul[1] +
                      (ijk[2] + off[2] - min_b[2]) * divb_mul[2];
      const Leaf* l = find_index(static_cast<size_t>(lin));
      if (l) out[found++] = l;
    }
    return found;
  }

  // Radius search over the voxel-centroid cloud (pclomp only), restating
  //   voxel_grid_covariance_omp.h:458-485 (radiusSearch -> pcl::KdTreeFLANN::radiusSearch, sorted results) and
  //   multi_voxel_grid_covariance_omp_impl.hpp:236-258.
  // PCL/FLANN pieces restated from their published behaviour: KdTreeFLANN passes float(radius*radius) to an exact
  // (eps = 0) FLANN KDTreeSingleIndex search with L2_Simple<float> (dist = dx*dx + dy*dy + dz*dz accumulated in float
  // in that order) and RadiusResultSet keeps a point iff dist < radius^2 (strict); results sorted by distance.
  // The kd-tree is restated as an exhaustive test over every cell whose centroid can lie inside the ball.
  // Searchable leaves: nr_points >= min_points at filter time, i.e. also the ones flagged -1 afterwards
  // (voxel_grid_covariance_omp_impl.hpp:294-322, multi_voxel_grid_covariance_omp_impl.hpp:346-359).
  //   mg_centroid = false: the cloud point is float(leaf.mean_) (pclomp, _omp_impl.hpp:301-303)
  //   mg_centroid = true : the cloud point is leaf.centroid_ (multigrid, multi_..._impl.hpp:221-225)
  void prepare_radius_search() {
    fast_.clear();
    if (flavour == FLAVOUR_PCLOMP) {
      for (auto& kv : leaves_map)
        if (kv.second.nr_points >= min_points_per_voxel || kv.second.nr_points == -1) fast_[kv.first] = &kv.second;
    } else {
      // svn: buildCentroidCloudAndIndexMap (voxel_grid_covariance_impl.hpp:383-436) takes every leaf left in the map (invalid ones
      // were erased), the cloud point is float(leaf.mean_); radiusSearch maps hits back through getLeaf (:505-555)
      leaves_robin.for_each([&](size_t k, const Leaf& l) { if (l.nr_points >= min_points_per_voxel) fast_[k] = &l; });
    }
  }
  void radius_search(const float* q, double radius, bool mg_centroid, std::vector<std::pair<float, const Leaf*>>& out) const {
    if (!built) return;
    const float r2 = static_cast<float>(radius * radius);
    const int reach = static_cast<int>(std::ceil(radius / static_cast<double>(leaf_size))) + 1;
    int c[3];
    for (int a = 0; a < 3; ++a) c[a] = static_cast<int>(std::floor(static_cast<double>(q[a]) / static_cast<double>(leaf_size)));
    for (int dz = -reach; dz <= reach; ++dz) {
      const int k = c[2] + dz; if (k < min_b[2] || k > max_b[2]) continue;
      for (int dy = -reach; dy <= reach; ++dy) {
        const int j = c[1] + dy; if (j < min_b[1] || j > max_b[1]) continue;
        for (int dx = -reach; dx <= reach; ++dx) {
          const int i = c[0] + dx; if (i < min_b[0] || i > max_b[0]) continue;
          const size_t lin = static_cast<size_t>((i - min_b[0]) * divb_mul[0] + (j - min_b[1]) * divb_mul[1] + (k - min_b[2]) * divb_mul[2]);
          auto it = fast_.find(lin);
          if (it == fast_.end()) continue;
          const Leaf* l = it->second;
          float pc[3];
          for (int a = 0; a < 3; ++a) pc[a] = mg_centroid ? l->centroid[a] : static_cast<float>(l->mean[a]);
          float d = 0.0f;
          for (int a = 0; a < 3; ++a) { const float diff = q[a] - pc[a]; d += diff * diff; }
          if (d < r2) out.emplace_back(d, l);
        }
      }
    }
  }
  std::unordered_map<size_t, const Leaf*> fast_;

  // Export every leaf that survives (pclomp: all leaves incl. flagged ones; svn: valid ones),
  // sorted by ascending linear index.
  void export_leaves(std::vector<size_t>& idx, std::vector<Leaf>& out) const {
    idx.clear(); out.clear();
    if (flavour == FLAVOUR_PCLOMP) {
      for (auto& kv : leaves_map) { idx.push_back(kv.first); out.push_back(kv.second); }
    } else {
      std::map<size_t, Leaf> tmp;
      leaves_robin.for_each([&](size_t k, const Leaf& l) { tmp[k] = l; });
      for (auto& kv : tmp) { idx.push_back(kv.first); out.push_back(kv.second); }
    }
  }
};

inline void finalize_leaf_pclomp(Leaf& leaf, int min_pts, double eig_mult) {
  // voxel_grid_covariance_omp_impl.hpp:286-360
  const double n = static_cast<double>(leaf.nr_points);
  for (int a = 0; a < 3; ++a) leaf.centroid[a] /= static_cast<float>(leaf.nr_points);
  double pt_sum[3] = {leaf.mean[0], leaf.mean[1], leaf.mean[2]};
  for (int a = 0; a < 3; ++a) leaf.mean[a] = leaf.mean[a] / n;
  if (leaf.nr_points < min_pts) return;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j)
      leaf.cov[3 * i + j] = (leaf.cov[3 * i + j] - 2.0 * (pt_sum[i] * leaf.mean[j])) / n + leaf.mean[i] * leaf.mean[j];
  const double f = (n - 1.0) / n;
  for (int i = 0; i < 9; ++i) leaf.cov[i] = leaf.cov[i] * f;
  double ev[3];
  eig3_sym(leaf.cov, ev, leaf.evecs);
  if (ev[0] < 0 || ev[1] < 0 || ev[2] <= 0) { leaf.nr_points = -1; return; }
  const double min_ev = eig_mult * ev[2];
  if (ev[0] < min_ev) {
    ev[0] = min_ev;
    if (ev[1] < min_ev) ev[1] = min_ev;
    // cov = evecs * diag(ev) * evecs.inverse()
    double vd[9], vinv[9];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) vd[3 * i + j] = leaf.evecs[3 * i + j] * ev[j];
    mat3_inverse(leaf.evecs, vinv);
    mat3_mul(vd, vinv, leaf.cov);
  }
  for (int a = 0; a < 3; ++a) leaf.evals[a] = ev[a];
  mat3_inverse(leaf.cov, leaf.icov);
  double mx = leaf.icov[0], mn = leaf.icov[0];
  for (int i = 1; i < 9; ++i) { mx = std::max(mx, leaf.icov[i]); mn = std::min(mn, leaf.icov[i]); }
  if (mx == std::numeric_limits<float>::infinity() || mn == -std::numeric_limits<float>::infinity()) leaf.nr_points = -1;
}

// returns false if the leaf must be erased (voxel_grid_covariance_impl.hpp:270-343)
inline bool finalize_leaf_svn(Leaf& leaf, int min_pts, double eig_mult) {
  if (leaf.nr_points < min_pts) return false;
  const double n = static_cast<double>(leaf.nr_points);
  for (int a = 0; a < 3; ++a) leaf.mean[a] = leaf.mean[a] / n;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) leaf.cov[3 * i + j] = (leaf.cov[3 * i + j] / n) - (leaf.mean[i] * leaf.mean[j]);
  if (n > 1) { const double f = n / (n - 1.0); for (int i = 0; i < 9; ++i) leaf.cov[i] = leaf.cov[i] * f; }
  else { for (int i = 0; i < 9; ++i) leaf.cov[i] = (i % 4 == 0) ? 1e-9 : 0.0; }
  double ev[3];
  eig3_sym(leaf.cov, ev, leaf.evecs);
  const double thr = 1e-12;
  if (ev[0] < 0 || ev[1] < 0 || ev[2] < thr) return false;
  const double min_ok = std::max(thr, ev[2] * eig_mult);
  bool recompose = false;
  if (ev[0] < min_ok) { ev[0] = min_ok; recompose = true; }
  if (ev[1] < min_ok) { ev[1] = min_ok; recompose = true; }
  for (int a = 0; a < 3; ++a) leaf.evals[a] = ev[a];
  if (recompose) {
    double vd[9], vt[9];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { vd[3 * i + j] = leaf.evecs[3 * i + j] * ev[j]; vt[3 * i + j] = leaf.evecs[3 * j + i]; }
    mat3_mul(vd, vt, leaf.cov);
  }
  mat3_inverse(leaf.cov, leaf.icov);
  double mx = 0;
  for (int i = 0; i < 9; ++i) { if (!std::isfinite(leaf.icov[i])) return false; mx = std::max(mx, std::fabs(leaf.icov[i])); }
  if (mx > 1e12) return false;
  return true;
}

inline int VoxelGrid::build(const float* xyz, size_t n, size_t sf, bool is_dense) {
  leaves_map.clear(); leaves_robin.clear(); built = false;
  if (n == 0) return -1;
  // pcl::getMinMax3D: dense clouds take every point, non-dense skip non-finite ones.
  float mn[3] = {std::numeric_limits<float>::max(), std::numeric_limits<float>::max(), std::numeric_limits<float>::max()};
  float mx[3] = {-mn[0], -mn[0], -mn[0]};
  for (size_t i = 0; i < n; ++i) {
    const float* p = xyz + i * sf;
    if (!is_dense && (!std::isfinite(p[0]) || !std::isfinite(p[1]) || !std::isfinite(p[2]))) continue;
    for (int a = 0; a < 3; ++a) { mn[a] = std::min(mn[a], p[a]); mx[a] = std::max(mx[a], p[a]); }
  }
  const float inv = inverse_leaf_size;
  const int64_t dx = static_cast<int64_t>((mx[0] - mn[0]) * inv) + 1;
  const int64_t dy = static_cast<int64_t>((mx[1] - mn[1]) * inv) + 1;
  const int64_t dz = static_cast<int64_t>((mx[2] - mn[2]) * inv) + 1;
  if (flavour == FLAVOUR_SVN && (dx < 0 || dy < 0 || dz < 0)) return -2;
  if ((dx * dy * dz) > static_cast<int64_t>(std::numeric_limits<int32_t>::max())) return -2;
  for (int a = 0; a < 3; ++a) {
    min_b[a] = static_cast<int>(std::floor(mn[a] * inv));
    max_b[a] = static_cast<int>(std::floor(mx[a] * inv));
    div_b[a] = max_b[a] - min_b[a] + 1;
  }
  divb_mul[0] = 1; divb_mul[1] = div_b[0]; divb_mul[2] = div_b[0] * div_b[1];
  built = true;
  for (size_t i = 0; i < n; ++i) {
    const float* p = xyz + i * sf;
    const bool finite = std::isfinite(p[0]) && std::isfinite(p[1]) && std::isfinite(p[2]);
    if (flavour == FLAVOUR_SVN) { if (!finite) continue; }       // voxel_grid_covariance_impl.hpp:219
    else if (!is_dense && !finite) continue;                     // voxel_grid_covariance_omp_impl.hpp:212-217
    int ijk[3];
    for (int a = 0; a < 3; ++a) {
      if (flavour == FLAVOUR_PCLOMP) ijk[a] = static_cast<int>(std::floor(p[a] * inv) - static_cast<float>(min_b[a]));
      else ijk[a] = static_cast<int>(std::floor(p[a] * inv) - min_b[a]);
    }
    const int idx = ijk[0] * divb_mul[0] + ijk[1] * divb_mul[1] + ijk[2] * divb_mul[2];
    Leaf& leaf = (flavour == FLAVOUR_PCLOMP) ? leaves_map[static_cast<size_t>(idx)] : leaves_robin[static_cast<size_t>(idx)];
    const double d[3] = {p[0], p[1], p[2]};
    for (int a = 0; a < 3; ++a) leaf.mean[a] += d[a];
    for (int a = 0; a < 3; ++a) leaf.centroid[a] += p[a];
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) leaf.cov[3 * r + c] += d[r] * d[c];
    ++leaf.nr_points;
  }
  if (flavour == FLAVOUR_PCLOMP) {
    for (auto& kv : leaves_map) finalize_leaf_pclomp(kv.second, min_points_per_voxel, min_covar_eigvalue_mult);
  } else {
    for (size_t key : leaves_robin.keys()) {
      Leaf& leaf = leaves_robin[key];
      if (!finalize_leaf_svn(leaf, min_points_per_voxel, min_covar_eigvalue_mult)) leaves_robin.erase(key);
    }
  }
  return 0;
}

}  // namespace oracle
