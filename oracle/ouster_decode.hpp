// ORACLE — TEST INFRASTRUCTURE ONLY (parity unpinned: the reference holds no golden vectors for this path).
// CPU restatement of the reference's scan producer, the step before setInputSource:
//   LidarCallback::Initialize        src/lidarcallback.cpp:191-341   (lookup tables, subset channels, vehicle box)
//   LidarCallback::DecodePacketRng19 src/lidarcallback.cpp:631-878   (RNG19_RFL8_SIG16_NIR16 packets -> LidarFrame)
//   LidarCallback::DecodePacketLegacy src/lidarcallback.cpp:376-629  (LEGACY packets: per-column frame ids, block status word)
//   LidarFrame                       include/dataframe.hpp           (the per-point arrays that are filled)
// The reference's AVX2 branch is compiled out (CMakeLists.txt:85-89 keeps -mavx2 -mfma commented), so the scalar
// branch (:799-869) is what runs: pt = range_m * x_1 + x_2 with two roundings (no FMA; this file is built with
// -ffp-contract=off).  Third-party arithmetic restated: Eigen::Matrix4d::inverse (cofactors / determinant) and the
// fixed-size 4x4 * 4x1 and 3x3 * 3x1 products as left-to-right dot products.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <memory>
#include <vector>

namespace oracle {

struct LidarConfig {
  int columns_per_frame = 2048, pixels_per_column = 128, columns_per_packet = 16, channel_stride = 1;
  float lidar_origin_to_beam_origin_mm = 0.f;
  float range_min = 1.0f, range_max = 200.0f;  // rfiltermin_/rfiltermax_ defaults, lidarcallback.hpp:76-77
  float z_min = -300.0f, z_max = 0.0f;         // zfiltermin_/zfiltermax_, lidarcallback.hpp:74-75
  int reflectivity_threshold = 0;
  bool legacy = false;                         // udp_profile_lidar == "LEGACY" (lidarcallback.cpp:198-203)
  float box_center[3] = {0, 0, 0}, box_dimensions[3] = {0, 0, 0};
  double Cb2s[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};  // body_to_lidar_rotation_ (row-major), Zero() by default like the reference
  double tb2s[3] = {0, 0, 0};
  std::vector<float> beam_azimuth_deg, beam_altitude_deg;
};

struct LidarFrameO {  // include/dataframe.hpp LidarFrame
  uint16_t frame_id = 0;
  double timestamp = std::numeric_limits<double>::lowest();
  double timestamp_end = std::numeric_limits<double>::lowest();
  double interframe_timedelta = std::numeric_limits<double>::lowest();
  uint32_t numberpoints = 0;
  std::vector<float> x, y, z, relative_timestamp;
  std::vector<uint16_t> c_id, m_id, reflectivity, signal, nir;
  std::vector<double> timestamp_points;
};

// general 4x4 inverse by cofactors (stand-in for Eigen::Matrix4d::inverse, lidarcallback.cpp:280)
inline bool inverse4(const double* m, double* inv) {
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

struct LidarDecoder {
  LidarConfig cfg;
  int subset_channels = 0;
  size_t block_size = 0, expected_size = 0;
  std::vector<float> x1, y1, z1, x2, y2, z2;  // x_1_ [m_id * pixels_per_column + c_id], x_2_ [m_id]
  float box_min[3], box_max[3];
  uint8_t refl_thr = 0;
  // decode state (lidarcallback.hpp:69-71)
  uint16_t frame_id_ = 0;
  uint32_t number_points_ = 0;
  double latest_timestamp_s = 0.0;
  std::unique_ptr<LidarFrameO> active;

  // constructor checks (:41-49) + Initialize (:191-341).  false = the reference would throw.
  bool init(const LidarConfig& c) {
    cfg = c;
    const int N = cfg.channel_stride;
    if (N != 1 && N != 2 && N != 4 && N != 8 && N != 16) return false;
    if (N > cfg.pixels_per_column) return false;
    if (static_cast<int>(cfg.beam_azimuth_deg.size()) != cfg.pixels_per_column || static_cast<int>(cfg.beam_altitude_deg.size()) != cfg.pixels_per_column)
      return false;
    if (cfg.legacy) {  // :198-203: no packet header / footer, 16-byte column header, 4-byte block status
      block_size = 16 + static_cast<size_t>(cfg.pixels_per_column) * 12 + 4;
      expected_size = static_cast<size_t>(cfg.columns_per_packet) * block_size;
    } else {
      block_size = 12 + static_cast<size_t>(cfg.pixels_per_column) * 12;                 // :193-197, :206
      expected_size = 32 + static_cast<size_t>(cfg.columns_per_packet) * block_size + 32;  // :207
    }
    subset_channels = (cfg.pixels_per_column + N - 1) / N;
    const int W = cfg.columns_per_frame, Hh = cfg.pixels_per_column;
    x1.assign(static_cast<size_t>(W) * Hh, 0.f); y1 = x1; z1 = x1;
    x2.assign(W, 0.f); y2 = x2; z2 = x2;
    std::vector<float> cos_alt(Hh), sin_alt(Hh);
    for (int i = 0; i < Hh; ++i) {  // :255-262
      const float alt_rad = cfg.beam_altitude_deg[i] * static_cast<float>(M_PI) / 180.0f;
      sin_alt[i] = std::sin(alt_rad);
      cos_alt[i] = std::cos(alt_rad);
    }
    double b2l[16] = {cfg.Cb2s[0], cfg.Cb2s[1], cfg.Cb2s[2], cfg.tb2s[0], cfg.Cb2s[3], cfg.Cb2s[4], cfg.Cb2s[5], cfg.tb2s[1],
                      cfg.Cb2s[6], cfg.Cb2s[7], cfg.Cb2s[8], cfg.tb2s[2], 0, 0, 0, 1};
    double l2b[16];
    if (!inverse4(b2l, l2b)) return false;  // :276-280 (Eigen would hand back non-finite values)
    for (int m = 0; m < W; ++m) {
      const float az = 2.0f * static_cast<float>(M_PI) * (1.0f - (static_cast<float>(m) / static_cast<float>(W)));  // :285
      const float ca = std::cos(az), sa = std::sin(az);
      const double off[4] = {static_cast<double>(cfg.lidar_origin_to_beam_origin_mm * 0.001f * ca),
                             static_cast<double>(cfg.lidar_origin_to_beam_origin_mm * 0.001f * sa), 0.0, 1.0};  // :289-293
      double t[3];
      for (int r = 0; r < 3; ++r) t[r] = ((l2b[4 * r] * off[0] + l2b[4 * r + 1] * off[1]) + l2b[4 * r + 2] * off[2]) + l2b[4 * r + 3] * off[3];
      x2[m] = static_cast<float>(t[0]); y2[m] = static_cast<float>(t[1]); z2[m] = static_cast<float>(t[2]);
      for (int ch = 0; ch < Hh; ++ch) {  // :300-322
        const float beam_az = cfg.beam_azimuth_deg[ch] * static_cast<float>(M_PI) / 180.0f;
        const float tot = az + beam_az;
        const float ct = std::cos(tot), st = std::sin(tot);
        const double d[3] = {static_cast<double>(cos_alt[ch] * ct), static_cast<double>(cos_alt[ch] * st), static_cast<double>(sin_alt[ch])};
        const size_t k = static_cast<size_t>(m) * Hh + ch;
        x1[k] = static_cast<float>((l2b[0] * d[0] + l2b[1] * d[1]) + l2b[2] * d[2]);
        y1[k] = static_cast<float>((l2b[4] * d[0] + l2b[5] * d[1]) + l2b[6] * d[2]);
        z1[k] = static_cast<float>((l2b[8] * d[0] + l2b[9] * d[1]) + l2b[10] * d[2]);
      }
    }
    for (int a = 0; a < 3; ++a) {  // :325-327
      const float half = cfg.box_dimensions[a] / 2.0f;
      box_min[a] = cfg.box_center[a] - half;
      box_max[a] = cfg.box_center[a] + half;
    }
    refl_thr = static_cast<uint8_t>(static_cast<float>(cfg.reflectivity_threshold));  // uint8_t = json.get<float>(), :157
    frame_id_ = 0; number_points_ = 0; latest_timestamp_s = 0.0;
    active.reset(new LidarFrameO());
    return true;
  }

  static uint16_t rd16(const uint8_t* p) { return static_cast<uint16_t>(p[0] | (p[1] << 8)); }
  static uint64_t rd64(const uint8_t* p) { uint64_t v = 0; for (int i = 7; i >= 0; --i) v = (v << 8) | p[i]; return v; }

  static uint32_t rd32(const uint8_t* p) { return static_cast<uint32_t>(p[0]) | (static_cast<uint32_t>(p[1]) << 8) | (static_cast<uint32_t>(p[2]) << 16) | (static_cast<uint32_t>(p[3]) << 24); }

  // DecodePacketLegacy (:376-629), scalar branch (:567-624; the AVX2 branch is compiled out).  The frame id travels in every
  // COLUMN header, so a frame can end in the middle of a packet -- and a second frame completed by the same packet overwrites
  // `completed_frame` (the first one is lost, like in the reference).
  std::unique_ptr<LidarFrameO> decode_packet_legacy(const uint8_t* packet, size_t size) {
    if (size != expected_size) return nullptr;  // :378-381
    std::unique_ptr<LidarFrameO> completed;
    double prev_frame_completed_latest_ts = 0.0;
    if (latest_timestamp_s > 0.0) prev_frame_completed_latest_ts = latest_timestamp_s;  // :385-388
    const int Hh = cfg.pixels_per_column;
    for (int col = 0; col < cfg.columns_per_packet; ++col) {
      const size_t block = static_cast<size_t>(col) * block_size;
      const uint64_t ts_ns = rd64(packet + block);
      const double col_ts = std::fmod(static_cast<double>(ts_ns) * 1e-9, 86400.0);  // :396
      if (col_ts < 0) continue;
      const uint16_t m_id = rd16(packet + block + 8);
      if (m_id >= static_cast<uint16_t>(cfg.columns_per_frame)) continue;  // :407-410
      const uint16_t fid = rd16(packet + block + 10);
      if (number_points_ == 0 && active->timestamp == 0.0) frame_id_ = fid;  // :416-418
      if (fid != frame_id_) {  // :420-431
        if (number_points_ > 0) {
          active->numberpoints = number_points_;
          active->timestamp_end = latest_timestamp_s;
          completed = std::move(active);
          active.reset(new LidarFrameO());
          prev_frame_completed_latest_ts = latest_timestamp_s;
        }
        number_points_ = 0;
        frame_id_ = fid;
      }
      latest_timestamp_s = col_ts;  // :433
      if (rd32(packet + block + 16 + static_cast<size_t>(Hh) * 12) != 0xFFFFFFFFu) continue;  // :435-441
      bool first_point = (number_points_ == 0);  // :443
      const size_t first_pixel = block + 16;
      for (uint16_t c_id = 0; c_id < static_cast<uint16_t>(Hh); c_id += cfg.channel_stride) {  // :568
        const size_t o = first_pixel + static_cast<size_t>(c_id) * 12;
        if (o + 11 >= size) continue;
        const uint32_t range_mm = rd32(packet + o) & 0x000FFFFFu;  // :576
        const float range_m = static_cast<float>(range_mm) * 0.001f;
        if (range_m < cfg.range_min || range_m > cfg.range_max || range_m == 0) continue;  // :579
        const uint8_t refl = packet[o + 4];
        const uint16_t sig = rd16(packet + o + 6), nir = rd16(packet + o + 8);
        const size_t k = static_cast<size_t>(m_id) * Hh + c_id;
        const float px = range_m * x1[k] + x2[m_id];  // :591-594
        const float py = range_m * y1[k] + y2[m_id];
        const float pz = range_m * z1[k] + z2[m_id];
        const bool in_box = px >= box_min[0] && px <= box_max[0] && py >= box_min[1] && py <= box_max[1] && pz >= box_min[2] && pz <= box_max[2];
        if (!in_box && ((pz >= cfg.z_min && pz <= cfg.z_max) || (refl >= refl_thr))) {  // :601
          const double rel = (active->numberpoints > 0 || number_points_ > 0) && active->timestamp > 0 ? std::max(0.0, col_ts - active->timestamp) : 0.0;
          active->x.push_back(px); active->y.push_back(py); active->z.push_back(pz);
          active->c_id.push_back(c_id); active->m_id.push_back(m_id);
          active->timestamp_points.push_back(col_ts);
          active->relative_timestamp.push_back(static_cast<float>(rel));
          active->reflectivity.push_back(refl); active->signal.push_back(sig); active->nir.push_back(nir);
          ++number_points_;
          if (first_point) {  // :618-624
            active->timestamp = col_ts;
            active->frame_id = frame_id_;
            active->interframe_timedelta = prev_frame_completed_latest_ts > 0.0 ? std::max(0.0, col_ts - prev_frame_completed_latest_ts) : 0.0;
            first_point = false;
          }
        }
      }
    }
    if (active) active->numberpoints = number_points_;  // :628-630
    return completed;
  }

  // DecodePacket (:344-351) dispatch + DecodePacketRng19 (:631-878).  Returns the completed frame or nullptr.
  std::unique_ptr<LidarFrameO> decode_packet(const uint8_t* packet, size_t size) {
    if (cfg.legacy) return decode_packet_legacy(packet, size);
    if (size != expected_size) return nullptr;  // :633-636
    std::unique_ptr<LidarFrameO> completed;
    if (rd16(packet) != 0x0001) return nullptr;  // :640-646
    const uint16_t fid = rd16(packet + 2);
    double prev_frame_completed_latest_ts = 0.0;
    if (number_points_ == 0 && active->timestamp == 0.0) frame_id_ = fid;  // :654-656
    if (fid != frame_id_) {  // :658-669
      if (number_points_ > 0) {
        active->numberpoints = number_points_;
        active->timestamp_end = latest_timestamp_s;
        completed = std::move(active);
        active.reset(new LidarFrameO());
        prev_frame_completed_latest_ts = latest_timestamp_s;
      }
      number_points_ = 0;
      frame_id_ = fid;
    }
    bool first_point = (number_points_ == 0);  // :671
    const int Hh = cfg.pixels_per_column;
    for (int col = 0; col < cfg.columns_per_packet; ++col) {
      const size_t block = 32 + static_cast<size_t>(col) * block_size;
      const uint64_t ts_ns = rd64(packet + block);
      const double col_ts = std::fmod(static_cast<double>(ts_ns) * 1e-9, 86400.0);  // :679
      if (col_ts < 0) continue;
      latest_timestamp_s = col_ts;  // :685
      const uint16_t m_id = rd16(packet + block + 8);
      if (m_id >= static_cast<uint16_t>(cfg.columns_per_frame)) continue;  // :691-694
      if (!(packet[block + 10] & 0x01)) continue;                           // :696-700
      const size_t first_pixel = block + 12;
      for (uint16_t c_id = 0; c_id < static_cast<uint16_t>(Hh); c_id += cfg.channel_stride) {  // :801
        const size_t o = first_pixel + static_cast<size_t>(c_id) * 12;
        if (o + 11 >= size) continue;
        const uint32_t range_mm = (static_cast<uint32_t>(packet[o]) | (static_cast<uint32_t>(packet[o + 1]) << 8) | (static_cast<uint32_t>(packet[o + 2]) << 16)) & 0x0007FFFFu;
        const float range_m = static_cast<float>(range_mm) * 0.001f;
        if (range_m < cfg.range_min || range_m > cfg.range_max || range_m == 0) continue;  // :813 (r_min_/r_max_ hold the same value for every channel)
        const uint8_t refl = packet[o + 4];
        const uint16_t sig = rd16(packet + o + 6), nir = rd16(packet + o + 8);
        const size_t k = static_cast<size_t>(m_id) * Hh + c_id;
        const float px = range_m * x1[k] + x2[m_id];  // :830-832, two roundings (no FMA)
        const float py = range_m * y1[k] + y2[m_id];
        const float pz = range_m * z1[k] + z2[m_id];
        const bool in_box = px >= box_min[0] && px <= box_max[0] && py >= box_min[1] && py <= box_max[1] && pz >= box_min[2] && pz <= box_max[2];
        if (!in_box && ((pz >= cfg.z_min && pz <= cfg.z_max) || (refl >= refl_thr))) {  // :839
          const double rel = (active->numberpoints > 0 || number_points_ > 0) && active->timestamp > 0 ? std::max(0.0, col_ts - active->timestamp) : 0.0;
          active->x.push_back(px); active->y.push_back(py); active->z.push_back(pz);
          active->c_id.push_back(c_id); active->m_id.push_back(m_id);
          active->timestamp_points.push_back(col_ts);
          active->relative_timestamp.push_back(static_cast<float>(rel));
          active->reflectivity.push_back(refl); active->signal.push_back(sig); active->nir.push_back(nir);
          ++number_points_;
          if (first_point) {  // :856-862
            active->timestamp = col_ts;
            active->frame_id = frame_id_;
            active->interframe_timedelta = prev_frame_completed_latest_ts > 0.0 ? std::max(0.0, col_ts - prev_frame_completed_latest_ts) : 0.0;
            first_point = false;
          }
        }
      }
    }
    if (active) active->numberpoints = number_points_;  // :873-875
    return completed;
  }
};

}  // namespace oracle
