// ORACLE — TEST INFRASTRUCTURE ONLY.  C entry points (ctypes) over the Ouster RNG19 decoder restatement
// (ouster_decode.hpp).  Only tests/ and bench.py's CPU legs may load this library.
#include "ouster_decode.hpp"

using namespace oracle;

namespace {
struct LidarOracle {
  LidarDecoder dec;
  std::vector<std::unique_ptr<LidarFrameO>> frames;  // completed frames, in completion order
};
}  // namespace

extern "C" {

// cfg_i: columns_per_frame, pixels_per_column, columns_per_packet, channel_stride, reflectivity_threshold, legacy profile (0 / 1)
// cfg_f: lidar_origin_to_beam_origin_mm, range_min, range_max, z_min, z_max, box_center[3], box_dimensions[3]
// cfg_d: Cb2s[9] row-major, tb2s[3]
void* orc_lidar_create(const int* cfg_i, const float* cfg_f, const double* cfg_d, const float* az_deg, const float* alt_deg) {
  LidarConfig c;
  c.columns_per_frame = cfg_i[0]; c.pixels_per_column = cfg_i[1]; c.columns_per_packet = cfg_i[2]; c.channel_stride = cfg_i[3];
  c.reflectivity_threshold = cfg_i[4]; c.legacy = cfg_i[5] != 0;
  c.lidar_origin_to_beam_origin_mm = cfg_f[0]; c.range_min = cfg_f[1]; c.range_max = cfg_f[2]; c.z_min = cfg_f[3]; c.z_max = cfg_f[4];
  for (int a = 0; a < 3; ++a) { c.box_center[a] = cfg_f[5 + a]; c.box_dimensions[a] = cfg_f[8 + a]; c.tb2s[a] = cfg_d[9 + a]; }
  for (int i = 0; i < 9; ++i) c.Cb2s[i] = cfg_d[i];
  if (c.pixels_per_column <= 0 || c.pixels_per_column > 4096) return nullptr;
  c.beam_azimuth_deg.assign(az_deg, az_deg + c.pixels_per_column);
  c.beam_altitude_deg.assign(alt_deg, alt_deg + c.pixels_per_column);
  LidarOracle* o = new LidarOracle();
  if (!o->dec.init(c)) { delete o; return nullptr; }
  return o;
}
void orc_lidar_destroy(void* h) { delete static_cast<LidarOracle*>(h); }
size_t orc_lidar_packet_bytes(void* h) { return static_cast<LidarOracle*>(h)->dec.expected_size; }
void orc_lidar_tables(void* h, float* x1, float* y1, float* z1, float* x2, float* y2, float* z2) {
  LidarDecoder& d = static_cast<LidarOracle*>(h)->dec;
  std::memcpy(x1, d.x1.data(), d.x1.size() * 4); std::memcpy(y1, d.y1.data(), d.y1.size() * 4); std::memcpy(z1, d.z1.data(), d.z1.size() * 4);
  std::memcpy(x2, d.x2.data(), d.x2.size() * 4); std::memcpy(y2, d.y2.data(), d.y2.size() * 4); std::memcpy(z2, d.z2.data(), d.z2.size() * 4);
}
// DecodePacket for n packets laid out every `pitch` bytes, each `size` bytes long; returns the number of frames completed so far.
long orc_lidar_decode(void* h, const uint8_t* packets, size_t n, size_t size, size_t pitch) {
  LidarOracle* o = static_cast<LidarOracle*>(h);
  for (size_t i = 0; i < n; ++i) {
    auto f = o->dec.decode_packet(packets + i * pitch, size);
    if (f) o->frames.push_back(std::move(f));
  }
  return static_cast<long>(o->frames.size());
}
// info: frame_id, timestamp, timestamp_end, interframe_timedelta, numberpoints
int orc_lidar_frame_info(void* h, long k, double* info) {
  LidarOracle* o = static_cast<LidarOracle*>(h);
  if (k < 0 || k >= static_cast<long>(o->frames.size())) return -1;
  const LidarFrameO& f = *o->frames[k];
  info[0] = f.frame_id; info[1] = f.timestamp; info[2] = f.timestamp_end; info[3] = f.interframe_timedelta; info[4] = f.numberpoints;
  return 0;
}
int orc_lidar_frame_get(void* h, long k, float* x, float* y, float* z, uint16_t* c_id, uint16_t* m_id, double* ts, float* rel, uint16_t* refl,
                        uint16_t* sig, uint16_t* nir) {
  LidarOracle* o = static_cast<LidarOracle*>(h);
  if (k < 0 || k >= static_cast<long>(o->frames.size())) return -1;
  const LidarFrameO& f = *o->frames[k];
  const size_t n = f.x.size();
  std::memcpy(x, f.x.data(), n * 4); std::memcpy(y, f.y.data(), n * 4); std::memcpy(z, f.z.data(), n * 4);
  std::memcpy(c_id, f.c_id.data(), n * 2); std::memcpy(m_id, f.m_id.data(), n * 2);
  std::memcpy(ts, f.timestamp_points.data(), n * 8); std::memcpy(rel, f.relative_timestamp.data(), n * 4);
  std::memcpy(refl, f.reflectivity.data(), n * 2); std::memcpy(sig, f.signal.data(), n * 2); std::memcpy(nir, f.nir.data(), n * 2);
  return 0;
}
void orc_lidar_clear_frames(void* h) { static_cast<LidarOracle*>(h)->frames.clear(); }

}  // extern "C"
