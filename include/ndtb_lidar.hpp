// ndtb_lidar.hpp -- header-only C++ facade over the scan-producer entry points of include/ndtb.h, with the class and
// method names of the reference's LidarCallback (include/lidarcallback.hpp:14-23, src/lidarcallback.cpp) and its
// LidarFrame (include/dataframe.hpp).  The JSON parsing of the reference's constructors (ParseMetadata / ParseParamdata,
// src/lidarcallback.cpp:62-188) stays with the caller: it fills an ndtb_lidar_config (the reference's private members,
// same meaning) and the two beam-angle arrays.  Decoding runs on the device (slam_sam_b200/csrc/lidar_decode.cu);
// DecodePacket returns the completed frame when the packet that starts the next revolution arrives, else nullptr --
// the reference's contract.  There is no CPU fallback: the constructor throws without a CUDA device.
#ifndef NDTB_LIDAR_HPP_
#define NDTB_LIDAR_HPP_

#include "ndtb.h"
#include <cstdint>
#include <deque>
#include <limits>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace ndtb_facade {

struct LidarFrame {  // include/dataframe.hpp LidarFrame (the members the decoder fills)
  uint16_t frame_id = 0;
  double timestamp = std::numeric_limits<double>::lowest();
  double timestamp_end = std::numeric_limits<double>::lowest();
  double interframe_timedelta = std::numeric_limits<double>::lowest();
  uint32_t numberpoints = 0;
  std::vector<float> x, y, z;
  std::vector<uint16_t> c_id, m_id;
  std::vector<double> timestamp_points;
  std::vector<float> relative_timestamp;
  std::vector<uint16_t> reflectivity, signal, nir;
  // not in the reference: the same points, packed (x, y, z, 1) in HBM, valid until the next completed frame
  const void* device_points = nullptr;

  void clear() {
    *this = LidarFrame();
  }
  void resize(size_t n) {
    x.resize(n); y.resize(n); z.resize(n); c_id.resize(n); m_id.resize(n); timestamp_points.resize(n); relative_timestamp.resize(n);
    reflectivity.resize(n); signal.resize(n); nir.resize(n);
  }
};

class LidarCallback {
 public:
  // host_arrays = false skips the device -> host copy of the per-point arrays (a registrar only needs device_points)
  LidarCallback(const ndtb_lidar_config& cfg, const std::vector<float>& beam_azimuth_deg, const std::vector<float>& beam_altitude_deg,
                int cuda_device = 0, bool host_arrays = true)
      : host_arrays_(host_arrays) {
    if (cfg.channel_stride == 0) throw std::runtime_error("Channel stride N must be positive");  // src/lidarcallback.cpp:41-49
    if (beam_azimuth_deg.size() != static_cast<size_t>(cfg.pixels_per_column) || beam_altitude_deg.size() != static_cast<size_t>(cfg.pixels_per_column))
      throw std::runtime_error("Beam azimuth/altitude angles missing or wrong size in JSON.");  // :213-216
    const int rc = ndtb_lidar_create(cuda_device, &cfg, beam_azimuth_deg.data(), beam_altitude_deg.data(), &h_);
    if (rc == NDTB_ERR_INVALID) throw std::runtime_error("Channel stride N must be one of 1, 2, 4, 8, or 16 and not exceed pixels_per_column_");
    if (rc != NDTB_OK) throw std::runtime_error("ndtb_lidar_create failed (no CUDA device? there is no CPU fallback)");
  }
  ~LidarCallback() { if (h_) ndtb_lidar_destroy(h_); }
  LidarCallback(const LidarCallback&) = delete;
  LidarCallback& operator=(const LidarCallback&) = delete;

  std::unique_ptr<LidarFrame> DecodePacket(const std::vector<uint8_t>& packet) { return DecodePackets(packet.data(), 1, packet.size()); }
  std::unique_ptr<LidarFrame> DecodePacketRng19(const std::vector<uint8_t>& packet) { return DecodePacket(packet); }

  // n packets laid out back to back (a replay buffer); returns the first frame they complete, further ones through NextFrame()
  std::unique_ptr<LidarFrame> DecodePackets(const uint8_t* packets, size_t n, size_t packet_bytes) {
    int closed = 0;
    const int rc = ndtb_lidar_push_packets(h_, packets, n, packet_bytes, packet_bytes, &closed);
    if (rc != NDTB_OK) throw std::runtime_error(std::string("ndtb_lidar_push_packets: ") + ndtb_lidar_last_error(h_));
    return closed ? NextFrame() : nullptr;
  }

  std::unique_ptr<LidarFrame> NextFrame() {
    ndtb_lidar_frame_info info;
    int available = 0;
    if (ndtb_lidar_pop_frame(h_, &info, &available) != NDTB_OK) throw std::runtime_error(std::string("ndtb_lidar_pop_frame: ") + ndtb_lidar_last_error(h_));
    if (!available) return nullptr;
    std::unique_ptr<LidarFrame> f = GetFrameFromPool();
    f->frame_id = info.frame_id; f->timestamp = info.timestamp; f->timestamp_end = info.timestamp_end;
    f->interframe_timedelta = info.interframe_timedelta; f->numberpoints = info.numberpoints;
    size_t n = 0;
    ndtb_lidar_frame_points(h_, &f->device_points, &n);
    if (host_arrays_) {
      f->resize(info.numberpoints);
      const int rc = ndtb_lidar_read_frame(h_, info.numberpoints, f->x.data(), f->y.data(), f->z.data(), f->c_id.data(), f->m_id.data(),
                                           f->timestamp_points.data(), f->relative_timestamp.data(), f->reflectivity.data(), f->signal.data(),
                                           f->nir.data());
      if (rc != NDTB_OK) throw std::runtime_error(std::string("ndtb_lidar_read_frame: ") + ndtb_lidar_last_error(h_));
    }
    return f;
  }

  void ReturnFrameToPool(std::unique_ptr<LidarFrame> frame) {  // src/lidarcallback.cpp:903-909
    if (frame) { frame->clear(); pool_.push_back(std::move(frame)); }
  }

  size_t expected_size() const { size_t n = 0; ndtb_lidar_packet_bytes(h_, &n); return n; }
  ndtb_lidar handle() const { return h_; }

 private:
  std::unique_ptr<LidarFrame> GetFrameFromPool() {  // :893-901
    if (!pool_.empty()) { auto f = std::move(pool_.front()); pool_.pop_front(); return f; }
    return std::unique_ptr<LidarFrame>(new LidarFrame());
  }
  ndtb_lidar h_ = nullptr;
  bool host_arrays_ = true;
  std::deque<std::unique_ptr<LidarFrame>> pool_;
};

}  // namespace ndtb_facade

#ifndef NDTB_FACADE_NO_GLOBAL_NAMES
using ndtb_facade::LidarCallback;  // the reference declares both at global scope
using ndtb_facade::LidarFrame;
#endif
#endif  // NDTB_LIDAR_HPP_
