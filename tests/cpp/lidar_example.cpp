// Drives include/ndtb_lidar.hpp the way the reference's drivers drive LidarCallback (one DecodePacket per UDP datagram,
// run/pipeline*.cpp).  Usage: lidar_example packets.bin columns pixels stride  (raw packets back to back; the synthetic
// sensor of tests/test_lidar_decode.py make_config(rotated=False)); prints one line per completed frame:
// "frame_id numberpoints timestamp timestamp_end interframe_timedelta sum(x) sum(c_id) sum(signal)".
#include "ndtb_lidar.hpp"
#include <cstdio>
#include <cstdlib>
#include <fstream>

int main(int argc, char** argv) {
  if (argc < 5) { std::fprintf(stderr, "usage: %s packets.bin columns pixels stride\n", argv[0]); return 2; }
  ndtb_lidar_config c;
  ndtb_lidar_default_config(&c);
  c.columns_per_frame = std::atoi(argv[2]); c.pixels_per_column = std::atoi(argv[3]); c.columns_per_packet = 16; c.channel_stride = std::atoi(argv[4]);
  c.lidar_origin_to_beam_origin_mm = 27.397f;
  c.range_min = 0.0f; c.range_max = 250.0f; c.z_min = -250.0f; c.z_max = 0.0f; c.reflectivity_threshold = 60;
  c.box_center[0] = -3.0f; c.box_center[1] = 0.0f; c.box_center[2] = 1.0f;
  c.box_dimensions[0] = 10.5f; c.box_dimensions[1] = 3.0f; c.box_dimensions[2] = 2.1f;
  const double R[9] = {-1, 0, 0, 0, -1, 0, 0, 0, 1}, t[3] = {0.135, 0.0, 0.1243};
  for (int i = 0; i < 9; ++i) c.body_to_lidar_rotation[i] = R[i];
  for (int i = 0; i < 3; ++i) c.body_to_lidar_translation[i] = t[i];
  const int H = c.pixels_per_column;
  std::vector<float> az(H), alt(H);
  const float pattern[4] = {2.09f, 0.71f, -0.69f, -2.06f};
  for (int i = 0; i < H; ++i) {
    az[i] = pattern[i % 4];
    // numpy.linspace(21.3, -21.9, H).astype(float32): start + i * step in double
    alt[i] = static_cast<float>(H > 1 ? 21.3 + static_cast<double>(i) * ((-21.9 - 21.3) / static_cast<double>(H - 1)) : 21.3);
  }
  if (H > 1) alt[H - 1] = static_cast<float>(-21.9);
  LidarCallback cb(c, az, alt);
  std::ifstream f(argv[1], std::ios::binary | std::ios::ate);
  if (!f) { std::fprintf(stderr, "cannot open %s\n", argv[1]); return 1; }
  std::vector<uint8_t> all(static_cast<size_t>(f.tellg()));
  f.seekg(0);
  f.read(reinterpret_cast<char*>(all.data()), static_cast<std::streamsize>(all.size()));
  const size_t size = cb.expected_size();
  for (size_t o = 0; o + size <= all.size(); o += size) {
    std::vector<uint8_t> packet(all.begin() + static_cast<std::ptrdiff_t>(o), all.begin() + static_cast<std::ptrdiff_t>(o + size));
    std::unique_ptr<LidarFrame> fr = cb.DecodePacket(packet);
    if (!fr) continue;
    double sx = 0; unsigned long long sc = 0, ss = 0;
    for (size_t i = 0; i < fr->numberpoints; ++i) { sx += fr->x[i]; sc += fr->c_id[i]; ss += fr->signal[i]; }
    std::printf("%u %u %.17g %.17g %.17g %.17g %llu %llu\n", fr->frame_id, fr->numberpoints, fr->timestamp, fr->timestamp_end, fr->interframe_timedelta, sx, sc, ss);
    cb.ReturnFrameToPool(std::move(fr));
  }
  return 0;
}
