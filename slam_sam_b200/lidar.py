"""Host-side mirror of the reference's LidarCallback (src/lidarcallback.cpp, include/lidarcallback.hpp) over the C ABI.

`LidarCallback.DecodePacket(packet)` keeps the reference's contract -- it returns the completed LidarFrame when the packet
that starts the next revolution arrives, else None -- but the decode runs on the device (lidar_decode.cu) and the frame's
points stay in HBM: `frame.device_points()` is what `ndtb_set_source_device` takes.  `DecodePackets` is the replay form
(a whole buffer of packets in one call).  There is no CPU fallback.

`encode_rng19_packets` writes synthetic RNG19_RFL8_SIG16_NIR16 packets (the layout DecodePacketRng19 parses, :631-812);
it exists for the tests and the bench, the reference has no encoder.
"""
import ctypes as C

import numpy as np

from . import ndtb
from .ndtb import LidarConfig, LidarFrameInfo, NdtbError, NDTB_OK, c_d, c_f, _ptr

c_u16 = C.POINTER(C.c_uint16)


def config_from_json(meta, param):
    """ParseMetadata / ParseParamdata (src/lidarcallback.cpp:62-188) for already-parsed JSON dicts.
    Returns (LidarConfig, beam_azimuth_deg, beam_altitude_deg)."""
    cfg = LidarConfig()
    ndtb.lib().ndtb_lidar_default_config(C.byref(cfg))
    fmt, par, beam = meta["lidar_data_format"], meta["config_params"], meta["beam_intrinsics"]
    if par["udp_profile_lidar"] not in ("RNG19_RFL8_SIG16_NIR16", "LEGACY"):
        raise ValueError("Unsupported udp_profile_lidar: " + par["udp_profile_lidar"])  # :205
    cfg.udp_profile_legacy = 1 if par["udp_profile_lidar"] == "LEGACY" else 0
    cfg.columns_per_frame = int(fmt["columns_per_frame"])
    cfg.pixels_per_column = int(fmt["pixels_per_column"])
    cfg.columns_per_packet = int(par["columns_per_packet"])
    cfg.lidar_origin_to_beam_origin_mm = float(beam["lidar_origin_to_beam_origin_mm"])
    lp = param["lidar_parameter"]
    if "tb2s" in lp:
        cfg.body_to_lidar_translation[:] = [float(v) for v in lp["tb2s"]]
    if "Cb2s" in lp:
        cfg.body_to_lidar_rotation[:] = [float(v) for v in lp["Cb2s"]]
    if "channelStride" in lp:
        cfg.channel_stride = int(lp["channelStride"])
    if "zAxisFilter" in lp:
        cfg.z_min, cfg.z_max = float(lp["zAxisFilter"][0]), float(lp["zAxisFilter"][1])
    if "reflectionThreshold" in lp:
        cfg.reflectivity_threshold = int(lp["reflectionThreshold"])
    if "rangeFilter" in lp:
        cfg.range_min, cfg.range_max = float(lp["rangeFilter"][0]), float(lp["rangeFilter"][1])
    box = lp.get("vehicleFilterBox")
    if isinstance(box, dict) and len(box.get("center", [])) == 3 and len(box.get("dimensions", [])) == 3:
        cfg.box_center[:] = [float(v) for v in box["center"]]
        cfg.box_dimensions[:] = [float(v) for v in box["dimensions"]]
    az = np.asarray(beam["beam_azimuth_angles"], np.float32)
    alt = np.asarray(beam["beam_altitude_angles"], np.float32)
    return cfg, az, alt


class LidarFrame:
    """The popped frame: scalar members of include/dataframe.hpp LidarFrame + lazy host copies of its arrays."""

    def __init__(self, owner, info):
        self._o = owner
        self.frame_id = info.frame_id
        self.timestamp = info.timestamp
        self.timestamp_end = info.timestamp_end
        self.interframe_timedelta = info.interframe_timedelta
        self.numberpoints = info.numberpoints
        self.packets = info.packets
        self._arrays = None

    def arrays(self):
        if self._arrays is None:
            n = self.numberpoints
            a = {"x": np.empty(n, np.float32), "y": np.empty(n, np.float32), "z": np.empty(n, np.float32), "c_id": np.empty(n, np.uint16),
                 "m_id": np.empty(n, np.uint16), "timestamp_points": np.empty(n, np.float64), "relative_timestamp": np.empty(n, np.float32),
                 "reflectivity": np.empty(n, np.uint16), "signal": np.empty(n, np.uint16), "nir": np.empty(n, np.uint16)}
            rc = self._o.L.ndtb_lidar_read_frame(self._o.h, n, _ptr(a["x"], c_f), _ptr(a["y"], c_f), _ptr(a["z"], c_f), _ptr(a["c_id"], c_u16),
                                                 _ptr(a["m_id"], c_u16), _ptr(a["timestamp_points"], c_d), _ptr(a["relative_timestamp"], c_f),
                                                 _ptr(a["reflectivity"], c_u16), _ptr(a["signal"], c_u16), _ptr(a["nir"], c_u16))
            self._o._check(rc, "ndtb_lidar_read_frame")
            self._arrays = a
        return self._arrays

    def device_points(self):
        """(device pointer to packed float4 (x, y, z, 1), n): valid until the next frame is popped."""
        p = C.c_void_p(); n = C.c_size_t()
        self._o._check(self._o.L.ndtb_lidar_frame_points(self._o.h, C.byref(p), C.byref(n)), "ndtb_lidar_frame_points")
        return p.value, n.value


class LidarCallback:
    def __init__(self, cfg, beam_azimuth_deg, beam_altitude_deg, device=0):
        self.L = ndtb.lib()
        self.h = C.c_void_p()
        self.cfg = cfg
        az = np.ascontiguousarray(beam_azimuth_deg, np.float32); alt = np.ascontiguousarray(beam_altitude_deg, np.float32)
        if az.size != cfg.pixels_per_column or alt.size != cfg.pixels_per_column:
            raise ValueError("Beam azimuth/altitude angles missing or wrong size")  # src/lidarcallback.cpp:213-216
        rc = self.L.ndtb_lidar_create(device, C.byref(cfg), _ptr(az, c_f), _ptr(alt, c_f), C.byref(self.h))
        if rc != NDTB_OK:
            self.h = None
            raise NdtbError(rc, "ndtb_lidar_create (bad channel stride / no CUDA device: there is no CPU fallback)")
        n = C.c_size_t()
        self.L.ndtb_lidar_packet_bytes(self.h, C.byref(n))
        self.expected_size = n.value

    @classmethod
    def from_json(cls, meta, param, device=0):
        return cls(*config_from_json(meta, param), device=device)

    def _check(self, rc, what):
        if rc != NDTB_OK:
            raise NdtbError(rc, f"{what}: {self.L.ndtb_lidar_last_error(self.h).decode()}")

    def close(self):
        if getattr(self, "h", None):
            self.L.ndtb_lidar_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def tables(self):
        W, Hh = self.cfg.columns_per_frame, self.cfg.pixels_per_column
        t = [np.empty(W * Hh, np.float32) for _ in range(3)] + [np.empty(W, np.float32) for _ in range(3)]
        self._check(self.L.ndtb_lidar_get_tables(self.h, *[_ptr(a, c_f) for a in t]), "ndtb_lidar_get_tables")
        return t

    def push(self, packets, n_packets=None, packet_bytes=None, pitch=None):
        """packets: uint8 array (n, packet_bytes) / bytes / int address (then n_packets and packet_bytes are required)."""
        if isinstance(packets, int):
            addr = packets
        else:
            a = np.frombuffer(packets, np.uint8) if isinstance(packets, (bytes, bytearray, memoryview)) else packets
            if a.ndim == 1:
                packet_bytes = packet_bytes or a.size
                a = a.reshape(-1, packet_bytes)
            assert a.flags.c_contiguous and a.dtype == np.uint8
            n_packets, packet_bytes = a.shape
            addr = a.ctypes.data
        closed = C.c_int(0)
        rc = self.L.ndtb_lidar_push_packets(self.h, addr, n_packets, packet_bytes, pitch or packet_bytes, C.byref(closed))
        self._check(rc, "ndtb_lidar_push_packets")
        return closed.value

    def pop(self):
        info = LidarFrameInfo(); avail = C.c_int(0)
        self._check(self.L.ndtb_lidar_pop_frame(self.h, C.byref(info), C.byref(avail)), "ndtb_lidar_pop_frame")
        return LidarFrame(self, info) if avail.value else None

    def sync(self):
        self._check(self.L.ndtb_lidar_sync(self.h), "ndtb_lidar_sync")

    def stats(self):
        n = C.c_uint64(); us = C.c_float()
        self.L.ndtb_lidar_get_stats(self.h, C.byref(n), C.byref(us))
        return n.value, us.value

    # --- the reference's method names -------------------------------------------------------------------------------
    def DecodePacket(self, packet):
        """LidarCallback::DecodePacket (src/lidarcallback.cpp:344-351): the completed frame or None."""
        return self.pop() if self.push(packet) else None

    DecodePacketRng19 = DecodePacket

    def DecodePackets(self, packets):
        """Replay form: every frame the packets complete, in order."""
        frames = []
        if self.push(packets):
            while True:
                f = self.pop()
                if f is None:
                    break
                f.arrays()  # a later pop recycles the slot
                frames.append(f)
        return frames


def encode_rng19_packets(range_mm, reflectivity, signal, nir, timestamps_ns, frame_ids, columns_per_packet=16, m_ids=None, status=None,
                         rng=None):
    """Synthetic RNG19_RFL8_SIG16_NIR16 packets.  range_mm etc.: (columns, pixels) arrays; timestamps_ns, frame_ids (and
    optionally m_ids, status): per column.  Columns are packed `columns_per_packet` at a time; a packet carries the frame
    id of its first column.  Bytes the decoder must ignore are filled with noise from `rng` (if given).
    Returns a uint8 array (n_packets, 32 + cpp * (12 + 12 * pixels) + 32)."""
    range_mm = np.asarray(range_mm, np.uint32)
    ncol, H = range_mm.shape
    assert ncol % columns_per_packet == 0
    n_pk = ncol // columns_per_packet
    block = 12 + 12 * H
    size = 32 + columns_per_packet * block + 32
    noise = (lambda shape: np.zeros(shape, np.uint8)) if rng is None else (lambda shape: rng.integers(0, 256, shape, dtype=np.uint8))
    out = noise((n_pk, size))
    m_ids = np.arange(ncol) if m_ids is None else np.asarray(m_ids)
    status = np.ones(ncol, np.uint8) if status is None else np.asarray(status, np.uint8)
    frame_ids = np.asarray(frame_ids, np.uint16)
    out[:, 0] = 1; out[:, 1] = 0  # packet_type 0x0001
    fid = frame_ids.reshape(n_pk, columns_per_packet)[:, 0]
    out[:, 2] = fid & 0xff; out[:, 3] = fid >> 8
    px = noise((ncol, H, 12))
    r = range_mm & 0x7FFFF
    px[:, :, 0] = r & 0xff
    px[:, :, 1] = (r >> 8) & 0xff
    px[:, :, 2] = (px[:, :, 2] & 0xF8) | ((r >> 16) & 0x07)  # the upper five bits of byte 2 are not range (mask 0x0007FFFF)
    px[:, :, 4] = np.asarray(reflectivity, np.uint8)
    px[:, :, 6:8] = np.ascontiguousarray(signal, np.uint16).astype("<u2").view(np.uint8).reshape(ncol, H, 2)
    px[:, :, 8:10] = np.ascontiguousarray(nir, np.uint16).astype("<u2").view(np.uint8).reshape(ncol, H, 2)
    cols = noise((ncol, block))
    cols[:, 0:8] = np.ascontiguousarray(timestamps_ns, np.uint64).astype("<u8").view(np.uint8).reshape(ncol, 8)
    cols[:, 8:10] = np.ascontiguousarray(m_ids).astype("<u2").view(np.uint8).reshape(ncol, 2)
    cols[:, 10] = status
    cols[:, 12:] = px.reshape(ncol, 12 * H)
    out[:, 32:32 + columns_per_packet * block] = cols.reshape(n_pk, columns_per_packet * block)
    return out


def encode_legacy_packets(range_mm, reflectivity, signal, nir, timestamps_ns, frame_ids, columns_per_packet=16, m_ids=None, status=None, rng=None):
    """Synthetic LEGACY packets (the layout DecodePacketLegacy parses, src/lidarcallback.cpp:391-497): no packet header; per
    column 16 header bytes {time stamp u64, measurement id u16, frame id u16, encoder u32}, 12 bytes per pixel {range u32 (20
    bits used), reflectivity u8, -, signal u16, nir u16, -} and a 4-byte block status (0xFFFFFFFF = valid).  The frame id is
    PER COLUMN: a frame may end in the middle of a packet.  `status`: per column, truthy = valid.
    Returns a uint8 array (n_packets, cpp * (16 + 12 * pixels + 4))."""
    range_mm = np.asarray(range_mm, np.uint32)
    ncol, H = range_mm.shape
    assert ncol % columns_per_packet == 0
    n_pk = ncol // columns_per_packet
    block = 16 + 12 * H + 4
    noise = (lambda shape: np.zeros(shape, np.uint8)) if rng is None else (lambda shape: rng.integers(0, 256, shape, dtype=np.uint8))
    m_ids = np.arange(ncol) if m_ids is None else np.asarray(m_ids)
    status = np.ones(ncol, np.uint8) if status is None else np.asarray(status, np.uint8)
    px = noise((ncol, H, 12))
    r = range_mm & 0xFFFFF
    px[:, :, 0] = r & 0xff
    px[:, :, 1] = (r >> 8) & 0xff
    px[:, :, 2] = (px[:, :, 2] & 0xF0) | ((r >> 16) & 0x0F)   # the upper twelve bits of the range word are not range (mask 0x000FFFFF)
    px[:, :, 4] = np.asarray(reflectivity, np.uint8)
    px[:, :, 6:8] = np.ascontiguousarray(signal, np.uint16).astype("<u2").view(np.uint8).reshape(ncol, H, 2)
    px[:, :, 8:10] = np.ascontiguousarray(nir, np.uint16).astype("<u2").view(np.uint8).reshape(ncol, H, 2)
    cols = noise((ncol, block))
    cols[:, 0:8] = np.ascontiguousarray(timestamps_ns, np.uint64).astype("<u8").view(np.uint8).reshape(ncol, 8)
    cols[:, 8:10] = np.ascontiguousarray(m_ids).astype("<u2").view(np.uint8).reshape(ncol, 2)
    cols[:, 10:12] = np.ascontiguousarray(frame_ids, np.uint16).astype("<u2").view(np.uint8).reshape(ncol, 2)
    cols[:, 16:16 + 12 * H] = px.reshape(ncol, 12 * H)
    good = status.astype(bool)
    cols[good, 16 + 12 * H:] = 0xFF
    bad = cols[~good, 16 + 12 * H:]
    bad[:, 0] &= 0xFE                                           # anything but 0xFFFFFFFF
    cols[~good, 16 + 12 * H:] = bad
    return cols.reshape(n_pk, columns_per_packet * block)
