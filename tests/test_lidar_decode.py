"""Scan producer: Ouster RNG19 packets -> LidarFrame, the reference's LidarCallback
(src/lidarcallback.cpp:191-341, 631-878).  CPU tests pin the oracle restatement (numpy re-derivation, state-machine cases,
golden file); GPU tests compare the device decoder with the oracle bit for bit through the C ABI."""
import ctypes as C
import os

import numpy as np
import pytest

from slam_sam_b200 import ndtb
from slam_sam_b200.lidar import LidarCallback, encode_rng19_packets

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "lidar_golden.npz")
ARRAYS = ("x", "y", "z", "c_id", "m_id", "timestamp_points", "relative_timestamp", "reflectivity", "signal", "nir")
SCALARS = ("frame_id", "timestamp", "timestamp_end", "interframe_timedelta", "numberpoints")


def make_config(cols=64, pixels=32, stride=1, rotated=False, **kw):
    """Synthetic sensor in the shape of config/lidar_meta_berlin.json + lidar_config_berlin.json (OS1-style beam fan)."""
    cfg = ndtb.LidarConfig()
    cfg.columns_per_frame, cfg.pixels_per_column, cfg.columns_per_packet, cfg.channel_stride = cols, pixels, 16, stride
    cfg.lidar_origin_to_beam_origin_mm = 27.397
    cfg.range_min, cfg.range_max = 0.0, 250.0
    cfg.z_min, cfg.z_max = -250.0, 0.0
    cfg.reflectivity_threshold = 60
    cfg.box_center[:] = [-3.0, 0.0, 1.0]
    cfg.box_dimensions[:] = [10.5, 3.0, 2.1]
    if rotated:
        a, b = np.deg2rad(12.5), np.deg2rad(-4.0)
        Rz = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]])
        Ry = np.array([[np.cos(b), 0, np.sin(b)], [0, 1, 0], [-np.sin(b), 0, np.cos(b)]])
        cfg.body_to_lidar_rotation[:] = list((Rz @ Ry).reshape(9))
        cfg.body_to_lidar_translation[:] = [0.135, -0.02, 0.1243]
    else:
        cfg.body_to_lidar_rotation[:] = [-1.0, 0, 0, 0, -1.0, 0, 0, 0, 1.0]
        cfg.body_to_lidar_translation[:] = [0.135, 0.0, 0.1243]
    for k, v in kw.items():
        setattr(cfg, k, v)
    alt = np.linspace(21.3, -21.9, pixels).astype(np.float32)
    az = np.tile(np.array([2.09, 0.71, -0.69, -2.06], np.float32), (pixels + 3) // 4)[:pixels]
    return cfg, az, alt


def make_stream(cfg, n_frames, seed, first_frame_id=7, t0_ns=1_762_300_000_000_000_000, noise=True, empty_frames=(), bad_columns=True):
    """Packets of n_frames revolutions (+ one packet of the next, which closes the last).  Returns uint8 (n_packets, size)."""
    rng = np.random.default_rng(seed)
    W, Hh, cpp = cfg.columns_per_frame, cfg.pixels_per_column, cfg.columns_per_packet
    out = []
    t = t0_ns
    for f in range(n_frames + 1):
        ncol = W if f < n_frames else cpp
        rng_mm = rng.integers(0, 1 << 19, (ncol, Hh)).astype(np.uint32)
        close = rng.random((ncol, Hh))
        rng_mm[close < 0.25] = rng.integers(0, 9000, int((close < 0.25).sum()))  # returns around the vehicle box
        rng_mm[close > 0.9] = 0                                                  # no return
        if f in empty_frames:
            rng_mm[:] = 0
        refl = rng.integers(0, 256, (ncol, Hh)); sig = rng.integers(0, 65536, (ncol, Hh)); nir = rng.integers(0, 65536, (ncol, Hh))
        ts = t + np.arange(ncol, dtype=np.uint64) * np.uint64(48_828)
        t = int(ts[-1]) + 48_828
        m_ids = np.arange(ncol) % W
        status = np.ones(ncol, np.uint8)
        if bad_columns and f < n_frames:
            status[rng.integers(0, ncol, 3)] = rng.integers(0, 256, 3) & 0xFE   # column_status bit 0 clear -> column skipped
            m_ids = m_ids.copy(); m_ids[rng.integers(0, ncol, 2)] = W + 5        # invalid measurement id -> column skipped
        fid = np.full(ncol, (first_frame_id + f) & 0xffff)
        out.append(encode_rng19_packets(rng_mm, refl, sig, nir, ts, fid, cpp, m_ids=m_ids, status=status, rng=rng if noise else None))
    return np.concatenate(out)


def assert_frames_equal(got, want):
    assert len(got) == len(want)
    for g, w in zip(got, want):
        for k in SCALARS:
            assert g[k] == w[k], k
        for k in ARRAYS:
            assert np.array_equal(g[k].view(np.uint8), w[k].view(np.uint8)), k


def numpy_decode_frame(cfg, tables, packets):
    """Independent vectorised re-derivation of one frame's points (all packets carry the same frame id)."""
    x1, y1, z1, x2, y2, z2 = tables
    W, Hh, cpp, N = cfg.columns_per_frame, cfg.pixels_per_column, cfg.columns_per_packet, cfg.channel_stride
    block = 12 + 12 * Hh
    cols = packets[:, 32:32 + cpp * block].reshape(-1, block)
    ts = cols[:, :8].copy().view("<u8")[:, 0].astype(np.float64) * 1e-9
    ts = np.fmod(ts, 86400.0)
    m_id = cols[:, 8:10].copy().view("<u2")[:, 0].astype(np.int64)
    ok = (m_id < W) & ((cols[:, 10] & 1) == 1)
    px = cols[:, 12:].reshape(-1, Hh, 12)[:, ::N]
    r_mm = (px[..., 0].astype(np.uint32) | (px[..., 1].astype(np.uint32) << 8) | (px[..., 2].astype(np.uint32) << 16)) & 0x7FFFF
    r = r_mm.astype(np.float32) * np.float32(0.001)
    mm = np.where(ok, m_id, 0)
    c_id = np.arange(0, Hh, N)
    k = mm[:, None] * Hh + c_id[None, :]
    X = r * x1[k] + x2[mm][:, None]; Y = r * y1[k] + y2[mm][:, None]; Z = r * z1[k] + z2[mm][:, None]
    half = np.array(cfg.box_dimensions[:], np.float32) / np.float32(2.0)
    lo = np.array(cfg.box_center[:], np.float32) - half; hi = np.array(cfg.box_center[:], np.float32) + half
    in_box = (X >= lo[0]) & (X <= hi[0]) & (Y >= lo[1]) & (Y <= hi[1]) & (Z >= lo[2]) & (Z <= hi[2])
    refl = px[..., 4]
    keep = ok[:, None] & ~((r < np.float32(cfg.range_min)) | (r > np.float32(cfg.range_max)) | (r == 0)) & ~in_box & \
        (((Z >= np.float32(cfg.z_min)) & (Z <= np.float32(cfg.z_max))) | (refl >= cfg.reflectivity_threshold))
    return dict(x=X[keep], y=Y[keep], z=Z[keep], c_id=np.broadcast_to(c_id, X.shape)[keep].astype(np.uint16),
                m_id=np.broadcast_to(m_id[:, None], X.shape)[keep].astype(np.uint16), timestamp_points=np.broadcast_to(ts[:, None], X.shape)[keep])


# ------------------------------------------------------------------------------------------------------ CPU: the oracle
@pytest.mark.parametrize("stride,rotated", [(1, False), (4, False), (2, True)])
def test_oracle_matches_numpy_rederivation(oracle, stride, rotated):
    cfg, az, alt = make_config(stride=stride, rotated=rotated)
    o = oracle.OracleLidar(cfg, az, alt)
    assert o.expected_size == 32 + 16 * (12 + 12 * 32) + 32
    pk = make_stream(cfg, 1, seed=3)
    frames = o.decode(pk)
    assert len(frames) == 1
    ref = numpy_decode_frame(cfg, o.tables(), pk[:-1])
    f = frames[0]
    assert f["numberpoints"] == len(ref["x"]) > 100
    for k in ref:
        assert np.array_equal(f[k], ref[k]), k
    # scalars: first point's column time, last column of the frame, first frame has no predecessor
    assert f["frame_id"] == 7 and f["timestamp"] == f["timestamp_points"][0] and f["interframe_timedelta"] == 0.0
    rel = np.maximum(0.0, f["timestamp_points"] - f["timestamp"]).astype(np.float32)
    assert np.array_equal(f["relative_timestamp"], rel)
    # the lookup tables are unit directions rotated into the body frame
    x1, y1, z1 = o.tables()[:3]
    assert np.allclose(x1 * x1 + y1 * y1 + z1 * z1, 1.0, atol=1e-6)


def test_oracle_frame_state_machine(oracle):
    cfg, az, alt = make_config()
    o = oracle.OracleLidar(cfg, az, alt)
    pk = make_stream(cfg, 4, seed=5, empty_frames=(1,))
    frames = o.decode(pk)
    # frame 1 has no returns: it is never emitted, and frame 2 then has no predecessor time (prev stays 0, :649-669)
    assert [f["frame_id"] for f in frames] == [7, 9, 10]
    assert frames[1]["interframe_timedelta"] == 0.0
    assert frames[2]["interframe_timedelta"] == pytest.approx(frames[2]["timestamp"] - frames[1]["timestamp_end"]) and frames[2]["interframe_timedelta"] > 0
    # packets of the wrong size or type are ignored (:633-646)
    o2 = oracle.OracleLidar(cfg, az, alt)
    bad = pk.copy(); bad[2, 0] = 2
    f2 = o2.decode(bad)
    assert f2[0]["numberpoints"] < frames[0]["numberpoints"]
    o3 = oracle.OracleLidar(cfg, az, alt)
    assert o3.decode(pk[:, :-1]) == []
    with pytest.raises(ValueError):
        oracle.OracleLidar(make_config(stride=3)[0], az, alt)  # the constructor throws (:44-46)


def test_oracle_golden_file(oracle):
    g = np.load(GOLDEN)
    cfg, az, alt = make_config(stride=int(g["stride"]), rotated=True)
    frames = oracle.OracleLidar(cfg, az, alt).decode(g["packets"])
    assert len(frames) == int(g["n_frames"])
    for i, f in enumerate(frames):
        for k in ARRAYS:
            assert np.array_equal(f[k].view(np.uint8), g[f"f{i}_{k}"].view(np.uint8)), k
        assert np.array_equal(np.array([f[k] for k in SCALARS], np.float64), g[f"f{i}_scalars"])


def test_lidar_create_needs_a_gpu_and_validates():
    import torch

    cfg, az, alt = make_config(stride=3)
    h = C.c_void_p()
    L = ndtb.lib()
    assert L.ndtb_lidar_create(0, C.byref(cfg), az.ctypes.data_as(ndtb.c_f), alt.ctypes.data_as(ndtb.c_f), C.byref(h)) == ndtb.NDTB_ERR_INVALID
    if not torch.cuda.is_available():
        with pytest.raises(ndtb.NdtbError):
            LidarCallback(*make_config())  # no CPU fallback
    assert C.sizeof(ndtb.LidarConfig) == 168 and C.sizeof(ndtb.LidarFrameInfo) == 32


# ------------------------------------------------------------------------------------------------- GPU: device decoder
def frames_as_dicts(frames):
    out = []
    for f in frames:
        d = dict(f.arrays())
        d.update({k: getattr(f, k) for k in SCALARS})
        out.append(d)
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("stride,rotated", [(1, False), (4, True), (16, False)])
def test_gpu_decoder_matches_oracle(oracle, stride, rotated):
    cfg, az, alt = make_config(stride=stride, rotated=rotated)
    o = oracle.OracleLidar(cfg, az, alt)
    cb = LidarCallback(cfg, az, alt)
    assert cb.expected_size == o.expected_size
    for a, b in zip(cb.tables(), o.tables()):
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    pk = make_stream(cfg, 5, seed=11 + stride, empty_frames=(2,))
    pk[9, 0] = 3  # a packet of another type in the middle of a frame
    want = o.decode(pk)
    got = frames_as_dicts(cb.DecodePackets(pk))           # one call
    assert_frames_equal(got, want)
    cb2 = LidarCallback(cfg, az, alt)                      # packet by packet, like the reference's caller
    got2 = []
    for i in range(pk.shape[0]):
        f = cb2.DecodePacket(pk[i])
        if f is not None:
            got2 += frames_as_dicts([f])
    assert_frames_equal(got2, want)
    assert cb2.DecodePacket(pk[0, :-1].copy()) is None     # wrong size: ignored


@pytest.mark.gpu
def test_gpu_decoder_full_size_and_hand_over_to_ndt(oracle):
    """2048 x 128 sensor with channelStride 4 (the reference's Berlin configuration) and a 1024 x 128 one with every
    channel; the popped frame's device points become the NDT source without leaving HBM."""
    for cols, stride in ((2048, 4), (1024, 1)):
        cfg, az, alt = make_config(cols=cols, pixels=128, stride=stride)
        o = oracle.OracleLidar(cfg, az, alt)
        cb = LidarCallback(cfg, az, alt)
        pk = make_stream(cfg, 2, seed=cols)
        want = o.decode(pk)
        assert cb.push(pk) == 2
        f0 = cb.pop()
        assert_frames_equal(frames_as_dicts([f0]), want[:1])
        f1 = cb.pop()
        assert_frames_equal(frames_as_dicts([f1]), want[1:])
        launches, us = cb.stats()
        assert launches == 2 and 0 < us < 5000
    # hand-over: device frame -> ndtb_set_source_device; derivatives equal those of the same points uploaded from the host
    from slam_sam_b200.registration import NormalDistributionsTransform

    a = f1.arrays()
    cloud = np.stack([a["x"], a["y"], a["z"]], 1)
    rng = np.random.default_rng(0)
    target = np.concatenate([cloud + rng.normal(0, 0.05, cloud.shape).astype(np.float32) for _ in range(4)])
    p = np.array([0.05, -0.02, 0.01, 0.001, -0.002, 0.003])
    res = []
    for mode in ("host", "device"):
        ndt = NormalDistributionsTransform()
        ndt.setResolution(2.0)
        ndt.setInputTarget(target)
        if mode == "host":
            ndt.setInputSource(cloud)
        else:
            ptr, n = f1.device_points()
            assert n == cloud.shape[0]
            rc = ndt.handle.L.ndtb_set_source_device(ndt.handle.h, ptr, n)
            assert rc == 0
        res.append(ndt.computeDerivatives(p))
    assert res[0][0] == res[1][0] and np.array_equal(res[0][1], res[1][1]) and np.array_equal(res[0][2], res[1][2])
    assert cb.pop() is None  # nothing else was closed; the popped frame went back to the pool


def edge_streams(cfg):
    """Packet streams around the corners of the frame state machine (src/lidarcallback.cpp:649-671, 856-862)."""
    out = {}
    out["frame_id_wraps_65535_to_0"] = make_stream(cfg, 3, seed=21, first_frame_id=65534)
    out["first_frame_id_is_0"] = make_stream(cfg, 2, seed=22, first_frame_id=0)
    a = make_stream(cfg, 2, seed=23, first_frame_id=40)
    n = cfg.columns_per_frame // cfg.columns_per_packet
    inter = np.concatenate([a[:n // 2], a[n:n + n // 2], a[n // 2:n], a[n + n // 2:]])  # A B A B: every change of id closes a frame
    out["interleaved_frame_ids"] = inter
    late = make_stream(cfg, 3, seed=24, first_frame_id=50, bad_columns=False)
    blk = 12 + 12 * cfg.pixels_per_column
    for c in range(cfg.columns_per_packet):  # frame 1's first packet has no returns: its first point comes from a later packet -> timedelta 0
        o = 32 + c * blk + 12
        late[n, o:o + 12 * cfg.pixels_per_column].reshape(-1, 12)[:, :3] = 0
    out["first_point_in_a_later_packet"] = late
    return out


def test_oracle_edge_streams(oracle):
    cfg, az, alt = make_config()
    s = edge_streams(cfg)
    f = oracle.OracleLidar(cfg, az, alt).decode(s["frame_id_wraps_65535_to_0"])
    assert [x["frame_id"] for x in f] == [65534, 65535, 0]
    f = oracle.OracleLidar(cfg, az, alt).decode(s["first_frame_id_is_0"])
    assert [x["frame_id"] for x in f] == [0, 1]
    f = oracle.OracleLidar(cfg, az, alt).decode(s["interleaved_frame_ids"])
    assert [x["frame_id"] for x in f] == [40, 41, 40, 41]
    f = oracle.OracleLidar(cfg, az, alt).decode(s["first_point_in_a_later_packet"])
    assert f[1]["interframe_timedelta"] == 0.0 and f[2]["interframe_timedelta"] > 0.0


@pytest.mark.gpu
def test_gpu_decoder_edge_streams(oracle):
    cfg, az, alt = make_config(stride=2)
    for name, pk in edge_streams(cfg).items():
        want = oracle.OracleLidar(cfg, az, alt).decode(pk)
        got = frames_as_dicts(LidarCallback(cfg, az, alt).DecodePackets(pk))
        assert len(want) >= 2, name
        assert_frames_equal(got, want)


@pytest.mark.gpu
def test_gpu_decoder_buffers_and_limits(oracle):
    import torch

    cfg, az, alt = make_config(stride=1)
    pk = make_stream(cfg, 2, seed=31)
    want = oracle.OracleLidar(cfg, az, alt).decode(pk)
    # pinned host memory (DMA straight from the caller's buffer) and a strided buffer (pitch > packet size)
    pinned = torch.from_numpy(pk).pin_memory()
    cb = LidarCallback(cfg, az, alt)
    assert cb.push(pinned.data_ptr(), pk.shape[0], pk.shape[1]) == 2
    got = []
    for _ in range(2):
        f = cb.pop(); got += frames_as_dicts([f])
    assert_frames_equal(got, want)
    wide = np.zeros((pk.shape[0], pk.shape[1] + 48), np.uint8)
    wide[:, :pk.shape[1]] = pk
    cb2 = LidarCallback(cfg, az, alt)
    assert cb2.push(wide.ctypes.data, pk.shape[0], pk.shape[1], pitch=wide.shape[1]) == 2
    got2 = []
    for _ in range(2):
        f = cb2.pop(); got2 += frames_as_dicts([f])
    assert_frames_equal(got2, want)
    # a "frame" longer than two revolutions: the extra packets are refused loudly, the decoder stays usable
    n = cfg.columns_per_frame // cfg.columns_per_packet
    same = np.concatenate([pk[:n]] * 3)
    cb3 = LidarCallback(cfg, az, alt)
    with pytest.raises(ndtb.NdtbError) as e:
        cb3.push(same)
    assert e.value.code == ndtb.NDTB_ERR_OVERFLOW
    assert cb3.push(pk[n:]) == 2 and cb3.pop().packets == 2 * n + 2 and cb3.pop().packets == n
    # nothing closed yet -> nothing to pop; empty push is a no-op
    cb4 = LidarCallback(cfg, az, alt)
    assert cb4.push(pk[:3]) == 0 and cb4.pop() is None and cb4.push(pk[:0]) == 0


# ------------------------------------------------------------------------------------------------------ LEGACY profile
def make_legacy_stream(cfg, frame_cols, seed, first_frame_id=11, t0_ns=1_762_300_000_000_000_000, empty_frames=(), bad_columns=True):
    """LEGACY packets (DecodePacketLegacy, src/lidarcallback.cpp:376-629) of frames with `frame_cols[f]` columns each -- the frame id
    travels in every column header, so frames whose length is not a multiple of the packet end in the middle of a packet --
    followed by the columns that close the last frame and pad the final packet."""
    from slam_sam_b200.lidar import encode_legacy_packets

    rng = np.random.default_rng(seed)
    W, Hh, cpp = cfg.columns_per_frame, cfg.pixels_per_column, cfg.columns_per_packet
    total = sum(frame_cols)
    pad = (-total) % cpp or cpp
    fid = np.concatenate([np.full(n, (first_frame_id + f) & 0xffff) for f, n in enumerate(list(frame_cols) + [pad])])
    ncol = total + pad
    rng_mm = rng.integers(0, 1 << 20, (ncol, Hh)).astype(np.uint32)
    close = rng.random((ncol, Hh))
    rng_mm[close < 0.3] = rng.integers(0, 9000, int((close < 0.3).sum()))
    rng_mm[close > 0.9] = 0
    start = 0
    for f, n in enumerate(frame_cols):
        if f in empty_frames:
            rng_mm[start:start + n] = 0
        start += n
    refl = rng.integers(0, 256, (ncol, Hh)); sig = rng.integers(0, 65536, (ncol, Hh)); nir = rng.integers(0, 65536, (ncol, Hh))
    ts = t0_ns + np.arange(ncol, dtype=np.uint64) * np.uint64(48_828)
    m_ids = np.arange(ncol) % W
    status = np.ones(ncol, np.uint8)
    if bad_columns:
        status[rng.integers(0, total, 5)] = 0                     # block status != 0xFFFFFFFF -> column skipped (:435-441)
        m_ids = m_ids.copy(); m_ids[rng.integers(0, total, 4)] = W + 9   # invalid measurement id -> the column leaves no trace (:407-410)
    return encode_legacy_packets(rng_mm, refl, sig, nir, ts, fid, cpp, m_ids=m_ids, status=status, rng=rng)


def test_oracle_legacy_state_machine(oracle):
    """The LEGACY restatement: packet size, frames cut at the column whose id differs, time stamps of the frame scalars."""
    cfg, az, alt = make_config(udp_profile_legacy=1)
    o = oracle.OracleLidar(cfg, az, alt)
    assert o.expected_size == 16 * (16 + 12 * 32 + 4)
    pk = make_legacy_stream(cfg, [64, 40, 64], seed=5, bad_columns=False)   # the second frame ends in the middle of a packet
    frames = o.decode(pk)
    assert [f["frame_id"] for f in frames] == [11, 12, 13]
    col_ts = np.fmod((1_762_300_000_000_000_000 + np.arange(64 + 40 + 64 + 8, dtype=np.uint64) * np.uint64(48_828)).astype(np.float64) * 1e-9, 86400.0)
    assert frames[0]["timestamp_end"] == col_ts[63] and frames[1]["timestamp_end"] == col_ts[103] and frames[2]["timestamp_end"] == col_ts[167]
    assert set(np.unique(frames[1]["timestamp_points"])) <= set(col_ts[64:104])
    assert frames[0]["interframe_timedelta"] == 0.0                         # nothing before the first packet
    # frame 1 starts at a packet boundary: prev = latest_timestamp_s at the start of its first packet
    assert frames[1]["interframe_timedelta"] == max(0.0, frames[1]["timestamp"] - col_ts[63])
    # frame 2 starts in the middle of a packet that completed frame 1: prev = latest_timestamp_s at that completion
    assert frames[2]["interframe_timedelta"] == max(0.0, frames[2]["timestamp"] - col_ts[103])
    assert o.decode(pk[:, :-1]) == []                                       # wrong size: every packet ignored


@pytest.mark.gpu
@pytest.mark.parametrize("stride,frame_cols", [(1, [64, 64, 64]), (2, [64, 40, 64, 24, 64]), (4, [50, 70, 64])])
def test_gpu_legacy_decoder_matches_oracle(oracle, stride, frame_cols):
    """LEGACY packets through the device decoder: frames that end at packet boundaries and in the middle of a packet, invalid
    measurement ids, bad block status, an empty frame; one call and packet by packet."""
    cfg, az, alt = make_config(stride=stride, udp_profile_legacy=1)
    o = oracle.OracleLidar(cfg, az, alt)
    pk = make_legacy_stream(cfg, frame_cols, seed=17 + stride, empty_frames=(1,) if len(frame_cols) > 3 else ())
    want = o.decode(pk)
    assert len(want) >= len(frame_cols) - 1 and all(f["numberpoints"] > 50 for f in want)
    cb = LidarCallback(cfg, az, alt)
    assert cb.expected_size == o.expected_size
    got = frames_as_dicts(cb.DecodePackets(pk))
    assert_frames_equal(got, want)
    cb2 = LidarCallback(cfg, az, alt)
    got2 = []
    for i in range(pk.shape[0]):
        f = cb2.DecodePacket(pk[i])
        if f is not None:
            got2 += frames_as_dicts([f])
    assert_frames_equal(got2, want)
