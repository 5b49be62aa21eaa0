"""Seeded synthetic inputs for the NDT hot path.

* two_plane_scene(): the reference's gtest scene, extern/svn_ndt/test/test_svn_ndt.cpp:44-131
  (two perpendicular 20 m planes on a 0.15 m lattice, target = ground-truth motion + N(0, 0.02)
  noise from std::default_random_engine(1337) / std::normal_distribution -- restated here for
  libstdc++: minstd_rand0 + polar Box-Muller).
* Scene / ouster_scan() / surface_map(): an "OS1-128-shaped" 128 x 1024 ray-cast scan and a dense
  surface-sampled map of the same procedural street scene (BASELINE.json configs 0-4).
All clouds are float32 (N, 4) arrays: x, y, z, pad -- the 16-byte stride of pcl::PointXYZ.
"""
import math

import numpy as np


# ----------------------------------------------------------------------------- SE(3) helpers
def so3_exp(w):
    w = np.asarray(w, np.float64)
    th = np.linalg.norm(w)
    W = np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]])
    if th * th <= np.finfo(np.float64).eps:
        return np.eye(3) + W + 0.5 * W @ W
    K = W / th
    return np.eye(3) + math.sin(th) * K + (2.0 * math.sin(th / 2) ** 2) * (K @ K)


def se3_exp(xi):
    """gtsam::Pose3::Expmap, xi = (omega, v) -> 4x4."""
    xi = np.asarray(xi, np.float64)
    w, v = xi[:3], xi[3:]
    R = so3_exp(w)
    th2 = float(w @ w)
    if th2 > np.finfo(np.float64).eps:
        wxv = np.cross(w, v)
        t = (wxv - R @ wxv + w * float(w @ v)) / th2
    else:
        t = v
    T = np.eye(4)
    T[:3, :3] = R
    T[:3, 3] = t
    return T


def rot_x(a):
    c, s = math.cos(a), math.sin(a)
    return np.array([[1, 0, 0], [0, c, -s], [0, s, c]], np.float64)


def rot_y(a):
    c, s = math.cos(a), math.sin(a)
    return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]], np.float64)


def rot_z(a):
    c, s = math.cos(a), math.sin(a)
    return np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]], np.float64)


def pose_matrix(xyz, rpy):
    """T = Trans * Rx(roll) * Ry(pitch) * Rz(yaw) -- the NDT parameterisation (ndt_omp.h:205-211)."""
    T = np.eye(4)
    T[:3, :3] = rot_x(rpy[0]) @ rot_y(rpy[1]) @ rot_z(rpy[2])
    T[:3, 3] = xyz
    return T


def pad4(xyz):
    out = np.zeros((xyz.shape[0], 4), np.float32)
    out[:, :3] = xyz
    return out


# ----------------------------------------------------------------------------- gtest scene
class _LibstdcxxNormal:
    """std::normal_distribution<double> over std::default_random_engine (minstd_rand0), libstdc++."""

    def __init__(self, seed, mean, std):
        self.x = seed % 2147483647 or 1
        self.mean, self.std = mean, std
        self.saved = None

    def _next(self):
        self.x = (self.x * 16807) % 2147483647
        return self.x

    def _canonical(self):
        R = 2147483646.0
        s = (self._next() - 1) + (self._next() - 1) * R
        r = s / (R * R)
        return r if r < 1.0 else math.nextafter(1.0, 0.0)

    def __call__(self):
        if self.saved is not None:
            ret, self.saved = self.saved, None
        else:
            while True:
                x = 2.0 * self._canonical() - 1.0
                y = 2.0 * self._canonical() - 1.0
                r2 = x * x + y * y
                if not (r2 > 1.0 or r2 == 0.0):
                    break
            mult = math.sqrt(-2 * math.log(r2) / r2)
            self.saved = x * mult
            ret = y * mult
        return ret * self.std + self.mean


def two_plane_scene(noise_std=0.02, seed=1337):
    """Returns dict(source, target, gt (4x4 f64), guess (4x4 f64)) per test_svn_ndt.cpp:44-131."""
    vals = []
    x = -10.0
    while x <= 10.0:
        vals.append(x)
        x += 0.15
    v = np.array(vals)
    n = len(v)
    src = np.zeros((2 * n * n, 3), np.float64)
    src[: n * n, 0] = np.repeat(v, n)
    src[: n * n, 1] = np.tile(v, n)
    src[n * n:, 0] = np.repeat(v, n)
    src[n * n:, 2] = np.tile(v, n)
    src = src.astype(np.float32)
    gt = np.eye(4)
    gt[:3, :3] = rot_z(0.2618) @ rot_y(0.0873)
    gt[:3, 3] = (0.5, 0.0, 0.3)
    noise = _LibstdcxxNormal(seed, 0.0, noise_std)
    p = src.astype(np.float64) @ gt[:3, :3].T + gt[:3, 3]
    nz = np.array([noise() for _ in range(p.size)]).reshape(p.shape)
    tgt = (p + nz).astype(np.float32)
    delta = np.array([0.05, -0.02, 0.04, 0.02, -0.01, 0.03])
    guess = gt @ se3_exp(-delta)
    return dict(source=pad4(src), target=pad4(tgt), gt=gt, guess=guess)


# ----------------------------------------------------------------------------- street scene
class Scene:
    """200 m x 200 m ground (z = 0 with +-0.05 m undulation), two facades at y = +-12 m (15 m tall),
    40 boxes and 30 poles (r = 0.15 m), all seeded."""

    def __init__(self, seed=1337, half=100.0):
        rng = np.random.Generator(np.random.PCG64(seed))
        self.half = half
        nb, npole = 40, 30
        c = np.stack([rng.uniform(-half * 0.9, half * 0.9, nb), rng.uniform(-10.5, 10.5, nb)], 1)
        s = rng.uniform(1.0, 4.0, (nb, 3))
        self.box_lo = np.concatenate([c - s[:, :2] / 2, np.zeros((nb, 1))], 1)
        self.box_hi = np.concatenate([c + s[:, :2] / 2, s[:, 2:3]], 1)
        # pilasters every 8 m on both facades (1 m wide, 0.6 m deep, full height): without relief a
        # straight street leaves the along-track translation unobservable and NDT's clamped Newton
        # step (ndt_omp_impl.hpp:963-965, no line search) oscillates instead of converging
        rng2 = np.random.Generator(np.random.PCG64(seed + 1))
        for side in (-1.0, 1.0):
            px = -half + 3.0 + np.cumsum(rng2.uniform(5.0, 9.0, int(2 * half / 5)))
            px = px[px < half - 3.0]
            w = rng2.uniform(0.8, 2.0, px.shape) / 2
            d = rng2.uniform(0.4, 1.5, px.shape)
            y0 = 12.0 - d if side > 0 else np.full_like(px, -12.0)
            y1 = np.full_like(px, 12.0) if side > 0 else -12.0 + d
            lo = np.stack([px - w, y0, np.zeros_like(px)], 1)
            hi = np.stack([px + w, y1, rng2.uniform(6.0, 15.0, px.shape)], 1)
            self.box_lo = np.concatenate([self.box_lo, lo], 0)
            self.box_hi = np.concatenate([self.box_hi, hi], 0)
        self.pole_c = np.stack([rng.uniform(-half * 0.9, half * 0.9, npole), rng.choice([-8.0, 8.0], npole) + rng.uniform(-0.5, 0.5, npole)], 1)
        self.pole_r = 0.15
        self.pole_h = rng.uniform(4.0, 8.0, npole)
        self.facade_y = 12.0
        self.facade_h = 15.0

    @staticmethod
    def ground_z(x, y):
        return 0.05 * np.sin(0.35 * x) * np.cos(0.27 * y)

    def cast(self, o, d, max_range=120.0):
        """o (3,), d (M,3) unit -> range (M,) with inf where nothing is hit within max_range."""
        M = d.shape[0]
        best = np.full(M, np.inf)
        with np.errstate(divide="ignore", invalid="ignore"):
            # ground: plane z = 0, then one fixed-point refinement onto the undulating surface
            t = np.where(d[:, 2] < -1e-9, -o[2] / d[:, 2], np.inf)
            for _ in range(2):
                gx, gy = o[0] + t * d[:, 0], o[1] + t * d[:, 1]
                gz = self.ground_z(np.nan_to_num(gx, posinf=0, neginf=0), np.nan_to_num(gy, posinf=0, neginf=0))
                t = np.where(d[:, 2] < -1e-9, (gz - o[2]) / d[:, 2], np.inf)
            ok = (np.abs(o[0] + t * d[:, 0]) <= self.half) & (np.abs(o[1] + t * d[:, 1]) <= self.half) & (t > 0)
            best = np.where(ok & (t < best), t, best)
            # facades y = +-12
            for ysign in (-1.0, 1.0):
                t = (ysign * self.facade_y - o[1]) / d[:, 1]
                z = o[2] + t * d[:, 2]
                x = o[0] + t * d[:, 0]
                ok = (t > 0) & (z >= 0) & (z <= self.facade_h) & (np.abs(x) <= self.half)
                best = np.where(ok & (t < best), t, best)
            # boxes (slab test)
            inv = 1.0 / d
            for lo, hi in zip(self.box_lo, self.box_hi):
                t1 = (lo - o) * inv
                t2 = (hi - o) * inv
                tn = np.nanmax(np.minimum(t1, t2), axis=1)
                tf = np.nanmin(np.maximum(t1, t2), axis=1)
                ok = (tn <= tf) & (tn > 0)
                best = np.where(ok & (tn < best), tn, best)
            # poles (vertical cylinders)
            a = d[:, 0] ** 2 + d[:, 1] ** 2
            for c, h in zip(self.pole_c, self.pole_h):
                oc = o[:2] - c
                b = 2 * (d[:, 0] * oc[0] + d[:, 1] * oc[1])
                cc = oc @ oc - self.pole_r ** 2
                disc = b * b - 4 * a * cc
                t = (-b - np.sqrt(np.maximum(disc, 0))) / (2 * a)
                z = o[2] + t * d[:, 2]
                ok = (disc > 0) & (t > 0) & (z >= 0) & (z <= h)
                best = np.where(ok & (t < best), t, best)
        best[best > max_range] = np.inf
        return best


def ouster_scan(scene, pose, beams=128, cols=1024, seed=0, noise=0.02, max_range=120.0, sensor_height=1.8):
    """One OS1-128-shaped scan (beams x cols rays) in the SENSOR frame, exactly beams*cols points:
    no-return rays are dropped and the cloud is re-padded by resampling returns (keeps N fixed)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    alt = np.deg2rad(np.linspace(-22.5, 22.5, beams))
    az = np.linspace(0, 2 * np.pi, cols, endpoint=False)
    A, Z = np.meshgrid(alt, az, indexing="ij")
    d_s = np.stack([np.cos(A) * np.cos(Z), np.cos(A) * np.sin(Z), np.sin(A)], -1).reshape(-1, 3)
    T = np.array(pose, np.float64)
    T_sensor = T.copy()
    T_sensor[2, 3] += sensor_height
    d_w = d_s @ T_sensor[:3, :3].T
    r = scene.cast(T_sensor[:3, 3], d_w, max_range)
    hit = np.isfinite(r)
    r = r + rng.normal(0.0, noise, r.shape)
    pts = d_s[hit] * r[hit, None]
    n_target = beams * cols
    if pts.shape[0] < n_target:
        extra = rng.integers(0, pts.shape[0], n_target - pts.shape[0])
        jitter = rng.normal(0.0, noise, (extra.shape[0], 3))
        pts = np.concatenate([pts, pts[extra] + jitter], 0)
    # sensor frame -> body frame (sensor sits sensor_height above the body origin)
    pts = pts + np.array([0, 0, sensor_height])
    return pad4(pts.astype(np.float32))


def surface_map(scene, n_points=1_000_000, seed=7, noise=0.02):
    """Dense map of the scene: points sampled on every surface with probability ~ area."""
    rng = np.random.Generator(np.random.PCG64(seed))
    h = scene.half
    areas = [4 * h * h, 2 * 2 * h * scene.facade_h]
    box_sz = scene.box_hi - scene.box_lo
    box_area = 2 * (box_sz[:, 0] * box_sz[:, 2] + box_sz[:, 1] * box_sz[:, 2]) + box_sz[:, 0] * box_sz[:, 1]
    pole_area = 2 * np.pi * scene.pole_r * scene.pole_h
    # boost small structures so they get NDT-grade density
    w = np.array([areas[0], areas[1], 4 * box_area.sum(), 40 * pole_area.sum()])
    cnt = np.floor(w / w.sum() * n_points).astype(int)
    cnt[0] += n_points - cnt.sum()
    out = []
    x, y = rng.uniform(-h, h, cnt[0]), rng.uniform(-h, h, cnt[0])
    out.append(np.stack([x, y, scene.ground_z(x, y)], 1))
    x, z = rng.uniform(-h, h, cnt[1]), rng.uniform(0, scene.facade_h, cnt[1])
    out.append(np.stack([x, rng.choice([-1.0, 1.0], cnt[1]) * scene.facade_y, z], 1))
    bi = rng.choice(len(box_area), cnt[2], p=box_area / box_area.sum())
    u = rng.uniform(0, 1, (cnt[2], 3))
    p = scene.box_lo[bi] + u * box_sz[bi]
    face = rng.integers(0, 5, cnt[2])  # 0,1: x faces; 2,3: y faces; 4: top
    p[face == 0, 0] = scene.box_lo[bi[face == 0], 0]
    p[face == 1, 0] = scene.box_hi[bi[face == 1], 0]
    p[face == 2, 1] = scene.box_lo[bi[face == 2], 1]
    p[face == 3, 1] = scene.box_hi[bi[face == 3], 1]
    p[face == 4, 2] = scene.box_hi[bi[face == 4], 2]
    out.append(p)
    pi = rng.choice(len(pole_area), cnt[3], p=pole_area / pole_area.sum())
    th = rng.uniform(0, 2 * np.pi, cnt[3])
    out.append(np.stack([scene.pole_c[pi, 0] + scene.pole_r * np.cos(th), scene.pole_c[pi, 1] + scene.pole_r * np.sin(th),
                         rng.uniform(0, 1, cnt[3]) * scene.pole_h[pi]], 1))
    pts = np.concatenate(out, 0) + rng.normal(0, noise, (n_points, 3))
    pts = pts[rng.permutation(n_points)]
    return pad4(pts.astype(np.float32))


def trajectory_pose(s, lateral=0.0):
    """Body pose (4x4) at arc-length s metres along the street (x axis), gentle yaw wobble."""
    yaw = 0.03 * math.sin(0.05 * s)
    return pose_matrix((-75.0 + s, lateral + 0.5 * math.sin(0.04 * s), 0.0), (0.0, 0.0, yaw))


def ouster_range_image(scene, pose, beams=128, cols=1024, seed=0, noise=0.02, max_range=120.0, sensor_height=1.8):
    """The same revolution as ouster_scan, but as the sensor reports it: a (cols, beams) image of ranges in millimetres
    (0 = no return) in Ouster's conventions -- measurement azimuth 2*pi*(1 - m/cols), beam 0 on top -- plus the sensor
    description (beam angles, body->lidar transform) with which the reference's LidarCallback lookup tables
    (src/lidarcallback.cpp:283-323) turn the image back into body-frame points.
    Returns (range_mm uint32 (cols, beams), beam_azimuth_deg, beam_altitude_deg, Cb2s (9,), tb2s (3,))."""
    rng = np.random.Generator(np.random.PCG64(seed))
    alt_deg = np.linspace(22.5, -22.5, beams).astype(np.float32)
    az_deg = np.zeros(beams, np.float32)
    m = np.arange(cols)
    theta = 2.0 * np.pi * (1.0 - m / cols)
    Z, A = np.meshgrid(theta, np.deg2rad(alt_deg.astype(np.float64)), indexing="ij")  # (cols, beams)
    d_s = np.stack([np.cos(A) * np.cos(Z), np.cos(A) * np.sin(Z), np.sin(A)], -1).reshape(-1, 3)
    T_sensor = np.array(pose, np.float64)
    T_sensor[2, 3] += sensor_height
    r = scene.cast(T_sensor[:3, 3], d_s @ T_sensor[:3, :3].T, max_range)
    hit = np.isfinite(r)
    r = np.where(hit, r + rng.normal(0.0, noise, r.shape), 0.0)
    range_mm = np.clip(np.rint(r * 1000.0), 0, (1 << 19) - 1).astype(np.uint32).reshape(cols, beams)
    Cb2s = np.eye(3).reshape(9)
    tb2s = np.array([0.0, 0.0, -sensor_height])  # lidar = body - (0, 0, h)
    return range_mm, az_deg, alt_deg, Cb2s, tb2s
