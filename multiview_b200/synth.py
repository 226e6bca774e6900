"""Seeded synthetic rig-calibration problems (BASELINE.json configs 1-5).

The reference ships no sample data, so the bench and the tests build their
inputs here.  Everything is vectorised numpy; the projection below is an independent restatement
of the camera model (reference camera_model/camera_params.cc:260-314,
rig_calibrator/bin/rig_calibrator.cc:322-390) used only to synthesise measurements.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

BLOCK_CAM, BLOCK_DEPTH_TRI, BLOCK_DEPTH_MESH, BLOCK_MESH_TRI, BLOCK_TRI = range(5)


# ------------------------------------------------------------------------------------------
# small vectorised algebra (quaternions stored x, y, z, w as in the 7-vector tx ty tz qx qy qz qw)
# ------------------------------------------------------------------------------------------
def quat_to_mat(q):
    q = q / np.linalg.norm(q, axis=-1, keepdims=True)
    x, y, z, w = q[..., 0], q[..., 1], q[..., 2], q[..., 3]
    m = np.empty(q.shape[:-1] + (3, 3))
    m[..., 0, 0] = 1 - 2 * (y * y + z * z); m[..., 0, 1] = 2 * (x * y - w * z); m[..., 0, 2] = 2 * (x * z + w * y)
    m[..., 1, 0] = 2 * (x * y + w * z); m[..., 1, 1] = 1 - 2 * (x * x + z * z); m[..., 1, 2] = 2 * (y * z - w * x)
    m[..., 2, 0] = 2 * (x * z - w * y); m[..., 2, 1] = 2 * (y * z + w * x); m[..., 2, 2] = 1 - 2 * (x * x + y * y)
    return m


def axis_angle_quat(axis, angle):
    axis = np.asarray(axis, dtype=np.float64)
    axis = axis / np.linalg.norm(axis, axis=-1, keepdims=True)
    half = 0.5 * np.asarray(angle)[..., None]
    return np.concatenate([axis * np.sin(half), np.cos(half)], axis=-1)


def quat_mul(a, b):
    ax, ay, az, aw = a[..., 0], a[..., 1], a[..., 2], a[..., 3]
    bx, by, bz, bw = b[..., 0], b[..., 1], b[..., 2], b[..., 3]
    return np.stack([aw * bx + ax * bw + ay * bz - az * by,
                     aw * by - ax * bz + ay * bw + az * bx,
                     aw * bz + ax * by - ay * bx + az * bw,
                     aw * bw - ax * bx - ay * by - az * bz], axis=-1)


def slerp(alpha, a, b):
    a = a / np.linalg.norm(a, axis=-1, keepdims=True)
    b = b / np.linalg.norm(b, axis=-1, keepdims=True)
    d = np.sum(a * b, axis=-1)
    sgn = np.where(d < 0, -1.0, 1.0)
    ad = np.minimum(np.abs(d), 1.0)
    theta = np.arccos(ad)
    st = np.sin(theta)
    small = st < 1e-12
    st = np.where(small, 1.0, st)
    s0 = np.where(small, 1 - alpha, np.sin((1 - alpha) * theta) / st)
    s1 = np.where(small, alpha, np.sin(alpha * theta) / st) * sgn
    return s0[..., None] * a + s1[..., None] * b


def rpc_num_dist_params(deg):
    return 2 * (deg + 1) * (deg + 2) - 2


def rpc_eval(u, coeffs, deg):
    """compute_rpc (camera_model/rpc_distortion.cc:116-160), vectorised over points."""
    x, y = u[..., 0], u[..., 1]
    vals = [np.zeros_like(x), np.ones_like(x), np.zeros_like(x), np.ones_like(x)]
    ci = 0
    for count in range(4):
        start = 1 if count in (1, 3) else 0
        for d in range(start, deg + 1):
            for i in range(d + 1):
                vals[count] = vals[count] + coeffs[ci] * x ** (d - i) * y ** i
                ci += 1
    return np.stack([vals[0] / vals[1], vals[2] / vals[3]], axis=-1)


@dataclass
class Sensor:
    name: str
    width: float
    height: float
    focal: float
    center: tuple
    dist: np.ndarray  # length 0 | 1 | 4 | 5 | 2n (RPC)

    @property
    def nd(self):
        return int(len(self.dist))


def distort(s: Sensor, f, c, dist, u):
    """Convert<UNDISTORTED_C, DISTORTED> (camera_params.cc:260-314, 412-415)."""
    h = np.array([s.width / 2.0, s.height / 2.0])
    nd = len(dist)
    if nd == 0:
        return u + c
    if nd == 1:
        n = u / f
        ru = np.linalg.norm(n, axis=-1)
        rd = np.arctan(ru * 2 * np.tan(dist[0] / 2)) / dist[0]
        conv = np.where(ru > 1e-5, rd / np.where(ru > 1e-5, ru, 1.0), 1.0)
        return c + conv[..., None] * n * f
    if nd in (4, 5):
        k1, k2, p1, p2 = dist[:4]
        k3 = dist[4] if nd == 5 else 0.0
        n = u / f
        x, y = n[..., 0], n[..., 1]
        r2 = x * x + y * y
        rad = 1 + k1 * r2 + k2 * r2 * r2 + k3 * r2 * r2 * r2
        dx = rad * x + 2 * p1 * x * y + p2 * (r2 + 2 * x * x)
        dy = rad * y + p1 * (r2 + 2 * y * y) + 2 * p2 * x * y
        return np.stack([dx, dy], axis=-1) * f + c
    n = nd // 2
    deg = int(round(np.sqrt(2.0 * n + 5.0) / 2.0 - 1.5))
    return h + rpc_eval(u, dist[:n], deg)


@dataclass
class Problem:
    """Flat description of one optimisation pass, exactly what the C ABI takes."""
    sensors: list
    no_rig: bool
    affine_depth: bool
    robust_threshold: float
    depth_tri_weight: float
    depth_mesh_weight: float
    # parameters (all float64, mutated in place by a solve)
    world_to_ref: np.ndarray
    ref_stamps: np.ndarray
    ref_to_cam: np.ndarray
    offsets: np.ndarray
    offset_lo: np.ndarray
    offset_hi: np.ndarray
    focal: np.ndarray
    center: np.ndarray
    distortion: np.ndarray
    depth_to_image: np.ndarray
    depth_scale: np.ndarray
    world_to_cam: np.ndarray
    xyz: np.ndarray
    # cameras
    cam_sensor: np.ndarray
    cam_beg: np.ndarray
    cam_end: np.ndarray
    cam_stamp: np.ndarray
    # observations (pid ascending, cid ascending)
    pid: np.ndarray
    cid: np.ndarray
    pix: np.ndarray
    has_depth: np.ndarray | None = None
    depth_xyz: np.ndarray | None = None
    has_mesh: np.ndarray | None = None
    mesh_xyz: np.ndarray | None = None
    # priors: kind -> (pid, ref_xyz, weight, robust_threshold)
    priors: dict = field(default_factory=dict)
    # float mask
    float_focal: np.ndarray | None = None
    float_center: np.ndarray | None = None
    float_dist: np.ndarray | None = None
    float_rig: np.ndarray | None = None
    float_d2i: np.ndarray | None = None
    float_poses: np.ndarray | None = None
    float_offsets: bool = False
    float_scale: bool = False
    # ground truth for tests (not passed to the solver)
    truth: dict = field(default_factory=dict)
    name: str = ""

    @property
    def num_obs(self):
        return int(len(self.pid))

    @property
    def num_landmarks(self):
        return int(len(self.xyz) // 3)

    def param_arrays(self):
        return dict(world_to_ref=self.world_to_ref, ref_to_cam=self.ref_to_cam, offsets=self.offsets,
                    focal=self.focal, center=self.center, distortion=self.distortion,
                    depth_to_image=self.depth_to_image, depth_scale=self.depth_scale,
                    world_to_cam=self.world_to_cam, xyz=self.xyz)

    def snapshot(self):
        return {k: v.copy() for k, v in self.param_arrays().items()}

    def restore(self, snap):
        for k, v in self.param_arrays().items():
            v[...] = snap[k]

    def keep_observations(self, keep: np.ndarray) -> "Problem":
        """Copy of the problem with only the observations where `keep` is true (tracks, cameras and parameters unchanged)."""
        import copy
        keep = np.asarray(keep, bool)
        q = copy.copy(self)
        q.pid = self.pid[keep].copy(); q.cid = self.cid[keep].copy()
        q.pix = self.pix.reshape(-1, 2)[keep].reshape(-1).copy()
        for nm, w in (("has_depth", 1), ("depth_xyz", 3), ("has_mesh", 1), ("mesh_xyz", 3)):
            a = getattr(self, nm)
            setattr(q, nm, None if a is None else a.reshape(-1, w)[keep].reshape(-1).copy())
        for nm in ("world_to_ref", "ref_to_cam", "offsets", "focal", "center", "distortion", "depth_to_image",
                   "depth_scale", "world_to_cam", "xyz"):
            setattr(q, nm, getattr(self, nm).copy())
        return q

    def shard(self, rank: int, world: int, locality: bool = True) -> "Problem":
        """Landmark shard for rank `rank` of `world`: ranges of tracks balanced by observation count; cameras and all
        non-landmark parameters are replicated.  With `locality` the tracks are first ordered by the
        earliest image that sees them, so that a rank's tracks touch ~1/world of the images (pass A then works on long
        per-image runs instead of `world` times shorter ones, and the rank's Schur blocks form one band)."""
        import copy
        P = self.num_landmarks
        counts = np.bincount(self.pid, minlength=P)
        start = np.concatenate([[0], np.cumsum(counts)])
        if locality and P > 0:
            first_cid = np.full(P, np.iinfo(np.int32).max, np.int64)
            np.minimum.at(first_cid, self.pid, self.cid)
            order = np.argsort(first_cid, kind="stable")
        else:
            order = np.arange(P)
        cum = np.concatenate([[0], np.cumsum(counts[order])])
        total = cum[-1]
        bounds = [int(np.searchsorted(cum, total * r / world, side="left")) for r in range(world)] + [P]
        bounds[0] = 0
        lo, hi = bounds[rank], bounds[rank + 1]
        tracks = order[lo:hi]                                   # global pid of the shard's tracks, in shard order
        n_t = counts[tracks]
        # observation indices of the shard, track-major in shard order (cid ascending inside a track is preserved)
        obs = (np.repeat(start[tracks], n_t) + (np.arange(int(n_t.sum())) - np.repeat(np.cumsum(n_t) - n_t, n_t))).astype(np.int64)
        newpid = np.full(P, -1, np.int64); newpid[tracks] = np.arange(hi - lo)
        q = copy.copy(self)
        q.xyz = self.xyz.reshape(-1, 3)[tracks].reshape(-1).copy()
        q.pid = np.repeat(np.arange(hi - lo), n_t).astype(np.int32)
        q.cid = self.cid[obs].copy()
        q.pix = self.pix.reshape(-1, 2)[obs].reshape(-1).copy()
        for nm, w in (("has_depth", 1), ("depth_xyz", 3), ("has_mesh", 1), ("mesh_xyz", 3)):
            a = getattr(self, nm)
            setattr(q, nm, None if a is None else a.reshape(-1, w)[obs].reshape(-1).copy())
        q.priors = {}
        for kind, (ppid, ref, w, th) in self.priors.items():
            np_ = newpid[ppid]
            sel = np.flatnonzero(np_ >= 0)
            sel = sel[np.argsort(np_[sel], kind="stable")]      # priors follow the shard's track order
            q.priors[kind] = (np_[sel].astype(np.int32), ref.reshape(-1, 3)[sel].reshape(-1).copy(), w, th)
        # replicated parameter blocks get their own copies so ranks do not alias in one process
        for nm in ("world_to_ref", "ref_to_cam", "offsets", "focal", "center", "distortion", "depth_to_image",
                   "depth_scale", "world_to_cam"):
            setattr(q, nm, getattr(self, nm).copy())
        q.truth = dict(self.truth, pid_range=(lo, hi), tracks=tracks)
        return q


def _pose7(t, q):
    return np.concatenate([t, q], axis=-1)


def world_to_cam_all(pr: Problem, world_to_ref=None, ref_to_cam=None, offsets=None, world_to_cam=None):
    """R (C,3,3), t (C,3) of every camera for given parameters (rig_calibrator.cc:362-390)."""
    w2r = (pr.world_to_ref if world_to_ref is None else world_to_ref).reshape(-1, 7)
    r2c = (pr.ref_to_cam if ref_to_cam is None else ref_to_cam).reshape(-1, 7)
    off = pr.offsets if offsets is None else offsets
    if pr.no_rig:
        w2c = (pr.world_to_cam if world_to_cam is None else world_to_cam).reshape(-1, 7)
        return quat_to_mat(w2c[:, 3:]), w2c[:, :3].copy()
    b, e, s = pr.cam_beg, pr.cam_end, pr.cam_sensor
    bs, es = pr.ref_stamps[b], pr.ref_stamps[e]
    direct = bs == es
    den = np.where(direct, 1.0, es - bs)
    alpha = np.where(direct, 0.0, ((pr.cam_stamp - bs) - off[s]) / den)
    qi = slerp(alpha, w2r[b, 3:], w2r[e, 3:])
    ti = (1 - alpha)[:, None] * w2r[b, :3] + alpha[:, None] * w2r[e, :3]
    Ri = quat_to_mat(qi)
    Rrc = quat_to_mat(r2c[s, 3:])
    R = np.where(direct[:, None, None], quat_to_mat(w2r[b, 3:]), Rrc @ Ri)
    t = np.where(direct[:, None], w2r[b, :3], np.einsum("cij,cj->ci", Rrc, ti) + r2c[s, :3])
    return R, t


def project(pr: Problem, R, t, xyz, pid, cid):
    X = xyz.reshape(-1, 3)[pid]
    Xc = np.einsum("nij,nj->ni", R[cid], X) + t[cid]
    out = np.empty((len(pid), 2))
    sens = pr.cam_sensor[cid]
    doff = np.concatenate([[0], np.cumsum([s.nd for s in pr.sensors])])
    for ti, s in enumerate(pr.sensors):
        m = sens == ti
        if not m.any():
            continue
        f = pr.focal[ti]
        u = f * (Xc[m, :2] / Xc[m, 2:3])
        out[m] = distort(s, f, pr.center[2 * ti:2 * ti + 2], pr.distortion[doff[ti]:doff[ti + 1]], u)
    return out, Xc


def default_sensors(kind: str):
    if kind == "pinhole":
        return [Sensor("cam", 640, 480, 600.0, (320.0, 240.0), np.zeros(0))]
    if kind == "rig3":
        return [Sensor("nav_cam", 1280, 960, 608.0, (640.5, 480.2), np.array([0.998])),
                Sensor("haz_cam", 224, 171, 210.0, (112.3, 85.1), np.array([-0.25, 0.08, 0.0012, -0.0008])),
                Sensor("sci_cam", 1336, 1002, 1000.0, (668.4, 500.9), np.array([0.06, -0.11, 0.0005, 0.0007]))]
    if kind == "rig3b":  # radtan with k3 on the third sensor
        s = default_sensors("rig3")
        s[2] = Sensor("sci_cam", 1336, 1002, 1000.0, (668.4, 500.9), np.array([0.06, -0.11, 0.0005, 0.0007, 0.02]))
        return s
    if kind == "rpc":
        deg = 3
        n = rpc_num_dist_params(deg)
        L = (deg + 1) * (deg + 2) // 2
        rng = np.random.default_rng(77)
        co = np.zeros(n)
        w, h, cx, cy = 1280.0, 960.0, 642.0, 478.5
        # num_x | den_x | num_y | den_y ; identity + offset (c - half) + mild higher-order terms
        co[0] = cx - w / 2; co[1] = 1.0
        co[L + (L - 1)] = cy - h / 2; co[L + (L - 1) + 2] = 1.0
        scale = np.array([1e-6] * 3 + [1e-9] * 4)  # degree 2 then degree 3 monomials (pixels^-k)
        co[3:L] += rng.normal(0, 1, L - 3) * scale
        co[L + (L - 1) + 3:2 * L + (L - 1)] += rng.normal(0, 1, L - 3) * scale
        co[L + 2:L + L - 1] += rng.normal(0, 1, L - 3) * scale * 1e-3   # den_x degree>=2
        co[2 * L + (L - 1) + 2:] += rng.normal(0, 1, L - 3) * scale * 1e-3
        und = co.copy()
        return [Sensor("rpc_cam", w, h, 900.0, (cx, cy), np.concatenate([co, und]))]
    raise ValueError(kind)


def reference_offset_bounds(ref_stamps, cam_sensor, cam_beg, cam_end, cam_stamp, offsets, max_change=1.0):
    """Bounds the reference puts on a floated ref_to_sensor_timestamp_offset (dense_map_utils.cc:1607-1652, applied at
    rig_calibrator.cc:1742-1745): every image must stay inside its bracket, the offset may move by at most
    --timestamp_offsets_max_change, and both ends are pulled in by 1e-5."""
    T = len(offsets)
    lo = np.full(T, -1.0e100); hi = np.full(T, 1.0e100)
    for t in range(T):
        sel = (cam_sensor == t) & (cam_beg != cam_end)
        if sel.any():
            lo[t] = max(lo[t], float(np.max(cam_stamp[sel] - ref_stamps[cam_end[sel]])))
            hi[t] = min(hi[t], float(np.min(cam_stamp[sel] - ref_stamps[cam_beg[sel]])))
    lo = np.maximum(lo, offsets - max_change) + 1.0e-5
    hi = np.minimum(hi, offsets + max_change) - 1.0e-5
    return lo, hi


def make_problem(*, seed=1, sensors="pinhole", num_ref=50, num_landmarks=10000, obs_per_track=4,
                 pixel_noise=0.5, pose_noise=(0.02, 0.2), rig_noise=(0.005, 0.1), xyz_noise=0.02,
                 outlier_frac=0.0, no_rig=False, depth=False, mesh=False, affine_depth=False,
                 depth_tri_weight=0.0, depth_mesh_weight=0.0, mesh_tri_weight=0.0, tri_weight=0.0,
                 tri_robust_threshold=0.1, robust_threshold=3.0,
                 float_poses=True, float_rig=True, float_intrinsics=(), float_d2i=False, float_scale=False,
                 float_offsets=False, intr_noise=0.0, ragged=False, name="") -> Problem:
    """Build one synthetic pass.  Geometry: the rig translates slowly along +x with a gentle yaw,
    landmarks are sampled in front of a randomly chosen camera and observed by `obs_per_track`
    time-adjacent cameras (banded co-visibility)."""
    rng = np.random.default_rng(seed)
    S = default_sensors(sensors) if isinstance(sensors, str) else sensors
    T = len(S)
    R = int(num_ref)
    # --- ground-truth trajectory (world -> ref) ---
    idx = np.arange(R)
    pos = np.stack([0.06 * idx, 0.15 * np.sin(0.05 * idx), 0.05 * np.cos(0.03 * idx)], axis=-1)  # ref centre in world
    yaw = 0.25 * np.sin(0.02 * idx) + 0.004 * idx * 0
    q_c2w = quat_mul(axis_angle_quat(np.tile([0.0, 1.0, 0.0], (R, 1)), yaw),
                     axis_angle_quat(np.tile([1.0, 0.0, 0.0], (R, 1)), 0.05 * np.sin(0.07 * idx)))
    q_w2r = q_c2w * np.array([-1.0, -1.0, -1.0, 1.0])
    Rw2r = quat_to_mat(q_w2r)
    t_w2r = -np.einsum("rij,rj->ri", Rw2r, pos)
    w2r_true = _pose7(t_w2r, q_w2r)
    ref_stamps = 1000.0 + 0.5 * idx  # 2 Hz
    # --- rig ---
    r2c_true = np.zeros((T, 7)); r2c_true[:, 6] = 1.0
    for t in range(1, T):
        ax = rng.normal(size=3)
        r2c_true[t, 3:] = axis_angle_quat(ax, np.deg2rad(rng.uniform(2, 5)))
        r2c_true[t, :3] = rng.uniform(-0.1, 0.1, 3)
    offsets_true = np.zeros(T)
    for t in range(1, T):
        offsets_true[t] = rng.uniform(-0.02, 0.02)
    # --- cameras: every ref frame, plus one image per non-ref sensor inside each bracket ---
    cs, cb, ce, ct = [np.zeros(R, np.int32)], [idx.astype(np.int32)], [idx.astype(np.int32)], [ref_stamps.copy()]
    for t in range(1, T):
        nb = R - 1
        frac = rng.uniform(0.15, 0.85, nb)
        st = ref_stamps[:-1] + frac * 0.5 + offsets_true[t]
        # avoid ties with ref stamps and across sensors
        cs.append(np.full(nb, t, np.int32)); cb.append(np.arange(nb, dtype=np.int32)); ce.append(np.arange(1, R, dtype=np.int32)); ct.append(st)
    cam_sensor = np.concatenate(cs); cam_beg = np.concatenate(cb); cam_end = np.concatenate(ce); cam_stamp = np.concatenate(ct)
    # order cameras by (ref-clock) time like lookupImages' sort (dense_map_utils.cc:1804)
    order = np.argsort(cam_stamp - offsets_true[cam_sensor], kind="stable")
    cam_sensor, cam_beg, cam_end, cam_stamp = cam_sensor[order], cam_beg[order], cam_end[order], cam_stamp[order]
    C = len(cam_sensor)

    focal = np.array([s.focal for s in S], dtype=np.float64)
    center = np.array([c for s in S for c in s.center], dtype=np.float64)
    distortion = np.concatenate([np.asarray(s.dist, dtype=np.float64) for s in S]) if sum(s.nd for s in S) else np.zeros(0)
    nd2i = 12 if affine_depth else 7
    d2i_true = np.zeros((T, nd2i))
    scale_true = np.ones(T)
    for t in range(T):
        q = axis_angle_quat(rng.normal(size=3), np.deg2rad(1.5))
        tt = rng.uniform(-0.02, 0.02, 3)
        if affine_depth:
            M = np.concatenate([quat_to_mat(q), tt[:, None]], axis=1)
            d2i_true[t] = M.reshape(-1)
        else:
            d2i_true[t] = np.concatenate([tt, q])
        scale_true[t] = 1.0 + 0.01 * t

    world_to_cam_true = np.zeros(0)
    truth_pr = Problem(S, False, affine_depth, robust_threshold, depth_tri_weight, depth_mesh_weight,
                       w2r_true.reshape(-1).copy(), ref_stamps, r2c_true.reshape(-1).copy(), offsets_true.copy(),
                       offsets_true - 1.0, offsets_true + 1.0, focal.copy(), center.copy(), distortion.copy(),
                       d2i_true.reshape(-1).copy(), scale_true.copy(), np.zeros(0), np.zeros(0),
                       cam_sensor, cam_beg, cam_end, cam_stamp, np.zeros(0, np.int32), np.zeros(0, np.int32),
                       np.zeros(0, np.float32))
    Rc, tc = world_to_cam_all(truth_pr)

    # --- landmarks and tracks ---
    P = int(num_landmarks)
    k = int(obs_per_track)
    if ragged:
        klen = rng.integers(2, k + 1, P)
    else:
        klen = np.full(P, k)
    span = min(C, max(k + 2, 2 * k))
    anchor = rng.integers(0, max(1, C - span + 1), P)
    # choose klen distinct cameras among `span` consecutive ones, per landmark (vectorised)
    keys = rng.random((P, span))
    pick = np.argsort(keys, axis=1)[:, :k]
    pick.sort(axis=1)
    mid = anchor + span // 2
    mid = np.minimum(mid, C - 1)
    # sample the landmark in the view cone of the middle camera, narrow enough that neighbours see it
    depth_z = rng.uniform(2.0, 6.0, P)
    half_fov = 0.18
    xn = rng.uniform(-half_fov, half_fov, P); yn = rng.uniform(-half_fov, half_fov, P)
    Xc_mid = np.stack([xn * depth_z, yn * depth_z, depth_z], axis=-1)
    X = np.einsum("pji,pj->pi", Rc[mid], Xc_mid - tc[mid])  # R^T (Xc - t)
    cols = np.arange(k)[None, :]
    valid = cols < klen[:, None]
    pid = np.repeat(np.arange(P, dtype=np.int32), k).reshape(P, k)[valid]
    cid = (anchor[:, None] + pick)[valid].astype(np.int32)
    truth_pr.xyz = X.reshape(-1).copy()
    pix_true, Xc = project(truth_pr, Rc, tc, truth_pr.xyz, pid, cid)
    N = len(pid)
    pix = pix_true + rng.normal(0, pixel_noise, (N, 2))
    if outlier_frac > 0:
        out = rng.random(N) < outlier_frac
        wh = np.array([[S[t].width, S[t].height] for t in cam_sensor[cid]])
        pix[out] = rng.random((int(out.sum()), 2)) * wh[out]
    pix32 = pix.astype(np.float32).reshape(-1)

    has_depth = depth_xyz = has_mesh = mesh_xyz = None
    priors = {}
    if depth:
        # depth clouds exist for sensor 1 (haz_cam) only: measured point in depth-sensor coordinates,
        # inverse of  Xc = scale * L * m + t  (rig_calibrator.cc:557-569)
        hd = (cam_sensor[cid] == min(1, T - 1)) & (rng.random(N) < 0.8)
        tsel = cam_sensor[cid]
        if affine_depth:
            Lm = d2i_true.reshape(T, 3, 4)[:, :, :3]; tt = d2i_true.reshape(T, 3, 4)[:, :, 3]
        else:
            Lm = quat_to_mat(d2i_true[:, 3:]); tt = d2i_true[:, :3]
        Linv = np.linalg.inv(Lm * scale_true[:, None, None])
        m = np.einsum("nij,nj->ni", Linv[tsel], Xc - tt[tsel]) + rng.normal(0, 0.005, (N, 3))
        m[~hd] = 0
        has_depth = hd.astype(np.uint8); depth_xyz = m.astype(np.float32).reshape(-1)
    if mesh:
        hm = rng.random(N) < 0.7
        mx = X[pid] + rng.normal(0, 0.01, (N, 3))
        mx = mx.astype(np.float32).astype(np.float64)  # BVH hits are float32
        has_mesh = hm.astype(np.uint8); mesh_xyz = mx.reshape(-1)
        if mesh_tri_weight > 0:
            sel = rng.random(P) < 0.8
            ppid = np.nonzero(sel)[0].astype(np.int32)
            ref = (X[ppid] + rng.normal(0, 0.01, (len(ppid), 3))).astype(np.float32).astype(np.float64)
            priors[BLOCK_MESH_TRI] = (ppid, ref.reshape(-1), float(mesh_tri_weight), float(robust_threshold))

    # --- perturbed initial guess ---
    def perturb_pose(p7, sig_t, sig_deg):
        p7 = p7.copy()
        n = len(p7)
        p7[:, :3] += rng.normal(0, sig_t, (n, 3))
        dq = axis_angle_quat(rng.normal(size=(n, 3)), np.deg2rad(sig_deg) * rng.normal(size=n))
        p7[:, 3:] = quat_mul(dq, p7[:, 3:])
        return p7
    w2r0 = perturb_pose(w2r_true, *pose_noise) if float_poses else w2r_true.copy()
    r2c0 = r2c_true.copy()
    if float_rig and T > 1:
        r2c0[1:] = perturb_pose(r2c_true[1:], *rig_noise)
    xyz0 = X + rng.normal(0, xyz_noise, X.shape)
    focal0, center0, dist0 = focal.copy(), center.copy(), distortion.copy()
    fm = np.zeros(T, np.uint8); cm = np.zeros(T, np.uint8); dm = np.zeros(T, np.uint8)
    for t in range(T):
        if "focal_length" in float_intrinsics:
            fm[t] = 1; focal0[t] *= 1 + intr_noise * rng.normal()
        if "optical_center" in float_intrinsics:
            cm[t] = 1; center0[2 * t:2 * t + 2] += intr_noise * 100 * rng.normal(size=2)
        if "distortion" in float_intrinsics and S[t].nd > 0:
            dm[t] = 1
    d2i0 = d2i_true.copy(); scale0 = scale_true.copy()
    if float_d2i and not affine_depth:
        d2i0 = perturb_pose(d2i_true, 0.003, 0.1)
    offsets0 = offsets_true.copy()
    if float_offsets:
        offsets0[1:] += rng.normal(0, 0.005, T - 1)
    world_to_cam0 = np.zeros(0)
    if no_rig:
        # each camera's own pose; init from the (perturbed) rig
        tmp = Problem(S, False, affine_depth, robust_threshold, 0, 0, w2r0.reshape(-1), ref_stamps, r2c0.reshape(-1), offsets0,
                      None, None, focal0, center0, dist0, d2i0.reshape(-1), scale0, np.zeros(0), np.zeros(0),
                      cam_sensor, cam_beg, cam_end, cam_stamp, pid, cid, pix32)
        R0, t0 = world_to_cam_all(tmp)
        world_to_cam0 = np.zeros((C, 7))
        world_to_cam0[:, :3] = t0
        # matrix -> quaternion via numpy (robust branchless: use the largest of w,x,y,z)
        world_to_cam0[:, 3:] = _mat_to_quat(R0)
        world_to_cam0 = world_to_cam0.reshape(-1)
    if tri_weight > 0:
        priors[BLOCK_TRI] = (np.arange(P, dtype=np.int32), xyz0.reshape(-1).copy(), float(tri_weight), float(tri_robust_threshold))

    off_lo, off_hi = reference_offset_bounds(ref_stamps, cam_sensor, cam_beg, cam_end, cam_stamp, offsets0)
    pm = np.zeros(T, np.uint8)
    if float_poses:
        pm[:] = 1 if no_rig else 0
        pm[0] = 1
    rm = np.zeros(T, np.uint8)
    if float_rig and not no_rig:
        rm[1:] = 1
    d2im = np.zeros(T, np.uint8)
    if float_d2i:
        d2im[min(1, T - 1)] = 1
    pr = Problem(S, bool(no_rig), bool(affine_depth), float(robust_threshold), float(depth_tri_weight), float(depth_mesh_weight),
                 np.ascontiguousarray(w2r0.reshape(-1)), ref_stamps, np.ascontiguousarray(r2c0.reshape(-1)), offsets0,
                 off_lo, off_hi, focal0, center0, dist0, np.ascontiguousarray(d2i0.reshape(-1)), scale0,
                 world_to_cam0, np.ascontiguousarray(xyz0.reshape(-1)),
                 cam_sensor.astype(np.int32), cam_beg.astype(np.int32), cam_end.astype(np.int32), cam_stamp.astype(np.float64),
                 pid.astype(np.int32), cid.astype(np.int32), pix32, has_depth, depth_xyz, has_mesh, mesh_xyz, priors,
                 fm, cm, dm, rm, d2im, pm, bool(float_offsets), bool(float_scale),
                 truth=dict(world_to_ref=w2r_true.reshape(-1), ref_to_cam=r2c_true.reshape(-1), xyz=X.reshape(-1),
                            focal=focal, center=center, distortion=distortion, offsets=offsets_true), name=name)
    return pr


def _mat_to_quat(Rm):
    n = len(Rm)
    q = np.empty((n, 4))
    tr = Rm[:, 0, 0] + Rm[:, 1, 1] + Rm[:, 2, 2]
    w = np.sqrt(np.maximum(0, 1 + tr)) / 2
    x = np.sqrt(np.maximum(0, 1 + Rm[:, 0, 0] - Rm[:, 1, 1] - Rm[:, 2, 2])) / 2
    y = np.sqrt(np.maximum(0, 1 - Rm[:, 0, 0] + Rm[:, 1, 1] - Rm[:, 2, 2])) / 2
    z = np.sqrt(np.maximum(0, 1 - Rm[:, 0, 0] - Rm[:, 1, 1] + Rm[:, 2, 2])) / 2
    x = np.copysign(x, Rm[:, 2, 1] - Rm[:, 1, 2]); y = np.copysign(y, Rm[:, 0, 2] - Rm[:, 2, 0]); z = np.copysign(z, Rm[:, 1, 0] - Rm[:, 0, 1])
    q[:, 0], q[:, 1], q[:, 2], q[:, 3] = x, y, z, w
    return q / np.linalg.norm(q, axis=1, keepdims=True)


# BASELINE.json configs ------------------------------------------------------------------------
def config(n: int, scale: float = 1.0, seed: int | None = None) -> Problem:
    """The five BASELINE.json configs; `scale` shrinks frame/landmark counts for CPU-sized tests."""
    s = lambda v, lo=4: max(lo, int(round(v * scale)))
    if n == 1:
        return make_problem(seed=seed or 1, sensors="pinhole", num_ref=s(50, 8), num_landmarks=s(10000, 50), obs_per_track=4,
                            float_poses=True, float_rig=False, name="config1_pinhole")
    if n == 2:
        return make_problem(seed=seed or 2, sensors="rig3", num_ref=s(300, 8), num_landmarks=s(60000, 60), obs_per_track=5,
                            outlier_frac=0.02, name="config2_rig3")
    if n == 3:
        return make_problem(seed=seed or 3, sensors="rig3", num_ref=s(100, 8), num_landmarks=s(20000, 60), obs_per_track=5,
                            outlier_frac=0.02, depth=True, mesh=True, depth_tri_weight=1000.0, depth_mesh_weight=50.0,
                            mesh_tri_weight=25.0, float_intrinsics=("focal_length", "optical_center", "distortion"),
                            float_d2i=True, intr_noise=1e-3, name="config3_depth_mesh")
    if n == 4:
        return make_problem(seed=seed or 4, sensors="rpc", num_ref=s(1000, 8), num_landmarks=s(200000, 60), obs_per_track=6,
                            float_rig=False, float_intrinsics=("focal_length", "distortion"), intr_noise=1e-3,
                            name="config4_rpc")
    if n == 5:
        return make_problem(seed=seed or 5, sensors="rig3", num_ref=s(2000, 8), num_landmarks=s(1000000, 60), obs_per_track=8,
                            outlier_frac=0.02, name="config5_scaling")
    raise ValueError(n)
