"""Per-track steps around a pass: triangulation, undistortion, outlier flagging.

CPU part: the oracle (oracle/tracks.h) against independent known answers -- cv2.undistortPoints (the very call the
reference makes for radtan, camera_params.cc:348), closed forms at 40 digits (tests/golden/undistort.npz from
tools/make_golden.py) and the exact recovery of noise-free landmarks.
GPU part: mvba_triangulate / mvba_flag_outliers / mvba_flag_exclusion through the C ABI against the oracle on the
same seeded passes: landmarks to 1e-9 relative, inlier flags and counters bit-exact."""
import ctypes as C
import os

import numpy as np
import pytest

from multiview_b200 import _abi, synth

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
dp = C.POINTER(C.c_double)


def _undist(lib, size, nd, f, c, dist, pix):
    out = np.zeros_like(pix)
    c = np.ascontiguousarray(c, np.float64); dist = np.ascontiguousarray(dist, np.float64)
    for i in range(len(pix)):
        p = np.ascontiguousarray(pix[i]); o = np.zeros(2)
        lib.oracle_undistort(C.c_double(size[0]), C.c_double(size[1]), C.c_int(nd), C.c_double(f), c.ctypes.data_as(dp),
                             dist.ctypes.data_as(dp), p.ctypes.data_as(dp), o.ctypes.data_as(dp))
        out[i] = o
    return out


@pytest.mark.parametrize("model", ["none", "fov", "radtan4", "radtan5", "rpc2"])
def test_undistortion_matches_independent_answers(oracle_lib, model):
    g = np.load(os.path.join(GOLD, "undistort.npz"))
    f, c, size, pix = float(g["focal"]), g["center"], g["size"], g["pix"]
    if model == "none":
        got = _undist(oracle_lib, size, 0, f, c, np.zeros(1), pix)
    elif model == "fov":
        got = _undist(oracle_lib, size, 1, f, c, np.array([float(g["fov_w"])]), pix)
    elif model == "rpc2":
        got = _undist(oracle_lib, size, len(g["rpc2_coeffs"]), f, c, g["rpc2_coeffs"], pix)
    else:
        got = _undist(oracle_lib, size, len(g[model + "_coeffs"]), f, c, g[model + "_coeffs"], pix)
    want = g[model]
    assert np.abs(got - want).max() <= 1e-9 * max(1.0, np.abs(want).max())


CASES = {
    "pinhole": dict(sensors="pinhole", num_ref=12, num_landmarks=300, obs_per_track=4, float_rig=False),
    "rig3": dict(sensors="rig3", num_ref=12, num_landmarks=400, obs_per_track=5, outlier_frac=0.03),
    "rig3_ragged": dict(sensors="rig3b", num_ref=10, num_landmarks=300, obs_per_track=6, ragged=True, outlier_frac=0.02),
    "rpc": dict(sensors="rpc", num_ref=10, num_landmarks=300, obs_per_track=6, float_rig=False),
    "no_rig": dict(sensors="rig3", num_ref=10, num_landmarks=300, obs_per_track=5, no_rig=True),
}


def test_noise_free_landmarks_are_recovered(oracle_lib):
    """Exact cameras and pixels: the reweighted linear triangulation returns the landmark (up to float32 pixels)."""
    pr = synth.make_problem(seed=11, sensors="pinhole", num_ref=12, num_landmarks=200, obs_per_track=4, float_rig=False,
                            pixel_noise=0.0, pose_noise=(0.0, 0.0), rig_noise=(0.0, 0.0), xyz_noise=0.0)
    truth = pr.xyz.reshape(-1, 3).copy()
    o = _abi.Binding(oracle_lib, "oracle_", pr)
    xyz, inl = o.triangulate()
    o.close()
    assert inl.all()
    assert np.abs(xyz - truth).max() <= 1e-3  # pixels are rounded to float32 (6e-5 px) and rays converge at a few degrees


def test_tracks_with_fewer_than_two_rays_lose_all_features(oracle_lib):
    pr = synth.make_problem(seed=12, **CASES["rig3"])
    o = _abi.Binding(oracle_lib, "oracle_", pr)
    inl = np.ones(pr.num_obs, np.uint8)
    first = np.flatnonzero(pr.pid == 7)
    inl[first[1:]] = 0                       # one ray left on track 7
    xyz0 = np.full(3 * pr.num_landmarks, -5.0)
    xyz, out = o.triangulate(inl, xyz0)
    o.close()
    assert not out[first].any()
    assert (xyz[7] == -5.0).all()            # untouched, like the reference's default-constructed entry
    others = pr.pid != 7
    assert out[others].all()


def test_outliers_never_return_and_gross_errors_are_flagged(oracle_lib):
    pr = synth.make_problem(seed=13, **CASES["rig3"])
    o = _abi.Binding(oracle_lib, "oracle_", pr)
    _, r = o.evaluate(False)
    kind, row0, src = o.residual_layout()
    cam = kind == 0
    norm = np.zeros(pr.num_obs)
    norm[src[cam]] = np.hypot(r[row0[cam]], r[row0[cam] + 1])
    start = np.ones(pr.num_obs, np.uint8); start[::9] = 0
    inl, counts = o.flag_outliers(0.0, 4.0, start)
    o.close()
    assert not inl[::9].any()
    assert (inl == ((norm <= 4.0) & (start == 1))).all()      # angle threshold 0: only the reprojection rule acts
    assert counts[0] == start.sum() and counts[1] == 0 and counts[2] == start.sum() and counts[3] == start.sum() - inl.sum()


# ---------------------------------------------------------------------------- GPU parity
@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_gpu_triangulation_matches_oracle(name, oracle_lib, product_lib):
    pr = synth.make_problem(seed=31, **CASES[name])
    rng = np.random.default_rng(5)
    start = (rng.random(pr.num_obs) > 0.15).astype(np.uint8)  # ragged inlier sets
    for p in range(0, pr.num_landmarks, 17):                  # and some tracks with one ray or none
        q = np.flatnonzero(pr.pid == p)
        start[q[(p // 17) % 2:]] = 0
        start[q[:(p // 17) % 2]] = 1
    xyz0 = rng.normal(size=3 * pr.num_landmarks)
    o = _abi.Binding(oracle_lib, "oracle_", pr); g = _abi.Binding(product_lib, "mvba_", pr)
    xo, io = o.triangulate(start, xyz0)
    xg, ig = g.triangulate(start, xyz0)
    o.close(); g.close()
    assert (io == ig).all()
    assert np.abs(xo - xg).max() <= 1e-9 * max(1.0, np.abs(xo).max())
    assert (io == 0).sum() > (start == 0).sum()               # the < 2 rays rule fired


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_gpu_outlier_flags_match_oracle(name, oracle_lib, product_lib):
    pr = synth.make_problem(seed=32, **CASES[name])
    rng = np.random.default_rng(6)
    start = (rng.random(pr.num_obs) > 0.1).astype(np.uint8)
    o = _abi.Binding(oracle_lib, "oracle_", pr); g = _abi.Binding(product_lib, "mvba_", pr)
    # thresholds placed inside the populated range so that both rules fire
    _, r = o.evaluate(False)
    kind, row0, _ = o.residual_layout()
    cam = kind == 0
    err = np.hypot(r[row0[cam]], r[row0[cam] + 1])
    max_err = float(np.percentile(err, 80))
    for min_angle in (0.0, 2.0):
        io, co = o.flag_outliers(min_angle, max_err, start)
        ig, cg = g.flag_outliers(min_angle, max_err, start)
        # a residual within 1e-9 of the threshold may legitimately fall on either side
        ambiguous = np.zeros(pr.num_obs, bool)
        _, _, src = o.residual_layout()
        ambiguous[src[cam]] = np.abs(err - max_err) <= 1e-9 * max_err
        assert ((io == ig) | ambiguous).all()
        if not ambiguous.any():
            assert (co == cg).all()
        assert not ig[start == 0].any()
    size = np.array([[s.width, s.height] for s in pr.sensors]); crop = np.round(size * 0.5)
    eo = o.flag_exclusion(size, crop, start); eg = g.flag_exclusion(size, crop, start)
    o.close(); g.close()
    assert (eo == eg).all() and eo.sum() <= start.sum() and not eg[start == 0].any()
