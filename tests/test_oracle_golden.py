"""CPU tests (no GPU): the oracle against the committed known-answer vectors in tests/golden/
(tools/make_golden.py: scipy Slerp, cv2.projectPoints, mpmath at 40 digits -- independent of the oracle)
and against the invariants SURVEY.md section 4 lists.  The reference itself ships no tests or fixtures for
this path, so these vectors are "restatement-derived": parity is UNPINNED by the reference."""
import ctypes as C
import os

import numpy as np
import pytest

from multiview_b200 import _abi, synth

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
D = C.POINTER(C.c_double)


def _d(a):
    return a.ctypes.data_as(D)


def test_pose_interpolation_matches_scipy(oracle_lib):
    g = np.load(os.path.join(GOLD, "pose_interp.npz"))
    out = np.zeros(12); ok = C.c_int(0)
    oracle_lib.oracle_world_to_cam.argtypes = [D, D, D, C.c_double, C.c_double, C.c_double, C.c_double, D, C.POINTER(C.c_int)]
    for i in range(len(g["bs"])):
        b, e, r = (np.ascontiguousarray(g[k][i]) for k in ("beg", "end", "r2c"))
        oracle_lib.oracle_world_to_cam(_d(b), _d(e), _d(r), g["bs"][i], g["es"][i], g["off"][i], g["cs"][i], _d(out), C.byref(ok))
        assert ok.value == 1
        assert np.abs(out - g["world_to_cam"][i]).max() < 1e-11 * max(1.0, np.abs(g["world_to_cam"][i]).max())
    # beg stamp == end stamp: T(beg); end pose, rig transform and offset are ignored (rig_calibrator.cc:371-374)
    for i in range(len(g["degenerate"])):
        b, e, r = (np.ascontiguousarray(g[k][i]) for k in ("beg", "end", "r2c"))
        oracle_lib.oracle_world_to_cam(_d(b), _d(e), _d(r), 5.0, 5.0, 0.3, 5.0, _d(out), C.byref(ok))
        assert np.abs(out - g["degenerate"][i]).max() < 1e-12 * max(1.0, np.abs(g["degenerate"][i]).max())


def test_interpolation_fraction_outside_bracket_is_flagged(oracle_lib):
    """LOG(FATAL) 'Out of bounds in interpolation' at rig_calibrator.cc:346."""
    oracle_lib.oracle_world_to_cam.argtypes = [D, D, D, C.c_double, C.c_double, C.c_double, C.c_double, D, C.POINTER(C.c_int)]
    p = np.array([0, 0, 0, 0, 0, 0, 1.0]); out = np.zeros(12); ok = C.c_int(1)
    oracle_lib.oracle_world_to_cam(_d(p), _d(p), _d(p), 10.0, 11.0, 0.0, 11.5, _d(out), C.byref(ok))
    assert ok.value == 0


def _distort(lib, w, h, f, c, dist, u):
    lib.oracle_distort.argtypes = [C.c_double, C.c_double, C.c_int, C.c_double, D, D, D, D]
    out = np.zeros_like(u)
    dist = np.ascontiguousarray(dist, dtype=np.float64); c = np.ascontiguousarray(c, dtype=np.float64)
    for i in range(len(u)):
        ui = np.ascontiguousarray(u[i]); oi = np.zeros(2)
        lib.oracle_distort(w, h, len(dist), f, _d(c), _d(dist if len(dist) else np.zeros(1)), _d(ui), _d(oi))
        out[i] = oi
    return out


@pytest.mark.parametrize("model", ["none", "fov", "radtan4", "radtan5", "rpc2", "rpc3"])
def test_distortion_models(oracle_lib, model):
    g = np.load(os.path.join(GOLD, "distortion.npz"))
    f, c, (w, h), u = float(g["focal"]), g["center"], g["size"], g["u"]
    dist = {"none": np.zeros(0), "fov": np.array([float(g["fov_w"])])}.get(model)
    if dist is None:
        dist = g[model + "_coeffs"]
    got = _distort(oracle_lib, w, h, f, c, dist, u)
    assert np.abs(got - g[model]).max() < 2e-10, np.abs(got - g[model]).max()   # pixels; cv2 itself is ~1e-13 here


def test_rpc_second_half_and_identity(oracle_lib):
    """Only the distortion half of the RPC block enters the residual (rpc_distortion.cc:272-293), and the
    identity coefficients reproduce the input (rpc_distortion.cc:303-318)."""
    g = np.load(os.path.join(GOLD, "distortion.npz"))
    f, c, (w, h), u = float(g["focal"]), g["center"], g["size"], g["u"][:20]
    co = g["rpc3_coeffs"].copy()
    a = _distort(oracle_lib, w, h, f, c, co, u)
    co[len(co) // 2:] = 7.0
    assert np.array_equal(a, _distort(oracle_lib, w, h, f, c, co, u))
    for deg in (1, 2, 3, 4):
        n = synth.rpc_num_dist_params(deg); L = (deg + 1) * (deg + 2) // 2
        ident = np.zeros(2 * n); ident[1] = 1.0; ident[L + (L - 1) + 2] = 1.0
        assert np.abs(_distort(oracle_lib, w, h, f, c, ident, u) - (u + np.array([w, h]) / 2)).max() < 1e-12
    assert [synth.rpc_num_dist_params(d) for d in (1, 2, 3, 4)] == [10, 22, 38, 58]


def test_rigid_array_round_trip(oracle_lib):
    """array_to_rigid_transform(rigid_transform_to_array(T)) == T (dense_map_utils.cc:159-178)."""
    rng = np.random.default_rng(3)
    oracle_lib.oracle_array_to_rigid.argtypes = [D, D]; oracle_lib.oracle_rigid_to_array.argtypes = [D, D]
    for _ in range(50):
        a = np.concatenate([rng.normal(size=3), rng.normal(size=4)])
        m = np.zeros(12); a2 = np.zeros(7); m2 = np.zeros(12)
        oracle_lib.oracle_array_to_rigid(_d(a), _d(m)); oracle_lib.oracle_rigid_to_array(_d(m), _d(a2)); oracle_lib.oracle_array_to_rigid(_d(a2), _d(m2))
        assert np.abs(m - m2).max() < 1e-14
        assert abs(np.linalg.norm(a2[3:]) - 1) < 1e-14
        R = m.reshape(3, 4)[:, :3]
        assert np.abs(R @ R.T - np.eye(3)).max() < 1e-14 and np.linalg.det(R) > 0


def test_cauchy_loss(oracle_lib):
    g = np.load(os.path.join(GOLD, "cauchy.npz"))
    oracle_lib.oracle_cauchy.argtypes = [C.c_double, C.c_double, D, D]
    for s, rho, sq in zip(g["s"], g["rho"], g["sqrt_rho1"]):
        a, b = C.c_double(0), C.c_double(0)
        oracle_lib.oracle_cauchy(float(g["a"]), float(s), C.byref(a), C.byref(b))
        assert abs(a.value - rho) <= 1e-14 * max(1.0, rho) and abs(b.value - sq) <= 1e-15
