"""CPU tests (no GPU): the C ABI surface of libmvba.so, the oracle's self-consistency (three Jacobian
routes, indexing replay, convergence to the synthetic truth) and the landmark sharding used by N>1 runs."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from multiview_b200 import _abi, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# ---------------------------------------------------------------------------------------------
# C ABI
# ---------------------------------------------------------------------------------------------
def _declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "mvba.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(mvba_[a-z0-9_]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol(product_lib):
    names = _declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(product_lib, n), f"libmvba.so does not export {n}"
    assert product_lib.mvba_abi_version() == 1


def test_struct_layouts_match_the_header(product_lib):
    """ctypes mirrors vs the sizes the C compiler gives the header structs."""
    import subprocess, tempfile
    src = '#include <stdio.h>\n#include "mvba.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu\\n",sizeof(mvba_config),sizeof(mvba_params),' \
          'sizeof(mvba_float_mask),sizeof(mvba_solve_options),sizeof(mvba_summary),sizeof(mvba_iteration));return 0;}\n'
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "s.c"), "w").write(src)
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), "-o", os.path.join(d, "s"), os.path.join(d, "s.c")])
        sizes = [int(x) for x in subprocess.check_output([os.path.join(d, "s")]).split()]
    mirrors = [_abi.Config, _abi.Params, _abi.FloatMask, _abi.SolveOptions, _abi.Summary, _abi.Iteration]
    assert sizes == [C.sizeof(m) for m in mirrors]


def test_options_init_matches_reference_settings(product_lib):
    """rig_calibrator.cc:1910-1918 + Ceres 1.14 defaults."""
    o = _abi.SolveOptions()
    assert product_lib.mvba_solve_options_init(C.byref(o)) == 0
    d = _abi.default_options()
    for k, _ in _abi.SolveOptions._fields_:
        assert getattr(o, k) == getattr(d, k), k
    assert (o.max_num_iterations, o.parameter_tolerance, o.function_tolerance, o.gradient_tolerance) == (20, 1e-12, 1e-16, 1e-16)
    assert (o.eta, o.max_linear_solver_iterations, o.initial_trust_region_radius, o.min_relative_decrease) == (0.1, 500, 1e4, 1e-3)


def test_compute_entry_points_fail_loudly_without_a_device(product_lib):
    """There is no CPU fallback: on a box without an sm_100 device every compute call returns MVBA_ENODEVICE."""
    if product_lib.mvba_device_count() > 0:
        pytest.skip("a CUDA device is present")
    pr = synth.make_problem(seed=1, sensors="pinhole", num_ref=8, num_landmarks=20, obs_per_track=3, float_rig=False)
    b = _abi.Binding(product_lib, "mvba_", pr)          # describing the problem needs no device
    with pytest.raises(_abi.MvbaError) as e:
        b.evaluate(False)
    assert "rc=-2" in str(e.value)
    with pytest.raises(_abi.MvbaError):
        b.solve(_abi.default_options(max_num_iterations=1))
    b.close()


def test_argument_validation(product_lib):
    cfg = _abi.Config(0, 1, 1, 1, 0, 0, 3.0, 0.0, 0.0)
    h = C.c_void_p()
    assert product_lib.mvba_create(C.byref(cfg), 0, C.byref(h)) == -1        # num_sensors must be positive
    cfg = _abi.Config(1, 1, 1, 1, 0, 0, 0.0, 0.0, 0.0)
    assert product_lib.mvba_create(C.byref(cfg), 0, C.byref(h)) == -1        # robust_threshold must be positive
    assert b"robust_threshold" in product_lib.mvba_last_error()
    cfg = _abi.Config(1, 1, 1, 1, 0, 0, 3.0, 0.0, 0.0)
    assert product_lib.mvba_create(C.byref(cfg), 0, C.byref(h)) == 0
    assert product_lib.mvba_set_sensor(h, 3, C.c_double(10), C.c_double(10), 0) == -1
    assert product_lib.mvba_set_sensor(h, 0, C.c_double(10), C.c_double(10), 500) == -1
    assert product_lib.mvba_destroy(h) == 0


def test_rpc_fit_argument_validation_and_no_device(product_lib):
    """mvba_rpc_fit rejects bad sizes before touching the device, and has no CPU fallback."""
    dp = C.POINTER(C.c_double)
    xy = np.zeros((4, 2)); c = np.zeros(_abi.rpc_num_params(1)); st = np.zeros(4)
    args = lambda n, deg, it: (C.c_int32(0), C.c_int64(n), xy.ctypes.data_as(dp), xy.ctypes.data_as(dp), C.c_int32(deg), C.c_int32(it),
                               C.c_double(1e-12), C.c_int32(0), c.ctypes.data_as(dp), st.ctypes.data_as(dp))
    for bad in (args(0, 1, 10), args(4, 0, 10), args(4, 7, 10), args(4, 1, -1)):
        assert product_lib.mvba_rpc_fit(*bad) == -1
        assert b"mvba_rpc_fit" in product_lib.mvba_last_error()
    assert product_lib.mvba_rpc_fit(C.c_int32(0), C.c_int64(4), None, xy.ctypes.data_as(dp), 1, 10, C.c_double(1e-12), 0, c.ctypes.data_as(dp), None) == -1
    if product_lib.mvba_device_count() == 0:
        assert product_lib.mvba_rpc_fit(*args(4, 1, 10)) == -2      # MVBA_ENODEVICE


# ---------------------------------------------------------------------------------------------
# oracle self-consistency
# ---------------------------------------------------------------------------------------------
CASES = {
    "rig3_all": dict(sensors="rig3b", num_ref=8, num_landmarks=60, obs_per_track=5, float_offsets=True, ragged=True,
                     float_intrinsics=("focal_length", "optical_center", "distortion"), outlier_frac=0.02),
    "depth": dict(sensors="rig3", num_ref=8, num_landmarks=60, obs_per_track=5, depth=True, mesh=True, depth_tri_weight=1000.0,
                  depth_mesh_weight=50.0, mesh_tri_weight=25.0, tri_weight=0.5, float_d2i=True, float_scale=True),
    "rpc": dict(sensors="rpc", num_ref=8, num_landmarks=60, obs_per_track=5, float_rig=False, float_intrinsics=("focal_length", "distortion")),
}


@pytest.mark.parametrize("name", sorted(CASES))
def test_three_jacobian_routes_agree(name, oracle_lib):
    """mode 0: Ceres-rule central differences (what the reference computes, rig_calibrator.cc:489-492);
    mode 1: forward-mode duals (the arbiter the CUDA kernels are held to); mode 2: long double differences."""
    pr = synth.make_problem(seed=7, **CASES[name])
    o = _abi.Binding(oracle_lib, "oracle_", pr)
    kind, _, _ = o.residual_layout()
    for k in np.unique(kind):
        n = int((kind == k).sum())
        _, Jn = o.evaluate_jacobian(int(k), 0, n, mode=0)
        _, Jd = o.evaluate_jacobian(int(k), 0, n, mode=1)
        _, Jl = o.evaluate_jacobian(int(k), 0, n, mode=2)
        scale = np.abs(Jd).max(axis=(1, 2), keepdims=True) + 1e-300
        # exact derivative vs long double central differences (themselves limited to ~1e-8 on the RPC
        # coefficient columns, whose entries reach 1e8 pixels per unit coefficient)
        assert (np.abs(Jd - Jl) / scale).max() < (1e-7 if name == "rpc" else 1e-9), k
        # the reference's own numeric Jacobian: step max(1e-6 |x|, sqrt(eps)); its error is what it is
        # (RPC denominators: a 1.5e-8 step on a 1e-9 coefficient whose column is ~1e8 px/unit leaves a few percent
        #  of truncation error in the reference's numeric Jacobian; documented, not a gate)
        assert (np.abs(Jn - Jd) / scale).max() < (0.1 if name == "rpc" else 5e-5), k
    o.close()


def _replay_layout(pr):
    """Residual-block order of rig_calibrator.cc:1614-1903 restated in Python."""
    kinds, rows, src = [], [], []
    row = 0
    pri = {k: dict(zip(v[0].tolist(), range(len(v[0])))) for k, v in pr.priors.items()}
    i = 0
    N = pr.num_obs
    for p in range(pr.num_landmarks):
        any_obs = False
        while i < N and pr.pid[i] == p:
            any_obs = True
            kinds.append(synth.BLOCK_CAM); rows.append(row); src.append(i); row += 2
            hd = pr.has_depth is not None and pr.has_depth[i]
            if pr.depth_tri_weight > 0 and hd:
                kinds.append(synth.BLOCK_DEPTH_TRI); rows.append(row); src.append(i); row += 3
            if pr.depth_mesh_weight > 0 and hd and pr.has_mesh is not None and pr.has_mesh[i]:
                kinds.append(synth.BLOCK_DEPTH_MESH); rows.append(row); src.append(i); row += 3
            i += 1
        for k in (synth.BLOCK_MESH_TRI, synth.BLOCK_TRI):
            if any_obs and k in pri and p in pri[k]:
                kinds.append(k); rows.append(row); src.append(pri[k][p]); row += 3
    return np.array(kinds, np.int32), np.array(rows, np.int64), np.array(src, np.int64), row


@pytest.mark.parametrize("name", sorted(CASES))
def test_indexing_replay_is_bit_exact(name, oracle_lib):
    pr = synth.make_problem(seed=9, **CASES[name])
    o = _abi.Binding(oracle_lib, "oracle_", pr)
    kind, row0, src = o.residual_layout()
    k2, r2, s2, nres = _replay_layout(pr)
    assert o.num_residuals() == (nres, len(k2))
    assert np.array_equal(kind, k2) and np.array_equal(row0, r2) and np.array_equal(src, s2)
    o.close()


def test_solve_recovers_the_synthetic_truth(oracle_lib):
    """Noise-free pinhole pass: the LM of the oracle must drive the cost to ~0 and the landmarks back to the truth
    up to the gauge (compared through reprojection, which is gauge invariant)."""
    pr = synth.make_problem(seed=4, sensors="pinhole", num_ref=10, num_landmarks=150, obs_per_track=4, float_rig=False, pixel_noise=0.0)
    o = _abi.Binding(oracle_lib, "oracle_", pr)
    s, log = o.solve(_abi.default_options(max_num_iterations=30), jac_mode=1)
    assert s.initial_cost > 1.0
    assert s.final_cost < 1e-6          # float32 rounding of the pixels leaves ~1e-5 px residuals
    c, r = o.evaluate(False)
    assert np.abs(r).max() < 1e-3
    o.close()


def test_residual_is_weighted_and_loss_free_in_evaluate(oracle_lib):
    """Problem::Evaluate with apply_loss_function = false (rig_calibrator.cc:1163-1166): cost = sum r^2 / 2."""
    pr = synth.make_problem(seed=2, **CASES["depth"])
    o = _abi.Binding(oracle_lib, "oracle_", pr)
    c0, r = o.evaluate(False)
    assert abs(c0 - 0.5 * float(r @ r)) <= 1e-12 * c0
    c1, _ = o.evaluate(True)
    assert c1 < c0
    o.close()


# ---------------------------------------------------------------------------------------------
# landmark sharding (N > 1 runs)
# ---------------------------------------------------------------------------------------------
def test_shards_partition_the_tracks_and_sum_to_the_whole(oracle_lib):
    pr = synth.make_problem(seed=6, **CASES["depth"])
    o = _abi.Binding(oracle_lib, "oracle_", pr)
    c_all, r_all = o.evaluate(True)
    o.close()
    world = 3
    costs, nres, seen = 0.0, 0, []
    for rank in range(world):
        q = pr.shard(rank, world)
        lo, hi = q.truth["pid_range"]
        seen.append((lo, hi))
        assert q.num_landmarks == hi - lo and (np.diff(q.pid) >= 0).all() and (q.pid >= 0).all() and (q.pid < q.num_landmarks).all()
        b = _abi.Binding(oracle_lib, "oracle_", q)
        c, r = b.evaluate(True)
        costs += c; nres += len(r)
        b.close()
    assert seen[0][0] == 0 and seen[-1][1] == pr.num_landmarks and all(seen[i][1] == seen[i + 1][0] for i in range(world - 1))
    assert nres == len(r_all) and abs(costs - c_all) <= 1e-12 * c_all


@pytest.mark.parametrize("locality", [False, True])
def test_both_sharding_modes_partition_the_tracks(locality, oracle_lib):
    """Input-order and image-ordered landmark ranges: every track on exactly one rank, observation counts balanced, costs add up.
    Image-ordered shards touch a fraction of the images each -- the case that needs the column layout to be the union over ranks."""
    pr = synth.make_problem(seed=12, sensors="rig3", num_ref=40, num_landmarks=1500, obs_per_track=5, outlier_frac=0.02)
    o = _abi.Binding(oracle_lib, "oracle_", pr)
    c_all, _ = o.evaluate(True)
    o.close()
    world = 4
    owner = np.full(pr.num_landmarks, -1)
    cost, nobs, images = 0.0, [], []
    for rank in range(world):
        q = pr.shard(rank, world, locality=locality)
        tracks = q.truth["tracks"]
        assert (owner[tracks] == -1).all()
        owner[tracks] = rank
        assert np.allclose(q.xyz.reshape(-1, 3), pr.xyz.reshape(-1, 3)[tracks])
        b = _abi.Binding(oracle_lib, "oracle_", q)
        c, _ = b.evaluate(True)
        b.close()
        cost += c; nobs.append(q.num_obs); images.append(len(np.unique(q.cid)))
    assert (owner >= 0).all() and abs(cost - c_all) <= 1e-12 * c_all
    assert max(nobs) - min(nobs) <= 2 * 5 + 0.02 * pr.num_obs
    n_img = len(pr.cam_sensor)
    if locality:
        assert max(images) < 0.75 * n_img          # a rank sees only part of the images ...
    else:
        assert min(images) > 0.9 * n_img           # ... while input-order ranges of random tracks see nearly all of them


def test_the_product_package_carries_no_checker_code():
    """The CPU oracle is test infrastructure: nothing under multiview_b200/ may call, link, include or import it (the one
    hook is _abi.register_checker, which oracle/binding.py calls from the tests' side)."""
    pkg = os.path.join(ROOT, "multiview_b200")
    pat = r'oracle_[a-z_]+\s*\(|["\'][^"\'\n]*liboracle|#include\s+"[^"]*oracle|^\s*import oracle|^\s*from oracle'
    hits = []
    for d, _, files in os.walk(pkg):
        for f in files:
            if not f.endswith((".py", ".cu", ".cuh", ".h", ".cc", ".cpp", "Makefile")):
                continue
            txt = open(os.path.join(d, f), errors="replace").read()
            for m in re.finditer(pat, txt, flags=re.M):
                hits.append((os.path.relpath(os.path.join(d, f), ROOT), m.group(0)))
    assert not hits, hits
