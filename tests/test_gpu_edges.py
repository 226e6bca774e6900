"""GPU parity, edge cases that have device code of their own ('edge cases with device code and no GPU test'):
  * an interpolation fraction outside [0, 1]  -> MVBA_EBRACKET, caller parameters untouched (rig_calibrator.cc:346 is a LOG(FATAL));
  * a floated time offset held by an ACTIVE bound (rig_calibrator.cc:1742-1745);
  * tracks wider than the tile window            -> the generic warp-per-track kernel next to the tile kernel;
  * windows wide enough for the 8- and 12/16-block-per-warp instantiations of both tile kernels.
Every case asserts through mvba_build_plan that the path it is meant to cover is the one that ran."""
import numpy as np
import pytest

from multiview_b200 import _abi, synth
from test_gpu_parity_solve import _param_error

pytestmark = pytest.mark.gpu


def _both(pr, oracle_lib, product_lib, iters, env=None, monkeypatch=None):
    snap = pr.snapshot()
    o = _abi.Binding(oracle_lib, "oracle_", pr)
    so, lo = o.solve(_abi.default_options(max_num_iterations=iters), jac_mode=1, num_threads=1)
    xo = pr.snapshot(); o.close()
    pr.restore(snap)
    if env:
        for k, v in env.items():
            monkeypatch.setenv(k, v)
    g = _abi.Binding(product_lib, "mvba_", pr)
    sg, lg = g.solve(_abi.default_options(max_num_iterations=iters))
    plan = g.build_plan()
    xg = pr.snapshot(); g.close()
    pr.restore(snap)
    return (so, lo, xo), (sg, lg, xg), plan


def _same_path(a, b, pr, ptol=5e-5):
    (so, lo, xo), (sg, lg, xg) = a, b
    assert abs(sg.initial_cost - so.initial_cost) <= 1e-9 * so.initial_cost
    assert sg.num_iterations == so.num_iterations
    for p, q in zip(lg, lo):
        assert p["step_accepted"] == q["step_accepted"], (p, q)
        assert abs(p["cost"] - q["cost"]) <= 1e-6 * q["cost"], (p, q)
    for k in xo:
        if xo[k].size:
            assert _param_error(k, xo[k], xg[k], pr) <= ptol, k


def test_out_of_bracket_is_an_error_and_leaves_the_parameters_alone(product_lib):
    pr = synth.make_problem(seed=5, sensors="rig3", num_ref=10, num_landmarks=300, obs_per_track=5, float_offsets=True)
    pr.offset_lo[:] = pr.offsets - 1.0; pr.offset_hi[:] = pr.offsets + 1.0   # bounds that do NOT protect the brackets
    pr.offsets[1] += 0.3                                                    # images of sensor 1 now fall outside their brackets
    snap = pr.snapshot()
    g = _abi.Binding(product_lib, "mvba_", pr)
    with pytest.raises(_abi.MvbaError, match="rc=-5"):
        g.evaluate(True)
    with pytest.raises(_abi.MvbaError, match="rc=-5"):
        g.solve(_abi.default_options(max_num_iterations=3))
    g.close()
    for k, v in pr.snapshot().items():
        assert np.array_equal(v, snap[k]), k


def test_a_step_that_leaves_a_bracket_returns_the_error_without_writing_back(product_lib):
    """In-bracket start, bounds too loose to protect the bracket, a huge trust region: some LM candidate leaves a bracket."""
    pr = synth.make_problem(seed=23, sensors="rig3", num_ref=12, num_landmarks=500, obs_per_track=5, float_offsets=True,
                            tri_weight=5.0, tri_robust_threshold=1.0)
    pr.offset_lo[:] = pr.offsets - 1.0; pr.offset_hi[:] = pr.offsets + 1.0
    snap = pr.snapshot()
    g = _abi.Binding(product_lib, "mvba_", pr)
    try:
        g.solve(_abi.default_options(max_num_iterations=60, eta=1e-8, max_trust_region_radius=1e7))
    except _abi.MvbaError as e:
        assert "rc=-5" in str(e)
        for k, v in pr.snapshot().items():
            assert np.array_equal(v, snap[k]), k
    else:
        pytest.skip("this start never left a bracket")
    finally:
        g.close()


def test_active_offset_bound(oracle_lib, product_lib):
    """The solve wants to take sensor 2's offset beyond a bound: the candidate is projected onto the bound (Ceres'
    ParameterBlock::Plus projection; the oracle does the same) and stays there."""
    pr = synth.make_problem(seed=23, sensors="rig3", num_ref=12, num_landmarks=500, obs_per_track=5, float_offsets=True)
    snap = pr.snapshot()
    o = _abi.Binding(oracle_lib, "oracle_", pr)
    o.solve(_abi.default_options(max_num_iterations=8), jac_mode=1, num_threads=1)  # where the free solve takes the offset
    o.close()
    free = pr.offsets[2]
    pr.restore(snap)
    # the bound sits a tenth of the way from the start to where the unconstrained solve goes
    bound = pr.offsets[2] + 0.1 * (free - pr.offsets[2])
    assert abs(free - pr.offsets[2]) > 1e-4
    if free > pr.offsets[2]: pr.offset_hi[2] = bound
    else: pr.offset_lo[2] = bound
    a, b, plan = _both(pr, oracle_lib, product_lib, 8)
    assert a[2]["offsets"][2] == bound, "the oracle never reached the bound: the case does not test what it says"
    assert b[2]["offsets"][2] == bound
    _same_path(a, b, pr)


def _with_wide_tracks(pr, n_wide):
    """Re-aims the observations of the first n_wide tracks at images spread over the whole sequence (gross outliers for the
    robust loss, but tracks whose pose blocks no window can hold)."""
    C = len(pr.cam_sensor)
    cid = pr.cid.copy()
    for p in range(n_wide):
        idx = np.flatnonzero(pr.pid == p)
        cid[idx] = np.linspace(0, C - 1, len(idx)).round().astype(np.int32)
    pr.cid = cid
    return pr


@pytest.mark.parametrize("staged", [True, False])
def test_tracks_wider_than_the_window_take_the_generic_kernel(staged, oracle_lib, product_lib, monkeypatch):
    pr = synth.make_problem(seed=9, sensors="rig3", num_ref=40, num_landmarks=900, obs_per_track=6, outlier_frac=0.02)
    pr = _with_wide_tracks(pr, 6)
    a, b, plan = _both(pr, oracle_lib, product_lib, 5, env={"MVBA_TILE_UNSTAGED": "0" if staged else "1"}, monkeypatch=monkeypatch)
    assert plan["pass_b"] == (2 if staged else 1) and plan["fallback_tracks"] >= 6, plan
    _same_path(a, b, pr)


@pytest.mark.parametrize("obs_per_track,intr,want", [(12, False, (2, 10)), (12, True, (1, 8)), (24, True, (1, 12))])
def test_wide_windows_use_the_larger_accumulator_instantiations(obs_per_track, intr, want, oracle_lib, product_lib):
    kw = dict(float_intrinsics=("focal_length",), intr_noise=1e-3) if intr else {}
    pr = synth.make_problem(seed=4, sensors="rig3", num_ref=40, num_landmarks=400, obs_per_track=obs_per_track, **kw)
    a, b, plan = _both(pr, oracle_lib, product_lib, 4)
    assert plan["pass_b"] == want[0] and plan["blocks_per_warp"] >= want[1] and plan["fallback_tracks"] == 0, plan
    _same_path(a, b, pr)


def test_depth_mesh_and_prior_blocks_on_the_fast_kernels(oracle_lib, product_lib):
    """Depth, depth-mesh, mesh and triangulation-prior blocks with every intrinsic and depth transform FIXED: the pass then takes
    the per-image accumulators of pass A, the staged tile kernel and the camera-major candidate cost (camcost_kernel evaluates the
    depth kinds itself, landmark priors go to residual_kernel with an offset) -- the kernels BASELINE config 5 runs, which itself
    has no such blocks."""
    pr = synth.make_problem(seed=23, sensors="rig3", num_ref=12, num_landmarks=500, obs_per_track=5, depth=True, mesh=True,
                            depth_tri_weight=1000.0, depth_mesh_weight=50.0, mesh_tri_weight=25.0, tri_weight=0.5)
    a, b, plan = _both(pr, oracle_lib, product_lib, 6)
    assert plan["pass_a_accumulators"] == 1 and plan["pass_b"] == 2, plan
    assert b[0].num_successful_steps >= 3
    _same_path(a, b, pr)
