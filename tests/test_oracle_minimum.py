"""The oracle's Ceres-1.14-style solve against an INDEPENDENT minimiser of an INDEPENDENTLY assembled objective.

The robust objective 1/2 sum_blocks rho_a(|r_block|^2), rho_a(s) = a^2 log(1 + s / a^2) (CauchyLoss as the reference attaches
it, rig_calibrator.cc:394-411, one scale for the reprojection / depth blocks and one per prior kind) is rebuilt here in numpy from
the RAW residuals and the residual layout; MINPACK's Levenberg-Marquardt (scipy.optimize.least_squares, method="lm", forward
differences) then (i) cannot improve the point the oracle converged to, and (ii) started from the same initial guess arrives at
the same cost.  This is the "converged cost within 1e-6" bar of the north star, held against something that shares no code
with the oracle's LM, Schur complement, PCG or Jacobians."""
import numpy as np
import pytest
from scipy.optimize import least_squares

from multiview_b200 import _abi, synth

CASES = {
    "rig3_cam_only": dict(sensors="rig3", num_ref=6, num_landmarks=40, obs_per_track=4, outlier_frac=0.05),
    "depth_mesh_priors": dict(sensors="rig3", num_ref=6, num_landmarks=40, obs_per_track=4, depth=True, mesh=True, depth_tri_weight=1000.0,
                              depth_mesh_weight=50.0, mesh_tri_weight=25.0, tri_weight=0.5, float_d2i=True, float_scale=True),
}


def objective(pr, b):
    kind, _, _ = b.residual_layout()
    _, nb = b.num_residuals()
    blk = np.repeat(np.arange(nb), np.where(kind == 0, 2, 3))
    a = np.full(nb, pr.robust_threshold)
    for k, (_, _, _, th) in pr.priors.items():
        a[kind == k] = th
    T = len(pr.sensors)
    nd2i = pr.depth_to_image.size // T
    free = [pr.world_to_ref, pr.xyz] + [pr.ref_to_cam[7 * t:7 * t + 7] for t in range(1, T) if pr.float_rig[t]]
    free += [pr.depth_to_image[nd2i * t:nd2i * (t + 1)] for t in range(T) if pr.float_d2i is not None and pr.float_d2i[t]]
    if pr.float_scale:
        free += [pr.depth_scale]

    def get():
        return np.concatenate([arr.ravel() for arr in free])

    def z(x):
        o = 0
        for arr in free:
            arr[...] = x[o:o + arr.size].reshape(arr.shape); o += arr.size
        _, r = b.evaluate(False)
        s = np.bincount(blk, r * r, minlength=nb)
        rho = a * a * np.log1p(s / (a * a))
        return r * np.sqrt(rho / np.maximum(s, 1e-300))[blk]      # |z_block|^2 = rho(|r_block|^2)
    return get, z


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_solve_reaches_the_minimum_minpack_finds(name, oracle_lib):
    pr = synth.make_problem(seed=21, **CASES[name])
    snap = pr.snapshot()
    b = _abi.Binding(oracle_lib, "oracle_", pr)
    get, z = objective(pr, b)
    x0 = get()
    c0, _ = b.evaluate(True)
    assert abs(0.5 * (z(x0) ** 2).sum() - c0) <= 1e-13 * c0          # the numpy objective IS the oracle's robust cost
    s, _ = b.solve(_abi.default_options(max_num_iterations=200), num_threads=1)
    assert s.final_cost < 0.5 * s.initial_cost
    kw = dict(method="lm", xtol=1e-15, ftol=1e-15, gtol=1e-15, max_nfev=200000)
    polished = least_squares(z, get(), **kw)
    assert s.final_cost - polished.cost <= 1e-8 * s.final_cost       # nothing left to gain at the oracle's point
    pr.restore(snap)
    fresh = least_squares(z, x0, **kw)
    assert abs(fresh.cost - s.final_cost) <= 1e-6 * s.final_cost      # and the same minimum from the same start
    b.close()
