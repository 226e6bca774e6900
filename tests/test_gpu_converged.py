"""GPU parity, part 3: the CONVERGED state of the device LM solve against the CPU oracle.

north_star: "the converged cost must match to 1e-6 relative and the optimised poses to 1e-6 m and 1e-6 rad".  That is only
well defined where the minimum is: the three hard families of test_gpu_parity_solve.py (depth + mesh with depth_to_image and
scale floating; every intrinsic floating on ragged tracks; floating time offsets) are run here in a WELL-POSED form --
the stay-near-initial landmark prior of the reference (XYZError with --tri_weight, rig_calibrator.cc:1886-1901) pins the
7-dof gauge, there are no gross outliers, max_trust_region_radius = 1e7 keeps the LM damping from vanishing along the
quaternion-norm null directions (the reference has no manifold, dense_map_utils.cc:171-178) -- and both sides iterate with
an almost exact Newton step (eta = 1e-8) for at most 120 iterations: both stop on the function tolerance after 60 to 90
(at 60 the cost still moves by 1e-9 relative per iteration and two oracle runs on 8 threads were seen 8.7e-7 apart in a pose;
at the tolerance stop they are within 3e-9).  tests/test_oracle_converged.py (CPU) shows that this state is path independent at the tolerances used here:
the oracle on 1 and on 8 threads (different summation orders, different CG iteration counts) agree to 5e-7 in the poses
(typically 1e-9) and to 5e-5 px in the intrinsics.

Tolerances: cost 1e-6 relative (north_star); poses, rig transforms, depth transforms 1e-6 m / 1e-6 rad, landmarks 5e-6 m
(2 to 6 m in front of the cameras they are held less tightly than the poses; the device solve has been seen 1.6e-6 m away on
single landmarks of the all-intrinsics family, varying from run to run with the order of its atomics),
time offsets 1e-6 s (north_star); focal length / optical centre / distortion coefficients 2e-3 in their native units
(pixels, dimensionless): the optical centre is the flattest direction of these problems (it trades against the camera
rotation) -- the oracle's own 1-vs-8-thread spread reaches 2e-5 px and the device solve, whose atomics change the summation
order from run to run, has been seen to land 7e-4 px away at equal cost (1e-6); all far below the 0.5 px pixel noise.
The floated time offsets carry the reference's bounds (every image stays inside its bracket, synth.reference_offset_bounds),
as rig_calibrator.cc:1742-1745 sets them; without them a long step along the weak offset direction leaves a bracket, which
is a LOG(FATAL) in the reference and MVBA_EBRACKET here."""
import numpy as np
import pytest

from multiview_b200 import _abi, synth
from test_gpu_parity_solve import CASES, _param_error

pytestmark = pytest.mark.gpu

FAMILIES = ("depth_mesh", "rig3_ragged_intr", "offsets")
OPTS = dict(max_num_iterations=120, eta=1e-8, max_trust_region_radius=1e7)
INTRINSIC = ("focal", "center", "distortion")


def well_posed(name, seed=23):
    kw = dict(CASES[name])
    kw.update(tri_weight=5.0, tri_robust_threshold=1.0, outlier_frac=0.0)
    if name == "offsets":
        # a time offset and the rig translation along the direction of motion are nearly the same direction for this slow,
        # almost straight trajectory: with both floating the minimum is a long valley that ends on an offset bound, where the
        # projected steps of either side only creep (radius shrinking every iteration) and no iteration count converges.
        # The rig is held, the offsets are then determined (to about 1 ms) and the converged state is well defined
        kw.update(float_rig=False)
    return synth.make_problem(seed=seed, **kw)


@pytest.mark.parametrize("name", FAMILIES)
def test_converged_state_matches_the_oracle(name, oracle_lib, product_lib):
    pr = well_posed(name)
    snap = pr.snapshot()
    o = _abi.Binding(oracle_lib, "oracle_", pr)
    so, lo = o.solve(_abi.default_options(**OPTS), jac_mode=1, num_threads=1)
    xo = pr.snapshot(); o.close()
    pr.restore(snap)
    g = _abi.Binding(product_lib, "mvba_", pr)
    sg, lg = g.solve(_abi.default_options(**OPTS))
    xg = pr.snapshot(); g.close()
    # both sides are at a stationary point of the same objective: the last step no longer changes the cost
    for s, log in ((so, lo), (sg, lg)):
        assert abs(log[-1]["cost_change"]) <= 1e-7 * s.final_cost, log[-1]
    assert abs(sg.initial_cost - so.initial_cost) <= 1e-9 * so.initial_cost
    assert abs(sg.final_cost - so.final_cost) <= 1e-6 * so.final_cost
    for k in xo:
        if xo[k].size:
            tol = 2e-3 if k in INTRINSIC else 5e-6 if k == "xyz" else 1e-6
            err = _param_error(k, xo[k], xg[k], pr)
            assert err <= tol, (k, err)
