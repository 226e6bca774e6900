"""Dev tool: the well-posed 'offsets' family of tests/test_gpu_converged.py, oracle vs device, with the iteration logs."""
import sys, os, ctypes
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from multiview_b200 import _abi, synth
import oracle.binding  # noqa: F401
from test_gpu_parity_solve import CASES, _param_error
kw = dict(CASES["offsets"]); kw.update(tri_weight=5.0, tri_robust_threshold=1.0, outlier_frac=0.0)
pr = synth.make_problem(seed=23, **kw)
OPTS = dict(max_num_iterations=int(sys.argv[1]) if len(sys.argv) > 1 else 60, eta=1e-8, max_trust_region_radius=1e7)
snap = pr.snapshot()
print("bounds", pr.offset_lo, pr.offset_hi, "start", pr.offsets)
olib = ctypes.CDLL(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "liboracle.so"))
o = _abi.Binding(olib, "oracle_", pr)
so, lo = o.solve(_abi.default_options(**OPTS), jac_mode=1, num_threads=1)
xo = pr.snapshot(); o.close(); pr.restore(snap)
g = _abi.Binding(_abi.load_product(), "mvba_", pr)
sg, lg = g.solve(_abi.default_options(**OPTS))
xg = pr.snapshot(); g.close()
print("final cost", so.final_cost, sg.final_cost, "offsets", xo["offsets"], xg["offsets"])
for k in xo:
    if xo[k].size: print(k, _param_error(k, xo[k], xg[k], pr))
for i, (a, b) in enumerate(zip(lo, lg)):
    print(i, "%.10e %.10e acc %d %d cg %d %d rad %.2e %.2e step %.2e %.2e" % (a["cost"], b["cost"], a["step_accepted"], b["step_accepted"], a["linear_iterations"], b["linear_iterations"], a["radius"], b["radius"], a["step_norm"], b["step_norm"]))
