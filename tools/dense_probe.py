"""Dev tool: repeat tests/test_gpu_dense.py's rig3_intr comparison and print the parameter differences dense-vs-PCG and dense-vs-oracle."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from multiview_b200 import synth, _abi
import oracle.binding  # noqa: F401
from test_gpu_dense import CASES, EXACT, DAMPED, _solve
from test_gpu_parity_solve import _param_error
olib = C.CDLL(os.path.join(ROOT, "oracle", "liboracle.so")); plib = _abi.load_product()
name = sys.argv[1] if len(sys.argv) > 1 else "rig3_intr"
pr = synth.make_problem(seed=17, **CASES[name]); snap = pr.snapshot()
so, lo, xo = _solve(olib, "oracle_", pr, snap, **EXACT, **DAMPED)
for rep in range(int(sys.argv[2]) if len(sys.argv) > 2 else 5):
    sd, ld, xd = _solve(plib, "mvba_", pr, snap, linear_solver=1, **DAMPED)
    sp, lp, xp = _solve(plib, "mvba_", pr, snap, **EXACT, **DAMPED)
    e = {k: (_param_error(k, xd[k], xp[k], pr), _param_error(k, xd[k], xo[k], pr), _param_error(k, xp[k], xo[k], pr)) for k in xd if xd[k].size}
    worst = max(e, key=lambda k: max(e[k]))
    print(rep, "worst", worst, ["%.2e" % v for v in e[worst]], "(dense-pcg, dense-oracle, pcg-oracle)", "cost rel", abs(ld[-1]["cost"] - lo[-1]["cost"]) / lo[-1]["cost"])
