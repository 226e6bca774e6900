"""Dev tool: repeat one parity case on the GPU and print per-iteration cost / CG iterations next to the oracle's."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from multiview_b200 import synth, _abi
import oracle.binding  # noqa: F401
from test_gpu_parity_solve import CASES
name = sys.argv[1]; reps = int(sys.argv[2]); iters = int(sys.argv[3])
olib = C.CDLL(os.path.join(_abi.ROOT, "oracle", "liboracle.so")); plib = _abi.load_product()
pr = synth.make_problem(seed=23, **CASES[name]); snap = pr.snapshot()
o = _abi.Binding(olib, "oracle_", pr); so, lo = o.solve(_abi.default_options(max_num_iterations=iters), jac_mode=1); o.close()
print("oracle", [(f"{r['cost']:.8e}", r['linear_iterations'], r['step_accepted']) for r in lo])
for k in range(reps):
    pr.restore(snap)
    g = _abi.Binding(plib, "mvba_", pr); sg, lg = g.solve(_abi.default_options(max_num_iterations=iters)); g.close()
    bad = [i for i, (a, b) in enumerate(zip(lg, lo)) if abs(a['cost'] - b['cost']) > 1e-6 * b['cost'] or a['linear_iterations'] != b['linear_iterations']]
    print(k, "first differing iteration", bad[:1], [(f"{r['cost']:.8e}", r['linear_iterations']) for r in lg][bad[0] - 1 if bad else 0:(bad[0] + 2 if bad else 1)])
