"""Dev tool: per-iteration logs of the CUDA solver and the CPU oracle on one tests/test_gpu_parity_solve.py case."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import ctypes as C
from multiview_b200 import synth, _abi
import oracle.binding  # noqa: F401
from test_gpu_parity_solve import CASES

def main():
    name = sys.argv[1]; iters = int(sys.argv[2]) if len(sys.argv) > 2 else 12
    seed = int(sys.argv[3]) if len(sys.argv) > 3 else 23
    olib = C.CDLL(os.path.join(_abi.ROOT, "oracle", "liboracle.so")); plib = _abi.load_product()
    pr = synth.make_problem(seed=seed, **CASES[name]); snap = pr.snapshot()
    o = _abi.Binding(olib, "oracle_", pr); so, lo = o.solve(_abi.default_options(max_num_iterations=iters), jac_mode=1); xo = pr.snapshot(); o.close()
    pr.restore(snap)
    g = _abi.Binding(plib, "mvba_", pr); sg, lg = g.solve(_abi.default_options(max_num_iterations=iters)); xg = pr.snapshot(); g.close()
    for i in range(min(len(lo), len(lg))):
        print(f"{i:3d} o {lo[i]['cost']:.12e} r={lo[i]['radius']:.3e} cg={lo[i]['linear_iterations']:3d} a={lo[i]['step_accepted']} |g|={lo[i]['gradient_max_norm']:.3e} | "
              f"g {lg[i]['cost']:.12e} r={lg[i]['radius']:.3e} cg={lg[i]['linear_iterations']:3d} a={lg[i]['step_accepted']} step={lg[i]['step_norm']:.3e}")
    for k in xo:
        if xo[k].size: print(f"   max|d {k}| = {np.abs(xo[k]-xg[k]).max():.3e}")
if __name__ == "__main__":
    main()
