"""Dev tool: run the CUDA solver and the CPU oracle on the same synthetic pass and print both logs."""
import ctypes as C, sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from multiview_b200 import synth, _abi
import oracle.binding  # noqa: F401

def main():
    cfgs = [int(a) for a in sys.argv[1].split(",")] if len(sys.argv) > 1 else [1, 2, 3, 4]
    scale = float(sys.argv[2]) if len(sys.argv) > 2 else 0.02
    iters = int(sys.argv[3]) if len(sys.argv) > 3 else 15
    eta = float(sys.argv[4]) if len(sys.argv) > 4 else 0.1
    pre = int(sys.argv[5]) if len(sys.argv) > 5 else 0
    olib = C.CDLL(os.path.join(_abi.ROOT, "oracle", "liboracle.so"))
    plib = _abi.load_product()
    for n in cfgs:
        pr = synth.config(n, scale=scale)
        snap = pr.snapshot()
        o = _abi.Binding(olib, "oracle_", pr)
        t = time.time(); so, lo = o.solve(_abi.default_options(max_num_iterations=iters, eta=eta), jac_mode=1); to = time.time() - t
        xo = pr.snapshot(); o.close()
        pr.restore(snap)
        g = _abi.Binding(plib, "mvba_", pr)
        t = time.time(); sg, lg = g.solve(_abi.default_options(max_num_iterations=iters, eta=eta, preconditioner=pre)); tg = time.time() - t
        xg = pr.snapshot(); g.close()
        print(f"== {pr.name} obs={pr.num_obs} P={pr.num_landmarks}")
        print(f"  oracle: {so.initial_cost:.9e} -> {so.final_cost:.9e} it={so.num_iterations} ok={so.num_successful_steps} cg={so.num_linear_iterations} {to:.2f}s")
        print(f"  cuda  : {sg.initial_cost:.9e} -> {sg.final_cost:.9e} it={sg.num_iterations} ok={sg.num_successful_steps} cg={sg.num_linear_iterations} {tg:.2f}s"
              f" dev={sg.seconds_device:.4f}s build={sg.seconds_build:.4f} lin={sg.seconds_linear:.4f} upd={sg.seconds_update:.4f} launches={sg.num_kernel_launches}")
        for i in range(min(len(lo), len(lg))):
            print(f"   {i:3d} o {lo[i]['cost']:.9e} r={lo[i]['radius']:.2e} cg={lo[i]['linear_iterations']:3d} a={lo[i]['step_accepted']} | "
                  f"g {lg[i]['cost']:.9e} r={lg[i]['radius']:.2e} cg={lg[i]['linear_iterations']:3d} a={lg[i]['step_accepted']} step={lg[i]['step_norm']:.2e}")
        for k in xo:
            if xo[k].size:
                print(f"   max|d {k}| = {np.abs(xo[k]-xg[k]).max():.3e}")

if __name__ == "__main__":
    main()
