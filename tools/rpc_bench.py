"""Times mvba_rpc_fit (GPU) against oracle_rpc_fit (CPU port of fitRpcDist, all host threads) on the fit that
updateRpcUndistortion runs per sensor between passes: num_samples x num_samples pairs, degree continuation up to `deg`,
100 LM iterations per degree.  Prints one JSON line; not part of bench.py's headline metric."""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from multiview_b200 import _abi  # noqa: E402
import oracle.binding  # noqa: F401
from test_rpc_fit import grid, radtan, rpc_eval  # noqa: E402

ns = int(sys.argv[1]) if len(sys.argv) > 1 else 400
deg = int(sys.argv[2]) if len(sys.argv) > 2 else 4
ns_cpu = int(sys.argv[3]) if len(sys.argv) > 3 else 120
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
olib = C.CDLL(os.path.join(root, "oracle", "liboracle.so"))
plib = _abi.load_product()
u = grid(ns); d = radtan(u)
_abi.rpc_fit(plib, "mvba_", u[:1000], d[:1000], 1)          # context creation, module load
t = time.time(); cg, sg = _abi.rpc_fit(plib, "mvba_", u, d, deg); t_gpu = time.time() - t
t = time.time(); ce, se = _abi.rpc_fit(plib, "mvba_", u, d, deg, jacobian_mode=1); t_gpu_exact = time.time() - t
uc = grid(ns_cpu); dc = radtan(uc)
t = time.time(); co, so = _abi.rpc_fit(olib, "oracle_", uc, dc, deg); t_cpu = time.time() - t
cg2, sg2 = _abi.rpc_fit(plib, "mvba_", uc, dc, deg)
print(json.dumps({"samples": len(u), "degree": deg, "gpu_s": t_gpu, "gpu_lm_iterations": sg["iterations"],
                  "gpu_ms_per_iteration": 1e3 * t_gpu / max(1, sg["iterations"]), "gpu_rms_px": float(np.sqrt(2 * sg["final_cost"] / len(u))),
                  "gpu_exact_s": t_gpu_exact, "gpu_exact_rms_px": float(np.sqrt(2 * se["final_cost"] / len(u))),
                  "cpu_samples": len(uc), "cpu_s": t_cpu, "cpu_lm_iterations": so["iterations"], "cpu_threads": olib.oracle_max_threads(),
                  "cpu_s_scaled_to_gpu_samples": t_cpu * len(u) / len(uc),
                  "same_sample_cost_rel_diff": abs(sg2["final_cost"] - so["final_cost"]) / so["final_cost"],
                  "same_sample_fn_diff_px": float(np.abs(rpc_eval(uc, cg2) - rpc_eval(uc, co)).max())}))
