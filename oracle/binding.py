"""ctypes binding of the CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE: only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / reference arm import this.  It subclasses the product's binder (same setters, same problem layout) and
registers itself for the ``oracle_`` symbol prefix, so `_abi.Binding(oracle_lib, "oracle_", problem)` keeps working in the tests
while multiview_b200/_abi.py itself carries no oracle code."""
import ctypes as C
import os

import numpy as np

from multiview_b200 import _abi
from multiview_b200._abi import _p, c_double_p

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class OracleBinding(_abi.Binding):
    def _create(self, cfg, device):
        self.lib.oracle_create.restype = C.c_void_p
        self.h = C.c_void_p(self.lib.oracle_create(C.byref(cfg)))

    def _evaluate_jacobian(self, kind, first, count, mode, r, J):
        self._ck(self.lib.oracle_evaluate_jacobian(self.h, C.c_int32(kind), C.c_int64(first), C.c_int64(count), C.c_int32(mode),
                                                   _p(r, c_double_p), _p(J, c_double_p)))

    def _solve(self, o, s, its, max_log, jac_mode, num_threads):
        self._ck(self.lib.oracle_solve(self.h, C.byref(o), C.byref(s), its, C.c_int32(max_log), C.c_int32(jac_mode), C.c_int32(num_threads)))


def _rpc_fit(lib, a, b, deg, num_iterations, parameter_tolerance, num_threads, c, st):
    return lib.oracle_rpc_fit(C.c_int64(len(a)), _p(a, c_double_p), _p(b, c_double_p), C.c_int32(deg), C.c_int32(num_iterations),
                              C.c_double(parameter_tolerance), C.c_int32(num_threads), _p(c, c_double_p), _p(st, c_double_p))


def compute_rpc(oracle_lib, pts, coeffs):
    """compute_rpc (rpc_distortion.cc:116-160) of the oracle at an array of points."""
    pts = np.ascontiguousarray(pts, np.float64); c = np.ascontiguousarray(coeffs, np.float64)
    out = np.zeros_like(pts)
    for i in range(len(pts)):
        oracle_lib.oracle_compute_rpc(_p(pts[i], c_double_p), _p(c, c_double_p), C.c_int32(len(c)), _p(out[i], c_double_p))
    return out


def load():
    return C.CDLL(os.path.join(ROOT, "oracle", "liboracle.so"))


_abi.register_checker("oracle_", OracleBinding, _rpc_fit)
