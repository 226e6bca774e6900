"""GPU: mvba_rpc_fit_static, the RPC fit for callers that may not allocate device memory (include/mvba.h: its scratch is one
file-scope `static __device__` array of the library; no counterpart in the reference),
against mvba_rpc_fit, the allocating entry point -- through the C ABI, on calls in a row that share the
array:
  1. 3600 samples, degree 3: both entry points launch the same grid, so the results must be the same BITS;
  2. 900 other samples, degree 2, straight after: what call 1 left in the array must not leak into this fit;
  3. 40000 samples, degree 4: more samples than MVBA_RPC_STATIC_MAX_PARTS CTAs would take at 64 each, the partial sums are
     folded in another order than the allocating entry point's -- compared with the tolerances tests/test_rpc_fit.py uses
     between two implementations of the fit (cost 1e-6 relative, fitted function 1e-4 px, coefficients 1e-4 relative);
  4. call 1 again: the same bits as the first time.
Sizes beyond MVBA_RPC_STATIC_MAX_N are refused with a message that names the limit.

The file name sorts last on purpose: a failure here must not hide other tests."""
import ctypes as C

import numpy as np
import pytest

from multiview_b200 import _abi
from test_rpc_fit import grid, radtan, rpc_eval

MAX_N = 65536      # MVBA_RPC_STATIC_MAX_N, include/mvba.h
MAX_PARTS = 96     # MVBA_RPC_STATIC_MAX_PARTS
dp = C.POINTER(C.c_double)


def _fit(lib, name, u, d, deg, jacobian_mode=1, num_iterations=100):
    u = np.ascontiguousarray(u, np.float64); d = np.ascontiguousarray(d, np.float64)
    c = np.zeros(_abi.rpc_num_params(deg)); st = np.zeros(4)
    rc = getattr(lib, name)(C.c_int32(0), C.c_int64(len(u)), u.ctypes.data_as(dp), d.ctypes.data_as(dp), C.c_int32(deg),
                            C.c_int32(num_iterations), C.c_double(1e-12), C.c_int32(jacobian_mode), c.ctypes.data_as(dp), st.ctypes.data_as(dp))
    return rc, c, st


def _same_fit(a, b, u, exact):
    (rca, ca, sa), (rcb, cb, sb) = a, b
    assert rca == 0 and rcb == 0
    if exact:
        assert np.array_equal(ca, cb) and np.array_equal(sa, sb)
        return
    assert abs(sa[0] - sb[0]) <= 1e-12 * sb[0]                       # initial cost: one pass over the samples
    assert abs(sa[1] - sb[1]) <= 1e-6 * sb[1]
    assert np.abs(rpc_eval(u, ca) - rpc_eval(u, cb)).max() <= 1e-4
    assert np.abs(ca - cb).max() <= 1e-4 * np.abs(cb).max()


@pytest.mark.gpu
def test_static_scratch_fit_equals_the_allocating_fit_on_calls_in_a_row(product_lib):
    lib = product_lib
    assert hasattr(lib, "mvba_rpc_fit_static")
    u1 = grid(60); d1 = radtan(u1)                                                   # 3600 samples: 57 CTAs on either path
    u2 = grid(30, half=(500.0, 300.0)); d2 = radtan(u2, k=(0.12, -0.03, -0.0007, 0.0011))
    u3 = grid(200); d3 = radtan(u3)                                                  # 40000 samples: 96 CTAs here, up to 296 there
    assert len(u1) <= 64 * MAX_PARTS < len(u3) <= MAX_N

    first = _fit(lib, "mvba_rpc_fit_static", u1, d1, 3)
    _same_fit(first, _fit(lib, "mvba_rpc_fit", u1, d1, 3), u1, exact=True)
    assert first[2][3] >= 1 and first[2][1] < first[2][0]                            # it did fit something

    second = _fit(lib, "mvba_rpc_fit_static", u2, d2, 2)                             # straight after, smaller, other data and degree
    _same_fit(second, _fit(lib, "mvba_rpc_fit", u2, d2, 2), u2, exact=True)

    third = _fit(lib, "mvba_rpc_fit_static", u3, d3, 4, jacobian_mode=0)
    _same_fit(third, _fit(lib, "mvba_rpc_fit", u3, d3, 4, jacobian_mode=0), u3, exact=False)

    again = _fit(lib, "mvba_rpc_fit_static", u1, d1, 3)
    _same_fit(again, first, u1, exact=True)


@pytest.mark.gpu
def test_static_scratch_fit_refuses_sizes_beyond_its_limit(product_lib):
    n = MAX_N + 1
    u = np.zeros((n, 2)); u[:, 0] = np.linspace(-1, 1, n); u[:, 1] = np.linspace(1, -1, n) ** 2
    rc, _, _ = _fit(product_lib, "mvba_rpc_fit_static", u, u.copy(), 1)
    assert rc == -1                                                                  # MVBA_EINVAL
    msg = product_lib.mvba_last_error().decode()
    assert "mvba_rpc_fit_static" in msg and str(MAX_N) in msg
    rc, c, st = _fit(product_lib, "mvba_rpc_fit_static", u[:MAX_N], u[:MAX_N].copy(), 1)   # the limit itself is served
    assert rc == 0 and st[1] <= 1e-20 * MAX_N


def test_static_scratch_fit_checks_its_arguments_before_it_touches_a_device(product_lib):
    """Runs without a GPU: the size limit and the argument checks come first, and there is no CPU fallback behind them."""
    u = np.zeros((4, 2))
    for n, deg, it in ((0, 1, 10), (4, 0, 10), (4, 7, 10), (4, 1, -1)):
        c = np.zeros(_abi.rpc_num_params(max(deg, 1))); st = np.zeros(4)
        rc = product_lib.mvba_rpc_fit_static(C.c_int32(0), C.c_int64(n), u.ctypes.data_as(dp), u.ctypes.data_as(dp), C.c_int32(deg), C.c_int32(it),
                                             C.c_double(1e-12), C.c_int32(0), c.ctypes.data_as(dp), st.ctypes.data_as(dp))
        assert rc == -1 and b"mvba_rpc_fit_static" in product_lib.mvba_last_error()
    big = np.zeros((MAX_N + 1, 2))
    rc, _, _ = _fit(product_lib, "mvba_rpc_fit_static", big, big, 1)
    assert rc == -1 and str(MAX_N).encode() in product_lib.mvba_last_error()
    if product_lib.mvba_device_count() == 0:
        rc, _, _ = _fit(product_lib, "mvba_rpc_fit_static", u, u, 1)
        assert rc == -2                                                              # MVBA_ENODEVICE
