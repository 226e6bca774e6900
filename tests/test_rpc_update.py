"""CameraParameters::updateRpcUndistortion (camera_model/camera_params.cc:214-254 -> fitRpcUndist, rpc_distortion.cc:658-719)
in the host mirror (multiview_b200/host/camera_params.h + rpc_pairs.h): after a pass that floated RPC distortion the reference
re-fits the UNDISTORTION half of the coefficient vector from the distortion half (rig_calibrator.cc:1946-1948).  The fit itself
is the device's mvba_rpc_fit, registered as a function pointer; on CPU the same member is driven with a Python callback that
forwards to the oracle's fit, so the sampling, the pair construction and the copy-back are covered without a GPU."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from multiview_b200 import _abi
from test_rpc_fit import grid, radtan, rpc_eval

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FITTER = C.CFUNCTYPE(C.c_int, C.c_int32, C.c_int64, C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_int32, C.c_int32, C.c_double,
                     C.c_int32, C.POINTER(C.c_double), C.POINTER(C.c_double))
dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)
SIZE = np.array([1280, 960], np.int32); UNDIST = np.array([1400, 1050], np.int32); CROP = np.array([1280, 960], np.int32)
FOCAL = np.array([600.0, 600.0]); CENTER = np.array([640.0, 480.0])
DEG = 2


def _host():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "multiview_b200", "host")])
    return C.CDLL(os.path.join(ROOT, "multiview_b200", "libmvba_host.so"))


def _dist_half(oracle_lib):
    u = grid(40, half=(700.0, 525.0))
    c, _ = _abi.rpc_fit(oracle_lib, "oracle_", u, radtan(u), DEG, num_threads=0)
    return c


def _update(host, dist, fitter):
    rc = host.mvh_update_rpc_undistortion(SIZE.ctypes.data_as(ip), FOCAL.ctypes.data_as(dp), CENTER.ctypes.data_as(dp), dist.ctypes.data_as(dp),
                                          len(dist), UNDIST.ctypes.data_as(ip), CROP.ctypes.data_as(ip), fitter)
    host.mvh_camera_last_error.restype = C.c_char_p
    return rc, (host.mvh_camera_last_error() or b"").decode()


def _round_trip_error(dist):
    n = len(dist) // 2
    u = grid(25, half=(500.0, 380.0))
    d = rpc_eval(u, dist[:n])
    return np.abs(rpc_eval(d, dist[n:]) - u).max()


def test_update_refits_the_undistortion_half(oracle_lib):
    host = _host()
    seen = {}

    @FITTER
    def via_oracle(device, n, in_xy, out_xy, deg, iters, ptol, jac_mode, coeffs, stats):
        seen.update(n=n, deg=deg, iters=iters, ptol=ptol)
        return oracle_lib.oracle_rpc_fit(C.c_int64(n), in_xy, out_xy, deg, 20, C.c_double(ptol), 0, coeffs, stats)   # 20 iterations keep the CPU test short

    half = _dist_half(oracle_lib)
    dist = np.concatenate([half, np.zeros_like(half)])
    dist[len(half) + 1] = 1.0; dist[len(half) + 2 * ((DEG + 1) * (DEG + 2) // 2) - 1 + 2] = 1.0     # identity undistortion to start from
    before = _round_trip_error(dist)
    rc, msg = _update(host, dist, C.cast(via_oracle, C.c_void_p))
    assert rc == 0, msg
    assert np.array_equal(dist[:len(half)], half)                      # the distortion half is input only
    assert seen["deg"] == DEG and seen["iters"] == 100 and seen["ptol"] == 1e-12   # fitRpcUndist's settings
    assert 100000 < seen["n"] <= 400 * 400                               # the 400 x 400 grid minus samples outside the crop box
    after = _round_trip_error(dist)
    assert before > 50.0 and after < 3.0, (before, after)    # a degree-2 ratio inverts a degree-2 ratio only approximately (1.4 px in the far corners)


def test_update_without_a_fitter_fails_loudly():
    host = _host()
    dist = np.zeros(2 * _abi.rpc_num_params(DEG)); dist[1] = 1.0
    rc, msg = _update(host, dist, None)
    assert rc == -1 and "no RPC fitter" in msg


@pytest.mark.gpu
def test_update_on_the_device_matches_the_oracle_driven_update(oracle_lib, product_lib):
    host = _host()

    @FITTER
    def via_oracle(device, n, in_xy, out_xy, deg, iters, ptol, jac_mode, coeffs, stats):
        return oracle_lib.oracle_rpc_fit(C.c_int64(n), in_xy, out_xy, deg, iters, C.c_double(ptol), 0, coeffs, stats)

    half = _dist_half(oracle_lib)
    a = np.concatenate([half, np.zeros_like(half)]); b = a.copy()
    rc, msg = _update(host, a, C.cast(via_oracle, C.c_void_p)); assert rc == 0, msg
    rc, msg = _update(host, b, C.cast(product_lib.mvba_rpc_fit, C.c_void_p)); assert rc == 0, msg
    assert _round_trip_error(b) < 3.0
    u = grid(25, half=(500.0, 380.0))
    assert np.abs(rpc_eval(u, a[len(half):]) - rpc_eval(u, b[len(half):])).max() <= 1e-3
