"""camera::CameraParameters mirror (multiview_b200/host/camera_params.h) against the committed golden vectors:
UNDISTORTED_C -> DISTORTED from tests/golden/distortion.npz (mpmath, cv2.projectPoints) and DISTORTED -> UNDISTORTED_C
from tests/golden/undistort.npz (cv2.undistortPoints, mpmath), plus the frame algebra of the ten declared conversions and
the five intrinsic matrices (camera_params.cc:374-451)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
RAW, DISTORTED, DISTORTED_C, UNDISTORTED, UNDISTORTED_C = range(5)
dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)


@pytest.fixture(scope="module")
def host():
    path = os.path.join(ROOT, "multiview_b200", "libmvba_host.so")
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "multiview_b200", "host")])
    lib = C.CDLL(path)
    lib.mvh_camera_last_error.restype = C.c_char_p
    return lib


def convert(lib, size, focal, center, dist, src, dst, pts, undist_size=None, crop=None, want_k=False):
    size = np.asarray(size, np.int32); focal = np.asarray(focal, np.float64) * np.ones(2); center = np.asarray(center, np.float64)
    dist = np.asarray(dist, np.float64); pts = np.ascontiguousarray(pts, np.float64); out = np.zeros_like(pts)
    us = None if undist_size is None else np.asarray(undist_size, np.int32); cr = None if crop is None else np.asarray(crop, np.int32)
    k = np.zeros(9); mf = C.c_double()
    rc = lib.mvh_camera_convert(size.ctypes.data_as(ip), focal.ctypes.data_as(dp), center.ctypes.data_as(dp),
                                dist.ctypes.data_as(dp) if dist.size else None, int(dist.size),
                                None if us is None else us.ctypes.data_as(ip), None if cr is None else cr.ctypes.data_as(ip),
                                src, dst, len(pts), pts.ctypes.data_as(dp), out.ctypes.data_as(dp), k.ctypes.data_as(dp), C.byref(mf))
    if rc != 0:
        raise RuntimeError(lib.mvh_camera_last_error().decode())
    return (out, k.reshape(3, 3), mf.value) if want_k else out


@pytest.mark.parametrize("model", ["none", "fov", "radtan4", "radtan5", "rpc2", "rpc3"])
def test_distortion_direction(host, model):
    g = np.load(os.path.join(GOLD, "distortion.npz"))
    f, c, size, u = float(g["focal"]), g["center"], g["size"].astype(int), g["u"]
    dist = {"none": [], "fov": [float(g["fov_w"])]}.get(model, None)
    if dist is None:
        dist = g[model + "_coeffs"]
    got = convert(host, size, f, c, dist, UNDISTORTED_C, DISTORTED, u)
    assert np.abs(got - g[model]).max() <= 1e-9 * max(1.0, np.abs(g[model]).max())


@pytest.mark.parametrize("model", ["none", "fov", "radtan4", "radtan5", "rpc2"])
def test_undistortion_direction(host, model):
    g = np.load(os.path.join(GOLD, "undistort.npz"))
    f, c, size, pix = float(g["focal"]), g["center"], g["size"].astype(int), g["pix"]
    dist = {"none": [], "fov": [float(g["fov_w"])]}.get(model, None)
    if dist is None:
        dist = g[model + "_coeffs"]
    got = convert(host, size, f, c, dist, DISTORTED, UNDISTORTED_C, pix)
    assert np.abs(got - g[model]).max() <= 1e-9 * max(1.0, np.abs(g[model]).max())


def test_frame_algebra_and_intrinsics(host):
    rng = np.random.default_rng(4)
    size, usize, crop = (1280, 960), (1500, 1200), (16, 8)
    focal, center, dist = (608.0, 610.5), (640.5, 480.2), [0.998]
    pts = rng.uniform(100, 900, (50, 2))
    kw = dict(undist_size=usize, crop=crop)
    a = convert(host, size, focal, center, dist, RAW, DISTORTED, pts, **kw)
    assert np.allclose(a, pts - crop, rtol=0, atol=0)
    assert np.allclose(convert(host, size, focal, center, dist, DISTORTED, RAW, a, **kw), pts, rtol=0, atol=1e-12)
    uc = convert(host, size, focal, center, dist, DISTORTED, UNDISTORTED_C, pts, **kw)
    u = convert(host, size, focal, center, dist, DISTORTED, UNDISTORTED, pts, **kw)
    assert np.abs(u - (uc + np.array(usize) / 2.0)).max() <= 1e-12
    assert np.abs(convert(host, size, focal, center, dist, UNDISTORTED, UNDISTORTED_C, u, **kw) - uc).max() <= 1e-10
    assert np.abs(convert(host, size, focal, center, dist, UNDISTORTED_C, UNDISTORTED, uc, **kw) - u).max() <= 1e-10
    dc = convert(host, size, focal, center, dist, UNDISTORTED_C, DISTORTED_C, uc, **kw)
    assert np.abs(dc + np.array(size) / 2.0 - pts).max() <= 1e-8            # FOV model: distort(undistort(p)) = p
    assert np.abs(convert(host, size, focal, center, dist, UNDISTORTED, DISTORTED, u, **kw) - pts).max() <= 1e-8
    assert np.abs(convert(host, size, focal, center, dist, DISTORTED_C, UNDISTORTED_C, pts - np.array(size) / 2.0, **kw) - uc).max() <= 1e-10
    want_c = {RAW: np.array(center) + crop, DISTORTED: np.array(center), DISTORTED_C: np.array(center) - np.array(size) / 2.0,
              UNDISTORTED: np.array(usize) / 2.0, UNDISTORTED_C: np.zeros(2)}
    pairs = {RAW: DISTORTED, DISTORTED: RAW, DISTORTED_C: UNDISTORTED_C, UNDISTORTED: UNDISTORTED_C, UNDISTORTED_C: UNDISTORTED}
    for frame, dst in pairs.items():
        _, k, mf = convert(host, size, focal, center, dist, frame, dst, pts[:1], want_k=True, **kw)
        assert mf == (focal[0] + focal[1]) / 2.0                                 # GetFocalLength() is the mean of fx, fy
        assert np.allclose(k, [[focal[0], 0, want_c[frame][0]], [0, focal[1], want_c[frame][1]], [0, 0, 1]], rtol=0, atol=1e-12)


def test_undeclared_conversion_and_bad_distortion_length_are_errors(host):
    with pytest.raises(RuntimeError, match="explicitly specified conversions"):
        convert(host, (640, 480), 600.0, (320, 240), [], RAW, UNDISTORTED, np.zeros((1, 2)))
    with pytest.raises(RuntimeError, match="irregular distortion vector size"):
        convert(host, (640, 480), 600.0, (320, 240), [0.1, 0.2, 0.3], UNDISTORTED_C, DISTORTED, np.zeros((1, 2)))


def test_rpc_sample_pairs_follow_gen_undist_dist_pairs(host):
    """genUndistDistPairs (rpc_distortion.cc:497-553) replayed in numpy on top of the UNDISTORTED -> DISTORTED conversion."""
    size, focal, center = (224, 171), 210.0, (112.25, 85.5)
    dist = [-0.25, 0.08, 0.0012, -0.0008]; usize, crop, ns = (300, 240), (200, 150), 25
    sz = np.asarray(size, np.int32); fo = np.asarray([focal, focal]); ce = np.asarray(center); di = np.asarray(dist)
    us = np.asarray(usize, np.int32); cr = np.asarray(crop, np.int32)
    u_out = np.zeros(2 * ns * ns); d_out = np.zeros(2 * ns * ns)
    n = host.mvh_gen_undist_dist_pairs(sz.ctypes.data_as(ip), fo.ctypes.data_as(dp), ce.ctypes.data_as(dp), di.ctypes.data_as(dp), 4,
                                       us.ctypes.data_as(ip), cr.ctypes.data_as(ip), ns, u_out.ctypes.data_as(dp), d_out.ctypes.data_as(dp))
    xs = (usize[0] - 1.0) * np.arange(ns) / (ns - 1.0); ys = (usize[1] - 1.0) * np.arange(ns) / (ns - 1.0)
    X, Y = np.meshgrid(xs, ys, indexing="ij")
    und = np.stack([X.ravel(), Y.ravel()], 1)
    dis = convert(host, size, focal, center, dist, UNDISTORTED, DISTORTED, und, undist_size=usize)
    keep = (np.abs(dis[:, 0] - size[0] / 2.0) <= crop[0] / 2.0) & (np.abs(dis[:, 1] - size[1] / 2.0) <= crop[1] / 2.0)
    assert 0 < keep.sum() < ns * ns and n == keep.sum()
    assert np.array_equal(u_out[:2 * n].reshape(-1, 2), und[keep] - np.asarray(usize) / 2.0)
    assert np.array_equal(d_out[:2 * n].reshape(-1, 2), dis[keep] - np.asarray(size) / 2.0)


def test_depth_value_lookup(host):
    """depthValue (dense_map_utils.cc:1362-1391): nearest pixel with C round(), zero triple = no measurement, hitting the far
    bound after rounding = no measurement, beyond it = the reference aborts."""
    rows, cols = 4, 6
    cloud = np.arange(rows * cols * 3, dtype=np.float32).reshape(rows, cols, 3) + 1.0
    cloud[2, 3] = 0.0
    fp, ucp = C.POINTER(C.c_float), C.POINTER(C.c_ubyte)
    pix = np.array([[0.0, 0.0], [2.4, 1.6], [2.5, 0.5], [3.2, 1.9], [5.49, 3.49], [5.5, 1.0], [1.0, 3.5], [0.49, 0.49]])
    has = np.zeros(len(pix), np.uint8); xyz = np.zeros((len(pix), 3))
    assert host.mvh_depth_values(cloud.ctypes.data_as(fp), rows, cols, len(pix), pix.ctypes.data_as(dp), has.ctypes.data_as(ucp), xyz.ctypes.data_as(dp)) == 0
    # round half away from zero: (2.5, 0.5) -> col 3, row 1
    want_rc = [(0, 0), (2, 2), (1, 3), None, (3, 5), None, None, (0, 0)]
    for i, rc in enumerate(want_rc):
        if rc is None:
            assert has[i] == 0 and (xyz[i] == 0).all()
        else:
            assert has[i] == 1 and np.array_equal(xyz[i], cloud[rc].astype(np.float64))
    bad = np.array([[6.6, 1.0]])
    assert host.mvh_depth_values(cloud.ctypes.data_as(fp), rows, cols, 1, bad.ctypes.data_as(dp), has.ctypes.data_as(ucp), xyz.ctypes.data_as(dp)) == -1
    assert b"Out of range in the depth cloud" in host.mvh_camera_last_error()
    # empty cloud
    assert host.mvh_depth_values(None, 0, 0, 1, pix.ctypes.data_as(dp), has.ctypes.data_as(ucp), xyz.ctypes.data_as(dp)) == 0 and has[0] == 0


def test_compact_observations_keeps_order_and_ids(host):
    rng = np.random.default_rng(5)
    n = 200
    pid = np.sort(rng.integers(0, 40, n)).astype(np.int32); cid = rng.integers(0, 12, n).astype(np.int32)
    pix = rng.uniform(0, 500, (n, 2)).astype(np.float32); inl = (rng.random(n) > 0.3).astype(np.uint8)
    opid = np.zeros(n, np.int32); ocid = np.zeros(n, np.int32); opix = np.zeros((n, 2), np.float32); src = np.zeros(n, np.int64)
    host.mvh_compact_observations.restype = C.c_longlong
    k = host.mvh_compact_observations(C.c_longlong(n), pid.ctypes.data_as(ip), cid.ctypes.data_as(ip), pix.ctypes.data_as(C.POINTER(C.c_float)),
                                      inl.ctypes.data_as(C.POINTER(C.c_ubyte)), opid.ctypes.data_as(ip), ocid.ctypes.data_as(ip),
                                      opix.ctypes.data_as(C.POINTER(C.c_float)), src.ctypes.data_as(C.POINTER(C.c_longlong)))
    sel = np.flatnonzero(inl)
    assert k == len(sel) and np.array_equal(src[:k], sel)
    assert np.array_equal(opid[:k], pid[sel]) and np.array_equal(ocid[:k], cid[sel]) and np.array_equal(opix[:k], pix[sel])


def test_host_pose_math_matches_scipy_golden_and_flags_bad_brackets(host):
    """multiview_b200/host/pose_math.h (array_to_rigid_transform, linearInterp, calc_world_to_cam_trans on the host, for the
    camera list a driver writes after the solve) against tests/golden/pose_interp.npz (scipy Slerp / Rotation)."""
    g = np.load(os.path.join(GOLD, "pose_interp.npz"))
    zero, one = np.array([0], np.int32), np.array([1], np.int32)
    out = np.zeros(12)

    def w2c(b, e, r, bs, es, off, cs):
        w2r = np.ascontiguousarray(np.concatenate([b, e])); stamps = np.array([bs, es]); cstamp = np.array([cs]); offs = np.array([off])
        r = np.ascontiguousarray(r)
        rc = host.mvh_world_to_cam_all(1, zero.ctypes.data_as(ip), zero.ctypes.data_as(ip), one.ctypes.data_as(ip), cstamp.ctypes.data_as(dp),
                                       stamps.ctypes.data_as(dp), w2r.ctypes.data_as(dp), r.ctypes.data_as(dp), offs.ctypes.data_as(dp), out.ctypes.data_as(dp))
        return rc, np.concatenate([out[:9].reshape(3, 3), out[9:, None]], axis=1).ravel()      # -> 3x4 row-major like the fixture

    for i in range(len(g["bs"])):
        rc, m = w2c(g["beg"][i], g["end"][i], g["r2c"][i], g["bs"][i], g["es"][i], g["off"][i], g["cs"][i])
        assert rc == 0 and np.abs(m - g["world_to_cam"][i]).max() < 1e-11 * max(1.0, np.abs(g["world_to_cam"][i]).max())
    for i in range(len(g["degenerate"])):                   # beg stamp == end stamp: the beg pose itself
        rc, m = w2c(g["beg"][i], g["end"][i], g["r2c"][i], 5.0, 5.0, 0.3, 5.0)
        assert rc == 0 and np.abs(m - g["degenerate"][i]).max() < 1e-12 * max(1.0, np.abs(g["degenerate"][i]).max())
    p = np.array([0, 0, 0, 0, 0, 0, 1.0])
    rc, _ = w2c(p, p, p, 10.0, 11.0, 0.0, 11.5)
    assert rc == -1 and b"Out of bounds in interpolation" in host.mvh_camera_last_error()
    # array -> transform -> array keeps a unit quaternion (up to the sign Eigen picks) and normalises a scaled one
    rng = np.random.default_rng(4)
    for _ in range(20):
        a = rng.normal(size=7); a[3:] /= np.linalg.norm(a[3:])
        back = np.zeros(7); scaled = a.copy(); scaled[3:] *= 3.0
        host.mvh_rigid_roundtrip(scaled.ctypes.data_as(dp), back.ctypes.data_as(dp))
        s = np.sign(back[3:] @ a[3:])
        assert np.abs(back[:3] - a[:3]).max() == 0 and np.abs(s * back[3:] - a[3:]).max() < 1e-14


def camera_from_file(lib, path, base_dir=""):
    out = np.zeros(13); dist = np.zeros(128)
    rc = lib.mvh_camera_from_file(str(path).encode(), base_dir.encode(), out.ctypes.data_as(dp), dist.ctypes.data_as(dp), 128)
    if rc != 0:
        raise RuntimeError(lib.mvh_camera_last_error().decode())
    return out, dist[:int(out[12])]


def test_calibration_file_constructor(host, tmp_path):
    """CameraParameters(calibration_file, base_dir), camera_params.cc:50-104: a file written by cv::FileStorage itself (the reference
    reads it with cv::FileStorage), then the fall-back key names in the layout OpenCV's calibration tool writes, the default
    undistorted size and crop, the base_dir fall-back and the missing-file error."""
    cv2 = pytest.importorskip("cv2")
    K = np.array([[608.25, 0, 640.5], [0, 607.75, 480.125], [0, 0, 1]])
    dist = np.array([[-0.25, 0.08, 0.0012, -0.0008, 0.011]])
    p = tmp_path / "cam.yml"
    fs = cv2.FileStorage(str(p), cv2.FILE_STORAGE_WRITE)
    for k, v in (("width", 1280), ("height", 960), ("undistorted_width", 1500), ("undistorted_height", 1100), ("crop_x", 12), ("crop_y", 7)):
        fs.write(k, v)
    fs.write("intrinsic_matrix", K); fs.write("distortion", dist)
    fs.release()
    out, d = camera_from_file(host, p)
    assert list(out[:8]) == [1280, 960, 1280, 960, 1500, 1100, 12, 7]
    assert list(out[8:12]) == [608.25, 607.75, 640.5, 480.125] and np.array_equal(d, dist.ravel())
    q = tmp_path / "calib.yaml"
    q.write_text("%YAML:1.0\n---\nimage_width: 224\nimage_height: 171\ncamera_matrix: !!opencv-matrix\n   rows: 3\n   cols: 3\n   dt: d\n"
                 "   data: [ 2.1000000000000000e+02, 0., 1.1230000000000000e+02, 0.,\n       2.0950000000000000e+02, 8.5099999999999994e+01, 0., 0., 1. ]\n"
                 "distortion_coefficients: !!opencv-matrix\n   rows: 1\n   cols: 1\n   dt: d\n   data: [ 9.9800000000000000e-01 ]\n")
    out, d = camera_from_file(host, q)
    assert list(out[:8]) == [224, 171, 224, 171, 224, 171, 0, 0]          # undistorted size defaults to the distorted one, crop to zero
    assert list(out[8:12]) == [210.0, 209.5, 112.3, 85.1] and list(d) == [0.998]
    out2, _ = camera_from_file(host, "calib.yaml", str(tmp_path))          # relative to the map file's directory
    assert np.array_equal(out, out2)
    with pytest.raises(RuntimeError, match="Calibration file does not exist: nowhere.yml"):
        camera_from_file(host, "nowhere.yml")


def test_generate_remap_maps_matches_opencv(host):
    """GenerateRemapMaps, camera_params.cc:361-371: Convert<UNDISTORTED, DISTORTED> at every pixel of the undistorted image is the map
    cv::initUndistortRectifyMap builds for the new camera matrix [f 0 w'/2; 0 f h'/2] (the undistorted frame has its centre at half its
    size).  Also at scale 0.5."""
    cv2 = pytest.importorskip("cv2")
    size = np.array([224, 171], np.int32); us = np.array([260, 200], np.int32)
    focal = np.array([210.0, 210.0]); center = np.array([112.3, 85.1]); dist = np.array([-0.25, 0.08, 0.0012, -0.0008])
    K = np.array([[focal[0], 0, center[0]], [0, focal[1], center[1]], [0, 0, 1]])
    newK = np.array([[focal[0], 0, us[0] / 2.0], [0, focal[1], us[1] / 2.0], [0, 0, 1]])
    want, _ = cv2.initUndistortRectifyMap(K, dist, np.eye(3), newK, (int(us[0]), int(us[1])), cv2.CV_32FC2)
    for scale in (1.0, 0.5):
        buf = np.zeros((int(us[1]), int(us[0]), 2), np.float32)
        rows, cols = C.c_int(), C.c_int()
        rc = host.mvh_generate_remap_maps(size.ctypes.data_as(ip), focal.ctypes.data_as(dp), center.ctypes.data_as(dp), dist.ctypes.data_as(dp), 4,
                                          us.ctypes.data_as(ip), C.c_double(scale), buf.ctypes.data_as(C.POINTER(C.c_float)), C.c_longlong(buf.size // 2),
                                          C.byref(rows), C.byref(cols))
        assert rc == 0 and (rows.value, cols.value) == (int(scale * us[1]), int(scale * us[0]))
        got = buf.ravel()[:2 * rows.value * cols.value].reshape(rows.value, cols.value, 2)
        step = int(round(1 / scale))
        assert np.abs(got / scale - want[::step, ::step]).max() < 2e-3
