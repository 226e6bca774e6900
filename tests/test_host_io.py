"""Host-side file formats (multiview_b200/host, C++): rig_config.txt, cameras.txt, NVM_V3.

The expected texts are laid out here, independently of the C++ code, the way the reference writers produce them
(dense_map_utils.cc:753-850, interest_point.cc:1405-1475: `key: values`, 17 significant digits in stream default
format, rotation row-major then translation) -- so a read -> write round trip must reproduce them byte for byte."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def host():
    path = os.path.join(ROOT, "multiview_b200", "libmvba_host.so")
    if not os.path.exists(path):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "multiview_b200", "host")])
    lib = C.CDLL(path)
    lib.mvh_last_error.restype = C.c_char_p
    return lib


def g17(v):
    return "%.17g" % v


def _sensor_text(name, f, c, dist, kind, size, crop, usize, r2s, d2i, off):
    j = lambda v: " ".join(g17(x) for x in v)
    return (f"\nsensor_name: {name}\nfocal_length: {g17(f)}\noptical_center: {j(c)}\ndistortion_coeffs: {j(dist)}\n"
            f"distortion_type: {kind}\nimage_size: {size[0]} {size[1]}\ndistorted_crop_size: {crop[0]} {crop[1]}\n"
            f"undistorted_image_size: {usize[0]} {usize[1]}\nref_to_sensor_transform: {j(r2s)}\n"
            f"depth_to_image_transform: {j(d2i)}\nref_to_sensor_timestamp_offset: {g17(off)}\n")


_I = [1, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0, 0]
_c, _s = np.cos(np.deg2rad(1.0)), np.sin(np.deg2rad(1.0))
RIG = ("ref_sensor_name: nav_cam\n"
       + _sensor_text("nav_cam", 608.04, [640.5, 480.2], [0.998], "fisheye", (1280, 960), (1200, 900), (1500, 1200), _I, _I, 0.0)
       + _sensor_text("haz_cam", 210.12345678901234, [112.25, 85.5], [-0.25, 0.08, 0.0012, -0.0008], "radtan", (224, 171), (224, 171),
                      (250, 200), [_c, -_s, 0, _s, _c, 0, 0, 0, 1, 0.05, -0.02, 0.003], [0.99, 0, 0, 0, 0.99, 0, 0, 0, 0.99, 0.001, 0.002, -0.003], 0.0159)
       + _sensor_text("sci_cam", 1000.0, [668, 501], [], "no_distortion", (1336, 1002), (1336, 1002), (1336, 1002),
                      [1, 0, 0, 0, 1, 0, 0, 0, 1, -0.1, 0, 0], _I, -0.0137))


def test_rig_config_round_trip_is_byte_exact(host, tmp_path):
    src = tmp_path / "rig_config.txt"; dst = tmp_path / "out.txt"
    # comments and blank lines are ignored by the reader
    src.write_text("# a comment\n" + RIG.replace("sensor_name: haz_cam", "sensor_name: haz_cam   # the depth camera"))
    n = C.c_int()
    assert host.mvh_rig_config_roundtrip(str(src).encode(), str(dst).encode(), 1, 1, C.byref(n)) == 0, host.mvh_last_error()
    assert n.value == 3
    assert dst.read_text() == RIG
    # model_rig = false writes identity rig transforms
    assert host.mvh_rig_config_roundtrip(str(src).encode(), str(dst).encode(), 1, 0, C.byref(n)) == 0
    lines = [l for l in dst.read_text().splitlines() if l.startswith("ref_to_sensor_transform:")]
    assert all(l == "ref_to_sensor_transform: 1 0 0 0 1 0 0 0 1 0 0 0" for l in lines)


@pytest.mark.parametrize("edit, message", [
    (("sensor_name: nav_cam", "sensor_name: other_cam"), "reference sensor must be the first"),
    (("distortion_type: fisheye", "distortion_type: radtan"), "distortion type must be: fisheye"),
    (("optical_center: 112.25 85.5", "optical_center: 112.25"), "incorrect number of values for: optical_center:"),
    (("image_size: 224 171\n", ""), "Could not read value for: image_size:"),
    (("ref_to_sensor_transform: 1 0 0 0 1 0 0 0 1 -0.10000000000000001 0 0", "ref_to_sensor_transform: 0 0 0 0 0 0 0 0 0 0 0 0"),
     "valid transforms among the sensors"),
])
def test_rig_config_validation(host, tmp_path, edit, message):
    src = tmp_path / "bad.txt"
    assert edit[0] in RIG
    src.write_text(RIG.replace(edit[0], edit[1], 1))
    assert host.mvh_rig_config_roundtrip(str(src).encode(), str(tmp_path / "o.txt").encode(), 1, 1, None) == -1
    assert message in host.mvh_last_error().decode()


def test_camera_list_round_trip(host, tmp_path):
    rng = np.random.default_rng(3)
    rows = []
    for i in range(5):
        vals = rng.normal(size=12)
        rows.append(f"data/{['nav_cam', 'haz_cam', 'sci_cam'][i % 3]}/{10000.5 + i:.1f}.jpg " + " ".join(g17(v) for v in vals))
    text = "# image_file world_to_image_transform\n" + "\n".join(rows) + "\n"
    src = tmp_path / "cameras.txt"; dst = tmp_path / "out.txt"
    src.write_text(text)
    n = C.c_int(); stamp = C.c_double(); sensor = C.create_string_buffer(64)
    assert host.mvh_camera_list_roundtrip(str(src).encode(), str(dst).encode(), C.byref(n), C.byref(stamp), sensor, 64) == 0
    assert n.value == 5 and stamp.value == 10000.5 and sensor.value == b"nav_cam"
    assert dst.read_text() == text
    src.write_text(text + "data/nav_cam/1.jpg 1 2 3\n")
    assert host.mvh_camera_list_roundtrip(str(src).encode(), str(dst).encode(), None, None, None, 0) == -1
    assert b"Expecting 12 values" in host.mvh_last_error()


def _quat_wxyz(R):
    from scipy.spatial.transform import Rotation
    x, y, z, w = Rotation.from_matrix(R).as_quat()
    q = np.array([w, x, y, z])
    return q if w >= 0 else -q


def test_nvm_export_and_round_trip(host, tmp_path):
    """Inlier export (optical-centre shift, per-camera renumbering, the `> 2` views rule) checked against a numpy replay,
    then NVM -> read -> write reproduces the file byte for byte."""
    from scipy.spatial.transform import Rotation
    rng = np.random.default_rng(9)
    C_, T, P = 6, 2, 8
    cam_sensor = np.array([0, 1, 0, 1, 0, 1], np.int32)
    focal = np.array([600.0, 210.5]); center = np.array([[320.25, 240.5], [112.0, 85.75]])
    Rm = Rotation.from_rotvec(rng.normal(size=(C_, 3)) * 0.3).as_matrix()
    t = rng.normal(size=(C_, 3))
    w2c = np.concatenate([Rm.reshape(C_, 9), t], axis=1)
    pid, cid = [], []
    for p in range(P):
        k = 2 + p % 4
        cs = np.sort(rng.choice(C_, size=k, replace=False))
        pid += [p] * k; cid += list(cs)
    pid = np.array(pid, np.int32); cid = np.array(cid, np.int32); N = len(pid)
    pix = rng.uniform(10, 200, (N, 2)).astype(np.float32)
    inlier = np.ones(N, np.uint8); inlier[3] = 0; inlier[-2] = 0
    xyz = rng.normal(size=(P, 3))
    names = tmp_path / "names.txt"
    image_files = [f"data/cam{cam_sensor[c]}/{1000 + c}.jpg" for c in range(C_)]
    names.write_text("\n".join(image_files) + "\n")
    out = tmp_path / "cameras.nvm"
    dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)
    rc = host.mvh_nvm_export(str(names).encode(), str(out).encode(), C_, cam_sensor.ctypes.data_as(ip), T, focal.ctypes.data_as(dp),
                             np.ascontiguousarray(center).ctypes.data_as(dp), np.ascontiguousarray(w2c).ctypes.data_as(dp), C.c_longlong(N),
                             pid.ctypes.data_as(ip), cid.ctypes.data_as(ip), pix.ctypes.data_as(C.POINTER(C.c_float)),
                             inlier.ctypes.data_as(C.POINTER(C.c_ubyte)), P, np.ascontiguousarray(xyz).ctypes.data_as(dp))
    assert rc == 0, host.mvh_last_error()
    toks = out.read_text().split()
    assert toks[0] == "NVM_V3" and int(toks[1]) == C_
    pos = 2
    for c in range(C_):
        rec = toks[pos:pos + 11]; pos += 11
        assert rec[0] == image_files[c] and float(rec[1]) == focal[cam_sensor[c]] and rec[9:] == ["0", "0"]
        q = np.array([float(v) for v in rec[2:6]]); ctr = np.array([float(v) for v in rec[6:9]])
        q = q if q[0] >= 0 else -q
        assert np.abs(q - _quat_wxyz(Rm[c])).max() < 1e-12
        assert np.abs(ctr - (-Rm[c].T @ t[c])).max() < 1e-12
    # numpy replay of the export
    fid_count = np.zeros(C_, int); tracks = []
    for p in range(P):
        cf = {}
        for i in np.flatnonzero(pid == p):
            if inlier[i]:
                cf[int(cid[i])] = (int(fid_count[cid[i]]), pix[i].astype(np.float64) - center[cam_sensor[cid[i]]]); fid_count[cid[i]] += 1
        if len(cf) > 2:
            tracks.append((p, cf))
    assert int(toks[pos]) == len(tracks) and 0 < len(tracks) < P; pos += 1
    for p, cf in tracks:
        assert [float(v) for v in toks[pos:pos + 3]] == list(xyz[p]) and toks[pos + 3:pos + 6] == ["0", "0", "0"] and int(toks[pos + 6]) == len(cf)
        pos += 7
        for c in sorted(cf):
            assert int(toks[pos]) == c and int(toks[pos + 1]) == cf[c][0]
            assert [float(toks[pos + 2]), float(toks[pos + 3])] == list(cf[c][1])
            pos += 4
    assert pos == len(toks)
    # read -> write: identical file (quaternion -> matrix -> quaternion is stable to the printed digits only if exact; compare numerically)
    out2 = tmp_path / "again.nvm"
    nc, npnt = C.c_int(), C.c_int()
    assert host.mvh_nvm_roundtrip(str(out).encode(), str(out2).encode(), C.byref(nc), C.byref(npnt)) == 0, host.mvh_last_error()
    assert nc.value == C_ and npnt.value == len(tracks)
    a, b = out.read_text().split(), out2.read_text().split()
    assert len(a) == len(b)
    for x, y in zip(a, b):
        if x != y:
            assert abs(float(x) - float(y)) <= 1e-12 * max(1.0, abs(float(x)))


# ---------------------------------------------------------------------------------------------
# pinned on the reference's own Python parsers (rig_calibrator/bin/rig_utils.py): tests/golden/host_formats.json holds what
# parseRigConfig and parse_images_and_camera_poses read back from files written by multiview_b200/host
# (tools/make_golden_hostio.py); the values must be the ones the files were composed from.
# ---------------------------------------------------------------------------------------------
def _check_rig_parse(parsed):
    want = [("nav_cam", 608.04, [640.5, 480.2], [0.998], "fisheye", (1280, 960), (1200, 900), (1500, 1200), _I, _I, 0.0),
            ("haz_cam", 210.12345678901234, [112.25, 85.5], [-0.25, 0.08, 0.0012, -0.0008], "radtan", (224, 171), (224, 171), (250, 200),
             [_c, -_s, 0, _s, _c, 0, 0, 0, 1, 0.05, -0.02, 0.003], [0.99, 0, 0, 0, 0.99, 0, 0, 0, 0.99, 0.001, 0.002, -0.003], 0.0159),
            ("sci_cam", 1000.0, [668, 501], [], "no_distortion", (1336, 1002), (1336, 1002), (1336, 1002), [1, 0, 0, 0, 1, 0, 0, 0, 1, -0.1, 0, 0], _I,
             -0.0137)]
    assert len(parsed) == 3
    f = lambda v: [float(x) for x in v]
    for cam, (name, fl, c, dist, kind, size, crop, usize, r2s, d2i, off) in zip(parsed, want):
        assert cam["sensor_name"] == name and cam["distortion_type"] == kind
        assert float(cam["focal_length"]) == fl and f(cam["optical_center"]) == list(c) and f(cam["distortion_coeffs"]) == list(dist)
        assert [int(v) for v in cam["image_size"]] == list(size) and [int(v) for v in cam["distorted_crop_size"]] == list(crop)
        assert [int(v) for v in cam["undistorted_image_size"]] == list(usize)
        assert f(cam["ref_to_sensor_transform"]) == [float(v) for v in r2s] and f(cam["depth_to_image_transform"]) == [float(v) for v in d2i]
        assert float(cam["ref_to_sensor_timestamp_offset"]) == off


def test_reference_parsers_read_our_files_golden():
    import json
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "host_formats.json")))
    _check_rig_parse(gold["rig_config"])
    rng = np.random.default_rng(3)
    rows = [rng.normal(size=12) for _ in range(5)]
    names = [f"data/{['nav_cam', 'haz_cam', 'sci_cam'][i % 3]}/{10000.5 + i:.1f}.jpg" for i in range(5)]
    for sensor, got in gold["camera_list"].items():
        idx = [i for i in range(5) if f"/{sensor}/" in names[i]]
        assert got["images"] == [names[i] for i in idx]
        for i, M in zip(idx, got["world_to_cam"]):
            M = np.array(M)
            assert np.array_equal(M[:3, :3].ravel(), rows[i][:9]) and np.array_equal(M[:3, 3], rows[i][9:]) and np.array_equal(M[3], [0, 0, 0, 1])


def test_reference_parser_reads_our_rig_config_live(host, tmp_path):
    ref_bin = "/root/reference/rig_calibrator/bin"
    if not os.path.exists(os.path.join(ref_bin, "rig_utils.py")):
        pytest.skip("the reference is not present on this box; tests/golden/host_formats.json covers this")
    import sys, warnings
    sys.path.insert(0, ref_bin)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        import rig_utils
    src = tmp_path / "in.txt"; dst = tmp_path / "rig_config.txt"
    src.write_text(RIG)
    assert host.mvh_rig_config_roundtrip(str(src).encode(), str(dst).encode(), 1, 1, None) == 0
    _check_rig_parse(rig_utils.parseRigConfig(str(dst)))
