"""The command line of examples/rig_calibrator (multiview_b200/host/cli_flags.h) against the reference's own source: the 43 flags with
their types and defaults, and every message of parameterValidation(), are held to tests/golden/cli_flags.json, which
tools/make_golden_cli.py extracts from rig_calibrator/bin/rig_calibrator.cc:135-316, 893-955.  Then gflags' syntax, the
'what floats' parsers (dense_map_utils.cc:54-152) through --dry_run, the rig transforms derived from the poses
(calc_rig_using_word_to_cam, rig_calibrator.cc:1190-1262) and the .ply reader behind --mesh.  All on CPU."""
import ctypes as C
import json
import os
import struct
import subprocess

import numpy as np
import pytest

from test_rig_pass import dataset, run, ROOT  # noqa: F401  (dataset is a fixture)

GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "cli_flags.json")))


@pytest.fixture(scope="module")
def host():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "multiview_b200", "host")])
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "examples")])
    return C.CDLL(os.path.join(ROOT, "multiview_b200", "libmvba_host.so"))


def cli(*args):
    return subprocess.run([os.path.join(ROOT, "examples", "rig_calibrator"), *args], capture_output=True, text=True, timeout=120)


def test_flag_table_is_the_references(host):
    buf = C.create_string_buffer(8192)
    n = host.mvh_flag_table(buf, len(buf))
    assert n == 43 == len(GOLD["flags"])
    for line, want in zip(buf.value.decode().splitlines(), GOLD["flags"]):
        name, kind, default = line.split(" ", 2)
        assert (name, kind) == (want["name"], want["type"])
        got = {"d": float, "i": int, "b": lambda v: v == "true", "s": json.loads}[kind](default)
        assert got == want["default"], name


# one command line per LOG(FATAL) of parameterValidation, in the reference's order
OK = ["--rig_config", "r.txt", "--nvm", "m.nvm", "--out_dir", "o"]
VALIDATION = [
    OK + ["--robust_threshold", "0"],
    OK + ["--bracket_len", "-1"],
    ["--rig_config", "r.txt", "--out_dir", "o", "--camera_poses", "c.txt"],
    OK + ["--timestamp_offsets_max_change", "-0.5"],
    OK + ["--min_triangulation_angle", "0"],
    OK + ["--depth_tri_weight", "-1"],
    OK + ["--mesh_tri_weight", "-1"],
    OK + ["--depth_mesh_weight", "-1"],
    OK + ["--tri_weight", "-1"],
    OK + ["--tri_weight", "1"],
    OK + ["--registration"],
    OK + ["--float_scale", "--affine_depth_to_image"],
    OK + ["--no_rig", "--float_timestamp_offsets"],
    ["--rig_config", "r.txt", "--nvm", "m.nvm"],
    OK + ["--no_rig", "--use_initial_rig_transforms"],
    OK + ["--out_texture_dir", "t"],
    ["--nvm", "m.nvm", "--out_dir", "o"],
    OK + ["--camera_poses", "c.txt"],
    ["--rig_config", "r.txt", "--out_dir", "o", "--num_overlaps", "3"],
]


def test_every_validation_message_of_the_reference(host):
    assert len(VALIDATION) == len(GOLD["validation_messages"]) == 19
    for args, msg in zip(VALIDATION, GOLD["validation_messages"]):
        out = cli(*args)
        assert out.returncode == 2, (args, out.stderr)
        assert msg.strip() in out.stderr, (args, msg, out.stderr)


def test_gflags_syntax(host, dataset):
    d, pr = dataset
    base = [f"--rig_config={d / 'rig_config.txt'}", "-nvm", str(d / "map.nvm"), "--out_dir", str(d), "--dry_run", "--bracket_len=1.0"]
    assert cli(*base).returncode == 0
    assert cli(*base, "--noverbose", "--save_nvm=false", "--float_scale=yes", "-num_iterations=3").returncode == 0
    bad = cli(*base, "--no_such_flag")
    assert bad.returncode == 1 and "unknown command line flag 'no_such_flag'" in bad.stderr
    bad = cli(*base, "--num_iterations", "ten")
    assert bad.returncode == 1 and "illegal value 'ten' specified for int32 flag 'num_iterations'" in bad.stderr
    bad = cli(*base, "--float_scale=maybe")
    assert bad.returncode == 1 and "illegal value 'maybe'" in bad.stderr
    bad = cli(*base, "--robust_threshold")
    assert bad.returncode == 1 and "missing its argument" in bad.stderr


def test_out_of_scope_flags_are_refused_not_ignored(host, dataset):
    d, _ = dataset
    base = ["--rig_config", str(d / "rig_config.txt"), "--nvm", str(d / "map.nvm"), "--out_dir", str(d), "--dry_run"]
    for extra, word in ((["--num_overlaps", "5"], "--num_overlaps"), (["--save_matches"], "--save_matches"), (["--export_to_voxblox"], "--export_to_voxblox"),
                        (["--save_transformed_depth_clouds"], "--save_transformed_depth_clouds"),
                        (["--registration", "--hugin_file", "a.pto", "--xyz_file", "b.txt"], "--registration"),
                        (["--mesh", "m.ply", "--out_texture_dir", "t"], "--out_texture_dir")):
        out = cli(*base, *extra)
        assert out.returncode == 2 and word in out.stderr and "outside the bundle-adjustment path" in out.stderr, (extra, out.stderr)


def test_what_floats(host, dataset):
    d, pr = dataset
    names = [s.name for s in pr.sensors]
    out = run(d, "--dry_run", "--bracket_len", "1.0", "--intrinsics_to_float", f"{names[1]}:focal_length,optical_center {names[2]}:distortion,focal_length",
              "--rig_transforms_to_float", f"{names[1]},{names[2]}", "--camera_poses_to_float", names[0], "--depth_to_image_transforms_to_float", names[1],
              "--float_timestamp_offsets", "--float_scale")
    assert out.returncode == 0, out.stderr
    fl = json.loads(out.stdout)["float"]
    assert fl == {"focal_length": [0, 1, 1], "optical_center": [0, 1, 0], "distortion": [0, 0, 1], "rig": [0, 1, 1], "depth_to_image": [0, 1, 0],
                  "poses": [1, 0, 0], "timestamp_offsets": 1, "scale": 1}
    # defaults: nothing floats (rig_calibrator.cc:1702-1752 with empty sets)
    fl = json.loads(run(d, "--dry_run", "--bracket_len", "1.0").stdout)["float"]
    assert all(v == [0, 0, 0] for k, v in fl.items() if isinstance(v, list)) and fl["timestamp_offsets"] == 0 and fl["scale"] == 0
    for args, msg in ((["--rig_transforms_to_float", names[0]], "Cannot float the rig transform from reference camera to itself."),
                      (["--camera_poses_to_float", "lidar"], "A specified sensor name is not among the known sensors. Offending camera: lidar"),
                      (["--intrinsics_to_float", f"{names[1]}:skew"], "Unexpected value when parsing intrinsics to float: skew"),
                      (["--intrinsics_to_float", "focal_length"], "Incorrectly set option for floating intrinsics.")):
        out = run(d, "--dry_run", "--bracket_len", "1.0", *args)
        assert out.returncode == 2 and msg in out.stderr, (args, out.stderr)


def test_rig_transforms_from_the_poses(host, dataset):
    """Without --use_initial_rig_transforms the rig comes from the poses of the map (rig_calibrator.cc:1364-1376).  The synthetic
    map's poses were made from the rig, so the derived rig is the rig (up to the text round trip and the sign of a quaternion)."""
    d, pr = dataset
    T = len(pr.sensors)
    got = np.array(json.loads(run(d, "--dry_run", "--bracket_len", "1.0").stdout)["ref_to_cam"]).reshape(T, 7)
    kept = np.array(json.loads(run(d, "--dry_run", "--bracket_len", "1.0", "--use_initial_rig_transforms").stdout)["ref_to_cam"]).reshape(T, 7)
    want = pr.ref_to_cam.reshape(T, 7)
    for a in (got, kept):
        for t in range(T):
            assert np.abs(a[t, :3] - want[t, :3]).max() < 1e-9
            assert min(np.abs(a[t, 3:] - want[t, 3:]).max(), np.abs(a[t, 3:] + want[t, 3:]).max()) < 1e-9
    assert not np.array_equal(got[1:], kept[1:])      # they did take different routes


def test_rig_from_poses_takes_the_middle_image_and_normalises_the_determinant(host):
    """calc_rig_using_word_to_cam: per matrix entry vals[vals.size() / 2] in image order (no sort), then linear / det^(1/3)."""
    from scipy.spatial.transform import Rotation
    rng = np.random.default_rng(5)
    dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)
    R = 4
    ref_stamps = np.arange(R, dtype=np.float64)
    w2r = np.zeros((R, 7)); w2r[:, :3] = rng.normal(size=(R, 3)); w2r[:, 3:] = Rotation.random(R, random_state=1).as_quat()
    cams = [(0, r, r, float(r)) for r in range(R)] + [(1, k, k + 1, k + 0.25 * (k + 1)) for k in range(3)]
    C_ = len(cams)
    sensor = np.array([c[0] for c in cams], np.int32); beg = np.array([c[1] for c in cams], np.int32); end = np.array([c[2] for c in cams], np.int32)
    stamp = np.array([c[3] for c in cams]); offsets = np.array([0.0, 0.05])
    true_rig = []
    w2c = np.zeros((C_, 12))
    from multiview_b200 import synth
    for c in range(C_):
        if sensor[c] == 0:
            Rm = Rotation.from_quat(w2r[beg[c], 3:]).as_matrix(); t = w2r[beg[c], :3]
        else:
            alpha = ((stamp[c] - ref_stamps[beg[c]]) - offsets[1]) / (ref_stamps[end[c]] - ref_stamps[beg[c]])
            from scipy.spatial.transform import Slerp
            sl = Slerp([0, 1], Rotation.from_quat([w2r[beg[c], 3:], w2r[end[c], 3:]]))
            Ri = sl([alpha]).as_matrix()[0]; ti = (1 - alpha) * w2r[beg[c], :3] + alpha * w2r[end[c], :3]
            Rr = Rotation.random(random_state=10 + c).as_matrix() * (1.0 + 0.1 * c); tr = rng.normal(size=3)    # a different rig for every image
            true_rig.append((Rr, tr))
            Rm = Rr @ Ri; t = Rr @ ti + tr
        w2c[c, :9] = Rm.ravel(); w2c[c, 9:] = t
    out = np.zeros((2, 12))
    rc = host.mvh_rig_from_poses(2, C_, sensor.ctypes.data_as(ip), beg.ctypes.data_as(ip), end.ctypes.data_as(ip), stamp.ctypes.data_as(dp),
                                 ref_stamps.ctypes.data_as(dp), np.ascontiguousarray(w2r).ctypes.data_as(dp), w2c.ctypes.data_as(dp), offsets.ctypes.data_as(dp),
                                 out.ctypes.data_as(dp))
    assert rc == 0
    assert np.array_equal(out[0], [1, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0, 0])
    Rr, tr = true_rig[len(true_rig) // 2]                 # the middle image of sensor 1
    assert np.abs(out[1, :9].reshape(3, 3) - Rr / np.cbrt(np.linalg.det(Rr))).max() < 1e-12
    assert np.abs(out[1, 9:] - tr).max() < 1e-12
    assert abs(np.linalg.det(out[1, :9].reshape(3, 3)) - 1.0) < 1e-12


def _read_ply(host, path):
    nv, nf = C.c_longlong(), C.c_longlong()
    assert host.mvh_read_ply(str(path).encode(), 0, None, 0, None, C.byref(nv), C.byref(nf)) == 0
    v = np.zeros((nv.value, 3), np.float32); f = np.zeros((nf.value, 3), np.uint32)
    assert host.mvh_read_ply(str(path).encode(), nv.value, v.ctypes.data_as(C.POINTER(C.c_float)), nf.value, f.ctypes.data_as(C.POINTER(C.c_uint)),
                             C.byref(nv), C.byref(nf)) == 0
    return v, f


def test_ply_reader_ascii_and_binary(host, tmp_path):
    verts = np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0.5], [0, 1, 0.25], [2, 2, 2]], np.float32)
    header = ("ply\nformat {fmt} 1.0\ncomment made by a test\nelement vertex 5\nproperty float x\nproperty float y\nproperty float z\n"
              "property uchar red\nproperty double confidence\nelement face 2\nproperty list uchar int vertex_indices\nend_header\n")
    a = tmp_path / "a.ply"
    a.write_text(header.format(fmt="ascii") + "".join(f"{x} {y} {z} 200 0.5\n" for x, y, z in verts) + "4 0 1 2 3\n3 1 4 2\n")
    b = tmp_path / "b.ply"
    with open(b, "wb") as f:
        f.write(header.format(fmt="binary_little_endian").encode())
        for x, y, z in verts:
            f.write(struct.pack("<fffBd", x, y, z, 200, 0.5))
        f.write(struct.pack("<Biiii", 4, 0, 1, 2, 3) + struct.pack("<Biii", 3, 1, 4, 2))
    for p in (a, b):
        v, f = _read_ply(host, p)
        assert np.array_equal(v, verts)
        assert np.array_equal(f, [[0, 1, 2], [0, 2, 3], [1, 4, 2]])        # the quad is split into a fan
    bad = tmp_path / "c.ply"
    bad.write_text(header.format(fmt="ascii") + "".join(f"{x} {y} {z} 200 0.5\n" for x, y, z in verts) + "3 0 1 9\n3 1 4 2\n")
    nv, nf = C.c_longlong(), C.c_longlong()
    assert host.mvh_read_ply(str(bad).encode(), 0, None, 0, None, C.byref(nv), C.byref(nf)) == -1
