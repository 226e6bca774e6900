"""examples/rig_calibrator.cc, the stand-alone driver with the reference's command line, on a CPU box: a synthetic rig is written out as rig_config.txt + NVM (image paths
carrying sensor and timestamp), the driver reads them back, brackets the images and assembles the pass; --dry_run prints it before
the first device call.  The assembled arrays must be the synthetic problem's own.  Without --dry_run on a box with no GPU the
driver must fail loudly at the first compute call (no CPU fallback)."""
import ctypes as C
import json
import os
import subprocess

import numpy as np
import pytest

from multiview_b200 import _abi, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)
g17 = lambda v: "%.17g" % v


def quat_to_rows(p7):
    from scipy.spatial.transform import Rotation
    return list(Rotation.from_quat(p7[3:]).as_matrix().ravel()) + list(p7[:3])


def write_dataset(d, pr):
    """rig_config.txt + map.nvm of a synthetic problem in directory d (image paths carry sensor and timestamp)."""
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "multiview_b200", "host")])
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "examples")])
    host = C.CDLL(os.path.join(ROOT, "multiview_b200", "libmvba_host.so"))
    S = pr.sensors; T = len(S)
    names = [s.name for s in S]
    j = lambda v: " ".join(g17(x) for x in v)
    txt = f"ref_sensor_name: {names[0]}\n"
    kind = lambda n: {0: "no_distortion", 1: "fisheye", 4: "radtan", 5: "radtan"}.get(n, "rpc")
    r2c = pr.ref_to_cam.reshape(T, 7)
    doff = np.concatenate([[0], np.cumsum([len(s.dist) for s in S])])
    for t, s in enumerate(S):
        txt += (f"\nsensor_name: {s.name}\nfocal_length: {g17(pr.focal[t])}\noptical_center: {j(pr.center[2 * t:2 * t + 2])}\n"
                f"distortion_coeffs: {j(pr.distortion[doff[t]:doff[t + 1]])}\ndistortion_type: {kind(len(s.dist))}\nimage_size: {int(s.width)} {int(s.height)}\n"
                f"distorted_crop_size: {int(s.width)} {int(s.height)}\nundistorted_image_size: {int(s.width) + 200} {int(s.height) + 200}\n"
                f"ref_to_sensor_transform: {j(quat_to_rows(r2c[t]))}\ndepth_to_image_transform: 1 0 0 0 1 0 0 0 1 0 0 0\n"
                f"ref_to_sensor_timestamp_offset: {g17(pr.offsets[t])}\n")
    (d / "rig_config.txt").write_text(txt)
    files = [f"data/{names[pr.cam_sensor[c]]}/{g17(pr.cam_stamp[c])}.jpg" for c in range(len(pr.cam_sensor))]
    perm = np.random.default_rng(0).permutation(len(files))            # the NVM lists the images in no particular order
    inv = np.argsort(perm)
    (d / "names.txt").write_text("\n".join(files[c] for c in perm) + "\n")
    Rm, tv = synth.world_to_cam_all(pr)
    w2c = np.ascontiguousarray(np.concatenate([Rm.reshape(-1, 9), tv], axis=1)[perm])
    cs = np.ascontiguousarray(pr.cam_sensor[perm].astype(np.int32))
    cid = np.ascontiguousarray(inv[pr.cid].astype(np.int32))            # camera index in the file's order
    order = np.lexsort((cid, pr.pid))                                     # nvm_from_inliers wants cid ascending inside a track
    pid = np.ascontiguousarray(pr.pid[order].astype(np.int32)); cid = np.ascontiguousarray(cid[order])
    pix = np.ascontiguousarray(pr.pix.reshape(-1, 2)[order].astype(np.float32))
    inl = np.ones(len(pid), np.uint8)
    focal = np.ascontiguousarray(pr.focal); center = np.ascontiguousarray(pr.center); xyz = np.ascontiguousarray(pr.xyz)
    rc = host.mvh_nvm_export(str(d / "names.txt").encode(), str(d / "map.nvm").encode(), len(files), cs.ctypes.data_as(ip), T, focal.ctypes.data_as(dp),
                             center.ctypes.data_as(dp), w2c.ctypes.data_as(dp), C.c_longlong(len(pid)), pid.ctypes.data_as(ip), cid.ctypes.data_as(ip),
                             pix.ctypes.data_as(C.POINTER(C.c_float)), inl.ctypes.data_as(C.POINTER(C.c_ubyte)), pr.num_landmarks, xyz.ctypes.data_as(dp))
    assert rc == 0
    return d, pr


@pytest.fixture(scope="module")
def dataset(tmp_path_factory):
    return write_dataset(tmp_path_factory.mktemp("rig_pass"), synth.make_problem(seed=31, sensors="rig3", num_ref=10, num_landmarks=120, obs_per_track=5))


def run(d, *extra, cwd=None):
    exe = os.path.join(ROOT, "examples", "rig_calibrator")
    return subprocess.run([exe, "--rig_config", str(d / "rig_config.txt"), "--nvm", str(d / "map.nvm"), "--out_dir", str(d / "out"), *extra],
                          capture_output=True, text=True, timeout=300, cwd=cwd)


def test_dry_run_assembles_the_synthetic_pass(dataset):
    d, pr = dataset
    out = run(d, "--dry_run", "--bracket_len", "1.0")
    assert out.returncode == 0, out.stderr
    got = json.loads(out.stdout)
    assert (got["sensors"], got["ref_frames"], got["cameras"]) == (len(pr.sensors), len(pr.ref_stamps), len(pr.cam_sensor))
    cams = np.array(got["cams"])
    assert np.array_equal(cams[:, 0], pr.cam_sensor) and np.array_equal(cams[:, 1], pr.cam_beg) and np.array_equal(cams[:, 2], pr.cam_end)
    assert np.array_equal(cams[:, 3], pr.cam_stamp)                        # timestamps survive the file name round trip exactly
    # all tracks have 5 > 2 views, so the NVM holds every feature; pixels went through float32 - centre + centre
    assert got["tracks"] == pr.num_landmarks and got["features"] == pr.num_obs
    ff = np.array(got["first_features"])
    assert np.array_equal(ff[:, 0], pr.pid[:6]) and np.array_equal(ff[:, 1], pr.cid[:6])
    assert np.abs(ff[:, 2:] - pr.pix.reshape(-1, 2)[:6]).max() < 1e-3
    lo, hi = np.array(got["offset_bounds"]).T
    assert (lo[1:] < pr.offsets[1:]).all() and (pr.offsets[1:] < hi[1:]).all()


def test_without_a_device_the_driver_fails_at_the_first_compute_call(dataset):
    lib = _abi.load_product()
    if lib.mvba_device_count() > 0:
        pytest.skip("a CUDA device is present")
    d, _ = dataset
    out = run(d, "--bracket_len", "1.0")
    assert out.returncode == 3 and "no CUDA device" in out.stderr


@pytest.mark.gpu
def test_the_assembled_pass_runs_on_the_device_and_follows_the_oracle_driven_pass(dataset, oracle_lib):
    """The whole driver on a GPU: files in -> bracketing -> [triangulate -> residual report -> solve -> report -> outlier flags] x 2 ->
    files out, against the same loop driven through the CPU oracle from the synthetic problem's own arrays
    (rig_calibrator.cc:1550-1985).  The NVM stores keypoints as text relative to the optical centre, so the driver's pixels
    differ from the problem's by float32 round-off (< 1e-4 px): costs are compared to 1e-4 relative, poses to 1e-5."""
    d, pr = dataset
    names = [s.name for s in pr.sensors]
    out = run(d, "--bracket_len", "1.0", "--calibrator_num_passes", "2", "--num_iterations", "10", "--use_initial_rig_transforms",
              "--rig_transforms_to_float", " ".join(names[1:]), "--camera_poses_to_float", names[0], "--min_triangulation_angle", "0.01")
    assert out.returncode == 0, out.stderr + out.stdout[-2000:]
    text = out.stdout
    assert text.count("The 25, 50, 75, and 100th percentile residual stats before opt") == 2
    assert text.count("The 25, 50, 75, and 100th percentile residual stats after opt") == 2
    for s in pr.sensors:
        assert f"{s.name}_pix_x: " in text
    import re
    got = [(int(m.group(1)), int(m.group(2)), float(m.group(3)), float(m.group(4)))
           for m in re.finditer(r"pass (\d+): (\d+) inlier features, cost (\S+) -> (\S+) in", text)]
    assert len(got) == 2
    # ---- the same loop through the oracle ----
    T = len(pr.sensors)
    img = np.array([[s.width, s.height] for s in pr.sensors], np.float64)
    o_all = _abi.Binding(oracle_lib, "oracle_", pr)
    inl = o_all.flag_exclusion(img, img)
    want = []
    for p in range(2):
        xyz, inl = o_all.triangulate(inl, pr.xyz)
        pr.xyz[:] = xyz.ravel()
        q = pr.keep_observations(inl == 1)
        o = _abi.Binding(oracle_lib, "oracle_", q)
        s, _ = o.solve(_abi.default_options(max_num_iterations=10), jac_mode=1, num_threads=1)
        o.close()
        for k, v in q.param_arrays().items():
            pr.param_arrays()[k][...] = v
        want.append((p, int((inl == 1).sum()), s.initial_cost, s.final_cost))
        inl, _ = o_all.flag_outliers(0.01, 25.0, inl)
    o_all.close()
    for g, w in zip(got, want):
        assert g[0] == w[0] and g[1] == w[1], (g, w)
        assert abs(g[2] - w[2]) <= 1e-4 * w[2] and abs(g[3] - w[3]) <= 1e-4 * w[3], (g, w)
    # ---- written camera list: world_to_cam of every image at the optimised parameters ----
    Rm, tv = synth.world_to_cam_all(pr)
    rows = [l.split() for l in (d / "out" / "cameras.txt").read_text().splitlines() if l and not l.startswith("#")]
    assert (d / "out" / "rig_config.txt").exists() and not (d / "out" / "cameras.nvm").exists()       # the NVM only with --save_nvm
    files = [f"data/{pr.sensors[pr.cam_sensor[c]].name}/{g17(pr.cam_stamp[c])}.jpg" for c in range(len(pr.cam_sensor))]
    by_file = {r[0]: np.array(r[1:], float) for r in rows}
    for c, f in enumerate(files):
        m = by_file[f]
        assert np.abs(m[:9] - Rm[c].ravel()).max() <= 1e-5 and np.abs(m[9:12] - tv[c]).max() <= 1e-5, f


@pytest.mark.gpu
def test_the_flags_reach_the_solver(dataset, oracle_lib):
    """One pass with the reference's 'what floats' flags and the stay-near-the-triangulation prior: reference-sensor poses, ONE rig
    transform, one focal length and the time offsets float (rig_calibrator.cc:1702-1752), --tri_weight adds an XYZError per track
    with an inlier feature against the pass's own triangulation (:1571-1580, 1886-1901).  The oracle solves the same pass
    assembled by hand from the synthetic problem."""
    d, _ = dataset
    pr = synth.make_problem(seed=31, sensors="rig3", num_ref=10, num_landmarks=120, obs_per_track=5)     # as the files were written
    names = [s.name for s in pr.sensors]
    out = run(d, "--bracket_len", "1.0", "--calibrator_num_passes", "1", "--num_iterations", "2", "--use_initial_rig_transforms",
              "--camera_poses_to_float", names[0], "--rig_transforms_to_float", names[1], "--intrinsics_to_float", f"{names[2]}:focal_length",
              "--float_timestamp_offsets", "--tri_weight", "0.5", "--tri_robust_threshold", "0.1", "--min_triangulation_angle", "0.01", "--save_nvm")
    assert out.returncode == 0, out.stderr + out.stdout[-2000:]
    assert "tri_x_m: " in out.stdout and (d / "out" / "cameras.nvm").exists()
    import re
    m = re.search(r"pass 0: (\d+) inlier features, cost (\S+) -> (\S+) in", out.stdout)
    got = (int(m.group(1)), float(m.group(2)), float(m.group(3)))
    T = len(pr.sensors)
    u8 = lambda v: np.array(v, np.uint8)
    pr.float_poses = u8([1, 0, 0]); pr.float_rig = u8([0, 1, 0]); pr.float_focal = u8([0, 0, 1]); pr.float_center = u8([0] * T); pr.float_dist = u8([0] * T)
    pr.float_offsets = True
    img = np.array([[s.width, s.height] for s in pr.sensors], np.float64)
    o_all = _abi.Binding(oracle_lib, "oracle_", pr)
    inl = o_all.flag_exclusion(img, img)
    xyz, inl = o_all.triangulate(inl, pr.xyz)
    o_all.close()
    pr.xyz[:] = xyz.ravel()
    q = pr.keep_observations(inl == 1)
    pids = np.unique(q.pid).astype(np.int32)
    q.priors = {synth.BLOCK_TRI: (pids, xyz.reshape(-1, 3)[pids].reshape(-1).copy(), 0.5, 0.1)}
    o = _abi.Binding(oracle_lib, "oracle_", q)
    s, _ = o.solve(_abi.default_options(max_num_iterations=2), jac_mode=1, num_threads=1)
    o.close()
    assert got[0] == int((inl == 1).sum())
    assert abs(got[1] - s.initial_cost) <= 1e-4 * s.initial_cost and abs(got[2] - s.final_cost) <= 1e-4 * s.final_cost, (got, s.initial_cost, s.final_cost)
    # the written rig: only the floated transform and the floated focal length moved
    rig_in = (d / "rig_config.txt").read_text().split("\n\n")
    rig_out = (d / "out" / "rig_config.txt").read_text().split("\n\n")
    field = lambda blk, key: [l for l in blk.splitlines() if l.startswith(key)][0]
    assert field(rig_out[2], "focal_length:") == field(rig_in[2], "focal_length:")
    assert field(rig_out[3], "focal_length:") != field(rig_in[3], "focal_length:")


def depth_and_mesh_inputs(d, pr):
    """Depth clouds (<image stem>.pc, readXyzImage's layout: rows, cols, channels as int32, then float32 xyz per pixel) for the
    images of sensor 1, holding the camera-frame position of every feature at the initial state, and a box mesh around the scene."""
    import struct
    from test_mesh import scene_box
    Rm, tv = synth.world_to_cam_all(pr)
    _, Xc = synth.project(pr, Rm, tv, pr.xyz, pr.pid, pr.cid)
    s1 = pr.sensors[1]
    W, H = int(s1.width), int(s1.height)
    clouds = {}
    px = pr.pix.reshape(-1, 2)
    for i in np.nonzero(pr.cam_sensor[pr.cid] == 1)[0]:
        col, row = int(np.round(px[i, 0])), int(np.round(px[i, 1]))     # np.round and std::round differ only at exact .5 of a float32 pixel
        assert 0 <= col < W and 0 <= row < H
        clouds.setdefault(int(pr.cid[i]), np.zeros((H, W, 3), np.float32))[row, col] = Xc[i].astype(np.float32)
    os.makedirs(d / "data" / pr.sensors[1].name, exist_ok=True)
    for c, img in clouds.items():
        with open(d / "data" / pr.sensors[1].name / (g17(pr.cam_stamp[c]) + ".pc"), "wb") as f:
            f.write(struct.pack("<iii", H, W, 3)); f.write(img.tobytes())
    has = np.zeros(pr.num_obs, np.uint8); dxyz = np.zeros((pr.num_obs, 3), np.float32)
    for i in np.nonzero(pr.cam_sensor[pr.cid] == 1)[0]:
        v = clouds[int(pr.cid[i])][int(np.round(px[i, 1])), int(np.round(px[i, 0]))]
        if np.any(v != 0):
            has[i] = 1; dxyz[i] = v
    (v, f), _, _ = scene_box(pr, 6)
    with open(d / "box.ply", "wb") as fh:
        fh.write((f"ply\nformat binary_little_endian 1.0\nelement vertex {len(v)}\nproperty float x\nproperty float y\nproperty float z\n"
                  f"element face {len(f)}\nproperty list uchar int vertex_indices\nend_header\n").encode())
        fh.write(v.astype("<f4").tobytes())
        for tri in f:
            fh.write(struct.pack("<Biii", 3, *[int(k) for k in tri]))
    return has, dxyz, v, f


@pytest.mark.gpu
@pytest.mark.parametrize("affine", [False, True])
def test_mesh_and_depth_flags(affine, dataset, oracle_lib):
    """--mesh (a .ply read by the driver, rays intersected on the device), depth clouds found beside the images, --depth_tri_weight,
    --depth_mesh_weight, --mesh_tri_weight, --depth_to_image_transforms_to_float and --float_scale: one pass against the oracle
    driven by hand through the same steps (rig_calibrator.cc:1582-1589, 1755-1840, 1857-1883)."""
    d, _ = dataset
    pr = synth.make_problem(seed=31, sensors="rig3", num_ref=10, num_landmarks=120, obs_per_track=5)
    names = [s.name for s in pr.sensors]
    # the fixture's rig_config.txt carries the identity as depth_to_image_transform for every sensor: the oracle's problem must too
    pr.depth_to_image[:] = np.tile([0, 0, 0, 0, 0, 0, 1.0], len(pr.sensors)); pr.depth_scale[:] = 1.0
    has, dxyz, verts, faces = depth_and_mesh_inputs(d, pr)
    if affine:   # --affine_depth_to_image: twelve free numbers per floated depth transform, the scale block constant (rig_calibrator.cc:1445, 1787)
        pr.affine_depth = True
        pr.depth_to_image = np.tile([1.0, 0, 0, 0, 0, 1.0, 0, 0, 0, 0, 1.0, 0], len(pr.sensors))
    out = run(d, "--bracket_len", "1.0", "--calibrator_num_passes", "1", "--num_iterations", "4", "--use_initial_rig_transforms",
              "--camera_poses_to_float", names[0], "--rig_transforms_to_float", " ".join(names[1:]), "--mesh", str(d / "box.ply"),
              "--depth_tri_weight", "200", "--depth_mesh_weight", "20", "--mesh_tri_weight", "10", "--depth_to_image_transforms_to_float", names[1],
              "--affine_depth_to_image" if affine else "--float_scale", "--min_triangulation_angle", "0.01", "--max_ray_dist", "50", cwd=str(d))
    assert out.returncode == 0, out.stderr + out.stdout[-2000:]
    for tag in ("depth_tri_x_m: ", "depth_mesh_x_m: ", "mesh_tri_x_m: "):
        assert tag in out.stdout, tag
    import re
    m = re.search(r"pass 0: (\d+) inlier features, cost (\S+) -> (\S+) in", out.stdout)
    got = (int(m.group(1)), float(m.group(2)), float(m.group(3)))
    T = len(pr.sensors)
    u8 = lambda v: np.array(v, np.uint8)
    pr.float_poses = u8([1, 0, 0]); pr.float_rig = u8([0, 1, 1]); pr.float_d2i = u8([0, 1, 0]); pr.float_scale = not affine
    pr.depth_tri_weight = 200.0; pr.depth_mesh_weight = 20.0
    img = np.array([[s.width, s.height] for s in pr.sensors], np.float64)
    o_all = _abi.Binding(oracle_lib, "oracle_", pr)
    inl = o_all.flag_exclusion(img, img)
    xyz, inl = o_all.triangulate(inl, pr.xyz)
    pr.xyz[:] = xyz.ravel()
    hm, mxyz, pmesh = o_all.mesh_intersect(verts, faces, 0.0, 50.0, inl)
    o_all.close()
    pr.has_depth = has; pr.depth_xyz = dxyz.reshape(-1).copy(); pr.has_mesh = hm; pr.mesh_xyz = mxyz.reshape(-1).copy()
    q = pr.keep_observations(inl == 1)
    pids = np.unique(q.pid).astype(np.int32)
    pids = pids[~np.all(pmesh[pids] == 1e100, axis=1)]
    q.priors = {synth.BLOCK_MESH_TRI: (pids, pmesh[pids].reshape(-1).copy(), 10.0, pr.robust_threshold)}
    o = _abi.Binding(oracle_lib, "oracle_", q)
    s, _ = o.solve(_abi.default_options(max_num_iterations=4), jac_mode=1, num_threads=1)
    o.close()
    assert got[0] == int((inl == 1).sum())
    assert abs(got[1] - s.initial_cost) <= 1e-4 * s.initial_cost and abs(got[2] - s.final_cost) <= 1e-4 * s.final_cost, (got, s.initial_cost, s.final_cost)


def no_rig_problem(pr):
    """The synthetic rig problem as --no_rig sees it: every image a camera of its own (sensor-major, time order inside a sensor,
    lookupImagesNoBrackets), its pose the one the map holds (here: the rig-derived pose of the initial state)."""
    import copy
    order = np.lexsort((pr.cam_stamp, pr.cam_sensor))
    inv = np.argsort(order)                                   # rig camera index -> no-rig camera index
    Rm, tv = synth.world_to_cam_all(pr)
    w2c = np.concatenate([tv, synth._mat_to_quat(Rm)], axis=1)[order]
    q = copy.copy(pr)
    q.no_rig = True
    q.world_to_cam = np.ascontiguousarray(w2c.reshape(-1))
    q.cam_sensor = np.ascontiguousarray(pr.cam_sensor[order]); q.cam_stamp = np.ascontiguousarray(pr.cam_stamp[order])
    idx = np.concatenate([np.arange((q.cam_sensor == t).sum()) for t in range(len(pr.sensors))]).astype(np.int32)
    q.cam_beg = idx.copy(); q.cam_end = idx.copy()
    cid = inv[pr.cid].astype(np.int32)
    o = np.lexsort((cid, pr.pid))
    q.pid = np.ascontiguousarray(pr.pid[o]); q.cid = np.ascontiguousarray(cid[o])
    q.pix = np.ascontiguousarray(pr.pix.reshape(-1, 2)[o].reshape(-1))
    for nm in ("world_to_ref", "ref_to_cam", "offsets", "focal", "center", "distortion", "depth_to_image", "depth_scale", "xyz"):
        setattr(q, nm, getattr(pr, nm).copy())
    T = len(pr.sensors)
    q.float_poses = np.ones(T, np.uint8); q.float_rig = np.zeros(T, np.uint8)
    return q


def test_no_rig_dry_run_lists_every_image_as_its_own_camera(dataset):
    d, _ = dataset
    pr = synth.make_problem(seed=31, sensors="rig3", num_ref=10, num_landmarks=120, obs_per_track=5)
    q = no_rig_problem(pr)
    out = run(d, "--dry_run", "--no_rig")
    assert out.returncode == 0, out.stderr
    got = json.loads(out.stdout)
    cams = np.array(got["cams"])
    assert np.array_equal(cams[:, 0], q.cam_sensor) and np.array_equal(cams[:, 1], q.cam_beg) and np.array_equal(cams[:, 2], q.cam_end)
    assert np.array_equal(cams[:, 3], q.cam_stamp)
    ff = np.array(got["first_features"])
    assert np.array_equal(ff[:, 0], q.pid[:6]) and np.array_equal(ff[:, 1], q.cid[:6])


@pytest.mark.gpu
def test_no_rig_pass(dataset, oracle_lib):
    """--no_rig: no bracketing, one free pose per image of the sensors named in --camera_poses_to_float (rig_calibrator.cc:1487-1492,
    1669-1730), identity rig transforms in the written rig_config.txt (dense_map_utils.cc:836-839)."""
    d, _ = dataset
    pr = synth.make_problem(seed=31, sensors="rig3", num_ref=10, num_landmarks=120, obs_per_track=5)
    names = [s.name for s in pr.sensors]
    out = run(d, "--no_rig", "--calibrator_num_passes", "1", "--num_iterations", "2", "--camera_poses_to_float", " ".join(names),
              "--min_triangulation_angle", "0.01")
    assert out.returncode == 0, out.stderr + out.stdout[-2000:]
    import re
    m = re.search(r"pass 0: (\d+) inlier features, cost (\S+) -> (\S+) in", out.stdout)
    got = (int(m.group(1)), float(m.group(2)), float(m.group(3)))
    q = no_rig_problem(pr)
    img = np.array([[s.width, s.height] for s in pr.sensors], np.float64)
    o_all = _abi.Binding(oracle_lib, "oracle_", q)
    inl = o_all.flag_exclusion(img, img)
    xyz, inl = o_all.triangulate(inl, q.xyz)
    o_all.close()
    q.xyz[:] = xyz.ravel()
    q2 = q.keep_observations(inl == 1)
    o = _abi.Binding(oracle_lib, "oracle_", q2)
    s, _ = o.solve(_abi.default_options(max_num_iterations=2), jac_mode=1, num_threads=1)
    o.close()
    assert got[0] == int((inl == 1).sum())
    assert abs(got[1] - s.initial_cost) <= 1e-4 * s.initial_cost and abs(got[2] - s.final_cost) <= 1e-4 * s.final_cost, (got, s.initial_cost, s.final_cost)
    rig_out = (d / "out" / "rig_config.txt").read_text()
    assert rig_out.count("ref_to_sensor_transform: 1 0 0 0 1 0 0 0 1 0 0 0") == len(names)


RPC_CASE = dict(seed=7, sensors="rpc", num_ref=12, num_landmarks=300, obs_per_track=6, float_rig=False,
                float_intrinsics=("focal_length", "distortion"), intr_noise=1e-3)


@pytest.mark.gpu
def test_rpc_distortion_floats_and_the_undistortion_half_is_refitted(tmp_path, oracle_lib):
    """--intrinsics_to_float 'cam:focal_length,distortion' on an RPC sensor: the distortion half of the model floats in the pass
    (camera_params.cc:309-312 inside the functor), and after the pass the undistortion half is re-fitted from it
    (rig_calibrator.cc:1940-1948 -> CameraParameters::updateRpcUndistortion) before rig_config.txt is written."""
    d, pr = write_dataset(tmp_path, synth.make_problem(**RPC_CASE))
    name = pr.sensors[0].name
    nd = len(pr.sensors[0].dist); half = nd // 2
    out = run(d, "--calibrator_num_passes", "1", "--num_iterations", "3", "--use_initial_rig_transforms", "--camera_poses_to_float", name,
              "--intrinsics_to_float", f"{name}:focal_length,distortion", "--min_triangulation_angle", "0.01")
    assert out.returncode == 0, out.stderr + out.stdout[-2000:]
    import re
    m = re.search(r"pass 0: (\d+) inlier features, cost (\S+) -> (\S+) in", out.stdout)
    got = (int(m.group(1)), float(m.group(2)), float(m.group(3)))
    q = synth.make_problem(**RPC_CASE)
    dist_in = q.distortion.copy()
    img = np.array([[s.width, s.height] for s in q.sensors], np.float64)
    o_all = _abi.Binding(oracle_lib, "oracle_", q)
    inl = o_all.flag_exclusion(img, img)
    xyz, inl = o_all.triangulate(inl, q.xyz)
    o_all.close()
    q.xyz[:] = xyz.ravel()
    q2 = q.keep_observations(inl == 1)
    o = _abi.Binding(oracle_lib, "oracle_", q2)
    s, _ = o.solve(_abi.default_options(max_num_iterations=3), jac_mode=1, num_threads=1)
    o.close()
    assert got[0] == int((inl == 1).sum())
    assert abs(got[1] - s.initial_cost) <= 1e-4 * s.initial_cost and abs(got[2] - s.final_cost) <= 1e-4 * s.final_cost, (got, s.initial_cost, s.final_cost)
    line = [l for l in (d / "out" / "rig_config.txt").read_text().splitlines() if l.startswith("distortion_coeffs:")][0]
    dist_out = np.array(line.split()[1:], float)
    assert len(dist_out) == nd
    assert np.abs(dist_out[:half] - q2.distortion[:half]).max() <= 1e-4 * max(1.0, np.abs(q2.distortion[:half]).max())   # the solved half (pixels went through the NVM's text)
    assert not np.array_equal(dist_out[:half], dist_in[:half]) and not np.array_equal(dist_out[half:], dist_in[half:])  # both halves moved
    # The re-fitted undistortion half against the same refit driven on the CPU (the host mirror with the oracle's RPC fit as its
    # fitter, capped at 20 LM iterations per degree to keep the test short; the driver ran the device fit with the reference's 100).
    # After three LM iterations the solved distortion is far from a smooth lens model, so the degree-3 inverse is only good to
    # several pixels for either fitter -- but far better than the stale half the pass started with.
    import oracle.binding as ob
    import test_rpc_update as tru
    host = tru._host()
    sz = q.sensors[0]
    rows = {l.split(":")[0]: l.split()[1:] for l in (d / "out" / "rig_config.txt").read_text().splitlines() if ":" in l}
    size = np.array([sz.width, sz.height], np.int32); und = np.array(rows["undistorted_image_size"], np.int32); crop = np.array(rows["distorted_crop_size"], np.int32)
    focal = np.array([float(rows["focal_length"][0])] * 2); center = np.array(rows["optical_center"], float)

    def fit(device, n, in_xy, out_xy, deg, iters, ptol, jmode, coeffs, stats):
        return oracle_lib.oracle_rpc_fit(C.c_int64(n), in_xy, out_xy, deg, 20, C.c_double(ptol), 0, coeffs, stats)
    cb = tru.FITTER(fit)
    ref = dist_out.copy(); ref[half:] = dist_in[half:]
    rc = host.mvh_update_rpc_undistortion(size.ctypes.data_as(ip), focal.ctypes.data_as(dp), center.ctypes.data_as(dp), ref.ctypes.data_as(dp), nd,
                                          und.ctypes.data_as(ip), crop.ctypes.data_as(ip), cb)
    assert rc == 0
    g = np.stack(np.meshgrid(np.linspace(-0.45, 0.45, 9) * sz.width, np.linspace(-0.45, 0.45, 9) * sz.height), -1).reshape(-1, 2)
    dg = ob.compute_rpc(oracle_lib, g, dist_out[:half])
    mine, theirs, stale = (ob.compute_rpc(oracle_lib, dg, h) for h in (dist_out[half:], ref[half:], dist_in[half:]))
    assert np.abs(mine - theirs).max() <= 0.05, np.abs(mine - theirs).max()
    assert np.abs(mine - g).max() < 0.5 * np.abs(stale - g).max()
