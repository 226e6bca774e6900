"""examples/rig_pass.cc, the stand-alone driver, on a CPU box: a synthetic rig is written out as rig_config.txt + NVM (image paths
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


@pytest.fixture(scope="module")
def dataset(tmp_path_factory):
    d = tmp_path_factory.mktemp("rig_pass")
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "multiview_b200", "host")])
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "examples")])
    host = C.CDLL(os.path.join(ROOT, "multiview_b200", "libmvba_host.so"))
    pr = synth.make_problem(seed=31, sensors="rig3", num_ref=10, num_landmarks=120, obs_per_track=5)
    S = pr.sensors; T = len(S)
    names = [s.name for s in S]
    j = lambda v: " ".join(g17(x) for x in v)
    txt = f"ref_sensor_name: {names[0]}\n"
    kind = {0: "no_distortion", 1: "fisheye", 4: "radtan", 5: "radtan"}
    r2c = pr.ref_to_cam.reshape(T, 7)
    for t, s in enumerate(S):
        txt += (f"\nsensor_name: {s.name}\nfocal_length: {g17(pr.focal[t])}\noptical_center: {j(pr.center[2 * t:2 * t + 2])}\n"
                f"distortion_coeffs: {j(s.dist)}\ndistortion_type: {kind[len(s.dist)]}\nimage_size: {int(s.width)} {int(s.height)}\n"
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


def run(d, *extra):
    exe = os.path.join(ROOT, "examples", "rig_pass")
    return subprocess.run([exe, "--rig_config", str(d / "rig_config.txt"), "--nvm", str(d / "map.nvm"), "--out_dir", str(d), *extra],
                          capture_output=True, text=True, timeout=300)


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
    out = run(d, "--bracket_len", "1.0", "--num_passes", "2", "--num_iterations", "10", "--float_rig")
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
    rows = [l.split() for l in (d / "cameras.txt").read_text().splitlines() if l and not l.startswith("#")]
    files = [f"data/{pr.sensors[pr.cam_sensor[c]].name}/{g17(pr.cam_stamp[c])}.jpg" for c in range(len(pr.cam_sensor))]
    by_file = {r[0]: np.array(r[1:], float) for r in rows}
    for c, f in enumerate(files):
        m = by_file[f]
        assert np.abs(m[:9] - Rm[c].ravel()).max() <= 1e-5 and np.abs(m[9:12] - tv[c]).max() <= 1e-5, f
