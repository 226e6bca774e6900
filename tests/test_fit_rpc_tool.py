"""examples/fit_rpc.cc, the reference's RPC fitting tool (rig_calibrator/bin/fit_rpc.cc:59-146): flags and validation messages on CPU,
the fit itself on the GPU -- a radtan sensor is replaced by a degree-3 RPC model that must reproduce it and invert itself."""
import os
import subprocess

import numpy as np
import pytest

from multiview_b200 import _abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SIZE, FOCAL, CENTER = (640, 480), 600.0, (321.5, 238.25)
RADTAN = (-0.05, 0.01, 0.0005, -0.0003)
CONFIG = (f"ref_sensor_name: cam\n\nsensor_name: cam\nfocal_length: {FOCAL}\noptical_center: {CENTER[0]} {CENTER[1]}\n"
          f"distortion_coeffs: {' '.join(map(str, RADTAN))}\ndistortion_type: radtan\nimage_size: {SIZE[0]} {SIZE[1]}\n"
          f"distorted_crop_size: {SIZE[0]} {SIZE[1]}\nundistorted_image_size: {SIZE[0] + 60} {SIZE[1] + 40}\n"
          "ref_to_sensor_transform: 1 0 0 0 1 0 0 0 1 0 0 0\ndepth_to_image_transform: 1 0 0 0 1 0 0 0 1 0 0 0\nref_to_sensor_timestamp_offset: 0\n")


@pytest.fixture(scope="module")
def tool():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "multiview_b200", "host")])
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "examples")])
    return os.path.join(ROOT, "examples", "fit_rpc")


def run(tool, *args):
    return subprocess.run([tool, *args], capture_output=True, text=True, timeout=300)


def test_validation_messages_of_the_reference(tool, tmp_path):
    cfg = tmp_path / "rig_config.txt"; cfg.write_text(CONFIG)
    ok = ["--camera_config", str(cfg), "--out_dir", str(tmp_path / "out"), "--rpc_degree", "3", "--num_samples", "50"]
    for drop, msg in ((0, "Camera config file was not specified."), (2, "Output camera config directory was not specified."),
                      (4, "The RPC degree must be positive."), (6, "The number of samples must be positive.")):
        out = run(tool, *(ok[:drop] + ok[drop + 2:]))
        assert out.returncode == 2 and msg in out.stderr, (drop, out.stderr)
    out = run(tool, *ok, "--no_such_flag=1")
    assert out.returncode == 1 and "unknown command line flag 'no_such_flag'" in out.stderr


def test_without_a_device_the_fit_fails_loudly(tool, tmp_path):
    if _abi.load_product().mvba_device_count() > 0:
        pytest.skip("a CUDA device is present")
    cfg = tmp_path / "rig_config.txt"; cfg.write_text(CONFIG)
    out = run(tool, "--camera_config", str(cfg), "--out_dir", str(tmp_path / "out"), "--rpc_degree=2", "-num_samples", "30")
    assert out.returncode == 3 and "no CUDA device" in out.stderr
    assert "Found " in out.stdout and "pixel correspondences" in out.stdout      # the host half ran


@pytest.mark.gpu
def test_radtan_sensor_becomes_an_rpc_model(tool, tmp_path):
    from test_camera_params import convert, host as _host_fixture, UNDISTORTED_C, DISTORTED_C  # noqa: F401
    import ctypes as C
    cfg = tmp_path / "rig_config.txt"; cfg.write_text(CONFIG)
    out = run(tool, "--camera_config", str(cfg), "--out_dir", str(tmp_path / "out"), "--rpc_degree", "3", "--num_samples", "80", "--num_iterations", "20")
    assert out.returncode == 0, out.stderr + out.stdout
    rows = {l.split(":")[0]: l.split()[1:] for l in (tmp_path / "out" / "rig_config.txt").read_text().splitlines() if ":" in l}
    assert rows["distortion_type"] == ["rpc"] and len(rows["distortion_coeffs"]) == 2 * _abi.rpc_num_params(3) == 76
    assert rows["focal_length"] == [str(FOCAL)] or float(rows["focal_length"][0]) == FOCAL
    rpc = np.array(rows["distortion_coeffs"], float)
    lib = C.CDLL(os.path.join(ROOT, "multiview_b200", "libmvba_host.so")); lib.mvh_camera_last_error.restype = C.c_char_p
    g = np.stack(np.meshgrid(np.linspace(-0.42, 0.42, 11) * SIZE[0], np.linspace(-0.42, 0.42, 11) * SIZE[1]), -1).reshape(-1, 2)
    want = convert(lib, SIZE, FOCAL, CENTER, RADTAN, UNDISTORTED_C, DISTORTED_C, g)
    got = convert(lib, SIZE, FOCAL, CENTER, rpc, UNDISTORTED_C, DISTORTED_C, g)
    assert np.abs(got - want).max() < 0.05, np.abs(got - want).max()
    back = convert(lib, SIZE, FOCAL, CENTER, rpc, DISTORTED_C, UNDISTORTED_C, got)
    assert np.abs(back - g).max() < 0.05, np.abs(back - g).max()
    import re
    assert float(re.search(r"Max distort_undistort error: (\S+)", out.stdout).group(1)) < 0.5
