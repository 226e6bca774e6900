"""Generates tests/golden/host_formats.json with the REFERENCE's own Python parsers (rig_calibrator/bin/rig_utils.py:
parseRigConfig, parse_images_and_camera_poses) applied to files written by multiview_b200/host (libmvba_host.so): what the
reference reads back from our writers' output.  Run in the container that has /root/reference."""
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, "/root/reference/rig_calibrator/bin")


def camera_list_text():
    import numpy as np
    rng = np.random.default_rng(3)
    rows = []
    for i in range(5):
        vals = rng.normal(size=12)
        rows.append(f"data/{['nav_cam', 'haz_cam', 'sci_cam'][i % 3]}/{10000.5 + i:.1f}.jpg " + " ".join("%.17g" % v for v in vals))
    return "# image_file world_to_image_transform\n" + "\n".join(rows) + "\n"


def main():
    import rig_utils                      # the reference's module, imported from where it lies
    import test_host_io as T
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "multiview_b200", "host")])
    host = C.CDLL(os.path.join(ROOT, "multiview_b200", "libmvba_host.so"))
    out = {}
    with tempfile.TemporaryDirectory() as d:
        src, dst = os.path.join(d, "in.txt"), os.path.join(d, "rig_config.txt")
        open(src, "w").write(T.RIG)
        assert host.mvh_rig_config_roundtrip(src.encode(), dst.encode(), 1, 1, None) == 0
        out["rig_config"] = rig_utils.parseRigConfig(dst)             # list of dicts of strings, as the reference returns them
        csrc, cdst = os.path.join(d, "cin.txt"), os.path.join(d, "cameras.txt")
        open(csrc, "w").write(camera_list_text())
        assert host.mvh_camera_list_roundtrip(csrc.encode(), cdst.encode(), None, None, None, 0) == 0
        out["camera_list"] = {}
        for sensor in ("nav_cam", "haz_cam", "sci_cam"):
            images, w2c = rig_utils.parse_images_and_camera_poses(cdst, sensor)
            out["camera_list"][sensor] = {"images": images, "world_to_cam": [m.tolist() for m in w2c]}
    json.dump(out, open(os.path.join(ROOT, "tests", "golden", "host_formats.json"), "w"), indent=1)
    print("sensors:", [c["sensor_name"] for c in out["rig_config"]], "images:", {k: len(v["images"]) for k, v in out["camera_list"].items()})


if __name__ == "__main__":
    main()
