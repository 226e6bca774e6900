"""evalResiduals / calc_residuals_stats (rig_calibrator.cc:753-789, 1157-1180): names and scales per residual row, raw
residuals divided by their weight, the 25/50/75/100 "percentiles" the reference prints (entries [0.25 n], [0.5 n], [0.75 n],
[n - 1] of the sorted absolute values) and the printed text.  CPU: the C++ host code against a numpy restatement on the
oracle's residuals.  GPU: the same report from the device residuals (mvba_evaluate) against the oracle's."""
import ctypes as C
import os

import numpy as np
import pytest

from multiview_b200 import _abi, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASE = dict(sensors="rig3", num_ref=12, num_landmarks=400, obs_per_track=5, depth=True, mesh=True, depth_tri_weight=1000.0,
            depth_mesh_weight=50.0, mesh_tri_weight=25.0, tri_weight=0.5, outlier_frac=0.02)
BASE = {1: "depth_tri", 2: "depth_mesh", 3: "mesh_tri", 4: "tri"}


def _host():
    import subprocess
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "multiview_b200", "host")])
    return C.CDLL(os.path.join(ROOT, "multiview_b200", "libmvba_host.so"))


def _weights(pr):
    w = {3: 0.0, 4: 0.0}
    for kind, (_, _, wt, _) in pr.priors.items():
        w[kind] = wt
    return np.array([pr.depth_tri_weight, pr.depth_mesh_weight, w[3], w[4]])


def report(host, pr, b, tag):
    kind, row0, src = b.residual_layout()
    _, r = b.evaluate(False)
    names = "\n".join(s.name for s in pr.sensors).encode()
    text = C.create_string_buffer(1 << 16); table = np.zeros(5 * 64)
    w = _weights(pr)
    dp, ip, lp = C.POINTER(C.c_double), C.POINTER(C.c_int), C.POINTER(C.c_longlong)
    n = host.mvh_residual_stats(C.c_longlong(len(kind)), kind.ctypes.data_as(ip), row0.ctypes.data_as(lp), src.ctypes.data_as(lp),
                                pr.cid.ctypes.data_as(ip), pr.cam_sensor.ctypes.data_as(ip), names, w.ctypes.data_as(dp),
                                C.c_longlong(len(r)), r.ctypes.data_as(dp), tag.encode(), text, len(text), table.ctypes.data_as(dp), len(table))
    assert n > 0
    return text.value.decode(), table[:5 * n].reshape(n, 5), (kind, row0, src, r, w)


def numpy_stats(pr, kind, row0, src, r, w):
    groups = {}
    for b in range(len(kind)):
        if kind[b] == 0:
            cam = pr.sensors[pr.cam_sensor[pr.cid[src[b]]]].name
            rows = [(cam + "_pix_x", 1.0), (cam + "_pix_y", 1.0)]
        else:
            rows = [(BASE[kind[b]] + ax, w[kind[b] - 1]) for ax in ("_x_m", "_y_m", "_z_m")]
        for a, (nm, sc) in enumerate(rows):
            groups.setdefault(nm, []).append(abs(r[row0[b] + a] / sc))
    out = []
    for nm in sorted(groups):
        v = np.sort(np.array(groups[nm])); n = len(v)
        out.append((nm, v[int(0.25 * n)], v[int(0.5 * n)], v[int(0.75 * n)], v[n - 1], n))
    return out


def test_residual_report_matches_a_numpy_restatement(oracle_lib):
    host = _host()
    pr = synth.make_problem(seed=3, **CASE)
    o = _abi.Binding(oracle_lib, "oracle_", pr)
    text, table, raw = report(host, pr, o, "before opt")
    o.close()
    want = numpy_stats(pr, *raw)
    assert len(want) == len(table) == 3 * 2 + 4 * 3          # three sensors x (x, y) and four 3-row kinds
    lines = text.strip().split("\n")
    assert lines[0] == "The 25, 50, 75, and 100th percentile residual stats before opt"
    for (nm, a, b, c, d, n), row, line in zip(want, table, lines[1:]):
        assert np.array_equal(row, [a, b, c, d, n])
        assert line == f"{nm}: {a:.5g} {b:.5g} {c:.5g} {d:.5g} ({n} residuals)"   # std::setprecision(5), default float format
    # the depth rows are in metres again: dividing by the weight undoes the scaling of the residual
    depth = dict((w[0], w) for w in want)["depth_tri_x_m"]
    assert depth[4] < 1.0


@pytest.mark.gpu
def test_device_residual_report_equals_the_oracle_report(oracle_lib, product_lib):
    host = _host()
    pr = synth.make_problem(seed=3, **CASE)
    o = _abi.Binding(oracle_lib, "oracle_", pr)
    text_o, table_o, _ = report(host, pr, o, "after opt")
    o.close()
    g = _abi.Binding(product_lib, "mvba_", pr)
    text_g, table_g, _ = report(host, pr, g, "after opt")
    g.close()
    assert np.array_equal(table_o[:, 4], table_g[:, 4])
    assert np.allclose(table_o[:, :4], table_g[:, :4], rtol=1e-9, atol=1e-12)
    assert [l.split(":")[0] for l in text_o.split("\n")] == [l.split(":")[0] for l in text_g.split("\n")]
