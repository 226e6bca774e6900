"""multiview_b200/host/bracket.h (lookupImage / lookupImagesAndBrackets, dense_map_utils.cc:702-739, 1441-1660): which image of
which sensor goes between which reference frames, and the timestamp-offset bounds -- against a hand-made timeline and against a
brute-force numpy selection on random timelines."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)


@pytest.fixture(scope="module")
def host():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "multiview_b200", "host")])
    lib = C.CDLL(os.path.join(ROOT, "multiview_b200", "libmvba_host.so"))
    lib.mvh_camera_last_error.restype = C.c_char_p
    return lib


def bracket(host, ref, stamps, offsets, bracket_len=0.6, max_change=100.0, clouds=None):
    T = len(stamps); R = len(ref)
    ref = np.ascontiguousarray(ref, np.float64); counts = np.array([len(s) for s in stamps], np.int32)
    flat = np.ascontiguousarray(np.concatenate(stamps), np.float64); offsets = np.ascontiguousarray(offsets, np.float64)
    cc = cf = None
    if clouds is not None:
        cc = np.array([len(s) for s in clouds], np.int32); cf = np.ascontiguousarray(np.concatenate(clouds), np.float64)
    n = R * T
    sensor = np.zeros(n, np.int32); beg = np.zeros(n, np.int32); end = np.zeros(n, np.int32); ts = np.zeros(n)
    img = np.zeros(n, np.int32); cld = np.zeros(n, np.int32); lo = np.zeros(T); hi = np.zeros(T)
    k = host.mvh_lookup_images_and_brackets(C.c_double(bracket_len), C.c_double(max_change), R, ref.ctypes.data_as(dp), T, counts.ctypes.data_as(ip),
                                            flat.ctypes.data_as(dp), None if cc is None else cc.ctypes.data_as(ip),
                                            None if cf is None else cf.ctypes.data_as(dp), offsets.ctypes.data_as(dp), sensor.ctypes.data_as(ip),
                                            beg.ctypes.data_as(ip), end.ctypes.data_as(ip), ts.ctypes.data_as(dp), img.ctypes.data_as(ip),
                                            cld.ctypes.data_as(ip), lo.ctypes.data_as(dp), hi.ctypes.data_as(dp))
    if k < 0:
        raise RuntimeError(host.mvh_camera_last_error().decode())
    return [(int(sensor[i]), int(beg[i]), int(end[i]), float(ts[i]), int(img[i]), int(cld[i])) for i in range(k)], lo, hi


def test_hand_made_timeline(host):
    ref = [10.0, 11.0, 12.0]
    haz = [10.3, 10.45, 11.0, 11.9, 12.0, 12.5]        # offset 0: 10.45 is closer to the middle of [10, 11) than 10.3; 11.0 sits ON a bracket start
    sci = [10.02, 11.6]                                # offset 0.02: bracket 0 is [10.02, 11.02); 11.6 lies in bracket 1 = [11.02, 12.02)
    cams, lo, hi = bracket(host, ref, [ref, haz, sci], [0.0, 0.0, 0.02], bracket_len=2.0, clouds=[[], [10.5, 11.95], []])
    assert cams == [(0, 0, 0, 10.0, 0, -1), (1, 0, 1, 10.45, 1, 0), (2, 0, 1, 10.02, 0, -1),
                    (0, 1, 1, 11.0, 1, -1), (1, 1, 2, 11.9, 3, 1), (2, 1, 2, 11.6, 1, -1),
                    (0, 2, 2, 12.0, 2, -1), (1, 2, 2, 12.0, 4, -1)]      # last frame pairs with itself: only an image exactly on it
    # bounds: max over images of (t - ref[end]) and min of (t - ref[beg]), clipped to offset +- max_change, shrunk by 1e-5
    assert np.isclose(lo[1], max(10.45 - 11, 11.9 - 12, 12.0 - 12.0) + 1e-5) and np.isclose(hi[1], min(10.45 - 10, 11.9 - 11, 12.0 - 12.0) - 1e-5)
    assert np.isclose(lo[2], max(10.02 - 11, 11.6 - 12) + 1e-5) and np.isclose(hi[2], min(10.02 - 10, 11.6 - 11) - 1e-5)
    # a bracket longer than bracket_len keeps an image only if the one closest to its middle sits exactly on its beginning: in [11, 12)
    # that is 11.9, not the 11.0 on the boundary, so the haz image of that bracket is dropped
    cams, _, _ = bracket(host, ref, [ref, haz, sci], [0.0, 0.0, 0.02], bracket_len=0.5)
    assert [c for c in cams if c[0] != 0] == [(2, 0, 1, 10.02, 0, -1), (1, 2, 2, 12.0, 4, -1)]
    with pytest.raises(RuntimeError, match="Cannot look up camera"):
        bracket(host, ref, [[10.0, 11.5, 12.0], haz, sci], [0.0, 0.0, 0.02])
    with pytest.raises(RuntimeError, match="chronological"):
        bracket(host, ref, [ref, [10.5, 10.2, 11.2], sci], [0.0, 0.0, 0.02])


def brute_force(ref, stamps, offsets, bracket_len):
    out = []
    R = len(ref)
    for b in range(R):
        e = b + 1 if b + 1 < R else b
        out.append((0, b, b, ref[b]))
        for t in range(1, len(stamps)):
            bts, ets = ref[b] + offsets[t], ref[e] + offsets[t]
            if e == b:
                ets = np.nextafter(ets, ets + 1.0)
            s = np.asarray(stamps[t])
            cand = s[(s >= bts) & (s < ets)]
            if len(cand) == 0:
                continue
            best = cand[np.argmin(np.abs(cand - (bts + ets) / 2.0))]       # argmin keeps the earliest on ties, like the strict '<'
            if best > bts and ets - bts > bracket_len:
                continue
            out.append((t, b, e, float(best)))
    return out


@pytest.mark.parametrize("seed", range(6))
def test_random_timelines_match_brute_force(host, seed):
    rng = np.random.default_rng(seed)
    R = 40
    ref = np.cumsum(rng.uniform(0.2, 1.2, R)) + 100.0
    offsets = [0.0, float(rng.uniform(-0.05, 0.05)), float(rng.uniform(-0.05, 0.05))]
    stamps = [ref]
    for t in (1, 2):
        s = np.sort(rng.uniform(ref[0] - 1.0, ref[-1] + 1.0, 90))
        s[::15] = ref[rng.integers(0, R, len(s[::15]))] + offsets[t]          # some images exactly on a bracket boundary
        stamps.append(np.sort(s))
    cams, lo, hi = bracket(host, ref, stamps, offsets, bracket_len=0.8, max_change=0.03)
    want = brute_force(ref, stamps, offsets, 0.8)
    assert [(c[0], c[1], c[2], c[3]) for c in cams] == want
    for c in cams:
        assert stamps[c[0]][c[4]] == c[3]
    for t in (1, 2):
        sel = [c for c in cams if c[0] == t]
        assert lo[t] >= offsets[t] - 0.03 and hi[t] <= offsets[t] + 0.03
        if sel:   # every kept image stays inside its bracket for any offset in the allowed range
            assert all(ref[c[1]] <= c[3] - hi[t] + 1e-4 and c[3] - lo[t] <= ref[c[2]] + 1e-4 for c in sel if c[1] != c[2])


def test_no_rig_lookup_takes_every_image(host):
    """lookupImagesNoBrackets: one camera per image, sensor-major, beg = end = position in the sensor's list; repeated timestamps are a
    look-up failure exactly as in the reference (the forward scan stops on the first of the two)."""
    stamps = [[1.0, 2.0, 3.5], [1.2, 2.2], [0.5]]
    counts = np.array([3, 2, 1], np.int32); flat = np.concatenate(stamps)
    n = 6
    sensor = np.zeros(n, np.int32); beg = np.zeros(n, np.int32); end = np.zeros(n, np.int32); ts = np.zeros(n); img = np.zeros(n, np.int32)
    args = lambda f: (3, counts.ctypes.data_as(ip), f.ctypes.data_as(dp), sensor.ctypes.data_as(ip), beg.ctypes.data_as(ip), end.ctypes.data_as(ip),
                      ts.ctypes.data_as(dp), img.ctypes.data_as(ip))
    assert host.mvh_lookup_images_no_brackets(*args(flat)) == 6
    assert list(sensor) == [0, 0, 0, 1, 1, 2] and list(beg) == [0, 1, 2, 0, 1, 0] and list(end) == list(beg) and list(img) == list(beg)
    assert list(ts) == [1.0, 2.0, 3.5, 1.2, 2.2, 0.5]
    dup = flat.copy(); dup[1] = 1.0
    assert host.mvh_lookup_images_no_brackets(*args(dup)) == 6 and list(img[:3]) == [0, 0, 2]   # equal stamps: the forward scan stops on the first of the two
    bad = flat.copy(); bad[1] = 0.5                                      # a stamp the forward scan has already passed cannot be looked up
    assert host.mvh_lookup_images_no_brackets(*args(bad)) == -1
