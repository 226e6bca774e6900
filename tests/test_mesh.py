"""feature rays against a triangle mesh (ray_mesh_intersect / meshTriangulations,
texture_processing.cc:1436-1479, 1566-1625; float32 triangle test of rayint/libs/acc/primitives.h).

CPU: the oracle (brute force over all triangles) against the closed-form ray / plane intersection, the distance window,
the bad_xyz convention and the per-track mean.  GPU: mvba_mesh_intersect (BVH traversal kernel) against the oracle."""
import ctypes as C

import numpy as np
import pytest

from multiview_b200 import _abi, synth

BAD = 1e100
CASE = dict(sensors="rig3", num_ref=10, num_landmarks=300, obs_per_track=5, outlier_frac=0.02)


def camera_centres(pr):
    R, t = synth.world_to_cam_all(pr)
    return -np.einsum("cji,cj->ci", R, t)


def box_mesh(lo, hi, n):
    """The six faces of an axis-aligned box, each an n x n grid of quads split into two triangles."""
    verts, faces = [], []
    for axis in range(3):
        u, v = [a for a in range(3) if a != axis]
        for val in (lo[axis], hi[axis]):
            base = len(verts)
            for i in range(n + 1):
                for j in range(n + 1):
                    p = np.zeros(3); p[axis] = val
                    p[u] = lo[u] + (hi[u] - lo[u]) * i / n; p[v] = lo[v] + (hi[v] - lo[v]) * j / n
                    verts.append(p)
            for i in range(n):
                for j in range(n):
                    a = base + i * (n + 1) + j
                    faces += [[a, a + 1, a + n + 2], [a, a + n + 2, a + n + 1]]
    return np.array(verts, np.float32), np.array(faces, np.uint32)


def scene_box(pr, n):
    pts = np.concatenate([pr.xyz.reshape(-1, 3), camera_centres(pr)])
    lo, hi = pts.min(0) - 1.0, pts.max(0) + 1.5
    return box_mesh(lo, hi, n), lo, hi


def test_oracle_matches_closed_form_box_exit(oracle_lib):
    """A ray that starts inside an axis-aligned box leaves it through exactly one face: the smallest positive slab distance."""
    pr = synth.make_problem(seed=41, **CASE)
    (verts, faces), lo, hi = scene_box(pr, 3)
    o = _abi.Binding(oracle_lib, "oracle_", pr)
    has, xyz, mean = o.mesh_intersect(verts, faces, 0.01, 1e4)
    o.close()
    assert has.all()
    ctr = camera_centres(pr)[pr.cid]
    ray = xyz - ctr
    dist = np.linalg.norm(ray, axis=1)
    ray /= dist[:, None]
    lo32, hi32 = lo.astype(np.float32).astype(np.float64), hi.astype(np.float32).astype(np.float64)
    with np.errstate(divide="ignore", invalid="ignore"):
        t = np.where(ray > 0, (hi32 - ctr) / ray, (lo32 - ctr) / ray)
    want = np.nanmin(np.where(np.isfinite(t) & (t > 0), t, np.inf), axis=1)
    assert np.abs(dist - want).max() <= 2e-6 * want.max()          # float32 hit distance
    # the hit lies along the pixel's ray: it re-projects onto the measured pixel
    R, tt = synth.world_to_cam_all(pr)
    xc = np.einsum("nij,nj->ni", R[pr.cid], xyz) + tt[pr.cid]
    assert (xc[:, 2] > 0).all()
    # per-track mean of the hits
    for p in (0, 7, pr.num_landmarks - 1):
        assert np.abs(mean[p] - xyz[pr.pid == p].mean(0)).max() <= 1e-12 * np.abs(mean[p]).max()


def test_distance_window_outliers_and_bad_xyz(oracle_lib):
    pr = synth.make_problem(seed=42, **CASE)
    (verts, faces), _, _ = scene_box(pr, 2)
    o = _abi.Binding(oracle_lib, "oracle_", pr)
    has, xyz, _ = o.mesh_intersect(verts, faces, 0.01, 1e4)
    dist = np.linalg.norm(xyz - camera_centres(pr)[pr.cid], axis=1)
    cut = float(np.median(dist))
    inl = np.ones(pr.num_obs, np.uint8); inl[::5] = 0
    has2, xyz2, mean2 = o.mesh_intersect(verts, faces, 0.01, cut, inl)
    o.close()
    near = dist <= cut * (1 - 1e-6); far = dist >= cut * (1 + 1e-6)
    assert has2[near & (inl == 1)].all() and not has2[far].any() and not has2[inl == 0].any()
    assert (xyz2[has2 == 0] == BAD).all()
    for p in range(pr.num_landmarks):
        sel = (pr.pid == p) & (has2 == 1)
        if not sel.any():
            assert (mean2[p] == BAD).all()


@pytest.mark.gpu
@pytest.mark.parametrize("sensors", ["pinhole", "rig3", "rpc"])
def test_gpu_mesh_intersection_matches_oracle(sensors, oracle_lib, product_lib):
    pr = synth.make_problem(seed=43, **dict(CASE, sensors=sensors, float_rig=sensors == "rig3"))
    (verts, faces), _, _ = scene_box(pr, 24)                          # 6912 triangles: a few BVH levels
    rng = np.random.default_rng(2)
    inl = (rng.random(pr.num_obs) > 0.1).astype(np.uint8)
    o = _abi.Binding(oracle_lib, "oracle_", pr); g = _abi.Binding(product_lib, "mvba_", pr)
    ho, xo, mo = o.mesh_intersect(verts, faces, 0.05, 30.0, inl)
    hg, xg, mg = g.mesh_intersect(verts, faces, 0.05, 30.0, inl)
    o.close(); g.close()
    assert ho.sum() > 0.5 * inl.sum()
    assert (ho == hg).all()
    scale = np.abs(xo[ho == 1]).max()
    assert np.abs(xo - xg)[ho == 1].max() <= 1e-6 * scale              # float32 hit distance, 1 ulp = 6e-8 relative
    assert (xg[hg == 0] == BAD).all()
    ok = (mo != BAD).all(axis=1)
    assert ((mg != BAD).all(axis=1) == ok).all() and np.abs(mo - mg)[ok].max() <= 1e-6 * scale
