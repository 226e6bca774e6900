"""Generates tests/golden/rayint.npz from the REFERENCE's own ray / mesh code (oracle/_ref/libref_rayint.so, built by
oracle/ref_build/Makefile from /root/reference/rayint/libs): meshes, float32 rays and the reference's hit flag, distance and
triangle index for each ray.  Run in the container that has /root/reference; the vectors travel, the reference does not."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def ref_lib():
    path = os.path.join(ROOT, "oracle", "_ref", "libref_rayint.so")
    if not os.path.exists(path):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle", "ref_build")])
    lib = C.CDLL(path)
    lib.ref_rayint_intersect.restype = C.c_int64
    return lib


def ref_intersect(lib, verts, faces, origin, direction, tmin, tmax, threads=1):
    fp, up = C.POINTER(C.c_float), C.POINTER(C.c_uint32)
    n = len(origin)
    hit = np.zeros(n, np.uint8); t = np.zeros(n, np.float32); idx = np.zeros(n, np.uint32)
    lib.ref_rayint_intersect(C.c_int64(len(verts)), verts.ctypes.data_as(fp), C.c_int64(len(faces)), faces.ctypes.data_as(up), C.c_int64(n),
                             origin.ctypes.data_as(fp), direction.ctypes.data_as(fp), C.c_float(tmin), C.c_float(tmax), C.c_int32(threads),
                             hit.ctypes.data_as(C.POINTER(C.c_uint8)), t.ctypes.data_as(fp), idx.ctypes.data_as(up))
    return hit, t, idx


def terrain(rng, n, size=8.0, rough=0.6):
    """A height-field mesh: (n + 1)^2 vertices, 2 n^2 triangles, with a few degenerate and duplicated triangles mixed in."""
    xs = np.linspace(-size, size, n + 1)
    X, Y = np.meshgrid(xs, xs, indexing="ij")
    Z = rough * rng.normal(size=X.shape) + 0.15 * X
    verts = np.stack([X.ravel(), Y.ravel(), Z.ravel()], 1).astype(np.float32)
    faces = []
    for i in range(n):
        for j in range(n):
            a = i * (n + 1) + j
            faces += [[a, a + 1, a + n + 2], [a, a + n + 2, a + n + 1]]
    faces += [[0, 0, 1], [5, 5, 5], faces[7], faces[11][::-1]]          # degenerate (zero normal), duplicate, flipped duplicate
    return verts, np.array(faces, np.uint32)


def rays(rng, n, height=6.0, spread=7.0):
    o = np.stack([rng.uniform(-spread, spread, n), rng.uniform(-spread, spread, n), rng.uniform(2.0, height, n)], 1)
    d = np.stack([rng.normal(size=n), rng.normal(size=n), -np.abs(rng.normal(size=n)) - 0.2], 1)
    d[: n // 10, 2] *= -1                                              # some rays leave upwards: no hit
    # the caller's normalisation, float32 (texture_processing.cc:1462-1466 normalises a Vec3f)
    o32 = o.astype(np.float32); d32 = d.astype(np.float32)
    d32 /= np.sqrt((d32 * d32).sum(1, dtype=np.float32))[:, None]
    # a share of rays aimed exactly at mesh vertices / edge midpoints: the barycentric eps = 1e-3 rule decides
    return np.ascontiguousarray(o32), np.ascontiguousarray(d32)


def main():
    lib = ref_lib()
    rng = np.random.default_rng(20261104)
    out = {}
    for k, (n, nr) in enumerate([(6, 400), (24, 1500), (48, 1500)]):
        verts, faces = terrain(rng, n)
        o, d = rays(rng, nr)
        # aim a tenth of the rays at vertices and edge midpoints of the mesh
        m = nr // 10
        tgt = verts[rng.integers(0, len(verts), m)].astype(np.float64)
        f = faces[rng.integers(0, len(faces) - 4, m)]
        tgt[m // 2:] = 0.5 * (verts[f[m // 2:, 0]].astype(np.float64) + verts[f[m // 2:, 1]].astype(np.float64))
        dd = (tgt - o[:m].astype(np.float64)).astype(np.float32)
        dd /= np.sqrt((dd * dd).sum(1, dtype=np.float32))[:, None]
        d[:m] = dd
        for tag, (tmin, tmax) in {"wide": (0.0, 1e4), "window": (3.0, 6.5)}.items():
            hit, t, idx = ref_intersect(lib, verts, faces, o, d, tmin, tmax)
            out[f"m{k}_{tag}_hit"] = hit; out[f"m{k}_{tag}_t"] = t; out[f"m{k}_{tag}_idx"] = idx
            out[f"m{k}_{tag}_range"] = np.array([tmin, tmax], np.float32)
            print(f"mesh {k} ({len(faces)} triangles) {tag}: {int(hit.sum())} of {nr} rays hit")
        out[f"m{k}_verts"] = verts; out[f"m{k}_faces"] = faces; out[f"m{k}_origin"] = o; out[f"m{k}_dir"] = d
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "rayint.npz"), **out)


if __name__ == "__main__":
    sys.exit(main())
