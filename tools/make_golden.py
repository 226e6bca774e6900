"""Generates tests/golden/*.npz: known-answer vectors for the oracle, from INDEPENDENT implementations
(scipy Rotation/Slerp, cv2.projectPoints, mpmath at 40 digits) -- not from the oracle and not from the
reference (the reference ships no tests or fixtures for this path and cannot be built here: SURVEY.md 8(c)).
Every vector is labelled "restatement-derived".  Run:  python tools/make_golden.py
"""
import os
import sys

import cv2
import mpmath as mp
import numpy as np
from scipy.spatial.transform import Rotation, Slerp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden")
rng = np.random.default_rng(20261104)


def pose_interp():
    """calc_world_to_cam_trans (rig_calibrator.cc:362-390) via scipy: world_to_cam = T(ref_to_cam) * interp(world_to_ref)."""
    n = 64
    beg = np.concatenate([rng.normal(0, 2, (n, 3)), rng.normal(size=(n, 4))], axis=1)   # un-normalised quaternions on purpose
    end = beg + np.concatenate([rng.normal(0, 0.2, (n, 3)), rng.normal(0, 0.15, (n, 4))], axis=1)
    end[::9, 3:] *= -1.0                                                                # antipodal representation: shortest arc
    r2c = np.concatenate([rng.normal(0, 0.1, (n, 3)), rng.normal(size=(n, 4))], axis=1)
    bs = 1000.0 + rng.uniform(0, 100, n)
    es = bs + rng.uniform(0.1, 1.0, n)
    off = rng.uniform(-0.02, 0.02, n)
    alpha = rng.uniform(0, 1, n)
    alpha[:4] = [0.0, 1.0, 0.5, 1e-9]
    cs = bs + off + alpha * (es - bs)
    out = np.zeros((n, 12))
    for i in range(n):
        a = ((cs[i] - bs[i]) - off[i]) / (es[i] - bs[i])
        R0 = Rotation.from_quat(beg[i, 3:]); R1 = Rotation.from_quat(end[i, 3:])
        Ri = Slerp([0.0, 1.0], Rotation.concatenate([R0, R1]))([a])[0].as_matrix()
        ti = (1 - a) * beg[i, :3] + a * end[i, :3]
        Rrc = Rotation.from_quat(r2c[i, 3:]).as_matrix()
        R = Rrc @ Ri
        t = Rrc @ ti + r2c[i, :3]
        out[i] = np.concatenate([R, t[:, None]], axis=1).reshape(-1)
    # degenerate bracket (beg stamp == end stamp): world_to_cam = T(beg), rig transform ignored (:371-374)
    deg = np.zeros((8, 12))
    for i in range(8):
        R = Rotation.from_quat(beg[i, 3:]).as_matrix()
        deg[i] = np.concatenate([R, beg[i, :3][:, None]], axis=1).reshape(-1)
    np.savez(os.path.join(OUT, "pose_interp.npz"), beg=beg, end=end, r2c=r2c, bs=bs, es=es, off=off, cs=cs, world_to_cam=out, degenerate=deg)


def fov_mp(u, f, c, w):
    """FOV / fisheye model (camera_params.cc:269-281) at 40 digits."""
    mp.mp.dps = 40
    nx, ny = mp.mpf(u[0]) / mp.mpf(f), mp.mpf(u[1]) / mp.mpf(f)
    ru = mp.sqrt(nx * nx + ny * ny)
    rd = mp.atan(ru * 2 * mp.tan(mp.mpf(w) / 2)) / mp.mpf(w)
    conv = rd / ru if ru > mp.mpf("1e-5") else mp.mpf(1)
    return [float(mp.mpf(c[0]) + conv * nx * mp.mpf(f)), float(mp.mpf(c[1]) + conv * ny * mp.mpf(f))]


def rpc_mp(u, coeffs, deg):
    """compute_rpc (rpc_distortion.cc:116-160) at 40 digits: num_x | den_x | num_y | den_y, monomials x^(d-i) y^i."""
    mp.mp.dps = 40
    x, y = mp.mpf(u[0]), mp.mpf(u[1])
    vals = [mp.mpf(0), mp.mpf(1), mp.mpf(0), mp.mpf(1)]
    ci = 0
    for cnt in range(4):
        for d in range(1 if cnt in (1, 3) else 0, deg + 1):
            for i in range(d + 1):
                vals[cnt] += mp.mpf(coeffs[ci]) * x ** (d - i) * y ** i
                ci += 1
    return [float(vals[0] / vals[1]), float(vals[2] / vals[3])]


def distortion():
    n = 200
    u = rng.uniform(-500, 500, (n, 2))
    u[:3] = [[0.0, 0.0], [1e-4, -2e-4], [3e-3, 0.0]]        # through the ru <= 1e-5 branch of the FOV model
    f, c, wh = 608.0, np.array([640.5, 480.2]), np.array([1280.0, 960.0])
    fov_w = 0.998
    fov = np.array([fov_mp(u[i], f, c, fov_w) for i in range(n)])
    # radtan 4 and 5 via OpenCV: projectPoints on (x, y, 1) with K = [[f,0,cx],[0,f,cy]] (camera_params.cc:282-308)
    K = np.array([[f, 0, c[0]], [0, f, c[1]], [0, 0, 1.0]])
    d4 = np.array([-0.25, 0.08, 0.0012, -0.0008]); d5 = np.array([0.06, -0.11, 0.0005, 0.0007, 0.02])
    pts = np.concatenate([u / f, np.ones((n, 1))], axis=1).reshape(-1, 1, 3)
    rad4 = cv2.projectPoints(pts, np.zeros(3), np.zeros(3), K, np.array([d4[0], d4[1], d4[2], d4[3], 0.0]))[0].reshape(n, 2)
    rad5 = cv2.projectPoints(pts, np.zeros(3), np.zeros(3), K, np.array([d5[0], d5[1], d5[2], d5[3], d5[4]]))[0].reshape(n, 2)
    # RPC degree 2 and 3; only the first half of the block enters the distortion (rpc_distortion.cc:272-293)
    rpc = {}
    for deg in (2, 3):
        nco = 2 * (deg + 1) * (deg + 2) - 2
        L = (deg + 1) * (deg + 2) // 2
        co = np.zeros(nco)
        co[0] = 2.5; co[1] = 1.0
        co[L + (L - 1)] = -1.5; co[L + (L - 1) + 2] = 1.0
        co += rng.normal(0, 1, nco) * np.where(np.arange(nco) % 7 == 0, 1e-7, 1e-9)
        und = rng.normal(size=nco)                                # second half: must not influence the result
        vals = np.array([rpc_mp(u[i], co, deg) for i in range(n)]) + wh / 2
        rpc[f"rpc{deg}_coeffs"] = np.concatenate([co, und]); rpc[f"rpc{deg}"] = vals
    np.savez(os.path.join(OUT, "distortion.npz"), u=u, focal=f, center=c, size=wh, none=u + c, fov_w=fov_w, fov=fov,
             radtan4_coeffs=d4, radtan4=rad4, radtan5_coeffs=d5, radtan5=rad5, **rpc)


def cauchy_and_lm_rules():
    """Ceres 1.14 CauchyLoss(a): rho = b log(1 + s/b), b = a^2; Corrector with rho'' <= 0 scales by sqrt(rho')."""
    s = np.concatenate([[0.0, 1e-12, 9.0], rng.uniform(0, 400, 29)])
    a = 3.0
    mp.mp.dps = 40
    rho = np.array([float(mp.mpf(a) ** 2 * mp.log(1 + mp.mpf(x) / mp.mpf(a) ** 2)) for x in s])
    sq = np.array([float(mp.sqrt(1 / (1 + mp.mpf(x) / mp.mpf(a) ** 2))) for x in s])
    np.savez(os.path.join(OUT, "cauchy.npz"), s=s, a=a, rho=rho, sqrt_rho1=sq)

def undistortion():
    """Convert<DISTORTED, UNDISTORTED_C> (camera_params.cc:316-355): radtan through cv2.undistortPoints (the call the
    reference makes), FOV in closed form with mpmath, RPC through the second half of the coefficient block.
    Own generator so that the older fixtures stay byte-identical."""
    r2 = np.random.default_rng(20261104)
    n = 200
    f, c, wh = 608.0, np.array([640.5, 480.2]), np.array([1280.0, 960.0])
    pix = np.stack([r2.uniform(40, 1240, n), r2.uniform(30, 930, n)], axis=1)
    pix[0] = c                       # the optical centre itself (rd = 0 branch of the FOV model)
    pix[1] = c + [3e-3, -2e-3]       # rd <= 1e-5
    K = np.array([[f, 0, c[0]], [0, f, c[1]], [0, 0, 1.0]])
    Pn = np.array([[f, 0, 0.0], [0, f, 0.0], [0, 0, 1.0]])   # undistorted intrinsics minus the undistorted half size
    d4 = np.array([-0.25, 0.08, 0.0012, -0.0008]); d5 = np.array([0.06, -0.11, 0.0005, 0.0007, 0.02])
    src = pix.reshape(-1, 1, 2).astype(np.float64)
    rad4 = cv2.undistortPoints(src, K, np.array([d4[0], d4[1], d4[2], d4[3], 0.0]), None, Pn).reshape(n, 2)
    rad5 = cv2.undistortPoints(src, K, d5, None, Pn).reshape(n, 2)
    fov_w = 0.998
    fov = np.zeros((n, 2))
    for i in range(n):
        nx, ny = (mp.mpf(pix[i, 0]) - mp.mpf(c[0])) / f, (mp.mpf(pix[i, 1]) - mp.mpf(c[1])) / f
        rd = mp.sqrt(nx * nx + ny * ny)
        conv = mp.mpf(1)
        if rd > mp.mpf("1e-5"):
            conv = mp.tan(rd * mp.mpf(fov_w)) / (2 * mp.tan(mp.mpf(fov_w) / 2)) / rd
        fov[i] = [float(conv * nx * f), float(conv * ny * f)]
    deg = 2
    nco = 2 * (deg + 1) * (deg + 2) - 2
    L = (deg + 1) * (deg + 2) // 2
    und = np.zeros(nco)
    und[1] = 1.0; und[L + (L - 1) + 2] = 1.0
    und += r2.normal(0, 1, nco) * np.where(np.arange(nco) % 5 == 0, 1e-7, 1e-9)
    dis = r2.normal(size=nco)        # first half: must not influence the undistortion
    rpc = np.array([rpc_mp(pix[i] - wh / 2, und, deg) for i in range(n)])
    np.savez(os.path.join(OUT, "undistort.npz"), pix=pix, focal=f, center=c, size=wh, none=pix - c, fov_w=fov_w, fov=fov,
             radtan4_coeffs=d4, radtan4=rad4, radtan5_coeffs=d5, radtan5=rad5, rpc2_coeffs=np.concatenate([dis, und]), rpc2=rpc)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    if len(sys.argv) > 1 and sys.argv[1] == "undistort":
        undistortion(); print("wrote undistort.npz"); sys.exit(0)
    pose_interp(); distortion(); cauchy_and_lm_rules(); undistortion()
    print("wrote", sorted(os.listdir(OUT)))
