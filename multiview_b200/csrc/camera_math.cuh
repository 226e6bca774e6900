// camera_math.cuh -- FP64 device math of the hot path: pose interpolation tables and the
// projection + lens distortion with analytic Jacobians.  Written from the reference's behaviour:
//   array_to_rigid_transform   rig_calibrator/src/dense_map_utils.cc:171-178
//   linearInterp (slerp)       rig_calibrator/src/dense_map_utils.cc:315-329
//   calc_world_to_cam_trans    rig_calibrator/bin/rig_calibrator.cc:322-390
//   DistortCentered / Convert  camera_model/camera_params.cc:260-314, 412-415
//   compute_rpc                camera_model/rpc_distortion.cc:116-160
// The reference differentiates numerically (DynamicNumericDiffCostFunction); the closed forms
// here are checked against forward-mode duals of the restated functors in tests/.
#pragma once
#include "common.cuh"

namespace mvba {

// q = (x y z w) storage order.  Rotation matrix of a unit quaternion, Eigen toRotationMatrix form.
__device__ __forceinline__ void quat_to_R(const double* q, double* R) {
  double x = q[0], y = q[1], z = q[2], w = q[3];
  double tx = 2 * x, ty = 2 * y, tz = 2 * z;
  double twx = tx * w, twy = ty * w, twz = tz * w;
  double txx = tx * x, txy = ty * x, txz = tz * x;
  double tyy = ty * y, tyz = tz * y, tzz = tz * z;
  R[0] = 1 - (tyy + tzz); R[1] = txy - twz; R[2] = txz + twy;
  R[3] = txy + twz; R[4] = 1 - (txx + tzz); R[5] = tyz - twx;
  R[6] = txz - twy; R[7] = tyz + twx; R[8] = 1 - (txx + tyy);
}

// G(q): 3x4 map from a quaternion perturbation to the left angular perturbation of R(q):
//   dR = [G dq]x R,  G = 2 [ w I + [v]x | -v ]   (columns x y z w).  G q = 0.
__device__ __forceinline__ void quat_G(const double* q, double* G) {
  double x = q[0], y = q[1], z = q[2], w = q[3];
  G[0] = 2 * w;  G[1] = -2 * z; G[2] = 2 * y;  G[3] = -2 * x;
  G[4] = 2 * z;  G[5] = 2 * w;  G[6] = -2 * x; G[7] = -2 * y;
  G[8] = -2 * y; G[9] = 2 * x;  G[10] = 2 * w; G[11] = -2 * z;
}

// Normalise q (4) -> qn, returns 1/|q|.
__device__ __forceinline__ double quat_normalize(const double* q, double* qn) {
  double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  double inv = 1.0 / n;
  qn[0] = q[0] * inv; qn[1] = q[1] * inv; qn[2] = q[2] * inv; qn[3] = q[3] * inv;
  return inv;
}

__device__ __forceinline__ void mat3_mul(const double* A, const double* B, double* C) {
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) C[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
}
__device__ __forceinline__ void mat3_vec(const double* A, const double* v, double* o) {
#pragma unroll
  for (int i = 0; i < 3; i++) o[i] = A[3 * i] * v[0] + A[3 * i + 1] * v[1] + A[3 * i + 2] * v[2];
}
// C(3x4) = A(3x3) * B(3x4)
__device__ __forceinline__ void mat3_mul34(const double* A, const double* B, double* C) {
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) C[4 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[4 + j] + A[3 * i + 2] * B[8 + j];
}

// d/dtheta of sin(k theta)/sin(theta); series for small theta avoids the cancellation.
__device__ __forceinline__ double dsinc_ratio(double k, double theta, double st, double ct) {
  if (theta < 1e-3) {
    double k2 = k * k, t2 = theta * theta;
    return k * (1 - k2) * theta * (1.0 / 3.0 + (7 - 3 * k2) * t2 / 90.0);
  }
  return (k * cos(k * theta) * st - sin(k * theta) * ct) / (st * st);
}

// Fills the table of camera `c` from the parameter vector x.  Returns false when the
// interpolation fraction leaves [0,1] (LOG(FATAL) at rig_calibrator.cc:346).
__device__ inline bool prep_camera(const View& v, const double* __restrict__ x, int c, CamTab& o) {
  const int p0 = v.cam_p0[c], p1 = v.cam_p1[c], t = v.cam_sensor[c];
  const double bs = v.cam_bs[c], es = v.cam_es[c], cs = v.cam_cs[c];
  const double* a0 = x + v.x_pose + 7 * p0;
  double q0[4], inv0 = quat_normalize(a0 + 3, q0);
  bool ok = true;
  if (bs == es) {  // ref camera, --no_rig, or degenerate bracket: world_to_cam = T(beg)
    quat_to_R(q0, o.R);
    o.t[0] = a0[0]; o.t[1] = a0[1]; o.t[2] = a0[2];
#pragma unroll
    for (int i = 0; i < 9; i++) o.Rrc[i] = (i % 4 == 0) ? 1.0 : 0.0;
    o.trc[0] = o.trc[1] = o.trc[2] = 0;
    double G[12];
    quat_G(q0, G);
#pragma unroll
    for (int i = 0; i < 12; i++) { o.W0[i] = G[i] * inv0; o.W1[i] = 0; o.Wrc[i] = 0; }
#pragma unroll
    for (int i = 0; i < 3; i++) { o.wa[i] = 0; o.dt[i] = 0; }
    o.alpha = 0; o.dalpha = 0; o.rig_on = 0; o.pad = 0;
    return true;
  }
  const double* a1 = x + (p1 >= 0 ? v.x_pose + 7 * p1 : v.x_ident);
  const double* ar = x + v.sens[t].x_rig;
  const double off = x[v.sens[t].x_off];
  double q1[4], inv1 = quat_normalize(a1 + 3, q1);
  double qr[4], invr = quat_normalize(ar + 3, qr);
  const double alpha = ((cs - bs) - off) / (es - bs);
  if (!(alpha >= 0.0 && alpha <= 1.0)) ok = false;
  // slerp(alpha, q0, q1), Eigen semantics (sign flip for the short arc, lerp when |dot| >= 1-eps)
  double d = q0[0] * q1[0] + q0[1] * q1[1] + q0[2] * q1[2] + q0[3] * q1[3];
  double sg = d < 0 ? -1.0 : 1.0;
  double cabs = fabs(d);
  double s0, s1, ks0, ks1, as0, as1;
  if (cabs >= 1.0 - 2.220446049250313e-16) {
    s0 = 1 - alpha; s1 = alpha; ks0 = ks1 = 0; as0 = -1; as1 = 1;
  } else {
    double th = acos(cabs), st = sin(th), ct = cabs;
    s0 = sin((1 - alpha) * th) / st; s1 = sin(alpha * th) / st;
    double dth = -1.0 / st;  // dtheta/dc
    ks0 = dsinc_ratio(1 - alpha, th, st, ct) * dth;
    ks1 = dsinc_ratio(alpha, th, st, ct) * dth;
    as0 = -th * cos((1 - alpha) * th) / st;
    as1 = th * cos(alpha * th) / st;
  }
  double q[4], kv[4], dqa[4];
#pragma unroll
  for (int i = 0; i < 4; i++) {
    q[i] = s0 * q0[i] + sg * s1 * q1[i];
    kv[i] = q0[i] * ks0 + sg * q1[i] * ks1;
    dqa[i] = q0[i] * as0 + sg * q1[i] * as1;
  }
  double Ri[9], G[12];
  quat_to_R(q, Ri);
  quat_G(q, G);
  // M0 = dq/da = s0 I + kv (sg q1)^T ; M1 = dq/db = sg s1 I + kv (sg q0)^T
  // Omega0 = G M0 N0, N0 = (I - q0 q0^T) inv0   =>  G M0 N0 = (G M0 - (G M0 q0) q0^T) inv0
  double Gk[3];
#pragma unroll
  for (int i = 0; i < 3; i++) Gk[i] = G[4 * i] * kv[0] + G[4 * i + 1] * kv[1] + G[4 * i + 2] * kv[2] + G[4 * i + 3] * kv[3];
  double Om0[12], Om1[12];
#pragma unroll
  for (int i = 0; i < 3; i++) {
    double m0[4], m1[4], m0q = 0, m1q = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      m0[j] = s0 * G[4 * i + j] + Gk[i] * sg * q1[j];
      m1[j] = sg * s1 * G[4 * i + j] + Gk[i] * sg * q0[j];
      m0q += m0[j] * q0[j]; m1q += m1[j] * q1[j];
    }
#pragma unroll
    for (int j = 0; j < 4; j++) { Om0[4 * i + j] = (m0[j] - m0q * q0[j]) * inv0; Om1[4 * i + j] = (m1[j] - m1q * q1[j]) * inv1; }
  }
  double wal[3];
#pragma unroll
  for (int i = 0; i < 3; i++) wal[i] = G[4 * i] * dqa[0] + G[4 * i + 1] * dqa[1] + G[4 * i + 2] * dqa[2] + G[4 * i + 3] * dqa[3];
  quat_to_R(qr, o.Rrc);
  o.trc[0] = ar[0]; o.trc[1] = ar[1]; o.trc[2] = ar[2];
  mat3_mul(o.Rrc, Ri, o.R);
  double ti[3], dti[3];
#pragma unroll
  for (int i = 0; i < 3; i++) { ti[i] = (1 - alpha) * a0[i] + alpha * a1[i]; dti[i] = a1[i] - a0[i]; }
  mat3_vec(o.Rrc, ti, o.t);
#pragma unroll
  for (int i = 0; i < 3; i++) o.t[i] += o.trc[i];
  mat3_mul34(o.Rrc, Om0, o.W0);
  mat3_mul34(o.Rrc, Om1, o.W1);
  mat3_vec(o.Rrc, wal, o.wa);
  mat3_vec(o.Rrc, dti, o.dt);
  double Gr[12];
  quat_G(qr, Gr);
#pragma unroll
  for (int i = 0; i < 12; i++) o.Wrc[i] = Gr[i] * invr;
  o.alpha = alpha; o.dalpha = -1.0 / (es - bs); o.rig_on = 1.0; o.pad = 0;
  return ok;
}

// Depth-to-image table of sensor t (rig_calibrator.cc:557-566): Mc = s L m + td.
__device__ inline void prep_depth(const View& v, const double* __restrict__ x, int t, DepthTab& o) {
  const double* a = x + v.sens[t].x_d2i;
  o.s = x[v.sens[t].x_scale];
  if (v.nd2i == 12) {
#pragma unroll
    for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) o.L[3 * r + c] = a[4 * r + c]; o.td[r] = a[4 * r + 3]; }
#pragma unroll
    for (int i = 0; i < 12; i++) o.Wd[i] = 0;
  } else {
    double q[4], inv = quat_normalize(a + 3, q), G[12];
    quat_to_R(q, o.L);
    quat_G(q, G);
#pragma unroll
    for (int i = 0; i < 12; i++) o.Wd[i] = G[i] * inv;
    o.td[0] = a[0]; o.td[1] = a[1]; o.td[2] = a[2];
  }
}

// Packed order of the RPC coefficients (rpc_distortion.cc:116-160): numerator of x (all monomials up to `deg`, degree-major,
// inside a degree by the power of y), denominator of x (the same without the constant term, which is 1), numerator of y,
// denominator of y.  rpc_tri(d) + i is the place of ux^(d-i) uy^i in a numerator block; a denominator block is one shorter.
struct RpcLayout { int o1, o2, o3, n; };   // first entry of den_x, num_y, den_y; number of coefficients of one half
__host__ __device__ __forceinline__ int rpc_tri(int d) { return d * (d + 1) / 2; }
__host__ __device__ __forceinline__ RpcLayout rpc_layout(int deg) {
  const int nn = (deg + 1) * (deg + 2) / 2;
  return RpcLayout{nn, 2 * nn - 1, 3 * nn - 1, 4 * nn - 2};
}

// One BracketedCamError evaluation (rig_calibrator.cc:447-482) given Xc = world_to_cam * X.
//   r = Convert<UNDISTORTED_C,DISTORTED>(f * Xc.xy / Xc.z) - meas
// Outputs: r[2]; A = dr/dXc (2x3 row-major); Jf = dr/dfocal (2); Jc: dr/dcenter = Jc * I2;
// Jd0/Jd1: dr/d(distortion) rows, nd_nz entries each (may be null).
template <bool WANT_J>
__device__ __forceinline__ void project_distort(const SensorDev& s, const double* __restrict__ x, const double* Xc,
                                                float2 meas, double* r, double* A, double* Jf, double* Jc,
                                                double* Jd0, double* Jd1, const double* __restrict__ fov) {
  const double f = x[s.x_focal];
  const double cx = x[s.x_center], cy = x[s.x_center + 1];
  const double iz = 1.0 / Xc[2];
  const double px = Xc[0] * iz, py = Xc[1] * iz;  // hnormalized()
  const double hx = s.half[0], hy = s.half[1];
  double dx, dy;                 // distorted pixel
  double dpx[2], dpy[2];         // d(dx,dy)/d(px), d(dx,dy)/d(py)   (columns)
  const double* dist = x + s.x_dist;
  if (s.nd == 0) {
    double ux = f * px, uy = f * py;
    dx = (ux + cx - hx) + hx; dy = (uy + cy - hy) + hy;
    if (WANT_J) { dpx[0] = f; dpx[1] = 0; dpy[0] = 0; dpy[1] = f; Jf[0] = px; Jf[1] = py; *Jc = 1.0; }
  } else if (s.nd == 1) {
    // FOV model, camera_params.cc:269-281
    // (the reference normalises (f p) / f; the quotient is p up to one rounding, and p is used directly)
    const double nx = px, ny = py;
    const double ru = sqrt(nx * nx + ny * ny);
    const double k = fov[0];             // 2 tan(w / 2), once per sensor and evaluation point (prep_cameras_kernel)
    const double th = 0.5 * k;
    const double at = atan(ru * k);
    const double rd = at * fov[1];       // 1 / w
    const double iru = 1.0 / ru;
    double conv = 1.0;
    if (ru > 1e-5) conv = rd * iru;
    dx = (cx - hx) + conv * (nx * f) + hx; dy = (cy - hy) + conv * (ny * f) + hy;
    if (WANT_J) {
      double dconv_dru = 0, dconv_dw = 0;
      if (ru > 1e-5) {
        const double iden = 1.0 / (1.0 + k * k * ru * ru), iw = fov[1];
        dconv_dru = (k * ru * iden - at) * (iw * iru * iru);
        double dk = 1.0 + th * th;  // d(2 tan(w/2))/dw
        dconv_dw = (dk * iden) * iw - at * (iw * iw * iru);
      }
      double g = (ru > 0) ? dconv_dru * iru : 0.0;
      dpx[0] = f * (conv + g * nx * nx); dpx[1] = f * g * nx * ny;
      dpy[0] = f * g * nx * ny; dpy[1] = f * (conv + g * ny * ny);
      Jf[0] = conv * nx; Jf[1] = conv * ny; *Jc = 1.0;
      if (Jd0) { Jd0[0] = f * nx * dconv_dw; Jd1[0] = f * ny * dconv_dw; }
    }
  } else if (s.nd == 4 || s.nd == 5) {
    // Tsai / radtan, camera_params.cc:282-308
    const double k1 = dist[0], k2 = dist[1], p1 = dist[2], p2 = dist[3];
    const double k3 = (s.nd == 5) ? dist[4] : 0.0;
    const double nx = px, ny = py;  // (f p) / f in the reference
    const double r2 = nx * nx + ny * ny;
    const double rad = 1.0 + k1 * r2 + k2 * r2 * r2 + k3 * r2 * r2 * r2;
    const double ex = rad * nx + (2.0 * p1 * nx * ny + p2 * (r2 + 2.0 * nx * nx));
    const double ey = rad * ny + (p1 * (r2 + 2.0 * ny * ny) + 2.0 * p2 * nx * ny);
    dx = ex * f + (cx - hx) + hx; dy = ey * f + (cy - hy) + hy;
    if (WANT_J) {
      const double drad = k1 + 2.0 * k2 * r2 + 3.0 * k3 * r2 * r2;
      dpx[0] = f * (rad + 2.0 * drad * nx * nx + 2.0 * p1 * ny + 6.0 * p2 * nx);
      dpx[1] = f * (2.0 * drad * nx * ny + 2.0 * p1 * nx + 2.0 * p2 * ny);
      dpy[0] = f * (2.0 * drad * nx * ny + 2.0 * p1 * nx + 2.0 * p2 * ny);
      dpy[1] = f * (rad + 2.0 * drad * ny * ny + 6.0 * p1 * ny + 2.0 * p2 * nx);
      Jf[0] = ex; Jf[1] = ey; *Jc = 1.0;
      if (Jd0) {
        Jd0[0] = f * r2 * nx; Jd1[0] = f * r2 * ny;
        Jd0[1] = f * r2 * r2 * nx; Jd1[1] = f * r2 * r2 * ny;
        Jd0[2] = f * 2.0 * nx * ny; Jd1[2] = f * (r2 + 2.0 * ny * ny);
        Jd0[3] = f * (r2 + 2.0 * nx * nx); Jd1[3] = f * 2.0 * nx * ny;
        if (s.nd == 5) { Jd0[4] = f * r2 * r2 * r2 * nx; Jd1[4] = f * r2 * r2 * r2 * ny; }
      }
    }
  } else {
    // RPC, camera_params.cc:309-312 -> compute_rpc (rpc_distortion.cc:116-160): no focal
    // normalisation and no optical-centre term.
    const int deg = s.rpc_deg;
    const double ux = f * px, uy = f * py;
    // The monomials ux^a uy^i are walked i-major with running products (no power tables: a dynamically indexed local array
    // would live in local memory); rpc_ci gives the coefficient's place in the reference's packed order.
    const RpcLayout L = rpc_layout(deg);
    double v0 = 0, v1 = 1, v2 = 0, v3 = 1, x0 = 0, x1 = 0, x2 = 0, x3 = 0, y0 = 0, y1 = 0, y2 = 0, y3 = 0;
    {
      double yi = 1.0, yi1 = 0.0;  // uy^i, uy^(i-1)
      for (int i = 0; i <= deg; i++) {
        double m = yi, mx = 0.0, my = yi1;  // ux^a uy^i, ux^(a-1) uy^i, ux^a uy^(i-1)
        for (int a = 0; a + i <= deg; a++) {
          const int o = rpc_tri(a + i) + i;
          const double c0 = dist[o], c2 = dist[L.o2 + o];
          const double c1 = (a + i > 0) ? dist[L.o1 + o - 1] : 0.0, c3 = (a + i > 0) ? dist[L.o3 + o - 1] : 0.0;
          v0 = fma(c0, m, v0); v1 = fma(c1, m, v1); v2 = fma(c2, m, v2); v3 = fma(c3, m, v3);
          if (WANT_J) {
            const double dxm = a * mx, dym = i * my;
            x0 = fma(c0, dxm, x0); x1 = fma(c1, dxm, x1); x2 = fma(c2, dxm, x2); x3 = fma(c3, dxm, x3);
            y0 = fma(c0, dym, y0); y1 = fma(c1, dym, y1); y2 = fma(c2, dym, y2); y3 = fma(c3, dym, y3);
          }
          mx = m; m *= ux; my *= ux;
        }
        yi1 = yi; yi *= uy;
      }
    }
    const double idx = 1.0 / v1, idy = 1.0 / v3;
    const double ox = v0 * idx, oy = v2 * idy;
    dx = ox + hx; dy = oy + hy;
    if (WANT_J) {
      // d(out)/d(u) then u = f p
      double dox_dux = (x0 - ox * x1) * idx, dox_duy = (y0 - ox * y1) * idx;
      double doy_dux = (x2 - oy * x3) * idy, doy_duy = (y2 - oy * y3) * idy;
      dpx[0] = f * dox_dux; dpx[1] = f * doy_dux;
      dpy[0] = f * dox_duy; dpy[1] = f * doy_duy;
      Jf[0] = dox_dux * px + dox_duy * py; Jf[1] = doy_dux * px + doy_duy * py;
      *Jc = 0.0;
      if (Jd0) {
        double yi = 1.0;
        for (int i = 0; i <= deg; i++) {
          double m = yi;
          for (int a = 0; a + i <= deg; a++) {
            const int o = rpc_tri(a + i) + i;
            Jd0[o] = m * idx; Jd1[o] = 0.0;
            Jd0[L.o2 + o] = 0.0; Jd1[L.o2 + o] = m * idy;
            if (a + i > 0) {
              Jd0[L.o1 + o - 1] = -ox * m * idx; Jd1[L.o1 + o - 1] = 0.0;
              Jd0[L.o3 + o - 1] = 0.0; Jd1[L.o3 + o - 1] = -oy * m * idy;
            }
            m *= ux;
          }
          yi *= uy;
        }
      }
    }
  }
  r[0] = dx - (double)meas.x;
  r[1] = dy - (double)meas.y;
  if (WANT_J) {
    // dp/dXc = [[iz, 0, -px iz], [0, iz, -py iz]]
    A[0] = dpx[0] * iz; A[1] = dpy[0] * iz; A[2] = -(dpx[0] * px + dpy[0] * py) * iz;
    A[3] = dpx[1] * iz; A[4] = dpy[1] * iz; A[5] = -(dpx[1] * px + dpy[1] * py) * iz;
  }
}

}  // namespace mvba
