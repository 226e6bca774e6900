// items.cuh -- evaluation of one residual block ("item") of any kind and of single Jacobian columns.
// Shared by the residual kernel, the Jacobian test hook and the fused normal-equation build.
//   CAM        BracketedCamError        reference rig_calibrator.cc:447-482
//   DEPTH_TRI  BracketedDepthError      reference rig_calibrator.cc:547-583
//   DEPTH_MESH BracketedDepthMeshError  reference rig_calibrator.cc:642-675
//   PRIOR      XYZError                 reference rig_calibrator.cc:721-727
#pragma once
#include "camera_math.cuh"

namespace mvba {

// Virtual f-column layout of an item (landmark columns are handled separately):
//   0-2 t0 | 3-6 q0 | 7-9 t1 | 10-13 q1 | 14-16 t_rc | 17-20 q_rc | 21 offset        (pose part, 22)
//   CAM:   22 focal | 23-24 centre | 25.. distortion (nd_nz)
//   DEPTH: 22.. depth_to_image (nd2i) | 22+nd2i scale
#define MVBA_VPOSE 22

struct Item {
  int kind, nr, cid, sensor, pid;
  double r[3];    // raw residual
  double A[9];    // nr x 3: d r / d Xc (CAM) or w R^T (depth kinds)
  double E[9];    // nr x 3: d r / d xyz (zero when the block has no landmark)
  double Z[3];    // R X (CAM) or Mc - t (depth)
  double Xc[3];   // camera-frame point the pose derivatives act on
  double Jf[3];   // CAM: dr/dfocal ; DEPTH: L m (unscaled depth point, for the scale column)
  double Jc;      // CAM: dr/dcentre = Jc I
  double m[3];    // DEPTH: raw depth measurement (affine depth_to_image columns)
  double loss_a;  // Cauchy scale of this block
  double weight;
};

// R (9) and t (3) of one camera image are the first 12 doubles of its 560-byte CamTab row: six 128-bit loads instead of
// twelve 64-bit ones (these gathers are divergent across a warp, so every request costs a sector).
__device__ __forceinline__ void load_rt(const CamTab* __restrict__ tab, int cid, double* rt) {
  const double2* q = reinterpret_cast<const double2*>(tab + cid);
#pragma unroll
  for (int k = 0; k < 6; k++) { const double2 v2 = __ldg(q + k); rt[2 * k] = v2.x; rt[2 * k + 1] = v2.y; }
}

__device__ __forceinline__ void load_camtab(const CamTab* __restrict__ tab, int cid, const double* X, double* Xc, double* Z, double* rt) {
  load_rt(tab, cid, rt);
#pragma unroll
  for (int i = 0; i < 3; i++) {
    Z[i] = rt[3 * i] * X[0] + rt[3 * i + 1] * X[1] + rt[3 * i + 2] * X[2];
    Xc[i] = Z[i] + rt[9 + i];
  }
}

// KEEP_R: R stays in registers across the projection (nine more live values) instead of being loaded a second time for E = A R:
// for kernels that are bound by L1 requests (the back-substitution: five divergent 128-bit loads less per observation).
template <bool WANT_J, bool WANT_E = true, bool KEEP_R = false>
__device__ inline void eval_cam_rec(const View& v, const double* __restrict__ x, const CamTab* __restrict__ tab, int cid, int pid, int sensor, float2 pix,
                                    Item& it, double* Jd0, double* Jd1) {
  it.kind = MVBA_BLOCK_CAM; it.nr = 2;
  it.cid = cid; it.pid = pid; it.sensor = sensor;
  const double* X = x + v.x_xyz + 3 * (int64_t)it.pid;
  double Xw[3] = {X[0], X[1], X[2]};
  double rt[12];
  load_camtab(tab, it.cid, Xw, it.Xc, it.Z, rt);
  project_distort<WANT_J>(v.sens[it.sensor], x, it.Xc, pix, it.r, it.A, it.Jf, &it.Jc, Jd0, Jd1, v.fov_tab + 2 * it.sensor);
  it.r[2] = 0;
  if (WANT_J && WANT_E) {
    if (!KEEP_R) {  // R again (L1 hit) rather than 9 registers held across the projection
      const double2* q = reinterpret_cast<const double2*>(tab + it.cid);
#pragma unroll
      for (int k = 0; k < 5; k++) { const double2 v2 = __ldg(q + k); rt[2 * k] = v2.x; rt[2 * k + 1] = v2.y; }
    }
#pragma unroll
    for (int a = 0; a < 2; a++)
#pragma unroll
      for (int b = 0; b < 3; b++)
        it.E[3 * a + b] = it.A[3 * a] * rt[b] + it.A[3 * a + 1] * rt[3 + b] + it.A[3 * a + 2] * rt[6 + b];
  }
  it.loss_a = v.robust_th; it.weight = 1.0;
}

template <bool WANT_J, bool WANT_E = true, bool KEEP_R = false>
__device__ inline void eval_cam(const View& v, const double* __restrict__ x, const CamTab* __restrict__ tab, int64_t i,
                                Item& it, double* Jd0, double* Jd1) {
  const int cid = v.cid[i];
  eval_cam_rec<WANT_J, WANT_E, KEEP_R>(v, x, tab, cid, v.pid[i], v.cam_sensor[cid], v.pix[i], it, Jd0, Jd1);
}

// The same block from the packed observation record (image, sensor, pixel in one 128-bit load) and a landmark the caller knows:
// for thread-per-track kernels, whose per-observation loads are all divergent.
template <bool WANT_J, bool WANT_E = true, bool KEEP_R = false>
__device__ inline void eval_cam_packed(const View& v, const double* __restrict__ x, const CamTab* __restrict__ tab, int64_t i, int pid,
                                       Item& it, double* Jd0, double* Jd1) {
  const int4 rc = __ldg(v.orec + i);
  eval_cam_rec<WANT_J, WANT_E, KEEP_R>(v, x, tab, rc.x, pid, rc.y, make_float2(__int_as_float(rc.z), __int_as_float(rc.w)), it, Jd0, Jd1);
}

// kind = DEPTH_TRI or DEPTH_MESH; j = index inside that kind.
__device__ inline void eval_depth(const View& v, const double* __restrict__ x, const CamTab* __restrict__ tab,
                                  const DepthTab* __restrict__ dtab, int kind, int64_t j, Item& it) {
  const int64_t i = (kind == MVBA_BLOCK_DEPTH_TRI) ? v.dtri_obs[j] : v.dmesh_obs[j];
  const double w = (kind == MVBA_BLOCK_DEPTH_TRI) ? v.w_dtri : v.w_dmesh;
  it.kind = kind; it.nr = 3;
  it.cid = v.cid[i]; it.pid = v.pid[i]; it.sensor = v.cam_sensor[it.cid];
  const DepthTab& d = dtab[it.sensor];
  const double m[3] = {(double)v.depth_xyz[3 * i], (double)v.depth_xyz[3 * i + 1], (double)v.depth_xyz[3 * i + 2]};
  double Mc[3];
  it.m[0] = m[0]; it.m[1] = m[1]; it.m[2] = m[2];
#pragma unroll
  for (int a = 0; a < 3; a++) {
    it.Jf[a] = d.L[3 * a] * m[0] + d.L[3 * a + 1] * m[1] + d.L[3 * a + 2] * m[2];
    Mc[a] = d.s * it.Jf[a] + d.td[a];
  }
  const double* R = tab[it.cid].R;
  const double* t = tab[it.cid].t;
#pragma unroll
  for (int a = 0; a < 3; a++) { it.Xc[a] = Mc[a]; it.Z[a] = Mc[a] - __ldg(t + a); }
  double Mw[3];
#pragma unroll
  for (int a = 0; a < 3; a++) Mw[a] = __ldg(R + a) * it.Z[0] + __ldg(R + 3 + a) * it.Z[1] + __ldg(R + 6 + a) * it.Z[2];  // R^T
  double ref[3];
  if (kind == MVBA_BLOCK_DEPTH_TRI) {
    const double* X = x + v.x_xyz + 3 * (int64_t)it.pid;
    ref[0] = X[0]; ref[1] = X[1]; ref[2] = X[2];
  } else {
    ref[0] = v.mesh_xyz[3 * i]; ref[1] = v.mesh_xyz[3 * i + 1]; ref[2] = v.mesh_xyz[3 * i + 2];
  }
#pragma unroll
  for (int a = 0; a < 3; a++) {
    it.r[a] = w * (ref[a] - Mw[a]);
#pragma unroll
    for (int b = 0; b < 3; b++) {
      it.A[3 * a + b] = w * __ldg(R + 3 * b + a);
      it.E[3 * a + b] = (kind == MVBA_BLOCK_DEPTH_TRI && a == b) ? w : 0.0;
    }
  }
  it.Jc = 0; it.loss_a = v.robust_th; it.weight = w;
}

__device__ inline void eval_prior(const View& v, const double* __restrict__ x, int64_t j, Item& it) {
  it.kind = MVBA_BLOCK_TRI; it.nr = 3; it.cid = -1; it.sensor = -1; it.pid = v.prior_pid[j];
  const double w = v.prior_w[j];
  const double* X = x + v.x_xyz + 3 * (int64_t)it.pid;
#pragma unroll
  for (int a = 0; a < 3; a++) {
    it.r[a] = w * (X[a] - v.prior_ref[3 * j + a]);
#pragma unroll
    for (int b = 0; b < 3; b++) { it.E[3 * a + b] = (a == b) ? w : 0.0; it.A[3 * a + b] = 0; }
    it.Z[a] = 0; it.Xc[a] = 0; it.Jf[a] = 0;
  }
  it.Jc = 0; it.loss_a = v.prior_th[j]; it.weight = w;
}

// d Xc / d (pose-part virtual column vc), 0 <= vc < 22.
__device__ __forceinline__ void dXc_col(const CamTab* __restrict__ ctp, const double* Z, const double* Xc, int vc, double* d) {
  const CamTab& ct = *ctp;
  if (vc < 3) {  // t0
    double s = 1.0 - __ldg(&ct.alpha);
    d[0] = s * __ldg(ct.Rrc + vc); d[1] = s * __ldg(ct.Rrc + 3 + vc); d[2] = s * __ldg(ct.Rrc + 6 + vc);
  } else if (vc < 7) {  // q0
    int j = vc - 3;
    double w[3] = {__ldg(ct.W0 + j), __ldg(ct.W0 + 4 + j), __ldg(ct.W0 + 8 + j)};
    cross3(w, Z, d);
  } else if (vc < 10) {  // t1
    int j = vc - 7;
    double s = __ldg(&ct.alpha);
    d[0] = s * __ldg(ct.Rrc + j); d[1] = s * __ldg(ct.Rrc + 3 + j); d[2] = s * __ldg(ct.Rrc + 6 + j);
  } else if (vc < 14) {  // q1
    int j = vc - 10;
    double w[3] = {__ldg(ct.W1 + j), __ldg(ct.W1 + 4 + j), __ldg(ct.W1 + 8 + j)};
    cross3(w, Z, d);
  } else if (vc < 17) {  // t_rc
    int j = vc - 14;
    double s = __ldg(&ct.rig_on);
    d[0] = j == 0 ? s : 0.0; d[1] = j == 1 ? s : 0.0; d[2] = j == 2 ? s : 0.0;
  } else if (vc < 21) {  // q_rc
    int j = vc - 17;
    double w[3] = {__ldg(ct.Wrc + j), __ldg(ct.Wrc + 4 + j), __ldg(ct.Wrc + 8 + j)};
    double y[3] = {Xc[0] - __ldg(ct.trc), Xc[1] - __ldg(ct.trc + 1), Xc[2] - __ldg(ct.trc + 2)};
    cross3(w, y, d);
  } else {  // offset
    double w[3] = {__ldg(ct.wa), __ldg(ct.wa + 1), __ldg(ct.wa + 2)};
    cross3(w, Z, d);
    double s = __ldg(&ct.dalpha);
    d[0] = s * (d[0] + __ldg(ct.dt)); d[1] = s * (d[1] + __ldg(ct.dt + 1)); d[2] = s * (d[2] + __ldg(ct.dt + 2));
  }
}

// Number of virtual f-columns of an item.
__device__ __forceinline__ int item_vwidth(const View& v, const Item& it) {
  if (it.kind == MVBA_BLOCK_CAM) return MVBA_VPOSE + 3 + v.sens[it.sensor].nd_nz;
  if (it.kind == MVBA_BLOCK_DEPTH_TRI || it.kind == MVBA_BLOCK_DEPTH_MESH) return MVBA_VPOSE + v.nd2i + 1;
  return 0;
}

// Reduced-system column of virtual column vc (-1: the parameter is constant).
__device__ __forceinline__ int item_gcol(const View& v, const Item& it, int vc) {
  const SensorDev& s = v.sens[it.sensor];
  if (vc < 7) { int b = v.pose_col[v.cam_p0[it.cid]]; return b < 0 ? -1 : b + vc; }
  if (vc < 14) { int p1 = v.cam_p1[it.cid]; if (p1 < 0) return -1; int b = v.pose_col[p1]; return b < 0 ? -1 : b + vc - 7; }
  if (vc < 21) return s.c_rig < 0 ? -1 : s.c_rig + vc - 14;
  if (vc == 21) return s.c_off;
  int k = vc - MVBA_VPOSE;
  if (it.kind == MVBA_BLOCK_CAM) {
    if (k == 0) return s.c_focal;
    if (k < 3) return s.c_center < 0 ? -1 : s.c_center + k - 1;
    return s.c_dist < 0 ? -1 : s.c_dist + k - 3;
  }
  if (k < v.nd2i) return s.c_d2i < 0 ? -1 : s.c_d2i + k;
  return s.c_scale;
}

// Column vc of the (un-corrected) Jacobian: out[0..nr).  Jd0/Jd1: CAM distortion rows.
__device__ __forceinline__ void item_jcol(const View& v, const Item& it, const CamTab* __restrict__ tab,
                                          const DepthTab* __restrict__ dtab, int vc, const double* Jd0,
                                          const double* Jd1, double* out) {
  if (vc < MVBA_VPOSE) {
    double d[3];
    dXc_col(tab + it.cid, it.Z, it.Xc, vc, d);
    for (int a = 0; a < it.nr; a++) out[a] = it.A[3 * a] * d[0] + it.A[3 * a + 1] * d[1] + it.A[3 * a + 2] * d[2];
    return;
  }
  int k = vc - MVBA_VPOSE;
  if (it.kind == MVBA_BLOCK_CAM) {
    if (k == 0) { out[0] = it.Jf[0]; out[1] = it.Jf[1]; }
    else if (k == 1) { out[0] = it.Jc; out[1] = 0; }
    else if (k == 2) { out[0] = 0; out[1] = it.Jc; }
    else { out[0] = Jd0[k - 3]; out[1] = Jd1[k - 3]; }
    return;
  }
  // depth kinds: dr = -A dMc
  const DepthTab& dt = dtab[it.sensor];
  double d[3];
  if (k < v.nd2i) {
    if (v.nd2i == 12) {  // affine: Mc_row = s sum_j A[row][j] m_j + A[row][3]
      int row = k >> 2, col = k & 3;
      d[0] = d[1] = d[2] = 0;
      double val = (col < 3) ? dt.s * it.m[col] : 1.0;
      if (row == 0) d[0] = val; else if (row == 1) d[1] = val; else d[2] = val;
    } else if (k < 3) {
      d[0] = k == 0; d[1] = k == 1; d[2] = k == 2;
    } else {
      int j = k - 3;
      double w[3] = {dt.Wd[j], dt.Wd[4 + j], dt.Wd[8 + j]};
      cross3(w, it.Jf, d);
      d[0] *= dt.s; d[1] *= dt.s; d[2] *= dt.s;
    }
  } else {
    d[0] = it.Jf[0]; d[1] = it.Jf[1]; d[2] = it.Jf[2];
  }
  for (int a = 0; a < 3; a++) out[a] = -(it.A[3 * a] * d[0] + it.A[3 * a + 1] * d[1] + it.A[3 * a + 2] * d[2]);
}

}  // namespace mvba
