// kernels_backsub.cuh -- landmark back-substitution and model cost change, one THREAD per track.
//
// Given the reduced-system solution y_f the landmark part is  y_e = Hinv (E'r - E'F y_f)  and the model
// cost change is  sum_rows jy (r - jy / 2),  jy = F y_f + E y_e  (step = -y; Ceres 1.14
// TrustRegionMinimizer).  F y_f per residual block is a DIRECTIONAL derivative: all pose / rig / offset
// columns act on the camera-frame point through
//     dXc = a_c + w_c x Z + w_rc x (Xc - t_rc)
// with three 3-vectors per camera image that depend only on y_f (dir_cameras_kernel), so no Jacobian
// column is ever formed here.  With s1 = sum fy.r, s2 = sum fy.fy, u = E'fy, H0 = E'E, g = E'r:
//     model change = s1 + y_e'g - s2/2 - y_e'u - y_e'H0 y_e/2.
#pragma once
#include "kernels_build.cuh"
#include "kernels_eval.cuh"

namespace mvba {

struct alignas(16) CamDir { double a[3], w[3], wrc[3], trc[3]; };  // 96 bytes: gathered per residual block as six 128-bit loads
static_assert(sizeof(CamDir) == 96, "CamDir layout");
struct SensDir { double td[3], wd[3], ds, dA[12]; };

__device__ __forceinline__ double dir_at(const double* __restrict__ yf, const double* __restrict__ sf, int gc) {
  return gc < 0 ? 0.0 : yf[gc] * (sf ? sf[gc] : 1.0);
}

// One thread per camera image (+ one per sensor): fold y_f into the direction tables.
__global__ void dir_cameras_kernel(View v, const CamTab* __restrict__ tab, const DepthTab* __restrict__ dtab,
                                   const double* __restrict__ yf, const double* __restrict__ sf, CamDir* __restrict__ cd,
                                   SensDir* __restrict__ sdir) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c < v.C) {
    const CamTab& ct = tab[c];
    const SensorDev& s = v.sens[v.cam_sensor[c]];
    const int p0 = v.cam_p0[c], p1 = v.cam_p1[c];
    const int c0 = v.pose_col[p0], c1 = p1 >= 0 ? v.pose_col[p1] : -1;
    double d0[7], d1[7], dr[7];
#pragma unroll
    for (int k = 0; k < 7; k++) {
      d0[k] = dir_at(yf, sf, c0 < 0 ? -1 : c0 + k);
      d1[k] = dir_at(yf, sf, c1 < 0 ? -1 : c1 + k);
      dr[k] = dir_at(yf, sf, s.c_rig < 0 ? -1 : s.c_rig + k);
    }
    const double doff = dir_at(yf, sf, s.c_off) * ct.dalpha;
    CamDir o;
#pragma unroll
    for (int i = 0; i < 3; i++) {
      o.a[i] = ct.Rrc[3 * i] * ((1.0 - ct.alpha) * d0[0] + ct.alpha * d1[0]) + ct.Rrc[3 * i + 1] * ((1.0 - ct.alpha) * d0[1] + ct.alpha * d1[1]) +
               ct.Rrc[3 * i + 2] * ((1.0 - ct.alpha) * d0[2] + ct.alpha * d1[2]) + ct.rig_on * dr[i] + doff * ct.dt[i];
      o.w[i] = ct.W0[4 * i] * d0[3] + ct.W0[4 * i + 1] * d0[4] + ct.W0[4 * i + 2] * d0[5] + ct.W0[4 * i + 3] * d0[6] +
               ct.W1[4 * i] * d1[3] + ct.W1[4 * i + 1] * d1[4] + ct.W1[4 * i + 2] * d1[5] + ct.W1[4 * i + 3] * d1[6] + doff * ct.wa[i];
      o.wrc[i] = ct.Wrc[4 * i] * dr[3] + ct.Wrc[4 * i + 1] * dr[4] + ct.Wrc[4 * i + 2] * dr[5] + ct.Wrc[4 * i + 3] * dr[6];
      o.trc[i] = ct.trc[i];
    }
    cd[c] = o;
  } else if (c < v.C + v.T) {
    const int t = c - v.C;
    const SensorDev& s = v.sens[t];
    const DepthTab& d = dtab[t];
    SensDir o;
    for (int k = 0; k < 12; k++) o.dA[k] = 0;
    for (int k = 0; k < 3; k++) { o.td[k] = 0; o.wd[k] = 0; }
    if (v.nd2i == 12) {
      for (int k = 0; k < 12; k++) o.dA[k] = dir_at(yf, sf, s.c_d2i < 0 ? -1 : s.c_d2i + k);
    } else {
      double dq[4];
      for (int k = 0; k < 3; k++) o.td[k] = dir_at(yf, sf, s.c_d2i < 0 ? -1 : s.c_d2i + k);
      for (int k = 0; k < 4; k++) dq[k] = dir_at(yf, sf, s.c_d2i < 0 ? -1 : s.c_d2i + 3 + k);
      for (int i = 0; i < 3; i++) o.wd[i] = d.Wd[4 * i] * dq[0] + d.Wd[4 * i + 1] * dq[1] + d.Wd[4 * i + 2] * dq[2] + d.Wd[4 * i + 3] * dq[3];
    }
    o.ds = dir_at(yf, sf, s.c_scale);
    sdir[t] = o;
  }
}

struct BacksubArgs {
  const double* x; const CamTab* tab; const DepthTab* dtab;
  const CamDir* cd; const SensDir* sdir;
  const double *jscale_f, *jscale_e;
  const double *Hinv, *ge, *yf;
  double* ye;      // 3 per landmark (scaled variables)
  double* lm;      // LM_MCC accumulates
  int lm_mcc;      // slot index
};

#define MVBA_MAXDZ 64  // nd_nz <= MVBA_MAX_DISTORTION / 2

// F y_f of one residual block (rows scaled by sqrt(rho')) from the direction tables.
template <bool DZ>
__device__ __forceinline__ void item_dir(const View& v, const BacksubArgs& a, const Item& it, double sr, const double* jd0,
                                         const double* jd1, double* fy) {
  double dd[12];  // a | w | wrc | trc of the block's camera image
  {
    const double2* q = reinterpret_cast<const double2*>(a.cd + it.cid);
#pragma unroll
    for (int k = 0; k < 6; k++) { const double2 v2 = __ldg(q + k); dd[2 * k] = v2.x; dd[2 * k + 1] = v2.y; }
  }
  const double y[3] = {it.Xc[0] - dd[9], it.Xc[1] - dd[10], it.Xc[2] - dd[11]};
  double dx[3], c1[3], c2[3];
  const double w[3] = {dd[3], dd[4], dd[5]}, wr[3] = {dd[6], dd[7], dd[8]};
  cross3(w, it.Z, c1);
  cross3(wr, y, c2);
#pragma unroll
  for (int i = 0; i < 3; i++) dx[i] = dd[i] + c1[i] + c2[i];
  const SensorDev& s = v.sens[it.sensor];
  if (it.kind == MVBA_BLOCK_CAM) {
    const double df = dir_at(a.yf, a.jscale_f, s.c_focal);
    const double dcx = dir_at(a.yf, a.jscale_f, s.c_center), dcy = dir_at(a.yf, a.jscale_f, s.c_center < 0 ? -1 : s.c_center + 1);
    fy[0] = it.A[0] * dx[0] + it.A[1] * dx[1] + it.A[2] * dx[2] + it.Jf[0] * df + it.Jc * dcx;
    fy[1] = it.A[3] * dx[0] + it.A[4] * dx[1] + it.A[5] * dx[2] + it.Jf[1] * df + it.Jc * dcy;
    fy[2] = 0;
    if (DZ && s.c_dist >= 0) {
      for (int k = 0; k < s.nd_nz; k++) {
        const double dk = dir_at(a.yf, a.jscale_f, s.c_dist + k);
        fy[0] += jd0[k] * dk; fy[1] += jd1[k] * dk;
      }
    }
    fy[0] *= sr; fy[1] *= sr;
  } else {
    const SensDir& sd = a.sdir[it.sensor];
    const DepthTab& dt = a.dtab[it.sensor];
    double dm[3], c3[3];
    const double wd[3] = {sd.wd[0], sd.wd[1], sd.wd[2]};
    cross3(wd, it.Jf, c3);
#pragma unroll
    for (int i = 0; i < 3; i++) {
      dm[i] = sd.td[i] + dt.s * c3[i] + sd.ds * it.Jf[i] +
              dt.s * (sd.dA[4 * i] * it.m[0] + sd.dA[4 * i + 1] * it.m[1] + sd.dA[4 * i + 2] * it.m[2]) + sd.dA[4 * i + 3];
      dx[i] -= dm[i];
    }
#pragma unroll
    for (int i = 0; i < 3; i++) fy[i] = sr * (it.A[3 * i] * dx[0] + it.A[3 * i + 1] * dx[1] + it.A[3 * i + 2] * dx[2]);
  }
}

// DZ: some sensor's distortion coefficients float (their Jacobian columns are needed, dynamically indexed -> local memory);
// without it the whole residual block stays in registers.
template <bool DZ, bool KEEP_R = false>
__global__ void __launch_bounds__(128, 4) backsub_kernel(View v, BacksubArgs a) {
  double mcc = 0;
  double jd[DZ ? 2 * MVBA_MAXDZ : 2];
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < v.P; p += gridDim.x * blockDim.x) {
    const int64_t r0[5] = {v.trk_obs0[p], v.trk_dtri0[p], v.trk_dmesh0[p], v.trk_pm0[p], v.trk_pt0[p]};
    const int64_t r1[5] = {v.trk_obs0[p + 1], v.trk_dtri0[p + 1], v.trk_dmesh0[p + 1], v.trk_pm0[p + 1], v.trk_pt0[p + 1]};
    if (r1[0] - r0[0] + r1[1] - r0[1] + r1[2] - r0[2] + r1[3] - r0[3] + r1[4] - r0[4] == 0) continue;
    double se[3] = {1, 1, 1};
    if (a.jscale_e) { se[0] = a.jscale_e[3 * (size_t)p]; se[1] = a.jscale_e[3 * (size_t)p + 1]; se[2] = a.jscale_e[3 * (size_t)p + 2]; }
    double u[3] = {0, 0, 0}, H0[6] = {0, 0, 0, 0, 0, 0}, s1 = 0, s2 = 0;
#pragma unroll
    for (int kind = 0; kind < 5; kind++) {
      for (int64_t j = r0[kind]; j < r1[kind]; j++) {
        Item it;
        if (kind == 0) eval_cam_packed<true, true, KEEP_R>(v, a.x, a.tab, j, p, it, DZ ? jd : nullptr, DZ ? jd + MVBA_MAXDZ : nullptr);
        else if (kind == 1) eval_depth(v, a.x, a.tab, a.dtab, MVBA_BLOCK_DEPTH_TRI, j, it);
        else if (kind == 2) eval_depth(v, a.x, a.tab, a.dtab, MVBA_BLOCK_DEPTH_MESH, j, it);
        else eval_prior(v, a.x, j, it);
        const int nr = kind == 0 ? 2 : 3;  // == it.nr, known per kind so that the rows are indexed statically
        double s = it.r[0] * it.r[0] + it.r[1] * it.r[1];
        if (nr == 3) s += it.r[2] * it.r[2];
        double hr, sr;
        cauchy(it.loss_a, s, &hr, &sr);
        double fy[3] = {0, 0, 0};
        if (kind < 3) item_dir<DZ>(v, a, it, sr, jd, jd + (DZ ? MVBA_MAXDZ : 0), fy);
#pragma unroll
        for (int r = 0; r < 3; r++) {
          if (r < nr) {
            const double e0 = sr * it.E[3 * r] * se[0], e1 = sr * it.E[3 * r + 1] * se[1], e2 = sr * it.E[3 * r + 2] * se[2];
            H0[0] += e0 * e0; H0[1] += e0 * e1; H0[2] += e0 * e2; H0[3] += e1 * e1; H0[4] += e1 * e2; H0[5] += e2 * e2;
            u[0] += e0 * fy[r]; u[1] += e1 * fy[r]; u[2] += e2 * fy[r];
            s1 += fy[r] * (sr * it.r[r]); s2 += fy[r] * fy[r];
          }
        }
      }
    }
    const double* Hi = a.Hinv + 6 * (size_t)p;
    const double* g = a.ge + 3 * (size_t)p;
    const double w0 = g[0] - u[0], w1 = g[1] - u[1], w2 = g[2] - u[2];
    const double ye[3] = {Hi[0] * w0 + Hi[1] * w1 + Hi[2] * w2, Hi[1] * w0 + Hi[3] * w1 + Hi[4] * w2, Hi[2] * w0 + Hi[4] * w1 + Hi[5] * w2};
    for (int k = 0; k < 3; k++) a.ye[3 * (size_t)p + k] = ye[k];
    const double yHy = ye[0] * (H0[0] * ye[0] + H0[1] * ye[1] + H0[2] * ye[2]) + ye[1] * (H0[1] * ye[0] + H0[3] * ye[1] + H0[4] * ye[2]) +
                       ye[2] * (H0[2] * ye[0] + H0[4] * ye[1] + H0[5] * ye[2]);
    mcc += s1 + (ye[0] * g[0] + ye[1] * g[1] + ye[2] * g[2]) - 0.5 * s2 - (ye[0] * u[0] + ye[1] * u[1] + ye[2] * u[2]) - 0.5 * yHy;
  }
  block_accumulate(mcc, a.lm + a.lm_mcc);
}

// Squared column norms of the landmark part of the (loss-corrected, unscaled) Jacobian: diag(E'E) per landmark, one thread
// per track.  Ceres' Jacobi scaling is computed once per solve from these (TrustRegionMinimizer::Init); F's share comes from pass A.
__global__ void __launch_bounds__(128, 4) colnorm_e_kernel(View v, const double* __restrict__ x, const CamTab* __restrict__ tab,
                                                           const DepthTab* __restrict__ dtab, double* __restrict__ diagE) {
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < v.P; p += gridDim.x * blockDim.x) {
    const int64_t r0[5] = {v.trk_obs0[p], v.trk_dtri0[p], v.trk_dmesh0[p], v.trk_pm0[p], v.trk_pt0[p]};
    const int64_t r1[5] = {v.trk_obs0[p + 1], v.trk_dtri0[p + 1], v.trk_dmesh0[p + 1], v.trk_pm0[p + 1], v.trk_pt0[p + 1]};
    double H[3] = {0, 0, 0};
#pragma unroll
    for (int kind = 0; kind < 5; kind++) {
      for (int64_t j = r0[kind]; j < r1[kind]; j++) {
        Item it;
        if (kind == 0) eval_cam_packed<true>(v, x, tab, j, p, it, nullptr, nullptr);
        else if (kind == 1) eval_depth(v, x, tab, dtab, MVBA_BLOCK_DEPTH_TRI, j, it);
        else if (kind == 2) eval_depth(v, x, tab, dtab, MVBA_BLOCK_DEPTH_MESH, j, it);
        else eval_prior(v, x, j, it);
        const int nr = kind == 0 ? 2 : 3;
        double s = it.r[0] * it.r[0] + it.r[1] * it.r[1];
        if (nr == 3) s += it.r[2] * it.r[2];
        const double sr = cauchy_sr(it.loss_a, s);
#pragma unroll
        for (int r = 0; r < 3; r++) {
          if (r < nr) {
#pragma unroll
            for (int b = 0; b < 3; b++) { const double e = sr * it.E[3 * r + b]; H[b] += e * e; }
          }
        }
      }
    }
    diagE[3 * (size_t)p] = H[0]; diagE[3 * (size_t)p + 1] = H[1]; diagE[3 * (size_t)p + 2] = H[2];
  }
}

}  // namespace mvba
