// kernels_camq.cuh -- pass A of the normal-equation build when no intrinsics / depth-transform column is free
// (every f column of a residual block is a pose, rig-transform or time-offset column): F'F, F'r, diag(F'F).
//
// Every such Jacobian column of a block seen by image c is  A d_c  with  d_c = u_c + w_c x Z :  two 3-vectors per
// (image, column) -- translations u = coef, rotations w = coef, rig rotation w = coef and u = coef x (t - t_rc)
// (because Xc - t_rc = Z + (t - t_rc)), time offset w = dalpha wa, u = dalpha dt (CamCoef rows, common.cuh).  Since
// A_r . (w x Z) = w . (Z x A_r), one Jacobian row of the block is
//     J_r = b_r' T_c,   b_r = sqrt(rho') [A_r ; Z x A_r]  (6 numbers per row),  T_c = [u ; w]  (6 x 22, per image)
// and the image's whole contribution is  F'F = T_c' (sum_r b_r b_r') T_c,  F'r = T_c' (sum_r b_r r_r).
//   camq_kernel     one thread per residual block, camera-major: residual + dr/dXc, then 27 FMAs per residual row
//                   into Q = sum b b' (21) and q = sum b r (6); one reduction per chunk of an image.  The work per
//                   observation no longer depends on how many columns the image has (was: 2 x 23 x 24 / 2).
//   camfold_kernel  one warp per image: T' Q T (at most 22 x 22) -> reduced system, same destinations as the staged-row
//                   kernel (kernels_camera.cuh), which remains the path for images with free intrinsics.
// Ceres 1.14 equivalent: the F'F + F'r part of SchurEliminator / ImplicitSchurComplement set-up.
#pragma once
#include "kernels_camera.cuh"

namespace mvba {

#define MVBA_CAMQ_N 28  // 21 + 6 accumulators per image, padded

struct CamQArgs {
  const double* x;
  const CamTab* tab;
  const DepthTab* dtab;
  const int32_t* chunk_cam;    // per chunk: camera image
  const int32_t* chunk_beg;    // nchunks + 1: range into citems
  const int32_t* citems;       // block codes, camera-major: [0,N) CAM, then DEPTH_TRI, then DEPTH_MESH
  const TSlot* crec;           // the same list as packed records: CAM blocks carry landmark and pixel (a = pid, b = kind << 24, px, py),
                               // the depth kinds their index inside the kind (a)
  int nchunks;
  double* Q;                   // C x MVBA_CAMQ_N, zeroed before the launch
};

__global__ void __launch_bounds__(256, 2) camq_kernel(View v, CamQArgs a) {
  __shared__ double red[8][MVBA_CAMQ_N];
  const int tid = threadIdx.x, lane = tid & 31, wib = tid >> 5;
  for (int ch = blockIdx.x; ch < a.nchunks; ch += gridDim.x) {
    const int cam = a.chunk_cam[ch];
    const int i0 = a.chunk_beg[ch], i1 = a.chunk_beg[ch + 1];
    double q[27];
#pragma unroll
    for (int k = 0; k < 27; k++) q[k] = 0;
    // the chunk's image: rotation, translation and sensor are the same for every block (broadcast loads, then registers)
    double Rm[9], tv[3];
    {
      const CamTab& ct = a.tab[cam];
#pragma unroll
      for (int k = 0; k < 9; k++) Rm[k] = __ldg(ct.R + k);
#pragma unroll
      for (int k = 0; k < 3; k++) tv[k] = __ldg(ct.t + k);
    }
    const int sensor = v.cam_sensor[cam];
    for (int i = i0 + tid; i < i1; i += 256) {
      const int4 raw = __ldg(reinterpret_cast<const int4*>(a.crec) + i);
      TSlot sl; sl.a = raw.x; sl.b = (uint32_t)raw.y; sl.px = __int_as_float(raw.z); sl.py = __int_as_float(raw.w);  // one coalesced 16-byte record per block: no index chain in front of the landmark gather
      const int kind = (sl.b >> 24) & 7;
      Item it;
      if (kind == 0) {
        const double* X = a.x + v.x_xyz + 3 * (int64_t)sl.a;
        const double X0 = X[0], X1 = X[1], X2 = X[2];
#pragma unroll
        for (int k = 0; k < 3; k++) {
          it.Z[k] = Rm[3 * k] * X0 + Rm[3 * k + 1] * X1 + Rm[3 * k + 2] * X2;
          it.Xc[k] = it.Z[k] + tv[k];
        }
        it.nr = 2; it.loss_a = v.robust_th;
        project_distort<true>(v.sens[sensor], a.x, it.Xc, make_float2(sl.px, sl.py), it.r, it.A, it.Jf, &it.Jc, nullptr, nullptr, v.fov_tab + 2 * sensor);
        it.r[2] = 0;
      } else {
        eval_depth(v, a.x, a.tab, a.dtab, kind == 1 ? MVBA_BLOCK_DEPTH_TRI : MVBA_BLOCK_DEPTH_MESH, sl.a, it);
      }
      double s = it.r[0] * it.r[0] + it.r[1] * it.r[1];
      if (it.nr == 3) s += it.r[2] * it.r[2];
      double hr, sr;
      cauchy(it.loss_a, s, &hr, &sr);
#pragma unroll
      for (int r = 0; r < 3; r++) {
        if (r < it.nr) {
          const double a0 = sr * it.A[3 * r], a1 = sr * it.A[3 * r + 1], a2 = sr * it.A[3 * r + 2];
          const double b[6] = {a0, a1, a2, it.Z[1] * a2 - it.Z[2] * a1, it.Z[2] * a0 - it.Z[0] * a2, it.Z[0] * a1 - it.Z[1] * a0};
          const double rr = sr * it.r[r];
          int k = 0;
#pragma unroll
          for (int i2 = 0; i2 < 6; i2++)
#pragma unroll
            for (int j2 = i2; j2 < 6; j2++) { q[k] = fma(b[i2], b[j2], q[k]); k++; }
#pragma unroll
          for (int i2 = 0; i2 < 6; i2++) q[21 + i2] = fma(b[i2], rr, q[21 + i2]);
        }
      }
    }
#pragma unroll
    for (int k = 0; k < 27; k++) {
      const double s = warp_sum(q[k]);
      if (lane == 0) red[wib][k] = s;
    }
    __syncthreads();
    if (tid < 27) {
      double s = 0;
#pragma unroll
      for (int w = 0; w < 8; w++) s += red[w][tid];
      if (s != 0.0) atomicAdd(a.Q + (size_t)cam * MVBA_CAMQ_N + tid, s);
    }
    __syncthreads();
  }
}

// Cost of a candidate point over the same camera-major records: the lanes of a warp share the image, so its rotation and
// translation are broadcast loads (the item-ordered residual_kernel gathers them per observation and is bound by L1 requests).
// Covers every CAM and depth block; blocks without an image (landmark priors) stay with residual_kernel.
__global__ void __launch_bounds__(256) camcost_kernel(View v, CamQArgs a, double* __restrict__ cost) {
  const int tid = threadIdx.x;
  double acc = 0;
  for (int ch = blockIdx.x; ch < a.nchunks; ch += gridDim.x) {
    const int cam = a.chunk_cam[ch];
    const int i0 = a.chunk_beg[ch], i1 = a.chunk_beg[ch + 1];
    double Rm[9], tv[3];
    {
      const CamTab& ct = a.tab[cam];
#pragma unroll
      for (int k = 0; k < 9; k++) Rm[k] = __ldg(ct.R + k);
#pragma unroll
      for (int k = 0; k < 3; k++) tv[k] = __ldg(ct.t + k);
    }
    const int sensor = v.cam_sensor[cam];
    for (int i = i0 + tid; i < i1; i += 256) {
      const int4 raw = __ldg(reinterpret_cast<const int4*>(a.crec) + i);
      const int kind = ((uint32_t)raw.y >> 24) & 7;
      double r[3], s;
      double loss_a = v.robust_th;
      if (kind == 0) {
        const double* X = a.x + v.x_xyz + 3 * (int64_t)raw.x;
        const double X0 = X[0], X1 = X[1], X2 = X[2];
        double Xc[3];
#pragma unroll
        for (int k = 0; k < 3; k++) Xc[k] = Rm[3 * k] * X0 + Rm[3 * k + 1] * X1 + Rm[3 * k + 2] * X2 + tv[k];
        double A[6], Jf[2], Jc;
        project_distort<false>(v.sens[sensor], a.x, Xc, make_float2(__int_as_float(raw.z), __int_as_float(raw.w)), r, A, Jf, &Jc, nullptr, nullptr, v.fov_tab + 2 * sensor);
        s = r[0] * r[0] + r[1] * r[1];
      } else {
        Item it;
        eval_depth(v, a.x, a.tab, a.dtab, kind == 1 ? MVBA_BLOCK_DEPTH_TRI : MVBA_BLOCK_DEPTH_MESH, raw.x, it);
        s = it.r[0] * it.r[0] + it.r[1] * it.r[1] + it.r[2] * it.r[2];
        loss_a = it.loss_a;
      }
      double hr, sr;
      cauchy(loss_a, s, &hr, &sr);
      acc += hr;
    }
  }
  block_accumulate(acc, cost);
}

struct CamFoldArgs {
  const CamTab* tab;
  const CamCoef* coef;
  const double* Q;
  const double* jscale_f;  // Jacobi column scaling (null: 1)
  int mode;                // 0 full, 1 squared column norms only
};

// index of (i <= j) in the packed upper triangle of a 6 x 6 matrix, row-major
__device__ __forceinline__ int tri6(int i, int j) { return i * 6 - i * (i - 1) / 2 + (j - i); }

__global__ void __launch_bounds__(128) camfold_kernel(View v, CamFoldArgs a, Reduced S) {
  __shared__ double sQ[4][36 + 6];
  __shared__ double sT[4][MVBA_VPOSE][6];   // scaled [u ; w] of every active column
  __shared__ double sQT[4][MVBA_VPOSE][6];  // Q T
  __shared__ int sgc[4][MVBA_VPOSE];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int cam = blockIdx.x * 4 + wib;
  if (cam >= v.C) return;  // warps are independent below (__syncwarp only)
  const double* Qg = a.Q + (size_t)cam * MVBA_CAMQ_N;
  double* Qs = sQ[wib];
  if (lane < 27) {
    const double val = Qg[lane];
    if (lane >= 21) Qs[36 + lane - 21] = val;
    else {
      int i = 0, k = lane;
      while (k >= 6 - i) { k -= 6 - i; i++; }
      const int j = i + k;
      Qs[6 * i + j] = val; Qs[6 * j + i] = val;
    }
  }
  // active columns in ascending virtual-column order (the order of the staged-row kernel)
  Item it; it.cid = cam; it.sensor = v.cam_sensor[cam]; it.kind = MVBA_BLOCK_CAM;
  const int gc = lane < MVBA_VPOSE ? item_gcol(v, it, lane) : -1;
  const unsigned act = __ballot_sync(0xffffffffu, gc >= 0);
  const int nact = __popc(act);
  if (nact == 0) return;
  if (gc >= 0) {
    const int pos = __popc(act & ((1u << lane) - 1));
    const int vc = lane;
    const double* cf = a.coef[cam].c[vc];
    double u[3] = {0, 0, 0}, w[3] = {0, 0, 0};
    const bool rot = (vc >= 3 && vc < 7) || (vc >= 10 && vc < 14) || vc >= 17;
    if (!rot) { u[0] = cf[0]; u[1] = cf[1]; u[2] = cf[2]; }
    else {
      w[0] = cf[0]; w[1] = cf[1]; w[2] = cf[2];
      if (vc == 21) { u[0] = cf[3]; u[1] = cf[4]; u[2] = cf[5]; }
      else if (vc >= 17) {  // rig rotation acts on Xc - t_rc = Z + (t - t_rc)
        const CamTab& ct = a.tab[cam];
        const double d[3] = {ct.t[0] - ct.trc[0], ct.t[1] - ct.trc[1], ct.t[2] - ct.trc[2]};
        cross3(w, d, u);
      }
    }
    const double sc = a.jscale_f ? a.jscale_f[gc] : 1.0;
    double* T = sT[wib][pos];
    T[0] = sc * u[0]; T[1] = sc * u[1]; T[2] = sc * u[2]; T[3] = sc * w[0]; T[4] = sc * w[1]; T[5] = sc * w[2];
    sgc[wib][pos] = gc;
  }
  __syncwarp();
  if (lane < nact) {
    const double* T = sT[wib][lane];
    double* QT = sQT[wib][lane];
    double gr = 0;
#pragma unroll
    for (int i = 0; i < 6; i++) {
      double s = 0;
#pragma unroll
      for (int j = 0; j < 6; j++) s = fma(Qs[6 * i + j], T[j], s);
      QT[i] = s;
      gr = fma(T[i], Qs[36 + i], gr);
    }
    const int gi = sgc[wib][lane];
    if (a.mode == 0 && gr != 0.0) { atomicAdd(S.g + gi, gr); atomicAdd(S.rhs + gi, gr); }
  }
  __syncwarp();
  const int npair = nact * (nact + 1) / 2;
  for (int idx = lane; idx < npair; idx += 32) {
    int i, j;
    tri_decode(idx, i, j);  // i <= j
    const double* Ti = sT[wib][i];
    const double* QTj = sQT[wib][j];
    double val = 0;
#pragma unroll
    for (int k = 0; k < 6; k++) val = fma(Ti[k], QTj[k], val);
    if (val == 0.0) continue;
    const int gi = sgc[wib][i], gj = sgc[wib][j];
    if (i == j) atomicAdd(S.diagF + gi, val);
    if (a.mode == 0) {
      addS(S, gi, gj, val);
      if (S.Fpp) addFdiag(S, gi, gj, val);
    }
  }
}

}  // namespace mvba
