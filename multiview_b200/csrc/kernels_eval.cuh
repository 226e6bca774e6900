// kernels_eval.cuh -- camera-table preparation, cost/residual evaluation and the stored-Jacobian
// kernel (north-star kernel (a): per-observation FP64 residual + analytic Jacobian).
#pragma once
#include "items.cuh"

namespace mvba {

// One thread per camera image (+ one per sensor for the depth tables).
__global__ void prep_cameras_kernel(View v, const double* __restrict__ x, CamTab* __restrict__ tab,
                                    DepthTab* __restrict__ dtab, int* __restrict__ err, CamCoef* __restrict__ coef,
                                    CamPack* __restrict__ pack) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c < v.C) {
    CamTab o;
    bool ok = prep_camera(v, x, c, o);
    tab[c] = o;
    if (coef) {
      CamCoef& q = coef[c];
      for (int k = 0; k < 3; k++)
        for (int a = 0; a < 3; a++) {
          q.c[k][a] = (1.0 - o.alpha) * o.Rrc[3 * a + k];
          q.c[7 + k][a] = o.alpha * o.Rrc[3 * a + k];
          q.c[14 + k][a] = (a == k) ? o.rig_on : 0.0;
        }
      for (int j = 0; j < 4; j++)
        for (int a = 0; a < 3; a++) { q.c[3 + j][a] = o.W0[4 * a + j]; q.c[10 + j][a] = o.W1[4 * a + j]; q.c[17 + j][a] = o.Wrc[4 * a + j]; }
      for (int a = 0; a < 3; a++) { q.c[21][a] = o.dalpha * o.wa[a]; q.c[22][a] = o.dalpha * o.dt[a]; }
      if (pack) {
        CamPack& k = pack[c];
        for (int i = 0; i < 9; i++) k.R[i] = o.R[i];
        for (int a = 0; a < 3; a++) { k.t[a] = o.t[a]; k.dl[a] = o.t[a] - o.trc[a]; }
        for (int r = 0; r < 23; r++) for (int a = 0; a < 3; a++) k.c[r][a] = q.c[r][a];
        k.sensor = v.cam_sensor[c]; k.pad0 = 0; k.pad1 = 0;
      }
    }
    if (!ok) atomicExch(err, MVBA_EBRACKET);
  } else if (c < v.C + v.T) {
    const int t = c - v.C;
    if (dtab != nullptr) {
      DepthTab d;
      prep_depth(v, x, t, d);
      dtab[t] = d;
    }
    if (v.sens[t].nd == 1) {  // FOV model constants of this evaluation point (camera_params.cc:269-281)
      const double w = x[v.sens[t].x_dist];
      v.fov_tab[2 * t] = 2.0 * tan(0.5 * w);
      v.fov_tab[2 * t + 1] = 1.0 / w;
    }
  }
}

// Block-level sum -> one atomicAdd per CTA.
__device__ __forceinline__ void block_accumulate(double val, double* out) {
  __shared__ double sm[32];
  val = warp_sum(val);
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) sm[w] = val;
  __syncthreads();
  if (w == 0) {
    val = lane < (blockDim.x >> 5) ? sm[lane] : 0.0;
    val = warp_sum(val);
    if (lane == 0) atomicAdd(out, val);
  }
  __syncthreads();
}

// Cost (and optionally raw residuals in reference row order) over every residual block.
// Item index space: [0,N) CAM, then DEPTH_TRI, DEPTH_MESH, PRIOR.
__global__ void __launch_bounds__(256) residual_kernel(View v, const double* __restrict__ x, const CamTab* __restrict__ tab,
                                                       const DepthTab* __restrict__ dtab, int apply_loss,
                                                       double* __restrict__ cost, double* __restrict__ residuals, int64_t first = 0) {
  const int64_t total = v.N + v.n_dtri + v.n_dmesh + v.n_prior;
  double acc = 0;
  for (int64_t i = first + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    Item it;
    int64_t row0;
    if (i < v.N) { eval_cam<false>(v, x, tab, i, it, nullptr, nullptr); row0 = v.obs_row0[i]; }
    else if (i < v.N + v.n_dtri) { int64_t j = i - v.N; eval_depth(v, x, tab, dtab, MVBA_BLOCK_DEPTH_TRI, j, it); row0 = v.dtri_row0[j]; }
    else if (i < v.N + v.n_dtri + v.n_dmesh) { int64_t j = i - v.N - v.n_dtri; eval_depth(v, x, tab, dtab, MVBA_BLOCK_DEPTH_MESH, j, it); row0 = v.dmesh_row0[j]; }
    else { int64_t j = i - v.N - v.n_dtri - v.n_dmesh; eval_prior(v, x, j, it); row0 = v.prior_row0[j]; }
    double s = 0;
    for (int a = 0; a < it.nr; a++) s += it.r[a] * it.r[a];
    if (apply_loss) { double hr, sr; cauchy(it.loss_a, s, &hr, &sr); acc += hr; }
    else acc += 0.5 * s;
    if (residuals) for (int a = 0; a < it.nr; a++) residuals[row0 + a] = it.r[a];
  }
  block_accumulate(acc, cost);
}

// Hook layout (include/mvba.h, mvba_evaluate_jacobian): columns in the reference's block order.
__device__ __forceinline__ int hook_col(const View& v, int kind, int vc) {
  if (vc < 21) return vc;  // beg7 end7 ref_to_cam7
  if (kind == MVBA_BLOCK_CAM) {
    if (vc == 21) return 24;        // offset after xyz3
    return 25 + (vc - MVBA_VPOSE);  // focal, centre, distortion
  }
  if (vc == 21) return (kind == MVBA_BLOCK_DEPTH_TRI) ? 21 + v.nd2i + 1 + 3 : 21 + v.nd2i + 1;
  return 21 + (vc - MVBA_VPOSE);    // depth_to_image, scale
}

// Stored Jacobian: one thread per block of `kind`; J row-major [count][rows][W], no loss correction,
// every parameter block included.  Also the timed "unfused" residual+Jacobian kernel of bench.py.
__global__ void __launch_bounds__(128) jacobian_store_kernel(View v, const double* __restrict__ x, const CamTab* __restrict__ tab,
                                                             const DepthTab* __restrict__ dtab, int kind, int64_t first,
                                                             int64_t count, int W, double* __restrict__ rout,
                                                             double* __restrict__ J) {
  int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= count) return;
  const int64_t j = first + q;
  Item it;
  const int rows = (kind == MVBA_BLOCK_CAM) ? 2 : 3;
  double* Jr = J + (size_t)q * rows * W;
  if (kind == MVBA_BLOCK_CAM) {
    // distortion rows are written in place (columns 28..)
    eval_cam<true>(v, x, tab, j, it, Jr + 28, Jr + W + 28);
  } else if (kind == MVBA_BLOCK_DEPTH_TRI || kind == MVBA_BLOCK_DEPTH_MESH) {
    eval_depth(v, x, tab, dtab, kind, j, it);
  } else {
    // priors of both kinds share one list; `first` indexes the merged list restricted by the host
    eval_prior(v, x, kind == MVBA_BLOCK_TRI ? v.n_pm + j : j, it);
  }
  for (int a = 0; a < rows; a++) rout[q * rows + a] = it.r[a];
  if (kind == MVBA_BLOCK_MESH_TRI || kind == MVBA_BLOCK_TRI) {
    for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) Jr[a * W + b] = it.E[3 * a + b];
    return;
  }
  const int vw = item_vwidth(v, it);
  const int ndz = (kind == MVBA_BLOCK_CAM) ? v.sens[it.sensor].nd_nz : 0;
  for (int vc = 0; vc < vw; vc++) {
    if (kind == MVBA_BLOCK_CAM && vc >= MVBA_VPOSE + 3) break;  // distortion already in place
    double col[3];
    item_jcol(v, it, tab, dtab, vc, nullptr, nullptr, col);
    int hc = hook_col(v, kind, vc);
    for (int a = 0; a < rows; a++) Jr[a * W + hc] = col[a];
  }
  if (kind == MVBA_BLOCK_CAM) {
    for (int a = 0; a < 2; a++) {
      for (int b = 0; b < 3; b++) Jr[a * W + 21 + b] = it.E[3 * a + b];
      for (int c = 28 + ndz; c < W; c++) Jr[a * W + c] = 0.0;  // RPC undistortion half / narrower sensors
    }
  } else if (kind == MVBA_BLOCK_DEPTH_TRI) {
    int c0 = 21 + v.nd2i + 1;
    for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) Jr[a * W + c0 + b] = it.E[3 * a + b];
  }
}

}  // namespace mvba
