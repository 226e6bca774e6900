// kernels_camera.cuh -- pass A of the normal-equation build: camera-major accumulation of F'F, F'r
// and diag(F'F).
//
// Every residual block with non-landmark ("f") columns belongs to exactly one camera image, and all
// blocks of one image share the same f parameter blocks (beg pose, end pose, rig transform, offset,
// intrinsics, depth transform).  So F'F restricted to one image is a small dense symmetric matrix
// (config 5: 21 x 21, RPC config 4: 46 x 46).  One CTA owns a chunk of one image's blocks:
//   phase 1  one thread per block: residual + analytic Jacobian, Cauchy-corrected and Jacobi-scaled,
//            staged as dense rows [J | r] in shared memory
//   phase 2  the rows are folded into 4x4 register patches of [J | r]' [J | r]  (thread = patch x row
//            group), two LDS.128 per 16 DFMA
//   flush    once per chunk: patches -> reduced system with atomics (no per-observation atomics, so
//            the shared rig / intrinsics columns are no longer a same-address hot spot)
// Ceres 1.14 equivalent: the F'F + F'r part of SchurEliminator / ImplicitSchurComplement set-up.
#pragma once
#include "kernels_build.cuh"

namespace mvba {

#define MVBA_CAM_THREADS 256
#define MVBA_CAM_MAXACT 88  // active f columns of one image (+ r column) must fit 22 blocks of 4

struct CamPassArgs {
  const double* x;
  const CamTab* tab;
  const DepthTab* dtab;
  const double* jscale_f;      // Jacobi column scaling (null: 1)
  const int32_t* chunk_cam;    // per chunk: camera image
  const int32_t* chunk_beg;    // nchunks + 1: range into citems
  const int32_t* citems;       // block codes, camera-major: [0,N) CAM, then DEPTH_TRI, then DEPTH_MESH
  int nchunks;
  int batch_items;             // blocks staged per round
  int ldj;                     // row stride of the staging buffer (doubles, multiple of 4)
  int mode;                    // 0 full, 1 squared column norms only
};

// Shared memory: Js[batch_items * 3][ldj] doubles | act_gc[MAXACT] | act_vcam[MAXACT] | act_vdep[MAXACT] ints
__host__ __device__ inline size_t cam_pass_smem(int batch_items, int ldj) {
  return (size_t)batch_items * 3 * ldj * 8 + 3 * MVBA_CAM_MAXACT * 4 + 16;
}

__global__ void __launch_bounds__(MVBA_CAM_THREADS) camera_pass_kernel(View v, CamPassArgs a, Reduced S) {
  extern __shared__ double smem_d[];
  double* Js = smem_d;
  int* act_gc = reinterpret_cast<int*>(Js + (size_t)a.batch_items * 3 * a.ldj);
  int* act_vcam = act_gc + MVBA_CAM_MAXACT;
  int* act_vdep = act_vcam + MVBA_CAM_MAXACT;
  __shared__ int s_nact, s_dist0, s_distn;
  const int tid = threadIdx.x;

  for (int ch = blockIdx.x; ch < a.nchunks; ch += gridDim.x) {
    const int cam = a.chunk_cam[ch];
    const int i0 = a.chunk_beg[ch], i1 = a.chunk_beg[ch + 1];
    const int sensor = v.cam_sensor[cam];
    const SensorDev& sd = v.sens[sensor];
    // ---- active column list of this image (camera-local virtual columns):
    //   0-21 pose part | 22 focal | 23-24 centre | 25.. distortion (nd_nz) | then depth_to_image (nd2i) | scale
    if (tid == 0) {
      Item it; it.cid = cam; it.sensor = sensor;
      int n = 0, d0 = -1, dn = 0;
      const int ncam = MVBA_VPOSE + 3 + sd.nd_nz, ndep = v.nd2i + 1;
      for (int cv = 0; cv < ncam + ndep; cv++) {
        int vcam = -1, vdep = -1;
        if (cv < MVBA_VPOSE) { vcam = cv; vdep = cv; }
        else if (cv < ncam) vcam = cv;
        else vdep = MVBA_VPOSE + (cv - ncam);
        int gc;
        if (vcam >= 0) { it.kind = MVBA_BLOCK_CAM; gc = item_gcol(v, it, vcam); }
        else { it.kind = MVBA_BLOCK_DEPTH_TRI; gc = item_gcol(v, it, vdep); }
        if (gc < 0) continue;
        if (vcam >= MVBA_VPOSE + 3) { if (d0 < 0) d0 = n; dn++; }
        if (n < MVBA_CAM_MAXACT - 1) { act_gc[n] = gc; act_vcam[n] = vcam; act_vdep[n] = vdep; n++; }
      }
      s_nact = n; s_dist0 = d0; s_distn = dn;
    }
    __syncthreads();
    const int nact = s_nact, dist0 = s_dist0;
    if (nact == 0) { __syncthreads(); continue; }
    const int ncol = nact + 1;  // + r column
    const int nb = (ncol + 3) >> 2;
    const int npatch = nb * (nb + 1) / 2;
    const int ngrp = max(1, MVBA_CAM_THREADS / npatch);
    const bool worker = tid < ngrp * npatch;
    const int grp = tid / npatch;
    int bi = 0, bj = 0;
    {
      int pp = tid % npatch;  // row-major over the upper block triangle
      int rowlen = nb;
      while (pp >= rowlen) { pp -= rowlen; rowlen--; bi++; }
      bj = bi + pp;
    }
    double acc[4][4];
#pragma unroll
    for (int ii = 0; ii < 4; ii++)
#pragma unroll
      for (int jj = 0; jj < 4; jj++) acc[ii][jj] = 0;

    const int rpi = (a.citems[i1 - 1] >= v.N) ? 3 : 2;  // rows per staged block (codes ascend inside an image)
    for (int b0 = i0; b0 < i1; b0 += a.batch_items) {
      const int cnt = min(a.batch_items, i1 - b0);
      // ---- phase 1: evaluate and stage ----
      for (int q = tid; q < cnt; q += MVBA_CAM_THREADS) {
        const int64_t code = a.citems[b0 + q];
        double* row0 = Js + (size_t)(rpi * q) * a.ldj;
        double* row1 = row0 + a.ldj;
        double* row2 = (rpi == 3) ? row1 + a.ldj : row1;  // CAM-only chunks have no third row
        Item it;
        const bool is_cam = code < v.N;
        if (is_cam) {
          // distortion rows are written in place and scaled below
          double* jd0 = dist0 >= 0 ? row0 + dist0 : nullptr;
          double* jd1 = dist0 >= 0 ? row1 + dist0 : nullptr;
          eval_cam<true>(v, a.x, a.tab, code, it, jd0, jd1);
        } else if (code < v.N + v.n_dtri) {
          eval_depth(v, a.x, a.tab, a.dtab, MVBA_BLOCK_DEPTH_TRI, code - v.N, it);
        } else {
          eval_depth(v, a.x, a.tab, a.dtab, MVBA_BLOCK_DEPTH_MESH, code - v.N - v.n_dtri, it);
        }
        double s = 0;
        for (int k = 0; k < it.nr; k++) s += it.r[k] * it.r[k];
        double hr, sr;
        cauchy(it.loss_a, s, &hr, &sr);
        for (int c = 0; c < nact; c++) {
          const int vc = is_cam ? act_vcam[c] : act_vdep[c];
          double col[3] = {0, 0, 0};
          const double sc = sr * (a.jscale_f ? a.jscale_f[act_gc[c]] : 1.0);
          if (vc >= 0) {
            if (is_cam && vc >= MVBA_VPOSE + 3) { col[0] = row0[c]; col[1] = row1[c]; }
            else item_jcol(v, it, a.tab, a.dtab, vc, nullptr, nullptr, col);
          }
          row0[c] = col[0] * sc; row1[c] = col[1] * sc;
          if (rpi == 3) row2[c] = (it.nr == 3) ? col[2] * sc : 0.0;
        }
        row0[nact] = sr * it.r[0]; row1[nact] = sr * it.r[1];
        if (rpi == 3) row2[nact] = (it.nr == 3) ? sr * it.r[2] : 0.0;
        for (int c = ncol; c < 4 * nb; c++) { row0[c] = 0; row1[c] = 0; if (rpi == 3) row2[c] = 0; }
      }
      __syncthreads();
      // ---- phase 2: [J | r]' [J | r] in register patches ----
      if (worker) {
        const int nrows = rpi * cnt;
        const double* pa = Js + 4 * bi;
        const double* pb = Js + 4 * bj;
        for (int r = grp; r < nrows; r += ngrp) {
          const double2* qa = reinterpret_cast<const double2*>(pa + (size_t)r * a.ldj);
          const double2* qb = reinterpret_cast<const double2*>(pb + (size_t)r * a.ldj);
          const double2 a01 = qa[0], a23 = qa[1], b01 = qb[0], b23 = qb[1];
          const double av[4] = {a01.x, a01.y, a23.x, a23.y};
          const double bv[4] = {b01.x, b01.y, b23.x, b23.y};
#pragma unroll
          for (int ii = 0; ii < 4; ii++)
#pragma unroll
            for (int jj = 0; jj < 4; jj++) acc[ii][jj] = fma(av[ii], bv[jj], acc[ii][jj]);
        }
      }
      __syncthreads();
    }
    // ---- flush ----
    if (worker) {
#pragma unroll
      for (int ii = 0; ii < 4; ii++)
#pragma unroll
        for (int jj = 0; jj < 4; jj++) {
          const int i = 4 * bi + ii, j = 4 * bj + jj;
          const double val = acc[ii][jj];
          acc[ii][jj] = 0;
          if (i > j || i >= nact || j > nact || val == 0.0) continue;
          const int gi = act_gc[i];
          if (j == nact) {
            if (a.mode == 0) { atomicAdd(S.g + gi, val); atomicAdd(S.rhs + gi, val); }
            continue;
          }
          const int gj = act_gc[j];
          if (i == j) atomicAdd(S.diagF + gi, val);
          if (a.mode == 0) {
            addS(S, gi, gj, val);
            if (S.Fpp) addFdiag(S, gi, gj, val);
          }
        }
    }
    __syncthreads();
  }
}

}  // namespace mvba
