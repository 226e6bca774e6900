// kernels_camera.cuh -- pass A of the normal-equation build: camera-major accumulation of F'F, F'r
// and diag(F'F).
//
// Every residual block with non-landmark ("f") columns belongs to exactly one camera image, and all
// blocks of one image share the same f parameter blocks (beg pose, end pose, rig transform, offset,
// intrinsics, depth transform).  So F'F restricted to one image is a small dense symmetric matrix
// (config 5: 21 x 21, RPC config 4: 46 x 46).  One CTA owns a chunk of one image's blocks:
//   phase 1  one thread per block: residual + analytic Jacobian, Cauchy-corrected and Jacobi-scaled,
//            staged as dense rows [J | r] in shared memory.  The pose / rig / offset columns are
//            A d  with  d = coef, coef x Z or coef x (Xc - t_rc)  for the image's coefficient rows
//            (CamCoef, staged once per chunk), so the per-block loop has no table loads and no branches
//            that differ between threads.
//   phase 2  [J | r]' [J | r]: images with <= 24 staged columns (poses + rig, no free intrinsics) run it on
//            the FP64 tensor cores (DMMA m8n8k4: the eight warps split the rows, every warp keeps the six
//            upper 8x8 blocks); wider images fold 4x4 register patches (thread = patch x row group).
//   flush    once per chunk: -> reduced system with atomics (no per-observation atomics, so
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
  const CamCoef* coef;         // per image: pose-part coefficient rows (prep_cameras_kernel)
  int batch_rows;              // rows of the staging buffer (blocks staged per round = batch_rows / rows per block, <= CTA size)
  int ldj;                     // row stride of the staging buffer (doubles; = 12 (mod 16) so that DMMA fragment loads are conflict free)
  int mode;                    // 0 full, 1 squared column norms only
};

// Shared memory: Js[batch_rows + 4][ldj] | ccoef[23][3] | ctrc[3] | cscale[MAXACT] doubles | act_gc | act_vcam | act_vdep ints
__host__ __device__ inline size_t cam_pass_smem(int batch_rows, int ldj) {
  return ((size_t)(batch_rows + 4) * ldj + 72 + MVBA_CAM_MAXACT) * 8 + 3 * MVBA_CAM_MAXACT * 4 + 16;
}

__global__ void __launch_bounds__(MVBA_CAM_THREADS, 2) camera_pass_kernel(View v, CamPassArgs a, Reduced S) {
  extern __shared__ double smem_d[];
  double* Js = smem_d;
  double* ccoef = Js + (size_t)(a.batch_rows + 4) * a.ldj;   // 23 x 3
  double* ctrc = ccoef + 69;
  double* cscale = ccoef + 72;
  int* act_gc = reinterpret_cast<int*>(cscale + MVBA_CAM_MAXACT);
  int* act_vcam = act_gc + MVBA_CAM_MAXACT;
  int* act_vdep = act_vcam + MVBA_CAM_MAXACT;
  __shared__ int s_nact, s_dist0;
  const int tid = threadIdx.x, lane = tid & 31, wib = tid >> 5;
  constexpr int NW = MVBA_CAM_THREADS / 32;

  for (int ch = blockIdx.x; ch < a.nchunks; ch += gridDim.x) {
    const int cam = a.chunk_cam[ch];
    const int i0 = a.chunk_beg[ch], i1 = a.chunk_beg[ch + 1];
    const int sensor = v.cam_sensor[cam];
    const SensorDev& sd = v.sens[sensor];
    // ---- active column list of this image (camera-local virtual columns):
    //   0-21 pose part | 22 focal | 23-24 centre | 25.. distortion (nd_nz) | then depth_to_image (nd2i) | scale
    if (tid == 0) {
      Item it; it.cid = cam; it.sensor = sensor;
      int n = 0, d0 = -1;
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
        if (vcam >= MVBA_VPOSE + 3 && d0 < 0) d0 = n;
        if (n < MVBA_CAM_MAXACT - 1) { act_gc[n] = gc; act_vcam[n] = vcam; act_vdep[n] = vdep; cscale[n] = a.jscale_f ? a.jscale_f[gc] : 1.0; n++; }
      }
      s_nact = n; s_dist0 = d0;
    }
    if (tid >= 32 && tid < 32 + 69) ccoef[tid - 32] = __ldg(&a.coef[cam].c[0][0] + (tid - 32));
    if (tid >= 128 && tid < 131) ctrc[tid - 128] = __ldg(a.tab[cam].trc + (tid - 128));
    __syncthreads();
    const int nact = s_nact, dist0 = s_dist0;
    if (nact == 0) { __syncthreads(); continue; }
    const int ncol = nact + 1;  // + r column
    const bool tensor = ncol <= 24 && (a.ldj & 15) == 12;
    const int wcols = tensor ? ((ncol + 7) & ~7) : ((ncol + 3) & ~3);  // zero-padded row width the fold reads
    // scalar fold: 4x4 patches over the upper block triangle, several row groups
    const int nb = (ncol + 3) >> 2;
    const int npatch = nb * (nb + 1) / 2;
    const int ngrp = max(1, MVBA_CAM_THREADS / npatch);
    const bool worker = !tensor && tid < ngrp * npatch;
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
    // tensor fold: blocks (0,0) (0,1) (0,2) (1,1) (1,2) (2,2) of 8x8, two accumulator doubles each
    double tc[6][2];
#pragma unroll
    for (int ii = 0; ii < 6; ii++) { tc[ii][0] = 0; tc[ii][1] = 0; }

    const int rpi = (a.citems[i1 - 1] >= v.N) ? 3 : 2;  // rows per staged block (codes ascend inside an image)
    const int bmax = min(MVBA_CAM_THREADS, a.batch_rows / rpi);
    const int batch = ((i1 - i0) + ((i1 - i0 + bmax - 1) / bmax) - 1) / ((i1 - i0 + bmax - 1) / bmax);  // equal rounds
    for (int b0 = i0; b0 < i1; b0 += batch) {
      const int cnt = min(batch, i1 - b0);
      // ---- phase 1: evaluate and stage ----
      if (tid < cnt) {
        const int q = tid;
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
        const bool three = it.nr == 3;
        double s = it.r[0] * it.r[0] + it.r[1] * it.r[1];
        if (three) s += it.r[2] * it.r[2];
        double hr, sr;
        cauchy(it.loss_a, s, &hr, &sr);
        const double A0[3] = {it.A[0], it.A[1], it.A[2]}, A1[3] = {it.A[3], it.A[4], it.A[5]};
        const double A2[3] = {three ? it.A[6] : 0.0, three ? it.A[7] : 0.0, three ? it.A[8] : 0.0};
        const double Z[3] = {it.Z[0], it.Z[1], it.Z[2]};
        const double y[3] = {it.Xc[0] - ctrc[0], it.Xc[1] - ctrc[1], it.Xc[2] - ctrc[2]};
        for (int c = 0; c < nact; c++) {
          const int vc = is_cam ? act_vcam[c] : act_vdep[c];
          double col[3] = {0, 0, 0};
          const double sc = sr * cscale[c];
          if (vc >= 0 && vc < MVBA_VPOSE) {
            // d Xc / d (pose-part column): coef (translations), coef x Z (pose rotations, offset), coef x y (rig rotation)
            const double* cf = ccoef + 3 * vc;
            double d0 = cf[0], d1 = cf[1], d2 = cf[2];
            const bool rot = (vc >= 3 && vc < 7) || (vc >= 10 && vc < 14) || vc >= 17;
            if (rot) {
              const double* w = (vc >= 17 && vc < 21) ? y : Z;
              const double e0 = d1 * w[2] - d2 * w[1], e1 = d2 * w[0] - d0 * w[2], e2 = d0 * w[1] - d1 * w[0];
              d0 = e0; d1 = e1; d2 = e2;
              if (vc == 21) { d0 += cf[3]; d1 += cf[4]; d2 += cf[5]; }
            }
            col[0] = A0[0] * d0 + A0[1] * d1 + A0[2] * d2;
            col[1] = A1[0] * d0 + A1[1] * d1 + A1[2] * d2;
            col[2] = A2[0] * d0 + A2[1] * d1 + A2[2] * d2;
          } else if (vc >= 0) {
            if (is_cam && vc >= MVBA_VPOSE + 3) { col[0] = row0[c]; col[1] = row1[c]; }
            else item_jcol(v, it, a.tab, a.dtab, vc, nullptr, nullptr, col);
          }
          row0[c] = col[0] * sc; row1[c] = col[1] * sc;
          if (rpi == 3) row2[c] = three ? col[2] * sc : 0.0;
        }
        row0[nact] = sr * it.r[0]; row1[nact] = sr * it.r[1];
        if (rpi == 3) row2[nact] = three ? sr * it.r[2] : 0.0;
        for (int c = ncol; c < wcols; c++) { row0[c] = 0; row1[c] = 0; if (rpi == 3) row2[c] = 0; }
      }
      const int nrows = rpi * cnt;
      if (tensor && tid < 4 * 32) {  // rows up to the next multiple of 4: the DMMA K dimension
        const int r = nrows + (tid >> 5), c = tid & 31;
        if (r < ((nrows + 3) & ~3) && c < wcols) Js[(size_t)r * a.ldj + c] = 0.0;
      }
      __syncthreads();
      // ---- phase 2: [J | r]' [J | r] ----
      if (tensor) {
        // a DMMA k-step is four staged rows; fragment f holds Js[r0 + lane%4][8 f + lane/4], which is both the A fragment
        // of block row f and the B fragment of block column f
        const int nb8 = wcols >> 3, gid = lane >> 2, tig = lane & 3;
        const double* base = Js + (size_t)tig * a.ldj + gid;
        const int nks = (nrows + 3) >> 2;
#pragma unroll 2
        for (int ks = wib; ks < nks; ks += NW) {
          const double* pr = base + (size_t)(4 * ks) * a.ldj;
          const double f0 = pr[0];
          const double f1 = nb8 > 1 ? pr[8] : 0.0;
          const double f2 = nb8 > 2 ? pr[16] : 0.0;
          dmma884(tc[0][0], tc[0][1], f0, f0);
          if (nb8 > 1) { dmma884(tc[1][0], tc[1][1], f0, f1); dmma884(tc[3][0], tc[3][1], f1, f1); }
          if (nb8 > 2) { dmma884(tc[2][0], tc[2][1], f0, f2); dmma884(tc[4][0], tc[4][1], f1, f2); dmma884(tc[5][0], tc[5][1], f2, f2); }
        }
      }
      if (worker) {
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
    if (tensor) {
      const int gid = lane >> 2, tig = lane & 3;
      const int tbi[6] = {0, 0, 0, 1, 1, 2}, tbj[6] = {0, 1, 2, 1, 2, 2};
#pragma unroll
      for (int blk = 0; blk < 6; blk++)
#pragma unroll
        for (int e = 0; e < 2; e++) {
          const int i = 8 * tbi[blk] + gid, j = 8 * tbj[blk] + 2 * tig + e;
          const double val = tc[blk][e];
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
