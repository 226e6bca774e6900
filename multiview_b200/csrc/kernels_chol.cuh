// kernels_chol.cuh -- direct solve of the reduced camera system: dense blocked Cholesky with the panel solve and the
// trailing update on the FP64 tensor cores (mma.sync m8n8k4 f64 = DMMA).  mvba_solve_options.linear_solver = 1; the
// reference's ITERATIVE_SCHUR path (PCG, kernels_pcg.cuh) stays the default.  (north_star: "FP64 DMMA tensor cores ...
// for the dense reduced-camera Cholesky")  Used for reduced systems of a few thousand columns
// (BASELINE configs 1 and 2) and as the exact reference the PCG is tested against.
//
// Layout: A is (N + 64) x N row-major, N = nf rounded up to the panel width 64; only the lower triangle is read or written.
// Row N carries the right-hand side: factoring [S b; b' .] leaves y = L^-1 b in that row, so no forward substitution is
// needed.  Per panel k (right-looking):
//   chol_diag_kernel    one CTA: Cholesky of the 64 x 64 diagonal block in shared memory, and its inverse
//   chol_panel_kernel   CTA per 64-row tile below: L_ik = A_ik L_kk^-T as a 64^3 GEMM on DMMA
//   chol_syrk_kernel    CTA per 64 x 64 tile of the trailing lower triangle: A_ij -= L_ik L_jk' on DMMA
// and one CTA walks the panels backwards for  L' x = y  (chol_backsolve_kernel; the inverses of the diagonal blocks are kept).
#pragma once
#include "kernels_pcg.cuh"

namespace mvba {

#define MVBA_CHOL_NB 64
#define MVBA_CHOL_LDS 68  // shared-memory row stride (doubles), = 4 (mod 16): conflict-free DMMA fragment loads

// Block-sparse reduced system (after damping and mirroring, kernels_pcg.cuh) -> dense lower triangle + right-hand side row.
// The matrix is zeroed before the launch.
__global__ void chol_expand_kernel(Reduced S, Bsr B, const double* __restrict__ rhs, double* __restrict__ A, int N) {
  const int nf = S.nf, npc = S.npc, nB = S.B, nb = S.nb;
  const int64_t nbr = B.row_ptr ? B.row_ptr[nb] : 0;  // block entries of the pose-pose part, mirrored list
  const int64_t total = nbr * 49 + (int64_t)npc * nB + (int64_t)nB * nB + nf + (N - nf);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t q = i;
    if (q < nbr * 49) {  // entry e of the block list belongs to the block row found by search
      const int e = (int)(q / 49), rc = (int)(q % 49);
      const int c = B.col[e], sl = B.slot[e];
      int lo = 0, hi = nb - 1;  // row u with row_ptr[u] <= e < row_ptr[u + 1]
      while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (B.row_ptr[mid] <= e) lo = mid; else hi = mid - 1; }
      const int u = lo;
      if (c > u) continue;  // lower triangle only: block row u, block column c <= u
      const int r = rc / 7, cc = rc % 7;
      const double val = sl >= 0 ? S.Spp[(size_t)sl * 49 + 7 * r + cc] : S.Spp[(size_t)(-sl - 1) * 49 + 7 * cc + r];
      const int gi = 7 * u + r, gj = 7 * c + cc;
      if (gj <= gi) A[(size_t)gi * N + gj] = val;
      continue;
    }
    q -= nbr * 49;
    if (q < (int64_t)npc * nB) {
      const int g = (int)(q / nB), b = (int)(q % nB);
      A[(size_t)(npc + b) * N + g] = S.Spb[(size_t)g * nB + b];
      continue;
    }
    q -= (int64_t)npc * nB;
    if (q < (int64_t)nB * nB) {
      const int b = (int)(q / nB), c = (int)(q % nB);
      if (c <= b) A[(size_t)(npc + b) * N + npc + c] = S.Sbb[(size_t)b * nB + c];
      continue;
    }
    q -= (int64_t)nB * nB;
    if (q < nf) { A[(size_t)N * N + q] = rhs[q]; continue; }
    q -= nf;
    A[(size_t)(nf + q) * N + nf + q] = 1.0;  // padding: identity
  }
}

// One CTA.  Linv: 64 x 64 row-major, lower triangular inverse of the factored diagonal block.
__global__ void __launch_bounds__(256) chol_diag_kernel(double* __restrict__ A, int N, int k, double* __restrict__ Linv) {
  extern __shared__ __align__(16) double chol_sm[];
  double (*a)[MVBA_CHOL_NB + 1] = reinterpret_cast<double (*)[MVBA_CHOL_NB + 1]>(chol_sm);
  double (*x)[MVBA_CHOL_NB + 1] = reinterpret_cast<double (*)[MVBA_CHOL_NB + 1]>(chol_sm + MVBA_CHOL_NB * (MVBA_CHOL_NB + 1));
  const int tid = threadIdx.x;
  for (int idx = tid; idx < 4096; idx += 256) {
    const int r = idx >> 6, c = idx & 63;
    a[r][c] = c <= r ? A[(size_t)(k + r) * N + k + c] : 0.0;
  }
  __syncthreads();
  for (int j = 0; j < MVBA_CHOL_NB; j++) {
    if (tid == 0) a[j][j] = sqrt(a[j][j]);  // not positive definite -> NaN, which the LM loop treats as an invalid step
    __syncthreads();
    const double id = 1.0 / a[j][j];
    if (tid > j && tid < MVBA_CHOL_NB) a[tid][j] *= id;
    __syncthreads();
    for (int idx = tid; idx < 4096; idx += 256) {
      const int r = idx >> 6, c = idx & 63;
      if (c > j && r >= c) a[r][c] -= a[r][j] * a[c][j];
    }
    __syncthreads();
  }
  // X = L^-1, one thread per column (forward substitution), four partial sums for instruction-level parallelism
  if (tid < MVBA_CHOL_NB) {
    const int c = tid;
    for (int i = 0; i < MVBA_CHOL_NB; i++) {
      double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
      if (i >= c) {
        int j = c;
        for (; j + 3 < i; j += 4) {
          s0 = fma(a[i][j], x[j][c], s0); s1 = fma(a[i][j + 1], x[j + 1][c], s1);
          s2 = fma(a[i][j + 2], x[j + 2][c], s2); s3 = fma(a[i][j + 3], x[j + 3][c], s3);
        }
        for (; j < i; j++) s0 = fma(a[i][j], x[j][c], s0);
        x[i][c] = ((i == c ? 1.0 : 0.0) - ((s0 + s1) + (s2 + s3))) / a[i][i];
      } else {
        x[i][c] = 0.0;
      }
    }
  }
  __syncthreads();
  for (int idx = tid; idx < 4096; idx += 256) {
    const int r = idx >> 6, c = idx & 63;
    if (c <= r) A[(size_t)(k + r) * N + k + c] = a[r][c];
    Linv[idx] = x[r][c];
  }
}

// 64 x 64 x 64 product on DMMA: acc (+/-)= P Q' with P, Q row-major in shared memory (row stride MVBA_CHOL_LDS); warp w owns the
// eight 8 x 8 blocks of block row w.
template <bool NEG>
__device__ __forceinline__ void chol_gemm64(const double* __restrict__ Ps, const double* __restrict__ Qs, double (*acc)[2], int wib, int lane) {
  const int gid = lane >> 2, tig = lane & 3;
  const double* pa = Ps + (size_t)(8 * wib + gid) * MVBA_CHOL_LDS + tig;
  const double* pb = Qs + (size_t)gid * MVBA_CHOL_LDS + tig;
#pragma unroll 4
  for (int k0 = 0; k0 < MVBA_CHOL_NB; k0 += 4) {
    const double av = NEG ? -pa[k0] : pa[k0];
#pragma unroll
    for (int bj = 0; bj < 8; bj++) dmma884(acc[bj][0], acc[bj][1], av, pb[(size_t)(8 * bj) * MVBA_CHOL_LDS + k0]);
  }
}

__device__ __forceinline__ void chol_load_tile(const double* __restrict__ src, int ld, double* __restrict__ dst, int tid) {
  for (int idx = tid; idx < 2048; idx += 256) {  // 64 rows x 32 double2
    const int r = idx >> 5, c2 = idx & 31;
    const double2 v2 = *reinterpret_cast<const double2*>(src + (size_t)r * ld + 2 * c2);
    *reinterpret_cast<double2*>(dst + (size_t)r * MVBA_CHOL_LDS + 2 * c2) = v2;
  }
}

// grid = row tiles below the diagonal block (the right-hand-side tile included): A_ik <- A_ik Linv'
__global__ void __launch_bounds__(256) chol_panel_kernel(double* __restrict__ A, int N, int k, const double* __restrict__ Linv) {
  extern __shared__ __align__(16) double chol_sm[];
  double* Ps = chol_sm;
  double* Qs = chol_sm + MVBA_CHOL_NB * MVBA_CHOL_LDS;
  const int tid = threadIdx.x, lane = tid & 31, wib = tid >> 5;
  double* tile = A + (size_t)(k + MVBA_CHOL_NB * (blockIdx.x + 1)) * N + k;
  chol_load_tile(tile, N, Ps, tid);
  chol_load_tile(Linv, MVBA_CHOL_NB, Qs, tid);
  __syncthreads();
  double acc[8][2];
#pragma unroll
  for (int i = 0; i < 8; i++) { acc[i][0] = 0; acc[i][1] = 0; }
  chol_gemm64<false>(Ps, Qs, acc, wib, lane);
  const int gid = lane >> 2, tig = lane & 3;
#pragma unroll
  for (int bj = 0; bj < 8; bj++)
    *reinterpret_cast<double2*>(tile + (size_t)(8 * wib + gid) * N + 8 * bj + 2 * tig) = make_double2(acc[bj][0], acc[bj][1]);
}

// grid = tiles (i >= j) of the trailing lower triangle after panel k, m = tile rows below the panel (the last one is the
// right-hand-side tile, which has no diagonal tile of its own): A_ij -= L_ik L_jk'
__global__ void __launch_bounds__(256) chol_syrk_kernel(double* __restrict__ A, int N, int k, int m) {
  extern __shared__ __align__(16) double chol_sm[];
  double* Ps = chol_sm;
  double* Qs = chol_sm + MVBA_CHOL_NB * MVBA_CHOL_LDS;
  const int tid = threadIdx.x, lane = tid & 31, wib = tid >> 5;
  int ti, tj;
  tri_decode(blockIdx.x, tj, ti);  // tj <= ti < m
  if (ti == m - 1 && tj == m - 1) return;
  const int r0 = k + MVBA_CHOL_NB * (ti + 1), c0 = k + MVBA_CHOL_NB * (tj + 1);
  chol_load_tile(A + (size_t)r0 * N + k, N, Ps, tid);
  chol_load_tile(A + (size_t)c0 * N + k, N, Qs, tid);
  __syncthreads();
  double* tile = A + (size_t)r0 * N + c0;
  const int gid = lane >> 2, tig = lane & 3;
  double acc[8][2];
#pragma unroll
  for (int bj = 0; bj < 8; bj++) {
    const double2 v2 = *reinterpret_cast<const double2*>(tile + (size_t)(8 * wib + gid) * N + 8 * bj + 2 * tig);
    acc[bj][0] = v2.x; acc[bj][1] = v2.y;
  }
  chol_gemm64<true>(Ps, Qs, acc, wib, lane);
#pragma unroll
  for (int bj = 0; bj < 8; bj++)
    *reinterpret_cast<double2*>(tile + (size_t)(8 * wib + gid) * N + 8 * bj + 2 * tig) = make_double2(acc[bj][0], acc[bj][1]);
}

// One CTA of 1024 threads: x = L^-T y, panels last to first; y is row N of A, Linv the kept inverses (panel-major).
__global__ void __launch_bounds__(1024) chol_backsolve_kernel(const double* __restrict__ A, int N, int nf, const double* __restrict__ Linv,
                                                               double* __restrict__ xout) {
  extern __shared__ __align__(16) double chol_sm[];
  double* xs = chol_sm;                      // N
  double* part = xs + N;                     // 16 x 64
  double* rs = part + 16 * MVBA_CHOL_NB;     // 64
  const int tid = threadIdx.x, c = tid & 63, g = tid >> 6;
  for (int i = tid; i < N; i += 1024) xs[i] = A[(size_t)N * N + i];
  __syncthreads();
  for (int k = N - MVBA_CHOL_NB; k >= 0; k -= MVBA_CHOL_NB) {
    double s = 0;
    for (int i = k + MVBA_CHOL_NB + g; i < N; i += 16) s = fma(A[(size_t)i * N + k + c], xs[i], s);
    part[g * MVBA_CHOL_NB + c] = s;
    __syncthreads();
    if (tid < MVBA_CHOL_NB) {
      double r = xs[k + tid];
#pragma unroll
      for (int gg = 0; gg < 16; gg++) r -= part[gg * MVBA_CHOL_NB + tid];
      rs[tid] = r;
    }
    __syncthreads();
    if (tid < MVBA_CHOL_NB) {
      const double* Li = Linv + (size_t)(k / MVBA_CHOL_NB) * 4096;
      double v = 0;
      for (int j = tid; j < MVBA_CHOL_NB; j++) v = fma(Li[j * MVBA_CHOL_NB + tid], rs[j], v);
      xs[k + tid] = v;
    }
    __syncthreads();
  }
  for (int i = tid; i < nf; i += 1024) xout[i] = xs[i];
}

#define MVBA_CHOL_SMEM_DIAG (2 * MVBA_CHOL_NB * (MVBA_CHOL_NB + 1) * 8)
#define MVBA_CHOL_SMEM_GEMM (2 * MVBA_CHOL_NB * MVBA_CHOL_LDS * 8)
#define MVBA_CHOL_MAX_N 8192

}  // namespace mvba
