// kernels_pcg.cuh -- LM damping of the reduced system, block-Jacobi preconditioner and the
// preconditioned conjugate-gradient iteration (Ceres 1.14 ConjugateGradientsSolver semantics:
// Q-based stopping with eta, residual refresh every 10 iterations), all on device.
#pragma once
#include "kernels_build.cuh"
#include "kernels_eval.cuh"
#include <cooperative_groups.h>

namespace mvba {

// Block-sparse symmetric structure for S x: each block row lists (col block, slot, transposed).
struct Bsr {
  const int32_t *row_ptr, *col, *slot;  // slot >= 0: Spp[slot] as stored; slot < 0: transpose of Spp[-slot-1]
  int nb;
};

// Scalars of one PCG solve; slots indexed by iteration to avoid re-zeroing races.
//  per iteration it: [0] rho = r.z   [1] pq = p.q   [2] Q1 = -x.(b+r)
struct PcgScal { double* it; int* done; int* iters; double* bnorm2; };

// --- finalize: damping, mirroring, preconditioner ------------------------------------------
// One thread per reduced column: D^2 = clamp(diag F'F)/radius added to the diagonal of S.
__global__ void damp_kernel(Reduced S, double radius, double min_diag, double max_diag, double* __restrict__ Dsq) {
  int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= S.nf) return;
  double d = fmin(fmax(S.diagF[g], min_diag), max_diag) / radius;
  Dsq[g] = d;
  if (g < S.npc) {
    int u = g / 7, r = g - 7 * u;
    int slot = S.pairtab[(size_t)u * S.nb + u];
    S.Spp[(size_t)slot * 49 + r * 8] += d;
    if (S.Fpp) S.Fpp[(size_t)u * 49 + r * 8] += d;
  } else {
    int b = g - S.npc;
    S.Sbb[(size_t)b * S.B + b] += d;
    if (S.Fbb) S.Fbb[(size_t)b * S.B + b] += d;
  }
}

// In-register inverse of a symmetric positive definite 7x7 given its upper triangle (row-major 49).
__device__ inline void inv7_upper(const double* __restrict__ src, double* __restrict__ dst) {
  double A[49];
#pragma unroll
  for (int i = 0; i < 7; i++)
#pragma unroll
    for (int j = 0; j < 7; j++) A[7 * i + j] = (j >= i) ? src[7 * i + j] : src[7 * j + i];
  // Gauss-Jordan without pivoting (SPD)
#pragma unroll
  for (int k = 0; k < 7; k++) {
    double piv = 1.0 / A[8 * k];
#pragma unroll
    for (int j = 0; j < 7; j++) A[7 * k + j] = (j == k) ? piv : A[7 * k + j] * piv;
#pragma unroll
    for (int i = 0; i < 7; i++) {
      if (i == k) continue;
      double f = A[7 * i + k];
#pragma unroll
      for (int j = 0; j < 7; j++) A[7 * i + j] = (j == k) ? -f * piv : A[7 * i + j] - f * A[7 * k + j];
    }
  }
#pragma unroll
  for (int i = 0; i < 49; i++) dst[i] = A[i];
}

// One thread per pose block: mirror the diagonal block of S and invert the preconditioner block.
__global__ void precond_pose_kernel(Reduced S, int use_F, double* __restrict__ Minv) {
  int u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= S.nb) return;
  int slot = S.pairtab[(size_t)u * S.nb + u];
  double* D = S.Spp + (size_t)slot * 49;
  for (int i = 0; i < 7; i++) for (int j = 0; j < i; j++) D[7 * i + j] = D[7 * j + i];
  inv7_upper(use_F ? S.Fpp + (size_t)u * 49 : D, Minv + (size_t)u * 49);
}

// One CTA: mirror Sbb, then invert the border block (B x B, SPD) by Gauss-Jordan in global memory.
// blk: parameter-block id of each border column; entries coupling different blocks are dropped so that
// the preconditioner has Ceres' block-diagonal partition (null: keep the border dense).
__global__ void __launch_bounds__(256) precond_border_kernel(Reduced S, int use_F, const int32_t* __restrict__ blk,
                                                              double* __restrict__ Minv, double* __restrict__ work) {
  const int B = S.B, tid = threadIdx.x, nt = blockDim.x;
  for (int i = tid; i < B * B; i += nt) { int r = i / B, c = i - r * B; if (c < r) S.Sbb[i] = S.Sbb[(size_t)c * B + r]; }
  __syncthreads();
  const double* src = use_F ? S.Fbb : S.Sbb;
  for (int i = tid; i < B * B; i += nt) {
    int r = i / B, c = i - r * B;
    double val = (c >= r) ? src[i] : src[(size_t)c * B + r];
    if (blk && blk[r] != blk[c]) val = 0.0;
    work[i] = val; Minv[i] = (r == c) ? 1.0 : 0.0;
  }
  __syncthreads();
  __shared__ double piv_s;
  for (int k = 0; k < B; k++) {
    if (tid == 0) piv_s = 1.0 / work[(size_t)k * B + k];
    __syncthreads();
    double piv = piv_s;
    for (int j = tid; j < B; j += nt) { work[(size_t)k * B + j] *= piv; Minv[(size_t)k * B + j] *= piv; }
    __syncthreads();
    for (int i = tid; i < B * B; i += nt) {
      int r = i / B, c = i - r * B;
      if (r == k) continue;
      double f = work[(size_t)r * B + k];
      if (f != 0.0) {
        Minv[i] -= f * Minv[(size_t)k * B + c];
        if (c != k) work[i] -= f * work[(size_t)k * B + c];
      }
    }
    __syncthreads();
    for (int r = tid; r < B; r += nt) if (r != k) work[(size_t)r * B + k] = 0.0;
    __syncthreads();
  }
}

// --- S x ----------------------------------------------------------------------------------
// One warp per pose block row (lanes over the row's blocks), then warps for the border rows.
// y must have its border part zeroed before the launch (pose rows add S_pb' x_pose atomically).
// Row w of S x for one warp (w < nb: pose block row, else border row w - nb).
// (x and y are deliberately not __restrict__: the fused PCG kernel reads vectors written earlier in the same launch,
// which must not go through the non-coherent read-only path.)
__device__ __forceinline__ void spmv_row(const Reduced& S, const Bsr& A, const double* x, double* y, int w, int lane) {
  const int B = S.B;
  if (w < A.nb) {
    double acc[7] = {0, 0, 0, 0, 0, 0, 0};
    for (int e = A.row_ptr[w] + lane; e < A.row_ptr[w + 1]; e += 32) {
      int c = A.col[e], sl = A.slot[e];
      const double* xb = x + 7 * c;
      double xv[7];
#pragma unroll
      for (int k = 0; k < 7; k++) xv[k] = xb[k];
      if (sl >= 0) {
        const double* M = S.Spp + (size_t)sl * 49;
#pragma unroll
        for (int r = 0; r < 7; r++)
#pragma unroll
          for (int k = 0; k < 7; k++) acc[r] += M[7 * r + k] * xv[k];
      } else {
        const double* M = S.Spp + (size_t)(-sl - 1) * 49;
#pragma unroll
        for (int k = 0; k < 7; k++)
#pragma unroll
          for (int r = 0; r < 7; r++) acc[r] += M[7 * k + r] * xv[k];
      }
    }
    // border columns of this block row, and their transpose contribution to the border rows
    double xr[7];
#pragma unroll
    for (int r = 0; r < 7; r++) xr[r] = x[7 * w + r];
    for (int b = lane; b < B; b += 32) {
      double xb = x[S.npc + b], t = 0;
#pragma unroll
      for (int r = 0; r < 7; r++) { double s = S.Spb[(size_t)(7 * w + r) * B + b]; acc[r] += s * xb; t += s * xr[r]; }
      atomicAdd(y + S.npc + b, t);
    }
#pragma unroll
    for (int r = 0; r < 7; r++) acc[r] = warp_sum(acc[r]);
    if (lane == 0) {
#pragma unroll
      for (int r = 0; r < 7; r++) y[7 * w + r] = acc[r];
    }
  } else {
    int b = w - A.nb;  // border row
    if (b >= B) return;
    double t = 0;
    for (int c = lane; c < B; c += 32) t += S.Sbb[(size_t)b * B + c] * x[S.npc + c];
    t = warp_sum(t);
    if (lane == 0) atomicAdd(y + S.npc + b, t);
  }
}

__global__ void __launch_bounds__(256) spmv_kernel(Reduced S, Bsr A, const double* __restrict__ x, double* __restrict__ y,
                                                    const int* __restrict__ done) {
  if (done && *done) return;
  const int lane = threadIdx.x & 31;
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (w < A.nb + S.B) spmv_row(S, A, x, y, w, lane);
}

// --- PCG steps ------------------------------------------------------------------------------
// init: x = 0, r = b, |b|^2 ; border of q zeroed for the first SpMV
__global__ void pcg_init_kernel(int nf, int npc, const double* __restrict__ b, double* __restrict__ x, double* __restrict__ r,
                                double* __restrict__ q, PcgScal sc) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  double v = 0;
  if (i < nf) { x[i] = 0; r[i] = b[i]; v = b[i] * b[i]; if (i >= npc) q[i] = 0; }
  block_accumulate(v, sc.bnorm2);
}

// z = M^-1 r (block Jacobi), rho_it = r.z.  One thread per reduced column.
__global__ void __launch_bounds__(256) pcg_precond_kernel(Reduced S, const double* __restrict__ Mpp, const double* __restrict__ Mbb,
                                                           const double* __restrict__ r, double* __restrict__ z, PcgScal sc, int it) {
  if (*sc.done) return;
  int g = blockIdx.x * blockDim.x + threadIdx.x;
  double v = 0;
  if (g < S.nf) {
    double s = 0;
    if (g < S.npc) {
      int u = g / 7, k = g - 7 * u;
      const double* M = Mpp + (size_t)u * 49 + 7 * k;
      const double* rb = r + 7 * u;
#pragma unroll
      for (int j = 0; j < 7; j++) s += M[j] * rb[j];
    } else {
      int b = g - S.npc;
      for (int j = 0; j < S.B; j++) s += Mbb[(size_t)b * S.B + j] * r[S.npc + j];
    }
    z[g] = s;
    v = s * r[g];
  }
  block_accumulate(v, sc.it + 4 * it);
}

// p = z + beta p   (beta = rho_it / rho_{it-1}); zero the border part of q for the SpMV.
__global__ void pcg_dir_kernel(int nf, int npc, const double* __restrict__ z, double* __restrict__ p, double* __restrict__ q,
                               PcgScal sc, int it) {
  if (*sc.done) return;
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nf) return;
  double rho = sc.it[4 * it];
  if (i == 0 && (rho == 0.0 || !isfinite(rho))) { *sc.done = 2; }
  double beta = (it == 1) ? 0.0 : rho / sc.it[4 * (it - 1)];
  p[i] = z[i] + beta * p[i];
  if (i >= npc) q[i] = 0;
}

// pq_it = p.q
__global__ void pcg_dot_kernel(int nf, const double* __restrict__ p, const double* __restrict__ q, PcgScal sc, int it) {
  if (*sc.done) return;
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  double v = (i < nf) ? p[i] * q[i] : 0.0;
  block_accumulate(v, sc.it + 4 * it + 1);
}

// x += alpha p ; r -= alpha q ; Q1_it = -x.(b + r)
__global__ void pcg_update_kernel(int nf, const double* __restrict__ b, const double* __restrict__ p, const double* __restrict__ q,
                                  double* __restrict__ x, double* __restrict__ r, PcgScal sc, int it) {
  if (*sc.done) return;
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  double pq = sc.it[4 * it + 1], rho = sc.it[4 * it];
  double v = 0;
  if (pq <= 0.0 || !isfinite(pq)) {  // "matrix is indefinite": stop, keep the last iterate
    if (i == 0) *sc.done = 3;
    return;
  }
  double alpha = rho / pq;
  if (i < nf) {
    double xi = x[i] + alpha * p[i];
    double ri = r[i] - alpha * q[i];
    x[i] = xi; r[i] = ri;
    v = -xi * (b[i] + ri);
  }
  block_accumulate(v, sc.it + 4 * it + 2);
}

// r = b - q (q = S x), Q1 recomputed (residual_reset_period = 10)
__global__ void pcg_refresh_kernel(int nf, const double* __restrict__ b, const double* __restrict__ q, const double* __restrict__ x,
                                   double* __restrict__ r, PcgScal sc, int it) {
  if (*sc.done) return;
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  double v = 0;
  if (i < nf) { double ri = b[i] - q[i]; r[i] = ri; v = -x[i] * (b[i] + ri); }
  block_accumulate(v, sc.it + 4 * it + 3);
}

// Single thread: Ceres' Q-test  zeta = it (Q1 - Q0) / Q1 < eta.
__global__ void pcg_check_kernel(PcgScal sc, int it, int refreshed, double eta, int min_it, int max_it) {
  if (*sc.done) return;
  double Q1 = refreshed ? sc.it[4 * it + 3] : sc.it[4 * it + 2];
  if (refreshed) sc.it[4 * it + 2] = Q1;
  double Q0 = (it == 1) ? 0.0 : sc.it[4 * (it - 1) + 2];
  double zeta = it * (Q1 - Q0) / Q1;
  *sc.iters = it;
  if (zeta < eta && it >= min_it) *sc.done = 1;
  else if (it >= max_it) *sc.done = 4;
}

// --- the whole PCG solve in ONE cooperative launch -------------------------------------------
// Same arithmetic and stopping rules as the step kernels above (they remain as the fallback when the device cannot
// co-schedule the grid), with grid-wide barriers instead of kernel boundaries and no host round trip until the solve is
// over.  THREE barriers per iteration:
//   P  per parameter block (8 lanes per pose block, one warp for the border): x += alpha p, r -= alpha q of the PREVIOUS
//      iteration, its Q, then z = M^-1 r (block local) and rho = r.z                                   | barrier
//      Ceres' Q test of the previous iteration, beta;  p = z + beta p                                  | barrier
//   S  q = S p fused with p.q (the border rows of q are completed with atomics, so their share of p.q is taken from the
//      terms as they are produced: p_b' (S_pb' p_u) per pose row, p_b (S_bb p_b)_b per border row)       | barrier
// The reduced system (a few MB) stays L2 resident throughout.

// Row w of S x as spmv_row, returning this lane's share of x'(S x).
// (z != null: x is the PREVIOUS direction and the vector multiplied is p = z + beta x, formed on the fly by every reader with the
// same fused multiply-add; the owner of a row also stores its part of p to pn.  This folds Ceres' direction update into the
// product and saves one grid barrier per CG iteration without changing a single rounding.)
__device__ __forceinline__ double pcg_dir(const double* z, const double* x, double beta, int i) { return z ? fma(beta, x[i], z[i]) : x[i]; }
__device__ __forceinline__ double spmv_row_dot(const Reduced& S, const Bsr& A, const double* x, double* y, int w, int lane,
                                               const double* z = nullptr, double beta = 0.0, double* pn = nullptr) {
  const int B = S.B;
  double dot = 0;
  if (w < A.nb) {
    double acc[7] = {0, 0, 0, 0, 0, 0, 0};
    for (int e = A.row_ptr[w] + lane; e < A.row_ptr[w + 1]; e += 32) {
      const int c = A.col[e], sl = A.slot[e];
      double xv[7];
#pragma unroll
      for (int k = 0; k < 7; k++) xv[k] = pcg_dir(z, x, beta, 7 * c + k);
      if (sl >= 0) {
        const double* M = S.Spp + (size_t)sl * 49;
#pragma unroll
        for (int r = 0; r < 7; r++)
#pragma unroll
          for (int k = 0; k < 7; k++) acc[r] += M[7 * r + k] * xv[k];
      } else {
        const double* M = S.Spp + (size_t)(-sl - 1) * 49;
#pragma unroll
        for (int k = 0; k < 7; k++)
#pragma unroll
          for (int r = 0; r < 7; r++) acc[r] += M[7 * k + r] * xv[k];
      }
    }
    double xr[7];
#pragma unroll
    for (int r = 0; r < 7; r++) xr[r] = pcg_dir(z, x, beta, 7 * w + r);
    for (int b = lane; b < B; b += 32) {
      const double xb = pcg_dir(z, x, beta, S.npc + b);
      double t = 0;
#pragma unroll
      for (int r = 0; r < 7; r++) { const double s = S.Spb[(size_t)(7 * w + r) * B + b]; acc[r] += s * xb; t += s * xr[r]; }
      atomicAdd(y + S.npc + b, t);
      dot += xb * t;                      // x_b (S_pb' x_u)_b
    }
#pragma unroll
    for (int r = 0; r < 7; r++) acc[r] = warp_sum(acc[r]);
    if (lane == 0) {
#pragma unroll
      for (int r = 0; r < 7; r++) { y[7 * w + r] = acc[r]; dot += xr[r] * acc[r]; if (pn) pn[7 * w + r] = xr[r]; }
    }
  } else {
    const int b = w - A.nb;
    if (b < B) {
      double t = 0;
      for (int c = lane; c < B; c += 32) t += S.Sbb[(size_t)b * B + c] * pcg_dir(z, x, beta, S.npc + c);
      t = warp_sum(t);
      if (lane == 0) { const double xo = pcg_dir(z, x, beta, S.npc + b); atomicAdd(y + S.npc + b, t); dot += xo * t; if (pn) pn[S.npc + b] = xo; }
    }
  }
  return dot;
}


// Row w of S x with the row's blocks RESIDENT in shared memory (fused PCG: every CTA keeps the blocks of its block rows, and
// their share of S_pb, for the whole solve; transposed mirror blocks are stored transposed, so the arithmetic of a row is
// the one of spmv_row_dot, term for term).  sm_pp: blocks of entries [e0, ...), 49 doubles each; sm_pb: rows 7 r0 ... of S_pb.
template <bool DOT>
__device__ __forceinline__ double spmv_row_sm(const Reduced& S, const Bsr& A, const double* sm_pp, const double* sm_pb, int e0, int r0,
                                              const double* x, double* y, int w, int lane,
                                              const double* z = nullptr, double beta = 0.0, double* pn = nullptr, double* qb_s = nullptr) {
  const int B = S.B;
  double dot = 0;
  double acc[7] = {0, 0, 0, 0, 0, 0, 0};
  for (int e = A.row_ptr[w] + lane; e < A.row_ptr[w + 1]; e += 32) {
    const int c = A.col[e];
    double xv[7];
#pragma unroll
    for (int k = 0; k < 7; k++) xv[k] = pcg_dir(z, x, beta, 7 * c + k);
    const double* M = sm_pp + (size_t)(e - e0) * 49;
#pragma unroll
    for (int r = 0; r < 7; r++)
#pragma unroll
      for (int k = 0; k < 7; k++) acc[r] += M[7 * r + k] * xv[k];
  }
  double xr[7];
#pragma unroll
  for (int r = 0; r < 7; r++) xr[r] = pcg_dir(z, x, beta, 7 * w + r);
  for (int b = lane; b < B; b += 32) {
    const double xb = pcg_dir(z, x, beta, S.npc + b);
    double t = 0;
#pragma unroll
    for (int r = 0; r < 7; r++) { const double s = sm_pb[(size_t)(7 * (w - r0) + r) * B + b]; acc[r] += s * xb; t += s * xr[r]; }
    // the border rows of y collect a term from EVERY pose row: summed per CTA in shared memory first (qb_s), one global atomic per
    // CTA and border row afterwards -- 2000 atomics on the same 14 addresses per product were 6 us of every CG iteration
    if (qb_s) atomicAdd(qb_s + b, t); else atomicAdd(y + S.npc + b, t);
    if (DOT) dot += xb * t;
  }
#pragma unroll
  for (int r = 0; r < 7; r++) acc[r] = warp_sum(acc[r]);
  if (lane == 0) {
#pragma unroll
    for (int r = 0; r < 7; r++) { y[7 * w + r] = acc[r]; if (DOT) dot += xr[r] * acc[r]; if (pn) pn[7 * w + r] = xr[r]; }
  }
  return dot;
}

#define MVBA_PCG_QB 192  // border columns whose per-CTA sums fit the shared array (more: global atomics per row, as before)

__global__ void __launch_bounds__(512) pcg_fused_kernel(Reduced S, Bsr A, const double* Mpp, const double* Mbb, const double* b, double* x,
                                                         double* r, double* z, double* p, double* q, PcgScal sc, double eta, int min_it,
                                                         int max_it, int rows_per_cta, double* p2) {
  namespace cg = cooperative_groups;
  cg::grid_group grid = cg::this_grid();
  const int nf = S.nf, npc = S.npc, B = S.B, nb = S.nb;
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x, nthr = gridDim.x * blockDim.x;
  const int lane = threadIdx.x & 31, gwarp = gtid >> 5, nwarp = nthr >> 5;
  const int sub = lane & 7, gbase = lane & ~7;
  // rows_per_cta > 0: this CTA owns block rows [r0, r1) and keeps their blocks in shared memory for the whole solve
  extern __shared__ __align__(16) double pcg_sm[];
  const int wib = threadIdx.x >> 5, nwib = blockDim.x >> 5;
  const int r0 = rows_per_cta > 0 ? min(nb, (int)blockIdx.x * rows_per_cta) : 0;
  const int r1 = rows_per_cta > 0 ? min(nb, r0 + rows_per_cta) : 0;
  const int e0 = rows_per_cta > 0 ? A.row_ptr[r0] : 0, e1 = rows_per_cta > 0 ? A.row_ptr[r1] : 0;
  const double* sm_pp = pcg_sm;
  const double* sm_pb = pcg_sm + (size_t)(e1 - e0) * 49;
  if (rows_per_cta > 0) {
    for (int i = threadIdx.x; i < (e1 - e0) * 49; i += blockDim.x) {
      const int e = e0 + i / 49, rc = i % 49, rr = rc / 7, kk = rc % 7;
      const int sl = A.slot[e];
      pcg_sm[i] = sl >= 0 ? S.Spp[(size_t)sl * 49 + rc] : S.Spp[(size_t)(-sl - 1) * 49 + 7 * kk + rr];
    }
    double* pb = pcg_sm + (size_t)(e1 - e0) * 49;
    for (int i = threadIdx.x; i < 7 * (r1 - r0) * B; i += blockDim.x) pb[i] = S.Spb[(size_t)7 * r0 * B + i];
  }
  // per-CTA sums of the border rows of a product (spmv_row_sm), flushed with one atomic per CTA and row
  __shared__ double qb_sm[MVBA_PCG_QB];
  double* qb_s = (rows_per_cta > 0 && B <= MVBA_PCG_QB) ? qb_sm : nullptr;
  if (qb_s) for (int i = threadIdx.x; i < B; i += blockDim.x) qb_s[i] = 0.0;
  // two direction buffers: po holds the current p, pn receives the next one inside the product (no separate update pass)
  double* po = p; double* pn = p2;
  for (int i = gtid; i < nf; i += nthr) { x[i] = 0; r[i] = b[i]; po[i] = 0; pn[i] = 0; if (i >= npc) q[i] = 0; }
  grid.sync();
  int done = 0, it_done = 0;
  double Q0 = 0.0, alpha = 0.0;
  bool pending = false;  // the x, r update of the previous iteration has not been applied yet
  for (int it = 1; it <= max_it + 1; it++) {
    // ---- P: pending update of iteration it - 1, its Q; z = M^-1 r and rho of iteration it ----
    double vq = 0, vrho = 0;
    for (int u0 = 4 * gwarp; u0 < nb; u0 += 4 * nwarp) {
      const int u = u0 + (lane >> 3);
      const bool act = u < nb && sub < 7;
      const int i = act ? 7 * u + sub : 0;
      double ri = 0;
      if (act) {
        ri = r[i];
        if (pending) {
          const double xi = x[i] + alpha * po[i];
          ri -= alpha * q[i];
          x[i] = xi; r[i] = ri;
          vq += -xi * (b[i] + ri);
        }
      }
      double zi = 0;
#pragma unroll
      for (int j = 0; j < 7; j++) {
        const double rj = __shfl_sync(0xffffffffu, ri, gbase + j);
        if (act) zi += Mpp[(size_t)u * 49 + 7 * sub + j] * rj;
      }
      if (act) { z[i] = zi; vrho += zi * ri; }
    }
    if (gwarp == nwarp - 1 && B > 0) {  // the border block: one warp
      for (int bb = lane; bb < B; bb += 32) {
        const int i = npc + bb;
        if (pending) {
          const double xi = x[i] + alpha * po[i], ri = r[i] - alpha * q[i];
          x[i] = xi; r[i] = ri;
          vq += -xi * (b[i] + ri);
        }
        q[i] = 0;  // the border rows of q are completed with atomics in the product below
      }
      __syncwarp();
      for (int bb = lane; bb < B; bb += 32) {
        double s = 0;
        for (int j = 0; j < B; j++) s += Mbb[(size_t)bb * B + j] * r[npc + j];
        z[npc + bb] = s;
        vrho += s * r[npc + bb];
      }
    }
    if (pending) block_accumulate(vq, sc.it + 4 * (it - 1) + 2);
    if (it <= max_it) block_accumulate(vrho, sc.it + 4 * it);
    pending = false;
    grid.sync();
    if (it > 1) {  // Ceres' Q test of the previous iteration: zeta = it (Q1 - Q0) / Q1 < eta
      const int pit = it - 1;
      const double Q1 = sc.it[4 * pit + 2];
      const double zeta = pit * (Q1 - Q0) / Q1;
      Q0 = Q1;
      it_done = pit;
      if (zeta < eta && pit >= min_it) { done = 1; break; }
      if (pit >= max_it) { done = 4; break; }
    }
    const double rho = sc.it[4 * it];
    if (rho == 0.0 || !isfinite(rho)) { done = 2; break; }
    const double beta = (it == 1) ? 0.0 : rho / sc.it[4 * (it - 1)];
    // ---- S: p = z + beta p (formed by every reader, stored by the row's owner), q = S p and p.q ----
    {
      double v = 0;
      if (rows_per_cta > 0) {
        for (int w = r0 + wib; w < r1; w += nwib) v += spmv_row_sm<true>(S, A, sm_pp, sm_pb, e0, r0, po, q, w, lane, z, beta, pn, qb_s);
        if (blockIdx.x == gridDim.x - 1) for (int w = nb + wib; w < nb + B; w += nwib) v += spmv_row_dot(S, A, po, q, w, lane, z, beta, pn);
      } else {
        for (int w = gwarp; w < nb + B; w += nwarp) v += spmv_row_dot(S, A, po, q, w, lane, z, beta, pn);
      }
      block_accumulate(v, sc.it + 4 * it + 1);  // (its barriers also order the shared-memory sums below)
      if (qb_s) for (int i = threadIdx.x; i < B; i += blockDim.x) { const double t = qb_s[i]; qb_s[i] = 0.0; if (t != 0.0) atomicAdd(q + npc + i, t); }
    }
    grid.sync();
    { double* t = po; po = pn; pn = t; }
    const double pq = sc.it[4 * it + 1];
    if (pq <= 0.0 || !isfinite(pq)) { done = 3; break; }  // "matrix is indefinite": keep the last iterate
    alpha = rho / pq;
    pending = true;
    if (it % 10 == 0) {  // residual_reset_period: x updated now, r = b - S x, Q recomputed from it
      for (int i = gtid; i < nf; i += nthr) { x[i] += alpha * po[i]; if (i >= npc) q[i] = 0; }
      grid.sync();
      if (rows_per_cta > 0) {
        for (int w = r0 + wib; w < r1; w += nwib) spmv_row_sm<false>(S, A, sm_pp, sm_pb, e0, r0, x, q, w, lane, nullptr, 0.0, nullptr, qb_s);
        if (qb_s) {
          __syncthreads();
          for (int i = threadIdx.x; i < B; i += blockDim.x) { const double t = qb_s[i]; qb_s[i] = 0.0; if (t != 0.0) atomicAdd(q + npc + i, t); }
        }
        if (blockIdx.x == gridDim.x - 1) for (int w = nb + wib; w < nb + B; w += nwib) spmv_row(S, A, x, q, w, lane);
      } else {
        for (int w = gwarp; w < nb + B; w += nwarp) spmv_row(S, A, x, q, w, lane);
      }
      grid.sync();
      double v = 0;
      for (int i = gtid; i < nf; i += nthr) { const double ri = b[i] - q[i]; r[i] = ri; v += -x[i] * (b[i] + ri); if (i >= npc) q[i] = 0; }
      block_accumulate(v, sc.it + 4 * it + 2);
      pending = false;
      grid.sync();
    }
  }
  if (gtid == 0) { *sc.done = done; *sc.iters = it_done; }
}

}  // namespace mvba
