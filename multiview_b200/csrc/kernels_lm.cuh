// kernels_lm.cuh -- landmark back-substitution, model cost change, parameter update and the
// Levenberg-Marquardt accept/reject + trust-region radius update, all on device
// (Ceres 1.14 TrustRegionMinimizer / LevenbergMarquardtStrategy semantics).
#pragma once
#include "kernels_pcg.cuh"
#include "kernels_backsub.cuh"

namespace mvba {

// Device-resident LM state (doubles) -- index constants.
enum {
  LM_COST = 0,        // cost at x
  LM_CAND = 1,        // cost at x + step                      (rank-local partial, slots 1..4 are
  LM_MCC = 2,         // model cost change                      summed across ranks in one allreduce)
  LM_STEP2E = 3,      // |step|^2, landmark part
  LM_XNORM2E = 4,     // |x|^2, landmark part
  LM_RADIUS = 5,
  LM_DECREASE = 6,    // decrease factor
  LM_GMAX = 7,        // gradient max norm (as double)
  LM_RELDEC = 8,
  LM_ACCEPT = 9,      // 1.0 when the step was accepted
  LM_TERM = 10,       // 0 continue, 1 parameter tol, 2 function tol, 3 gradient tol, 4 min radius, 5 too many invalid steps
  LM_STEP2 = 11,      // |step|^2, replicated (non-landmark) part
  LM_XNORM2 = 12,     // |x|^2, replicated part
  LM_INVALID = 13,    // consecutive invalid steps (model cost change not positive)
  LM_NSLOTS = 16
};

// Jacobi scaling from squared column norms: 1 / (1 + sqrt(n2)).
__global__ void jscale_kernel(int64_t n, const double* __restrict__ n2, double* __restrict__ sc) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) sc[i] = 1.0 / (1.0 + sqrt(n2[i]));
}

// max |g_i / scale_i| over the f part and the landmark part -> *out (bit pattern max, out zeroed before).  `extra`: values
// that enter the maximum as they are (multi-GPU: the landmark-part maxima of all ranks, which travel in the summed buffer).
__global__ void gradmax_kernel(int nf, int64_t ne, const double* __restrict__ g, const double* __restrict__ ge,
                               const double* __restrict__ sf, const double* __restrict__ se, const double* __restrict__ extra, int nextra,
                               double* __restrict__ out) {
  double v = 0;
  // grid-stride (the launch bounds the grid): the atomics of all CTAs land on ONE address, and thousands of them serialise
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nf + ne + nextra; i += (int64_t)gridDim.x * blockDim.x) {
    if (i < nf) v = fmax(v, fabs(g[i] / (sf ? sf[i] : 1.0)));
    else if (i < nf + ne) { int64_t k = i - nf; v = fmax(v, fabs(ge[k] / (se ? se[k] : 1.0))); }
    else v = fmax(v, extra[i - nf - ne]);
  }
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  __shared__ double wmax[32];  // one atomic per CTA: a per-warp atomic on ONE address serialises ~100 k updates at 3 M landmarks
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) wmax[w] = v;
  __syncthreads();
  if (w == 0) {
    v = lane < (int)(blockDim.x >> 5) ? wmax[lane] : 0.0;
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    if (lane == 0 && v > 0) atomicMax(reinterpret_cast<unsigned long long*>(out), (unsigned long long)__double_as_longlong(v));
  }
}

// x_new = x - scale * y  over every parameter scalar; accumulates |step|^2 and |x|^2 (free only).
__global__ void __launch_bounds__(256) update_kernel(int nx, int nf, const int32_t* __restrict__ xcol, const double* __restrict__ x,
                                                      const double* __restrict__ yf, const double* __restrict__ ye,
                                                      const double* __restrict__ sf, const double* __restrict__ se,
                                                      const double* __restrict__ lo, const double* __restrict__ hi,
                                                      double* __restrict__ xnew, double* __restrict__ lm) {
  // grid-stride (the launch bounds the grid: every CTA ends in four atomics on the same four addresses); the two parts by parameter kind
  double part[4] = {0.0, 0.0, 0.0, 0.0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nx; i += gridDim.x * blockDim.x) {
    const int c = xcol[i];
    double xi = x[i], xn = xi;
    if (c >= 0) {
      double d = (c < nf) ? -yf[c] * (sf ? sf[c] : 1.0) : -ye[c - nf] * (se ? se[c - nf] : 1.0);
      xn = xi + d;
      if (lo) xn = fmin(fmax(xn, lo[i]), hi[i]);
      d = xn - xi;
      if (c >= nf) { part[2] += d * d; part[3] += xi * xi; } else { part[0] += d * d; part[1] += xi * xi; }
    }
    xnew[i] = xn;
  }
  __shared__ double sm4[4][8];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 4; k++) { part[k] = warp_sum(part[k]); if (lane == 0) sm4[k][w] = part[k]; }
  __syncthreads();
  if (threadIdx.x < 32) {
    const int k = lane >> 3, j = lane & 7;  // 4 sums x 8 warps
    double t = sm4[k][j];
    t += __shfl_down_sync(0xffffffffu, t, 4, 8); t += __shfl_down_sync(0xffffffffu, t, 2, 8); t += __shfl_down_sync(0xffffffffu, t, 1, 8);
    const int slot = k == 0 ? LM_STEP2 : k == 1 ? LM_XNORM2 : k == 2 ? LM_STEP2E : LM_XNORM2E;
    if (j == 0 && t != 0.0) atomicAdd(lm + slot, t);
  }
}

struct LmOpts { double min_rel_dec, param_tol, func_tol, max_radius, min_radius, grad_tol; };

// Single thread: accept/reject and radius update (LevenbergMarquardtStrategy::StepAccepted/Rejected).
__global__ void lm_decide_kernel(double* __restrict__ lm, LmOpts o, mvba_iteration* __restrict__ rec, int cg_iters_dummy) {
  double cost = lm[LM_COST], cand = lm[LM_CAND], mcc = lm[LM_MCC];
  double step = sqrt(lm[LM_STEP2] + lm[LM_STEP2E]), xn = sqrt(lm[LM_XNORM2] + lm[LM_XNORM2E]);
  double radius = lm[LM_RADIUS], dec = lm[LM_DECREASE];
  int accept = 0, term = 0;
  double change = 0, rel = 0;
  double invalid = lm[LM_INVALID];
  if (lm[LM_GMAX] <= o.grad_tol) {
    // the gradient at x (this iteration's build) is within tolerance: Ceres stops right after the step that led here
    // (TrustRegionMinimizer::HandleSuccessfulStep); the step computed meanwhile is dropped, x stays
    term = 3;
  } else if (!(mcc > 0) || !isfinite(mcc)) {
    radius /= dec; dec *= 2;
    invalid += 1.0;
    if (invalid >= 5.0) term = 5;  // max_num_consecutive_invalid_steps (Ceres default)
  } else {
    invalid = 0.0;
    if (!isfinite(cand)) cand = 1.7976931348623157e308;
    change = cost - cand;
    rel = change / mcc;
    if (step <= o.param_tol * (xn + o.param_tol)) term = 1;
    else if (fabs(change) <= o.func_tol * cost) term = 2;
    if (rel > o.min_rel_dec) {
      accept = 1;
      cost = cand;
      double t = 2.0 * rel - 1.0;
      radius = radius / fmax(1.0 / 3.0, 1.0 - t * t * t);
      radius = fmin(o.max_radius, radius);
      dec = 2.0;
    } else {
      radius /= dec; dec *= 2;
    }
  }
  if (radius < o.min_radius && term == 0) term = 4;
  lm[LM_COST] = cost; lm[LM_RADIUS] = radius; lm[LM_DECREASE] = dec; lm[LM_RELDEC] = rel;
  lm[LM_ACCEPT] = accept; lm[LM_TERM] = term; lm[LM_INVALID] = invalid;
  rec->cost = cost; rec->cost_change = change; rec->step_norm = step; rec->relative_decrease = rel; rec->radius = radius;
  rec->step_accepted = accept; rec->gradient_max_norm = lm[LM_GMAX]; rec->linear_iterations = term;
}

// x <- x_new when the step was accepted.
__global__ void commit_kernel(int nx, const double* __restrict__ lm, const double* __restrict__ xnew, double* __restrict__ x) {
  if (lm[LM_ACCEPT] == 0.0) return;
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nx) x[i] = xnew[i];
}

}  // namespace mvba
