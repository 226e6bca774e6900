// kernels_lm.cuh -- landmark back-substitution, model cost change, parameter update and the
// Levenberg-Marquardt accept/reject + trust-region radius update, all on device
// (Ceres 1.14 TrustRegionMinimizer / LevenbergMarquardtStrategy semantics).
#pragma once
#include "kernels_pcg.cuh"

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
  LM_TERM = 10,       // 0 continue, 1 parameter tol, 2 function tol, 3 gradient tol, 4 min radius
  LM_STEP2 = 11,      // |step|^2, replicated (non-landmark) part
  LM_XNORM2 = 12,     // |x|^2, replicated part
  LM_NSLOTS = 16
};

// Jacobi scaling from squared column norms: 1 / (1 + sqrt(n2)).
__global__ void jscale_kernel(int64_t n, const double* __restrict__ n2, double* __restrict__ sc) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) sc[i] = 1.0 / (1.0 + sqrt(n2[i]));
}

// max |g_i / scale_i| over the f part and the landmark part -> LM_GMAX (bit pattern max).
__global__ void gradmax_kernel(int nf, int64_t ne, const double* __restrict__ g, const double* __restrict__ ge,
                               const double* __restrict__ sf, const double* __restrict__ se, double* __restrict__ lm) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  double v = 0;
  if (i < nf) v = fabs(g[i] / (sf ? sf[i] : 1.0));
  else if (i < nf + ne) { int64_t k = i - nf; v = fabs(ge[k] / (se ? se[k] : 1.0)); }
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  if ((threadIdx.x & 31) == 0 && v > 0) atomicMax(reinterpret_cast<unsigned long long*>(lm + LM_GMAX), (unsigned long long)__double_as_longlong(v));
}

// Back-substitution (one warp per track): y_e = Hinv (E'r - E'F y_f) and the model cost change
//   -(J step)'(r + J step / 2),  step = -y.  Jacobians are recomputed, not read.
struct BacksubArgs {
  const double* x; const CamTab* tab; const DepthTab* dtab;
  const double *jscale_f, *jscale_e;
  const double *Hinv, *ge, *yf;
  double* ye;      // 3 per landmark (scaled variables)
  double* lm;      // LM_MCC accumulates
  double* fy;      // scratch: 3 doubles per item, per warp region of size maxitems
  int maxitems, maxdz;
};

__global__ void __launch_bounds__(256) backsub_kernel(View v, BacksubArgs a) {
  extern __shared__ double smem_d[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  // per warp: stash[CHUNK] | Jd[CHUNK][2][maxdz] | fy[maxitems][3] | rr[maxitems][3] | Es[maxitems][9]
  const size_t per_warp = sizeof(Stash) * MVBA_CHUNK / 8 + (size_t)MVBA_CHUNK * 2 * a.maxdz + (size_t)a.maxitems * 15;
  double* base = smem_d + per_warp * wib;
  Stash* stash = reinterpret_cast<Stash*>(base);
  double* Jd = base + sizeof(Stash) * MVBA_CHUNK / 8;
  double* fy = Jd + (size_t)MVBA_CHUNK * 2 * a.maxdz;
  double* rr = fy + 3 * (size_t)a.maxitems;
  double* Es = rr + 3 * (size_t)a.maxitems;
  double mcc = 0;
  for (int p = blockIdx.x * wpb + wib; p < v.P; p += gridDim.x * wpb) {
    const int64_t o0 = v.trk_obs0[p], o1 = v.trk_obs0[p + 1];
    const int64_t d0 = v.trk_dtri0[p], d1 = v.trk_dtri0[p + 1];
    const int64_t m0 = v.trk_dmesh0[p], m1 = v.trk_dmesh0[p + 1];
    const int64_t pm0 = v.trk_pm0[p], pm1 = v.trk_pm0[p + 1];
    const int64_t pt0 = v.trk_pt0[p], pt1 = v.trk_pt0[p + 1];
    const int n_obs = (int)(o1 - o0), n_dt = (int)(d1 - d0), n_dm = (int)(m1 - m0), n_pm = (int)(pm1 - pm0), n_pt = (int)(pt1 - pt0);
    const int nitems = n_obs + n_dt + n_dm + n_pm + n_pt;
    if (nitems == 0) continue;
    double se[3] = {1, 1, 1};
    if (a.jscale_e) { se[0] = a.jscale_e[3 * (size_t)p]; se[1] = a.jscale_e[3 * (size_t)p + 1]; se[2] = a.jscale_e[3 * (size_t)p + 2]; }
    double u[3] = {0, 0, 0};  // E' F y_f (lane 0 holds the total after each item)
    for (int c0 = 0; c0 < nitems; c0 += MVBA_CHUNK) {
      const int cnt = min(MVBA_CHUNK, nitems - c0);
      if (lane < cnt) {
        int q = c0 + lane;
        Item it;
        double* jd0 = Jd + (size_t)lane * 2 * a.maxdz;
        if (q < n_obs) eval_cam<true>(v, a.x, a.tab, o0 + q, it, jd0, jd0 + a.maxdz);
        else if (q < n_obs + n_dt) eval_depth(v, a.x, a.tab, a.dtab, MVBA_BLOCK_DEPTH_TRI, d0 + (q - n_obs), it);
        else if (q < n_obs + n_dt + n_dm) eval_depth(v, a.x, a.tab, a.dtab, MVBA_BLOCK_DEPTH_MESH, m0 + (q - n_obs - n_dt), it);
        else if (q < n_obs + n_dt + n_dm + n_pm) eval_prior(v, a.x, pm0 + (q - n_obs - n_dt - n_dm), it);
        else eval_prior(v, a.x, pt0 + (q - n_obs - n_dt - n_dm - n_pm), it);
        double s = 0;
        for (int k = 0; k < it.nr; k++) s += it.r[k] * it.r[k];
        double hr, sr;
        cauchy(it.loss_a, s, &hr, &sr);
        Stash& st = stash[lane];
        st.kind = it.kind; st.nr = it.nr; st.cid = it.cid; st.sensor = it.sensor; st.sr = sr; st.Jc = it.Jc;
#pragma unroll
        for (int k = 0; k < 3; k++) { st.r[k] = k < it.nr ? sr * it.r[k] : 0.0; st.Z[k] = it.Z[k]; st.Xc[k] = it.Xc[k]; st.Jf[k] = it.Jf[k]; st.m[k] = it.m[k]; }
#pragma unroll
        for (int k = 0; k < 9; k++) { st.A[k] = it.A[k]; st.E[k] = (k / 3 < it.nr) ? sr * it.E[k] * se[k % 3] : 0.0; }
#pragma unroll
        for (int k = 0; k < 3; k++) rr[3 * q + k] = st.r[k];
#pragma unroll
        for (int k = 0; k < 9; k++) Es[9 * q + k] = st.E[k];
      }
      __syncwarp();
      for (int j = 0; j < cnt; j++) {
        const Stash& st = stash[j];
        double f[3] = {0, 0, 0};
        if (st.kind != MVBA_BLOCK_MESH_TRI && st.kind != MVBA_BLOCK_TRI) {
          Item it;
          it.kind = st.kind; it.nr = st.nr; it.cid = st.cid; it.sensor = st.sensor; it.Jc = st.Jc;
#pragma unroll
          for (int k = 0; k < 3; k++) { it.Z[k] = st.Z[k]; it.Xc[k] = st.Xc[k]; it.Jf[k] = st.Jf[k]; it.m[k] = st.m[k]; }
#pragma unroll
          for (int k = 0; k < 9; k++) it.A[k] = st.A[k];
          const int vw = item_vwidth(v, it);
          const double* jd0 = Jd + (size_t)j * 2 * a.maxdz;
          for (int vc = lane; vc < vw; vc += 32) {
            int gc = item_gcol(v, it, vc);
            if (gc < 0) continue;
            double col[3] = {0, 0, 0};
            item_jcol(v, it, a.tab, a.dtab, vc, jd0, jd0 + a.maxdz, col);
            double sc = st.sr * (a.jscale_f ? a.jscale_f[gc] : 1.0) * a.yf[gc];
            for (int r = 0; r < st.nr; r++) f[r] += col[r] * sc;
          }
#pragma unroll
          for (int k = 0; k < 3; k++) f[k] = warp_sum(f[k]);
        }
        if (lane == 0) {
          int q = c0 + j;
          for (int k = 0; k < 3; k++) fy[3 * q + k] = f[k];
          for (int r = 0; r < st.nr; r++) for (int b = 0; b < 3; b++) u[b] += st.E[3 * r + b] * f[r];
        }
      }
      __syncwarp();
    }
    // lane 0 finishes the landmark
    if (lane == 0) {
      const double* Hi = a.Hinv + 6 * (size_t)p;
      const double* g = a.ge + 3 * (size_t)p;
      double w0 = g[0] - u[0], w1 = g[1] - u[1], w2 = g[2] - u[2];
      double ye[3] = {Hi[0] * w0 + Hi[1] * w1 + Hi[2] * w2, Hi[1] * w0 + Hi[3] * w1 + Hi[4] * w2, Hi[2] * w0 + Hi[4] * w1 + Hi[5] * w2};
      for (int k = 0; k < 3; k++) a.ye[3 * (size_t)p + k] = ye[k];
      for (int q = 0; q < nitems; q++) {
        for (int r = 0; r < 3; r++) {
          double jy = fy[3 * q + r] + Es[9 * q + 3 * r] * ye[0] + Es[9 * q + 3 * r + 1] * ye[1] + Es[9 * q + 3 * r + 2] * ye[2];
          double ms = -jy;  // J step
          mcc -= ms * (rr[3 * q + r] + 0.5 * ms);
        }
      }
    }
    __syncwarp();
  }
  mcc = warp_sum(mcc);
  if (lane == 0) atomicAdd(a.lm + LM_MCC, mcc);
}

// x_new = x - scale * y  over every parameter scalar; accumulates |step|^2 and |x|^2 (free only).
__global__ void __launch_bounds__(256) update_kernel(int nx, int nf, const int32_t* __restrict__ xcol, const double* __restrict__ x,
                                                      const double* __restrict__ yf, const double* __restrict__ ye,
                                                      const double* __restrict__ sf, const double* __restrict__ se,
                                                      const double* __restrict__ lo, const double* __restrict__ hi,
                                                      double* __restrict__ xnew, double* __restrict__ lm) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  double s2 = 0, x2 = 0;
  if (i < nx) {
    int c = xcol[i];
    double xi = x[i], xn = xi;
    if (c >= 0) {
      double d = (c < nf) ? -yf[c] * (sf ? sf[c] : 1.0) : -ye[c - nf] * (se ? se[c - nf] : 1.0);
      xn = xi + d;
      if (lo) xn = fmin(fmax(xn, lo[i]), hi[i]);
      d = xn - xi;
      s2 = d * d; x2 = xi * xi;
    }
    xnew[i] = xn;
  }
  // blocks are homogeneous enough: split the two parts by parameter kind
  const bool is_e = (i < nx) && xcol[i] >= nf;
  block_accumulate(is_e ? 0.0 : s2, lm + LM_STEP2);
  block_accumulate(is_e ? 0.0 : x2, lm + LM_XNORM2);
  block_accumulate(is_e ? s2 : 0.0, lm + LM_STEP2E);
  block_accumulate(is_e ? x2 : 0.0, lm + LM_XNORM2E);
}

struct LmOpts { double min_rel_dec, param_tol, func_tol, max_radius, min_radius; };

// Single thread: accept/reject and radius update (LevenbergMarquardtStrategy::StepAccepted/Rejected).
__global__ void lm_decide_kernel(double* __restrict__ lm, LmOpts o, mvba_iteration* __restrict__ rec, int cg_iters_dummy) {
  double cost = lm[LM_COST], cand = lm[LM_CAND], mcc = lm[LM_MCC];
  double step = sqrt(lm[LM_STEP2] + lm[LM_STEP2E]), xn = sqrt(lm[LM_XNORM2] + lm[LM_XNORM2E]);
  double radius = lm[LM_RADIUS], dec = lm[LM_DECREASE];
  int accept = 0, term = 0;
  double change = 0, rel = 0;
  if (!(mcc > 0) || !isfinite(mcc)) {
    radius /= dec; dec *= 2;
  } else {
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
  lm[LM_ACCEPT] = accept; lm[LM_TERM] = term;
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
