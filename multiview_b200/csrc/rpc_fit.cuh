// rpc_fit.cuh -- SURVEY.md 8(f) row N4: the RPC (ratio of polynomials) fit that rig_calibrator runs between passes
// whenever RPC distortion floats, and that fit_rpc runs once per sensor.
//   RpcFitError / fitCurrDegRPC          camera_model/rpc_distortion.cc:366-420, 556-617
//   fitRpcDist / fitRpcUndist            rpc_distortion.cc:619-728 (degree continuation from the identity model)
//   init_as_identity / increment_degree  rpc_distortion.cc:300-358
// The reference differentiates numerically (DynamicNumericDiffCostFunction, central differences, Ceres step rule
// delta = max(|c| 1e-6, sqrt(eps))).  For a ratio N / D the central difference has a closed form, which is what the kernel
// evaluates so that the trajectory of the fit follows the reference's and not that of the exact derivative:
//     d/d num_k :  m_k / D                              (N is linear in the coefficient: the difference quotient is exact)
//     d/d den_k : -N m_k / (D^2 - (delta_k m_k)^2)      (-> -N m_k / D^2 as delta -> 0; jacobian_mode 1 uses delta = 0)
// x and y residuals share no coefficient, so J'J is two (2T - 1)^2 blocks, T = (deg + 1)(deg + 2) / 2.
// One kernel pass per LM iteration: every CTA stages 64 samples' Jacobian rows in shared memory and accumulates its part of
// both normal-equation blocks in 4x4 register tiles; per-CTA partial sums are written out and folded in a fixed order by a second
// kernel (bit-reproducible).  The (2T - 1)-sized Cholesky solve of the LM step runs on the host between launches.
#pragma once
#include <cmath>
#include <vector>

#include "handle.cuh"

namespace mvba {

constexpr int RPC_MAX_DEG = 6;
constexpr int RPC_MAX_T = (RPC_MAX_DEG + 1) * (RPC_MAX_DEG + 2) / 2;  // 28 monomials
constexpr int RPC_MAX_H = 2 * RPC_MAX_T - 1;                          // 55 coefficients per coordinate
constexpr int RPC_LD = 56;                                            // padded row length (multiple of 4)
constexpr int RPC_CHUNK = 64;                                         // samples staged per step
constexpr int RPC_TILES = (RPC_LD / 4) * (RPC_LD / 4 + 1) / 2;        // 105 upper-triangular 4x4 tiles
constexpr int RPC_THREADS = 256;

struct RpcArgs {
  int64_t N;
  const double* in_xy; const double* out_xy;
  int deg, T, nh;              // nh = 2T - 1
  double coef[2 * RPC_MAX_H];  // [x: num(T) den(T-1) | y: num(T) den(T-1)]
  double delta[2 * RPC_MAX_H]; // central-difference step of every coefficient (0: exact derivative)
  int want_jac;
};

// per-CTA partial output: [JtJ x (nh*nh, upper tiles only touched) | JtJ y | g x | g y | cost]
__host__ __device__ inline size_t rpc_part_len(int nh) { return 2 * (size_t)RPC_LD * RPC_LD + 2 * RPC_LD + 8; }

__global__ void __launch_bounds__(RPC_THREADS) rpc_accum_kernel(RpcArgs a, double* __restrict__ part) {
  extern __shared__ double sm[];
  double* Jx = sm;                              // [RPC_CHUNK][RPC_LD]
  double* Jy = Jx + RPC_CHUNK * RPC_LD;
  double* rx = Jy + RPC_CHUNK * RPC_LD;         // [RPC_CHUNK]
  double* ry = rx + RPC_CHUNK;
  __shared__ double red[RPC_THREADS / 32];
  const int tid = threadIdx.x, nh = a.nh, T = a.T;
  // tile ownership: threads 0..104 own tile (ti <= tj) of x, threads 128..232 of y
  const int half = tid >> 7, tl = tid & 127;
  int ti = -1, tj = -1;
  if (tl < RPC_TILES) { int r = 0, t = tl; while (t >= RPC_LD / 4 - r) { t -= RPC_LD / 4 - r; r++; } ti = r; tj = r + t; }
  double acc[16];
#pragma unroll
  for (int k = 0; k < 16; k++) acc[k] = 0.0;
  double gacc = 0.0, cost = 0.0;   // threads tl < nh of each half own g[tl]
  const int64_t per = (a.N + gridDim.x - 1) / gridDim.x;
  const int64_t s0 = blockIdx.x * per, s1 = min(a.N, s0 + per);
  for (int64_t base = s0; base < s1; base += RPC_CHUNK) {
    const int cnt = (int)min((int64_t)RPC_CHUNK, s1 - base);
    __syncthreads();
    if (tid < RPC_CHUNK) {
      double* jx = Jx + tid * RPC_LD; double* jy = Jy + tid * RPC_LD;
      if (tid < cnt) {
        const double x = a.in_xy[2 * (base + tid)], y = a.in_xy[2 * (base + tid) + 1];
        double powx[RPC_MAX_DEG + 1], powy[RPC_MAX_DEG + 1], m[RPC_MAX_T];
        double vx = 1.0, vy = 1.0;
#pragma unroll
        for (int d = 0; d <= RPC_MAX_DEG; d++) { powx[d] = vx; vx *= x; powy[d] = vy; vy *= y; }
        int k = 0;
        for (int d = 0; d <= a.deg; d++) for (int i = 0; i <= d; i++) m[k++] = powx[d - i] * powy[i];
        // compute_rpc's accumulation order: coefficient times powx times powy, summed in storage order
        double v[4] = {0.0, 1.0, 0.0, 1.0};
        int ci = 0;
        for (int c = 0; c < 4; c++) {
          int kk = (c & 1) ? 1 : 0;
          for (int d = kk; d <= a.deg; d++) for (int i = 0; i <= d; i++) { v[c] = v[c] + a.coef[ci] * powx[d - i] * powy[i]; ci++; }
        }
        const double fx = v[0] / v[1], fy = v[2] / v[3];
        rx[tid] = fx - a.out_xy[2 * (base + tid)]; ry[tid] = fy - a.out_xy[2 * (base + tid) + 1];
        if (a.want_jac) {
          for (int q = 0; q < T; q++) { jx[q] = m[q] / v[1]; jy[q] = m[q] / v[3]; }
          for (int q = 1; q < T; q++) {
            const double ex = a.delta[T + q - 1] * m[q], ey = a.delta[nh + T + q - 1] * m[q];
            jx[T + q - 1] = -v[0] * m[q] / (v[1] * v[1] - ex * ex);
            jy[T + q - 1] = -v[2] * m[q] / (v[3] * v[3] - ey * ey);
          }
          for (int q = nh; q < RPC_LD; q++) { jx[q] = 0.0; jy[q] = 0.0; }
        }
      } else {
        rx[tid] = 0.0; ry[tid] = 0.0;
        for (int q = 0; q < RPC_LD; q++) { jx[q] = 0.0; jy[q] = 0.0; }
      }
    }
    __syncthreads();
    if (tid < RPC_CHUNK) cost += rx[tid] * rx[tid] + ry[tid] * ry[tid];
    if (a.want_jac) {
      const double* J = half ? Jy : Jx;
      if (ti >= 0) {
        for (int s = 0; s < RPC_CHUNK; s++) {
          const double4 u = *reinterpret_cast<const double4*>(J + s * RPC_LD + 4 * ti);
          const double4 w = *reinterpret_cast<const double4*>(J + s * RPC_LD + 4 * tj);
          acc[0] += u.x * w.x; acc[1] += u.x * w.y; acc[2] += u.x * w.z; acc[3] += u.x * w.w;
          acc[4] += u.y * w.x; acc[5] += u.y * w.y; acc[6] += u.y * w.z; acc[7] += u.y * w.w;
          acc[8] += u.z * w.x; acc[9] += u.z * w.y; acc[10] += u.z * w.z; acc[11] += u.z * w.w;
          acc[12] += u.w * w.x; acc[13] += u.w * w.y; acc[14] += u.w * w.z; acc[15] += u.w * w.w;
        }
      }
    }
    if (a.want_jac && tl < nh) {
      const double* J = half ? Jy : Jx;
      const double* r = half ? ry : rx;
      double g = 0.0;
      for (int s = 0; s < RPC_CHUNK; s++) g += J[s * RPC_LD + tl] * r[s];
      gacc += g;
    }
  }
  // write this CTA's partial sums
  double* out = part + blockIdx.x * rpc_part_len(nh);
  if (a.want_jac) {
    double* M = out + (size_t)half * RPC_LD * RPC_LD;
    if (ti >= 0) {
#pragma unroll
      for (int r = 0; r < 4; r++)
#pragma unroll
        for (int c = 0; c < 4; c++) M[(4 * ti + r) * RPC_LD + 4 * tj + c] = acc[4 * r + c];
    }
    if (tl < RPC_LD) out[2 * RPC_LD * RPC_LD + half * RPC_LD + tl] = tl < nh ? gacc : 0.0;
  }
  // cost: block reduction in a fixed order
  for (int o = 16; o > 0; o >>= 1) cost += __shfl_down_sync(0xffffffffu, cost, o);
  if ((tid & 31) == 0) red[tid >> 5] = cost;
  __syncthreads();
  if (tid == 0) { double c = 0.0; for (int w = 0; w < RPC_THREADS / 32; w++) c += red[w]; out[2 * RPC_LD * RPC_LD + 2 * RPC_LD] = c; }
}

// Fold the per-CTA partial sums in CTA order (deterministic).
__global__ void rpc_fold_kernel(const double* __restrict__ part, int nparts, size_t len, double* __restrict__ out) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= len) return;
  double s = 0.0;
  for (int p = 0; p < nparts; p++) s += part[p * len + i];
  out[i] = s;
}

// ---------------------------------------------------------------- host side: Levenberg-Marquardt around the kernel pass
struct RpcFitter {
  int64_t N = 0;
  DBuf<double> d_in, d_out, d_part, d_sum;
  std::vector<double> h_sum;
  cudaStream_t stream = nullptr;
  int grid = 1, launches = 0;

  int init(int64_t n, const double* in_xy, const double* out_xy, int sm_count) {
    N = n;
    MVBA_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    MVBA_CUDA(d_in.alloc(2 * (size_t)n)); MVBA_CUDA(d_out.alloc(2 * (size_t)n));
    MVBA_CUDA(cudaMemcpyAsync(d_in.p, in_xy, 16 * (size_t)n, cudaMemcpyHostToDevice, stream));
    MVBA_CUDA(cudaMemcpyAsync(d_out.p, out_xy, 16 * (size_t)n, cudaMemcpyHostToDevice, stream));
    grid = (int)std::max<int64_t>(1, std::min<int64_t>(2 * (int64_t)sm_count, (n + RPC_CHUNK - 1) / RPC_CHUNK));
    const size_t len = rpc_part_len(0);
    MVBA_CUDA(d_part.alloc(len * grid)); MVBA_CUDA(d_part.zero(stream)); MVBA_CUDA(d_sum.alloc(len));
    h_sum.assign(len, 0.0);
    MVBA_CUDA(cudaFuncSetAttribute(rpc_accum_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem()));
    return MVBA_OK;
  }
  // Caller-owned scratch instead of allocations (mvba_rpc_fit_static): `scratch` holds scratch_len doubles of DEVICE memory laid
  // out as [in 2n | out 2n | per-CTA partial sums | folded sums]; nothing is allocated and nothing is freed, the stream is the
  // calling thread's default stream.  The number of CTAs is capped by what the scratch holds.
  bool borrowed = false;
  static size_t static_len(int64_t n, int parts) { return 4 * (size_t)n + ((size_t)parts + 1) * rpc_part_len(0); }
  int init_static(int64_t n, const double* in_xy, const double* out_xy, int sm_count, double* scratch, size_t scratch_len, int max_parts) {
    N = n; borrowed = true; stream = cudaStreamPerThread;
    grid = (int)std::max<int64_t>(1, std::min<int64_t>(std::min<int64_t>(2 * (int64_t)sm_count, max_parts), (n + RPC_CHUNK - 1) / RPC_CHUNK));
    const size_t len = rpc_part_len(0);
    if (static_len(n, grid) > scratch_len) return fail(MVBA_EINVAL, "rpc fit: the static scratch area is too small for this many samples");
    d_in.p = scratch; d_in.n = 2 * (size_t)n;
    d_out.p = d_in.p + d_in.n; d_out.n = 2 * (size_t)n;
    d_part.p = d_out.p + d_out.n; d_part.n = len * grid;
    d_sum.p = d_part.p + d_part.n; d_sum.n = len;
    MVBA_CUDA(cudaMemcpyAsync(d_in.p, in_xy, 16 * (size_t)n, cudaMemcpyHostToDevice, stream));
    MVBA_CUDA(cudaMemcpyAsync(d_out.p, out_xy, 16 * (size_t)n, cudaMemcpyHostToDevice, stream));
    MVBA_CUDA(d_part.zero(stream));  // the area keeps the previous call's sums; tiles the kernel never writes must read as zero
    MVBA_CUDA(d_sum.zero(stream));
    h_sum.assign(len, 0.0);
    MVBA_CUDA(cudaFuncSetAttribute(rpc_accum_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem()));
    return MVBA_OK;
  }
  ~RpcFitter() {
    if (borrowed) { d_in.p = d_out.p = d_part.p = d_sum.p = nullptr; return; }  // not ours: the DBuf destructors must not free them
    if (stream) cudaStreamDestroy(stream);
  }
  static size_t smem() { return (2 * (size_t)RPC_CHUNK * RPC_LD + 2 * RPC_CHUNK) * sizeof(double); }

  // cost, J'J (two nh x nh blocks, full symmetric) and J'r at coefficients c (reference order)
  int pass(const std::vector<double>& c, int deg, int jacobian_mode, double* cost, std::vector<double>* JtJ, std::vector<double>* g) {
    RpcArgs a{};
    a.N = N; a.in_xy = d_in.p; a.out_xy = d_out.p; a.deg = deg; a.T = (deg + 1) * (deg + 2) / 2; a.nh = 2 * a.T - 1; a.want_jac = JtJ != nullptr;
    const int nh = a.nh;
    const double min_step = std::sqrt(2.220446049250313e-16);
    for (int i = 0; i < 2 * nh; i++) { a.coef[i] = c[i]; a.delta[i] = jacobian_mode == 0 ? std::max(std::fabs(c[i]) * 1e-6, min_step) : 0.0; }
    const size_t len = rpc_part_len(nh);
    rpc_accum_kernel<<<grid, RPC_THREADS, smem(), stream>>>(a, d_part.p);
    rpc_fold_kernel<<<(unsigned)((len + 255) / 256), 256, 0, stream>>>(d_part.p, grid, len, d_sum.p);
    launches += 2;
    MVBA_CUDA(cudaMemcpyAsync(h_sum.data(), d_sum.p, len * sizeof(double), cudaMemcpyDeviceToHost, stream));
    MVBA_CUDA(cudaStreamSynchronize(stream));
    *cost = 0.5 * h_sum[2 * RPC_LD * RPC_LD + 2 * RPC_LD];
    if (JtJ) {
      JtJ->assign(2 * (size_t)nh * nh, 0.0); g->assign(2 * (size_t)nh, 0.0);
      for (int h = 0; h < 2; h++) {
        const double* M = h_sum.data() + (size_t)h * RPC_LD * RPC_LD;
        for (int i = 0; i < nh; i++) {
          for (int j = i; j < nh; j++) { (*JtJ)[(size_t)h * nh * nh + i * nh + j] = M[i * RPC_LD + j]; (*JtJ)[(size_t)h * nh * nh + j * nh + i] = M[i * RPC_LD + j]; }
          (*g)[h * nh + i] = h_sum[2 * RPC_LD * RPC_LD + h * RPC_LD + i];
        }
      }
    }
    return MVBA_OK;
  }
};

// Cholesky solve of the n x n system A y = b (row-major, A destroyed, b overwritten with y)
inline bool rpc_chol_solve(double* A, double* b, int n) {
  for (int j = 0; j < n; j++) {
    double d = A[j * n + j];
    for (int k = 0; k < j; k++) d -= A[j * n + k] * A[j * n + k];
    if (!(d > 0)) return false;
    d = std::sqrt(d); A[j * n + j] = d;
    for (int i = j + 1; i < n; i++) {
      double s = A[i * n + j];
      for (int k = 0; k < j; k++) s -= A[i * n + k] * A[j * n + k];
      A[i * n + j] = s / d;
    }
  }
  for (int i = 0; i < n; i++) { double s = b[i]; for (int k = 0; k < i; k++) s -= A[i * n + k] * b[k]; b[i] = s / A[i * n + i]; }
  for (int i = n - 1; i >= 0; i--) { double s = b[i]; for (int k = i + 1; k < n; k++) s -= A[k * n + i] * b[k]; b[i] = s / A[i * n + i]; }
  return true;
}

struct RpcFitStats { double initial_cost = 0, final_cost = 0; int iterations = 0, successful = 0; };

// fitCurrDegRPC (rpc_distortion.cc:556-617) with Ceres 1.14 trust-region rules: Jacobi scaling from the first linearisation,
// LM diagonal clamp(diag, 1e-6, 1e32) / radius, exact step, tolerances tested before acceptance, radius update by step quality.
inline int rpc_fit_curr_degree(RpcFitter& F, int deg, int num_iterations, double parameter_tolerance, int jacobian_mode,
                               std::vector<double>& c, RpcFitStats* st) {
  const int T = (deg + 1) * (deg + 2) / 2, nh = 2 * T - 1, n = 2 * nh;
  std::vector<double> JtJ, g, JtJ2, g2, scale(n), A((size_t)nh * nh), As((size_t)nh * nh), y(n), gs(n), cand(n);
  double cost = 0;
  int rc = F.pass(c, deg, jacobian_mode, &cost, &JtJ, &g); if (rc) return rc;
  st->initial_cost = cost;
  for (int h = 0; h < 2; h++) for (int i = 0; i < nh; i++) scale[h * nh + i] = 1.0 / (1.0 + std::sqrt(JtJ[(size_t)h * nh * nh + i * nh + i]));
  auto gmax = [&]() { double m = 0; for (int i = 0; i < n; i++) m = std::max(m, std::fabs(g[i])); return m; };
  double radius = 1e4, decrease = 2.0;
  bool done = gmax() <= 1e-16;
  int iter = 0;
  while (!done && iter < num_iterations) {
    iter++;
    double mcc = 0;
    bool ok = true;
    for (int h = 0; h < 2 && ok; h++) {
      const double* M = &JtJ[(size_t)h * nh * nh];
      const double* sc = &scale[h * nh];
      for (int i = 0; i < nh; i++) {
        for (int j = 0; j < nh; j++) As[(size_t)i * nh + j] = M[i * nh + j] * sc[i] * sc[j];
        gs[h * nh + i] = g[h * nh + i] * sc[i];
        y[h * nh + i] = gs[h * nh + i];
      }
      A = As;
      for (int i = 0; i < nh; i++) A[(size_t)i * nh + i] += std::min(std::max(As[(size_t)i * nh + i], 1e-6), 1e32) / radius;
      ok = rpc_chol_solve(A.data(), &y[h * nh], nh);
      if (ok)
        for (int i = 0; i < nh; i++) {
          double Ad = 0;
          for (int j = 0; j < nh; j++) Ad += As[(size_t)i * nh + j] * y[h * nh + j];
          mcc += y[h * nh + i] * gs[h * nh + i] - 0.5 * y[h * nh + i] * Ad;
        }
    }
    if (!ok || !(mcc > 0) || !std::isfinite(mcc)) {
      radius /= decrease; decrease *= 2;
    } else {
      double sn = 0, xn = 0;
      for (int i = 0; i < n; i++) { const double d = -y[i] * scale[i]; cand[i] = c[i] + d; sn += d * d; xn += c[i] * c[i]; }
      double cand_cost = 0;
      rc = F.pass(cand, deg, jacobian_mode, &cand_cost, &JtJ2, &g2); if (rc) return rc;
      if (!std::isfinite(cand_cost)) cand_cost = 1.7976931348623157e308;
      if (std::sqrt(sn) <= parameter_tolerance * (std::sqrt(xn) + parameter_tolerance)) break;
      const double change = cost - cand_cost;
      if (std::fabs(change) <= 1e-16 * cost) break;
      const double rel = change / mcc;
      if (rel > 1e-3) {
        c = cand; cost = cand_cost; st->successful++;
        JtJ.swap(JtJ2); g.swap(g2);
        if (gmax() <= 1e-16) done = true;
        const double t = 2.0 * rel - 1.0;
        radius = std::min(1e16, radius / std::max(1.0 / 3.0, 1.0 - t * t * t));
        decrease = 2.0;
      } else {
        radius /= decrease; decrease *= 2;
      }
    }
    if (radius < 1e-32) break;
  }
  st->final_cost = cost; st->iterations += iter;
  return MVBA_OK;
}

// fitRpcDist's degree continuation (rpc_distortion.cc:640-652); c comes back in the reference's packed order
inline int rpc_fit_all(RpcFitter& F, int rpc_degree, int num_iterations, double parameter_tolerance, int jacobian_mode,
                       std::vector<double>& c, RpcFitStats* st) {
  c.assign(10, 0.0); c[1] = 1.0; c[5 + 2] = 1.0;  // init_as_identity at degree 1: num_x = x, num_y = y
  double first = 0;
  for (int d = 1; d <= rpc_degree; d++) {
    if (d >= 2) {  // increment_degree: r + 2 zero coefficients appended to each of the four polynomials
      const int Tp = d * (d + 1) / 2, add = d + 1;
      std::vector<double> out;
      int pos = 0;
      for (int poly = 0; poly < 4; poly++) {
        const int len = (poly % 2 == 0) ? Tp : Tp - 1;
        out.insert(out.end(), c.begin() + pos, c.begin() + pos + len);
        out.insert(out.end(), add, 0.0);
        pos += len;
      }
      c.swap(out);
    }
    RpcFitStats s{};
    int rc = rpc_fit_curr_degree(F, d, num_iterations, parameter_tolerance, jacobian_mode, c, &s); if (rc) return rc;
    if (d == 1) first = s.initial_cost;
    st->iterations += s.iterations; st->successful += s.successful; st->final_cost = s.final_cost;
  }
  st->initial_cost = first;
  return MVBA_OK;
}

}  // namespace mvba
