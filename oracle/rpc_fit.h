// oracle/rpc_fit.h -- CPU restatement (TEST INFRASTRUCTURE ONLY, never linked into the product) of the RPC fit that runs
// between rig_calibrator passes whenever RPC distortion floats:
//   RpcFitError                   camera_model/rpc_distortion.cc:366-420 (DynamicNumericDiffCostFunction, central differences)
//   fitCurrDegRPC                 rpc_distortion.cc:556-617  (one parameter block, no loss, ITERATIVE_SCHUR, 1e-16 tolerances)
//   fitRpcDist / fitRpcUndist     rpc_distortion.cc:619-728  (degree continuation from the identity model)
//   init_as_identity / increment_degree   rpc_distortion.cc:300-370
// Ceres 1.14 is not vendored in the reference; what is restated here is its published behaviour for this problem:
//   * numeric_diff.h: delta = max(|x| 1e-6, sqrt(eps)), (f(x + delta) - f(x - delta)) / (2 delta), per parameter;
//   * trust_region_minimizer.cc: Jacobi scaling 1 / (1 + ||col||) fixed at the first linearisation, Levenberg-Marquardt
//     diagonal clamp(diag J'J, 1e-6, 1e32) / radius, radius 1e4, step quality and radius rules, parameter / function /
//     gradient tolerances checked in that order BEFORE the step is accepted;
//   * the single parameter block is the whole elimination set, so ITERATIVE_SCHUR has an empty reduced system and the step
//     is the back-substitution (E'E + D'D)^-1 E'b: a dense Cholesky solve.
// Parity unpinned by the reference (no tests or fixtures there); pinned in tests/ on exactly representable rational models.
#pragma once
#include <cmath>
#include <limits>
#include <vector>

namespace oracle {

inline int rpc_num_params(int deg) { return 2 * (deg + 1) * (deg + 2) - 2; }

inline void rpc_init_identity(std::vector<double>& c, int deg) {
  c.assign(rpc_num_params(deg), 0.0);
  int num_len = (int)(c.size() + 2) / 4, den_len = num_len - 1;
  c[1] = 1.0;                          // num_x[1]
  c[num_len + den_len + 2] = 1.0;      // num_y[2]
}

// increment_degree: every one of the four polynomials gains the monomials of the next degree with zero coefficients
inline void rpc_increment_degree(std::vector<double>& c) {
  int n = (int)c.size(), num_len = (n + 2) / 4, den_len = num_len - 1, deg = rpc_degree(n);
  int add = deg + 2;
  std::vector<double> out;
  int pos = 0;
  for (int poly = 0; poly < 4; poly++) {
    int len = (poly % 2 == 0) ? num_len : den_len;
    for (int i = 0; i < len; i++) out.push_back(c[pos + i]);
    for (int i = 0; i < add; i++) out.push_back(0.0);
    pos += len;
  }
  c.swap(out);
}

// In-place Cholesky solve of the symmetric positive definite n x n system A y = b (A row-major, destroyed).
inline bool chol_solve(std::vector<double>& A, std::vector<double>& b, int n) {
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

struct RpcFitStats { double initial_cost, final_cost; int iterations, successful; };

// fitCurrDegRPC: min over c of 1/2 sum |rpc(in, c) - out|^2
inline void rpc_fit_curr_degree(int64_t N, const double* in_xy, const double* out_xy, int num_iterations, double parameter_tolerance,
                                std::vector<double>& c, RpcFitStats* st) {
  const int n = (int)c.size();
  const double min_step = std::sqrt(std::numeric_limits<double>::epsilon());
  std::vector<double> JtJ((size_t)n * n), g(n), scale(n, 0.0), A((size_t)n * n), rhs(n), cand(n);
  auto cost_at = [&](const std::vector<double>& cc) {
    double s = 0;
#pragma omp parallel for reduction(+ : s) schedule(static)
    for (int64_t i = 0; i < N; i++) {
      double v[2];
      compute_rpc(in_xy + 2 * i, cc.data(), n, v);
      double rx = v[0] - out_xy[2 * i], ry = v[1] - out_xy[2 * i + 1];
      s += rx * rx + ry * ry;
    }
    return 0.5 * s;
  };
  auto linearize = [&]() {
    std::fill(JtJ.begin(), JtJ.end(), 0.0); std::fill(g.begin(), g.end(), 0.0);
#pragma omp parallel
    {
      std::vector<double> lJ((size_t)n * n, 0.0), lg(n, 0.0), J(2 * (size_t)n), x(c);
#pragma omp for schedule(static)
      for (int64_t i = 0; i < N; i++) {
        double v[2], vp[2], vm[2];
        compute_rpc(in_xy + 2 * i, x.data(), n, v);
        double r[2] = {v[0] - out_xy[2 * i], v[1] - out_xy[2 * i + 1]};
        for (int j = 0; j < n; j++) {
          double xj = x[j], delta = std::max(std::fabs(xj) * 1e-6, min_step);
          x[j] = xj + delta; compute_rpc(in_xy + 2 * i, x.data(), n, vp);
          x[j] = xj - delta; compute_rpc(in_xy + 2 * i, x.data(), n, vm);
          x[j] = xj;
          double inv = 1.0 / (2.0 * delta);
          J[j] = ((vp[0] - out_xy[2 * i]) - (vm[0] - out_xy[2 * i])) * inv;
          J[n + j] = ((vp[1] - out_xy[2 * i + 1]) - (vm[1] - out_xy[2 * i + 1])) * inv;
        }
        for (int k = 0; k < 2; k++) {
          const double* Jr = &J[(size_t)k * n];
          for (int a = 0; a < n; a++) {
            if (Jr[a] == 0.0) continue;
            lg[a] += Jr[a] * r[k];
            for (int b = a; b < n; b++) lJ[(size_t)a * n + b] += Jr[a] * Jr[b];
          }
        }
      }
#pragma omp critical
      { for (size_t q = 0; q < lJ.size(); q++) JtJ[q] += lJ[q]; for (int a = 0; a < n; a++) g[a] += lg[a]; }
    }
    for (int a = 0; a < n; a++) for (int b = 0; b < a; b++) JtJ[(size_t)a * n + b] = JtJ[(size_t)b * n + a];
  };
  double cost = cost_at(c);
  st->initial_cost = cost; st->iterations = 0; st->successful = 0;
  linearize();
  for (int a = 0; a < n; a++) scale[a] = 1.0 / (1.0 + std::sqrt(JtJ[(size_t)a * n + a]));
  auto gmax = [&]() { double m = 0; for (int a = 0; a < n; a++) m = std::max(m, std::fabs(g[a])); return m; };
  double radius = 1e4, decrease = 2.0;
  bool done = gmax() <= 1e-16;
  int iter = 0;
  while (!done && iter < num_iterations) {
    iter++;
    for (int a = 0; a < n; a++) {
      for (int b = 0; b < n; b++) A[(size_t)a * n + b] = JtJ[(size_t)a * n + b] * scale[a] * scale[b];
      rhs[a] = g[a] * scale[a];
    }
    std::vector<double> As(A), gs(rhs);
    for (int a = 0; a < n; a++) A[(size_t)a * n + a] += std::min(std::max(As[(size_t)a * n + a], 1e-6), 1e32) / radius;
    bool ok = chol_solve(A, rhs, n);
    double mcc = 0;
    if (ok) {
      // step = -y; model cost change = -(step' g + step' A step / 2)
      for (int a = 0; a < n; a++) {
        double Ad = 0;
        for (int b = 0; b < n; b++) Ad += As[(size_t)a * n + b] * rhs[b];
        mcc += rhs[a] * gs[a] - 0.5 * rhs[a] * Ad;
      }
    }
    if (!ok || !(mcc > 0) || !std::isfinite(mcc)) {
      radius /= decrease; decrease *= 2;
    } else {
      double sn = 0, xn = 0;
      for (int a = 0; a < n; a++) { double d = -rhs[a] * scale[a]; cand[a] = c[a] + d; sn += d * d; xn += c[a] * c[a]; }
      double cand_cost = cost_at(cand);
      if (!std::isfinite(cand_cost)) cand_cost = std::numeric_limits<double>::max();
      if (std::sqrt(sn) <= parameter_tolerance * (std::sqrt(xn) + parameter_tolerance)) break;
      double change = cost - cand_cost;
      if (std::fabs(change) <= 1e-16 * cost) break;
      double rel = change / mcc;
      if (rel > 1e-3) {
        c = cand; cost = cand_cost; st->successful++;
        linearize();
        if (gmax() <= 1e-16) done = true;
        double t = 2.0 * rel - 1.0;
        radius = std::min(1e16, radius / std::max(1.0 / 3.0, 1.0 - t * t * t));
        decrease = 2.0;
      } else {
        radius /= decrease; decrease *= 2;
      }
    }
    if (radius < 1e-32) break;
  }
  st->final_cost = cost; st->iterations += iter;
}

// fitRpcDist's loop (rpc_distortion.cc:640-652): degree 1 from the identity, then one degree at a time
inline void rpc_fit(int64_t N, const double* in_xy, const double* out_xy, int deg, int num_iterations, double parameter_tolerance,
                    std::vector<double>& c, RpcFitStats* st) {
  rpc_init_identity(c, 1);
  RpcFitStats s{}; int total = 0, succ = 0; double first = 0;
  for (int d = 1; d <= deg; d++) {
    if (d >= 2) rpc_increment_degree(c);
    rpc_fit_curr_degree(N, in_xy, out_xy, num_iterations, parameter_tolerance, c, &s);
    if (d == 1) first = s.initial_cost;
    total += s.iterations; succ += s.successful;
  }
  st->initial_cost = first; st->final_cost = s.final_cost; st->iterations = total; st->successful = succ;
}

}  // namespace oracle
