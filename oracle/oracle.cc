// oracle/oracle.cc -- TEST INFRASTRUCTURE ONLY (checker + timed CPU baseline).  PARITY UNPINNED
// (see functors.h).  Restates, on the host and without dependencies:
//   * the problem assembly of one rig_calibrator pass: residual-block order, parameter-block
//     identity and the constancy rules (reference rig_calibrator/bin/rig_calibrator.cc:1611-1903);
//   * Ceres 1.14 behaviour the reference relies on (third party, pinned ceres-solver=1.14.0 in
//     conda/linux_deps_env.yaml; restated from its documented algorithm, not verifiable here):
//     DynamicNumericDiffCostFunction central differences, the loss-function Corrector for
//     CauchyLoss, TrustRegionMinimizer + LevenbergMarquardtStrategy, ITERATIVE_SCHUR =
//     ImplicitSchurComplement + ConjugateGradientsSolver with the JACOBI preconditioner.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library.
#include <omp.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <limits>
#include <vector>

#include "../include/mvba.h"
#include "functors.h"
#include "tracks.h"
#include "rpc_fit.h"

namespace oracle {

struct PBlock {
  double* ptr = nullptr;
  int size = 0;
  bool constant = true;
  bool used = false;   // appears in at least one residual block (else Ceres never sees it)
  bool is_e = false;   // landmark block -> eliminated (Ceres' automatic Schur ordering)
  int off = -1;        // offset in the e- or f- part of the free vector
  bool bounded = false;
  double lo = 0, hi = 0;
};

struct RBlock {
  int kind, np, nres, sensor;
  int pb[8];
  double meas[3], aux[3], stamps[3];  // stamps: beg, end, cam
  double weight, loss_a;
  int64_t src;
  int64_t row0;
};

struct Problem {
  mvba_config cfg{};
  std::vector<SensorConst> sensors;
  std::vector<int> dist_off;
  mvba_params prm{};
  bool have_params = false;
  std::vector<int32_t> cam_sensor, cam_beg, cam_end;
  std::vector<double> cam_stamp;
  int64_t nobs = 0;
  std::vector<int32_t> opid, ocid;
  std::vector<float> opix, odepth;
  std::vector<uint8_t> ohasd, ohasm;
  std::vector<double> omesh;
  struct Prior { std::vector<int32_t> pid; std::vector<double> ref; double w = 0, th = 0; } prior[2];
  std::vector<uint8_t> m_focal, m_center, m_dist, m_rig, m_d2i, m_poses;
  int float_offsets = 0, float_scale = 0;
  bool have_mask = false;

  double identity[7] = {0, 0, 0, 0, 0, 0, 1};  // identity_vec, rig_calibrator.cc:1419-1421
  double dist_placeholder = 0;                 // distortion_placeholder, :1608
  // built
  bool built = false;
  std::vector<PBlock> pbs;
  std::vector<RBlock> rbs;
  int64_t nres = 0;
  int nf = 0, ne = 0;  // scalar sizes of the free f / e parts
  int nd2i() const { return cfg.affine_depth_to_image ? 12 : 7; }
};

static int pb_pose(const Problem& P, int i) { return i; }  // world_to_ref[r] or world_to_cam[c]

static void build(Problem& P) {
  const mvba_config& c = P.cfg;
  const int T = c.num_sensors;
  const int npose = c.no_rig ? c.num_cameras : c.num_ref_frames;
  P.pbs.clear(); P.rbs.clear();
  auto add = [&](double* ptr, int size) { PBlock b; b.ptr = ptr; b.size = size; P.pbs.push_back(b); return (int)P.pbs.size() - 1; };
  for (int i = 0; i < npose; i++) add((c.no_rig ? P.prm.world_to_cam : P.prm.world_to_ref) + 7 * i, 7);
  const int id_identity = add(P.identity, 7);
  std::vector<int> id_rig(T), id_off(T), id_focal(T), id_center(T), id_dist(T), id_d2i(T), id_scale(T);
  for (int t = 0; t < T; t++) {
    id_rig[t] = add(P.prm.ref_to_cam + 7 * t, 7);
    id_off[t] = add(P.prm.offsets + t, 1);
    id_focal[t] = add(P.prm.focal + t, 1);
    id_center[t] = add(P.prm.center + 2 * t, 2);
    id_dist[t] = P.sensors[t].nd > 0 ? add(P.prm.distortion + P.dist_off[t], P.sensors[t].nd) : add(&P.dist_placeholder, 0);
    id_d2i[t] = add(P.prm.depth_to_image ? P.prm.depth_to_image + P.nd2i() * t : nullptr, P.nd2i());
    id_scale[t] = add(P.prm.depth_scale ? P.prm.depth_scale + t : nullptr, 1);
  }
  const int id_xyz0 = (int)P.pbs.size();
  for (int p = 0; p < c.num_landmarks; p++) { int id = add(P.prm.xyz + 3 * p, 3); P.pbs[id].is_e = true; }

  // Constancy: rig_calibrator.cc:1702-1752, 1787-1793, 1834-1839 (defaults to constant).
  auto m = [&](const std::vector<uint8_t>& v, int t) { return !v.empty() && v[t] != 0; };
  for (int i = 0; i < npose; i++) {
    int s = c.no_rig ? P.cam_sensor[i] : 0;
    P.pbs[pb_pose(P, i)].constant = !m(P.m_poses, s);
  }
  for (int t = 0; t < T; t++) {
    P.pbs[id_rig[t]].constant = !(m(P.m_rig, t) && t != 0 && !c.no_rig);
    bool off_free = P.float_offsets && t != 0 && !c.no_rig;
    P.pbs[id_off[t]].constant = !off_free;
    if (off_free && P.prm.offset_lo && P.prm.offset_hi) {
      P.pbs[id_off[t]].bounded = true; P.pbs[id_off[t]].lo = P.prm.offset_lo[t]; P.pbs[id_off[t]].hi = P.prm.offset_hi[t];
    }
    P.pbs[id_focal[t]].constant = !m(P.m_focal, t);
    P.pbs[id_center[t]].constant = !m(P.m_center, t);
    P.pbs[id_dist[t]].constant = !(m(P.m_dist, t) && P.sensors[t].nd > 0);
    P.pbs[id_d2i[t]].constant = !m(P.m_d2i, t);
    P.pbs[id_scale[t]].constant = !(P.float_scale && !c.affine_depth_to_image);
  }
  for (int p = 0; p < c.num_landmarks; p++) P.pbs[id_xyz0 + p].constant = false;  // landmarks always float

  // Residual blocks in reference order, rig_calibrator.cc:1614-1903.
  int64_t row = 0;
  size_t prior_it[2] = {0, 0};
  auto add_priors = [&](int pid) {
    for (int k = 0; k < 2; k++) {
      auto& pr = P.prior[k];
      while (prior_it[k] < pr.pid.size() && pr.pid[prior_it[k]] < pid) prior_it[k]++;  // tolerate gaps
      if (prior_it[k] < pr.pid.size() && pr.pid[prior_it[k]] == pid) {
        RBlock b{}; b.kind = k == 0 ? MVBA_BLOCK_MESH_TRI : MVBA_BLOCK_TRI; b.np = 1; b.nres = 3; b.sensor = -1;
        b.pb[0] = id_xyz0 + pid;
        for (int i = 0; i < 3; i++) b.aux[i] = pr.ref[3 * prior_it[k] + i];
        b.weight = pr.w; b.loss_a = pr.th; b.src = (int64_t)prior_it[k]; b.row0 = row; row += 3;
        P.rbs.push_back(b);
        prior_it[k]++;
      }
    }
  };
  int64_t i = 0;
  while (i < P.nobs) {
    int pid = P.opid[i];
    for (; i < P.nobs && P.opid[i] == pid; i++) {
      int cid = P.ocid[i];
      int t = P.cam_sensor[cid];
      double bs, es, cs = P.cam_stamp[cid];
      int pbeg, pend;
      if (!c.no_rig) {
        int b = P.cam_beg[cid], e = P.cam_end[cid];
        pbeg = pb_pose(P, b);
        pend = (t == 0 || b == e) ? id_identity : pb_pose(P, e);  // :1642-1645
        bs = P.prm.ref_stamps[b]; es = P.prm.ref_stamps[e];
      } else {
        bs = es = cs; pbeg = pb_pose(P, cid); pend = id_identity;   // :1653-1666
      }
      RBlock b{}; b.kind = MVBA_BLOCK_CAM; b.np = 8; b.nres = 2; b.sensor = t;
      b.pb[0] = pbeg; b.pb[1] = pend; b.pb[2] = id_rig[t]; b.pb[3] = id_xyz0 + pid; b.pb[4] = id_off[t];
      b.pb[5] = id_focal[t]; b.pb[6] = id_center[t]; b.pb[7] = id_dist[t];
      b.meas[0] = P.opix[2 * i]; b.meas[1] = P.opix[2 * i + 1];
      b.stamps[0] = bs; b.stamps[1] = es; b.stamps[2] = cs;
      b.weight = 1.0; b.loss_a = c.robust_threshold; b.src = i; b.row0 = row; row += 2;
      P.rbs.push_back(b);
      bool has_d = !P.ohasd.empty() && P.ohasd[i];
      if (c.depth_tri_weight > 0 && has_d) {  // :1755-1794
        RBlock d{}; d.kind = MVBA_BLOCK_DEPTH_TRI; d.np = 7; d.nres = 3; d.sensor = t;
        d.pb[0] = pbeg; d.pb[1] = pend; d.pb[2] = id_rig[t]; d.pb[3] = id_d2i[t]; d.pb[4] = id_scale[t];
        d.pb[5] = id_xyz0 + pid; d.pb[6] = id_off[t];
        for (int k = 0; k < 3; k++) d.meas[k] = P.odepth[3 * i + k];
        d.stamps[0] = bs; d.stamps[1] = es; d.stamps[2] = cs;
        d.weight = c.depth_tri_weight; d.loss_a = c.robust_threshold; d.src = i; d.row0 = row; row += 3;
        P.rbs.push_back(d);
      }
      bool has_m = !P.ohasm.empty() && P.ohasm[i];
      if (c.depth_mesh_weight > 0 && has_m && has_d) {  // :1797-1840
        RBlock d{}; d.kind = MVBA_BLOCK_DEPTH_MESH; d.np = 6; d.nres = 3; d.sensor = t;
        d.pb[0] = pbeg; d.pb[1] = pend; d.pb[2] = id_rig[t]; d.pb[3] = id_d2i[t]; d.pb[4] = id_scale[t];
        d.pb[5] = id_off[t];
        for (int k = 0; k < 3; k++) { d.meas[k] = P.odepth[3 * i + k]; d.aux[k] = P.omesh[3 * i + k]; }
        d.stamps[0] = bs; d.stamps[1] = es; d.stamps[2] = cs;
        d.weight = c.depth_mesh_weight; d.loss_a = c.robust_threshold; d.src = i; d.row0 = row; row += 3;
        P.rbs.push_back(d);
      }
    }
    add_priors(pid);
  }
  P.nres = row;
  for (auto& pb : P.pbs) pb.used = false;
  for (auto& rb : P.rbs) for (int k = 0; k < rb.np; k++) P.pbs[rb.pb[k]].used = true;
  // Free-vector layout: f part (everything that is not a landmark) then e part.
  P.nf = 0; P.ne = 0;
  for (auto& pb : P.pbs) {
    pb.off = -1;
    if (pb.constant || !pb.used || pb.size == 0) continue;
    if (pb.is_e) { pb.off = P.ne; P.ne += pb.size; } else { pb.off = P.nf; P.nf += pb.size; }
  }
  P.built = true;
}

template <class T> static bool eval_block(const Problem& P, const RBlock& b, const T* const* p, T* r) {
  switch (b.kind) {
    case MVBA_BLOCK_CAM: return bracketed_cam_error<T>(p, b.meas, b.stamps[0], b.stamps[1], b.stamps[2], P.sensors[b.sensor], r);
    case MVBA_BLOCK_DEPTH_TRI: return bracketed_depth_error<T>(p, P.nd2i(), b.weight, b.meas, b.stamps[0], b.stamps[1], b.stamps[2], r);
    case MVBA_BLOCK_DEPTH_MESH: return bracketed_depth_mesh_error<T>(p, P.nd2i(), b.weight, b.meas, b.aux, b.stamps[0], b.stamps[1], b.stamps[2], r);
    default: return xyz_error<T>(p, b.weight, b.aux, r);
  }
}

static bool eval_block_d(const Problem& P, const RBlock& b, double* r) {
  const double* p[8];
  for (int k = 0; k < b.np; k++) p[k] = P.pbs[b.pb[k]].ptr;
  return eval_block<double>(P, b, p, r);
}

// Jacobian of one block w.r.t. the listed parameter blocks (all, or only the non-constant ones).
// mode 0: Ceres 1.14 DynamicNumericDiffCostFunction, CENTRAL, relative_step_size 1e-6,
//         min step sqrt(eps)  (the "reference Jacobian");
// mode 1: forward-mode dual numbers (exact derivative of the restated functor);
// mode 2: long double central differences, step 1e-7*max(1,|x|) (independent arbiter).
// J is row-major nres x ncols with ncols = sum of sizes of the selected blocks, in block order.
static void jac_block(const Problem& P, const RBlock& b, int mode, bool skip_constant, double* r, double* J, int ncols) {
  const int nr = b.nres;
  if (mode == 1) {
    int n = 0;
    for (int k = 0; k < b.np; k++) if (!(skip_constant && P.pbs[b.pb[k]].constant)) n += P.pbs[b.pb[k]].size;
    Dual::n() = n;
    std::vector<Dual> store;
    store.reserve(200);
    const Dual* p[8];
    int col = 0;
    size_t pos[8];
    for (int k = 0; k < b.np; k++) {
      const PBlock& pb = P.pbs[b.pb[k]];
      pos[k] = store.size();
      bool sel = !(skip_constant && pb.constant);
      for (int j = 0; j < pb.size; j++) { Dual d(pb.ptr[j]); if (sel) d.d[col++] = 1.0; store.push_back(d); }
      if (pb.size == 0) store.push_back(Dual(0.0));
    }
    for (int k = 0; k < b.np; k++) p[k] = &store[pos[k]];
    Dual rr[3];
    eval_block<Dual>(P, b, p, rr);
    for (int i = 0; i < nr; i++) { r[i] = rr[i].v; for (int j = 0; j < n; j++) J[i * ncols + j] = rr[i].d[j]; }
    Dual::n() = 0;
    return;
  }
  if (mode == 2) {
    std::vector<long double> store;
    size_t pos[8];
    for (int k = 0; k < b.np; k++) { const PBlock& pb = P.pbs[b.pb[k]]; pos[k] = store.size(); for (int j = 0; j < pb.size; j++) store.push_back(pb.ptr[j]); if (pb.size == 0) store.push_back(0); }
    const long double* p[8];
    for (int k = 0; k < b.np; k++) p[k] = &store[pos[k]];
    long double r0[3], rp[3], rm[3];
    eval_block<long double>(P, b, p, r0);
    for (int i = 0; i < nr; i++) r[i] = (double)r0[i];
    int col = 0;
    for (int k = 0; k < b.np; k++) {
      const PBlock& pb = P.pbs[b.pb[k]];
      if (skip_constant && pb.constant) continue;
      for (int j = 0; j < pb.size; j++, col++) {
        long double x = store[pos[k] + j];
        long double h = 1e-7L * std::max(1.0L, std::fabs(x));
        // RPC coefficients multiply monomials of pixel coordinates up to ~1e8: keep the
        // denominator perturbation small (the Ceres-rule step is NOT small there).
        if (b.kind == MVBA_BLOCK_CAM && k == 7 && pb.size > 5) h = 1e-13L;
        store[pos[k] + j] = x + h; eval_block<long double>(P, b, p, rp);
        store[pos[k] + j] = x - h; eval_block<long double>(P, b, p, rm);
        store[pos[k] + j] = x;
        for (int i = 0; i < nr; i++) J[i * ncols + col] = (double)((rp[i] - rm[i]) / (2 * h));
      }
    }
    return;
  }
  // mode 0, Ceres rule
  std::vector<double> store;
  size_t pos[8];
  for (int k = 0; k < b.np; k++) { const PBlock& pb = P.pbs[b.pb[k]]; pos[k] = store.size(); for (int j = 0; j < pb.size; j++) store.push_back(pb.ptr[j]); if (pb.size == 0) store.push_back(0); }
  const double* p[8];
  for (int k = 0; k < b.np; k++) p[k] = &store[pos[k]];
  eval_block<double>(P, b, p, r);
  const double min_step = std::sqrt(std::numeric_limits<double>::epsilon());
  double rp[3], rm[3];
  int col = 0;
  for (int k = 0; k < b.np; k++) {
    const PBlock& pb = P.pbs[b.pb[k]];
    if (skip_constant && pb.constant) continue;
    for (int j = 0; j < pb.size; j++, col++) {
      double x = store[pos[k] + j];
      double delta = std::max(min_step, std::fabs(x) * 1e-6);
      store[pos[k] + j] = x + delta; eval_block<double>(P, b, p, rp);
      store[pos[k] + j] = x - delta; eval_block<double>(P, b, p, rm);
      store[pos[k] + j] = x;
      double inv = 1.0 / (2.0 * delta);
      for (int i = 0; i < nr; i++) J[i * ncols + col] = (rp[i] - rm[i]) * inv;
    }
  }
}

// Ceres CauchyLoss(a): rho(s) = b log(1 + s/b), b = a^2.  Corrector for rho'' <= 0: scale residual
// and Jacobian rows by sqrt(rho').  (rig_calibrator.cc:394-411 only ever selects "cauchy".)
static inline void cauchy(double a, double s, double* rho, double* sqrt_rho1) {
  double b = a * a, c = 1.0 / b;
  double sum = 1.0 + s * c;
  *rho = b * std::log(sum);
  *sqrt_rho1 = std::sqrt(1.0 / sum);
}

// ----------------------------------------------------------------------------------------
// Linearised problem: per residual block, corrected residual and corrected Jacobian for the
// non-constant parameter blocks, split into the e (landmark) part and the f part.
// ----------------------------------------------------------------------------------------
struct LinBlock {
  int nres, eoff;       // eoff = -1 when the block has no landmark
  int nfb;              // number of free f blocks
  int foff[8], fsz[8], fcol[8];
  int ncols;            // total free columns (e first, then f blocks)
  size_t j0;            // offset into J storage
  double r[3];
};
struct Lin {
  std::vector<LinBlock> lb;
  std::vector<double> J;
  double cost = 0;
};

static double evaluate(const Problem& P, bool apply_loss, double* residuals) {
  double cost = 0;
  const int64_t nb = (int64_t)P.rbs.size();
#pragma omp parallel for reduction(+ : cost) schedule(static)
  for (int64_t i = 0; i < nb; i++) {
    const RBlock& b = P.rbs[i];
    double r[3];
    eval_block_d(P, b, r);
    double s = 0;
    for (int k = 0; k < b.nres; k++) s += r[k] * r[k];
    if (apply_loss) { double rho, sr; cauchy(b.loss_a, s, &rho, &sr); cost += 0.5 * rho; }
    else cost += 0.5 * s;
    if (residuals) for (int k = 0; k < b.nres; k++) residuals[b.row0 + k] = r[k];
  }
  return cost;
}

static void linearize(const Problem& P, int mode, Lin& L) {
  const int64_t nb = (int64_t)P.rbs.size();
  if ((int64_t)L.lb.size() != nb) {
    L.lb.resize(nb);
    size_t j0 = 0;
    for (int64_t i = 0; i < nb; i++) {
      const RBlock& b = P.rbs[i];
      LinBlock& lb = L.lb[i];
      lb.nres = b.nres; lb.eoff = -1; lb.nfb = 0; lb.ncols = 0;
      for (int k = 0; k < b.np; k++) {
        const PBlock& pb = P.pbs[b.pb[k]];
        if (pb.constant || pb.size == 0) continue;
        lb.ncols += pb.size;
      }
      lb.j0 = j0; j0 += (size_t)lb.nres * lb.ncols;
    }
    L.J.resize(j0);
  }
  double cost = 0;
#pragma omp parallel for reduction(+ : cost) schedule(dynamic, 256)
  for (int64_t i = 0; i < nb; i++) {
    const RBlock& b = P.rbs[i];
    LinBlock& lb = L.lb[i];
    double Jt[3 * kMaxDual];
    double r[3];
    jac_block(P, b, mode, true, r, Jt, lb.ncols);
    double s = 0;
    for (int k = 0; k < b.nres; k++) s += r[k] * r[k];
    double rho, sr; cauchy(b.loss_a, s, &rho, &sr);
    cost += 0.5 * rho;
    // reorder columns: e block first, then f blocks in block order
    lb.nfb = 0; lb.eoff = -1;
    int col = 0, ecol = -1, esz = 0;
    int fc[8], fs[8], fo[8], nfb = 0;
    for (int k = 0; k < b.np; k++) {
      const PBlock& pb = P.pbs[b.pb[k]];
      if (pb.constant || pb.size == 0) continue;
      if (pb.is_e) { ecol = col; esz = pb.size; lb.eoff = pb.off; }
      else { fc[nfb] = col; fs[nfb] = pb.size; fo[nfb] = pb.off; nfb++; }
      col += pb.size;
    }
    double* J = &L.J[lb.j0];
    int out = 0;
    int dst_of_src[kMaxDual];
    if (ecol >= 0) { for (int j = 0; j < esz; j++) dst_of_src[ecol + j] = out++; }
    lb.nfb = nfb;
    for (int q = 0; q < nfb; q++) { lb.fcol[q] = out; lb.fsz[q] = fs[q]; lb.foff[q] = fo[q]; for (int j = 0; j < fs[q]; j++) dst_of_src[fc[q] + j] = out++; }
    for (int k = 0; k < b.nres; k++) {
      lb.r[k] = sr * r[k];
      for (int j = 0; j < lb.ncols; j++) J[k * lb.ncols + dst_of_src[j]] = sr * Jt[k * lb.ncols + j];
    }
  }
  L.cost = cost;
}

// ----------------------------------------------------------------------------------------
// Ceres-1.14-style Levenberg-Marquardt with ITERATIVE_SCHUR (implicit Schur + CG, JACOBI).
// ----------------------------------------------------------------------------------------
struct Solver {
  const Problem& P;
  Lin L;
  int nf, ne, npts;
  std::vector<std::vector<int64_t>> track_blocks;  // residual blocks per landmark (e block index)
  std::vector<int64_t> free_blocks;                // blocks without a landmark
  std::vector<double> scale_f, scale_e;            // Jacobi scaling
  std::vector<double> Dsq_f, Dsq_e;                // LM diagonal squared (scaled space)
  std::vector<double> Hinv;                        // 9 per landmark: (E'E + D_e^2)^-1
  std::vector<double> ge;                          // 3 per landmark: E' r
  std::vector<double> pre;                         // block-diagonal preconditioner inverse, per f param block
  std::vector<int> fb_off, fb_sz; std::vector<size_t> fb_pre;  // f parameter blocks
  explicit Solver(const Problem& p) : P(p) {
    nf = P.nf; ne = P.ne; npts = ne / 3;
    track_blocks.resize(npts);
    for (auto& pb : P.pbs) if (pb.off >= 0 && !pb.is_e) { fb_off.push_back(pb.off); fb_sz.push_back(pb.size); }
    size_t o = 0; for (size_t i = 0; i < fb_sz.size(); i++) { fb_pre.push_back(o); o += (size_t)fb_sz[i] * fb_sz[i]; }
    pre.resize(o);
  }
  void index_blocks() {
    for (auto& v : track_blocks) v.clear();
    free_blocks.clear();
    for (int64_t i = 0; i < (int64_t)L.lb.size(); i++) {
      if (L.lb[i].eoff >= 0) track_blocks[L.lb[i].eoff / 3].push_back(i); else free_blocks.push_back(i);
    }
  }
  static bool inv3(const double* A, double* out) {
    double c00 = A[4] * A[8] - A[5] * A[7], c01 = A[5] * A[6] - A[3] * A[8], c02 = A[3] * A[7] - A[4] * A[6];
    double det = A[0] * c00 + A[1] * c01 + A[2] * c02;
    double id = 1.0 / det;
    out[0] = c00 * id; out[1] = (A[2] * A[7] - A[1] * A[8]) * id; out[2] = (A[1] * A[5] - A[2] * A[4]) * id;
    out[3] = c01 * id; out[4] = (A[0] * A[8] - A[2] * A[6]) * id; out[5] = (A[2] * A[3] - A[0] * A[5]) * id;
    out[6] = c02 * id; out[7] = (A[1] * A[6] - A[0] * A[7]) * id; out[8] = (A[0] * A[4] - A[1] * A[3]) * id;
    return std::isfinite(id);
  }
  // column squared norms of the (currently stored) Jacobian
  void col_sqnorm(std::vector<double>& cf, std::vector<double>& ce) {
    cf.assign(nf, 0.0); ce.assign(ne, 0.0);
    for (auto& lb : L.lb) {
      const double* J = &L.J[lb.j0];
      for (int k = 0; k < lb.nres; k++) {
        const double* Jr = J + k * lb.ncols;
        if (lb.eoff >= 0) for (int j = 0; j < 3; j++) ce[lb.eoff + j] += Jr[j] * Jr[j];
        for (int q = 0; q < lb.nfb; q++) for (int j = 0; j < lb.fsz[q]; j++) cf[lb.foff[q] + j] += Jr[lb.fcol[q] + j] * Jr[lb.fcol[q] + j];
      }
    }
  }
  void apply_scaling() {
    for (auto& lb : L.lb) {
      double* J = &L.J[lb.j0];
      for (int k = 0; k < lb.nres; k++) {
        double* Jr = J + k * lb.ncols;
        if (lb.eoff >= 0) for (int j = 0; j < 3; j++) Jr[j] *= scale_e[lb.eoff + j];
        for (int q = 0; q < lb.nfb; q++) for (int j = 0; j < lb.fsz[q]; j++) Jr[lb.fcol[q] + j] *= scale_f[lb.foff[q] + j];
      }
    }
  }
  void gradient(std::vector<double>& gf, std::vector<double>& gE) {
    gf.assign(nf, 0.0); gE.assign(ne, 0.0);
    for (auto& lb : L.lb) {
      const double* J = &L.J[lb.j0];
      for (int k = 0; k < lb.nres; k++) {
        const double* Jr = J + k * lb.ncols;
        if (lb.eoff >= 0) for (int j = 0; j < 3; j++) gE[lb.eoff + j] += Jr[j] * lb.r[k];
        for (int q = 0; q < lb.nfb; q++) for (int j = 0; j < lb.fsz[q]; j++) gf[lb.foff[q] + j] += Jr[lb.fcol[q] + j] * lb.r[k];
      }
    }
  }
  // ImplicitSchurComplement::Init: landmark block inverses, block diagonal of F'F + D^2, rhs.
  void init_schur(std::vector<double>& rhs) {
    std::fill(pre.begin(), pre.end(), 0.0);
    // block diagonal of F'F: need map from foff to f block index
    std::vector<int> blk_of_off(nf, -1);
    for (size_t i = 0; i < fb_off.size(); i++) blk_of_off[fb_off[i]] = (int)i;
    for (auto& lb : L.lb) {
      const double* J = &L.J[lb.j0];
      for (int q = 0; q < lb.nfb; q++) {
        int bi = blk_of_off[lb.foff[q]]; int n = lb.fsz[q]; double* M = &pre[fb_pre[bi]];
        for (int k = 0; k < lb.nres; k++) { const double* Jr = J + k * lb.ncols + lb.fcol[q]; for (int a = 0; a < n; a++) for (int b2 = 0; b2 < n; b2++) M[a * n + b2] += Jr[a] * Jr[b2]; }
      }
    }
    for (size_t i = 0; i < fb_off.size(); i++) {
      int n = fb_sz[i]; double* M = &pre[fb_pre[i]];
      for (int a = 0; a < n; a++) M[a * n + a] += Dsq_f[fb_off[i] + a];
      invert_spd(M, n);
    }
    Hinv.assign((size_t)npts * 9, 0.0); ge.assign((size_t)npts * 3, 0.0);
    rhs.assign(nf, 0.0);
    // F' r
    for (auto& lb : L.lb) {
      const double* J = &L.J[lb.j0];
      for (int k = 0; k < lb.nres; k++) for (int q = 0; q < lb.nfb; q++) for (int j = 0; j < lb.fsz[q]; j++) rhs[lb.foff[q] + j] += J[k * lb.ncols + lb.fcol[q] + j] * lb.r[k];
    }
    for (int p = 0; p < npts; p++) {
      double H[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, g[3] = {0, 0, 0};
      for (int64_t bi : track_blocks[p]) {
        const LinBlock& lb = L.lb[bi]; const double* J = &L.J[lb.j0];
        for (int k = 0; k < lb.nres; k++) { const double* Jr = J + k * lb.ncols; for (int a = 0; a < 3; a++) { g[a] += Jr[a] * lb.r[k]; for (int b2 = 0; b2 < 3; b2++) H[a * 3 + b2] += Jr[a] * Jr[b2]; } }
      }
      for (int a = 0; a < 3; a++) H[a * 4] += Dsq_e[3 * p + a];
      inv3(H, &Hinv[9 * p]);
      for (int a = 0; a < 3; a++) ge[3 * p + a] = g[a];
      // rhs -= F'E Hinv E'r
      double v[3];
      for (int a = 0; a < 3; a++) v[a] = Hinv[9 * p + 3 * a] * g[0] + Hinv[9 * p + 3 * a + 1] * g[1] + Hinv[9 * p + 3 * a + 2] * g[2];
      for (int64_t bi : track_blocks[p]) {
        const LinBlock& lb = L.lb[bi]; const double* J = &L.J[lb.j0];
        for (int k = 0; k < lb.nres; k++) {
          const double* Jr = J + k * lb.ncols;
          double ev = Jr[0] * v[0] + Jr[1] * v[1] + Jr[2] * v[2];
          for (int q = 0; q < lb.nfb; q++) for (int j = 0; j < lb.fsz[q]; j++) rhs[lb.foff[q] + j] -= Jr[lb.fcol[q] + j] * ev;
        }
      }
    }
  }
  static void invert_spd(double* M, int n) {  // in-place inverse by Gauss-Jordan with partial pivoting
    std::vector<double> A(M, M + n * n), I(n * n, 0.0);
    for (int i = 0; i < n; i++) I[i * n + i] = 1;
    for (int c = 0; c < n; c++) {
      int piv = c; for (int r = c + 1; r < n; r++) if (std::fabs(A[r * n + c]) > std::fabs(A[piv * n + c])) piv = r;
      if (piv != c) for (int j = 0; j < n; j++) { std::swap(A[c * n + j], A[piv * n + j]); std::swap(I[c * n + j], I[piv * n + j]); }
      double d = 1.0 / A[c * n + c];
      for (int j = 0; j < n; j++) { A[c * n + j] *= d; I[c * n + j] *= d; }
      for (int r = 0; r < n; r++) if (r != c) { double f = A[r * n + c]; if (f != 0) for (int j = 0; j < n; j++) { A[r * n + j] -= f * A[c * n + j]; I[r * n + j] -= f * I[c * n + j]; } }
    }
    std::copy(I.begin(), I.end(), M);
  }
  // y = S x  (ImplicitSchurComplement::RightMultiply)
  void schur_mul(const std::vector<double>& x, std::vector<double>& y) {
    const int nt = omp_get_max_threads();
    std::vector<std::vector<double>> ys(nt, std::vector<double>(nf, 0.0));
#pragma omp parallel
    {
      std::vector<double>& yl = ys[omp_get_thread_num()];
#pragma omp for schedule(dynamic, 512)
      for (int p = 0; p < npts; p++) {
        double u[3] = {0, 0, 0};
        int nb = (int)track_blocks[p].size();
        static thread_local std::vector<double> fxs;
        if ((int)fxs.size() < 3 * nb) fxs.resize(3 * nb);
        double(*fxp)[3] = reinterpret_cast<double(*)[3]>(fxs.data());
        for (int t = 0; t < nb; t++) {
          const LinBlock& lb = L.lb[track_blocks[p][t]]; const double* J = &L.J[lb.j0];
          for (int k = 0; k < lb.nres; k++) {
            const double* Jr = J + k * lb.ncols; double s = 0;
            for (int q = 0; q < lb.nfb; q++) for (int j = 0; j < lb.fsz[q]; j++) s += Jr[lb.fcol[q] + j] * x[lb.foff[q] + j];
            fxp[t][k] = s;
            for (int a = 0; a < 3; a++) u[a] += Jr[a] * s;
          }
        }
        double v[3];
        for (int a = 0; a < 3; a++) v[a] = Hinv[9 * p + 3 * a] * u[0] + Hinv[9 * p + 3 * a + 1] * u[1] + Hinv[9 * p + 3 * a + 2] * u[2];
        for (int t = 0; t < nb; t++) {
          const LinBlock& lb = L.lb[track_blocks[p][t]]; const double* J = &L.J[lb.j0];
          for (int k = 0; k < lb.nres; k++) {
            const double* Jr = J + k * lb.ncols;
            double w = fxp[t][k] - (Jr[0] * v[0] + Jr[1] * v[1] + Jr[2] * v[2]);
            for (int q = 0; q < lb.nfb; q++) for (int j = 0; j < lb.fsz[q]; j++) yl[lb.foff[q] + j] += Jr[lb.fcol[q] + j] * w;
          }
        }
      }
#pragma omp for schedule(static)
      for (int64_t t = 0; t < (int64_t)free_blocks.size(); t++) {
        const LinBlock& lb = L.lb[free_blocks[t]]; const double* J = &L.J[lb.j0];
        for (int k = 0; k < lb.nres; k++) {
          const double* Jr = J + k * lb.ncols; double s = 0;
          for (int q = 0; q < lb.nfb; q++) for (int j = 0; j < lb.fsz[q]; j++) s += Jr[lb.fcol[q] + j] * x[lb.foff[q] + j];
          for (int q = 0; q < lb.nfb; q++) for (int j = 0; j < lb.fsz[q]; j++) yl[lb.foff[q] + j] += Jr[lb.fcol[q] + j] * s;
        }
      }
    }
    for (int j = 0; j < nf; j++) { double s = Dsq_f[j] * x[j]; for (int t = 0; t < nt; t++) s += ys[t][j]; y[j] = s; }
  }
  void precond(const std::vector<double>& r, std::vector<double>& z) {
    for (size_t i = 0; i < fb_off.size(); i++) {
      int n = fb_sz[i]; const double* M = &pre[fb_pre[i]];
      for (int a = 0; a < n; a++) { double s = 0; for (int b2 = 0; b2 < n; b2++) s += M[a * n + b2] * r[fb_off[i] + b2]; z[fb_off[i] + a] = s; }
    }
  }
  // ConjugateGradientsSolver::Solve (Ceres 1.14), r_tolerance = -1, Q-based stopping.
  int cg(const std::vector<double>& b, std::vector<double>& x, double q_tol, int max_it, int min_it) {
    std::vector<double> r(nf), z(nf), p(nf), q(nf), tmp(nf);
    auto dot = [&](const std::vector<double>& a, const std::vector<double>& c) { double s = 0; for (int i = 0; i < nf; i++) s += a[i] * c[i]; return s; };
    std::fill(x.begin(), x.end(), 0.0);
    double norm_b = std::sqrt(dot(b, b));
    if (norm_b == 0) return 0;
    r = b;
    double rho = 1.0, Q0 = 0.0;
    int it;
    for (it = 1; it <= max_it; it++) {
      precond(r, z);
      double last_rho = rho;
      rho = dot(r, z);
      if (rho == 0 || !std::isfinite(rho)) break;
      if (it == 1) p = z;
      else { double beta = rho / last_rho; if (beta == 0 || !std::isfinite(beta)) break; for (int i = 0; i < nf; i++) p[i] = z[i] + beta * p[i]; }
      schur_mul(p, q);
      double pq = dot(p, q);
      if (pq <= 0 || !std::isfinite(pq)) break;
      double alpha = rho / pq;
      if (!std::isfinite(alpha)) break;
      for (int i = 0; i < nf; i++) x[i] += alpha * p[i];
      if (it % 10 == 0) { schur_mul(x, tmp); for (int i = 0; i < nf; i++) r[i] = b[i] - tmp[i]; }  // residual_reset_period
      else for (int i = 0; i < nf; i++) r[i] -= alpha * q[i];
      double Q1 = 0; for (int i = 0; i < nf; i++) Q1 -= x[i] * (b[i] + r[i]);
      double zeta = it * (Q1 - Q0) / Q1;
      if (zeta < q_tol && it >= min_it) break;
      Q0 = Q1;
    }
    return std::min(it, max_it);
  }
};

static int solve(Problem& P, const mvba_solve_options& o, mvba_summary* S, mvba_iteration* its, int max_its, int jac_mode) {
  auto t0 = std::chrono::steady_clock::now();
  Solver sv(P);
  double setup_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();  // one-time structure analysis
  const int nf = sv.nf, ne = sv.ne;
  // free vector gather/scatter
  auto gather = [&](std::vector<double>& xf, std::vector<double>& xe) {
    xf.resize(nf); xe.resize(ne);
    for (auto& pb : P.pbs) if (pb.off >= 0) for (int j = 0; j < pb.size; j++) (pb.is_e ? xe : xf)[pb.off + j] = pb.ptr[j];
  };
  auto scatter = [&](const std::vector<double>& xf, const std::vector<double>& xe) {
    for (auto& pb : P.pbs) if (pb.off >= 0) for (int j = 0; j < pb.size; j++) {
      double v = (pb.is_e ? xe : xf)[pb.off + j];
      if (pb.bounded) v = std::min(std::max(v, pb.lo), pb.hi);
      pb.ptr[j] = v;
    }
  };
  std::vector<double> xf, xe, cf, ce, gf, gE;
  gather(xf, xe);
  linearize(P, jac_mode, sv.L);
  {
    auto t1 = std::chrono::steady_clock::now();
    sv.index_blocks();
    setup_s += std::chrono::duration<double>(std::chrono::steady_clock::now() - t1).count();
  }
  int njac = 1, nres = 0, nlin = 0, nsucc = 0;
  double cost = sv.L.cost;
  S->initial_cost = cost;
  sv.scale_f.assign(nf, 1.0); sv.scale_e.assign(ne, 1.0);
  auto gmax_unscaled = [&]() {
    // gradient in the original (unscaled) variables = g_scaled / scale
    double g = 0;
    for (int i = 0; i < nf; i++) g = std::max(g, std::fabs(gf[i] / sv.scale_f[i]));
    for (int i = 0; i < ne; i++) g = std::max(g, std::fabs(gE[i] / sv.scale_e[i]));
    return g;
  };
  if (o.jacobi_scaling) {
    sv.col_sqnorm(cf, ce);
    for (int i = 0; i < nf; i++) sv.scale_f[i] = 1.0 / (1.0 + std::sqrt(cf[i]));
    for (int i = 0; i < ne; i++) sv.scale_e[i] = 1.0 / (1.0 + std::sqrt(ce[i]));
    sv.apply_scaling();
  }
  sv.gradient(gf, gE);
  double gmax = gmax_unscaled();
  double radius = o.initial_trust_region_radius, decrease = 2.0;
  int term = MVBA_TERM_NO_CONVERGENCE;
  int iter = 0;
  if (its && max_its > 0) { its[0] = mvba_iteration{cost, 0, gmax, 0, 0, radius, 0, 1}; }
  if (gmax <= o.gradient_tolerance) term = MVBA_TERM_CONVERGENCE;
  std::vector<double> rhs(nf), df(nf), de(ne), xf2, xe2;
  int invalid_steps = 0;
  while (term == MVBA_TERM_NO_CONVERGENCE && iter < o.max_num_iterations) {
    iter++;
    sv.col_sqnorm(cf, ce);
    sv.Dsq_f.resize(nf); sv.Dsq_e.resize(ne);
    for (int i = 0; i < nf; i++) sv.Dsq_f[i] = std::min(std::max(cf[i], o.min_lm_diagonal), o.max_lm_diagonal) / radius;
    for (int i = 0; i < ne; i++) sv.Dsq_e[i] = std::min(std::max(ce[i], o.min_lm_diagonal), o.max_lm_diagonal) / radius;
    sv.init_schur(rhs);
    int cgit = sv.cg(rhs, df, o.eta, o.max_linear_solver_iterations, o.min_linear_solver_iterations);
    nlin += cgit;
    // back substitution: y_e = Hinv (E'r - E'F y_f);   step = -y
    for (int p = 0; p < sv.npts; p++) {
      double u[3] = {sv.ge[3 * p], sv.ge[3 * p + 1], sv.ge[3 * p + 2]};
      for (int64_t bi : sv.track_blocks[p]) {
        const LinBlock& lb = sv.L.lb[bi]; const double* J = &sv.L.J[lb.j0];
        for (int k = 0; k < lb.nres; k++) { const double* Jr = J + k * lb.ncols; double s = 0; for (int q = 0; q < lb.nfb; q++) for (int j = 0; j < lb.fsz[q]; j++) s += Jr[lb.fcol[q] + j] * df[lb.foff[q] + j]; for (int a = 0; a < 3; a++) u[a] -= Jr[a] * s; }
      }
      for (int a = 0; a < 3; a++) de[3 * p + a] = sv.Hinv[9 * p + 3 * a] * u[0] + sv.Hinv[9 * p + 3 * a + 1] * u[1] + sv.Hinv[9 * p + 3 * a + 2] * u[2];
    }
    for (int i = 0; i < nf; i++) df[i] = -df[i];
    for (int i = 0; i < ne; i++) de[i] = -de[i];
    // model cost change = -(J d)'(r + J d / 2)
    double mcc = 0;
    for (auto& lb : sv.L.lb) {
      const double* J = &sv.L.J[lb.j0];
      for (int k = 0; k < lb.nres; k++) {
        const double* Jr = J + k * lb.ncols; double s = 0;
        if (lb.eoff >= 0) for (int j = 0; j < 3; j++) s += Jr[j] * de[lb.eoff + j];
        for (int q = 0; q < lb.nfb; q++) for (int j = 0; j < lb.fsz[q]; j++) s += Jr[lb.fcol[q] + j] * df[lb.foff[q] + j];
        mcc -= s * (lb.r[k] + 0.5 * s);
      }
    }
    bool accepted = false;
    double cost_change = 0, rel = 0, step_norm = 0;
    if (!(mcc > 0) || !std::isfinite(mcc)) {
      radius /= decrease; decrease *= 2;  // invalid step
      if (++invalid_steps >= 5) { term = MVBA_TERM_FAILURE; }  // max_num_consecutive_invalid_steps, Ceres default
    } else {
      invalid_steps = 0;
      xf2 = xf; xe2 = xe;
      double sn = 0, xn = 0;
      for (int i = 0; i < nf; i++) { double d = df[i] * sv.scale_f[i]; xf2[i] += d; sn += d * d; xn += xf[i] * xf[i]; }
      for (int i = 0; i < ne; i++) { double d = de[i] * sv.scale_e[i]; xe2[i] += d; sn += d * d; xn += xe[i] * xe[i]; }
      step_norm = std::sqrt(sn); double x_norm = std::sqrt(xn);
      scatter(xf2, xe2);
      double cand = evaluate(P, true, nullptr); nres++;
      if (!std::isfinite(cand)) cand = std::numeric_limits<double>::max();
      cost_change = cost - cand;
      rel = cost_change / mcc;
      if (step_norm <= o.parameter_tolerance * (x_norm + o.parameter_tolerance)) { term = MVBA_TERM_CONVERGENCE; }
      else if (std::fabs(cost_change) <= o.function_tolerance * cost) { term = MVBA_TERM_CONVERGENCE; }
      if (rel > o.min_relative_decrease) {
        accepted = true; nsucc++;
        gather(xf, xe);  // picks up bound projection
        cost = cand;
        if (term == MVBA_TERM_NO_CONVERGENCE) {
          linearize(P, jac_mode, sv.L); njac++;
          if (o.jacobi_scaling) sv.apply_scaling();
          sv.gradient(gf, gE);
          gmax = gmax_unscaled();
          if (gmax <= o.gradient_tolerance) term = MVBA_TERM_CONVERGENCE;
        }
        double t = 2.0 * rel - 1.0;
        radius = radius / std::max(1.0 / 3.0, 1.0 - t * t * t);
        radius = std::min(o.max_trust_region_radius, radius);
        decrease = 2.0;
      } else {
        scatter(xf, xe);
        radius /= decrease; decrease *= 2;
      }
    }
    if (its && iter < max_its) its[iter] = mvba_iteration{cost, cost_change, gmax, step_norm, rel, radius, cgit, accepted ? 1 : 0};
    if (o.verbose) std::printf("oracle %3d cost %.12e change %.3e |g| %.3e |step| %.3e rho %.3e radius %.3e cg %d %s\n", iter, cost, cost_change, gmax, step_norm, rel, radius, cgit, accepted ? "ok" : "rejected");
    if (radius < o.min_trust_region_radius) { term = MVBA_TERM_CONVERGENCE; break; }
  }
  scatter(xf, xe);
  std::memset(S, 0, sizeof(*S));
  S->initial_cost = its && max_its > 0 ? its[0].cost : S->initial_cost;
  S->final_cost = cost; S->num_iterations = iter; S->num_successful_steps = nsucc; S->num_linear_iterations = nlin;
  S->termination = term; S->num_jacobian_evals = njac; S->num_residual_evals = nres;
  S->final_gradient_max_norm = gmax; S->final_radius = radius;
  S->seconds_total = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  S->seconds_build = setup_s;  // (oracle only) Solver construction + block indexing, so that bench.py can time the LM loop alone
  return 0;
}

}  // namespace oracle

using oracle::Problem;

extern "C" {

void* oracle_create(const mvba_config* cfg) {
  Problem* P = new Problem();
  P->cfg = *cfg;
  P->sensors.assign(cfg->num_sensors, oracle::SensorConst{{0, 0}, 0});
  P->dist_off.assign(cfg->num_sensors, 0);
  return P;
}
void oracle_destroy(void* h) { delete (Problem*)h; }

int oracle_set_sensor(void* h, int32_t t, double w, double hgt, int32_t nd) {
  Problem* P = (Problem*)h;
  P->sensors[t].half[0] = w / 2.0; P->sensors[t].half[1] = hgt / 2.0; P->sensors[t].nd = nd;  // camera_params.cc:109
  int o = 0; for (int i = 0; i < P->cfg.num_sensors; i++) { P->dist_off[i] = o; o += P->sensors[i].nd; }
  P->built = false;
  return 0;
}
int oracle_set_params(void* h, const mvba_params* p) { Problem* P = (Problem*)h; P->prm = *p; P->have_params = true; P->built = false; return 0; }
int oracle_set_cameras(void* h, int32_t n, const int32_t* sensor, const int32_t* beg, const int32_t* end, const double* stamp) {
  Problem* P = (Problem*)h;
  P->cam_sensor.assign(sensor, sensor + n); P->cam_beg.assign(beg, beg + n); P->cam_end.assign(end, end + n); P->cam_stamp.assign(stamp, stamp + n);
  P->built = false; return 0;
}
int oracle_set_observations(void* h, int64_t n, const int32_t* pid, const int32_t* cid, const float* pix, const uint8_t* has_depth,
                            const float* depth_xyz, const uint8_t* has_mesh, const double* mesh_xyz) {
  Problem* P = (Problem*)h;
  P->nobs = n; P->opid.assign(pid, pid + n); P->ocid.assign(cid, cid + n); P->opix.assign(pix, pix + 2 * n);
  P->ohasd.clear(); P->odepth.clear(); P->ohasm.clear(); P->omesh.clear();
  if (has_depth && depth_xyz) { P->ohasd.assign(has_depth, has_depth + n); P->odepth.assign(depth_xyz, depth_xyz + 3 * n); }
  if (has_mesh && mesh_xyz) { P->ohasm.assign(has_mesh, has_mesh + n); P->omesh.assign(mesh_xyz, mesh_xyz + 3 * n); }
  P->built = false; return 0;
}
int oracle_set_point_priors(void* h, int32_t kind, int64_t n, const int32_t* pid, const double* ref, double w, double th) {
  Problem* P = (Problem*)h;
  auto& pr = P->prior[kind == MVBA_BLOCK_MESH_TRI ? 0 : 1];
  pr.pid.assign(pid, pid + n); pr.ref.assign(ref, ref + 3 * n); pr.w = w; pr.th = th;
  P->built = false; return 0;
}
int oracle_set_float_mask(void* h, const mvba_float_mask* m) {
  Problem* P = (Problem*)h; int T = P->cfg.num_sensors;
  auto cp = [&](std::vector<uint8_t>& d, const uint8_t* s) { if (s) d.assign(s, s + T); else d.clear(); };
  cp(P->m_focal, m->focal); cp(P->m_center, m->center); cp(P->m_dist, m->distortion); cp(P->m_rig, m->rig); cp(P->m_d2i, m->depth_to_image); cp(P->m_poses, m->poses);
  P->float_offsets = m->float_timestamp_offsets; P->float_scale = m->float_scale; P->have_mask = true; P->built = false; return 0;
}
static void ensure(Problem* P) { if (!P->built) oracle::build(*P); }
int oracle_num_residuals(void* h, int64_t* nres, int64_t* nblocks) { Problem* P = (Problem*)h; ensure(P); if (nres) *nres = P->nres; if (nblocks) *nblocks = (int64_t)P->rbs.size(); return 0; }
int oracle_residual_layout(void* h, int32_t* kind, int64_t* row0, int64_t* src) {
  Problem* P = (Problem*)h; ensure(P);
  for (size_t i = 0; i < P->rbs.size(); i++) { if (kind) kind[i] = P->rbs[i].kind; if (row0) row0[i] = P->rbs[i].row0; if (src) src[i] = P->rbs[i].src; }
  return 0;
}
int oracle_num_free(void* h, int32_t* nf, int32_t* ne) { Problem* P = (Problem*)h; ensure(P); *nf = P->nf; *ne = P->ne; return 0; }
int oracle_evaluate(void* h, int32_t apply_loss, double* cost, double* residuals) {
  Problem* P = (Problem*)h; ensure(P);
  double c = oracle::evaluate(*P, apply_loss != 0, residuals);
  if (cost) *cost = c;
  return 0;
}
static int kind_width(Problem* P, int kind) {
  int ndmax = 0; for (auto& s : P->sensors) ndmax = std::max(ndmax, s.nd);
  switch (kind) {
    case MVBA_BLOCK_CAM: return 28 + ndmax;
    case MVBA_BLOCK_DEPTH_TRI: return 21 + P->nd2i() + 1 + 3 + 1;
    case MVBA_BLOCK_DEPTH_MESH: return 21 + P->nd2i() + 1 + 1;
    default: return 3;
  }
}
int oracle_jacobian_width(void* h, int32_t kind, int32_t* w) { *w = kind_width((Problem*)h, kind); return 0; }
// Same layout as mvba_evaluate_jacobian; mode as in jac_block.
int oracle_evaluate_jacobian(void* h, int32_t kind, int64_t first, int64_t count, int32_t mode, double* residuals, double* jac) {
  Problem* P = (Problem*)h; ensure(P);
  const int W = kind_width(P, kind);
  std::vector<int64_t> idx;
  for (size_t i = 0; i < P->rbs.size(); i++) if (P->rbs[i].kind == kind) idx.push_back((int64_t)i);
  if (first < 0 || first + count > (int64_t)idx.size()) return -1;
#pragma omp parallel for schedule(dynamic, 64)
  for (int64_t q = 0; q < count; q++) {
    const oracle::RBlock& b = P->rbs[idx[first + q]];
    int n = 0; for (int k = 0; k < b.np; k++) n += P->pbs[b.pb[k]].size;
    double r[3]; std::vector<double> J((size_t)3 * std::max(n, 1));
    oracle::jac_block(*P, b, mode, false, r, J.data(), n);
    for (int i = 0; i < b.nres; i++) {
      if (residuals) residuals[q * b.nres + i] = r[i];
      double* out = jac + ((size_t)q * b.nres + i) * W;
      for (int j = 0; j < W; j++) out[j] = 0;
      for (int j = 0; j < n; j++) out[j] = J[(size_t)i * n + j];
    }
  }
  return 0;
}
int oracle_solve(void* h, const mvba_solve_options* o, mvba_summary* s, mvba_iteration* its, int32_t max_its, int32_t jac_mode, int32_t num_threads) {
  Problem* P = (Problem*)h; ensure(P);
  if (num_threads > 0) omp_set_num_threads(num_threads);
  return oracle::solve(*P, *o, s, its, max_its, jac_mode);
}
// Timed CPU baseline legs: seconds for one Jacobian evaluation (linearize) over all blocks and
// for one cost-only evaluation, on num_threads host threads.
int oracle_time_linearize(void* h, int32_t jac_mode, int32_t num_threads, double* sec_jac, double* sec_cost) {
  Problem* P = (Problem*)h; ensure(P);
  if (num_threads > 0) omp_set_num_threads(num_threads);
  oracle::Lin L;
  auto t0 = std::chrono::steady_clock::now();
  oracle::linearize(*P, jac_mode, L);
  auto t1 = std::chrono::steady_clock::now();
  oracle::evaluate(*P, true, nullptr);
  auto t2 = std::chrono::steady_clock::now();
  *sec_jac = std::chrono::duration<double>(t1 - t0).count();
  *sec_cost = std::chrono::duration<double>(t2 - t1).count();
  return 0;
}
int oracle_max_threads(void) { return omp_get_max_threads(); }
void oracle_cauchy(double a, double s, double* rho, double* sqrt_rho1) { oracle::cauchy(a, s, rho, sqrt_rho1); }
// array <-> transform helpers exported for the host-logic tests.
void oracle_array_to_rigid(const double* a, double* m12) { auto A = oracle::array_to_rigid<double>(a); for (int i = 0; i < 3; i++) { for (int j = 0; j < 3; j++) m12[4 * i + j] = A.m[i][j]; m12[4 * i + 3] = A.t[i]; } }
void oracle_rigid_to_array(const double* m12, double* a) { oracle::Aff<double> A; for (int i = 0; i < 3; i++) { for (int j = 0; j < 3; j++) A.m[i][j] = m12[4 * i + j]; A.t[i] = m12[4 * i + 3]; } oracle::rigid_to_array(A, a); }
void oracle_world_to_cam(const double* beg, const double* end, const double* r2c, double bs, double es, double off, double cs, double* m12, int* ok) {
  bool k = true; auto A = oracle::calc_world_to_cam_trans<double>(beg, end, r2c, bs, es, off, cs, &k);
  for (int i = 0; i < 3; i++) { for (int j = 0; j < 3; j++) m12[4 * i + j] = A.m[i][j]; m12[4 * i + 3] = A.t[i]; } *ok = k;
}
void oracle_distort(double w, double hgt, int nd, double f, const double* c, const double* dist, const double* u, double* out) {
  oracle::SensorConst s{{w / 2, hgt / 2}, nd}; double dummy = 0;
  oracle::distort_undist_c_to_dist<double>(s, f, c, nd > 0 ? dist : &dummy, u, out);
}

// ---- per-track steps around a pass; same signatures as the mvba_ entry points ----
// world_to_cam of the image of CAM block b at the current parameters (calc_world_to_cam_trans, rig_calibrator.cc:362-390).
static oracle::Aff<double> block_world_to_cam(const Problem& P, const oracle::RBlock& b) {
  bool ok = true;
  return oracle::calc_world_to_cam_trans<double>(P.pbs[b.pb[0]].ptr, P.pbs[b.pb[1]].ptr, P.pbs[b.pb[2]].ptr, b.stamps[0], b.stamps[1],
                                                 P.pbs[b.pb[4]].ptr[0], b.stamps[2], &ok);
}
// CAM blocks in observation order (b.src = observation index)
static std::vector<const oracle::RBlock*> cam_blocks(const Problem& P) {
  std::vector<const oracle::RBlock*> out((size_t)P.nobs, nullptr);
  for (auto& b : P.rbs) if (b.kind == MVBA_BLOCK_CAM) out[(size_t)b.src] = &b;
  return out;
}
// multiViewTriangulation, interest_point.cc:649-722
int oracle_triangulate(void* h, uint8_t* inlier, double* xyz) {
  Problem* P = (Problem*)h; ensure(P);
  auto cb = cam_blocks(*P);
  int64_t i = 0;
  while (i < P->nobs) {
    const int pid = P->opid[i];
    int64_t j = i;
    while (j < P->nobs && P->opid[j] == pid) j++;
    std::vector<oracle::View34> views;
    for (int64_t q = i; q < j; q++) {
      if (!inlier[q]) continue;
      const oracle::RBlock& b = *cb[q];
      const oracle::SensorConst& s = P->sensors[b.sensor];
      const double f = P->pbs[b.pb[5]].ptr[0];
      const double* c = P->pbs[b.pb[6]].ptr;
      double dummy = 0;
      const double* dist = s.nd > 0 ? P->pbs[b.pb[7]].ptr : &dummy;
      const double dc[2] = {b.meas[0] - s.half[0], b.meas[1] - s.half[1]};  // Convert<DISTORTED, DISTORTED_C>
      oracle::View34 v;
      oracle::undistort_centered(s, f, c, dist, dc, v.p);
      oracle::Aff<double> w2c = block_world_to_cam(*P, b);
      for (int a = 0; a < 3; a++) {  // P = K [R | t], K = diag(f, f, 1)   (P_From_KRt)
        const double k = a < 2 ? f : 1.0;
        for (int e = 0; e < 3; e++) v.P[a][e] = k * w2c.m[a][e];
        v.P[a][3] = k * w2c.t[a];
      }
      views.push_back(v);
    }
    if (views.size() < 2) {
      for (int64_t q = i; q < j; q++) inlier[q] = 0;
    } else {
      double X[3];
      oracle::triangulate_views(views, X);
      for (int a = 0; a < 3; a++) xyz[3 * (size_t)pid + a] = X[a];
      if (!std::isfinite(X[0]) || !std::isfinite(X[1]) || !std::isfinite(X[2]))
        for (int64_t q = i; q < j; q++) inlier[q] = 0;
    }
    i = j;
  }
  return 0;
}
// flagOutliersByTriAngleAndReprojErr, rig_calibrator.cc:1045-1154 (residuals = evalResiduals at the current parameters)
int oracle_flag_outliers(void* h, double min_triangulation_angle, double max_reprojection_error, uint8_t* inlier, int64_t counts[4]) {
  Problem* P = (Problem*)h; ensure(P);
  auto cb = cam_blocks(*P);
  std::vector<double> res((size_t)P->nres);
  oracle::evaluate(*P, false, res.data());
  int64_t n_total = 0, n_angle = 0;
  auto centre = [&](int64_t q, double c[3]) {  // world_to_cam^-1 * 0
    oracle::Aff<double> inv = oracle::aff_inverse(block_world_to_cam(*P, *cb[q]));
    const double z[3] = {0, 0, 0};
    oracle::aff_apply(inv, z, c);
  };
  int64_t i = 0;
  while (i < P->nobs) {
    const int pid = P->opid[i];
    int64_t j = i;
    while (j < P->nobs && P->opid[j] == pid) j++;
    const double* X = P->prm.xyz + 3 * (size_t)pid;
    double max_angle = 0.0;
    for (int64_t q1 = i; q1 < j; q1++) {
      if (!inlier[q1]) continue;
      n_total++;
      double c1[3], r1[3];
      centre(q1, c1);
      for (int a = 0; a < 3; a++) r1[a] = X[a] - c1[a];
      double n1 = std::sqrt(r1[0] * r1[0] + r1[1] * r1[1] + r1[2] * r1[2]);
      if (n1 > 0) for (int a = 0; a < 3; a++) r1[a] /= n1;
      for (int64_t q2 = i; q2 < j; q2++) {
        if (P->ocid[q2] <= P->ocid[q1] || !inlier[q2]) continue;
        double c2[3], r2[3];
        centre(q2, c2);
        for (int a = 0; a < 3; a++) r2[a] = X[a] - c2[a];
        double n2 = std::sqrt(r2[0] * r2[0] + r2[1] * r2[1] + r2[2] * r2[2]);
        if (n2 > 0) for (int a = 0; a < 3; a++) r2[a] /= n2;
        double ang = (180.0 / M_PI) * std::acos(r1[0] * r2[0] + r1[1] * r2[1] + r1[2] * r2[2]);
        if (std::isnan(ang) || std::isinf(ang)) continue;
        max_angle = std::max(max_angle, ang);
      }
    }
    if (!(max_angle >= min_triangulation_angle))
      for (int64_t q = i; q < j; q++) if (inlier[q]) { inlier[q] = 0; n_angle++; }
    i = j;
  }
  int64_t n_total2 = 0, n_reproj = 0;
  for (int64_t q = 0; q < P->nobs; q++) {
    if (!inlier[q]) continue;
    n_total2++;
    const double rx = res[(size_t)cb[q]->row0], ry = res[(size_t)cb[q]->row0 + 1];
    const bool good = std::sqrt(rx * rx + ry * ry) <= max_reprojection_error;
    if (!good) { inlier[q] = 0; n_reproj++; }
  }
  if (counts) { counts[0] = n_total; counts[1] = n_angle; counts[2] = n_total2; counts[3] = n_reproj; }
  return 0;
}
// flagOutlierByExclusionDist, rig_calibrator.cc:1003-1038
int oracle_flag_exclusion(void* h, const double* image_size, const double* crop_size, uint8_t* inlier) {
  Problem* P = (Problem*)h;
  for (int64_t q = 0; q < P->nobs; q++) {
    const int t = P->cam_sensor[P->ocid[q]];
    if (std::abs((double)P->opix[2 * q] - image_size[2 * t] / 2.0) > crop_size[2 * t] / 2.0 ||
        std::abs((double)P->opix[2 * q + 1] - image_size[2 * t + 1] / 2.0) > crop_size[2 * t + 1] / 2.0)
      inlier[q] = 0;
  }
  return 0;
}
// meshTriangulations, texture_processing.cc:1566-1625
int oracle_mesh_intersect(void* h, int64_t num_vertices, const float* vertices, int64_t num_faces, const uint32_t* faces, double min_ray_dist,
                          double max_ray_dist, const uint8_t* inlier, uint8_t* has_mesh, double* mesh_xyz, double* pid_mesh_xyz) {
  Problem* P = (Problem*)h; ensure(P);
  auto cb = cam_blocks(*P);
  const double bad = 1e100;
  int64_t i = 0;
  while (i < P->nobs) {
    const int pid = P->opid[i];
    int64_t j = i;
    while (j < P->nobs && P->opid[j] == pid) j++;
    double avg[3] = {0, 0, 0}; int n = 0;
    for (int64_t q = i; q < j; q++) {
      has_mesh[q] = 0;
      for (int a = 0; a < 3; a++) mesh_xyz[3 * q + a] = bad;
      if (!inlier[q]) continue;
      const oracle::RBlock& b = *cb[q];
      const oracle::SensorConst& s = P->sensors[b.sensor];
      const double f = P->pbs[b.pb[5]].ptr[0];
      double dummy = 0;
      const double dc[2] = {b.meas[0] - s.half[0], b.meas[1] - s.half[1]};
      double u[2], out[3];
      oracle::undistort_centered(s, f, P->pbs[b.pb[6]].ptr, s.nd > 0 ? P->pbs[b.pb[7]].ptr : &dummy, dc, u);
      if (oracle::ray_mesh_brute(u, f, block_world_to_cam(*P, b), vertices, faces, num_faces, min_ray_dist, max_ray_dist, out)) {
        has_mesh[q] = 1;
        for (int a = 0; a < 3; a++) { mesh_xyz[3 * q + a] = out[a]; avg[a] += out[a]; }
        n++;
      }
    }
    if (pid_mesh_xyz) for (int a = 0; a < 3; a++) pid_mesh_xyz[3 * (size_t)pid + a] = n >= 1 ? avg[a] / n : bad;
    i = j;
  }
  return 0;
}
// UndistortCentered of one DISTORTED pixel, exported for the golden-vector tests (cv2.undistortPoints).
void oracle_undistort(double w, double hgt, int nd, double f, const double* c, const double* dist, const double* pix, double* out) {
  oracle::SensorConst s{{w / 2, hgt / 2}, nd}; double dummy = 0;
  const double dc[2] = {pix[0] - w / 2, pix[1] - hgt / 2};
  oracle::undistort_centered(s, f, c, nd > 0 ? dist : &dummy, dc, out);
}
// fitRpcDist-style fit (rpc_fit.h): coeffs[2 (deg + 1)(deg + 2) - 2]; stats = {initial cost, final cost, iterations, successful steps}
int oracle_rpc_fit(int64_t n, const double* in_xy, const double* out_xy, int32_t deg, int32_t num_iterations, double parameter_tolerance,
                   int32_t num_threads, double* coeffs, double* stats) {
  if (n <= 0 || deg < 1 || deg > 7) return -1;
  int prev = omp_get_max_threads();
  if (num_threads > 0) omp_set_num_threads(num_threads);
  std::vector<double> c; oracle::RpcFitStats st{};
  oracle::rpc_fit(n, in_xy, out_xy, deg, num_iterations, parameter_tolerance, c, &st);
  omp_set_num_threads(prev);
  for (size_t i = 0; i < c.size(); i++) coeffs[i] = c[i];
  if (stats) { stats[0] = st.initial_cost; stats[1] = st.final_cost; stats[2] = st.iterations; stats[3] = st.successful; }
  return 0;
}
// float32 rays against a mesh: the restated triangle test over all triangles (checked against the reference's BVHTree in tests/)
int64_t oracle_ray_mesh(int64_t nv, const float* verts, int64_t nf, const uint32_t* faces, int64_t n, const float* origin, const float* dir,
                        float tmin, float tmax, uint8_t* hit, float* t_out) {
  int64_t hits = 0;
#pragma omp parallel for reduction(+ : hits) schedule(static)
  for (int64_t i = 0; i < n; i++) {
    float best;
    bool ok = oracle::ray_mesh_nearest(origin + 3 * i, dir + 3 * i, tmin, tmax, verts, faces, nf, &best);
    hit[i] = ok ? 1 : 0; t_out[i] = best; hits += ok;
  }
  return hits;
}
// acc::intersect(ray, tri) restated (tracks.h ray_tri) for ONE triangle: 1 and the distance when the hit is accepted
int oracle_ray_tri(const float* origin, const float* dir, float tmin, float tmax, const float* a, const float* b, const float* c, float* t_out) {
  float t = 0;
  bool ok = oracle::ray_tri(origin, dir, tmin, tmax, a, b, c, &t);
  *t_out = t;
  return ok ? 1 : 0;
}
void oracle_compute_rpc(const double* p, const double* coeffs, int32_t n, double* out) { oracle::compute_rpc(p, coeffs, n, out); }
}
