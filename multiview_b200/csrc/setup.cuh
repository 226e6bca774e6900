// setup.cuh -- host logic that turns the caller's description of one pass into the device problem:
// residual-block order and row layout (reference rig_calibrator.cc:1614-1903), parameter-block
// constancy (:1702-1752, 1787-1793, 1834-1839) and the reduced-system column numbering.
#pragma once
#include <thread>
#include <algorithm>
#include <cmath>

#include "handle.cuh"
#include "kernels_eval.cuh"

namespace mvba {

static inline int rpc_degree_of(int num_dist) {  // rpc_distortion.cc:43-45
  return (int)std::round(std::sqrt(2.0 * num_dist + 5.0) / 2.0 - 1.5);
}

int allreduce_flags(mvba_problem* h, std::vector<uint8_t>& flags);  // solver.cuh: element-wise OR across ranks
int rank_agree(mvba_problem* h, int rc);                               // solver.cuh: collective agreement on a return code
// Scope guard for a function with rank-local early returns ahead of a collective: pass() is the agreement point on the
// normal path; leaving the scope without it (an early `return fail(...)`) reports the failure to the waiting peers.
struct AgreePoint {
  mvba_problem* h; bool passed = false;
  explicit AgreePoint(mvba_problem* hh) : h(hh) {}
  int pass() { passed = true; return rank_agree(h, MVBA_OK); }
  ~AgreePoint() { if (!passed) rank_agree(h, MVBA_EINVAL); }
};

void solver_destroy(mvba_problem* h);                                  // solver.cuh

inline int setup(mvba_problem* h) {
  if (h->built) return MVBA_OK;
  AgreePoint agree_layout(h);  // rank-local validation failures below must not leave the peers alone in the collective
  solver_destroy(h);  // structure tables of an earlier layout
  if (!h->have_params || !h->have_cams || !h->have_obs)
    return fail(MVBA_ESTATE, "mvba: set_params, set_cameras and set_observations must precede evaluation");
  const mvba_config& c = h->cfg;
  const int T = c.num_sensors, C = c.num_cameras, P = c.num_landmarks;
  const int npose = c.no_rig ? C : c.num_ref_frames;
  const int nd2i = h->nd2i();
  if (T > MVBA_MAX_SENSORS) return fail(MVBA_EINVAL, "mvba: too many sensors");
  if ((int)h->cam_sensor.size() != C) return fail(MVBA_EINVAL, "mvba: camera count mismatch");
  View& v = h->view;
  std::memset(&v, 0, sizeof(v));
  v.T = T; v.R = c.num_ref_frames; v.C = C; v.P = P; v.npose = npose; v.no_rig = c.no_rig; v.nd2i = nd2i;
  v.N = h->nobs; v.robust_th = c.robust_threshold; v.w_dtri = c.depth_tri_weight; v.w_dmesh = c.depth_mesh_weight;

  // ---- validate observations: pid ascending, indices in range ----
  for (int64_t i = 0; i < h->nobs; i++) {
    int p = h->opid[i], q = h->ocid[i];
    if (p < 0 || p >= P || q < 0 || q >= C) return fail(MVBA_EINVAL, "mvba: observation index out of range");
    if (i > 0 && h->opid[i - 1] > p) return fail(MVBA_EINVAL, "mvba: observations must be sorted by pid");
  }
  const bool have_depth = !h->ohasd.empty();
  const bool have_mesh = !h->ohasm.empty();

  // ---- cameras: pose block indices and bracketing stamps (rig_calibrator.cc:1631-1666) ----
  h->h_cam_p0.assign(C, 0); h->h_cam_p1.assign(C, -1); h->h_cam_bs.assign(C, 0); h->h_cam_es.assign(C, 0);
  for (int q = 0; q < C; q++) {
    int t = h->cam_sensor[q];
    if (t < 0 || t >= T) return fail(MVBA_EINVAL, "mvba: camera sensor index out of range");
    if (!c.no_rig) {
      int b = h->cam_beg[q], e = h->cam_end[q];
      if (b < 0 || b >= npose || e < 0 || e >= npose) return fail(MVBA_EINVAL, "mvba: bracketing index out of range");
      h->h_cam_p0[q] = b;
      h->h_cam_p1[q] = (t == 0 || b == e) ? -1 : e;
      h->h_cam_bs[q] = h->prm.ref_stamps[b]; h->h_cam_es[q] = h->prm.ref_stamps[e];
    } else {
      h->h_cam_p0[q] = q; h->h_cam_p1[q] = -1;
      h->h_cam_bs[q] = h->h_cam_es[q] = h->cam_stamp[q];
    }
  }

  // ---- which blocks are used by at least one residual ----
  std::vector<uint8_t> cam_used(C, 0), pose_used(npose, 0), sens_used(T, 0), sens_depth(T, 0), lm_used(P, 0);
  for (int64_t i = 0; i < h->nobs; i++) { cam_used[h->ocid[i]] = 1; lm_used[h->opid[i]] = 1; }
  for (int q = 0; q < C; q++) if (cam_used[q]) {
    pose_used[h->h_cam_p0[q]] = 1;
    if (h->h_cam_p1[q] >= 0) pose_used[h->h_cam_p1[q]] = 1;
    sens_used[h->cam_sensor[q]] = 1;
  }

  // ---- residual layout in reference order ----
  h->h_obs_row0.assign(h->nobs, 0);
  h->h_dtri_obs.clear(); h->h_dtri_row0.clear(); h->h_dmesh_obs.clear(); h->h_dmesh_row0.clear();
  h->h_trk_obs0.assign(P + 1, 0); h->h_trk_dtri0.assign(P + 1, 0); h->h_trk_dmesh0.assign(P + 1, 0);
  h->h_trk_pm0.assign(P + 1, 0); h->h_trk_pt0.assign(P + 1, 0);
  const int64_t n_pm = (int64_t)h->prior[0].pid.size(), n_pt = (int64_t)h->prior[1].pid.size();
  h->h_prior_pid.resize(n_pm + n_pt); h->h_prior_ref.resize(3 * (n_pm + n_pt));
  h->h_prior_w.resize(n_pm + n_pt); h->h_prior_th.resize(n_pm + n_pt); h->h_prior_row0.assign(n_pm + n_pt, 0);
  for (int k = 0; k < 2; k++) {
    const auto& pr = h->prior[k];
    int64_t base = k == 0 ? 0 : n_pm;
    for (size_t j = 0; j < pr.pid.size(); j++) {
      if (pr.pid[j] < 0 || pr.pid[j] >= P) return fail(MVBA_EINVAL, "mvba: prior pid out of range");
      if (j > 0 && pr.pid[j - 1] >= pr.pid[j]) return fail(MVBA_EINVAL, "mvba: priors must have strictly ascending pid");
      h->h_prior_pid[base + j] = pr.pid[j];
      for (int a = 0; a < 3; a++) h->h_prior_ref[3 * (base + j) + a] = pr.ref[3 * j + a];
      h->h_prior_w[base + j] = pr.w; h->h_prior_th[base + j] = pr.th;
    }
  }
  int64_t row = 0, blocks = 0, i = 0;
  size_t pit[2] = {0, 0};
  for (int p = 0; p < P; p++) {
    h->h_trk_obs0[p] = i; h->h_trk_dtri0[p] = (int64_t)h->h_dtri_obs.size(); h->h_trk_dmesh0[p] = (int64_t)h->h_dmesh_obs.size();
    h->h_trk_pm0[p] = (int64_t)pit[0]; h->h_trk_pt0[p] = n_pm + (int64_t)pit[1];
    bool any = false;
    for (; i < h->nobs && h->opid[i] == p; i++) {
      any = true;
      h->h_obs_row0[i] = row; row += 2; blocks++;
      bool hd = have_depth && h->ohasd[i];
      if (c.depth_tri_weight > 0 && hd) { h->h_dtri_obs.push_back(i); h->h_dtri_row0.push_back(row); row += 3; blocks++; sens_depth[h->cam_sensor[h->ocid[i]]] = 1; }
      bool hm = have_mesh && h->ohasm[i];
      if (c.depth_mesh_weight > 0 && hm && hd) { h->h_dmesh_obs.push_back(i); h->h_dmesh_row0.push_back(row); row += 3; blocks++; sens_depth[h->cam_sensor[h->ocid[i]]] = 1; }
    }
    for (int k = 0; k < 2; k++) {
      const auto& pr = h->prior[k];
      if (pit[k] < pr.pid.size() && pr.pid[pit[k]] == p) {
        if (!any) return fail(MVBA_EINVAL, "mvba: prior on a track without inlier observations (rig_calibrator.cc:1845-1855)");
        h->h_prior_row0[(k == 0 ? 0 : n_pm) + (int64_t)pit[k]] = row; row += 3; blocks++;
        pit[k]++;
      }
    }
  }
  if (i != h->nobs) return fail(MVBA_EINVAL, "mvba: observations not grouped by pid");
  h->h_trk_obs0[P] = i; h->h_trk_dtri0[P] = (int64_t)h->h_dtri_obs.size(); h->h_trk_dmesh0[P] = (int64_t)h->h_dmesh_obs.size();
  h->h_trk_pm0[P] = n_pm; h->h_trk_pt0[P] = n_pm + n_pt;
  h->nres = row; h->nblocks = blocks;
  v.n_dtri = (int64_t)h->h_dtri_obs.size(); v.n_dmesh = (int64_t)h->h_dmesh_obs.size();
  v.n_prior = n_pm + n_pt; v.n_pm = n_pm;

  // Multi-rank: the column layout must be the same on every rank, so "used by at least one residual" means used on ANY rank
  // (a landmark shard that never sees an image would otherwise drop that image's pose block and shift every later column).
  { int rc = agree_layout.pass(); if (rc) return rc; }
  AgreePoint agree_end(h);
  if (h->nranks > 1) {
    std::vector<uint8_t> flags;
    flags.insert(flags.end(), pose_used.begin(), pose_used.end());
    flags.insert(flags.end(), sens_used.begin(), sens_used.end());
    flags.insert(flags.end(), sens_depth.begin(), sens_depth.end());
    int rc = allreduce_flags(h, flags); if (rc) return rc;
    std::copy(flags.begin(), flags.begin() + npose, pose_used.begin());
    std::copy(flags.begin() + npose, flags.begin() + npose + T, sens_used.begin());
    std::copy(flags.begin() + npose + T, flags.end(), sens_depth.begin());
  }

  // ---- parameter vector layout ----
  int o = 0;
  v.x_pose = o; o += 7 * npose;
  v.x_ident = o; o += 7;
  int doff = 0;
  for (int t = 0; t < T; t++) {
    SensorDev& s = v.sens[t];
    const SensorHost& sh = h->sensors[t];
    s.half[0] = sh.w / 2.0; s.half[1] = sh.h / 2.0;  // camera_params.cc:109
    s.nd = sh.nd; s.dist_off = doff; doff += sh.nd;
    s.rpc_deg = 0; s.nd_nz = sh.nd;
    if (sh.nd > 5) {
      if (sh.nd % 2) return fail(MVBA_EINVAL, "mvba: RPC distortion needs an even number of coefficients");
      int n = sh.nd / 2, deg = rpc_degree_of(n);
      if (deg <= 0 || deg > 6 || 2 * (deg + 1) * (deg + 2) - 2 != n) return fail(MVBA_EINVAL, "mvba: irregular distortion vector size");
      s.rpc_deg = deg; s.nd_nz = n;
    } else if (sh.nd == 2 || sh.nd == 3) {
      return fail(MVBA_EINVAL, "mvba: irregular distortion vector size");
    }
    s.x_rig = o; o += 7; s.x_off = o; o += 1; s.x_focal = o; o += 1; s.x_center = o; o += 2;
    s.x_dist = o; o += sh.nd; s.x_d2i = o; o += nd2i; s.x_scale = o; o += 1;
  }
  v.x_xyz = o; o += 3 * P;
  v.nx = o;

  // ---- constancy and reduced-system columns ----
  auto m = [&](const std::vector<uint8_t>& a, int t) { return !a.empty() && a[t] != 0; };
  h->h_pose_col.assign(npose, -1);
  int col = 0;
  for (int p = 0; p < npose; p++) {
    int s = c.no_rig ? h->cam_sensor[p] : 0;
    if (m(h->m_poses, s) && pose_used[p]) { h->h_pose_col[p] = col; col += 7; }
  }
  v.npose_cols = col;
  h->bounded = false;
  for (int t = 0; t < T; t++) {
    SensorDev& s = v.sens[t];
    auto take = [&](bool free_, int width) { if (!free_ || width == 0) return -1; int r = col; col += width; return r; };
    bool u = sens_used[t] != 0, ud = sens_depth[t] != 0;
    s.c_rig = take(u && m(h->m_rig, t) && t != 0 && !c.no_rig, 7);
    s.c_off = take(u && h->float_offsets && t != 0 && !c.no_rig, 1);
    s.c_focal = take(u && m(h->m_focal, t), 1);
    s.c_center = take(u && m(h->m_center, t), 2);
    s.c_dist = take(u && m(h->m_dist, t) && s.nd > 0, s.nd);
    s.c_d2i = take(ud && m(h->m_d2i, t), nd2i);
    s.c_scale = take(ud && h->float_scale && !c.affine_depth_to_image, 1);
    if (s.c_off >= 0 && h->prm.offset_lo && h->prm.offset_hi) h->bounded = true;
  }
  v.nf = col; v.B = col - v.npose_cols;
  // parameter scalar -> free index, bounds
  h->h_xcol.assign(v.nx, -1);
  h->h_lo.assign(v.nx, -INFINITY); h->h_hi.assign(v.nx, INFINITY);
  for (int p = 0; p < npose; p++) if (h->h_pose_col[p] >= 0) for (int k = 0; k < 7; k++) h->h_xcol[v.x_pose + 7 * p + k] = h->h_pose_col[p] + k;
  for (int t = 0; t < T; t++) {
    const SensorDev& s = v.sens[t];
    auto put = [&](int xo, int c0, int w) { if (c0 >= 0) for (int k = 0; k < w; k++) h->h_xcol[xo + k] = c0 + k; };
    put(s.x_rig, s.c_rig, 7); put(s.x_off, s.c_off, 1); put(s.x_focal, s.c_focal, 1); put(s.x_center, s.c_center, 2);
    put(s.x_dist, s.c_dist, s.nd); put(s.x_d2i, s.c_d2i, nd2i); put(s.x_scale, s.c_scale, 1);
    if (s.c_off >= 0 && h->bounded) { h->h_lo[s.x_off] = h->prm.offset_lo[t]; h->h_hi[s.x_off] = h->prm.offset_hi[t]; }
  }
  for (int p = 0; p < P; p++) if (lm_used[p]) for (int k = 0; k < 3; k++) h->h_xcol[v.x_xyz + 3 * p + k] = v.nf + 3 * p + k;

  // ---- upload the static problem ----
  MVBA_CUDA(cudaSetDevice(h->device));
  MVBA_CUDA(h->d_cam_sensor.upload(h->cam_sensor)); MVBA_CUDA(h->d_cam_p0.upload(h->h_cam_p0)); MVBA_CUDA(h->d_cam_p1.upload(h->h_cam_p1));
  MVBA_CUDA(h->d_cam_bs.upload(h->h_cam_bs)); MVBA_CUDA(h->d_cam_es.upload(h->h_cam_es)); MVBA_CUDA(h->d_cam_cs.upload(h->cam_stamp));
  MVBA_CUDA(h->d_pose_col.upload(h->h_pose_col));
  MVBA_CUDA(h->d_pid.upload(h->opid)); MVBA_CUDA(h->d_cid.upload(h->ocid)); MVBA_CUDA(h->d_pix.upload(h->opix));
  {
    std::vector<int4> orec(h->ocid.size());
    for (size_t i = 0; i < orec.size(); i++) {
      int4 r; r.x = h->ocid[i]; r.y = h->cam_sensor[r.x];
      std::memcpy(&r.z, &h->opix[2 * i], 4); std::memcpy(&r.w, &h->opix[2 * i + 1], 4);
      orec[i] = r;
    }
    MVBA_CUDA(h->d_orec.upload(orec));
  }
  MVBA_CUDA(h->d_obs_row0.upload(h->h_obs_row0));
  MVBA_CUDA(h->d_dtri_obs.upload(h->h_dtri_obs)); MVBA_CUDA(h->d_dtri_row0.upload(h->h_dtri_row0));
  MVBA_CUDA(h->d_dmesh_obs.upload(h->h_dmesh_obs)); MVBA_CUDA(h->d_dmesh_row0.upload(h->h_dmesh_row0));
  if (v.n_dtri + v.n_dmesh > 0) { MVBA_CUDA(h->d_depth.upload(h->odepth)); }
  if (v.n_dmesh > 0) { MVBA_CUDA(h->d_mesh.upload(h->omesh)); }
  MVBA_CUDA(h->d_prior_pid.upload(h->h_prior_pid)); MVBA_CUDA(h->d_prior_ref.upload(h->h_prior_ref));
  MVBA_CUDA(h->d_prior_w.upload(h->h_prior_w)); MVBA_CUDA(h->d_prior_th.upload(h->h_prior_th)); MVBA_CUDA(h->d_prior_row0.upload(h->h_prior_row0));
  MVBA_CUDA(h->d_trk_obs0.upload(h->h_trk_obs0)); MVBA_CUDA(h->d_trk_dtri0.upload(h->h_trk_dtri0)); MVBA_CUDA(h->d_trk_dmesh0.upload(h->h_trk_dmesh0));
  MVBA_CUDA(h->d_trk_pm0.upload(h->h_trk_pm0)); MVBA_CUDA(h->d_trk_pt0.upload(h->h_trk_pt0));
  MVBA_CUDA(h->d_x.alloc(v.nx)); MVBA_CUDA(h->d_xnew.alloc(v.nx));
  MVBA_CUDA(h->d_tab.alloc(C)); MVBA_CUDA(h->d_dtab.alloc(T)); MVBA_CUDA(h->d_coef.alloc(C));
  MVBA_CUDA(h->d_fov.alloc(2 * (size_t)T)); v.fov_tab = h->d_fov.p;
  MVBA_CUDA(h->d_scalars.alloc(64)); MVBA_CUDA(h->d_err.alloc(1));
  v.cam_sensor = h->d_cam_sensor.p; v.cam_p0 = h->d_cam_p0.p; v.cam_p1 = h->d_cam_p1.p;
  v.cam_bs = h->d_cam_bs.p; v.cam_es = h->d_cam_es.p; v.cam_cs = h->d_cam_cs.p; v.pose_col = h->d_pose_col.p;
  v.pid = h->d_pid.p; v.cid = h->d_cid.p; v.pix = reinterpret_cast<const float2*>(h->d_pix.p); v.orec = h->d_orec.p; v.obs_row0 = h->d_obs_row0.p;
  v.dtri_obs = h->d_dtri_obs.p; v.dmesh_obs = h->d_dmesh_obs.p; v.dtri_row0 = h->d_dtri_row0.p; v.dmesh_row0 = h->d_dmesh_row0.p;
  v.depth_xyz = h->d_depth.p; v.mesh_xyz = h->d_mesh.p;
  v.prior_pid = h->d_prior_pid.p; v.prior_ref = h->d_prior_ref.p; v.prior_w = h->d_prior_w.p; v.prior_th = h->d_prior_th.p; v.prior_row0 = h->d_prior_row0.p;
  v.trk_obs0 = h->d_trk_obs0.p; v.trk_dtri0 = h->d_trk_dtri0.p; v.trk_dmesh0 = h->d_trk_dmesh0.p; v.trk_pm0 = h->d_trk_pm0.p; v.trk_pt0 = h->d_trk_pt0.p;
  if (h->h_x_pinned) { cudaHostUnregister(h->h_x.data()); h->h_x_pinned = false; }
  h->h_x.assign(v.nx, 0.0);
  // page-lock the staging vector: the parameter upload / write-back of every mvba_solve then runs at full PCIe rate
  if (v.nx > 0 && cudaHostRegister(h->h_x.data(), sizeof(double) * v.nx, cudaHostRegisterDefault) == cudaSuccess) h->h_x_pinned = true;
  else cudaGetLastError();
  h->built = true;
  return agree_end.pass();
}

// Caller arrays -> staging -> device parameter vector.
// memcpy of a large array on a few host threads (the landmark block of the parameter vector is tens of MB per solve)
inline void big_copy(void* dst, const void* src, size_t bytes) {
  const size_t min_chunk = (size_t)2 << 20;
  int nt = (int)std::min<size_t>(8, bytes / min_chunk);
  if (nt <= 1) { std::memcpy(dst, src, bytes); return; }
  std::vector<std::thread> th;
  const size_t per = ((bytes / nt) + 63) & ~(size_t)63;
  for (int t = 0; t < nt; t++) {
    const size_t o = (size_t)t * per;
    if (o >= bytes) break;
    const size_t len = std::min(per, bytes - o);
    th.emplace_back([=]() { std::memcpy((char*)dst + o, (const char*)src + o, len); });
  }
  for (auto& t : th) t.join();
}

inline int upload_params(mvba_problem* h) {
  const View& v = h->view;
  const mvba_params& p = h->prm;
  const int T = v.T, nd2i = v.nd2i;
  double* x = h->h_x.data();
  const double* poses = h->cfg.no_rig ? p.world_to_cam : p.world_to_ref;
  if (!poses || !p.xyz || !p.ref_to_cam || !p.offsets || !p.focal || !p.center)
    return fail(MVBA_EINVAL, "mvba: a required parameter array is NULL");
  std::memcpy(x + v.x_pose, poses, sizeof(double) * 7 * v.npose);
  const double ident[7] = {0, 0, 0, 0, 0, 0, 1};  // identity_vec, rig_calibrator.cc:1419-1421
  std::memcpy(x + v.x_ident, ident, sizeof(ident));
  for (int t = 0; t < T; t++) {
    const SensorDev& s = v.sens[t];
    std::memcpy(x + s.x_rig, p.ref_to_cam + 7 * t, 7 * sizeof(double));
    x[s.x_off] = p.offsets[t]; x[s.x_focal] = p.focal[t];
    x[s.x_center] = p.center[2 * t]; x[s.x_center + 1] = p.center[2 * t + 1];
    if (s.nd > 0) {
      if (!p.distortion) return fail(MVBA_EINVAL, "mvba: distortion array is NULL");
      std::memcpy(x + s.x_dist, p.distortion + s.dist_off, s.nd * sizeof(double));
    }
    if (p.depth_to_image) std::memcpy(x + s.x_d2i, p.depth_to_image + nd2i * t, nd2i * sizeof(double));
    else { for (int k = 0; k < nd2i; k++) x[s.x_d2i + k] = 0; if (nd2i == 7) x[s.x_d2i + 6] = 1; }
    x[s.x_scale] = p.depth_scale ? p.depth_scale[t] : 1.0;
  }
  big_copy(x + v.x_xyz, p.xyz, sizeof(double) * 3 * v.P);
  MVBA_CUDA(cudaMemcpyAsync(h->d_x.p, x, sizeof(double) * v.nx, cudaMemcpyHostToDevice, h->stream));
  return MVBA_OK;
}

// Device parameter vector -> caller arrays (ceres::Solve updates parameters in place).
inline int download_params(mvba_problem* h) {
  const View& v = h->view;
  const mvba_params& p = h->prm;
  double* x = h->h_x.data();
  MVBA_CUDA(cudaMemcpyAsync(x, h->d_x.p, sizeof(double) * v.nx, cudaMemcpyDeviceToHost, h->stream));
  MVBA_CUDA(cudaStreamSynchronize(h->stream));
  double* poses = h->cfg.no_rig ? p.world_to_cam : p.world_to_ref;
  std::memcpy(poses, x + v.x_pose, sizeof(double) * 7 * v.npose);
  for (int t = 0; t < v.T; t++) {
    const SensorDev& s = v.sens[t];
    std::memcpy(p.ref_to_cam + 7 * t, x + s.x_rig, 7 * sizeof(double));
    p.offsets[t] = x[s.x_off]; p.focal[t] = x[s.x_focal];
    p.center[2 * t] = x[s.x_center]; p.center[2 * t + 1] = x[s.x_center + 1];
    if (s.nd > 0) std::memcpy(p.distortion + s.dist_off, x + s.x_dist, s.nd * sizeof(double));
    if (p.depth_to_image) std::memcpy(p.depth_to_image + v.nd2i * t, x + s.x_d2i, v.nd2i * sizeof(double));
    if (p.depth_scale) p.depth_scale[t] = x[s.x_scale];
  }
  big_copy(p.xyz, x + v.x_xyz, sizeof(double) * 3 * v.P);
  return MVBA_OK;
}

inline int check_device_error(mvba_problem* h) {
  int err = 0;
  MVBA_CUDA(cudaMemcpyAsync(&err, h->d_err.p, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
  MVBA_CUDA(cudaStreamSynchronize(h->stream));
  if (err == MVBA_EBRACKET) return fail(MVBA_EBRACKET, "Out of bounds in interpolation. (rig_calibrator.cc:346)");
  if (err) return fail(err, "mvba: device-side failure");
  return MVBA_OK;
}

inline int run_prep(mvba_problem* h, const double* x) {
  const View& v = h->view;
  int n = v.C + v.T;
  prep_cameras_kernel<<<(n + 127) / 128, 128, 0, h->stream>>>(v, x, h->d_tab.p, h->d_dtab.p, h->d_err.p, h->d_coef.p, h->d_pack.p);
  h->launches++;
  return MVBA_OK;
}

}  // namespace mvba
