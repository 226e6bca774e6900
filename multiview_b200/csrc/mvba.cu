// mvba.cu -- the C ABI of libmvba.so (include/mvba.h).  Single translation unit; all device code
// is hand-written CUDA for sm_100a.  No CPU fallback and nothing from oracle/ is reachable here.
#include <algorithm>
#include <array>
#include <chrono>
#include <cstdlib>
#include <mutex>

#include "handle.cuh"
#include "setup.cuh"
#include "solver.cuh"
#include "kernels_tracks.cuh"
#include "kernels_mesh.cuh"
#include "rpc_fit.cuh"

namespace mvba {
thread_local std::string g_last_error;

static int ensure_device(int device) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0) { cudaGetLastError(); return fail(MVBA_ENODEVICE, "mvba: no CUDA device is visible (there is no CPU fallback)"); }
  if (device < 0 || device >= n) return fail(MVBA_ENODEVICE, "mvba: CUDA device ordinal out of range");
  cudaDeviceProp p;
  if (cudaGetDeviceProperties(&p, device) != cudaSuccess) return fail(MVBA_ENODEVICE, "mvba: cannot query the CUDA device");
  if (p.major < 10) return fail(MVBA_ENODEVICE, "mvba: libmvba.so carries sm_100a code only; device is sm_" + std::to_string(p.major) + std::to_string(p.minor));
  return MVBA_OK;
}

static int ready(mvba_problem* h) {
  if (!h) return fail(MVBA_EINVAL, "mvba: null handle");
  if (!h->stream) {  // the device is looked at once per handle: cudaGetDeviceProperties costs milliseconds, every call of the ABI comes through here
    int rc = ensure_device(h->device);
    if (rc) return rc;
  }
  MVBA_CUDA(cudaSetDevice(h->device));
  if (!h->stream) {
    MVBA_CUDA(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    cudaDeviceProp p; MVBA_CUDA(cudaGetDeviceProperties(&p, h->device));
    h->sm_count = p.multiProcessorCount;
  }
  if (h->built) return MVBA_OK;
  auto t0 = std::chrono::steady_clock::now();
  int rc2 = setup(h);
  h->seconds_setup = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  return rc2;
}

}  // namespace mvba

using namespace mvba;

extern "C" {

const char* mvba_last_error(void) { return g_last_error.c_str(); }
int mvba_abi_version(void) { return MVBA_ABI_VERSION; }
int mvba_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  int ok = 0;
  for (int i = 0; i < n; i++) { cudaDeviceProp p; if (cudaGetDeviceProperties(&p, i) == cudaSuccess && p.major >= 10) ok++; }
  return ok;
}

int mvba_create(const mvba_config* cfg, int device, mvba_problem** out) {
  if (!cfg || !out) return fail(MVBA_EINVAL, "mvba_create: null argument");
  if (cfg->num_sensors <= 0 || cfg->num_sensors > MVBA_MAX_SENSORS || cfg->num_cameras < 0 || cfg->num_landmarks < 0 || cfg->num_ref_frames < 0)
    return fail(MVBA_EINVAL, "mvba_create: bad sizes");
  if (!(cfg->robust_threshold > 0)) return fail(MVBA_EINVAL, "mvba_create: robust_threshold must be positive");
  mvba_problem* h = new mvba_problem();
  h->cfg = *cfg; h->device = device;
  h->sensors.assign(cfg->num_sensors, SensorHost());
  *out = h;
  return MVBA_OK;
}

int mvba_destroy(mvba_problem* h) {
  if (!h) return MVBA_OK;
  if (h->solver) { cudaSetDevice(h->device); solver_destroy(h); }
  if (h->h_x_pinned) { cudaSetDevice(h->device); cudaHostUnregister(h->h_x.data()); h->h_x_pinned = false; }
  if (h->stream) { cudaSetDevice(h->device); cudaStreamDestroy(h->stream); }
  delete h;
  return MVBA_OK;
}

int mvba_set_sensor(mvba_problem* h, int32_t t, double w, double hgt, int32_t nd) {
  if (!h || t < 0 || t >= h->cfg.num_sensors) return fail(MVBA_EINVAL, "mvba_set_sensor: bad sensor index");
  if (nd < 0 || nd > MVBA_MAX_DISTORTION) return fail(MVBA_EINVAL, "mvba_set_sensor: bad distortion length");
  h->sensors[t].w = w; h->sensors[t].h = hgt; h->sensors[t].nd = nd; h->built = false;
  return MVBA_OK;
}

int mvba_set_params(mvba_problem* h, const mvba_params* p) {
  if (!h || !p) return fail(MVBA_EINVAL, "mvba_set_params: null argument");
  h->prm = *p; h->have_params = true; h->built = false;
  return MVBA_OK;
}

int mvba_set_cameras(mvba_problem* h, int32_t n, const int32_t* sensor, const int32_t* beg, const int32_t* end, const double* stamp) {
  if (!h || n != h->cfg.num_cameras || (n > 0 && (!sensor || !stamp))) return fail(MVBA_EINVAL, "mvba_set_cameras: bad argument");
  if (!h->cfg.no_rig && n > 0 && (!beg || !end)) return fail(MVBA_EINVAL, "mvba_set_cameras: bracketing indices required with a rig");
  h->cam_sensor.assign(sensor, sensor + n); h->cam_stamp.assign(stamp, stamp + n);
  if (beg && end) { h->cam_beg.assign(beg, beg + n); h->cam_end.assign(end, end + n); } else { h->cam_beg.assign(n, 0); h->cam_end.assign(n, 0); }
  h->have_cams = true; h->built = false;
  return MVBA_OK;
}

int mvba_set_observations(mvba_problem* h, int64_t n, const int32_t* pid, const int32_t* cid, const float* pix,
                          const uint8_t* has_depth, const float* depth_xyz, const uint8_t* has_mesh, const double* mesh_xyz) {
  if (!h || n < 0 || (n > 0 && (!pid || !cid || !pix))) return fail(MVBA_EINVAL, "mvba_set_observations: bad argument");
  h->nobs = n; h->opid.assign(pid, pid + n); h->ocid.assign(cid, cid + n); h->opix.assign(pix, pix + 2 * n);
  h->ohasd.clear(); h->odepth.clear(); h->ohasm.clear(); h->omesh.clear();
  if (has_depth && depth_xyz) { h->ohasd.assign(has_depth, has_depth + n); h->odepth.assign(depth_xyz, depth_xyz + 3 * n); }
  if (has_mesh && mesh_xyz) { h->ohasm.assign(has_mesh, has_mesh + n); h->omesh.assign(mesh_xyz, mesh_xyz + 3 * n); }
  h->have_obs = true; h->built = false;
  return MVBA_OK;
}

int mvba_set_point_priors(mvba_problem* h, int32_t kind, int64_t n, const int32_t* pid, const double* ref, double w, double th) {
  if (!h || (kind != MVBA_BLOCK_MESH_TRI && kind != MVBA_BLOCK_TRI) || n < 0 || (n > 0 && (!pid || !ref)))
    return fail(MVBA_EINVAL, "mvba_set_point_priors: bad argument");
  if (n > 0 && !(th > 0)) return fail(MVBA_EINVAL, "mvba_set_point_priors: robust threshold must be positive");
  auto& pr = h->prior[kind == MVBA_BLOCK_MESH_TRI ? 0 : 1];
  pr.pid.assign(pid, pid + n); pr.ref.assign(ref, ref + 3 * n); pr.w = w; pr.th = th; h->built = false;
  return MVBA_OK;
}

int mvba_set_float_mask(mvba_problem* h, const mvba_float_mask* m) {
  if (!h || !m) return fail(MVBA_EINVAL, "mvba_set_float_mask: null argument");
  int T = h->cfg.num_sensors;
  auto cp = [&](std::vector<uint8_t>& d, const uint8_t* s) { if (s) d.assign(s, s + T); else d.clear(); };
  cp(h->m_focal, m->focal); cp(h->m_center, m->center); cp(h->m_dist, m->distortion); cp(h->m_rig, m->rig);
  cp(h->m_d2i, m->depth_to_image); cp(h->m_poses, m->poses);
  h->float_offsets = m->float_timestamp_offsets; h->float_scale = m->float_scale; h->have_mask = true; h->built = false;
  return MVBA_OK;
}

int mvba_num_residuals(mvba_problem* h, int64_t* nres, int64_t* nblocks) {
  int rc = ready(h); if (rc) return rc;
  if (nres) *nres = h->nres;
  if (nblocks) *nblocks = h->nblocks;
  return MVBA_OK;
}

int mvba_residual_layout(mvba_problem* h, int32_t* kind, int64_t* row0, int64_t* src) {
  int rc = ready(h); if (rc) return rc;
  // merge the per-kind lists back into creation order by row0
  struct Ent { int64_t row0; int32_t kind; int64_t src; };
  std::vector<Ent> e; e.reserve(h->nblocks);
  for (int64_t i = 0; i < h->nobs; i++) e.push_back({h->h_obs_row0[i], MVBA_BLOCK_CAM, i});
  for (size_t j = 0; j < h->h_dtri_obs.size(); j++) e.push_back({h->h_dtri_row0[j], MVBA_BLOCK_DEPTH_TRI, h->h_dtri_obs[j]});
  for (size_t j = 0; j < h->h_dmesh_obs.size(); j++) e.push_back({h->h_dmesh_row0[j], MVBA_BLOCK_DEPTH_MESH, h->h_dmesh_obs[j]});
  int64_t n_pm = h->view.n_pm;
  for (int64_t j = 0; j < h->view.n_prior; j++) e.push_back({h->h_prior_row0[j], j < n_pm ? MVBA_BLOCK_MESH_TRI : MVBA_BLOCK_TRI, j < n_pm ? j : j - n_pm});
  std::sort(e.begin(), e.end(), [](const Ent& a, const Ent& b) { return a.row0 < b.row0; });
  for (size_t i = 0; i < e.size(); i++) { if (kind) kind[i] = e[i].kind; if (row0) row0[i] = e[i].row0; if (src) src[i] = e[i].src; }
  return MVBA_OK;
}

int mvba_upload(mvba_problem* h) {
  int rc = ready(h); if (rc) return rc;
  rc = upload_params(h); if (rc) return rc;
  MVBA_CUDA(cudaStreamSynchronize(h->stream));
  return MVBA_OK;
}

int mvba_evaluate(mvba_problem* h, int32_t apply_loss, double* cost, double* residuals) {
  int rc = ready(h); if (rc) return rc;
  rc = upload_params(h); if (rc) return rc;
  const View& v = h->view;
  MVBA_CUDA(cudaMemsetAsync(h->d_err.p, 0, sizeof(int), h->stream));
  MVBA_CUDA(cudaMemsetAsync(h->d_scalars.p, 0, sizeof(double), h->stream));
  run_prep(h, h->d_x.p);
  double* dres = nullptr;
  if (residuals) { MVBA_CUDA(h->d_resid.alloc((size_t)h->nres)); dres = h->d_resid.p; }
  int64_t total = v.N + v.n_dtri + v.n_dmesh + v.n_prior;
  int grid = (int)std::min<int64_t>((total + 255) / 256, (int64_t)h->sm_count * 8);
  if (grid < 1) grid = 1;
  residual_kernel<<<grid, 256, 0, h->stream>>>(v, h->d_x.p, h->d_tab.p, h->d_dtab.p, apply_loss, h->d_scalars.p, dres);
  h->launches++;
  MVBA_CUDA(cudaGetLastError());
  double c = 0;
  MVBA_CUDA(cudaMemcpyAsync(&c, h->d_scalars.p, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  if (residuals && h->nres > 0) MVBA_CUDA(cudaMemcpyAsync(residuals, dres, sizeof(double) * h->nres, cudaMemcpyDeviceToHost, h->stream));
  rc = check_device_error(h); if (rc) return rc;
  if (cost) *cost = c;
  return MVBA_OK;
}

static int kind_width(mvba_problem* h, int kind) {
  int ndmax = 0;
  for (auto& s : h->sensors) ndmax = std::max(ndmax, s.nd);
  switch (kind) {
    case MVBA_BLOCK_CAM: return 28 + ndmax;
    case MVBA_BLOCK_DEPTH_TRI: return 21 + h->nd2i() + 1 + 3 + 1;
    case MVBA_BLOCK_DEPTH_MESH: return 21 + h->nd2i() + 1 + 1;
    default: return 3;
  }
}

int mvba_jacobian_width(mvba_problem* h, int32_t kind, int32_t* w) {
  if (!h || !w || kind < 0 || kind > MVBA_BLOCK_TRI) return fail(MVBA_EINVAL, "mvba_jacobian_width: bad argument");
  *w = kind_width(h, kind);
  return MVBA_OK;
}

int mvba_evaluate_jacobian(mvba_problem* h, int32_t kind, int64_t first, int64_t count, double* residuals, double* jac) {
  int rc = ready(h); if (rc) return rc;
  if (kind < 0 || kind > MVBA_BLOCK_TRI || !jac) return fail(MVBA_EINVAL, "mvba_evaluate_jacobian: bad argument");
  const View& v = h->view;
  int64_t avail = kind == MVBA_BLOCK_CAM ? v.N : kind == MVBA_BLOCK_DEPTH_TRI ? v.n_dtri : kind == MVBA_BLOCK_DEPTH_MESH ? v.n_dmesh
                  : kind == MVBA_BLOCK_MESH_TRI ? v.n_pm : v.n_prior - v.n_pm;
  if (first < 0 || count < 0 || first + count > avail) return fail(MVBA_EINVAL, "mvba_evaluate_jacobian: range out of bounds");
  if (count == 0) return MVBA_OK;
  rc = upload_params(h); if (rc) return rc;
  MVBA_CUDA(cudaMemsetAsync(h->d_err.p, 0, sizeof(int), h->stream));
  run_prep(h, h->d_x.p);
  const int W = kind_width(h, kind), rows = kind == MVBA_BLOCK_CAM ? 2 : 3;
  DBuf<double> dJ, dr;
  MVBA_CUDA(dJ.alloc((size_t)count * rows * W)); MVBA_CUDA(dr.alloc((size_t)count * rows));
  MVBA_CUDA(dJ.zero(h->stream));
  jacobian_store_kernel<<<(unsigned)((count + 127) / 128), 128, 0, h->stream>>>(v, h->d_x.p, h->d_tab.p, h->d_dtab.p, kind, first, count, W, dr.p, dJ.p);
  h->launches++;
  MVBA_CUDA(cudaGetLastError());
  MVBA_CUDA(cudaMemcpyAsync(jac, dJ.p, sizeof(double) * count * rows * W, cudaMemcpyDeviceToHost, h->stream));
  if (residuals) MVBA_CUDA(cudaMemcpyAsync(residuals, dr.p, sizeof(double) * count * rows, cudaMemcpyDeviceToHost, h->stream));
  return check_device_error(h);
}

int mvba_bench_jacobian(mvba_problem* h, int32_t repeats, float* ms_total) {
  int rc = ready(h); if (rc) return rc;
  const View& v = h->view;
  if (v.N == 0 || repeats <= 0) return fail(MVBA_EINVAL, "mvba_bench_jacobian: nothing to do");
  const int W = kind_width(h, MVBA_BLOCK_CAM);
  DBuf<double> dJ, dr;
  MVBA_CUDA(dJ.alloc((size_t)v.N * 2 * W)); MVBA_CUDA(dr.alloc((size_t)v.N * 2));
  run_prep(h, h->d_x.p);
  cudaEvent_t e0, e1;
  MVBA_CUDA(cudaEventCreate(&e0)); MVBA_CUDA(cudaEventCreate(&e1));
  MVBA_CUDA(cudaEventRecord(e0, h->stream));
  for (int i = 0; i < repeats; i++) {
    jacobian_store_kernel<<<(unsigned)((v.N + 127) / 128), 128, 0, h->stream>>>(v, h->d_x.p, h->d_tab.p, h->d_dtab.p, MVBA_BLOCK_CAM, 0, v.N, W, dr.p, dJ.p);
    h->launches++;
  }
  MVBA_CUDA(cudaEventRecord(e1, h->stream));
  MVBA_CUDA(cudaEventSynchronize(e1));
  MVBA_CUDA(cudaEventElapsedTime(ms_total, e0, e1));
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  MVBA_CUDA(cudaGetLastError());
  return MVBA_OK;
}

int mvba_solve_options_init(mvba_solve_options* o) {
  if (!o) return fail(MVBA_EINVAL, "mvba_solve_options_init: null argument");
  // rig_calibrator.cc:1910-1918 + Ceres 1.14 Solver::Options defaults
  o->max_num_iterations = 20; o->parameter_tolerance = 1e-12; o->function_tolerance = 1e-16; o->gradient_tolerance = 1e-16;
  o->initial_trust_region_radius = 1e4; o->max_trust_region_radius = 1e16; o->min_trust_region_radius = 1e-32;
  o->min_relative_decrease = 1e-3; o->min_lm_diagonal = 1e-6; o->max_lm_diagonal = 1e32; o->eta = 0.1;
  o->max_linear_solver_iterations = 500; o->min_linear_solver_iterations = 0; o->jacobi_scaling = 1;
  o->linear_solver = 0; o->preconditioner = 0; o->verbose = 0;
  return MVBA_OK;
}

int mvba_solve(mvba_problem* h, const mvba_solve_options* o, mvba_summary* s, mvba_iteration* its, int32_t max_its) {
  auto t_in = std::chrono::steady_clock::now();
  int rc = ready(h); if (rc) return rc;
  if (std::getenv("MVBA_TIMING")) std::fprintf(stderr, "[mvba timing] ready %.3f ms\n", std::chrono::duration<double>(std::chrono::steady_clock::now() - t_in).count() * 1e3);
  if (!o || !s) return fail(MVBA_EINVAL, "mvba_solve: null argument");
  if (o->linear_solver != 0 && o->linear_solver != 1) return fail(MVBA_EINVAL, "mvba_solve: linear_solver must be 0 (PCG on the reduced camera system) or 1 (dense Cholesky)");
  if (o->preconditioner < 0 || o->preconditioner > 2) return fail(MVBA_EINVAL, "mvba_solve: preconditioner must be 0, 1 or 2");
  return solver_solve(h, *o, s, its, max_its);
}

int mvba_host_seconds(mvba_problem* h, double* seconds_setup, double* seconds_prepare) {
  if (!h) return fail(MVBA_EINVAL, "mvba_host_seconds: null handle");
  if (seconds_setup) *seconds_setup = h->seconds_setup;
  if (seconds_prepare) *seconds_prepare = h->seconds_prepare;
  return MVBA_OK;
}

int mvba_build_plan(mvba_problem* h, int32_t plan[8]) {
  if (!h || !plan) return fail(MVBA_EINVAL, "mvba_build_plan: null argument");
  if (!h->solver) return fail(MVBA_ESTATE, "mvba_build_plan: no solve has been planned on this handle");
  const Solver* s = h->solver;
  plan[0] = s->cam_fast ? 1 : 0;
  plan[1] = s->ntiles == 0 ? 0 : s->tile2 ? 2 : 1;
  plan[2] = s->tile2 ? s->t2_maxs : s->tile_pr;
  plan[3] = s->tile_K; plan[4] = s->ntiles; plan[5] = s->nrounds; plan[6] = s->nfallback;
  plan[7] = s->tile2 ? s->t2_L8max : s->tile_L8max;
  return MVBA_OK;
}

int mvba_bench_build(mvba_problem* h, int32_t repeats, float* ms_total) {
  int rc = ready(h); if (rc) return rc;
  return solver_bench_build(h, repeats, ms_total);
}

// ---- per-track steps around a pass ----
static int tracks_begin(mvba_problem* h, const uint8_t* inlier, mvba::DBuf<uint8_t>& d_in) {
  int rc = ready(h); if (rc) return rc;
  if (!inlier) return fail(MVBA_EINVAL, "mvba: null inlier array");
  rc = upload_params(h); if (rc) return rc;
  MVBA_CUDA(cudaMemsetAsync(h->d_err.p, 0, sizeof(int), h->stream));
  run_prep(h, h->d_x.p);
  MVBA_CUDA(d_in.alloc((size_t)std::max<int64_t>(h->view.N, 1)));
  if (h->view.N > 0) MVBA_CUDA(cudaMemcpyAsync(d_in.p, inlier, (size_t)h->view.N, cudaMemcpyHostToDevice, h->stream));
  return MVBA_OK;
}

int mvba_triangulate(mvba_problem* h, uint8_t* inlier, double* xyz) {
  mvba::DBuf<uint8_t> d_in; mvba::DBuf<double> d_xyz;
  int rc = tracks_begin(h, inlier, d_in); if (rc) return rc;
  if (!xyz) return fail(MVBA_EINVAL, "mvba_triangulate: null xyz array");
  const View& v = h->view;
  if (v.P == 0) return MVBA_OK;
  MVBA_CUDA(d_xyz.alloc((size_t)3 * v.P));
  MVBA_CUDA(cudaMemcpyAsync(d_xyz.p, xyz, sizeof(double) * 3 * v.P, cudaMemcpyHostToDevice, h->stream));
  const int grid = std::max(1, std::min((v.P + 127) / 128, h->sm_count * 16));
  triangulate_kernel<<<grid, 128, 0, h->stream>>>(v, h->d_x.p, h->d_tab.p, d_in.p, d_xyz.p);
  h->launches++;
  MVBA_CUDA(cudaGetLastError());
  MVBA_CUDA(cudaMemcpyAsync(xyz, d_xyz.p, sizeof(double) * 3 * v.P, cudaMemcpyDeviceToHost, h->stream));
  if (v.N > 0) MVBA_CUDA(cudaMemcpyAsync(inlier, d_in.p, (size_t)v.N, cudaMemcpyDeviceToHost, h->stream));
  return check_device_error(h);
}

int mvba_flag_outliers(mvba_problem* h, double min_triangulation_angle, double max_reprojection_error, uint8_t* inlier, int64_t counts[4]) {
  mvba::DBuf<uint8_t> d_in; mvba::DBuf<unsigned long long> d_cnt;
  int rc = tracks_begin(h, inlier, d_in); if (rc) return rc;
  const View& v = h->view;
  MVBA_CUDA(d_cnt.alloc(4)); MVBA_CUDA(d_cnt.zero(h->stream));
  if (v.P > 0) {
    const int grid = std::max(1, std::min((v.P + 127) / 128, h->sm_count * 16));
    flag_angle_reproj_kernel<<<grid, 128, 0, h->stream>>>(v, h->d_x.p, h->d_tab.p, min_triangulation_angle, max_reprojection_error, d_in.p, d_cnt.p);
    h->launches++;
    MVBA_CUDA(cudaGetLastError());
  }
  unsigned long long c[4] = {0, 0, 0, 0};
  MVBA_CUDA(cudaMemcpyAsync(c, d_cnt.p, sizeof(c), cudaMemcpyDeviceToHost, h->stream));
  if (v.N > 0) MVBA_CUDA(cudaMemcpyAsync(inlier, d_in.p, (size_t)v.N, cudaMemcpyDeviceToHost, h->stream));
  rc = check_device_error(h); if (rc) return rc;
  if (counts) for (int i = 0; i < 4; i++) counts[i] = (int64_t)c[i];
  return MVBA_OK;
}

int mvba_flag_exclusion(mvba_problem* h, const double* image_size, const double* crop_size, uint8_t* inlier) {
  mvba::DBuf<uint8_t> d_in; mvba::DBuf<double> d_sz;
  int rc = tracks_begin(h, inlier, d_in); if (rc) return rc;
  if (!image_size || !crop_size) return fail(MVBA_EINVAL, "mvba_flag_exclusion: null size array");
  const View& v = h->view;
  std::vector<double> sz((size_t)4 * v.T);
  for (int i = 0; i < 2 * v.T; i++) { sz[i] = crop_size[i]; sz[2 * v.T + i] = image_size[i]; }
  MVBA_CUDA(d_sz.upload(sz));
  if (v.N > 0) {
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((v.N + 255) / 256, (int64_t)h->sm_count * 16));
    flag_exclusion_kernel<<<grid, 256, 0, h->stream>>>(v, d_sz.p, d_sz.p + 2 * v.T, d_in.p);
    h->launches++;
    MVBA_CUDA(cudaGetLastError());
    MVBA_CUDA(cudaMemcpyAsync(inlier, d_in.p, (size_t)v.N, cudaMemcpyDeviceToHost, h->stream));
  }
  return check_device_error(h);
}

// Median-split hierarchy over the triangles (host): nodes in depth-first order, triangles reordered leaf by leaf.
// Returns the depth of the tree (root = 1): the traversal kernel keeps one stack entry per level.
static int build_bvh(const std::vector<mvba::MeshTri>& in, std::vector<mvba::BvhNode>& nodes, std::vector<mvba::MeshTri>& out) {
  const int n = (int)in.size();
  std::vector<int> idx(n);
  for (int i = 0; i < n; i++) idx[i] = i;
  std::vector<std::array<float, 3>> cen(n);
  for (int i = 0; i < n; i++) for (int a = 0; a < 3; a++) cen[i][a] = (in[i].a[a] + in[i].b[a] + in[i].c[a]) / 3.0f;
  struct Job { int node, first, last, depth; };
  std::vector<Job> todo;
  int max_depth = 0;
  nodes.clear(); out.clear();
  nodes.push_back(mvba::BvhNode{});
  todo.push_back({0, 0, n, 1});
  while (!todo.empty()) {
    const Job j = todo.back(); todo.pop_back();
    max_depth = std::max(max_depth, j.depth);
    mvba::BvhNode nd{};
    float clo[3], chi[3];
    for (int a = 0; a < 3; a++) { nd.lo[a] = clo[a] = 3.4e38f; nd.hi[a] = chi[a] = -3.4e38f; }
    for (int k = j.first; k < j.last; k++) {
      const mvba::MeshTri& t = in[idx[k]];
      for (int a = 0; a < 3; a++) {
        nd.lo[a] = std::min(nd.lo[a], std::min(t.a[a], std::min(t.b[a], t.c[a])));
        nd.hi[a] = std::max(nd.hi[a], std::max(t.a[a], std::max(t.b[a], t.c[a])));
        clo[a] = std::min(clo[a], cen[idx[k]][a]); chi[a] = std::max(chi[a], cen[idx[k]][a]);
      }
    }
    nd.left = nd.right = -1; nd.first = 0; nd.count = 0;
    int axis = 0;
    for (int a = 1; a < 3; a++) if (chi[a] - clo[a] > chi[axis] - clo[axis]) axis = a;
    if (j.last - j.first <= 4 || !(chi[axis] > clo[axis])) {
      nd.first = (int)out.size(); nd.count = j.last - j.first;
      for (int k = j.first; k < j.last; k++) out.push_back(in[idx[k]]);
    } else {
      const int mid = (j.first + j.last) / 2;
      std::nth_element(idx.begin() + j.first, idx.begin() + mid, idx.begin() + j.last, [&](int p, int q) { return cen[p][axis] < cen[q][axis]; });
      nd.left = (int)nodes.size(); nodes.push_back(mvba::BvhNode{});
      nd.right = (int)nodes.size(); nodes.push_back(mvba::BvhNode{});
      todo.push_back({nd.left, j.first, mid, j.depth + 1});
      todo.push_back({nd.right, mid, j.last, j.depth + 1});
    }
    nodes[j.node] = nd;
  }
  return max_depth;
}

int mvba_mesh_intersect(mvba_problem* h, int64_t num_vertices, const float* vertices, int64_t num_faces, const uint32_t* faces,
                        double min_ray_dist, double max_ray_dist, const uint8_t* inlier, uint8_t* has_mesh, double* mesh_xyz, double* pid_mesh_xyz) {
  mvba::DBuf<uint8_t> d_in, d_has; mvba::DBuf<double> d_xyz, d_mean; mvba::DBuf<mvba::BvhNode> d_nodes; mvba::DBuf<mvba::MeshTri> d_tris;
  int rc = tracks_begin(h, inlier, d_in); if (rc) return rc;
  if (!vertices || !faces || num_faces < 1 || num_vertices < 3 || !has_mesh || !mesh_xyz) return fail(MVBA_EINVAL, "mvba_mesh_intersect: bad argument");
  std::vector<mvba::MeshTri> tris((size_t)num_faces), sorted;
  for (int64_t f = 0; f < num_faces; f++) {
    for (int k = 0; k < 3; k++) if (faces[3 * f + k] >= (uint64_t)num_vertices) return fail(MVBA_EINVAL, "mvba_mesh_intersect: face index out of range");
    for (int a = 0; a < 3; a++) {
      tris[f].a[a] = vertices[3 * (size_t)faces[3 * f] + a]; tris[f].b[a] = vertices[3 * (size_t)faces[3 * f + 1] + a];
      tris[f].c[a] = vertices[3 * (size_t)faces[3 * f + 2] + a];
    }
  }
  std::vector<mvba::BvhNode> nodes;
  const int depth = build_bvh(tris, nodes, sorted);
  if (depth + 2 > MVBA_BVH_STACK) return fail(MVBA_EINVAL, "mvba_mesh_intersect: BVH deeper than the traversal stack");
  MVBA_CUDA(d_nodes.upload(nodes)); MVBA_CUDA(d_tris.upload(sorted));
  const View& v = h->view;
  MVBA_CUDA(d_has.alloc((size_t)std::max<int64_t>(v.N, 1))); MVBA_CUDA(d_xyz.alloc((size_t)3 * std::max<int64_t>(v.N, 1)));
  MVBA_CUDA(d_mean.alloc((size_t)3 * std::max(v.P, 1)));
  if (v.N > 0) {
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((v.N + 127) / 128, (int64_t)h->sm_count * 16));
    mesh_intersect_kernel<<<grid, 128, 0, h->stream>>>(v, h->d_x.p, h->d_tab.p, d_nodes.p, d_tris.p, min_ray_dist, max_ray_dist, d_in.p, d_has.p, d_xyz.p, h->d_err.p);
    const int grid2 = std::max(1, std::min((v.P + 127) / 128, h->sm_count * 16));
    mesh_track_mean_kernel<<<grid2, 128, 0, h->stream>>>(v, d_has.p, d_xyz.p, d_mean.p);
    h->launches += 2;
    MVBA_CUDA(cudaGetLastError());
    MVBA_CUDA(cudaMemcpyAsync(has_mesh, d_has.p, (size_t)v.N, cudaMemcpyDeviceToHost, h->stream));
    MVBA_CUDA(cudaMemcpyAsync(mesh_xyz, d_xyz.p, sizeof(double) * 3 * v.N, cudaMemcpyDeviceToHost, h->stream));
    if (pid_mesh_xyz) MVBA_CUDA(cudaMemcpyAsync(pid_mesh_xyz, d_mean.p, sizeof(double) * 3 * v.P, cudaMemcpyDeviceToHost, h->stream));
  } else if (pid_mesh_xyz) {
    for (size_t i = 0; i < (size_t)3 * v.P; i++) pid_mesh_xyz[i] = MVBA_BAD_XYZ;  // no feature, no hit: every track gets bad_xyz
  }
  return check_device_error(h);
}

int mvba_nccl_unique_id(uint8_t id128[128]) { return comm_unique_id(id128); }
int mvba_comm_init(mvba_problem* h, int32_t rank, int32_t nranks, const uint8_t id128[128]) {
  if (!h) return fail(MVBA_EINVAL, "mvba_comm_init: null handle");
  int rc = ready(h); if (rc) return rc;
  return comm_init(h, rank, nranks, id128);
}

int mvba_rpc_fit(int32_t device, int64_t n, const double* in_xy, const double* out_xy, int32_t rpc_degree, int32_t num_iterations,
                 double parameter_tolerance, int32_t jacobian_mode, double* coeffs, double stats[4]) {
  if (n < 1 || !in_xy || !out_xy || !coeffs || rpc_degree < 1 || rpc_degree > mvba::RPC_MAX_DEG || num_iterations < 0)
    return fail(MVBA_EINVAL, "mvba_rpc_fit: bad argument");
  int rc = ensure_device(device); if (rc) return rc;
  MVBA_CUDA(cudaSetDevice(device));
  cudaDeviceProp p; MVBA_CUDA(cudaGetDeviceProperties(&p, device));
  mvba::RpcFitter F;
  rc = F.init(n, in_xy, out_xy, p.multiProcessorCount); if (rc) return rc;
  std::vector<double> c; mvba::RpcFitStats st;
  rc = mvba::rpc_fit_all(F, rpc_degree, num_iterations, parameter_tolerance, jacobian_mode, c, &st); if (rc) return rc;
  for (size_t i = 0; i < c.size(); i++) coeffs[i] = c[i];
  if (stats) { stats[0] = st.initial_cost; stats[1] = st.final_cost; stats[2] = st.iterations; stats[3] = st.successful; }
  return MVBA_OK;
}

}  // extern "C"

// ---- mvba_rpc_fit_static: the same fit with every byte of device scratch in one file-scope array (no allocation) ----
// Layout (doubles): [in 2n | out 2n | per-CTA partial normal equations | folded sums], RpcFitter::init_static.
namespace {
constexpr size_t kRpcStaticLen = 4 * (size_t)MVBA_RPC_STATIC_MAX_N + ((size_t)MVBA_RPC_STATIC_MAX_PARTS + 1) * (2 * (size_t)mvba::RPC_LD * mvba::RPC_LD + 2 * mvba::RPC_LD + 8);
constexpr int kRpcStaticMaxDevices = 64;
}  // namespace
static __device__ __align__(256) double g_rpc_static_scratch[kRpcStaticLen];  // ~7.1 MB on every device that loads the library
// In host code the name above is a host-side stand-in, not a device address: the address a kernel may use comes from
// cudaGetSymbolAddress, asked once per device with that device current.
static double* g_rpc_static_ptr[kRpcStaticMaxDevices];
static int g_rpc_static_sms[kRpcStaticMaxDevices];
static std::mutex g_rpc_static_mutex;

extern "C" {

int mvba_rpc_fit_static(int32_t device, int64_t n, const double* in_xy, const double* out_xy, int32_t rpc_degree, int32_t num_iterations,
                        double parameter_tolerance, int32_t jacobian_mode, double* coeffs, double stats[4]) {
  if (n < 1 || !in_xy || !out_xy || !coeffs || rpc_degree < 1 || rpc_degree > mvba::RPC_MAX_DEG || num_iterations < 0)
    return fail(MVBA_EINVAL, "mvba_rpc_fit_static: bad argument");
  if (n > MVBA_RPC_STATIC_MAX_N)
    return fail(MVBA_EINVAL, "mvba_rpc_fit_static: " + std::to_string(n) + " samples exceed the static scratch area (MVBA_RPC_STATIC_MAX_N = " +
                                 std::to_string(MVBA_RPC_STATIC_MAX_N) + "); use mvba_rpc_fit");
  if (device < 0 || device >= kRpcStaticMaxDevices) return fail(MVBA_ENODEVICE, "mvba_rpc_fit_static: CUDA device ordinal out of range");
  std::lock_guard<std::mutex> lock(g_rpc_static_mutex);  // one user of the array at a time within this process
  if (!g_rpc_static_ptr[device]) {                        // set-up, once per device
    int rc = ensure_device(device); if (rc) return rc;
    MVBA_CUDA(cudaSetDevice(device));
    double* ptr = nullptr;
    MVBA_CUDA(cudaGetSymbolAddress((void**)&ptr, g_rpc_static_scratch));
    int sms = 0; MVBA_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    g_rpc_static_sms[device] = sms; g_rpc_static_ptr[device] = ptr;
  }
  MVBA_CUDA(cudaSetDevice(device));
  mvba::RpcFitter F;
  int rc = F.init_static(n, in_xy, out_xy, g_rpc_static_sms[device], g_rpc_static_ptr[device], kRpcStaticLen, MVBA_RPC_STATIC_MAX_PARTS); if (rc) return rc;
  std::vector<double> c; mvba::RpcFitStats st;
  rc = mvba::rpc_fit_all(F, rpc_degree, num_iterations, parameter_tolerance, jacobian_mode, c, &st); if (rc) return rc;
  MVBA_CUDA(cudaStreamSynchronize(F.stream));
  for (size_t i = 0; i < c.size(); i++) coeffs[i] = c[i];
  if (stats) { stats[0] = st.initial_cost; stats[1] = st.final_cost; stats[2] = st.iterations; stats[3] = st.successful; }
  return MVBA_OK;
}

}  // extern "C"
