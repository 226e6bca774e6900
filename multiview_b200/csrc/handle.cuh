// handle.cuh -- host-side state of one mvba_problem: caller pointers, the device copy of the
// problem (SoA), the reference-order residual layout and the reduced-system structure.
#pragma once
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "common.cuh"

namespace mvba {

extern thread_local std::string g_last_error;
inline int fail(int code, const std::string& msg) { g_last_error = msg; return code; }

#define MVBA_CUDA(call)                                                                            \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess)                                                                         \
      return mvba::fail(MVBA_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e_));          \
  } while (0)

// Owning device buffer.
template <class T> struct DBuf {
  T* p = nullptr;
  size_t n = 0;
  ~DBuf() { release(); }
  void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
  cudaError_t alloc(size_t count) {
    if (count == n && p) return cudaSuccess;
    release();
    n = count;
    if (count == 0) return cudaSuccess;
    return cudaMalloc((void**)&p, count * sizeof(T));
  }
  cudaError_t upload(const std::vector<T>& h) {
    cudaError_t e = alloc(h.size());
    if (e != cudaSuccess || h.empty()) return e;
    return cudaMemcpy(p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice);
  }
  cudaError_t zero(cudaStream_t s) { return n ? cudaMemsetAsync(p, 0, n * sizeof(T), s) : cudaSuccess; }
};

struct SensorHost { double w = 0, h = 0; int nd = 0; };

struct Solver;  // solver.cuh

}  // namespace mvba

struct mvba_problem {
  mvba_config cfg{};
  int device = 0;
  cudaStream_t stream = nullptr;
  std::vector<mvba::SensorHost> sensors;
  mvba_params prm{};
  bool have_params = false, have_cams = false, have_obs = false, have_mask = false;
  // host copies of the static problem
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

  // ---- built by setup() ----
  bool built = false;
  mvba::View view{};
  int64_t nres = 0, nblocks = 0;
  std::vector<int64_t> h_obs_row0, h_dtri_obs, h_dtri_row0, h_dmesh_obs, h_dmesh_row0, h_prior_row0;
  std::vector<int32_t> h_prior_pid; std::vector<double> h_prior_ref, h_prior_w, h_prior_th;
  std::vector<int64_t> h_trk_obs0, h_trk_dtri0, h_trk_dmesh0, h_trk_pm0, h_trk_pt0;
  std::vector<int32_t> h_cam_p0, h_cam_p1, h_pose_col;
  std::vector<double> h_cam_bs, h_cam_es;
  std::vector<double> h_x;             // staging for the parameter vector (page-locked once sized: setup.cuh)
  bool h_x_pinned = false;
  std::vector<double> h_lo, h_hi;      // per-scalar bounds (only offsets are ever bounded)
  std::vector<int32_t> h_xcol;         // parameter scalar -> free index (f cols, then nf + 3*pid) or -1
  bool bounded = false;
  // device
  mvba::DBuf<int32_t> d_cam_sensor, d_cam_p0, d_cam_p1, d_pose_col, d_pid, d_cid, d_prior_pid;
  mvba::DBuf<double> d_cam_bs, d_cam_es, d_cam_cs, d_mesh, d_prior_ref, d_prior_w, d_prior_th;
  mvba::DBuf<float> d_pix, d_depth;
  mvba::DBuf<int4> d_orec;
  mvba::DBuf<int64_t> d_obs_row0, d_dtri_obs, d_dtri_row0, d_dmesh_obs, d_dmesh_row0, d_prior_row0;
  mvba::DBuf<int64_t> d_trk_obs0, d_trk_dtri0, d_trk_dmesh0, d_trk_pm0, d_trk_pt0;
  mvba::DBuf<double> d_x, d_xnew;
  mvba::DBuf<mvba::CamTab> d_tab;
  mvba::DBuf<mvba::DepthTab> d_dtab;
  mvba::DBuf<double> d_fov;
  mvba::DBuf<mvba::CamCoef> d_coef;
  mvba::DBuf<mvba::CamPack> d_pack;   // allocated when the staged tile kernel is planned (solver.cuh), else null
  mvba::DBuf<double> d_scalars;   // small scratch of device scalars
  mvba::DBuf<int> d_err;
  mvba::DBuf<double> d_resid;     // residual rows scratch
  int sm_count = 148;
  mvba::Solver* solver = nullptr;
  int64_t launches = 0;
  double seconds_setup = 0, seconds_prepare = 0;  // host wall clock of setup() and solver_prepare() (one-time structure analysis)
  // multi-GPU
  void* nccl_comm = nullptr;
  int rank = 0, nranks = 1;

  int nd2i() const { return cfg.affine_depth_to_image ? 12 : 7; }
};

namespace mvba {
int setup(mvba_problem* h);            // builds layout + uploads the static problem
int upload_params(mvba_problem* h);    // caller arrays -> d_x
int download_params(mvba_problem* h);  // d_x -> caller arrays
int run_prep(mvba_problem* h, const double* x);
}  // namespace mvba
