// common.cuh -- device-side view of one rig_calibrator optimisation pass and small FP64 helpers.
// Part of libmvba.so (product).  Nothing here may include or link anything under oracle/.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/mvba.h"

#define MVBA_MAX_SENSORS 16

namespace mvba {

// Per-camera table written by prep_cameras_kernel (one thread per camera image) and read by every
// per-observation kernel.  All transcendental work of the pose interpolation
// (reference rig_calibrator.cc:322-390, dense_map_utils.cc:315-329) happens once per camera here,
// so the per-observation kernels are pure FMA + distortion.
//   Xc = R X + t,   Z = R X
//   dXc/dt0 = (1-alpha) Rrc      dXc/dq0_j = W0[:,j] x Z
//   dXc/dt1 = alpha Rrc          dXc/dq1_j = W1[:,j] x Z
//   dXc/dtrc = rig_on I          dXc/dqrc_j = Wrc[:,j] x (Xc - trc)
//   dXc/doffset = dalpha (wa x Z + dt)
struct CamTab {
  double R[9], t[3];
  double Rrc[9], trc[3];
  double W0[12], W1[12], Wrc[12];  // 3x4 row-major, quaternion columns in storage order x y z w
  double wa[3], dt[3];
  double alpha, dalpha, rig_on, pad;
};
static_assert(sizeof(CamTab) == 70 * 8, "CamTab layout");

// Pose-part Jacobian coefficients of one camera image, one 3-vector per row (prep_cameras_kernel):
//   rows 0-2 (1-alpha) Rrc[:,k] | 3-6 W0[:,j] | 7-9 alpha Rrc[:,k] | 10-13 W1[:,j] | 14-16 rig_on e_k | 17-20 Wrc[:,j]
//   | 21 dalpha wa | 22 dalpha dt.   Virtual pose column vc = row (rows 21 and 22 both belong to the offset column).
// With M = (sr A)' E and N = [Z]x M, Nrc = [Xc - trc]x M the block's F'E rows are coef' M (translations, row 22),
// coef' N (rotations of the bracketing poses, row 21) and coef' Nrc (rig rotation).
struct CamCoef { double c[23][3]; };

// Per-image data the tile kernel stages in shared memory (kernels_tile2.cuh; written by prep_cameras_kernel next to CamTab / CamCoef).
struct alignas(16) CamPack {
  double R[9], t[3];
  double dl[3];      // t - t_rc: the rig rotation acts on Xc - t_rc = Z + dl
  double c[23][3];   // CamCoef rows
  int sensor, pad0;
  double pad1;
};
static_assert(sizeof(CamPack) == 688 && sizeof(CamPack) % 16 == 0, "CamPack layout");

// One residual block in a host-packed list (the rounds of kernels_tile2.cuh, the camera-major list of kernels_camq.cuh): CAM blocks carry everything the evaluation needs (a = landmark, unused here: the
// landmark comes from the track slot), other kinds their index inside the kind's array.
//   b = image | kind << 24 | track slot << 27    (kind 0 CAM, 1 DEPTH_TRI, 2 DEPTH_MESH, 3 MESH_TRI prior, 4 TRI prior)
struct alignas(16) TSlot { int32_t a; uint32_t b; float px, py; };
static_assert(sizeof(TSlot) == 16, "TSlot layout");

// Per-sensor depth-to-image table (prep_sensors_kernel): Mc = s L m + td
struct DepthTab {
  double L[9], td[3], s;
  double Wd[12];  // rigid: dMc/dq_j = s * Wd[:,j] x (L m)
};

struct SensorDev {
  double half[2];
  int nd;        // distortion block length (0 | 1 | 4 | 5 | 2n RPC)
  int dist_off;  // offset of the block inside the distortion array
  int rpc_deg;   // >0 for RPC
  int nd_nz;     // number of distortion columns with a non-zero Jacobian (nd, or nd/2 for RPC)
  // offsets of this sensor's parameter blocks inside the device parameter vector x
  int x_rig, x_off, x_focal, x_center, x_dist, x_d2i, x_scale;
  // first reduced-system column of each block, -1 when the block is constant
  int c_rig, c_off, c_focal, c_center, c_dist, c_d2i, c_scale;
};

// Everything a kernel needs, passed by value (fits the 4 KB kernel parameter space).
struct View {
  int T, R, C, P, npose, no_rig, nd2i;
  int64_t N;        // CAM blocks (observations)
  int64_t n_dtri, n_dmesh, n_prior, n_pm;  // n_pm: MESH_TRI priors (stored first), rest are TRI
  int nf, npose_cols, B;  // reduced system: nf = npose_cols + B
  double robust_th, w_dtri, w_dmesh;
  // parameter vector layout
  int x_pose, x_ident, x_xyz, nx;
  SensorDev sens[MVBA_MAX_SENSORS];
  // cameras
  const int32_t *cam_sensor, *cam_p0, *cam_p1;  // p0/p1: pose block index (p1 = -1: identity_vec)
  const double *cam_bs, *cam_es, *cam_cs;
  const int32_t* pose_col;  // npose: first reduced column or -1
  double* fov_tab;          // 2 per sensor, written by prep_cameras_kernel: 2 tan(w/2) and 1/w of a fisheye (FOV) sensor
  // observations (track-major, reference order)
  const int32_t *pid, *cid;
  const float2* pix;
  const int4* orec;  // per observation {image, sensor, pixel x, pixel y (float bits)}: what cid / cam_sensor / pix hold, in one 128-bit load
  const int64_t* obs_row0;  // first residual row of the CAM block
  // depth-tri / depth-mesh blocks: index of the originating observation
  const int64_t *dtri_obs, *dmesh_obs, *dtri_row0, *dmesh_row0;
  const float* depth_xyz;   // 3 per observation (only where used)
  const double* mesh_xyz;   // 3 per observation
  // priors
  const int32_t* prior_pid; const double* prior_ref; const double *prior_w, *prior_th; const int64_t* prior_row0;
  // per-track CSR over items
  const int64_t* trk_obs0;      // P+1: first observation of each track
  const int64_t* trk_dtri0;     // P+1
  const int64_t* trk_dmesh0;    // P+1
  const int64_t* trk_pm0;       // P+1: MESH_TRI priors of each track
  const int64_t* trk_pt0;       // P+1: TRI priors of each track (indices into the merged list)
};

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ void cross3(const double* a, const double* b, double* o) {
  o[0] = a[1] * b[2] - a[2] * b[1];
  o[1] = a[2] * b[0] - a[0] * b[2];
  o[2] = a[0] * b[1] - a[1] * b[0];
}

// D(8x8) += A(8x4) B(4x8) on the FP64 tensor cores: a = A[lane/4][lane%4], b = B[lane%4][lane/4],
// c0, c1 = C[lane/4][2 (lane%4) + {0, 1}].
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// 3x3 symmetric inverse from the 6 unique entries (xx xy xz yy yz zz); returns false when singular.
__device__ __forceinline__ bool inv_sym3(const double* h, double* o) {
  double c00 = h[3] * h[5] - h[4] * h[4];
  double c01 = h[2] * h[4] - h[1] * h[5];
  double c02 = h[1] * h[4] - h[2] * h[3];
  double det = h[0] * c00 + h[1] * c01 + h[2] * c02;
  double id = 1.0 / det;
  o[0] = c00 * id; o[1] = c01 * id; o[2] = c02 * id;
  o[3] = (h[0] * h[5] - h[2] * h[2]) * id;
  o[4] = (h[1] * h[2] - h[0] * h[4]) * id;
  o[5] = (h[0] * h[3] - h[1] * h[1]) * id;
  return isfinite(id);
}

// Ceres CauchyLoss(a) + Corrector for rho'' <= 0 (reference rig_calibrator.cc:394-411 always picks
// "cauchy"): cost 0.5 * b log(1 + s/b); residual and Jacobian rows scaled by sqrt(rho').
__device__ __forceinline__ void cauchy(double a, double s, double* half_rho, double* sqrt_rho1) {
  double b = a * a;
  double sum = 1.0 + s / b;
  *half_rho = 0.5 * b * log(sum);
  *sqrt_rho1 = rsqrt(sum);  // sqrt(rho'(s)) = (1 + s / b)^(-1/2): one reciprocal square root instead of a division and a root
}

// Corrector scale alone (the build needs the cost itself only at the first point of a solve: no logarithm otherwise).
__device__ __forceinline__ double cauchy_sr(double a, double s) {
  double b = a * a;
  return rsqrt(1.0 + s / b);
}

}  // namespace mvba
