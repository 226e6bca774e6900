// kernels_tracks.cuh -- the per-track steps that run around every optimisation pass of rig_calibrator
// (SURVEY.md 8(f) rows N1, N2), one THREAD per track, observations read from the same track-major SoA
// arrays the solver uses:
//   triangulate_kernel     multiViewTriangulation        reference rig_calibrator/src/interest_point.cc:649-722
//                          + Triangulate                 interest_point.cc:398-423
//                          + openMVG Triangulation::compute (3 reweighted linear passes)
//                                                         openMVG/multiview/triangulation_nview.cc:87-146
//                          + Convert<DISTORTED, UNDISTORTED_C>  camera_model/camera_params.cc:316-355
//   flag_angle_reproj_kernel  flagOutliersByTriAngleAndReprojErr  rig_calibrator/bin/rig_calibrator.cc:1045-1154
//   flag_exclusion_kernel     flagOutlierByExclusionDist           rig_calibrator.cc:1003-1038
// The inlier flags are one byte per observation (the reference's pid_cid_fid_inlier map); a flag that is 0 on
// entry stays 0 ("outliers never return", rig_calibrator.cc:1528-1532).
#pragma once
#include "items.cuh"

namespace mvba {

// UndistortCentered (camera_params.cc:316-355) of a DISTORTED pixel: returns the undistorted, centred pixel.
//   none:   d_c - (c - half)
//   FOV:    n = (d_c - (c - half)) / f, rd = |n|, ru = tan(rd w) / (2 tan(w / 2)), u = (rd > 1e-5 ? ru / rd : 1) n f
//   radtan: cv::undistortPoints -- 5 fixed-point iterations on the normalised point, then x f
//   RPC:    compute_rpc(d_c, second half of the coefficient block)  (rpc_distortion.cc:362-366)
__device__ inline void undistort_centered(const SensorDev& s, const double* __restrict__ x, double px, double py, double* u) {
  const double f = x[s.x_focal];
  const double cx = x[s.x_center], cy = x[s.x_center + 1];
  const double dcx = px - s.half[0], dcy = py - s.half[1];  // Convert<DISTORTED, DISTORTED_C>
  const double* dist = x + s.x_dist;
  if (s.nd == 0) {
    u[0] = dcx - (cx - s.half[0]); u[1] = dcy - (cy - s.half[1]);
  } else if (s.nd == 1) {
    const double w = dist[0];
    const double nx = (dcx - (cx - s.half[0])) / f, ny = (dcy - (cy - s.half[1])) / f;
    const double rd = sqrt(nx * nx + ny * ny);
    const double ru = tan(rd * w) / (2.0 * tan(w / 2.0));
    double conv = 1.0;
    if (rd > 1e-5) conv = ru / rd;
    u[0] = conv * (nx * f); u[1] = conv * (ny * f);
  } else if (s.nd == 4 || s.nd == 5) {
    const double k1 = dist[0], k2 = dist[1], p1 = dist[2], p2 = dist[3], k3 = s.nd == 5 ? dist[4] : 0.0;
    // src = d_c + half; normalised with the DISTORTED intrinsics (focal, optical centre)
    const double x0 = ((dcx + s.half[0]) - cx) / f, y0 = ((dcy + s.half[1]) - cy) / f;
    double xx = x0, yy = y0;
    for (int it = 0; it < 5; it++) {
      const double r2 = xx * xx + yy * yy;
      const double icdist = 1.0 / (1.0 + ((k3 * r2 + k2) * r2 + k1) * r2);
      const double ddx = 2.0 * p1 * xx * yy + p2 * (r2 + 2.0 * xx * xx);
      const double ddy = p1 * (r2 + 2.0 * yy * yy) + 2.0 * p2 * xx * yy;
      xx = (x0 - ddx) * icdist; yy = (y0 - ddy) * icdist;
    }
    // re-projected with the UNDISTORTED intrinsics (focal, undistorted half size), minus that half size
    u[0] = xx * f; u[1] = yy * f;
  } else {
    const int deg = s.rpc_deg;
    const double* co = dist + s.nd / 2;
    double powx[8], powy[8];
    powx[0] = 1; powy[0] = 1;
    for (int d = 1; d <= deg; d++) { powx[d] = powx[d - 1] * dcx; powy[d] = powy[d - 1] * dcy; }
    double val[4] = {0, 1, 0, 1};
    int ci = 0;
    for (int cnt = 0; cnt < 4; cnt++)
      for (int d = (cnt & 1); d <= deg; d++)
        for (int i = 0; i <= d; i++, ci++) val[cnt] += co[ci] * powx[d - i] * powy[i];
    u[0] = val[0] / val[1]; u[1] = val[2] / val[3];
  }
}

// Eigen's 3x3 inverse (cofactors / determinant), general matrix.
__device__ __forceinline__ void inv3(const double* m, double* o) {
  const double c00 = m[4] * m[8] - m[5] * m[7], c01 = m[5] * m[6] - m[3] * m[8], c02 = m[3] * m[7] - m[4] * m[6];
  const double det = m[0] * c00 + m[1] * c01 + m[2] * c02;
  const double id = 1.0 / det;
  o[0] = c00 * id; o[1] = (m[2] * m[7] - m[1] * m[8]) * id; o[2] = (m[1] * m[5] - m[2] * m[4]) * id;
  o[3] = c01 * id; o[4] = (m[0] * m[8] - m[2] * m[6]) * id; o[5] = (m[2] * m[3] - m[0] * m[5]) * id;
  o[6] = c02 * id; o[7] = (m[1] * m[6] - m[0] * m[7]) * id; o[8] = (m[0] * m[4] - m[1] * m[3]) * id;
}

__device__ __forceinline__ unsigned long long warp_sum_u64(unsigned long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// One thread per track.  P = K [R | t], K = diag(f, f, 1); three iterations of the reweighted normal equations.
__global__ void triangulate_kernel(View v, const double* __restrict__ x, const CamTab* __restrict__ tab, uint8_t* __restrict__ inlier,
                                   double* __restrict__ xyz_out) {
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < v.P; p += gridDim.x * blockDim.x) {
    const int64_t o0 = v.trk_obs0[p], o1 = v.trk_obs0[p + 1];
    int nviews = 0;
    for (int64_t i = o0; i < o1; i++) nviews += inlier[i] ? 1 : 0;
    if (nviews < 2) {  // interest_point.cc:687-698: fewer than two rays, every feature of the track becomes an outlier
      for (int64_t i = o0; i < o1; i++) inlier[i] = 0;
      continue;       // xyz stays as the caller initialised it (the reference leaves a default-constructed vector)
    }
    double X[3] = {0, 0, 0};
    for (int it = 0; it < 3; it++) {
      double AtA[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, Atb[3] = {0, 0, 0};
      for (int64_t i = o0; i < o1; i++) {
        if (!inlier[i]) continue;
        const int c = v.cid[i];
        const SensorDev& sd = v.sens[v.cam_sensor[c]];
        const double f = x[sd.x_focal];
        const float2 pix = v.pix[i];
        double u[2];
        undistort_centered(sd, x, (double)pix.x, (double)pix.y, u);
        const double* R = tab[c].R;
        const double* t = tab[c].t;
        // rows of P
        const double P0[4] = {f * R[0], f * R[1], f * R[2], f * t[0]}, P1[4] = {f * R[3], f * R[4], f * R[5], f * t[1]};
        const double P2[4] = {R[6], R[7], R[8], t[2]};
        double w = 1.0;
        if (it > 0) w = 1.0 / (P2[0] * X[0] + P2[1] * X[1] + P2[2] * X[2] + P2[3]);  // weights[i] = 1 / z of the previous pass
        double v1[3], v2[3];
#pragma unroll
        for (int j = 0; j < 3; j++) {
          v1[j] = w * (P0[j] - u[0] * P2[j]);
          v2[j] = w * (P1[j] - u[1] * P2[j]);
          Atb[j] += w * (v1[j] * (u[0] * P2[3] - P0[3]) + v2[j] * (u[1] * P2[3] - P1[3]));
        }
#pragma unroll
        for (int k = 0; k < 3; k++)
#pragma unroll
          for (int j = 0; j <= k; j++) {
            const double val = v1[j] * v1[k] + v2[j] * v2[k];
            AtA[3 * j + k] += val;
            if (j < k) AtA[3 * k + j] += val;
          }
      }
      double inv[9];
      inv3(AtA, inv);
      X[0] = inv[0] * Atb[0] + inv[1] * Atb[1] + inv[2] * Atb[2];
      X[1] = inv[3] * Atb[0] + inv[4] * Atb[1] + inv[5] * Atb[2];
      X[2] = inv[6] * Atb[0] + inv[7] * Atb[1] + inv[8] * Atb[2];
    }
    xyz_out[3 * (size_t)p] = X[0]; xyz_out[3 * (size_t)p + 1] = X[1]; xyz_out[3 * (size_t)p + 2] = X[2];
    if (!isfinite(X[0]) || !isfinite(X[1]) || !isfinite(X[2]))  // interest_point.cc:704-717
      for (int64_t i = o0; i < o1; i++) inlier[i] = 0;
  }
}

// One thread per track: largest angle between any two inlier rays (camera centre -> xyz), tracks below the minimum
// lose all their features; then every remaining feature whose pixel residual norm exceeds the threshold is dropped.
// counts: [0] features seen by the angle step, [1] removed by it, [2] seen by the reprojection step, [3] removed by it.
__global__ void flag_angle_reproj_kernel(View v, const double* __restrict__ x, const CamTab* __restrict__ tab, double min_angle_deg,
                                         double max_reproj_err, uint8_t* __restrict__ inlier, unsigned long long* __restrict__ counts) {
  unsigned long long c0 = 0, c1 = 0, c2 = 0, c3 = 0;
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < v.P; p += gridDim.x * blockDim.x) {
    const int64_t o0 = v.trk_obs0[p], o1 = v.trk_obs0[p + 1];
    const double* Xp = x + v.x_xyz + 3 * (size_t)p;
    const double X[3] = {Xp[0], Xp[1], Xp[2]};
    double max_angle = 0.0;
    int nin = 0;
    for (int64_t i = o0; i < o1; i++) {
      if (!inlier[i]) continue;
      nin++;
      const int ci = v.cid[i];
      const double* R = tab[ci].R; const double* t = tab[ci].t;
      // camera centre = world_to_cam^-1 * 0 = -R' t
      double r1[3];
#pragma unroll
      for (int a = 0; a < 3; a++) r1[a] = X[a] + (R[a] * t[0] + R[3 + a] * t[1] + R[6 + a] * t[2]);
      const double n1 = sqrt(r1[0] * r1[0] + r1[1] * r1[1] + r1[2] * r1[2]);
      if (n1 > 0) { r1[0] /= n1; r1[1] /= n1; r1[2] /= n1; }  // Eigen normalize(): unchanged when the norm is zero
      for (int64_t j = o0; j < o1; j++) {
        const int cj = v.cid[j];
        if (cj <= ci || !inlier[j]) continue;
        const double* R2 = tab[cj].R; const double* t2 = tab[cj].t;
        double r2[3];
#pragma unroll
        for (int a = 0; a < 3; a++) r2[a] = X[a] + (R2[a] * t2[0] + R2[3 + a] * t2[1] + R2[6 + a] * t2[2]);
        const double n2 = sqrt(r2[0] * r2[0] + r2[1] * r2[1] + r2[2] * r2[2]);
        if (n2 > 0) { r2[0] /= n2; r2[1] /= n2; r2[2] /= n2; }
        const double ang = (180.0 / 3.14159265358979323846) * acos(r1[0] * r2[0] + r1[1] * r2[1] + r1[2] * r2[2]);
        if (isnan(ang) || isinf(ang)) continue;
        max_angle = fmax(max_angle, ang);
      }
    }
    c0 += nin;
    if (!(max_angle >= min_angle_deg)) {
      for (int64_t i = o0; i < o1; i++) if (inlier[i]) { inlier[i] = 0; c1++; }
      continue;
    }
    for (int64_t i = o0; i < o1; i++) {
      if (!inlier[i]) continue;
      c2++;
      Item it;
      eval_cam<false>(v, x, tab, i, it, nullptr, nullptr);
      const bool good = sqrt(it.r[0] * it.r[0] + it.r[1] * it.r[1]) <= max_reproj_err;  // NaN is never an inlier
      if (!good) { inlier[i] = 0; c3++; }
    }
  }
  c0 = warp_sum_u64(c0); c1 = warp_sum_u64(c1); c2 = warp_sum_u64(c2); c3 = warp_sum_u64(c3);
  if ((threadIdx.x & 31) == 0) {
    if (c0) atomicAdd(counts, c0);
    if (c1) atomicAdd(counts + 1, c1);
    if (c2) atomicAdd(counts + 2, c2);
    if (c3) atomicAdd(counts + 3, c3);
  }
}

// One thread per observation: features outside the centred crop box of their sensor are outliers.
__global__ void flag_exclusion_kernel(View v, const double* __restrict__ crop /* 2 per sensor */, const double* __restrict__ size /* 2 per sensor */,
                                      uint8_t* __restrict__ inlier) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < v.N; i += (int64_t)gridDim.x * blockDim.x) {
    const int t = v.cam_sensor[v.cid[i]];
    const float2 pix = v.pix[i];
    // integer image sizes: dist_size / 2.0 and dist_crop_size / 2.0 (rig_calibrator.cc:1031-1034)
    if (fabs((double)pix.x - size[2 * t] / 2.0) > crop[2 * t] / 2.0 || fabs((double)pix.y - size[2 * t + 1] / 2.0) > crop[2 * t + 1] / 2.0)
      inlier[i] = 0;
  }
}

}  // namespace mvba
