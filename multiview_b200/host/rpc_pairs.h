// rpc_pairs.h -- sample pairs for the RPC fit (reference camera_model/rpc_distortion.cc:497-553, genUndistDistPairs):
// a num_samples x num_samples grid over the UNDISTORTED image, pushed through the camera's own distortion model, kept when
// the distorted pixel lies inside the centred crop box, both ends centred with the half sizes camera_params.cc uses.
// The pairs feed mvba_rpc_fit (include/mvba.h): fitRpcDist fits undist -> dist, fitRpcUndist dist -> undist.
#pragma once
#include <cmath>
#include <vector>

#include "camera_params.h"

namespace mvba_host {

inline void gen_undist_dist_pairs(int num_samples, const camera::CameraParameters& cam, std::vector<double>* undist_centered_xy,
                                  std::vector<double>* dist_centered_xy) {
  undist_centered_xy->clear(); dist_centered_xy->clear();
  const camera::Vector2i dist_size = cam.GetDistortedSize(), crop = cam.GetDistortedCropSize(), undist_size = cam.GetUndistortedSize();
  const camera::Vector2d dist_half = cam.GetDistortedHalfSize(), undist_half = cam.GetUndistortedHalfSize();
  for (int ix = 0; ix < num_samples; ix++) {
    const double x = (undist_size[0] - 1.0) * ix / (num_samples - 1.0);
    for (int iy = 0; iy < num_samples; iy++) {
      const double y = (undist_size[1] - 1.0) * iy / (num_samples - 1.0);
      camera::Vector2d u(x, y), d;
      cam.Convert<camera::UNDISTORTED, camera::DISTORTED>(u, &d);
      if (std::abs(d[0] - dist_size[0] / 2.0) > crop[0] / 2.0 || std::abs(d[1] - dist_size[1] / 2.0) > crop[1] / 2.0) continue;
      undist_centered_xy->push_back(u[0] - undist_half[0]); undist_centered_xy->push_back(u[1] - undist_half[1]);
      dist_centered_xy->push_back(d[0] - dist_half[0]); dist_centered_xy->push_back(d[1] - dist_half[1]);
    }
  }
}

}  // namespace mvba_host

// CameraParameters::updateRpcUndistortion (camera_params.cc:214-254 -> fitRpcUndist, rpc_distortion.cc:658-719): samples of the
// undistorted image pushed through the RPC DISTORTION half, then an RPC of the same degree is fitted the other way round
// (distorted -> undistorted) on the device and stored as the second half of the coefficient vector.  num_threads is the
// reference's Ceres thread count and has no meaning here.
inline void camera::CameraParameters::updateRpcUndistortion(int /*num_threads*/) {
  const int num_samples = 400, num_iterations = 100;
  const double parameter_tolerance = 1e-12;
  if (distortion_coeffs_.size() % 2 != 0) throw std::runtime_error("Must have an even number of RPC distortion coefficients.");
  if (!rpc_fitter()) throw std::runtime_error("updateRpcUndistortion: no RPC fitter registered (CameraParameters::SetRpcFitter(mvba_rpc_fit))");
  const int num_dist = (int)distortion_coeffs_.size() / 2;
  const int deg = rpc_degree(num_dist);
  std::vector<double> undist, dist;
  mvba_host::gen_undist_dist_pairs(num_samples, *this, &undist, &dist);
  for (size_t i = 0; i + 1 < undist.size(); i += 2) {  // the pairs of the fit: RPC distortion of the centred undistorted sample
    const Vector2d d = compute_rpc(Vector2d(undist[i], undist[i + 1]), distortion_coeffs_.data(), num_dist);
    dist[i] = d[0]; dist[i + 1] = d[1];
  }
  std::vector<double> coeffs((size_t)num_dist, 0.0);
  double stats[4] = {0, 0, 0, 0};
  const int rc = rpc_fitter()(0, (int64_t)(undist.size() / 2), dist.data(), undist.data(), deg, num_iterations, parameter_tolerance, 0, coeffs.data(), stats);
  if (rc != 0) throw std::runtime_error("updateRpcUndistortion: the device RPC fit failed (code " + std::to_string(rc) + ")");
  for (int i = 0; i < num_dist; i++) distortion_coeffs_[(size_t)num_dist + i] = coeffs[(size_t)i];
  can_undistort_ = true;
}
