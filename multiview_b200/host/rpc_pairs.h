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
