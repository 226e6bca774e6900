// pass_utils.h -- small host-side steps of one rig_calibrator pass that sit between the device calls of include/mvba.h.
//   depth_value            reference rig_calibrator/src/dense_map_utils.cc:1362-1391 (depthValue: nearest-pixel lookup in a
//                          camera's depth cloud, rows x cols x 3 float32, (0, 0, 0) = no measurement)
//   compact_observations   the `if (!inlier) continue` of the residual loops (rig_calibrator/bin/rig_calibrator.cc:1660-1668,
//                          1744-1746): the observation arrays of the next pass hold the inlier features only, in the same
//                          order; landmark ids are kept, so a track that lost every feature simply has no residual
#pragma once
#include <cmath>
#include <cstdint>
#include <stdexcept>
#include <vector>

namespace mvba_host {

// Returns true and fills xyz when the cloud has a valid measurement at round(dist_ip); false (xyz = 0) otherwise.
inline bool depth_value(const float* cloud, int rows, int cols, const double dist_ip[2], double xyz[3]) {
  xyz[0] = xyz[1] = xyz[2] = 0.0;
  if (cols == 0 && rows == 0) return false;  // empty cloud
  const int col = (int)std::round(dist_ip[0]), row = (int)std::round(dist_ip[1]);
  if (col < 0 || row < 0 || col > cols || row > rows) throw std::runtime_error("Out of range in the depth cloud.");
  if (col == cols || row == rows) return false;  // after rounding one may hit the bound
  const float* v = cloud + 3 * ((size_t)row * cols + col);
  if (v[0] == 0.0f && v[1] == 0.0f && v[2] == 0.0f) return false;
  xyz[0] = v[0]; xyz[1] = v[1]; xyz[2] = v[2];
  return true;
}

struct Observations {
  std::vector<int32_t> pid, cid;
  std::vector<float> pix;             // 2 per observation
  std::vector<uint8_t> has_depth, has_mesh;
  std::vector<double> depth_xyz, mesh_xyz;  // 3 per observation (may be empty)
  std::vector<int64_t> source;        // index of the feature in the uncompacted arrays
};

inline Observations compact_observations(int64_t n, const int32_t* pid, const int32_t* cid, const float* pix, const uint8_t* has_depth,
                                         const double* depth_xyz, const uint8_t* has_mesh, const double* mesh_xyz, const uint8_t* inlier) {
  Observations o;
  for (int64_t i = 0; i < n; i++) {
    if (!inlier[i]) continue;
    o.pid.push_back(pid[i]); o.cid.push_back(cid[i]); o.pix.push_back(pix[2 * i]); o.pix.push_back(pix[2 * i + 1]); o.source.push_back(i);
    if (has_depth) { o.has_depth.push_back(has_depth[i]); for (int a = 0; a < 3; a++) o.depth_xyz.push_back(depth_xyz[3 * i + a]); }
    if (has_mesh) { o.has_mesh.push_back(has_mesh[i]); for (int a = 0; a < 3; a++) o.mesh_xyz.push_back(mesh_xyz[3 * i + a]); }
  }
  return o;
}

}  // namespace mvba_host
