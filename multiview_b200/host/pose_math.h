// pose_math.h -- host-side pose algebra of the rig parameterisation, for what a driver does AFTER mvba_solve: turning the
// optimised world_to_ref / ref_to_cam / timestamp-offset arrays into one world_to_cam transform per image (the camera list the
// reference writes, rig_calibrator.cc:1924-1967 + dense_map_utils.cc:753-776).
//   rigid_transform_to_array / array_to_rigid_transform   rig_calibrator/src/dense_map_utils.cc:148-178
//   linearInterp                                          dense_map_utils.cc:315-329 (Eigen slerp: shortest arc, lerp when |dot| >= 1 - eps)
//   calc_interp_world_to_ref / calc_world_to_cam_trans    rig_calibrator/bin/rig_calibrator.cc:322-390
// Transforms are mvba_host::Affine (R row-major, then t), the layout of the text formats in rig_io.h.
#pragma once
#include <cmath>
#include <limits>
#include <stdexcept>
#include <string>
#include <vector>

#include "rig_io.h"

namespace mvba_host {

// [tx ty tz qx qy qz qw] -> transform; the quaternion is normalised first (Eigen Quaterniond::normalized())
inline Affine array_to_rigid(const double* a) {
  double q[4] = {a[6], a[3], a[4], a[5]};  // w x y z
  const double n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  for (double& v : q) v /= n;
  Affine out;
  quaternion_to_rotation(q, out.data());
  out[9] = a[0]; out[10] = a[1]; out[11] = a[2];
  return out;
}

inline void rigid_to_array(const Affine& t, double* a) {
  double q[4];
  rotation_to_quaternion(t.data(), q);
  a[0] = t[9]; a[1] = t[10]; a[2] = t[11];
  a[3] = q[1]; a[4] = q[2]; a[5] = q[3]; a[6] = q[0];
}

// C = A * B (apply B first)
inline Affine compose(const Affine& A, const Affine& B) {
  Affine C;
  for (int i = 0; i < 3; i++) {
    for (int j = 0; j < 3; j++) C[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
    C[9 + i] = A[3 * i] * B[9] + A[3 * i + 1] * B[10] + A[3 * i + 2] * B[11] + A[9 + i];
  }
  return C;
}

inline Affine linear_interp(double alpha, const Affine& a0, const Affine& a1) {
  double q0[4], q1[4], q[4];
  rotation_to_quaternion(a0.data(), q0);
  rotation_to_quaternion(a1.data(), q1);
  const double d = q0[0] * q1[0] + q0[1] * q1[1] + q0[2] * q1[2] + q0[3] * q1[3], ad = std::fabs(d);
  double s0, s1;
  if (ad >= 1.0 - std::numeric_limits<double>::epsilon()) {
    s0 = 1.0 - alpha; s1 = alpha;
  } else {
    const double theta = std::acos(ad), st = std::sin(theta);
    s0 = std::sin((1.0 - alpha) * theta) / st;
    s1 = std::sin(alpha * theta) / st;
  }
  if (d < 0.0) s1 = -s1;
  for (int k = 0; k < 4; k++) q[k] = s0 * q0[k] + s1 * q1[k];
  Affine out;
  quaternion_to_rotation(q, out.data());
  for (int i = 0; i < 3; i++) out[9 + i] = (1.0 - alpha) * a0[9 + i] + alpha * a1[9 + i];
  return out;
}

// world_to_cam of one image from the parameter arrays (7 doubles each).  beg == end (reference sensor, or --no_rig): the pose
// itself.  Throws where the reference aborts (rig_calibrator.cc:346).
inline Affine calc_world_to_cam(const double* beg_world_to_ref, const double* end_world_to_ref, const double* ref_to_cam, double beg_stamp,
                                double end_stamp, double ref_to_cam_offset, double cam_stamp) {
  if (beg_stamp == end_stamp) return array_to_rigid(beg_world_to_ref);
  const double alpha = ((cam_stamp - beg_stamp) - ref_to_cam_offset) / (end_stamp - beg_stamp);
  if (alpha < 0.0 || alpha > 1.0) throw std::runtime_error("Out of bounds in interpolation.");
  const Affine w2r = linear_interp(alpha, array_to_rigid(beg_world_to_ref), array_to_rigid(end_world_to_ref));
  return compose(array_to_rigid(ref_to_cam), w2r);
}

// Every image of a pass, in the SoA form mvba_set_cameras takes (sensor, bracketing ref frames, timestamp).
inline std::vector<Affine> world_to_cam_all(int num_cameras, const int* cam_sensor, const int* beg, const int* end, const double* cam_stamp,
                                            const double* ref_stamps, const double* world_to_ref, const double* ref_to_cam, const double* offsets) {
  std::vector<Affine> out((size_t)num_cameras);
  for (int c = 0; c < num_cameras; c++) {
    const int t = cam_sensor[c];
    out[c] = calc_world_to_cam(world_to_ref + 7 * beg[c], world_to_ref + 7 * end[c], ref_to_cam + 7 * t, ref_stamps[beg[c]], ref_stamps[end[c]],
                               offsets[t], cam_stamp[c]);
  }
  return out;
}

// Eigen::Affine3d::inverse() (Affine mode: general inverse of the 3x3 part, t' = -L^-1 t)
inline Affine inverse(const Affine& A) {
  const double a = A[0], b = A[1], c = A[2], d = A[3], e = A[4], f = A[5], g = A[6], h = A[7], i = A[8];
  const double det = a * (e * i - f * h) - b * (d * i - f * g) + c * (d * h - e * g);
  Affine B;
  B[0] = (e * i - f * h) / det; B[1] = (c * h - b * i) / det; B[2] = (b * f - c * e) / det;
  B[3] = (f * g - d * i) / det; B[4] = (a * i - c * g) / det; B[5] = (c * d - a * f) / det;
  B[6] = (d * h - e * g) / det; B[7] = (b * g - a * h) / det; B[8] = (a * e - b * d) / det;
  for (int r = 0; r < 3; r++) B[9 + r] = -(B[3 * r] * A[9] + B[3 * r + 1] * A[10] + B[3 * r + 2] * A[11]);
  return B;
}

// calc_rig_using_word_to_cam, rig_calibrator.cc:1190-1262: the rig transforms when --use_initial_rig_transforms is off.  Per image
// ref_to_cam = world_to_cam * interp(world_to_ref)^-1 (identity for the reference sensor); per sensor and per matrix entry the
// MIDDLE element of that list in image order (vals[vals.size() / 2]: the reference calls it a median but does not sort), then the
// 3x3 part is divided by the cube root of its determinant.  Throws where the reference aborts (a sensor without images).
inline std::vector<Affine> calc_rig_using_world_to_cam(int num_sensors, int num_cameras, const int* cam_sensor, const int* beg, const int* end,
                                                       const double* cam_stamp, const double* ref_stamps, const double* world_to_ref,
                                                       const std::vector<Affine>& world_to_cam, const double* offsets) {
  if ((size_t)num_cameras != world_to_cam.size()) throw std::runtime_error("There must be as many world to cam transforms as metadata sets for them.\n");
  std::vector<std::vector<Affine>> per(num_sensors);
  double ident7[7] = {0, 0, 0, 0, 0, 0, 1};
  for (int c = 0; c < num_cameras; c++) {
    const int t = cam_sensor[c];
    if (t == 0) { per[t].push_back(identity_affine()); continue; }
    const Affine w2r = calc_world_to_cam(world_to_ref + 7 * beg[c], world_to_ref + 7 * end[c], ident7, ref_stamps[beg[c]], ref_stamps[end[c]],
                                         offsets[t], cam_stamp[c]);
    per[t].push_back(compose(world_to_cam[c], inverse(w2r)));
  }
  std::vector<Affine> out(num_sensors);
  for (int t = 0; t < num_sensors; t++) {
    if (per[t].empty()) throw std::runtime_error("No poses were found for rig sensor with id: " + std::to_string(t) + "\n");
    Affine m = per[t][per[t].size() / 2];   // entry by entry the same element
    const double det = m[0] * (m[4] * m[8] - m[5] * m[7]) - m[1] * (m[3] * m[8] - m[5] * m[6]) + m[2] * (m[3] * m[7] - m[4] * m[6]);
    const double s = std::pow(det, 1.0 / 3.0);
    for (int k = 0; k < 9; k++) m[k] /= s;
    out[t] = m;
  }
  return out;
}

}  // namespace mvba_host
