// rig_io.h -- host-side (C++17, dependency free) readers and writers of the three text formats the rig_calibrator
// path exchanges with its user, byte-compatible with the reference:
//   rig_config.txt   writer  rig_calibrator/src/dense_map_utils.cc:779-850   reader :855-1057
//   cameras.txt      writer  dense_map_utils.cc:753-776                     reader rig_calibrator/src/interest_point.cc:1673-1738
//                    (sensor = parent directory name, timestamp = atof(basename): interest_point.cc:1619-1627)
//   NVM_V3           writer  interest_point.cc:1405-1475 (+ the inlier export :1338-1401)   reader :1240-1330
// Numbers are printed with 17 significant digits in the C++ stream default ("%.17g") format.
// Errors are reported by exception (std::runtime_error) where the reference calls LOG(FATAL).
#pragma once
#include <array>
#include <map>
#include <string>
#include <vector>

namespace mvba_host {

using Affine = std::array<double, 12>;  // file order: R row-major (9) then t (3)   (transform_utils.cc:30-72)
inline Affine identity_affine() { return Affine{1, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0, 0}; }

struct SensorConfig {
  std::string name;
  double focal_length = 0;                 // one scalar; the functors expand it to (f, f)
  double optical_center[2] = {0, 0};
  std::vector<double> distortion;          // 0 | 1 (fisheye) | 4-5 (radtan) | > 5 (rpc: distortion half + undistortion half)
  int image_size[2] = {0, 0};
  int distorted_crop_size[2] = {0, 0};
  int undistorted_image_size[2] = {0, 0};
  Affine ref_to_sensor = identity_affine();
  Affine depth_to_image = identity_affine();
  double ref_to_sensor_timestamp_offset = 0;
};

struct RigConfig {
  std::string ref_sensor_name;
  std::vector<SensorConfig> sensors;       // sensors[0] is the reference sensor
};

const char* distortion_type_name(size_t num_coeffs);  // no_distortion | fisheye | radtan | rpc
std::string affine_to_str(const Affine& a);
std::string format_double(double v);                    // precision(17) stream formatting

// have_rig_transforms: refuse an all-zero transform and require identity for the reference sensor (dense_map_utils.cc:1033-1049)
RigConfig read_rig_config(const std::string& path, bool have_rig_transforms);
// model_rig = false writes identity rig transforms ("write something valid", dense_map_utils.cc:836-839)
void write_rig_config(const std::string& path, const RigConfig& rig, bool model_rig);

struct CameraEntry {
  std::string image_file;
  Affine world_to_cam = identity_affine();
  std::string sensor_name;   // parent directory of image_file
  double timestamp = 0;      // atof(basename)
};
std::vector<CameraEntry> read_camera_list(const std::string& path);
void write_camera_list(const std::string& path, const std::vector<CameraEntry>& cams);

struct Nvm {
  std::vector<std::string> image_files;
  std::vector<double> focal_lengths;
  std::vector<Affine> world_to_cam;                          // rotation from the quaternion, t = -R c
  std::vector<std::vector<std::array<double, 2>>> keypoints; // per camera, relative to the optical centre
  std::vector<std::map<int, int>> pid_to_cid_fid;
  std::vector<std::array<double, 3>> xyz;
};
Nvm read_nvm(const std::string& path);
void write_nvm(const std::string& path, const Nvm& nvm);

// The inlier export of the reference (interest_point.cc:1338-1401): keypoints are shifted by the optical centre of their
// sensor, features re-numbered per camera in track order, and only tracks with MORE THAN TWO inlier views are kept
// (the reference's `> 2`, despite its comment).  Observations are track-major (pid ascending, cid ascending).
Nvm nvm_from_inliers(const std::vector<std::string>& image_files, const std::vector<int>& cam_sensor, const std::vector<double>& sensor_focal,
                     const std::vector<std::array<double, 2>>& sensor_center, const std::vector<Affine>& world_to_cam,
                     const std::vector<int>& pid, const std::vector<int>& cid, const std::vector<float>& pix,
                     const std::vector<unsigned char>& inlier, const std::vector<double>& xyz);

// rotation matrix (row-major) <-> quaternion (w, x, y, z), Eigen 3.4 conventions
void rotation_to_quaternion(const double* R, double* q_wxyz);
void quaternion_to_rotation(const double* q_wxyz, double* R);

}  // namespace mvba_host
