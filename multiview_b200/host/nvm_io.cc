// nvm_io.cc -- NVM_V3 reader / writer and the inlier export (see rig_io.h), plus the C test shims of the host library.
#include <cmath>
#include <cstring>
#include <fstream>
#include <stdexcept>

#include "rig_io.h"

namespace mvba_host {

Nvm read_nvm(const std::string& path) {
  std::ifstream f(path, std::ios::in);
  if (!f.good()) throw std::runtime_error("Unable to open file: " + path);
  std::string token;
  f >> token;
  if (token != "NVM_V3") throw std::runtime_error("File doesn't start with NVM token");
  long ncam = 0;
  f >> ncam;
  if (ncam < 1) throw std::runtime_error("NVM file is missing cameras");
  Nvm n;
  n.image_files.resize(ncam); n.focal_lengths.resize(ncam); n.world_to_cam.resize(ncam); n.keypoints.resize(ncam);
  for (long c = 0; c < ncam; c++) {
    double q[4], ctr[3], d1, d2;
    f >> n.image_files[c] >> n.focal_lengths[c] >> q[0] >> q[1] >> q[2] >> q[3] >> ctr[0] >> ctr[1] >> ctr[2] >> d1 >> d2;
    double R[9];
    quaternion_to_rotation(q, R);
    Affine& a = n.world_to_cam[c];
    for (int i = 0; i < 9; i++) a[i] = R[i];
    for (int i = 0; i < 3; i++) a[9 + i] = -(R[3 * i] * ctr[0] + R[3 * i + 1] * ctr[1] + R[3 * i + 2] * ctr[2]);  // t = -R c
  }
  long npts = 0;
  f >> npts;
  if (npts < 1) throw std::runtime_error("The NVM file has no triangulated points.");
  n.pid_to_cid_fid.resize(npts); n.xyz.resize(npts);
  for (long p = 0; p < npts; p++) {
    int color[3]; long nm = 0;
    f >> n.xyz[p][0] >> n.xyz[p][1] >> n.xyz[p][2] >> color[0] >> color[1] >> color[2] >> nm;
    for (long m = 0; m < nm; m++) {
      long cid, fid; double x, y;
      f >> cid >> fid >> x >> y;
      if (!f.good() || cid < 0 || cid >= ncam || fid < 0) throw std::runtime_error("Unable to correctly read PID: " + std::to_string(p));
      n.pid_to_cid_fid[p][(int)cid] = (int)fid;
      auto& kp = n.keypoints[cid];
      if ((long)kp.size() <= fid) kp.resize(fid + 1, {0.0, 0.0});
      kp[fid] = {x, y};
    }
    if (!f.good()) throw std::runtime_error("Unable to correctly read PID: " + std::to_string(p));
  }
  return n;
}

void write_nvm(const std::string& path, const Nvm& n) {
  if (n.image_files.size() != n.keypoints.size()) throw std::runtime_error("Unequal number of filenames and keypoints");
  if (n.pid_to_cid_fid.size() != n.xyz.size()) throw std::runtime_error("Unequal number of pid_to_cid_fid and xyz measurements");
  if (n.image_files.size() != n.world_to_cam.size()) throw std::runtime_error("Unequal number of filename and camera transforms");
  std::ofstream f(path, std::ios::out);
  if (!f.is_open()) throw std::runtime_error("Cannot open file for writing: " + path);
  f << "NVM_V3\n";
  f << n.image_files.size() << "\n";
  for (size_t c = 0; c < n.image_files.size(); c++) {
    const Affine& a = n.world_to_cam[c];
    double q[4];
    rotation_to_quaternion(a.data(), q);
    // camera centre = -R^-1 t; Eigen inverts the 3x3 generally, here R' is used (R is a rotation)
    double ctr[3];
    for (int i = 0; i < 3; i++) ctr[i] = -(a[i] * a[9] + a[3 + i] * a[10] + a[6 + i] * a[11]);
    f << n.image_files[c] << " " << format_double(n.focal_lengths[c]) << " " << format_double(q[0]) << " " << format_double(q[1]) << " "
      << format_double(q[2]) << " " << format_double(q[3]) << " " << format_double(ctr[0]) << " " << format_double(ctr[1]) << " "
      << format_double(ctr[2]) << " " << "0 0\n";
  }
  f << n.pid_to_cid_fid.size() << "\n";
  for (size_t p = 0; p < n.pid_to_cid_fid.size(); p++) {
    const auto& m = n.pid_to_cid_fid[p];
    if (m.size() <= 1) throw std::runtime_error("PID " + std::to_string(p) + " has " + std::to_string(m.size()) + " measurements");
    f << format_double(n.xyz[p][0]) << " " << format_double(n.xyz[p][1]) << " " << format_double(n.xyz[p][2]) << " 0 0 0 " << m.size();
    for (auto& cf : m) {
      const auto& kp = n.keypoints[cf.first][cf.second];
      f << " " << cf.first << " " << cf.second << " " << format_double(kp[0]) << " " << format_double(kp[1]);
    }
    f << "\n";
  }
}

Nvm nvm_from_inliers(const std::vector<std::string>& image_files, const std::vector<int>& cam_sensor, const std::vector<double>& sensor_focal,
                     const std::vector<std::array<double, 2>>& sensor_center, const std::vector<Affine>& world_to_cam,
                     const std::vector<int>& pid, const std::vector<int>& cid, const std::vector<float>& pix,
                     const std::vector<unsigned char>& inlier, const std::vector<double>& xyz) {
  Nvm n;
  const size_t C = image_files.size();
  n.image_files = image_files; n.world_to_cam = world_to_cam;
  n.focal_lengths.resize(C); n.keypoints.resize(C);
  for (size_t c = 0; c < C; c++) n.focal_lengths[c] = sensor_focal[cam_sensor[c]];
  size_t i = 0;
  while (i < pid.size()) {
    size_t j = i;
    while (j < pid.size() && pid[j] == pid[i]) j++;
    std::map<int, int> cf;
    for (size_t q = i; q < j; q++) {
      if (!inlier[q]) continue;
      const int c = cid[q], t = cam_sensor[c];
      // keypoints enter the file relative to the optical centre; every inlier feature is added, kept track or not
      n.keypoints[c].push_back({(double)pix[2 * q] - sensor_center[t][0], (double)pix[2 * q + 1] - sensor_center[t][1]});
      cf[c] = (int)n.keypoints[c].size() - 1;
    }
    if (cf.size() > 2) {  // sic: more than two views (interest_point.cc:1393)
      n.pid_to_cid_fid.push_back(cf);
      n.xyz.push_back({xyz[3 * (size_t)pid[i]], xyz[3 * (size_t)pid[i] + 1], xyz[3 * (size_t)pid[i] + 2]});
    }
    i = j;
  }
  return n;
}

}  // namespace mvba_host

// ---- C shims used by tests/test_host_io.py (ctypes): each returns 0 or -1 with the message in mvh_last_error() ----
static thread_local std::string g_err;
extern "C" {
const char* mvh_last_error() { return g_err.c_str(); }
int mvh_rig_config_roundtrip(const char* in_path, const char* out_path, int have_rig_transforms, int model_rig, int* num_sensors) {
  try {
    auto rig = mvba_host::read_rig_config(in_path, have_rig_transforms != 0);
    if (num_sensors) *num_sensors = (int)rig.sensors.size();
    mvba_host::write_rig_config(out_path, rig, model_rig != 0);
    return 0;
  } catch (const std::exception& e) { g_err = e.what(); return -1; }
}
int mvh_camera_list_roundtrip(const char* in_path, const char* out_path, int* n, double* first_stamp, char* first_sensor, int cap) {
  try {
    auto cams = mvba_host::read_camera_list(in_path);
    if (n) *n = (int)cams.size();
    if (!cams.empty()) {
      if (first_stamp) *first_stamp = cams[0].timestamp;
      if (first_sensor && cap > 0) { std::strncpy(first_sensor, cams[0].sensor_name.c_str(), cap - 1); first_sensor[cap - 1] = 0; }
    }
    mvba_host::write_camera_list(out_path, cams);
    return 0;
  } catch (const std::exception& e) { g_err = e.what(); return -1; }
}
int mvh_nvm_roundtrip(const char* in_path, const char* out_path, int* num_cams, int* num_points) {
  try {
    auto nvm = mvba_host::read_nvm(in_path);
    if (num_cams) *num_cams = (int)nvm.image_files.size();
    if (num_points) *num_points = (int)nvm.pid_to_cid_fid.size();
    mvba_host::write_nvm(out_path, nvm);
    return 0;
  } catch (const std::exception& e) { g_err = e.what(); return -1; }
}
// Inlier export: image names are "<sensor>/<10000 + c>.jpg"-style strings supplied by the caller, one per line in names_path.
int mvh_nvm_export(const char* names_path, const char* out_path, int C, const int* cam_sensor, int T, const double* focal, const double* center,
                   const double* world_to_cam12, long long N, const int* pid, const int* cid, const float* pix, const unsigned char* inlier,
                   int P, const double* xyz) {
  try {
    std::vector<std::string> names;
    { std::ifstream f(names_path); std::string s; while (std::getline(f, s)) if (!s.empty()) names.push_back(s); }
    if ((int)names.size() != C) throw std::runtime_error("mvh_nvm_export: expected one image name per camera");
    std::vector<int> cs(cam_sensor, cam_sensor + C);
    std::vector<double> fo(focal, focal + T);
    std::vector<std::array<double, 2>> ce(T);
    for (int t = 0; t < T; t++) ce[t] = {center[2 * t], center[2 * t + 1]};
    std::vector<mvba_host::Affine> w2c(C);
    for (int c = 0; c < C; c++) for (int k = 0; k < 12; k++) w2c[c][k] = world_to_cam12[12 * c + k];
    auto nvm = mvba_host::nvm_from_inliers(names, cs, fo, ce, w2c, std::vector<int>(pid, pid + N), std::vector<int>(cid, cid + N),
                                           std::vector<float>(pix, pix + 2 * N), std::vector<unsigned char>(inlier, inlier + N),
                                           std::vector<double>(xyz, xyz + 3 * (size_t)P));
    mvba_host::write_nvm(out_path, nvm);
    return 0;
  } catch (const std::exception& e) { g_err = e.what(); return -1; }
}
}
