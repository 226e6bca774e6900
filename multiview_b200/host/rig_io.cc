// rig_io.cc -- see rig_io.h.  Written against the reference's file formats, not its code structure: one tokenising
// line reader shared by the three formats, plain structs instead of Eigen / OpenCV types.
#include "rig_io.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <sstream>
#include <stdexcept>

namespace mvba_host {

std::string format_double(double v) {
  char buf[64];
  std::snprintf(buf, sizeof(buf), "%.17g", v);
  return buf;
}

std::string affine_to_str(const Affine& a) {
  std::string s;
  for (int i = 0; i < 12; i++) { if (i) s += ' '; s += format_double(a[i]); }
  return s;
}

const char* distortion_type_name(size_t n) {
  if (n == 0) return "no_distortion";
  if (n == 1) return "fisheye";
  if (n == 4 || n == 5) return "radtan";
  if (n > 5) return "rpc";
  throw std::runtime_error("Expecting 0, 1, 4, 5, or more distortion coefficients. Got: " + std::to_string(n) + ".");
}

namespace {

// "tag: v0 v1 ..." lines; '#' starts a comment; blank lines are skipped.
struct TagReader {
  std::ifstream f;
  explicit TagReader(const std::string& path) : f(path, std::ios::binary) {
    if (!f.is_open()) throw std::runtime_error("Cannot open file for reading: " + path);
  }
  // next non-empty line split into tag + value tokens; false at end of file
  bool next(std::string& tag, std::vector<std::string>& vals) {
    std::string line;
    while (std::getline(f, line)) {
      const size_t hash = line.find('#');
      if (hash != std::string::npos) line.erase(hash);
      std::istringstream iss(line);
      if (!(iss >> tag)) continue;
      vals.clear();
      std::string v;
      while (iss >> v) vals.push_back(v);
      return true;
    }
    return false;
  }
  std::vector<std::string> strings(const std::string& want, int n) {
    std::string tag; std::vector<std::string> vals;
    if (!next(tag, vals) || tag != want) throw std::runtime_error("Could not read value for: " + want);
    if (n >= 0 && (int)vals.size() != n) throw std::runtime_error("Read an incorrect number of values for: " + want);
    return vals;
  }
  std::vector<double> numbers(const std::string& want, int n) {
    std::string tag; std::vector<std::string> vals;
    if (!next(tag, vals) || tag != want) throw std::runtime_error("Could not read value for: " + want);
    std::vector<double> out;
    for (auto& v : vals) {  // like `iss >> double`: stop at the first token that is not a number
      char* end = nullptr;
      double d = std::strtod(v.c_str(), &end);
      if (end == v.c_str() || *end) break;
      out.push_back(d);
    }
    if (n >= 0 && (int)out.size() != n) throw std::runtime_error("Read an incorrect number of values for: " + want);
    return out;
  }
};

Affine to_affine(const std::vector<double>& v) {
  if (v.size() != 12) throw std::runtime_error("An affine transform must have 12 parameters.");
  Affine a; for (int i = 0; i < 12; i++) a[i] = v[i];
  return a;
}

}  // namespace

RigConfig read_rig_config(const std::string& path, bool have_rig_transforms) {
  TagReader r(path);
  RigConfig rig;
  rig.ref_sensor_name = r.strings("ref_sensor_name:", 1)[0];
  for (;;) {
    std::string tag; std::vector<std::string> vals;
    if (!r.next(tag, vals)) break;  // no more sensors
    if (tag != "sensor_name:" || vals.size() != 1) break;
    SensorConfig s;
    s.name = vals[0];
    const bool first = rig.sensors.empty();
    if ((first && s.name != rig.ref_sensor_name) || (!first && s.name == rig.ref_sensor_name))
      throw std::runtime_error("The reference sensor must be the first sensor specified in the rig configuration.");
    s.focal_length = r.numbers("focal_length:", 1)[0];
    auto c = r.numbers("optical_center:", 2);
    s.optical_center[0] = c[0]; s.optical_center[1] = c[1];
    s.distortion = r.numbers("distortion_coeffs:", -1);
    const size_t nd = s.distortion.size();
    if (nd != 0 && nd != 1 && nd != 4 && nd < 5) throw std::runtime_error("Expecting 0, 1, 4, 5, or more distortion coefficients.");
    const std::string type = r.strings("distortion_type:", 1)[0];
    if (type != distortion_type_name(nd))
      throw std::runtime_error(std::string("With ") + std::to_string(nd) + " distortion coefficients, distortion type must be: " + distortion_type_name(nd));
    auto sz = r.numbers("image_size:", 2);
    s.image_size[0] = (int)sz[0]; s.image_size[1] = (int)sz[1];
    sz = r.numbers("distorted_crop_size:", 2);
    s.distorted_crop_size[0] = (int)sz[0]; s.distorted_crop_size[1] = (int)sz[1];
    sz = r.numbers("undistorted_image_size:", 2);
    s.undistorted_image_size[0] = (int)sz[0]; s.undistorted_image_size[1] = (int)sz[1];
    s.ref_to_sensor = to_affine(r.numbers("ref_to_sensor_transform:", 12));
    if (have_rig_transforms) {
      bool all_zero = true;
      for (double v : s.ref_to_sensor) all_zero = all_zero && v == 0.0;
      if (all_zero) throw std::runtime_error("Failed to read valid transforms among the sensors on the rig");
    }
    s.depth_to_image = to_affine(r.numbers("depth_to_image_transform:", 12));
    s.ref_to_sensor_timestamp_offset = r.numbers("ref_to_sensor_timestamp_offset:", 1)[0];
    rig.sensors.push_back(s);
  }
  if (have_rig_transforms && !rig.sensors.empty() && rig.sensors[0].ref_to_sensor != identity_affine())
    throw std::runtime_error("The transform from the reference sensor to itself must be the identity.");
  return rig;
}

void write_rig_config(const std::string& path, const RigConfig& rig, bool model_rig) {
  std::ofstream f(path, std::ios::binary);
  if (!f.is_open()) throw std::runtime_error("Cannot open file for writing: " + path);
  if (rig.sensors.empty() || rig.sensors[0].name != rig.ref_sensor_name)
    throw std::runtime_error("The reference camera id must be always 0.");
  f << "ref_sensor_name: " << rig.ref_sensor_name << "\n";
  for (const SensorConfig& s : rig.sensors) {
    f << "\n";
    f << "sensor_name: " << s.name << "\n";
    f << "focal_length: " << format_double(s.focal_length) << "\n";
    f << "optical_center: " << format_double(s.optical_center[0]) << " " << format_double(s.optical_center[1]) << "\n";
    f << "distortion_coeffs: ";
    for (size_t d = 0; d < s.distortion.size(); d++) { f << format_double(s.distortion[d]); if (d + 1 < s.distortion.size()) f << " "; }
    f << "\n";
    f << "distortion_type: " << distortion_type_name(s.distortion.size()) << "\n";
    f << "image_size: " << s.image_size[0] << ' ' << s.image_size[1] << "\n";
    f << "distorted_crop_size: " << s.distorted_crop_size[0] << ' ' << s.distorted_crop_size[1] << "\n";
    f << "undistorted_image_size: " << s.undistorted_image_size[0] << ' ' << s.undistorted_image_size[1] << "\n";
    f << "ref_to_sensor_transform: " << affine_to_str(model_rig ? s.ref_to_sensor : identity_affine()) << "\n";
    f << "depth_to_image_transform: " << affine_to_str(s.depth_to_image) << "\n";
    f << "ref_to_sensor_timestamp_offset: " << format_double(s.ref_to_sensor_timestamp_offset) << "\n";
  }
}

// ---------------------------------------------------------------------------------------------------- camera list
static void split_image_path(const std::string& file, std::string& sensor, double& stamp) {
  // .../<sensor>/<timestamp>.<ext>
  const size_t slash = file.find_last_of('/');
  const std::string base = slash == std::string::npos ? file : file.substr(slash + 1);
  const std::string dir = slash == std::string::npos ? std::string() : file.substr(0, slash);
  const size_t s2 = dir.find_last_of('/');
  sensor = s2 == std::string::npos ? dir : dir.substr(s2 + 1);
  stamp = std::atof(base.c_str());
}

std::vector<CameraEntry> read_camera_list(const std::string& path) {
  std::ifstream f(path, std::ios::binary);
  if (!f.is_open()) throw std::runtime_error("Cannot open file for reading: " + path);
  std::vector<CameraEntry> out;
  std::string line;
  while (std::getline(f, line)) {
    if (line.empty() || line[0] == '#') continue;
    std::istringstream iss(line);
    CameraEntry e;
    if (!(iss >> e.image_file)) throw std::runtime_error("Cannot parse the image file in: " + path);
    int count = 0;
    double v;
    while (count < 12 && (iss >> v)) e.world_to_cam[count++] = v;
    if (count != 12) throw std::runtime_error("Expecting 12 values for the transform on line:\n" + line);
    split_image_path(e.image_file, e.sensor_name, e.timestamp);
    out.push_back(e);
  }
  return out;
}

void write_camera_list(const std::string& path, const std::vector<CameraEntry>& cams) {
  std::ofstream f(path, std::ios::binary);
  if (!f.is_open()) throw std::runtime_error("Cannot open file for writing: " + path);
  f << "# image_file world_to_image_transform\n";
  for (const CameraEntry& e : cams) f << e.image_file << " " << affine_to_str(e.world_to_cam) << "\n";
}

// ---------------------------------------------------------------------------------------------------- quaternions
void rotation_to_quaternion(const double* m, double* q) {  // Eigen 3.4 quaternionbase_assign_impl<Matrix3>
  auto M = [&](int r, int c) { return m[3 * r + c]; };
  double t = M(0, 0) + M(1, 1) + M(2, 2);
  double w, v[3];
  if (t > 0.0) {
    t = std::sqrt(t + 1.0);
    w = 0.5 * t;
    t = 0.5 / t;
    v[0] = (M(2, 1) - M(1, 2)) * t; v[1] = (M(0, 2) - M(2, 0)) * t; v[2] = (M(1, 0) - M(0, 1)) * t;
  } else {
    int i = 0;
    if (M(1, 1) > M(0, 0)) i = 1;
    if (M(2, 2) > M(i, i)) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    t = std::sqrt(M(i, i) - M(j, j) - M(k, k) + 1.0);
    v[i] = 0.5 * t;
    t = 0.5 / t;
    w = (M(k, j) - M(j, k)) * t;
    v[j] = (M(j, i) + M(i, j)) * t;
    v[k] = (M(k, i) + M(i, k)) * t;
  }
  q[0] = w; q[1] = v[0]; q[2] = v[1]; q[3] = v[2];
}

void quaternion_to_rotation(const double* q, double* R) {  // Eigen toRotationMatrix(): no normalisation
  const double w = q[0], x = q[1], y = q[2], z = q[3];
  const double tx = 2 * x, ty = 2 * y, tz = 2 * z;
  const double twx = tx * w, twy = ty * w, twz = tz * w, txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y, tyz = tz * y, tzz = tz * z;
  R[0] = 1 - (tyy + tzz); R[1] = txy - twz; R[2] = txz + twy;
  R[3] = txy + twz; R[4] = 1 - (txx + tzz); R[5] = tyz - twx;
  R[6] = txz - twy; R[7] = tyz + twx; R[8] = 1 - (txx + tyy);
}

}  // namespace mvba_host
