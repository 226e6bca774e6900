// camera_params.h -- dependency-free mirror of camera::CameraParameters (reference camera_model/camera_params.h:59-214,
// camera_model/camera_params.cc): same class name, namespace, member names, frame enum and conversion set, with a small
// 2-vector type standing in for Eigen::Vector2d / Vector2i.  The device code consumes the POD snapshot
// `pod()` of it; mvba_set_sensor / mvba_params carry the same numbers.
//   frames         camera_params.h:42-57
//   conversions    camera_params.cc:374-418 (the ten declared pairs)
//   intrinsics     camera_params.cc:420-451
//   DistortCentered / UndistortCentered  camera_params.cc:260-355 (radtan undistortion: the five fixed-point
//                  iterations of cv::undistortPoints, which is what the reference calls)
//   RPC            rpc_distortion.cc:116-160, 272-293 (distortion half, then undistortion half)
//   calibration-file constructor  camera_params.cc:50-104 (cv::FileStorage there; here a reader of the same YAML files:
//                  scalars and !!opencv-matrix nodes with dt: d / f / i, which is what OpenCV writes for these keys)
//   GenerateRemapMaps  camera_params.cc:361-371 (cv::Mat CV_32FC2 there; here rows x cols interleaved float pairs)
//   updateRpcUndistortion  camera_params.cc:214-254, defined in rpc_pairs.h (the fit is SURVEY.md 8(f) N4, on the device)
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <fstream>
#include <map>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace camera {

template <class T> struct Vec2 {
  T v[2];
  Vec2() : v{0, 0} {}
  Vec2(T a, T b) : v{a, b} {}
  T& operator[](int i) { return v[i]; }
  const T& operator[](int i) const { return v[i]; }
  Vec2 operator+(const Vec2& o) const { return Vec2(v[0] + o.v[0], v[1] + o.v[1]); }
  Vec2 operator-(const Vec2& o) const { return Vec2(v[0] - o.v[0], v[1] - o.v[1]); }
  bool operator==(const Vec2& o) const { return v[0] == o.v[0] && v[1] == o.v[1]; }
};
using Vector2d = Vec2<double>;
using Vector2i = Vec2<int>;
struct Matrix3d { double m[3][3]; double operator()(int r, int c) const { return m[r][c]; } };

enum { RAW, DISTORTED, DISTORTED_C, UNDISTORTED, UNDISTORTED_C };

// rpc_distortion.cc:43-57
inline int rpc_degree(int num_dist) { return (int)std::round(std::sqrt(2.0 * num_dist + 5.0) / 2.0 - 1.5); }
inline int rpc_num_dist_params(int deg) { return 2 * (deg + 1) * (deg + 2) - 2; }

// compute_rpc, rpc_distortion.cc:116-160: num_x | den_x | num_y | den_y over monomials x^(d-i) y^i; denominators start at
// degree 1 with an implicit leading 1.
inline Vector2d compute_rpc(const Vector2d& p, const double* coeffs, int n) {
  const int deg = rpc_degree(n);
  if (rpc_num_dist_params(deg) != n) throw std::runtime_error("Incorrect number of RPC coefficients.");
  double powx[16], powy[16];
  powx[0] = 1; powy[0] = 1;
  for (int d = 1; d <= deg; d++) { powx[d] = powx[d - 1] * p[0]; powy[d] = powy[d - 1] * p[1]; }
  double val[4] = {0, 1, 0, 1};
  int ci = 0;
  for (int cnt = 0; cnt < 4; cnt++)
    for (int d = (cnt & 1); d <= deg; d++)
      for (int i = 0; i <= d; i++) val[cnt] += coeffs[ci++] * powx[d - i] * powy[i];
  return Vector2d(val[0] / val[1], val[2] / val[3]);
}

// What the kernels read for one sensor.
struct CameraPod { double f, c[2], half[2]; int nd; double dist[128]; };

class CameraParameters {
 public:
  CameraParameters() = delete;
  CameraParameters(Vector2i const& image_size, Vector2d const& focal_length, Vector2d const& optical_center,
                   std::vector<double> const& distortion = std::vector<double>()) {
    SetDistortedSize(image_size);
    SetDistortedCropSize(image_size);
    SetUndistortedSize(image_size);
    focal_length_ = focal_length;
    optical_offset_ = optical_center;
    crop_offset_ = Vector2i(0, 0);
    SetDistortion(distortion);
  }

  // camera_params.cc:50-104: an OpenCV calibration file (YAML).  Keys with their fall-backs, in the reference's order: width / height
  // (image_width / image_height), intrinsic_matrix (camera_matrix), distortion (distortion_coefficients), undistorted_width /
  // undistorted_height (default: the distorted size), crop_x / crop_y (default 0).
  explicit CameraParameters(std::string const& calibration_file, std::string const& base_dir = "") {
    std::ifstream f(calibration_file);
    if (!f.good() && base_dir != "") f.open(base_dir + "/" + calibration_file);   // it may be relative to the map file's directory
    if (!f.good()) throw std::runtime_error("Calibration file does not exist: " + calibration_file);
    std::map<std::string, double> scalars;
    std::map<std::string, std::vector<double>> matrices;
    std::map<std::string, std::pair<int, int>> shapes;
    read_opencv_yaml(f, scalars, matrices, shapes);
    auto scalar = [&](const char* k) { auto it = scalars.find(k); return it == scalars.end() ? 0 : (int)it->second; };
    Vector2i size(scalar("width"), scalar("height"));
    if (size[0] == 0 || size[1] == 0) size = Vector2i(scalar("image_width"), scalar("image_height"));
    SetDistortedSize(size);
    SetDistortedCropSize(size);
    const char* km = matrices.count("intrinsic_matrix") && !matrices["intrinsic_matrix"].empty() ? "intrinsic_matrix" : "camera_matrix";
    if (!matrices.count(km) || shapes[km].first < 2 || shapes[km].second < 3) throw std::runtime_error("No camera matrix in calibration file: " + calibration_file);
    const std::vector<double>& K = matrices[km];
    const int kc = shapes[km].second;
    optical_offset_ = Vector2d(K[2], K[kc + 2]);
    focal_length_ = Vector2d(K[0], K[kc + 1]);
    const char* kd = matrices.count("distortion") && !matrices["distortion"].empty() ? "distortion" : "distortion_coefficients";
    SetDistortion(matrices.count(kd) ? matrices[kd] : std::vector<double>());
    size = Vector2i(scalar("undistorted_width"), scalar("undistorted_height"));
    if (!(size[0] > 0 && size[1] > 0)) size = distorted_image_size_;
    SetUndistortedSize(size);
    crop_offset_ = Vector2i(scalar("crop_x"), scalar("crop_y"));
  }

  // camera_params.cc:361-371: for every pixel of the (scaled) undistorted image the distorted-image position it samples,
  // the map cv::remap takes.  remap_map holds rows x cols pairs (x, y) as float32, row-major, like the reference's CV_32FC2 matrix;
  // rows = int(scale * undistorted height), cols = int(scale * undistorted width) (cv::Mat::create truncates the same way).
  void GenerateRemapMaps(std::vector<float>* remap_map, int* rows, int* cols, double scale = 1.0) const {
    *rows = (int)(scale * undistorted_image_size_[1]); *cols = (int)(scale * undistorted_image_size_[0]);
    remap_map->assign(2 * (size_t)*rows * (size_t)*cols, 0.0f);
    Vector2d distorted;
    for (int y = 0; y < *rows; y++)
      for (int x = 0; x < *cols; x++) {
        Convert<UNDISTORTED, DISTORTED>(Vector2d(x / scale, y / scale), &distorted);
        (*remap_map)[2 * ((size_t)y * *cols + x)] = (float)(scale * distorted[0]);
        (*remap_map)[2 * ((size_t)y * *cols + x) + 1] = (float)(scale * distorted[1]);
      }
  }

  template <int SRC, int DEST> void Convert(Vector2d const& input, Vector2d* output) const {
    if (SRC == RAW && DEST == DISTORTED) *output = input - Vector2d(crop_offset_[0], crop_offset_[1]);
    else if (SRC == DISTORTED && DEST == RAW) *output = input + Vector2d(crop_offset_[0], crop_offset_[1]);
    else if (SRC == UNDISTORTED_C && DEST == DISTORTED_C) DistortCentered(input, output);
    else if (SRC == DISTORTED_C && DEST == UNDISTORTED_C) UndistortCentered(input, output);
    else if (SRC == UNDISTORTED && DEST == UNDISTORTED_C) *output = input - undistorted_half_size_;
    else if (SRC == UNDISTORTED_C && DEST == UNDISTORTED) *output = input + undistorted_half_size_;
    else if (SRC == DISTORTED && DEST == UNDISTORTED) { UndistortCentered(input - distorted_half_size_, output); *output = *output + undistorted_half_size_; }
    else if (SRC == UNDISTORTED && DEST == DISTORTED) { Vector2d c; DistortCentered(input - undistorted_half_size_, &c); *output = c + distorted_half_size_; }
    else if (SRC == DISTORTED && DEST == UNDISTORTED_C) UndistortCentered(input - distorted_half_size_, output);
    else if (SRC == UNDISTORTED_C && DEST == DISTORTED) { DistortCentered(input, output); *output = *output + distorted_half_size_; }
    else throw std::runtime_error("Please use the explicitly specified conversions by using the correct enum.");
  }

  template <int FRAME> Matrix3d GetIntrinsicMatrix() const {
    Matrix3d k = {{{focal_length_[0], 0, 0}, {0, focal_length_[1], 0}, {0, 0, 1}}};
    Vector2d c(0, 0);
    if (FRAME == RAW) c = optical_offset_ + Vector2d(crop_offset_[0], crop_offset_[1]);
    else if (FRAME == DISTORTED) c = optical_offset_;
    else if (FRAME == DISTORTED_C) c = optical_offset_ - distorted_half_size_;
    else if (FRAME == UNDISTORTED) c = undistorted_half_size_;
    k.m[0][2] = c[0]; k.m[1][2] = c[1];
    return k;
  }

  void SetDistortedSize(Vector2i const& s) { distorted_image_size_ = s; distorted_half_size_ = Vector2d(s[0] / 2.0, s[1] / 2.0); }
  const Vector2i& GetDistortedSize() const { return distorted_image_size_; }
  const Vector2d& GetDistortedHalfSize() const { return distorted_half_size_; }
  void SetDistortedCropSize(Vector2i const& s) { distorted_crop_size_ = s; }
  const Vector2i& GetDistortedCropSize() const { return distorted_crop_size_; }
  void SetUndistortedSize(Vector2i const& s) { undistorted_image_size_ = s; undistorted_half_size_ = Vector2d(s[0] / 2.0, s[1] / 2.0); }
  const Vector2i& GetUndistortedSize() const { return undistorted_image_size_; }
  const Vector2d& GetUndistortedHalfSize() const { return undistorted_half_size_; }
  void SetCropOffset(Vector2i const& c) { crop_offset_ = c; }
  const Vector2i& GetCropOffset() const { return crop_offset_; }
  void SetOpticalOffset(Vector2d const& o) { optical_offset_ = o; }
  const Vector2d& GetOpticalOffset() const { return optical_offset_; }
  void SetFocalLength(Vector2d const& f) { focal_length_ = f; }
  double GetFocalLength() const { return (focal_length_[0] + focal_length_[1]) / 2.0; }  // the MEAN of fx, fy
  const Vector2d& GetFocalVector() const { return focal_length_; }

  // 0: none, 1: FOV, 4 | 5: Tsai / radtan, more: RPC (distortion half + undistortion half)
  void SetDistortion(std::vector<double> const& d) {
    distortion_coeffs_ = d;
    distortion_precalc1_ = distortion_precalc2_ = distortion_precalc3_ = 0;
    can_undistort_ = true;
    const size_t n = d.size();
    if (n == 1) { distortion_precalc1_ = 1 / d[0]; distortion_precalc2_ = 2 * std::tan(d[0] / 2); }
    else if (n > 5) {
      if (n % 2 != 0 || rpc_num_dist_params(rpc_degree((int)n / 2)) != (int)n / 2)
        throw std::runtime_error("Recieved irregular distortion vector size. Size = " + std::to_string(n));
    } else if (n != 0 && n != 4 && n != 5) {
      throw std::runtime_error("Recieved irregular distortion vector size. Size = " + std::to_string(n));
    }
  }
  const std::vector<double>& GetDistortion() const { return distortion_coeffs_; }
  // camera_params.cc:214-254: re-fit the undistortion half of an RPC model from its distortion half (fitRpcUndist,
  // rpc_distortion.cc:658-719: 400 x 400 samples, 100 LM iterations per degree, parameter tolerance 1e-12).  The fit itself
  // runs on the device: the entry point mvba_rpc_fit (include/mvba.h) is registered once with SetRpcFitter, so that this
  // header keeps no link-time dependency on libmvba.so.  Defined in rpc_pairs.h (needs the sample-pair generator).
  typedef int (*RpcFitter)(int32_t device, int64_t n, const double* in_xy, const double* out_xy, int32_t rpc_degree, int32_t num_iterations,
                           double parameter_tolerance, int32_t jacobian_mode, double* coeffs, double stats[4]);
  static RpcFitter& rpc_fitter() { static RpcFitter f = nullptr; return f; }
  static void SetRpcFitter(RpcFitter f) { rpc_fitter() = f; }
  inline void updateRpcUndistortion(int num_threads);
  // RPCLensDistortion::set_can_undistort (rig_calibrator.cc:1601-1605): undistortion is refused until the model is re-fitted
  void set_can_undistort(bool v) { can_undistort_ = v; }

  friend bool operator==(CameraParameters const& A, CameraParameters const& B) {
    return A.crop_offset_ == B.crop_offset_ && A.distorted_image_size_ == B.distorted_image_size_ &&
           A.distorted_crop_size_ == B.distorted_crop_size_ && A.undistorted_image_size_ == B.undistorted_image_size_ &&
           A.distorted_half_size_ == B.distorted_half_size_ && A.focal_length_ == B.focal_length_ &&
           A.optical_offset_ == B.optical_offset_ && A.distortion_coeffs_ == B.distortion_coeffs_ &&
           A.distortion_precalc1_ == B.distortion_precalc1_ && A.distortion_precalc2_ == B.distortion_precalc2_ &&
           A.distortion_precalc3_ == B.distortion_precalc3_;
  }

  CameraPod pod() const {
    CameraPod p{};
    p.f = GetFocalLength(); p.c[0] = optical_offset_[0]; p.c[1] = optical_offset_[1];
    p.half[0] = distorted_half_size_[0]; p.half[1] = distorted_half_size_[1];
    p.nd = (int)distortion_coeffs_.size();
    if (p.nd > 128) throw std::runtime_error("distortion block longer than MVBA_MAX_DISTORTION");
    for (int i = 0; i < p.nd; i++) p.dist[i] = distortion_coeffs_[i];
    return p;
  }

 private:
  // The subset of OpenCV's YAML persistence these files use: "key: number" and
  //   key: !!opencv-matrix / rows: R / cols: C / dt: d / data: [ v, v, ... ]  (the data list may span lines).
  static void read_opencv_yaml(std::istream& in, std::map<std::string, double>& scalars, std::map<std::string, std::vector<double>>& matrices,
                               std::map<std::string, std::pair<int, int>>& shapes) {
    std::string line, current;
    bool in_data = false;
    auto numbers = [](std::string text, std::vector<double>& out) {
      for (char& c : text) if (c == ',' || c == '[' || c == ']') c = ' ';
      std::istringstream ss(text);
      std::string tok;
      while (ss >> tok) { char* end = nullptr; const double v = std::strtod(tok.c_str(), &end); if (end != tok.c_str()) out.push_back(v); }
    };
    while (std::getline(in, line)) {
      if (!line.empty() && line.back() == '\r') line.pop_back();
      if (line.empty() || line[0] == '%' || line.rfind("---", 0) == 0) continue;
      if (in_data) { numbers(line, matrices[current]); if (line.find(']') != std::string::npos) in_data = false; continue; }
      const size_t colon = line.find(':');
      if (colon == std::string::npos) continue;
      const size_t k0 = line.find_first_not_of(" \t");
      std::string key = line.substr(k0, colon - k0), val = line.substr(colon + 1);
      const size_t v0 = val.find_first_not_of(" \t");
      val = v0 == std::string::npos ? "" : val.substr(v0);
      if (k0 == 0 && val.rfind("!!opencv-matrix", 0) == 0) { current = key; matrices[current].clear(); shapes[current] = {0, 0}; continue; }
      if (k0 > 0 && !current.empty()) {   // a field of the current matrix node
        if (key == "rows") shapes[current].first = std::atoi(val.c_str());
        else if (key == "cols") shapes[current].second = std::atoi(val.c_str());
        else if (key == "data") { numbers(val, matrices[current]); in_data = val.find(']') == std::string::npos; }
        continue;
      }
      current.clear();
      char* end = nullptr;
      const double v = std::strtod(val.c_str(), &end);
      if (end != val.c_str()) scalars[key] = v;
    }
  }

  void DistortCentered(Vector2d const& u, Vector2d* out) const {
    const size_t n = distortion_coeffs_.size();
    const Vector2d shift = optical_offset_ - distorted_half_size_;
    if (n == 0) { *out = u + shift; return; }
    if (n > 5) { *out = compute_rpc(u, distortion_coeffs_.data(), (int)n / 2); return; }
    const double nx = u[0] / focal_length_[0], ny = u[1] / focal_length_[1];
    if (n == 1) {
      const double ru = std::sqrt(nx * nx + ny * ny);
      const double rd = std::atan(ru * distortion_precalc2_) * distortion_precalc1_;
      const double conv = ru > 1e-5 ? rd / ru : 1;
      *out = shift + Vector2d(conv * (nx * focal_length_[0]), conv * (ny * focal_length_[1]));
      return;
    }
    const double k1 = distortion_coeffs_[0], k2 = distortion_coeffs_[1], p1 = distortion_coeffs_[2], p2 = distortion_coeffs_[3];
    const double k3 = n == 5 ? distortion_coeffs_[4] : 0;
    const double r2 = nx * nx + ny * ny;
    const double radial = 1 + k1 * r2 + k2 * r2 * r2 + k3 * r2 * r2 * r2;
    const double dx = radial * nx + (2 * p1 * nx * ny + p2 * (r2 + 2 * nx * nx));
    const double dy = radial * ny + (p1 * (r2 + 2 * ny * ny) + 2 * p2 * nx * ny);
    *out = Vector2d(dx * focal_length_[0], dy * focal_length_[1]) + shift;
  }

  void UndistortCentered(Vector2d const& d, Vector2d* out) const {
    const size_t n = distortion_coeffs_.size();
    const Vector2d shift = optical_offset_ - distorted_half_size_;
    if (n == 0) { *out = d - shift; return; }
    if (n > 5) {
      if (!can_undistort_) throw std::runtime_error("The RPC model is not ready for undistortion");
      *out = compute_rpc(d, distortion_coeffs_.data() + n / 2, (int)n / 2);
      return;
    }
    if (n == 1) {
      const double nx = (d[0] - shift[0]) / focal_length_[0], ny = (d[1] - shift[1]) / focal_length_[1];
      const double rd = std::sqrt(nx * nx + ny * ny);
      const double ru = std::tan(rd * distortion_coeffs_[0]) / distortion_precalc2_;
      const double conv = rd > 1e-5 ? ru / rd : 1.0;
      *out = Vector2d(conv * (nx * focal_length_[0]), conv * (ny * focal_length_[1]));
      return;
    }
    // cv::undistortPoints(src = d + half, K_dist, dist, R = I, P = K_undist) - undistorted half size
    const double k1 = distortion_coeffs_[0], k2 = distortion_coeffs_[1], p1 = distortion_coeffs_[2], p2 = distortion_coeffs_[3];
    const double k3 = n == 5 ? distortion_coeffs_[4] : 0;
    const double x0 = ((d[0] + distorted_half_size_[0]) - optical_offset_[0]) / focal_length_[0];
    const double y0 = ((d[1] + distorted_half_size_[1]) - optical_offset_[1]) / focal_length_[1];
    double x = x0, y = y0;
    for (int j = 0; j < 5; j++) {
      const double r2 = x * x + y * y;
      const double icdist = 1.0 / (1.0 + ((k3 * r2 + k2) * r2 + k1) * r2);
      const double ddx = 2 * p1 * x * y + p2 * (r2 + 2 * x * x), ddy = p1 * (r2 + 2 * y * y) + 2 * p2 * x * y;
      x = (x0 - ddx) * icdist; y = (y0 - ddy) * icdist;
    }
    *out = Vector2d(x * focal_length_[0], y * focal_length_[1]);
  }

  Vector2i crop_offset_, distorted_image_size_, distorted_crop_size_, undistorted_image_size_;
  Vector2d distorted_half_size_, undistorted_half_size_, focal_length_, optical_offset_;
  std::vector<double> distortion_coeffs_;
  double distortion_precalc1_ = 0, distortion_precalc2_ = 0, distortion_precalc3_ = 0;
  bool can_undistort_ = true;
};

}  // namespace camera
