// oracle/functors.h -- TEST INFRASTRUCTURE ONLY (checker + CPU baseline). Never linked into the
// product library.  PARITY UNPINNED: the reference ships no tests, golden vectors or fixtures
// for this path, and Ceres/Eigen are absent so the reference cannot be
// built here.  This file is a dependency-free CPU restatement of the reference's cost
// functors, each function citing the reference file:line it follows.  Templated on the scalar so
// that the same code runs on double (values, Ceres-rule numeric differentiation), long double
// and on forward-mode dual numbers (the arbiter for the analytic Jacobians of the CUDA path).
#pragma once
#include <cmath>
#include <cstddef>

namespace oracle {

// ----------------------------------------------------------------------------------------
// Forward-mode dual number with a runtime number of active partials.
// ----------------------------------------------------------------------------------------
constexpr int kMaxDual = 168;  // 28 + 128 distortion + slack
struct Dual {
  double v;
  double d[kMaxDual];
  static int& n() { static thread_local int n_ = 0; return n_; }
  Dual() : v(0) { for (int i = 0; i < n(); i++) d[i] = 0; }
  Dual(double x) : v(x) { for (int i = 0; i < n(); i++) d[i] = 0; }  // NOLINT
  Dual(const Dual& o) : v(o.v) { for (int i = 0; i < n(); i++) d[i] = o.d[i]; }
  Dual& operator=(const Dual& o) { v = o.v; for (int i = 0; i < n(); i++) d[i] = o.d[i]; return *this; }
};
inline Dual operator+(const Dual& a, const Dual& b) { Dual r; r.v = a.v + b.v; for (int i = 0; i < Dual::n(); i++) r.d[i] = a.d[i] + b.d[i]; return r; }
inline Dual operator-(const Dual& a, const Dual& b) { Dual r; r.v = a.v - b.v; for (int i = 0; i < Dual::n(); i++) r.d[i] = a.d[i] - b.d[i]; return r; }
inline Dual operator-(const Dual& a) { Dual r; r.v = -a.v; for (int i = 0; i < Dual::n(); i++) r.d[i] = -a.d[i]; return r; }
inline Dual operator*(const Dual& a, const Dual& b) { Dual r; r.v = a.v * b.v; for (int i = 0; i < Dual::n(); i++) r.d[i] = a.d[i] * b.v + a.v * b.d[i]; return r; }
inline Dual operator/(const Dual& a, const Dual& b) { Dual r; double inv = 1.0 / b.v; r.v = a.v * inv; for (int i = 0; i < Dual::n(); i++) r.d[i] = (a.d[i] - r.v * b.d[i]) * inv; return r; }
inline Dual& operator+=(Dual& a, const Dual& b) { a = a + b; return a; }
inline Dual& operator-=(Dual& a, const Dual& b) { a = a - b; return a; }
inline Dual& operator*=(Dual& a, const Dual& b) { a = a * b; return a; }
inline Dual& operator/=(Dual& a, const Dual& b) { a = a / b; return a; }
inline bool operator<(const Dual& a, const Dual& b) { return a.v < b.v; }
inline bool operator>(const Dual& a, const Dual& b) { return a.v > b.v; }
inline bool operator<=(const Dual& a, const Dual& b) { return a.v <= b.v; }
inline bool operator>=(const Dual& a, const Dual& b) { return a.v >= b.v; }
inline bool operator==(const Dual& a, const Dual& b) { return a.v == b.v; }
inline Dual chain(const Dual& a, double f, double df) { Dual r; r.v = f; for (int i = 0; i < Dual::n(); i++) r.d[i] = df * a.d[i]; return r; }
inline Dual sqrt(const Dual& a) { double s = std::sqrt(a.v); return chain(a, s, 0.5 / s); }
inline Dual sin(const Dual& a) { return chain(a, std::sin(a.v), std::cos(a.v)); }
inline Dual cos(const Dual& a) { return chain(a, std::cos(a.v), -std::sin(a.v)); }
inline Dual tan(const Dual& a) { double t = std::tan(a.v); return chain(a, t, 1.0 + t * t); }
inline Dual atan(const Dual& a) { return chain(a, std::atan(a.v), 1.0 / (1.0 + a.v * a.v)); }
inline Dual acos(const Dual& a) { return chain(a, std::acos(a.v), -1.0 / std::sqrt(1.0 - a.v * a.v)); }
inline Dual abs(const Dual& a) { return a.v < 0 ? -a : a; }
inline double value(const Dual& a) { return a.v; }
inline double value(double a) { return a; }
inline double value(long double a) { return (double)a; }
using std::abs; using std::acos; using std::atan; using std::cos; using std::sin; using std::sqrt; using std::tan;

template <class T> struct Eps;  // machine epsilon of the scalar the slerp threshold is taken in
template <> struct Eps<double> { static double get() { return 2.220446049250313e-16; } };
template <> struct Eps<Dual> { static double get() { return 2.220446049250313e-16; } };
template <> struct Eps<long double> { static long double get() { return 2.220446049250313e-16L; } };  // keep double's branch

// ----------------------------------------------------------------------------------------
// Minimal fixed-size algebra reproducing the Eigen 3.4 semantics the reference relies on.
// ----------------------------------------------------------------------------------------
template <class T> struct Quat { T w, x, y, z; };
template <class T> struct Aff { T m[3][3]; T t[3]; };  // Eigen::Affine3d: linear() and translation()

// Eigen::Quaternion::normalize() as used by array_to_rigid_transform (dense_map_utils.cc:174-175).
template <class T> void quat_normalize(Quat<T>& q) {
  T n = sqrt(q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z);
  q.w = q.w / n; q.x = q.x / n; q.y = q.y / n; q.z = q.z / n;
}

// Eigen 3.4 QuaternionBase::toRotationMatrix (no re-normalisation).
template <class T> void quat_to_mat(const Quat<T>& q, T m[3][3]) {
  T tx = T(2.0) * q.x, ty = T(2.0) * q.y, tz = T(2.0) * q.z;
  T twx = tx * q.w, twy = ty * q.w, twz = tz * q.w;
  T txx = tx * q.x, txy = ty * q.x, txz = tz * q.x;
  T tyy = ty * q.y, tyz = tz * q.y, tzz = tz * q.z;
  m[0][0] = T(1.0) - (tyy + tzz); m[0][1] = txy - twz; m[0][2] = txz + twy;
  m[1][0] = txy + twz; m[1][1] = T(1.0) - (txx + tzz); m[1][2] = tyz - twx;
  m[2][0] = txz - twy; m[2][1] = tyz + twx; m[2][2] = T(1.0) - (txx + tyy);
}

// Eigen 3.4 quaternion-from-rotation-matrix (Quaternion(const Matrix3&)), used by
// rigid_transform_to_array (dense_map_utils.cc:152) and linearInterp (:317-318).
template <class T> Quat<T> mat_to_quat(const T m[3][3]) {
  Quat<T> q;
  T t = m[0][0] + m[1][1] + m[2][2];
  if (t > T(0.0)) {
    t = sqrt(t + T(1.0));
    q.w = T(0.5) * t;
    t = T(0.5) / t;
    q.x = (m[2][1] - m[1][2]) * t;
    q.y = (m[0][2] - m[2][0]) * t;
    q.z = (m[1][0] - m[0][1]) * t;
  } else {
    int i = 0;
    if (m[1][1] > m[0][0]) i = 1;
    if (m[2][2] > m[i][i]) i = 2;
    int j = (i + 1) % 3, k = (j + 1) % 3;
    t = sqrt(m[i][i] - m[j][j] - m[k][k] + T(1.0));
    T v[3];
    v[i] = T(0.5) * t;
    t = T(0.5) / t;
    q.w = (m[k][j] - m[j][k]) * t;
    v[j] = (m[j][i] + m[i][j]) * t;
    v[k] = (m[k][i] + m[i][k]) * t;
    q.x = v[0]; q.y = v[1]; q.z = v[2];
  }
  return q;
}

// Eigen 3.4 QuaternionBase::slerp: shortest arc through the sign of the dot product, linear
// blend when |dot| >= 1 - epsilon.
template <class T> Quat<T> slerp(T t, const Quat<T>& a, const Quat<T>& b) {
  T one = T(1.0) - T(Eps<T>::get());
  T d = a.w * b.w + a.x * b.x + a.y * b.y + a.z * b.z;
  T absD = abs(d);
  T s0, s1;
  if (absD >= one) {
    s0 = T(1.0) - t; s1 = t;
  } else {
    T theta = acos(absD);
    T sinTheta = sin(theta);
    s0 = sin((T(1.0) - t) * theta) / sinTheta;
    s1 = sin(t * theta) / sinTheta;
  }
  if (d < T(0.0)) s1 = -s1;
  Quat<T> r;
  r.w = s0 * a.w + s1 * b.w; r.x = s0 * a.x + s1 * b.x;
  r.y = s0 * a.y + s1 * b.y; r.z = s0 * a.z + s1 * b.z;
  return r;
}

// array_to_rigid_transform, dense_map_utils.cc:171-178.
template <class T> Aff<T> array_to_rigid(const T* a) {
  Aff<T> r;
  Quat<T> q{a[6], a[3], a[4], a[5]};
  quat_normalize(q);
  quat_to_mat(q, r.m);
  r.t[0] = a[0]; r.t[1] = a[1]; r.t[2] = a[2];
  return r;
}
// rigid_transform_to_array, dense_map_utils.cc:148-157.
template <class T> void rigid_to_array(const Aff<T>& a, T* arr) {
  for (int i = 0; i < 3; i++) arr[i] = a.t[i];
  Quat<T> q = mat_to_quat(a.m);
  arr[3] = q.x; arr[4] = q.y; arr[5] = q.z; arr[6] = q.w;
}
// array_to_affine_transform, dense_map_utils.cc:194-208: 12 values, row-major rows 0..2 of the 4x4.
template <class T> Aff<T> array_to_affine(const T* a) {
  Aff<T> r;
  for (int row = 0; row < 3; row++) {
    for (int col = 0; col < 3; col++) r.m[row][col] = a[4 * row + col];
    r.t[row] = a[4 * row + 3];
  }
  return r;
}
template <class T> Aff<T> aff_mul(const Aff<T>& A, const Aff<T>& B) {  // A * B
  Aff<T> r;
  for (int i = 0; i < 3; i++) {
    for (int j = 0; j < 3; j++) r.m[i][j] = A.m[i][0] * B.m[0][j] + A.m[i][1] * B.m[1][j] + A.m[i][2] * B.m[2][j];
    r.t[i] = A.m[i][0] * B.t[0] + A.m[i][1] * B.t[1] + A.m[i][2] * B.t[2] + A.t[i];
  }
  return r;
}
template <class T> void aff_apply(const Aff<T>& A, const T x[3], T y[3]) {
  T o[3];
  for (int i = 0; i < 3; i++) o[i] = A.m[i][0] * x[0] + A.m[i][1] * x[1] + A.m[i][2] * x[2] + A.t[i];
  for (int i = 0; i < 3; i++) y[i] = o[i];
}
// Eigen Transform<double,3,Affine>::inverse(): general 3x3 inverse of the linear part by
// cofactors, translation = -inv * t.  Used at rig_calibrator.cc:572 and :667.
template <class T> Aff<T> aff_inverse(const Aff<T>& A) {
  const T(*m)[3] = A.m;
  T c00 = m[1][1] * m[2][2] - m[1][2] * m[2][1];
  T c10 = m[1][2] * m[2][0] - m[1][0] * m[2][2];
  T c20 = m[1][0] * m[2][1] - m[1][1] * m[2][0];
  T det = m[0][0] * c00 + m[0][1] * c10 + m[0][2] * c20;
  T id = T(1.0) / det;
  Aff<T> r;
  r.m[0][0] = c00 * id; r.m[1][0] = c10 * id; r.m[2][0] = c20 * id;
  r.m[0][1] = (m[0][2] * m[2][1] - m[0][1] * m[2][2]) * id;
  r.m[1][1] = (m[0][0] * m[2][2] - m[0][2] * m[2][0]) * id;
  r.m[2][1] = (m[0][1] * m[2][0] - m[0][0] * m[2][1]) * id;
  r.m[0][2] = (m[0][1] * m[1][2] - m[0][2] * m[1][1]) * id;
  r.m[1][2] = (m[0][2] * m[1][0] - m[0][0] * m[1][2]) * id;
  r.m[2][2] = (m[0][0] * m[1][1] - m[0][1] * m[1][0]) * id;
  for (int i = 0; i < 3; i++) r.t[i] = -(r.m[i][0] * A.t[0] + r.m[i][1] * A.t[1] + r.m[i][2] * A.t[2]);
  return r;
}

// linearInterp, dense_map_utils.cc:315-329.
template <class T> Aff<T> linear_interp(T alpha, const Aff<T>& a0, const Aff<T>& a1) {
  Quat<T> r0 = mat_to_quat(a0.m), r1 = mat_to_quat(a1.m);
  Aff<T> r;
  for (int i = 0; i < 3; i++) r.t[i] = (T(1.0) - alpha) * a0.t[i] + alpha * a1.t[i];
  quat_to_mat(slerp(alpha, r0, r1), r.m);
  return r;
}

// calc_interp_world_to_ref, rig_calibrator.cc:322-353.  *ok = false replaces LOG(FATAL) (:346).
template <class T>
Aff<T> calc_interp_world_to_ref(const T* beg, const T* end, double beg_stamp, double end_stamp, T offset,
                                double cam_stamp, bool* ok) {
  Aff<T> b = array_to_rigid(beg), e = array_to_rigid(end);
  if (end_stamp == beg_stamp) return b;
  T alpha = (T(cam_stamp - beg_stamp) - offset) / T(end_stamp - beg_stamp);
  if (alpha < T(0.0) || alpha > T(1.0)) *ok = false;
  return linear_interp(alpha, b, e);
}

// calc_world_to_cam_trans, rig_calibrator.cc:362-390.
template <class T>
Aff<T> calc_world_to_cam_trans(const T* beg, const T* end, const T* ref_to_cam, double beg_stamp,
                               double end_stamp, T offset, double cam_stamp, bool* ok) {
  if (beg_stamp == end_stamp) return array_to_rigid(beg);
  Aff<T> r2c = array_to_rigid(ref_to_cam);
  Aff<T> w2r = calc_interp_world_to_ref(beg, end, beg_stamp, end_stamp, offset, cam_stamp, ok);
  return aff_mul(r2c, w2r);
}

// rpc_degree / num_dist_params, rpc_distortion.cc:43-49.
inline int rpc_degree(int num_dist) { return (int)std::round(std::sqrt(2.0 * num_dist + 5.0) / 2.0 - 1.5); }
inline int rpc_num_dist_params(int deg) { return 2 * (deg + 1) * (deg + 2) - 2; }

// compute_rpc, rpc_distortion.cc:116-160.  coeffs has n = num_dist_params(deg) entries.
template <class T> void compute_rpc(const T p[2], const T* coeffs, int n, T out[2]) {
  int deg = rpc_degree(n);
  T powx[16], powy[16];
  T vx = T(1.0), vy = T(1.0);
  for (int d = 0; d <= deg; d++) { powx[d] = vx; vx = vx * p[0]; powy[d] = vy; vy = vy * p[1]; }
  T vals[4] = {T(0.0), T(1.0), T(0.0), T(1.0)};
  int ci = 0;
  for (int count = 0; count < 4; count++) {
    int start = (count == 1 || count == 3) ? 1 : 0;
    for (int d = start; d <= deg; d++)
      for (int i = 0; i <= d; i++) { vals[count] = vals[count] + coeffs[ci] * powx[d - i] * powy[i]; ci++; }
  }
  out[0] = vals[0] / vals[1];
  out[1] = vals[2] / vals[3];
}

// Per-sensor constants the functor copies from camera::CameraParameters (rig_calibrator.cc:460).
struct SensorConst { double half[2]; int nd; };

// CameraParameters::Convert<UNDISTORTED_C, DISTORTED> = DistortCentered + distorted_half_size
// (camera_params.cc:412-415, 260-314), with SetFocalLength((f,f)), SetOpticalOffset,
// SetDistortion (:149-208) applied first as the functor does (rig_calibrator.cc:460-467).
template <class T>
void distort_undist_c_to_dist(const SensorConst& s, T f, const T c[2], const T* dist, const T u[2], T out[2]) {
  T dc[2];
  T h[2] = {T(s.half[0]), T(s.half[1])};
  if (s.nd == 0) {
    dc[0] = u[0] + c[0] - h[0]; dc[1] = u[1] + c[1] - h[1];
  } else if (s.nd == 1) {
    T pre1 = T(1.0) / dist[0];                   // distortion_precalc1_, camera_params.cc:186
    T pre2 = T(2.0) * tan(dist[0] / T(2.0));     // distortion_precalc2_, :188
    T nx = u[0] / f, ny = u[1] / f;
    T ru = sqrt(nx * nx + ny * ny);
    T rd = atan(ru * pre2) * pre1;
    T conv = T(1.0);
    if (ru > T(1e-5)) conv = rd / ru;
    dc[0] = (c[0] - h[0]) + conv * (nx * f);
    dc[1] = (c[1] - h[1]) + conv * (ny * f);
  } else if (s.nd == 4 || s.nd == 5) {
    T k1 = dist[0], k2 = dist[1], p1 = dist[2], p2 = dist[3];
    T k3 = T(0.0);
    if (s.nd == 5) k3 = dist[4];
    T nx = u[0] / f, ny = u[1] / f;
    T r2 = nx * nx + ny * ny;
    T radial = T(1.0) + k1 * r2 + k2 * r2 * r2 + k3 * r2 * r2 * r2;
    T dx = radial * nx, dy = radial * ny;
    dx = dx + (T(2.0) * p1 * nx * ny + p2 * (r2 + T(2.0) * nx * nx));
    dy = dy + (p1 * (r2 + T(2.0) * ny * ny) + T(2.0) * p2 * nx * ny);
    dc[0] = dx * f + (c[0] - h[0]);
    dc[1] = dy * f + (c[1] - h[1]);
  } else {
    // RPC: first half of the block is the distortion model (rpc_distortion.cc:272-293, 360-362).
    compute_rpc(u, dist, s.nd / 2, dc);
  }
  out[0] = dc[0] + h[0]; out[1] = dc[1] + h[1];
}

// BracketedCamError::operator(), rig_calibrator.cc:447-482.
// p: beg7 end7 ref_to_cam7 xyz3 offset1 focal1 center2 dist[nd]
template <class T>
bool bracketed_cam_error(const T* const* p, const double meas[2], double beg_stamp, double end_stamp,
                         double cam_stamp, const SensorConst& s, T* r) {
  bool ok = true;
  Aff<T> w2c = calc_world_to_cam_trans(p[0], p[1], p[2], beg_stamp, end_stamp, p[4][0], cam_stamp, &ok);
  T X[3] = {p[3][0], p[3][1], p[3][2]};
  aff_apply(w2c, X, X);
  T f = p[5][0];
  T u[2] = {f * (X[0] / X[2]), f * (X[1] / X[2])};
  T d[2];
  distort_undist_c_to_dist(s, f, p[6], p[7], u, d);
  r[0] = d[0] - T(meas[0]);
  r[1] = d[1] - T(meas[1]);
  return ok;
}

// Shared tail of BracketedDepthError / BracketedDepthMeshError (rig_calibrator.cc:557-572, 652-667).
template <class T>
void depth_to_world(const Aff<T>& w2c, const T* d2i, int nd2i, T scale, const double meas[3], T M[3]) {
  Aff<T> D = (nd2i == 12) ? array_to_affine(d2i) : array_to_rigid(d2i);
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) D.m[i][j] = D.m[i][j] * scale;
  T m[3] = {T(meas[0]), T(meas[1]), T(meas[2])};
  aff_apply(D, m, M);
  Aff<T> inv = aff_inverse(w2c);
  aff_apply(inv, M, M);
}

// BracketedDepthError::operator(), rig_calibrator.cc:547-583.
// p: beg7 end7 ref_to_cam7 depth_to_image[7|12] scale1 xyz3 offset1
template <class T>
bool bracketed_depth_error(const T* const* p, int nd2i, double weight, const double meas[3], double beg_stamp,
                           double end_stamp, double cam_stamp, T* r) {
  bool ok = true;
  Aff<T> w2c = calc_world_to_cam_trans(p[0], p[1], p[2], beg_stamp, end_stamp, p[6][0], cam_stamp, &ok);
  T M[3];
  depth_to_world(w2c, p[3], nd2i, p[4][0], meas, M);
  for (int i = 0; i < 3; i++) r[i] = T(weight) * (p[5][i] - M[i]);
  return ok;
}

// BracketedDepthMeshError::operator(), rig_calibrator.cc:642-675.
// p: beg7 end7 ref_to_cam7 depth_to_image[7|12] scale1 offset1
template <class T>
bool bracketed_depth_mesh_error(const T* const* p, int nd2i, double weight, const double meas[3],
                                const double mesh[3], double beg_stamp, double end_stamp, double cam_stamp,
                                T* r) {
  bool ok = true;
  Aff<T> w2c = calc_world_to_cam_trans(p[0], p[1], p[2], beg_stamp, end_stamp, p[5][0], cam_stamp, &ok);
  T M[3];
  depth_to_world(w2c, p[3], nd2i, p[4][0], meas, M);
  for (int i = 0; i < 3; i++) r[i] = T(weight) * (T(mesh[i]) - M[i]);
  return ok;
}

// XYZError::operator(), rig_calibrator.cc:721-727.
template <class T> bool xyz_error(const T* const* p, double weight, const double ref[3], T* r) {
  for (int i = 0; i < 3; i++) r[i] = T(weight) * (p[0][i] - T(ref[i]));
  return true;
}

}  // namespace oracle
