// camera_shim.cc -- C test shims over camera::CameraParameters (tests/test_camera_params.py).
#include <cstring>
#include <string>

#include "camera_params.h"
#include "rpc_pairs.h"
#include "pass_utils.h"
#include "pose_math.h"
#include "bracket.h"
#include "residual_stats.h"
#include "cli_flags.h"
#include "ply_io.h"

static thread_local std::string g_cam_err;

template <int S, int D> static void conv_all(const camera::CameraParameters& cp, int n, const double* in, double* out) {
  for (int i = 0; i < n; i++) {
    camera::Vector2d o;
    cp.Convert<S, D>(camera::Vector2d(in[2 * i], in[2 * i + 1]), &o);
    out[2 * i] = o[0]; out[2 * i + 1] = o[1];
  }
}

extern "C" {
const char* mvh_camera_last_error() { return g_cam_err.c_str(); }
// frames: 0 RAW, 1 DISTORTED, 2 DISTORTED_C, 3 UNDISTORTED, 4 UNDISTORTED_C
int mvh_camera_convert(const int* size, const double* focal, const double* center, const double* dist, int nd, const int* undist_size,
                       const int* crop_offset, int src, int dst, int n, const double* in, double* out, double* k9, double* mean_focal) {
  try {
    using namespace camera;
    CameraParameters cp(Vector2i(size[0], size[1]), Vector2d(focal[0], focal[1]), Vector2d(center[0], center[1]),
                        std::vector<double>(dist, dist + nd));
    if (undist_size) cp.SetUndistortedSize(Vector2i(undist_size[0], undist_size[1]));
    if (crop_offset) cp.SetCropOffset(Vector2i(crop_offset[0], crop_offset[1]));
    const int key = 10 * src + dst;
    switch (key) {
      case 10 * RAW + DISTORTED: conv_all<RAW, DISTORTED>(cp, n, in, out); break;
      case 10 * DISTORTED + RAW: conv_all<DISTORTED, RAW>(cp, n, in, out); break;
      case 10 * UNDISTORTED_C + DISTORTED_C: conv_all<UNDISTORTED_C, DISTORTED_C>(cp, n, in, out); break;
      case 10 * DISTORTED_C + UNDISTORTED_C: conv_all<DISTORTED_C, UNDISTORTED_C>(cp, n, in, out); break;
      case 10 * UNDISTORTED + UNDISTORTED_C: conv_all<UNDISTORTED, UNDISTORTED_C>(cp, n, in, out); break;
      case 10 * UNDISTORTED_C + UNDISTORTED: conv_all<UNDISTORTED_C, UNDISTORTED>(cp, n, in, out); break;
      case 10 * DISTORTED + UNDISTORTED: conv_all<DISTORTED, UNDISTORTED>(cp, n, in, out); break;
      case 10 * UNDISTORTED + DISTORTED: conv_all<UNDISTORTED, DISTORTED>(cp, n, in, out); break;
      case 10 * DISTORTED + UNDISTORTED_C: conv_all<DISTORTED, UNDISTORTED_C>(cp, n, in, out); break;
      case 10 * UNDISTORTED_C + DISTORTED: conv_all<UNDISTORTED_C, DISTORTED>(cp, n, in, out); break;
      default: throw std::runtime_error("Please use the explicitly specified conversions by using the correct enum.");
    }
    if (k9) {
      Matrix3d k = src == RAW ? cp.GetIntrinsicMatrix<RAW>() : src == DISTORTED ? cp.GetIntrinsicMatrix<DISTORTED>()
                 : src == DISTORTED_C ? cp.GetIntrinsicMatrix<DISTORTED_C>() : src == UNDISTORTED ? cp.GetIntrinsicMatrix<UNDISTORTED>()
                 : cp.GetIntrinsicMatrix<UNDISTORTED_C>();
      for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) k9[3 * r + c] = k(r, c);
    }
    if (mean_focal) *mean_focal = cp.GetFocalLength();
    return 0;
  } catch (const std::exception& e) { g_cam_err = e.what(); return -1; }
}
// genUndistDistPairs: returns the number of pairs kept (<= num_samples^2), -1 on error; the outputs hold 2 doubles per pair
int mvh_gen_undist_dist_pairs(const int* size, const double* focal, const double* center, const double* dist, int nd, const int* undist_size,
                              const int* crop_size, int num_samples, double* undist_centered, double* dist_centered) {
  try {
    using namespace camera;
    CameraParameters cp(Vector2i(size[0], size[1]), Vector2d(focal[0], focal[1]), Vector2d(center[0], center[1]),
                        std::vector<double>(dist, dist + nd));
    if (undist_size) cp.SetUndistortedSize(Vector2i(undist_size[0], undist_size[1]));
    if (crop_size) cp.SetDistortedCropSize(Vector2i(crop_size[0], crop_size[1]));
    std::vector<double> u, d;
    mvba_host::gen_undist_dist_pairs(num_samples, cp, &u, &d);
    for (size_t i = 0; i < u.size(); i++) { undist_centered[i] = u[i]; dist_centered[i] = d[i]; }
    return (int)(u.size() / 2);
  } catch (const std::exception& e) { g_cam_err = e.what(); return -1; }
}
// depthValue for n pixels of one cloud: has[i] / xyz[3 i]; -1 when a pixel is out of range (the reference aborts)
int mvh_depth_values(const float* cloud, int rows, int cols, int n, const double* dist_ip, unsigned char* has, double* xyz) {
  try {
    for (int i = 0; i < n; i++) has[i] = mvba_host::depth_value(cloud, rows, cols, dist_ip + 2 * i, xyz + 3 * i) ? 1 : 0;
    return 0;
  } catch (const std::exception& e) { g_cam_err = e.what(); return -1; }
}
// compact_observations: returns the number of inliers; outputs sized for n
long long mvh_compact_observations(long long n, const int* pid, const int* cid, const float* pix, const unsigned char* inlier, int* opid, int* ocid,
                                   float* opix, long long* source) {
  mvba_host::Observations o = mvba_host::compact_observations(n, pid, cid, pix, nullptr, nullptr, nullptr, nullptr, inlier);
  for (size_t i = 0; i < o.pid.size(); i++) { opid[i] = o.pid[i]; ocid[i] = o.cid[i]; opix[2 * i] = o.pix[2 * i]; opix[2 * i + 1] = o.pix[2 * i + 1]; source[i] = o.source[i]; }
  return (long long)o.pid.size();
}
// world_to_cam of every image (12 doubles each: R row-major, t); -1 when an image falls outside its bracket
int mvh_world_to_cam_all(int num_cameras, const int* cam_sensor, const int* beg, const int* end, const double* cam_stamp, const double* ref_stamps,
                         const double* world_to_ref, const double* ref_to_cam, const double* offsets, double* out12) {
  try {
    std::vector<mvba_host::Affine> w = mvba_host::world_to_cam_all(num_cameras, cam_sensor, beg, end, cam_stamp, ref_stamps, world_to_ref, ref_to_cam, offsets);
    for (int c = 0; c < num_cameras; c++) for (int k = 0; k < 12; k++) out12[12 * c + k] = w[c][k];
    return 0;
  } catch (const std::exception& e) { g_cam_err = e.what(); return -1; }
}
// rigid_transform_to_array(array_to_rigid_transform(a)): the round trip the copy-back performs
void mvh_rigid_roundtrip(const double* a7, double* out7) { mvba_host::rigid_to_array(mvba_host::array_to_rigid(a7), out7); }
// lookup_images_and_brackets over flat arrays: stamps of all sensors concatenated with per-sensor counts.  Outputs are sized by the caller
// for num_ref * num_sensors entries; returns the number of cameras, -1 on error.
int mvh_lookup_images_and_brackets(double bracket_len, double max_change, int num_ref, const double* ref_stamps, int num_sensors, const int* counts,
                                   const double* stamps, const int* cloud_counts, const double* cloud_stamps, const double* offsets, int* sensor,
                                   int* beg, int* end, double* timestamp, int* image_index, int* cloud_index, double* min_off, double* max_off) {
  try {
    std::vector<std::vector<double>> img(num_sensors), cld;
    size_t o = 0;
    for (int t = 0; t < num_sensors; t++) { img[t].assign(stamps + o, stamps + o + counts[t]); o += counts[t]; }
    if (cloud_counts) { cld.resize(num_sensors); o = 0; for (int t = 0; t < num_sensors; t++) { cld[t].assign(cloud_stamps + o, cloud_stamps + o + cloud_counts[t]); o += cloud_counts[t]; } }
    mvba_host::Bracketing b = mvba_host::lookup_images_and_brackets(bracket_len, max_change, std::vector<double>(ref_stamps, ref_stamps + num_ref), img, cld,
                                                                    std::vector<double>(offsets, offsets + num_sensors));
    for (size_t i = 0; i < b.cams.size(); i++) {
      sensor[i] = b.cams[i].sensor; beg[i] = b.cams[i].beg_ref_index; end[i] = b.cams[i].end_ref_index; timestamp[i] = b.cams[i].timestamp;
      image_index[i] = b.cams[i].image_index; cloud_index[i] = b.cams[i].cloud_index;
    }
    for (int t = 0; t < num_sensors; t++) { min_off[t] = b.min_timestamp_offset[t]; max_off[t] = b.max_timestamp_offset[t]; }
    return (int)b.cams.size();
  } catch (const std::exception& e) { g_cam_err = e.what(); return -1; }
}
// lookup_images_no_brackets: outputs sized for the total number of images; returns the number of cameras, -1 on error
int mvh_lookup_images_no_brackets(int num_sensors, const int* counts, const double* stamps, int* sensor, int* beg, int* end, double* timestamp, int* image_index) {
  try {
    std::vector<std::vector<double>> img(num_sensors);
    size_t o = 0;
    for (int t = 0; t < num_sensors; t++) { img[t].assign(stamps + o, stamps + o + counts[t]); o += counts[t]; }
    mvba_host::Bracketing b = mvba_host::lookup_images_no_brackets(img, {});
    for (size_t i = 0; i < b.cams.size(); i++) {
      sensor[i] = b.cams[i].sensor; beg[i] = b.cams[i].beg_ref_index; end[i] = b.cams[i].end_ref_index; timestamp[i] = b.cams[i].timestamp;
      image_index[i] = b.cams[i].image_index;
    }
    return (int)b.cams.size();
  } catch (const std::exception& e) { g_cam_err = e.what(); return -1; }
}
// residual report of one evaluation (residual_stats.h): cam_names is a '\n'-separated list, one name per sensor; text receives the
// reference's printout (at most text_cap - 1 characters), table 5 numbers per name (q25 q50 q75 q100 count) in the printed order;
// returns the number of names, -1 on error
int mvh_residual_stats(long long num_blocks, const int* kind, const long long* row0, const long long* src, const int* cid, const int* cam_sensor,
                       const char* cam_names, const double* weights4, long long num_residuals, const double* residuals, const char* tag,
                       char* text, int text_cap, double* table, int table_cap) {
  try {
    std::vector<std::string> names_t;
    { std::string cur; for (const char* p = cam_names; ; p++) { if (*p == '\n' || *p == 0) { names_t.push_back(cur); cur.clear(); if (*p == 0) break; } else cur.push_back(*p); } }
    std::vector<int32_t> k32(kind, kind + num_blocks), cid32, cs32;
    std::vector<int64_t> r64(row0, row0 + num_blocks), s64(src, src + num_blocks);
    int64_t maxobs = -1; for (long long b = 0; b < num_blocks; b++) if (kind[b] == 0) maxobs = std::max<int64_t>(maxobs, src[b]);
    cid32.assign(cid, cid + maxobs + 1);
    int maxcam = -1; for (int c : cid32) maxcam = std::max(maxcam, c);
    cs32.assign(cam_sensor, cam_sensor + maxcam + 1);
    std::vector<std::string> names; std::vector<double> scales;
    mvba_host::residual_names_and_scales(num_blocks, k32.data(), r64.data(), s64.data(), cid32.data(), cs32.data(), names_t, weights4, num_residuals, names, scales);
    std::vector<mvba_host::ResidualStat> st = mvba_host::calc_residuals_stats(std::vector<double>(residuals, residuals + num_residuals), names, scales);
    const std::string out = mvba_host::format_residual_stats(st, tag);
    if (text && text_cap > 0) { const size_t n = std::min<size_t>(out.size(), (size_t)text_cap - 1); std::memcpy(text, out.data(), n); text[n] = 0; }
    for (size_t i = 0; i < st.size() && (int)(5 * i + 4) < table_cap; i++) {
      table[5 * i] = st[i].q25; table[5 * i + 1] = st[i].q50; table[5 * i + 2] = st[i].q75; table[5 * i + 3] = st[i].q100; table[5 * i + 4] = (double)st[i].count;
    }
    return (int)st.size();
  } catch (const std::exception& e) { g_cam_err = e.what(); return -1; }
}
// CameraParameters::updateRpcUndistortion through a caller-supplied fitter (the address of mvba_rpc_fit, or of the oracle's
// oracle_rpc_fit in tests): dist[nd] is updated in place (undistortion half); returns 0, -1 on error
int mvh_update_rpc_undistortion(const int* size, const double* focal, const double* center, double* dist, int nd, const int* undist_size,
                                const int* crop_size, void* fitter) {
  try {
    using namespace camera;
    CameraParameters cp(Vector2i(size[0], size[1]), Vector2d(focal[0], focal[1]), Vector2d(center[0], center[1]), std::vector<double>(dist, dist + nd));
    if (undist_size) cp.SetUndistortedSize(Vector2i(undist_size[0], undist_size[1]));
    if (crop_size) cp.SetDistortedCropSize(Vector2i(crop_size[0], crop_size[1]));
    CameraParameters::SetRpcFitter(reinterpret_cast<CameraParameters::RpcFitter>(fitter));
    cp.updateRpcUndistortion(1);
    const std::vector<double>& d = cp.GetDistortion();
    for (int i = 0; i < nd; i++) dist[i] = d[(size_t)i];
    return 0;
  } catch (const std::exception& e) { g_cam_err = e.what(); return -1; }
}
// ---- tests/test_cli.py ----
// read_ply: counts, then (when the buffers are large enough) the arrays
int mvh_read_ply(const char* path, long long cap_vertices, float* vertices, long long cap_faces, unsigned* faces, long long* num_vertices, long long* num_faces) {
  try {
    mvba_host::TriangleMesh m = mvba_host::read_ply(path);
    *num_vertices = (long long)m.vertices.size() / 3; *num_faces = (long long)m.faces.size() / 3;
    if (*num_vertices <= cap_vertices && vertices) std::memcpy(vertices, m.vertices.data(), m.vertices.size() * sizeof(float));
    if (*num_faces <= cap_faces && faces) std::memcpy(faces, m.faces.data(), m.faces.size() * sizeof(unsigned));
    return 0;
  } catch (const std::exception& e) { g_cam_err = e.what(); return -1; }
}
// calc_rig_using_word_to_cam; world_to_cam as 12 doubles per image (R row-major, t), out_ref_to_cam 12 per sensor
int mvh_rig_from_poses(int num_sensors, int num_cameras, const int* cam_sensor, const int* beg, const int* end, const double* cam_stamp,
                       const double* ref_stamps, const double* world_to_ref7, const double* world_to_cam12, const double* offsets, double* out_ref_to_cam12) {
  try {
    std::vector<mvba_host::Affine> w2c((size_t)num_cameras);
    for (int c = 0; c < num_cameras; c++) std::memcpy(w2c[c].data(), world_to_cam12 + 12 * (size_t)c, 12 * sizeof(double));
    auto r = mvba_host::calc_rig_using_world_to_cam(num_sensors, num_cameras, cam_sensor, beg, end, cam_stamp, ref_stamps, world_to_ref7, w2c, offsets);
    for (int t = 0; t < num_sensors; t++) std::memcpy(out_ref_to_cam12 + 12 * (size_t)t, r[t].data(), 12 * sizeof(double));
    return 0;
  } catch (const std::exception& e) { g_cam_err = e.what(); return -1; }
}
// the flag table as text: one "name type default" line per flag, in declaration order
int mvh_flag_table(char* out, int cap) {
  mvba_host::Flags f;
  std::string s;
  for (const mvba_host::FlagSpec& k : mvba_host::flag_table()) {
    const char* base = reinterpret_cast<const char*>(&f) + k.offset;
    std::string v;
    if (k.type == 'd') v = mvba_host::format_double(*reinterpret_cast<const double*>(base));
    else if (k.type == 'i') v = std::to_string(*reinterpret_cast<const int*>(base));
    else if (k.type == 'b') v = *reinterpret_cast<const bool*>(base) ? "true" : "false";
    else v = "\"" + *reinterpret_cast<const std::string*>(base) + "\"";
    s += std::string(k.name) + " " + k.type + " " + v + "\n";
  }
  if ((int)s.size() + 1 > cap) return -1;
  std::memcpy(out, s.c_str(), s.size() + 1);
  return (int)mvba_host::flag_table().size();
}
// ---- tests/test_camera_params.py: the calibration-file constructor and GenerateRemapMaps ----
// out16: width height crop_w crop_h undist_w undist_h crop_x crop_y fx fy cx cy nd, then up to 3 distortion values are not enough: dist
// receives GetDistortion() (cap_dist values at most)
int mvh_camera_from_file(const char* file, const char* base_dir, double* out13, double* dist, int cap_dist) {
  try {
    camera::CameraParameters cp(std::string(file), std::string(base_dir ? base_dir : ""));
    const camera::Vector2i a = cp.GetDistortedSize(), b = cp.GetDistortedCropSize(), c = cp.GetUndistortedSize(), d = cp.GetCropOffset();
    const camera::Vector2d f = cp.GetFocalVector(), o = cp.GetOpticalOffset();
    const double v[13] = {(double)a[0], (double)a[1], (double)b[0], (double)b[1], (double)c[0], (double)c[1], (double)d[0], (double)d[1], f[0], f[1], o[0], o[1],
                          (double)cp.GetDistortion().size()};
    std::memcpy(out13, v, sizeof(v));
    for (size_t i = 0; i < cp.GetDistortion().size() && (int)i < cap_dist; i++) dist[i] = cp.GetDistortion()[i];
    return 0;
  } catch (const std::exception& e) { g_cam_err = e.what(); return -1; }
}
int mvh_generate_remap_maps(const int* size, const double* focal, const double* center, const double* dist, int nd, const int* undist_size, double scale,
                            float* map, long long cap_pairs, int* rows, int* cols) {
  try {
    using namespace camera;
    CameraParameters cp(Vector2i(size[0], size[1]), Vector2d(focal[0], focal[1]), Vector2d(center[0], center[1]), std::vector<double>(dist, dist + nd));
    if (undist_size) cp.SetUndistortedSize(Vector2i(undist_size[0], undist_size[1]));
    std::vector<float> m;
    cp.GenerateRemapMaps(&m, rows, cols, scale);
    if ((long long)m.size() / 2 > cap_pairs) { g_cam_err = "map buffer too small"; return -1; }
    std::memcpy(map, m.data(), m.size() * sizeof(float));
    return 0;
  } catch (const std::exception& e) { g_cam_err = e.what(); return -1; }
}
}
