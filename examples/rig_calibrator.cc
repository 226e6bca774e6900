// examples/rig_calibrator.cc -- the reference's rig_calibrator command line over the pieces of this repository: rig_config.txt + an NVM map
// in, refined rig_config.txt / cameras.txt (/ cameras.nvm) out.  Same 43 flags, defaults and validation messages as
// rig_calibrator/bin/rig_calibrator.cc:135-316, 893-955 (multiview_b200/host/cli_flags.h), same order of work as its main()
// (:1265-2030): readRigConfig, readNvm, lookupImages, rig transforms from the poses unless --use_initial_rig_transforms,
// parameter vectors, flagOutlierByExclusionDist, then per pass multiViewTriangulation, meshTriangulations, the residual blocks with the
// reference's constancy rules (:1702-1752, 1787-1793), evalResiduals before / after, the solve, copy-back (+ updateRpcUndistortion),
// flagOutliersByTriAngleAndReprojErr; finally saveCameraPoses, writeRigConfig and, with --save_nvm, writeNvm.  Every device step goes
// through include/mvba.h, every host step through multiview_b200/host.
// No image is opened: features and tracks come from the NVM file, sensors from the parent directory of each image path, timestamps
// from the file name (interest_point.cc:1619-1627); a depth cloud is read from <image stem>.pc when that file exists (:1664-1669).
// Flags whose work lies outside the bundle-adjustment path (--num_overlaps, --no_nvm_matches, --camera_poses, --registration,
// --out_texture_dir, --save_matches, --export_to_voxblox, --save_transformed_depth_clouds) are parsed and validated, then refused.
// Accepted without effect: --num_opt_threads (Ceres' thread count; it is passed on to updateRpcUndistortion like the reference does),
// --num_match_threads, --initial_max_reprojection_error (the matcher's), --verbose, --hugin_file / --xyz_file without --registration.
// One switch of its own: --dry_run stops before the first device call and prints the assembled pass (tests/test_rig_pass.py on a CPU box).
// Exit codes: 1 command-line syntax (gflags), 2 a failed validation / input error (where the reference's LOG(FATAL) aborts), 3 a failed
// device call (there is no CPU fallback).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <filesystem>
#include <fstream>
#include <map>
#include <string>
#include <vector>

#include "../include/mvba.h"
#include "../multiview_b200/host/bracket.h"
#include "../multiview_b200/host/cli_flags.h"
#include "../multiview_b200/host/pass_utils.h"
#include "../multiview_b200/host/ply_io.h"
#include "../multiview_b200/host/pose_math.h"
#include "../multiview_b200/host/residual_stats.h"
#include "../multiview_b200/host/rpc_pairs.h"
#include "../multiview_b200/host/rig_io.h"

using namespace mvba_host;

static void die(const std::string& m, int code = 2) { std::fprintf(stderr, "rig_calibrator: %s\n", m.c_str()); std::exit(code); }
#define CK(call) do { if ((call) != MVBA_OK) { std::fprintf(stderr, "rig_calibrator: %s failed: %s\n", #call, mvba_last_error()); std::exit(3); } } while (0)

static void split_image_path(const std::string& file, std::string& sensor, double& stamp) {  // interest_point.cc:1619-1627
  size_t s1 = file.find_last_of('/');
  std::string dir = s1 == std::string::npos ? "" : file.substr(0, s1), base = s1 == std::string::npos ? file : file.substr(s1 + 1);
  size_t s0 = dir.find_last_of('/');
  sensor = s0 == std::string::npos ? dir : dir.substr(s0 + 1);
  if (base.empty() || base[0] < '0' || base[0] > '9') die("Image name (without directory) must start with digits. Got: " + base + "\n");
  stamp = std::atof(base.c_str());
}

struct DepthCloud { int rows = 0, cols = 0; std::vector<float> xyz; };
static bool read_xyz_image(const std::string& file, DepthCloud& c) {  // readXyzImage, interest_point.cc:1585-1609
  std::ifstream f(file, std::ios::binary);
  if (!f.is_open()) return false;
  int channels = 0;
  f.read((char*)&c.rows, 4); f.read((char*)&c.cols, 4); f.read((char*)&channels, 4);
  if (!f || c.rows < 0 || c.cols < 0 || channels < 0 || channels > 3) die("Cannot read depth cloud: " + file);
  c.xyz.assign(3 * (size_t)c.rows * c.cols, 0.0f);
  for (size_t px = 0; px < (size_t)c.rows * c.cols; px++) f.read((char*)&c.xyz[3 * px], 4 * channels);
  if (!f) die("Cannot read depth cloud: " + file);
  return true;
}

int main(int argc, char** argv) {
  Flags F;
  std::set<std::string> extra;
  try { extra = parse_command_line(argc, argv, F, {"dry_run"}); } catch (const std::exception& e) { die(e.what(), 1); }
  try { parameter_validation(F); } catch (const std::exception& e) { die(e.what()); }
  for (const std::string& f : out_of_scope(F)) die(f + " is outside the bundle-adjustment path this build covers");
  const bool dry = extra.count("dry_run") > 0;

  RigConfig rig; Nvm nvm;
  try { rig = read_rig_config(F.rig_config, F.use_initial_rig_transforms); nvm = read_nvm(F.nvm); } catch (const std::exception& e) { die(e.what()); }
  const int T = (int)rig.sensors.size(), NI = (int)nvm.image_files.size();
  std::vector<std::string> cam_names;
  for (auto& s : rig.sensors) cam_names.push_back(s.name);

  // ---- what floats (rig_calibrator.cc:1423-1441) ----
  std::vector<std::set<std::string>> intr_float;
  std::set<std::string> rig_float, pose_float, d2i_float;
  try {
    intr_float = parse_intrinsics_to_float(F.intrinsics_to_float, cam_names);
    rig_float = parse_rig_transforms_to_float(cam_names, 0, F.rig_transforms_to_float);
    pose_float = parse_camera_names(cam_names, F.camera_poses_to_float);
    d2i_float = parse_camera_names(cam_names, F.depth_to_image_transforms_to_float);
  } catch (const std::exception& e) { die(e.what()); }

  // ---- images per sensor in time order (readNvm, interest_point.cc:1744-1790), depth clouds beside them ----
  std::vector<std::vector<std::pair<double, int>>> per_sensor(T);   // (timestamp, nvm image index)
  for (int i = 0; i < NI; i++) {
    std::string name; double stamp;
    split_image_path(nvm.image_files[i], name, stamp);
    int t = -1;
    for (int k = 0; k < T; k++) if (rig.sensors[k].name == name) t = k;
    if (t < 0) die("Could not determine sensor name from image path: " + nvm.image_files[i] + "\n");
    for (auto& p : per_sensor[t]) if (p.first == stamp) die("Duplicate timestamp " + format_double(stamp) + " for sensor id " + std::to_string(t) + "\n");
    per_sensor[t].push_back({stamp, i});
  }
  std::vector<std::vector<double>> stamps(T), cloud_stamps(T);
  std::vector<std::vector<DepthCloud>> clouds(T);
  std::vector<double> offsets(T);
  bool have_clouds = false;
  for (int t = 0; t < T; t++) {
    std::sort(per_sensor[t].begin(), per_sensor[t].end());
    for (auto& p : per_sensor[t]) {
      stamps[t].push_back(p.first);
      const std::string& img = nvm.image_files[p.second];
      const size_t dot = img.find_last_of('.'), slash = img.find_last_of('/');
      const std::string pc = (dot != std::string::npos && (slash == std::string::npos || dot > slash) ? img.substr(0, dot) : img) + ".pc";
      DepthCloud c;
      if (!dry && read_xyz_image(pc, c)) { cloud_stamps[t].push_back(p.first); clouds[t].push_back(std::move(c)); have_clouds = true; }
    }
    offsets[t] = rig.sensors[t].ref_to_sensor_timestamp_offset;
  }
  const std::vector<double> ref_stamps = stamps[0];
  Bracketing br;
  try {
    const std::vector<std::vector<double>> no_clouds;
    br = F.no_rig ? lookup_images_no_brackets(stamps, have_clouds ? cloud_stamps : no_clouds)
                  : lookup_images(F.bracket_len, F.timestamp_offsets_max_change, ref_stamps, stamps, have_clouds ? cloud_stamps : no_clouds, offsets);
  } catch (const std::exception& e) { die(e.what()); }
  const int C = (int)br.cams.size(), R = (int)ref_stamps.size();
  std::vector<int32_t> cam_sensor(C), cam_beg(C), cam_end(C);
  std::vector<double> cam_stamp(C);
  std::vector<int> cam_of_image(NI, -1), cam_cloud(C, -1);
  std::vector<std::string> cam_files(C);
  std::vector<Affine> world_to_cam(C);
  for (int c = 0; c < C; c++) {
    const BracketedImage& b = br.cams[c];
    cam_sensor[c] = b.sensor; cam_beg[c] = b.beg_ref_index; cam_end[c] = b.end_ref_index; cam_stamp[c] = b.timestamp;
    const int img = per_sensor[b.sensor][b.image_index].second;
    cam_of_image[img] = c; cam_files[c] = nvm.image_files[img]; world_to_cam[c] = nvm.world_to_cam[img];
    cam_cloud[c] = b.cloud_index;
  }
  std::vector<int> ib(cam_beg.begin(), cam_beg.end()), ie(cam_end.begin(), cam_end.end()), is(cam_sensor.begin(), cam_sensor.end());

  // ---- parameter vectors (rig_calibrator.cc:1347-1492) ----
  std::vector<double> world_to_ref(7 * (size_t)R), ref_to_cam(7 * (size_t)T), focal(T), center(2 * (size_t)T), dist, scale(T, 1.0), world_to_cam_vec;
  for (int r = 0; r < R; r++) rigid_to_array(nvm.world_to_cam[per_sensor[0][r].second], &world_to_ref[7 * (size_t)r]);
  if (!F.use_initial_rig_transforms && !F.no_rig) {   // the rig from the poses the map came with (:1364-1376)
    try {
      std::vector<Affine> r2c = calc_rig_using_world_to_cam(T, C, is.data(), ib.data(), ie.data(), cam_stamp.data(), ref_stamps.data(), world_to_ref.data(),
                                                            world_to_cam, offsets.data());
      for (int t = 0; t < T; t++) rig.sensors[t].ref_to_sensor = r2c[t];
    } catch (const std::exception& e) { die(e.what()); }
  }
  const int nd2i = F.affine_depth_to_image ? MVBA_NUM_AFFINE_PARAMS : MVBA_NUM_RIGID_PARAMS;
  std::vector<double> d2i((size_t)nd2i * T), img_size(2 * (size_t)T), crop_size(2 * (size_t)T);
  for (int t = 0; t < T; t++) {
    const SensorConfig& s = rig.sensors[t];
    rigid_to_array(s.ref_to_sensor, &ref_to_cam[7 * (size_t)t]);
    // the scale of depth_to_image is held apart (:1451-1458): cube root of the determinant
    Affine d = s.depth_to_image;
    const double det = d[0] * (d[4] * d[8] - d[5] * d[7]) - d[1] * (d[3] * d[8] - d[5] * d[6]) + d[2] * (d[3] * d[7] - d[4] * d[6]);
    scale[t] = std::pow(det, 1.0 / 3.0);
    for (int k = 0; k < 9; k++) d[k] /= scale[t];
    if (F.affine_depth_to_image) {       // affine_transform_to_array, dense_map_utils.cc:181-192: rows of [L | t]
      for (int r = 0; r < 3; r++) { for (int k = 0; k < 3; k++) d2i[12 * (size_t)t + 4 * r + k] = d[3 * r + k]; d2i[12 * (size_t)t + 4 * r + 3] = d[9 + r]; }
    } else {
      rigid_to_array(d, &d2i[7 * (size_t)t]);
    }
    focal[t] = s.focal_length; center[2 * t] = s.optical_center[0]; center[2 * t + 1] = s.optical_center[1];
    dist.insert(dist.end(), s.distortion.begin(), s.distortion.end());
    img_size[2 * t] = s.image_size[0]; img_size[2 * t + 1] = s.image_size[1];
    crop_size[2 * t] = s.distorted_crop_size[0]; crop_size[2 * t + 1] = s.distorted_crop_size[1];
  }
  if (F.no_rig) {
    world_to_cam_vec.resize(7 * (size_t)C);
    for (int c = 0; c < C; c++) rigid_to_array(world_to_cam[c], &world_to_cam_vec[7 * (size_t)c]);
  }

  // ---- features, track-major, cid ascending inside a track (the std::map order of pid_to_cid_fid); NVM keypoints are relative to the
  // optical centre (appendMatchesFromNvm, interest_point.cc:1800-1850) ----
  const int P = (int)nvm.pid_to_cid_fid.size();
  if (P == 0) die("No interest points were found. Must specify either --nvm or positive --num_overlaps.\n");
  std::vector<int32_t> opid, ocid; std::vector<float> pix;
  for (int p = 0; p < P; p++) {
    std::vector<std::pair<int, int>> cf;
    for (auto& kv : nvm.pid_to_cid_fid[p]) if (cam_of_image[kv.first] >= 0) cf.push_back({cam_of_image[kv.first], kv.second});
    std::sort(cf.begin(), cf.end());
    for (size_t k = 0; k < cf.size(); k++) {
      int img = -1;
      for (auto& kv : nvm.pid_to_cid_fid[p]) if (cam_of_image[kv.first] == cf[k].first) img = kv.first;
      const int t = cam_sensor[cf[k].first];
      opid.push_back(p); ocid.push_back(cf[k].first);
      pix.push_back((float)(nvm.keypoints[img][cf[k].second][0] + center[2 * t]));
      pix.push_back((float)(nvm.keypoints[img][cf[k].second][1] + center[2 * t + 1]));
    }
  }
  const int64_t N = (int64_t)opid.size();
  std::vector<double> xyz(3 * (size_t)P);
  for (int p = 0; p < P; p++) for (int a = 0; a < 3; a++) xyz[3 * (size_t)p + a] = nvm.xyz[p][a];

  // ---- float mask: the SetParameterBlockConstant rules of :1702-1752, 1787-1793 ----
  std::vector<uint8_t> f_focal(T), f_center(T), f_dist(T), f_rig(T), f_d2i(T), f_pose(T);
  for (int t = 0; t < T; t++) {
    f_focal[t] = intr_float[t].count("focal_length") > 0; f_center[t] = intr_float[t].count("optical_center") > 0;
    f_dist[t] = intr_float[t].count("distortion") > 0;
    f_rig[t] = rig_float.count(cam_names[t]) > 0; f_d2i[t] = d2i_float.count(cam_names[t]) > 0; f_pose[t] = pose_float.count(cam_names[t]) > 0;
  }

  std::printf("{\"sensors\": %d, \"ref_frames\": %d, \"cameras\": %d, \"tracks\": %d, \"features\": %lld,\n \"cams\": [", T, R, C, P, (long long)N);
  for (int c = 0; c < C; c++) std::printf("%s[%d, %d, %d, %.17g]", c ? ", " : "", cam_sensor[c], cam_beg[c], cam_end[c], cam_stamp[c]);
  std::printf("],\n \"offset_bounds\": [");
  for (int t = 0; t < T; t++) std::printf("%s[%.17g, %.17g]", t ? ", " : "", br.min_timestamp_offset[t], br.max_timestamp_offset[t]);
  std::printf("],\n \"first_features\": [");
  for (int64_t i = 0; i < std::min<int64_t>(N, 6); i++) std::printf("%s[%d, %d, %.9g, %.9g]", i ? ", " : "", opid[i], ocid[i], pix[2 * i], pix[2 * i + 1]);
  std::printf("],\n \"ref_to_cam\": [");
  for (size_t k = 0; k < ref_to_cam.size(); k++) std::printf("%s%.17g", k ? ", " : "", ref_to_cam[k]);
  std::printf("],\n \"float\": {");
  const std::pair<const char*, const std::vector<uint8_t>*> masks[6] = {{"focal_length", &f_focal}, {"optical_center", &f_center}, {"distortion", &f_dist},
                                                                        {"rig", &f_rig}, {"depth_to_image", &f_d2i}, {"poses", &f_pose}};
  for (int m = 0; m < 6; m++) {
    std::printf("%s\"%s\": [", m ? ", " : "", masks[m].first);
    for (int t = 0; t < T; t++) std::printf("%s%d", t ? ", " : "", (int)(*masks[m].second)[t]);
    std::printf("]");
  }
  std::printf(", \"timestamp_offsets\": %d, \"scale\": %d}}\n", (int)F.float_timestamp_offsets, (int)F.float_scale);
  if (dry) return 0;

  // ---- mesh (loadMeshBuildTree, :1298-1299) and the depth measurement of every feature (depthValue, :1757) ----
  TriangleMesh mesh;
  if (F.mesh != "") { try { mesh = read_ply(F.mesh); } catch (const std::exception& e) { die(e.what()); } }
  std::vector<uint8_t> has_depth; std::vector<double> depth_xyz;
  if (have_clouds) {
    has_depth.assign((size_t)N, 0); depth_xyz.assign(3 * (size_t)N, 0.0);
    for (int64_t i = 0; i < N; i++) {
      const int c = ocid[i], t = cam_sensor[c];
      if (cam_cloud[c] < 0) continue;
      const DepthCloud& dc = clouds[t][cam_cloud[c]];
      const double ip[2] = {pix[2 * i], pix[2 * i + 1]};
      try { has_depth[i] = depth_value(dc.xyz.data(), dc.rows, dc.cols, ip, &depth_xyz[3 * i]); } catch (const std::exception& e) { die(e.what()); }
    }
  }

  // ---- device side ----
  mvba_config cfg = {};
  cfg.num_sensors = T; cfg.num_ref_frames = R; cfg.num_cameras = C; cfg.num_landmarks = P;
  cfg.no_rig = F.no_rig; cfg.affine_depth_to_image = F.affine_depth_to_image;
  cfg.robust_threshold = F.robust_threshold; cfg.depth_tri_weight = F.depth_tri_weight; cfg.depth_mesh_weight = F.mesh != "" ? F.depth_mesh_weight : 0.0;
  mvba_params prm = {};
  prm.world_to_ref = world_to_ref.data(); prm.ref_stamps = ref_stamps.data(); prm.ref_to_cam = ref_to_cam.data(); prm.offsets = offsets.data();
  prm.offset_lo = br.min_timestamp_offset.data(); prm.offset_hi = br.max_timestamp_offset.data();
  prm.focal = focal.data(); prm.center = center.data(); prm.distortion = dist.data(); prm.depth_to_image = d2i.data(); prm.depth_scale = scale.data();
  prm.world_to_cam = F.no_rig ? world_to_cam_vec.data() : nullptr;
  prm.xyz = xyz.data();
  mvba_float_mask mask = {f_focal.data(), f_center.data(), f_dist.data(), f_rig.data(), f_d2i.data(), f_pose.data(), F.float_timestamp_offsets, F.float_scale};
  auto make = [&](int64_t n, const int32_t* pid, const int32_t* cid, const float* px, const uint8_t* hd, const float* dxyz, const uint8_t* hm, const double* mxyz) {
    mvba_problem* h = nullptr;
    CK(mvba_create(&cfg, 0, &h));
    for (int t = 0; t < T; t++) CK(mvba_set_sensor(h, t, img_size[2 * t], img_size[2 * t + 1], (int)rig.sensors[t].distortion.size()));
    CK(mvba_set_params(h, &prm));
    CK(mvba_set_cameras(h, C, cam_sensor.data(), F.no_rig ? nullptr : cam_beg.data(), F.no_rig ? nullptr : cam_end.data(), cam_stamp.data()));
    CK(mvba_set_observations(h, n, pid, cid, px, hd, dxyz, hm, mxyz));
    CK(mvba_set_float_mask(h, &mask));
    return h;
  };
  mvba_problem* h_all = make(N, opid.data(), ocid.data(), pix.data(), nullptr, nullptr, nullptr, nullptr);   // all features: triangulation, mesh, outlier flags
  std::vector<uint8_t> inl((size_t)N, 1);
  CK(mvba_flag_exclusion(h_all, img_size.data(), crop_size.data(), inl.data()));
  std::vector<uint8_t> has_mesh; std::vector<double> mesh_xyz, pid_mesh_xyz;
  bool rpc_refit_registered = false;
  for (int pass = 0; pass < F.calibrator_num_passes; pass++) {
    std::printf("\nOptimization pass %d / %d\n", pass + 1, F.calibrator_num_passes);
    CK(mvba_triangulate(h_all, inl.data(), xyz.data()));
    const std::vector<double> xyz_orig = xyz;
    if (F.mesh != "") {
      has_mesh.assign((size_t)N, 0); mesh_xyz.assign(3 * (size_t)N, 0.0); pid_mesh_xyz.assign(3 * (size_t)P, 0.0);
      CK(mvba_mesh_intersect(h_all, (int64_t)mesh.vertices.size() / 3, mesh.vertices.data(), (int64_t)mesh.faces.size() / 3, mesh.faces.data(),
                             F.min_ray_dist, F.max_ray_dist, inl.data(), has_mesh.data(), mesh_xyz.data(), pid_mesh_xyz.data()));
    }
    Observations o = compact_observations(N, opid.data(), ocid.data(), pix.data(), have_clouds ? has_depth.data() : nullptr, have_clouds ? depth_xyz.data() : nullptr,
                                          F.mesh != "" ? has_mesh.data() : nullptr, F.mesh != "" ? mesh_xyz.data() : nullptr, inl.data());
    std::vector<float> dxyz32(o.depth_xyz.begin(), o.depth_xyz.end());
    mvba_problem* h = make((int64_t)o.pid.size(), o.pid.data(), o.cid.data(), o.pix.data(), have_clouds ? o.has_depth.data() : nullptr,
                           have_clouds ? dxyz32.data() : nullptr, F.mesh != "" ? o.has_mesh.data() : nullptr, F.mesh != "" ? o.mesh_xyz.data() : nullptr);
    // XYZError blocks for tracks with at least one inlier feature (:1845-1901)
    std::vector<int32_t> tri_pid, mesh_pid; std::vector<double> tri_ref, mesh_ref;
    for (size_t k = 0; k < o.pid.size(); k++) {
      const int p = o.pid[k];
      if (k > 0 && o.pid[k - 1] == p) continue;
      if (F.mesh != "" && F.mesh_tri_weight > 0 && !(pid_mesh_xyz[3 * (size_t)p] == 1.0e+100 && pid_mesh_xyz[3 * (size_t)p + 1] == 1.0e+100 && pid_mesh_xyz[3 * (size_t)p + 2] == 1.0e+100)) {
        mesh_pid.push_back(p); for (int a = 0; a < 3; a++) mesh_ref.push_back(pid_mesh_xyz[3 * (size_t)p + a]);
      }
      if (F.tri_weight > 0.0) { tri_pid.push_back(p); for (int a = 0; a < 3; a++) tri_ref.push_back(xyz_orig[3 * (size_t)p + a]); }
    }
    if (!mesh_pid.empty()) CK(mvba_set_point_priors(h, MVBA_BLOCK_MESH_TRI, (int64_t)mesh_pid.size(), mesh_pid.data(), mesh_ref.data(), F.mesh_tri_weight, F.robust_threshold));
    if (!tri_pid.empty()) CK(mvba_set_point_priors(h, MVBA_BLOCK_TRI, (int64_t)tri_pid.size(), tri_pid.data(), tri_ref.data(), F.tri_weight, F.tri_robust_threshold));
    mvba_solve_options so; mvba_solve_options_init(&so);
    so.max_num_iterations = F.num_iterations; so.parameter_tolerance = F.parameter_tolerance; so.verbose = 1;
    mvba_summary sum;
    // residual report before and after the solve (evalResiduals, rig_calibrator.cc:1907 and 1970)
    auto report = [&](const char* tag) {
      int64_t nres = 0, nblk = 0;
      CK(mvba_num_residuals(h, &nres, &nblk));
      std::vector<int32_t> kind((size_t)nblk); std::vector<int64_t> row0((size_t)nblk), src((size_t)nblk);
      CK(mvba_residual_layout(h, kind.data(), row0.data(), src.data()));
      std::vector<double> res((size_t)nres);
      double cost = 0;
      CK(mvba_evaluate(h, 0, &cost, res.data()));
      std::vector<std::string> names; std::vector<double> scales;
      const double w4[4] = {F.depth_tri_weight, F.depth_mesh_weight, F.mesh_tri_weight, F.tri_weight};
      residual_names_and_scales(nblk, kind.data(), row0.data(), src.data(), o.cid.data(), cam_sensor.data(), cam_names, w4, nres, names, scales);
      std::fputs(format_residual_stats(calc_residuals_stats(res, names, scales), tag).c_str(), stdout);
    };
    report("before opt");
    CK(mvba_solve(h, &so, &sum, nullptr, 0));
    // RPC sensors whose distortion floated: the undistortion half is re-fitted from the new distortion half (:1946-1948 ->
    // CameraParameters::updateRpcUndistortion); the fit runs on the device (mvba_rpc_fit)
    size_t d0 = 0;
    for (int t = 0; t < T; t++) {
      SensorConfig& sc = rig.sensors[t];
      const size_t nd = sc.distortion.size();
      if (f_dist[t] && nd > 5) {
        if (!rpc_refit_registered) { camera::CameraParameters::SetRpcFitter(&mvba_rpc_fit); rpc_refit_registered = true; }
        try {
          std::vector<double> dv(dist.begin() + d0, dist.begin() + d0 + nd);
          camera::CameraParameters cp(camera::Vector2i(sc.image_size[0], sc.image_size[1]), camera::Vector2d(focal[t], focal[t]),
                                      camera::Vector2d(center[2 * t], center[2 * t + 1]), dv);
          cp.SetUndistortedSize(camera::Vector2i(sc.undistorted_image_size[0], sc.undistorted_image_size[1]));
          cp.SetDistortedCropSize(camera::Vector2i(sc.distorted_crop_size[0], sc.distorted_crop_size[1]));
          cp.updateRpcUndistortion(F.num_opt_threads);
          dv = cp.GetDistortion();
          std::copy(dv.begin(), dv.end(), dist.begin() + d0);
        } catch (const std::exception& e) { die(e.what()); }
      }
      d0 += nd;
    }
    report("after opt");
    std::printf("pass %d: %zu inlier features, cost %.6e -> %.6e in %d iterations\n", pass, o.pid.size(), sum.initial_cost, sum.final_cost, sum.num_iterations);
    CK(mvba_destroy(h));
    int64_t counts[4];
    CK(mvba_flag_outliers(h_all, F.min_triangulation_angle, F.max_reprojection_error, inl.data(), counts));
    std::printf("Removed %lld outlier features with small angle of convergence, out of %lld\nRemoved %lld outlier features using reprojection error, out of %lld\n",
                (long long)counts[1], (long long)counts[0], (long long)counts[3], (long long)counts[2]);
  }
  CK(mvba_destroy(h_all));

  // ---- copy-back and outputs (rig_calibrator.cc:1924-1967, 1987-2021; dense_map_utils.cc:753-850; interest_point.cc:1338-1475) ----
  size_t d0 = 0;
  for (int t = 0; t < T; t++) {
    SensorConfig& sc = rig.sensors[t];
    sc.ref_to_sensor = array_to_rigid(&ref_to_cam[7 * (size_t)t]); sc.ref_to_sensor_timestamp_offset = offsets[t];
    sc.focal_length = focal[t]; sc.optical_center[0] = center[2 * t]; sc.optical_center[1] = center[2 * t + 1];
    std::copy(dist.begin() + d0, dist.begin() + d0 + sc.distortion.size(), sc.distortion.begin()); d0 += sc.distortion.size();
    Affine d;
    if (F.affine_depth_to_image) { for (int r = 0; r < 3; r++) { for (int k = 0; k < 3; k++) d[3 * r + k] = d2i[12 * (size_t)t + 4 * r + k]; d[9 + r] = d2i[12 * (size_t)t + 4 * r + 3]; } }
    else d = array_to_rigid(&d2i[7 * (size_t)t]);
    for (int k = 0; k < 9; k++) d[k] *= scale[t];   // the scale goes back in (:1987-1989)
    sc.depth_to_image = d;
  }
  std::vector<Affine> w2c(C);
  if (F.no_rig) for (int c = 0; c < C; c++) w2c[c] = array_to_rigid(&world_to_cam_vec[7 * (size_t)c]);
  else w2c = world_to_cam_all(C, is.data(), ib.data(), ie.data(), cam_stamp.data(), ref_stamps.data(), world_to_ref.data(), ref_to_cam.data(), offsets.data());
  std::vector<CameraEntry> entries(C);
  for (int c = 0; c < C; c++) { entries[c].image_file = cam_files[c]; entries[c].world_to_cam = w2c[c]; }
  try {
    std::filesystem::create_directories(F.out_dir);   // dense_map::createDir, as saveCameraPoses / writeRigConfig do
    write_camera_list(F.out_dir + "/cameras.txt", entries);
    write_rig_config(F.out_dir + "/rig_config.txt", rig, !F.no_rig);
    if (F.save_nvm) {
      std::vector<std::array<double, 2>> ctr(T);
      for (int t = 0; t < T; t++) ctr[t] = {center[2 * t], center[2 * t + 1]};
      std::vector<int> vpid(opid.begin(), opid.end()), vcid(ocid.begin(), ocid.end());
      write_nvm(F.out_dir + "/cameras.nvm", nvm_from_inliers(cam_files, is, focal, ctr, w2c, vpid, vcid, pix, inl, xyz));
    }
  } catch (const std::exception& e) { die(e.what()); }
  return 0;
}
