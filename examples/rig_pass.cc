// examples/rig_pass.cc -- a stand-alone driver over the pieces of this repository: rig_config.txt + an NVM map in, refined
// rig_config.txt / cameras.txt / NVM out.  It follows the pass structure of the reference's rig_calibrator
// (rig_calibrator/bin/rig_calibrator.cc:1332 lookupImages, :1524 exclusion, :1538-1548 triangulation, :1611-1919 one Ceres pass,
// :1969-1983 outlier flagging, :1924-1967 copy-back) with every device step done through include/mvba.h and every host step
// through multiview_b200/host.  No image is opened: features and tracks come from the NVM file (detection and matching are out
// of scope), sensors from the parent directory of each image path, timestamps from the file name, as in the reference.
//   rig_pass --rig_config rig.txt --nvm map.nvm --out_dir out [--num_passes 2] [--num_iterations 20] [--float_rig]
//            [--bracket_len 0.6] [--timestamp_offsets_max_change 1] [--min_triangulation_angle 0.01]
//            [--max_reprojection_error 25] [--robust_threshold 3] [--dry_run]
// --dry_run stops before the first device call and prints the assembled pass (what tests/test_rig_pass.py checks on a CPU box).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "../include/mvba.h"
#include "../multiview_b200/host/bracket.h"
#include "../multiview_b200/host/pass_utils.h"
#include "../multiview_b200/host/pose_math.h"
#include "../multiview_b200/host/residual_stats.h"
#include "../multiview_b200/host/rpc_pairs.h"
#include "../multiview_b200/host/rig_io.h"

using namespace mvba_host;

static void die(const std::string& m) { std::fprintf(stderr, "rig_pass: %s\n", m.c_str()); std::exit(2); }
#define CK(call) do { if ((call) != MVBA_OK) { std::fprintf(stderr, "rig_pass: %s failed: %s\n", #call, mvba_last_error()); std::exit(3); } } while (0)

static void split_image_path(const std::string& file, std::string& sensor, double& stamp) {  // interest_point.cc:1619-1627
  size_t s1 = file.find_last_of('/');
  std::string dir = s1 == std::string::npos ? "" : file.substr(0, s1), base = s1 == std::string::npos ? file : file.substr(s1 + 1);
  size_t s0 = dir.find_last_of('/');
  sensor = s0 == std::string::npos ? dir : dir.substr(s0 + 1);
  stamp = std::atof(base.c_str());
}

int main(int argc, char** argv) {
  std::map<std::string, std::string> opt = {{"num_passes", "2"}, {"num_iterations", "20"}, {"bracket_len", "0.6"}, {"timestamp_offsets_max_change", "1"},
                                            {"min_triangulation_angle", "0.01"}, {"max_reprojection_error", "25"}, {"robust_threshold", "3"}};
  for (int i = 1; i < argc; i++) {
    std::string a = argv[i];
    if (a.rfind("--", 0) != 0) die("unexpected argument " + a);
    a = a.substr(2);
    if (a == "dry_run" || a == "float_rig") opt[a] = "1";
    else if (i + 1 < argc) opt[a] = argv[++i];
    else die("missing value for --" + a);
  }
  for (const char* k : {"rig_config", "nvm", "out_dir"}) if (!opt.count(k)) die(std::string("--") + k + " is required");
  const bool dry = opt.count("dry_run") > 0;

  RigConfig rig; Nvm nvm;
  try { rig = read_rig_config(opt["rig_config"], true); nvm = read_nvm(opt["nvm"]); } catch (const std::exception& e) { die(e.what()); }
  const int T = (int)rig.sensors.size(), NI = (int)nvm.image_files.size();

  // ---- images per sensor in time order, reference frames, brackets (dense_map_utils.cc:1441-1660) ----
  std::vector<std::vector<std::pair<double, int>>> per_sensor(T);   // (timestamp, nvm image index)
  for (int i = 0; i < NI; i++) {
    std::string name; double stamp;
    split_image_path(nvm.image_files[i], name, stamp);
    int t = -1;
    for (int k = 0; k < T; k++) if (rig.sensors[k].name == name) t = k;
    if (t < 0) die("image " + nvm.image_files[i] + " belongs to no sensor of the rig");
    per_sensor[t].push_back({stamp, i});
  }
  std::vector<std::vector<double>> stamps(T);
  std::vector<double> offsets(T);
  for (int t = 0; t < T; t++) {
    std::sort(per_sensor[t].begin(), per_sensor[t].end());
    for (auto& p : per_sensor[t]) stamps[t].push_back(p.first);
    offsets[t] = rig.sensors[t].ref_to_sensor_timestamp_offset;
  }
  const std::vector<double>& ref_stamps = stamps[0];
  Bracketing br;
  try { br = lookup_images(std::atof(opt["bracket_len"].c_str()), std::atof(opt["timestamp_offsets_max_change"].c_str()), ref_stamps, stamps, {}, offsets); }
  catch (const std::exception& e) { die(e.what()); }
  const int C = (int)br.cams.size(), R = (int)ref_stamps.size();
  std::vector<int32_t> cam_sensor(C), cam_beg(C), cam_end(C);
  std::vector<double> cam_stamp(C);
  std::vector<int> cam_of_image(NI, -1);
  std::vector<std::string> cam_files(C);
  for (int c = 0; c < C; c++) {
    const BracketedImage& b = br.cams[c];
    cam_sensor[c] = b.sensor; cam_beg[c] = b.beg_ref_index; cam_end[c] = b.end_ref_index; cam_stamp[c] = b.timestamp;
    const int img = per_sensor[b.sensor][b.image_index].second;
    cam_of_image[img] = c; cam_files[c] = nvm.image_files[img];
  }

  // ---- parameter vectors (rig_calibrator.cc:1401-1492) ----
  std::vector<double> world_to_ref(7 * (size_t)R), ref_to_cam(7 * (size_t)T), focal(T), center(2 * (size_t)T), dist, d2i(7 * (size_t)T), scale(T, 1.0);
  for (int r = 0; r < R; r++) rigid_to_array(nvm.world_to_cam[per_sensor[0][r].second], &world_to_ref[7 * (size_t)r]);
  std::vector<double> img_size(2 * (size_t)T), crop_size(2 * (size_t)T);
  for (int t = 0; t < T; t++) {
    const SensorConfig& s = rig.sensors[t];
    rigid_to_array(s.ref_to_sensor, &ref_to_cam[7 * (size_t)t]);
    rigid_to_array(s.depth_to_image, &d2i[7 * (size_t)t]);
    focal[t] = s.focal_length; center[2 * t] = s.optical_center[0]; center[2 * t + 1] = s.optical_center[1];
    dist.insert(dist.end(), s.distortion.begin(), s.distortion.end());
    img_size[2 * t] = s.image_size[0]; img_size[2 * t + 1] = s.image_size[1];
    crop_size[2 * t] = s.distorted_crop_size[0]; crop_size[2 * t + 1] = s.distorted_crop_size[1];
  }

  // ---- features, track-major, cid ascending inside a track (the std::map order of pid_to_cid_fid); NVM keypoints are relative to the
  // optical centre (interest_point.cc:1289-1296) ----
  const int P = (int)nvm.pid_to_cid_fid.size();
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
  std::printf("{\"sensors\": %d, \"ref_frames\": %d, \"cameras\": %d, \"tracks\": %d, \"features\": %lld,\n \"cams\": [", T, R, C, P, (long long)N);
  for (int c = 0; c < C; c++) std::printf("%s[%d, %d, %d, %.17g]", c ? ", " : "", cam_sensor[c], cam_beg[c], cam_end[c], cam_stamp[c]);
  std::printf("],\n \"offset_bounds\": [");
  for (int t = 0; t < T; t++) std::printf("%s[%.17g, %.17g]", t ? ", " : "", br.min_timestamp_offset[t], br.max_timestamp_offset[t]);
  std::printf("],\n \"first_features\": [");
  for (int64_t i = 0; i < std::min<int64_t>(N, 6); i++) std::printf("%s[%d, %d, %.9g, %.9g]", i ? ", " : "", opid[i], ocid[i], pix[2 * i], pix[2 * i + 1]);
  std::printf("]}\n");
  if (dry) return 0;

  // ---- device side ----
  mvba_config cfg = {};
  cfg.num_sensors = T; cfg.num_ref_frames = R; cfg.num_cameras = C; cfg.num_landmarks = P;
  cfg.robust_threshold = std::atof(opt["robust_threshold"].c_str());
  mvba_params prm = {};
  prm.world_to_ref = world_to_ref.data(); prm.ref_stamps = ref_stamps.data(); prm.ref_to_cam = ref_to_cam.data(); prm.offsets = offsets.data();
  prm.offset_lo = br.min_timestamp_offset.data(); prm.offset_hi = br.max_timestamp_offset.data();
  prm.focal = focal.data(); prm.center = center.data(); prm.distortion = dist.data(); prm.depth_to_image = d2i.data(); prm.depth_scale = scale.data();
  prm.xyz = xyz.data();
  std::vector<uint8_t> f_none(T, 0), f_pose(T, 1), f_rig(T, opt.count("float_rig") ? 1 : 0);
  mvba_float_mask mask = {f_none.data(), f_none.data(), f_none.data(), f_rig.data(), f_none.data(), f_pose.data(), 0, 0};
  auto make = [&](int64_t n, const int32_t* pid, const int32_t* cid, const float* px) {
    mvba_problem* h = nullptr;
    CK(mvba_create(&cfg, 0, &h));
    for (int t = 0; t < T; t++) CK(mvba_set_sensor(h, t, img_size[2 * t], img_size[2 * t + 1], (int)rig.sensors[t].distortion.size()));
    CK(mvba_set_params(h, &prm));
    CK(mvba_set_cameras(h, C, cam_sensor.data(), cam_beg.data(), cam_end.data(), cam_stamp.data()));
    CK(mvba_set_observations(h, n, pid, cid, px, nullptr, nullptr, nullptr, nullptr));
    CK(mvba_set_float_mask(h, &mask));
    return h;
  };
  mvba_problem* h_all = make(N, opid.data(), ocid.data(), pix.data());   // all features: triangulation and outlier flags
  std::vector<uint8_t> inl((size_t)N, 1);
  CK(mvba_flag_exclusion(h_all, img_size.data(), crop_size.data(), inl.data()));
  const int passes = std::atoi(opt["num_passes"].c_str());
  for (int pass = 0; pass < passes; pass++) {
    CK(mvba_triangulate(h_all, inl.data(), xyz.data()));
    Observations o = compact_observations(N, opid.data(), ocid.data(), pix.data(), nullptr, nullptr, nullptr, nullptr, inl.data());
    mvba_problem* h = make((int64_t)o.pid.size(), o.pid.data(), o.cid.data(), o.pix.data());
    mvba_solve_options so; mvba_solve_options_init(&so);
    so.max_num_iterations = std::atoi(opt["num_iterations"].c_str()); so.verbose = 1;
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
      std::vector<std::string> names, cam_names; std::vector<double> scales;
      for (int t = 0; t < T; t++) cam_names.push_back(rig.sensors[t].name);
      const double w4[4] = {0, 0, 0, 0};
      residual_names_and_scales(nblk, kind.data(), row0.data(), src.data(), o.cid.data(), cam_sensor.data(), cam_names, w4, nres, names, scales);
      std::fputs(format_residual_stats(calc_residuals_stats(res, names, scales), tag).c_str(), stdout);
    };
    report("before opt");
    CK(mvba_solve(h, &so, &sum, nullptr, 0));
    report("after opt");
    std::printf("pass %d: %zu inlier features, cost %.6e -> %.6e in %d iterations\n", pass, o.pid.size(), sum.initial_cost, sum.final_cost, sum.num_iterations);
    CK(mvba_destroy(h));
    int64_t counts[4];
    CK(mvba_flag_outliers(h_all, std::atof(opt["min_triangulation_angle"].c_str()), std::atof(opt["max_reprojection_error"].c_str()), inl.data(), counts));
    std::printf("Removed %lld outlier features with small angle of convergence, out of %lld\nRemoved %lld outlier features using reprojection error, out of %lld\n",
                (long long)counts[1], (long long)counts[0], (long long)counts[3], (long long)counts[2]);
  }
  CK(mvba_destroy(h_all));

  // ---- RPC sensors: the undistortion half of the model is re-fitted from its distortion half after the passes
  //      (rig_calibrator.cc:1946-1948 -> CameraParameters::updateRpcUndistortion); the fit runs on the device (mvba_rpc_fit) ----
  if (opt.count("refit_rpc_undistortion")) {
    camera::CameraParameters::SetRpcFitter(&mvba_rpc_fit);
    for (int t = 0; t < T; t++) {
      SensorConfig& sc = rig.sensors[t];
      if (sc.distortion.size() <= 5) continue;
      try {
        camera::CameraParameters cp(camera::Vector2i(sc.image_size[0], sc.image_size[1]), camera::Vector2d(focal[t], focal[t]),
                                    camera::Vector2d(center[2 * t], center[2 * t + 1]), sc.distortion);
        cp.SetUndistortedSize(camera::Vector2i(sc.undistorted_image_size[0], sc.undistorted_image_size[1]));
        cp.SetDistortedCropSize(camera::Vector2i(sc.distorted_crop_size[0], sc.distorted_crop_size[1]));
        cp.updateRpcUndistortion(1);
        sc.distortion = cp.GetDistortion();
      } catch (const std::exception& e) { die(e.what()); }
    }
  }

  // ---- copy-back and outputs (rig_calibrator.cc:1924-1967; dense_map_utils.cc:753-850; interest_point.cc:1338-1475) ----
  for (int t = 0; t < T; t++) { rig.sensors[t].ref_to_sensor = array_to_rigid(&ref_to_cam[7 * (size_t)t]); rig.sensors[t].ref_to_sensor_timestamp_offset = offsets[t]; }
  std::vector<int> ib(cam_beg.begin(), cam_beg.end()), ie(cam_end.begin(), cam_end.end()), is(cam_sensor.begin(), cam_sensor.end());
  std::vector<Affine> w2c = world_to_cam_all(C, is.data(), ib.data(), ie.data(), cam_stamp.data(), ref_stamps.data(), world_to_ref.data(), ref_to_cam.data(), offsets.data());
  std::vector<CameraEntry> entries(C);
  for (int c = 0; c < C; c++) { entries[c].image_file = cam_files[c]; entries[c].world_to_cam = w2c[c]; }
  std::vector<std::array<double, 2>> ctr(T);
  for (int t = 0; t < T; t++) ctr[t] = {center[2 * t], center[2 * t + 1]};
  std::vector<int> vpid(opid.begin(), opid.end()), vcid(ocid.begin(), ocid.end());
  try {
    write_rig_config(opt["out_dir"] + "/rig_config.txt", rig, true);
    write_camera_list(opt["out_dir"] + "/cameras.txt", entries);
    write_nvm(opt["out_dir"] + "/cameras.nvm", nvm_from_inliers(cam_files, is, focal, ctr, w2c, vpid, vcid, pix, inl, xyz));
  } catch (const std::exception& e) { die(e.what()); }
  return 0;
}
