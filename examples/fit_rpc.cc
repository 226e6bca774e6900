// examples/fit_rpc.cc -- the reference's fit_rpc tool (rig_calibrator/bin/fit_rpc.cc:59-146) over this repository: replace the lens
// distortion model of the FIRST sensor of a rig configuration by an RPC model of a given degree fitted to it.  Same flags, defaults
// and messages as the reference (--rpc_degree, --num_samples, --num_iterations 20, --num_opt_threads 16, --parameter_tolerance 1e-12,
// --camera_config, --out_dir, --verbose); same steps: readRigConfig, fitRpcDist (rpc_distortion.cc:622-656: num_samples^2 undistorted /
// distorted pixel pairs inside the crop box, degree continuation from the identity model), fitRpcUndist (:658-719: the same undistorted
// samples pushed through the fitted distortion, fitted the other way round), evalRpcDistUndist (the two residual reports),
// writeRigConfig.  The fits run on the device through mvba_rpc_fit (include/mvba.h); there is no CPU fallback.
// Exit codes as in rig_calibrator.cc: 1 command-line syntax, 2 a failed validation / input error, 3 a failed device call.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <filesystem>
#include <string>
#include <vector>

#include "../include/mvba.h"
#include "../multiview_b200/host/rpc_pairs.h"
#include "../multiview_b200/host/rig_io.h"

using namespace mvba_host;

static void die(const std::string& m, int code = 2) { std::fprintf(stderr, "fit_rpc: %s\n", m.c_str()); std::exit(code); }

int main(int argc, char** argv) {
  int rpc_degree = -1, num_samples = -1, num_iterations = 20, num_opt_threads = 16;
  double parameter_tolerance = 1e-12;
  std::string camera_config, out_dir;
  bool verbose = false;
  for (int i = 1; i < argc; i++) {   // gflags syntax: --flag=value, --flag value, -flag, --noverbose
    std::string a = argv[i];
    if (a.size() < 2 || a[0] != '-') die("unexpected argument " + a, 1);
    a = a.substr(a[1] == '-' ? 2 : 1);
    std::string name = a, value; bool has = false;
    const size_t eq = a.find('=');
    if (eq != std::string::npos) { name = a.substr(0, eq); value = a.substr(eq + 1); has = true; }
    if (name == "verbose") { verbose = !has || (value != "false" && value != "0"); continue; }
    if (name == "noverbose") { verbose = false; continue; }
    if (!has) { if (i + 1 >= argc) die("ERROR: flag '" + name + "' is missing its argument", 1); value = argv[++i]; }
    char* end = nullptr;
    auto bad = [&]() { die("ERROR: illegal value '" + value + "' specified for flag '" + name + "'", 1); };
    if (name == "rpc_degree") { rpc_degree = (int)std::strtol(value.c_str(), &end, 0); if (*end || value.empty()) bad(); }
    else if (name == "num_samples") { num_samples = (int)std::strtol(value.c_str(), &end, 0); if (*end || value.empty()) bad(); }
    else if (name == "num_iterations") { num_iterations = (int)std::strtol(value.c_str(), &end, 0); if (*end || value.empty()) bad(); }
    else if (name == "num_opt_threads") { num_opt_threads = (int)std::strtol(value.c_str(), &end, 0); if (*end || value.empty()) bad(); }
    else if (name == "parameter_tolerance") { parameter_tolerance = std::strtod(value.c_str(), &end); if (*end || value.empty()) bad(); }
    else if (name == "camera_config") camera_config = value;
    else if (name == "out_dir") out_dir = value;
    else die("ERROR: unknown command line flag '" + name + "'", 1);
  }
  (void)num_opt_threads; (void)verbose;
  if (camera_config.empty()) die("Camera config file was not specified.");
  if (out_dir.empty()) die("Output camera config directory was not specified.");
  if (rpc_degree <= 0) die("The RPC degree must be positive.");
  if (num_samples <= 0) die("The number of samples must be positive.");
  if (rpc_degree > 6) die("RPC degrees above 6 are not supported by mvba_rpc_fit.");

  RigConfig rig;
  try { rig = read_rig_config(camera_config, true); } catch (const std::exception& e) { die(e.what()); }
  if (rig.sensors.empty()) die("No sensors in " + camera_config);
  SensorConfig& sc = rig.sensors[0];
  std::printf("Focal length is %.17g %.17g\n", sc.focal_length, sc.focal_length);
  std::vector<double> undist, dist;
  try {
    camera::CameraParameters cam(camera::Vector2i(sc.image_size[0], sc.image_size[1]), camera::Vector2d(sc.focal_length, sc.focal_length),
                                 camera::Vector2d(sc.optical_center[0], sc.optical_center[1]), sc.distortion);
    cam.SetUndistortedSize(camera::Vector2i(sc.undistorted_image_size[0], sc.undistorted_image_size[1]));
    cam.SetDistortedCropSize(camera::Vector2i(sc.distorted_crop_size[0], sc.distorted_crop_size[1]));
    gen_undist_dist_pairs(num_samples, cam, &undist, &dist);
  } catch (const std::exception& e) { die(e.what()); }
  const int64_t n = (int64_t)undist.size() / 2;
  std::printf("Found %lld pixel correspondences between undistorted and distorted images within set bounds.\n", (long long)n);
  if (n == 0) die("No pixel correspondences: check the image sizes and the crop box.");
  const int half = camera::rpc_num_dist_params(rpc_degree);
  std::vector<double> cd((size_t)half), cu((size_t)half), dist_rpc(dist.size());
  double stats[4];
  std::printf("\nComputing RPC distortion.\n");
  if (mvba_rpc_fit(0, n, undist.data(), dist.data(), rpc_degree, num_iterations, parameter_tolerance, 0, cd.data(), stats) != MVBA_OK)
    die(std::string("mvba_rpc_fit failed: ") + mvba_last_error(), 3);
  std::printf("RPC distortion fit: cost %.6e -> %.6e in %d iterations\n", stats[0], stats[1], (int)stats[2]);
  for (int64_t i = 0; i < n; i++) {   // fitRpcUndist: the undistorted samples through the FITTED distortion
    const camera::Vector2d d = camera::compute_rpc(camera::Vector2d(undist[2 * i], undist[2 * i + 1]), cd.data(), half);
    dist_rpc[2 * i] = d[0]; dist_rpc[2 * i + 1] = d[1];
  }
  std::printf("\nComputing RPC undistortion.\n");
  if (mvba_rpc_fit(0, n, dist_rpc.data(), undist.data(), rpc_degree, num_iterations, parameter_tolerance, 0, cu.data(), stats) != MVBA_OK)
    die(std::string("mvba_rpc_fit failed: ") + mvba_last_error(), 3);
  std::printf("RPC undistortion fit: cost %.6e -> %.6e in %d iterations\n", stats[0], stats[1], (int)stats[2]);
  // evalRpcDistUndist (rpc_distortion.cc:730-790): the fitted distortion against the model it replaces, and the round trip
  double e_dist = 0, e_round = 0;
  for (int64_t i = 0; i < n; i++) {
    const camera::Vector2d u(undist[2 * i], undist[2 * i + 1]);
    const camera::Vector2d d = camera::compute_rpc(u, cd.data(), half), b = camera::compute_rpc(d, cu.data(), half);
    e_dist = std::max(e_dist, std::hypot(d[0] - dist[2 * i], d[1] - dist[2 * i + 1]));
    e_round = std::max(e_round, std::hypot(b[0] - u[0], b[1] - u[1]));
  }
  std::printf("Max error of the RPC distortion against the input model: %.6g\nMax distort_undistort error: %.6g\n", e_dist, e_round);
  sc.distortion = cd;
  sc.distortion.insert(sc.distortion.end(), cu.begin(), cu.end());
  try {
    std::filesystem::create_directories(out_dir);
    write_rig_config(out_dir + "/rig_config.txt", rig, true);
  } catch (const std::exception& e) { die(e.what()); }
  return 0;
}
