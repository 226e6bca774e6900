// cli_flags.h -- the command line of the reference's rig_calibrator, without gflags / glog.
//   * the 43 flags of rig_calibrator/bin/rig_calibrator.cc:135-316, same names, types and default values, in that order;
//   * gflags' syntax: --flag=value, --flag value, -flag, booleans as --flag / --noflag / --flag=true|false;
//   * parameterValidation (rig_calibrator.cc:893-955): same tests in the same order with the same messages, raised as
//     std::runtime_error where the reference's LOG(FATAL) aborts;
//   * parse_intrinsics_to_float / parse_camera_names / parse_rig_transforms_to_float (dense_map_utils.cc:54-152).
// Flags whose work lies outside the bundle-adjustment path (feature matching, registration, texturing, voxblox, depth-cloud
// export: SURVEY.md section 2 "out of scope") are parsed and validated like the others; out_of_scope() names the ones a
// command line actually uses so that the driver can refuse them instead of ignoring them.
#pragma once
#include <cstdlib>
#include <cstring>
#include <map>
#include <set>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace mvba_host {

struct Flags {
  // rig_calibrator.cc:135-316, in declaration order
  double robust_threshold = 3.0;
  int num_iterations = 20;
  double bracket_len = 0.6;
  std::string intrinsics_to_float, rig_transforms_to_float, camera_poses_to_float, depth_to_image_transforms_to_float;
  bool float_scale = false, float_timestamp_offsets = false;
  double timestamp_offsets_max_change = 1.0;
  double depth_tri_weight = 0.0;
  std::string mesh;
  double mesh_tri_weight = 0.0, depth_mesh_weight = 0.0, tri_weight = 0.0, tri_robust_threshold = 0.0;
  bool affine_depth_to_image = false;
  int calibrator_num_passes = 2;
  double initial_max_reprojection_error = 300.0, max_reprojection_error = 25.0, min_triangulation_angle = 0.5;
  std::string out_texture_dir;
  double min_ray_dist = 0.0, max_ray_dist = 100.0;
  bool registration = false;
  std::string hugin_file, xyz_file;
  double parameter_tolerance = 1e-12;
  int num_opt_threads = 16, num_match_threads = 8;
  bool no_rig = false;
  std::string out_dir, rig_config, nvm;
  bool no_nvm_matches = false;
  std::string camera_poses;
  int num_overlaps = 0;
  bool use_initial_rig_transforms = false, save_nvm = false, save_matches = false, export_to_voxblox = false,
       save_transformed_depth_clouds = false, verbose = false;
};

struct FlagSpec {
  const char* name;
  char type;      // 'd' double, 'i' int32, 's' string, 'b' bool
  size_t offset;  // into Flags
};

#define MVBA_FLAG(n, t) {#n, t, offsetof(Flags, n)}
inline const std::vector<FlagSpec>& flag_table() {
  static const std::vector<FlagSpec> tab = {
      MVBA_FLAG(robust_threshold, 'd'), MVBA_FLAG(num_iterations, 'i'), MVBA_FLAG(bracket_len, 'd'), MVBA_FLAG(intrinsics_to_float, 's'),
      MVBA_FLAG(rig_transforms_to_float, 's'), MVBA_FLAG(camera_poses_to_float, 's'), MVBA_FLAG(depth_to_image_transforms_to_float, 's'),
      MVBA_FLAG(float_scale, 'b'), MVBA_FLAG(float_timestamp_offsets, 'b'), MVBA_FLAG(timestamp_offsets_max_change, 'd'),
      MVBA_FLAG(depth_tri_weight, 'd'), MVBA_FLAG(mesh, 's'), MVBA_FLAG(mesh_tri_weight, 'd'), MVBA_FLAG(depth_mesh_weight, 'd'),
      MVBA_FLAG(tri_weight, 'd'), MVBA_FLAG(tri_robust_threshold, 'd'), MVBA_FLAG(affine_depth_to_image, 'b'),
      MVBA_FLAG(calibrator_num_passes, 'i'), MVBA_FLAG(initial_max_reprojection_error, 'd'), MVBA_FLAG(max_reprojection_error, 'd'),
      MVBA_FLAG(min_triangulation_angle, 'd'), MVBA_FLAG(out_texture_dir, 's'), MVBA_FLAG(min_ray_dist, 'd'), MVBA_FLAG(max_ray_dist, 'd'),
      MVBA_FLAG(registration, 'b'), MVBA_FLAG(hugin_file, 's'), MVBA_FLAG(xyz_file, 's'), MVBA_FLAG(parameter_tolerance, 'd'),
      MVBA_FLAG(num_opt_threads, 'i'), MVBA_FLAG(num_match_threads, 'i'), MVBA_FLAG(no_rig, 'b'), MVBA_FLAG(out_dir, 's'),
      MVBA_FLAG(rig_config, 's'), MVBA_FLAG(nvm, 's'), MVBA_FLAG(no_nvm_matches, 'b'), MVBA_FLAG(camera_poses, 's'),
      MVBA_FLAG(num_overlaps, 'i'), MVBA_FLAG(use_initial_rig_transforms, 'b'), MVBA_FLAG(save_nvm, 'b'), MVBA_FLAG(save_matches, 'b'),
      MVBA_FLAG(export_to_voxblox, 'b'), MVBA_FLAG(save_transformed_depth_clouds, 'b'), MVBA_FLAG(verbose, 'b')};
  return tab;
}
#undef MVBA_FLAG

inline bool parse_bool_text(const std::string& v, bool& out) {  // gflags: true/false, t/f, yes/no, y/n, 1/0 (case-insensitive)
  std::string s;
  for (char c : v) s += (char)std::tolower((unsigned char)c);
  if (s == "1" || s == "t" || s == "true" || s == "y" || s == "yes") { out = true; return true; }
  if (s == "0" || s == "f" || s == "false" || s == "n" || s == "no") { out = false; return true; }
  return false;
}

inline void set_flag(Flags& f, const FlagSpec& s, const std::string& value) {
  char* base = reinterpret_cast<char*>(&f) + s.offset;
  char* end = nullptr;
  const std::string bad = std::string("ERROR: illegal value '") + value + "' specified for " +
                          (s.type == 'd' ? "double" : s.type == 'i' ? "int32" : s.type == 'b' ? "bool" : "string") + " flag '" + s.name + "'";
  switch (s.type) {
    case 'd': { double v = std::strtod(value.c_str(), &end); if (value.empty() || *end) throw std::runtime_error(bad); *reinterpret_cast<double*>(base) = v; break; }
    case 'i': { long v = std::strtol(value.c_str(), &end, 0); if (value.empty() || *end || v != (int)v) throw std::runtime_error(bad); *reinterpret_cast<int*>(base) = (int)v; break; }
    case 'b': { bool v; if (!parse_bool_text(value, v)) throw std::runtime_error(bad); *reinterpret_cast<bool*>(base) = v; break; }
    default: *reinterpret_cast<std::string*>(base) = value;
  }
}

// google::ParseCommandLineFlags(&argc, &argv, true).  `extra` lists additional boolean switches of the caller (this repository's
// driver has --dry_run); the ones present are returned.  Non-flag arguments are returned in `positional`.
inline std::set<std::string> parse_command_line(int argc, char** argv, Flags& f, const std::set<std::string>& extra = {},
                                                std::vector<std::string>* positional = nullptr) {
  std::set<std::string> seen_extra;
  const auto& tab = flag_table();
  auto find = [&](const std::string& n) -> const FlagSpec* { for (auto& s : tab) if (n == s.name) return &s; return nullptr; };
  for (int i = 1; i < argc; i++) {
    std::string a = argv[i];
    if (a == "--") { for (i++; i < argc; i++) if (positional) positional->push_back(argv[i]); break; }
    if (a.size() < 2 || a[0] != '-') { if (positional) positional->push_back(a); continue; }
    a = a.substr(a[1] == '-' ? 2 : 1);
    std::string name = a, value; bool has_value = false;
    size_t eq = a.find('=');
    if (eq != std::string::npos) { name = a.substr(0, eq); value = a.substr(eq + 1); has_value = true; }
    if (extra.count(name)) { seen_extra.insert(name); continue; }
    const FlagSpec* s = find(name);
    if (!s && !has_value && name.rfind("no", 0) == 0) {   // --noflag for a boolean
      const FlagSpec* b = find(name.substr(2));
      if (b && b->type == 'b') { set_flag(f, *b, "false"); continue; }
    }
    if (!s) throw std::runtime_error("ERROR: unknown command line flag '" + name + "'");
    if (s->type == 'b') { set_flag(f, *s, has_value ? value : "true"); continue; }
    if (!has_value) {
      if (i + 1 >= argc) throw std::runtime_error("ERROR: flag '" + name + "' is missing its argument");
      value = argv[++i];
    }
    set_flag(f, *s, value);
  }
  return seen_extra;
}

// parameterValidation(), rig_calibrator.cc:893-955: same order, same text.
inline void parameter_validation(const Flags& F) {
  auto fatal = [](const char* m) { throw std::runtime_error(m); };
  if (F.robust_threshold <= 0.0) fatal("The robust threshold must be positive.\n");
  if (F.bracket_len <= 0.0) fatal("Bracket length must be positive.");
  if (F.num_overlaps < 1 && (F.nvm == "" || F.no_nvm_matches))
    fatal("No nvm file was specified or it is not desired to read its matches. "
          "Then must set a positive --num_overlaps to be able to find new interest point matches.");
  if (F.timestamp_offsets_max_change < 0) fatal("The timestamp offsets must be non-negative.");
  if (F.min_triangulation_angle <= 0.0) fatal("The min triangulation angle must be positive.\n");
  if (F.depth_tri_weight < 0.0) fatal("The depth weight must non-negative.\n");
  if (F.mesh_tri_weight < 0.0) fatal("The mesh weight must non-negative.\n");
  if (F.depth_mesh_weight < 0.0) fatal("The depth mesh weight must non-negative.\n");
  if (F.tri_weight < 0.0) fatal("The triangulation weight must non-negative.\n");
  if (F.tri_weight > 0.0 && F.tri_robust_threshold <= 0.0) fatal("The triangulation robust threshold must be positive.\n");
  if (F.registration && (F.xyz_file.empty() || F.hugin_file.empty()))
    fatal("In order to register the map, the hugin and xyz file must be specified.");
  if (F.float_scale && F.affine_depth_to_image)
    fatal("The options --float_scale and --affine_depth_to_image should not be used "
          "together. If the latter is used, the scale is always floated.\n");
  if (F.no_rig && F.float_timestamp_offsets) fatal("Cannot float timestamps with option --no_rig.\n");
  if (F.out_dir == "") fatal("The output directory was not specified.\n");
  if (F.use_initial_rig_transforms && F.no_rig) fatal("Cannot use initial rig transforms if not modeling the rig.\n");
  if (F.out_texture_dir != "" && F.mesh == "") fatal("Cannot project camera images onto a mesh if a mesh was not provided.\n");
  if (F.rig_config == "") fatal("Must specify the initial rig configuration via --rig_config.\n");
  if (F.camera_poses != "" && F.nvm != "") fatal("Cannot specify both --nvm and --camera_poses.\n");
  if (F.camera_poses == "" && F.nvm == "") fatal("Must specify the cameras via --nvm or --camera_poses.\n");
}

// Flags of this command line that ask for work outside the bundle-adjustment path; empty when the driver can honour all of it.
inline std::vector<std::string> out_of_scope(const Flags& F) {
  std::vector<std::string> v;
  if (F.num_overlaps > 0) v.push_back("--num_overlaps (feature detection and matching)");
  if (F.no_nvm_matches) v.push_back("--no_nvm_matches (feature detection and matching)");
  if (F.camera_poses != "") v.push_back("--camera_poses (a camera list has no matches; needs --num_overlaps)");
  if (F.registration) v.push_back("--registration (control-point registration)");
  if (F.out_texture_dir != "") v.push_back("--out_texture_dir (texturing)");
  if (F.save_matches) v.push_back("--save_matches (match files)");
  if (F.export_to_voxblox) v.push_back("--export_to_voxblox");
  if (F.save_transformed_depth_clouds) v.push_back("--save_transformed_depth_clouds");
  return v;
}

inline void replace_separators_with_space(std::string& s) {  // dense_map_utils.cc:42-46: backslash, colon, comma and white space
  for (char& c : s) if (c == '\\' || c == ':' || c == ',' || c == '\t' || c == '\r' || c == '\n') c = ' ';
}

// dense_map_utils.cc:54-98
inline std::vector<std::set<std::string>> parse_intrinsics_to_float(const std::string& str, const std::vector<std::string>& cam_names) {
  std::string input = str;
  replace_separators_with_space(input);
  std::istringstream iss(input);
  std::string curr_cam, val;
  std::map<std::string, std::set<std::string>> local;
  while (iss >> val) {
    bool is_cam = false;
    for (auto& n : cam_names) if (val == n) { curr_cam = val; is_cam = true; break; }
    if (is_cam) continue;
    if (val != "focal_length" && val != "optical_center" && val != "distortion")
      throw std::runtime_error("Unexpected value when parsing intrinsics to float: " + val + "\n");
    if (curr_cam == "") throw std::runtime_error("Incorrectly set option for floating intrinsics.\n");
    local[curr_cam].insert(val);
  }
  std::vector<std::set<std::string>> out(cam_names.size());
  for (size_t t = 0; t < cam_names.size(); t++) out[t] = local[cam_names[t]];
  return out;
}

// dense_map_utils.cc:102-134
inline std::set<std::string> parse_camera_names(const std::vector<std::string>& cam_names, const std::string& str) {
  std::string input = str;
  replace_separators_with_space(input);
  std::istringstream iss(input);
  std::string val;
  std::set<std::string> out;
  while (iss >> val) {
    bool known = false;
    for (auto& n : cam_names) if (val == n) { known = true; break; }
    if (!known) throw std::runtime_error("Error: A specified sensor name is not among the known sensors. Offending camera: " + val + "\n");
    out.insert(val);
  }
  return out;
}

// dense_map_utils.cc:140-152
inline std::set<std::string> parse_rig_transforms_to_float(const std::vector<std::string>& cam_names, int ref_cam_type, const std::string& str) {
  std::set<std::string> out = parse_camera_names(cam_names, str);
  for (auto& n : out) if (n == cam_names[ref_cam_type]) throw std::runtime_error("Cannot float the rig transform from reference camera to itself.\n");
  return out;
}

}  // namespace mvba_host
