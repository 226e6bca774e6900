// residual_stats.h -- the residual report rig_calibrator prints before and after every optimisation pass.
//   residual names / scales   reference rig_calibrator/bin/rig_calibrator.cc:1691-1694, 1769-1774, 1818-1823, 1877-1882,
//                             1895-1900 (one name and one scale per residual ROW, in block order)
//   evalResiduals             :1157-1180  raw residuals (no loss), each divided by its scale (the block's weight)
//   calc_residuals_stats      :753-789    per name: sorted |residual|, entries [0.25 n], [0.5 n], [0.75 n], [n - 1]
// The residual vector comes from mvba_evaluate(apply_loss = 0) and the block list from mvba_residual_layout
// (include/mvba.h), both in the reference's row order; nothing here touches the device.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <map>
#include <sstream>
#include <iomanip>
#include <stdexcept>
#include <string>
#include <vector>

namespace mvba_host {

struct ResidualStat { std::string name; double q25 = 0, q50 = 0, q75 = 0, q100 = 0; int64_t count = 0; };

// kind / row0 / src: mvba_residual_layout (kind 0 CAM .. 4 TRI, first row, index of the source observation or prior).
// cid: image of every observation; cam_sensor: sensor of every image; weights: depth_tri, depth_mesh, mesh_tri, tri.
inline void residual_names_and_scales(int64_t num_blocks, const int32_t* kind, const int64_t* row0, const int64_t* src, const int32_t* cid,
                                      const int32_t* cam_sensor, const std::vector<std::string>& cam_names, const double weights[4],
                                      int64_t num_residuals, std::vector<std::string>& names, std::vector<double>& scales) {
  names.assign((size_t)num_residuals, std::string());
  scales.assign((size_t)num_residuals, 1.0);
  static const char* base[5] = {"", "depth_tri", "depth_mesh", "mesh_tri", "tri"};
  static const char* axis[3] = {"_x_m", "_y_m", "_z_m"};
  for (int64_t b = 0; b < num_blocks; b++) {
    const int k = kind[b];
    const int64_t r = row0[b];
    if (k == 0) {
      const std::string& cam = cam_names.at((size_t)cam_sensor[cid[src[b]]]);
      if (r + 2 > num_residuals) throw std::runtime_error("residual layout does not fit the residual vector");
      names[(size_t)r] = cam + "_pix_x"; names[(size_t)r + 1] = cam + "_pix_y";
    } else if (k >= 1 && k <= 4) {
      if (r + 3 > num_residuals) throw std::runtime_error("residual layout does not fit the residual vector");
      for (int a = 0; a < 3; a++) { names[(size_t)r + a] = std::string(base[k]) + axis[a]; scales[(size_t)r + a] = weights[k - 1]; }
    } else {
      throw std::runtime_error("unknown residual block kind");
    }
  }
}

// calc_residuals_stats on residuals / scales; the result is ordered by name like the reference's std::map.
inline std::vector<ResidualStat> calc_residuals_stats(const std::vector<double>& residuals, const std::vector<std::string>& names,
                                                      const std::vector<double>& scales) {
  if (residuals.size() != names.size()) throw std::runtime_error("There must be as many residuals as residual names.");
  if (residuals.size() != scales.size()) throw std::runtime_error("There must be as many residual values as residual scales.");
  std::map<std::string, std::vector<double>> stats;
  for (size_t i = 0; i < residuals.size(); i++) stats[names[i]].push_back(std::fabs(residuals[i] / scales[i]));
  std::vector<ResidualStat> out;
  for (auto& kv : stats) {
    std::vector<double>& vals = kv.second;
    std::sort(vals.begin(), vals.end());
    const int len = (int)vals.size();
    ResidualStat s;
    s.name = kv.first; s.count = len;
    if (len > 0) {
      s.q25 = vals[(size_t)(int)(0.25 * len)]; s.q50 = vals[(size_t)(int)(0.50 * len)];
      s.q75 = vals[(size_t)(int)(0.75 * len)]; s.q100 = vals[(size_t)(len - 1)];
    }
    out.push_back(s);
  }
  return out;
}

// The text the reference writes to stdout (rig_calibrator.cc:768-788).
inline std::string format_residual_stats(const std::vector<ResidualStat>& stats, const std::string& tag) {
  std::ostringstream os;
  os << "The 25, 50, 75, and 100th percentile residual stats " << tag << std::endl;
  for (const ResidualStat& s : stats) {
    if (s.count == 0) os << s.name << ": " << "none";
    else os << std::setprecision(5) << s.name << ": " << s.q25 << ' ' << s.q50 << ' ' << s.q75 << ' ' << s.q100;
    os << " (" << s.count << " residuals)" << std::endl;
  }
  return os.str();
}

}  // namespace mvba_host
