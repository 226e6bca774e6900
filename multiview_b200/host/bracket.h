// bracket.h -- which image of every sensor goes with which pair of reference frames: the host step that produces the
// (sensor, beg_ref_index, end_ref_index, timestamp) arrays mvba_set_cameras takes and the timestamp-offset bounds
// mvba_params.offset_lo / offset_hi carry.  Timestamps only; pixels and clouds stay with the caller (each result names the index of the
// chosen image / cloud in the caller's per-sensor lists).
//   lookupImage               reference rig_calibrator/src/dense_map_utils.cc:702-739 (first message at or after a time, scanning forward)
//   lookupImagesAndBrackets   dense_map_utils.cc:1441-1660
// For every consecutive pair of reference frames [beg, end) (the last frame pairs with itself) and every non-reference sensor, the
// image whose timestamp, in that sensor's clock, falls in [ref[beg] + offset, ref[end] + offset) and lies closest to the middle of
// the bracket is kept, provided the bracket is not longer than bracket_len (unless the image sits exactly on its beginning).
#pragma once
#include <algorithm>
#include <cmath>
#include <stdexcept>
#include <vector>

namespace mvba_host {

// First entry at or after `desired` scanning forward from start_pos (which is left at the last entry looked at); found_time is that
// entry's time even when the scan fails, as in the reference.
inline bool lookup_time(double desired, const std::vector<double>& stamps, int& start_pos, double& found_time) {
  found_time = -1.0;
  double prev = -1.0;
  for (int pos = start_pos; pos < (int)stamps.size(); pos++) {
    start_pos = pos;
    found_time = stamps[pos];
    if (found_time < prev) throw std::runtime_error("Found images not in chronological order.");
    prev = found_time;
    if (found_time >= desired) return true;
  }
  return false;
}

struct BracketedImage {
  int sensor = 0;
  double timestamp = 0;       // in the sensor's own clock
  double ref_timestamp = 0;   // timestamp - ref_to_cam_timestamp_offset
  int beg_ref_index = 0, end_ref_index = 0;
  int image_index = -1;       // position in image_stamps[sensor]
  int cloud_index = -1;       // position in cloud_stamps[sensor] of the first cloud at or after the image (-1: none looked up)
  double cloud_timestamp = -1.0;
};

struct Bracketing {
  std::vector<BracketedImage> cams;
  std::vector<double> min_timestamp_offset, max_timestamp_offset;  // per sensor; (-1e100, 1e100) for the reference sensor
};

inline Bracketing lookup_images_and_brackets(double bracket_len, double timestamp_offsets_max_change, const std::vector<double>& ref_timestamps,
                                             const std::vector<std::vector<double>>& image_stamps,
                                             const std::vector<std::vector<double>>& cloud_stamps,  // empty: no depth data
                                             const std::vector<double>& ref_to_cam_timestamp_offsets) {
  const int R = (int)ref_timestamps.size(), T = (int)image_stamps.size();
  Bracketing out;
  out.min_timestamp_offset.assign(T, -1.0e+100);
  out.max_timestamp_offset.assign(T, 1.0e+100);
  std::vector<int> image_start(T, 0), cloud_start(T, 0);
  for (int beg = 0; beg < R; beg++) {
    int end = beg + 1;
    const bool last = end == R;
    if (last) end = beg;
    for (int t = 0; t < T; t++) {
      BracketedImage cam;
      if (t == 0) {
        cam.sensor = 0; cam.timestamp = ref_timestamps[beg]; cam.ref_timestamp = cam.timestamp; cam.beg_ref_index = beg; cam.end_ref_index = beg;
        double found = -1.0;
        if (!lookup_time(cam.timestamp, image_stamps[0], image_start[0], found) || found != cam.timestamp)
          throw std::runtime_error("Cannot look up camera at time " + std::to_string(cam.timestamp) + ".");
        cam.image_index = image_start[0];
      } else {
        const double off = ref_to_cam_timestamp_offsets[t];
        const double beg_ts = ref_timestamps[beg] + off;
        double end_ts = ref_timestamps[end] + off;
        if (end_ts == beg_ts && last) end_ts = std::nextafter(end_ts, end_ts + 1.0);
        if (end_ts <= beg_ts) throw std::runtime_error("Ref timestamps must be in strictly increasing order.");
        const double mid = (beg_ts + end_ts) / 2.0;
        int pos = image_start[t], best_index = -1;
        double cur = beg_ts, best_dist = 1.0e+100, best_time = -1.0, found = -1.0;
        while (true) {
          if (found > end_ts) break;
          if (!lookup_time(cur, image_stamps[t], pos, found)) break;
          const bool inside = beg_ts <= found && found < end_ts;
          const double dist = std::abs(found - mid);
          if (dist < best_dist && inside) { best_dist = dist; best_time = found; image_start[t] = pos; best_index = pos; }
          cur = std::nextafter(found, found + 1.0);
        }
        if (best_time < 0.0) continue;                                           // nothing in this bracket
        if (best_time > beg_ts && end_ts - beg_ts > bracket_len) continue;       // bracket too long
        if (best_time < beg_ts || best_time >= end_ts) continue;
        cam.sensor = t; cam.timestamp = best_time; cam.ref_timestamp = best_time - off; cam.beg_ref_index = beg; cam.end_ref_index = end;
        cam.image_index = best_index;
        out.min_timestamp_offset[t] = std::max(out.min_timestamp_offset[t], cam.timestamp - ref_timestamps[end]);
        out.max_timestamp_offset[t] = std::min(out.max_timestamp_offset[t], cam.timestamp - ref_timestamps[beg]);
      }
      if (!cloud_stamps.empty()) {
        const bool ok = lookup_time(cam.timestamp, cloud_stamps[t], cloud_start[t], cam.cloud_timestamp);
        cam.cloud_index = ok ? cloud_start[t] : -1;
      }
      out.cams.push_back(cam);
    }
  }
  for (int t = 1; t < T; t++) {
    out.min_timestamp_offset[t] = std::max(out.min_timestamp_offset[t], ref_to_cam_timestamp_offsets[t] - timestamp_offsets_max_change) + 1.0e-5;
    out.max_timestamp_offset[t] = std::min(out.max_timestamp_offset[t], ref_to_cam_timestamp_offsets[t] + timestamp_offsets_max_change) - 1.0e-5;
  }
  return out;
}

// lookupImagesNoBrackets (dense_map_utils.cc:1664-1743), the --no_rig case: every image of every sensor is a camera of its own, beg = end =
// its position in the sensor's list, no timestamp offset, no offset bounds.
inline Bracketing lookup_images_no_brackets(const std::vector<std::vector<double>>& image_stamps, const std::vector<std::vector<double>>& cloud_stamps) {
  const int T = (int)image_stamps.size();
  Bracketing out;
  out.min_timestamp_offset.assign(T, -1.0e+100);
  out.max_timestamp_offset.assign(T, 1.0e+100);
  std::vector<int> image_start(T, 0), cloud_start(T, 0);
  for (int t = 0; t < T; t++)
    for (int k = 0; k < (int)image_stamps[t].size(); k++) {
      BracketedImage cam;
      cam.sensor = t; cam.timestamp = image_stamps[t][k]; cam.ref_timestamp = cam.timestamp; cam.beg_ref_index = k; cam.end_ref_index = k;
      double found = -1.0;
      if (!lookup_time(cam.timestamp, image_stamps[t], image_start[t], found) || found != cam.timestamp)
        throw std::runtime_error("Cannot look up camera at time " + std::to_string(cam.timestamp) + ".");
      cam.image_index = image_start[t];
      if (!cloud_stamps.empty()) {
        const bool ok = lookup_time(cam.timestamp, cloud_stamps[t], cloud_start[t], cam.cloud_timestamp);
        cam.cloud_index = ok ? cloud_start[t] : -1;
      }
      out.cams.push_back(cam);
    }
  return out;
}

// lookupImages (dense_map_utils.cc:1745-1805) for the rig case: brackets, at least one image per sensor, cameras in order of their
// reference-clock time (the reference calls std::sort there; a stable sort is used so that equal times keep the bracket order).
inline Bracketing lookup_images(double bracket_len, double timestamp_offsets_max_change, const std::vector<double>& ref_timestamps,
                                const std::vector<std::vector<double>>& image_stamps, const std::vector<std::vector<double>>& cloud_stamps,
                                const std::vector<double>& ref_to_cam_timestamp_offsets) {
  Bracketing b = lookup_images_and_brackets(bracket_len, timestamp_offsets_max_change, ref_timestamps, image_stamps, cloud_stamps, ref_to_cam_timestamp_offsets);
  std::vector<int> count(image_stamps.size(), 0);
  for (const BracketedImage& c : b.cams) count[c.sensor]++;
  for (int n : count) if (n == 0) throw std::runtime_error("Could not find images for all sensors. Cannot continue.");
  std::stable_sort(b.cams.begin(), b.cams.end(), [](const BracketedImage& x, const BracketedImage& y) { return x.ref_timestamp < y.ref_timestamp; });
  return b;
}

}  // namespace mvba_host
