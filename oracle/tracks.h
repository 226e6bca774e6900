// oracle/tracks.h -- CPU restatement (TEST INFRASTRUCTURE ONLY, never linked into the product) of the per-track steps
// that run around every rig_calibrator pass:
//   UndistortCentered            camera_model/camera_params.cc:316-355   (cv::undistortPoints of OpenCV 3.2 restated:
//                                 5 fixed-point iterations, modules/imgproc/src/undistort.cpp cvUndistortPoints)
//   Triangulation::compute       openMVG/multiview/triangulation_nview.cc:87-146 (iter = 3, triangulation_nview.hpp:67)
//   Triangulate                  rig_calibrator/src/interest_point.cc:398-423
//   multiViewTriangulation       interest_point.cc:649-722
//   flagOutlierByExclusionDist   rig_calibrator/bin/rig_calibrator.cc:1003-1038
//   flagOutliersByTriAngleAndReprojErr  rig_calibrator.cc:1045-1154
// Included by oracle.cc (needs its Problem / RBlock types).  Triangulation / undistortion / flags: parity unpinned by the reference
// (no tests or fixtures there); pinned against cv2.undistortPoints and a noise-free recovery property in tests/.  The ray /
// triangle test IS pinned on the reference: its header-only rayint code compiles here (oracle/ref_build -> oracle/_ref), and
// tests/test_rayint_pin.py holds ray_tri to it bit for bit (live library + tests/golden/rayint.npz).
#pragma once

namespace oracle {

inline void undistort_centered(const SensorConst& s, double f, const double c[2], const double* dist, const double dc[2], double u[2]) {
  if (s.nd == 0) {
    u[0] = dc[0] - (c[0] - s.half[0]); u[1] = dc[1] - (c[1] - s.half[1]);
  } else if (s.nd == 1) {
    double nx = (dc[0] - (c[0] - s.half[0])) / f, ny = (dc[1] - (c[1] - s.half[1])) / f;
    double rd = std::sqrt(nx * nx + ny * ny);
    double pre2 = 2.0 * std::tan(dist[0] / 2.0);
    double ru = std::tan(rd * dist[0]) / pre2;
    double conv = 1.0;
    if (rd > 1e-5) conv = ru / rd;
    u[0] = conv * (nx * f); u[1] = conv * (ny * f);
  } else if (s.nd == 4 || s.nd == 5) {
    // src = distorted_c + distorted_half_size; K_dist = (f, optical centre); K_undist = (f, undistorted_half_size);
    // result minus undistorted_half_size
    double k[5] = {dist[0], dist[1], dist[2], dist[3], s.nd == 5 ? dist[4] : 0.0};
    double x0 = ((dc[0] + s.half[0]) - c[0]) / f, y0 = ((dc[1] + s.half[1]) - c[1]) / f;
    double x = x0, y = y0;
    for (int j = 0; j < 5; j++) {
      double r2 = x * x + y * y;
      double icdist = 1.0 / (1.0 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
      double dx = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x);
      double dy = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y;
      x = (x0 - dx) * icdist; y = (y0 - dy) * icdist;
    }
    u[0] = x * f; u[1] = y * f;
  } else {
    compute_rpc(dc, dist + s.nd / 2, s.nd / 2, u);
  }
}

struct View34 { double P[3][4]; double p[2]; };

// openMVG::Triangulation::compute(3)
inline void triangulate_views(const std::vector<View34>& views, double X[3]) {
  const int n = (int)views.size();
  std::vector<double> weights(n, 1.0);
  for (int it = 0; it < 3; it++) {
    double AtA[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}}, Atb[3] = {0, 0, 0};
    for (int i = 0; i < n; i++) {
      const auto& P = views[i].P; const double* p = views[i].p;
      const double w = weights[i];
      double v1[3], v2[3];
      for (int j = 0; j < 3; j++) {
        v1[j] = w * (P[0][j] - p[0] * P[2][j]);
        v2[j] = w * (P[1][j] - p[1] * P[2][j]);
        Atb[j] += w * (v1[j] * (p[0] * P[2][3] - P[0][3]) + v2[j] * (p[1] * P[2][3] - P[1][3]));
      }
      for (int k = 0; k < 3; k++)
        for (int j = 0; j <= k; j++) {
          double v = v1[j] * v1[k] + v2[j] * v2[k];
          AtA[j][k] += v;
          if (j < k) AtA[k][j] += v;
        }
    }
    // Eigen 3x3 inverse: cofactors over the determinant
    const double (*m)[3] = AtA;
    double c00 = m[1][1] * m[2][2] - m[1][2] * m[2][1], c01 = m[1][2] * m[2][0] - m[1][0] * m[2][2], c02 = m[1][0] * m[2][1] - m[1][1] * m[2][0];
    double det = m[0][0] * c00 + m[0][1] * c01 + m[0][2] * c02, id = 1.0 / det;
    double inv[3][3] = {{c00 * id, (m[0][2] * m[2][1] - m[0][1] * m[2][2]) * id, (m[0][1] * m[1][2] - m[0][2] * m[1][1]) * id},
                        {c01 * id, (m[0][0] * m[2][2] - m[0][2] * m[2][0]) * id, (m[0][2] * m[1][0] - m[0][0] * m[1][2]) * id},
                        {c02 * id, (m[0][1] * m[2][0] - m[0][0] * m[2][1]) * id, (m[0][0] * m[1][1] - m[0][1] * m[1][0]) * id}};
    for (int a = 0; a < 3; a++) X[a] = inv[a][0] * Atb[0] + inv[a][1] * Atb[1] + inv[a][2] * Atb[2];
    for (int i = 0; i < n; i++) {
      const auto& P = views[i].P;
      double z = P[2][0] * X[0] + P[2][1] * X[1] + P[2][2] * X[2] + P[2][3];
      weights[i] = 1.0 / z;
    }
  }
}

// ---- ray / mesh (rayint/libs/acc/primitives.h:100-136, float32 exactly as written there; compiled with -ffp-contract=off) ----
inline float dot3f(const float* a, const float* b) { float s = 0.0f; s = s + a[0] * b[0]; s = s + a[1] * b[1]; s = s + a[2] * b[2]; return s; }

inline bool ray_tri(const float* o, const float* d, float tmin, float tmax, const float* A, const float* B, const float* C, float* t_out) {
  float v0[3], v1[3], oa[3];
  for (int i = 0; i < 3; i++) { v0[i] = B[i] - A[i]; v1[i] = C[i] - A[i]; oa[i] = o[i] - A[i]; }
  float n[3] = {v1[1] * v0[2] - v1[2] * v0[1], v1[2] * v0[0] - v1[0] * v0[2], v1[0] * v0[1] - v1[1] * v0[0]};
  float len = std::sqrt(dot3f(n, n));
  if (len < std::numeric_limits<float>::epsilon()) return false;
  for (int i = 0; i < 3; i++) n[i] = n[i] / len;
  double cosine = dot3f(n, d);
  if (std::abs(cosine) < std::numeric_limits<double>::epsilon()) return false;
  float t = -dot3f(n, oa) / cosine;
  if (t < tmin || tmax < t) return false;
  float v2[3];
  for (int i = 0; i < 3; i++) v2[i] = oa[i] + t * d[i];
  float d00 = dot3f(v0, v0), d01 = dot3f(v0, v1), d11 = dot3f(v1, v1), d20 = dot3f(v2, v0), d21 = dot3f(v2, v1);
  double denom = d00 * d11 - d01 * d01;
  float b1 = (d11 * d20 - d01 * d21) / denom;
  float b2 = (d00 * d21 - d01 * d20) / denom;
  float b0 = 1.0f - b1 - b2;
  *t_out = t;
  const float eps = 1e-3;
  return -eps <= b0 && b0 <= 1.0f + eps && -eps <= b1 && b1 <= 1.0f + eps && -eps <= b2 && b2 <= 1.0f + eps;
}

// BVHTree::intersect (rayint/libs/acc/bvh_tree.h:351-405) without the hierarchy: the nearest accepted hit over ALL triangles, tmax
// shrinking with every hit as the traversal does.  Pinned against the reference's own BVHTree (oracle/_ref/libref_rayint.so,
// tests/golden/rayint.npz).
inline bool ray_mesh_nearest(const float* o, const float* d, float tmin, float tmax, const float* verts, const uint32_t* faces, int64_t nfaces,
                             float* best_out) {
  float best = std::numeric_limits<float>::infinity();
  for (int64_t k = 0; k < nfaces; k++) {
    float t;
    if (ray_tri(o, d, tmin, tmax, verts + 3 * (size_t)faces[3 * k], verts + 3 * (size_t)faces[3 * k + 1], verts + 3 * (size_t)faces[3 * k + 2], &t) &&
        !(t > best)) { best = t; tmax = t; }
  }
  *best_out = best;
  return best < std::numeric_limits<float>::infinity();
}

// ray_mesh_intersect (texture_processing.cc:1436-1479) without an acceleration structure: the nearest accepted hit over
// ALL triangles, which is what the BVH traversal returns.
inline bool ray_mesh_brute(const double u[2], double f, const Aff<double>& w2c, const float* verts, const uint32_t* faces, int64_t nfaces,
                           double min_dist, double max_dist, double out[3]) {
  double ray[3] = {u[0] / f, u[1] / f, 1.0};
  double rn = std::sqrt(ray[0] * ray[0] + ray[1] * ray[1] + ray[2] * ray[2]);
  for (int a = 0; a < 3; a++) ray[a] /= rn;
  Aff<double> c2w = aff_inverse(w2c);
  double wr[3], ctr[3] = {c2w.t[0], c2w.t[1], c2w.t[2]};
  for (int a = 0; a < 3; a++) wr[a] = c2w.m[a][0] * ray[0] + c2w.m[a][1] * ray[1] + c2w.m[a][2] * ray[2];
  float o[3] = {(float)ctr[0], (float)ctr[1], (float)ctr[2]}, d[3] = {(float)wr[0], (float)wr[1], (float)wr[2]};
  float dn = std::sqrt(dot3f(d, d));
  for (int a = 0; a < 3; a++) d[a] = d[a] / dn;
  float best;
  if (!ray_mesh_nearest(o, d, (float)min_dist, (float)max_dist, verts, faces, nfaces, &best)) return false;
  for (int a = 0; a < 3; a++) out[a] = ctr[a] + (double)best * wr[a];
  return true;
}

}  // namespace oracle
