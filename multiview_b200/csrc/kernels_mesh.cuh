// kernels_mesh.cuh -- SURVEY.md 8(f) row N3: camera rays against a triangle mesh, one THREAD per observation.
//   ray_mesh_intersect   reference rig_calibrator/src/texture_processing.cc:1436-1479
//   meshTriangulations   texture_processing.cc:1566-1625 (per-feature hit + per-track mean)
//   triangle / box tests rayint/libs/acc/primitives.h:86-136 (float32, the reference's arithmetic is kept operation by
//                        operation: sequential float dot products, the float / double mix of the barycentric test, eps = 1e-3)
//   traversal            rayint/libs/acc/bvh_tree.h:369-405 (nearer child first, tmax shrinks with every hit)
// The hierarchy itself is ours (host-built median split, <= 4 triangles per leaf).  The nearest accepted hit does not depend on
// how the boxes are nested, EXCEPT for rays within the test's eps = 1e-3 barycentric band of an edge: such a hit may lie outside
// its triangle's box, and then which of the neighbouring triangles answers depends on the tree -- in the reference too
// (tests/test_rayint_pin.py measures it against the reference's own BVHTree: ~1 % of random rays, <= 2e-3 relative in distance).
// No graphics API is involved.
#pragma once
#include "kernels_tracks.cuh"

namespace mvba {

struct BvhNode { float lo[3], hi[3]; int left, right, first, count; };  // leaf: left < 0, triangles [first, first + count)
struct MeshTri { float a[3], b[3], c[3]; };

// float arithmetic without FMA contraction, in the reference's association order
__device__ __forceinline__ float dot3f(const float* a, const float* b) {
  return __fadd_rn(__fadd_rn(__fadd_rn(0.0f, __fmul_rn(a[0], b[0])), __fmul_rn(a[1], b[1])), __fmul_rn(a[2], b[2]));
}
__device__ __forceinline__ float stdminf(float a, float b) { return (b < a) ? b : a; }  // std::min / std::max, NaN behaviour included
__device__ __forceinline__ float stdmaxf(float a, float b) { return (a < b) ? b : a; }

// acc::intersect(ray, aabb): slab test that ignores NaN slabs
__device__ __forceinline__ bool ray_box(const float* o, const float* d, float tmin, float tmax, const BvhNode& n, float* tnear) {
  const float inf = __int_as_float(0x7f800000);
#pragma unroll
  for (int i = 0; i < 3; i++) {
    const float t1 = __fdiv_rn(__fsub_rn(n.lo[i], o[i]), d[i]);
    const float t2 = __fdiv_rn(__fsub_rn(n.hi[i], o[i]), d[i]);
    tmin = stdmaxf(tmin, stdminf(stdminf(t1, t2), inf));
    tmax = stdminf(tmax, stdmaxf(stdmaxf(t1, t2), -inf));
  }
  *tnear = tmin;
  return tmax >= stdmaxf(tmin, 0.0f);
}

// acc::intersect(ray, tri)
__device__ __forceinline__ bool ray_tri(const float* o, const float* d, float tmin, float tmax, const MeshTri& tr, float* t_out) {
  float v0[3], v1[3], oa[3];
#pragma unroll
  for (int i = 0; i < 3; i++) { v0[i] = __fsub_rn(tr.b[i], tr.a[i]); v1[i] = __fsub_rn(tr.c[i], tr.a[i]); oa[i] = __fsub_rn(o[i], tr.a[i]); }
  float nrm[3] = {__fsub_rn(__fmul_rn(v1[1], v0[2]), __fmul_rn(v1[2], v0[1])), __fsub_rn(__fmul_rn(v1[2], v0[0]), __fmul_rn(v1[0], v0[2])),
                  __fsub_rn(__fmul_rn(v1[0], v0[1]), __fmul_rn(v1[1], v0[0]))};  // v1.cross(v0)
  const float len = __fsqrt_rn(dot3f(nrm, nrm));
  if (len < 1.1920929e-07f) return false;
#pragma unroll
  for (int i = 0; i < 3; i++) nrm[i] = __fdiv_rn(nrm[i], len);
  const double cosine = (double)dot3f(nrm, d);
  if (fabs(cosine) < 2.220446049250313e-16) return false;
  const float t = (float)((double)(-dot3f(nrm, oa)) / cosine);
  if (t < tmin || tmax < t) return false;
  float v2[3];
#pragma unroll
  for (int i = 0; i < 3; i++) v2[i] = __fadd_rn(oa[i], __fmul_rn(t, d[i]));
  const float d00 = dot3f(v0, v0), d01 = dot3f(v0, v1), d11 = dot3f(v1, v1), d20 = dot3f(v2, v0), d21 = dot3f(v2, v1);
  const double denom = (double)__fsub_rn(__fmul_rn(d00, d11), __fmul_rn(d01, d01));
  const float b1 = (float)((double)__fsub_rn(__fmul_rn(d11, d20), __fmul_rn(d01, d21)) / denom);
  const float b2 = (float)((double)__fsub_rn(__fmul_rn(d00, d21), __fmul_rn(d01, d20)) / denom);
  const float b0 = __fsub_rn(__fsub_rn(1.0f, b1), b2);
  *t_out = t;
  const float eps = 1e-3f;
  return -eps <= b0 && b0 <= __fadd_rn(1.0f, eps) && -eps <= b1 && b1 <= __fadd_rn(1.0f, eps) && -eps <= b2 && b2 <= __fadd_rn(1.0f, eps);
}

#define MVBA_BAD_XYZ 1e100  // bad_xyz, rig_calibrator.cc:1546
#define MVBA_BVH_STACK 64   // traversal stack entries; the host refuses deeper trees (a median-split tree over 2^31 triangles has 30 levels)

__global__ void mesh_intersect_kernel(View v, const double* __restrict__ x, const CamTab* __restrict__ tab, const BvhNode* __restrict__ nodes,
                                      const MeshTri* __restrict__ tris, double min_ray_dist, double max_ray_dist,
                                      const uint8_t* __restrict__ inlier, uint8_t* __restrict__ has_mesh, double* __restrict__ mesh_xyz, int* __restrict__ err) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < v.N; i += (int64_t)gridDim.x * blockDim.x) {
    double out[3] = {MVBA_BAD_XYZ, MVBA_BAD_XYZ, MVBA_BAD_XYZ};
    uint8_t hit_any = 0;
    if (inlier[i]) {
      const int c = v.cid[i];
      const SensorDev& sd = v.sens[v.cam_sensor[c]];
      const float2 pix = v.pix[i];
      double u[2];
      undistort_centered(sd, x, (double)pix.x, (double)pix.y, u);
      const double f = x[sd.x_focal];
      double ray[3] = {u[0] / f, u[1] / f, 1.0};
      const double rn = sqrt(ray[0] * ray[0] + ray[1] * ray[1] + ray[2] * ray[2]);
      ray[0] /= rn; ray[1] /= rn; ray[2] /= rn;
      // cam_to_world = world_to_cam.inverse(): general 3x3 inverse of the linear part, translation = -inv t
      double Ri[9];
      inv3(tab[c].R, Ri);
      const double* t = tab[c].t;
      double ctr[3], wr[3];
#pragma unroll
      for (int a = 0; a < 3; a++) {
        ctr[a] = -(Ri[3 * a] * t[0] + Ri[3 * a + 1] * t[1] + Ri[3 * a + 2] * t[2]);
        wr[a] = Ri[3 * a] * ray[0] + Ri[3 * a + 1] * ray[1] + Ri[3 * a + 2] * ray[2];
      }
      float o[3] = {(float)ctr[0], (float)ctr[1], (float)ctr[2]}, d[3] = {(float)wr[0], (float)wr[1], (float)wr[2]};
      const float dn = __fsqrt_rn(dot3f(d, d));
      d[0] = __fdiv_rn(d[0], dn); d[1] = __fdiv_rn(d[1], dn); d[2] = __fdiv_rn(d[2], dn);
      const float tmin = (float)min_ray_dist;
      float tmax = (float)max_ray_dist;
      float best = __int_as_float(0x7f800000);
      int stack[MVBA_BVH_STACK], sp = 0;
      stack[sp++] = 0;
      while (sp > 0) {
        const BvhNode& n = nodes[stack[--sp]];
        if (n.left >= 0) {
          float tl, tr;
          const bool hl = ray_box(o, d, tmin, tmax, nodes[n.left], &tl), hr = ray_box(o, d, tmin, tmax, nodes[n.right], &tr);
          if (hl && hr) {
            if (tl < tr) { stack[sp++] = n.right; stack[sp++] = n.left; } else { stack[sp++] = n.left; stack[sp++] = n.right; }
          } else {
            if (hr) stack[sp++] = n.right;
            if (hl) stack[sp++] = n.left;
          }
          if (sp > MVBA_BVH_STACK - 2) { atomicExch(err, MVBA_EINVAL); break; }  // cannot happen for a tree the host accepted; never truncate silently
        } else {
          for (int k = n.first; k < n.first + n.count; k++) {
            float tt;
            if (ray_tri(o, d, tmin, tmax, tris[k], &tt) && !(tt > best)) { best = tt; tmax = tt; }
          }
        }
      }
      if (best < __int_as_float(0x7f800000)) {
        hit_any = 1;
        const double dist = (double)best;
        out[0] = ctr[0] + dist * wr[0]; out[1] = ctr[1] + dist * wr[1]; out[2] = ctr[2] + dist * wr[2];
      }
    }
    has_mesh[i] = hit_any;
    mesh_xyz[3 * i] = out[0]; mesh_xyz[3 * i + 1] = out[1]; mesh_xyz[3 * i + 2] = out[2];
  }
}

// Per track: mean of the features' mesh hits, bad_xyz when there is none.
__global__ void mesh_track_mean_kernel(View v, const uint8_t* __restrict__ has_mesh, const double* __restrict__ mesh_xyz, double* __restrict__ pid_mesh_xyz) {
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < v.P; p += gridDim.x * blockDim.x) {
    double s[3] = {0, 0, 0};
    int n = 0;
    for (int64_t i = v.trk_obs0[p]; i < v.trk_obs0[p + 1]; i++)
      if (has_mesh[i]) { s[0] += mesh_xyz[3 * i]; s[1] += mesh_xyz[3 * i + 1]; s[2] += mesh_xyz[3 * i + 2]; n++; }
#pragma unroll
    for (int a = 0; a < 3; a++) pid_mesh_xyz[3 * (size_t)p + a] = n >= 1 ? s[a] / n : MVBA_BAD_XYZ;
  }
}

}  // namespace mvba
