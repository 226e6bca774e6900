// oracle/ref_build/ref_rayint.cc -- TEST INFRASTRUCTURE ONLY.  A C entry point around the REFERENCE's own ray / mesh code,
// compiled from the headers where they lie under /root/reference/rayint/libs (acc/bvh_tree.h, acc/primitives.h,
// math/vector.h: header-only, no external dependency) into oracle/_ref/libref_rayint.so.  No reference source is copied into
// this repository; this file only instantiates acc::BVHTree<unsigned int, math::Vec3f> the way
// rig_calibrator/include/rig_calibrator/texture_processing.h:57 and texture_processing.cc:286, 1462-1479 do.
#include <cstdint>
#include <limits>
#include <memory>
#include <vector>

#include <acc/bvh_tree.h>
#include <math/vector.h>

typedef acc::BVHTree<unsigned int, math::Vec3f> BVHTree;

extern "C" {

// rays: origin[3 n], dir[3 n] (float32, dir already normalised by the caller as ray_mesh_intersect does); the hit flag, the
// float distance and the struck triangle of every ray come back.  Returns the number of hits.
int64_t ref_rayint_intersect(int64_t nv, const float* verts, int64_t nf, const uint32_t* faces, int64_t n, const float* origin, const float* dir,
                             float tmin, float tmax, int32_t build_threads, uint8_t* hit_out, float* t_out, uint32_t* idx_out) {
  std::vector<math::Vec3f> vertices((size_t)nv);
  for (int64_t i = 0; i < nv; i++) vertices[i] = math::Vec3f(verts[3 * i], verts[3 * i + 1], verts[3 * i + 2]);
  std::vector<unsigned int> f(faces, faces + 3 * nf);
  BVHTree tree(f, vertices, build_threads > 0 ? build_threads : 1);
  int64_t hits = 0;
  for (int64_t i = 0; i < n; i++) {
    BVHTree::Ray ray;
    ray.origin = math::Vec3f(origin[3 * i], origin[3 * i + 1], origin[3 * i + 2]);
    ray.dir = math::Vec3f(dir[3 * i], dir[3 * i + 1], dir[3 * i + 2]);
    ray.tmin = tmin; ray.tmax = tmax;
    BVHTree::Hit hit;
    const bool ok = tree.intersect(ray, &hit);
    hit_out[i] = ok ? 1 : 0; t_out[i] = ok ? hit.t : std::numeric_limits<float>::infinity(); idx_out[i] = ok ? hit.idx : 0xffffffffu;
    hits += ok;
  }
  return hits;
}

}
