// TEST INFRASTRUCTURE ONLY.  C wrapper over the reference's own CPU twins of the shader's
// intersection routines: TestIntersection::RayTriangle and
// TestIntersection::RayAabb, cave/engine/math/intersection.cpp:18-70, compiled in place.
// Ray keeps m_dist private with no accessor; the wrapper opens the class for reading.
#include <cstring>
#include <string_view>

#define private public
#include "engine/math/ray.h"
#undef private
#include "engine/math/aabb.h"
#include "engine/math/intersection.h"

namespace cave {
void ReportErrorImpl(std::string_view, std::string_view, int, std::string_view, std::string_view) {}
}  // namespace cave

using cave::Vector3f;

// rays: n x 6 (start xyz, end xyz); tris: n x 9; out: n x 2 (hit, m_dist after the call; m_dist starts at 1)
extern "C" __attribute__((visibility("default"))) void cave_ref_ray_triangle(int n, const float* rays, const float* tris, float* out) {
    for (int i = 0; i < n; ++i) {
        const float* r = rays + 6 * i;
        const float* t = tris + 9 * i;
        cave::Ray ray(Vector3f(r[0], r[1], r[2]), Vector3f(r[3], r[4], r[5]));
        const bool hit = cave::TestIntersection::RayTriangle(Vector3f(t[0], t[1], t[2]), Vector3f(t[3], t[4], t[5]), Vector3f(t[6], t[7], t[8]), ray);
        out[2 * i] = hit ? 1.0f : 0.0f;
        out[2 * i + 1] = ray.m_dist;
    }
}

// boxes: n x 6 (min xyz, max xyz); out: n x 2 (hit, m_dist after the call = tmin on a hit)
extern "C" __attribute__((visibility("default"))) void cave_ref_ray_aabb(int n, const float* rays, const float* boxes, float* out) {
    for (int i = 0; i < n; ++i) {
        const float* r = rays + 6 * i;
        const float* b = boxes + 6 * i;
        cave::Ray ray(Vector3f(r[0], r[1], r[2]), Vector3f(r[3], r[4], r[5]));
        cave::AABB box(Vector3f(b[0], b[1], b[2]), Vector3f(b[3], b[4], b[5]));
        const bool hit = cave::TestIntersection::RayAabb(box, ray);
        out[2 * i] = hit ? 1.0f : 0.0f;
        out[2 * i + 1] = ray.m_dist;
    }
}
