// TEST INFRASTRUCTURE ONLY.  C wrapper over the reference's own TransformComponent
// (cave/engine/scene/transform_component.cpp, compiled in place by oracle/Makefile) restating
// the ancestor walk of UpdateHierarchy (cave/engine/systems/ecs_systems.cpp:139-171), so the
// world matrices the product's scene ingest computes can be checked bit for bit against glm.
#include <cstdio>
#include <cstring>
#include <string_view>

#include "engine/scene/transform_component.h"

namespace cave {
// declared at cave/engine/core/error/error_macros.h:25-29
void ReportErrorImpl(std::string_view, std::string_view, int, std::string_view, std::string_view) {}
}  // namespace cave

static cave::TransformComponent make(const float* trs) {
    cave::TransformComponent c;
    c.SetTranslation(cave::Vector3f(trs[0], trs[1], trs[2]));
    c.SetRotation(cave::Vector4f(trs[3], trs[4], trs[5], trs[6]));
    c.SetScale(cave::Vector3f(trs[7], trs[8], trs[9]));
    return c;
}

// chain = n x 10 floats (translation xyz, rotation xyzw, scale xyz): the entity itself first,
// then its parent, grandparent, ...  out16 = column-major world matrix.
extern "C" __attribute__((visibility("default"))) void cave_ref_world_matrix(int n, const float* chain, float* out16) {
    cave::Matrix4x4f world = make(chain).GetLocalMatrix();                                  // ecs_systems.cpp:149
    for (int k = 1; k < n; ++k) world = make(chain + 10 * k).GetLocalMatrix() * world;      // :156
    std::memcpy(out16, &world[0][0], 64);
}
