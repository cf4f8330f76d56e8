// TEST INFRASTRUCTURE ONLY -- C wrapper around the UNMODIFIED reference BVH
// builder, compiled from /root/reference in place (see oracle/Makefile):
//   cave/engine/renderer/path_tracer/bvh_accel.cpp  (BvhAccel::Construct :265-288,
//                                                    FillGpuBvhAccel :237-263)
//   cave/engine/math/box.cpp                        (Box<3>::SurfaceArea :7-17)
// Output: oracle/_ref/libcave_ref_bvh.so.  Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may load it.
#include <chrono>
#include <cstdio>

#include "engine/math/matrix.h"
#include "engine/renderer/path_tracer/bvh_accel.h"

namespace cave {

static thread_local char s_last_error[512];

// declared at cave/engine/core/error/error_macros.h:25-29; the engine's real
// implementation lives in core/error/error_macros.cpp which drags the logger in.
void ReportErrorImpl(std::string_view p_function, std::string_view p_file, int p_line,
                     std::string_view p_error, std::string_view p_optional_message) {
    std::snprintf(s_last_error, sizeof(s_last_error), "%.*s (%.*s:%d) %.*s %.*s",
                  (int)p_function.size(), p_function.data(), (int)p_file.size(), p_file.data(), p_line,
                  (int)p_error.size(), p_error.data(), (int)p_optional_message.size(), p_optional_message.data());
}

}  // namespace cave

static_assert(sizeof(cave::GpuPtBvh) == 48, "GpuPtBvh layout");

extern "C" {

__attribute__((visibility("default"))) const char* caveref_last_error() { return cave::s_last_error; }

// Returns node count (2*tris-1) or -1 if the reference trapped (degenerate
// geometry: DEV_ASSERT / CRASH_NOW_MSG), -2 if `cap` is too small.
// ms[0] = Construct wall ms, ms[1] = FillGpuBvhAccel wall ms (may be null).
__attribute__((visibility("default"))) int caveref_bvh_build(const uint32_t* p_indices, int p_index_count,
                                                              const float* p_positions, int p_vertex_count,
                                                              void* p_out_nodes, int p_cap, double* p_ms) {
    using clk = std::chrono::steady_clock;
    std::vector<uint32_t> indices(p_indices, p_indices + p_index_count);
    std::vector<cave::Vector3f> vertices(p_vertex_count);
    for (int i = 0; i < p_vertex_count; ++i) {
        vertices[i] = cave::Vector3f(p_positions[3 * i], p_positions[3 * i + 1], p_positions[3 * i + 2]);
    }
    try {
        auto t0 = clk::now();
        auto root = cave::BvhAccel::Construct(indices, vertices);
        auto t1 = clk::now();
        std::vector<cave::GpuPtBvh> out;
        root->FillGpuBvhAccel(out);
        auto t2 = clk::now();
        if (p_ms) {
            p_ms[0] = std::chrono::duration<double, std::milli>(t1 - t0).count();
            p_ms[1] = std::chrono::duration<double, std::milli>(t2 - t1).count();
        }
        if ((int)out.size() > p_cap) {
            return -2;
        }
        // _padding is uninitialised in the reference (bvh_accel.cpp:241); zero it so
        // byte compares are meaningful on the 40 defined bytes + zero padding.
        for (auto& n : out) {
            n._padding = cave::Vector2f(0.0f, 0.0f);
        }
        std::memcpy(p_out_nodes, out.data(), out.size() * sizeof(cave::GpuPtBvh));
        return (int)out.size();
    } catch (const CaveRefTrap&) {
        return -1;
    }
}

// glm::inverse exactly as the uploader calls it (path_tracer.cpp:178); column-major 16 floats.
__attribute__((visibility("default"))) void caveref_mat4_inverse(const float* p_m, float* p_out) {
    glm::mat4 m;
    std::memcpy(&m[0][0], p_m, 64);
    glm::mat4 inv = glm::inverse(m);
    std::memcpy(p_out, &inv[0][0], 64);
}
}
