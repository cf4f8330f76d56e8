// TEST INFRASTRUCTURE ONLY -- the reference's shader logic compiled as C++.
// cave/shader/shared_path_tracer.h (WangHash .. RayColor, :60-239) is included
// verbatim from /root/reference (through a build-dir copy whose only change is
// the MAX_BOUNCE define, see gen_shim.py).  The compute entry point
// (cave/shader/glsl/path_tracer.cs.glsl:11-60) is GLSL-only syntax and is
// restated line for line in caveref_trace_render() below.
// Output: oracle/_ref/libcave_ref_trace_mb<K>.so.
#include <atomic>
#include <chrono>
#include <thread>

#include "engine/math/matrix.h"
#include "engine/math/vector.h"

namespace cave {

// In the shaders the cbuffer members are globals; in C++ the cbuffer is a struct,
// so the one member the shared header reads (shared_path_tracer.h:159) is
// provided here.
static int c_ptObjectCount;

// GLSL/HLSL built-ins the header relies on.
static inline Vector3f reflect(const Vector3f& p_i, const Vector3f& p_n) {
    return p_i - 2.0f * dot(p_n, p_i) * p_n;
}
#define mul(a, b) ((a) * (b))

template<int A, int B, int C>
static inline Vector3f normalize(const Swizzle3<float, 4, A, B, C, -1>& p_swizzle) {
    return normalize(Vector3f(p_swizzle));
}

}  // namespace cave

#include "shared_path_tracer.h"

namespace cave {

// The header declares these as unbounded extern arrays (:17-21); they must be
// real arrays, so fixed-capacity BSS is used (pages are touched only when filled).
constexpr size_t CAP_VERTS = 6u << 20;
constexpr size_t CAP_TRIS = 3u << 20;
constexpr size_t CAP_NODES = 6u << 20;
constexpr size_t CAP_MESHES = 4096;
constexpr size_t CAP_MATS = 4096;
GpuPtVertex GlobalPtVertices[CAP_VERTS];
GpuPtIndex GlobalPtIndices[CAP_TRIS];
GpuPtBvh GlobalPtBvhs[CAP_NODES];
GpuPtMesh GlobalPtMeshes[CAP_MESHES];
GpuPtMaterial GlobalPtMaterials[CAP_MATS];

static_assert(sizeof(GpuPtVertex) == 32 && sizeof(GpuPtIndex) == 16 && sizeof(GpuPtBvh) == 48);
static_assert(sizeof(GpuPtMesh) == 144 && sizeof(GpuPtMaterial) == 32);

}  // namespace cave

#define CAVEREF_API extern "C" __attribute__((visibility("default")))

CAVEREF_API int caveref_trace_max_bounce() { return MAX_BOUNCE; }

// Arrays are in the reference's GPU layouts (structured_buffer.hlsl.h:24-61).
CAVEREF_API int caveref_trace_load(const void* p_vertices, int p_nv, const void* p_indices, int p_nt,
                                   const void* p_bvhs, int p_nb, const void* p_meshes, int p_nm,
                                   const void* p_materials, int p_nmat) {
    using namespace cave;
    if ((size_t)p_nv > CAP_VERTS || (size_t)p_nt > CAP_TRIS || (size_t)p_nb > CAP_NODES ||
        (size_t)p_nm > CAP_MESHES || (size_t)p_nmat > CAP_MATS) {
        return -1;
    }
    std::memcpy(GlobalPtVertices, p_vertices, (size_t)p_nv * sizeof(GpuPtVertex));
    std::memcpy(GlobalPtIndices, p_indices, (size_t)p_nt * sizeof(GpuPtIndex));
    std::memcpy(GlobalPtBvhs, p_bvhs, (size_t)p_nb * sizeof(GpuPtBvh));
    std::memcpy(GlobalPtMeshes, p_meshes, (size_t)p_nm * sizeof(GpuPtMesh));
    std::memcpy(GlobalPtMaterials, p_materials, (size_t)p_nmat * sizeof(GpuPtMaterial));
    c_ptObjectCount = p_nm;
    return 0;
}

struct CaverefCamera {
    float position[3];
    float fov_degree;
    float forward[3];
    float right[3];
    float up[3];
};

// One invocation of main() (path_tracer.cs.glsl:11-60) for pixel (x, y).
// Returns the number of path segments (HitScene calls) the invocation made,
// recovered from how far RayColor advanced the RNG: 3 Random() per hit
// (shared_path_tracer.h:219-221) and none on a miss.
static int ShadePixel(const CaverefCamera& p_cam, int p_x, int p_y, int p_w, int p_h, int p_frame_index,
                      int p_scene_dirty, float* p_pixel) {
    using namespace cave;
    Vector2f fPixelCoords((float)p_x, (float)p_y);

    uint seed = uint(uint(p_x) * uint(1973) + uint(p_y) * uint(9277) + uint(p_frame_index) * uint(26699)) | uint(1);

    float j0 = Random(seed);
    float j1 = Random(seed);
    Vector2f jitter = Vector2f(j0, j1) - 0.5f;

    Vector3f rayDir;
    {
        Vector2f dims((float)p_w, (float)p_h);
        Vector2f uvJitter = (fPixelCoords + jitter) / dims;
        Vector2f screen = 2.0f * uvJitter - 1.0f;

        Vector2f resolution((float)p_w, (float)p_h);
        float aspectRatio = resolution.x / resolution.y;
        screen.y /= aspectRatio;
        float halfFov = p_cam.fov_degree;
        float camDistance = std::tan(halfFov * MY_PI / 180.0f);
        Vector3f right(p_cam.right[0], p_cam.right[1], p_cam.right[2]);
        Vector3f up(p_cam.up[0], p_cam.up[1], p_cam.up[2]);
        Vector3f forward(p_cam.forward[0], p_cam.forward[1], p_cam.forward[2]);
        // mat3(right, up, forward) * vec3(screen, camDistance)
        rayDir = right * screen.x + up * screen.y + forward * camDistance;
        rayDir = normalize(rayDir);
    }

    Ray ray;
    ray.origin = Vector3f(p_cam.position[0], p_cam.position[1], p_cam.position[2]);
    ray.direction = rayDir;
    ray.invDir = 1.0f / ray.direction;
    ray.t = RAY_T_MAX;

    const uint seed_before = seed;
    Vector3f radiance = RayColor(ray, seed);

    int hits = 0;
    {
        uint s = seed_before;
        while (s != seed && hits < MAX_BOUNCE) {
            WangHash(s);
            WangHash(s);
            WangHash(s);
            ++hits;
        }
    }
    const int segments = hits < MAX_BOUNCE ? hits + 1 : MAX_BOUNCE;

    float final_color[4] = { radiance.x, radiance.y, radiance.z, 1.0f };
    if (p_scene_dirty == 0) {
        const float* accumulated = p_pixel;
        float weight = accumulated[3];
        Vector3f new_color = radiance + weight * Vector3f(accumulated[0], accumulated[1], accumulated[2]);
        float new_weight = accumulated[3] + 1.0f;
        new_color /= new_weight;
        final_color[0] = new_color.x;
        final_color[1] = new_color.y;
        final_color[2] = new_color.z;
        final_color[3] = new_weight;
    }
    p_pixel[0] = final_color[0];
    p_pixel[1] = final_color[1];
    p_pixel[2] = final_color[2];
    p_pixel[3] = final_color[3];
    return segments;
}

// Renders frames [first_frame, first_frame + n_frames) over the pixel rectangle
// [x0,x1) x [y0,y1) of a W x H image into `accum` (W*H float4, row-major, pixel
// (x,y) at y*W+x exactly as imageStore addresses it).  The first frame is a
// "dirty" frame (overwrite) iff reset_first != 0.  Rows are dealt to `threads`
// std::threads.  Returns total path segments; *ms = wall milliseconds.
CAVEREF_API long long caveref_trace_render(const CaverefCamera* p_cam, int p_w, int p_h, int p_x0, int p_y0,
                                           int p_x1, int p_y1, int p_first_frame, int p_n_frames,
                                           int p_reset_first, float* p_accum, int p_threads, double* p_ms) {
    std::atomic<long long> total{ 0 };
    std::atomic<int> next_row{ p_y0 };
    auto t0 = std::chrono::steady_clock::now();
    auto worker = [&]() {
        long long local = 0;
        for (;;) {
            const int y = next_row.fetch_add(1);
            if (y >= p_y1) {
                break;
            }
            for (int x = p_x0; x < p_x1; ++x) {
                float* pixel = p_accum + 4 * ((size_t)y * p_w + x);
                for (int f = 0; f < p_n_frames; ++f) {
                    const int dirty = (f == 0 && p_reset_first) ? 1 : 0;
                    local += ShadePixel(*p_cam, x, y, p_w, p_h, p_first_frame + f, dirty, pixel);
                }
            }
        }
        total += local;
    };
    if (p_threads <= 1) {
        worker();
    } else {
        std::vector<std::thread> pool;
        for (int i = 0; i < p_threads; ++i) {
            pool.emplace_back(worker);
        }
        for (auto& t : pool) {
            t.join();
        }
    }
    auto t1 = std::chrono::steady_clock::now();
    if (p_ms) {
        *p_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    }
    return total.load();
}
