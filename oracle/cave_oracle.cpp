// TEST INFRASTRUCTURE ONLY -- see cave_oracle.h.  CPU restatement of the
// reference's algorithms; every function cites the reference lines it follows.
// All arithmetic is IEEE fp32 with no contraction (built with -ffp-contract=off),
// i.e. what the GLSL/HLSL shader and the engine's C++ builder compute.
#include "cave_oracle.h"

#include <atomic>
#include <chrono>
#include <cmath>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>

#define CO_API extern "C" __attribute__((visibility("default")))

namespace {

struct V3 {
    float x, y, z;
    float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
inline V3 operator+(V3 a, V3 b) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
inline V3 operator-(V3 a, V3 b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
inline V3 operator*(V3 a, V3 b) { return { a.x * b.x, a.y * b.y, a.z * b.z }; }
inline V3 operator*(float s, V3 a) { return { s * a.x, s * a.y, s * a.z }; }
inline V3 operator*(V3 a, float s) { return { a.x * s, a.y * s, a.z * s }; }
// cave/engine/math/detail/common.h:50-58 -- note the argument order on ties.
inline float fmin_ref(float a, float b) { return a < b ? a : b; }
inline float fmax_ref(float a, float b) { return a > b ? a : b; }
// cave/engine/math/detail/vector_math.h:513-523
inline float dot(V3 a, V3 b) {
    V3 t = a * b;
    float r = t.x + t.y;
    r += t.z;
    return r;
}
// vector_math.h:539-545
inline V3 cross(V3 a, V3 b) { return { a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y }; }
// vector_math.h:527-536
inline V3 normalize(V3 a) {
    const float inv = 1.0f / std::sqrt(dot(a, a));
    return a * inv;
}
// vector_math.h:507-510
inline V3 lerp(V3 x, V3 y, float s) { return (1.0f - s) * x + s * y; }

// cave/engine/math/box.h:11-73 (Box<3>)
struct Box {
    V3 mn{ std::numeric_limits<float>::infinity(), std::numeric_limits<float>::infinity(), std::numeric_limits<float>::infinity() };
    V3 mx{ -std::numeric_limits<float>::infinity(), -std::numeric_limits<float>::infinity(), -std::numeric_limits<float>::infinity() };
    void expand(V3 p) {
        mn = { fmin_ref(mn.x, p.x), fmin_ref(mn.y, p.y), fmin_ref(mn.z, p.z) };
        mx = { fmax_ref(mx.x, p.x), fmax_ref(mx.y, p.y), fmax_ref(mx.z, p.z) };
    }
    void unite(const Box& o) {
        mn = { fmin_ref(mn.x, o.mn.x), fmin_ref(mn.y, o.mn.y), fmin_ref(mn.z, o.mn.z) };
        mx = { fmax_ref(mx.x, o.mx.x), fmax_ref(mx.y, o.mx.y), fmax_ref(mx.z, o.mx.z) };
    }
    // box.h:33-49: zero-extent axes are padded by +-1e-4
    void make_valid() {
        const V3 size = mx - mn;
        if (size.x == 0.0f) { mn.x -= 0.0001f; mx.x += 0.0001f; }
        if (size.y == 0.0f) { mn.y -= 0.0001f; mx.y += 0.0001f; }
        if (size.z == 0.0f) { mn.z -= 0.0001f; mx.z += 0.0001f; }
    }
    bool valid() const { return !(mn.x >= mx.x) && !(mn.y >= mx.y) && !(mn.z >= mx.z); }
    // cave/engine/math/box.cpp:7-17
    float surface_area() const {
        if (!valid()) return 0.0f;
        const V3 s = mx - mn;
        const float sx = std::fabs(s.x), sy = std::fabs(s.y), sz = std::fabs(s.z);
        return 2.0f * (sx * sy + sx * sz + sy * sz);
    }
    V3 center() const { return 0.5f * (mn + mx); }
};

// bvh_accel.cpp:76-87 -- ties keep the lower axis
inline int dominant_axis(const Box& b) {
    const V3 span = b.mx - b.mn;
    int axis = 0;
    if (span[axis] < span.y) axis = 1;
    if (span[axis] < span.z) axis = 2;
    return axis;
}

struct Builder {
    const uint32_t* tri;  // 3 per triangle
    const V3* pos;
    std::vector<Box> tri_box;
    std::vector<V3> centroid;
    co_bvh* out;
    int cap;
    int counter = 0;
    bool failed = false;

    // One call == one BvhBuilder::ConstructHelper (bvh_accel.cpp:115-219).  `miss`
    // is what DiscoverIdx (:221-235) would find by walking the parents: the
    // nearest ancestor's right child created after this node.
    void construct(std::vector<uint32_t>& ids, int depth, int miss) {
        if (failed) return;
        if (depth > 64 || ids.empty() || counter >= cap) {  // :117-120 trap; empty set -> DEV_ASSERT :139
            failed = true;
            return;
        }
        const int n = (int)ids.size();
        Box box;  // AABBFromTriangles :89-99
        for (uint32_t t : ids) {
            box.expand(pos[tri[3 * t]]);
            box.expand(pos[tri[3 * t + 1]]);
            box.expand(pos[tri[3 * t + 2]]);
        }
        box.make_valid();
        const int me = counter++;  // creation order == pre-order == array slot (:124, :237-263)
        co_bvh& node = out[me];
        std::memset(&node, 0, sizeof(node));
        node.min[0] = box.mn.x; node.min[1] = box.mn.y; node.min[2] = box.mn.z;
        node.max[0] = box.mx.x; node.max[1] = box.mx.y; node.max[2] = box.mx.z;
        node.miss_idx = miss;
        node.triangle_index = -1;
        if (n == 1) {  // :128-132
            node.leaf = 1;
            node.triangle_index = (int)ids[0];
            node.hit_idx = miss;  // :234 leaf: hit == miss
            return;
        }
        const float parent_sa = box.surface_area();
        std::vector<uint32_t> left, right;
        if (n <= 4 || parent_sa == 0.0f) {  // :138-142 -> SplitByAxis :101-113
            if (!(parent_sa > 0.0f)) {
                failed = true;
                return;
            }
            const int axis = dominant_axis(box);
            // std::sort on <=4 elements is libstdc++'s insertion sort, which is stable;
            // key = AABB centre (BvhSorter :38-55), not the centroid.
            std::vector<uint32_t> sorted = ids;
            for (int i = 1; i < n; ++i) {
                const uint32_t v = sorted[i];
                const float key = tri_box[v].center()[axis];
                int j = i - 1;
                while (j >= 0 && key < tri_box[sorted[j]].center()[axis]) {
                    sorted[j + 1] = sorted[j];
                    --j;
                }
                sorted[j + 1] = v;
            }
            const int mid = n / 2;
            left.assign(sorted.begin(), sorted.begin() + mid);
            right.assign(sorted.begin() + mid, sorted.end());
        } else {
            constexpr int kBuckets = 12;
            Box cbox;  // :151-156
            for (uint32_t t : ids) cbox.expand(centroid[t]);
            cbox.make_valid();
            const int axis = dominant_axis(cbox);
            const float tmin = cbox.mn[axis];
            const float tmax = cbox.mx[axis];
            int count[kBuckets] = {};
            Box bbox[kBuckets];
            for (uint32_t t : ids) {  // :162-169
                const float tmp = ((centroid[t][axis] - tmin) * kBuckets) / (tmax - tmin);
                int slot = (int)tmp;
                slot = slot < 0 ? 0 : (slot > kBuckets - 1 ? kBuckets - 1 : slot);
                ++count[slot];
                bbox[slot].unite(tri_box[t]);
            }
            float costs[kBuckets - 1];
            for (int i = 0; i < kBuckets - 1; ++i) {  // :171-186
                Box b0, b1;
                int c0 = 0, c1 = 0;
                for (int j = 0; j <= i; ++j) { b0.unite(bbox[j]); c0 += count[j]; }
                for (int j = i + 1; j < kBuckets; ++j) { b1.unite(bbox[j]); c1 += count[j]; }
                costs[i] = 0.125f + (c0 * b0.surface_area() + c1 * b1.surface_area()) / parent_sa;
            }
            int split = 0;  // :188-196 first strict minimum
            float best = costs[0];
            for (int i = 0; i < kBuckets - 1; ++i) {
                if (costs[i] < best) { split = i; best = costs[i]; }
            }
            for (uint32_t t : ids) {  // :203-214 -- note the different association
                float tmp = (centroid[t][axis] - tmin) / (tmax - tmin);
                tmp *= kBuckets;
                int slot = (int)tmp;
                slot = slot < 0 ? 0 : (slot > kBuckets - 1 ? kBuckets - 1 : slot);
                (slot <= split ? left : right).push_back(t);
            }
        }
        // the right child is created only after the whole left subtree, so its index is
        // not known yet: construct the left subtree with a placeholder miss link and patch.
        const int left_index = counter;
        node.hit_idx = left_index;  // :234
        const int patch_from = counter;
        construct(left, depth + 1, /*miss=*/-2);
        if (failed) return;
        const int right_index = counter;
        for (int i = patch_from; i < right_index; ++i) {
            if (out[i].miss_idx == -2) out[i].miss_idx = right_index;
            if (out[i].hit_idx == -2) out[i].hit_idx = right_index;
        }
        construct(right, depth + 1, miss);
    }
};

}  // namespace

CO_API int caveoracle_bvh_build(const uint32_t* indices, int index_count, const float* positions, int vertex_count,
                                co_bvh* out_nodes, int cap) {
    if (index_count <= 0 || index_count % 3 != 0 || !indices || !positions || !out_nodes) return -3;
    const int n = index_count / 3;
    if (cap < 2 * n - 1) return -2;
    for (int i = 0; i < index_count; ++i) {
        if (indices[i] >= (uint32_t)vertex_count) return -3;
    }
    Builder b;
    b.tri = indices;
    b.pos = reinterpret_cast<const V3*>(positions);
    b.out = out_nodes;
    b.cap = cap;
    b.tri_box.resize(n);
    b.centroid.resize(n);
    for (int i = 0; i < n; ++i) {  // BvhBuilder ctor :57-74
        const V3 a = b.pos[indices[3 * i]], bb = b.pos[indices[3 * i + 1]], c = b.pos[indices[3 * i + 2]];
        b.centroid[i] = (1.0f / 3.0f) * (a + bb + c);
        Box box;
        box.expand(a);
        box.expand(bb);
        box.expand(c);
        box.make_valid();
        b.tri_box[i] = box;
    }
    std::vector<uint32_t> ids(n);
    for (int i = 0; i < n; ++i) ids[i] = (uint32_t)i;
    b.construct(ids, 0, -1);
    if (b.failed) return -1;
    return b.counter;
}

// glm::inverse for mat4 (thirdparty/glm/glm/detail/func_matrix.inl:388-446, glm
// commit 2d4c4b4), cofactor expansion; m[c*4+r] column-major.  The grouping of
// every product/sum follows glm so the fp32 result is identical.
CO_API void caveoracle_mat4_inverse(const float* m, float* out) {
#define M(c, r) m[(c) * 4 + (r)]
    const float c00 = M(2, 2) * M(3, 3) - M(3, 2) * M(2, 3);
    const float c02 = M(1, 2) * M(3, 3) - M(3, 2) * M(1, 3);
    const float c03 = M(1, 2) * M(2, 3) - M(2, 2) * M(1, 3);
    const float c04 = M(2, 1) * M(3, 3) - M(3, 1) * M(2, 3);
    const float c06 = M(1, 1) * M(3, 3) - M(3, 1) * M(1, 3);
    const float c07 = M(1, 1) * M(2, 3) - M(2, 1) * M(1, 3);
    const float c08 = M(2, 1) * M(3, 2) - M(3, 1) * M(2, 2);
    const float c10 = M(1, 1) * M(3, 2) - M(3, 1) * M(1, 2);
    const float c11 = M(1, 1) * M(2, 2) - M(2, 1) * M(1, 2);
    const float c12 = M(2, 0) * M(3, 3) - M(3, 0) * M(2, 3);
    const float c14 = M(1, 0) * M(3, 3) - M(3, 0) * M(1, 3);
    const float c15 = M(1, 0) * M(2, 3) - M(2, 0) * M(1, 3);
    const float c16 = M(2, 0) * M(3, 2) - M(3, 0) * M(2, 2);
    const float c18 = M(1, 0) * M(3, 2) - M(3, 0) * M(1, 2);
    const float c19 = M(1, 0) * M(2, 2) - M(2, 0) * M(1, 2);
    const float c20 = M(2, 0) * M(3, 1) - M(3, 0) * M(2, 1);
    const float c22 = M(1, 0) * M(3, 1) - M(3, 0) * M(1, 1);
    const float c23 = M(1, 0) * M(2, 1) - M(2, 0) * M(1, 1);
    const float f0[4] = { c00, c00, c02, c03 }, f1[4] = { c04, c04, c06, c07 }, f2[4] = { c08, c08, c10, c11 };
    const float f3[4] = { c12, c12, c14, c15 }, f4[4] = { c16, c16, c18, c19 }, f5[4] = { c20, c20, c22, c23 };
    const float v0[4] = { M(1, 0), M(0, 0), M(0, 0), M(0, 0) }, v1[4] = { M(1, 1), M(0, 1), M(0, 1), M(0, 1) };
    const float v2[4] = { M(1, 2), M(0, 2), M(0, 2), M(0, 2) }, v3[4] = { M(1, 3), M(0, 3), M(0, 3), M(0, 3) };
    float inv[16];
    for (int r = 0; r < 4; ++r) {
        const float sa = (r & 1) ? -1.0f : 1.0f, sb = -sa;
        inv[0 * 4 + r] = (v1[r] * f0[r] - v2[r] * f1[r] + v3[r] * f2[r]) * sa;
        inv[1 * 4 + r] = (v0[r] * f0[r] - v2[r] * f3[r] + v3[r] * f4[r]) * sb;
        inv[2 * 4 + r] = (v0[r] * f1[r] - v1[r] * f3[r] + v3[r] * f5[r]) * sa;
        inv[3 * 4 + r] = (v0[r] * f2[r] - v1[r] * f4[r] + v2[r] * f5[r]) * sb;
    }
    const float d0 = M(0, 0) * inv[0], d1 = M(0, 1) * inv[4], d2 = M(0, 2) * inv[8], d3 = M(0, 3) * inv[12];
    const float det = (d0 + d1) + (d2 + d3);
    const float inv_det = 1.0f / det;
    for (int i = 0; i < 16; ++i) out[i] = inv[i] * inv_det;
#undef M
}

namespace {

struct Ray {
    V3 origin;
    float t;
    V3 direction;
    V3 inv_dir;
};

// shared_path_tracer.h:60-67
inline uint32_t wang_hash(uint32_t& s) {
    s = (s ^ 61u) ^ (s >> 16);
    s *= 9u;
    s = s ^ (s >> 4);
    s *= 0x27d4eb2du;
    s = s ^ (s >> 15);
    return s;
}
// :70-72 (float(h) rounds to 24 bits; /2^32 is exact; can return 1.0)
inline float rnd(uint32_t& s) { return (float)wang_hash(s) * (1.0f / 4294967296.0f); }
// :75-82
inline V3 random_unit_vector(uint32_t& s) {
    const float z = rnd(s) * 2.0f - 1.0f;
    const float a = rnd(s) * 6.283185307179586f;
    const float r = std::sqrt(1.0f - z * z);
    const float x = r * std::cos(a);
    const float y = r * std::sin(a);
    return { x, y, z };
}

// glm mat4 * vec4 (thirdparty/glm/glm/detail/type_mat4x4.inl:562-573):
// (col0*x + col1*y) + (col2*z + col3*w), xyz only.
inline V3 mat_mul(const float* m, V3 v, float w) {
    V3 r;
    r.x = (m[0] * v.x + m[4] * v.y) + (m[8] * v.z + m[12] * w);
    r.y = (m[1] * v.x + m[5] * v.y) + (m[9] * v.z + m[13] * w);
    r.z = (m[2] * v.x + m[6] * v.y) + (m[10] * v.z + m[14] * w);
    return r;
}

struct Scene {
    const co_vertex* vertices;
    const co_index* indices;
    const co_bvh* bvhs;
    const co_mesh* meshes;
    int mesh_count;
    const co_material* materials;
};

constexpr float kEpsilon = 1.1920929e-7f;  // :24
constexpr float kRayTMax = 999999999.0f;   // :27

// HitTriangle :88-130 (Moeller-Trumbore, back faces culled by det < EPSILON)
inline bool hit_triangle(const Scene& sc, Ray& ray, int tri_id, float& out_u, float& out_v) {
    const int* idx = sc.indices[tri_id].tri;
    const V3 A = *reinterpret_cast<const V3*>(sc.vertices[idx[0]].position);
    const V3 B = *reinterpret_cast<const V3*>(sc.vertices[idx[1]].position);
    const V3 C = *reinterpret_cast<const V3*>(sc.vertices[idx[2]].position);
    const V3 AB = B - A;
    const V3 AC = C - A;
    const V3 P = cross(ray.direction, AC);
    const float det = dot(AB, P);
    if (det < kEpsilon) return false;
    const float inv_det = 1.0f / det;
    const V3 AO = ray.origin - A;
    const V3 Q = cross(AO, AB);
    const float u = dot(AO, P) * inv_det;
    const float v = dot(ray.direction, Q) * inv_det;
    if (u < 0.0f || v < 0.0f || u + v > 1.0f) return false;
    const float t = dot(AC, Q) * inv_det;
    if (t >= ray.t || t < kEpsilon) return false;
    ray.t = t;
    out_u = u;
    out_v = v;
    return true;
}

// HitBvh :133-149; min/max written as the C++ header evaluates them
// (a<b?a:b / a>b?a:b, common.h:50-58) -- differs from IEEE minNum only on NaN.
inline bool hit_bvh(const Ray& ray, const float* bmin, const float* bmax) {
    const V3 mn{ bmin[0], bmin[1], bmin[2] }, mx{ bmax[0], bmax[1], bmax[2] };
    const V3 t0 = (mn - ray.origin) * ray.inv_dir;
    const V3 t1 = (mx - ray.origin) * ray.inv_dir;
    const V3 ts{ fmin_ref(t0.x, t1.x), fmin_ref(t0.y, t1.y), fmin_ref(t0.z, t1.z) };
    const V3 tb{ fmax_ref(t0.x, t1.x), fmax_ref(t0.y, t1.y), fmax_ref(t0.z, t1.z) };
    const float tmin = fmax_ref(ts.x, fmax_ref(ts.y, ts.z));
    const float tmax = fmin_ref(tb.x, fmin_ref(tb.y, tb.z));
    return (tmin < tmax) && (ray.t > tmin);
}

struct Hit {
    int mesh = -1, tri = -1;
    float u = 0.0f, v = 0.0f;
};

// HitScene :152-189: every mesh, stackless threaded walk, fixed order.
inline Hit hit_scene(const Scene& sc, Ray& ray, long long* cnt) {
    Hit res;
    for (int mesh_id = 0; mesh_id < sc.mesh_count; ++mesh_id) {
        const co_mesh& mesh = sc.meshes[mesh_id];
        Ray local;
        local.origin = mat_mul(mesh.transform_inv, ray.origin, 1.0f);
        local.direction = mat_mul(mesh.transform_inv, ray.direction, 0.0f);
        local.inv_dir = { 1.0f / local.direction.x, 1.0f / local.direction.y, 1.0f / local.direction.z };
        local.t = ray.t;
        ++cnt[3];
        int bvh_index = mesh.root_bvh_id;
        while (bvh_index >= 0) {
            const co_bvh& bvh = sc.bvhs[bvh_index];
            ++cnt[1];
            if (hit_bvh(local, bvh.min, bvh.max)) {
                if (bvh.triangle_index != -1) {
                    float u, v;
                    ++cnt[2];
                    if (hit_triangle(sc, local, bvh.triangle_index, u, v)) {
                        res.tri = bvh.triangle_index;
                        res.mesh = mesh_id;
                        res.u = u;
                        res.v = v;
                        ray.t = local.t;
                    }
                }
                bvh_index = bvh.hit_idx;
            } else {
                bvh_index = bvh.miss_idx;
            }
        }
    }
    return res;
}

// RayColor :191-239 with MAX_BOUNCE as a run-time value.
inline V3 ray_color(const Scene& sc, Ray& ray, uint32_t& state, int max_bounce, long long* cnt) {
    V3 radiance{ 0, 0, 0 };
    V3 throughput{ 1, 1, 1 };
    for (int i = 0; i < max_bounce; ++i) {
        ++cnt[0];
        const Hit hit = hit_scene(sc, ray, cnt);
        if (hit.mesh >= 0 && hit.tri >= 0) {
            ray.origin = ray.origin + ray.t * ray.direction;
            ray.t = kRayTMax;
            const co_mesh& mesh = sc.meshes[hit.mesh];
            const int* idx = sc.indices[hit.tri].tri;
            const V3 n1 = *reinterpret_cast<const V3*>(sc.vertices[idx[0]].normal);
            const V3 n2 = *reinterpret_cast<const V3*>(sc.vertices[idx[1]].normal);
            const V3 n3 = *reinterpret_cast<const V3*>(sc.vertices[idx[2]].normal);
            V3 n = n1 + hit.u * (n2 - n1) + hit.v * (n3 - n1);
            n = normalize(mat_mul(mesh.transform, n, 0.0f));
            const co_material& mat = sc.materials[mesh.material_id];
            const V3 diffuse_color{ mat.base_color[0], mat.base_color[1], mat.base_color[2] };
            const float reflect_chance = rnd(state) > mat.metallic ? 0.0f : 1.0f;
            const V3 diffuse_dir = normalize(n + random_unit_vector(state));
            V3 reflect_dir = ray.direction - 2.0f * dot(n, ray.direction) * n;  // GLSL reflect(I, N)
            reflect_dir = normalize(lerp(reflect_dir, diffuse_dir, mat.roughness * mat.roughness));
            ray.direction = normalize(lerp(diffuse_dir, reflect_dir, reflect_chance));
            ray.inv_dir = { 1.0f / ray.direction.x, 1.0f / ray.direction.y, 1.0f / ray.direction.z };
            radiance = radiance + V3{ mat.emissive[0], mat.emissive[1], mat.emissive[2] } * throughput;
            throughput = throughput * diffuse_color;
        } else {
            radiance = radiance + V3{ 0.3f, 0.3f, 0.3f } * throughput;
            break;
        }
    }
    return radiance;
}

// main() path_tracer.cs.glsl:11-60
inline void shade_pixel(const Scene& sc, const co_camera& cam, float cam_distance, int max_bounce, int x, int y, int w,
                        int h, int frame_index, int dirty, float* pixel, long long* cnt) {
    uint32_t seed = ((uint32_t)x * 1973u + (uint32_t)y * 9277u + (uint32_t)frame_index * 26699u) | 1u;
    const float j0 = rnd(seed) - 0.5f;
    const float j1 = rnd(seed) - 0.5f;
    const float u = ((float)x + j0) / (float)w;
    const float v = ((float)y + j1) / (float)h;
    const float sx = 2.0f * u - 1.0f;
    float sy = 2.0f * v - 1.0f;
    const float aspect = (float)w / (float)h;
    sy /= aspect;
    const V3 right{ cam.right[0], cam.right[1], cam.right[2] };
    const V3 up{ cam.up[0], cam.up[1], cam.up[2] };
    const V3 forward{ cam.forward[0], cam.forward[1], cam.forward[2] };
    Ray ray;
    ray.origin = { cam.position[0], cam.position[1], cam.position[2] };
    ray.direction = normalize(right * sx + up * sy + forward * cam_distance);
    ray.inv_dir = { 1.0f / ray.direction.x, 1.0f / ray.direction.y, 1.0f / ray.direction.z };
    ray.t = kRayTMax;
    const V3 radiance = ray_color(sc, ray, seed, max_bounce, cnt);
    if (dirty) {
        pixel[0] = radiance.x; pixel[1] = radiance.y; pixel[2] = radiance.z; pixel[3] = 1.0f;
    } else {
        const float weight = pixel[3];
        const float nw = weight + 1.0f;
        pixel[0] = (radiance.x + weight * pixel[0]) / nw;
        pixel[1] = (radiance.y + weight * pixel[1]) / nw;
        pixel[2] = (radiance.z + weight * pixel[2]) / nw;
        pixel[3] = nw;
    }
}

}  // namespace

CO_API long long caveoracle_trace_render(const co_vertex* vertices, const co_index* indices, const co_bvh* bvhs,
                                         const co_mesh* meshes, int mesh_count, const co_material* materials,
                                         const co_camera* camera, int max_bounce, int width, int height, int x0,
                                         int y0, int x1, int y1, int first_frame, int n_frames, int reset_first,
                                         float* accum, int threads, long long* counters, double* ms) {
    const Scene sc{ vertices, indices, bvhs, meshes, mesh_count, materials };
    // camDistance = tan(fovDegree * PI / 180): the full fov, as the shader has it (:31-32)
    const float cam_distance = std::tan(camera->fov_degree * 3.141592653589793f / 180.0f);
    std::atomic<int> next_row{ y0 };
    std::atomic<long long> total[CO_N_COUNTERS];
    for (auto& t : total) t = 0;
    const auto t0 = std::chrono::steady_clock::now();
    auto worker = [&]() {
        long long cnt[CO_N_COUNTERS] = {};
        for (;;) {
            const int y = next_row.fetch_add(1);
            if (y >= y1) break;
            for (int x = x0; x < x1; ++x) {
                float* pixel = accum + 4 * ((size_t)y * width + x);
                for (int f = 0; f < n_frames; ++f) {
                    shade_pixel(sc, *camera, cam_distance, max_bounce, x, y, width, height, first_frame + f,
                                (f == 0 && reset_first) ? 1 : 0, pixel, cnt);
                }
            }
        }
        for (int i = 0; i < CO_N_COUNTERS; ++i) total[i] += cnt[i];
    };
    if (threads <= 1) {
        worker();
    } else {
        std::vector<std::thread> pool;
        for (int i = 0; i < threads; ++i) pool.emplace_back(worker);
        for (auto& t : pool) t.join();
    }
    const auto t1 = std::chrono::steady_clock::now();
    if (ms) *ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    if (counters) {
        for (int i = 0; i < CO_N_COUNTERS; ++i) counters[i] = total[i].load();
    }
    return total[0].load();
}

// post_process.ps.glsl:72-76
CO_API void caveoracle_tonemap(const float* accum, float* out, long long pixel_count) {
    for (long long i = 0; i < pixel_count; ++i) {
        for (int c = 0; c < 3; ++c) {
            float v = accum[4 * i + c];
            v = v / (v + 1.0f);
            out[4 * i + c] = std::pow(v, 1.0f / 2.2f);
        }
        out[4 * i + 3] = 1.0f;
    }
}
