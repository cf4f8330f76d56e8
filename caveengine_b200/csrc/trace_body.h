// Path-trace pass: RNG, camera ray, ray/BVH traversal, ray/triangle test, BSDF bounce
// loop and the progressive accumulator -- the device restatement of
//   cave/shader/glsl/path_tracer.cs.glsl:11-60   (compute entry)
//   cave/shader/shared_path_tracer.h:60-239      (WangHash .. RayColor)
// Two traversals share every other line of code:
//   * hit_scene_reference(): HitScene (:152-189) literally -- every mesh, stackless
//     hit/miss-link walk over the 48-byte GpuPtBvh array.  Parity instrument.
//   * hit_scene_fast(): the production traversal -- TLAS over instances, near-first
//     ordered descent with a short stack over 32-byte sibling-pair nodes, leaves of up
//     to 4 pre-gathered triangles.  Same closest hit (see cavept.h CAVEPT_TRACE_*).
// The ray/triangle arithmetic is one shared function evaluated in the reference's
// operation order without FMA contraction, so both traversals produce bit-identical
// (t, u, v) for the same (ray, triangle).
#pragma once
#include "cv_common.h"

namespace cavept {

constexpr float kEpsilon = 1.1920929e-7f;  // EPSILON, shared_path_tracer.h:24
constexpr float kRayTMax = 999999999.0f;   // RAY_T_MAX, :27
#ifndef CV_LEAF_MAX
#define CV_LEAF_MAX 2
#endif
constexpr int kFastLeafMax = CV_LEAF_MAX;            // leaves of the production tree hold at most this many triangles (measured on B200: 2 beats 4 and 1)
#ifndef CV_TRI_HINT
#define CV_TRI_HINT 1   // L1 hint of the triangle / normal fetches (cv_common.h ld256_hint)
#endif
#ifndef CV_NODE_HINT
#define CV_NODE_HINT 2  // L1 hint of the sibling-pair fetch (measured on B200, config 3: triangles no_allocate +1.7 %, nodes evict_last +0.2 %; evict_first on triangles and a streaming accumulator: no gain)
#endif
#ifndef CV_STACK_SIZE
#define CV_STACK_SIZE 40
#endif
constexpr int kStackSize = CV_STACK_SIZE;  // per-lane entries in shared memory (20 KB per 128-thread block; 40 costs nothing against 32 on B200, 48 costs 8 %: measured)

// 32-byte node of the production layout.  Siblings are adjacent (a pair is 64 bytes).
//   a = the node's own reference (see make_leaf_ref / make_instance_ref below)
//   b = subtree triangle count (informational)
struct alignas(16) FastNode {
    float mn[3];
    int a;
    float mx[3];
    int b;
};
static_assert(sizeof(FastNode) == 32, "FastNode");

// What the tracer actually fetches: a sibling PAIR in 32 bytes = ONE 256-bit load (one L1
// wavefront per lane instead of two).  The twelve box planes are 15-bit coordinates on a
// per-axis grid over the scene box (cell = extent / 32767, rounded outward at build time), two
// per word; r0/r1 are the children's references (interior = index of the child's own QPair).
//   q[0] = c0.min.x | c0.max.x << 16   q[1] = c0.min.y | c0.max.y << 16   q[2] = c0.min.z | c0.max.z << 16
//   q[3..5] the same for child 1
// A coordinate g decodes with one PRMT to the float 0.5 + g / 65536 (bits 0x3F000000 | g << 8),
// and the slab planes are t = v * I + O with I = 65536 / d_grid, O = -(32768 + o_grid) / d_grid.
// Min and max of an axis share a word, so the PRMT selector itself picks the plane the ray
// enters through (near) or leaves through (far) from the sign of the direction: no min/max
// per axis in the slab test.
struct alignas(32) QPair {
    uint32_t q[6];
    int r0, r1;
};
static_assert(sizeof(QPair) == 32, "QPair");

// Four-wide node (CAVEPT_KERNEL_WIDE4): record k is sibling pair k with interior children replaced by THEIR pairs,
// largest box first, until there are four children or only leaves (wbvh_body.h wb_body_q4).  Same 15-bit grid and
// plane packing as QPair (child c: q[3c .. 3c+2]); r[c] = child reference (interior = index of the child's own record =
// its pair index, so QPair k and Q4Node k root the same subtree; leaf reference; kRefSentinel with an inverted box
// for an unused slot).  64 bytes = two 256-bit loads from one line: half the dependent fetch rounds of the pair walk.
struct alignas(64) Q4Node {
    uint32_t q[12];
    int r[4];
};
static_assert(sizeof(Q4Node) == 64, "Q4Node");
// A four-wide step defers up to three children, so a walk can need 1.5 x the tree's levels of stack; entries past
// the shared-memory stack go to a small per-thread spill array (never touched on the scenes measured).
constexpr int kWideSpill = 64;
constexpr int kWideMaxLevels = ((kStackSize + kWideSpill - 4) * 2) / 3;  // deeper trees use the pair walk
constexpr float kGridCells = 32767.0f;
constexpr uint32_t kSelLo = 0x7104u, kSelHi = 0x7324u, kSelFlip = 0x0220u;  // __byte_perm selectors: low / high half -> bytes 1..2

// Pre-gathered triangle, 3 x float4.  In `ftris`: v0 = world-space A, v1 = AB = B - A,
// v2 = AC = C - A (the two subtractions HitTriangle starts with, :94-95, done once at build
// time -- same fp32 operations, same bits); v0.w = instance, v1.w = scene-global triangle id.
// In `fnormals`: the three vertex normals times the instance's world matrix (w = 0, not
// normalised: the shader interpolates first, transforms, then normalises, :206-210 -- the
// transform is linear, so transforming the corners first is the same up to fp32 rounding and
// bit-identical for an identity instance); v0.w = instance, v1.w = material id.  Records are padded
// to 64 bytes so one triangle is two 256-bit loads from a single cache line.
struct alignas(64) FastTri {
    F4 v0, v1, v2, pad;
};
static_assert(sizeof(FastTri) == 64, "FastTri");

// Texture coordinates of a slot's triangle, in the slot's own vertex order (a mirroring instance swaps B and C).
struct alignas(32) FastUv {
    float uv[6];
    float pad[2];
};
static_assert(sizeof(FastUv) == 32, "FastUv");
struct TexInfo {
    int offset;  // first texel in tex_data
    int width, height;
    int pad;
};

struct TraceParams {
    // reference-layout arrays (what BindPathTracerData binds, path_tracer.cpp:324-331)
    const GpuPtVertex* vertices;
    const GpuPtIndex* indices;
    const GpuPtBvh* bvhs;
    const GpuPtMesh* meshes;
    const GpuPtMaterial* materials;
    int mesh_count;  // c_ptObjectCount
    // production layout
    const FastNode* fnodes;  // fp32 boxes as built (read-back / tests); the tracer reads qpairs
    const QPair* qpairs;
    const Q4Node* q4nodes;                // four-wide records over the same tree (CAVEPT_KERNEL_WIDE4); may be null otherwise
    float grid_origin[3], grid_scale[3];  // grid coordinate = (world - origin) * scale, in [0, 32767]
    const FastTri* ftris;
    const FastTri* fnormals;  // per FastTri slot: the three world-transformed vertex normals, v0.w = instance, v1.w = material
    int root_ref;             // node reference of the root
    // PerFrameConstantBuffer (cbuffer.hlsl.h:193-207)
    float cam_pos[3], cam_right[3], cam_up[3], cam_fwd[3];
    float cam_distance;  // tan(c_cameraFovDegree * PI / 180), path_tracer.cs.glsl:31-32
    int width, height;
    int tile_x0, tile_y0, tile_x1, tile_y1;  // pixel rectangle a launch covers (c_tileOffset, cbuffer.hlsl.h:183); all 0 = the full image
    int max_bounce;
    int first_frame, n_frames;
    int reset_first;  // c_sceneDirty for the first frame of the launch
    int accum_sums;   // 0: rgb = running mean, a = n (path_tracer.cs.glsl:46-59); 1: rgb = sum, a = n (multi-GPU group mode)
    F4* accum;
    unsigned long long* counters;  // [0] segments, [1] node visits, [2] triangle tests
    // textured materials / environment map (cavept.h "textured materials ..."; the kTex instantiation only)
    const FastUv* fuvs;        // per FastTri slot: the three vertices' texture coordinates; null: no mesh has any
    const TexInfo* tex_table;  // one entry per texture of the set
    const F4* tex_data;        // all texels, RGBA32F
    const int* mat_maps;       // per material: base-colour map, metal-rough map (texture index or -1)
    const F4* env;             // equirectangular environment map (rgb, w unused); null: the constant sky
    int env_w, env_h;
};

struct Ray {
    F3 origin;
    float t;
    F3 direction;
    F3 inv_dir;
};

struct Hit {
    int mesh;
    int tri;
    float u, v;
};

// ---- RNG, shared_path_tracer.h:60-82 -------------------------------------------------
CV_HD uint32_t wang_hash(uint32_t& s) {
    s = (s ^ 61u) ^ (s >> 16);
    s *= 9u;
    s = s ^ (s >> 4);
    s *= 0x27d4eb2du;
    s = s ^ (s >> 15);
    return s;
}
CV_HD float rnd(uint32_t& s) { return (float)wang_hash(s) * (1.0f / 4294967296.0f); }
CV_HD F3 random_unit_vector(uint32_t& s) {
    const float z = rnd(s) * 2.0f - 1.0f;
    const float a = rnd(s) * 6.283185307179586f;
    const float r = cv_sqrt(1.0f - z * z);
    float sn, cs;
#if defined(__CUDA_ARCH__)
    sincosf(a, &sn, &cs);
#else
    sn = sinf(a);
    cs = cosf(a);
#endif
    return f3(r * cs, r * sn, z);
}

// glm mat4 * vec4, xyz of the result (thirdparty/glm/glm/detail/type_mat4x4.inl:562-573)
CV_HD F3 mat_mul(const float* m, F3 v, float w) {
    return f3((m[0] * v.x + m[4] * v.y) + (m[8] * v.z + m[12] * w), (m[1] * v.x + m[5] * v.y) + (m[9] * v.z + m[13] * w),
              (m[2] * v.x + m[6] * v.y) + (m[10] * v.z + m[14] * w));
}
CV_HD F3 recip3(F3 d) { return f3(cv_rcp(d.x), cv_rcp(d.y), cv_rcp(d.z)); }

// ---- HitTriangle, :88-130 ------------------------------------------------------------
CV_HD bool hit_triangle_edges(F3 A, F3 AB, F3 AC, F3 origin, F3 direction, float& ray_t, float& out_u, float& out_v);
CV_HD bool hit_triangle(F3 A, F3 B, F3 C, F3 origin, F3 direction, float& ray_t, float& out_u, float& out_v) {
    return hit_triangle_edges(A, B - A, C - A, origin, direction, ray_t, out_u, out_v);
}
CV_HD bool hit_triangle_edges(F3 A, F3 AB, F3 AC, F3 origin, F3 direction, float& ray_t, float& out_u, float& out_v) {
    const F3 P = cross3(direction, AC);
    const float det = dot3(AB, P);
    if (det < kEpsilon) return false;  // back faces culled
    const float inv_det = cv_rcp(det);
    const F3 AO = origin - A;
    const F3 Q = cross3(AO, AB);
    const float u = dot3(AO, P) * inv_det;
    const float v = dot3(direction, Q) * inv_det;
    if (u < 0.0f || v < 0.0f || u + v > 1.0f) return false;
    const float t = dot3(AC, Q) * inv_det;
    if (t >= ray_t || t < kEpsilon) return false;
    ray_t = t;
    out_u = u;
    out_v = v;
    return true;
}

// ---- HitBvh, :133-149 ---------------------------------------------------------------
CV_HD void slab_interval_reference(const Ray& r, F3 mn, F3 mx, float& tmin, float& tmax) {
    const F3 t0 = (mn - r.origin) * r.inv_dir;
    const F3 t1 = (mx - r.origin) * r.inv_dir;
    const F3 ts = f3(min_ref(t0.x, t1.x), min_ref(t0.y, t1.y), min_ref(t0.z, t1.z));
    const F3 tb = f3(max_ref(t0.x, t1.x), max_ref(t0.y, t1.y), max_ref(t0.z, t1.z));
    tmin = max_ref(ts.x, max_ref(ts.y, ts.z));
    tmax = min_ref(tb.x, min_ref(tb.y, tb.z));
}
CV_HD bool hit_bvh_reference(const Ray& r, F3 mn, F3 mx) {
    float tmin, tmax;
    slab_interval_reference(r, mn, mx, tmin, tmax);
    return (tmin < tmax) && (r.t > tmin);
}

// Unit probe of the two routines above on caller-supplied rays, triangles and boxes (SURVEY.md
// section 8 row a20: cross-check against the engine's CPU twins TestIntersection::RayTriangle /
// RayAabb, cave/engine/math/intersection.cpp:18-70).  rays: origin xyz, direction xyz, t;
// out: triangle hit, t, u, v, box hit, tmin, tmax, 0.
CV_HD void body_probe(const float* rays, const float* tris, const float* boxes, float* out, int i) {
    const float* r = rays + 7 * (size_t)i;
    const float* t = tris + 9 * (size_t)i;
    const float* b = boxes + 6 * (size_t)i;
    Ray ray;
    ray.origin = f3(r[0], r[1], r[2]);
    ray.direction = f3(r[3], r[4], r[5]);
    ray.inv_dir = recip3(ray.direction);
    ray.t = r[6];
    float ray_t = ray.t, u = 0.0f, v = 0.0f;
    const bool tri_hit = hit_triangle(f3(t[0], t[1], t[2]), f3(t[3], t[4], t[5]), f3(t[6], t[7], t[8]), ray.origin, ray.direction, ray_t, u, v);
    float tmin, tmax;
    slab_interval_reference(ray, f3(b[0], b[1], b[2]), f3(b[3], b[4], b[5]), tmin, tmax);
    float* o = out + 8 * (size_t)i;
    o[0] = tri_hit ? 1.0f : 0.0f;
    o[1] = ray_t;
    o[2] = u;
    o[3] = v;
    o[4] = hit_bvh_reference(ray, f3(b[0], b[1], b[2]), f3(b[3], b[4], b[5])) ? 1.0f : 0.0f;
    o[5] = tmin;
    o[6] = tmax;
    o[7] = 0.0f;
}

// ---- HitScene, :152-189, literal -----------------------------------------------------
template<bool kCount>
CV_D Hit hit_scene_reference(const TraceParams& p, Ray& ray, unsigned long long* cnt) {
    Hit res;
    res.mesh = -1;
    res.tri = -1;
    res.u = 0.0f;
    res.v = 0.0f;
    for (int mesh_id = 0; mesh_id < p.mesh_count; ++mesh_id) {
        const GpuPtMesh& mesh = p.meshes[mesh_id];
        Ray local;
        local.origin = mat_mul(mesh.transformInv, ray.origin, 1.0f);
        local.direction = mat_mul(mesh.transformInv, ray.direction, 0.0f);
        local.inv_dir = recip3(local.direction);
        local.t = ray.t;
        int bvh_index = mesh.rootBvhId;
        while (bvh_index >= 0) {
            const GpuPtBvh& bvh = p.bvhs[bvh_index];
            if (kCount) ++cnt[1];
            if (hit_bvh_reference(local, f3(bvh.min[0], bvh.min[1], bvh.min[2]), f3(bvh.max[0], bvh.max[1], bvh.max[2]))) {
                const int tri = bvh.triangleIndex;
                if (tri != -1) {
                    if (kCount) ++cnt[2];
                    const int* idx = p.indices[tri].tri;
                    const GpuPtVertex& va = p.vertices[idx[0]];
                    const GpuPtVertex& vb = p.vertices[idx[1]];
                    const GpuPtVertex& vc = p.vertices[idx[2]];
                    float u, v;
                    if (hit_triangle(f3(va.position[0], va.position[1], va.position[2]), f3(vb.position[0], vb.position[1], vb.position[2]),
                                     f3(vc.position[0], vc.position[1], vc.position[2]), local.origin, local.direction, local.t, u, v)) {
                        res.tri = tri;
                        res.mesh = mesh_id;
                        res.u = u;
                        res.v = v;
                        ray.t = local.t;
                    }
                }
                bvh_index = bvh.hitIdx;
            } else {
                bvh_index = bvh.missIdx;
            }
        }
    }
    return res;
}

// ---- production traversal ---------------------------------------------------------------
// One world-space tree over every (instance, triangle) pair (wbvh_body.h).  Node references
// (one int): interior = index of the QPair holding the node's two children;
// leaf = sign bit | (count - 1) << 29 | first FastTri slot.  (In the fp32 FastNode array the
// builder emits, an interior reference is the index of the child pair's first record = 2 * QPair index.)
constexpr int kRefSentinel = 0x7fffffff;  // bottom of the stack: nothing left to traverse
CV_HD bool ref_is_interior(int r) { return (uint32_t)r < (uint32_t)kRefSentinel; }
CV_HD int make_leaf_ref(int first_slot, int count) { return (int)(0x80000000u | ((uint32_t)(count - 1) << 29) | (uint32_t)first_slot); }

#if defined(__CUDA_ARCH__)
#define CV_WARP_ANY(pred) __any_sync(0xffffffffu, (pred))
#define CV_WARP_COUNT(pred) __popc(__ballot_sync(0xffffffffu, (pred)))
#elif defined(CV_WARP_EMU)  // tests/emu/emu.cpp: 32 host threads per warp, votes through a barrier
bool warp_emu_any(bool pred);
int warp_emu_count(bool pred);
int warp_emu_draw(int* counter, bool want);
#define CV_WARP_ANY(pred) cavept::warp_emu_any(pred)
#define CV_WARP_COUNT(pred) cavept::warp_emu_count(pred)
#else  // the emulator steps one lane at a time: a warp of one
#define CV_WARP_ANY(pred) (pred)
#define CV_WARP_COUNT(pred) ((pred) ? 1 : 0)
#endif

// Per-lane traversal state of the production kernel.  The stack lives in shared memory,
// lane-strided ([depth][thread]), so any mix of depths is bank-conflict free.
struct Walk {
    int* stack;
    int stride;
    int* spill;      // four-wide walk only: entries kStackSize.. of the stack (per-thread local memory)
    int sp;          // entries in shared memory; entry 0 holds the sentinel
    int top;         // the top of the stack lives in a register: a pop never waits for shared memory
    int ref;         // node reference being processed (interior, leaf, or the sentinel)
    int q0, q1;      // postponed leaves (0 = empty); q0 is the older
    F3 inv, oinv;    // slab planes: t = v * inv + oinv for a decoded grid coordinate v (see QPair)
    uint32_t near_x, near_y, near_z;  // PRMT selectors of the entry plane per axis (exit = ^ kSelFlip)
    int hit_slot;
    float hit_u, hit_v;
};

// 1/d for the slab test.  Not IEEE-critical: boxes are padded (wbvh_body.h) and a zero
// component is replaced by a huge finite slope so that 0 * inf never makes a NaN.
CV_HD float slab_recip(float d) {
    const float a = fabsf(d) > 1e-30f ? d : (f2u(d) >> 31 ? -1e-30f : 1e-30f);
    return rcp_approx(a);
}

CV_D void walk_begin(const TraceParams& p, Walk& w, const Ray& ray) {
    w.stack[0] = kRefSentinel;
    w.sp = 1;
    w.top = kRefSentinel;
    w.ref = p.root_ref;
    w.q0 = 0;
    w.q1 = 0;
    w.hit_slot = -1;
    w.hit_u = 0.0f;
    w.hit_v = 0.0f;
    // the ray in grid coordinates; t is unchanged because origin and direction scale together
    const F3 og = f3((ray.origin.x - p.grid_origin[0]) * p.grid_scale[0], (ray.origin.y - p.grid_origin[1]) * p.grid_scale[1], (ray.origin.z - p.grid_origin[2]) * p.grid_scale[2]);
    const F3 ig = f3(slab_recip(ray.direction.x * p.grid_scale[0]), slab_recip(ray.direction.y * p.grid_scale[1]), slab_recip(ray.direction.z * p.grid_scale[2]));
    w.inv = 65536.0f * ig;
    w.oinv = f3(-((32768.0f + og.x) * ig.x), -((32768.0f + og.y) * ig.y), -((32768.0f + og.z) * ig.z));
    w.near_x = (f2u(ig.x) >> 31) ? kSelHi : kSelLo;
    w.near_y = (f2u(ig.y) >> 31) ? kSelHi : kSelLo;
    w.near_z = (f2u(ig.z) >> 31) ? kSelHi : kSelLo;
}

CV_D void walk_push(Walk& w, int r) {
    w.stack[w.sp * w.stride] = w.top;
    ++w.sp;
    w.top = r;
}
CV_D int walk_pop(Walk& w) {
    const int r = w.top;
    w.top = w.stack[(--w.sp) * w.stride];  // consumed by the NEXT pop or push: the latency is hidden
    return r;
}

// 15-bit grid coordinate (low or high half of `w`, picked by the selector) -> the float
// 0.5 + g / 65536: one PRMT
CV_HD float q_sel(uint32_t w, uint32_t sel) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(__byte_perm(w, 0x3F000000u, sel));
#else
    return u2f(0x3F000000u | ((sel == kSelLo ? (w & 0xffffu) : (w >> 16)) << 8));
#endif
}

// A leaf held in `ref` moves to the queue when there is room; the lane pops its next node.
CV_D void walk_park_leaf(Walk& w) {
    if (w.ref < 0 && w.q1 == 0) {
        if (w.q0 == 0) w.q0 = w.ref; else w.q1 = w.ref;
        w.ref = walk_pop(w);
    }
}

// One interior step: fetch the sibling pair (QPair), slab-test both boxes against [0, ray.t], descend
// into the nearer hit child and push the other.  Same interval logic as HitBvh (:133-149);
// boxes entirely behind the origin or beyond the current hit are rejected.
template<bool kCount, bool kStream = false>
CV_D void walk_step(const TraceParams& p, Walk& w, float ray_t, unsigned long long* cnt) {
    const F8 c = ld256_hint<kStream ? CV_NODE_HINT : 0>(p.qpairs + w.ref);  // the whole pair: one 256-bit load
    if (kCount) cnt[1] += 2;
    const uint32_t far_x = w.near_x ^ kSelFlip, far_y = w.near_y ^ kSelFlip, far_z = w.near_z ^ kSelFlip;
    const uint32_t q0 = f2u(c.a), q1 = f2u(c.b), q2 = f2u(c.c), q3 = f2u(c.d), q4 = f2u(c.e), q5 = f2u(c.f);
    const float ax = cv_fma(q_sel(q0, w.near_x), w.inv.x, w.oinv.x), bx = cv_fma(q_sel(q0, far_x), w.inv.x, w.oinv.x);
    const float ay = cv_fma(q_sel(q1, w.near_y), w.inv.y, w.oinv.y), by = cv_fma(q_sel(q1, far_y), w.inv.y, w.oinv.y);
    const float az = cv_fma(q_sel(q2, w.near_z), w.inv.z, w.oinv.z), bz = cv_fma(q_sel(q2, far_z), w.inv.z, w.oinv.z);
    const float cx = cv_fma(q_sel(q3, w.near_x), w.inv.x, w.oinv.x), dx = cv_fma(q_sel(q3, far_x), w.inv.x, w.oinv.x);
    const float cy = cv_fma(q_sel(q4, w.near_y), w.inv.y, w.oinv.y), dy = cv_fma(q_sel(q4, far_y), w.inv.y, w.oinv.y);
    const float cz = cv_fma(q_sel(q5, w.near_z), w.inv.z, w.oinv.z), dz = cv_fma(q_sel(q5, far_z), w.inv.z, w.oinv.z);
    const float n0 = fmaxf(fmaxf(ax, ay), fmaxf(az, 0.0f)), f0 = fminf(fminf(bx, by), fminf(bz, ray_t));
    const float n1 = fmaxf(fmaxf(cx, cy), fmaxf(cz, 0.0f)), f1 = fminf(fminf(dx, dy), fminf(dz, ray_t));
    const bool h0 = n0 <= f0, h1 = n1 <= f1;
    const int r0 = (int)f2u(c.g), r1 = (int)f2u(c.h);
    const bool second_first = h1 && (!h0 || n1 < n0);
    // Branch-free bookkeeping (the three outcomes would otherwise be three divergent paths, each
    // costing its own issue slots): spill the register top to the free slot above the stack and
    // read the entry below it unconditionally, then select.  While a lane is stepping the sentinel
    // is still stacked, so sp >= 1 and both addresses are its own.
    const bool both = h0 && h1, none = !(h0 || h1);
    const int near_ref = second_first ? r1 : r0, far_ref = second_first ? r0 : r1;
    w.stack[w.sp * w.stride] = w.top;
    const int below = w.stack[(w.sp - 1) * w.stride];
    w.ref = none ? w.top : near_ref;
    w.top = both ? far_ref : (none ? below : w.top);
    w.sp += both ? 1 : (none ? -1 : 0);
}

// ---- four-wide walk (Q4Node) ------------------------------------------------------------------------------
CV_D void wide_store(Walk& w, int idx, int v) {
    if (idx < kStackSize) w.stack[idx * w.stride] = v; else w.spill[idx - kStackSize] = v;
}
CV_D int wide_load(const Walk& w, int idx) { return idx < kStackSize ? w.stack[idx * w.stride] : w.spill[idx - kStackSize]; }
CV_D int walk_pop_wide(Walk& w) {
    const int r = w.top;
    --w.sp;
    w.top = wide_load(w, w.sp);
    return r;
}
CV_D void walk_park_leaf_wide(Walk& w) {
    if (w.ref < 0 && w.q1 == 0) {
        if (w.q0 == 0) w.q0 = w.ref; else w.q1 = w.ref;
        w.ref = walk_pop_wide(w);
    }
}
// compare-exchange of (entry distance, reference): afterwards a is the nearer
CV_HD void wide_order(float& ka, int& ra, float& kb, int& rb) {
    const bool swap = kb < ka;
    const float k0 = swap ? kb : ka, k1 = swap ? ka : kb;
    const int r0 = swap ? rb : ra, r1 = swap ? ra : rb;
    ka = k0; ra = r0; kb = k1; rb = r1;
}
// One four-wide step: two 256-bit loads, four slab tests (same planes and selectors as walk_step), the hit children
// sorted by entry distance; descend into the nearest and defer the others farthest-first, so pops come nearest-first.
template<bool kCount>
CV_D void walk_step_wide(const TraceParams& p, Walk& w, float ray_t, unsigned long long* cnt) {
    const Q4Node* node = p.q4nodes + w.ref;
    const F8 a = ld256(node);            // q[0..7]
    const F8 b = ld256(&node->q[8]);     // q[8..11], r[0..3]
    if (kCount) cnt[1] += 4;
    const uint32_t far_x = w.near_x ^ kSelFlip, far_y = w.near_y ^ kSelFlip, far_z = w.near_z ^ kSelFlip;
    const uint32_t qw[12] = { f2u(a.a), f2u(a.b), f2u(a.c), f2u(a.d), f2u(a.e), f2u(a.f), f2u(a.g), f2u(a.h), f2u(b.a), f2u(b.b), f2u(b.c), f2u(b.d) };
    float key[4];
    int ref[4] = { (int)f2u(b.e), (int)f2u(b.f), (int)f2u(b.g), (int)f2u(b.h) };
    int n_hit = 0;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const float nx = cv_fma(q_sel(qw[3 * c], w.near_x), w.inv.x, w.oinv.x), fx = cv_fma(q_sel(qw[3 * c], far_x), w.inv.x, w.oinv.x);
        const float ny = cv_fma(q_sel(qw[3 * c + 1], w.near_y), w.inv.y, w.oinv.y), fy = cv_fma(q_sel(qw[3 * c + 1], far_y), w.inv.y, w.oinv.y);
        const float nz = cv_fma(q_sel(qw[3 * c + 2], w.near_z), w.inv.z, w.oinv.z), fz = cv_fma(q_sel(qw[3 * c + 2], far_z), w.inv.z, w.oinv.z);
        const float tn = fmaxf(fmaxf(nx, ny), fmaxf(nz, 0.0f)), tf = fminf(fminf(fx, fy), fminf(fz, ray_t));
        const bool h = tn <= tf;
        key[c] = h ? tn : CV_INF;
        n_hit += h ? 1 : 0;
    }
    wide_order(key[0], ref[0], key[1], ref[1]);
    wide_order(key[2], ref[2], key[3], ref[3]);
    wide_order(key[0], ref[0], key[2], ref[2]);
    wide_order(key[1], ref[1], key[3], ref[3]);
    wide_order(key[1], ref[1], key[2], ref[2]);
    // stack: old top spilled to entry sp, then the deferred children farthest first, the nearest of them in `top`
    const int up1 = n_hit == 4 ? ref[3] : ref[2];
    int below;
    if (w.sp + 2 < kStackSize) {  // the usual case: every slot touched is in shared memory; branch-free as in walk_step
        w.stack[w.sp * w.stride] = w.top;
        below = w.stack[(w.sp - 1) * w.stride];
        w.stack[(w.sp + 1) * w.stride] = up1;
        w.stack[(w.sp + 2) * w.stride] = ref[2];
    } else {
        wide_store(w, w.sp, w.top);
        below = wide_load(w, w.sp - 1);
        if (n_hit >= 3) wide_store(w, w.sp + 1, up1);
        if (n_hit == 4) wide_store(w, w.sp + 2, ref[2]);
    }
    const int old_top = w.top;
    w.ref = n_hit > 0 ? ref[0] : old_top;
    w.top = n_hit >= 2 ? ref[1] : (n_hit == 1 ? old_top : below);
    w.sp += n_hit - 1;
}

// One triangle of the oldest postponed leaf.  World-space vertices: for an identity
// instance this is HitScene's arithmetic bit for bit; for a transformed instance the
// reference intersects in mesh space with the inverse-transformed ray (:160-166), which is
// the same (t, u, v) up to fp32 rounding of the transform.
template<bool kCount, bool kStream = false>
CV_D void walk_test_triangle(const TraceParams& p, Walk& w, const Ray& ray, float& ray_t, unsigned long long* cnt) {
    const int leaf = w.q0;
    const int slot = leaf & 0x1fffffff;
    const int more = (leaf >> 29) & 3;
    // kStream (chosen per launch by the host): a many-sample launch over a scene far larger than L1 is dominated by
    // incoherent bounce segments -- a triangle record will not be touched again before it is evicted, so it is fetched
    // without allocating an L1 line, and the sibling pairs are marked evict-last (+1.8 % on config 3).  Small scenes
    // live in L1 and the camera rays of a one-sample launch hit the same triangles: those launches keep the plain
    // loads (the hint applied everywhere cost config 2 a factor 2.7 and config 5 10 %; choosing per lane by bounce, or
    // per launch by a run-time branch around the loads, cost config 3 7 %: all measured -- hence a template parameter).
    const F8 t0 = ld256_hint<kStream ? CV_TRI_HINT : 0>(&p.ftris[slot]);     // A.xyz, instance, AB.xyz, triangle id
    const F8 t1 = ld256_hint<kStream ? CV_TRI_HINT : 0>(&p.ftris[slot].v2);  // AC.xyz (same cache line)
    if (kCount) ++cnt[2];
    float u, v;
    if (hit_triangle_edges(f3(t0.a, t0.b, t0.c), f3(t0.e, t0.f, t0.g), f3(t1.a, t1.b, t1.c), ray.origin, ray.direction, ray_t, u, v)) {
        w.hit_slot = slot;
        w.hit_u = u;
        w.hit_v = v;
    }
    if (more > 0) {
        w.q0 = make_leaf_ref(slot + 1, more);
    } else {
        w.q0 = w.q1;
        w.q1 = 0;
    }
}

// ---- one RayColor iteration after the scene query (:198-235).  Returns true when the
//      path ends (miss, or MAX_BOUNCE segments traced). ---------------------------------
// `n` is the interpolated normal already multiplied by the instance's world matrix (:206-210,
// before the normalize); `mat` the instance's material.
CV_D bool shade_hit(const TraceParams& p, bool is_hit, F3 n, const GpuPtMaterial& mat, Ray& ray, uint32_t& state, F3& radiance, F3& throughput, int& bounce) {
    if (is_hit) {
        ray.origin = ray.origin + ray.t * ray.direction;
        ray.t = kRayTMax;
        n = normalize3(n);
        const float reflect_chance = rnd(state) > mat.metallic ? 0.0f : 1.0f;
        const F3 diffuse_dir = normalize3(n + random_unit_vector(state));
        F3 reflect_dir = ray.direction - 2.0f * dot3(n, ray.direction) * n;
        reflect_dir = normalize3(lerp3(reflect_dir, diffuse_dir, mat.roughness * mat.roughness));
        ray.direction = normalize3(lerp3(diffuse_dir, reflect_dir, reflect_chance));
        ray.inv_dir = recip3(ray.direction);
        radiance = radiance + f3(mat.emissive[0], mat.emissive[1], mat.emissive[2]) * throughput;
        throughput = throughput * f3(mat.baseColor[0], mat.baseColor[1], mat.baseColor[2]);
        ++bounce;
        return bounce >= p.max_bounce;
    }
    radiance = radiance + f3(0.3f, 0.3f, 0.3f) * throughput;
    return true;
}

CV_D bool shade_segment(const TraceParams& p, const Hit& hit, Ray& ray, uint32_t& state, F3& radiance, F3& throughput, int& bounce) {
    const bool is_hit = hit.mesh >= 0 && hit.tri >= 0;
    F3 n = f3(0, 0, 0);
    GpuPtMaterial mat = {};
    if (is_hit) {
        const int* idx = p.indices[hit.tri].tri;
        const GpuPtVertex& va = p.vertices[idx[0]];
        const GpuPtVertex& vb = p.vertices[idx[1]];
        const GpuPtVertex& vc = p.vertices[idx[2]];
        const F3 n1 = f3(va.normal[0], va.normal[1], va.normal[2]);
        const F3 n2 = f3(vb.normal[0], vb.normal[1], vb.normal[2]);
        const F3 n3 = f3(vc.normal[0], vc.normal[1], vc.normal[2]);
        const GpuPtMesh& mesh = p.meshes[hit.mesh];
        n = mat_mul(mesh.transform, n1 + hit.u * (n2 - n1) + hit.v * (n3 - n1), 0.0f);
        mat = p.materials[mesh.materialId];
    }
    return shade_hit(p, is_hit, n, mat, ray, state, radiance, throughput, bounce);
}

// ---- textured shading: the reference's switched-off RayColor (glsl/path_tracer.cs.glsl:62-112) ------------------
// bilinear, repeat wrap: texel centres at (i + 0.5) / size
CV_HD int tex_wrap(int i, int n) {
    i %= n;
    return i < 0 ? i + n : i;
}
CV_D F4 tex_sample(const F4* texels, int w, int h, float u, float v) {
    const float x = u * (float)w - 0.5f, y = v * (float)h - 0.5f;
    const float fx0 = floorf(x), fy0 = floorf(y);
    const float fx = x - fx0, fy = y - fy0;
    const int x0 = tex_wrap((int)fx0, w), y0 = tex_wrap((int)fy0, h), x1 = tex_wrap((int)fx0 + 1, w), y1 = tex_wrap((int)fy0 + 1, h);
    const F4 a = texels[(size_t)y0 * w + x0], b = texels[(size_t)y0 * w + x1], c = texels[(size_t)y1 * w + x0], d = texels[(size_t)y1 * w + x1];
    F4 r;
    r.x = (a.x + fx * (b.x - a.x)) + fy * ((c.x + fx * (d.x - c.x)) - (a.x + fx * (b.x - a.x)));
    r.y = (a.y + fx * (b.y - a.y)) + fy * ((c.y + fx * (d.y - c.y)) - (a.y + fx * (b.y - a.y)));
    r.z = (a.z + fx * (b.z - a.z)) + fy * ((c.z + fx * (d.z - c.z)) - (a.z + fx * (b.z - a.z)));
    r.w = (a.w + fx * (b.w - a.w)) + fy * ((c.w + fx * (d.w - c.w)) - (a.w + fx * (b.w - a.w)));
    return r;
}
// equirectangular lookup of a unit direction (sample_spherical_map as the engine writes it, hlsl/to_cube_map.ps.hlsl:9-15)
CV_D F3 env_sample(const TraceParams& p, F3 d) {
    const float u = atan2f(d.z, d.x) * 0.1591f + 0.5f;
    const float sy = d.y < -1.0f ? -1.0f : (d.y > 1.0f ? 1.0f : d.y);
    const float v = asinf(sy) * 0.3183f + 0.5f;
    const F4 c = tex_sample(p.env, p.env_w, p.env_h, u, v);
    return f3(c.x, c.y, c.z);
}
// One iteration of that RayColor after the scene query.  slot / (u, v) identify the hit; `mat` is the instance's
// material and is overridden by its maps.  Returns true when the path ends.
CV_D bool shade_hit_textured(const TraceParams& p, bool is_hit, int slot, float hu, float hv, int material, F3 n, GpuPtMaterial mat, Ray& ray, uint32_t& state, F3& radiance,
                             F3& throughput, int& bounce) {
    if (!is_hit) {
        const F3 sky = p.env ? env_sample(p, normalize3(ray.direction)) : f3(0.3f, 0.3f, 0.3f);
        radiance = radiance + sky * throughput;
        return true;
    }
    const int base_map = p.mat_maps ? p.mat_maps[2 * material] : -1, mr_map = p.mat_maps ? p.mat_maps[2 * material + 1] : -1;
    if (p.fuvs && (base_map >= 0 || mr_map >= 0)) {
        const FastUv& t = p.fuvs[slot];
        const float tu = t.uv[0] + hu * (t.uv[2] - t.uv[0]) + hv * (t.uv[4] - t.uv[0]);
        const float tv = t.uv[1] + hu * (t.uv[3] - t.uv[1]) + hv * (t.uv[5] - t.uv[1]);
        if (base_map >= 0) {
            const TexInfo ti = p.tex_table[base_map];
            const F4 c = tex_sample(p.tex_data + ti.offset, ti.width, ti.height, tu, tv);
            if (c.w <= 0.01f) {  // alpha cut-out: `continue` -- the ray goes on from the hit point, the bounce is spent (:78-80)
                ray.origin = ray.origin + ray.t * ray.direction;
                ray.t = kRayTMax;
                ++bounce;
                return bounce >= p.max_bounce;
            }
            mat.baseColor[0] = c.x;
            mat.baseColor[1] = c.y;
            mat.baseColor[2] = c.z;
        }
        if (mr_map >= 0) {
            const TexInfo ti = p.tex_table[mr_map];
            const F4 c = tex_sample(p.tex_data + ti.offset, ti.width, ti.height, tu, tv);
            mat.metallic = c.z;
            mat.roughness = c.y;
        }
    }
    return shade_hit(p, true, n, mat, ray, state, radiance, throughput, bounce);
}

// ---- main(), path_tracer.cs.glsl:11-44: seed, jitter, camera ray -----------------------
CV_D void start_path(const TraceParams& p, int x, int y, int frame_index, Ray& ray, uint32_t& seed) {
    seed = ((uint32_t)x * 1973u + (uint32_t)y * 9277u + (uint32_t)frame_index * 26699u) | 1u;
    const float j0 = rnd(seed) - 0.5f;
    const float j1 = rnd(seed) - 0.5f;
    const float u = cv_div((float)x + j0, (float)p.width);
    const float v = cv_div((float)y + j1, (float)p.height);
    const float sx = 2.0f * u - 1.0f;
    float sy = 2.0f * v - 1.0f;
    const float aspect = cv_div((float)p.width, (float)p.height);
    sy = cv_div(sy, aspect);
    const F3 right = f3(p.cam_right[0], p.cam_right[1], p.cam_right[2]);
    const F3 up = f3(p.cam_up[0], p.cam_up[1], p.cam_up[2]);
    const F3 fwd = f3(p.cam_fwd[0], p.cam_fwd[1], p.cam_fwd[2]);
    ray.origin = f3(p.cam_pos[0], p.cam_pos[1], p.cam_pos[2]);
    ray.direction = normalize3(right * sx + up * sy + fwd * p.cam_distance);
    ray.inv_dir = recip3(ray.direction);
    ray.t = kRayTMax;
}

// ---- accumulate, path_tracer.cs.glsl:46-59 ----------------------------------------------
CV_HD void accumulate(F4& acc, F3 radiance, bool dirty) {
    if (dirty) {
        acc.x = radiance.x;
        acc.y = radiance.y;
        acc.z = radiance.z;
        acc.w = 1.0f;
    } else {
        const float weight = acc.w;
        const float nw = weight + 1.0f;
        acc.x = cv_div(radiance.x + weight * acc.x, nw);
        acc.y = cv_div(radiance.y + weight * acc.y, nw);
        acc.z = cv_div(radiance.z + weight * acc.z, nw);
        acc.w = nw;
    }
}

// (sum, n) instead of (mean, n): what every GPU of a group keeps so that ONE sum-reduce composites the frame
CV_HD void accumulate_sum(F4& acc, F3 radiance, bool dirty) {
    acc.x = dirty ? radiance.x : acc.x + radiance.x;
    acc.y = dirty ? radiance.y : acc.y + radiance.y;
    acc.z = dirty ? radiance.z : acc.z + radiance.z;
    acc.w = dirty ? 1.0f : acc.w + 1.0f;
}
// One code path for both modes: mean = (radiance + w * mean) / (w + 1) as the shader writes it; sum = (radiance + 1 * sum) / 1,
// which is radiance + sum exactly.  (A run-time branch between two inlined bodies here cost 13 % of a one-sample launch
// on config 5 -- the accumulator load lost its place in the schedule; measured, profiles/r02z2_call.sh.)
CV_HD void accumulate_mode(const int accum_sums, F4& acc, F3 radiance, bool dirty) {
    if (dirty) {
        acc.x = radiance.x;
        acc.y = radiance.y;
        acc.z = radiance.z;
        acc.w = 1.0f;
    } else {
        const float weight = acc.w;
        const float nw = weight + 1.0f;
        const float m = accum_sums ? 1.0f : weight, d = accum_sums ? 1.0f : nw;
        acc.x = cv_div(radiance.x + m * acc.x, d);
        acc.y = cv_div(radiance.y + m * acc.y, d);
        acc.z = cv_div(radiance.z + m * acc.z, d);
        acc.w = nw;
    }
}

// One pixel, frames [first_frame, first_frame + n_frames), REFERENCE_WALK traversal.  The
// segment loop is flat: a lane whose path ends starts its next sample at once.
// Returns the number of segments traced; visit counters go to cnt[1..2] when kCount.
template<bool kCount>
CV_D unsigned int body_trace_pixel_reference(const TraceParams& p, int x, int y, unsigned long long* cnt) {
    if (x >= p.width || y >= p.height || p.n_frames <= 0) return 0u;
    const size_t pix = (size_t)y * (size_t)p.width + (size_t)x;
    F4 acc = p.accum[pix];
    Ray ray;
    uint32_t seed;
    int f = 0;
    int bounce = 0;
    unsigned int segments = 0;
    F3 radiance = f3(0.0f, 0.0f, 0.0f), throughput = f3(1.0f, 1.0f, 1.0f);
    start_path(p, x, y, p.first_frame, ray, seed);
    for (;;) {
        const Hit hit = hit_scene_reference<kCount>(p, ray, cnt);
        ++segments;
        if (shade_segment(p, hit, ray, seed, radiance, throughput, bounce)) {
            accumulate_mode(p.accum_sums, acc, radiance, f == 0 && p.reset_first != 0);
            if (++f == p.n_frames) break;
            radiance = f3(0.0f, 0.0f, 0.0f);
            throughput = f3(1.0f, 1.0f, 1.0f);
            bounce = 0;
            start_path(p, x, y, p.first_frame + f, ray, seed);
        }
    }
    p.accum[pix] = acc;
    return segments;
}

// Production kernel body: ONE loop per warp in which every lane is descending the tree,
// has triangles waiting, or has finished its segment -- whatever sample and bounce it is on.
// Each trip of the loop the warp votes and runs the unit of work most lanes can take part in:
//   A  one interior step.  A lane that reaches a leaf parks it (two-entry queue) and keeps
//      descending -- speculative traversal: the pending triangles can only shorten ray.t, so
//      the extra nodes visited never change the closest hit;
//   B  one triangle of the oldest parked leaf, once at least as many lanes wait on triangles
//      as can still step;
//   C  shade + next segment (or next sample) for lanes whose traversal is complete, in batches
//      of >= kShadeBatch lanes so the long shading sequence is not replayed per lane.
// The order in which ONE lane visits its nodes and triangles does not depend on its
// neighbours, so images are deterministic.
// (Tried and measured slower on B200: a separate batched phase for accumulate + next sample --
// the lanes waiting for it cost more step-phase utilisation than the batching saves.)
// A vote costs ~45 warp instructions at full width, a step 55: every vote is therefore followed
// by several steps (6 in the quorum loop, 4 in the minority branch; a lane that cannot step idles
// through the rest of the group), and shading waits for 24 lanes -- the long shading sequence at
// 10 of 32 lanes was 16 % of the kernel.  ncu source-page counts before/after: r01g profiles.
// The groups only change WHEN a lane's next visit happens, never which visits it makes.
#ifndef CV_STEP_QUORUM
#define CV_STEP_QUORUM 20
#endif
#ifndef CV_TIGHT_STEPS
#define CV_TIGHT_STEPS 6
#endif
#ifndef CV_MINORITY_STEPS
#define CV_MINORITY_STEPS 4
#endif
#ifndef CV_PARK_EVERY
#define CV_PARK_EVERY 2  // parked leaves are collected before every second step of a group (a lane that reaches a leaf in between idles one step; the 13-instruction collect is full-width work)
#endif
#ifndef CV_TRI_PER_VOTE
#define CV_TRI_PER_VOTE 1
#endif
#ifndef CV_SHADE_BATCH
#define CV_SHADE_BATCH 24
#endif
constexpr int kStepQuorum = CV_STEP_QUORUM;  // of 32 lanes: keep stepping while at least this many can
constexpr int kShadeBatch = CV_SHADE_BATCH;

// ---- job queue of the persistent launch (kQueue): a job = one pixel for all n_frames of the launch, numbered
//      tile-major (8x4 pixel tiles: the 32 jobs a warp draws together are one coherent tile) and handed out by ONE
//      warp-aggregated atomicAdd on a global counter.  A lane whose pixel is finished draws the next job instead of
//      idling until its warp's slowest lane ends, so a launch has no tail however few frames it carries. --------------
struct TraceJobs {
    int tiles_x;         // 8x4 tiles across the launch rectangle
    int n_jobs;          // tiles_x * tiles_y * 32
    int x0, y0, x1, y1;  // launch rectangle in pixels
};
CV_HD TraceJobs trace_jobs(const TraceParams& p) {
    TraceJobs j;
    j.x0 = p.tile_x1 > 0 ? p.tile_x0 : 0;
    j.y0 = p.tile_x1 > 0 ? p.tile_y0 : 0;
    j.x1 = p.tile_x1 > 0 ? p.tile_x1 : p.width;
    j.y1 = p.tile_x1 > 0 ? p.tile_y1 : p.height;
    j.tiles_x = (j.x1 - j.x0 + 7) / 8;
    const int tiles_y = (j.y1 - j.y0 + 3) / 4;
    j.n_jobs = p.n_frames > 0 ? j.tiles_x * tiles_y * 32 : 0;
    return j;
}
// job -> pixel; false for the padding of a ragged rectangle
CV_HD bool trace_job_pixel(const TraceJobs& j, int job, int& x, int& y) {
    const int tile = job >> 5, in = job & 31;
    x = j.x0 + (tile % j.tiles_x) * 8 + (in & 7);
    y = j.y0 + (tile / j.tiles_x) * 4 + (in >> 3);
    return x < j.x1 && y < j.y1;
}
// Warp-aggregated draw: the lanes with `want` get consecutive job numbers (0x7fffffff for the others).
CV_D int trace_draw_jobs(int* counter, bool want) {
#if defined(__CUDA_ARCH__)
    const unsigned m = __ballot_sync(0xffffffffu, want);
    if (m == 0u) return 0x7fffffff;
    const int lane = (int)(threadIdx.x & 31u);
    const int leader = __ffs(m) - 1;
    int base = 0;
    if (lane == leader) base = atomicAdd(counter, __popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    return want ? base + __popc(m & ((1u << lane) - 1u)) : 0x7fffffff;
#elif defined(CV_WARP_EMU)
    return warp_emu_draw(counter, want);
#else
    if (!want) return 0x7fffffff;
    const int j = *counter;
    *counter = j + 1;
    return j;
#endif
}

// kQueue = false: the lane owns pixel (x, y) for the whole launch (grid of pixel tiles).
// kQueue = true:  persistent launch; (x, y) is the lane's first job (x >= width: none yet) and finished pixels are
//                 replaced from the job queue.
//
// kWide: the four-wide walk over Q4Node records (walk_step_wide) instead of the pair walk; everything else is shared.
#ifndef CV_WIDE_TIGHT_STEPS
#define CV_WIDE_TIGHT_STEPS 3
#endif
#ifndef CV_WIDE_MINORITY_STEPS
#define CV_WIDE_MINORITY_STEPS 2
#endif
#ifndef CV_WIDE_PARK_EVERY
#define CV_WIDE_PARK_EVERY 1
#endif
template<bool kCount, bool kWide, bool kStream = false>
CV_D void walk_step_any(const TraceParams& p, Walk& w, float ray_t, unsigned long long* cnt) {
    if (kWide) walk_step_wide<kCount>(p, w, ray_t, cnt); else walk_step<kCount, kStream>(p, w, ray_t, cnt);
}
template<bool kWide>
CV_D void walk_park_any(Walk& w) {
    if (kWide) walk_park_leaf_wide(w); else walk_park_leaf(w);
}

template<bool kCount, int kQuorum = kStepQuorum, int kBatch = kShadeBatch, bool kQueue = false, bool kWide = false, bool kTex = false, bool kStream = false>
CV_D unsigned int body_trace_pixel_fast(const TraceParams& p, int x, int y, int* stack, int stride, unsigned long long* cnt, int* job_counter = nullptr, int* spill = nullptr) {
    constexpr int kTight = kWide ? CV_WIDE_TIGHT_STEPS : CV_TIGHT_STEPS;
    constexpr int kMinority = kWide ? CV_WIDE_MINORITY_STEPS : CV_MINORITY_STEPS;
    constexpr int kParkEvery = kWide ? CV_WIDE_PARK_EVERY : CV_PARK_EVERY;
    const bool valid = x < p.width && y < p.height && p.n_frames > 0;
    size_t pix = valid ? (size_t)y * (size_t)p.width + (size_t)x : 0;
    Ray ray;
    uint32_t seed = 1u;
    int f = 0;
    int bounce = 0;
    unsigned int segments = 0;
    F3 radiance = f3(0.0f, 0.0f, 0.0f), throughput = f3(1.0f, 1.0f, 1.0f);
    bool done = kQueue ? false : !valid;
    bool vacant = kQueue && !valid;  // kQueue: the lane holds no pixel and draws a job at the next shading phase
    bool queue_open = kQueue;        // warp-uniform
    Walk w;
    w.stack = stack;
    w.stride = stride;
    w.spill = spill;
    ray.origin = f3(0, 0, 0);
    ray.direction = f3(0, 0, 1);
    ray.inv_dir = f3(0, 0, 1);
    ray.t = kRayTMax;
    if (valid) start_path(p, x, y, p.first_frame, ray, seed);
    walk_begin(p, w, ray);
    if (!valid) w.ref = kRefSentinel;
    for (;;) {
        // ---- A, the tight loop: one ballot per step while a quorum can step -----------------
        for (;;) {
            walk_park_any<kWide>(w);
            const bool can_step = ref_is_interior(w.ref);
            if (CV_WARP_COUNT(can_step) < kQuorum) break;
            if (can_step) walk_step_any<kCount, kWide, kStream>(p, w, ray.t, cnt);
            // second step of the trip without a vote of its own (measured +2 %): the lanes that can
            // still step do, the others would have idled through it anyway
#pragma unroll
            for (int k = 1; k < kTight; ++k) {
                if ((k % kParkEvery) == 0) walk_park_any<kWide>(w);
                if (ref_is_interior(w.ref)) walk_step_any<kCount, kWide, kStream>(p, w, ray.t, cnt);
            }
        }
        // ---- the vote ------------------------------------------------------------------------
        const bool can_step = ref_is_interior(w.ref);
        const bool has_tri = w.q0 != 0;
        const bool need_shade = !done && !can_step && !has_tri;  // ref is the sentinel: a parked-full lane has_tri
        const int n_step = CV_WARP_COUNT(can_step);
        const int n_tri = CV_WARP_COUNT(has_tri);
        const int n_shade = CV_WARP_COUNT(need_shade);
        if (n_step + n_tri + n_shade == 0) break;  // every lane has rendered all its frames
        if (n_tri > 0 && n_tri >= n_step) {
            if (has_tri) walk_test_triangle<kCount, kStream>(p, w, ray, ray.t, cnt);
#pragma unroll
            for (int k = 1; k < CV_TRI_PER_VOTE; ++k) {
                if (w.q0 != 0) walk_test_triangle<kCount, kStream>(p, w, ray, ray.t, cnt);
            }
        } else if (n_shade >= kBatch || n_step == 0) {
            bool new_sample = false, want_job = false;
            if (need_shade) {
                if (kQueue && vacant) {
                    want_job = true;
                } else {
                    ++segments;
                    const bool is_hit = w.hit_slot >= 0;
                    F3 n = f3(0, 0, 0);
                    GpuPtMaterial mat = {};
                    if (is_hit) {
                        // world-space vertex normals and the material id, pre-gathered at build time
                        const F8 a = ld256_hint<kStream ? CV_TRI_HINT : 0>(&p.fnormals[w.hit_slot]);     // n1.xyz, instance, n2.xyz, material
                        const F8 b = ld256_hint<kStream ? CV_TRI_HINT : 0>(&p.fnormals[w.hit_slot].v2);  // n3.xyz
                        const F3 n1 = f3(a.a, a.b, a.c), n2 = f3(a.e, a.f, a.g), n3 = f3(b.a, b.b, b.c);
                        n = n1 + w.hit_u * (n2 - n1) + w.hit_v * (n3 - n1);
                        const F8 m = ld256(&p.materials[(int)f2u(a.h)]);
                        mat.baseColor[0] = m.a; mat.baseColor[1] = m.b; mat.baseColor[2] = m.c; mat.roughness = m.d;
                        mat.emissive[0] = m.e; mat.emissive[1] = m.f; mat.emissive[2] = m.g; mat.metallic = m.h;
                    }
                    int material = 0;
                    if (kTex && is_hit) material = (int)f2u(ld256(&p.fnormals[w.hit_slot]).h);
                    const bool ended = kTex ? shade_hit_textured(p, is_hit, w.hit_slot, w.hit_u, w.hit_v, material, n, mat, ray, seed, radiance, throughput, bounce)
                                            : shade_hit(p, is_hit, n, mat, ray, seed, radiance, throughput, bounce);
                    if (ended) {
                        F4 acc = p.accum[pix];
                        accumulate_mode(p.accum_sums, acc, radiance, f == 0 && p.reset_first != 0);
                        p.accum[pix] = acc;
                        new_sample = true;
                        if (++f == p.n_frames) {
                            if (kQueue) want_job = true; else done = true;
                        }
                    }
                }
            }
            if (kQueue) {
                // want_job: this lane has no pixel to go on with (it was vacant, or its pixel's last frame was just accumulated)
                if (queue_open && CV_WARP_ANY(want_job)) {  // warp-uniform: the lanes that need a pixel draw the next jobs together
                    const TraceJobs jobs = trace_jobs(p);  // recomputed from the launch constants: nothing of it stays live in the hot loop
                    const int job = trace_draw_jobs(job_counter, want_job);
                    if (want_job && job < jobs.n_jobs && trace_job_pixel(jobs, job, x, y)) {
                        pix = (size_t)y * (size_t)p.width + (size_t)x;
                        f = 0;
                        new_sample = true;
                        want_job = false;  // served
                    }
                    if (CV_WARP_ANY(want_job && job >= jobs.n_jobs)) queue_open = false;
                }
                vacant = want_job;                      // unserved (padding job, or the queue ran out): draws again at the next shading phase ...
                if (want_job && !queue_open) done = true;  // ... unless the queue is empty: the lane retires
            }
            if (need_shade) {
                if (new_sample && !done && !(kQueue && vacant)) {
                    radiance = f3(0.0f, 0.0f, 0.0f);
                    throughput = f3(1.0f, 1.0f, 1.0f);
                    bounce = 0;
                    start_path(p, x, y, p.first_frame + f, ray, seed);
                }
                walk_begin(p, w, ray);
                if (done || (kQueue && vacant)) w.ref = kRefSentinel;
            }
        } else {
            if (can_step) walk_step_any<kCount, kWide, kStream>(p, w, ray.t, cnt);
#pragma unroll
            for (int k = 1; k < kMinority; ++k) {
                if ((k % kParkEvery) == 0) walk_park_any<kWide>(w);
                if (ref_is_interior(w.ref)) walk_step_any<kCount, kWide, kStream>(p, w, ray.t, cnt);
            }
        }
    }
    return segments;
}

}  // namespace cavept
