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
constexpr int kFastLeafMax = 4;            // subtrees of <= 4 triangles are median-split in the reference (:138); they become leaves here
constexpr int kStackSize = 32;  // per-lane entries in shared memory (16 KB per 128-thread block)

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

// Pre-gathered triangle: 3 x float4; w lanes carry the scene-global triangle id and padding.
struct alignas(16) FastTri {
    F4 v0, v1, v2;
};
static_assert(sizeof(FastTri) == 48, "FastTri");

struct TraceParams {
    // reference-layout arrays (what BindPathTracerData binds, path_tracer.cpp:324-331)
    const GpuPtVertex* vertices;
    const GpuPtIndex* indices;
    const GpuPtBvh* bvhs;
    const GpuPtMesh* meshes;
    const GpuPtMaterial* materials;
    int mesh_count;  // c_ptObjectCount
    // production layout
    const FastNode* fnodes;
    const FastTri* ftris;
    const FastTri* fnormals;    // per FastTri slot: the three vertex normals (shading)
    const int* inst_fast_root;  // per instance: node reference of its mesh's root
    int tlas_root;              // node reference of the TLAS root
    // PerFrameConstantBuffer (cbuffer.hlsl.h:193-207)
    float cam_pos[3], cam_right[3], cam_up[3], cam_fwd[3];
    float cam_distance;  // tan(c_cameraFovDegree * PI / 180), path_tracer.cs.glsl:31-32
    int width, height;
    int max_bounce;
    int first_frame, n_frames;
    int reset_first;  // c_sceneDirty for the first frame of the launch
    F4* accum;
    unsigned long long* counters;  // [0] segments, [1] node visits, [2] triangle tests
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
CV_HD F3 recip3(F3 d) { return f3(cv_div(1.0f, d.x), cv_div(1.0f, d.y), cv_div(1.0f, d.z)); }

// ---- HitTriangle, :88-130 ------------------------------------------------------------
CV_HD bool hit_triangle(F3 A, F3 B, F3 C, F3 origin, F3 direction, float& ray_t, float& out_u, float& out_v) {
    const F3 AB = B - A;
    const F3 AC = C - A;
    const F3 P = cross3(direction, AC);
    const float det = dot3(AB, P);
    if (det < kEpsilon) return false;  // back faces culled
    const float inv_det = cv_div(1.0f, det);
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
CV_HD bool hit_bvh_reference(const Ray& r, F3 mn, F3 mx) {
    const F3 t0 = (mn - r.origin) * r.inv_dir;
    const F3 t1 = (mx - r.origin) * r.inv_dir;
    const F3 ts = f3(min_ref(t0.x, t1.x), min_ref(t0.y, t1.y), min_ref(t0.z, t1.z));
    const F3 tb = f3(max_ref(t0.x, t1.x), max_ref(t0.y, t1.y), max_ref(t0.z, t1.z));
    const float tmin = max_ref(ts.x, max_ref(ts.y, ts.z));
    const float tmax = min_ref(tb.x, min_ref(tb.y, tb.z));
    return (tmin < tmax) && (r.t > tmin);
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
// Node references (one int): interior = index of the child pair (children at ref, ref+1);
// leaf = sign bit | (count-1) << 29 | first FastTri slot inside a BLAS, sign bit | instance
// index inside the TLAS.  A node's own reference is stored in its record, so one 64-byte
// fetch of a sibling pair yields both boxes and both references.
constexpr int kRefSentinel = 0x7fffffff;  // bottom of the stack
constexpr int kRefDeferredPop = 0x7ffffffe;
CV_HD bool ref_is_interior(int r) { return (uint32_t)r < (uint32_t)kRefDeferredPop; }
CV_HD int make_leaf_ref(int first_slot, int count) { return (int)(0x80000000u | ((uint32_t)(count - 1) << 29) | (uint32_t)first_slot); }
CV_HD int make_instance_ref(int instance) { return (int)(0x80000000u | (uint32_t)instance); }

#if defined(__CUDA_ARCH__)
#define CV_WARP_ANY(pred) __any_sync(0xffffffffu, (pred))
#define CV_WARP_COUNT(pred) __popc(__ballot_sync(0xffffffffu, (pred)))
#else  // the emulator steps one lane at a time: a warp of one
#define CV_WARP_ANY(pred) (pred)
#define CV_WARP_COUNT(pred) ((pred) ? 1 : 0)
#endif

// Slab test of the production traversal.  Same interval logic as HitBvh (:133-149) with
// two result-neutral changes: boxes entirely behind the origin are rejected, and the exit
// distance is widened by 2 ulp so that rounding differences to the reference's
// (min - o) * inv form can only ever add candidate triangles, never lose one.
CV_HD bool slab_fast(F3 inv, F3 oinv, float ray_t, const FastNode& n, float& t_entry) {
    const float ax = cv_fma(n.mn[0], inv.x, oinv.x), bx = cv_fma(n.mx[0], inv.x, oinv.x);
    const float ay = cv_fma(n.mn[1], inv.y, oinv.y), by = cv_fma(n.mx[1], inv.y, oinv.y);
    const float az = cv_fma(n.mn[2], inv.z, oinv.z), bz = cv_fma(n.mx[2], inv.z, oinv.z);
    const float tmin = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
    const float tmax = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz)) * 1.0000003f;
    t_entry = tmin;
    return (tmin <= tmax) && (ray_t > tmin) && (tmax >= 0.0f);
}

// Per-lane traversal state of the production kernel.
struct Walk {
    int* stack;     // lane-strided: entry k of this lane is stack[k * stride]
    int stride;
    int sp;         // entry 0 holds the sentinel
    int ref;        // current node reference
    int leaf;       // postponed BLAS leaf (0 = none)
    int instance;   // -1 while in the TLAS
    int blas_sp;    // stack depth at instance entry (0 = not inside an instance)
    F3 o, d, inv, oinv;  // active ray: world ray in the TLAS, mesh-local ray inside an instance
    int hit_instance, hit_slot;
    float hit_u, hit_v;
};

CV_D void walk_set_ray(Walk& w, F3 o, F3 d) {
    w.o = o;
    w.d = d;
    w.inv = recip3(d);
    w.oinv = f3(-(o.x * w.inv.x), -(o.y * w.inv.y), -(o.z * w.inv.z));
}

CV_D void walk_begin(const TraceParams& p, Walk& w, const Ray& ray) {
    w.stack[0] = kRefSentinel;
    w.sp = 1;
    w.ref = p.tlas_root;
    w.leaf = 0;
    w.instance = -1;
    w.blas_sp = 0;
    w.hit_instance = -1;
    w.hit_slot = -1;
    w.hit_u = 0.0f;
    w.hit_v = 0.0f;
    walk_set_ray(w, ray.origin, ray.direction);
}

// Pop the next reference.  Leaving an instance restores the world ray; that is postponed
// (kRefDeferredPop) while a leaf of this instance is still waiting to be intersected.
CV_D void walk_pop(Walk& w, const Ray& ray) {
    if (w.sp == w.blas_sp) {
        if (w.leaf < 0) {
            w.ref = kRefDeferredPop;
            return;
        }
        walk_set_ray(w, ray.origin, ray.direction);
        w.instance = -1;
        w.blas_sp = 0;
    }
    w.ref = w.stack[(--w.sp) * w.stride];
}

template<bool kCount>
CV_D void walk_intersect_leaf(const TraceParams& p, Walk& w, int leaf_ref, Ray& ray, unsigned long long* cnt) {
    const int first = leaf_ref & 0x1fffffff;
    const int count = ((leaf_ref >> 29) & 3) + 1;
    for (int k = 0; k < count; ++k) {
        const FastTri tr = p.ftris[first + k];
        if (kCount) ++cnt[2];
        float u, v;
        if (hit_triangle(f3(tr.v0.x, tr.v0.y, tr.v0.z), f3(tr.v1.x, tr.v1.y, tr.v1.z), f3(tr.v2.x, tr.v2.y, tr.v2.z), w.o, w.d, ray.t, u, v)) {
            w.hit_slot = first + k;
            w.hit_instance = w.instance;
            w.hit_u = u;
            w.hit_v = v;
        }
    }
}

// One interior step (or instance entry) for a lane whose reference is actionable.
template<bool kCount>
CV_D void walk_step(const TraceParams& p, Walk& w, const Ray& ray, unsigned long long* cnt) {
    if (w.ref < 0) {  // TLAS leaf: enter the instance (HitScene :160-166)
        w.instance = w.ref & 0x7fffffff;
        const GpuPtMesh& mesh = p.meshes[w.instance];
        walk_set_ray(w, mat_mul(mesh.transformInv, ray.origin, 1.0f), mat_mul(mesh.transformInv, ray.direction, 0.0f));
        w.blas_sp = w.sp;
        w.ref = p.inst_fast_root[w.instance];
    } else {
        const FastNode c0 = p.fnodes[w.ref];
        const FastNode c1 = p.fnodes[w.ref + 1];
        if (kCount) cnt[1] += 2;
        float t0, t1;
        const bool h0 = slab_fast(w.inv, w.oinv, ray.t, c0, t0);
        const bool h1 = slab_fast(w.inv, w.oinv, ray.t, c1, t1);
        if (h0 && h1) {
            const bool swap = t1 < t0;
            w.ref = swap ? c1.a : c0.a;
            w.stack[(w.sp++) * w.stride] = swap ? c0.a : c1.a;
        } else if (h0 || h1) {
            w.ref = h0 ? c0.a : c1.a;
        } else {
            walk_pop(w, ray);
        }
    }
    // postpone the first BLAS leaf and keep descending while other lanes still search
    if (w.ref < 0 && w.instance >= 0 && w.leaf == 0 && w.sp > w.blas_sp) {
        w.leaf = w.ref;
        w.ref = w.stack[(--w.sp) * w.stride];
    }
}

// ---- one RayColor iteration after the scene query (:198-235).  Returns true when the
//      path ends (miss, or MAX_BOUNCE segments traced). ---------------------------------
CV_D bool shade_hit(const TraceParams& p, bool is_hit, int mesh_id, F3 n1, F3 n2, F3 n3, float u, float v, Ray& ray, uint32_t& state, F3& radiance,
                    F3& throughput, int& bounce) {
    if (is_hit) {
        ray.origin = ray.origin + ray.t * ray.direction;
        ray.t = kRayTMax;
        const GpuPtMesh& mesh = p.meshes[mesh_id];
        F3 n = n1 + u * (n2 - n1) + v * (n3 - n1);
        n = normalize3(mat_mul(mesh.transform, n, 0.0f));
        const GpuPtMaterial& mat = p.materials[mesh.materialId];
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
    F3 n1 = f3(0, 0, 0), n2 = n1, n3 = n1;
    if (is_hit) {
        const int* idx = p.indices[hit.tri].tri;
        const GpuPtVertex& va = p.vertices[idx[0]];
        const GpuPtVertex& vb = p.vertices[idx[1]];
        const GpuPtVertex& vc = p.vertices[idx[2]];
        n1 = f3(va.normal[0], va.normal[1], va.normal[2]);
        n2 = f3(vb.normal[0], vb.normal[1], vb.normal[2]);
        n3 = f3(vc.normal[0], vc.normal[1], vc.normal[2]);
    }
    return shade_hit(p, is_hit, hit.mesh, n1, n2, n3, hit.u, hit.v, ray, state, radiance, throughput, bounce);
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
            accumulate(acc, radiance, f == 0 && p.reset_first != 0);
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

// Production kernel body: ONE loop per warp in which every lane is either descending the
// BVH, intersecting a leaf, or shading -- whatever sample and bounce it is on.
//   phase A  interior steps; a lane that reaches a leaf postpones it and keeps descending
//            while any other lane is still searching (speculative while-while traversal);
//   phase B  every lane intersects its postponed / current leaf;
//   phase C  lanes whose traversal finished shade and start their next segment -- done in
//            batches (>= kShadeBatch lanes, or nobody left traversing) so the shading code
//            is not replayed for one lane at a time.
// Lanes of a warp never wait for the slowest ray of a segment, only for a shading batch.
constexpr int kShadeBatch = 8;

template<bool kCount>
CV_D unsigned int body_trace_pixel_fast(const TraceParams& p, int x, int y, int* stack, int stride, unsigned long long* cnt) {
    const bool valid = x < p.width && y < p.height && p.n_frames > 0;
    const size_t pix = valid ? (size_t)y * (size_t)p.width + (size_t)x : 0;
    Ray ray;
    uint32_t seed = 1u;
    int f = 0;
    int bounce = 0;
    unsigned int segments = 0;
    F3 radiance = f3(0.0f, 0.0f, 0.0f), throughput = f3(1.0f, 1.0f, 1.0f);
    bool done = !valid;
    Walk w;
    w.stack = stack;
    w.stride = stride;
    ray.origin = f3(0, 0, 0);
    ray.direction = f3(0, 0, 1);
    ray.inv_dir = f3(0, 0, 1);
    ray.t = kRayTMax;
    if (valid) start_path(p, x, y, p.first_frame, ray, seed);
    walk_begin(p, w, ray);
    if (!valid) w.ref = kRefSentinel;
    for (;;) {
        // ---- phase A -------------------------------------------------------------------
        for (;;) {
            const bool can_step = ref_is_interior(w.ref) || (w.ref < 0 && w.instance < 0);
            if (!CV_WARP_ANY(can_step && w.leaf == 0)) break;  // nobody is searching any more
            if (can_step) walk_step<kCount>(p, w, ray, cnt);
        }
        // ---- phase B -------------------------------------------------------------------
        if (w.leaf < 0) {
            walk_intersect_leaf<kCount>(p, w, w.leaf, ray, cnt);
            w.leaf = 0;
        }
        if (w.ref < 0 && w.instance >= 0) {
            walk_intersect_leaf<kCount>(p, w, w.ref, ray, cnt);
            walk_pop(w, ray);
        } else if (w.ref == kRefDeferredPop) {
            walk_pop(w, ray);
        }
        // ---- phase C -------------------------------------------------------------------
        const bool need_shade = !done && w.ref == kRefSentinel && w.leaf == 0;
        const int n_need = CV_WARP_COUNT(need_shade);
        const int n_walking = CV_WARP_COUNT(!done && !need_shade);
        if (need_shade && (n_need >= kShadeBatch || n_walking == 0)) {
            ++segments;
            const bool is_hit = w.hit_slot >= 0;
            F3 n1 = f3(0, 0, 0), n2 = n1, n3 = n1;
            if (is_hit) {
                const FastTri nr = p.fnormals[w.hit_slot];
                n1 = f3(nr.v0.x, nr.v0.y, nr.v0.z);
                n2 = f3(nr.v1.x, nr.v1.y, nr.v1.z);
                n3 = f3(nr.v2.x, nr.v2.y, nr.v2.z);
            }
            if (shade_hit(p, is_hit, w.hit_instance, n1, n2, n3, w.hit_u, w.hit_v, ray, seed, radiance, throughput, bounce)) {
                F4 acc = p.accum[pix];
                accumulate(acc, radiance, f == 0 && p.reset_first != 0);
                p.accum[pix] = acc;
                if (++f == p.n_frames) {
                    done = true;
                } else {
                    radiance = f3(0.0f, 0.0f, 0.0f);
                    throughput = f3(1.0f, 1.0f, 1.0f);
                    bounce = 0;
                    start_path(p, x, y, p.first_frame + f, ray, seed);
                }
            }
            if (!done) walk_begin(p, w, ray);
        }
        if (!CV_WARP_ANY(!done)) break;
    }
    return segments;
}

}  // namespace cavept
