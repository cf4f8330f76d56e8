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
constexpr int kStackSize = 64;

// 32-byte node of the production layout.  Siblings are adjacent (pair at a, a+1).
//   interior: a = index of the left child, b = 0
//   leaf:     a = first FastTri slot,      b = triangle count (1..4)
//   TLAS leaf: a = instance index,          b = -1
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
    const int* inst_fast_root;  // per instance: FastNode index of its mesh's root
    int tlas_root;
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

// Slab test of the production traversal: same interval logic as HitBvh plus the
// (result-neutral) rejection of boxes entirely behind the origin; returns entry t.
CV_HD bool slab_fast(F3 o, F3 inv, float ray_t, const float* mn, const float* mx, float& t_entry) {
    const float ax = (mn[0] - o.x) * inv.x, bx = (mx[0] - o.x) * inv.x;
    const float ay = (mn[1] - o.y) * inv.y, by = (mx[1] - o.y) * inv.y;
    const float az = (mn[2] - o.z) * inv.z, bz = (mx[2] - o.z) * inv.z;
    const float tmin = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
    const float tmax = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
    t_entry = tmin;
    return (tmin <= tmax) && (ray_t > tmin) && (tmax >= 0.0f);
}

// ---- production traversal ---------------------------------------------------------------
template<bool kCount>
CV_D Hit hit_scene_fast(const TraceParams& p, Ray& ray, unsigned long long* cnt) {
    Hit res;
    res.mesh = -1;
    res.tri = -1;
    res.u = 0.0f;
    res.v = 0.0f;
    int stack[kStackSize];
    int sp = 0;
    // active ray = world ray while in the TLAS, mesh-local ray while inside an instance
    F3 o = ray.origin, d = ray.direction, inv = ray.inv_dir;
    int cur_instance = -1;
    int node = p.tlas_root;  // TLAS nodes live at indices >= tlas_base == tlas_root (root is the first TLAS node)
    const int tlas_base = p.tlas_root;
    for (;;) {
        if (node >= tlas_base && cur_instance >= 0) {  // popped back into the TLAS
            o = ray.origin;
            d = ray.direction;
            inv = ray.inv_dir;
            cur_instance = -1;
        }
        const FastNode nd = p.fnodes[node];
        bool descend = false;
        if (nd.b == 0) {  // interior: test both children, go near first
            const FastNode c0 = p.fnodes[nd.a];
            const FastNode c1 = p.fnodes[nd.a + 1];
            if (kCount) cnt[1] += 2;
            float t0, t1;
            const bool h0 = slab_fast(o, inv, ray.t, c0.mn, c0.mx, t0);
            const bool h1 = slab_fast(o, inv, ray.t, c1.mn, c1.mx, t1);
            if (h0 && h1) {
                const bool swap = t1 < t0;
                node = swap ? nd.a + 1 : nd.a;
                if (sp < kStackSize) stack[sp++] = swap ? nd.a : nd.a + 1;
                descend = true;
            } else if (h0 || h1) {
                node = h0 ? nd.a : nd.a + 1;
                descend = true;
            }
        } else if (nd.b > 0) {  // BLAS leaf
            for (int k = 0; k < nd.b; ++k) {
                const FastTri tr = p.ftris[nd.a + k];
                if (kCount) ++cnt[2];
                float u, v;
                if (hit_triangle(f3(tr.v0.x, tr.v0.y, tr.v0.z), f3(tr.v1.x, tr.v1.y, tr.v1.z), f3(tr.v2.x, tr.v2.y, tr.v2.z), o, d, ray.t, u, v)) {
                    res.tri = (int)f2u(tr.v0.w);
                    res.mesh = cur_instance;
                    res.u = u;
                    res.v = v;
                }
            }
        } else {  // TLAS leaf: enter the instance (HitScene :160-166)
            cur_instance = nd.a;
            const GpuPtMesh& mesh = p.meshes[cur_instance];
            o = mat_mul(mesh.transformInv, ray.origin, 1.0f);
            d = mat_mul(mesh.transformInv, ray.direction, 0.0f);
            inv = recip3(d);
            node = p.inst_fast_root[cur_instance];
            // the mesh root's own box was already accepted through the TLAS (a padded superset)
            descend = true;
        }
        if (!descend) {
            if (sp == 0) break;
            node = stack[--sp];
        }
    }
    return res;
}

// ---- one RayColor iteration after the scene query (:198-235).  Returns true when the
//      path ends (miss, or MAX_BOUNCE segments traced). ---------------------------------
CV_D bool shade_segment(const TraceParams& p, const Hit& hit, Ray& ray, uint32_t& state, F3& radiance, F3& throughput, int& bounce) {
    if (hit.mesh >= 0 && hit.tri >= 0) {
        ray.origin = ray.origin + ray.t * ray.direction;
        ray.t = kRayTMax;
        const GpuPtMesh& mesh = p.meshes[hit.mesh];
        const int* idx = p.indices[hit.tri].tri;
        const GpuPtVertex& va = p.vertices[idx[0]];
        const GpuPtVertex& vb = p.vertices[idx[1]];
        const GpuPtVertex& vc = p.vertices[idx[2]];
        const F3 n1 = f3(va.normal[0], va.normal[1], va.normal[2]);
        const F3 n2 = f3(vb.normal[0], vb.normal[1], vb.normal[2]);
        const F3 n3 = f3(vc.normal[0], vc.normal[1], vc.normal[2]);
        F3 n = n1 + hit.u * (n2 - n1) + hit.v * (n3 - n1);
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

// One pixel, frames [first_frame, first_frame + n_frames).  The segment loop is flat: a
// lane whose path ends starts its next sample at once, so the lanes of a warp stay in
// the traversal code together whatever bounce each of them is on.
// Returns the number of segments traced; visit counters go to cnt[1..2] when kCount.
template<bool kFast, bool kCount>
CV_D unsigned int body_trace_pixel(const TraceParams& p, int x, int y, unsigned long long* cnt) {
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
        const Hit hit = kFast ? hit_scene_fast<kCount>(p, ray, cnt) : hit_scene_reference<kCount>(p, ray, cnt);
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

}  // namespace cavept
