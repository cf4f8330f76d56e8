// Production path-trace kernel body, second generation: persistent warps, a global job queue and
// TWO pixel streams per lane ("ray pool" of two records per thread in shared memory).
//
// What it replaces: HitScene + RayColor + the compute entry (cave/shader/shared_path_tracer.h:152-239,
// cave/shader/glsl/path_tracer.cs.glsl:11-60), like body_trace_pixel_fast (trace_body.h).  Same tree
// (QPair / FastTri), same walk_step / walk_test_triangle / shade_hit / start_path / accumulate, so every
// pixel receives bit-identical arithmetic; only WHO executes it and WHEN changes:
//
//   * job queue (persistent threads): a job = one pixel for all n_frames of the launch.  Jobs are
//     numbered tile-major (8x4 pixel tiles, so the 32 jobs a warp draws together are one coherent tile)
//     and handed out by one warp-aggregated atomicAdd on a global counter.  A lane whose pixel is
//     finished draws the next job instead of idling until the warp's slowest lane ends: no launch tail.
//     One lane runs all frames of a pixel in order, so the running mean folds in the reference's order.
//   * two records per lane: each lane owns two independent pixel streams.  The traversal state (ray,
//     stack pointer, parked leaves) of the ACTIVE one lives in registers; everything else (path
//     throughput/radiance/seed, and the other stream's ray or finished hit) lives in shared memory.
//     When the active traversal completes, the lane files the hit as PENDING and switches to its other
//     stream at once if that one has a FRESH ray -- it does not wait for the warp to shade.  Shading runs
//     on the PENDING records when (almost) every lane has one, i.e. at full width, and turns them back into
//     FRESH rays (next bounce, next frame, or the next job).  This is the ray compaction of a wavefront
//     tracer done in place: the idle slots a finished ray leaves in a warp are refilled from the lane's
//     own second stream instead of by moving rays between lanes.
//
// Record layout in shared memory (floats / int bits), [field][slot][thread] so every access is
// conflict-free: 0-2 ray origin, 3-5 ray direction, 6 t of the finished segment, 7 hit FastTri slot (-1 miss),
// 8-9 u v, 10-12 throughput, 13-15 radiance, 16 RNG state, 17 pixel index, 18 frame-in-launch | bounce << 24.
#pragma once
#include "trace_body.h"

namespace cavept {

constexpr int kPoolFields = 19;
constexpr int kPoolSlots = 2;
constexpr int kPoolWordsPerThread = kPoolFields * kPoolSlots;

enum : uint32_t { POOL_EMPTY = 0u, POOL_FRESH = 1u, POOL_ACTIVE = 2u, POOL_PENDING = 3u };

#ifndef CV_POOL_QUORUM
#define CV_POOL_QUORUM 24
#endif
#ifndef CV_POOL_SHADE_BATCH
#define CV_POOL_SHADE_BATCH 28
#endif
#ifndef CV_POOL_IDLE_MAX
#define CV_POOL_IDLE_MAX 6  // shade as soon as this many lanes have nothing left to traverse (both records PENDING)
#endif
#ifndef CV_POOL_REFILL_BATCH
#define CV_POOL_REFILL_BATCH 1  // lanes with a finished pixel that must wait before the warp draws new jobs
#endif
#ifndef CV_POOL_TIGHT_STEPS
#define CV_POOL_TIGHT_STEPS 6
#endif
#ifndef CV_POOL_MINORITY_STEPS
#define CV_POOL_MINORITY_STEPS 4
#endif

// the job queue is trace_body.h's (TraceJobs / trace_jobs / trace_job_pixel / trace_draw_jobs)
using PoolJobs = TraceJobs;
CV_HD PoolJobs pool_jobs(const TraceParams& p) { return trace_jobs(p); }
CV_HD int pool_job_pixel(const TraceParams& p, const PoolJobs& j, int job, int& x, int& y) { return trace_job_pixel(j, job, x, y) ? y * p.width + x : -1; }
CV_D int pool_draw_jobs(int* counter, bool want) { return trace_draw_jobs(counter, want); }

#define CV_REC(field, slot) rec[((field) * kPoolSlots + (slot)) * stride]

CV_D void pool_store_ray(float* rec, int stride, int s, const Ray& ray) {
    CV_REC(0, s) = ray.origin.x; CV_REC(1, s) = ray.origin.y; CV_REC(2, s) = ray.origin.z;
    CV_REC(3, s) = ray.direction.x; CV_REC(4, s) = ray.direction.y; CV_REC(5, s) = ray.direction.z;
}
CV_D void pool_load_ray(const float* rec, int stride, int s, Ray& ray) {
    ray.origin = f3(CV_REC(0, s), CV_REC(1, s), CV_REC(2, s));
    ray.direction = f3(CV_REC(3, s), CV_REC(4, s), CV_REC(5, s));
    ray.t = kRayTMax;
}

// A new pixel job in slot s: first sample's camera ray, empty path state.  Returns the new slot state.
CV_D uint32_t pool_begin_job(const TraceParams& p, const PoolJobs& jobs, float* rec, int stride, int s, int job) {
    if (job >= jobs.n_jobs) return POOL_EMPTY;
    int x, y;
    const int pix = pool_job_pixel(p, jobs, job, x, y);
    if (pix < 0) return POOL_EMPTY;  // padding of a ragged rectangle: the slot draws again at the next refill
    Ray ray;
    uint32_t seed;
    start_path(p, x, y, p.first_frame, ray, seed);
    pool_store_ray(rec, stride, s, ray);
    CV_REC(10, s) = 1.0f; CV_REC(11, s) = 1.0f; CV_REC(12, s) = 1.0f;
    CV_REC(13, s) = 0.0f; CV_REC(14, s) = 0.0f; CV_REC(15, s) = 0.0f;
    CV_REC(16, s) = u2f(seed);
    CV_REC(17, s) = u2f((uint32_t)pix);
    CV_REC(18, s) = u2f(0u);
    return POOL_FRESH;
}

// Shade the PENDING record of slot s (RayColor's loop body after the scene query, :198-235), accumulate
// when the path ends, and leave the slot FRESH (next segment / next frame) or EMPTY (pixel finished: the
// caller refills).  Returns true when the pixel's last frame was accumulated.
CV_D bool pool_shade_slot(const TraceParams& p, float* rec, int stride, int s) {
    Ray ray;
    ray.origin = f3(CV_REC(0, s), CV_REC(1, s), CV_REC(2, s));
    ray.direction = f3(CV_REC(3, s), CV_REC(4, s), CV_REC(5, s));
    ray.t = CV_REC(6, s);
    const int hit_slot = (int)f2u(CV_REC(7, s));
    const float hu = CV_REC(8, s), hv = CV_REC(9, s);
    F3 throughput = f3(CV_REC(10, s), CV_REC(11, s), CV_REC(12, s));
    F3 radiance = f3(CV_REC(13, s), CV_REC(14, s), CV_REC(15, s));
    uint32_t seed = f2u(CV_REC(16, s));
    const uint32_t pix = f2u(CV_REC(17, s));
    const uint32_t fb = f2u(CV_REC(18, s));
    int f = (int)(fb & 0xffffffu), bounce = (int)(fb >> 24);
    const bool is_hit = hit_slot >= 0;
    F3 n = f3(0, 0, 0);
    GpuPtMaterial mat = {};
    if (is_hit) {
        const F8 a = ld256(&p.fnormals[hit_slot]);     // n1.xyz, instance, n2.xyz, material
        const F8 b = ld256(&p.fnormals[hit_slot].v2);  // n3.xyz
        const F3 n1 = f3(a.a, a.b, a.c), n2 = f3(a.e, a.f, a.g), n3 = f3(b.a, b.b, b.c);
        n = n1 + hu * (n2 - n1) + hv * (n3 - n1);
        const F8 m = ld256(&p.materials[(int)f2u(a.h)]);
        mat.baseColor[0] = m.a; mat.baseColor[1] = m.b; mat.baseColor[2] = m.c; mat.roughness = m.d;
        mat.emissive[0] = m.e; mat.emissive[1] = m.f; mat.emissive[2] = m.g; mat.metallic = m.h;
    }
    bool finished = false;
    if (shade_hit(p, is_hit, n, mat, ray, seed, radiance, throughput, bounce)) {
        F4 acc = p.accum[pix];
        accumulate_mode(p.accum_sums, acc, radiance, f == 0 && p.reset_first != 0);
        p.accum[pix] = acc;
        ++f;
        if (f == p.n_frames) {
            finished = true;
        } else {
            radiance = f3(0.0f, 0.0f, 0.0f);
            throughput = f3(1.0f, 1.0f, 1.0f);
            bounce = 0;
            const int x = (int)(pix % (uint32_t)p.width), y = (int)(pix / (uint32_t)p.width);
            start_path(p, x, y, p.first_frame + f, ray, seed);
        }
    }
    if (!finished) {
        pool_store_ray(rec, stride, s, ray);
        CV_REC(10, s) = throughput.x; CV_REC(11, s) = throughput.y; CV_REC(12, s) = throughput.z;
        CV_REC(13, s) = radiance.x; CV_REC(14, s) = radiance.y; CV_REC(15, s) = radiance.z;
        CV_REC(16, s) = u2f(seed);
        CV_REC(18, s) = u2f((uint32_t)f | ((uint32_t)bounce << 24));
    }
    return finished;
}

// Slot states, two bits per slot in one register (the shade / refill loops index them by slot number).
CV_HD uint32_t pool_state(uint32_t st, int s) { return (st >> (2 * s)) & 3u; }
CV_HD uint32_t pool_with_state(uint32_t st, int s, uint32_t v) { return (st & ~(3u << (2 * s))) | (v << (2 * s)); }
CV_HD bool pool_any_slot(uint32_t st, uint32_t v) { return pool_state(st, 0) == v || pool_state(st, 1) == v; }

// One warp (one lane in the emulator) of the persistent kernel.  `stack` and `rec` point at this thread's column.
// Returns the number of segments this lane shaded.  Every long code sequence (start_path, walk_begin, the
// shading) is instantiated ONCE: the body is a single loop whose stages are entered on warp-uniform votes.
template<bool kCount>
CV_D unsigned int body_trace_pool(const TraceParams& p, int* stack, float* rec, int stride, int* job_counter, unsigned long long* cnt) {
    const PoolJobs jobs = pool_jobs(p);
    unsigned int segments = 0;
    uint32_t st = 0u;        // both slots EMPTY
    bool queue_open = true;  // warp-uniform: false once a draw came back past the last job
    bool refill = true;      // warp-uniform: draw jobs for the EMPTY slots at the top of the next trip
    int a = 0;               // the slot whose traversal is in registers
    bool active = false;
    Ray ray;
    ray.origin = f3(0, 0, 0);
    ray.direction = f3(0, 0, 1);
    ray.inv_dir = f3(0, 0, 1);
    ray.t = kRayTMax;
    Walk w;
    w.stack = stack;
    w.stride = stride;
    w.sp = 1;
    w.top = kRefSentinel;
    w.ref = kRefSentinel;
    w.q0 = w.q1 = 0;
    w.inv = w.oinv = f3(0, 0, 0);
    w.near_x = w.near_y = w.near_z = kSelLo;
    w.hit_slot = -1;
    w.hit_u = w.hit_v = 0.0f;
    for (;;) {
        // ---- refill: EMPTY slots draw the next jobs (pixels finish rarely: once per n_frames samples) -----
        if (refill && queue_open) {
#pragma unroll 1
            for (int s = 0; s < kPoolSlots; ++s) {
                const bool want = pool_state(st, s) == POOL_EMPTY;
                if (!CV_WARP_ANY(want)) continue;
                const int job = pool_draw_jobs(job_counter, want);
                if (want) st = pool_with_state(st, s, pool_begin_job(p, jobs, rec, stride, s, job));
                if (CV_WARP_ANY(want && job >= jobs.n_jobs)) queue_open = false;
            }
        }
        refill = false;
        // ---- A, the tight loop: one ballot per group of steps while a quorum can step ---------------
        for (;;) {
            walk_park_leaf(w);
            const bool can_step = ref_is_interior(w.ref);
            if (CV_WARP_COUNT(can_step) < CV_POOL_QUORUM) break;
            if (can_step) walk_step<kCount>(p, w, ray.t, cnt);
#pragma unroll
            for (int k = 1; k < CV_POOL_TIGHT_STEPS; ++k) {
                if ((k % CV_PARK_EVERY) == 0) walk_park_leaf(w);
                if (ref_is_interior(w.ref)) walk_step<kCount>(p, w, ray.t, cnt);
            }
        }
        // ---- retire + switch: a completed traversal files its hit as PENDING; a lane without a traversal
        //      picks up a FRESH record of its own (its other stream, or what the last shading left) ----------
        {
            const bool has_tri0 = w.q0 != 0;
            const bool fin = active && !ref_is_interior(w.ref) && !has_tri0;
            const bool start = (!active || fin) && (pool_state(st, 0) == POOL_FRESH || pool_state(st, 1) == POOL_FRESH);
            if (CV_WARP_ANY(fin || start)) {
                if (fin) {
                    CV_REC(6, a) = ray.t;
                    CV_REC(7, a) = u2f((uint32_t)w.hit_slot);
                    CV_REC(8, a) = w.hit_u;
                    CV_REC(9, a) = w.hit_v;
                    st = pool_with_state(st, a, POOL_PENDING);
                    active = false;
                    w.ref = kRefSentinel;
                }
                if (start) {
                    a = pool_state(st, 0) == POOL_FRESH ? 0 : 1;
                    pool_load_ray(rec, stride, a, ray);
                    walk_begin(p, w, ray);
                    st = pool_with_state(st, a, POOL_ACTIVE);
                    active = true;
                }
                if (CV_WARP_COUNT(ref_is_interior(w.ref)) >= CV_POOL_QUORUM) continue;  // enough lanes are stepping again
            }
        }
        // ---- the vote ---------------------------------------------------------------------------------
        const bool step_now = ref_is_interior(w.ref);
        const bool has_tri = w.q0 != 0;
        const bool pending = pool_any_slot(st, POOL_PENDING);
        const int n_step = CV_WARP_COUNT(step_now);
        const int n_tri = CV_WARP_COUNT(has_tri);
        const int n_pending = CV_WARP_COUNT(pending);
        if (n_step + n_tri + n_pending == 0) {
            // nothing in flight (a FRESH record would have been started above): draw more jobs, or the launch is done
            if (!queue_open || !CV_WARP_ANY(pool_any_slot(st, POOL_EMPTY))) break;
            refill = true;
            continue;
        }
        if (n_tri > 0 && n_tri >= n_step) {
            if (has_tri) walk_test_triangle<kCount>(p, w, ray, ray.t, cnt);
        } else if (n_pending >= CV_POOL_SHADE_BATCH || CV_WARP_COUNT(pending && !active) >= CV_POOL_IDLE_MAX || n_step == 0) {
            // ---- C: shade every PENDING record ------------------------------------------------------------
#pragma unroll 1
            for (int s = 0; s < kPoolSlots; ++s) {
                const bool mine = pool_state(st, s) == POOL_PENDING;
                if (!CV_WARP_ANY(mine)) continue;
                if (mine) {
                    ++segments;
                    st = pool_with_state(st, s, pool_shade_slot(p, rec, stride, s) ? POOL_EMPTY : POOL_FRESH);
                }
            }
            // finished pixels (and the padding slots of a ragged rectangle) draw again once enough of them wait
            refill = CV_WARP_COUNT(pool_any_slot(st, POOL_EMPTY)) >= CV_POOL_REFILL_BATCH;
        } else {
            if (step_now) walk_step<kCount>(p, w, ray.t, cnt);
#pragma unroll
            for (int k = 1; k < CV_POOL_MINORITY_STEPS; ++k) {
                if ((k % CV_PARK_EVERY) == 0) walk_park_leaf(w);
                if (ref_is_interior(w.ref)) walk_step<kCount>(p, w, ray.t, cnt);
            }
        }
    }
    return segments;
}

#undef CV_REC

}  // namespace cavept
