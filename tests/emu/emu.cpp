// TEST CODE -- serial host emulation of the cavept kernel bodies.
// The kernel bodies under caveengine_b200/csrc/*_body.h are plain functions of a thread
// index; this file steps them one index at a time with the same host-side level loop the
// product uses (bvh_parity_driver.h, compact_body.h), so the algorithms can be checked
// against the oracle in a container without a GPU.  Built by tests/emu_loader.py with
// g++ -ffp-contract=off.  The product never links this.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "bvh_parity_driver.h"
#include "host_math.h"
#include "trace_body.h"
#include "trace_pool_body.h"
#include "wbvh_body.h"

using namespace cavept;

struct EmuExec {
    unsigned long long launches = 0;
    template<class F>
    void launch(int n, F f) {
        for (int i = 0; i < n; ++i) f(i);
        ++launches;
    }
    void scan_u8(const unsigned char* in, int* out, int n) {
        int run = 0;
        for (int i = 0; i < n; ++i) { out[i] = run; run += in[i]; }
        out[n] = run;
    }
    void scan_i32(const int* in, int* out, int n) {
        int run = 0;
        for (int i = 0; i < n; ++i) { out[i] = run; run += in[i]; }
        out[n] = run;
    }
    void sort_pairs(uint32_t* keys, int* vals, uint32_t*, int*, int*, int n, int) {
        std::vector<std::pair<uint32_t, int>> v(n);
        for (int i = 0; i < n; ++i) v[i] = { keys[i], vals[i] };
        std::stable_sort(v.begin(), v.end(), [](const std::pair<uint32_t, int>& a, const std::pair<uint32_t, int>& b) { return a.first < b.first; });
        for (int i = 0; i < n; ++i) { keys[i] = v[i].first; vals[i] = v[i].second; }
    }
    void wb_prim(const WBuild& b) { launch(b.n_prims, FnWbPrim{ b }); }
    void parity_reduce(const BuildArrays& a) { launch(a.n_tris, FnReduce{ a }); }
    void parity_bin(const BuildArrays& a) { launch(a.n_tris, FnBin{ a }); }
    void parity_split(const BuildArrays& a, int n_items) { launch(n_items, FnSplit{ a, n_items }); }
    void parity_subtrees(const BuildArrays& a, int n_roots) {
        for (int r = 0; r < n_roots; ++r) parity_subtree_serial(a, r);
        ++launches;
    }
    void wb_bin(const WBuild& b) { launch(b.n_prims, FnWbBin{ b }); }
    void wb_split(const WBuild& b, int level, int bound) { launch(bound, FnWbSplit{ b, level, true }); }
    void wb_subtrees(const WBuild& b, int n_roots) {
        for (int r = 0; r < n_roots; ++r) wb_subtree_serial(b, r);
        ++launches;
    }
    void wb_partition(const WBuild& b) {
        launch(b.n_prims, FnWbFlag{ b });
        scan_u8(b.flags, b.scan, b.n_prims);
        launch(b.n_prims, FnWbScatter{ b });
    }
    void upload(void* dst, const void* src, size_t bytes) { std::memcpy(dst, src, bytes); }
    void download(void* dst, const void* src, size_t bytes) { std::memcpy(dst, src, bytes); }
    void fill(void* dst, int byte, size_t bytes) { std::memset(dst, byte, bytes); }
};

#define EMU_API extern "C" __attribute__((visibility("default")))

// Largest item the parity builder's subtree phase takes (0 = level loop only); tests build both ways.
static int g_parity_sub_max = kParitySubMax;
EMU_API void emu_set_parity_sub_max(int n) { g_parity_sub_max = n < 0 ? kParitySubMax : n; }

// Builds the scene-global GpuPtBvh array for meshes whose triangles occupy
// [tri_off[m], tri_off[m+1]) of `indices` (vertex ids already scene-global).
// Also returns per-node parent / subtree-count arrays.  rc 0 ok, -1 degenerate.
EMU_API int emu_bvh_build(const GpuPtVertex* vertices, const GpuPtIndex* indices, int n_tris, const int* tri_off, int n_meshes, GpuPtBvh* out_nodes,
                          int* out_parent, int* out_count, int* out_levels, int* out_error_bits) {
    const int n_nodes = 2 * n_tris - n_meshes;
    std::vector<F4> bmin(n_tris), bmax(n_tris), cent(n_tris);
    std::vector<int> order(n_tris), order_next(n_tris), node_of(n_tris), node_of_next(n_tris), scan(n_tris + 1), counters(8);
    std::vector<unsigned char> flags(n_tris);
    const size_t items_cap = (size_t)std::max(n_tris, n_meshes) + 2;
    std::vector<BuildItem> items(items_cap), items_next(items_cap), scratch(n_meshes), sub_roots(items_cap);
    std::vector<ItemAccum> accum(items_cap);
    const int bins_cap = n_tris / 5 + n_meshes + 2;
    std::vector<ItemBins> bins(bins_cap);
    BuildArrays a{};
    a.vertices = vertices;
    a.indices = indices;
    a.n_tris = n_tris;
    a.tri_bmin = bmin.data();
    a.tri_bmax = bmax.data();
    a.tri_centroid = cent.data();
    a.order = order.data();
    a.order_next = order_next.data();
    a.node_of = node_of.data();
    a.node_of_next = node_of_next.data();
    a.flags = flags.data();
    a.scan = scan.data();
    a.items = items.data();
    a.items_next = items_next.data();
    a.accum = accum.data();
    a.bins = bins.data();
    a.counters = counters.data();
    a.bins_capacity = bins_cap;
    a.sub_roots = g_parity_sub_max > 0 ? sub_roots.data() : nullptr;
    a.sub_capacity = (int)items_cap;
    a.sub_max = g_parity_sub_max;
    a.nodes = out_nodes;
    a.node_parent = out_parent;
    a.node_count = out_count;
    std::memset(out_nodes, 0xCD, sizeof(GpuPtBvh) * (size_t)n_nodes);
    EmuExec ex;
    BuildReport rep{};
    const int rc = build_bvh_parity(ex, a, tri_off, tri_off, n_meshes, scratch.data(), &rep);
    if (out_levels) *out_levels = rep.levels;
    if (out_error_bits) *out_error_bits = rep.error_bits;
    return rc;
}

// The four-wide records (Q4Node, n_prims + 8 of them) the next emu_wbvh_build / emu_wbvh_refit writes and emu_render
// (fast == 4) walks; null = not built.  A binding instead of one more argument on three entry points.
static Q4Node* g_q4nodes = nullptr;
EMU_API void emu_bind_q4(Q4Node* q4nodes) { g_q4nodes = q4nodes; }
// textured shading inputs (trace_body.h shade_hit_textured); used by emu_render mode 6
static const FastUv* g_fuvs = nullptr;
static const TexInfo* g_tex_table = nullptr;
static const F4* g_tex_data = nullptr;
static const int* g_mat_maps = nullptr;
static const F4* g_env = nullptr;
static int g_env_w = 0, g_env_h = 0;
EMU_API void emu_bind_textures(const FastUv* fuvs, const TexInfo* table, const F4* texels, const int* mat_maps, const F4* env, int env_w, int env_h) {
    g_fuvs = fuvs;
    g_tex_table = table;
    g_tex_data = texels;
    g_mat_maps = mat_maps;
    g_env = env;
    g_env_w = env_w;
    g_env_h = env_h;
}
// tex_sample (trace_body.h) at n points: uv = 2 floats each, out = 4 floats each
EMU_API void emu_tex_sample(const F4* texels, int w, int h, const float* uv, int n, float* out) {
    for (int i = 0; i < n; ++i) {
        const F4 c = tex_sample(texels, w, h, uv[2 * i], uv[2 * i + 1]);
        out[4 * i] = c.x;
        out[4 * i + 1] = c.y;
        out[4 * i + 2] = c.z;
        out[4 * i + 3] = c.w;
    }
}
static unsigned char* g_pair_depth = nullptr;  // n_prims + 8 bytes, written by the SAH build, read by the refit
EMU_API void emu_bind_pair_depth(unsigned char* p) { g_pair_depth = p; }

// Production (world-space) BVH exactly as cavept_api.cu builds it.  inst_prim_off has
// n_inst + 1 entries, inst_tri_off n_inst.  fnodes needs 2 * (n_prims + 8) entries, qpairs n_prims + 8, ftris /
// fnormals n_prims.  Returns 0 ok; report = {levels, n_pairs, root_ref, error_bits}.
EMU_API int emu_wbvh_build(const GpuPtVertex* vertices, const GpuPtIndex* indices, const GpuPtMesh* instances, int n_inst, const int* inst_prim_off,
                           const int* inst_tri_off, int force_median /* 2: LBVH builder, 3: its balanced fallback, 4: hybrid (radix tree + SAH treelets) */, FastNode* fnodes, QPair* qpairs, FastTri* ftris, FastTri* fnormals, int* report, float* out_pad,
                           float* out_grid) {
    const int P = inst_prim_off[n_inst];
    const size_t items_cap = (size_t)P / 2 + 8, pairs_cap = (size_t)P + 8;
    std::vector<F4> bmin(P), bmax(P);
    std::vector<int> order(P), order_next(P), node_of(P), node_of_next(P), scan(P + 1), counters(cavept::kWCounters);
    std::vector<unsigned char> flags(P);
    const int sub_cap = P / 128 + 64;
    std::vector<WItem> items(items_cap), items_next(items_cap), sub_roots(sub_cap);
    const size_t bins_cap = (size_t)P / (cavept::kWDirectMax + 1) + 2;
    std::vector<WBins> bins(bins_cap);
    std::vector<uint32_t> root_acc(12);
    WBuild b{};
    b.vertices = vertices;
    b.indices = indices;
    b.instances = instances;
    b.inst_prim_off = inst_prim_off;
    b.inst_tri_off = inst_tri_off;
    b.n_inst = n_inst;
    b.n_prims = P;
    b.force_median = (force_median == 1 || force_median == 3) ? 1 : 0;  // 3: LBVH with the balanced fallback forced
    std::vector<uint32_t> keys(P), keys_tmp(P);
    std::vector<F4> nbox_mn(pairs_cap), nbox_mx(pairs_cap);
    b.keys = keys.data();
    b.keys_tmp = keys_tmp.data();
    b.sort_hist = nullptr;
    b.nbox_mn = nbox_mn.data();
    b.nbox_mx = nbox_mx.data();
    b.bmin = bmin.data();
    b.bmax = bmax.data();
    b.order = order.data();
    b.order_next = order_next.data();
    b.node_of = node_of.data();
    b.node_of_next = node_of_next.data();
    b.flags = flags.data();
    b.scan = scan.data();
    b.items = items.data();
    b.items_next = items_next.data();
    b.bins = bins.data();
    b.items_capacity = (int)items_cap;
    b.bins_capacity = (int)bins_cap;
    b.sub_roots = sub_roots.data();
    b.sub_capacity = sub_cap;
    b.pairs_capacity = (int)pairs_cap;
    b.counters = counters.data();
    b.root_acc = root_acc.data();
    b.fnodes = fnodes;
    b.qpairs = qpairs;
    b.q4nodes = g_q4nodes;
    std::vector<unsigned char> depth_scratch(pairs_cap);
    b.pair_depth = g_pair_depth ? g_pair_depth : depth_scratch.data();
    b.ftris = ftris;
    b.fnormals = fnormals;
    EmuExec ex;
    WReport rep{};
    const int rc = force_median >= 2 ? build_lbvh(ex, b, &rep, force_median == 4) : build_wbvh(ex, b, &rep);
    report[0] = rep.levels;
    report[1] = rep.n_pairs;
    report[2] = wb_quantised_ref(rep.root_ref);
    report[3] = rep.error_bits;
    if (force_median < 2) {
        for (int L = 0; L < rep.levels && L < 124; ++L) report[4 + L] = rep.level_items[L];  // report has 4 + 124 ints
    }
    if (out_pad) *out_pad = rep.pad;
    if (out_grid) {
        for (int k = 0; k < 3; ++k) {
            out_grid[k] = rep.grid_origin[k];
            out_grid[3 + k] = rep.grid_scale[k];
        }
    }
    return rc;
}

// Refit of an emu_wbvh_build tree after the instance transforms changed (wbvh_body.h refit_wbvh), as
// cavept_update_instances drives it.  n_pairs / n_levels = WReport of the build; out = scene box + area sum.
EMU_API void emu_wbvh_refit(const GpuPtVertex* vertices, const GpuPtIndex* indices, const GpuPtMesh* instances, int n_inst, int n_prims, float pad, const float* grid,
                            int n_pairs, int n_levels, const unsigned char* moved, FastNode* fnodes, QPair* qpairs, FastTri* ftris, FastTri* fnormals, float* out) {
    WBuild b{};
    b.vertices = vertices;
    b.indices = indices;
    b.instances = instances;
    b.n_inst = n_inst;
    b.n_prims = n_prims;
    b.pad = pad;
    b.fnodes = fnodes;
    b.qpairs = qpairs;
    b.q4nodes = g_q4nodes;
    b.ftris = ftris;
    b.fnormals = fnormals;
    for (int k = 0; k < 3; ++k) {
        b.grid_origin[k] = grid[k];
        b.grid_scale[k] = grid[3 + k];
    }
    b.pair_depth = g_pair_depth;
    float area = 0.0f;
    EmuExec ex;
    refit_wbvh(ex, b, n_pairs, n_levels, moved, &area, out);
}

EMU_API void emu_mat4_inverse(const float* m, float* out) { mat4_inverse(m, out); }

struct EmuCamera {
    float position[3];
    float fov_degree;
    float forward[3];
    float right[3];
    float up[3];
};

static TraceParams emu_params(const GpuPtVertex* vertices, const GpuPtIndex* indices, const GpuPtBvh* bvhs, const GpuPtMesh* meshes, int mesh_count,
                              const GpuPtMaterial* materials, const FastNode* fnodes, const QPair* qpairs, const float* grid, const FastTri* ftris, const FastTri* fnormals, int root_ref,
                              const EmuCamera* cam, int max_bounce, int width, int height, int first_frame, int n_frames, int reset_first, float* accum) {
    TraceParams p{};
    p.vertices = vertices;
    p.indices = indices;
    p.bvhs = bvhs;
    p.meshes = meshes;
    p.materials = materials;
    p.mesh_count = mesh_count;
    p.fnodes = fnodes;
    p.qpairs = qpairs;
    p.q4nodes = g_q4nodes;
    for (int k = 0; k < 3; ++k) {
        p.grid_origin[k] = grid[k];
        p.grid_scale[k] = grid[3 + k];
    }
    p.ftris = ftris;
    p.fnormals = fnormals;
    p.root_ref = root_ref;
    std::memcpy(p.cam_pos, cam->position, 12);
    std::memcpy(p.cam_right, cam->right, 12);
    std::memcpy(p.cam_up, cam->up, 12);
    std::memcpy(p.cam_fwd, cam->forward, 12);
    p.cam_distance = std::tan(cam->fov_degree * 3.141592653589793f / 180.0f);
    p.width = width;
    p.height = height;
    p.max_bounce = max_bounce;
    p.first_frame = first_frame;
    p.n_frames = n_frames;
    p.reset_first = reset_first;
    p.accum = reinterpret_cast<F4*>(accum);
    return p;
}

EMU_API long long emu_render(const GpuPtVertex* vertices, const GpuPtIndex* indices, const GpuPtBvh* bvhs, const GpuPtMesh* meshes, int mesh_count,
                             const GpuPtMaterial* materials, const FastNode* fnodes, const QPair* qpairs, const float* grid, const FastTri* ftris, const FastTri* fnormals, int root_ref,
                             const EmuCamera* cam, int max_bounce, int width, int height, int first_frame, int n_frames, int reset_first, float* accum,
                             int fast, unsigned long long* counters) {
    TraceParams p = emu_params(vertices, indices, bvhs, meshes, mesh_count, materials, fnodes, qpairs, grid, ftris, fnormals, root_ref, cam, max_bounce, width, height, first_frame,
                               n_frames, reset_first, accum);
    unsigned long long cnt[3] = { 0, 0, 0 };
    p.counters = cnt;
    if (fast == 6) {
        p.fuvs = g_fuvs;
        p.tex_table = g_tex_table;
        p.tex_data = g_tex_data;
        p.mat_maps = g_mat_maps;
        p.env = g_env;
        p.env_w = g_env_w;
        p.env_h = g_env_h;
    }
    long long segs = 0;
    int stack[kStackSize];
    if (fast == 2) {  // the persistent job-queue body (trace_pool_body.h): one emulated lane draws every job, two at a time
        float rec[kPoolWordsPerThread];
        int job_counter = 0;
        segs = body_trace_pool<true>(p, stack, rec, 1, &job_counter, cnt);
        if (counters) {
            counters[0] = (unsigned long long)segs;
            counters[1] = cnt[1];
            counters[2] = cnt[2];
        }
        return segs;
    }
    if (fast == 3) {  // the per-pixel body as a persistent lane drawing jobs from the queue (k_trace_queue)
        int job_counter = 0;
        const TraceJobs jobs = trace_jobs(p);
        const int job = trace_draw_jobs(&job_counter, true);
        int x = p.width, y = 0;
        if (job < jobs.n_jobs) {
            if (!trace_job_pixel(jobs, job, x, y)) x = p.width;
        }
        segs = body_trace_pixel_fast<true, kStepQuorum, kShadeBatch, true>(p, x, y, stack, 1, cnt, &job_counter);
        if (counters) {
            counters[0] = (unsigned long long)segs;
            counters[1] = cnt[1];
            counters[2] = cnt[2];
        }
        return segs;
    }
    int spill[kWideSpill];
    for (int y = 0; y < height; ++y) {
        for (int x = 0; x < width; ++x) {
            if (fast == 4) segs += body_trace_pixel_fast<true, kStepQuorum, kShadeBatch, false, true>(p, x, y, stack, 1, cnt, nullptr, spill);  // four-wide walk (k_trace_wide)
            else if (fast == 6) segs += body_trace_pixel_fast<true, kStepQuorum, kShadeBatch, false, false, true>(p, x, y, stack, 1, cnt);  // textured instantiation
            else segs += fast ? body_trace_pixel_fast<true>(p, x, y, stack, 1, cnt) : body_trace_pixel_reference<true>(p, x, y, cnt);
        }
    }
    if (counters) {
        counters[0] = (unsigned long long)segs;
        counters[1] = cnt[1];
        counters[2] = cnt[2];
    }
    return segs;
}

// Closest primary-ray hit per pixel (frame `frame`): mode 0 = every triangle slot tested in turn with the kernel's own
// triangle arithmetic (no tree: the ground truth a traversal may not lose), 1 = the production walk.
// out_slot[pixel] = FastTri slot or -1, out_t[pixel] = ray.t after the query.
static void emu_walk_one(const TraceParams& p, Walk& w, Ray& ray) {
    walk_begin(p, w, ray);
    for (;;) {
        walk_park_leaf(w);
        if (ref_is_interior(w.ref)) walk_step<false>(p, w, ray.t, nullptr);
        else if (w.q0 != 0) walk_test_triangle<false>(p, w, ray, ray.t, nullptr);
        else break;
    }
}
EMU_API void emu_primary_hits(const FastNode* fnodes, const QPair* qpairs, const float* grid, const FastTri* ftris, int n_slots, int root_ref, const EmuCamera* cam,
                              int width, int height, int frame, int mode, int* out_slot, float* out_t) {
    TraceParams p = emu_params(nullptr, nullptr, nullptr, nullptr, 0, nullptr, fnodes, qpairs, grid, ftris, nullptr, root_ref, cam, 1, width, height, frame, 1, 1, nullptr);
    int stack[kStackSize];
    for (int y = 0; y < height; ++y) {
        for (int x = 0; x < width; ++x) {
            Ray ray;
            uint32_t seed;
            start_path(p, x, y, frame, ray, seed);
            int slot = -1;
            if (mode == 0) {
                for (int s = 0; s < n_slots; ++s) {
                    float u, v;
                    const FastTri& t = ftris[s];
                    if (hit_triangle_edges(f3(t.v0.x, t.v0.y, t.v0.z), f3(t.v1.x, t.v1.y, t.v1.z), f3(t.v2.x, t.v2.y, t.v2.z), ray.origin, ray.direction, ray.t, u, v)) slot = s;
                }
            } else {
                Walk w;
                w.stack = stack;
                w.stride = 1;
                w.spill = nullptr;
                emu_walk_one(p, w, ray);
                slot = w.hit_slot;
            }
            out_slot[(size_t)y * width + x] = slot;
            out_t[(size_t)y * width + x] = ray.t;
        }
    }
}

#if defined(CV_WARP_EMU)
// ---- warps of 32 host threads: the votes (__any_sync / __ballot_sync) and the warp-aggregated job draw go through a
//      barrier, so the warp-level scheduling of the persistent kernel (who draws, who retires, when the loop ends)
//      runs with real divergence between lanes.  A thread outside any warp (the serial entry points) is a warp of one.
#include <atomic>
#include <barrier>
#include <thread>
namespace {
struct EmuWarp {
    std::barrier<> bar{ 32 };
    std::atomic<unsigned> word[3];
    int bcast = 0;
    EmuWarp() {
        for (auto& w : word) w.store(0u);
    }
};
thread_local EmuWarp* tl_warp = nullptr;
thread_local int tl_lane = 0;
thread_local unsigned tl_round = 0;
unsigned emu_ballot(bool pred) {
    EmuWarp* w = tl_warp;
    const unsigned r = tl_round++ % 3u;
    if (tl_lane == 0) w->word[(r + 1u) % 3u].store(0u);
    if (pred) w->word[r].fetch_or(1u << tl_lane);
    w->bar.arrive_and_wait();
    return w->word[r].load();
}
}  // namespace
namespace cavept {
bool warp_emu_any(bool pred) { return tl_warp ? emu_ballot(pred) != 0u : pred; }
int warp_emu_count(bool pred) { return tl_warp ? __builtin_popcount(emu_ballot(pred)) : (pred ? 1 : 0); }
int warp_emu_draw(int* counter, bool want) {
    if (!tl_warp) {
        if (!want) return 0x7fffffff;
        return (*counter)++;
    }
    const unsigned m = emu_ballot(want);
    if (m == 0u) return 0x7fffffff;
    const int leader = __builtin_ctz(m);
    if (tl_lane == leader) tl_warp->bcast = __atomic_fetch_add(counter, __builtin_popcount(m), __ATOMIC_RELAXED);
    tl_warp->bar.arrive_and_wait();
    const int base = tl_warp->bcast;
    tl_warp->bar.arrive_and_wait();  // nobody overwrites bcast before everyone has read it
    return want ? base + __builtin_popcount(m & ((1u << tl_lane) - 1u)) : 0x7fffffff;
}
}  // namespace cavept

// k_trace_queue with n_warps concurrent warps of 32 threads each; returns segments, -1 if the job counter ended wrong
EMU_API long long emu_render_warps(const GpuPtVertex* vertices, const GpuPtIndex* indices, const GpuPtBvh* bvhs, const GpuPtMesh* meshes, int mesh_count,
                                   const GpuPtMaterial* materials, const FastNode* fnodes, const QPair* qpairs, const float* grid, const FastTri* ftris, const FastTri* fnormals,
                                   int root_ref, const EmuCamera* cam, int max_bounce, int width, int height, int first_frame, int n_frames, int reset_first, float* accum,
                                   int n_warps, const int* tile, unsigned long long* counters) {
    TraceParams p = emu_params(vertices, indices, bvhs, meshes, mesh_count, materials, fnodes, qpairs, grid, ftris, fnormals, root_ref, cam, max_bounce, width, height, first_frame,
                               n_frames, reset_first, accum);
    if (tile) {
        p.tile_x0 = tile[0];
        p.tile_y0 = tile[1];
        p.tile_x1 = tile[2];
        p.tile_y1 = tile[3];
    }
    unsigned long long shared_cnt[3] = { 0, 0, 0 };
    p.counters = shared_cnt;
    int job_counter = 0;
    std::atomic<long long> segs{ 0 };
    std::atomic<unsigned long long> visits[3];
    for (auto& v : visits) v.store(0);
    std::vector<EmuWarp> warps(n_warps);
    std::vector<std::thread> threads;
    for (int wi = 0; wi < n_warps; ++wi) {
        for (int lane = 0; lane < 32; ++lane) {
            threads.emplace_back([&, wi, lane]() {
                tl_warp = &warps[wi];
                tl_lane = lane;
                tl_round = 0;
                int stack[kStackSize];
                unsigned long long cnt[3] = { 0, 0, 0 };
                const TraceJobs jobs = trace_jobs(p);
                const int job = trace_draw_jobs(&job_counter, true);
                int x = p.width, y = 0;
                if (job < jobs.n_jobs) {
                    if (!trace_job_pixel(jobs, job, x, y)) x = p.width;
                }
                segs += body_trace_pixel_fast<true, kStepQuorum, kShadeBatch, true>(p, x, y, stack, 1, cnt, &job_counter);
                for (int k = 0; k < 3; ++k) visits[k] += cnt[k];
                tl_warp = nullptr;
            });
        }
    }
    for (auto& t : threads) t.join();
    if (counters) {
        counters[0] = (unsigned long long)segs.load();
        counters[1] = visits[1].load();
        counters[2] = visits[2].load();
    }
    return segs.load();
}
#endif  // CV_WARP_EMU

// the product's HitTriangle / HitBvh bodies on caller-supplied cases
EMU_API void emu_probe(int n, const float* rays, const float* tris, const float* boxes, float* out) {
    for (int i = 0; i < n; ++i) cavept::body_probe(rays, tris, boxes, out, i);
}
