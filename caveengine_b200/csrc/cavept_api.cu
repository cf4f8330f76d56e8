// C ABI of libcavept (include/cavept.h): context, scene upload, BVH build, render,
// read-back.  Host logic mirrors cave/engine/renderer/path_tracer/path_tracer.cpp;
// all geometry/tracing work is done by the kernels in kernels.cu.  No CPU path.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "cavept_ctx.h"
#include "host_math.h"

namespace {

// GpuPtMaterial from MaterialComponent, path_tracer.cpp:140-145
GpuPtMaterial make_gpu_material(const cavept_material& m) {
    GpuPtMaterial g;
    for (int k = 0; k < 3; ++k) {
        g.baseColor[k] = m.base_color[k];
        g.emissive[k] = g.baseColor[k] * m.emissive;
    }
    g.roughness = m.roughness;
    g.metallic = m.metallic;
    return g;
}

// UpdateAccelStructure (path_tracer.cpp:120-201): materials de-duplicated in mesh-key
// order, instance array with world matrix + glm::inverse.
// material maps in GpuPtMaterial-slot order (what the kernel indexes with GpuPtMesh::materialId)
int upload_material_maps(cavept_ctx* ctx) {
    if (!ctx->have_maps) return CAVEPT_OK;
    std::vector<int> maps(2 * ctx->gpu_material_src.size() + 2, -1);
    for (size_t g = 0; g < ctx->gpu_material_src.size(); ++g) {
        const size_t m = (size_t)ctx->gpu_material_src[g];
        if (2 * m + 1 < ctx->api_mat_maps.size()) {
            maps[2 * g] = ctx->api_mat_maps[2 * m];
            maps[2 * g + 1] = ctx->api_mat_maps[2 * m + 1];
        }
    }
    CU(ctx, ctx->d_mat_maps.reserve(maps.size()));
    ctx->ex.upload(ctx->d_mat_maps.ptr, maps.data(), maps.size() * sizeof(int));
    CU(ctx, cudaStreamSynchronize(ctx->ex.stream));
    return check_exec(ctx, "material maps");
}

int rebuild_instance_arrays(cavept_ctx* ctx) {
    std::map<int, int> lookup;  // material id -> GpuPtMaterial slot
    ctx->gpu_materials.clear();
    ctx->gpu_material_src.clear();
    for (size_t mesh_id = 0; mesh_id < ctx->meshes.size(); ++mesh_id) {  // std::map<Guid, MeshData> iteration == ascending key
        const MeshRecord& rec = ctx->meshes[mesh_id];
        if (rec.built < 0) continue;
        if (lookup.count(rec.first_material)) continue;
        lookup[rec.first_material] = (int)ctx->gpu_materials.size();
        ctx->gpu_materials.push_back(make_gpu_material(ctx->materials[rec.first_material]));
        ctx->gpu_material_src.push_back(rec.first_material);
    }
    ctx->gpu_meshes.resize(ctx->instances.size());
    for (size_t i = 0; i < ctx->instances.size(); ++i) {
        const cavept_instance& in = ctx->instances[i];
        const MeshRecord& rec = ctx->meshes[in.mesh];
        GpuPtMesh& g = ctx->gpu_meshes[i];
        std::memset(&g, 0, sizeof(g));
        std::memcpy(g.transform, in.transform, sizeof(g.transform));
        mat4_inverse(g.transform, g.transformInv);
        g.rootBvhId = rec.node_off;
        g.materialId = lookup[rec.first_material];
    }
    CU(ctx, ctx->d_meshes.reserve(ctx->gpu_meshes.size()));
    CU(ctx, ctx->d_materials.reserve(ctx->gpu_materials.size()));
    ctx->ex.upload(ctx->d_meshes.ptr, ctx->gpu_meshes.data(), ctx->gpu_meshes.size() * sizeof(GpuPtMesh));
    ctx->ex.upload(ctx->d_materials.ptr, ctx->gpu_materials.data(), ctx->gpu_materials.size() * sizeof(GpuPtMaterial));
    CU(ctx, cudaStreamSynchronize(ctx->ex.stream));
    if (int e = check_exec(ctx, "instance upload")) return e;
    return upload_material_maps(ctx);
}

// Production BVH: one world-space tree over every (instance, triangle) pair (wbvh_body.h).
// Needs d_vertices / d_indices / d_meshes; rerun whenever geometry or a transform changes.
int build_world(cavept_ctx* ctx) {
    const int n_inst = (int)ctx->instances.size();
    std::vector<int> prim_off(n_inst + 1), tri_off(n_inst);
    long long total = 0;
    for (int i = 0; i < n_inst; ++i) {
        const MeshRecord& rec = ctx->meshes[ctx->instances[i].mesh];
        prim_off[i] = (int)total;
        tri_off[i] = rec.tri_off;
        total += rec.tri_count;
        if (total > (1ll << 28)) return fail(ctx, CAVEPT_ERR_CAPACITY, "more than 2^28 instanced triangles");
    }
    prim_off[n_inst] = (int)total;
    const int P = (int)total;
    ctx->n_prims = P;
    // interior items of a level are disjoint ranges of >= kFastLeafMax + 1 prims; only those of more than kWDirectMax prims own arena bins
    const size_t items_cap = (size_t)P / (kFastLeafMax + 1) + 8, pairs_cap = (size_t)P + 8, bins_cap = (size_t)P / (kWDirectMax + 1) + 2, sub_cap = (size_t)P / (ctx->cfg.bvh_mode == CAVEPT_BVH_HYBRID ? 16 : 128) + 64;  // hybrid: a treelet next to a large sibling can be a sliver
    CU(ctx, ctx->d_inst_prim_off.reserve(n_inst + 1));
    CU(ctx, ctx->d_inst_tri_off.reserve(n_inst));
    CU(ctx, ctx->d_prim_bmin.reserve(P));
    CU(ctx, ctx->d_prim_bmax.reserve(P));
    CU(ctx, ctx->d_w_order.reserve(P));
    CU(ctx, ctx->d_w_order_next.reserve(P));
    CU(ctx, ctx->d_w_node_of.reserve(P));
    CU(ctx, ctx->d_w_node_of_next.reserve(P));
    CU(ctx, ctx->d_w_flags.reserve(P));
    CU(ctx, ctx->d_w_scan.reserve((size_t)P + 1));
    CU(ctx, ctx->d_scan_tmp.reserve((size_t)P / kScanTile + 4));
    const bool hybrid = ctx->cfg.bvh_mode == CAVEPT_BVH_HYBRID;  // the radix tree above, SAH treelets below
    const bool lbvh = ctx->cfg.bvh_mode == CAVEPT_BVH_LBVH || hybrid;
    if (lbvh) {
        CU(ctx, ctx->d_w_keys.reserve(P));
        CU(ctx, ctx->d_w_keys_tmp.reserve(P));
        CU(ctx, ctx->d_w_sort_hist.reserve(2 * sort_hist_entries(P)));
        CU(ctx, ctx->d_w_nbox_mn.reserve(pairs_cap));
        CU(ctx, ctx->d_w_nbox_mx.reserve(pairs_cap));
        CU(ctx, ctx->d_scan_tmp.reserve(std::max((size_t)P, sort_hist_entries(P)) / kScanTile + 4));
        if (hybrid) {  // the subtree launch's queue and the depth tags it writes; it works out of shared memory otherwise
            CU(ctx, ctx->d_pair_depth.reserve(pairs_cap));
            CU(ctx, ctx->d_w_sub_roots.reserve(sub_cap));
        }
    } else {
        CU(ctx, ctx->d_w_items.reserve(items_cap));
        CU(ctx, ctx->d_w_items_next.reserve(items_cap));
        CU(ctx, ctx->d_w_bins.reserve(bins_cap));
        CU(ctx, ctx->d_pair_depth.reserve(pairs_cap));
        CU(ctx, ctx->d_w_sub_roots.reserve(sub_cap));
    }
    CU(ctx, ctx->d_w_counters.reserve(kWCounters));
    CU(ctx, ctx->d_w_root_acc.reserve(12));
    CU(ctx, ctx->d_fnodes.reserve(2 * pairs_cap));
    CU(ctx, ctx->d_qpairs.reserve(pairs_cap));
    const bool want_q4 = ctx->kernel_variant == CAVEPT_KERNEL_WIDE4;  // the four-wide records exist only while that variant is selected
    if (want_q4) CU(ctx, ctx->d_q4nodes.reserve(pairs_cap));
    CU(ctx, ctx->d_ftris.reserve(P));
    CU(ctx, ctx->d_fnormals.reserve(P));
    ctx->ex.scan_tmp = ctx->d_scan_tmp.ptr;
    ctx->ex.upload(ctx->d_inst_prim_off.ptr, prim_off.data(), prim_off.size() * sizeof(int));
    ctx->ex.upload(ctx->d_inst_tri_off.ptr, tri_off.data(), tri_off.size() * sizeof(int));
    CU(ctx, cudaStreamSynchronize(ctx->ex.stream));

    WBuild b{};
    b.vertices = ctx->d_vertices.ptr;
    b.indices = ctx->d_indices.ptr;
    b.instances = ctx->d_meshes.ptr;
    b.inst_prim_off = ctx->d_inst_prim_off.ptr;
    b.inst_tri_off = ctx->d_inst_tri_off.ptr;
    b.n_inst = n_inst;
    b.n_prims = P;
    b.bmin = ctx->d_prim_bmin.ptr;
    b.bmax = ctx->d_prim_bmax.ptr;
    b.order = ctx->d_w_order.ptr;
    b.order_next = ctx->d_w_order_next.ptr;
    b.node_of = ctx->d_w_node_of.ptr;
    b.node_of_next = ctx->d_w_node_of_next.ptr;
    b.flags = ctx->d_w_flags.ptr;
    b.scan = ctx->d_w_scan.ptr;
    b.items = ctx->d_w_items.ptr;
    b.items_next = ctx->d_w_items_next.ptr;
    b.bins = ctx->d_w_bins.ptr;
    b.items_capacity = (int)items_cap;
    b.pairs_capacity = (int)pairs_cap;
    b.bins_capacity = (int)bins_cap;
    b.pair_depth = ctx->d_pair_depth.ptr;
    b.sub_roots = ctx->d_w_sub_roots.ptr;
    b.sub_capacity = (int)sub_cap;
    b.counters = ctx->d_w_counters.ptr;
    b.root_acc = ctx->d_w_root_acc.ptr;
    b.keys = ctx->d_w_keys.ptr;
    b.keys_tmp = ctx->d_w_keys_tmp.ptr;
    b.sort_hist = ctx->d_w_sort_hist.ptr;
    b.nbox_mn = ctx->d_w_nbox_mn.ptr;
    b.nbox_mx = ctx->d_w_nbox_mx.ptr;
    b.fnodes = ctx->d_fnodes.ptr;
    b.qpairs = ctx->d_qpairs.ptr;
    b.q4nodes = want_q4 ? ctx->d_q4nodes.ptr : nullptr;
    b.ftris = ctx->d_ftris.ptr;
    b.fnormals = ctx->d_fnormals.ptr;

    const unsigned long long launches_before = ctx->ex.launches;
    CU(ctx, cudaEventRecord(ctx->ev0, ctx->ex.stream));
    WReport rep{};
    int rc = lbvh ? run_build_lbvh(ctx->ex, b, &rep, hybrid) : run_build_wbvh(ctx->ex, b, &rep);
    if (!lbvh && rc == 0 && rep.levels + 2 > kStackSize) {  // deeper than the traversal stack: rebuild balanced (median splits)
        b.force_median = 1;
        rc = run_build_wbvh(ctx->ex, b, &rep);
    }
    CU(ctx, cudaEventRecord(ctx->ev1, ctx->ex.stream));
    CU(ctx, cudaStreamSynchronize(ctx->ex.stream));
    if (int e = check_exec(ctx, "world bvh build")) return e;
    if (rc != 0) return fail(ctx, CAVEPT_ERR_CAPACITY, "world BVH build failed (bits %d)", rep.error_bits);
    if (rep.levels + 2 > kStackSize) return fail(ctx, CAVEPT_ERR_CAPACITY, "world BVH depth %d exceeds the traversal stack (%d)", rep.levels, kStackSize);
    float ms = 0.0f;
    CU(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    ctx->stats.last_fast_build_ms = ms;
    ctx->stats.fast_build_levels = rep.levels;
    ctx->stats.fast_build_launches = (int)(ctx->ex.launches - launches_before);
    ctx->n_pairs = rep.n_pairs;
    ctx->root_ref = wb_quantised_ref(rep.root_ref);
    for (int k = 0; k < 3; ++k) {
        ctx->grid_origin[k] = rep.grid_origin[k];
        ctx->grid_scale[k] = rep.grid_scale[k];
    }
    ctx->fast_levels = rep.levels;
    ctx->q4_ready = want_q4;
    ctx->uvs_dirty = true;  // slots were re-ordered
    ctx->stats.fast_node_count = 2 * rep.n_pairs;
    // what a later refit needs (SAH builder: the pairs of a level are one index range)
    ctx->refit_ready = !lbvh && rep.levels > 0;
    ctx->world_pad = rep.pad;
    ctx->world_extent = rep.pad * 1048576.0f;
    ctx->area_baseline = -1.0f;
    return CAVEPT_OK;
}

// Instance transforms changed, topology kept: re-transform the moved instances' triangle slots and recompute every
// box bottom-up (wbvh_body.h refit_wbvh).  Returns 1 when the refitted tree is kept, 0 when the caller must rebuild
// (scene left the quantisation grid / the padding's scale, or the tree's surface-area sum grew past kRefitQuality),
// < 0 on error.
constexpr float kRefitQuality = 1.3f;
int refit_world(cavept_ctx* ctx, const std::vector<unsigned char>& moved) {
    if (!ctx->refit_ready || ctx->n_pairs <= 0) return 0;
    const int n_inst = (int)ctx->instances.size();
    if (cudaSuccess != ctx->d_moved.reserve(n_inst) || cudaSuccess != ctx->d_area_sum.reserve(1)) return -fail(ctx, CAVEPT_ERR_CUDA, "refit scratch");
    WBuild b{};
    b.vertices = ctx->d_vertices.ptr;
    b.indices = ctx->d_indices.ptr;
    b.instances = ctx->d_meshes.ptr;
    b.n_inst = n_inst;
    b.n_prims = ctx->n_prims;
    b.pad = ctx->world_pad;
    b.fnodes = ctx->d_fnodes.ptr;
    b.qpairs = ctx->d_qpairs.ptr;
    b.q4nodes = ctx->q4_ready ? ctx->d_q4nodes.ptr : nullptr;
    b.ftris = ctx->d_ftris.ptr;
    b.fnormals = ctx->d_fnormals.ptr;
    for (int k = 0; k < 3; ++k) {
        b.grid_origin[k] = ctx->grid_origin[k];
        b.grid_scale[k] = ctx->grid_scale[k];
    }
    const int n_levels = ctx->fast_levels;
    b.pair_depth = ctx->d_pair_depth.ptr;
    float out[7];
    if (ctx->area_baseline < 0.0f) {  // first update after a build: the built tree's own surface-area sum (nothing moved: the pass reproduces the tree)
        ctx->ex.fill(ctx->d_moved.ptr, 0, (size_t)n_inst);
        run_refit_wbvh(ctx->ex, b, ctx->n_pairs, n_levels, ctx->d_moved.ptr, ctx->d_area_sum.ptr, out);
        ctx->area_baseline = out[6];
    }
    ctx->ex.upload(ctx->d_moved.ptr, moved.data(), (size_t)n_inst);
    run_refit_wbvh(ctx->ex, b, ctx->n_pairs, n_levels, ctx->d_moved.ptr, ctx->d_area_sum.ptr, out);
    if (check_exec(ctx, "world bvh refit")) return -CAVEPT_ERR_CUDA;
    float extent = 0.0f;
    bool inside = true;
    for (int c = 0; c < 3; ++c) {
        const float lo = out[c], hi = out[3 + c];
        extent = std::fmax(extent, std::fmax(std::fabs(lo), std::fabs(hi)));
        const float g_lo = (lo - ctx->grid_origin[c]) * ctx->grid_scale[c], g_hi = (hi - ctx->grid_origin[c]) * ctx->grid_scale[c];
        if (!(g_lo >= 1.0f && g_hi <= kGridCells - 1.0f)) inside = false;  // outward rounding must not meet the clamp at 0 / 32767
    }
    // the padding was derived from the old extent (pad = extent * 2^-20 against ~2^-23 of slab-test rounding: 8x slack), and
    // the refitted root box itself carries the pad, so a scene that has not grown reads extent * (1 + 2^-20) here
    if (!inside || extent > ctx->world_extent * 1.001f + 1e-30f) return 0;
    if (!(out[6] <= kRefitQuality * ctx->area_baseline)) return 0;
    return 1;
}

int build_all(cavept_ctx* ctx) {
    const int n_built = (int)ctx->built_order.size();
    const int T = ctx->n_tris;
    const int n_nodes = ctx->n_nodes;
    std::vector<int> tri_off(n_built + 1), vert_off(n_built + 1);
    for (int k = 0; k < n_built; ++k) {
        const MeshRecord& rec = ctx->meshes[ctx->built_order[k]];
        tri_off[k] = rec.tri_off;
        vert_off[k] = rec.vert_off;
    }
    tri_off[n_built] = T;
    vert_off[n_built] = ctx->n_verts;
    CU(ctx, ctx->d_tri_off.reserve(n_built + 1));
    CU(ctx, ctx->d_vert_off.reserve(n_built + 1));
    ctx->ex.upload(ctx->d_tri_off.ptr, tri_off.data(), tri_off.size() * sizeof(int));
    ctx->ex.upload(ctx->d_vert_off.ptr, vert_off.data(), vert_off.size() * sizeof(int));

    // raw -> reference layouts
    launch_assemble_vertices(ctx->ex, ctx->d_raw_pos.ptr, ctx->d_raw_nrm.ptr, ctx->d_vertices.ptr, ctx->n_verts);
    launch_assemble_indices(ctx->ex, ctx->d_raw_idx.ptr, ctx->d_tri_off.ptr, ctx->d_vert_off.ptr, n_built, ctx->d_indices.ptr, T);
    ctx->nodes_resident = false;
    ctx->stats.last_build_ms = 0.0f;
    ctx->stats.build_levels = 0;
    ctx->stats.build_launches = 0;
    if (ctx->cfg.bvh_mode != CAVEPT_BVH_PARITY) {  // production tree only: no reference-layout GpuPtBvh array
        CU(ctx, cudaStreamSynchronize(ctx->ex.stream));
        return check_exec(ctx, "scene assembly");
    }

    const size_t items_cap = (size_t)std::max(T, n_built) + 2;
    const size_t bins_cap = (size_t)T / 5 + (size_t)n_built + 2;
    CU(ctx, ctx->d_tri_bmin.reserve(T));
    CU(ctx, ctx->d_tri_bmax.reserve(T));
    CU(ctx, ctx->d_tri_centroid.reserve(T));
    CU(ctx, ctx->d_order.reserve(T));
    CU(ctx, ctx->d_order_next.reserve(T));
    CU(ctx, ctx->d_node_of.reserve(T));
    CU(ctx, ctx->d_node_of_next.reserve(T));
    CU(ctx, ctx->d_flags.reserve(T));
    CU(ctx, ctx->d_scan.reserve((size_t)std::max(T, n_nodes) + 1));
    CU(ctx, ctx->d_scan_tmp.reserve((size_t)std::max(T, n_nodes) / kScanTile + 4));
    CU(ctx, ctx->d_counters.reserve(8));
    CU(ctx, ctx->d_items.reserve(items_cap));
    CU(ctx, ctx->d_items_next.reserve(items_cap));
    CU(ctx, ctx->d_accum_items.reserve(items_cap));
    CU(ctx, ctx->d_bins.reserve(bins_cap));
    CU(ctx, ctx->d_sub_items.reserve(items_cap));
    CU(ctx, ctx->d_nodes.reserve(n_nodes));
    CU(ctx, ctx->d_node_parent.reserve(n_nodes));
    CU(ctx, ctx->d_node_count.reserve(n_nodes));
    ctx->ex.scan_tmp = ctx->d_scan_tmp.ptr;

    BuildArrays a{};
    a.vertices = ctx->d_vertices.ptr;
    a.indices = ctx->d_indices.ptr;
    a.n_tris = T;
    a.tri_bmin = ctx->d_tri_bmin.ptr;
    a.tri_bmax = ctx->d_tri_bmax.ptr;
    a.tri_centroid = ctx->d_tri_centroid.ptr;
    a.order = ctx->d_order.ptr;
    a.order_next = ctx->d_order_next.ptr;
    a.node_of = ctx->d_node_of.ptr;
    a.node_of_next = ctx->d_node_of_next.ptr;
    a.flags = ctx->d_flags.ptr;
    a.scan = ctx->d_scan.ptr;
    a.items = ctx->d_items.ptr;
    a.items_next = ctx->d_items_next.ptr;
    a.accum = ctx->d_accum_items.ptr;
    a.bins = ctx->d_bins.ptr;
    a.counters = ctx->d_counters.ptr;
    a.bins_capacity = (int)bins_cap;
    a.sub_roots = ctx->d_sub_items.ptr;
    a.sub_capacity = (int)items_cap;
    a.sub_max = kParitySubMax;
    a.nodes = ctx->d_nodes.ptr;
    a.node_parent = ctx->d_node_parent.ptr;
    a.node_count = ctx->d_node_count.ptr;

    const unsigned long long launches_before = ctx->ex.launches;
    CU(ctx, cudaEventRecord(ctx->ev0, ctx->ex.stream));
    std::vector<BuildItem> scratch(n_built);
    BuildReport report{};
    const int rc = run_build_bvh_parity(ctx->ex, a, tri_off.data(), ctx->d_tri_off.ptr, n_built, scratch.data(), &report);
    CU(ctx, cudaEventRecord(ctx->ev1, ctx->ex.stream));
    ctx->build_levels = report.levels;
    if (int e = check_exec(ctx, "bvh build")) return e;
    if (rc != 0) {
        return fail(ctx, CAVEPT_ERR_DEGENERATE_GEOMETRY, "BVH build failed (bits %d:%s%s%s%s) -- the reference traps on this input (bvh_accel.cpp:117-120,138-141)",
                    report.error_bits, (report.error_bits & BUILD_ERR_DEPTH) ? " depth>64" : "", (report.error_bits & BUILD_ERR_EMPTY_SIDE) ? " empty-split-side" : "",
                    (report.error_bits & BUILD_ERR_BAD_AREA) ? " zero-area" : "", (report.error_bits & BUILD_ERR_BINS) ? " bins" : "");
    }

    CU(ctx, cudaStreamSynchronize(ctx->ex.stream));
    float ms_build = 0.0f;
    CU(ctx, cudaEventElapsedTime(&ms_build, ctx->ev0, ctx->ev1));
    ctx->stats.last_build_ms = ms_build;
    ctx->stats.build_levels = report.levels;
    ctx->stats.build_launches = (int)(ctx->ex.launches - launches_before);
    ctx->nodes_resident = true;
    return CAVEPT_OK;
}

}  // namespace

extern "C" {

CAVEPT_API int cavept_create(const cavept_config* config, cavept_ctx** out_ctx) {
    if (!config || !out_ctx) return CAVEPT_ERR_INVALID_ARG;
    *out_ctx = nullptr;
    if (config->width <= 0 || config->height <= 0 || config->max_bounce <= 0) return CAVEPT_ERR_INVALID_ARG;
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0) return CAVEPT_ERR_CUDA;
    cavept_ctx* ctx = new cavept_ctx();
    ctx->cfg = *config;
    if (config->device >= 0) {
        ctx->device = config->device;
    } else if (cudaGetDevice(&ctx->device) != cudaSuccess) {
        delete ctx;
        return CAVEPT_ERR_CUDA;
    }
    bool ok = cudaSetDevice(ctx->device) == cudaSuccess;
    ok = ok && cudaStreamCreateWithFlags(&ctx->ex.stream, cudaStreamNonBlocking) == cudaSuccess;
    ok = ok && cudaEventCreate(&ctx->ev0) == cudaSuccess && cudaEventCreate(&ctx->ev1) == cudaSuccess && cudaEventCreate(&ctx->ev2) == cudaSuccess && cudaEventCreate(&ctx->ev3) == cudaSuccess;
    const size_t n_pixels = (size_t)config->width * config->height;
    ok = ok && ctx->d_accum.reserve(n_pixels) == cudaSuccess && ctx->d_stats.reserve(4) == cudaSuccess && ctx->d_job_counter.reserve(4) == cudaSuccess;
    ok = ok && cudaDeviceGetAttribute(&ctx->n_sms, cudaDevAttrMultiProcessorCount, ctx->device) == cudaSuccess;
    ok = ok && cudaMemsetAsync(ctx->d_accum.ptr, 0, n_pixels * sizeof(F4), ctx->ex.stream) == cudaSuccess;
    ok = ok && cudaMemsetAsync(ctx->d_stats.ptr, 0, 4 * sizeof(unsigned long long), ctx->ex.stream) == cudaSuccess;
    ok = ok && cudaStreamSynchronize(ctx->ex.stream) == cudaSuccess;
    if (!ok) {
        cavept_destroy(ctx);
        return CAVEPT_ERR_CUDA;
    }
    *out_ctx = ctx;
    return CAVEPT_OK;
}

CAVEPT_API void cavept_destroy(cavept_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->group) cavept_group_destroy(ctx);
    if (ctx->ex.stream) cudaStreamSynchronize(ctx->ex.stream);
    ctx->d_vertices.release(); ctx->d_indices.release(); ctx->d_nodes.release(); ctx->d_meshes.release(); ctx->d_materials.release();
    ctx->d_fnodes.release(); ctx->d_qpairs.release(); ctx->d_q4nodes.release(); ctx->d_ftris.release(); ctx->d_fnormals.release(); ctx->d_inst_prim_off.release(); ctx->d_inst_tri_off.release();
    ctx->d_prim_bmin.release(); ctx->d_prim_bmax.release(); ctx->d_w_order.release(); ctx->d_w_order_next.release(); ctx->d_w_node_of.release();
    ctx->d_w_node_of_next.release(); ctx->d_w_scan.release(); ctx->d_w_counters.release(); ctx->d_w_flags.release(); ctx->d_w_items.release();
    ctx->d_w_items_next.release(); ctx->d_w_bins.release(); ctx->d_w_root_acc.release();
    ctx->d_w_keys.release(); ctx->d_w_keys_tmp.release(); ctx->d_w_sort_hist.release(); ctx->d_w_nbox_mn.release(); ctx->d_w_nbox_mx.release();
    ctx->d_raw_pos.release(); ctx->d_raw_nrm.release(); ctx->d_raw_idx.release(); ctx->d_tri_off.release(); ctx->d_vert_off.release();
    ctx->d_tri_bmin.release(); ctx->d_tri_bmax.release(); ctx->d_tri_centroid.release();
    ctx->d_order.release(); ctx->d_order_next.release(); ctx->d_node_of.release(); ctx->d_node_of_next.release();
    ctx->d_scan.release(); ctx->d_scan_tmp.release(); ctx->d_counters.release(); ctx->d_flags.release();
    ctx->d_items.release(); ctx->d_items_next.release(); ctx->d_sub_items.release(); ctx->d_accum_items.release(); ctx->d_bins.release();
    ctx->d_node_parent.release(); ctx->d_node_count.release();
    ctx->d_accum.release(); ctx->d_tone.release(); ctx->d_tone8.release(); ctx->d_stats.release(); ctx->d_job_counter.release();
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    if (ctx->ev2) cudaEventDestroy(ctx->ev2);
    if (ctx->ev3) cudaEventDestroy(ctx->ev3);
    if (ctx->ex.stream) cudaStreamDestroy(ctx->ex.stream);
    delete ctx;
}

CAVEPT_API const char* cavept_last_error(const cavept_ctx* ctx) { return ctx ? ctx->last_error.c_str() : "null context"; }

CAVEPT_API int cavept_get_image_size(const cavept_ctx* ctx, int* width, int* height) {
    if (!ctx) return CAVEPT_ERR_INVALID_ARG;
    if (width) *width = ctx->cfg.width;
    if (height) *height = ctx->cfg.height;
    return CAVEPT_OK;
}

CAVEPT_API int cavept_set_mode(cavept_ctx* ctx, int mode) {
    if (!ctx) return CAVEPT_ERR_INVALID_ARG;
    if (mode != CAVEPT_MODE_NONE && mode != CAVEPT_MODE_INTERACTIVE && mode != CAVEPT_MODE_TILED) {
        return fail(ctx, CAVEPT_ERR_INVALID_ARG, "invalid mode %d", mode);  // reference: CRASH_NOW_MSG("Invalid mode"), path_tracer.cpp:59-61
    }
    ctx->mode = mode;
    return CAVEPT_OK;
}

CAVEPT_API int cavept_is_active(const cavept_ctx* ctx) {
    // PathTracer::IsActive, path_tracer.cpp:312-322
    return (ctx && ctx->mode == CAVEPT_MODE_INTERACTIVE && ctx->scene_resident) ? 1 : 0;
}

CAVEPT_API int cavept_upload_scene(cavept_ctx* ctx, const cavept_mesh* meshes, int mesh_count, const cavept_instance* instances, int instance_count,
                                   const cavept_material* materials, int material_count) {
    CAVEPT_RANGE();
    if (!ctx) return CAVEPT_ERR_INVALID_ARG;
    if (!meshes || !instances || !materials || mesh_count <= 0 || instance_count <= 0 || material_count <= 0) {
        return fail(ctx, CAVEPT_ERR_INVALID_ARG, "upload_scene: null or empty input");
    }
    if (int e = select_device(ctx)) return e;
    ctx->scene_resident = false;
    ctx->have_uvs = ctx->have_maps = false;  // texture coordinates and material maps belong to the scene they were set on
    ctx->meshes.assign(mesh_count, MeshRecord());
    ctx->built_order.clear();
    long long n_verts = 0, n_tris = 0;
    for (int i = 0; i < instance_count; ++i) {  // CreateAccelStructure: each unique mesh once, first appearance order
        const cavept_instance& in = instances[i];
        if (in.mesh < 0 || in.mesh >= mesh_count || in.material < 0 || in.material >= material_count) {
            return fail(ctx, CAVEPT_ERR_INVALID_ARG, "instance %d references mesh %d / material %d out of range", i, in.mesh, in.material);
        }
        MeshRecord& rec = ctx->meshes[in.mesh];
        if (rec.built >= 0) continue;
        const cavept_mesh& m = meshes[in.mesh];
        if (!m.positions || !m.normals || !m.indices || m.vertex_count <= 0 || m.index_count <= 0 || m.index_count % 3 != 0) {
            return fail(ctx, CAVEPT_ERR_INVALID_ARG, "mesh %d: bad arrays (index_count %% 3 != 0 is a DEV_ASSERT in the reference, bvh_accel.cpp:269)", in.mesh);
        }
        for (int k = 0; k < m.index_count; ++k) {
            if (m.indices[k] >= (uint32_t)m.vertex_count) return fail(ctx, CAVEPT_ERR_INVALID_ARG, "mesh %d: index %d out of range", in.mesh, k);
        }
        rec.built = (int)ctx->built_order.size();
        rec.vertex_count = m.vertex_count;
        rec.tri_count = m.index_count / 3;
        rec.vert_off = (int)n_verts;
        rec.tri_off = (int)n_tris;
        rec.node_off = (int)(2 * n_tris - rec.built);  // sum of (2 n_i - 1) over the meshes before
        rec.first_material = in.material;
        ctx->built_order.push_back(in.mesh);
        n_verts += m.vertex_count;
        n_tris += rec.tri_count;
        if (n_tris > (1ll << 28) || n_verts > (1ll << 29)) return fail(ctx, CAVEPT_ERR_CAPACITY, "scene too large");
    }
    const int n_built = (int)ctx->built_order.size();
    ctx->n_verts = (int)n_verts;
    ctx->n_tris = (int)n_tris;
    ctx->n_nodes = (int)(2 * n_tris - n_built);
    ctx->instances.assign(instances, instances + instance_count);
    ctx->materials.assign(materials, materials + material_count);

    CU(ctx, ctx->d_raw_pos.reserve((size_t)n_verts * 3));
    CU(ctx, ctx->d_raw_nrm.reserve((size_t)n_verts * 3));
    CU(ctx, ctx->d_raw_idx.reserve((size_t)n_tris * 3));
    CU(ctx, ctx->d_vertices.reserve(n_verts));
    CU(ctx, ctx->d_indices.reserve(n_tris));
    for (int k = 0; k < n_built; ++k) {
        const MeshRecord& rec = ctx->meshes[ctx->built_order[k]];
        const cavept_mesh& m = meshes[ctx->built_order[k]];
        ctx->ex.upload(ctx->d_raw_pos.ptr + (size_t)rec.vert_off * 3, m.positions, (size_t)m.vertex_count * 3 * sizeof(float));
        ctx->ex.upload(ctx->d_raw_nrm.ptr + (size_t)rec.vert_off * 3, m.normals, (size_t)m.vertex_count * 3 * sizeof(float));
        ctx->ex.upload(ctx->d_raw_idx.ptr + (size_t)rec.tri_off * 3, m.indices, (size_t)m.index_count * sizeof(uint32_t));
    }
    CU(ctx, cudaStreamSynchronize(ctx->ex.stream));  // caller's buffers are free to go after this point
    if (int e = check_exec(ctx, "scene upload")) return e;
    if (int e = build_all(ctx)) return e;
    if (int e = rebuild_instance_arrays(ctx)) return e;
    if (int e = build_world(ctx)) return e;
    ctx->scene_resident = true;
    ctx->dirty = true;
    ctx->stats.triangle_count = ctx->n_tris;
    ctx->stats.node_count = ctx->n_nodes;
    ctx->stats.instance_count = instance_count;
    return CAVEPT_OK;
}

CAVEPT_API int cavept_update_instances(cavept_ctx* ctx, const cavept_instance* instances, int instance_count) {
    CAVEPT_RANGE();
    if (!ctx || !instances) return CAVEPT_ERR_INVALID_ARG;
    if (!ctx->scene_resident) return fail(ctx, CAVEPT_ERR_NOT_READY, "no scene resident");
    if (instance_count != (int)ctx->instances.size()) return fail(ctx, CAVEPT_ERR_INVALID_ARG, "instance count changed (%d != %zu): re-upload the scene", instance_count, ctx->instances.size());
    for (int i = 0; i < instance_count; ++i) {
        if (instances[i].mesh != ctx->instances[i].mesh) return fail(ctx, CAVEPT_ERR_INVALID_ARG, "instance %d changed mesh: re-upload the scene", i);
    }
    if (int e = select_device(ctx)) return e;
    std::vector<unsigned char> moved(instance_count, 0);
    int n_moved = 0;
    for (int i = 0; i < instance_count; ++i) {
        if (std::memcmp(ctx->instances[i].transform, instances[i].transform, sizeof(float) * 16) != 0) {
            moved[i] = 1;
            ++n_moved;
            std::memcpy(ctx->instances[i].transform, instances[i].transform, sizeof(float) * 16);
        }
    }
    ctx->dirty = true;  // SCENE_DIRTY_WORLD: the reference re-uploads M x 144 bytes and restarts the accumulation (path_tracer.cpp:162-199)
    ctx->stats.last_update_instances_ms = 0.0f;
    ctx->stats.last_update_was_refit = 1;
    if (n_moved == 0) return CAVEPT_OK;
    CU(ctx, cudaEventRecord(ctx->ev2, ctx->ex.stream));
    if (int e = rebuild_instance_arrays(ctx)) return e;
    int kept = ctx->refit_enabled ? refit_world(ctx, moved) : 0;
    if (kept < 0) return -kept;
    if (kept == 0) {
        if (int e = build_world(ctx)) return e;
    }
    CU(ctx, cudaEventRecord(ctx->ev3, ctx->ex.stream));
    CU(ctx, cudaEventSynchronize(ctx->ev3));
    float ms = 0.0f;
    CU(ctx, cudaEventElapsedTime(&ms, ctx->ev2, ctx->ev3));
    ctx->stats.last_update_instances_ms = ms;
    ctx->stats.last_update_was_refit = kept;
    return CAVEPT_OK;
}

CAVEPT_API int cavept_update_materials(cavept_ctx* ctx, const cavept_material* materials, int material_count) {
    CAVEPT_RANGE();
    if (!ctx || !materials) return CAVEPT_ERR_INVALID_ARG;
    if (!ctx->scene_resident) return fail(ctx, CAVEPT_ERR_NOT_READY, "no scene resident");
    if (material_count != (int)ctx->materials.size()) return fail(ctx, CAVEPT_ERR_INVALID_ARG, "material count changed");
    if (int e = select_device(ctx)) return e;
    ctx->materials.assign(materials, materials + material_count);
    if (int e = rebuild_instance_arrays(ctx)) return e;
    ctx->dirty = true;
    return CAVEPT_OK;
}

CAVEPT_API int cavept_rebuild_mesh(cavept_ctx* ctx, int mesh, const float* positions, const float* normals) {
    CAVEPT_RANGE();
    if (!ctx || !positions) return CAVEPT_ERR_INVALID_ARG;
    if (!ctx->scene_resident) return fail(ctx, CAVEPT_ERR_NOT_READY, "no scene resident");
    if (mesh < 0 || mesh >= (int)ctx->meshes.size() || ctx->meshes[mesh].built < 0) return fail(ctx, CAVEPT_ERR_INVALID_ARG, "mesh %d is not resident", mesh);
    if (int e = select_device(ctx)) return e;
    const MeshRecord& rec = ctx->meshes[mesh];
    ctx->ex.upload(ctx->d_raw_pos.ptr + (size_t)rec.vert_off * 3, positions, (size_t)rec.vertex_count * 3 * sizeof(float));
    if (normals) ctx->ex.upload(ctx->d_raw_nrm.ptr + (size_t)rec.vert_off * 3, normals, (size_t)rec.vertex_count * 3 * sizeof(float));
    CU(ctx, cudaStreamSynchronize(ctx->ex.stream));
    ctx->scene_resident = false;
    if (int e = build_all(ctx)) return e;  // all meshes share one level-synchronous pass
    if (int e = rebuild_instance_arrays(ctx)) return e;
    if (int e = build_world(ctx)) return e;
    ctx->scene_resident = true;
    ctx->dirty = true;
    return CAVEPT_OK;
}

CAVEPT_API int cavept_get_mesh_bvh(cavept_ctx* ctx, int mesh, void* out, int capacity_nodes) {
    if (!ctx || !out) return -CAVEPT_ERR_INVALID_ARG;
    if (!ctx->scene_resident || mesh < 0 || mesh >= (int)ctx->meshes.size() || ctx->meshes[mesh].built < 0) return -fail(ctx, CAVEPT_ERR_NOT_READY, "mesh %d is not resident", mesh);
    if (!ctx->nodes_resident) return -fail(ctx, CAVEPT_ERR_NOT_READY, "the reference-layout BVH is only built in CAVEPT_BVH_PARITY mode");
    if (select_device(ctx)) return -CAVEPT_ERR_CUDA;
    const MeshRecord& rec = ctx->meshes[mesh];
    const int n = 2 * rec.tri_count - 1;
    if (capacity_nodes < n) return -fail(ctx, CAVEPT_ERR_CAPACITY, "need %d nodes", n);
    GpuPtBvh* o = static_cast<GpuPtBvh*>(out);
    if (cudaMemcpyAsync(o, ctx->d_nodes.ptr + rec.node_off, (size_t)n * sizeof(GpuPtBvh), cudaMemcpyDeviceToHost, ctx->ex.stream) != cudaSuccess ||
        cudaStreamSynchronize(ctx->ex.stream) != cudaSuccess) {
        return -fail(ctx, CAVEPT_ERR_CUDA, "bvh read-back failed");
    }
    for (int i = 0; i < n; ++i) {  // undo AppendBvhs (path_tracer.cpp:90-118): back to what FillGpuBvhAccel emits
        if (o[i].hitIdx >= 0) o[i].hitIdx -= rec.node_off;
        if (o[i].missIdx >= 0) o[i].missIdx -= rec.node_off;
        if (o[i].triangleIndex >= 0) o[i].triangleIndex -= rec.tri_off;
    }
    return n;
}

CAVEPT_API int cavept_get_scene_counts(cavept_ctx* ctx, int* vertices, int* triangles, int* nodes, int* instances, int* materials) {
    if (!ctx) return CAVEPT_ERR_INVALID_ARG;
    if (!ctx->scene_resident) return fail(ctx, CAVEPT_ERR_NOT_READY, "no scene resident");
    if (vertices) *vertices = ctx->n_verts;
    if (triangles) *triangles = ctx->n_tris;
    if (nodes) *nodes = ctx->n_nodes;
    if (instances) *instances = (int)ctx->gpu_meshes.size();
    if (materials) *materials = (int)ctx->gpu_materials.size();
    return CAVEPT_OK;
}

CAVEPT_API int cavept_get_scene_arrays(cavept_ctx* ctx, void* v, void* i, void* b, void* m, void* mat) {
    if (!ctx) return CAVEPT_ERR_INVALID_ARG;
    if (!ctx->scene_resident) return fail(ctx, CAVEPT_ERR_NOT_READY, "no scene resident");
    if (int e = select_device(ctx)) return e;
    if (v) ctx->ex.download(v, ctx->d_vertices.ptr, (size_t)ctx->n_verts * sizeof(GpuPtVertex));
    if (i) ctx->ex.download(i, ctx->d_indices.ptr, (size_t)ctx->n_tris * sizeof(GpuPtIndex));
    if (b) {
        if (!ctx->nodes_resident) return fail(ctx, CAVEPT_ERR_NOT_READY, "the reference-layout BVH is only built in CAVEPT_BVH_PARITY mode");
        ctx->ex.download(b, ctx->d_nodes.ptr, (size_t)ctx->n_nodes * sizeof(GpuPtBvh));
    }
    if (m) ctx->ex.download(m, ctx->d_meshes.ptr, ctx->gpu_meshes.size() * sizeof(GpuPtMesh));
    if (mat) ctx->ex.download(mat, ctx->d_materials.ptr, ctx->gpu_materials.size() * sizeof(GpuPtMaterial));
    return check_exec(ctx, "scene read-back");
}

// ---- textured materials, alpha cut-out, environment map (the reference's `#if 0` RayColor, glsl/path_tracer.cs.glsl:62-112) ----
CAVEPT_API int cavept_set_mesh_uvs(cavept_ctx* ctx, int mesh, const float* uvs) {
    CAVEPT_RANGE();
    if (!ctx || !uvs) return CAVEPT_ERR_INVALID_ARG;
    if (!ctx->scene_resident) return fail(ctx, CAVEPT_ERR_NOT_READY, "no scene resident");
    if (mesh < 0 || mesh >= (int)ctx->meshes.size()) return fail(ctx, CAVEPT_ERR_INVALID_ARG, "mesh %d out of range", mesh);
    const MeshRecord& rec = ctx->meshes[mesh];
    if (rec.built < 0) return CAVEPT_OK;  // no instance uses it: nothing to texture
    if (int e = select_device(ctx)) return e;
    if (!ctx->have_uvs) {
        CU(ctx, ctx->d_uvs.reserve((size_t)ctx->n_verts * 2));
        ctx->ex.fill(ctx->d_uvs.ptr, 0, (size_t)ctx->n_verts * 2 * sizeof(float));
        ctx->have_uvs = true;
    }
    ctx->ex.upload(ctx->d_uvs.ptr + (size_t)rec.vert_off * 2, uvs, (size_t)rec.vertex_count * 2 * sizeof(float));
    CU(ctx, cudaStreamSynchronize(ctx->ex.stream));  // the caller's array is free to go
    ctx->uvs_dirty = true;
    ctx->dirty = true;
    return check_exec(ctx, "mesh uvs");
}

CAVEPT_API int cavept_upload_textures(cavept_ctx* ctx, const cavept_texture* textures, int texture_count) {
    CAVEPT_RANGE();
    if (!ctx || texture_count < 0 || (texture_count > 0 && !textures)) return CAVEPT_ERR_INVALID_ARG;
    if (int e = select_device(ctx)) return e;
    ctx->dirty = true;
    ctx->n_textures = 0;
    ctx->have_maps = false;
    if (texture_count == 0) return CAVEPT_OK;
    std::vector<TexInfo> table(texture_count);
    size_t total = 0;
    for (int t = 0; t < texture_count; ++t) {
        const cavept_texture& x = textures[t];
        if (!x.pixels || x.width <= 0 || x.height <= 0 || (x.format != CAVEPT_TEX_RGBA8 && x.format != CAVEPT_TEX_RGBA32F))
            return fail(ctx, CAVEPT_ERR_INVALID_ARG, "texture %d: bad size, format or pointer", t);
        if (total + (size_t)x.width * x.height > (size_t)0x7fffffff) return fail(ctx, CAVEPT_ERR_CAPACITY, "more than 2^31 texels");
        table[t].offset = (int)total;
        table[t].width = x.width;
        table[t].height = x.height;
        table[t].pad = 0;
        total += (size_t)x.width * x.height;
    }
    std::vector<F4> texels(total);
    for (int t = 0; t < texture_count; ++t) {
        const cavept_texture& x = textures[t];
        F4* dst = texels.data() + table[t].offset;
        const size_t n = (size_t)x.width * x.height;
        if (x.format == CAVEPT_TEX_RGBA32F) {
            std::memcpy(dst, x.pixels, n * sizeof(F4));
        } else {
            const unsigned char* src = static_cast<const unsigned char*>(x.pixels);
            for (size_t i = 0; i < n; ++i) {
                dst[i].x = (float)src[4 * i] / 255.0f;
                dst[i].y = (float)src[4 * i + 1] / 255.0f;
                dst[i].z = (float)src[4 * i + 2] / 255.0f;
                dst[i].w = (float)src[4 * i + 3] / 255.0f;
            }
        }
    }
    CU(ctx, ctx->d_tex_data.reserve(total));
    CU(ctx, ctx->d_tex_table.reserve(texture_count));
    ctx->ex.upload(ctx->d_tex_data.ptr, texels.data(), total * sizeof(F4));
    ctx->ex.upload(ctx->d_tex_table.ptr, table.data(), table.size() * sizeof(TexInfo));
    CU(ctx, cudaStreamSynchronize(ctx->ex.stream));
    ctx->n_textures = texture_count;
    return check_exec(ctx, "texture upload");
}

CAVEPT_API int cavept_set_material_maps(cavept_ctx* ctx, const int* base_color_maps, const int* material_maps, int material_count) {
    CAVEPT_RANGE();
    if (!ctx || !base_color_maps || !material_maps) return CAVEPT_ERR_INVALID_ARG;
    if (!ctx->scene_resident) return fail(ctx, CAVEPT_ERR_NOT_READY, "no scene resident");
    if (material_count != (int)ctx->materials.size()) return fail(ctx, CAVEPT_ERR_INVALID_ARG, "material count %d != %zu", material_count, ctx->materials.size());
    std::vector<int> maps(2 * (size_t)material_count);
    bool any = false;
    for (int m = 0; m < material_count; ++m) {
        const int a = base_color_maps[m], b = material_maps[m];
        if (a >= ctx->n_textures || b >= ctx->n_textures) return fail(ctx, CAVEPT_ERR_INVALID_ARG, "material %d names a texture outside the set of %d", m, ctx->n_textures);
        maps[2 * m] = a < 0 ? -1 : a;
        maps[2 * m + 1] = b < 0 ? -1 : b;
        any = any || a >= 0 || b >= 0;
    }
    if (int e = select_device(ctx)) return e;
    ctx->api_mat_maps = maps;
    ctx->have_maps = any;
    ctx->dirty = true;
    return upload_material_maps(ctx);
}

CAVEPT_API int cavept_set_env_map(cavept_ctx* ctx, const float* rgb, int width, int height) {
    CAVEPT_RANGE();
    if (!ctx) return CAVEPT_ERR_INVALID_ARG;
    ctx->dirty = true;
    if (!rgb) {
        ctx->env_w = ctx->env_h = 0;
        return CAVEPT_OK;
    }
    if (width <= 0 || height <= 0) return fail(ctx, CAVEPT_ERR_INVALID_ARG, "environment map %d x %d", width, height);
    if (int e = select_device(ctx)) return e;
    const size_t n = (size_t)width * height;
    std::vector<F4> texels(n);
    for (size_t i = 0; i < n; ++i) {
        texels[i].x = rgb[3 * i];
        texels[i].y = rgb[3 * i + 1];
        texels[i].z = rgb[3 * i + 2];
        texels[i].w = 1.0f;
    }
    CU(ctx, ctx->d_env.reserve(n));
    ctx->ex.upload(ctx->d_env.ptr, texels.data(), n * sizeof(F4));
    CU(ctx, cudaStreamSynchronize(ctx->ex.stream));
    ctx->env_w = width;
    ctx->env_h = height;
    return check_exec(ctx, "environment map");
}

CAVEPT_API int cavept_get_fast_tree(cavept_ctx* ctx, void* out_records, int capacity_records) {
    if (!ctx || !out_records) return -CAVEPT_ERR_INVALID_ARG;
    if (!ctx->scene_resident) return -fail(ctx, CAVEPT_ERR_NOT_READY, "no scene resident");
    const int n = 2 * ctx->n_pairs;
    if (capacity_records < n) return -fail(ctx, CAVEPT_ERR_INVALID_ARG, "capacity %d < %d records", capacity_records, n);
    if (int e = select_device(ctx)) return -e;
    ctx->ex.download(out_records, ctx->d_fnodes.ptr, (size_t)n * sizeof(FastNode));
    if (int e = check_exec(ctx, "fast tree read-back")) return -e;
    return n;
}

CAVEPT_API int cavept_set_camera(cavept_ctx* ctx, const float position[3], const float right[3], const float up[3], const float forward[3], float fov_degree) {
    if (!ctx || !position || !right || !up || !forward) return CAVEPT_ERR_INVALID_ARG;
    std::memcpy(ctx->cam_pos, position, 12);
    std::memcpy(ctx->cam_right, right, 12);
    std::memcpy(ctx->cam_up, up, 12);
    std::memcpy(ctx->cam_fwd, forward, 12);
    ctx->cam_fov = fov_degree;
    ctx->dirty = true;  // SCENE_DIRTY_CAMERA
    return CAVEPT_OK;
}

CAVEPT_API int cavept_reset(cavept_ctx* ctx) {
    if (!ctx) return CAVEPT_ERR_INVALID_ARG;
    ctx->dirty = true;
    return CAVEPT_OK;
}

// One launch over the pixel rectangle [x0, x0+w) x [y0, y0+h); w == 0 means the whole image.
static int render_rect(cavept_ctx* ctx, int x0, int y0, int w, int h, int first_frame_index, int n_frames) {
    if (!ctx) return CAVEPT_ERR_INVALID_ARG;
    if (n_frames <= 0 || n_frames > (1 << 24)) return fail(ctx, CAVEPT_ERR_INVALID_ARG, "n_frames must be in [1, 2^24]");
    const bool tiled = w > 0;
    const bool active = cavept_is_active(ctx) || (tiled && ctx->mode == CAVEPT_MODE_TILED && ctx->scene_resident);
    if (!active) return fail(ctx, CAVEPT_ERR_NOT_READY, "path tracer not active (mode INTERACTIVE and a resident scene are required)");
    if (ctx->cfg.trace_mode == CAVEPT_TRACE_REFERENCE_WALK && !ctx->nodes_resident) return fail(ctx, CAVEPT_ERR_NOT_READY, "CAVEPT_TRACE_REFERENCE_WALK needs the reference-layout BVH (CAVEPT_BVH_PARITY)");
    if (int e = select_device(ctx)) return e;
    TraceParams p{};
    p.vertices = ctx->d_vertices.ptr;
    p.indices = ctx->d_indices.ptr;
    p.bvhs = ctx->d_nodes.ptr;
    p.meshes = ctx->d_meshes.ptr;
    p.materials = ctx->d_materials.ptr;
    p.mesh_count = (int)ctx->gpu_meshes.size();
    p.fnodes = ctx->d_fnodes.ptr;
    p.qpairs = ctx->d_qpairs.ptr;
    p.q4nodes = ctx->q4_ready ? ctx->d_q4nodes.ptr : nullptr;
    for (int k = 0; k < 3; ++k) {
        p.grid_origin[k] = ctx->grid_origin[k];
        p.grid_scale[k] = ctx->grid_scale[k];
    }
    p.ftris = ctx->d_ftris.ptr;
    p.fnormals = ctx->d_fnormals.ptr;
    p.root_ref = ctx->root_ref;
    std::memcpy(p.cam_pos, ctx->cam_pos, 12);
    std::memcpy(p.cam_right, ctx->cam_right, 12);
    std::memcpy(p.cam_up, ctx->cam_up, 12);
    std::memcpy(p.cam_fwd, ctx->cam_fwd, 12);
    p.cam_distance = std::tan(ctx->cam_fov * 3.141592653589793f / 180.0f);  // path_tracer.cs.glsl:31-32 (the full angle, as the shader has it)
    p.width = ctx->cfg.width;
    p.height = ctx->cfg.height;
    if (tiled) {
        p.tile_x0 = x0, p.tile_y0 = y0, p.tile_x1 = x0 + w, p.tile_y1 = y0 + h;
        if (ctx->dirty) {  // c_sceneDirty with tiles: an emptied accumulator (weight 0) makes every tile's first sample an overwrite
            CU(ctx, cudaMemsetAsync(ctx->d_accum.ptr, 0, (size_t)ctx->cfg.width * ctx->cfg.height * sizeof(F4), ctx->ex.stream));
            ctx->dirty = false;
        }
    }
    p.max_bounce = ctx->cfg.max_bounce;
    p.first_frame = first_frame_index;
    p.n_frames = n_frames;
    p.reset_first = ctx->dirty ? 1 : 0;
    p.accum_sums = ctx->accum_sums;
    p.accum = ctx->d_accum.ptr;
    p.counters = ctx->d_stats.ptr;
    const bool fast = ctx->cfg.trace_mode == CAVEPT_TRACE_FAST;
    // textured materials / environment map: the launch takes the per-pixel kernel's textured instantiation
    const bool use_maps = ctx->have_maps && ctx->have_uvs && ctx->n_textures > 0;
    const bool textured = fast && (use_maps || ctx->env_w > 0);
    if (textured) {
        if (use_maps && (ctx->uvs_dirty || ctx->d_fuvs.cap < (size_t)ctx->n_prims)) {
            CU(ctx, ctx->d_fuvs.reserve(ctx->n_prims));
            launch_emit_uvs(ctx->ex, ctx->d_ftris.ptr, ctx->d_indices.ptr, ctx->d_meshes.ptr, ctx->d_uvs.ptr, ctx->d_fuvs.ptr, ctx->n_prims);
            ctx->uvs_dirty = false;
        }
        p.fuvs = use_maps ? ctx->d_fuvs.ptr : nullptr;
        p.tex_table = ctx->d_tex_table.ptr;
        p.tex_data = ctx->d_tex_data.ptr;
        p.mat_maps = use_maps ? ctx->d_mat_maps.ptr : nullptr;
        p.env = ctx->env_w > 0 ? ctx->d_env.ptr : nullptr;
        p.env_w = ctx->env_w;
        p.env_h = ctx->env_h;
    }
    int stack_entries = 0, grid_blocks = 0;
    if (fast && ctx->kernel_variant == CAVEPT_KERNEL_POOL) {
        stack_entries = ctx->fast_levels + 2 < 4 ? 4 : ctx->fast_levels + 2;
        if (ctx->pool_blocks_depth != stack_entries) {
            ctx->pool_blocks_per_sm = trace_pool_blocks_per_sm(stack_entries, false);
            ctx->pool_blocks_depth = stack_entries;
        }
        if (ctx->pool_blocks_per_sm <= 0) return fail(ctx, CAVEPT_ERR_CAPACITY, "k_trace_pool does not fit an SM at stack depth %d", stack_entries);
        grid_blocks = ctx->n_sms * ctx->pool_blocks_per_sm;
    }
    CU(ctx, cudaEventRecord(ctx->ev0, ctx->ex.stream));
    if (textured) launch_trace(ctx->ex, p, true, ctx->count_visits, true);
    else if (grid_blocks > 0) launch_trace_pool(ctx->ex, p, ctx->count_visits, stack_entries, ctx->d_job_counter.ptr, grid_blocks);
    else if (fast && ctx->kernel_variant == CAVEPT_KERNEL_QUEUE) launch_trace_queue(ctx->ex, p, ctx->count_visits, ctx->d_job_counter.ptr, ctx->n_sms);
    else if (fast && ctx->kernel_variant == CAVEPT_KERNEL_WIDE4 && ctx->q4_ready && ctx->fast_levels <= kWideMaxLevels && ctx->n_pairs > 0) launch_trace_wide(ctx->ex, p, ctx->count_visits);
    else {
        // triangle + normal records beyond what any L1 keeps, and a launch of several samples (mostly incoherent bounce
        // segments): the instantiation with the L1 hints on its gathers (trace_body.h walk_test_triangle)
        const bool stream = fast && (size_t)ctx->n_prims * 2 * sizeof(FastTri) > ((size_t)4 << 20) && n_frames >= 4;
        launch_trace(ctx->ex, p, fast, ctx->count_visits, false, stream);
    }
    CU(ctx, cudaEventRecord(ctx->ev1, ctx->ex.stream));
    ctx->dirty = false;
    return check_exec(ctx, "render launch");
}

CAVEPT_API int cavept_render(cavept_ctx* ctx, int first_frame_index, int n_frames) {
    CAVEPT_RANGE();
    return render_rect(ctx, 0, 0, 0, 0, first_frame_index, n_frames);
}

CAVEPT_API int cavept_render_tile(cavept_ctx* ctx, int x0, int y0, int w, int h, int first_frame_index, int n_frames) {
    CAVEPT_RANGE();
    if (!ctx) return CAVEPT_ERR_INVALID_ARG;
    if (x0 < 0 || y0 < 0 || w <= 0 || h <= 0 || x0 > ctx->cfg.width - w || y0 > ctx->cfg.height - h) {
        return fail(ctx, CAVEPT_ERR_INVALID_ARG, "tile (%d,%d)+(%d,%d) is not inside the %dx%d image", x0, y0, w, h, ctx->cfg.width, ctx->cfg.height);
    }
    return render_rect(ctx, x0, y0, w, h, first_frame_index, n_frames);
}

CAVEPT_API int cavept_sync(cavept_ctx* ctx) {
    CAVEPT_RANGE();
    if (!ctx) return CAVEPT_ERR_INVALID_ARG;
    if (int e = select_device(ctx)) return e;
    if (ctx->group) return cavept_group_sync(ctx);  // polls the communicator: a failed peer is an error code, not a hang
    CU(ctx, cudaStreamSynchronize(ctx->ex.stream));
    return CAVEPT_OK;
}

CAVEPT_API int cavept_read_accum(cavept_ctx* ctx, float* out, size_t bytes) {
    CAVEPT_RANGE();
    if (!ctx || !out) return CAVEPT_ERR_INVALID_ARG;
    const size_t need = (size_t)ctx->cfg.width * ctx->cfg.height * sizeof(F4);
    if (bytes < need) return fail(ctx, CAVEPT_ERR_CAPACITY, "need %zu bytes", need);
    if (int e = select_device(ctx)) return e;
    ctx->ex.download(out, ctx->d_accum.ptr, need);
    return check_exec(ctx, "accumulator read-back");
}

CAVEPT_API int cavept_write_accum(cavept_ctx* ctx, const float* in, size_t bytes) {
    if (!ctx || !in) return CAVEPT_ERR_INVALID_ARG;
    const size_t need = (size_t)ctx->cfg.width * ctx->cfg.height * sizeof(F4);
    if (bytes < need) return fail(ctx, CAVEPT_ERR_CAPACITY, "need %zu bytes", need);
    if (int e = select_device(ctx)) return e;
    ctx->ex.upload(ctx->d_accum.ptr, in, need);
    CU(ctx, cudaStreamSynchronize(ctx->ex.stream));
    ctx->dirty = false;
    return check_exec(ctx, "accumulator upload");
}

CAVEPT_API int cavept_read_tonemapped_rgba8(cavept_ctx* ctx, uint8_t* out, size_t bytes) {
    CAVEPT_RANGE();
    if (!ctx || !out) return CAVEPT_ERR_INVALID_ARG;
    const size_t n = (size_t)ctx->cfg.width * ctx->cfg.height;
    if (bytes < n * 4) return fail(ctx, CAVEPT_ERR_CAPACITY, "need %zu bytes", n * 4);
    if (ctx->accum_sums) return fail(ctx, CAVEPT_ERR_NOT_READY, "the accumulator holds sums (group mode): leave the group first, or read cavept_read_composite");
    if (int e = select_device(ctx)) return e;
    CU(ctx, ctx->d_tone8.reserve(n * 4));
    launch_tonemap_rgba8(ctx->ex, ctx->d_accum.ptr, ctx->d_tone8.ptr, (int)n);
    ctx->ex.download(out, ctx->d_tone8.ptr, n * 4);
    return check_exec(ctx, "tone pass");
}

CAVEPT_API int cavept_read_tonemapped_rgba16f(cavept_ctx* ctx, uint16_t* out, size_t bytes) {
    if (!ctx || !out) return CAVEPT_ERR_INVALID_ARG;
    CAVEPT_RANGE();
    const size_t n = (size_t)ctx->cfg.width * ctx->cfg.height;
    if (bytes < n * 8) return fail(ctx, CAVEPT_ERR_CAPACITY, "need %zu bytes", n * 8);
    if (ctx->accum_sums) return fail(ctx, CAVEPT_ERR_NOT_READY, "the accumulator holds sums (group mode): tone-map the composite on the destination rank after cavept_group_destroy, or read cavept_read_composite");
    if (int e = select_device(ctx)) return e;
    CU(ctx, ctx->d_tone16.reserve(n * 4));
    launch_tonemap_f16(ctx->ex, ctx->d_accum.ptr, ctx->d_tone16.ptr, (int)n);
    ctx->ex.download(out, ctx->d_tone16.ptr, n * 8);
    return check_exec(ctx, "tone pass (rgba16f)");
}

// The tone-mapped frame as a binary PPM (P6, 8 bits per channel, row 0 = top as the accumulator stores it): the
// headless stand-in for presenting RG_RES_POST_PROCESS (graphics_manager.cpp:485-493).
CAVEPT_API int cavept_write_ppm(cavept_ctx* ctx, const char* path) {
    if (!ctx || !path) return CAVEPT_ERR_INVALID_ARG;
    CAVEPT_RANGE();
    const size_t n = (size_t)ctx->cfg.width * ctx->cfg.height;
    ctx->host_rgba8.resize(n * 4);
    if (int e = cavept_read_tonemapped_rgba8(ctx, ctx->host_rgba8.data(), ctx->host_rgba8.size())) return e;
    FILE* f = std::fopen(path, "wb");
    if (!f) return fail(ctx, CAVEPT_ERR_INVALID_ARG, "cannot open %s for writing", path);
    std::fprintf(f, "P6\n%d %d\n255\n", ctx->cfg.width, ctx->cfg.height);
    std::vector<unsigned char> row((size_t)ctx->cfg.width * 3);
    bool ok = true;
    for (int y = 0; y < ctx->cfg.height && ok; ++y) {
        const unsigned char* src = ctx->host_rgba8.data() + (size_t)y * ctx->cfg.width * 4;
        for (int x = 0; x < ctx->cfg.width; ++x) {
            row[3 * x] = src[4 * x];
            row[3 * x + 1] = src[4 * x + 1];
            row[3 * x + 2] = src[4 * x + 2];
        }
        ok = std::fwrite(row.data(), 1, row.size(), f) == row.size();
    }
    ok = (std::fclose(f) == 0) && ok;
    return ok ? CAVEPT_OK : fail(ctx, CAVEPT_ERR_INVALID_ARG, "short write to %s", path);
}

// Device time of the last cavept_render without blocking: CAVEPT_ERR_NOT_READY while the launch is still running.
CAVEPT_API int cavept_last_render_ms(cavept_ctx* ctx, float* out_ms) {
    if (!ctx || !out_ms) return CAVEPT_ERR_INVALID_ARG;
    if (int e = select_device(ctx)) return e;
    const cudaError_t q = cudaEventQuery(ctx->ev1);
    if (q == cudaErrorNotReady) return CAVEPT_ERR_NOT_READY;
    if (q != cudaSuccess) {
        (void)cudaGetLastError();
        return fail(ctx, CAVEPT_ERR_NOT_READY, "no render has been launched yet");
    }
    float ms = 0.0f;
    if (cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1) != cudaSuccess) {
        (void)cudaGetLastError();
        return fail(ctx, CAVEPT_ERR_NOT_READY, "no render has been launched yet");
    }
    *out_ms = ms;
    return CAVEPT_OK;
}

CAVEPT_API int cavept_read_tonemapped_rgba32f(cavept_ctx* ctx, float* out, size_t bytes) {
    if (!ctx || !out) return CAVEPT_ERR_INVALID_ARG;
    const size_t n = (size_t)ctx->cfg.width * ctx->cfg.height;
    if (bytes < n * sizeof(F4)) return fail(ctx, CAVEPT_ERR_CAPACITY, "need %zu bytes", n * sizeof(F4));
    if (int e = select_device(ctx)) return e;
    CU(ctx, ctx->d_tone.reserve(n));
    launch_tonemap_f32(ctx->ex, ctx->d_accum.ptr, ctx->d_tone.ptr, (int)n);
    ctx->ex.download(out, ctx->d_tone.ptr, n * sizeof(F4));
    return check_exec(ctx, "tone pass");
}

CAVEPT_API int cavept_get_stats(cavept_ctx* ctx, cavept_stats* out) {
    if (!ctx || !out) return CAVEPT_ERR_INVALID_ARG;
    if (int e = select_device(ctx)) return e;
    unsigned long long c[4] = { 0, 0, 0, 0 };
    ctx->ex.download(c, ctx->d_stats.ptr, sizeof(c));
    if (int e = check_exec(ctx, "stats read-back")) return e;
    ctx->stats.segments = c[0];
    ctx->stats.node_visits = c[1];
    ctx->stats.triangle_tests = c[2];
    ctx->stats.kernel_launches = ctx->ex.launches;
    float ms = 0.0f;
    if (cudaEventQuery(ctx->ev1) == cudaSuccess && cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1) == cudaSuccess) ctx->stats.last_render_ms = ms;
    *out = ctx->stats;
    return CAVEPT_OK;
}

CAVEPT_API int cavept_reset_stats(cavept_ctx* ctx) {
    if (!ctx) return CAVEPT_ERR_INVALID_ARG;
    if (int e = select_device(ctx)) return e;
    ctx->ex.fill(ctx->d_stats.ptr, 0, 4 * sizeof(unsigned long long));
    return check_exec(ctx, "stats reset");
}

CAVEPT_API int cavept_set_count_visits(cavept_ctx* ctx, int enable) {
    if (!ctx) return CAVEPT_ERR_INVALID_ARG;
    ctx->count_visits = enable != 0;
    return CAVEPT_OK;
}

CAVEPT_API int cavept_set_refit(cavept_ctx* ctx, int enable) {
    if (!ctx) return CAVEPT_ERR_INVALID_ARG;
    ctx->refit_enabled = enable != 0;
    return CAVEPT_OK;
}

CAVEPT_API int cavept_set_kernel_variant(cavept_ctx* ctx, int variant) {
    if (!ctx) return CAVEPT_ERR_INVALID_ARG;
    if (variant != CAVEPT_KERNEL_POOL && variant != CAVEPT_KERNEL_PIXEL && variant != CAVEPT_KERNEL_QUEUE && variant != CAVEPT_KERNEL_WIDE4) return fail(ctx, CAVEPT_ERR_INVALID_ARG, "invalid kernel variant %d", variant);
    ctx->kernel_variant = variant;
    if (variant == CAVEPT_KERNEL_WIDE4 && ctx->scene_resident && !ctx->q4_ready && ctx->n_pairs > 0) {
        // build the four-wide records over the resident tree (wbvh_body.h wb_body_q4)
        if (int e = select_device(ctx)) return e;
        CU(ctx, ctx->d_q4nodes.reserve((size_t)ctx->n_prims + 8));
        WBuild b{};
        b.fnodes = ctx->d_fnodes.ptr;
        b.q4nodes = ctx->d_q4nodes.ptr;
        for (int k = 0; k < 3; ++k) {
            b.grid_origin[k] = ctx->grid_origin[k];
            b.grid_scale[k] = ctx->grid_scale[k];
        }
        run_build_q4(ctx->ex, b, ctx->n_pairs);
        if (int e = check_exec(ctx, "four-wide records")) return e;
        ctx->q4_ready = true;
    }
    return CAVEPT_OK;
}

CAVEPT_API void* cavept_stream(cavept_ctx* ctx) { return ctx ? (void*)ctx->ex.stream : nullptr; }
CAVEPT_API void* cavept_accum_device_ptr(cavept_ctx* ctx) { return ctx ? (void*)ctx->d_accum.ptr : nullptr; }

CAVEPT_API int cavept_accum_to_sums(cavept_ctx* ctx) {
    if (!ctx) return CAVEPT_ERR_INVALID_ARG;
    if (int e = select_device(ctx)) return e;
    launch_accum_scale(ctx->ex, ctx->d_accum.ptr, ctx->cfg.width * ctx->cfg.height, true);
    return check_exec(ctx, "accum_to_sums");
}
CAVEPT_API int cavept_accum_from_sums(cavept_ctx* ctx) {
    if (!ctx) return CAVEPT_ERR_INVALID_ARG;
    if (int e = select_device(ctx)) return e;
    launch_accum_scale(ctx->ex, ctx->d_accum.ptr, ctx->cfg.width * ctx->cfg.height, false);
    return check_exec(ctx, "accum_from_sums");
}

}  // extern "C"

CAVEPT_API void* cavept_host_alloc(size_t bytes) {
    void* p = nullptr;
    if (bytes == 0 || cudaHostAlloc(&p, bytes, cudaHostAllocPortable) != cudaSuccess) {
        (void)cudaGetLastError();
        return nullptr;
    }
    return p;
}

CAVEPT_API void cavept_host_free(void* p) {
    if (p) (void)cudaFreeHost(p);
}

CAVEPT_API int cavept_probe_intersections(cavept_ctx* ctx, int count, const float* rays, const float* triangles, const float* boxes, float* out) {
    if (!ctx || !rays || !triangles || !boxes || !out || count < 0) return CAVEPT_ERR_INVALID_ARG;
    if (int e = select_device(ctx)) return e;
    if (count == 0) return CAVEPT_OK;
    DevBuf<float> r, t, b, o;
    CU(ctx, r.reserve((size_t)count * 7));
    CU(ctx, t.reserve((size_t)count * 9));
    CU(ctx, b.reserve((size_t)count * 6));
    CU(ctx, o.reserve((size_t)count * 8));
    ctx->ex.upload(r.ptr, rays, (size_t)count * 7 * sizeof(float));
    ctx->ex.upload(t.ptr, triangles, (size_t)count * 9 * sizeof(float));
    ctx->ex.upload(b.ptr, boxes, (size_t)count * 6 * sizeof(float));
    launch_probe(ctx->ex, r.ptr, t.ptr, b.ptr, o.ptr, count);
    CU(ctx, cudaMemcpyAsync(out, o.ptr, (size_t)count * 8 * sizeof(float), cudaMemcpyDeviceToHost, ctx->ex.stream));
    CU(ctx, cudaStreamSynchronize(ctx->ex.stream));
    return check_exec(ctx, "probe_intersections");
}

CAVEPT_API int cavept_measure_gather(cavept_ctx* ctx, size_t working_set_bytes, int iters, double* out_gathers_per_s) {
    if (!ctx || !out_gathers_per_s || working_set_bytes < (1u << 20) || working_set_bytes > ((size_t)8 << 30) || iters < 1) return CAVEPT_ERR_INVALID_ARG;
    if (int e = select_device(ctx)) return e;
    DevBuf<unsigned char> buf;
    CU(ctx, buf.reserve(working_set_bytes));
    CU(ctx, cudaMemsetAsync(buf.ptr, 0, working_set_bytes, ctx->ex.stream));
    double lane_gathers = 0.0;
    const float ms = run_gather_probe(ctx->ex, buf.ptr, working_set_bytes, iters, ctx->ev2, ctx->ev3, ctx->n_sms, &lane_gathers);
    if (int e = check_exec(ctx, "gather probe")) return e;
    *out_gathers_per_s = ms > 0.0f ? lane_gathers / (ms * 1e-3) : 0.0;
    return CAVEPT_OK;
}

CAVEPT_API int cavept_sort_pairs(cavept_ctx* ctx, uint32_t* keys, int32_t* values, int count, int key_bits) {
    if (!ctx || !keys || !values || count < 0 || key_bits < 1 || key_bits > 32) return CAVEPT_ERR_INVALID_ARG;
    if (int e = select_device(ctx)) return e;
    if (count == 0) return CAVEPT_OK;
    DevBuf<uint32_t> k, kt;
    DevBuf<int> v, vt, hist, tmp;
    CU(ctx, k.reserve(count));
    CU(ctx, kt.reserve(count));
    CU(ctx, v.reserve(count));
    CU(ctx, vt.reserve(count));
    CU(ctx, hist.reserve(2 * sort_hist_entries(count)));
    CU(ctx, tmp.reserve(sort_hist_entries(count) / kScanTile + 4));
    struct Restore {  // the context's scan scratch comes back on every return path
        CudaExec& ex;
        int* saved;
        ~Restore() { ex.scan_tmp = saved; }
    } restore{ ctx->ex, ctx->ex.scan_tmp };
    ctx->ex.scan_tmp = tmp.ptr;
    ctx->ex.upload(k.ptr, keys, (size_t)count * sizeof(uint32_t));
    ctx->ex.upload(v.ptr, values, (size_t)count * sizeof(int));
    ctx->ex.sort_pairs(k.ptr, v.ptr, kt.ptr, vt.ptr, hist.ptr, count, key_bits);
    CU(ctx, cudaMemcpyAsync(keys, k.ptr, (size_t)count * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->ex.stream));
    CU(ctx, cudaMemcpyAsync(values, v.ptr, (size_t)count * sizeof(int), cudaMemcpyDeviceToHost, ctx->ex.stream));
    CU(ctx, cudaStreamSynchronize(ctx->ex.stream));
    return check_exec(ctx, "sort_pairs");
}
