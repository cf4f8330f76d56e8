// Private to libcavept: the context behind the C ABI (include/cavept.h) and the helpers every
// translation unit of the host layer shares.  Not installed; nothing here is part of the ABI.
#pragma once
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>  // header-only: ranges cost nothing unless a profiler injects itself

#include "../../include/cavept.h"
#include "kernels.h"

using namespace cavept;

// Device buffer that only grows.  RAII: freed with its owner (a context member, or a temporary of one API call,
// so an early error return cannot leak).  reserve() allocates the larger block BEFORE releasing the old one: a
// failed grow leaves the buffer as it was.  Contents are not carried over.
template<class T>
struct DevBuf {
    T* ptr = nullptr;
    size_t cap = 0;  // elements
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    ~DevBuf() { release(); }
    cudaError_t reserve(size_t n) {
        if (n <= cap) return cudaSuccess;
        const size_t want = n + n / 8 + 16;
        T* fresh = nullptr;
        cudaError_t e = cudaMalloc(&fresh, want * sizeof(T));
        if (e != cudaSuccess) {
            (void)cudaGetLastError();
            if (ptr == nullptr) return e;
            // not enough room for both: give the old block back and try once more
            cudaFree(ptr);
            ptr = nullptr;
            cap = 0;
            e = cudaMalloc(&fresh, want * sizeof(T));
            if (e != cudaSuccess) return e;
        } else if (ptr) {
            cudaFree(ptr);
        }
        ptr = fresh;
        cap = want;
        return cudaSuccess;
    }
    void release() {
        if (ptr) cudaFree(ptr);
        ptr = nullptr;
        cap = 0;
    }
};

// One NVTX range per API call (SURVEY.md section 5 "tracing"): shows up in Nsight timelines as cavept_<call>.
struct ApiRange {
    explicit ApiRange(const char* name) { nvtxRangePushA(name); }
    ~ApiRange() { nvtxRangePop(); }
    ApiRange(const ApiRange&) = delete;
    ApiRange& operator=(const ApiRange&) = delete;
};
#define CAVEPT_RANGE() ApiRange nvtx_range_(__func__)

struct MeshRecord {
    int built = -1;  // order of first appearance among instances, or -1 if unreferenced
    int vertex_count = 0, tri_count = 0;
    int vert_off = 0, tri_off = 0, node_off = 0;
    int first_material = 0;  // material of the first instance using this mesh (path_tracer.cpp:227-231)
};

struct cavept_group_state;  // cavept_group.cu: NCCL communicator + rank of a context that joined a multi-GPU group

struct cavept_ctx {
    cavept_config cfg{};
    int device = 0;
    int mode = CAVEPT_MODE_NONE;
    bool scene_resident = false;
    bool dirty = true;
    bool count_visits = false;
    std::string last_error;
    CudaExec ex;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr, ev3 = nullptr;

    // scene bookkeeping
    std::vector<MeshRecord> meshes;
    std::vector<int> built_order;  // built index -> input mesh id
    std::vector<cavept_instance> instances;
    std::vector<cavept_material> materials;
    std::vector<GpuPtMesh> gpu_meshes;
    std::vector<GpuPtMaterial> gpu_materials;
    int n_verts = 0, n_tris = 0, n_nodes = 0, n_prims = 0, n_pairs = 0, root_ref = 0, fast_levels = 0;
    bool nodes_resident = false;  // reference-layout GpuPtBvh array built (CAVEPT_BVH_PARITY)

    // reference-layout arrays
    DevBuf<GpuPtVertex> d_vertices;
    DevBuf<GpuPtIndex> d_indices;
    DevBuf<GpuPtBvh> d_nodes;
    DevBuf<GpuPtMesh> d_meshes;
    DevBuf<GpuPtMaterial> d_materials;
    // production layout
    DevBuf<FastNode> d_fnodes;
    DevBuf<QPair> d_qpairs;
    // textured materials / environment map (cavept.h "textured materials ...")
    DevBuf<float> d_uvs;          // 2 floats per scene-global vertex (zero where a mesh has none)
    DevBuf<FastUv> d_fuvs;        // per FastTri slot, re-emitted after a build when uvs_dirty
    DevBuf<F4> d_tex_data;
    DevBuf<TexInfo> d_tex_table;
    DevBuf<int> d_mat_maps;       // 2 per GpuPtMaterial slot (the de-duplicated order the kernel indexes)
    std::vector<int> api_mat_maps;     // 2 per caller-side material, as cavept_set_material_maps received them
    std::vector<int> gpu_material_src; // caller-side material of each GpuPtMaterial slot
    DevBuf<F4> d_env;
    int env_w = 0, env_h = 0;
    int n_textures = 0;
    bool have_uvs = false, uvs_dirty = false, have_maps = false;
    DevBuf<Q4Node> d_q4nodes;
    bool q4_ready = false;  // four-wide records built over the resident tree (CAVEPT_KERNEL_WIDE4 only)
    float grid_origin[3] = { 0, 0, 0 }, grid_scale[3] = { 1, 1, 1 };
    DevBuf<FastTri> d_ftris, d_fnormals;
    DevBuf<int> d_inst_prim_off, d_inst_tri_off;
    DevBuf<F4> d_prim_bmin, d_prim_bmax;
    DevBuf<int> d_w_order, d_w_order_next, d_w_node_of, d_w_node_of_next, d_w_scan, d_w_counters;
    DevBuf<unsigned char> d_w_flags;
    DevBuf<WItem> d_w_items, d_w_items_next;
    DevBuf<WBins> d_w_bins;
    DevBuf<uint32_t> d_w_root_acc, d_w_keys, d_w_keys_tmp;
    DevBuf<int> d_w_sort_hist;
    DevBuf<F4> d_w_nbox_mn, d_w_nbox_mx;
    // upload staging
    DevBuf<float> d_raw_pos, d_raw_nrm;
    DevBuf<uint32_t> d_raw_idx;
    DevBuf<int> d_tri_off, d_vert_off;
    // build scratch
    DevBuf<F4> d_tri_bmin, d_tri_bmax, d_tri_centroid;
    DevBuf<int> d_order, d_order_next, d_node_of, d_node_of_next, d_scan, d_scan_tmp, d_counters;
    DevBuf<unsigned char> d_flags;
    DevBuf<BuildItem> d_items, d_items_next, d_sub_items;  // d_sub_items: roots queued for the parity builder's subtree phase
    DevBuf<ItemAccum> d_accum_items;
    DevBuf<ItemBins> d_bins;
    DevBuf<int> d_node_parent, d_node_count;
    // image
    DevBuf<F4> d_accum, d_tone;
    DevBuf<unsigned char> d_tone8;
    DevBuf<unsigned short> d_tone16;
    std::vector<unsigned char> host_rgba8;  // staging of cavept_write_ppm
    DevBuf<unsigned long long> d_stats;
    // refit of the production tree on instance updates (wbvh_body.h refit_wbvh): SAH builder only
    bool refit_ready = false, refit_enabled = true;
    DevBuf<WItem> d_w_sub_roots;         // SAH builder: items of <= kWSubMax prims queued for the subtree phase
    DevBuf<unsigned char> d_pair_depth;  // SAH tree: depth of the interior node owning sibling pair k (the refit's schedule)
    float world_pad = 0.0f, world_extent = 0.0f;  // box padding of the built tree and the largest |coordinate| it was derived from
    float area_baseline = -1.0f;         // surface-area sum of the freshly built tree (SAH cost proxy), computed on first use
    DevBuf<unsigned char> d_moved;
    DevBuf<float> d_area_sum;
    DevBuf<F4> d_composite;      // group mode: the composited frame (mean, n) on the destination rank(s)
    cavept_group_state* group = nullptr;
    DevBuf<int> d_job_counter;   // job queue of the persistent trace kernel (trace_pool_body.h)
    int kernel_variant = CAVEPT_KERNEL_PIXEL;  // measured fastest (profiles/r02c: pixel 3775, queue 3331, pool 2533 M segments/s at 256 spp)
    int accum_sums = 0;          // group mode: the accumulator holds (sum, n)
    int n_sms = 0;
    int pool_blocks_depth = -1, pool_blocks_per_sm = 0;  // occupancy of k_trace_pool, cached per stack depth

    float cam_pos[3] = { 0, 0, 0 }, cam_right[3] = { 1, 0, 0 }, cam_up[3] = { 0, 1, 0 }, cam_fwd[3] = { 0, 0, 1 };
    float cam_fov = 50.0f;

    cavept_stats stats{};
    int build_levels = 0;
};

inline int fail(cavept_ctx* ctx, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (ctx) ctx->last_error = buf;
    return code;
}

#define CU(ctx, call)                                                                                   \
    do {                                                                                                \
        cudaError_t e_ = (call);                                                                        \
        if (e_ != cudaSuccess) return fail(ctx, CAVEPT_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); \
    } while (0)

inline int check_exec(cavept_ctx* ctx, const char* where) {
    if (ctx->ex.first_error != cudaSuccess) {
        const cudaError_t e = ctx->ex.first_error;
        ctx->ex.first_error = cudaSuccess;
        return fail(ctx, CAVEPT_ERR_CUDA, "%s: %s", where, cudaGetErrorString(e));
    }
    return CAVEPT_OK;
}

inline int select_device(cavept_ctx* ctx) {
    CU(ctx, cudaSetDevice(ctx->device));
    return CAVEPT_OK;
}

// cavept_group.cu
int cavept_group_sync(cavept_ctx* ctx);  // stream wait that polls ncclCommGetAsyncError
