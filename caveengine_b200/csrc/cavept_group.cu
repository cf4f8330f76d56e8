// Multi-GPU frame compositor behind the C ABI (include/cavept.h, "multi-GPU group"; SURVEY.md section 8b
// "Threading": "multi-GPU group object wraps N ctxs + NCCL comm", section 8e).  New relative to the reference,
// which renders on one GPU (cave/engine/renderer/graphics_manager.cpp:421-428 runs the single path-tracer pass).
//
// One process per GPU; each process's context joins a group: cavept_group_create() makes the NCCL communicator
// and switches the context to (sum, n) accumulation, so that compositing the frame is ONE ncclReduce /
// ncclAllReduce of W*H float4 sums over NVLink / NVSwitch followed by one normalising pass into a separate
// composite buffer (the local sums stay untouched: progressive accumulation carries on and compositing again later
// is valid).  The image-tile split (each GPU renders a band of rows) is one grouped ncclBroadcast per band, in place.
//
// NCCL is bound at run time (dlopen of libnccl.so.2 -- the copy already loaded in the process, e.g. PyTorch's, or
// the system one): libcavept.so has no link-time dependency on it and single-GPU users never load it.  The few
// NCCL declarations used are restated below (public API of NCCL 2.x, nccl.h) so the build needs no NCCL headers.
#include <dlfcn.h>
#include <unistd.h>

#include <cstring>
#include <mutex>

#include "cavept_ctx.h"

namespace {

// ---- the slice of nccl.h this file uses (ABI-stable across NCCL 2.x) ---------------------------------
typedef struct ncclComm* ncclComm_t;
typedef struct {
    char internal[128];
} ncclUniqueId;
typedef int ncclResult_t;  // ncclSuccess = 0, ncclInProgress = 7
constexpr int kNcclFloat32 = 7, kNcclSum = 0;
static_assert(sizeof(ncclUniqueId) == CAVEPT_GROUP_ID_BYTES, "ncclUniqueId");

struct NcclApi {
    void* handle = nullptr;
    std::string where;
    ncclResult_t (*GetVersion)(int*) = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*CommAbort)(ncclComm_t) = nullptr;
    ncclResult_t (*CommGetAsyncError)(ncclComm_t, ncclResult_t*) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*Reduce)(const void*, void*, size_t, int, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
};

NcclApi g_nccl;
std::mutex g_nccl_mutex;

template<class F>
bool bind(void* h, const char* name, F& fn) {
    fn = reinterpret_cast<F>(dlsym(h, name));
    return fn != nullptr;
}

// Loads NCCL once per process.  Order: $CAVEPT_NCCL_LIBRARY, then "libnccl.so.2" (which resolves to a copy the
// process already mapped -- PyTorch's bundled one -- before the system search path), then "libnccl.so".
const NcclApi* nccl_api(std::string* err) {
    std::lock_guard<std::mutex> lock(g_nccl_mutex);
    if (g_nccl.handle) return &g_nccl;
    const char* candidates[3] = { getenv("CAVEPT_NCCL_LIBRARY"), "libnccl.so.2", "libnccl.so" };
    void* h = nullptr;
    for (const char* c : candidates) {
        if (!c || !*c) continue;
        h = dlopen(c, RTLD_NOW | RTLD_GLOBAL);
        if (h) {
            g_nccl.where = c;
            break;
        }
    }
    if (!h) {
        if (err) *err = std::string("NCCL not found (set CAVEPT_NCCL_LIBRARY): ") + (dlerror() ? dlerror() : "dlopen failed");
        return nullptr;
    }
    NcclApi a;
    a.handle = h;
    a.where = g_nccl.where;
    const bool ok = bind(h, "ncclGetVersion", a.GetVersion) && bind(h, "ncclGetUniqueId", a.GetUniqueId) && bind(h, "ncclCommInitRank", a.CommInitRank) &&
                    bind(h, "ncclCommDestroy", a.CommDestroy) && bind(h, "ncclCommAbort", a.CommAbort) && bind(h, "ncclCommGetAsyncError", a.CommGetAsyncError) &&
                    bind(h, "ncclGetErrorString", a.GetErrorString) && bind(h, "ncclReduce", a.Reduce) && bind(h, "ncclAllReduce", a.AllReduce) &&
                    bind(h, "ncclBroadcast", a.Broadcast) && bind(h, "ncclGroupStart", a.GroupStart) && bind(h, "ncclGroupEnd", a.GroupEnd);
    if (!ok) {
        if (err) *err = "the NCCL library found lacks a required symbol";
        dlclose(h);
        return nullptr;
    }
    g_nccl = a;
    return &g_nccl;
}

int fail_nccl(cavept_ctx* ctx, const NcclApi* api, const char* what, ncclResult_t r) {
    return fail(ctx, CAVEPT_ERR_NCCL, "%s: %s", what, api && api->GetErrorString ? api->GetErrorString(r) : "NCCL error");
}

}  // namespace

struct cavept_group_state {
    const NcclApi* api = nullptr;
    ncclComm_t comm = nullptr;
    int rank = 0, nranks = 1;
    int nccl_version = 0;
    bool composite_valid = false;
};

#define NC(ctx, g, call)                                                      \
    do {                                                                      \
        ncclResult_t r_ = (call);                                             \
        if (r_ != 0) return fail_nccl(ctx, (g)->api, #call, r_);              \
    } while (0)

// Frames [first, first + n) split into nranks contiguous blocks that differ by at most one frame.
static void frame_block(int first, int n, int rank, int nranks, int* out_first, int* out_count) {
    const int base = n / nranks, rem = n % nranks;
    *out_count = base + (rank < rem ? 1 : 0);
    *out_first = first + rank * base + (rank < rem ? rank : rem);
}
// Rows of the image in whole multiples of `align` (the trace kernels work on 8x4 / 16x8 pixel tiles), as even
// as possible; the count may be 0 when there are more ranks than row groups.
static void row_band(int height, int rank, int nranks, int align, int* y0, int* h) {
    const int units = (height + align - 1) / align;
    const int base = units / nranks, rem = units % nranks;
    const int u0 = rank * base + (rank < rem ? rank : rem);
    const int u1 = u0 + base + (rank < rem ? 1 : 0);
    const int a = u0 * align < height ? u0 * align : height, b = u1 * align < height ? u1 * align : height;
    *y0 = a;
    *h = b - a;
}

extern "C" {

CAVEPT_API int cavept_group_unique_id(void* out_id, size_t bytes) {
    if (!out_id || bytes < CAVEPT_GROUP_ID_BYTES) return CAVEPT_ERR_INVALID_ARG;
    std::string err;
    const NcclApi* api = nccl_api(&err);
    if (!api) return CAVEPT_ERR_NCCL;
    ncclUniqueId id;
    if (api->GetUniqueId(&id) != 0) return CAVEPT_ERR_NCCL;
    std::memcpy(out_id, &id, sizeof(id));
    return CAVEPT_OK;
}

CAVEPT_API int cavept_group_create(cavept_ctx* ctx, int rank, int nranks, const void* unique_id, size_t bytes) {
    if (!ctx) return CAVEPT_ERR_INVALID_ARG;
    if (nranks < 1 || rank < 0 || rank >= nranks) return fail(ctx, CAVEPT_ERR_INVALID_ARG, "group_create: rank %d of %d", rank, nranks);
    if (ctx->group) return fail(ctx, CAVEPT_ERR_INVALID_ARG, "the context already belongs to a group");
    if (nranks > 1 && (!unique_id || bytes < CAVEPT_GROUP_ID_BYTES)) return fail(ctx, CAVEPT_ERR_INVALID_ARG, "group_create: a %d-byte unique id from cavept_group_unique_id is required", CAVEPT_GROUP_ID_BYTES);
    if (int e = select_device(ctx)) return e;
    cavept_group_state* g = new cavept_group_state();
    g->rank = rank;
    g->nranks = nranks;
    if (nranks > 1) {  // a group of one needs no communicator (and does not load NCCL)
        std::string err;
        g->api = nccl_api(&err);
        if (!g->api) {
            delete g;
            return fail(ctx, CAVEPT_ERR_NCCL, "%s", err.c_str());
        }
        g->api->GetVersion(&g->nccl_version);
        ncclUniqueId id;
        std::memcpy(&id, unique_id, sizeof(id));
        const ncclResult_t r = g->api->CommInitRank(&g->comm, nranks, id, rank);
        if (r != 0) {
            const int rc = fail_nccl(ctx, g->api, "ncclCommInitRank", r);
            delete g;
            return rc;
        }
    }
    // the accumulator switches from (mean, n) to (sum, n): what is already there is converted once
    const size_t n_pixels = (size_t)ctx->cfg.width * ctx->cfg.height;
    if (!ctx->accum_sums) launch_accum_scale(ctx->ex, ctx->d_accum.ptr, (int)n_pixels, true);
    ctx->accum_sums = 1;
    ctx->group = g;
    return check_exec(ctx, "group_create");
}

CAVEPT_API int cavept_group_info(const cavept_ctx* ctx, int* rank, int* nranks, int* nccl_version) {
    if (!ctx || !ctx->group) return CAVEPT_ERR_NOT_READY;
    if (rank) *rank = ctx->group->rank;
    if (nranks) *nranks = ctx->group->nranks;
    if (nccl_version) *nccl_version = ctx->group->nccl_version;
    return CAVEPT_OK;
}

CAVEPT_API int cavept_group_destroy(cavept_ctx* ctx) {
    if (!ctx) return CAVEPT_ERR_INVALID_ARG;
    cavept_group_state* g = ctx->group;
    if (!g) return CAVEPT_OK;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->ex.stream);
    int rc = CAVEPT_OK;
    if (g->comm) {
        const ncclResult_t r = g->api->CommDestroy(g->comm);
        if (r != 0) rc = fail_nccl(ctx, g->api, "ncclCommDestroy", r);
    }
    // back to the reference's (mean, n) accumulator
    launch_accum_scale(ctx->ex, ctx->d_accum.ptr, ctx->cfg.width * ctx->cfg.height, false);
    ctx->accum_sums = 0;
    ctx->group = nullptr;
    delete g;
    return rc != CAVEPT_OK ? rc : check_exec(ctx, "group_destroy");
}

CAVEPT_API int cavept_group_frame_block(int first_frame, int n_frames, int rank, int nranks, int* out_first, int* out_count) {
    if (nranks < 1 || rank < 0 || rank >= nranks || n_frames < 0 || !out_first || !out_count) return CAVEPT_ERR_INVALID_ARG;
    frame_block(first_frame, n_frames, rank, nranks, out_first, out_count);
    return CAVEPT_OK;
}

CAVEPT_API int cavept_group_row_band(int height, int rank, int nranks, int align, int* out_y0, int* out_rows) {
    if (nranks < 1 || rank < 0 || rank >= nranks || height < 0 || align < 1 || !out_y0 || !out_rows) return CAVEPT_ERR_INVALID_ARG;
    row_band(height, rank, nranks, align, out_y0, out_rows);
    return CAVEPT_OK;
}

// Sample split: every rank rendered its own c_frameIndex block into (sum, n).  One collective on the context's
// stream, then the normalising pass (sum / n) into the composite buffer of the destination rank(s).
CAVEPT_API int cavept_group_composite(cavept_ctx* ctx, int dst_rank) {
    if (!ctx) return CAVEPT_ERR_INVALID_ARG;
    cavept_group_state* g = ctx->group;
    if (!g) return fail(ctx, CAVEPT_ERR_NOT_READY, "the context is not in a group (cavept_group_create)");
    if (dst_rank < -1 || dst_rank >= g->nranks) return fail(ctx, CAVEPT_ERR_INVALID_ARG, "composite: destination rank %d of %d", dst_rank, g->nranks);
    if (int e = select_device(ctx)) return e;
    const size_t n_pixels = (size_t)ctx->cfg.width * ctx->cfg.height;
    const bool mine = dst_rank < 0 || dst_rank == g->rank;
    if (mine) CU(ctx, ctx->d_composite.reserve(n_pixels));
    if (g->comm) {
        if (dst_rank < 0) NC(ctx, g, g->api->AllReduce(ctx->d_accum.ptr, ctx->d_composite.ptr, n_pixels * 4, kNcclFloat32, kNcclSum, g->comm, ctx->ex.stream));
        else NC(ctx, g, g->api->Reduce(ctx->d_accum.ptr, mine ? ctx->d_composite.ptr : nullptr, n_pixels * 4, kNcclFloat32, kNcclSum, dst_rank, g->comm, ctx->ex.stream));
        ++ctx->stats.collectives;
    } else {
        CU(ctx, cudaMemcpyAsync(ctx->d_composite.ptr, ctx->d_accum.ptr, n_pixels * sizeof(F4), cudaMemcpyDeviceToDevice, ctx->ex.stream));
    }
    if (mine) {
        launch_accum_scale(ctx->ex, ctx->d_composite.ptr, (int)n_pixels, false);
        g->composite_valid = true;
    }
    return check_exec(ctx, "group_composite");
}

// Image-tile split: rank r rendered rows cavept_group_row_band(height, r, nranks, align) with cavept_render_tile.
// One grouped broadcast per band, in place in the accumulator: afterwards every rank holds the whole image
// (no arithmetic in the exchange: the result is bit-exact).
CAVEPT_API int cavept_group_gather_bands(cavept_ctx* ctx, int align) {
    if (!ctx) return CAVEPT_ERR_INVALID_ARG;
    cavept_group_state* g = ctx->group;
    if (!g) return fail(ctx, CAVEPT_ERR_NOT_READY, "the context is not in a group (cavept_group_create)");
    if (align < 1) return fail(ctx, CAVEPT_ERR_INVALID_ARG, "gather_bands: align %d", align);
    if (!g->comm) return CAVEPT_OK;
    if (int e = select_device(ctx)) return e;
    const int W = ctx->cfg.width, H = ctx->cfg.height;
    NC(ctx, g, g->api->GroupStart());
    for (int r = 0; r < g->nranks; ++r) {
        int y0, h;
        row_band(H, r, g->nranks, align, &y0, &h);
        if (h <= 0) continue;
        float* band = reinterpret_cast<float*>(ctx->d_accum.ptr + (size_t)y0 * W);
        const ncclResult_t rr = g->api->Broadcast(band, band, (size_t)h * W * 4, kNcclFloat32, r, g->comm, ctx->ex.stream);
        if (rr != 0) {
            g->api->GroupEnd();
            return fail_nccl(ctx, g->api, "ncclBroadcast", rr);
        }
    }
    NC(ctx, g, g->api->GroupEnd());
    ++ctx->stats.collectives;
    return CAVEPT_OK;
}

CAVEPT_API int cavept_read_composite(cavept_ctx* ctx, float* out, size_t bytes) {
    if (!ctx || !out) return CAVEPT_ERR_INVALID_ARG;
    if (!ctx->group || !ctx->group->composite_valid) return fail(ctx, CAVEPT_ERR_NOT_READY, "no composite on this rank (cavept_group_composite with this rank as destination)");
    const size_t need = (size_t)ctx->cfg.width * ctx->cfg.height * sizeof(F4);
    if (bytes < need) return fail(ctx, CAVEPT_ERR_CAPACITY, "need %zu bytes", need);
    if (int e = select_device(ctx)) return e;
    if (int e = cavept_sync(ctx)) return e;  // polls the communicator: a failed collective surfaces here, not as a hang
    ctx->ex.download(out, ctx->d_composite.ptr, need);
    return check_exec(ctx, "composite read-back");
}

CAVEPT_API void* cavept_composite_device_ptr(cavept_ctx* ctx) { return (ctx && ctx->group && ctx->group->composite_valid) ? (void*)ctx->d_composite.ptr : nullptr; }

}  // extern "C"

// Called by cavept_sync when the context is in a group: waits for the stream while polling
// ncclCommGetAsyncError, so a peer failure aborts the communicator and returns CAVEPT_ERR_NCCL instead of hanging.
int cavept_group_sync(cavept_ctx* ctx) {
    cavept_group_state* g = ctx->group;
    if (!g || !g->comm) {
        CU(ctx, cudaStreamSynchronize(ctx->ex.stream));
        return CAVEPT_OK;
    }
    for (unsigned spins = 0;; ++spins) {
        const cudaError_t q = cudaStreamQuery(ctx->ex.stream);
        if (q == cudaSuccess) break;
        if (q != cudaErrorNotReady) return fail(ctx, CAVEPT_ERR_CUDA, "cudaStreamQuery: %s", cudaGetErrorString(q));
        ncclResult_t async = 0;
        const ncclResult_t r = g->api->CommGetAsyncError(g->comm, &async);
        if (r != 0 || (async != 0 && async != 7 /* ncclInProgress */)) {
            const int rc = fail_nccl(ctx, g->api, "asynchronous NCCL error (communicator aborted)", r != 0 ? r : async);
            g->api->CommAbort(g->comm);
            g->comm = nullptr;
            return rc;
        }
        if (spins > 64) usleep(20);
    }
    ncclResult_t async = 0;
    if (g->api->CommGetAsyncError(g->comm, &async) == 0 && async != 0 && async != 7) return fail_nccl(ctx, g->api, "asynchronous NCCL error", async);
    return CAVEPT_OK;
}
