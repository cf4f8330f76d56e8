// cavept_sort_pairs_static: the device radix sort of cavept_sort_pairs for callers that cannot allocate (code under
// CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// The sort's scratch -- the ping-pong key / value buffers, the per-(digit, block) histograms with their scanned
// offsets, and the scan's block sums -- lives in ONE file-scope __device__ array below instead of cudaMalloc'd buffers.
// Every GPU that loads the library holds the array (kStaticBytes, about 8.5 MB) until the process exits.
//
// In host code the array's NAME is a host-side stand-in, not a device address: its device address is fetched with
// cudaGetSymbolAddress, once per device, with that device current, and the kernels are handed that pointer.
//
// Every call on a device shares the array.  Nothing in it has to start at zero (each pass writes every histogram
// slot, offset, block sum and ping-pong entry before reading it), so nothing is cleared between calls; what a call
// leaves behind is never read by the next.  Calls that OVERLAP on one GPU (two streams, two contexts, two threads) race
// on it: callers serialise them.
#include <mutex>

#include "cavept_ctx.h"

namespace {

constexpr int kMaxPairs = CAVEPT_STATIC_SORT_MAX_PAIRS;
constexpr size_t align256(size_t b) { return (b + 255) & ~(size_t)255; }
constexpr size_t kHistEntries = (size_t)256 * ((kMaxPairs + kSortTile - 1) / kSortTile) + 1;  // == sort_hist_entries(kMaxPairs)
constexpr size_t kOffKeys = 0;
constexpr size_t kOffVals = kOffKeys + align256((size_t)kMaxPairs * sizeof(uint32_t));
constexpr size_t kOffHist = kOffVals + align256((size_t)kMaxPairs * sizeof(int));
constexpr size_t kOffScan = kOffHist + align256(2 * kHistEntries * sizeof(int));
constexpr size_t kStaticBytes = kOffScan + align256((kHistEntries / kScanTile + 4) * sizeof(int));

static __device__ __align__(256) unsigned char g_sort_scratch[kStaticBytes];

constexpr int kMaxDevices = 64;
unsigned char* g_scratch_ptr[kMaxDevices] = {};  // device address of g_sort_scratch on each device, fetched on first use
std::mutex g_scratch_mutex;

// The array's address on `device` (which the caller has made current).  Not an allocation: the array came with the module.
cudaError_t scratch_on_device(int device, unsigned char** out) {
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (g_scratch_ptr[device] == nullptr) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_sort_scratch);
        if (e != cudaSuccess) return e;
        g_scratch_ptr[device] = static_cast<unsigned char*>(p);
    }
    *out = g_scratch_ptr[device];
    return cudaSuccess;
}

}  // namespace

extern "C" CAVEPT_API int cavept_sort_pairs_static(cavept_ctx* ctx, uint32_t* d_keys, int32_t* d_values, int count, int key_bits) {
    CAVEPT_RANGE();
    if (!ctx || !d_keys || !d_values || count < 0 || key_bits < 1 || key_bits > 32) return fail(ctx, CAVEPT_ERR_INVALID_ARG, "sort_pairs_static: bad argument");
    if (count > kMaxPairs)
        return fail(ctx, CAVEPT_ERR_CAPACITY, "sort_pairs_static: %d pairs exceed the static scratch's maximum of %d (CAVEPT_STATIC_SORT_MAX_PAIRS); use cavept_sort_pairs", count,
                    kMaxPairs);
    if (ctx->device < 0 || ctx->device >= kMaxDevices) return fail(ctx, CAVEPT_ERR_CAPACITY, "sort_pairs_static: device ordinal %d beyond %d", ctx->device, kMaxDevices - 1);
    if (int e = select_device(ctx)) return e;
    if (count <= 1) return CAVEPT_OK;
    unsigned char* base = nullptr;
    CU(ctx, scratch_on_device(ctx->device, &base));
    CudaExec ex;  // the context's stream, the static array's block sums: the context's own scan scratch is not touched
    ex.stream = ctx->ex.stream;
    ex.scan_tmp = reinterpret_cast<int*>(base + kOffScan);
    ex.sort_pairs(d_keys, d_values, reinterpret_cast<uint32_t*>(base + kOffKeys), reinterpret_cast<int*>(base + kOffVals), reinterpret_cast<int*>(base + kOffHist), count, key_bits);
    ctx->ex.launches += ex.launches;
    ctx->ex.note(ex.first_error);
    return check_exec(ctx, "sort_pairs_static");
}
