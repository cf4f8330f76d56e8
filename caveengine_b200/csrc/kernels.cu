// __global__ wrappers for the kernel bodies, the scan primitive, and launchers.
// Compiled for sm_100a only, with -fmad=false (see cv_common.h on float semantics).
#include "kernels.h"
#if defined(CV_PARITY_SUB_PROFILE) || defined(CV_SUB_PROFILE)
#include <cstdio>
#endif
#ifdef CV_SUB_PROFILE
#define SUBT(k) { const long long t1_ = clock64(); prof[k] += t1_ - t0_; t0_ = t1_; }
#else
#define SUBT(k)
#endif

#include <cuda_fp16.h>

namespace cavept {

// ---- generic element-wise launch -------------------------------------------------------
template<class F>
__global__ void __launch_bounds__(256) k_for_each(F f, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) f(i);
}

template<class F>
void CudaExec::launch(int n, F f) {
    if (n <= 0) return;
    k_for_each<F><<<(n + 255) / 256, 256, 0, stream>>>(f, n);
    note(cudaGetLastError());
    ++launches;
}

void CudaExec::upload(void* dst, const void* src, size_t bytes) {
    if (bytes) note(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, stream));
    // pageable sources are staged before the call returns; pinned sources must stay live
    // until the stream reaches the copy -- callers that reuse pinned buffers synchronise.
}
void CudaExec::download(void* dst, const void* src, size_t bytes) {
    if (bytes) note(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, stream));
    note(cudaStreamSynchronize(stream));
}
void CudaExec::fill(void* dst, int byte, size_t bytes) {
    if (bytes) note(cudaMemsetAsync(dst, byte, bytes, stream));
}

// ---- exclusive scan (reduce, scan of block sums, apply) ---------------------------------
// Tiles of kScanTile = 256 threads x 8 consecutive elements.  out has n + 1 entries.
constexpr int kScanThreads = 256;
constexpr int kScanPerThread = kScanTile / kScanThreads;

__device__ __forceinline__ int block_exclusive_scan(int v, int* total) {
    // exclusive scan of one int per thread over a 256-thread block
    __shared__ int warp_sums[kScanThreads / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        int s = lane < kScanThreads / 32 ? warp_sums[lane] : 0;
#pragma unroll
        for (int d = 1; d < kScanThreads / 32; d <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, s, d);
            if (lane >= d) s += t;
        }
        if (lane < kScanThreads / 32) warp_sums[lane] = s;  // inclusive over warps
    }
    __syncthreads();
    const int warp_offset = warp == 0 ? 0 : warp_sums[warp - 1];
    *total = warp_sums[kScanThreads / 32 - 1];
    const int result = warp_offset + incl - v;
    __syncthreads();  // warp_sums is reused by the next call
    return result;
}

template<class T>
__global__ void __launch_bounds__(kScanThreads) k_scan_tile_sums(const T* __restrict__ in, int* __restrict__ sums, int n) {
    const int base = blockIdx.x * kScanTile + threadIdx.x * kScanPerThread;
    int s = 0;
#pragma unroll
    for (int k = 0; k < kScanPerThread; ++k) {
        const int i = base + k;
        if (i < n) s += (int)in[i];
    }
    int total;
    block_exclusive_scan(s, &total);
    if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

// single block: sums[0..nb) -> exclusive scan in place, sums[nb] = grand total
__global__ void __launch_bounds__(kScanThreads) k_scan_sums(int* sums, int nb) {
    __shared__ int carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < nb; base += kScanThreads) {
        const int i = base + threadIdx.x;
        const int v = i < nb ? sums[i] : 0;
        int total;
        const int ex = block_exclusive_scan(v, &total);
        const int carry = carry_s;
        if (i < nb) sums[i] = carry + ex;
        __syncthreads();
        if (threadIdx.x == 0) carry_s = carry + total;
        __syncthreads();
    }
    if (threadIdx.x == 0) sums[nb] = carry_s;
}

template<class T>
__global__ void __launch_bounds__(kScanThreads) k_scan_apply(const T* __restrict__ in, const int* __restrict__ sums, int* __restrict__ out, int n, int nb) {
    const int base = blockIdx.x * kScanTile + threadIdx.x * kScanPerThread;
    int v[kScanPerThread];
    int s = 0;
#pragma unroll
    for (int k = 0; k < kScanPerThread; ++k) {
        const int i = base + k;
        v[k] = i < n ? (int)in[i] : 0;
        s += v[k];
    }
    int total;
    int run = block_exclusive_scan(s, &total) + sums[blockIdx.x];
#pragma unroll
    for (int k = 0; k < kScanPerThread; ++k) {
        const int i = base + k;
        if (i < n) out[i] = run;
        run += v[k];
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) out[n] = sums[nb];
}

template<class T>
static void scan_impl(CudaExec& ex, const T* in, int* out, int n) {
    if (n <= 0) return;
    const int nb = (n + kScanTile - 1) / kScanTile;
    k_scan_tile_sums<T><<<nb, kScanThreads, 0, ex.stream>>>(in, ex.scan_tmp, n);
    k_scan_sums<<<1, kScanThreads, 0, ex.stream>>>(ex.scan_tmp, nb);
    k_scan_apply<T><<<nb, kScanThreads, 0, ex.stream>>>(in, ex.scan_tmp, out, n, nb);
    ex.note(cudaGetLastError());
    ex.launches += 3;
}
void CudaExec::scan_u8(const unsigned char* in, int* out, int n) { scan_impl<unsigned char>(*this, in, out, n); }
void CudaExec::scan_i32(const int* in, int* out, int n) { scan_impl<int>(*this, in, out, n); }

// ---- radix sort (Morton keys of the LBVH builder) -------------------------------------------
// Stable LSD sort, 8 bits per pass.  A block owns a tile of kSortTile = 256 x 16 consecutive
// keys; warp w owns 512 of them and visits them in order (round r, lane l), so the rank of a
// key among equal digits is (equal digits in earlier warps) + (in earlier rounds of this warp)
// + (in lower lanes of this round, from match.any).  Per pass: block histograms -> exclusive
// scan over (digit, block) -> ranked scatter.
constexpr int kSortThreads = 256;
constexpr int kSortRounds = kSortTile / kSortThreads;  // 16
constexpr int kSortWarps = kSortThreads / 32;

__global__ void __launch_bounds__(kSortThreads) k_sort_hist(const uint32_t* __restrict__ keys, int* __restrict__ hist, int n, int shift, int nb) {
    __shared__ int h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const int base = blockIdx.x * kSortTile;
#pragma unroll 4
    for (int r = 0; r < kSortRounds; ++r) {
        const int e = base + r * kSortThreads + threadIdx.x;
        if (e < n) atomicAdd(&h[(keys[e] >> shift) & 255u], 1);
    }
    __syncthreads();
    hist[threadIdx.x * nb + blockIdx.x] = h[threadIdx.x];
}

__global__ void __launch_bounds__(kSortThreads) k_sort_scatter(const uint32_t* __restrict__ keys, const int* __restrict__ vals, uint32_t* __restrict__ out_keys,
                                                               int* __restrict__ out_vals, const int* __restrict__ offsets, int n, int shift, int nb) {
    __shared__ int cnt[kSortWarps][256];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < kSortWarps * 256; i += kSortThreads) (&cnt[0][0])[i] = 0;
    __syncthreads();
    const int base = blockIdx.x * kSortTile + warp * (32 * kSortRounds);
    uint32_t key[kSortRounds];
    int val[kSortRounds], rank[kSortRounds];
    const unsigned lt = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < kSortRounds; ++r) {
        const int e = base + r * 32 + lane;
        const bool valid = e < n;
        key[r] = valid ? keys[e] : 0xffffffffu;  // padding sorts (and ranks) after every real key of the tile
        val[r] = valid ? vals[e] : 0;
        const int d = valid ? (int)((key[r] >> shift) & 255u) : 255;
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        const int before = cnt[warp][d];
        __syncwarp();
        if ((peers & lt) == 0u) cnt[warp][d] = before + __popc(peers);
        __syncwarp();
        rank[r] = before + __popc(peers & lt);
    }
    __syncthreads();
    {  // per digit: exclusive prefix over the warps, plus the block's global offset for that digit
        const int d = threadIdx.x;
        int run = offsets[d * nb + blockIdx.x];
#pragma unroll
        for (int w = 0; w < kSortWarps; ++w) {
            const int c = cnt[w][d];
            cnt[w][d] = run;
            run += c;
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kSortRounds; ++r) {
        const int e = base + r * 32 + lane;
        if (e < n) {
            const int d = (int)((key[r] >> shift) & 255u);
            const int pos = cnt[warp][d] + rank[r];
            out_keys[pos] = key[r];
            out_vals[pos] = val[r];
        }
    }
}

void CudaExec::sort_pairs(uint32_t* keys, int* vals, uint32_t* keys_tmp, int* vals_tmp, int* hist, int n, int bits) {
    if (n <= 1) return;
    const int nb = (n + kSortTile - 1) / kSortTile;
    int passes = (bits + 7) / 8;
    if (passes & 1) ++passes;  // an even number of passes leaves the result in keys / vals
    for (int pass = 0; pass < passes; ++pass) {
        const int shift = 8 * pass;
        k_sort_hist<<<nb, kSortThreads, 0, stream>>>(keys, hist, n, shift, nb);
        ++launches;
        int* offsets = hist + sort_hist_entries(n);
        scan_i32(hist, offsets, 256 * nb);
        k_sort_scatter<<<nb, kSortThreads, 0, stream>>>(keys, vals, keys_tmp, vals_tmp, offsets, n, shift, nb);
        ++launches;
        note(cudaGetLastError());
        uint32_t* tk = keys; keys = keys_tmp; keys_tmp = tk;
        int* tv = vals; vals = vals_tmp; vals_tmp = tv;
    }
}

// ---- parity builder: the two per-triangle passes with atomics (bvh_parity_body.h body_reduce / body_bin are the
//      per-element statements).  At the upper levels every triangle of a block belongs to the same item: the block folds
//      into a private copy in shared memory and flushes it with one atomic per word, instead of every warp hitting the
//      item's accumulator (a 262k-triangle mesh: 8192 warps x 18 atomics on the same 18 words).  Min / max / add are
//      order-free: same bits out.  A block that straddles items runs the per-element body.
__global__ void __launch_bounds__(256) k_parity_reduce(const BuildArrays a) {
    __shared__ ItemAccum s_acc;
    const int tid = threadIdx.x, i = blockIdx.x * 256 + tid;
    const int w0 = a.node_of[blockIdx.x * 256];
    const int w = i < a.n_tris ? a.node_of[i] : w0;
    if (!__syncthreads_and(w == w0)) {
        body_reduce(a, i);
        return;
    }
    if (w0 < 0) return;
    if (tid < 3) {
        s_acc.mn[tid] = ~0ull;
        s_acc.mx[tid] = 0ull;
        s_acc.cmn[tid] = 0xffffffffu;
        s_acc.cmx[tid] = 0u;
    }
    __syncthreads();
    if (i < a.n_tris) body_reduce_to(a, i, w0, s_acc);
    __syncthreads();
    ItemAccum& acc = a.accum[w0];
    if (tid < 3) atomicMin(&acc.mn[tid], s_acc.mn[tid]);
    else if (tid < 6) atomicMax(&acc.mx[tid - 3], s_acc.mx[tid - 3]);
    else if (tid < 9) atomicMin(&acc.cmn[tid - 6], s_acc.cmn[tid - 6]);
    else if (tid < 12) atomicMax(&acc.cmx[tid - 9], s_acc.cmx[tid - 9]);
}
void CudaExec::parity_reduce(const BuildArrays& a) {
    if (a.n_tris <= 0) return;
    k_parity_reduce<<<(a.n_tris + 255) / 256, 256, 0, stream>>>(a);
    note(cudaGetLastError());
    ++launches;
}
__global__ void __launch_bounds__(256) k_parity_bin(const BuildArrays a) {
    __shared__ ItemBins s_bins;
    const int tid = threadIdx.x, i = blockIdx.x * 256 + tid;
    const int w0 = a.node_of[blockIdx.x * 256];
    const int w = i < a.n_tris ? a.node_of[i] : w0;
    if (!__syncthreads_and(w == w0)) {
        body_bin(a, i);
        return;
    }
    if (w0 < 0) return;
    const BuildItem& it = a.items[w0];
    if (it.kind != ITEM_SAH) return;
    if (tid < kBuckets) {
        s_bins.count[tid] = 0;
        for (int c = 0; c < 3; ++c) {
            s_bins.mn[tid][c] = 0xffffffffu;
            s_bins.mx[tid][c] = 0u;
        }
    }
    __syncthreads();
    if (i < a.n_tris) body_bin_to(a, i, it, s_bins);
    __syncthreads();
    if (tid < kBuckets && s_bins.count[tid] > 0) {
        ItemBins& b = a.bins[it.bin_slot];
        atomicAdd(&b.count[tid], s_bins.count[tid]);
        for (int c = 0; c < 3; ++c) {
            atomicMin(&b.mn[tid][c], s_bins.mn[tid][c]);
            atomicMax(&b.mx[tid][c], s_bins.mx[tid][c]);
        }
    }
}
void CudaExec::parity_bin(const BuildArrays& a) {
    if (a.n_tris <= 0) return;
    k_parity_bin<<<(a.n_tris + 255) / 256, 256, 0, stream>>>(a);
    note(cudaGetLastError());
    ++launches;
}

// ---- parity builder, second phase: one block builds the whole subtree of each queued item (bvh_parity_body.h
//      parity_subtree_serial is the serial statement the emulator runs).  The passes are the level loop's own bodies over
//      the block's element range and its private slices of the item / accumulator / bucket arrays, with __syncthreads()
//      where the level loop has kernel boundaries; the "goes left" scan is block-local in shared memory.
#ifndef CV_PARITY_SUB_THREADS
#define CV_PARITY_SUB_THREADS 512
#endif
constexpr int kParSubThreads = CV_PARITY_SUB_THREADS;
#ifndef CV_PARITY_SUB_MINBLOCKS
#define CV_PARITY_SUB_MINBLOCKS 2
#endif
// The fold pass of a subtree build.  Elements of an item are contiguous in `order`, so a warp's lanes fold by a segmented
// shuffle scan (five steps, no divergence however many items the warp straddles) and the last lane of each segment sends
// the atomics -- body_reduce's match.any groups would run one after the other, up to 32 per warp at the deep levels.
// Same operands, order-free min / max: same accumulator bits.
__device__ __forceinline__ void par_sub_reduce(const BuildArrays& b, int start, int end, int tid, int lane) {
    for (int base = start; base < end; base += kParSubThreads) {
        const int i = base + tid;
        const int w = i < end ? b.node_of[i] : -1;
        unsigned long long kmn[3] = { ~0ull, ~0ull, ~0ull }, kmx[3] = { 0ull, 0ull, 0ull };
        uint32_t cmn[3] = { 0xffffffffu, 0xffffffffu, 0xffffffffu }, cmx[3] = { 0u, 0u, 0u };
        if (w >= 0) {
            uint32_t ck[3];
            body_element_keys(b, i, kmn, kmx, ck);
#pragma unroll
            for (int k = 0; k < 3; ++k) cmn[k] = cmx[k] = ck[k];
        }
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int wo = __shfl_up_sync(0xffffffffu, w, d);
            const bool take = lane >= d && wo == w;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const unsigned long long on = __shfl_up_sync(0xffffffffu, kmn[k], d), ox = __shfl_up_sync(0xffffffffu, kmx[k], d);
                const uint32_t cn = __shfl_up_sync(0xffffffffu, cmn[k], d), cx = __shfl_up_sync(0xffffffffu, cmx[k], d);
                if (take) {
                    kmn[k] = on < kmn[k] ? on : kmn[k];
                    kmx[k] = ox > kmx[k] ? ox : kmx[k];
                    cmn[k] = cn < cmn[k] ? cn : cmn[k];
                    cmx[k] = cx > cmx[k] ? cx : cmx[k];
                }
            }
        }
        const int wn = __shfl_down_sync(0xffffffffu, w, 1);
        if (w >= 0 && (lane == 31 || wn != w)) {
            ItemAccum& acc = b.accum[w];
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                atomicMin(&acc.mn[k], kmn[k]);
                atomicMax(&acc.mx[k], kmx[k]);
                atomicMin(&acc.cmn[k], cmn[k]);
                atomicMax(&acc.cmx[k], cmx[k]);
            }
        }
    }
}
// The bucket pass of a subtree build: one atomic per word and element, no grouping -- inside a subtree at most sub_max / 12
// elements meet on a word, and the lanes of a warp scatter over up to 12 buckets per item (grouping them runs the groups
// one after the other).  count / min / max are order-free: same bucket bits.
__device__ __forceinline__ void par_sub_bin(const BuildArrays& b, int start, int end, int tid) {
    for (int i = start + tid; i < end; i += kParSubThreads) {
        const int w = b.node_of[i];
        if (w < 0) continue;
        const BuildItem& it = b.items[w];
        if (it.kind != ITEM_SAH) continue;
        int g;
        const int slot = body_bin_slot(b, i, it, g);
        const F4 bmn = b.tri_bmin[g], bmx = b.tri_bmax[g];
        ItemBins& bb = b.bins[it.bin_slot];
        atomicAdd(&bb.count[slot], 1);
        atomicMin(&bb.mn[slot][0], ordered_key(bmn.x));
        atomicMin(&bb.mn[slot][1], ordered_key(bmn.y));
        atomicMin(&bb.mn[slot][2], ordered_key(bmn.z));
        atomicMax(&bb.mx[slot][0], ordered_key(bmx.x));
        atomicMax(&bb.mx[slot][1], ordered_key(bmx.y));
        atomicMax(&bb.mx[slot][2], ordered_key(bmx.z));
    }
}
// The split pass of a subtree build: sixteen lanes per item.  Lane j < 12 holds bucket j of the item's record (seven
// words); the buckets are passed round by shuffle in bucket order and lane i < 11 folds bucket j into the left side of
// candidate i when j <= i, else into its right side -- the folds of body_split_cost in the same order with the same
// operands (same cost bits), without every lane holding the whole 84-word record in registers.  Then every lane repeats
// the serial selection over the eleven costs (first strict minimum, :190-193).
// `list` holds the level's bucketed items (most items of the deep levels are not: one look per thread finds them).
__device__ __forceinline__ void par_split_item(const BuildArrays& b, BuildItem& it, int lane);
__device__ __forceinline__ void par_sub_split(const BuildArrays& b, const int* list, int n_list, int tid, int lane) {
    for (int q = tid >> 4; q < n_list; q += kParSubThreads / 16) par_split_item(b, b.items[list[q]], lane);
}
// one bucketed item on the sixteen lanes of a half warp (all sixteen call it)
__device__ __forceinline__ void par_split_item(const BuildArrays& b, BuildItem& it, int lane) {
    const unsigned hmask = 0xffffu << (lane & 16);
    const int sub = lane & 15, half = lane & 16;
    {
        const ItemBins& rec = b.bins[it.bin_slot];
        int cnt = 0;
        F3 bmn = f3(CV_INF, CV_INF, CV_INF), bmx = f3(-CV_INF, -CV_INF, -CV_INF);
        if (sub < kBuckets) {
            cnt = rec.count[sub];
            if (cnt != 0) {  // an untouched bucket decodes to (+inf, -inf) == Box::MakeInvalid (box.h:19-22)
                bmn = f3(ordered_value(rec.mn[sub][0]), ordered_value(rec.mn[sub][1]), ordered_value(rec.mn[sub][2]));
                bmx = f3(ordered_value(rec.mx[sub][0]), ordered_value(rec.mx[sub][1]), ordered_value(rec.mx[sub][2]));
            }
        }
        F3 mn0 = f3(CV_INF, CV_INF, CV_INF), mx0 = f3(-CV_INF, -CV_INF, -CV_INF);
        F3 mn1 = mn0, mx1 = mx0;
        int c0 = 0, c1 = 0;
#pragma unroll
        for (int j = 0; j < kBuckets; ++j) {
            const int src = half + j;
            const int cj = __shfl_sync(hmask, cnt, src);
            const F3 nj = f3(__shfl_sync(hmask, bmn.x, src), __shfl_sync(hmask, bmn.y, src), __shfl_sync(hmask, bmn.z, src));
            const F3 xj = f3(__shfl_sync(hmask, bmx.x, src), __shfl_sync(hmask, bmx.y, src), __shfl_sync(hmask, bmx.z, src));
            if (j <= sub) {
                mn0 = f3(min_ref(mn0.x, nj.x), min_ref(mn0.y, nj.y), min_ref(mn0.z, nj.z));
                mx0 = f3(max_ref(mx0.x, xj.x), max_ref(mx0.y, xj.y), max_ref(mx0.z, xj.z));
                c0 += cj;
            } else {
                mn1 = f3(min_ref(mn1.x, nj.x), min_ref(mn1.y, nj.y), min_ref(mn1.z, nj.z));
                mx1 = f3(max_ref(mx1.x, xj.x), max_ref(mx1.y, xj.y), max_ref(mx1.z, xj.z));
                c1 += cj;
            }
        }
        const float cost = 0.125f + cv_div((float)c0 * surface_area(mn0, mx0) + (float)c1 * surface_area(mn1, mx1), it.parent_sa);
        int split = 0;
        float best = 0.0f;
#pragma unroll
        for (int i = 0; i < kBuckets - 1; ++i) {
            const float c = __shfl_sync(hmask, cost, half + i);
            if (i == 0 || c < best) {
                split = i;
                best = c;
            }
        }
        if (sub == 0) it.split = split;
    }
}
// the level loop's split pass on the same sixteen-lane statement (the per-item serial body_split prices the eleven
// candidates one after the other on one thread: 11 us per level for a few hundred items)
__global__ void __launch_bounds__(256) k_parity_split(const BuildArrays a, int n_items) {
    const int w = (int)((blockIdx.x * 256u + threadIdx.x) >> 4);
    if (w >= n_items) return;  // uniform over a half warp
    BuildItem& it = a.items[w];
    if (it.kind != ITEM_SAH) return;
    par_split_item(a, it, threadIdx.x & 31);
}
void CudaExec::parity_split(const BuildArrays& a, int n_items) {
    if (n_items <= 0) return;
    k_parity_split<<<(n_items * 16 + 255) / 256, 256, 0, stream>>>(a, n_items);
    note(cudaGetLastError());
    ++launches;
}
__global__ void __launch_bounds__(kParSubThreads, CV_PARITY_SUB_MINBLOCKS) k_parity_subtrees(const BuildArrays ga, int n_roots) {
    __shared__ int s_scan[kParitySubMax + 1];
    __shared__ int s_counters[kParityCounters];
    __shared__ int s_warp[kParSubThreads / 32];
    __shared__ int s_list[kParitySubMax / 5 + 1], s_n_list;  // the level's bucketed items (each has at least 5 triangles)
    const int r = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (r >= n_roots) return;
    const BuildItem root = ga.sub_roots[r];
    const int start = root.start, n = root.count, end = start + n;
    BuildArrays b;
    parity_subtree_view(ga, root, b, s_counters, s_scan - start);
    if (tid == 0) {
        parity_subtree_setup(b, root);
        s_n_list = 0;
    }
    for (int i = start + tid; i < end; i += kParSubThreads) b.node_of[i] = 0;
    __syncthreads();
    const int per = (n + kParSubThreads - 1) / kParSubThreads;
    const int j0 = min(tid * per, n), j1 = min(j0 + per, n);
    int n_items = 1, levels = 0, err = 0;
#ifdef CV_PARITY_SUB_PROFILE
    long long ph[9] = { 0 }, t0 = clock64(), t1;
#define PH(k) t1 = clock64(); ph[k] += t1 - t0; t0 = t1;
#else
#define PH(k)
#endif
    while (n_items > 0) {
        for (int w = tid; w < n_items; w += kParSubThreads) body_init_accum(b, w, n_items);
        __syncthreads();
        PH(0)
        par_sub_reduce(b, start, end, tid, lane);
        __syncthreads();
        PH(1)
        for (int w = tid; w < n_items; w += kParSubThreads) body_node(b, w, n_items);
        __syncthreads();
        PH(2)
        par_sub_bin(b, start, end, tid);
        __syncthreads();
        PH(3)
        for (int w = tid; w < n_items; w += kParSubThreads) {
            if (b.items[w].kind == ITEM_SAH) s_list[atomicAdd(&s_n_list, 1)] = w;
        }
        __syncthreads();
        par_sub_split(b, s_list, s_n_list, tid, lane);
        __syncthreads();
        PH(4)
        {  // flags of a thread's contiguous chunk, then the exclusive scan over the block's range
            int sum = 0;
            for (int j = j0; j < j1; ++j) {
                body_flag(b, start + j);
                sum += b.flags[start + j];
            }
            int inc = sum;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, inc, d);
                if (lane >= d) inc += v;
            }
            if (lane == 31) s_warp[warp] = inc;
            __syncthreads();
        PH(5)
            int base = 0;
            for (int k = 0; k < warp; ++k) base += s_warp[k];
            int run = base + inc - sum;
            for (int j = j0; j < j1; ++j) {
                s_scan[j] = run;
                run += b.flags[start + j];
            }
            if (tid == kParSubThreads - 1) s_scan[n] = base + inc;
        }
        __syncthreads();
        PH(6)
        for (int w = tid; w < n_items; w += kParSubThreads) body_children(b, w, n_items);
        __syncthreads();
        PH(7)
        for (int i = start + tid; i < end; i += kParSubThreads) body_scatter(b, i);
        __syncthreads();
        PH(8)
        ++levels;
        n_items = s_counters[0];
        err = s_counters[2];
        __syncthreads();
        PH(8)
        if (tid == 0) {
            s_counters[0] = 0;
            s_counters[1] = 0;
            s_n_list = 0;
        }
        swap_ptr(b.items, b.items_next);
        swap_ptr(b.order, b.order_next);
        swap_ptr(b.node_of, b.node_of_next);
        if (err != 0 || root.depth + levels > kMaxDepth + 2) break;
        __syncthreads();
        PH(8)
    }
#ifdef CV_PARITY_SUB_PROFILE
    if (tid == 0 && (r % 64) == 0) printf("root %d n %d levels %d cycles init %lld reduce %lld node %lld bin %lld split %lld flag %lld scan %lld children %lld scatter+ %lld\n", r, n, levels, ph[0], ph[1], ph[2], ph[3], ph[4], ph[5], ph[6], ph[7], ph[8]);
#endif
    if (tid == 0) parity_subtree_account(ga, root, levels, err | ((n_items > 0 && err == 0) ? BUILD_ERR_DEPTH : 0));
}
void CudaExec::parity_subtrees(const BuildArrays& a, int n_roots) {
    if (n_roots <= 0) return;
    k_parity_subtrees<<<n_roots, kParSubThreads, 0, stream>>>(a, n_roots);
    note(cudaGetLastError());
    ++launches;
}

// ---- prim pass of the production builders: per-block reduction of the 12 bounds, then 12
//      atomics per block (the generic per-index body would issue 12 per warp on the same words)
__global__ void __launch_bounds__(256) k_wb_prim(const WBuild b) {
    __shared__ uint32_t red[12][8];
    uint32_t k[12];
#pragma unroll
    for (int c = 0; c < 12; ++c) k[c] = ((c / 3) % 2 == 0) ? 0xffffffffu : 0u;
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < b.n_prims; p += gridDim.x * blockDim.x) {
        F3 mn, mx;
        wb_prim_box(b, p, mn, mx);
        const F3 c = 0.5f * (mn + mx);
        const float v[12] = { mn.x, mn.y, mn.z, mx.x, mx.y, mx.z, c.x, c.y, c.z, c.x, c.y, c.z };
#pragma unroll
        for (int q = 0; q < 12; ++q) {
            const uint32_t o = ordered_key(v[q]);
            k[q] = ((q / 3) % 2 == 0) ? min(k[q], o) : max(k[q], o);
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int q = 0; q < 12; ++q) {
        const uint32_t r = ((q / 3) % 2 == 0) ? __reduce_min_sync(0xffffffffu, k[q]) : __reduce_max_sync(0xffffffffu, k[q]);
        if (lane == 0) red[q][warp] = r;
    }
    __syncthreads();
    if (threadIdx.x < 12) {
        const int q = threadIdx.x;
        uint32_t r = red[q][0];
        for (int w = 1; w < 8; ++w) r = ((q / 3) % 2 == 0) ? min(r, red[q][w]) : max(r, red[q][w]);
        if ((q / 3) % 2 == 0) atomicMin(&b.root_acc[q], r); else atomicMax(&b.root_acc[q], r);
    }
}
void CudaExec::wb_prim(const WBuild& b) {
    if (b.n_prims <= 0) return;
    const int blocks = min((b.n_prims + 255) / 256, 148 * 8);
    k_wb_prim<<<blocks, 256, 0, stream>>>(b);
    note(cudaGetLastError());
    ++launches;
}

// ---- SAH split of the production builder: one WARP per item that owns arena bins (wbvh_body.h wb_body_split is
//      the per-item serial statement the emulator runs; same bins in, same winner out: prefix / suffix unions are
//      exact min / max, the cost expression is the same three fp32 operations, and the serial "first strictly
//      smaller cost over slots 0..2, planes 0..14" is the lexicographic minimum of (cost bits, slot * 16 + plane)).
//      Lane k (and k + 16, redundantly) holds bin k of the axis slot being swept; the bins are cleared as read. -----
__device__ __forceinline__ WBox shfl_box(unsigned mask, const WBox& v, int src, bool up_down, int delta) {
    WBox r;
    if (up_down) {
        r.mn = f3(__shfl_up_sync(mask, v.mn.x, delta, 16), __shfl_up_sync(mask, v.mn.y, delta, 16), __shfl_up_sync(mask, v.mn.z, delta, 16));
        r.mx = f3(__shfl_up_sync(mask, v.mx.x, delta, 16), __shfl_up_sync(mask, v.mx.y, delta, 16), __shfl_up_sync(mask, v.mx.z, delta, 16));
    } else if (delta > 0) {
        r.mn = f3(__shfl_down_sync(mask, v.mn.x, delta, 16), __shfl_down_sync(mask, v.mn.y, delta, 16), __shfl_down_sync(mask, v.mn.z, delta, 16));
        r.mx = f3(__shfl_down_sync(mask, v.mx.x, delta, 16), __shfl_down_sync(mask, v.mx.y, delta, 16), __shfl_down_sync(mask, v.mx.z, delta, 16));
    } else {
        r.mn = f3(__shfl_sync(mask, v.mn.x, src, 16), __shfl_sync(mask, v.mn.y, src, 16), __shfl_sync(mask, v.mn.z, src, 16));
        r.mx = f3(__shfl_sync(mask, v.mx.x, src, 16), __shfl_sync(mask, v.mx.y, src, 16), __shfl_sync(mask, v.mx.z, src, 16));
    }
    return r;
}
// ---- bin pass of the level loop (wbvh_body.h wb_body_bin is the per-prim statement).  The items the level loop keeps
//      have more than kWSubMax prims, so the 256 consecutive positions of a block fall into at most two of them: the
//      block bins into shared memory and flushes its non-empty bins with one global atomic per word, instead of
//      21 global atomics per prim.  (A third item in a block -- only when the subtree queue overflowed -- goes
//      straight to global memory.)
__global__ void __launch_bounds__(256) k_wb_bin(const WBuild b) {
    __shared__ WBin s_bins[2][kWAxes * kWBins];
    __shared__ int s_w[2];
    const int tid = threadIdx.x, i = blockIdx.x * 256 + tid;
    for (int q = tid; q < 2 * kWAxes * kWBins * 8; q += 256) reinterpret_cast<uint32_t*>(&s_bins[0][0])[q] = 0u;
    if (tid < 2) s_w[tid] = -1;
    __syncthreads();
    if (i < b.n_prims) {
        const int w = b.node_of[i];
        if (w >= 0) {
            const WItem& it = b.items[w];
            if (it.axis >= 0 && it.bins_slot >= 0) {
                int slot = -1;
                int old = atomicCAS(&s_w[0], -1, w);
                if (old == -1 || old == w) {
                    slot = 0;
                } else {
                    old = atomicCAS(&s_w[1], -1, w);
                    if (old == -1 || old == w) slot = 1;
                }
                const int p = b.order[i];
                const F4 mn = b.bmin[p], mx = b.bmax[p];
#pragma unroll
                for (int a = 0; a < kWAxes; ++a) {
                    const int k = a * kWBins + wb_bin_of(it, wb_slot_axis(it, a), mn, mx);
                    WBin& bin = slot >= 0 ? s_bins[slot][k] : b.bins[it.bins_slot].b[k];
                    // lanes of the warp that land in the same bin combine first (match.any + redux.sync): one atomic per word and group
                    const CvGroup g = cv_group_by(slot >= 0 ? slot * (kWAxes * kWBins) + k : 2 * kWAxes * kWBins + it.bins_slot * (kWAxes * kWBins) + k);
                    cv_group_add_i32(g, reinterpret_cast<int*>(&bin.count), 1);
                    cv_group_max_u32(g, &bin.nmn[0], ~ordered_key(mn.x));
                    cv_group_max_u32(g, &bin.nmn[1], ~ordered_key(mn.y));
                    cv_group_max_u32(g, &bin.nmn[2], ~ordered_key(mn.z));
                    cv_group_max_u32(g, &bin.mx[0], ordered_key(mx.x));
                    cv_group_max_u32(g, &bin.mx[1], ordered_key(mx.y));
                    cv_group_max_u32(g, &bin.mx[2], ordered_key(mx.z));
                }
            }
        }
    }
    __syncthreads();
    if (tid < 2 * kWAxes * kWBins) {
        const int slot = tid / (kWAxes * kWBins), k = tid % (kWAxes * kWBins);
        const int w = s_w[slot];
        const WBin v = s_bins[slot][k];
        if (w >= 0 && v.count > 0u) {
            WBin& bin = b.bins[b.items[w].bins_slot].b[k];
            atomicAdd(&bin.count, v.count);
            atomicMax(&bin.nmn[0], v.nmn[0]);
            atomicMax(&bin.nmn[1], v.nmn[1]);
            atomicMax(&bin.nmn[2], v.nmn[2]);
            atomicMax(&bin.mx[0], v.mx[0]);
            atomicMax(&bin.mx[1], v.mx[1]);
            atomicMax(&bin.mx[2], v.mx[2]);
        }
    }
}
void CudaExec::wb_bin(const WBuild& b) {
    if (b.n_prims <= 0) return;
    k_wb_bin<<<(b.n_prims + 255) / 256, 256, 0, stream>>>(b);
    note(cudaGetLastError());
    ++launches;
}

// The winner of one item's bins: best = (cost bits << 8) | (slot * 16 + plane), ~0 when no plane has two non-empty sides;
// box / cnt = the two sides of the winning plane.  All lanes return the same values.
struct WSweep {
    unsigned long long best;
    WBox box[2];
    int cnt[2];
};
template<bool kClear>
__device__ __forceinline__ WSweep wb_warp_sweep(WBin* item_bins, int lane) {
    constexpr unsigned kAll = 0xffffffffu;
    const int k = lane & 15;
    WSweep r;
    r.best = ~0ull;
    r.box[0] = r.box[1] = wb_empty();
    r.cnt[0] = r.cnt[1] = 0;
    for (int a = 0; a < kWAxes; ++a) {
        WBin* src = &item_bins[a * kWBins + k];
        const WBin raw = *src;
        if (kClear) {
            __syncwarp();
            if (lane < 16) {
                WBin z = {};
                *src = z;
            }
        }
        const WLBin mine = wb_lbin_of(raw);
        int lc = mine.count, rc = mine.count;
        WBox lb = mine.box, rb = mine.box;
#pragma unroll
        for (int d = 1; d < 16; d <<= 1) {
            const int cu = __shfl_up_sync(kAll, lc, d, 16), cd = __shfl_down_sync(kAll, rc, d, 16);
            const WBox bu = shfl_box(kAll, lb, 0, true, d), bd = shfl_box(kAll, rb, 0, false, d);
            if (k >= d) {
                lc += cu;
                wb_grow(lb, bu.mn, bu.mx);
            }
            if (k + d < 16) {
                rc += cd;
                wb_grow(rb, bd.mn, bd.mx);
            }
        }
        const float ra = rc > 0 ? wb_half_area(rb) : 0.0f;
        const int rn = __shfl_down_sync(kAll, rc, 1, 16);          // suffix count / area of bins k + 1 .. 15
        const float rarea = __shfl_down_sync(kAll, ra, 1, 16);
        const bool valid = k < kWBins - 1 && lc > 0 && rn > 0;
        const float cost = (float)lc * wb_half_area(lb) + (float)rn * rarea;
        unsigned long long key = valid ? (((unsigned long long)f2u(cost)) << 8) | (unsigned long long)(a * kWBins + k) : ~0ull;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            const unsigned long long o = __shfl_xor_sync(kAll, key, d);
            key = o < key ? o : key;
        }
        if (key < r.best) {  // warp-uniform
            r.best = key;
            const int kk = (int)(key & 15ull);
            r.cnt[0] = __shfl_sync(kAll, lc, kk, 16);
            r.box[0] = shfl_box(kAll, lb, kk, false, 0);
            r.cnt[1] = __shfl_sync(kAll, rc, kk + 1, 16);
            r.box[1] = shfl_box(kAll, rb, kk + 1, false, 0);
        }
    }
    return r;
}
__device__ __noinline__ void wb_split_serial(const WBuild& b, int w, int level) { wb_body_split(b, w, level, true); }
__global__ void __launch_bounds__(256) k_wb_split_binned(const WBuild b, int level) {
    const int w = (int)((blockIdx.x * 256u + threadIdx.x) >> 5);
    const int lane = threadIdx.x & 31;
    if (w >= b.level_items[level]) return;  // warp-uniform
    WItem& it = b.items[w];
    if (it.axis < 0 || it.bins_slot < 0) {
        // a median split, or an item of <= kWDirectMax prims (in the level loop only when the subtree queue overflowed):
        // the per-item serial statement, on one lane
        if (lane == 0) wb_split_serial(b, w, level);
        return;
    }
    WSweep r = wb_warp_sweep<true>(&b.bins[it.bins_slot].b[0], lane);
    if (lane == 0) {
        if (r.best == ~0ull) {
            it.axis = -1;  // every prim in one bin on every axis: median split
        } else {
            const int key = (int)(r.best & 0xffull);
            it.axis = wb_slot_axis(it, key / kWBins);
            it.split = key % kWBins;
        }
        wb_finish_split(b, it, level, r.box, r.cnt);
    }
}
void CudaExec::wb_split(const WBuild& b, int level, int bound) {
    if (bound <= 0) return;
    k_wb_split_binned<<<(bound * 32 + 255) / 256, 256, 0, stream>>>(b, level);  // one warp per item of the level
    note(cudaGetLastError());
    ++launches;
}

// ---- partition of a level: left flags computed inside the scan's first kernel, the stable scatter inside its last
//      (the emulator runs wb_body_flag / scan / wb_body_scatter).  The scatter needs the scan value at the START of a
//      prim's item, which may lie in an earlier tile: the first kernel leaves (tile, count within the tile) in the item.
__global__ void __launch_bounds__(kScanThreads) k_wb_part_sums(const WBuild b, int* __restrict__ sums) {
    const int n = b.n_prims;
    const int base = blockIdx.x * kScanTile + threadIdx.x * kScanPerThread;
    unsigned char f[kScanPerThread];
    int s = 0;
#pragma unroll
    for (int k = 0; k < kScanPerThread; ++k) {
        const int i = base + k;
        f[k] = i < n ? wb_flag_of(b, i) : 0;
        if (i < n) b.flags[i] = f[k];
        s += f[k];
    }
    int total;
    int run = block_exclusive_scan(s, &total);
    if (threadIdx.x == 0) sums[blockIdx.x] = total;
#pragma unroll
    for (int k = 0; k < kScanPerThread; ++k) {
        const int i = base + k;
        if (i < n) {
            const int w = b.node_of[i];
            if (w >= 0 && b.items[w].start == i) {
                b.items[w].scan_tile = blockIdx.x;
                b.items[w].scan_local = run;
            }
        }
        run += f[k];
    }
}
__global__ void __launch_bounds__(kScanThreads) k_wb_part_scatter(const WBuild b, const int* __restrict__ sums) {
    const int n = b.n_prims;
    const int base = blockIdx.x * kScanTile + threadIdx.x * kScanPerThread;
    unsigned char f[kScanPerThread];
    int s = 0;
#pragma unroll
    for (int k = 0; k < kScanPerThread; ++k) {
        const int i = base + k;
        f[k] = i < n ? b.flags[i] : 0;
        s += f[k];
    }
    int total;
    int run = block_exclusive_scan(s, &total) + sums[blockIdx.x];
#pragma unroll
    for (int k = 0; k < kScanPerThread; ++k) {
        const int i = base + k;
        if (i < n) {
            const int w = b.node_of[i];
            if (w < 0) {
                b.order_next[i] = b.order[i];
                b.node_of_next[i] = -1;
            } else {
                const WItem& it = b.items[w];
                wb_scatter_to(b, i, w, f[k], run, sums[it.scan_tile] + it.scan_local);
            }
        }
        run += f[k];
    }
}
void CudaExec::wb_partition(const WBuild& b) {
    if (b.n_prims <= 0) return;
    const int nb = (b.n_prims + kScanTile - 1) / kScanTile;
    k_wb_part_sums<<<nb, kScanThreads, 0, stream>>>(b, scan_tmp);
    k_scan_sums<<<1, kScanThreads, 0, stream>>>(scan_tmp, nb);
    k_wb_part_scatter<<<nb, kScanThreads, 0, stream>>>(b, scan_tmp);
    note(cudaGetLastError());
    launches += 3;
}

// ---- subtree phase of the SAH builder: one block per queued item of <= kWSubMax prims builds the item's whole
//      subtree out of SHARED memory (wbvh_body.h wb_subtree_serial is the serial statement the emulator runs; the
//      tree is the same multiset of records: same bins, same sweep, same stable partition).  The block loads the
//      item's prim boxes once; per level every WARP takes items of the level in turn and does all of an item's
//      work itself -- bins in its own 1.5 KB of shared memory (shared-memory atomics), the warp-wide sweep, the two
//      children, the stable partition by ballot prefix -- so a level costs one block barrier and no global traffic
//      beyond the sibling pair it emits.
#ifndef CV_SUB_THREADS
#define CV_SUB_THREADS 256
#endif
#ifndef CV_SUB_MINBLOCKS
#define CV_SUB_MINBLOCKS 3
#endif
constexpr int kSubThreads = CV_SUB_THREADS;
constexpr int kSubWarps = kSubThreads / 32;
constexpr int kSubItems = kWSubMax / (kFastLeafMax + 1) + 1;
struct SItem {
    unsigned short start, count;  // positions relative to the subtree's first position
    int pair;
    short child[2];               // index of each child in the next level, -1: leaf
    unsigned short n_left;
    signed char axis;             // as WItem::axis
    unsigned char split;
    unsigned char depth;
    unsigned char spare[3];
    float box_mn[3], box_mx[3], cscale[3];
};
struct SubShared {
    F4 pmn[kWSubMax], pmx[kWSubMax];        // prim boxes by local prim id
    int gid[kWSubMax];                      // the prim's global index (what `order` holds)
    unsigned short ord[2][kWSubMax];        // local prim id at each position, ping-pong
    short item_of[2][kWSubMax];             // item of each position (-1: settled in a leaf)
    SItem items[2][kSubItems];
    WBin bins[kSubWarps][kWAxes * kWBins];  // one set per warp
    int n_items[2];
    int lv_items[kWSubLevels + 2];
    int error;
    int pair_next;  // hybrid builder (WItem::pair_lo >= 0): the next free slot of the subtree's own pair range; -1: pairs come from the global counter
};
__device__ __forceinline__ int sub_slot_axis(int item_axis, int a) { return kWAxes == 1 ? item_axis : a; }

// One thread: the item's two children (leaf references or items of the next level) and its sibling pair.
__device__ __forceinline__ void sub_emit_children(SubShared& S, const WBuild& gb, const SItem& it, int nxt, int start, const WBox box[2], const int cnt[2], int child[2]) {
    const int s0 = it.start;
    for (int s = 0; s < 2; ++s) {
        const int cs = s == 0 ? s0 : s0 + cnt[0];
        int ref;
        child[s] = -1;
        if (cnt[s] <= kFastLeafMax) {
            ref = make_leaf_ref(start + cs, cnt[s]);
        } else {
            const int pair = S.pair_next >= 0 ? atomicAdd(&S.pair_next, 1) : atomicAdd(&gb.counters[1], 1);
            const int slot = atomicAdd(&S.n_items[nxt], 1);
            if (pair >= gb.pairs_capacity || slot >= kSubItems) {
                atomicOr(&gb.counters[2], WB_ERR_CAPACITY);
                ref = make_leaf_ref(start + cs, 1);
            } else {
                ref = 2 * pair;
                gb.pair_depth[pair] = (unsigned char)(it.depth + 1);
                child[s] = slot;
                SItem ch;
                ch.start = (unsigned short)cs;
                ch.count = (unsigned short)cnt[s];
                ch.pair = pair;
                ch.child[0] = ch.child[1] = -1;
                ch.n_left = 0;
                ch.split = 0;
                ch.depth = (unsigned char)(it.depth + 1);
                ch.box_mn[0] = box[s].mn.x; ch.box_mn[1] = box[s].mn.y; ch.box_mn[2] = box[s].mn.z;
                ch.box_mx[0] = box[s].mx.x; ch.box_mx[1] = box[s].mx.y; ch.box_mx[2] = box[s].mx.z;
                ch.axis = (signed char)wb_binning_setup(ch.box_mn, ch.box_mx, ch.depth, gb.force_median, ch.cscale);
                S.items[nxt][slot] = ch;
            }
        }
        FastNode o;
        o.mn[0] = box[s].mn.x; o.mn[1] = box[s].mn.y; o.mn[2] = box[s].mn.z;
        o.mx[0] = box[s].mx.x; o.mx[1] = box[s].mx.y; o.mx[2] = box[s].mx.z;
        o.a = ref;
        o.b = cnt[s];
        gb.fnodes[2 * it.pair + s] = o;
    }
}
__device__ __forceinline__ void sub_median(const SItem& it, WBox box[2], int cnt[2]) {
    cnt[0] = it.count / 2;
    cnt[1] = it.count - cnt[0];
    for (int s = 0; s < 2; ++s) {
        box[s].mn = f3(it.box_mn[0], it.box_mn[1], it.box_mn[2]);
        box[s].mx = f3(it.box_mx[0], it.box_mx[1], it.box_mx[2]);
    }
}
// A small item (<= kSubSerialMax prims), all of it in ONE thread, registers only.  Same winner as the binned sweep:
// the planes of an axis that separate the prims differently are exactly k = bin of some prim (k < 15) -- every other
// plane repeats the partition of the largest such k below it at a higher index, which the sweep's "first strictly
// smaller cost" never prefers -- so the thread evaluates those <= count candidates per axis directly from the prims
// (count^2 box unions per axis, boxes read from shared memory) and keeps the lexicographic minimum of (cost, axis
// slot * 16 + k).  Unions are exact min / max, the cost is the same three fp32 operations: same bits as the sweep.
#ifndef CV_SUB_SERIAL_SWEEP
#define CV_SUB_SERIAL_SWEEP 1
#endif
#ifndef CV_SUB_HALF_ITEMS
#define CV_SUB_HALF_ITEMS 1
#endif
constexpr int kSubSerialMax = 16;
__device__ void sub_item_serial(SubShared& S, const WBuild& gb, int cur, int nxt, int w, int start) {
    const SItem& it = S.items[cur][w];
    const int s0 = it.start, c = it.count;
    int axis = it.axis, split = 0;
    WBox box[2];
    int cnt[2] = { 0, 0 };
    box[0] = box[1] = wb_empty();
    if (axis >= 0) {
        float best_cost = CV_INF;
        int best_key = -1;
        for (int a = 0; a < kWAxes; ++a) {
            const int ax = sub_slot_axis(axis, a);
            unsigned long long packed = 0ull;  // 4 bits per prim: its bin on this axis
            for (int j = 0; j < c; ++j) {
                const int id = S.ord[cur][s0 + j];
                packed |= (unsigned long long)wb_bin_at(it.box_mn, it.cscale, ax, S.pmn[id], S.pmx[id]) << (4 * j);
            }
#if CV_SUB_SERIAL_SWEEP
            // The same candidates at O(count) unions per axis instead of count^2: the prims of bin k are found by a
            // nibble-equality mask over `packed`; a descending pass leaves the half area of "every prim of a bin > k"
            // for each k, an ascending pass grows "every prim of a bin <= k" and prices plane k wherever some prim has
            // bin k.  Unions are exact min / max (order-free), the cost is the same expression on the same boxes, and
            // planes come in ascending (axis slot, k) order with a strict "<": the same lexicographic minimum.
            const unsigned long long ones = 0x1111111111111111ull, live = c >= 16 ? ~0ull : ((1ull << (4 * c)) - 1ull);
            float right_area[kWBins];
            {
                WBox r = wb_empty();
#pragma unroll
                for (int k = kWBins - 1; k >= 1; --k) {
                    unsigned long long x = packed ^ (ones * (unsigned long long)k);
                    x |= x >> 1;
                    x |= x >> 2;
                    for (unsigned long long m = ~x & ones & live; m; m &= m - 1) {
                        const int id = S.ord[cur][s0 + ((__ffsll((long long)m) - 1) >> 2)];
                        const F4 mn = S.pmn[id], mx = S.pmx[id];
                        wb_grow(r, f3(mn.x, mn.y, mn.z), f3(mx.x, mx.y, mx.z));
                    }
                    right_area[k] = wb_half_area(r);  // read only where the right side is not empty
                }
            }
            {
                WBox l = wb_empty();
                int nl = 0;
#pragma unroll
                for (int k = 0; k < kWBins - 1; ++k) {
                    unsigned long long x = packed ^ (ones * (unsigned long long)k);
                    x |= x >> 1;
                    x |= x >> 2;
                    unsigned long long m = ~x & ones & live;
                    if (m == 0ull) continue;  // no prim has bin k: not a candidate
                    for (; m; m &= m - 1) {
                        const int id = S.ord[cur][s0 + ((__ffsll((long long)m) - 1) >> 2)];
                        const F4 mn = S.pmn[id], mx = S.pmx[id];
                        wb_grow(l, f3(mn.x, mn.y, mn.z), f3(mx.x, mx.y, mx.z));
                        ++nl;
                    }
                    if (nl == c) continue;  // nothing on the right
                    const float cost = (float)nl * wb_half_area(l) + (float)(c - nl) * right_area[k + 1];
                    if (cost < best_cost) {
                        best_cost = cost;
                        best_key = a * kWBins + k;
                    }
                }
            }
#else
            for (int j = 0; j < c; ++j) {
                const int k = (int)((packed >> (4 * j)) & 15ull);
                if (k >= kWBins - 1) continue;
                const int key = a * kWBins + k;
                WBox l = wb_empty(), r = wb_empty();
                int nl = 0;
                for (int i = 0; i < c; ++i) {
                    const int id = S.ord[cur][s0 + i];
                    const F4 mn = S.pmn[id], mx = S.pmx[id];
                    if ((int)((packed >> (4 * i)) & 15ull) <= k) {
                        wb_grow(l, f3(mn.x, mn.y, mn.z), f3(mx.x, mx.y, mx.z));
                        ++nl;
                    } else {
                        wb_grow(r, f3(mn.x, mn.y, mn.z), f3(mx.x, mx.y, mx.z));
                    }
                }
                if (nl == c) continue;  // nothing on the right
                const float cost = (float)nl * wb_half_area(l) + (float)(c - nl) * wb_half_area(r);
                if (cost < best_cost || (cost == best_cost && key < best_key)) {
                    best_cost = cost;
                    best_key = key;
                    box[0] = l;
                    box[1] = r;
                    cnt[0] = nl;
                    cnt[1] = c - nl;
                }
            }
#endif
        }
        if (best_key < 0) {
            axis = -1;
        } else {
            axis = sub_slot_axis(axis, best_key / kWBins);
            split = best_key % kWBins;
#if CV_SUB_SERIAL_SWEEP
            // the two sides of the winning plane (the sweep kept areas only)
            for (int j = 0; j < c; ++j) {
                const int id = S.ord[cur][s0 + j];
                const F4 mn = S.pmn[id], mx = S.pmx[id];
                const int side = wb_bin_at(it.box_mn, it.cscale, axis, mn, mx) <= split ? 0 : 1;
                wb_grow(box[side], f3(mn.x, mn.y, mn.z), f3(mx.x, mx.y, mx.z));
                ++cnt[side];
            }
#endif
        }
    }
    if (axis < 0) sub_median(it, box, cnt);
    int child[2];
    sub_emit_children(S, gb, it, nxt, start, box, cnt, child);
    int left_run = 0, right_run = 0;
    for (int j = 0; j < c; ++j) {
        const int id = S.ord[cur][s0 + j];
        const bool flag = axis >= 0 ? wb_bin_at(it.box_mn, it.cscale, axis, S.pmn[id], S.pmx[id]) <= split : j < cnt[0];
        const int dst = flag ? s0 + left_run++ : s0 + cnt[0] + right_run++;
        S.ord[nxt][dst] = (unsigned short)id;
        S.item_of[nxt][dst] = (short)(flag ? child[0] : child[1]);
    }
}
// The same small item on a GROUP of G = 4 / 8 / 16 adjacent lanes (all G call it; G >= count): lane j < count holds prim j and prices the
// plane "bin of prim j" on each axis over all the item's prims -- count unions per lane instead of count^2 in one thread,
// every lane in step (the boxes are read from shared memory at one address: a broadcast).  The half warp then takes the
// lexicographic minimum of (cost, axis slot * 16 + k) by shuffles -- the order sub_item_serial's "cost < best || (cost ==
// best && key < best_key)" realises -- and the winning lane's two boxes and counts go to every lane; lane 0 emits the
// children; the stable partition is a ballot prefix over the sixteen lanes.  Same candidates, same unions (exact min /
// max), same cost expression: the same tree.
template<int G>
__device__ __forceinline__ void sub_item_group(SubShared& S, const WBuild& gb, int cur, int nxt, int w, int start, int lane) {
    const int sub = lane & (G - 1), half = lane & ~(G - 1);  // `half`: the group's first lane
    const unsigned hmask = ((1u << G) - 1u) << half;
    const SItem& it = S.items[cur][w];
    const int s0 = it.start, c = it.count;
    const bool mine = sub < c;
    int axis = it.axis, split = 0;
    WBox box[2];
    int cnt[2] = { 0, 0 };
    box[0] = box[1] = wb_empty();
    const int my_id = mine ? S.ord[cur][s0 + sub] : 0;
    int my_bin[kWAxes];
#pragma unroll
    for (int a = 0; a < kWAxes; ++a) my_bin[a] = 0;
    if (axis >= 0) {
        const F4 pmn = S.pmn[my_id], pmx = S.pmx[my_id];
        float best_cost = CV_INF;
        int best_key = 0x7fffffff, best_nl = 0;
        WBox best_l = wb_empty(), best_r = wb_empty();
#pragma unroll
        for (int a = 0; a < kWAxes; ++a) {
            const int ax = sub_slot_axis(axis, a);
            const int k = mine ? wb_bin_at(it.box_mn, it.cscale, ax, pmn, pmx) : kWBins;  // a lane without a prim prices nothing
            my_bin[a] = k;
            WBox l = wb_empty(), r = wb_empty();
            int nl = 0;
            for (int i = 0; i < c; ++i) {
                const int ki = __shfl_sync(hmask, k, half + i);
                const int id = S.ord[cur][s0 + i];
                const F4 mn = S.pmn[id], mx = S.pmx[id];
                if (ki <= k) {
                    wb_grow(l, f3(mn.x, mn.y, mn.z), f3(mx.x, mx.y, mx.z));
                    ++nl;
                } else {
                    wb_grow(r, f3(mn.x, mn.y, mn.z), f3(mx.x, mx.y, mx.z));
                }
            }
            if (mine && k < kWBins - 1 && nl != c) {
                const float cost = (float)nl * wb_half_area(l) + (float)(c - nl) * wb_half_area(r);
                const int key = a * kWBins + k;
                if (cost < best_cost || (cost == best_cost && key < best_key)) {
                    best_cost = cost;
                    best_key = key;
                    best_nl = nl;
                    best_l = l;
                    best_r = r;
                }
            }
        }
        // the half warp's minimum of (cost, key); then everything of the lane that holds it
        float m_cost = best_cost;
        int m_key = best_key;
#pragma unroll
        for (int d = G / 2; d >= 1; d >>= 1) {
            const float oc = __shfl_xor_sync(hmask, m_cost, d);
            const int ok = __shfl_xor_sync(hmask, m_key, d);
            if (oc < m_cost || (oc == m_cost && ok < m_key)) {
                m_cost = oc;
                m_key = ok;
            }
        }
        if (m_key == 0x7fffffff) {
            axis = -1;  // no plane separates the prims: median split
        } else {
            const unsigned holders = __ballot_sync(hmask, best_cost == m_cost && best_key == m_key) & hmask;
            const int src = __ffs(holders) - 1;  // lane index within the warp
            box[0].mn = f3(__shfl_sync(hmask, best_l.mn.x, src), __shfl_sync(hmask, best_l.mn.y, src), __shfl_sync(hmask, best_l.mn.z, src));
            box[0].mx = f3(__shfl_sync(hmask, best_l.mx.x, src), __shfl_sync(hmask, best_l.mx.y, src), __shfl_sync(hmask, best_l.mx.z, src));
            box[1].mn = f3(__shfl_sync(hmask, best_r.mn.x, src), __shfl_sync(hmask, best_r.mn.y, src), __shfl_sync(hmask, best_r.mn.z, src));
            box[1].mx = f3(__shfl_sync(hmask, best_r.mx.x, src), __shfl_sync(hmask, best_r.mx.y, src), __shfl_sync(hmask, best_r.mx.z, src));
            cnt[0] = __shfl_sync(hmask, best_nl, src);
            cnt[1] = c - cnt[0];
            axis = sub_slot_axis(axis, m_key / kWBins);
            split = m_key % kWBins;
        }
    }
    if (axis < 0) sub_median(it, box, cnt);
    int child[2] = { -1, -1 };
    if (sub == 0) sub_emit_children(S, gb, it, nxt, start, box, cnt, child);
    child[0] = __shfl_sync(hmask, child[0], half);
    child[1] = __shfl_sync(hmask, child[1], half);
    // stable partition: lefts keep their order, rights keep theirs
    bool flag = false;
    if (mine) {
        if (axis >= 0) {
            int kb = my_bin[0];
#pragma unroll
            for (int a = 1; a < kWAxes; ++a) kb = (kWAxes > 1 && axis == a) ? my_bin[a] : kb;
            flag = kb <= split;
        } else {
            flag = sub < cnt[0];
        }
    }
    const unsigned gbits = (1u << G) - 1u;
    const unsigned lefts = (__ballot_sync(hmask, flag) >> half) & gbits, valid = (__ballot_sync(hmask, mine) >> half) & gbits;
    if (mine) {
        const unsigned lt = (1u << sub) - 1u;
        const int dst = flag ? s0 + __popc(lefts & lt) : s0 + cnt[0] + __popc(~lefts & valid & lt);
        S.ord[nxt][dst] = (unsigned short)my_id;
        S.item_of[nxt][dst] = (short)(flag ? child[0] : child[1]);
    }
}
// A larger item, all of it in ONE warp: bins in the warp's own shared memory, the warp-wide sweep, partition by ballot prefix.
__device__ void sub_item_warp(SubShared& S, const WBuild& gb, int cur, int nxt, int w, int start, int warp, int lane, long long* prof) {
#ifdef CV_SUB_PROFILE
    long long t0_ = clock64();
#endif
    const SItem& it = S.items[cur][w];
    WBin* wb = S.bins[warp];
    const int s0 = it.start, c = it.count;
    int axis = it.axis, split = 0;
    WBox box[2];
    int cnt[2] = { 0, 0 };
    box[0] = box[1] = wb_empty();
    if (axis >= 0) {
        for (int q = lane; q < kWAxes * kWBins * 8; q += 32) reinterpret_cast<uint32_t*>(wb)[q] = 0u;
        __syncwarp();
        for (int j = lane; j < c; j += 32) {
            const int id = S.ord[cur][s0 + j];
            const F4 mn = S.pmn[id], mx = S.pmx[id];
#pragma unroll
            for (int a = 0; a < kWAxes; ++a) {
                WBin& bin = wb[a * kWBins + wb_bin_at(it.box_mn, it.cscale, sub_slot_axis(axis, a), mn, mx)];
                atomicAdd(&bin.count, 1u);
                atomicMax(&bin.nmn[0], ~ordered_key(mn.x));
                atomicMax(&bin.nmn[1], ~ordered_key(mn.y));
                atomicMax(&bin.nmn[2], ~ordered_key(mn.z));
                atomicMax(&bin.mx[0], ordered_key(mx.x));
                atomicMax(&bin.mx[1], ordered_key(mx.y));
                atomicMax(&bin.mx[2], ordered_key(mx.z));
            }
        }
        __syncwarp();
        SUBT(0)
        const WSweep sw = wb_warp_sweep<false>(wb, lane);
        __syncwarp();
        SUBT(1)
        if (sw.best == ~0ull) {
            axis = -1;  // every prim in one bin on every axis: median split
        } else {
            const int key = (int)(sw.best & 0xffull);
            axis = sub_slot_axis(axis, key / kWBins);
            split = key % kWBins;
            box[0] = sw.box[0];
            box[1] = sw.box[1];
            cnt[0] = sw.cnt[0];
            cnt[1] = sw.cnt[1];
        }
    }
    if (axis < 0) sub_median(it, box, cnt);
    int child[2] = { -1, -1 };
    if (lane == 0) sub_emit_children(S, gb, it, nxt, start, box, cnt, child);
    child[0] = __shfl_sync(0xffffffffu, child[0], 0);
    child[1] = __shfl_sync(0xffffffffu, child[1], 0);
    SUBT(2)
    // stable partition of the item's positions: lefts keep their order, rights keep theirs
    int left_run = 0, right_run = 0;
    for (int base = 0; base < c; base += 32) {
        const int j = base + lane;
        const bool valid = j < c;
        const int id = valid ? S.ord[cur][s0 + j] : 0;
        bool flag;
        if (axis >= 0) flag = valid && wb_bin_at(it.box_mn, it.cscale, axis, S.pmn[id], S.pmx[id]) <= split;
        else flag = valid && j < cnt[0];
        const unsigned m = __ballot_sync(0xffffffffu, flag), mv = __ballot_sync(0xffffffffu, valid);
        const unsigned lt = (1u << lane) - 1u;
        if (valid) {
            const int dst = flag ? s0 + left_run + __popc(m & lt) : s0 + cnt[0] + right_run + __popc(~m & mv & lt);
            S.ord[nxt][dst] = (unsigned short)id;
            S.item_of[nxt][dst] = (short)(flag ? child[0] : child[1]);
        }
        left_run += __popc(m);
        right_run += __popc(~m & mv);
    }
    __syncwarp();
    SUBT(3)
}

__global__ void __launch_bounds__(kSubThreads, CV_SUB_MINBLOCKS) k_wb_subtrees(const WBuild gb, int n_roots) {
    extern __shared__ __align__(32) unsigned char s_raw[];
    SubShared& S = *reinterpret_cast<SubShared*>(s_raw);
    const int r = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (r >= n_roots) return;
    const WItem root = gb.sub_roots[r];
    const int n = root.count, start = root.start;
    for (int j = tid; j < n; j += kSubThreads) {
        const int p = gb.order[start + j];
        S.gid[j] = p;
        S.pmn[j] = gb.bmin[p];
        S.pmx[j] = gb.bmax[p];
        S.ord[0][j] = (unsigned short)j;
        S.item_of[0][j] = 0;
    }
    if (tid == 0) {
        SItem it;
        it.start = 0;
        it.count = (unsigned short)n;
        it.pair = root.pair;
        it.child[0] = it.child[1] = -1;
        it.n_left = 0;
        it.axis = (signed char)root.axis;
        it.split = 0;
        it.depth = (unsigned char)root.depth;
        for (int c = 0; c < 3; ++c) {
            it.box_mn[c] = root.box_mn[c];
            it.box_mx[c] = root.box_mx[c];
            it.cscale[c] = root.cscale[c];
        }
        S.items[0][0] = it;
        S.n_items[0] = 1;
        S.n_items[1] = 0;
        S.error = 0;
        S.pair_next = root.pair_lo;
    }
    for (int l = tid; l < kWSubLevels + 2; l += kSubThreads) S.lv_items[l] = 0;
    __syncthreads();
    int cur = 0, level = 0;
    long long prof[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
#ifdef CV_SUB_PROFILE
    long long t0_ = clock64();
    const long long tstart_ = t0_;
#endif
    for (;; ++level, cur ^= 1) {
        const int ni = S.n_items[cur];
        if (ni == 0) break;
        if (level >= kWSubLevels) {
            if (tid == 0) S.error = 1;
            break;
        }
        const int nxt = cur ^ 1;
        if (tid == 0) S.lv_items[level] = ni;
        // positions already settled in a leaf belong to no item: carry them over
        for (int j = tid; j < n; j += kSubThreads) {
            if (S.item_of[cur][j] < 0) {
                S.ord[nxt][j] = S.ord[cur][j];
                S.item_of[nxt][j] = -1;
            }
        }
        for (int w = warp; w < ni; w += kSubWarps) {  // warp-uniform
            if (S.items[cur][w].count > kSubSerialMax) sub_item_warp(S, gb, cur, nxt, w, start, warp, lane, prof);
        }
        SUBT(4)
#if CV_SUB_HALF_ITEMS
        // small items: sixteen lanes each (sub_item_group<16>); one thread per item (sub_item_serial) waits on one lane's chain
        // of count^2 unions.  (Groups of 8 and 4 lanes for items of <= 8 / <= 4 prims, three passes: measured 2-18 % slower.)
        for (int w = tid >> 4; w < ni; w += kSubThreads / 16) {  // uniform over a half warp
            if (S.items[cur][w].count <= kSubSerialMax) sub_item_group<16>(S, gb, cur, nxt, w, start, lane);
        }
#else
        for (int w = tid; w < ni; w += kSubThreads) {
            if (S.items[cur][w].count <= kSubSerialMax) sub_item_serial(S, gb, cur, nxt, w, start);
        }
#endif
        SUBT(5)
        __syncthreads();
        SUBT(6)
        if (tid == 0) S.n_items[cur] = 0;  // becomes the level after next
        __syncthreads();
    }
    __syncthreads();
#ifdef CV_SUB_PROFILE
    if ((r % 97) == 0 && (tid == 0 || tid == 224)) printf("root %d n %d levels %d warp %d total %lld | in warp items: bin %lld sweep %lld emit %lld partition %lld | loop: warp-items(+carry) %lld serial-items %lld barrier %lld\n", r, n, level, warp, clock64() - tstart_, prof[0], prof[1], prof[2], prof[3], prof[4], prof[5], prof[6]);
#endif
    for (int j = tid; j < n; j += kSubThreads) gb.order[start + j] = S.gid[S.ord[cur][j]];
    if (tid == 0) {
        if (S.error) atomicOr(&gb.counters[2], WB_ERR_CAPACITY);
        wb_subtree_account(gb, root, level, S.lv_items);
    }
}
void CudaExec::wb_subtrees(const WBuild& b, int n_roots) {
    if (n_roots <= 0) return;
    note(cudaFuncSetAttribute(k_wb_subtrees, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SubShared)));
    k_wb_subtrees<<<n_roots, kSubThreads, sizeof(SubShared), stream>>>(b, n_roots);
    note(cudaGetLastError());
    ++launches;
}

// ---- scene assembly -------------------------------------------------------------------
struct FnAssembleVertices {
    const float* positions;
    const float* normals;
    GpuPtVertex* out;
    __device__ void operator()(int i) const {
        GpuPtVertex v;
        v.position[0] = positions[3 * i];
        v.position[1] = positions[3 * i + 1];
        v.position[2] = positions[3 * i + 2];
        v.pad1 = 0;
        v.normal[0] = normals[3 * i];
        v.normal[1] = normals[3 * i + 1];
        v.normal[2] = normals[3 * i + 2];
        v.pad2 = 0;
        out[i] = v;
    }
};
struct FnAssembleIndices {
    const uint32_t* raw;
    const int* tri_off;
    const int* vert_off;
    int n_meshes;
    GpuPtIndex* out;
    __device__ void operator()(int i) const {
        int lo = 0, hi = n_meshes;
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (tri_off[mid] <= i) lo = mid; else hi = mid;
        }
        const int off = vert_off[lo];  // AppendIndices: dest.tri = source.tri + p_vertex_count
        GpuPtIndex o;
        o.tri[0] = (int)raw[3 * i] + off;
        o.tri[1] = (int)raw[3 * i + 1] + off;
        o.tri[2] = (int)raw[3 * i + 2] + off;
        o.pad1 = 0;
        out[i] = o;
    }
};
void launch_assemble_vertices(CudaExec& ex, const float* positions, const float* normals, GpuPtVertex* out, int n_vertices) {
    ex.launch(n_vertices, FnAssembleVertices{ positions, normals, out });
}
void launch_assemble_indices(CudaExec& ex, const uint32_t* raw_indices, const int* tri_off, const int* vert_off, int n_meshes, GpuPtIndex* out, int n_tris) {
    ex.launch(n_tris, FnAssembleIndices{ raw_indices, tri_off, vert_off, n_meshes, out });
}

int run_build_bvh_parity(CudaExec& ex, BuildArrays a, const int* tri_off_host, const int* d_tri_off, int n_meshes, BuildItem* host_scratch, BuildReport* report) {
    return build_bvh_parity(ex, a, tri_off_host, d_tri_off, n_meshes, host_scratch, report);
}
// per FastTri slot: the texture coordinates of its triangle's vertices, in the slot's vertex order (wb_tri_box / wb_emit_slot
// swap B and C for a mirroring instance)
struct FnEmitUvs {
    const FastTri* ftris;
    const GpuPtIndex* indices;
    const GpuPtMesh* instances;
    const float* uvs;  // 2 floats per scene-global vertex
    FastUv* out;
    __device__ void operator()(int s) const {
        const int tri = (int)f2u(ftris[s].v1.w), inst = (int)f2u(ftris[s].v0.w);
        const float* m = instances[inst].transform;
        const float det = m[0] * (m[5] * m[10] - m[9] * m[6]) - m[4] * (m[1] * m[10] - m[9] * m[2]) + m[8] * (m[1] * m[6] - m[5] * m[2]);
        const int* idx = indices[tri].tri;
        const int i0 = idx[0], i1 = det < 0.0f ? idx[2] : idx[1], i2 = det < 0.0f ? idx[1] : idx[2];
        FastUv o;
        o.uv[0] = uvs[2 * i0]; o.uv[1] = uvs[2 * i0 + 1];
        o.uv[2] = uvs[2 * i1]; o.uv[3] = uvs[2 * i1 + 1];
        o.uv[4] = uvs[2 * i2]; o.uv[5] = uvs[2 * i2 + 1];
        o.pad[0] = o.pad[1] = 0.0f;
        out[s] = o;
    }
};
void launch_emit_uvs(CudaExec& ex, const FastTri* ftris, const GpuPtIndex* indices, const GpuPtMesh* instances, const float* uvs, FastUv* out, int n_slots) {
    ex.launch(n_slots, FnEmitUvs{ ftris, indices, instances, uvs, out });
}

int run_build_wbvh(CudaExec& ex, WBuild b, WReport* report) { return build_wbvh(ex, b, report); }
int run_build_lbvh(CudaExec& ex, WBuild b, WReport* report, bool hybrid) { return build_lbvh(ex, b, report, hybrid); }
void run_build_q4(CudaExec& ex, WBuild b, int n_pairs) { ex.launch(n_pairs, FnWbQ4{ b, n_pairs }); }
void run_refit_wbvh(CudaExec& ex, WBuild b, int n_pairs, int n_levels, const unsigned char* d_moved, float* d_area_sum, float out[7]) {
    refit_wbvh(ex, b, n_pairs, n_levels, d_moved, d_area_sum, out);
}

// ---- path-trace pass ------------------------------------------------------------------
// Block = 128 threads = a 16 x 8 pixel tile; each warp covers an 8 x 4 sub-tile so the
// primary rays of a warp are coherent.  Replaces Dispatch(ceil(W/16), ceil(H/16), 1) of a
// 16x16 work group (cave/engine/render_graph/common_passes.cpp:990-997).
template<bool kFast, bool kCount, int kQuorum = kStepQuorum, int kBatch = kShadeBatch, bool kTex = false, bool kStream = false>
__global__ void __launch_bounds__(128, kTex ? 6 : 8) k_trace(const TraceParams p) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int x = p.tile_x0 + blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
    const int y = p.tile_y0 + blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
    if (p.tile_x1 > 0 && (x >= p.tile_x1 || y >= p.tile_y1)) x = p.width;  // outside the tile: the body treats x == width as no pixel
    unsigned long long cnt[3] = { 0ull, 0ull, 0ull };
    __shared__ int s_stack[kFast ? kStackSize * 128 : 1];  // [depth][thread]: conflict-free whatever each lane's depth
    unsigned int segs = kFast ? body_trace_pixel_fast<kCount, kQuorum, kBatch, false, false, kTex, kStream>(p, x, y, s_stack + threadIdx.x, 128, cnt) : body_trace_pixel_reference<kCount>(p, x, y, cnt);
    segs = __reduce_add_sync(0xffffffffu, segs);
    if (lane == 0 && segs) atomicAdd(&p.counters[0], (unsigned long long)segs);
    if (kCount) {
#pragma unroll
        for (int k = 1; k < 3; ++k) {
            unsigned long long v = cnt[k];
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xffffffffu, v, d);
            if (lane == 0 && v) atomicAdd(&p.counters[k], v);
        }
    }
}

// ---- the per-pixel body over the four-wide records (CAVEPT_KERNEL_WIDE4): same grid, same stack, half the
//      dependent node fetches per segment.
template<bool kCount>
__global__ void __launch_bounds__(128, 8) k_trace_wide(const TraceParams p) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int x = p.tile_x0 + blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
    const int y = p.tile_y0 + blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
    if (p.tile_x1 > 0 && (x >= p.tile_x1 || y >= p.tile_y1)) x = p.width;
    unsigned long long cnt[3] = { 0ull, 0ull, 0ull };
    __shared__ int s_stack[kStackSize * 128];
    int spill[kWideSpill];  // local memory; touched only when a walk defers more than kStackSize nodes
    unsigned int segs = body_trace_pixel_fast<kCount, kStepQuorum, kShadeBatch, false, true>(p, x, y, s_stack + threadIdx.x, 128, cnt, nullptr, spill);
    segs = __reduce_add_sync(0xffffffffu, segs);
    if (lane == 0 && segs) atomicAdd(&p.counters[0], (unsigned long long)segs);
    if (kCount) {
#pragma unroll
        for (int k = 1; k < 3; ++k) {
            unsigned long long v = cnt[k];
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xffffffffu, v, d);
            if (lane == 0 && v) atomicAdd(&p.counters[k], v);
        }
    }
}
void launch_trace_wide(CudaExec& ex, const TraceParams& p, bool count_visits) {
    const int tw = p.tile_x1 > 0 ? p.tile_x1 - p.tile_x0 : p.width, th = p.tile_x1 > 0 ? p.tile_y1 - p.tile_y0 : p.height;
    const dim3 grid((tw + 15) / 16, (th + 7) / 8, 1), block(128, 1, 1);
    if (count_visits) k_trace_wide<true><<<grid, block, 0, ex.stream>>>(p);
    else k_trace_wide<false><<<grid, block, 0, ex.stream>>>(p);
    ex.note(cudaGetLastError());
    ++ex.launches;
}

// ---- second-generation production kernel: persistent warps + job queue + two pixel streams per lane
//      (trace_pool_body.h).  Grid = SMs x resident blocks; dynamic shared memory = the lane-strided
//      traversal stack ([stack_entries][128] ints) followed by the record pool ([38][128] floats).
#ifndef CV_POOL_BLOCKS
#define CV_POOL_BLOCKS 6
#endif
template<bool kCount>
__global__ void __launch_bounds__(128, CV_POOL_BLOCKS) k_trace_pool(const TraceParams p, int stack_entries, int* job_counter) {
    extern __shared__ __align__(16) int s_pool[];
    int* stack = s_pool + threadIdx.x;
    float* rec = reinterpret_cast<float*>(s_pool + stack_entries * 128) + threadIdx.x;
    unsigned long long cnt[3] = { 0ull, 0ull, 0ull };
    unsigned int segs = body_trace_pool<kCount>(p, stack, rec, 128, job_counter, cnt);
    const int lane = threadIdx.x & 31;
    segs = __reduce_add_sync(0xffffffffu, segs);
    if (lane == 0 && segs) atomicAdd(&p.counters[0], (unsigned long long)segs);
    if (kCount) {
#pragma unroll
        for (int k = 1; k < 3; ++k) {
            unsigned long long v = cnt[k];
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xffffffffu, v, d);
            if (lane == 0 && v) atomicAdd(&p.counters[k], v);
        }
    }
}

// ---- the per-pixel body as a persistent launch with the job queue (CAVEPT_KERNEL_QUEUE): same registers, same
//      fixed 40-entry stack and 8 blocks per SM as k_trace<fast>; the grid is SMs x 8 blocks whatever the image.
template<bool kCount>
__global__ void __launch_bounds__(128, 8) k_trace_queue(const TraceParams p, int* job_counter) {
    const int lane = threadIdx.x & 31;
    __shared__ int s_stack[kStackSize * 128];
    const TraceJobs jobs = trace_jobs(p);
    const int job = trace_draw_jobs(job_counter, true);
    int x = p.width, y = 0;  // x == width: no pixel yet (past the last job, or the padding of a ragged rectangle)
    if (job < jobs.n_jobs) {
        if (!trace_job_pixel(jobs, job, x, y)) x = p.width;
    }
    unsigned long long cnt[3] = { 0ull, 0ull, 0ull };
    unsigned int segs = body_trace_pixel_fast<kCount, kStepQuorum, kShadeBatch, true>(p, x, y, s_stack + threadIdx.x, 128, cnt, job_counter);
    segs = __reduce_add_sync(0xffffffffu, segs);
    if (lane == 0 && segs) atomicAdd(&p.counters[0], (unsigned long long)segs);
    if (kCount) {
#pragma unroll
        for (int k = 1; k < 3; ++k) {
            unsigned long long v = cnt[k];
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xffffffffu, v, d);
            if (lane == 0 && v) atomicAdd(&p.counters[k], v);
        }
    }
}

void launch_trace_queue(CudaExec& ex, const TraceParams& p, bool count_visits, int* job_counter, int n_sms) {
    const TraceJobs jobs = trace_jobs(p);
    int blocks = (jobs.n_jobs + 127) / 128;
    if (blocks > n_sms * 8) blocks = n_sms * 8;
    if (blocks < 1) blocks = 1;
    ex.note(cudaMemsetAsync(job_counter, 0, sizeof(int), ex.stream));
    if (count_visits) k_trace_queue<true><<<blocks, 128, 0, ex.stream>>>(p, job_counter);
    else k_trace_queue<false><<<blocks, 128, 0, ex.stream>>>(p, job_counter);
    ex.note(cudaGetLastError());
    ++ex.launches;
}

size_t trace_pool_smem_bytes(int stack_entries) { return (size_t)(stack_entries + kPoolWordsPerThread) * 128 * sizeof(int); }

// Resident blocks per SM of k_trace_pool at this stack depth (occupancy query, cached per depth by the caller).
int trace_pool_blocks_per_sm(int stack_entries, bool count_visits) {
    int n = 0;
    const size_t smem = trace_pool_smem_bytes(stack_entries);
    const cudaError_t e = count_visits ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_trace_pool<true>, 128, smem)
                                       : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_trace_pool<false>, 128, smem);
    return e == cudaSuccess ? n : 0;
}

void launch_trace_pool(CudaExec& ex, const TraceParams& p, bool count_visits, int stack_entries, int* job_counter, int grid_blocks) {
    const size_t smem = trace_pool_smem_bytes(stack_entries);
    const PoolJobs jobs = pool_jobs(p);
    int blocks = (jobs.n_jobs / 2 + 127) / 128;  // every thread starts with two jobs
    if (blocks > grid_blocks) blocks = grid_blocks;
    if (blocks < 1) blocks = 1;
    ex.note(cudaMemsetAsync(job_counter, 0, sizeof(int), ex.stream));
    if (count_visits) k_trace_pool<true><<<blocks, 128, smem, ex.stream>>>(p, stack_entries, job_counter);
    else k_trace_pool<false><<<blocks, 128, smem, ex.stream>>>(p, stack_entries, job_counter);
    ex.note(cudaGetLastError());
    ++ex.launches;
}

void launch_trace(CudaExec& ex, const TraceParams& p, bool fast, bool count_visits, bool textured, bool stream) {
    const int tw = p.tile_x1 > 0 ? p.tile_x1 - p.tile_x0 : p.width, th = p.tile_x1 > 0 ? p.tile_y1 - p.tile_y0 : p.height;
    const dim3 grid((tw + 15) / 16, (th + 7) / 8, 1), block(128, 1, 1);
    if (fast && textured) {  // the instantiation with texture / environment-map shading (trace_body.h shade_hit_textured)
        if (count_visits) k_trace<true, true, kStepQuorum, kShadeBatch, true><<<grid, block, 0, ex.stream>>>(p);
        else k_trace<true, false, kStepQuorum, kShadeBatch, true><<<grid, block, 0, ex.stream>>>(p);
    } else if (fast && stream) {  // many-sample launch over a large scene: L1 hints on the gathers
        if (count_visits) k_trace<true, true, kStepQuorum, kShadeBatch, false, true><<<grid, block, 0, ex.stream>>>(p);
        else k_trace<true, false, kStepQuorum, kShadeBatch, false, true><<<grid, block, 0, ex.stream>>>(p);
    } else if (fast) {
        if (count_visits) k_trace<true, true><<<grid, block, 0, ex.stream>>>(p);
        else k_trace<true, false><<<grid, block, 0, ex.stream>>>(p);
    } else {
        if (count_visits) k_trace<false, true><<<grid, block, 0, ex.stream>>>(p);
        else k_trace<false, false><<<grid, block, 0, ex.stream>>>(p);
    }
    ex.note(cudaGetLastError());
    ++ex.launches;
}

// ---- gather probe: the roof the trace kernel is measured against -------------------------------------------
// Every lane of every warp fetches one random 32-byte record (ONE LDG.E.256, the QPair / FastTri access of
// walk_step) from a working set of `lines` 128-byte lines, in a dependent chain like a traversal.  With the working
// set = the scene's node + triangle bytes this is the rate at which this GPU can serve the kernel's gathers from L2
// (64 KB: from L1).  bench.py calls it untimed, in the same run as the measurement.
__global__ void __launch_bounds__(128) k_gather_probe(const char* __restrict__ base, unsigned lines, int iters, float* __restrict__ out) {
    unsigned s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    float acc = 0.0f;
    for (int i = 0; i < iters; ++i) {
        s = s * 1664525u + 1013904223u;
        const unsigned line = (unsigned)(((unsigned long long)(s >> 4) * lines) >> 28);  // uniform in [0, lines)
        const F8 a = ld256(base + (size_t)line * 128 + ((s & 3u) << 5));
        acc += a.a + a.h;
        s += __float_as_uint(acc) & 1u;  // the next address depends on the data: a dependent chain
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
float run_gather_probe(CudaExec& ex, const void* base, size_t bytes, int iters, cudaEvent_t e0, cudaEvent_t e1, int n_sms, double* lane_gathers) {
    const int blocks = n_sms * 8;
    const unsigned lines = (unsigned)(bytes / 128);
    float* out = reinterpret_cast<float*>(const_cast<void*>(base));  // results land in the first 600 KB of the working set itself
    k_gather_probe<<<blocks, 128, 0, ex.stream>>>(reinterpret_cast<const char*>(base), lines, iters / 8 + 1, out);  // warm
    ex.note(cudaEventRecord(e0, ex.stream));
    k_gather_probe<<<blocks, 128, 0, ex.stream>>>(reinterpret_cast<const char*>(base), lines, iters, out);
    ex.note(cudaEventRecord(e1, ex.stream));
    ex.note(cudaGetLastError());
    ex.note(cudaEventSynchronize(e1));
    ex.launches += 2;
    float ms = 0.0f;
    ex.note(cudaEventElapsedTime(&ms, e0, e1));
    *lane_gathers = (double)blocks * 128.0 * (double)iters;
    return ms;
}

struct FnProbe {
    const float *rays, *tris, *boxes;
    float* out;
    __device__ void operator()(int i) const { body_probe(rays, tris, boxes, out, i); }
};
void launch_probe(CudaExec& ex, const float* rays, const float* tris, const float* boxes, float* out, int n) { ex.launch(n, FnProbe{ rays, tris, boxes, out }); }

// ---- tone pass: c/(c+1), pow(c, 1/2.2) (cave/shader/glsl/post_process.ps.glsl:72-76) ----
__device__ __forceinline__ float tone(float c) {
    c = __fdiv_rn(c, c + 1.0f);
    return powf(c, 1.0f / 2.2f);
}
struct FnToneRgba8 {
    const F4* accum;
    uchar4* out;
    __device__ void operator()(int i) const {
        const F4 a = accum[i];
        uchar4 o;
        o.x = (unsigned char)(fminf(fmaxf(tone(a.x), 0.0f), 1.0f) * 255.0f + 0.5f);
        o.y = (unsigned char)(fminf(fmaxf(tone(a.y), 0.0f), 1.0f) * 255.0f + 0.5f);
        o.z = (unsigned char)(fminf(fmaxf(tone(a.z), 0.0f), 1.0f) * 255.0f + 0.5f);
        o.w = 255;
        out[i] = o;
    }
};
struct FnToneF32 {
    const F4* accum;
    F4* out;
    __device__ void operator()(int i) const {
        const F4 a = accum[i];
        F4 o;
        o.x = tone(a.x);
        o.y = tone(a.y);
        o.z = tone(a.z);
        o.w = 1.0f;
        out[i] = o;
    }
};
// RT_FMT_TONE = R16G16B16A16_FLOAT (cave/engine/render_graph/render_graph_defines.h:64): the format the reference's tone
// target has; round-to-nearest-even halves of the fp32 tone values, alpha 1.0
struct FnToneF16 {
    const F4* accum;
    uint2* out;
    __device__ void operator()(int i) const {
        const F4 a = accum[i];
        const unsigned r = __half_as_ushort(__float2half_rn(tone(a.x))), g = __half_as_ushort(__float2half_rn(tone(a.y)));
        const unsigned b = __half_as_ushort(__float2half_rn(tone(a.z)));
        out[i] = make_uint2(r | (g << 16), b | (0x3C00u << 16));
    }
};
void launch_tonemap_f16(CudaExec& ex, const F4* accum, unsigned short* out, int n_pixels) { ex.launch(n_pixels, FnToneF16{ accum, reinterpret_cast<uint2*>(out) }); }

void launch_tonemap_rgba8(CudaExec& ex, const F4* accum, unsigned char* out, int n_pixels) {
    ex.launch(n_pixels, FnToneRgba8{ accum, reinterpret_cast<uchar4*>(out) });
}
void launch_tonemap_f32(CudaExec& ex, const F4* accum, F4* out, int n_pixels) { ex.launch(n_pixels, FnToneF32{ accum, out }); }

// (mean, n) <-> (sum, n) for the multi-GPU sum-reduce
struct FnAccumScale {
    F4* accum;
    bool to_sums;
    __device__ void operator()(int i) const {
        F4 a = accum[i];
        if (to_sums) {
            a.x *= a.w;
            a.y *= a.w;
            a.z *= a.w;
        } else if (a.w > 0.0f) {
            a.x = __fdiv_rn(a.x, a.w);
            a.y = __fdiv_rn(a.y, a.w);
            a.z = __fdiv_rn(a.z, a.w);
        }
        accum[i] = a;
    }
};
void launch_accum_scale(CudaExec& ex, F4* accum, int n_pixels, bool to_sums) { ex.launch(n_pixels, FnAccumScale{ accum, to_sums }); }

}  // namespace cavept
