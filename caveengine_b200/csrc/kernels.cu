// __global__ wrappers for the kernel bodies, the scan primitive, and launchers.
// Compiled for sm_100a only, with -fmad=false (see cv_common.h on float semantics).
#include "kernels.h"

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
__global__ void __launch_bounds__(256) k_wb_split_binned(const WBuild b, int level) {
    constexpr unsigned kAll = 0xffffffffu;
    const int w = (int)((blockIdx.x * 256u + threadIdx.x) >> 5);
    const int lane = threadIdx.x & 31, k = lane & 15;
    if (w >= b.counters[kWLevelBase + level]) return;  // warp-uniform
    WItem& it = b.items[w];
    if (it.axis < 0 || it.bins_slot < 0) return;  // median items and small items: FnWbSplit, one thread each
    unsigned long long best = ~0ull;
    WBox best_box[2];
    int best_cnt[2] = { 0, 0 };
    best_box[0] = best_box[1] = wb_empty();
    for (int a = 0; a < kWAxes; ++a) {
        WBin* src = &b.bins[it.bins_slot].b[a * kWBins + k];
        const WBin raw = *src;
        __syncwarp();
        if (lane < 16) {
            WBin z = {};
            *src = z;
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
        if (key < best) {  // warp-uniform
            best = key;
            const int kk = (int)(key & 15ull);
            best_cnt[0] = __shfl_sync(kAll, lc, kk, 16);
            best_box[0] = shfl_box(kAll, lb, kk, false, 0);
            best_cnt[1] = __shfl_sync(kAll, rc, kk + 1, 16);
            best_box[1] = shfl_box(kAll, rb, kk + 1, false, 0);
        }
    }
    if (lane == 0) {
        if (best == ~0ull) {
            it.axis = -1;  // every prim in one bin on every axis: median split
        } else {
            const int key = (int)(best & 0xffull);
            it.axis = wb_slot_axis(it, key / kWBins);
            it.split = key % kWBins;
        }
        wb_finish_split(b, it, level, best_box, best_cnt);
    }
}
void CudaExec::wb_split(const WBuild& b, int level, int bound) {
    if (bound <= 0) return;
    // arena items: at most bins_capacity of them, and at most `bound` items altogether
    const int warps = bound < b.bins_capacity ? bound : b.bins_capacity;
    k_wb_split_binned<<<(warps * 32 + 255) / 256, 256, 0, stream>>>(b, level);
    note(cudaGetLastError());
    ++launches;
    launch(bound, FnWbSplit{ b, level, false });
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
int run_build_wbvh(CudaExec& ex, WBuild b, WReport* report) { return build_wbvh(ex, b, report); }
int run_build_lbvh(CudaExec& ex, WBuild b, WReport* report) { return build_lbvh(ex, b, report); }
void run_build_q4(CudaExec& ex, WBuild b, int n_pairs) { ex.launch(n_pairs, FnWbQ4{ b, n_pairs }); }
void run_refit_wbvh(CudaExec& ex, WBuild b, const int* level_start, int n_levels, const unsigned char* d_moved, float* d_area_sum, float out[7]) {
    refit_wbvh(ex, b, level_start, n_levels, d_moved, d_area_sum, out);
}

// ---- path-trace pass ------------------------------------------------------------------
// Block = 128 threads = a 16 x 8 pixel tile; each warp covers an 8 x 4 sub-tile so the
// primary rays of a warp are coherent.  Replaces Dispatch(ceil(W/16), ceil(H/16), 1) of a
// 16x16 work group (cave/engine/render_graph/common_passes.cpp:990-997).
template<bool kFast, bool kCount, int kQuorum = kStepQuorum, int kBatch = kShadeBatch>
__global__ void __launch_bounds__(128, 8) k_trace(const TraceParams p) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int x = p.tile_x0 + blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
    const int y = p.tile_y0 + blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
    if (p.tile_x1 > 0 && (x >= p.tile_x1 || y >= p.tile_y1)) x = p.width;  // outside the tile: the body treats x == width as no pixel
    unsigned long long cnt[3] = { 0ull, 0ull, 0ull };
    __shared__ int s_stack[kFast ? kStackSize * 128 : 1];  // [depth][thread]: conflict-free whatever each lane's depth
    unsigned int segs = kFast ? body_trace_pixel_fast<kCount, kQuorum, kBatch>(p, x, y, s_stack + threadIdx.x, 128, cnt) : body_trace_pixel_reference<kCount>(p, x, y, cnt);
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

void launch_trace(CudaExec& ex, const TraceParams& p, bool fast, bool count_visits) {
    const int tw = p.tile_x1 > 0 ? p.tile_x1 - p.tile_x0 : p.width, th = p.tile_x1 > 0 ? p.tile_y1 - p.tile_y0 : p.height;
    const dim3 grid((tw + 15) / 16, (th + 7) / 8, 1), block(128, 1, 1);
    if (fast) {
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
