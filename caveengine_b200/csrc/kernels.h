// Device-side entry points of libcavept (implemented in kernels.cu), called by the
// C-ABI layer in cavept_api.cu.  Everything here launches on an explicit stream.
#pragma once
#include <cuda_runtime.h>

#include "bvh_parity_driver.h"
#include "trace_body.h"
#include "trace_pool_body.h"
#include "wbvh_body.h"

namespace cavept {

struct CudaExec {
    cudaStream_t stream = nullptr;
    unsigned long long launches = 0;
    int* scan_tmp = nullptr;  // block sums, >= n / kScanTile + 2 ints
    cudaError_t first_error = cudaSuccess;

    template<class F>
    void launch(int n, F f);
    void scan_u8(const unsigned char* in, int* out, int n);
    void scan_i32(const int* in, int* out, int n);
    // stable LSD radix sort of (key, value) pairs by the low `bits` of the key; result in keys/vals.
    // tmp buffers hold n entries each, hist needs 2 * sort_hist_entries(n) ints.
    void sort_pairs(uint32_t* keys, int* vals, uint32_t* keys_tmp, int* vals_tmp, int* hist, int n, int bits);
    void parity_reduce(const BuildArrays& a);  // parity builder: AABB / centroid folds, block-private where a block's triangles share an item
    void parity_bin(const BuildArrays& a);     // parity builder: bucket counts and boxes, same
    void parity_split(const BuildArrays& a, int n_items);     // parity builder: the 11 candidate costs of every bucketed item, sixteen lanes per item
    void parity_subtrees(const BuildArrays& a, int n_roots);  // parity builder, second phase: one block builds the whole subtree of each queued item
    void wb_prim(const WBuild& b);  // world AABBs + scene / centroid bounds, block-reduced
    void wb_bin(const WBuild& b);                          // bin pass of a level, block-privatised in shared memory
    void wb_split(const WBuild& b, int level, int bound);  // SAH split of every item of a level: a warp per arena item, a thread per small item
    void wb_partition(const WBuild& b);                    // left flags + exclusive scan + stable scatter in three launches
    void wb_subtrees(const WBuild& b, int n_roots);        // second phase: one block builds the whole subtree of each queued item
    void upload(void* dst, const void* src, size_t bytes);
    void download(void* dst, const void* src, size_t bytes);
    void fill(void* dst, int byte, size_t bytes);
    void note(cudaError_t e) {
        if (first_error == cudaSuccess && e != cudaSuccess) first_error = e;
    }
};

constexpr int kScanTile = 2048;
constexpr int kSortTile = 4096;  // keys per block of the radix sort: 256 threads x 16
inline size_t sort_hist_entries(int n) { return (size_t)256 * (size_t)((n + kSortTile - 1) / kSortTile) + 1; }

// scene assembly: raw MeshAsset arrays -> reference GPU layouts
// (ConstructMesh + AppendVertices/AppendIndices, path_tracer.cpp:27-47, :72-88)
void launch_assemble_vertices(CudaExec& ex, const float* positions, const float* normals, GpuPtVertex* out, int n_vertices);
void launch_assemble_indices(CudaExec& ex, const uint32_t* raw_indices, const int* tri_off, const int* vert_off, int n_meshes, GpuPtIndex* out, int n_tris);

int run_build_bvh_parity(CudaExec& ex, BuildArrays a, const int* tri_off_host, const int* d_tri_off, int n_meshes, BuildItem* host_scratch, BuildReport* report);
int run_build_wbvh(CudaExec& ex, WBuild b, WReport* report);
int run_build_lbvh(CudaExec& ex, WBuild b, WReport* report, bool hybrid = false);
void run_build_q4(CudaExec& ex, WBuild b, int n_pairs);  // Q4Node records over a finished pair tree
void run_refit_wbvh(CudaExec& ex, WBuild b, int n_pairs, int n_levels, const unsigned char* d_moved, float* d_area_sum, float out[7]);

void launch_trace(CudaExec& ex, const TraceParams& p, bool fast, bool count_visits, bool textured = false, bool stream = false);
void launch_emit_uvs(CudaExec& ex, const FastTri* ftris, const GpuPtIndex* indices, const GpuPtMesh* instances, const float* uvs, FastUv* out, int n_slots);
// the per-pixel body over the four-wide records (CAVEPT_KERNEL_WIDE4; p.q4nodes must be set)
void launch_trace_wide(CudaExec& ex, const TraceParams& p, bool count_visits);
// the per-pixel body as a persistent launch drawing pixel jobs from a global queue (CAVEPT_KERNEL_QUEUE)
void launch_trace_queue(CudaExec& ex, const TraceParams& p, bool count_visits, int* job_counter, int n_sms);
// persistent job-queue kernel (trace_pool_body.h); stack_entries = tree levels + 2, grid_blocks = SMs x resident blocks
size_t trace_pool_smem_bytes(int stack_entries);
int trace_pool_blocks_per_sm(int stack_entries, bool count_visits);
void launch_trace_pool(CudaExec& ex, const TraceParams& p, bool count_visits, int stack_entries, int* job_counter, int grid_blocks);
// random 32-byte LDG.E.256 gathers over `bytes` of device memory; returns the device time of `iters` gathers per lane
float run_gather_probe(CudaExec& ex, const void* base, size_t bytes, int iters, cudaEvent_t e0, cudaEvent_t e1, int n_sms, double* lane_gathers);
void launch_probe(CudaExec& ex, const float* rays, const float* tris, const float* boxes, float* out, int n);
void launch_tonemap_rgba8(CudaExec& ex, const F4* accum, unsigned char* out, int n_pixels);
void launch_tonemap_f16(CudaExec& ex, const F4* accum, unsigned short* out_rgba16f, int n_pixels);
void launch_tonemap_f32(CudaExec& ex, const F4* accum, F4* out, int n_pixels);
void launch_accum_scale(CudaExec& ex, F4* accum, int n_pixels, bool to_sums);

}  // namespace cavept
