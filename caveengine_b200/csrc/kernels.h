// Device-side entry points of libcavept (implemented in kernels.cu), called by the
// C-ABI layer in cavept_api.cu.  Everything here launches on an explicit stream.
#pragma once
#include <cuda_runtime.h>

#include "bvh_parity_driver.h"
#include "trace_body.h"
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
    void upload(void* dst, const void* src, size_t bytes);
    void download(void* dst, const void* src, size_t bytes);
    void fill(void* dst, int byte, size_t bytes);
    void note(cudaError_t e) {
        if (first_error == cudaSuccess && e != cudaSuccess) first_error = e;
    }
};

constexpr int kScanTile = 2048;

// scene assembly: raw MeshAsset arrays -> reference GPU layouts
// (ConstructMesh + AppendVertices/AppendIndices, path_tracer.cpp:27-47, :72-88)
void launch_assemble_vertices(CudaExec& ex, const float* positions, const float* normals, GpuPtVertex* out, int n_vertices);
void launch_assemble_indices(CudaExec& ex, const uint32_t* raw_indices, const int* tri_off, const int* vert_off, int n_meshes, GpuPtIndex* out, int n_tris);

int run_build_bvh_parity(CudaExec& ex, BuildArrays a, const int* tri_off_host, const int* d_tri_off, int n_meshes, BuildItem* host_scratch, BuildReport* report);
int run_build_wbvh(CudaExec& ex, WBuild b, WReport* report);

void launch_trace(CudaExec& ex, const TraceParams& p, bool fast, bool count_visits);
void launch_tonemap_rgba8(CudaExec& ex, const F4* accum, unsigned char* out, int n_pixels);
void launch_tonemap_f32(CudaExec& ex, const F4* accum, F4* out, int n_pixels);
void launch_accum_scale(CudaExec& ex, F4* accum, int n_pixels, bool to_sums);

}  // namespace cavept
