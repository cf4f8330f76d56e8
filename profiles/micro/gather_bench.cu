// Microbenchmark: cost of divergent gathers through L1 on B200 (one 128-byte line per lane).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather_bench gather_bench.cu && ./gather_bench
// Modes: 0 = 1 x LDG.128 per lane, 1 = 2 x LDG.128 (same line), 2 = 4 x LDG.128 (same 64 B),
//        3 = 1 x LDG.256, 4 = 2 x LDG.256 (same line), 5 = 2 x LDG.256 with lane pairs sharing a line,
//        6 = 4 x LDG.128 with lane quads sharing a 64 B record
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>

struct alignas(32) F8 { float v[8]; };
__device__ __forceinline__ F8 ld256(const void* p) {
    F8 r;
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7]) : "l"(p));
    return r;
}
__device__ __forceinline__ float4 ld128(const void* p) { return __ldg(reinterpret_cast<const float4*>(p)); }

template<int MODE>
__global__ void __launch_bounds__(128) k(const char* base, unsigned mask_lines, int iters, float* out) {
    unsigned s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    const int lane = threadIdx.x & 31;
    float acc = 0.f;
    for (int i = 0; i < iters; ++i) {
        s = s * 1664525u + 1013904223u;
        unsigned line = (s >> 8) & mask_lines;            // random 128-byte line
        if (MODE == 5) line = __shfl_sync(0xffffffffu, line, lane & ~1);
        if (MODE == 6) line = __shfl_sync(0xffffffffu, line, lane & ~3);
        const char* p = base + (size_t)line * 128;
        if (MODE == 0) { float4 a = ld128(p); acc += a.x + a.w; }
        if (MODE == 1) { float4 a = ld128(p), b = ld128(p + 16); acc += a.x + b.w; }
        if (MODE == 2) { float4 a = ld128(p), b = ld128(p + 16), c = ld128(p + 32), d = ld128(p + 48); acc += a.x + b.w + c.y + d.z; }
        if (MODE == 3) { F8 a = ld256(p); acc += a.v[0] + a.v[7]; }
        if (MODE == 4) { F8 a = ld256(p), b = ld256(p + 32); acc += a.v[0] + b.v[7]; }
        if (MODE == 5) { F8 a = ld256(p + (lane & 1) * 32), b = ld256(p + 64 + (lane & 1) * 32); acc += a.v[0] + b.v[7]; }
        if (MODE == 6) { float4 a = ld128(p + (lane & 3) * 16); acc += a.x + a.w; }
        s += __float_as_uint(acc) & 1u;                    // dependent chain like a traversal
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template<int MODE>
void run(const char* d, size_t bytes, float* out, const char* name) {
    const int blocks = 148 * 8, iters = 2000;
    const unsigned mask = (unsigned)(bytes / 128 - 1);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<blocks, 128>>>(d, mask, 100, out);
    cudaEventRecord(e0);
    k<MODE><<<blocks, 128>>>(d, mask, iters, out);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double lane_iters = (double)blocks * 128 * iters;
    const double clk = 1.965e9;
    printf("%-44s ws %6zu KB: %7.3f ms  %7.2f G lane-gathers/s  %6.2f SM-cycles per warp-iteration\n", name, bytes >> 10, ms, lane_iters / ms / 1e6,
           ms * 1e-3 * clk * 148 / (lane_iters / 32));
}

int main() {
    const size_t maxb = 64u << 20;
    char* d; float* out;
    cudaMalloc(&d, maxb); cudaMemset(d, 0, maxb); cudaMalloc(&out, 148 * 8 * 128 * 4);
    for (size_t ws : { (size_t)64 << 10, (size_t)4 << 20, (size_t)32 << 20 }) {
        run<0>(d, ws, out, "1 x LDG.128 per lane");
        run<1>(d, ws, out, "2 x LDG.128 same line");
        run<2>(d, ws, out, "4 x LDG.128 same 64 B");
        run<3>(d, ws, out, "1 x LDG.256 per lane");
        run<4>(d, ws, out, "2 x LDG.256 same line (64 B)");
        run<5>(d, ws, out, "2 x LDG.256, lane pairs share a line (2x64B)");
        run<6>(d, ws, out, "1 x LDG.128, lane quads share 64 B");
    }
    return 0;
}
