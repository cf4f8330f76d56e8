// Shared definitions for the cavept kernels.
//
// Kernel bodies are written as `CV_HD` functions of a global thread index so that
// the exact same source can be (a) wrapped by __global__ kernels in kernels.cu and
// (b) stepped serially by the g++-built emulator under tests/emu/, which lets the
// algorithms be checked against the oracle in a container without a GPU.  The
// emulator is test code; the product has no CPU path.
//
// Float semantics: device code is compiled with -fmad=false and host emulation with
// -ffp-contract=off, so every `a * b + c` below is two rounded operations, exactly
// as the reference's C++ builder (x86-64, no FMA) and its fp32 shader evaluate it.
// Where the fast kernel wants a fused multiply-add it calls cv_fma() explicitly.
#pragma once

#include <stdint.h>

#if defined(__CUDACC__)
#define CV_HD __host__ __device__ __forceinline__
#define CV_D __device__ __forceinline__
#else
#define CV_HD inline
#define CV_D inline
#include <cmath>
#include <cstring>
#endif

namespace cavept {

// ---- reference GPU layouts: cave/shader/structured_buffer.hlsl.h:24-61 -------------
struct GpuPtBvh {  // 48 B
    float min[3];
    int missIdx;
    float max[3];
    int hitIdx;
    float pad[2];
    int leaf;
    int triangleIndex;
};
struct GpuPtVertex {  // 32 B
    float position[3];
    int pad1;
    float normal[3];
    int pad2;
};
struct GpuPtIndex {  // 16 B
    int tri[3];
    int pad1;
};
struct GpuPtMesh {  // 144 B
    float transform[16];
    float transformInv[16];
    float pad[2];
    int rootBvhId;
    int materialId;
};
struct GpuPtMaterial {  // 32 B
    float baseColor[3];
    float roughness;
    float emissive[3];
    float metallic;
};
static_assert(sizeof(GpuPtBvh) == 48 && sizeof(GpuPtVertex) == 32 && sizeof(GpuPtIndex) == 16, "layout");
static_assert(sizeof(GpuPtMesh) == 144 && sizeof(GpuPtMaterial) == 32, "layout");

struct F3 {
    float x, y, z;
};
struct alignas(16) F4 {
    float x, y, z, w;
};

CV_HD F3 f3(float x, float y, float z) {
    F3 r;
    r.x = x;
    r.y = y;
    r.z = z;
    return r;
}
CV_HD F3 operator+(F3 a, F3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
CV_HD F3 operator-(F3 a, F3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
CV_HD F3 operator*(F3 a, F3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
CV_HD F3 operator*(float s, F3 a) { return f3(s * a.x, s * a.y, s * a.z); }
CV_HD F3 operator*(F3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
CV_HD float comp(F3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

// cave/engine/math/detail/common.h:50-58 -- (a < b ? a : b), (a > b ? a : b)
CV_HD float min_ref(float a, float b) { return a < b ? a : b; }
CV_HD float max_ref(float a, float b) { return a > b ? a : b; }

// cave/engine/math/detail/vector_math.h:513-545
CV_HD float dot3(F3 a, F3 b) {
    float r = a.x * b.x + a.y * b.y;
    r += a.z * b.z;
    return r;
}
CV_HD F3 cross3(F3 a, F3 b) { return f3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); }

CV_HD float cv_sqrt(float x) {
#if defined(__CUDA_ARCH__)
    return __fsqrt_rn(x);
#else
    return sqrtf(x);
#endif
}
CV_HD float cv_div(float a, float b) {
#if defined(__CUDA_ARCH__)
    return __fdiv_rn(a, b);
#else
    return a / b;
#endif
}
// 1/x, correctly rounded: the same bits as cv_div(1.0f, x) (both are the IEEE quotient), fewer instructions
CV_HD float cv_rcp(float x) {
#if defined(__CUDA_ARCH__)
    return __frcp_rn(x);
#else
    return 1.0f / x;
#endif
}
CV_HD float cv_fma(float a, float b, float c) {
#if defined(__CUDA_ARCH__)
    return __fmaf_rn(a, b, c);
#else
    return fmaf(a, b, c);
#endif
}
CV_HD F3 normalize3(F3 a) {  // vector_math.h:527-536
    const float inv = cv_rcp(cv_sqrt(dot3(a, a)));
    return a * inv;
}
CV_HD F3 lerp3(F3 x, F3 y, float s) { return (1.0f - s) * x + s * y; }  // vector_math.h:507-510

CV_HD uint32_t f2u(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    uint32_t u;
    memcpy(&u, &f, 4);
    return u;
#endif
}
CV_HD float u2f(uint32_t u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f;
    memcpy(&f, &u, 4);
    return f;
#endif
}

#define CV_INF (u2f(0x7f800000u))

// 32-byte record fetched with ONE 256-bit load (LDG.E.256, new with sm_100): half the L1
// wavefronts of two 128-bit loads when every lane reads a different cache line.
struct alignas(32) F8 {
    float a, b, c, d, e, f, g, h;
};
CV_HD F8 ld256(const void* p) {
    F8 r;
#if defined(__CUDA_ARCH__)
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(r.a), "=f"(r.b), "=f"(r.c), "=f"(r.d), "=f"(r.e), "=f"(r.f), "=f"(r.g), "=f"(r.h)
                 : "l"(p));
#else
    memcpy(&r, p, 32);
#endif
    return r;
}
// the same load with an L1 eviction hint: kHint 1 = L1::no_allocate (records read once per visit: triangles, normals),
// 2 = L1::evict_last (records many lanes come back to: the upper levels of the tree)
template<int kHint>
CV_HD F8 ld256_hint(const void* p) {
    F8 r;
#if defined(__CUDA_ARCH__)
    if (kHint == 1) {
        asm volatile("ld.global.nc.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=f"(r.a), "=f"(r.b), "=f"(r.c), "=f"(r.d), "=f"(r.e), "=f"(r.f), "=f"(r.g), "=f"(r.h)
                     : "l"(p));
    } else if (kHint == 2) {
        asm volatile("ld.global.nc.L1::evict_last.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=f"(r.a), "=f"(r.b), "=f"(r.c), "=f"(r.d), "=f"(r.e), "=f"(r.f), "=f"(r.g), "=f"(r.h)
                     : "l"(p));
    } else {
        r = ld256(p);
    }
#else
    memcpy(&r, p, 32);
#endif
    return r;
}
// 1/x to ~1 ulp (MUFU.RCP): for slopes whose error the padded boxes absorb, never for results.
CV_HD float rcp_approx(float x) {
#if defined(__CUDA_ARCH__)
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
#else
    return 1.0f / x;
#endif
}

// Order-preserving float -> uint32 map with -0 and +0 sharing one key.
CV_HD uint32_t ordered_key(float v) {
    uint32_t b = f2u(v);
    if ((b << 1) == 0u) b = 0u;
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
CV_HD float ordered_value(uint32_t k) { return u2f((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k); }

// ---- atomics: real ones on the device, plain serial updates in the emulator --------
CV_D void cv_atomic_min_u64(unsigned long long* p, unsigned long long v) {
#if defined(__CUDA_ARCH__)
    atomicMin(p, v);
#else
    if (v < *p) *p = v;
#endif
}
CV_D void cv_atomic_max_u64(unsigned long long* p, unsigned long long v) {
#if defined(__CUDA_ARCH__)
    atomicMax(p, v);
#else
    if (v > *p) *p = v;
#endif
}
CV_D void cv_atomic_min_u32(uint32_t* p, uint32_t v) {
#if defined(__CUDA_ARCH__)
    atomicMin(p, v);
#else
    if (v < *p) *p = v;
#endif
}
CV_D void cv_atomic_max_u32(uint32_t* p, uint32_t v) {
#if defined(__CUDA_ARCH__)
    atomicMax(p, v);
#else
    if (v > *p) *p = v;
#endif
}
CV_D int cv_atomic_add_i32(int* p, int v) {
#if defined(__CUDA_ARCH__)
    return atomicAdd(p, v);
#else
    int o = *p;
    *p = o + v;
    return o;
#endif
}
CV_D void cv_atomic_or_i32(int* p, int v) {
#if defined(__CUDA_ARCH__)
    atomicOr(p, v);
#else
    *p |= v;
#endif
}
CV_D void cv_atomic_max_i32(int* p, int v) {
#if defined(__CUDA_ARCH__)
    atomicMax(p, v);
#else
    if (v > *p) *p = v;
#endif
}
CV_D void cv_atomic_add_u64(unsigned long long* p, unsigned long long v) {
#if defined(__CUDA_ARCH__)
    atomicAdd(p, v);
#else
    *p += v;
#endif
}

CV_D void cv_threadfence() {
#if defined(__CUDA_ARCH__)
    __threadfence();
#endif
}
// load that bypasses the (non-coherent) L1: for data another SM wrote earlier in the same launch
CV_D F4 cv_load_cg(const F4* p) {
#if defined(__CUDA_ARCH__)
    const float4 v = __ldcg(reinterpret_cast<const float4*>(p));
    F4 r;
    r.x = v.x; r.y = v.y; r.z = v.z; r.w = v.w;
    return r;
#else
    return *p;
#endif
}
CV_HD int cv_clz(uint32_t v) {
#if defined(__CUDA_ARCH__)
    return __clz((int)v);
#else
    return v ? __builtin_clz(v) : 32;
#endif
}

// ---- warp-aggregated atomics ------------------------------------------------------------
// Lanes of a warp that pass the same key combine their operands with one redux.sync and a
// single lane issues the atomic: up to 32x fewer atomics on the hot addresses of the
// builders' top levels.  The emulator runs a warp of one.
struct CvGroup {
    unsigned mask;
    bool leader;
};
CV_D CvGroup cv_group_by(int key) {
    CvGroup g;
#if defined(__CUDA_ARCH__)
    g.mask = __match_any_sync(__activemask(), key);
    g.leader = (__ffs(g.mask) - 1) == (int)(threadIdx.x & 31u);
#else
    (void)key;
    g.mask = 1u;
    g.leader = true;
#endif
    return g;
}
CV_D void cv_group_min_u32(const CvGroup& g, uint32_t* p, uint32_t v) {
#if defined(__CUDA_ARCH__)
    v = __reduce_min_sync(g.mask, v);
    if (g.leader) atomicMin(p, v);
#else
    if (v < *p) *p = v;
#endif
}
CV_D void cv_group_max_u32(const CvGroup& g, uint32_t* p, uint32_t v) {
#if defined(__CUDA_ARCH__)
    v = __reduce_max_sync(g.mask, v);
    if (g.leader) atomicMax(p, v);
#else
    if (v > *p) *p = v;
#endif
}
// 64-bit min/max over the group with two 32-bit redux.sync: the high words first, then the low words of
// the lanes that hold the winning high word
CV_D void cv_group_min_u64(const CvGroup& g, unsigned long long* p, unsigned long long v) {
#if defined(__CUDA_ARCH__)
    const uint32_t hi = (uint32_t)(v >> 32), lo = (uint32_t)v;
    const uint32_t mhi = __reduce_min_sync(g.mask, hi);
    const uint32_t mlo = __reduce_min_sync(g.mask, hi == mhi ? lo : 0xffffffffu);
    if (g.leader) atomicMin(p, ((unsigned long long)mhi << 32) | mlo);
#else
    if (v < *p) *p = v;
#endif
}
CV_D void cv_group_max_u64(const CvGroup& g, unsigned long long* p, unsigned long long v) {
#if defined(__CUDA_ARCH__)
    const uint32_t hi = (uint32_t)(v >> 32), lo = (uint32_t)v;
    const uint32_t mhi = __reduce_max_sync(g.mask, hi);
    const uint32_t mlo = __reduce_max_sync(g.mask, hi == mhi ? lo : 0u);
    if (g.leader) atomicMax(p, ((unsigned long long)mhi << 32) | mlo);
#else
    if (v > *p) *p = v;
#endif
}
CV_D void cv_group_add_i32(const CvGroup& g, int* p, int v) {
#if defined(__CUDA_ARCH__)
    v = __reduce_add_sync(g.mask, v);
    if (g.leader) atomicAdd(p, v);
#else
    *p += v;
#endif
}

}  // namespace cavept
