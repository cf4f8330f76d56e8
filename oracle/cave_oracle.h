/* TEST INFRASTRUCTURE ONLY -- CPU restatement ("port") of CaveEngine's path-tracer
 * hot path, in fp32 exactly as the shader computes it.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library; the product (libcavept.so) never links, loads or calls it.
 *
 * Parity status: PINNED against the reference's own sources compiled in place
 * (oracle/_ref, see oracle/Makefile) -- tests/test_oracle_vs_ref.py -- and against
 * the golden fixtures under tests/golden/ that oracle/_ref generated
 * (tests/golden/make_golden.py).  The reference ships no tests or golden vectors
 * for this path, so those are the only pins that exist.
 *
 * Struct layouts are the reference's GPU layouts,
 * cave/shader/structured_buffer.hlsl.h:24-61.
 */
#ifndef CAVE_ORACLE_H_
#define CAVE_ORACLE_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct { float min[3]; int miss_idx; float max[3]; int hit_idx; float pad[2]; int leaf; int triangle_index; } co_bvh;      /* GpuPtBvh 48 B */
typedef struct { float position[3]; int pad1; float normal[3]; int pad2; } co_vertex;                                             /* GpuPtVertex 32 B */
typedef struct { int tri[3]; int pad1; } co_index;                                                                                 /* GpuPtIndex 16 B */
typedef struct { float transform[16]; float transform_inv[16]; float pad[2]; int root_bvh_id; int material_id; } co_mesh;        /* GpuPtMesh 144 B */
typedef struct { float base_color[3]; float roughness; float emissive[3]; float metallic; } co_material;                          /* GpuPtMaterial 32 B */
typedef struct { float position[3]; float fov_degree; float forward[3]; float right[3]; float up[3]; } co_camera;

/* counters[0]=path segments (HitScene calls), [1]=node visits (HitBvh calls),
 * [2]=triangle tests (HitTriangle calls), [3]=mesh-loop iterations */
enum { CO_N_COUNTERS = 4 };

/* BvhAccel::Construct + FillGpuBvhAccel (bvh_accel.cpp:265-288, :237-263).
 * Returns node count, -1 degenerate geometry (where the reference traps),
 * -2 capacity, -3 bad arguments. */
int caveoracle_bvh_build(const uint32_t* indices, int index_count, const float* positions, int vertex_count,
                         co_bvh* out_nodes, int cap);

/* glm::inverse(mat4) as called at path_tracer.cpp:178 (column-major). */
void caveoracle_mat4_inverse(const float* m, float* out);

/* main() of path_tracer.cs.glsl:11-60 + shared_path_tracer.h:60-239 over a pixel
 * rectangle, frames [first_frame, first_frame+n_frames); first frame overwrites
 * iff reset_first.  Returns segments; counters may be NULL. */
long long caveoracle_trace_render(const co_vertex* vertices, const co_index* indices, const co_bvh* bvhs,
                                  const co_mesh* meshes, int mesh_count, const co_material* materials,
                                  const co_camera* camera, int max_bounce, int width, int height, int x0, int y0,
                                  int x1, int y1, int first_frame, int n_frames, int reset_first, float* accum,
                                  int threads, long long* counters, double* ms);

/* post_process.ps.glsl:72-76: Reinhard c/(c+1) then pow(c, 1/2.2); rgb of a
 * float4 accumulator -> float4 (alpha 1). */
void caveoracle_tonemap(const float* accum, float* out, long long pixel_count);

#ifdef __cplusplus
}
#endif
#endif
