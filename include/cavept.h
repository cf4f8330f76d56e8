/* cavept -- C ABI of the B200-native drop-in for CaveEngine's progressive path
 * tracer pass.  Plain pointers and sizes only; no C++/torch/CUDA types.  Every
 * entry point cites the reference interface it replaces (paths relative to the
 * CaveEngine repository root).
 *
 * The library is CUDA-only: there is no CPU code path.  Every call that needs the
 * device returns CAVEPT_ERR_CUDA when no usable sm_100 device is present.
 *
 * Threading: a context is externally synchronised (the reference runs the whole
 * path on the main thread, cave/engine/runtime/application.cpp:269-272).  Distinct
 * contexts (one per GPU / per process) are independent.
 *
 * Ownership: the caller owns every host pointer; the library copies during the call
 * and never retains host pointers.  The context owns all device memory.
 */
#ifndef CAVEPT_H_
#define CAVEPT_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define CAVEPT_API __attribute__((visibility("default")))
#else
#define CAVEPT_API
#endif

typedef struct cavept_ctx cavept_ctx;

/* Return codes.  The reference traps (DEV_ASSERT / CRASH_NOW_MSG,
 * cave/engine/core/error/error_macros.h:88-125) where this ABI returns an error. */
enum {
    CAVEPT_OK = 0,
    CAVEPT_ERR_INVALID_ARG = 1,
    CAVEPT_ERR_DEGENERATE_GEOMETRY = 2, /* bvh_accel.cpp:117-120 (depth > 64), :138-141 (empty split side) */
    CAVEPT_ERR_CUDA = 3,
    CAVEPT_ERR_NOT_READY = 4, /* render/read before a scene is resident */
    CAVEPT_ERR_CAPACITY = 5,
    CAVEPT_ERR_NCCL = 6 /* multi-GPU group: NCCL not loadable, a collective failed, or a peer died (cavept_sync polls) */
};

/* enum class PathTracerMode, cave/engine/renderer/path_tracer_render_system.h:10-14 */
enum { CAVEPT_MODE_NONE = 0, CAVEPT_MODE_INTERACTIVE = 1, CAVEPT_MODE_TILED = 2 };

/* Which traversal the render launch uses.
 * REFERENCE_WALK reproduces HitScene (cave/shader/shared_path_tracer.h:152-189)
 * literally on the device: every mesh, stackless hit/miss-link walk over the 48-byte
 * GpuPtBvh array (needs CAVEPT_BVH_PARITY).  FAST is the production kernel: one
 * world-space tree over every (instance, triangle) pair, near-first ordered traversal
 * with a shared-memory stack over 32-byte nodes, pre-gathered world-space triangles.
 * Both return the same closest hit (up to exact-tie order, SURVEY.md section 8a note ii,
 * and fp32 rounding of the instance transform for non-identity instances). */
enum { CAVEPT_TRACE_FAST = 0, CAVEPT_TRACE_REFERENCE_WALK = 1 };

/* Which acceleration structures a scene upload builds (always on the device).
 * PARITY: the reference-layout per-mesh BVHs, bit-exact with BvhBuilder::ConstructHelper
 * (cave/engine/renderer/path_tracer/bvh_accel.cpp:115-219) -- what cavept_get_mesh_bvh /
 * cavept_get_scene_arrays read back and REFERENCE_WALK traverses -- AND the production tree.
 * FAST: the production tree only (binned-SAH world-space builder), for renders and
 * per-frame rebuilds (the reference never rebuilds: path_tracer.cpp:69).
 * LBVH: the production tree only, built by Morton codes + radix sort + a bottom-up AABB
 * reduction (several times faster to build than FAST, somewhat slower to trace): for
 * geometry that changes every frame (BASELINE config 5).
 * HYBRID: the production tree only: LBVH's radix tree down to treelets of at most 768 triangles, each of which is then
 * rebuilt by the binned-SAH builder's subtree launch (one thread block per treelet, out of shared memory).  Nearly FAST's
 * tree quality (config 3's scene: 74.3 node records per segment against 72.6 / LBVH 80.3) at a build cost that does not
 * pay one pass over all triangles per level: for large static scenes (several million triangles). */
enum { CAVEPT_BVH_PARITY = 0, CAVEPT_BVH_FAST = 1, CAVEPT_BVH_LBVH = 2, CAVEPT_BVH_HYBRID = 3 };

/* Which production kernel CAVEPT_TRACE_FAST launches.  Same tree, same per-pixel arithmetic in the same per-pixel order:
 * POOL, PIXEL and QUEUE produce bit-identical images (tests/test_gpu_parity.py).
 * PIXEL (default): one thread per pixel for the whole launch on a grid of pixel tiles; a warp is an 8x4 tile for all
 *        frames of the launch.  Measured fastest on B200.
 * QUEUE: persistent warps; a lane owns one pixel at a time for all frames of the launch and draws the next pixel from
 *        a global job queue when it is done (no launch tail, but a warp's lanes drift onto different tiles: 12 % slower
 *        at 256 frames per launch);  csrc/trace_body.h body_trace_pixel_fast<kQueue>.
 * POOL:  persistent warps, job queue, TWO pixel streams per lane with the inactive one parked in shared memory
 *        (csrc/trace_pool_body.h) -- measured slower on B200, kept selectable as the experiment it is.
 * WIDE4: PIXEL's launch shape over four-wide node records (each sibling pair with its interior children replaced by their
 *        own pairs: 64 bytes, two loads from one line, four slab tests, children sorted by entry distance): half the
 *        dependent node fetches per segment.  Same closest hits; a lane's visit order differs, so an image can differ
 *        from the pair walk's where two triangles tie in t (tests/test_gpu_parity.py states the bound).  Measured 11 %
 *        slower than PIXEL on B200 (two loads from one line cost two L1 wavefronts); its records are built when the
 *        variant is selected, not before. */
enum { CAVEPT_KERNEL_POOL = 0, CAVEPT_KERNEL_PIXEL = 1, CAVEPT_KERNEL_QUEUE = 2, CAVEPT_KERNEL_WIDE4 = 3 };

typedef struct cavept_config {
    int device;     /* CUDA ordinal; -1 = current device */
    int width;      /* accumulator size; reference default 1920x1080 (renderer/graphics_dvars.h:3) */
    int height;
    int max_bounce; /* MAX_BOUNCE, a #define 5 in the reference (shared_path_tracer.h:25); run-time here */
    int trace_mode; /* CAVEPT_TRACE_* */
    int bvh_mode;   /* CAVEPT_BVH_* */
} cavept_config;

/* MeshAsset fields the uploader reads (cave/engine/assets/mesh_asset.h:44-46). */
typedef struct cavept_mesh {
    const float* positions;  /* vertex_count * 3, tightly packed */
    const float* normals;    /* vertex_count * 3 */
    const uint32_t* indices; /* index_count, 3 per triangle */
    int vertex_count;
    int index_count;
} cavept_mesh;

/* One MeshRendererComponent + TransformComponent
 * (cave/engine/scene/mesh_renderer_component.h:21-44, transform_component.h:26,41).
 * As in the reference (path_tracer.cpp:227-231) the material of the FIRST instance
 * that references a mesh is used by every instance of that mesh. */
typedef struct cavept_instance {
    int mesh;
    int material;
    float transform[16]; /* world matrix, column-major (glm::mat4) */
} cavept_instance;

/* MaterialComponent fields (cave/engine/scene/material_component.h:12-21); the device
 * copy is GpuPtMaterial with emissive = base_color * emissive (path_tracer.cpp:140-145). */
typedef struct cavept_material {
    float base_color[3];
    float roughness;
    float metallic;
    float emissive;
} cavept_material;

typedef struct cavept_stats {
    uint64_t segments;        /* path segments (HitScene calls) traced since the last reset of stats */
    uint64_t node_visits;     /* filled only when CAVEPT_STATS_COUNT_VISITS was requested */
    uint64_t triangle_tests;
    uint64_t kernel_launches; /* our __global__ launches since context creation */
    float last_build_ms;      /* device time of the last reference-layout BVH build (all meshes), CUDA events; 0 in CAVEPT_BVH_FAST */
    float last_fast_build_ms; /* device time of the last production (world-space) BVH build */
    float last_render_ms;     /* device time of the last cavept_render, CUDA events */
    int triangle_count;
    int node_count;           /* GpuPtBvh nodes in the reference layout */
    int fast_node_count;      /* 32-byte nodes of the production layout */
    int instance_count;
    int build_levels;
    int build_launches;
    int fast_build_levels;
    int fast_build_launches;
    float last_update_instances_ms; /* device time of the last cavept_update_instances (re-transform + refit, or rebuild) */
    int last_update_was_refit;      /* 1: the last instance update refitted the tree; 0: it fell back to a rebuild */
    uint64_t collectives;           /* NCCL collectives enqueued by the multi-GPU group calls */
} cavept_stats;

/* ---- lifetime; replaces the file-static PathTracer in
 *      cave/engine/renderer/path_tracer_render_system.cpp:8-10 ------------------- */
CAVEPT_API int cavept_create(const cavept_config* config, cavept_ctx** out_ctx);
CAVEPT_API void cavept_destroy(cavept_ctx* ctx);
CAVEPT_API const char* cavept_last_error(const cavept_ctx* ctx);
CAVEPT_API int cavept_get_image_size(const cavept_ctx* ctx, int* width, int* height); /* either may be NULL */

/* ---- SetPathTracerMode / IsPathTracerActive,
 *      path_tracer_render_system.h:17-18, path_tracer.cpp:312-322 ---------------- */
CAVEPT_API int cavept_set_mode(cavept_ctx* ctx, int mode);
CAVEPT_API int cavept_is_active(const cavept_ctx* ctx); /* 1 iff mode == INTERACTIVE and a BVH is resident */

/* ---- RequestPathTracerUpdate -> PathTracer::Update -> CreateAccelStructure
 *      (path_tracer.cpp:49-70, :203-310): builds every referenced mesh's BVH on the
 *      device, concatenates the scene-global arrays (:72-118) and uploads instances
 *      and materials (:120-201).  Resets the accumulator. --------------------------- */
CAVEPT_API int cavept_upload_scene(cavept_ctx* ctx, const cavept_mesh* meshes, int mesh_count,
                                   const cavept_instance* instances, int instance_count,
                                   const cavept_material* materials, int material_count);
/* The two dirty-driven re-uploads of UpdateAccelStructure (path_tracer.cpp:156-158, :197-199).
 * Instance count/mesh ids must match the uploaded scene; both mark the scene dirty.
 * cavept_update_instances does NOT rebuild: the reference's per-mesh trees live in mesh space, so there a moved
 * instance costs M x 144 bytes; the production tree is in world space, so the moved instances' triangle slots are
 * re-transformed and every box is refitted bottom-up, one launch per tree level (csrc/wbvh_body.h refit_wbvh; ~0.1 ms
 * at 262k triangles).  The refit is kept when the scene still fits the tree's quantisation grid and the surface-area
 * sum of the tree stayed within 1.3x of the freshly built tree's; otherwise (and always with CAVEPT_BVH_LBVH, whose
 * rebuild is as cheap) the tree is rebuilt.  cavept_stats.last_update_was_refit / last_update_instances_ms report which
 * and how long.  cavept_set_refit(ctx, 0) forces the rebuild (measurement / A-B). */
CAVEPT_API int cavept_update_instances(cavept_ctx* ctx, const cavept_instance* instances, int instance_count);
CAVEPT_API int cavept_update_materials(cavept_ctx* ctx, const cavept_material* materials, int material_count);
CAVEPT_API int cavept_set_refit(cavept_ctx* ctx, int enable);
/* New vertex positions (and normals, may be NULL) for one mesh + BVH rebuild on the
 * device (BASELINE config 5; beyond the reference, path_tracer.cpp:69 "@TODO"). */
CAVEPT_API int cavept_rebuild_mesh(cavept_ctx* ctx, int mesh, const float* positions, const float* normals);

/* ---- textured materials, alpha cut-out and an HDR environment map: the half of the reference's path tracer that is
 *      switched off in its source (`#if 0`: cave/shader/glsl/path_tracer.cs.glsl:62-112, the material / texture set-up
 *      cave/engine/renderer/path_tracer/path_tracer.cpp:254-273), built here as that code states it:
 *        hit: base-colour map present -> sample it at the interpolated UV; alpha <= 0.01 -> the segment passes straight
 *             through (origin moves to the hit point, one bounce is consumed, no random numbers are drawn); otherwise the
 *             sampled rgb IS the diffuse colour (not multiplied by the material's albedo, :76-84); metal-rough map
 *             present -> metallic = its .b, roughness = its .g (:88-92);
 *        miss: radiance += environment(equirectangular uv of the direction) * throughput instead of the constant 0.3 sky (:103-105).
 *      Textures are sampled bilinearly with repeat wrap (GL's default `texture()` on a mip-less image).  The reference
 *      cannot run this path (its structs no longer exist), so there is nothing to pin it against: parity unpinned.  With no
 *      texture and no environment map set, none of this is compiled into the launch (the default kernels are unchanged).
 *      Only the per-pixel FAST kernel has the textured instantiation. ------------------------------------------------ */
typedef struct cavept_texture {
    const void* pixels; /* width * height texels, row 0 first */
    int width, height;
    int format;         /* CAVEPT_TEX_RGBA8 (unorm) or CAVEPT_TEX_RGBA32F */
} cavept_texture;
enum { CAVEPT_TEX_RGBA8 = 0, CAVEPT_TEX_RGBA32F = 1 };
/* Per-vertex texture coordinates of one mesh (vertex_count * 2 floats; MeshAsset::texcoords_0, mesh_asset.h:47). */
CAVEPT_API int cavept_set_mesh_uvs(cavept_ctx* ctx, int mesh, const float* uvs);
/* Replaces the texture set (count 0 clears it and every material's maps). */
CAVEPT_API int cavept_upload_textures(cavept_ctx* ctx, const cavept_texture* textures, int texture_count);
/* Per material: index of its base-colour map and of its metal-rough map in the texture set, -1 = none
 * (MaterialComponent::TEXTURE_BASE / TEXTURE_METALLIC_ROUGHNESS, path_tracer.cpp:270-272). */
CAVEPT_API int cavept_set_material_maps(cavept_ctx* ctx, const int* base_color_maps, const int* material_maps, int material_count);
/* Equirectangular environment map, width * height * 3 floats (c_hdrEnvMap); NULL restores the constant sky. */
CAVEPT_API int cavept_set_env_map(cavept_ctx* ctx, const float* rgb, int width, int height);

/* ---- parity read-backs of what BindPathTracerData binds (path_tracer.cpp:324-331):
 *      the reference-layout arrays, structured_buffer.hlsl.h:24-61 ------------------ */
CAVEPT_API int cavept_get_mesh_bvh(cavept_ctx* ctx, int mesh, void* out_gpu_pt_bvh, int capacity_nodes); /* mesh-local indices, as FillGpuBvhAccel emits; returns count or <0 */
CAVEPT_API int cavept_get_scene_counts(cavept_ctx* ctx, int* vertices, int* triangles, int* nodes, int* instances, int* materials);
CAVEPT_API int cavept_get_scene_arrays(cavept_ctx* ctx, void* gpu_pt_vertices, void* gpu_pt_indices, void* gpu_pt_bvhs, void* gpu_pt_meshes, void* gpu_pt_materials); /* any may be NULL */
/* The production (world-space) tree as the builder emitted it, before quantisation: 32-byte records
 * {float min[3]; int ref; float max[3]; int count}, siblings adjacent (record 2k, 2k+1 = sibling pair k; an interior
 * ref = index of the child pair's first record, a leaf ref = sign bit | (count-1) << 29 | first triangle slot).
 * No counterpart in the reference (HitScene loops over meshes, shared_path_tracer.h:159); read back so that tests can
 * compare the device-built tree with the serial statement of the same builder.  Returns the record count or < 0. */
CAVEPT_API int cavept_get_fast_tree(cavept_ctx* ctx, void* out_records, int capacity_records);

/* ---- per-frame constants, PerFrameConstantBuffer c_camera* (cave/shader/cbuffer.hlsl.h:193-207),
 *      filled at cave/engine/runtime/render_system.cpp:89-93.  Marks the scene dirty
 *      (SCENE_DIRTY_CAMERA, cave/engine/scene/scene.cpp:60-63). -------------------- */
CAVEPT_API int cavept_set_camera(cavept_ctx* ctx, const float position[3], const float right[3], const float up[3],
                                 const float forward[3], float fov_degree);
/* c_sceneDirty = 1 for the next rendered frame (render_system.cpp:125): the next frame
 * overwrites the accumulator instead of averaging into it (path_tracer.cs.glsl:46-57). */
CAVEPT_API int cavept_reset(cavept_ctx* ctx);

/* ---- PathTracerPassFunc (cave/engine/render_graph/common_passes.cpp:977-999): one
 *      Dispatch per frame in the reference; here n_frames consecutive c_frameIndex
 *      values [first_frame_index, +n_frames) in one launch, accumulated with the
 *      reference's running mean.  Asynchronous on the context's stream. ------------- */
CAVEPT_API int cavept_render(cavept_ctx* ctx, int first_frame_index, int n_frames);
CAVEPT_API int cavept_sync(cavept_ctx* ctx);
/* PathTracerMode::TILED (path_tracer_render_system.h:13; "Not implemented" in the reference,
 * path_tracer.cpp:53-55) with c_tileOffset (cave/shader/cbuffer.hlsl.h:183): the same pass over the
 * pixel rectangle [x0, x0+w) x [y0, y0+h) only.  Seeds use the image coordinates, so the union of
 * tiles of one frame equals cavept_render of that frame bit for bit.  Allowed in mode INTERACTIVE
 * or TILED.  After cavept_reset the first tile launch empties the whole accumulator (weight 0), so
 * every tile's first sample overwrites, as c_sceneDirty does for a full frame.  Also the unit of
 * the image-tile split across GPUs. */
CAVEPT_API int cavept_render_tile(cavept_ctx* ctx, int x0, int y0, int w, int h, int first_frame_index, int n_frames);

/* ---- read-back.  The reference leaves RG_RES_PATHTRACER (RGBA32F, rgb = running
 *      mean, a = sample count; common_passes.cpp:1001-1009) on the GPU; this is the
 *      headless equivalent.  Pixel (x, y) at index y*width + x, as imageStore writes. */
CAVEPT_API int cavept_read_accum(cavept_ctx* ctx, float* out_rgba32f, size_t bytes);
/* Tone pass: c/(c+1) then pow(c, 1/2.2) (cave/shader/glsl/post_process.ps.glsl:72-76). */
CAVEPT_API int cavept_read_tonemapped_rgba8(cavept_ctx* ctx, uint8_t* out_rgba8, size_t bytes);
CAVEPT_API int cavept_read_tonemapped_rgba32f(cavept_ctx* ctx, float* out_rgba32f, size_t bytes);
/* The same in the reference's tone-target format, RT_FMT_TONE = R16G16B16A16_FLOAT
 * (cave/engine/render_graph/render_graph_defines.h:64, common_passes.cpp:1011-1039): 4 IEEE halves per pixel. */
CAVEPT_API int cavept_read_tonemapped_rgba16f(cavept_ctx* ctx, uint16_t* out_rgba16f, size_t bytes);
/* The tone-mapped frame as a binary PPM (P6): the headless stand-in for presenting RG_RES_POST_PROCESS. */
CAVEPT_API int cavept_write_ppm(cavept_ctx* ctx, const char* path);
/* Restore a saved accumulator (checkpoint/resume; the accumulator is the only state). */
CAVEPT_API int cavept_write_accum(cavept_ctx* ctx, const float* rgba32f, size_t bytes);

/* ---- measurement and multi-GPU plumbing ------------------------------------------ */
CAVEPT_API int cavept_get_stats(cavept_ctx* ctx, cavept_stats* out);
CAVEPT_API int cavept_reset_stats(cavept_ctx* ctx);
/* Device time of the last cavept_render WITHOUT blocking (cavept_get_stats synchronises the stream):
 * CAVEPT_ERR_NOT_READY while that launch is still running or before the first one. */
CAVEPT_API int cavept_last_render_ms(cavept_ctx* ctx, float* out_ms);
CAVEPT_API int cavept_set_count_visits(cavept_ctx* ctx, int enable); /* instrumented kernels: node visits / triangle tests */
CAVEPT_API int cavept_set_kernel_variant(cavept_ctx* ctx, int variant); /* CAVEPT_KERNEL_*; measurement / A-B comparison */
CAVEPT_API void* cavept_stream(cavept_ctx* ctx);            /* cudaStream_t the context launches on */
CAVEPT_API void* cavept_accum_device_ptr(cavept_ctx* ctx);  /* float4[width*height] */
/* Compositor support: converts the accumulator in place between
 * (mean, n) and (sum, n) so ranks can be combined with one sum-reduce of float4. */
CAVEPT_API int cavept_accum_to_sums(cavept_ctx* ctx);
CAVEPT_API int cavept_accum_from_sums(cavept_ctx* ctx);
/* ---- multi-GPU group: the frame compositor (SURVEY.md section 8b "Threading": "multi-GPU group object wraps N ctxs +
 *      NCCL comm"; section 8e).  New relative to the reference, which runs the pass on one GPU
 *      (cave/engine/renderer/graphics_manager.cpp:421-428).  One process (or thread) per GPU, each with its own
 *      context; rank 0 makes an id with cavept_group_unique_id and hands the 128 bytes to the other ranks by whatever
 *      channel the host application has (the engine's own IPC, a file, MPI ...); every rank then calls
 *      cavept_group_create (collective).  From then on the context accumulates (sum, n) instead of (mean, n), so a
 *      frame is composited by ONE ncclReduce / ncclAllReduce of width*height float4 over NVLink, on the context's
 *      stream, into a separate composite buffer (mean, n) on the destination rank(s); the local sums are untouched
 *      (progressive accumulation continues, compositing again later is valid).  NCCL is loaded at run time
 *      (dlopen "libnccl.so.2"; override with $CAVEPT_NCCL_LIBRARY); single-GPU users never load it.
 *      Errors: CAVEPT_ERR_NCCL; cavept_sync() polls ncclCommGetAsyncError while it waits, aborts the communicator
 *      and returns CAVEPT_ERR_NCCL if a peer failed, instead of hanging. ------------------------------------ */
#define CAVEPT_GROUP_ID_BYTES 128
CAVEPT_API int cavept_group_unique_id(void* out_id, size_t bytes);
CAVEPT_API int cavept_group_create(cavept_ctx* ctx, int rank, int nranks, const void* unique_id, size_t bytes); /* nranks == 1: no communicator, no NCCL */
CAVEPT_API int cavept_group_destroy(cavept_ctx* ctx);  /* back to (mean, n) */
CAVEPT_API int cavept_group_info(const cavept_ctx* ctx, int* rank, int* nranks, int* nccl_version);
/* Sample split (the default): rank r renders frames cavept_group_frame_block(first, n, r, nranks) of the
 * full image; then cavept_group_composite(ctx, dst) -- dst = a rank, or -1 for every rank (all-reduce). */
CAVEPT_API int cavept_group_frame_block(int first_frame, int n_frames, int rank, int nranks, int* out_first, int* out_count);
CAVEPT_API int cavept_group_composite(cavept_ctx* ctx, int dst_rank);
CAVEPT_API int cavept_read_composite(cavept_ctx* ctx, float* out_rgba32f, size_t bytes); /* (mean, n), synchronises */
CAVEPT_API void* cavept_composite_device_ptr(cavept_ctx* ctx);                            /* NULL until this rank holds a composite */
/* Image-tile split (latency-oriented alternative): rank r renders rows cavept_group_row_band(height, r, nranks, align)
 * with cavept_render_tile; cavept_group_gather_bands broadcasts every band in place (bit-exact: no arithmetic). */
CAVEPT_API int cavept_group_row_band(int height, int rank, int nranks, int align, int* out_y0, int* out_rows);
CAVEPT_API int cavept_group_gather_bands(cavept_ctx* ctx, int align);

/* Roofline probe: the rate (lane-gathers per second) at which this GPU serves random 32-byte LDG.E.256 gathers -- the
 * node / triangle access of the trace kernel -- from a working set of `working_set_bytes` (the scene's node + triangle
 * bytes: from L2; 64 KB..1 MB: from L1), `iters` dependent gathers per lane on a full grid.  bench.py reports the
 * trace kernel's achieved gather rate against this, measured in the same run. */
CAVEPT_API int cavept_measure_gather(cavept_ctx* ctx, size_t working_set_bytes, int iters, double* out_gathers_per_s);

/* The device radix sort the LBVH builder orders its Morton codes with, on host arrays (copied in,
 * sorted on the device by the low `key_bits` of the keys, copied back): stable, so the result is
 * defined bit for bit.  Exposed so the sort can be checked on its own at full size. */
CAVEPT_API int cavept_sort_pairs(cavept_ctx* ctx, uint32_t* keys, int32_t* values, int count, int key_bits);
/* The same sort for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * `d_keys` / `d_values` are DEVICE arrays of `count` entries, sorted in place, asynchronously on the context's stream
 * (cavept_sync, or order your own work behind cavept_stream).  The call allocates nothing: its scratch (ping-pong
 * buffers, digit histograms, scan sums) is one static __device__ array of about 8.5 MB inside the library, held by every
 * GPU that loads the library until the process exits.  count > CAVEPT_STATIC_SORT_MAX_PAIRS returns
 * CAVEPT_ERR_CAPACITY (cavept_last_error says so).  Every call on a GPU shares that array: calls that overlap on the GPU
 * (different streams / contexts / threads) race on it -- serialise them.  Nothing in the array needs to start at zero;
 * each pass overwrites what it reads, so back-to-back calls of different sizes are independent.  Same result, bit for
 * bit, as cavept_sort_pairs (csrc/cavept_static.cu). */
#define CAVEPT_STATIC_SORT_MAX_PAIRS (1 << 20)
CAVEPT_API int cavept_sort_pairs_static(cavept_ctx* ctx, uint32_t* d_keys, int32_t* d_values, int count, int key_bits);

/* ---- asset ingest: the engine's own files -> the flat arrays of
 *      cavept_upload_scene.  Host-side parsing only; nothing here touches the device.
 *      .mesh  = MeshAsset::SerializeBinary (cave/engine/assets/mesh_asset.cpp:45-61) over Archive's
 *               "u64 count + raw elements" vectors (cave/engine/core/io/archive.h:56-75);
 *      .scene = the entity YAML of Scene serialisation (cave/engine/scene/scene.cpp:370,437); world
 *               matrices follow TransformComponent::GetLocalMatrix (scene/transform_component.cpp:16-21)
 *               and UpdateHierarchy (systems/ecs_systems.cpp:139-171) in glm's fp32 operation order;
 *               one instance per entity with MeshRendererComponent + TransformComponent, material =
 *               materials[0] (path_tracer.cpp:227-231); mesh GUIDs resolve through `*.mesh.meta`
 *               side-cars found under resource_root (NULL = the scene file's directory).  Renderers
 *               whose mesh does not resolve are skipped, as DEV_VERIFY does (path_tracer.cpp:216),
 *               and counted.  `err` (may be NULL) receives a message on failure. ------------------- */
/* Page-locked host memory for per-frame inputs (BASELINE config 5 streams 30 MB of vertex positions per
 * frame): arrays handed to cavept_rebuild_mesh / cavept_upload_scene from such a buffer are DMA-copied at
 * PCIe rate instead of being staged through the driver's bounce buffer.  Ordinary pointers keep working. */
CAVEPT_API void* cavept_host_alloc(size_t bytes);
CAVEPT_API void cavept_host_free(void* ptr);

/* Unit probe: runs the device's HitTriangle (shared_path_tracer.h:88-130) and
 * HitBvh (:133-149) on `count` caller-supplied cases, for the cross-check against the engine's CPU twins
 * TestIntersection::RayTriangle / RayAabb (cave/engine/math/intersection.cpp:18-70).
 * rays: count x 7 (origin xyz, direction xyz, t); triangles: count x 9 (A, B, C); boxes: count x 6 (min, max);
 * out: count x 8 (triangle hit 0/1, t after the test, u, v, box hit 0/1, slab tmin, slab tmax, 0). */
CAVEPT_API int cavept_probe_intersections(cavept_ctx* ctx, int count, const float* rays, const float* triangles, const float* boxes, float* out);

typedef struct cavept_scene_file cavept_scene_file;
CAVEPT_API int cavept_scene_file_open(const char* scene_path, const char* resource_root, cavept_scene_file** out, char* err, size_t err_capacity);
CAVEPT_API int cavept_mesh_file_open(const char* mesh_path, cavept_scene_file** out, char* err, size_t err_capacity); /* one mesh, identity instance, default material */
CAVEPT_API void cavept_scene_file_close(cavept_scene_file* scene);
CAVEPT_API int cavept_scene_file_counts(const cavept_scene_file* scene, int* meshes, int* instances, int* materials, int* skipped_renderers);
CAVEPT_API const cavept_mesh* cavept_scene_file_meshes(const cavept_scene_file* scene);       /* valid until close */
CAVEPT_API const cavept_instance* cavept_scene_file_instances(const cavept_scene_file* scene);
CAVEPT_API const cavept_material* cavept_scene_file_materials(const cavept_scene_file* scene);
CAVEPT_API int cavept_upload_scene_file(cavept_ctx* ctx, const cavept_scene_file* scene);     /* == cavept_upload_scene on the arrays above */
/* Writes a mesh in the engine's .mesh layout (one subset, RENDERABLE, no tangents/uvs/skin). */
CAVEPT_API int cavept_mesh_file_write(const char* path, const cavept_mesh* mesh);

#ifdef __cplusplus
}
#endif
#endif /* CAVEPT_H_ */
