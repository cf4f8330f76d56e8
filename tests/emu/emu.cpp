// TEST CODE -- serial host emulation of the cavept kernel bodies.
// The kernel bodies under caveengine_b200/csrc/*_body.h are plain functions of a thread
// index; this file steps them one index at a time with the same host-side level loop the
// product uses (bvh_parity_driver.h, compact_body.h), so the algorithms can be checked
// against the oracle in a container without a GPU.  Built by tests/emu_loader.py with
// g++ -ffp-contract=off.  The product never links this.
#include <cstdlib>
#include <cstring>
#include <vector>

#include "bvh_parity_driver.h"
#include "compact_body.h"
#include "tlas_host.h"
#include "trace_body.h"

using namespace cavept;

struct EmuExec {
    unsigned long long launches = 0;
    template<class F>
    void launch(int n, F f) {
        for (int i = 0; i < n; ++i) f(i);
        ++launches;
    }
    void scan_u8(const unsigned char* in, int* out, int n) {
        int run = 0;
        for (int i = 0; i < n; ++i) { out[i] = run; run += in[i]; }
        out[n] = run;
    }
    void scan_i32(const int* in, int* out, int n) {
        int run = 0;
        for (int i = 0; i < n; ++i) { out[i] = run; run += in[i]; }
        out[n] = run;
    }
    void upload(void* dst, const void* src, size_t bytes) { std::memcpy(dst, src, bytes); }
    void download(void* dst, const void* src, size_t bytes) { std::memcpy(dst, src, bytes); }
    void fill(void* dst, int byte, size_t bytes) { std::memset(dst, byte, bytes); }
};

#define EMU_API extern "C" __attribute__((visibility("default")))

// Builds the scene-global GpuPtBvh array for meshes whose triangles occupy
// [tri_off[m], tri_off[m+1]) of `indices` (vertex ids already scene-global).
// Also returns per-node parent / subtree-count arrays.  rc 0 ok, -1 degenerate.
EMU_API int emu_bvh_build(const GpuPtVertex* vertices, const GpuPtIndex* indices, int n_tris, const int* tri_off, int n_meshes, GpuPtBvh* out_nodes,
                          int* out_parent, int* out_count, int* out_levels, int* out_error_bits) {
    const int n_nodes = 2 * n_tris - n_meshes;
    std::vector<F4> bmin(n_tris), bmax(n_tris), cent(n_tris);
    std::vector<int> order(n_tris), order_next(n_tris), node_of(n_tris), node_of_next(n_tris), scan(n_tris + 1), counters(4);
    std::vector<unsigned char> flags(n_tris);
    const size_t items_cap = (size_t)std::max(n_tris, n_meshes) + 2;
    std::vector<BuildItem> items(items_cap), items_next(items_cap), scratch(n_meshes);
    std::vector<ItemAccum> accum(items_cap);
    const int bins_cap = n_tris / 5 + n_meshes + 2;
    std::vector<ItemBins> bins(bins_cap);
    BuildArrays a{};
    a.vertices = vertices;
    a.indices = indices;
    a.n_tris = n_tris;
    a.tri_bmin = bmin.data();
    a.tri_bmax = bmax.data();
    a.tri_centroid = cent.data();
    a.order = order.data();
    a.order_next = order_next.data();
    a.node_of = node_of.data();
    a.node_of_next = node_of_next.data();
    a.flags = flags.data();
    a.scan = scan.data();
    a.items = items.data();
    a.items_next = items_next.data();
    a.accum = accum.data();
    a.bins = bins.data();
    a.counters = counters.data();
    a.bins_capacity = bins_cap;
    a.nodes = out_nodes;
    a.node_parent = out_parent;
    a.node_count = out_count;
    std::memset(out_nodes, 0xCD, sizeof(GpuPtBvh) * (size_t)n_nodes);
    EmuExec ex;
    BuildReport rep{};
    const int rc = build_bvh_parity(ex, a, tri_off, tri_off, n_meshes, scratch.data(), &rep);
    if (out_levels) *out_levels = rep.levels;
    if (out_error_bits) *out_error_bits = rep.error_bits;
    return rc;
}

// Production layout from the arrays emu_bvh_build returned.  fnodes needs
// n_nodes + n_meshes + 2*n_instances + 8 entries, ftris n_tris.  Returns pair count.
EMU_API int emu_compact(const GpuPtBvh* nodes, const int* parent, const int* count, int n_nodes, const GpuPtVertex* vertices, const GpuPtIndex* indices,
                        FastNode* fnodes, FastTri* ftris, FastTri* fnormals, int* mesh_root_ref) {
    std::vector<int> is_pair(n_nodes), pair_rank(n_nodes + 1), leaf_tris(n_nodes), tri_base(n_nodes + 1);
    CompactArrays c{};
    c.nodes = nodes;
    c.node_parent = parent;
    c.node_count = count;
    c.n_nodes = n_nodes;
    c.vertices = vertices;
    c.indices = indices;
    c.is_pair = is_pair.data();
    c.pair_rank = pair_rank.data();
    c.leaf_tris = leaf_tris.data();
    c.tri_base = tri_base.data();
    c.fnodes = fnodes;
    c.ftris = ftris;
    c.fnormals = fnormals;
    c.mesh_root_ref = mesh_root_ref;
    EmuExec ex;
    return run_compact(ex, c);
}

// TLAS exactly as cavept_api.cu builds it: root boxes = first node of each built mesh.
// mesh_of_instance[i] = built-mesh index; mesh_root_node[k] = GpuPtBvh slot of mesh k's root.
// Writes TLAS nodes at fnodes[base..]; returns number of TLAS nodes, root reference in *root_ref.
EMU_API int emu_tlas(const GpuPtBvh* nodes, const GpuPtMesh* meshes, int n_instances, const int* mesh_of_instance, const int* mesh_root_node, int base,
                     FastNode* fnodes, int* root_ref) {
    std::vector<WorldBox> boxes(n_instances);
    for (int i = 0; i < n_instances; ++i) {
        const GpuPtBvh& r = nodes[mesh_root_node[mesh_of_instance[i]]];
        boxes[i] = transform_box(meshes[i].transform, r.min, r.max);
    }
    std::vector<FastNode> tlas;
    build_tlas(boxes, base, tlas, root_ref);
    std::memcpy(fnodes + base, tlas.data(), tlas.size() * sizeof(FastNode));
    return (int)tlas.size();
}

EMU_API void emu_mat4_inverse(const float* m, float* out) { mat4_inverse(m, out); }

struct EmuCamera {
    float position[3];
    float fov_degree;
    float forward[3];
    float right[3];
    float up[3];
};

EMU_API long long emu_render(const GpuPtVertex* vertices, const GpuPtIndex* indices, const GpuPtBvh* bvhs, const GpuPtMesh* meshes, int mesh_count,
                             const GpuPtMaterial* materials, const FastNode* fnodes, const FastTri* ftris, const FastTri* fnormals, const int* inst_fast_root, int tlas_root,
                             const EmuCamera* cam, int max_bounce, int width, int height, int first_frame, int n_frames, int reset_first, float* accum,
                             int fast, unsigned long long* counters) {
    TraceParams p{};
    p.vertices = vertices;
    p.indices = indices;
    p.bvhs = bvhs;
    p.meshes = meshes;
    p.materials = materials;
    p.mesh_count = mesh_count;
    p.fnodes = fnodes;
    p.ftris = ftris;
    p.fnormals = fnormals;
    p.inst_fast_root = inst_fast_root;
    p.tlas_root = tlas_root;
    std::memcpy(p.cam_pos, cam->position, 12);
    std::memcpy(p.cam_right, cam->right, 12);
    std::memcpy(p.cam_up, cam->up, 12);
    std::memcpy(p.cam_fwd, cam->forward, 12);
    p.cam_distance = std::tan(cam->fov_degree * 3.141592653589793f / 180.0f);
    p.width = width;
    p.height = height;
    p.max_bounce = max_bounce;
    p.first_frame = first_frame;
    p.n_frames = n_frames;
    p.reset_first = reset_first;
    p.accum = reinterpret_cast<F4*>(accum);
    unsigned long long cnt[3] = { 0, 0, 0 };
    p.counters = cnt;
    long long segs = 0;
    int stack[kStackSize];
    for (int y = 0; y < height; ++y) {
        for (int x = 0; x < width; ++x) {
            segs += fast ? body_trace_pixel_fast<true>(p, x, y, stack, 1, cnt) : body_trace_pixel_reference<true>(p, x, y, cnt);
        }
    }
    if (counters) {
        counters[0] = (unsigned long long)segs;
        counters[1] = cnt[1];
        counters[2] = cnt[2];
    }
    return segs;
}
