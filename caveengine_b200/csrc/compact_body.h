// Production layout ("compact") derived on the device from the parity-mode BVH.
//
// The reference tree has one triangle per leaf and median-splits every set of <= 4
// triangles (bvh_accel.cpp:138-142).  The production layout keeps the SAH part of the
// topology bit-for-bit (same boxes) and collapses every subtree of <= kFastLeafMax
// triangles into one leaf whose triangles are pre-gathered in pre-order:
//   * 32-byte nodes, the two children of a node adjacent (one 64-byte fetch per step);
//   * triangles as 3 x float4 (48 B) instead of the reference's 16 B index fetch plus
//     three dependent 32 B vertex gathers (112 B).
// Node j of the reference array is "kept" iff it is a root or its parent holds more
// than kFastLeafMax triangles.  Positions: the children of the r-th kept interior node
// (r = exclusive scan in pre-order) sit at 2r and 2r+1.  A mesh root has no record of its
// own (its box is never tested: the TLAS holds a padded superset); only its node
// reference is kept, in mesh_root_ref[m].
#pragma once
#include "cv_common.h"
#include "trace_body.h"

namespace cavept {

struct CompactArrays {
    const GpuPtBvh* nodes;
    const int* node_parent;  // >= 0: parent slot; -1 - m: root of the m-th built mesh
    const int* node_count;
    int n_nodes;
    const GpuPtVertex* vertices;
    const GpuPtIndex* indices;
    int* is_pair;     // per node: 1 iff kept interior
    int* pair_rank;   // exclusive scan of is_pair (n_nodes + 1)
    int* leaf_tris;   // per node: triangle count iff kept leaf else 0
    int* tri_base;    // exclusive scan of leaf_tris (n_nodes + 1)
    FastNode* fnodes;
    FastTri* ftris;
    FastTri* fnormals;    // per FastTri slot: the triangle's three vertex normals
    int* mesh_root_ref;   // per built mesh: node reference of its root
    int n_pairs;          // filled by the driver after the scan
};

CV_HD bool compact_kept(const CompactArrays& c, int j) {
    const int parent = c.node_parent[j];
    return parent < 0 || c.node_count[parent] > kFastLeafMax;
}

CV_D void body_compact_classify(const CompactArrays& c, int j) {
    if (j >= c.n_nodes) return;
    const bool kept = compact_kept(c, j);
    const int count = c.node_count[j];
    c.is_pair[j] = (kept && count > kFastLeafMax) ? 1 : 0;
    c.leaf_tris[j] = (kept && count <= kFastLeafMax) ? count : 0;
}

CV_D void body_compact_emit(const CompactArrays& c, int j) {
    if (j >= c.n_nodes) return;
    if (!compact_kept(c, j)) return;
    const int parent = c.node_parent[j];
    const int count = c.node_count[j];
    int ref;
    if (count > kFastLeafMax) {
        ref = 2 * c.pair_rank[j];
    } else {
        ref = make_leaf_ref(c.tri_base[j], count);
        // the subtree of j occupies slots [j, j + 2*count - 1) in pre-order
        int k = 0;
        for (int s = j; s < j + 2 * count - 1; ++s) {
            const int tri = c.nodes[s].triangleIndex;
            if (tri < 0) continue;
            const int* idx = c.indices[tri].tri;
            const GpuPtVertex va = c.vertices[idx[0]], vb = c.vertices[idx[1]], vc = c.vertices[idx[2]];
            FastTri t;
            t.v0.x = va.position[0]; t.v0.y = va.position[1]; t.v0.z = va.position[2]; t.v0.w = u2f((uint32_t)tri);
            t.v1.x = vb.position[0]; t.v1.y = vb.position[1]; t.v1.z = vb.position[2]; t.v1.w = 0.0f;
            t.v2.x = vc.position[0]; t.v2.y = vc.position[1]; t.v2.z = vc.position[2]; t.v2.w = 0.0f;
            c.ftris[c.tri_base[j] + k] = t;
            t.v0.x = va.normal[0]; t.v0.y = va.normal[1]; t.v0.z = va.normal[2];
            t.v1.x = vb.normal[0]; t.v1.y = vb.normal[1]; t.v1.z = vb.normal[2];
            t.v2.x = vc.normal[0]; t.v2.y = vc.normal[1]; t.v2.z = vc.normal[2];
            c.fnormals[c.tri_base[j] + k] = t;
            ++k;
        }
    }
    if (parent < 0) {
        c.mesh_root_ref[-1 - parent] = ref;
        return;
    }
    const GpuPtBvh src = c.nodes[j];
    FastNode out;
    out.mn[0] = src.min[0]; out.mn[1] = src.min[1]; out.mn[2] = src.min[2];
    out.mx[0] = src.max[0]; out.mx[1] = src.max[1]; out.mx[2] = src.max[2];
    out.a = ref;
    out.b = count;
    c.fnodes[2 * c.pair_rank[parent] + (j == parent + 1 ? 0 : 1)] = out;
}

struct FnCompactClassify {
    CompactArrays c;
    CV_D void operator()(int j) const { body_compact_classify(c, j); }
};
struct FnCompactEmit {
    CompactArrays c;
    CV_D void operator()(int j) const { body_compact_emit(c, j); }
};

// Runs the compaction; returns the number of sibling pairs (2*pairs nodes written).
template<class Exec>
int run_compact(Exec& ex, CompactArrays c) {
    ex.launch(c.n_nodes, FnCompactClassify{ c });
    ex.scan_i32(c.is_pair, c.pair_rank, c.n_nodes);
    ex.scan_i32(c.leaf_tris, c.tri_base, c.n_nodes);
    int n_pairs = 0;
    ex.download(&n_pairs, c.pair_rank + c.n_nodes, sizeof(int));
    c.n_pairs = n_pairs;
    ex.launch(c.n_nodes, FnCompactEmit{ c });
    return n_pairs;
}

}  // namespace cavept
