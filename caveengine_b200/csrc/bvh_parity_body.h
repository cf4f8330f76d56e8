// Parity-mode BVH builder: a level-synchronous restatement of the reference's
// depth-first recursive builder, BvhBuilder::ConstructHelper
// (cave/engine/renderer/path_tracer/bvh_accel.cpp:115-219), that emits the exact
// bytes FillGpuBvhAccel (:237-263) + AppendBvhs (path_tracer.cpp:90-118) produce.
//
// How a recursion becomes level-synchronous without changing a single bit:
//  * Every subtree over n triangles has exactly 2n-1 nodes (1 triangle per leaf,
//    :128-132), and FillGpuBvhAccel writes nodes in creation (pre-) order, so a node
//    at array slot i with n_left triangles in its left child has children at i+1 and
//    i+2*n_left.  Slots are therefore known top-down; no recursion order is needed.
//  * DiscoverIdx (:221-235): miss(left child) = slot of the right sibling,
//    miss(right child) = miss(parent), miss(root) = -1; hit = left child, or miss for
//    a leaf.
//  * The triangle set of a node is an ordered list in the reference (stable
//    partition :203-214, insertion sort :101-113).  Here all lists of one level are
//    the segments of one array `order`, and the stable partition is an exclusive scan
//    of the "goes left" flags.
//  * Node AABBs are min/max folds; `a < b ? a : b` (common.h:50-58) keeps the LATER
//    operand on ties, which only matters for +0 vs -0.  The fold is done with 64-bit
//    atomic min/max on keys (ordered value, sequence position, sign bit) so the result
//    is the same bit pattern the sequential fold yields.
//  * All float expressions are written in the reference's association order and the
//    file is compiled without FMA contraction.
//
// Every meshes' BVH is built in the same pass: the roots are the level-0 items.
#pragma once
#include "cv_common.h"

namespace cavept {

constexpr int kBuckets = 12;     // BUCKED_MAX, bvh_accel.cpp:144
constexpr int kMaxDepth = 64;    // :117
constexpr float kBoxMinSize = 0.0001f;  // Box::BOX_MIN_SIZE, cave/engine/math/box.h:12
#ifndef CV_PARITY_SUB_MAX
#define CV_PARITY_SUB_MAX 1024
#endif
constexpr int kParitySubMax = CV_PARITY_SUB_MAX;  // an item of at most this many triangles leaves the level loop: one thread block builds its whole subtree (parity_subtree_*)

enum { ITEM_LEAF = 0, ITEM_SMALL = 1, ITEM_SAH = 2 };
enum { BUILD_ERR_DEPTH = 1, BUILD_ERR_EMPTY_SIDE = 2, BUILD_ERR_BAD_AREA = 4, BUILD_ERR_BINS = 8 };

struct BuildItem {
    int start, count;  // segment of `order`
    int node;          // scene-global GpuPtBvh slot
    int miss;          // scene-global miss link, or -1
    int parent;        // scene-global parent slot; -1 - m for the root of the m-th mesh
    int depth;
    int kind;
    int axis;
    float tmin, tmax, parent_sa;
    int split;
    int n_left;
    int child_slot;    // index of the left child's item in the next level, or -1: that child left the level loop (queued subtree root)
    int child_slot_r;  // the same for the right child
    int bin_slot;
    int tri_offset;    // first scene-global triangle of this item's mesh (AppendBvhs offset)
};

struct ItemAccum {  // fold state of AABBFromTriangles (:89-99) and the centroid box (:151-156)
    unsigned long long mn[3], mx[3];
    uint32_t cmn[3], cmx[3];
};

struct ItemBins {  // BucketInfo[12], :145-149
    int count[kBuckets];
    uint32_t mn[kBuckets][3];
    uint32_t mx[kBuckets][3];
};

struct BuildArrays {
    // scene-global inputs (reference layouts)
    const GpuPtVertex* vertices;
    const GpuPtIndex* indices;  // vertex ids already offset per mesh (AppendIndices, path_tracer.cpp:80-88)
    int n_tris;
    // per-triangle precompute (BvhBuilder ctor :57-74)
    F4* tri_bmin;
    F4* tri_bmax;
    F4* tri_centroid;
    // level state
    int* order;        // triangle ids, segmented by item
    int* order_next;
    int* node_of;      // item index of each element, -1 once its leaf is written
    int* node_of_next;
    unsigned char* flags;
    int* scan;         // exclusive scan of flags, n_tris + 1 entries
    BuildItem* items;
    BuildItem* items_next;
    ItemAccum* accum;
    ItemBins* bins;
    int* counters;     // [0] next-level item count, [1] bin slots used, [2] error bits, [3] subtree roots queued, [4] deepest level a subtree build reached + 1
    int bins_capacity;
    // second phase (parity_subtree_*): children of at most sub_max triangles are queued here instead of entering the next
    // level; null inside a subtree build and when the phase is switched off
    BuildItem* sub_roots;
    int sub_capacity;
    int sub_max;
    // outputs
    GpuPtBvh* nodes;
    int* node_parent;  // per node: parent slot (scene-global), or -1 - m for the root of mesh m
    int* node_count;   // per node: triangles in its subtree
};

CV_HD F3 load_pos(const GpuPtVertex* v, int i) { return f3(v[i].position[0], v[i].position[1], v[i].position[2]); }

// Box<3>::MakeValid, cave/engine/math/box.h:33-49
CV_HD void make_valid(F3& mn, F3& mx) {
    const F3 size = mx - mn;
    if (size.x == 0.0f) { mn.x -= kBoxMinSize; mx.x += kBoxMinSize; }
    if (size.y == 0.0f) { mn.y -= kBoxMinSize; mx.y += kBoxMinSize; }
    if (size.z == 0.0f) { mn.z -= kBoxMinSize; mx.z += kBoxMinSize; }
}

// Box<3>::SurfaceArea, cave/engine/math/box.cpp:7-17 (0 for any box with min >= max on an axis)
CV_HD float surface_area(F3 mn, F3 mx) {
    if (mn.x >= mx.x || mn.y >= mx.y || mn.z >= mx.z) return 0.0f;
    const F3 s = mx - mn;
    const float sx = fabsf(s.x), sy = fabsf(s.y), sz = fabsf(s.z);
    return 2.0f * (sx * sy + sx * sz + sy * sz);
}

// DominantAxis, bvh_accel.cpp:76-87
CV_HD int dominant_axis(F3 mn, F3 mx) {
    const F3 span = mx - mn;
    int axis = 0;
    if (comp(span, axis) < span.y) axis = 1;
    if (comp(span, axis) < span.z) axis = 2;
    return axis;
}

// ---- K1: per-triangle AABB + centroid (BvhBuilder ctor, :57-74) ---------------------
CV_D void body_tri_prep(const BuildArrays& a, int g) {
    if (g >= a.n_tris) return;
    const int* t = a.indices[g].tri;
    const F3 p0 = load_pos(a.vertices, t[0]), p1 = load_pos(a.vertices, t[1]), p2 = load_pos(a.vertices, t[2]);
    const F3 c = (1.0f / 3.0f) * ((p0 + p1) + p2);
    F3 mn = f3(CV_INF, CV_INF, CV_INF), mx = f3(-CV_INF, -CV_INF, -CV_INF);
    const F3 p[3] = { p0, p1, p2 };
    for (int k = 0; k < 3; ++k) {  // ExpandPoint, box.h:51-54
        mn = f3(min_ref(mn.x, p[k].x), min_ref(mn.y, p[k].y), min_ref(mn.z, p[k].z));
        mx = f3(max_ref(mx.x, p[k].x), max_ref(mx.y, p[k].y), max_ref(mx.z, p[k].z));
    }
    make_valid(mn, mx);
    F4 o;
    o.x = mn.x; o.y = mn.y; o.z = mn.z; o.w = 0.0f;
    a.tri_bmin[g] = o;
    o.x = mx.x; o.y = mx.y; o.z = mx.z;
    a.tri_bmax[g] = o;
    o.x = c.x; o.y = c.y; o.z = c.z;
    a.tri_centroid[g] = o;
}

CV_HD unsigned long long min_key(float v, uint32_t pos) {
    return ((unsigned long long)ordered_key(v) << 32) | (unsigned long long)(((0x7fffffffu - pos) << 1) | (f2u(v) >> 31));
}
CV_HD unsigned long long max_key(float v, uint32_t pos) {
    return ((unsigned long long)ordered_key(v) << 32) | (unsigned long long)((pos << 1) | (f2u(v) >> 31));
}
CV_HD float key_value(unsigned long long k) {
    const float v = ordered_value((uint32_t)(k >> 32));
    return (v == 0.0f) ? u2f(((uint32_t)k & 1u) << 31) : v;
}

CV_D void body_init_accum(const BuildArrays& a, int w, int n_items) {
    if (w >= n_items) return;
    ItemAccum& acc = a.accum[w];
    for (int k = 0; k < 3; ++k) {
        acc.mn[k] = ~0ull;
        acc.mx[k] = 0ull;
        acc.cmn[k] = 0xffffffffu;
        acc.cmx[k] = 0u;
    }
}

// ---- per element: fold vertices into the item's AABB and centroid box ---------------
// `acc` is the item's accumulator, or a block's private copy of it (kernels.cu k_parity_reduce): the folds are min / max
// of keys, so where they are combined does not change the result
// the element's own fold operands: (value, position, sign) keys of its three vertices per axis, already folded over
// the three, and the ordered keys of its centroid
CV_D void body_element_keys(const BuildArrays& a, int i, unsigned long long kmn[3], unsigned long long kmx[3], uint32_t ck[3]) {
    const int g = a.order[i];
    const int* t = a.indices[g].tri;
    for (int k = 0; k < 3; ++k) {
        const F3 p = load_pos(a.vertices, t[k]);
        const uint32_t pos = (uint32_t)i * 3u + (uint32_t)k;
        const unsigned long long n0 = min_key(p.x, pos), n1 = min_key(p.y, pos), n2 = min_key(p.z, pos);
        const unsigned long long x0 = max_key(p.x, pos), x1 = max_key(p.y, pos), x2 = max_key(p.z, pos);
        if (k == 0) {
            kmn[0] = n0; kmn[1] = n1; kmn[2] = n2;
            kmx[0] = x0; kmx[1] = x1; kmx[2] = x2;
        } else {
            kmn[0] = n0 < kmn[0] ? n0 : kmn[0]; kmn[1] = n1 < kmn[1] ? n1 : kmn[1]; kmn[2] = n2 < kmn[2] ? n2 : kmn[2];
            kmx[0] = x0 > kmx[0] ? x0 : kmx[0]; kmx[1] = x1 > kmx[1] ? x1 : kmx[1]; kmx[2] = x2 > kmx[2] ? x2 : kmx[2];
        }
    }
    const F4 c = a.tri_centroid[g];
    ck[0] = ordered_key(c.x);
    ck[1] = ordered_key(c.y);
    ck[2] = ordered_key(c.z);
}
CV_D void body_reduce_to(const BuildArrays& a, int i, int w, ItemAccum& acc) {
    unsigned long long kmn[3], kmx[3];
    uint32_t ck[3];
    body_element_keys(a, i, kmn, kmx, ck);
    // lanes of the warp that fold into the same node combine first (top levels: all 32 hit one node)
    const CvGroup grp = cv_group_by(w);
    for (int k = 0; k < 3; ++k) {
        cv_group_min_u64(grp, &acc.mn[k], kmn[k]);
        cv_group_max_u64(grp, &acc.mx[k], kmx[k]);
    }
    for (int k = 0; k < 3; ++k) {
        cv_group_min_u32(grp, &acc.cmn[k], ck[k]);
        cv_group_max_u32(grp, &acc.cmx[k], ck[k]);
    }
}
CV_D void body_reduce(const BuildArrays& a, int i) {
    if (i >= a.n_tris) return;
    const int w = a.node_of[i];
    if (w < 0) return;
    body_reduce_to(a, i, w, a.accum[w]);
}

// ---- per item: write the node, classify (:128-142), prepare the SAH pass ------------
// The item and its accumulator are copied in first and the item is written back once: through the references every field
// read had to wait for the store before it (the compiler cannot tell the arrays apart), a chain of L2 round trips.
CV_D void body_node_of(const BuildArrays& a, BuildItem& it, const ItemAccum& acc);
CV_D void body_node(const BuildArrays& a, int w, int n_items) {
    if (w >= n_items) return;
    BuildItem it = a.items[w];
    const ItemAccum acc = a.accum[w];
    body_node_of(a, it, acc);
    a.items[w] = it;
}
CV_D void body_node_of(const BuildArrays& a, BuildItem& it, const ItemAccum& acc) {
    F3 mn = f3(key_value(acc.mn[0]), key_value(acc.mn[1]), key_value(acc.mn[2]));
    F3 mx = f3(key_value(acc.mx[0]), key_value(acc.mx[1]), key_value(acc.mx[2]));
    make_valid(mn, mx);

    GpuPtBvh node;
    node.min[0] = mn.x; node.min[1] = mn.y; node.min[2] = mn.z;
    node.max[0] = mx.x; node.max[1] = mx.y; node.max[2] = mx.z;
    node.missIdx = it.miss;
    node.pad[0] = 0.0f;
    node.pad[1] = 0.0f;
    a.node_parent[it.node] = it.parent;
    a.node_count[it.node] = it.count;

    if (it.count == 1) {
        node.hitIdx = it.miss;  // leaf: hit == miss, :234
        node.leaf = 1;
        node.triangleIndex = a.order[it.start];  // already scene-global (AppendBvhs adds the triangle offset)
        a.nodes[it.node] = node;
        a.node_of[it.start] = -1;
        it.kind = ITEM_LEAF;
        it.n_left = 0;
        return;
    }
    node.hitIdx = it.node + 1;
    node.leaf = 0;
    node.triangleIndex = -1;
    a.nodes[it.node] = node;

    const float parent_sa = surface_area(mn, mx);
    if (it.count <= 4 || parent_sa == 0.0f) {  // SplitByAxis, :101-113
        if (!(parent_sa > 0.0f)) cv_atomic_or_i32(&a.counters[2], BUILD_ERR_BAD_AREA);
        const int axis = dominant_axis(mn, mx);
        // libstdc++ std::sort on < 16 elements is a (stable) insertion sort; key = AABB centre (:38-55)
        // (every index below is static once unrolled: keys and ids stay in registers and the loads go out together)
        int ids[4] = { 0, 0, 0, 0 };
        float keys[4] = { 0.0f, 0.0f, 0.0f, 0.0f };
        const int n = it.count <= 4 ? it.count : 4;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int k = 0; k < 4; ++k) {
            if (k < n) {
                const int g = a.order[it.start + k];
                const F4 bmn = a.tri_bmin[g], bmx = a.tri_bmax[g];
                const F3 center = 0.5f * (f3(bmn.x, bmn.y, bmn.z) + f3(bmx.x, bmx.y, bmx.z));
                ids[k] = g;
                keys[k] = comp(center, axis);
            }
        }
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int k = 1; k < 4; ++k) {  // insert element k into the sorted prefix: it moves down while it is strictly smaller
            bool moving = k < n;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
            for (int j = k - 1; j >= 0; --j) {
                moving = moving && keys[j + 1] < keys[j];
                if (moving) {
                    const float tk = keys[j]; keys[j] = keys[j + 1]; keys[j + 1] = tk;
                    const int ti = ids[j]; ids[j] = ids[j + 1]; ids[j + 1] = ti;
                }
            }
        }
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int k = 0; k < 4; ++k) {
            if (k < n) a.order[it.start + k] = ids[k];
        }
        it.kind = ITEM_SMALL;
        it.n_left = it.count / 2;
        return;
    }

    F3 cmn = f3(ordered_value(acc.cmn[0]), ordered_value(acc.cmn[1]), ordered_value(acc.cmn[2]));
    F3 cmx = f3(ordered_value(acc.cmx[0]), ordered_value(acc.cmx[1]), ordered_value(acc.cmx[2]));
    make_valid(cmn, cmx);  // :156
    it.kind = ITEM_SAH;
    it.axis = dominant_axis(cmn, cmx);
    it.tmin = comp(cmn, it.axis);
    it.tmax = comp(cmx, it.axis);
    it.parent_sa = parent_sa;
    const int slot = cv_atomic_add_i32(&a.counters[1], 1);
    if (slot >= a.bins_capacity) {
        cv_atomic_or_i32(&a.counters[2], BUILD_ERR_BINS);
        it.kind = ITEM_SMALL;  // keeps the pass well-defined; the error bit fails the build
        it.n_left = it.count / 2;
        return;
    }
    it.bin_slot = slot;
    ItemBins& b = a.bins[slot];
    for (int k = 0; k < kBuckets; ++k) {
        b.count[k] = 0;
        for (int c = 0; c < 3; ++c) {
            b.mn[k][c] = 0xffffffffu;
            b.mx[k][c] = 0u;
        }
    }
}

CV_HD int clamp_bucket(int s) { return s < 0 ? 0 : (s > kBuckets - 1 ? kBuckets - 1 : s); }

// ---- per element of SAH items: bucket counts and boxes (:162-169) -------------------
// `b` is the item's bucket record, or a block's private copy of it (kernels.cu k_parity_bin)
// the bucket of element i of a bucketed item, first slot formula (:162-169), and its triangle
CV_D int body_bin_slot(const BuildArrays& a, int i, const BuildItem& it, int& g) {
    g = a.order[i];
    const F4 c = a.tri_centroid[g];
    const float cv = it.axis == 0 ? c.x : (it.axis == 1 ? c.y : c.z);
    const float tmp = cv_div((cv - it.tmin) * (float)kBuckets, it.tmax - it.tmin);
    return clamp_bucket((int)tmp);
}
CV_D void body_bin_to(const BuildArrays& a, int i, const BuildItem& it, ItemBins& b) {
    int g;
    const int slot = body_bin_slot(a, i, it, g);
    const F4 bmn = a.tri_bmin[g], bmx = a.tri_bmax[g];
    // lanes of the warp that land in the same bucket combine first (match.any + redux.sync): one atomic per word and
    // group instead of seven per triangle -- count / min / max are order-free, so the bucket contents are the same bits
    const CvGroup grp = cv_group_by(it.bin_slot * kBuckets + slot);
    cv_group_add_i32(grp, &b.count[slot], 1);
    cv_group_min_u32(grp, &b.mn[slot][0], ordered_key(bmn.x));
    cv_group_min_u32(grp, &b.mn[slot][1], ordered_key(bmn.y));
    cv_group_min_u32(grp, &b.mn[slot][2], ordered_key(bmn.z));
    cv_group_max_u32(grp, &b.mx[slot][0], ordered_key(bmx.x));
    cv_group_max_u32(grp, &b.mx[slot][1], ordered_key(bmx.y));
    cv_group_max_u32(grp, &b.mx[slot][2], ordered_key(bmx.z));
}
CV_D void body_bin(const BuildArrays& a, int i) {
    if (i >= a.n_tris) return;
    const int w = a.node_of[i];
    if (w < 0) return;
    const BuildItem& it = a.items[w];
    if (it.kind != ITEM_SAH) return;
    body_bin_to(a, i, it, a.bins[it.bin_slot]);
}

// ---- per SAH item: the 11 candidate costs and the first strict minimum (:171-196) ---
// Cost of splitting after bucket i (:173-193): both sides folded in bucket order.  Every word of the record is read
// first and every index below is static once unrolled, so the record lives in registers: read bucket by bucket between
// stores to a local array, each load waited for the store before it (84 dependent round trips to L2 per item).
CV_D float body_split_cost(const ItemBins& src, int i, float parent_sa) {
    ItemBins b;
    {
        constexpr int kWords = (int)(sizeof(ItemBins) / sizeof(uint32_t));
        const uint32_t* in = reinterpret_cast<const uint32_t*>(&src);
        uint32_t* out = reinterpret_cast<uint32_t*>(&b);
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int q = 0; q < kWords; ++q) out[q] = in[q];
    }
    F3 mn0 = f3(CV_INF, CV_INF, CV_INF), mx0 = f3(-CV_INF, -CV_INF, -CV_INF);
    F3 mn1 = mn0, mx1 = mx0;
    int c0 = 0, c1 = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int j = 0; j < kBuckets; ++j) {
        const int cnt = b.count[j];
        // an untouched bucket decodes to (+inf, -inf) == Box::MakeInvalid (box.h:19-22)
        F3 bmn = f3(ordered_value(b.mn[j][0]), ordered_value(b.mn[j][1]), ordered_value(b.mn[j][2]));
        F3 bmx = f3(ordered_value(b.mx[j][0]), ordered_value(b.mx[j][1]), ordered_value(b.mx[j][2]));
        if (cnt == 0) {
            bmn = f3(CV_INF, CV_INF, CV_INF);
            bmx = f3(-CV_INF, -CV_INF, -CV_INF);
        }
        if (j <= i) {
            mn0 = f3(min_ref(mn0.x, bmn.x), min_ref(mn0.y, bmn.y), min_ref(mn0.z, bmn.z));
            mx0 = f3(max_ref(mx0.x, bmx.x), max_ref(mx0.y, bmx.y), max_ref(mx0.z, bmx.z));
            c0 += cnt;
        } else {
            mn1 = f3(min_ref(mn1.x, bmn.x), min_ref(mn1.y, bmn.y), min_ref(mn1.z, bmn.z));
            mx1 = f3(max_ref(mx1.x, bmx.x), max_ref(mx1.y, bmx.y), max_ref(mx1.z, bmx.z));
            c1 += cnt;
        }
    }
    return 0.125f + cv_div((float)c0 * surface_area(mn0, mx0) + (float)c1 * surface_area(mn1, mx1), parent_sa);
}
CV_D void body_split(const BuildArrays& a, int w, int n_items) {
    if (w >= n_items) return;
    BuildItem& it = a.items[w];
    if (it.kind != ITEM_SAH) return;
    const ItemBins& b = a.bins[it.bin_slot];
    int split = 0;
    float best = 0.0f;
    for (int i = 0; i < kBuckets - 1; ++i) {
        const float cost = body_split_cost(b, i, it.parent_sa);
        if (i == 0 || cost < best) {
            split = i;
            best = cost;
        }
    }
    it.split = split;
}

// ---- per element: "goes left" flag; SAH items use the second slot formula (:203-214)
CV_D void body_flag(const BuildArrays& a, int i) {
    if (i >= a.n_tris) return;
    const int w = a.node_of[i];
    unsigned char f = 0;
    if (w >= 0) {
        const BuildItem& it = a.items[w];
        if (it.kind == ITEM_SAH) {
            const int g = a.order[i];
            const F4 c = a.tri_centroid[g];
            const float cv = it.axis == 0 ? c.x : (it.axis == 1 ? c.y : c.z);
            float tmp = cv_div(cv - it.tmin, it.tmax - it.tmin);
            tmp *= (float)kBuckets;
            f = clamp_bucket((int)tmp) <= it.split ? 1 : 0;
        } else if (it.kind == ITEM_SMALL) {
            f = (i - it.start) < it.n_left ? 1 : 0;  // already sorted in place
        }
    }
    a.flags[i] = f;
}

// ---- per item: emit the two children into the next level ---------------------------
CV_D void body_children_of(const BuildArrays& a, BuildItem& it);
CV_D void body_children(const BuildArrays& a, int w, int n_items) {
    if (w >= n_items) return;
    BuildItem it = a.items[w];
    if (it.kind == ITEM_LEAF) return;
    body_children_of(a, it);
    a.items[w] = it;
}
CV_D void body_children_of(const BuildArrays& a, BuildItem& it) {
    const int n_left = a.scan[it.start + it.count] - a.scan[it.start];
    it.n_left = n_left;
    if (n_left <= 0 || n_left >= it.count) {  // empty side: the reference recurses into an empty set and traps (:139)
        cv_atomic_or_i32(&a.counters[2], BUILD_ERR_EMPTY_SIDE);
        it.kind = ITEM_LEAF;
        return;
    }
    if (it.depth + 1 > kMaxDepth) {  // :117-120
        cv_atomic_or_i32(&a.counters[2], BUILD_ERR_DEPTH);
        it.kind = ITEM_LEAF;
        return;
    }
    BuildItem l, r;
    l.start = it.start;
    l.count = n_left;
    l.node = it.node + 1;
    l.miss = it.node + 2 * n_left;
    r.start = it.start + n_left;
    r.count = it.count - n_left;
    r.node = it.node + 2 * n_left;
    r.miss = it.miss;
    l.parent = r.parent = it.node;
    l.depth = r.depth = it.depth + 1;
    l.kind = r.kind = ITEM_LEAF;
    l.axis = r.axis = 0;
    l.tmin = r.tmin = l.tmax = r.tmax = l.parent_sa = r.parent_sa = 0.0f;
    l.split = r.split = l.n_left = r.n_left = 0;
    l.child_slot = r.child_slot = l.child_slot_r = r.child_slot_r = l.bin_slot = r.bin_slot = -1;
    l.tri_offset = r.tri_offset = it.tri_offset;
    // a child small enough to be a closed problem for one thread block leaves the level loop (its elements are parked:
    // node_of = -1, carried along by body_scatter until the subtree phase picks the range up)
    const bool queue_l = a.sub_roots && l.count <= a.sub_max, queue_r = a.sub_roots && r.count <= a.sub_max;
    int n_queued = (queue_l ? 1 : 0) + (queue_r ? 1 : 0);
    if (n_queued) {
        int q = cv_atomic_add_i32(&a.counters[3], n_queued);
        if (q + n_queued > a.sub_capacity) {
            cv_atomic_or_i32(&a.counters[2], BUILD_ERR_BINS);
            it.kind = ITEM_LEAF;
            return;
        }
        if (queue_l) a.sub_roots[q++] = l;
        if (queue_r) a.sub_roots[q] = r;
    }
    int slot = n_queued < 2 ? cv_atomic_add_i32(&a.counters[0], 2 - n_queued) : 0;
    it.child_slot = it.child_slot_r = -1;
    if (!queue_l) {
        it.child_slot = slot;
        a.items_next[slot++] = l;
    }
    if (!queue_r) {
        it.child_slot_r = slot;
        a.items_next[slot] = r;
    }
}

// ---- per element: stable partition into the children's segments --------------------
CV_D void body_scatter(const BuildArrays& a, int i) {
    if (i >= a.n_tris) return;
    const int w = a.node_of[i];
    if (w < 0) {
        a.node_of_next[i] = -1;
        a.order_next[i] = a.order[i];
        return;
    }
    const BuildItem& it = a.items[w];
    if (it.kind == ITEM_LEAF) {  // only reachable on a failed build
        a.node_of_next[i] = -1;
        a.order_next[i] = a.order[i];
        return;
    }
    const int lefts_before = a.scan[i] - a.scan[it.start];
    int dst, child;
    if (a.flags[i]) {
        dst = it.start + lefts_before;
        child = it.child_slot;
    } else {
        dst = it.start + it.n_left + ((i - it.start) - lefts_before);
        child = it.child_slot_r;
    }
    a.order_next[dst] = a.order[i];
    a.node_of_next[dst] = child;
}

// ---- second phase: subtree builds ---------------------------------------------------------------------------------
// An item of at most sub_max triangles is a closed problem: its elements are one contiguous range of `order` that
// nothing else touches, its nodes one contiguous range of slots.  One thread block runs the SAME per-element / per-item
// passes over that range, level after level, with block barriers in place of kernel boundaries (kernels.cu
// k_parity_subtrees), so the deepest ~15 levels of every tree cost one launch instead of eleven each.  Same passes,
// same operands, order-free folds: same bytes.  Scratch: by then the level loop's item / accumulator / bucket arrays
// are dead, and each root owns the slice of them its element range addresses -- items and accumulators
// [start, start + count) (a level has at most `count` items), buckets [ceil(start / 5), +floor(count / 5)) (a bucketed
// item has at least 5 triangles).  `counters` and `scan_base` are the block's own (scan_base[i] is addressed with the
// global element index i in [start, start + count]).
constexpr int kParityCounters = 5;
CV_HD void parity_subtree_view(const BuildArrays& ga, const BuildItem& root, BuildArrays& b, int* counters, int* scan_base) {
    b = ga;
    b.sub_roots = nullptr;
    b.items = ga.items + root.start;
    b.items_next = ga.items_next + root.start;
    b.accum = ga.accum + root.start;
    b.bins = ga.bins + (root.start + 4) / 5;
    b.bins_capacity = root.count / 5;
    b.counters = counters;
    b.scan = scan_base;
}
// one thread: the root becomes item 0 of the private level 0
CV_D void parity_subtree_setup(const BuildArrays& b, const BuildItem& root) {
    b.items[0] = root;
    b.counters[0] = 0;
    b.counters[1] = 0;
    b.counters[2] = 0;
}
// after a subtree's last level (one thread): error bits and depth into the global counters
CV_D void parity_subtree_account(const BuildArrays& ga, const BuildItem& root, int levels, int error_bits) {
    if (error_bits) cv_atomic_or_i32(&ga.counters[2], error_bits);
    cv_atomic_max_i32(&ga.counters[4], root.depth + levels);
}
// the serial statement (emulator); the device runs k_parity_subtrees, one block per root, same passes
CV_D void parity_subtree_serial(const BuildArrays& ga, int r) {
    const BuildItem root = ga.sub_roots[r];
    int counters[kParityCounters] = { 0, 0, 0, 0, 0 };
    BuildArrays b;
    parity_subtree_view(ga, root, b, counters, ga.scan);  // subtrees run one after the other here: the global scan array serves each in turn
    parity_subtree_setup(b, root);
    const int start = root.start, end = root.start + root.count;
    for (int i = start; i < end; ++i) b.node_of[i] = 0;
    int n_items = 1, levels = 0;
    while (n_items > 0) {
        for (int w = 0; w < n_items; ++w) body_init_accum(b, w, n_items);
        for (int i = start; i < end; ++i) body_reduce(b, i);
        for (int w = 0; w < n_items; ++w) body_node(b, w, n_items);
        for (int i = start; i < end; ++i) body_bin(b, i);
        for (int w = 0; w < n_items; ++w) body_split(b, w, n_items);
        int run = 0;
        for (int i = start; i < end; ++i) {
            body_flag(b, i);
            b.scan[i] = run;
            run += b.flags[i];
        }
        b.scan[end] = run;
        for (int w = 0; w < n_items; ++w) body_children(b, w, n_items);
        for (int i = start; i < end; ++i) body_scatter(b, i);
        ++levels;
        n_items = counters[0];
        counters[0] = 0;
        counters[1] = 0;
        BuildItem* ti = b.items; b.items = b.items_next; b.items_next = ti;
        int* to = b.order; b.order = b.order_next; b.order_next = to;
        int* tn = b.node_of; b.node_of = b.node_of_next; b.node_of_next = tn;
        if (counters[2] != 0 || root.depth + levels > kMaxDepth + 2) break;
    }
    parity_subtree_account(ga, root, levels, counters[2] | (n_items > 0 && counters[2] == 0 ? BUILD_ERR_DEPTH : 0));
}

}  // namespace cavept
