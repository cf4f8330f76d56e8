// Production BVH builder: one world-space tree over every (instance, triangle) pair.
//
// The reference has no such structure: HitScene (cave/shader/shared_path_tracer.h:152-189)
// loops over every mesh and walks each mesh's own tree.  The closest hit of that loop is
// the closest hit over all (instance, triangle) pairs, so ONE tree over all pairs returns
// the same hit while visiting far fewer nodes.  Triangles are
// stored in world space: for an identity instance that is HitScene's arithmetic bit for bit;
// for a transformed instance the reference intersects in mesh space with the
// inverse-transformed ray (:160-166) -- the same (t, u, v) up to fp32 rounding of the transform.
//
// Construction is top-down binned SAH, level-synchronous (all nodes of a level in the same
// launches), 16 bins on the axis of largest centroid extent:
//   prim pass   world AABB of every pair, scene bounds + centroid bounds (atomics)
//   per level   clear bins | bin (count, AABB, centroid bounds per bin) | split (SAH sweep,
//               child boxes straight from the bins = the bottom-up AABB reduction is folded
//               into the binning) | flag | exclusive scan | stable scatter
//   emit        leaves are contiguous ranges of the final order: pre-gather 3 x float4
//               positions and normals per slot
// Nodes are the 32-byte FastNode records of trace_body.h, siblings adjacent.
//
// Boxes are padded by pad = 2^-20 * (largest |coordinate| in the scene) so that the
// fused-multiply-add slab test of the tracer (walk_step, trace_body.h) is conservative.
#pragma once
#include "cv_common.h"
#include "trace_body.h"

namespace cavept {

constexpr int kWBins = 16;
#ifndef CV_WB_AXES
#define CV_WB_AXES 3
#endif
#ifndef CV_WB_AXES3_MIN
#define CV_WB_AXES3_MIN 0
#endif
constexpr int kWAxes = CV_WB_AXES;  // 3: every axis is binned and the SAH sweep takes the cheapest plane of all three; 1: the axis of largest centroid extent only
constexpr int kWLevelBase = 4;     // counters[4 + L] = interior items of level L (the level loop never reads them back one by one)
constexpr int kWMaxLevels = 124;
constexpr int kWBinBase = kWLevelBase + kWMaxLevels + 4;  // counters[kWBinBase + L] = bin-arena slots handed out for level L
constexpr int kWCounters = kWBinBase + kWMaxLevels + 4;   // ints in the counter block
constexpr int kWMedianDepth = 48;  // safety net: positional median splits below this depth

#ifndef CV_SUB_MAX
#define CV_SUB_MAX 768
#endif
constexpr int kWSubMax = CV_SUB_MAX;   // an item of at most this many prims leaves the level-synchronous phase: its whole subtree is built by one thread block (wb_subtree_*)
constexpr int kWSubLevels = 64;  // levels a subtree may have below its root
constexpr int kWDirectMax = 8;  // items of at most this many prims are split by one thread straight from the prims: no global bins, no atomics

// One SAH bin.  Cleared = all zero: the min corner is kept as the COMPLEMENT of its ordered key (folded with
// atomicMax), so the split pass clears what it has read with plain zero stores and the bin arena never needs a
// clearing pass of its own (it is zeroed once per build).
struct alignas(32) WBin {
    uint32_t count;
    uint32_t nmn[3];  // ~ordered_key(AABB min)
    uint32_t mx[3];   // ordered_key(AABB max)
    uint32_t spare;
};
static_assert(sizeof(WBin) == 32, "WBin");
struct WBins {
    WBin b[kWAxes * kWBins];  // [axis slot][bin]; with kWAxes == 1 the single slot holds the axis of largest extent
};

struct WItem {
    int start, count;  // segment of `order`
    int pair;          // children go to fnodes[2*pair], fnodes[2*pair + 1]
    int depth;
    int axis;          // >= 0: SAH split (after the split pass: the axis of the chosen plane); -1: positional median split
    float cmin[3], cscale[3];  // per axis: bin = (box centre - cmin) * cscale (cscale 0: no extent, everything in bin 0)
    int split;         // left child = bins [0, split]
    int n_left;
    int child_item[2];  // index of each child in the next level, -1 if it is a leaf
    int bins_slot;      // index of this item's WBins in the arena (items of more than kWDirectMax prims), -1: none
    int scan_tile, scan_local;  // partition pass: tile holding `start` and the exclusive count of left flags before it in that tile
    float box_mn[3], box_mx[3];
    int pair_lo;        // a queued subtree root of the hybrid builder: its subtree's pairs come from [pair_lo, ...), the dead radix-tree slots of its own range; -1: from the global counter
};

enum { WB_ERR_CAPACITY = 1 };

struct WBuild {
    // inputs: the reference-layout scene arrays
    const GpuPtVertex* vertices;
    const GpuPtIndex* indices;
    const GpuPtMesh* instances;
    const int* inst_prim_off;  // n_inst + 1
    const int* inst_tri_off;   // per instance: scene-global index of its mesh's first triangle
    int n_inst;
    int n_prims;
    float pad;
    int force_median;
    // per prim
    F4* bmin;  // xyz world AABB min, w = bits of the scene-global triangle id
    F4* bmax;  // xyz world AABB max, w = bits of the instance id
    int* order;
    int* order_next;
    int* node_of;
    int* node_of_next;
    unsigned char* flags;
    int* scan;
    WItem* items;
    WItem* items_next;
    WBins* bins;
    int items_capacity;
    int pairs_capacity;
    int bins_capacity;   // WBins records in the arena: at most n_prims / (kWDirectMax + 1) items of a level own one
    int* counters;       // global: [0] deepest level reached by a subtree build + 1, [1] pairs allocated, [2] error bits, [3] subtree roots queued
    int* level_items;    // [L] = interior items of level L -- the level-synchronous phase: counters + kWLevelBase; a subtree build: block-local, L counted from its root
    int* level_bins;     // [L] = bin-arena slots handed out for level L (same two homes)
    WItem* sub_roots;    // level-synchronous phase: queue of items of <= kWSubMax prims (null inside a subtree build)
    int sub_capacity;
    // LBVH builder only
    uint32_t* keys;      // Morton code per prim, then sorted
    uint32_t* keys_tmp;
    int* sort_hist;      // 2 * sort_hist_entries(n_prims)
    F4* nbox_mn;         // box of the interior node with heap index h (written bottom-up)
    F4* nbox_mx;
    int* node_first;     // radix tree: slot range of internal node i
    int* node_last;
    int* node_split;     // last slot of its left child
    int* node_parent;
    int* node_count;     // children that are leaves, then arrivals of the bottom-up pass
    float morton_min[3], morton_scale[3];
    uint32_t* root_acc;  // [0..2] scene min keys, [3..5] max keys, [6..8] centroid min keys, [9..11] centroid max keys
    // outputs
    FastNode* fnodes;
    QPair* qpairs;  // quantised copy of fnodes, what the tracer fetches
    Q4Node* q4nodes;  // four-wide records over the same tree (null: not built)
    unsigned char* pair_depth;  // SAH builder: depth of the interior node that owns sibling pair k (root pair: 0); the refit walks the pairs deepest first
    float grid_origin[3], grid_scale[3];
    FastTri* ftris;
    FastTri* fnormals;
};

CV_HD int wb_find_instance(const int* off, int n, int p) {
    int lo = 0, hi = n;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (off[mid] <= p) lo = mid; else hi = mid;
    }
    return lo;
}

// world-space position of a mesh-space vertex (glm column-major world matrix)
CV_HD F3 wb_world(const float* m, const float* v) {
    return f3(m[0] * v[0] + m[4] * v[1] + m[8] * v[2] + m[12], m[1] * v[0] + m[5] * v[1] + m[9] * v[2] + m[13], m[2] * v[0] + m[6] * v[1] + m[10] * v[2] + m[14]);
}

// ---- prim pass ---------------------------------------------------------------------------
// world AABB of (scene-global triangle, instance), unpadded
CV_D void wb_tri_box(const WBuild& b, int tri, int inst, F3& mn, F3& mx) {
    const float* m = b.instances[inst].transform;
    const int* t = b.indices[tri].tri;
    mn = f3(CV_INF, CV_INF, CV_INF);
    mx = f3(-CV_INF, -CV_INF, -CV_INF);
    for (int k = 0; k < 3; ++k) {
        const F3 w = wb_world(m, b.vertices[t[k]].position);
        mn = f3(fminf(mn.x, w.x), fminf(mn.y, w.y), fminf(mn.z, w.z));
        mx = f3(fmaxf(mx.x, w.x), fmaxf(mx.y, w.y), fmaxf(mx.z, w.z));
    }
}
CV_D void wb_prim_box(const WBuild& b, int p, F3& mn, F3& mx) {
    const int inst = wb_find_instance(b.inst_prim_off, b.n_inst, p);
    const int tri = b.inst_tri_off[inst] + (p - b.inst_prim_off[inst]);
    wb_tri_box(b, tri, inst, mn, mx);
    F4 o;
    o.x = mn.x; o.y = mn.y; o.z = mn.z; o.w = u2f((uint32_t)tri);
    b.bmin[p] = o;
    o.x = mx.x; o.y = mx.y; o.z = mx.z; o.w = u2f((uint32_t)inst);
    b.bmax[p] = o;
}
// per-index form (the emulator's; the device runs k_wb_prim, which reduces per block first)
CV_D void wb_body_prim(const WBuild& b, int p) {
    if (p >= b.n_prims) return;
    F3 mn, mx;
    wb_prim_box(b, p, mn, mx);
    const F3 c = 0.5f * (mn + mx);
    const CvGroup g = cv_group_by(0);
    cv_group_min_u32(g, &b.root_acc[0], ordered_key(mn.x));
    cv_group_min_u32(g, &b.root_acc[1], ordered_key(mn.y));
    cv_group_min_u32(g, &b.root_acc[2], ordered_key(mn.z));
    cv_group_max_u32(g, &b.root_acc[3], ordered_key(mx.x));
    cv_group_max_u32(g, &b.root_acc[4], ordered_key(mx.y));
    cv_group_max_u32(g, &b.root_acc[5], ordered_key(mx.z));
    cv_group_min_u32(g, &b.root_acc[6], ordered_key(c.x));
    cv_group_min_u32(g, &b.root_acc[7], ordered_key(c.y));
    cv_group_min_u32(g, &b.root_acc[8], ordered_key(c.z));
    cv_group_max_u32(g, &b.root_acc[9], ordered_key(c.x));
    cv_group_max_u32(g, &b.root_acc[10], ordered_key(c.y));
    cv_group_max_u32(g, &b.root_acc[11], ordered_key(c.z));
}

// pad the boxes (b.pad is known once the scene bounds are), identity order, everything in item 0
CV_D void wb_body_init(const WBuild& b, int p) {
    if (p >= b.n_prims) return;
    F4 mn = b.bmin[p], mx = b.bmax[p];
    mn.x -= b.pad; mn.y -= b.pad; mn.z -= b.pad;
    mx.x += b.pad; mx.y += b.pad; mx.z += b.pad;
    b.bmin[p] = mn;
    b.bmax[p] = mx;
    b.order[p] = p;
    b.node_of[p] = 0;
}

// ---- per level ----------------------------------------------------------------------------
// Binning set-up of an item from its box: the bins of every axis span the box (measured on the config scenes: the
// node count per segment is within 0.5 % of binning over the exact centroid bounds, which cost 18 more atomics per prim).
CV_HD int wb_binning_setup(const float box_mn[3], const float box_mx[3], int depth, int force_median, float cscale[3]) {
    const F3 ext = f3(box_mx[0] - box_mn[0], box_mx[1] - box_mn[1], box_mx[2] - box_mn[2]);
    int axis = 0;
    if (ext.y > comp(ext, axis)) axis = 1;
    if (ext.z > comp(ext, axis)) axis = 2;
    const float e = comp(ext, axis);
    for (int c = 0; c < 3; ++c) cscale[c] = comp(ext, c) > 0.0f ? cv_div((float)kWBins, comp(ext, c)) : 0.0f;
    // no extent (or the safety net): split the list in the middle
    return (!(e > 0.0f) || force_median || depth >= kWMedianDepth) ? -1 : axis;
}
CV_HD void wb_item_setup(WItem& it, int force_median) {
    for (int c = 0; c < 3; ++c) it.cmin[c] = it.box_mn[c];
    it.axis = wb_binning_setup(it.box_mn, it.box_mx, it.depth, force_median, it.cscale);
}

CV_HD int wb_bin_at(const float* cmin, const float* cscale, int axis, F4 mn, F4 mx) {
    const float c = axis == 0 ? 0.5f * (mn.x + mx.x) : (axis == 1 ? 0.5f * (mn.y + mx.y) : 0.5f * (mn.z + mx.z));
    const int k = (int)((c - cmin[axis]) * cscale[axis]);
    return k < 0 ? 0 : (k > kWBins - 1 ? kWBins - 1 : k);
}
CV_HD int wb_bin_of(const WItem& it, int axis, F4 mn, F4 mx) { return wb_bin_at(it.cmin, it.cscale, axis, mn, mx); }
// axis binned in slot a of an item's bins
CV_HD int wb_slot_axis(const WItem& it, int a) { return kWAxes == 1 ? it.axis : a; }

// one thread per prim: count and AABB into the bins of its item (items of more than kWDirectMax prims only)
CV_D void wb_body_bin(const WBuild& b, int i) {
    if (i >= b.n_prims) return;
    const int w = b.node_of[i];
    if (w < 0) return;
    const WItem& it = b.items[w];
    if (it.axis < 0 || it.bins_slot < 0) return;
    const int p = b.order[i];
    const F4 mn = b.bmin[p], mx = b.bmax[p];
    for (int a = 0; a < kWAxes; ++a) {
        const int k = a * kWBins + wb_bin_of(it, wb_slot_axis(it, a), mn, mx);
        WBin& bin = b.bins[it.bins_slot].b[k];
        const CvGroup g = cv_group_by(it.bins_slot * (kWAxes * kWBins) + k);
        cv_group_add_i32(g, reinterpret_cast<int*>(&bin.count), 1);
        cv_group_max_u32(g, &bin.nmn[0], ~ordered_key(mn.x));
        cv_group_max_u32(g, &bin.nmn[1], ~ordered_key(mn.y));
        cv_group_max_u32(g, &bin.nmn[2], ~ordered_key(mn.z));
        cv_group_max_u32(g, &bin.mx[0], ordered_key(mx.x));
        cv_group_max_u32(g, &bin.mx[1], ordered_key(mx.y));
        cv_group_max_u32(g, &bin.mx[2], ordered_key(mx.z));
    }
}

struct WBox {
    F3 mn, mx;
};
CV_HD WBox wb_empty() {
    WBox r;
    r.mn = f3(CV_INF, CV_INF, CV_INF);
    r.mx = f3(-CV_INF, -CV_INF, -CV_INF);
    return r;
}
CV_HD void wb_grow(WBox& a, F3 mn, F3 mx) {
    a.mn = f3(fminf(a.mn.x, mn.x), fminf(a.mn.y, mn.y), fminf(a.mn.z, mn.z));
    a.mx = f3(fmaxf(a.mx.x, mx.x), fmaxf(a.mx.y, mx.y), fmaxf(a.mx.z, mx.z));
}
CV_HD float wb_half_area(const WBox& a) {
    const F3 e = a.mx - a.mn;
    return e.x * e.y + e.y * e.z + e.z * e.x;
}

// The 16 bins of one axis as the sweep sees them (an empty bin is the empty box)
struct WLBin {
    int count;
    WBox box;
};
CV_HD WLBin wb_lbin_of(const WBin& s) {
    WLBin r;
    r.count = (int)s.count;
    r.box = wb_empty();
    if (r.count > 0) {
        r.box.mn = f3(ordered_value(~s.nmn[0]), ordered_value(~s.nmn[1]), ordered_value(~s.nmn[2]));
        r.box.mx = f3(ordered_value(s.mx[0]), ordered_value(s.mx[1]), ordered_value(s.mx[2]));
    }
    return r;
}
// SAH sweep over the 15 planes of one axis slot: cost(k) = n_left * area_left + n_right * area_right; a plane with an
// empty side is no candidate.  Updates (best_cost, best_key = slot * 16 + plane) on a strictly smaller cost, so over
// slots 0..2 and planes 0..14 in order the winner is the lexicographic minimum of (cost, key) -- which is how the
// device's warp-wide version (k_wb_split_binned, kernels.cu) states it.
CV_HD void wb_sweep(const WLBin* lb, int slot, float& best_cost, int& best_key) {
    float right_area[kWBins];
    int right_cnt[kWBins];
    WBox acc = wb_empty();
    int n = 0;
    for (int k = kWBins - 1; k >= 1; --k) {
        wb_grow(acc, lb[k].box.mn, lb[k].box.mx);
        n += lb[k].count;
        right_area[k] = n > 0 ? wb_half_area(acc) : 0.0f;
        right_cnt[k] = n;
    }
    acc = wb_empty();
    n = 0;
    for (int k = 0; k < kWBins - 1; ++k) {
        wb_grow(acc, lb[k].box.mn, lb[k].box.mx);
        n += lb[k].count;
        if (n == 0 || right_cnt[k + 1] == 0) continue;
        const float cost = (float)n * wb_half_area(acc) + (float)right_cnt[k + 1] * right_area[k + 1];
        if (cost < best_cost) {
            best_cost = cost;
            best_key = slot * kWBins + k;
        }
    }
}

// Second half of a split, shared by every path: the item's two children (boxes and counts known) become leaf
// references or interior items of the next level; the sibling pair is written.  box[] / cnt[] are ignored for a
// median split (axis < 0).
CV_D void wb_finish_split(const WBuild& b, WItem& it, int level, WBox box[2], int cnt[2]) {
    if (it.axis < 0) {
        cnt[0] = it.count / 2;
        cnt[1] = it.count - cnt[0];
        for (int s = 0; s < 2; ++s) {
            box[s].mn = f3(it.box_mn[0], it.box_mn[1], it.box_mn[2]);
            box[s].mx = f3(it.box_mx[0], it.box_mx[1], it.box_mx[2]);
        }
    }
    it.n_left = cnt[0];
    for (int s = 0; s < 2; ++s) {
        const int start = s == 0 ? it.start : it.start + cnt[0];
        int ref;
        it.child_item[s] = -1;
        if (cnt[s] <= kFastLeafMax) {
            ref = make_leaf_ref(start, cnt[s]);
        } else {
            const int pair = cv_atomic_add_i32(&b.counters[1], 1);
            // a child of at most kWSubMax prims is queued for a subtree build instead of entering the next level
            // (its prims stay where they are: node_of = -1 from here on); a full queue just means it stays in the level loop
            const int sub = (b.sub_roots && cnt[s] <= kWSubMax) ? cv_atomic_add_i32(&b.counters[3], 1) : b.sub_capacity;
            const bool queued = sub < b.sub_capacity;
            const int slot = queued ? 0 : cv_atomic_add_i32(&b.level_items[level + 1], 1);
            const int bins_slot = (!queued && cnt[s] > kWDirectMax) ? cv_atomic_add_i32(&b.level_bins[level + 1], 1) : -1;
            if (pair >= b.pairs_capacity || slot >= b.items_capacity || bins_slot >= b.bins_capacity) {
                cv_atomic_or_i32(&b.counters[2], WB_ERR_CAPACITY);
                ref = make_leaf_ref(start, 1);
            } else {
                ref = 2 * pair;
                b.pair_depth[pair] = (unsigned char)(it.depth + 1);
                if (!queued) it.child_item[s] = slot;
                WItem c;
                c.start = start;
                c.count = cnt[s];
                c.pair = pair;
                c.depth = it.depth + 1;
                c.split = c.n_left = 0;
                c.child_item[0] = c.child_item[1] = -1;
                c.bins_slot = bins_slot;
                c.scan_tile = c.scan_local = 0;
                c.pair_lo = -1;
                c.box_mn[0] = box[s].mn.x; c.box_mn[1] = box[s].mn.y; c.box_mn[2] = box[s].mn.z;
                c.box_mx[0] = box[s].mx.x; c.box_mx[1] = box[s].mx.y; c.box_mx[2] = box[s].mx.z;
                wb_item_setup(c, b.force_median);
                if (queued) b.sub_roots[sub] = c; else b.items_next[slot] = c;
            }
        }
        FastNode o;
        o.mn[0] = box[s].mn.x; o.mn[1] = box[s].mn.y; o.mn[2] = box[s].mn.z;
        o.mx[0] = box[s].mx.x; o.mx[1] = box[s].mx.y; o.mx[2] = box[s].mx.z;
        o.a = ref;
        o.b = cnt[s];
        b.fnodes[2 * it.pair + s] = o;
    }
}

// One thread per item: SAH sweep over every binned axis (the bins come from the arena, which is cleared as it is
// read, or -- items of at most kWDirectMax prims -- straight from the item's prims), child boxes from the winning
// axis' bins (the bottom-up AABB reduction is folded into the binning), then wb_finish_split.
// with_arena_items = false: items that own arena bins are left to the device's warp-wide kernel.
CV_D void wb_body_split(const WBuild& b, int w, int level, bool with_arena_items) {
    const int n_items = b.level_items[level];
    if (w >= n_items) return;
    WItem& it = b.items[w];
    if (!with_arena_items && it.axis >= 0 && it.bins_slot >= 0) return;
    WBox box[2];
    int cnt[2] = { 0, 0 };
    box[0] = box[1] = wb_empty();
    if (it.axis >= 0) {
        WLBin lb[kWBins], best_lb[kWBins];
        float best_cost = CV_INF;
        int best_key = -1;
        for (int a = 0; a < kWAxes; ++a) {
            if (it.bins_slot >= 0) {
                WBin* src = &b.bins[it.bins_slot].b[a * kWBins];
                for (int k = 0; k < kWBins; ++k) {
                    lb[k] = wb_lbin_of(src[k]);
                    WBin z = {};
                    src[k] = z;
                }
            } else {
                for (int k = 0; k < kWBins; ++k) {
                    lb[k].count = 0;
                    lb[k].box = wb_empty();
                }
                const int axis = wb_slot_axis(it, a);
                for (int j = 0; j < it.count; ++j) {
                    const int p = b.order[it.start + j];
                    const F4 mn = b.bmin[p], mx = b.bmax[p];
                    WLBin& d = lb[wb_bin_of(it, axis, mn, mx)];
                    ++d.count;
                    wb_grow(d.box, f3(mn.x, mn.y, mn.z), f3(mx.x, mx.y, mx.z));
                }
            }
            const int before = best_key;
            wb_sweep(lb, a, best_cost, best_key);
            if (best_key != before) {
                for (int k = 0; k < kWBins; ++k) best_lb[k] = lb[k];
            }
        }
        if (best_key < 0) {
            it.axis = -1;  // every prim in one bin on every axis (coincident centres): median split
        } else {
            it.axis = wb_slot_axis(it, best_key / kWBins);
            it.split = best_key % kWBins;
            for (int k = 0; k < kWBins; ++k) {
                if (best_lb[k].count == 0) continue;
                const int s = k <= it.split ? 0 : 1;
                wb_grow(box[s], best_lb[k].box.mn, best_lb[k].box.mx);
                cnt[s] += best_lb[k].count;
            }
        }
    }
    wb_finish_split(b, it, level, box, cnt);
}

// does prim position i go to the left child of its item?
CV_D unsigned char wb_flag_of(const WBuild& b, int i) {
    const int w = b.node_of[i];
    if (w < 0) return 0;
    const WItem& it = b.items[w];
    if (it.axis >= 0) {
        const int p = b.order[i];
        return wb_bin_of(it, it.axis, b.bmin[p], b.bmax[p]) <= it.split ? 1 : 0;
    }
    return (i - it.start) < it.n_left ? 1 : 0;
}
CV_D void wb_body_flag(const WBuild& b, int i) {
    if (i >= b.n_prims) return;
    b.flags[i] = wb_flag_of(b, i);
}

// stable partition: scan_i = left flags before position i (whole array), scan_start = the same at the item's start
CV_D void wb_scatter_to(const WBuild& b, int i, int w, unsigned char flag, int scan_i, int scan_start) {
    const WItem& it = b.items[w];
    const int lefts_before = scan_i - scan_start;
    int dst, side;
    if (flag) {
        dst = it.start + lefts_before;
        side = 0;
    } else {
        dst = it.start + it.n_left + ((i - it.start) - lefts_before);
        side = 1;
    }
    b.order_next[dst] = b.order[i];
    b.node_of_next[dst] = it.child_item[side];
}
CV_D void wb_body_scatter(const WBuild& b, int i) {
    if (i >= b.n_prims) return;
    const int w = b.node_of[i];
    if (w < 0) {
        b.order_next[i] = b.order[i];
        b.node_of_next[i] = -1;
        return;
    }
    wb_scatter_to(b, i, w, b.flags[i], b.scan[i], b.scan[b.items[w].start]);
}

// ---- subtree builds: the second phase of the SAH builder ------------------------------------------------------
// An item of <= kWSubMax prims is a closed problem: its prims are one contiguous range of `order` that nothing else
// touches.  One thread block builds its whole subtree with the same per-prim / per-item passes as the level loop
// (bin -> split -> flag -> scan -> scatter), block barriers in place of kernel boundaries, so the ~15 deepest levels of
// the tree cost one launch instead of six each.  Scratch: the level loop's own item and bin arrays are dead by then and
// are carved up by prim range (a subtree over `count` prims has at most count / 3 items and count / 9 arena bins per level).  Pairs come from the same global
// counter and carry their owner's depth (pair_depth), which is all the refit needs.
// wb_subtree_view: a block's private view of the build (no writes); wb_subtree_setup: the view + the root item in
// items[0] and the level-0 counters (one thread).
CV_HD void wb_subtree_view(const WBuild& gb, const WItem& root, WBuild& b, int* lv_items, int* lv_bins) {
    b = gb;
    b.sub_roots = nullptr;
    b.level_items = lv_items;
    b.level_bins = lv_bins;
    // [ceil(start / m), ceil(start / m) + floor(count / m)) with m = the fewest prims of an interior item (3) resp. of an
    // arena item (9): disjoint for disjoint prim ranges, and a level of the subtree never has more such items
    constexpr int mi = kFastLeafMax + 1, mb = kWDirectMax + 1;
    const int ibase = (root.start + mi - 1) / mi, bbase = (root.start + mb - 1) / mb;
    b.items = gb.items + ibase;
    b.items_next = gb.items_next + ibase;
    b.items_capacity = root.count / mi;
    b.bins = gb.bins + bbase;
    b.bins_capacity = root.count / mb;
}
CV_D WItem wb_subtree_setup(const WBuild& gb, int r, WBuild& b, int* lv_items, int* lv_bins) {
    WItem root = gb.sub_roots[r];
    wb_subtree_view(gb, root, b, lv_items, lv_bins);
    root.bins_slot = root.count > kWDirectMax ? 0 : -1;
    b.items[0] = root;
    lv_items[0] = 1;
    lv_bins[0] = root.bins_slot >= 0 ? 1 : 0;
    return root;
}
// after a subtree's last level: its statistics into the global counters, its prims back into the buffer the emit pass reads
CV_D void wb_subtree_account(const WBuild& gb, const WItem& root, int levels, const int* lv_items) {
    for (int l = 0; l < levels && root.depth + l < kWMaxLevels; ++l) cv_atomic_add_i32(&gb.counters[kWLevelBase + root.depth + l], lv_items[l]);
#if defined(__CUDA_ARCH__)
    atomicMax(&gb.counters[0], root.depth + levels);
#else
    if (gb.counters[0] < root.depth + levels) gb.counters[0] = root.depth + levels;
#endif
}
// the serial statement (emulator); the device runs k_wb_subtrees (kernels.cu), one block per root, same passes
CV_D void wb_subtree_serial(const WBuild& gb, int r) {
    WBuild b;
    int lv_items[kWSubLevels + 2] = { 0 }, lv_bins[kWSubLevels + 2] = { 0 };
    const WItem root = wb_subtree_setup(gb, r, b, lv_items, lv_bins);
    int own[4] = { 0, root.pair_lo, 0, 0 };  // hybrid builder: this subtree's pairs come from its own slot range
    if (root.pair_lo >= 0) b.counters = own;
    const int start = root.start, end = root.start + root.count;
    for (int i = start; i < end; ++i) b.node_of[i] = 0;
    int level = 0;
    for (; lv_items[level] > 0; ++level) {
        if (level >= kWSubLevels) {
            cv_atomic_or_i32(&gb.counters[2], WB_ERR_CAPACITY);
            break;
        }
        for (int i = start; i < end; ++i) wb_body_bin(b, i);
        for (int w = 0; w < lv_items[level]; ++w) wb_body_split(b, w, level, true);
        int run = 0;
        for (int i = start; i < end; ++i) {
            b.flags[i] = wb_flag_of(b, i);
            b.scan[i] = run;
            run += b.flags[i];
        }
        for (int i = start; i < end; ++i) wb_body_scatter(b, i);
        WItem* ti = b.items; b.items = b.items_next; b.items_next = ti;
        int* to = b.order; b.order = b.order_next; b.order_next = to;
        int* tn = b.node_of; b.node_of = b.node_of_next; b.node_of_next = tn;
    }
    if (b.order != gb.order) {
        for (int i = start; i < end; ++i) gb.order[i] = b.order[i];
    }
    if (root.pair_lo >= 0 && own[2] != 0) cv_atomic_or_i32(&gb.counters[2], own[2]);
    wb_subtree_account(gb, root, level, lv_items);
}

// ---- emit: slot s of the final order holds prim order[s] --------------------------------------
// World-space vertices (the same wb_world() values the boxes were built from) and the vertex
// normals times the world matrix.  A mirroring instance (det < 0) flips the winding the reference's mesh-space
// back-face test sees (shared_path_tracer.h:106), so its triangles are stored as (A, C, B).
CV_D void wb_emit_slot(const WBuild& b, int s, uint32_t tri, uint32_t inst);
CV_D void wb_body_emit(const WBuild& b, int s) {
    if (s >= b.n_prims) return;
    const int p = b.order[s];
    wb_emit_slot(b, s, f2u(b.bmin[p].w), f2u(b.bmax[p].w));
}
CV_D void wb_emit_slot(const WBuild& b, int s, uint32_t tri, uint32_t inst) {
    const float* m = b.instances[inst].transform;
    const float det = m[0] * (m[5] * m[10] - m[9] * m[6]) - m[4] * (m[1] * m[10] - m[9] * m[2]) + m[8] * (m[1] * m[6] - m[5] * m[2]);
    const int* idx = b.indices[tri].tri;
    const int i1 = det < 0.0f ? idx[2] : idx[1], i2 = det < 0.0f ? idx[1] : idx[2];
    const GpuPtVertex va = b.vertices[idx[0]], vb = b.vertices[i1], vc = b.vertices[i2];
    const F3 wa = wb_world(m, va.position), wb = wb_world(m, vb.position), wc = wb_world(m, vc.position);
    const F3 ab = wb - wa, ac = wc - wa;  // HitTriangle's first two lines (shared_path_tracer.h:94-95)
    FastTri t;
    t.v0.x = wa.x; t.v0.y = wa.y; t.v0.z = wa.z; t.v0.w = u2f(inst);
    t.v1.x = ab.x; t.v1.y = ab.y; t.v1.z = ab.z; t.v1.w = u2f(tri);
    t.v2.x = ac.x; t.v2.y = ac.y; t.v2.z = ac.z; t.v2.w = 0.0f;
    t.pad.x = t.pad.y = t.pad.z = t.pad.w = 0.0f;
    b.ftris[s] = t;
    const F3 na = mat_mul(m, f3(va.normal[0], va.normal[1], va.normal[2]), 0.0f);  // the shader's own transform (:210)
    const F3 nb = mat_mul(m, f3(vb.normal[0], vb.normal[1], vb.normal[2]), 0.0f);
    const F3 nc = mat_mul(m, f3(vc.normal[0], vc.normal[1], vc.normal[2]), 0.0f);
    t.v0.x = na.x; t.v0.y = na.y; t.v0.z = na.z;
    t.v1.x = nb.x; t.v1.y = nb.y; t.v1.z = nb.z; t.v1.w = u2f((uint32_t)b.instances[inst].materialId);
    t.v2.x = nc.x; t.v2.y = nc.y; t.v2.z = nc.z;
    b.fnormals[s] = t;
}

// ---- quantise: FastNode pair k -> QPair k --------------------------------------------------
// Grid coordinates are rounded OUTWARD (floor for min planes, ceil for max planes) with a 1/64
// cell margin that covers the fp32 rounding of this conversion and of the tracer's grid-space
// ray set-up (each below 2^-8 cell), so a quantised box always contains the fp32 box.
CV_HD uint32_t wb_grid_lo(float v, float origin, float scale) {
    float g = floorf((v - origin) * scale - 0.015625f);
    g = g < 0.0f ? 0.0f : (g > kGridCells ? kGridCells : g);
    return (uint32_t)(int)g;
}
CV_HD uint32_t wb_grid_hi(float v, float origin, float scale) {
    float g = ceilf((v - origin) * scale + 0.015625f);
    g = g < 0.0f ? 0.0f : (g > kGridCells ? kGridCells : g);
    return (uint32_t)(int)g;
}
CV_HD int wb_quantised_ref(int r) { return ref_is_interior(r) ? (r >> 1) : r; }

CV_D void wb_body_quantise(const WBuild& b, int k, int n_pairs) {
    if (k >= n_pairs) return;
    QPair o;
    for (int s = 0; s < 2; ++s) {
        const FastNode n = b.fnodes[2 * k + s];
        uint32_t lo[3], hi[3];
        for (int c = 0; c < 3; ++c) {
            lo[c] = wb_grid_lo(n.mn[c], b.grid_origin[c], b.grid_scale[c]);
            hi[c] = wb_grid_hi(n.mx[c], b.grid_origin[c], b.grid_scale[c]);
        }
        for (int c = 0; c < 3; ++c) o.q[3 * s + c] = lo[c] | (hi[c] << 16);
        if (s == 0) o.r0 = wb_quantised_ref(n.a); else o.r1 = wb_quantised_ref(n.a);
    }
    b.qpairs[k] = o;
}

// ---- four-wide records: Q4Node k = pair k with interior children replaced by their own pairs, largest box first,
//      until there are four children or only leaves.  One independent thread per pair: every pair gets a record, so
//      an interior child's reference is simply its pair index, and QPair k / Q4Node k root the same subtree.  Records
//      of unused pair slots (the LBVH builder leaves holes) are never referenced; their garbage is bounds-checked.
CV_D void wb_body_q4(const WBuild& b, int k, int n_pairs) {
    if (k >= n_pairs) return;
    FastNode ch[4];
    ch[0] = b.fnodes[2 * k];
    ch[1] = b.fnodes[2 * k + 1];
    int n = 2;
    while (n < 4) {
        int best = -1;
        float best_area = -1.0f;
        for (int i = 0; i < n; ++i) {
            const int r = ch[i].a;
            if (!ref_is_interior(r) || (uint32_t)r + 1u >= 2u * (uint32_t)n_pairs) continue;
            const float ex = ch[i].mx[0] - ch[i].mn[0], ey = ch[i].mx[1] - ch[i].mn[1], ez = ch[i].mx[2] - ch[i].mn[2];
            const float area = ex * ey + ey * ez + ez * ex;
            if (area > best_area) {
                best_area = area;
                best = i;
            }
        }
        if (best < 0) break;
        const int r = ch[best].a;
        FastNode c0 = b.fnodes[r], c1 = b.fnodes[r + 1];
        // replace the expanded child in place, its sibling goes to the end (dynamic indexing kept out of registers' way)
        for (int i = 0; i < 4; ++i) {
            if (i == best) ch[i] = c0;
            if (i == n) ch[i] = c1;
        }
        ++n;
    }
    Q4Node o;
    for (int c = 0; c < 4; ++c) {
        if (c < n) {
            for (int a = 0; a < 3; ++a)
                o.q[3 * c + a] = wb_grid_lo(ch[c].mn[a], b.grid_origin[a], b.grid_scale[a]) | (wb_grid_hi(ch[c].mx[a], b.grid_origin[a], b.grid_scale[a]) << 16);
            o.r[c] = wb_quantised_ref(ch[c].a);
        } else {
            for (int a = 0; a < 3; ++a) o.q[3 * c + a] = 32767u;  // min plane 32767, max plane 0: never hit
            o.r[c] = kRefSentinel;
        }
    }
    b.q4nodes[k] = o;
}
struct FnWbQ4 { WBuild b; int n; CV_D void operator()(int i) const { wb_body_q4(b, i, n); } };

// ---- refit: instance transforms changed, topology kept ---------------------------
// The reference re-uploads M x 144 bytes of GpuPtMesh on SCENE_DIRTY_WORLD and never touches a BVH
// (cave/engine/renderer/path_tracer/path_tracer.cpp:162-199): its per-mesh trees live in mesh space.  The production
// tree is in world space, so a moved instance's triangles are re-transformed in their slots (same arithmetic as
// wb_body_emit) and every box is recomputed bottom-up, one launch per level: the SAH builder allocates the sibling
// pairs of level L+1 in one contiguous range after those of level L, so "all pairs of a level" is an index range.
// A leaf's box is the union of its triangles' padded world boxes, an interior record's the union of its child pair
// -- exactly what the builder's bins produced, so refitting an unmoved scene reproduces the tree bit for bit.
// The host keeps the result only if the moved scene still fits the quantisation grid and the tree's surface-area
// sum (the SAH cost proxy) stayed within a factor of the freshly built tree's; otherwise it rebuilds.
CV_D void wb_body_retransform(const WBuild& b, int s, const unsigned char* moved) {
    if (s >= b.n_prims) return;
    const uint32_t inst = f2u(b.ftris[s].v0.w);
    if (!moved[inst]) return;
    wb_emit_slot(b, s, f2u(b.ftris[s].v1.w), inst);
}

CV_D void wb_body_refit_pair(const WBuild& b, int k, int n_pairs, int depth, float* area_sum) {
    QPair q;
    float area = 0.0f;
    const bool mine = k < n_pairs && b.pair_depth[k] == (unsigned char)depth;  // every thread of the launch reaches the warp reduction below
    for (int s = 0; mine && s < 2; ++s) {
        FastNode n = b.fnodes[2 * k + s];
        WBox box = wb_empty();
        if (ref_is_interior(n.a)) {
            const FastNode c0 = b.fnodes[n.a], c1 = b.fnodes[n.a + 1];  // the child pair's two records (refitted by the previous, deeper launch)
            wb_grow(box, f3(c0.mn[0], c0.mn[1], c0.mn[2]), f3(c0.mx[0], c0.mx[1], c0.mx[2]));
            wb_grow(box, f3(c1.mn[0], c1.mn[1], c1.mn[2]), f3(c1.mx[0], c1.mx[1], c1.mx[2]));
        } else {
            const int first = n.a & 0x1fffffff, count = ((n.a >> 29) & 3) + 1;
            for (int t = 0; t < count; ++t) {
                const FastTri& tri = b.ftris[first + t];
                F3 mn, mx;
                wb_tri_box(b, (int)f2u(tri.v1.w), (int)f2u(tri.v0.w), mn, mx);
                wb_grow(box, f3(mn.x - b.pad, mn.y - b.pad, mn.z - b.pad), f3(mx.x + b.pad, mx.y + b.pad, mx.z + b.pad));
            }
        }
        n.mn[0] = box.mn.x; n.mn[1] = box.mn.y; n.mn[2] = box.mn.z;
        n.mx[0] = box.mx.x; n.mx[1] = box.mx.y; n.mx[2] = box.mx.z;
        b.fnodes[2 * k + s] = n;
        area += wb_half_area(box);
        for (int c = 0; c < 3; ++c) q.q[3 * s + c] = wb_grid_lo(n.mn[c], b.grid_origin[c], b.grid_scale[c]) | (wb_grid_hi(n.mx[c], b.grid_origin[c], b.grid_scale[c]) << 16);
        if (s == 0) q.r0 = wb_quantised_ref(n.a); else q.r1 = wb_quantised_ref(n.a);
    }
    if (mine) b.qpairs[k] = q;
    if (area_sum) {
#if defined(__CUDA_ARCH__)
        // one atomic per warp: the active lanes of a k_for_each launch are a contiguous run starting at lane 0 (lanes whose pair
        // belongs to another depth carry 0)
        const unsigned mask = __activemask();
        const int lane = (int)(threadIdx.x & 31u);
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            const float o = __shfl_down_sync(mask, area, d);
            if (lane + d < 32 && ((mask >> (lane + d)) & 1u)) area += o;
        }
        if (lane == 0 && area != 0.0f) atomicAdd(area_sum, area);
#else
        *area_sum += area;
#endif
    }
}

struct FnWbRetransform { WBuild b; const unsigned char* moved; CV_D void operator()(int i) const { wb_body_retransform(b, i, moved); } };
struct FnWbRefitPair { WBuild b; int n_pairs; int depth; float* area_sum; CV_D void operator()(int i) const { wb_body_refit_pair(b, i, n_pairs, depth, area_sum); } };

// One launch per depth, deepest first, over all pairs: a pair is refitted in the launch of its owner's depth
// (b.pair_depth), after the pairs of its children.  out[0..5] = refitted scene box (union of the root pair),
// out[6] = surface-area sum of all records.
template<class Exec>
void refit_wbvh(Exec& ex, WBuild b, int n_pairs, int n_levels, const unsigned char* d_moved, float* d_area_sum, float out[7]) {
    ex.fill(d_area_sum, 0, sizeof(float));
    ex.launch(b.n_prims, FnWbRetransform{ b, d_moved });
    for (int L = n_levels - 1; L >= 0; --L) ex.launch(n_pairs, FnWbRefitPair{ b, n_pairs, L, d_area_sum });
    if (b.q4nodes) ex.launch(n_pairs, FnWbQ4{ b, n_pairs });
    FastNode root[2];
    ex.download(root, b.fnodes, sizeof(root));
    for (int c = 0; c < 3; ++c) {
        out[c] = root[0].mn[c] < root[1].mn[c] ? root[0].mn[c] : root[1].mn[c];
        out[3 + c] = root[0].mx[c] > root[1].mx[c] ? root[0].mx[c] : root[1].mx[c];
    }
    ex.download(&out[6], d_area_sum, sizeof(float));
}

// ---- LBVH builder: Morton codes + radix sort + binary radix tree + bottom-up AABB reduction ------
// For scenes rebuilt every frame (BASELINE config 5; the reference never rebuilds,
// path_tracer.cpp:69).  Prims are sorted along the 30-bit Morton curve of their box centres.
// Hierarchy: the binary radix tree over the sorted codes (every split falls on the highest
// differing bit, i.e. on a cell boundary of the curve; equal codes are split by index bits),
// every internal node found independently from its neighbours' common-prefix lengths; ranges of
// <= 2 slots are leaves.  Boxes: one bottom-up pass in which the second child to arrive at a
// node folds both child boxes, writes the node's QPair + FastNode records and climbs on.
// Fallback when that tree is deeper than the traversal stack (measured height): the balanced
// binary partition of the sorted sequence -- node j of level k owns slots [j*N >> k, (j+1)*N >> k),
// heap-indexed, depth ceil(log2 N) whatever the geometry (boxes overlap more: measured 3.4x
// slower to trace on the 5 M-triangle sheet, which is why it is only the fallback).
CV_HD uint32_t wb_expand10(uint32_t v) {  // 10 bits -> every third bit
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
CV_D void wb_body_morton(const WBuild& b, int p) {
    if (p >= b.n_prims) return;
    const F4 mn = b.bmin[p], mx = b.bmax[p];
    const float c[3] = { 0.5f * (mn.x + mx.x), 0.5f * (mn.y + mx.y), 0.5f * (mn.z + mx.z) };
    uint32_t q[3];
    for (int k = 0; k < 3; ++k) {
        float g = (c[k] - b.morton_min[k]) * b.morton_scale[k];
        g = g < 0.0f ? 0.0f : (g > 1023.0f ? 1023.0f : g);
        q[k] = (uint32_t)(int)g;
    }
    b.keys[p] = (wb_expand10(q[0]) << 2) | (wb_expand10(q[1]) << 1) | wb_expand10(q[2]);
}

// common-prefix length of the sorted keys at slots i and j (ties broken by the slot index)
CV_HD int wb_delta(const uint32_t* keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const uint32_t a = keys[i], c = keys[j];
    return a == c ? 32 + cv_clz((uint32_t)i ^ (uint32_t)j) : cv_clz(a ^ c);
}

// Internal node i of the radix tree: its slot range, split and parent links of its children.
//   node_first/node_last[i] = range; node_split[i] = last slot of the left child;
//   node_parent[child internal node] = i; node_count[i] = how many of its children are leaves.
CV_D void wb_body_radix_node(const WBuild& b, int i) {
    const int n = b.n_prims;
    if (i >= n - 1) return;
    const uint32_t* keys = b.keys;
    const int d = wb_delta(keys, n, i, i + 1) - wb_delta(keys, n, i, i - 1) < 0 ? -1 : 1;
    const int dmin = wb_delta(keys, n, i, i - d);
    int lmax = 2;
    while (wb_delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2) {
        if (wb_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    }
    const int j = i + l * d;
    const int dnode = wb_delta(keys, n, i, j);
    int s = 0;
    for (int t = (l + 1) / 2;; t = (t + 1) / 2) {
        if (wb_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
        if (t <= 1) break;
    }
    const int split = i + s * d + (d < 0 ? -1 : 0);
    const int first = i < j ? i : j, last = i < j ? j : i;
    b.node_first[i] = first;
    b.node_last[i] = last;
    b.node_split[i] = split;
    int leaves = 0;
    if (split - first + 1 <= kFastLeafMax) ++leaves; else b.node_parent[split] = i;
    if (last - split <= kFastLeafMax) ++leaves; else b.node_parent[split + 1] = i;
    b.node_count[i] = leaves | (leaves << 8);  // low byte counts arrivals on top of the leaves; the high byte stays
    if (i == 0) b.node_parent[0] = -1;
}

// child `side` of internal node i: box, reference, triangle count, height
CV_D void wb_radix_child(const WBuild& b, int i, int side, WBox& box, int& ref, int& count, int& height) {
    const int split = b.node_split[i];
    const int lo = side == 0 ? b.node_first[i] : split + 1, hi = side == 0 ? split : b.node_last[i];
    count = hi - lo + 1;
    if (count <= kFastLeafMax) {
        box = wb_empty();
        for (int s = lo; s <= hi; ++s) {
            const int p = b.order[s];
            const F4 mn = b.bmin[p], mx = b.bmax[p];
            wb_grow(box, f3(mn.x, mn.y, mn.z), f3(mx.x, mx.y, mx.z));
        }
        ref = make_leaf_ref(lo, count);
        height = 0;
    } else {
        const int c = side == 0 ? split : split + 1;
        const F4 mn = cv_load_cg(&b.nbox_mn[c]), mx = cv_load_cg(&b.nbox_mx[c]);
        box.mn = f3(mn.x, mn.y, mn.z);
        box.mx = f3(mx.x, mx.y, mx.z);
        ref = 2 * c;
        height = (int)f2u(mn.w);
    }
}

CV_D void wb_write_pair(const WBuild& b, int h, const WBox box[2], const int ref[2], const int count[2]) {
    QPair q;
    for (int s = 0; s < 2; ++s) {
        FastNode o;
        o.mn[0] = box[s].mn.x; o.mn[1] = box[s].mn.y; o.mn[2] = box[s].mn.z;
        o.mx[0] = box[s].mx.x; o.mx[1] = box[s].mx.y; o.mx[2] = box[s].mx.z;
        o.a = ref[s];
        o.b = count[s];
        b.fnodes[2 * h + s] = o;
        for (int c = 0; c < 3; ++c) q.q[3 * s + c] = wb_grid_lo(o.mn[c], b.grid_origin[c], b.grid_scale[c]) | (wb_grid_hi(o.mx[c], b.grid_origin[c], b.grid_scale[c]) << 16);
        if (s == 0) q.r0 = wb_quantised_ref(ref[s]); else q.r1 = wb_quantised_ref(ref[s]);
    }
    b.qpairs[h] = q;
}

// Bottom-up: starts at every node whose two children are leaves; at each parent the first
// arrival stops, the second (node_count reaches 2) has both child boxes and continues.
// The climb is the divergent part (ncu: 2.9 of 32 lanes active per instruction), so it carries
// only the box union and the height; the sibling-pair records (fp32 FastNode pair + quantised
// QPair, two thirds of the old loop body) are written afterwards by wb_body_radix_pairs, one
// fully converged thread per node.
// The climb from node i, whose subtree box/height the caller holds.  After `max_hops` hops the
// climber parks itself in `list_out` (its box is in nbox[i]) and the next launch carries on with
// the survivors packed 32 to a warp: a climb is a chain of dependent memory round trips, and what
// limits the kernel is how many climbers are resident, not instruction issue -- one launch over all
// nodes ends with warps that hold one or two live lanes each (ncu: 2.9 of 32).
CV_D void wb_radix_climb(const WBuild& b, int i, WBox mine, int height, int max_hops, int* list_out, int* count_out) {
    for (int hop = 0;; ++hop) {
        F4 o;
        o.x = mine.mn.x; o.y = mine.mn.y; o.z = mine.mn.z; o.w = u2f((uint32_t)height);
        b.nbox_mn[i] = o;
        o.x = mine.mx.x; o.y = mine.mx.y; o.z = mine.mx.z; o.w = 0.0f;
        b.nbox_mx[i] = o;
        const int parent = b.node_parent[i];
        if (parent < 0) return;
        if (hop >= max_hops) {
            list_out[cv_atomic_add_i32(count_out, 1)] = i;
            return;
        }
        // everything that does not depend on the handshake is requested before it: the parent's range
        // and the box of its leaf children (wb_body_radix_leafbox, an earlier launch)
        const int split = b.node_split[parent], first = b.node_first[parent], last = b.node_last[parent];
        const F4 pmn = b.nbox_mn[parent], pmx = b.nbox_mx[parent];
        cv_threadfence();
        if ((cv_atomic_add_i32(&b.node_count[parent], 1) & 0xff) != 1) return;
        cv_threadfence();
        wb_grow(mine, f3(pmn.x, pmn.y, pmn.z), f3(pmx.x, pmx.y, pmx.z));  // empty (+inf, -inf) when the parent has no leaf child
        const bool i_am_left = i == split;  // interior children are `split` (left) and `split + 1`
        const int sib = i_am_left ? split + 1 : split;
        const int sib_count = i_am_left ? last - split : split - first + 1;
        int hgt = 0;
        if (sib_count > kFastLeafMax) {  // an interior sibling: its climber got there first, the box is final
            const F4 smn = cv_load_cg(&b.nbox_mn[sib]), smx = cv_load_cg(&b.nbox_mx[sib]);
            wb_grow(mine, f3(smn.x, smn.y, smn.z), f3(smx.x, smx.y, smx.z));
            hgt = (int)f2u(smn.w);
        }
        height = 1 + (height > hgt ? height : hgt);
        i = parent;
    }
}

// Before the climb: every interior node folds the prim boxes of its LEAF children into nbox (empty if it has
// none) -- converged, no dependencies; the climb then never gathers prims.
CV_D void wb_body_radix_leafbox(const WBuild& b, int i) {
    if (i >= b.n_prims - 1) return;
    const int first = b.node_first[i], last = b.node_last[i], split = b.node_split[i];
    if (last - first + 1 <= kFastLeafMax) return;
    WBox box = wb_empty();
    if (split - first + 1 <= kFastLeafMax) {
        for (int s = first; s <= split; ++s) {
            const int p = b.order[s];
            const F4 mn = b.bmin[p], mx = b.bmax[p];
            wb_grow(box, f3(mn.x, mn.y, mn.z), f3(mx.x, mx.y, mx.z));
        }
    }
    if (last - split <= kFastLeafMax) {
        for (int s = split + 1; s <= last; ++s) {
            const int p = b.order[s];
            const F4 mn = b.bmin[p], mx = b.bmax[p];
            wb_grow(box, f3(mn.x, mn.y, mn.z), f3(mx.x, mx.y, mx.z));
        }
    }
    F4 o;
    o.x = box.mn.x; o.y = box.mn.y; o.z = box.mn.z; o.w = u2f(0u);
    b.nbox_mn[i] = o;
    o.x = box.mx.x; o.y = box.mx.y; o.z = box.mx.z; o.w = 0.0f;
    b.nbox_mx[i] = o;
}

CV_D void wb_body_radix_boxes(const WBuild& b, int i, int max_hops, int* list_out, int* count_out) {
    if (i >= b.n_prims - 1) return;
    if (b.node_last[i] - b.node_first[i] + 1 <= kFastLeafMax) return;  // folded into its parent as a leaf
    if ((b.node_count[i] >> 8) != 2) return;  // start only where both children are leaves
    const F4 mn = b.nbox_mn[i], mx = b.nbox_mx[i];  // both children folded by wb_body_radix_leafbox
    WBox mine;
    mine.mn = f3(mn.x, mn.y, mn.z);
    mine.mx = f3(mx.x, mx.y, mx.z);
    wb_radix_climb(b, i, mine, 1, max_hops, list_out, count_out);
}

// a later round: climber k of the previous round's list
CV_D void wb_body_radix_resume(const WBuild& b, int k, const int* list_in, const int* count_in, int max_hops, int* list_out, int* count_out) {
    if (k >= *count_in) return;  // written by the previous launch
    const int i = list_in[k];
    const F4 mn = cv_load_cg(&b.nbox_mn[i]), mx = cv_load_cg(&b.nbox_mx[i]);
    WBox mine;
    mine.mn = f3(mn.x, mn.y, mn.z);
    mine.mx = f3(mx.x, mx.y, mx.z);
    wb_radix_climb(b, i, mine, (int)f2u(mn.w), max_hops, list_out, count_out);
}

// After the climb every interior node's box is final: write each node's sibling pair.
CV_D void wb_body_radix_pairs(const WBuild& b, int i) {
    if (i >= b.n_prims - 1) return;
    if (b.node_last[i] - b.node_first[i] + 1 <= kFastLeafMax) return;
    WBox box[2];
    int ref[2], count[2], hgt[2];
    wb_radix_child(b, i, 0, box[0], ref[0], count[0], hgt[0]);
    wb_radix_child(b, i, 1, box[1], ref[1], count[1], hgt[1]);
    wb_write_pair(b, i, box, ref, count);
}

// ---- hybrid builder (CAVEPT_BVH_HYBRID): the radix tree above, binned-SAH subtrees below -------------------------
// After the radix tree is complete (boxes, pairs): every interior node of at most kWSubMax prims whose parent has more
// is the root of a treelet that the SAH builder's subtree launch rebuilds from the prims (wb_subtree_* / k_wb_subtrees).
// The treelet keeps the radix node's pair slot, so the record its parent holds stays valid; the pairs of its subtree
// go to slots [first + 1, last - 1] -- the count - 2 radix nodes below it, dead from here on (a radix subtree over
// [first, last] owns exactly those indices besides its own, which is `first` or `last`; the other end belongs to an
// ancestor) -- and a binary tree over `count` prims needs at most count - 2 pairs below its root.  The pair array keeps
// its size and the top of the tree costs what the LBVH costs, not one pass over all prims per level.
CV_D void wb_body_hybrid_root(const WBuild& b, int i) {
    if (i >= b.n_prims - 1) return;
    const int first = b.node_first[i], last = b.node_last[i], count = last - first + 1;
    if (count <= kFastLeafMax || count > kWSubMax) return;
    int parent = i == 0 ? -1 : b.node_parent[i];
    if (parent >= 0 && b.node_last[parent] - b.node_first[parent] + 1 <= kWSubMax) return;  // inside a treelet
    int depth = 0;
    for (int a = parent; a >= 0; a = b.node_parent[a]) ++depth;
    const int q = cv_atomic_add_i32(&b.counters[3], 1);
    if (q >= b.sub_capacity) {
        cv_atomic_or_i32(&b.counters[2], WB_ERR_CAPACITY);
        return;
    }
    WItem it;
    it.start = first;
    it.count = count;
    it.pair = i;
    it.depth = depth;
    it.split = it.n_left = 0;
    it.child_item[0] = it.child_item[1] = -1;
    it.bins_slot = -1;
    it.scan_tile = it.scan_local = 0;
    it.pair_lo = first + 1;
    const F4 mn = b.nbox_mn[i], mx = b.nbox_mx[i];
    it.box_mn[0] = mn.x; it.box_mn[1] = mn.y; it.box_mn[2] = mn.z;
    it.box_mx[0] = mx.x; it.box_mx[1] = mx.y; it.box_mx[2] = mx.z;
    wb_item_setup(it, b.force_median);
    b.pair_depth[i] = (unsigned char)depth;
    b.sub_roots[q] = it;
}

CV_HD int wb_lbvh_lo(int n, int k, int j) { return (int)(((long long)j * (long long)n) >> k); }

// box of the node (k, j) seen from its parent: a leaf's box is folded from its prims, an
// interior node's was written by the previous (deeper) launch
CV_D void wb_lbvh_child(const WBuild& b, int k, int j, WBox& box, int& ref, int& count) {
    const int lo = wb_lbvh_lo(b.n_prims, k, j), hi = wb_lbvh_lo(b.n_prims, k, j + 1);
    count = hi - lo;
    if (count <= kFastLeafMax) {
        box = wb_empty();
        for (int s = lo; s < hi; ++s) {
            const int p = b.order[s];
            const F4 mn = b.bmin[p], mx = b.bmax[p];
            wb_grow(box, f3(mn.x, mn.y, mn.z), f3(mx.x, mx.y, mx.z));
        }
        ref = make_leaf_ref(lo, count);
    } else {
        const int h = (1 << k) - 1 + j;
        const F4 mn = b.nbox_mn[h], mx = b.nbox_mx[h];
        box.mn = f3(mn.x, mn.y, mn.z);
        box.mx = f3(mx.x, mx.y, mx.z);
        ref = 2 * h;  // FastNode-array convention: index of the child pair's first record
    }
}

CV_D void wb_body_lbvh_level(const WBuild& b, int j, int k) {
    if (j >= (1 << k)) return;
    const int lo = wb_lbvh_lo(b.n_prims, k, j), hi = wb_lbvh_lo(b.n_prims, k, j + 1);
    if (hi - lo <= kFastLeafMax) return;  // a leaf: its parent folds it
    const int h = (1 << k) - 1 + j;
    WBox box[2];
    int ref[2], count[2];
    wb_lbvh_child(b, k + 1, 2 * j, box[0], ref[0], count[0]);
    wb_lbvh_child(b, k + 1, 2 * j + 1, box[1], ref[1], count[1]);
    wb_write_pair(b, h, box, ref, count);
    wb_grow(box[0], box[1].mn, box[1].mx);
    F4 o;
    o.x = box[0].mn.x; o.y = box[0].mn.y; o.z = box[0].mn.z; o.w = 0.0f;
    b.nbox_mn[h] = o;
    o.x = box[0].mx.x; o.y = box[0].mx.y; o.z = box[0].mx.z;
    b.nbox_mx[h] = o;
}

struct FnWbPrim { WBuild b; CV_D void operator()(int i) const { wb_body_prim(b, i); } };
struct FnWbInit { WBuild b; CV_D void operator()(int i) const { wb_body_init(b, i); } };
struct FnWbBin { WBuild b; CV_D void operator()(int i) const { wb_body_bin(b, i); } };
struct FnWbSplit { WBuild b; int level; bool with_arena_items; CV_D void operator()(int i) const { wb_body_split(b, i, level, with_arena_items); } };
struct FnWbFlag { WBuild b; CV_D void operator()(int i) const { wb_body_flag(b, i); } };
struct FnWbScatter { WBuild b; CV_D void operator()(int i) const { wb_body_scatter(b, i); } };
struct FnWbEmit { WBuild b; CV_D void operator()(int i) const { wb_body_emit(b, i); } };
struct FnWbMorton { WBuild b; CV_D void operator()(int i) const { wb_body_morton(b, i); } };
struct FnWbRadixNode { WBuild b; CV_D void operator()(int i) const { wb_body_radix_node(b, i); } };
struct FnWbRadixBoxes { WBuild b; int max_hops; int* list_out; int* count_out; CV_D void operator()(int i) const { wb_body_radix_boxes(b, i, max_hops, list_out, count_out); } };
struct FnWbRadixResume { WBuild b; const int* list_in; const int* count_in; int max_hops; int* list_out; int* count_out; CV_D void operator()(int i) const { wb_body_radix_resume(b, i, list_in, count_in, max_hops, list_out, count_out); } };
struct FnWbRadixLeafBox { WBuild b; CV_D void operator()(int i) const { wb_body_radix_leafbox(b, i); } };
struct FnWbRadixPairs { WBuild b; CV_D void operator()(int i) const { wb_body_radix_pairs(b, i); } };
struct FnWbHybridRoot { WBuild b; CV_D void operator()(int i) const { wb_body_hybrid_root(b, i); } };
struct FnWbLbvhLevel { WBuild b; int k; CV_D void operator()(int i) const { wb_body_lbvh_level(b, i, k); } };
struct FnWbQuantise { WBuild b; int n; CV_D void operator()(int i) const { wb_body_quantise(b, i, n); } };

struct WReport {
    int levels;
    int n_pairs;
    int root_ref;
    int error_bits;
    float scene_mn[3], scene_mx[3];
    float pad;
    float grid_origin[3], grid_scale[3];
    int level_items[kWMaxLevels + 4];  // SAH builder: interior nodes (= sibling pairs owned) per level; the refit walks these ranges
};

// Common first part of both production builders: world AABB of every prim, scene and centroid
// bounds, the box padding and the quantisation grid, identity order.  Returns 1 when the whole
// scene fits one leaf (done), 0 to continue.  acc[6..11] = centroid bounds keys.
template<class Exec>
int wb_prologue(Exec& ex, WBuild& b, WReport* rep, uint32_t acc[12]) {
    rep->levels = 0;
    rep->n_pairs = 0;
    rep->error_bits = 0;
    rep->root_ref = kRefSentinel;
    if (b.n_prims <= 0) return 1;
    for (int k = 0; k < 12; ++k) acc[k] = ((k / 3) % 2 == 0) ? 0xffffffffu : 0u;
    ex.upload(b.root_acc, acc, 12 * sizeof(uint32_t));
    ex.wb_prim(b);
    ex.download(acc, b.root_acc, 12 * sizeof(uint32_t));
    float scale = 0.0f;
    for (int k = 0; k < 3; ++k) {
        rep->scene_mn[k] = ordered_value(acc[k]);
        rep->scene_mx[k] = ordered_value(acc[3 + k]);
        const float a0 = rep->scene_mn[k] < 0 ? -rep->scene_mn[k] : rep->scene_mn[k], a1 = rep->scene_mx[k] < 0 ? -rep->scene_mx[k] : rep->scene_mx[k];
        scale = scale > a0 ? scale : a0;
        scale = scale > a1 ? scale : a1;
    }
    b.pad = scale * (1.0f / 1048576.0f) + 1e-30f;
    rep->pad = b.pad;
    for (int k = 0; k < 3; ++k) {  // the quantisation grid spans the padded scene box plus two spare cells on each side,
        // so outward rounding never meets the clamp at 0 / 32767
        const float lo = rep->scene_mn[k] - b.pad, hi = rep->scene_mx[k] + b.pad;
        const float gs = (kGridCells - 4.0f) / (hi - lo);
        b.grid_scale[k] = rep->grid_scale[k] = gs;
        b.grid_origin[k] = rep->grid_origin[k] = lo - 2.0f / gs;
    }
    ex.launch(b.n_prims, FnWbInit{ b });
    if (b.n_prims <= kFastLeafMax) {
        rep->root_ref = make_leaf_ref(0, b.n_prims);
        ex.launch(b.n_prims, FnWbEmit{ b });
        return 1;
    }
    return 0;
}

// Host-side level loop (Exec = CudaExec on the device, EmuExec in the test emulator).  Per level:
//   bin        one pass over the prims: count + AABB into the arena bins of items of more than kWDirectMax prims
//   split      per item: SAH sweep over all binned axes, children created with their binning set-up (Exec::wb_split:
//              the emulator runs wb_body_split per item; the device runs a warp per arena item + a thread per small item)
//   partition  left flags -> exclusive scan -> stable scatter (Exec::wb_partition: three serial passes in the
//              emulator; on the device the flags are computed inside the scan's first kernel and the scatter inside its last)
template<class Exec>
int build_wbvh(Exec& ex, WBuild b, WReport* rep) {
    uint32_t acc[12];
    if (wb_prologue(ex, b, rep, acc)) return 0;
    WItem root;
    root.start = 0;
    root.count = b.n_prims;
    root.pair = 0;
    root.depth = 0;
    root.split = root.n_left = 0;
    root.child_item[0] = root.child_item[1] = -1;
    root.bins_slot = b.n_prims > kWDirectMax ? 0 : -1;
    root.scan_tile = root.scan_local = 0;
    root.pair_lo = -1;
    for (int k = 0; k < 3; ++k) {
        root.box_mn[k] = rep->scene_mn[k] - b.pad;
        root.box_mx[k] = rep->scene_mx[k] + b.pad;
    }
    wb_item_setup(root, b.force_median);
    ex.upload(b.items, &root, sizeof(root));
    ex.fill(b.bins, 0, (size_t)b.bins_capacity * sizeof(WBins));  // once: the split pass clears what it reads
    ex.fill(b.pair_depth, 0, 1);                                  // the root owns pair 0
    // counters: [1] pairs allocated, [2] error bits, [kWLevelBase + L] interior items of level L, [kWBinBase + L]
    // arena slots of level L.  The level loop runs kLevelsPerSync levels per device->host round trip: launches over
    // items are sized by an upper bound (items are disjoint ranges of >= kFastLeafMax + 1 prims, and at most 2^L of
    // them) and read the true count on the device; a level past the last one finds no items and no prims to move.
    constexpr int kLevelsPerSync = 4;
    int counters[kWCounters] = { 0, 1, 0, 0, 1 };
    counters[kWBinBase] = 1;
    ex.upload(b.counters, counters, sizeof(counters));
    b.level_items = b.counters + kWLevelBase;
    b.level_bins = b.counters + kWBinBase;
    const int items_bound = b.n_prims / (kFastLeafMax + 1) + 1;
    rep->root_ref = 0;
    int level = 0;
    if (b.n_prims > kWSubMax || !b.sub_roots) {  // level-synchronous phase: items of more than kWSubMax prims
        for (;;) {
            // the level loop cannot end before items of kWSubMax prims exist: no read-back until then, every two levels after
            int burst = kLevelsPerSync;
            if (b.sub_roots) {
                burst = 2;
                if (level == 0) {
                    burst = 1;
                    while (((long long)kWSubMax << burst) < (long long)b.n_prims) ++burst;
                }
            }
            for (int k = 0; k < burst; ++k, ++level) {
                const long long pow2 = level < 30 ? (1ll << level) : (1ll << 30);
                const int bound = (int)(pow2 < items_bound ? pow2 : items_bound);
                ex.wb_bin(b);
                ex.wb_split(b, level, bound);
                ex.wb_partition(b);
                WItem* ti = b.items; b.items = b.items_next; b.items_next = ti;
                int* to = b.order; b.order = b.order_next; b.order_next = to;
                int* tn = b.node_of; b.node_of = b.node_of_next; b.node_of_next = tn;
            }
            ex.download(counters, b.counters, sizeof(counters));
            if (counters[2] != 0) {
                rep->error_bits = counters[2];
                return -1;
            }
            if (counters[kWLevelBase + level] == 0) break;
            if (level + 32 >= kWMaxLevels) {
                rep->error_bits = WB_ERR_CAPACITY;
                return -1;
            }
        }
    } else {  // the whole scene is one subtree
        ex.upload(b.sub_roots, &root, sizeof(root));
        counters[3] = 1;
        counters[kWLevelBase] = 0;  // the subtree build accounts for its own levels
        ex.upload(b.counters, counters, sizeof(counters));
    }
    // subtree phase: one block per queued item
    const int n_roots = counters[3] < b.sub_capacity ? counters[3] : b.sub_capacity;
    if (n_roots > 0) {
        ex.wb_subtrees(b, n_roots);
        ex.download(counters, b.counters, sizeof(counters));
        if (counters[2] != 0) {
            rep->error_bits = counters[2];
            return -1;
        }
    }
    rep->n_pairs = counters[1];
    rep->levels = 0;
    while (rep->levels < kWMaxLevels && counters[kWLevelBase + rep->levels] > 0) {
        rep->level_items[rep->levels] = counters[kWLevelBase + rep->levels];
        ++rep->levels;
    }
    ex.launch(b.n_prims, FnWbEmit{ b });
    ex.launch(rep->n_pairs, FnWbQuantise{ b, rep->n_pairs });
    if (b.q4nodes) ex.launch(rep->n_pairs, FnWbQ4{ b, rep->n_pairs });
    return 0;
}

// LBVH build (see wb_body_morton): prologue, Morton codes, radix sort, one launch per level
// bottom-up, emit.  No device->host round trip after the prologue.
// hybrid = true: the radix tree's treelets of at most kWSubMax prims are rebuilt by the SAH subtree launch (wb_body_hybrid_root)
template<class Exec>
int build_lbvh(Exec& ex, WBuild b, WReport* rep, bool hybrid = false) {
    uint32_t acc[12];
    if (wb_prologue(ex, b, rep, acc)) return 0;
    for (int k = 0; k < 3; ++k) {
        const float lo = ordered_value(acc[6 + k]), hi = ordered_value(acc[9 + k]);
        b.morton_min[k] = lo;
        b.morton_scale[k] = hi > lo ? 1024.0f / (hi - lo) : 0.0f;
    }
    ex.launch(b.n_prims, FnWbMorton{ b });
    ex.sort_pairs(b.keys, b.order, b.keys_tmp, b.order_next, b.sort_hist, b.n_prims, 30);
    // radix tree; its height comes back with the root box
    b.node_first = b.node_of;
    b.node_last = b.node_of_next;
    b.node_split = b.order_next;  // free once the sort is done
    b.node_parent = b.scan;
    b.node_count = reinterpret_cast<int*>(b.keys_tmp);
    ex.launch(b.n_prims - 1, FnWbRadixNode{ b });
    {   // bottom-up boxes in rounds of kClimbHops hops; survivors are re-packed between rounds (see wb_radix_climb)
        constexpr int kClimbHops = 3, kClimbRounds = 8;
        int* lists[2] = { reinterpret_cast<int*>(b.keys), reinterpret_cast<int*>(b.flags) };  // both free now; n and n/4 ints: a round's survivors are <= n/(3r+1)
        int* counts = reinterpret_cast<int*>(b.root_acc);                                      // 12 ints, consumed by the prologue
        ex.fill(counts, 0, 12 * sizeof(int));
        const int n_nodes = b.n_prims - 1;
        ex.launch(n_nodes, FnWbRadixLeafBox{ b });
        ex.launch(n_nodes, FnWbRadixBoxes{ b, kClimbHops, lists[0], counts + 0 });
        for (int r = 1; r < kClimbRounds; ++r) {
            const bool last = r == kClimbRounds - 1;
            ex.launch(n_nodes / (kClimbHops * r + 1) + 1,
                      FnWbRadixResume{ b, lists[(r - 1) & 1], counts + (r - 1), last ? 0x7fffffff : kClimbHops, lists[r & 1], counts + r });
        }
    }
    ex.launch(b.n_prims - 1, FnWbRadixPairs{ b });
    F4 root_mn;
    ex.download(&root_mn, b.nbox_mn, sizeof(F4));
    const int height = (int)f2u(root_mn.w);
    rep->levels = height;
    rep->n_pairs = b.n_prims - 1;  // internal-node slots in use (collapsed leaves are holes)
    rep->root_ref = 0;
    bool too_deep = height + 2 > kStackSize;
    if (hybrid && !too_deep && !b.force_median) {
        int counters[kWCounters] = { 0 };
        ex.upload(b.counters, counters, sizeof(counters));
        b.level_items = b.counters + kWLevelBase;
        b.level_bins = b.counters + kWBinBase;
        ex.launch(b.n_prims - 1, FnWbHybridRoot{ b });
        ex.download(counters, b.counters, sizeof(counters));
        if (counters[2] != 0) {
            rep->error_bits = counters[2];
            return -1;
        }
        ex.wb_subtrees(b, counters[3]);
        ex.download(counters, b.counters, sizeof(counters));
        if (counters[2] != 0) {
            rep->error_bits = counters[2];
            return -1;
        }
        // a treelet may be deeper than the radix subtree it replaces: the height is what the subtree builds report
        if (counters[0] > rep->levels) rep->levels = counters[0];
        too_deep = rep->levels + 2 > kStackSize;
        if (!too_deep) ex.launch(rep->n_pairs, FnWbQuantise{ b, rep->n_pairs });
    }
    if (too_deep || b.force_median) {
        int k_last = 0;  // deepest level that still has an interior node (sizes at a level differ by <= 1)
        while ((int)(((long long)b.n_prims + (1ll << (k_last + 1)) - 1) >> (k_last + 1)) > kFastLeafMax) ++k_last;
        for (int k = k_last; k >= 0; --k) ex.launch(1 << k, FnWbLbvhLevel{ b, k });
        rep->levels = k_last + 1;
        rep->n_pairs = 1 << (k_last + 1);  // heap slots in use (leaf positions are holes)
    }
    if (b.q4nodes) ex.launch(rep->n_pairs, FnWbQ4{ b, rep->n_pairs });
    ex.launch(b.n_prims, FnWbEmit{ b });
    return 0;
}

}  // namespace cavept
