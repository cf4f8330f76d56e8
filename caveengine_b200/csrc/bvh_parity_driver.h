// Host-side level loop of the parity builder, written against an "Exec" policy so the
// same orchestration runs on the device (CudaExec in kernels.cu) and in the test-only
// serial emulator (tests/emu).  Exec provides:
//   template<class F> void launch(int n, F f);                 // f(i) for i in [0, n)
//   void scan_u8(const unsigned char* in, int* out, int n);    // exclusive, out[n] = total
//   void scan_i32(const int* in, int* out, int n);
//   void upload(void* dst, const void* src, size_t bytes);
//   void download(void* dst, const void* src, size_t bytes);   // synchronises
//   void fill(void* dst, int byte, size_t bytes);
//   void parity_reduce(const BuildArrays&); void parity_bin(const BuildArrays&);   // body_reduce / body_bin over all triangles
//   void parity_split(const BuildArrays&, int n_items);         // body_split over the level's items
//   void parity_subtrees(const BuildArrays&, int n_roots);     // second phase: one block (or the serial statement) per queued root
#pragma once
#include <stddef.h>

#include "bvh_parity_body.h"

namespace cavept {

struct FnTriPrep { BuildArrays a; CV_D void operator()(int i) const { body_tri_prep(a, i); } };
struct FnInitAccum { BuildArrays a; int n; CV_D void operator()(int i) const { body_init_accum(a, i, n); } };
struct FnReduce { BuildArrays a; CV_D void operator()(int i) const { body_reduce(a, i); } };
struct FnNode { BuildArrays a; int n; CV_D void operator()(int i) const { body_node(a, i, n); } };
struct FnBin { BuildArrays a; CV_D void operator()(int i) const { body_bin(a, i); } };
struct FnSplit { BuildArrays a; int n; CV_D void operator()(int i) const { body_split(a, i, n); } };
struct FnFlag { BuildArrays a; CV_D void operator()(int i) const { body_flag(a, i); } };
struct FnChildren { BuildArrays a; int n; CV_D void operator()(int i) const { body_children(a, i, n); } };
struct FnScatter { BuildArrays a; CV_D void operator()(int i) const { body_scatter(a, i); } };

// order[i] = i, node_of[i] = index of the mesh whose triangle range contains i
struct FnInitOrder {
    BuildArrays a;
    const int* tri_off;  // n_meshes + 1 ascending offsets
    int n_meshes;
    CV_D void operator()(int i) const {
        int lo = 0, hi = n_meshes;  // last m with tri_off[m] <= i
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (tri_off[mid] <= i) lo = mid; else hi = mid;
        }
        a.order[i] = i;
        a.node_of[i] = lo;
    }
};

struct BuildReport {
    int levels;
    int error_bits;
    int launches;
};

template<class T>
CV_HD void swap_ptr(T*& x, T*& y) {
    T* t = x;
    x = y;
    y = t;
}

// tri_off_host: n_meshes + 1 offsets into the scene-global triangle array.  Mesh m's
// nodes occupy [2*tri_off[m] - m, ...) -- the concatenation CreateAccelStructure makes
// (path_tracer.cpp:233-243).  d_tri_off is the same array on the Exec's side.
template<class Exec>
int build_bvh_parity(Exec& ex, BuildArrays a, const int* tri_off_host, const int* d_tri_off, int n_meshes, BuildItem* host_items_scratch,
                     BuildReport* report) {
    report->levels = 0;
    report->error_bits = 0;
    if (a.n_tris <= 0 || n_meshes <= 0) return 0;
    ex.launch(a.n_tris, FnTriPrep{ a });
    ex.launch(a.n_tris, FnInitOrder{ a, d_tri_off, n_meshes });
    for (int m = 0; m < n_meshes; ++m) {
        BuildItem r;
        r.start = tri_off_host[m];
        r.count = tri_off_host[m + 1] - tri_off_host[m];
        r.node = 2 * tri_off_host[m] - m;
        r.miss = -1;
        r.parent = -1 - m;
        r.depth = 0;
        r.kind = ITEM_LEAF;
        r.axis = 0;
        r.tmin = r.tmax = r.parent_sa = 0.0f;
        r.split = r.n_left = 0;
        r.child_slot = r.child_slot_r = r.bin_slot = -1;
        r.tri_offset = tri_off_host[m];
        host_items_scratch[m] = r;
    }
    ex.upload(a.items, host_items_scratch, sizeof(BuildItem) * (size_t)n_meshes);
    int n_items = n_meshes;
    int counters[kParityCounters] = { 0, 0, 0, 0, 0 };
    ex.upload(a.counters, counters, sizeof(counters));
    while (n_items > 0) {
        ex.launch(n_items, FnInitAccum{ a, n_items });
        ex.parity_reduce(a);
        ex.launch(n_items, FnNode{ a, n_items });
        ex.parity_bin(a);
        ex.parity_split(a, n_items);
        ex.launch(a.n_tris, FnFlag{ a });
        ex.scan_u8(a.flags, a.scan, a.n_tris);
        ex.launch(n_items, FnChildren{ a, n_items });
        ex.launch(a.n_tris, FnScatter{ a });
        ex.download(counters, a.counters, sizeof(counters));
        ++report->levels;
        if (counters[2] != 0) {
            report->error_bits = counters[2];
            return -1;
        }
        n_items = counters[0];
        counters[0] = 0;
        counters[1] = 0;
        ex.upload(a.counters, counters, 2 * sizeof(int));
        swap_ptr(a.items, a.items_next);
        swap_ptr(a.order, a.order_next);
        swap_ptr(a.node_of, a.node_of_next);
        if (report->levels > kMaxDepth + 2) {
            report->error_bits = BUILD_ERR_DEPTH;
            return -1;
        }
    }
    // second phase: every queued item's whole subtree in one launch (a.order is the buffer the last level scattered into)
    const int n_roots = counters[3];
    if (n_roots > 0) {
        ex.parity_subtrees(a, n_roots);
        ex.download(counters, a.counters, sizeof(counters));
        if (counters[4] > report->levels) report->levels = counters[4];
        if (counters[2] != 0) {
            report->error_bits = counters[2];
            return -1;
        }
    }
    return 0;
}

}  // namespace cavept
