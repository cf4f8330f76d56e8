// Top-level tree over instances, built on the host from each instance's world-space
// box.  The reference has no TLAS: HitScene loops over every mesh
// (cave/shader/shared_path_tracer.h:159).  Instance count is the number of
// MeshRendererComponents (tens to a few thousand), so this is O(M log M) host work
// done where the reference's CPU-side UpdateAccelStructure runs
// (cave/engine/renderer/path_tracer/path_tracer.cpp:162-199); the result only prunes
// and orders instance visits, it never changes which triangle is the closest hit.
#pragma once
#include <algorithm>
#include <cmath>
#include <vector>

#include "trace_body.h"

namespace cavept {

struct WorldBox {
    float mn[3], mx[3];
};

// Box of the 8 corners of [mn, mx] under a column-major 4x4, padded so that float
// rounding in the reference's local-space tests can never see a hit outside it.
inline WorldBox transform_box(const float* m, const float* mn, const float* mx) {
    WorldBox w;
    for (int k = 0; k < 3; ++k) {
        w.mn[k] = INFINITY;
        w.mx[k] = -INFINITY;
    }
    float amax = 0.0f;
    for (int c = 0; c < 8; ++c) {
        const float x = (c & 1) ? mx[0] : mn[0], y = (c & 2) ? mx[1] : mn[1], z = (c & 4) ? mx[2] : mn[2];
        for (int k = 0; k < 3; ++k) {
            const float v = m[0 + k] * x + m[4 + k] * y + m[8 + k] * z + m[12 + k];
            w.mn[k] = std::min(w.mn[k], v);
            w.mx[k] = std::max(w.mx[k], v);
            amax = std::max(amax, std::fabs(v));
        }
    }
    const float pad = 1e-4f * amax + 1e-5f;
    for (int k = 0; k < 3; ++k) {
        w.mn[k] -= pad;
        w.mx[k] += pad;
    }
    return w;
}

namespace detail {
inline void tlas_fill(const std::vector<WorldBox>& boxes, std::vector<int>& ids, int lo, int hi, int slot, int base, std::vector<FastNode>& out, int depth,
                      int& max_depth) {
    FastNode n;
    for (int k = 0; k < 3; ++k) {
        n.mn[k] = INFINITY;
        n.mx[k] = -INFINITY;
    }
    float cmn[3] = { INFINITY, INFINITY, INFINITY }, cmx[3] = { -INFINITY, -INFINITY, -INFINITY };
    for (int i = lo; i < hi; ++i) {
        const WorldBox& b = boxes[ids[i]];
        for (int k = 0; k < 3; ++k) {
            n.mn[k] = std::min(n.mn[k], b.mn[k]);
            n.mx[k] = std::max(n.mx[k], b.mx[k]);
            const float c = 0.5f * (b.mn[k] + b.mx[k]);
            cmn[k] = std::min(cmn[k], c);
            cmx[k] = std::max(cmx[k], c);
        }
    }
    max_depth = std::max(max_depth, depth);
    if (hi - lo == 1) {
        n.a = make_instance_ref(ids[lo]);
        n.b = 1;
        out[slot] = n;
        return;
    }
    int axis = 0;
    if (cmx[1] - cmn[1] > cmx[axis] - cmn[axis]) axis = 1;
    if (cmx[2] - cmn[2] > cmx[axis] - cmn[axis]) axis = 2;
    const int mid = (lo + hi) / 2;
    std::stable_sort(ids.begin() + lo, ids.begin() + hi, [&](int x, int y) {
        return boxes[x].mn[axis] + boxes[x].mx[axis] < boxes[y].mn[axis] + boxes[y].mx[axis];
    });
    const int pair = (int)out.size();
    out.resize(out.size() + 2);
    n.a = base + pair;  // interior reference = index of the child pair
    n.b = hi - lo;
    out[slot] = n;
    tlas_fill(boxes, ids, lo, mid, pair, base, out, depth + 1, max_depth);
    tlas_fill(boxes, ids, mid, hi, pair + 1, base, out, depth + 1, max_depth);
}
}  // namespace detail

// Child pairs are written to out[0..]; references are absolute (base + position in out), so
// `base` must be even-aligned with the FastNode array position out[0] will be copied to.
// The root has no record: its node reference is returned in *root_ref.  Returns depth.
inline int build_tlas(const std::vector<WorldBox>& boxes, int base, std::vector<FastNode>& out, int* root_ref) {
    out.clear();
    out.resize(1);  // scratch slot 0 receives the root record; callers skip it
    std::vector<int> ids(boxes.size());
    for (size_t i = 0; i < ids.size(); ++i) ids[i] = (int)i;
    int max_depth = 0;
    // out[0] is a scratch root record; pairs start at out[1], hence base - 1 below
    detail::tlas_fill(boxes, ids, 0, (int)boxes.size(), 0, base - 1, out, 0, max_depth);
    *root_ref = out[0].a;
    out.erase(out.begin());
    return max_depth;
}

// glm::inverse(mat4) restated for the instance array (path_tracer.cpp:178); same
// cofactor grouping as thirdparty/glm/glm/detail/func_matrix.inl:388-446 so the fp32
// result matches the reference bit for bit.  Column-major, m[c*4 + r].
inline void mat4_inverse(const float* m, float* out) {
    auto at = [&](int c, int r) { return m[c * 4 + r]; };
    const float s00 = at(2, 2) * at(3, 3) - at(3, 2) * at(2, 3), s02 = at(1, 2) * at(3, 3) - at(3, 2) * at(1, 3), s03 = at(1, 2) * at(2, 3) - at(2, 2) * at(1, 3);
    const float s04 = at(2, 1) * at(3, 3) - at(3, 1) * at(2, 3), s06 = at(1, 1) * at(3, 3) - at(3, 1) * at(1, 3), s07 = at(1, 1) * at(2, 3) - at(2, 1) * at(1, 3);
    const float s08 = at(2, 1) * at(3, 2) - at(3, 1) * at(2, 2), s10 = at(1, 1) * at(3, 2) - at(3, 1) * at(1, 2), s11 = at(1, 1) * at(2, 2) - at(2, 1) * at(1, 2);
    const float s12 = at(2, 0) * at(3, 3) - at(3, 0) * at(2, 3), s14 = at(1, 0) * at(3, 3) - at(3, 0) * at(1, 3), s15 = at(1, 0) * at(2, 3) - at(2, 0) * at(1, 3);
    const float s16 = at(2, 0) * at(3, 2) - at(3, 0) * at(2, 2), s18 = at(1, 0) * at(3, 2) - at(3, 0) * at(1, 2), s19 = at(1, 0) * at(2, 2) - at(2, 0) * at(1, 2);
    const float s20 = at(2, 0) * at(3, 1) - at(3, 0) * at(2, 1), s22 = at(1, 0) * at(3, 1) - at(3, 0) * at(1, 1), s23 = at(1, 0) * at(2, 1) - at(2, 0) * at(1, 1);
    const float fac[6][4] = { { s00, s00, s02, s03 }, { s04, s04, s06, s07 }, { s08, s08, s10, s11 },
                              { s12, s12, s14, s15 }, { s16, s16, s18, s19 }, { s20, s20, s22, s23 } };
    float vec[4][4];
    for (int k = 0; k < 4; ++k) {
        vec[k][0] = at(1, k);
        vec[k][1] = vec[k][2] = vec[k][3] = at(0, k);
    }
    float cof[16];
    for (int r = 0; r < 4; ++r) {
        const float sign_a = (r & 1) ? -1.0f : 1.0f;
        cof[0 + r] = (vec[1][r] * fac[0][r] - vec[2][r] * fac[1][r] + vec[3][r] * fac[2][r]) * sign_a;
        cof[4 + r] = (vec[0][r] * fac[0][r] - vec[2][r] * fac[3][r] + vec[3][r] * fac[4][r]) * -sign_a;
        cof[8 + r] = (vec[0][r] * fac[1][r] - vec[1][r] * fac[3][r] + vec[3][r] * fac[5][r]) * sign_a;
        cof[12 + r] = (vec[0][r] * fac[2][r] - vec[1][r] * fac[4][r] + vec[2][r] * fac[5][r]) * -sign_a;
    }
    const float det = (at(0, 0) * cof[0] + at(0, 1) * cof[4]) + (at(0, 2) * cof[8] + at(0, 3) * cof[12]);
    const float inv_det = 1.0f / det;
    for (int i = 0; i < 16; ++i) out[i] = cof[i] * inv_det;
}

}  // namespace cavept
