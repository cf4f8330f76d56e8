// Host-side math of the scene upload (the work the reference does on the CPU in
// PathTracer::UpdateAccelStructure, cave/engine/renderer/path_tracer/path_tracer.cpp:162-199).
#pragma once

namespace cavept {

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
