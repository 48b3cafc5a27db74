// ff_hostmath.h -- host-side 4x4 transform and camera matrices in the reference's (glm 0.9.9.7) operation order.
//
// The model matrix of a geometry (utilities.h:179-189: T * Rx * Ry * Rz * S, degrees) and the camera's
// inverse(view) * inverse(projection) (utilities.h:299-317, kernel.cu:203) are computed ONCE on the host here and
// handed to the kernels; the reference recomputes two 4x4 inverses per thread.  The order of the floating-point
// operations follows glm's (ext/matrix_transform.inl:10-119, ext/matrix_clip_space.inl:372-389,
// detail/func_matrix.inl:294-352, detail/type_mat4x4.inl:630-648) so that baked vertices and camera rays have
// the same bits as the reference's.
#pragma once
#include <cmath>
#include "ff_math.h"

namespace ffh {
using ffm::M4;
using ffm::V3;
using ffm::V4;

inline V4 v4(float x, float y, float z, float w) { V4 r; r.x = x; r.y = y; r.z = z; r.w = w; return r; }
inline V4 scale4(V4 a, float s) { return v4(a.x * s, a.y * s, a.z * s, a.w * s); }
inline V4 add4(V4 a, V4 b) { return v4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline V4 sub4(V4 a, V4 b) { return v4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
inline V4 mul4(V4 a, V4 b) { return v4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
inline float& at(M4& m, int col, int row) { return (&m.c[col].x)[row]; }
inline float at(const M4& m, int col, int row) { return (&m.c[col].x)[row]; }

inline M4 identity() {
    M4 m;
    m.c[0] = v4(1, 0, 0, 0); m.c[1] = v4(0, 1, 0, 0); m.c[2] = v4(0, 0, 1, 0); m.c[3] = v4(0, 0, 0, 1);
    return m;
}
// column j of A*B = ((A0*b.x + A1*b.y) + A2*b.z) + A3*b.w
inline M4 matmul(const M4& a, const M4& b) {
    M4 r;
    for (int j = 0; j < 4; ++j) {
        const V4 bj = b.c[j];
        r.c[j] = add4(add4(add4(scale4(a.c[0], bj.x), scale4(a.c[1], bj.y)), scale4(a.c[2], bj.z)), scale4(a.c[3], bj.w));
    }
    return r;
}
// M * v = (c0*x + c1*y) + (c2*z + c3*w)
inline V4 matvec(const M4& m, V4 v) {
    return add4(add4(scale4(m.c[0], v.x), scale4(m.c[1], v.y)), add4(scale4(m.c[2], v.z), scale4(m.c[3], v.w)));
}
inline M4 transposed(const M4& m) {
    M4 r;
    for (int c = 0; c < 4; ++c)
        for (int rr = 0; rr < 4; ++rr) at(r, c, rr) = at(m, rr, c);
    return r;
}
// cofactor expansion in glm's grouping: 18 2x2 sub-determinants, six factor vectors, sign masks, 1/det last
inline M4 inverted(const M4& M) {
    auto e = [&](int c, int r) { return at(M, c, r); };
    const float c00 = e(2, 2) * e(3, 3) - e(3, 2) * e(2, 3), c02 = e(1, 2) * e(3, 3) - e(3, 2) * e(1, 3), c03 = e(1, 2) * e(2, 3) - e(2, 2) * e(1, 3);
    const float c04 = e(2, 1) * e(3, 3) - e(3, 1) * e(2, 3), c06 = e(1, 1) * e(3, 3) - e(3, 1) * e(1, 3), c07 = e(1, 1) * e(2, 3) - e(2, 1) * e(1, 3);
    const float c08 = e(2, 1) * e(3, 2) - e(3, 1) * e(2, 2), c10 = e(1, 1) * e(3, 2) - e(3, 1) * e(1, 2), c11 = e(1, 1) * e(2, 2) - e(2, 1) * e(1, 2);
    const float c12 = e(2, 0) * e(3, 3) - e(3, 0) * e(2, 3), c14 = e(1, 0) * e(3, 3) - e(3, 0) * e(1, 3), c15 = e(1, 0) * e(2, 3) - e(2, 0) * e(1, 3);
    const float c16 = e(2, 0) * e(3, 2) - e(3, 0) * e(2, 2), c18 = e(1, 0) * e(3, 2) - e(3, 0) * e(1, 2), c19 = e(1, 0) * e(2, 2) - e(2, 0) * e(1, 2);
    const float c20 = e(2, 0) * e(3, 1) - e(3, 0) * e(2, 1), c22 = e(1, 0) * e(3, 1) - e(3, 0) * e(1, 1), c23 = e(1, 0) * e(2, 1) - e(2, 0) * e(1, 1);
    const V4 f0 = v4(c00, c00, c02, c03), f1 = v4(c04, c04, c06, c07), f2 = v4(c08, c08, c10, c11);
    const V4 f3 = v4(c12, c12, c14, c15), f4 = v4(c16, c16, c18, c19), f5 = v4(c20, c20, c22, c23);
    const V4 a0 = v4(e(1, 0), e(0, 0), e(0, 0), e(0, 0)), a1 = v4(e(1, 1), e(0, 1), e(0, 1), e(0, 1));
    const V4 a2 = v4(e(1, 2), e(0, 2), e(0, 2), e(0, 2)), a3 = v4(e(1, 3), e(0, 3), e(0, 3), e(0, 3));
    const V4 i0 = add4(sub4(mul4(a1, f0), mul4(a2, f1)), mul4(a3, f2));
    const V4 i1 = add4(sub4(mul4(a0, f0), mul4(a2, f3)), mul4(a3, f4));
    const V4 i2 = add4(sub4(mul4(a0, f1), mul4(a1, f3)), mul4(a3, f5));
    const V4 i3 = add4(sub4(mul4(a0, f2), mul4(a1, f4)), mul4(a2, f5));
    const V4 sa = v4(+1, -1, +1, -1), sb = v4(-1, +1, -1, +1);
    M4 inv;
    inv.c[0] = mul4(i0, sa); inv.c[1] = mul4(i1, sb); inv.c[2] = mul4(i2, sa); inv.c[3] = mul4(i3, sb);
    const V4 row0 = v4(inv.c[0].x, inv.c[1].x, inv.c[2].x, inv.c[3].x);
    const V4 d0 = mul4(M.c[0], row0);
    const float det = (d0.x + d0.y) + (d0.z + d0.w);
    const float rdet = 1.0f / det;
    for (int c = 0; c < 4; ++c) inv.c[c] = scale4(inv.c[c], rdet);
    return inv;
}
inline float to_radians(float deg) { return deg * 0.01745329251994329576923690768489f; }
inline M4 translation(V3 t) {
    const M4 m = identity();
    M4 r = m;
    r.c[3] = add4(add4(add4(scale4(m.c[0], t.x), scale4(m.c[1], t.y)), scale4(m.c[2], t.z)), m.c[3]);
    return r;
}
inline M4 rotation(float angle, V3 axis_in) {
    const M4 m = identity();
    const float c = std::cos(angle), s = std::sin(angle);
    const V3 a = ffm::normalize(axis_in);
    const V3 t = ffm::mul(a, 1.0f - c);
    const float r00 = c + t.x * a.x, r01 = t.x * a.y + s * a.z, r02 = t.x * a.z - s * a.y;
    const float r10 = t.y * a.x - s * a.z, r11 = c + t.y * a.y, r12 = t.y * a.z + s * a.x;
    const float r20 = t.z * a.x + s * a.y, r21 = t.z * a.y - s * a.x, r22 = c + t.z * a.z;
    M4 r;
    r.c[0] = add4(add4(scale4(m.c[0], r00), scale4(m.c[1], r01)), scale4(m.c[2], r02));
    r.c[1] = add4(add4(scale4(m.c[0], r10), scale4(m.c[1], r11)), scale4(m.c[2], r12));
    r.c[2] = add4(add4(scale4(m.c[0], r20), scale4(m.c[1], r21)), scale4(m.c[2], r22));
    r.c[3] = m.c[3];
    return r;
}
inline M4 scaling(V3 s) {
    const M4 m = identity();
    M4 r;
    r.c[0] = scale4(m.c[0], s.x); r.c[1] = scale4(m.c[1], s.y); r.c[2] = scale4(m.c[2], s.z); r.c[3] = m.c[3];
    return r;
}
// Geometry ctor, utilities.h:179-189
inline void model_from_trs(const float pos[3], const float rot_deg[3], const float scl[3], M4& model, M4& inverse) {
    const M4 T = translation(ffm::v3(pos[0], pos[1], pos[2]));
    M4 R = rotation(to_radians(rot_deg[0]), ffm::v3(1, 0, 0));
    R = matmul(R, rotation(to_radians(rot_deg[1]), ffm::v3(0, 1, 0)));
    R = matmul(R, rotation(to_radians(rot_deg[2]), ffm::v3(0, 0, 1)));
    const M4 S = scaling(ffm::v3(scl[0], scl[1], scl[2]));
    model = matmul(matmul(T, R), S);
    inverse = inverted(model);
}
inline M4 look_at_rh(V3 eye, V3 center, V3 up) {
    const V3 f = ffm::normalize(ffm::sub(center, eye));
    const V3 s = ffm::normalize(ffm::cross(f, up));
    const V3 u = ffm::cross(s, f);
    M4 r = identity();
    r.c[0].x = s.x; r.c[1].x = s.y; r.c[2].x = s.z;
    r.c[0].y = u.x; r.c[1].y = u.y; r.c[2].y = u.z;
    r.c[0].z = -f.x; r.c[1].z = -f.y; r.c[2].z = -f.z;
    r.c[3].x = -ffm::dot(s, eye); r.c[3].y = -ffm::dot(u, eye); r.c[3].z = ffm::dot(f, eye);
    return r;
}
inline M4 perspective_fov_rh_no(float fov, float width, float height, float z_near, float z_far) {
    const float h = std::cos(0.5f * fov) / std::sin(0.5f * fov);
    const float w = h * height / width;
    M4 r;
    for (int c = 0; c < 4; ++c) r.c[c] = v4(0, 0, 0, 0);
    r.c[0].x = w;
    r.c[1].y = h;
    r.c[2].z = -(z_far + z_near) / (z_far - z_near);
    r.c[2].w = -1.0f;
    r.c[3].z = -(2.0f * z_far * z_near) / (z_far - z_near);
    return r;
}
// Camera::UpdateBasisAxis, utilities.h:407-418
inline void camera_basis(const float world_up[3], float yaw_deg, float pitch_deg, V3& forward, V3& right, V3& up) {
    V3 front;
    // The reference writes unqualified cos / sin on floats at global scope (utilities.h:411-413).  Re-hosted with g++ (the
    // pin: oracle/_ref), those are the C library's double functions: float arguments are widened, the PRODUCT is formed in
    // double and rounded to float once.  std::cos(float) * std::cos(float) differs from that in the last bit for about one
    // direction in five (tests/golden/ref_camera_walk.npz).
    const double ry = (double)to_radians(yaw_deg), rp = (double)to_radians(pitch_deg);
    front.x = (float)(::cos(ry) * ::cos(rp));
    front.y = (float)::sin(rp);
    front.z = (float)(::sin(ry) * ::cos(rp));
    forward = ffm::normalize(front);
    right = ffm::normalize(ffm::cross(forward, ffm::v3(world_up[0], world_up[1], world_up[2])));
    up = ffm::normalize(ffm::cross(right, forward));
}
}  // namespace ffh
