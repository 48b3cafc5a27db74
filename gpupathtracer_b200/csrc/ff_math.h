// ff_math.h -- fp32 arithmetic with ONE correctly-rounded IEEE operation per operator, identical on
// host (g++ -ffp-contract=off) and device (explicit *_rn intrinsics, so nvcc never contracts to FMA).
//
// The expression ORDER follows glm 0.9.9.7 as the reference uses it (dot: (x+y)+z, cross, normalize
// = v * (1/sqrt(dot)), distance = length(p1-p0); external/include/glm-0.9.9.7/detail/
// func_geometric.inl:8-21,48-88), because hit IDs must match the reference's intersectTriangle /
// intersectRays (kernel.cu:35-176) bit for bit.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define FF_HD __host__ __device__ __forceinline__
#else
#define FF_HD inline
#endif

namespace ffm {

#if defined(__CUDA_ARCH__)
FF_HD float fadd(float a, float b) { return __fadd_rn(a, b); }
FF_HD float fsub(float a, float b) { return __fsub_rn(a, b); }
FF_HD float fmul(float a, float b) { return __fmul_rn(a, b); }
FF_HD float fdiv(float a, float b) { return __fdiv_rn(a, b); }
FF_HD float fsqrt(float a) { return __fsqrt_rn(a); }
FF_HD float ffma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
#else
FF_HD float fadd(float a, float b) { return a + b; }
FF_HD float fsub(float a, float b) { return a - b; }
FF_HD float fmul(float a, float b) { return a * b; }
FF_HD float fdiv(float a, float b) { return a / b; }
FF_HD float fsqrt(float a) { return sqrtf(a); }
FF_HD float ffma(float a, float b, float c) { return fmaf(a, b, c); }
#endif

struct V3 { float x, y, z; };
struct V4 { float x, y, z, w; };
struct M4 { V4 c[4]; };  // column-major: c[col].row

FF_HD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
FF_HD V3 add(V3 a, V3 b) { return v3(fadd(a.x, b.x), fadd(a.y, b.y), fadd(a.z, b.z)); }
FF_HD V3 sub(V3 a, V3 b) { return v3(fsub(a.x, b.x), fsub(a.y, b.y), fsub(a.z, b.z)); }
FF_HD V3 mul(V3 a, float s) { return v3(fmul(a.x, s), fmul(a.y, s), fmul(a.z, s)); }
FF_HD V3 mul3(V3 a, V3 b) { return v3(fmul(a.x, b.x), fmul(a.y, b.y), fmul(a.z, b.z)); }
FF_HD V3 neg(V3 a) { return v3(-a.x, -a.y, -a.z); }
FF_HD float dot(V3 a, V3 b) { return fadd(fadd(fmul(a.x, b.x), fmul(a.y, b.y)), fmul(a.z, b.z)); }
FF_HD V3 cross(V3 x, V3 y) {
    return v3(fsub(fmul(x.y, y.z), fmul(y.y, x.z)), fsub(fmul(x.z, y.x), fmul(y.z, x.x)),
              fsub(fmul(x.x, y.y), fmul(y.x, x.y)));
}
FF_HD float inversesqrt(float x) { return fdiv(1.0f, fsqrt(x)); }
FF_HD V3 normalize(V3 v) { return mul(v, inversesqrt(dot(v, v))); }
FF_HD float length(V3 v) { return fsqrt(dot(v, v)); }
FF_HD float distance(V3 p0, V3 p1) { return length(sub(p1, p0)); }

// The reference's ray/triangle test (kernel.cu:35-108) on a ray whose direction is already
// normalised.  Returns true on an accepted hit and sets t (distance along d).  Comparison senses are
// kept exactly (NaNs fall through the same way).
FF_HD bool intersect_triangle(V3 v0, V3 v1, V3 v2, V3 o, V3 d, float& t_out) {
    // Written without early returns (on the GPU divergent exits cost more than the arithmetic they skip); the
    // accept condition is the conjunction of the reference's tests with their comparison senses unchanged.
    const float EPSILON = 0.000001f;
    V3 edge1 = sub(v1, v0), edge2 = sub(v2, v0);
    V3 N = cross(edge1, edge2);
    const bool back = dot(d, N) > 0.0f;              // kernel.cu:48-51
    V3 pvec = cross(d, edge2);
    float det = dot(edge1, pvec);
    const bool flat = det < EPSILON;                 // kernel.cu:57
    V3 tvec = sub(o, v0);
    float u = dot(tvec, pvec);
    const bool out_u = u < 0.0f || u > det;          // kernel.cu:64
    V3 qvec = cross(tvec, edge1);
    float v = dot(d, qvec);
    const bool out_v = v < 0.0f || fadd(u, v) > det;  // kernel.cu:71
    float t = dot(edge2, qvec);
    float invDet = fdiv(1.0f, det);  // kernel.cu:77 divides in double; rounds to the same float
    t = fmul(t, invDet);
    t_out = t;
    return !back && !flat && !out_u && !out_v && t > EPSILON;  // kernel.cu:97
}
// The barycentrics intersectTriangle computes and then drops (u *= invDet; v *= invDet, kernel.cu:80-81), same operations
// in the same order as above; (u, v) weight v1 and v2, 1 - u - v weights v0.
FF_HD void triangle_barycentrics(V3 v0, V3 v1, V3 v2, V3 o, V3 d, float& u_out, float& v_out) {
    V3 edge1 = sub(v1, v0), edge2 = sub(v2, v0);
    V3 pvec = cross(d, edge2);
    float det = dot(edge1, pvec);
    V3 tvec = sub(o, v0);
    float u = dot(tvec, pvec);
    V3 qvec = cross(tvec, edge1);
    float v = dot(d, qvec);
    float invDet = fdiv(1.0f, det);
    u_out = fmul(u, invDet);
    v_out = fmul(v, invDet);
}
// Closest-hit key of setIntersection (kernel.cu:113-115) for a baked (M = I) triangle.
FF_HD float hit_key(V3 o, V3 d, float t, V3& point) {
    point = add(o, mul(d, t));
    return distance(point, o);
}

// ------------------------------------------------------------------ Philox4x32-10 (Salmon et al. 2011)
struct U4 { uint32_t x, y, z, w; };
FF_HD U4 philox4x32_10(U4 ctr, uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int r = 0; r < 10; ++r) {
#if defined(__CUDA_ARCH__)
        uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
        uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
#else
        uint64_t p0 = (uint64_t)M0 * ctr.x, p1 = (uint64_t)M1 * ctr.z;
        uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0, hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
#endif
        U4 n;
        n.x = hi1 ^ ctr.y ^ k0;
        n.y = lo1;
        n.z = hi0 ^ ctr.w ^ k1;
        n.w = lo0;
        ctr = n;
        k0 += W0;
        k1 += W1;
    }
    return ctr;
}
FF_HD float u01(uint32_t u) { return fmul((float)(u >> 8), 5.9604644775390625e-08f); }  // [0,1)

// sin/cos(2 pi u), u in [0,1): quadrant reduction + fixed polynomials in correctly-rounded fma, so
// host and device agree bit for bit (library sin/cos do not).  DESIGN.md section 5.
FF_HD void sincos2pi(float u, float& s, float& c) {
    float x = fmul(u, 4.0f);
    float q = rintf(x);  // round-half-even
    float r = fsub(x, q);
    float r2 = fmul(r, r);
    const float S0 = 1.57079632679489662f, S1 = -0.645964097506246254f, S2 = 0.0796926262461670451f,
                S3 = -0.00468175413531868810f, S4 = 0.000160441184787359822f;
    const float C0 = 1.0f, C1 = -1.23370055013616983f, C2 = 0.253669507901048014f, C3 = -0.0208634807633529609f,
                C4 = 0.000919260274839426580f;
    float sp = ffma(S4, r2, S3); sp = ffma(sp, r2, S2); sp = ffma(sp, r2, S1); sp = ffma(sp, r2, S0);
    float sr = fmul(sp, r);
    float cp = ffma(C4, r2, C3); cp = ffma(cp, r2, C2); cp = ffma(cp, r2, C1); cp = ffma(cp, r2, C0);
    int qi = (int)q & 3;
    if (qi == 0) { s = sr; c = cp; }
    else if (qi == 1) { s = cp; c = -sr; }
    else if (qi == 2) { s = -sr; c = -cp; }
    else { s = -cp; c = sr; }
}
// CosineSampleHemisphere (utilities.h:46-55) with sin/cos(2 pi u2) from sincos2pi.
FF_HD V3 cosine_sample_hemisphere(float u1, float u2) {
    float r = fsqrt(u1);
    float s, c;
    sincos2pi(u2, s, c);
    float z = fsqrt(fmaxf(0.0f, fsub(1.0f, u1)));
    return v3(fmul(r, c), fmul(r, s), z);
}
// Orthonormal basis around a unit normal (Duff et al. 2017), one rounded op per operator.
FF_HD void onb(V3 n, V3& t, V3& b) {
    float sign = copysignf(1.0f, n.z);
    float a = fdiv(-1.0f, fadd(sign, n.z));
    float bb = fmul(fmul(n.x, n.y), a);
    t = v3(fadd(1.0f, fmul(sign, fmul(fmul(n.x, n.x), a))), fmul(sign, bb), fmul(-sign, n.x));
    b = v3(bb, fadd(sign, fmul(fmul(n.y, n.y), a)), -n.y);
}

}  // namespace ffm
