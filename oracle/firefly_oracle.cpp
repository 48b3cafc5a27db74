// TEST INFRASTRUCTURE ONLY -- CPU restatement ("port") of the reference's hot path.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library; the product (gpupathtracer_b200/) never does and has no CPU fallback.
//
// What is restated, and from where (paths relative to /root/reference):
//   * intersectPlane / intersectTriangle / setIntersection / intersectRays
//       PathTracer/FireflyEngine/kernel.cu:8-32, 35-108, 110-125, 127-176
//   * camera ray generation and Camera::UpdateBasisAxis / Get*Matrix
//       kernel.cu:186-205, PathTracer/FireflyEngine/utilities.h:299-317, 407-418
//   * Geometry ctor (model = T * Rx * Ry * Rz * S, inverse)        utilities.h:176-189
//   * shade as coded (abs(normal) -> uchar3)                        kernel.cu:178-184, 207-215
//   * CosineSampleHemisphere / diffuse + emitter BXDF intent        utilities.h:46-55, 90-139
//   * glm 0.9.9.7 expression order (vendored by the reference under external/include/glm-0.9.9.7):
//       dot/cross/normalize/length/distance  detail/func_geometric.inl:8-21,48-88
//       mat4*vec4, mat4*mat4                 detail/type_mat4x4.inl:536-575,630-648
//       inverse(mat4), transpose             detail/func_matrix.inl:170-196,294-352
//       translate/rotate/scale/lookAtRH      ext/matrix_transform.inl:10-119
//       perspectiveFovRH_NO                  ext/matrix_clip_space.inl:372-389
//       radians                              detail/func_trigonometric.inl:9-14
//
// Parity status: PINNED.  This restatement is checked bit-for-bit against the reference's own text
// compiled from /root/reference (oracle/_ref, see oracle/ref_shim.cpp) by tests/test_oracle_golden.py (the *_vs_ref tests)
// when that library is present, and against the golden vectors under tests/golden/ that were
// generated from it (tools/make_golden.py).  The reference itself ships no tests or vectors.
//
// Everything here is plain scalar fp32 with one IEEE operation per source operator; build with
// -O2 -ffp-contract=off -fno-fast-math so the compiler never fuses or re-associates.
//
// New text with no reference counterpart (the reference has no bounce loop, no usable RNG, no
// acceleration structure): Philox4x32-10, sincos2pi, the ONB, and the "as-intended" path tracer
// (SURVEY.md section 8 row a7).  Their exact operation sequence is the specification that the CUDA
// integrators are held to; it is written down in DESIGN.md section 5.
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

namespace {

// ---------------------------------------------------------------- glm-order vector/matrix maths
struct V3 { float x, y, z; };
struct V4 { float x, y, z, w; };
struct M4 { V4 c[4]; };  // column-major like glm: c[col].row

inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
inline V3 add(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 sub(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 mul(V3 a, float s) { return V3{a.x * s, a.y * s, a.z * s}; }
inline V3 neg(V3 a) { return V3{-a.x, -a.y, -a.z}; }
// func_geometric.inl:48-53  tmp = a*b; tmp.x + tmp.y + tmp.z
inline float dot(V3 a, V3 b) { float tx = a.x * b.x, ty = a.y * b.y, tz = a.z * b.z; return (tx + ty) + tz; }
// func_geometric.inl:68-78
inline V3 cross(V3 x, V3 y) {
    return V3{x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y};
}
inline float inversesqrt(float x) { return 1.0f / std::sqrt(x); }  // func_exponential.inl:136-139
inline V3 normalize(V3 v) { return mul(v, inversesqrt(dot(v, v))); }  // func_geometric.inl:82-88
inline float length(V3 v) { return std::sqrt(dot(v, v)); }
inline float distance(V3 p0, V3 p1) { return length(sub(p1, p0)); }  // func_geometric.inl:16-21

inline V4 add(V4 a, V4 b) { return V4{a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
inline V4 sub(V4 a, V4 b) { return V4{a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w}; }
inline V4 mul(V4 a, V4 b) { return V4{a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w}; }
inline V4 mul(V4 a, float s) { return V4{a.x * s, a.y * s, a.z * s, a.w * s}; }
inline V4 splat(float s) { return V4{s, s, s, s}; }
inline V4 v4(V3 v, float w) { return V4{v.x, v.y, v.z, w}; }
inline V3 xyz(V4 v) { return V3{v.x, v.y, v.z}; }
inline float dot4(V4 a, V4 b) { V4 t = mul(a, b); return (t.x + t.y) + (t.z + t.w); }
inline V4 normalize4(V4 v) { return mul(v, inversesqrt(dot4(v, v))); }
inline float& el(V4& v, int i) { return (&v.x)[i]; }
inline float el(const V4& v, int i) { return (&v.x)[i]; }

inline M4 identity() {
    M4 m;
    m.c[0] = V4{1, 0, 0, 0}; m.c[1] = V4{0, 1, 0, 0}; m.c[2] = V4{0, 0, 1, 0}; m.c[3] = V4{0, 0, 0, 1};
    return m;
}
// type_mat4x4.inl:536-575
inline V4 mulMV(const M4& m, V4 v) {
    V4 Mul0 = mul(m.c[0], splat(v.x));
    V4 Mul1 = mul(m.c[1], splat(v.y));
    V4 Add0 = add(Mul0, Mul1);
    V4 Mul2 = mul(m.c[2], splat(v.z));
    V4 Mul3 = mul(m.c[3], splat(v.w));
    V4 Add1 = add(Mul2, Mul3);
    return add(Add0, Add1);
}
// type_mat4x4.inl:630-648
inline M4 mulMM(const M4& a, const M4& b) {
    M4 r;
    for (int j = 0; j < 4; ++j)
        r.c[j] = add(add(add(mul(a.c[0], b.c[j].x), mul(a.c[1], b.c[j].y)), mul(a.c[2], b.c[j].z)), mul(a.c[3], b.c[j].w));
    return r;
}
inline M4 transpose(const M4& m) {
    M4 r;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) el(r.c[i], j) = el(m.c[j], i);
    return r;
}
// func_matrix.inl:294-352
M4 inverse(const M4& M) {
#define m(i, j) el(M.c[i], j)
    float Coef00 = m(2, 2) * m(3, 3) - m(3, 2) * m(2, 3);
    float Coef02 = m(1, 2) * m(3, 3) - m(3, 2) * m(1, 3);
    float Coef03 = m(1, 2) * m(2, 3) - m(2, 2) * m(1, 3);
    float Coef04 = m(2, 1) * m(3, 3) - m(3, 1) * m(2, 3);
    float Coef06 = m(1, 1) * m(3, 3) - m(3, 1) * m(1, 3);
    float Coef07 = m(1, 1) * m(2, 3) - m(2, 1) * m(1, 3);
    float Coef08 = m(2, 1) * m(3, 2) - m(3, 1) * m(2, 2);
    float Coef10 = m(1, 1) * m(3, 2) - m(3, 1) * m(1, 2);
    float Coef11 = m(1, 1) * m(2, 2) - m(2, 1) * m(1, 2);
    float Coef12 = m(2, 0) * m(3, 3) - m(3, 0) * m(2, 3);
    float Coef14 = m(1, 0) * m(3, 3) - m(3, 0) * m(1, 3);
    float Coef15 = m(1, 0) * m(2, 3) - m(2, 0) * m(1, 3);
    float Coef16 = m(2, 0) * m(3, 2) - m(3, 0) * m(2, 2);
    float Coef18 = m(1, 0) * m(3, 2) - m(3, 0) * m(1, 2);
    float Coef19 = m(1, 0) * m(2, 2) - m(2, 0) * m(1, 2);
    float Coef20 = m(2, 0) * m(3, 1) - m(3, 0) * m(2, 1);
    float Coef22 = m(1, 0) * m(3, 1) - m(3, 0) * m(1, 1);
    float Coef23 = m(1, 0) * m(2, 1) - m(2, 0) * m(1, 1);
    V4 Fac0{Coef00, Coef00, Coef02, Coef03}, Fac1{Coef04, Coef04, Coef06, Coef07}, Fac2{Coef08, Coef08, Coef10, Coef11};
    V4 Fac3{Coef12, Coef12, Coef14, Coef15}, Fac4{Coef16, Coef16, Coef18, Coef19}, Fac5{Coef20, Coef20, Coef22, Coef23};
    V4 Vec0{m(1, 0), m(0, 0), m(0, 0), m(0, 0)}, Vec1{m(1, 1), m(0, 1), m(0, 1), m(0, 1)};
    V4 Vec2{m(1, 2), m(0, 2), m(0, 2), m(0, 2)}, Vec3{m(1, 3), m(0, 3), m(0, 3), m(0, 3)};
    V4 Inv0 = add(sub(mul(Vec1, Fac0), mul(Vec2, Fac1)), mul(Vec3, Fac2));
    V4 Inv1 = add(sub(mul(Vec0, Fac0), mul(Vec2, Fac3)), mul(Vec3, Fac4));
    V4 Inv2 = add(sub(mul(Vec0, Fac1), mul(Vec1, Fac3)), mul(Vec3, Fac5));
    V4 Inv3 = add(sub(mul(Vec0, Fac2), mul(Vec1, Fac4)), mul(Vec2, Fac5));
    V4 SignA{+1, -1, +1, -1}, SignB{-1, +1, -1, +1};
    M4 Inv;
    Inv.c[0] = mul(Inv0, SignA); Inv.c[1] = mul(Inv1, SignB); Inv.c[2] = mul(Inv2, SignA); Inv.c[3] = mul(Inv3, SignB);
    V4 Row0{Inv.c[0].x, Inv.c[1].x, Inv.c[2].x, Inv.c[3].x};
    V4 Dot0 = mul(M.c[0], Row0);
    float Dot1 = (Dot0.x + Dot0.y) + (Dot0.z + Dot0.w);
    float OneOverDeterminant = 1.0f / Dot1;
    for (int i = 0; i < 4; ++i) Inv.c[i] = mul(Inv.c[i], OneOverDeterminant);
    return Inv;
#undef m
}
inline float radians(float deg) { return deg * 0.01745329251994329576923690768489f; }
// ext/matrix_transform.inl:10-15
inline M4 translate(const M4& m, V3 v) {
    M4 r = m;
    r.c[3] = add(add(add(mul(m.c[0], v.x), mul(m.c[1], v.y)), mul(m.c[2], v.z)), m.c[3]);
    return r;
}
// ext/matrix_transform.inl:18-46
M4 rotate(const M4& m, float angle, V3 v) {
    float c = std::cos(angle), s = std::sin(angle);
    V3 axis = normalize(v);
    V3 temp = mul(axis, (1.0f - c));  // (T(1)-c) * axis : scalar*vec multiplies each component by the scalar
    float R00 = c + temp.x * axis.x, R01 = temp.x * axis.y + s * axis.z, R02 = temp.x * axis.z - s * axis.y;
    float R10 = temp.y * axis.x - s * axis.z, R11 = c + temp.y * axis.y, R12 = temp.y * axis.z + s * axis.x;
    float R20 = temp.z * axis.x + s * axis.y, R21 = temp.z * axis.y - s * axis.x, R22 = c + temp.z * axis.z;
    M4 r;
    r.c[0] = add(add(mul(m.c[0], R00), mul(m.c[1], R01)), mul(m.c[2], R02));
    r.c[1] = add(add(mul(m.c[0], R10), mul(m.c[1], R11)), mul(m.c[2], R12));
    r.c[2] = add(add(mul(m.c[0], R20), mul(m.c[1], R21)), mul(m.c[2], R22));
    r.c[3] = m.c[3];
    return r;
}
inline M4 scale(const M4& m, V3 v) {
    M4 r;
    r.c[0] = mul(m.c[0], v.x); r.c[1] = mul(m.c[1], v.y); r.c[2] = mul(m.c[2], v.z); r.c[3] = m.c[3];
    return r;
}
// utilities.h:179-189
void model_matrices(V3 pos, V3 rotDeg, V3 scl, M4& model, M4& inv) {
    M4 T = translate(identity(), pos);
    M4 R = rotate(identity(), radians(rotDeg.x), v3(1, 0, 0));
    R = mulMM(R, rotate(identity(), radians(rotDeg.y), v3(0, 1, 0)));
    R = mulMM(R, rotate(identity(), radians(rotDeg.z), v3(0, 0, 1)));
    M4 S = scale(identity(), scl);
    model = mulMM(mulMM(T, R), S);
    inv = inverse(model);
}
// ext/matrix_transform.inl:99-119
M4 lookAtRH(V3 eye, V3 center, V3 up) {
    V3 f = normalize(sub(center, eye));
    V3 s = normalize(cross(f, up));
    V3 u = cross(s, f);
    M4 r = identity();
    r.c[0].x = s.x; r.c[1].x = s.y; r.c[2].x = s.z;
    r.c[0].y = u.x; r.c[1].y = u.y; r.c[2].y = u.z;
    r.c[0].z = -f.x; r.c[1].z = -f.y; r.c[2].z = -f.z;
    r.c[3].x = -dot(s, eye); r.c[3].y = -dot(u, eye); r.c[3].z = dot(f, eye);
    return r;
}
// ext/matrix_clip_space.inl:372-389
M4 perspectiveFovRH_NO(float fov, float width, float height, float zNear, float zFar) {
    float rad = fov;
    float h = std::cos(0.5f * rad) / std::sin(0.5f * rad);
    float w = h * height / width;
    M4 r;
    for (int i = 0; i < 4; ++i) r.c[i] = V4{0, 0, 0, 0};
    r.c[0].x = w;
    r.c[1].y = h;
    r.c[2].z = -(zFar + zNear) / (zFar - zNear);
    r.c[2].w = -1.0f;
    r.c[3].z = -(2.0f * zFar * zNear) / (zFar - zNear);
    return r;
}

// ---------------------------------------------------------------- reference data model
struct Tri { V3 v0, v1, v2; };
struct Hit {  // utilities.h:57-66
    V3 point{0, 0, 0}, normal{0, 0, 0};
    float t = 0.f;
    bool hit = false;
    int geom = -1, tri = -1;
};
enum GType { SPHERE = 0, PLANE = 1, TRIANGLEMESH = 2 };  // utilities.h:141-146
// BXDF utilities.h:77-88; type = BXDFTyp utilities.h:68-75 (0 EMITTER, 1 DIFFUSE, 2 MIRROR, 3 GLASS)
struct Material {
    int type; V3 albedo, emissive; float intensity;
    V3 specular{1, 1, 1}; float ior = 1.5f; V3 transmittance{1, 1, 1};   // m_specularColor, m_refractiveIndex, m_transmittanceColor
};
struct Geom {
    int type;
    M4 model, inv;
    V3 normal{0, 0, 1};  // utilities.h:230
    float radius = 1.0f;  // Geometry::m_sphereRadius (utilities.h:227), used by intersectSphere
    std::vector<Tri> tris;
    // shading attributes the reference's loader keeps and its tracer never reads; empty when absent
    std::vector<V3> normals;          // 3 per triangle, object space (Triangle::m_n0..2, utilities.h:168-170)
    std::vector<float> uvs;           // 6 per triangle (Triangle::m_uv0..2, utilities.h:163-166)
    std::vector<int> face_mat;        // per triangle: index into face_mats, -1 = the geometry's material (utilities.h:837)
    std::vector<Material> face_mats;
};
struct Scene {
    std::vector<Geom> geoms;
    std::vector<Material> mats;
};
struct Ray { V3 o, d; };

// kernel.cu:8-32
bool intersectPlane(const Geom& plane, const Ray& ray, Hit& out) {
    float denom = dot(plane.normal, ray.d);
    if ((double)std::fabs(denom) > 1e-7) {  // float compared against the double literal 1e-7
        V3 p0l0 = neg(ray.o);
        float t = dot(p0l0, plane.normal) / denom;
        V3 P = add(ray.o, mul(ray.d, t));
        if (!(P.x >= -0.5f && P.x <= 0.5f && P.y >= -0.5f && P.y <= 0.5f)) return false;
        if (t > 0.0f) { out.t = t; out.point = P; out.normal = plane.normal; return true; }
        return false;
    }
    return false;
}
// GeometryType::SPHERE: the reference declares it (utilities.h:143, m_sphereRadius utilities.h:227) and leaves the
// branch empty (kernel.cu:166-169).  New text in the style of intersectPlane: object-space sphere of radius r at the
// origin, ENTERING root only (front faces only, like intersectTriangle's culling: a ray that starts inside sees
// nothing), accepted when t > 0; the normal returned is the object-space point (not normalised, like setIntersection
// leaves normals).
bool intersectSphere(const Geom& sph, const Ray& ray, Hit& out) {
    float b = dot(ray.o, ray.d);
    float c = dot(ray.o, ray.o) - sph.radius * sph.radius;
    float disc = b * b - c;
    if (!(disc > 0.0f)) return false;
    float t = -b - std::sqrt(disc);
    if (t > 0.0f) { out.t = t; out.point = add(ray.o, mul(ray.d, t)); out.normal = out.point; return true; }
    return false;
}
// kernel.cu:35-108
bool intersectTriangle(const Tri& tri, const Ray& ray, Hit& out) {
    const float EPSILON = 0.000001;
    V3 edge1 = sub(tri.v1, tri.v0), edge2 = sub(tri.v2, tri.v0);
    V3 Normal = cross(edge1, edge2);
    if (dot(ray.d, Normal) > 0) return false;
    V3 pvec = cross(ray.d, edge2);
    float det = dot(edge1, pvec);
    if (det < EPSILON) return false;
    V3 tvec = sub(ray.o, tri.v0);
    float u = dot(tvec, pvec);
    if (u < 0.0f || u > det) return false;
    V3 qvec = cross(tvec, edge1);
    float v = dot(ray.d, qvec);
    if (v < 0.0f || u + v > det) return false;
    float t = dot(edge2, qvec);
    float invDet = (float)(1.0 / (double)det);  // kernel.cu:77: double divide rounded to float
    t *= invDet;
    if (t > EPSILON) {
        out.point = add(ray.o, mul(ray.d, t));
        out.t = t;
        out.normal = normalize(cross(edge1, edge2));
        return true;
    }
    return false;
}
// kernel.cu:110-125
bool setIntersection(float& tMax, Hit& out, const Hit& local, const M4& model, const Ray& ray) {
    V3 worldPOI = xyz(mulMV(model, v4(local.point, 1.0f)));
    float d = distance(worldPOI, ray.o);
    if (d < tMax) {
        out.normal = xyz(mulMV(inverse(transpose(model)), v4(local.normal, 0.f)));
        out.point = worldPOI;
        out.t = d;
        out.hit = true;
        tMax = d;
        return true;
    }
    return false;
}
// kernel.cu:127-176
Hit intersectRays(const Ray& ray, const Scene& sc) {
    Hit out;
    float tMax = INFINITY;
    for (int i = 0; i < (int)sc.geoms.size(); ++i) {
        const Geom& g = sc.geoms[i];
        Ray os{xyz(mulMV(g.inv, v4(ray.o, 1.f))), xyz(normalize4(mulMV(g.inv, v4(ray.d, 0.f))))};
        Hit local;
        if (g.type == TRIANGLEMESH) {
            for (int j = 0; j < (int)g.tris.size(); ++j)
                if (intersectTriangle(g.tris[j], os, local))
                    if (setIntersection(tMax, out, local, g.model, ray)) { out.geom = i; out.tri = j; }
        } else if (g.type == PLANE) {
            if (intersectPlane(g, os, local))
                if (setIntersection(tMax, out, local, g.model, ray)) { out.geom = i; out.tri = -1; }
        } else if (g.type == SPHERE) {  // an empty branch in the reference (kernel.cu:166-169); see intersectSphere
            if (intersectSphere(g, os, local))
                if (setIntersection(tMax, out, local, g.model, ray)) { out.geom = i; out.tri = -1; }
        }
    }
    return out;
}

// ---------------------------------------------------------------- camera (utilities.h:269-317,407-418)
struct Camera {
    V3 pos, up, right, forward, worldUp;
    float yaw, pitch, width, height, fov, nearClip, farClip;
    void updateBasis() {
        V3 front;
        // utilities.h:411-413 call the global (double) cos / sin on floats: products in double, one rounding to float
        const double ry = (double)radians(yaw), rp = (double)radians(pitch);
        front.x = (float)(::cos(ry) * ::cos(rp));
        front.y = (float)::sin(rp);
        front.z = (float)(::sin(ry) * ::cos(rp));
        forward = normalize(front);
        right = normalize(cross(forward, worldUp));
        up = normalize(cross(right, forward));
    }
    M4 invView() const { return inverse(lookAtRH(pos, add(pos, forward), up)); }
    M4 invProj() const { return inverse(perspectiveFovRH_NO(radians(fov), width, height, nearClip, farClip)); }
};
Camera make_camera(const float pos[3], const float wup[3], float yaw, float pitch, float fov, float w, float h,
                   float nearc, float farc) {
    Camera c;
    c.pos = v3(pos[0], pos[1], pos[2]); c.worldUp = v3(wup[0], wup[1], wup[2]);
    c.yaw = yaw; c.pitch = pitch; c.fov = fov; c.width = w; c.height = h; c.nearClip = nearc; c.farClip = farc;
    c.updateBasis();
    return c;
}
// kernel.cu:198-205 with pixel coordinates as floats (px = x, or x + jitter in AA mode)
inline Ray gen_ray(const Camera& cam, const M4& invVP, float px, float py) {
    Ray r;
    r.o = cam.pos;
    float Px = (px / cam.width) * 2.f - 1.f;
    float Py = 1.f - (py / cam.height) * 2.f;
    V4 ndc = mul(V4{Px, Py, 1.f, 1.f}, cam.farClip);
    V3 look = xyz(mulMV(invVP, ndc));
    r.d = normalize(sub(look, r.o));
    return r;
}

// ---------------------------------------------------------------- Philox4x32-10 (Salmon et al. 2011)
// constants as in /usr/local/cuda/include/curand_philox4x32_x.h:88-91
struct U4 { uint32_t x, y, z, w; };
inline U4 philox4x32_10(U4 ctr, uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)M0 * ctr.x, p1 = (uint64_t)M1 * ctr.z;
        U4 n;
        n.x = (uint32_t)(p1 >> 32) ^ ctr.y ^ k0;
        n.y = (uint32_t)p1;
        n.z = (uint32_t)(p0 >> 32) ^ ctr.w ^ k1;
        n.w = (uint32_t)p0;
        ctr = n;
        k0 += W0; k1 += W1;
    }
    return ctr;
}
inline float u01(uint32_t u) { return (float)(u >> 8) * 5.9604644775390625e-08f; }  // (u>>8) * 2^-24, in [0,1)

// ---------------------------------------------------------------- exact-op helpers shared by spec
// sincos of 2*pi*u for u in [0,1): quadrant reduction + fixed odd/even polynomials evaluated with
// correctly rounded fmaf, so CPU and GPU (__fmaf_rn) agree bit for bit.  DESIGN.md section 5.
inline void sincos2pi(float u, float& s, float& c) {
    float x = u * 4.0f;            // exact
    float q = std::nearbyintf(x);  // round-half-even, 0..4
    float r = x - q;               // exact, in [-0.5, 0.5]; angle = r * pi/2
    float r2 = r * r;
    // sin(pi/2 r) = r * (S0 + S1 r^2 + S2 r^4 + S3 r^6), cos(pi/2 r) = C0 + C1 r^2 + C2 r^4 + C3 r^6 + C4 r^8
    const float S0 = 1.57079632679489662f, S1 = -0.645964097506246254f, S2 = 0.0796926262461670451f,
                S3 = -0.00468175413531868810f, S4 = 0.000160441184787359822f;
    const float C0 = 1.0f, C1 = -1.23370055013616983f, C2 = 0.253669507901048014f, C3 = -0.0208634807633529609f,
                C4 = 0.000919260274839426580f;
    float sp = std::fmaf(S4, r2, S3); sp = std::fmaf(sp, r2, S2); sp = std::fmaf(sp, r2, S1); sp = std::fmaf(sp, r2, S0);
    float sr = sp * r;
    float cp = std::fmaf(C4, r2, C3); cp = std::fmaf(cp, r2, C2); cp = std::fmaf(cp, r2, C1); cp = std::fmaf(cp, r2, C0);
    int qi = (int)q & 3;
    switch (qi) {
        case 0: s = sr; c = cp; break;
        case 1: s = cp; c = -sr; break;
        case 2: s = -sr; c = -cp; break;
        default: s = -cp; c = sr; break;
    }
}
// utilities.h:46-55 with cos/sin(2 pi u2) replaced by sincos2pi (library sin/cos differ CPU vs GPU)
inline V3 cosineSampleHemisphere(float u1, float u2) {
    float r = std::sqrt(u1);
    float s, c;
    sincos2pi(u2, s, c);
    float z = std::sqrt(std::fmax(0.0f, 1.0f - u1));
    return V3{r * c, r * s, z};
}
// branch-light orthonormal basis (Duff et al. 2017), one IEEE op per operator
inline void onb(V3 n, V3& t, V3& b) {
    float sign = std::copysign(1.0f, n.z);
    float a = -1.0f / (sign + n.z);
    float bb = (n.x * n.y) * a;
    t = V3{1.0f + sign * ((n.x * n.x) * a), sign * bb, -sign * n.x};
    b = V3{bb, sign + (n.y * n.y) * a, -n.y};
}

// ---------------------------------------------------------------- flattened ("baked") scene
// World-space triangles: v_w = vec3(M * vec4(v_obj, 1)), identity transform afterwards.
struct Flat {
    std::vector<Tri> tris;
    std::vector<int> geom, tri;  // source numbering (geometryIndex, triangleIndex)
    std::vector<V3> wnormal;     // 3 per flat triangle: inverse(transpose(M)) * (n, 0), not normalised; zero when the geometry has none
};
void flatten(const Scene& sc, Flat& f) {
    for (int i = 0; i < (int)sc.geoms.size(); ++i) {
        const Geom& g = sc.geoms[i];
        if (g.type != TRIANGLEMESH) continue;
        for (int j = 0; j < (int)g.tris.size(); ++j) {
            Tri w;
            w.v0 = xyz(mulMV(g.model, v4(g.tris[j].v0, 1.f)));
            w.v1 = xyz(mulMV(g.model, v4(g.tris[j].v1, 1.f)));
            w.v2 = xyz(mulMV(g.model, v4(g.tris[j].v2, 1.f)));
            f.tris.push_back(w); f.geom.push_back(i); f.tri.push_back(j);
            for (int c = 0; c < 3; ++c) {
                V3 n{0, 0, 0};
                if (!g.normals.empty()) n = xyz(mulMV(inverse(transpose(g.model)), v4(g.normals[3 * (size_t)j + c], 0.f)));
                f.wnormal.push_back(n);
            }
        }
    }
}
// Material of triangle `tri` of geometry `geom` (tri < 0: analytic primitive).
inline const Material& material_of(const Scene& sc, int geom, int tri) {
    const Geom& g = sc.geoms[geom];
    if (tri >= 0 && !g.face_mat.empty() && g.face_mat[tri] >= 0) return g.face_mats[g.face_mat[tri]];
    return sc.mats[geom];
}
// The barycentrics intersectTriangle computes and drops (kernel.cu:53-81): u weights v1, v weights v2.
inline void barycentrics(const Tri& tri, const Ray& ray, float& u, float& v) {
    V3 edge1 = sub(tri.v1, tri.v0), edge2 = sub(tri.v2, tri.v0);
    V3 pvec = cross(ray.d, edge2);
    float det = dot(edge1, pvec);
    V3 tvec = sub(ray.o, tri.v0);
    u = dot(tvec, pvec);
    V3 qvec = cross(tvec, edge1);
    v = dot(ray.d, qvec);
    float invDet = (float)(1.0 / (double)det);
    u *= invDet;
    v *= invDet;
}
// Shading attributes of a triangle hit; GPU twin: surface_attributes in ff_render.cu.
struct SurfAttr { bool smooth = false; V3 ns{0, 0, 0}; float u = 0, v = 0, tu = 0, tv = 0; };
inline SurfAttr surface_attributes(const Scene& sc, const Flat& f, int flat, const Ray& ray) {
    SurfAttr a;
    if (flat < 0) return a;
    const Geom& g = sc.geoms[f.geom[flat]];
    Ray os{ray.o, normalize(ray.d)};
    barycentrics(f.tris[flat], os, a.u, a.v);
    float w0 = (1.0f - a.u) - a.v;
    if (!g.normals.empty()) {
        const V3* n = &f.wnormal[3 * (size_t)flat];
        V3 ns = add(add(mul(n[0], w0), mul(n[1], a.u)), mul(n[2], a.v));
        if (dot(ns, ns) > 0.0f) { a.smooth = true; a.ns = normalize(ns); }
    }
    if (!g.uvs.empty()) {
        const float* t = &g.uvs[6 * (size_t)f.tri[flat]];
        a.tu = (t[0] * w0 + t[2] * a.u) + t[4] * a.v;
        a.tv = (t[1] * w0 + t[3] * a.u) + t[5] * a.v;
    }
    return a;
}
// intersectRays on the baked scene: every geometry has M = I, so the object-space ray is
// (o, normalize(d)) and the closest-hit key is distance(o + d'*t, o)  (kernel.cu:113-115,138).
// Planes keep their own transform.  Iteration order = (geometry, triangle) order, strict <.
struct FlatHit { bool hit = false; int geom = -1, tri = -1, flat = -1; float t = INFINITY; V3 point{0,0,0}, normal{0,0,0}; };
FlatHit intersectFlat(const Ray& ray, const Scene& sc, const Flat& f) {
    FlatHit out;
    float tMax = INFINITY;
    Ray os{ray.o, normalize(ray.d)};
    size_t k = 0;
    for (int i = 0; i < (int)sc.geoms.size(); ++i) {
        const Geom& g = sc.geoms[i];
        if (g.type == TRIANGLEMESH) {
            for (int j = 0; j < (int)g.tris.size(); ++j, ++k) {
                Hit local;
                if (intersectTriangle(f.tris[k], os, local)) {
                    float d = distance(local.point, ray.o);
                    if (d < tMax) {
                        tMax = d; out.hit = true; out.geom = i; out.tri = j; out.flat = (int)k; out.t = d;
                        out.point = local.point; out.normal = local.normal;
                    }
                }
            }
        } else if (g.type == PLANE) {
            Ray ps{xyz(mulMV(g.inv, v4(ray.o, 1.f))), xyz(normalize4(mulMV(g.inv, v4(ray.d, 0.f))))};
            Hit local;
            if (intersectPlane(g, ps, local)) {
                V3 worldPOI = xyz(mulMV(g.model, v4(local.point, 1.0f)));
                float d = distance(worldPOI, ray.o);
                if (d < tMax) {
                    tMax = d; out.hit = true; out.geom = i; out.tri = -1; out.flat = -1; out.t = d; out.point = worldPOI;
                    out.normal = xyz(mulMV(inverse(transpose(g.model)), v4(local.normal, 0.f)));
                }
            }
        } else if (g.type == SPHERE) {
            Ray ps{xyz(mulMV(g.inv, v4(ray.o, 1.f))), xyz(normalize4(mulMV(g.inv, v4(ray.d, 0.f))))};
            Hit local;
            if (intersectSphere(g, ps, local)) {
                V3 worldPOI = xyz(mulMV(g.model, v4(local.point, 1.0f)));
                float d = distance(worldPOI, ray.o);
                if (d < tMax) {
                    tMax = d; out.hit = true; out.geom = i; out.tri = -1; out.flat = -1; out.t = d; out.point = worldPOI;
                    out.normal = xyz(mulMV(inverse(transpose(g.model)), v4(local.normal, 0.f)));
                }
            }
        }
    }
    return out;
}

template <class F>
void parallel_for(size_t n, int nthreads, F f) {
    if (nthreads <= 1 || n < 2) { f((size_t)0, n); return; }
    std::vector<std::thread> th;
    size_t chunk = (n + nthreads - 1) / nthreads;
    for (int t = 0; t < nthreads; ++t) {
        size_t a = (size_t)t * chunk, b = std::min(n, a + chunk);
        if (a >= b) break;
        th.emplace_back([=] { f(a, b); });
    }
    for (auto& t : th) t.join();
}

struct OScene {
    Scene sc;
    Flat flat;
    std::vector<int> lights;  // flat indices of the triangles of emitter meshes, in flat order
    bool flat_valid = false;
    const Flat& get_flat() {
        if (!flat_valid) {
            flat = Flat(); flatten(sc, flat); flat_valid = true;
            lights.clear();
            for (size_t k = 0; k < flat.tris.size(); ++k) if (material_of(sc, flat.geom[k], flat.tri[k]).type == 0) lights.push_back((int)k);
        }
        return flat;
    }
    const std::vector<int>& get_lights() { get_flat(); return lights; }
};

// Shadow-ray query: is any front-facing triangle (or plane) hit with closest-hit key < tmax?  Same test and
// key as intersectFlat, existence instead of minimum, so the answer does not depend on the scan order.
bool occludedFlat(const Ray& ray, float tmax, const Scene& sc, const Flat& f) {
    Ray os{ray.o, normalize(ray.d)};
    for (size_t k = 0; k < f.tris.size(); ++k) {
        Hit local;
        if (intersectTriangle(f.tris[k], os, local) && distance(local.point, ray.o) < tmax) return true;
    }
    for (const Geom& g : sc.geoms) {
        if (g.type == TRIANGLEMESH) continue;
        Ray ps{xyz(mulMV(g.inv, v4(ray.o, 1.f))), xyz(normalize4(mulMV(g.inv, v4(ray.d, 0.f))))};
        Hit local;
        const bool hit = g.type == PLANE ? intersectPlane(g, ps, local) : intersectSphere(g, ps, local);
        if (hit && distance(xyz(mulMV(g.model, v4(local.point, 1.0f))), ray.o) < tmax) return true;
    }
    return false;
}

// The as-intended path tracer.  One path.
struct RenderParams {
    int spp, max_bounces; uint32_t seed_lo, seed_hi; int first_sample; int aa; float background[3]; float ray_eps; int nee;
};
inline V3 mul3(V3 a, V3 b) { return V3{a.x * b.x, a.y * b.y, a.z * b.z}; }

// Next-event estimation (the north star's shadow-ray stage): one light sample per diffuse
// shading event.  Lights = the triangles of emitter meshes, in flat order.  A triangle is picked uniformly, a point
// uniformly on it; the emitter radiates from its front side only, the side the closest-hit test can see.
// Returns false when the sample carries nothing; otherwise the shadow ray (origin `so`, unnormalised direction
// `sd`, limit `tmax`) and the radiance `contrib` to add when it is unoccluded.
bool sample_light(const OScene& os, const Flat& f, const std::vector<int>& lights, uint32_t pixel, uint32_t sample, int b,
                  const RenderParams& rp, V3 so, V3 n, V3 ng, bool smooth, V3 beta, V3& sd, float& tmax, V3& contrib) {
    const uint32_t nl = (uint32_t)lights.size();
    if (nl == 0) return false;
    U4 r = philox4x32_10(U4{pixel, sample, (uint32_t)b, 2u}, rp.seed_lo, rp.seed_hi);
    uint32_t k = (uint32_t)(u01(r.x) * (float)nl);
    if (k > nl - 1) k = nl - 1;
    const int fi = lights[k];
    const Tri& t = f.tris[fi];
    float su = std::sqrt(u01(r.y));
    float b0 = 1.0f - su, b1 = u01(r.z) * su;
    float b2 = (1.0f - b0) - b1;
    V3 y = add(add(mul(t.v0, b0), mul(t.v1, b1)), mul(t.v2, b2));
    V3 N = cross(sub(t.v1, t.v0), sub(t.v2, t.v0));
    float len = length(N);
    sd = sub(y, so);
    float dist2 = dot(sd, sd);
    float dist = std::sqrt(dist2);
    float cosx = dot(n, sd) / dist;
    float cosl = (-dot(N, sd) / dist) / len;
    if (!(cosx > 0.0f && cosl > 0.0f && dist2 > 0.0f)) return false;
    if (smooth && !(dot(ng, sd) > 0.0f)) return false;  // below the real surface: the shadow ray would leave through culled back faces
    float w = (((cosx * cosl) / dist2) * (0.5f * len)) * ((float)nl * 0.318309886183790672f);
    const Material& m = material_of(os.sc, f.geom[fi], f.tri[fi]);
    contrib = mul(mul3(beta, mul(m.emissive, m.intensity)), w);
    tmax = dist * 0.999f;
    return true;
}

V3 trace_path(OScene& os, const Camera& cam, const M4& invVP, int x, int y, uint32_t sample, const RenderParams& rp,
              long* ray_count, long* shadow_count = nullptr) {
    const Flat& f = os.get_flat();
    const std::vector<int>& lights = os.get_lights();
    uint32_t pixel = (uint32_t)(y * (int)cam.width + x);
    float px = (float)x, py = (float)y;
    if (rp.aa) {
        U4 j = philox4x32_10(U4{pixel, sample, 0u, 1u}, rp.seed_lo, rp.seed_hi);
        px = px + u01(j.z); py = py + u01(j.w);
    }
    Ray ray = gen_ray(cam, invVP, px, py);
    V3 beta{1, 1, 1}, L{0, 0, 0};
    bool specular = false;  // the previous event was a mirror / glass bounce
    V3 bg{rp.background[0], rp.background[1], rp.background[2]};
    for (int b = 0; b <= rp.max_bounces; ++b) {
        FlatHit h = intersectFlat(ray, os.sc, f);
        if (ray_count) ++*ray_count;
        if (!h.hit) { L = add(L, mul3(beta, bg)); break; }
        const Material& m = material_of(os.sc, h.geom, h.tri);
        const SurfAttr sa = surface_attributes(os.sc, f, h.flat, ray);
        if (m.type == 0) {
            // with NEE the direct light of sampled emitters (mesh emitters) arrives through the shadow rays; a hit found
            // by sampling the BSDF counts only for the camera ray and after a specular event (which has no light sample)
            const bool sampled = rp.nee && h.tri >= 0 && !lights.empty();
            if (!sampled || b == 0 || specular) L = add(L, mul3(beta, mul(m.emissive, m.intensity)));
            break;
        }
        if (b == rp.max_bounces) break;
        V3 n = normalize(h.normal);
        if (dot(n, ray.d) > 0.0f) n = neg(n);
        // shading normal: interpolated corner normal when the geometry carries normals, turned to the side of n; ray
        // offsets and sidedness stay with the geometric normal n
        V3 ns = n;
        if (sa.smooth) { ns = sa.ns; if (dot(ns, n) < 0.0f) ns = neg(ns); }
        U4 r = philox4x32_10(U4{pixel, sample, (uint32_t)b, 0u}, rp.seed_lo, rp.seed_hi);
        if (m.type == 2 || m.type == 3) {
            // MIRROR / GLASS (BXDFTyp utilities.h:72-73; the reference declares the fields, utilities.h:84-88, and no
            // code).  Only front faces are ever hit (kernel.cu:48-51), so a ray meets a glass surface from outside only:
            // one refracting interface, air -> ior; Schlick's Fresnel picks reflection with probability F.
            V3 dn = normalize(ray.d);
            float cosi = -dot(dn, ns);
            if (sa.smooth && !(cosi > 0.0f)) { ns = n; cosi = -dot(dn, n); }
            bool reflect = true;
            V3 d;
            if (m.type == 3) {
                float eta = 1.0f / m.ior;
                float r0 = (1.0f - m.ior) / (1.0f + m.ior);
                r0 = r0 * r0;
                float c1 = 1.0f - cosi;
                float F = r0 + (1.0f - r0) * ((c1 * c1) * (c1 * c1) * c1);
                float sin2t = (eta * eta) * (1.0f - cosi * cosi);
                if (!(sin2t > 1.0f) && !(u01(r.z) < F)) {
                    reflect = false;
                    d = add(mul(dn, eta), mul(ns, eta * cosi - std::sqrt(1.0f - sin2t)));
                    beta = mul3(beta, m.transmittance);
                    ray.o = sub(h.point, mul(n, rp.ray_eps));
                }
            }
            if (reflect) {
                d = add(dn, mul(ns, 2.0f * cosi));
                if (sa.smooth) { float c = dot(d, n); if (!(c > 0.0f)) d = sub(d, mul(n, 2.0f * c)); }
                beta = mul3(beta, m.specular);
                ray.o = add(h.point, mul(n, rp.ray_eps));
            }
            ray.d = d;
            specular = true;
            continue;
        }
        specular = false;
        beta = mul3(beta, m.albedo);
        V3 w = cosineSampleHemisphere(u01(r.x), u01(r.y));
        V3 t, bt;
        onb(ns, t, bt);
        V3 d = add(add(mul(t, w.x), mul(bt, w.y)), mul(ns, w.z));
        if (sa.smooth) { float c = dot(d, n); if (!(c > 0.0f)) d = sub(d, mul(n, 2.0f * c)); }
        ray.o = add(h.point, mul(n, rp.ray_eps));
        ray.d = d;
        if (rp.nee) {
            V3 sd, contrib; float tmax;
            if (sample_light(os, f, lights, pixel, sample, b, rp, ray.o, ns, n, sa.smooth, beta, sd, tmax, contrib)) {
                if (shadow_count) ++*shadow_count;
                if (!occludedFlat(Ray{ray.o, sd}, tmax, os.sc, f)) L = add(L, contrib);
            }
        }
    }
    return L;
}

}  // namespace

// =================================================================== C ABI (ctypes from tests/bench)
extern "C" {

int orc_hardware_threads() { return (int)std::thread::hardware_concurrency(); }
void* orc_scene_new() { return new OScene(); }
void orc_scene_free(void* s) { delete (OScene*)s; }

// mat15 = {type (BXDFTyp: 0 emitter, 1 diffuse, 2 mirror, 3 glass), albedo rgb, emissive rgb, intensity, specular rgb, ior,
//          transmittance rgb} or NULL
static void push_material(OScene* os, const float* m15) {
    Material m;
    m.type = 1; m.albedo = v3(0.8f, 0.8f, 0.8f); m.emissive = v3(0, 0, 0); m.intensity = 0.f;
    if (m15) {
        m.type = (int)m15[0]; m.albedo = v3(m15[1], m15[2], m15[3]); m.emissive = v3(m15[4], m15[5], m15[6]); m.intensity = m15[7];
        m.specular = v3(m15[8], m15[9], m15[10]); m.ior = m15[11]; m.transmittance = v3(m15[12], m15[13], m15[14]);
    }
    os->sc.mats.push_back(m);
}
int orc_scene_add_mesh(void* s, const float* xyz9, long n_tris, const float pos[3], const float rot[3],
                       const float scale[3], const float* mat8) {
    OScene* os = (OScene*)s;
    Geom g;
    g.type = TRIANGLEMESH;
    model_matrices(v3(pos[0], pos[1], pos[2]), v3(rot[0], rot[1], rot[2]), v3(scale[0], scale[1], scale[2]), g.model, g.inv);
    g.tris.resize((size_t)n_tris);
    for (long i = 0; i < n_tris; ++i) {
        const float* p = xyz9 + 9 * i;
        g.tris[i] = Tri{v3(p[0], p[1], p[2]), v3(p[3], p[4], p[5]), v3(p[6], p[7], p[8])};
    }
    os->sc.geoms.push_back(std::move(g));
    push_material(os, mat8);
    os->flat_valid = false;
    return (int)os->sc.geoms.size() - 1;
}
int orc_scene_add_plane(void* s, const float pos[3], const float rot[3], const float scale[3], const float* mat8) {
    OScene* os = (OScene*)s;
    Geom g;
    g.type = PLANE;
    model_matrices(v3(pos[0], pos[1], pos[2]), v3(rot[0], rot[1], rot[2]), v3(scale[0], scale[1], scale[2]), g.model, g.inv);
    os->sc.geoms.push_back(std::move(g));
    push_material(os, mat8);
    os->flat_valid = false;
    return (int)os->sc.geoms.size() - 1;
}
int orc_scene_add_sphere(void* s, float radius, const float pos[3], const float rot[3], const float scale[3], const float* mat8) {
    OScene* os = (OScene*)s;
    Geom g;
    g.type = SPHERE;
    g.radius = radius;
    model_matrices(v3(pos[0], pos[1], pos[2]), v3(rot[0], rot[1], rot[2]), v3(scale[0], scale[1], scale[2]), g.model, g.inv);
    os->sc.geoms.push_back(std::move(g));
    push_material(os, mat8);
    os->flat_valid = false;
    return (int)os->sc.geoms.size() - 1;
}
static Material material_from15(const float* m15) {
    Material m;
    m.type = (int)m15[0]; m.albedo = v3(m15[1], m15[2], m15[3]); m.emissive = v3(m15[4], m15[5], m15[6]); m.intensity = m15[7];
    m.specular = v3(m15[8], m15[9], m15[10]); m.ior = m15[11]; m.transmittance = v3(m15[12], m15[13], m15[14]);
    return m;
}
// shading attributes of a mesh geometry (NULL clears): normals9 / uvs6 per triangle; face_mat indexes mats15 rows, -1 = own
void orc_scene_set_normals(void* s, int gi, const float* n9) {
    OScene* os = (OScene*)s;
    Geom& g = os->sc.geoms[gi];
    g.normals.clear();
    if (n9) for (size_t c = 0; c < 3 * g.tris.size(); ++c) g.normals.push_back(v3(n9[3 * c], n9[3 * c + 1], n9[3 * c + 2]));
    os->flat_valid = false;
}
void orc_scene_set_uvs(void* s, int gi, const float* uv6) {
    OScene* os = (OScene*)s;
    Geom& g = os->sc.geoms[gi];
    g.uvs.clear();
    if (uv6) g.uvs.assign(uv6, uv6 + 6 * g.tris.size());
    os->flat_valid = false;
}
void orc_scene_set_face_materials(void* s, int gi, const int* face_mat, const float* mats15, int n_mats) {
    OScene* os = (OScene*)s;
    Geom& g = os->sc.geoms[gi];
    g.face_mat.clear(); g.face_mats.clear();
    if (face_mat) {
        g.face_mat.assign(face_mat, face_mat + g.tris.size());
        for (int k = 0; k < n_mats; ++k) g.face_mats.push_back(material_from15(mats15 + 15 * k));
    }
    os->flat_valid = false;
}
void orc_geometry_matrices(void* s, int gi, float* model16, float* inv16) {
    OScene* os = (OScene*)s;
    std::memcpy(model16, &os->sc.geoms[gi].model, 64);
    std::memcpy(inv16, &os->sc.geoms[gi].inv, 64);
}
long orc_flat_count(void* s) { return (long)((OScene*)s)->get_flat().tris.size(); }
void orc_flat_triangles(void* s, float* xyz9, int* geom, int* tri) {
    const Flat& f = ((OScene*)s)->get_flat();
    for (size_t i = 0; i < f.tris.size(); ++i) {
        if (xyz9) std::memcpy(xyz9 + 9 * i, &f.tris[i], 36);
        if (geom) geom[i] = f.geom[i];
        if (tri) tri[i] = f.tri[i];
    }
}

// mode 0: the reference's full object-space path (kernel.cu:127-176); mode 1: baked/flattened scene.
void orc_intersect(void* s, int mode, const float* rays6, long n, unsigned char* out_hit, int* out_geom, int* out_tri,
                   float* out_t, float* out_point3, float* out_normal3, int nthreads) {
    OScene* os = (OScene*)s;
    if (mode == 1) os->get_flat();
    parallel_for((size_t)n, nthreads, [=](size_t a, size_t b) {
        for (size_t i = a; i < b; ++i) {
            Ray ray{v3(rays6[6 * i], rays6[6 * i + 1], rays6[6 * i + 2]), v3(rays6[6 * i + 3], rays6[6 * i + 4], rays6[6 * i + 5])};
            bool hit; int g, t; float tt; V3 p, nn;
            if (mode == 0) {
                Hit h = intersectRays(ray, os->sc);
                hit = h.hit; g = hit ? h.geom : -1; t = (hit && os->sc.geoms[h.geom].type == TRIANGLEMESH) ? h.tri : -1;
                tt = hit ? h.t : INFINITY; p = h.point; nn = h.normal;
            } else {
                FlatHit h = intersectFlat(ray, os->sc, os->flat);
                hit = h.hit; g = h.geom; t = h.tri; tt = h.t; p = h.point; nn = h.normal;
            }
            if (out_hit) out_hit[i] = hit ? 1 : 0;
            if (out_geom) out_geom[i] = g;
            if (out_tri) out_tri[i] = t;
            if (out_t) out_t[i] = tt;
            if (out_point3) { out_point3[3 * i] = hit ? p.x : 0.f; out_point3[3 * i + 1] = hit ? p.y : 0.f; out_point3[3 * i + 2] = hit ? p.z : 0.f; }
            if (out_normal3) { out_normal3[3 * i] = hit ? nn.x : 0.f; out_normal3[3 * i + 1] = hit ? nn.y : 0.f; out_normal3[3 * i + 2] = hit ? nn.z : 0.f; }
        }
    });
}
int orc_intersect_triangle(const float v9[9], const float o[3], const float d[3], float* t, float* p3, float* n3) {
    Tri tri{v3(v9[0], v9[1], v9[2]), v3(v9[3], v9[4], v9[5]), v3(v9[6], v9[7], v9[8])};
    Ray ray{v3(o[0], o[1], o[2]), v3(d[0], d[1], d[2])};
    Hit h;
    bool ok = intersectTriangle(tri, ray, h);
    if (ok) { *t = h.t; p3[0] = h.point.x; p3[1] = h.point.y; p3[2] = h.point.z; n3[0] = h.normal.x; n3[1] = h.normal.y; n3[2] = h.normal.z; }
    return ok ? 1 : 0;
}
double orc_time_intersect(void* s, int mode, const float* rays6, long n, int nthreads, int reps, int* out_geom,
                          int* out_tri, float* out_t) {
    double best = 1e30;
    for (int r = 0; r < reps; ++r) {
        auto t0 = std::chrono::steady_clock::now();
        orc_intersect(s, mode, rays6, n, nullptr, out_geom, out_tri, out_t, nullptr, nullptr, nthreads);
        double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        if (dt < best) best = dt;
    }
    return best;
}

void orc_camera_rays(const float pos[3], const float world_up[3], float yaw, float pitch, float fov, int width,
                     int height, float near_clip, float far_clip, float* rays6, float* basis9) {
    Camera cam = make_camera(pos, world_up, yaw, pitch, fov, (float)width, (float)height, near_clip, far_clip);
    if (basis9) {
        basis9[0] = cam.forward.x; basis9[1] = cam.forward.y; basis9[2] = cam.forward.z;
        basis9[3] = cam.right.x; basis9[4] = cam.right.y; basis9[5] = cam.right.z;
        basis9[6] = cam.up.x; basis9[7] = cam.up.y; basis9[8] = cam.up.z;
    }
    if (!rays6) return;
    M4 invVP = mulMM(cam.invView(), cam.invProj());
    for (int y = 0; y < height; ++y)
        for (int x = 0; x < width; ++x) {
            Ray r = gen_ray(cam, invVP, (float)x, (float)y);
            float* o = rays6 + 6 * ((size_t)y * width + x);
            o[0] = r.o.x; o[1] = r.o.y; o[2] = r.o.z; o[3] = r.d.x; o[4] = r.d.y; o[5] = r.d.z;
        }
}
void orc_camera_inverse_matrices(const float pos[3], const float world_up[3], float yaw, float pitch, float fov,
                                 int width, int height, float near_clip, float far_clip, float* inv_view16,
                                 float* inv_proj16) {
    Camera cam = make_camera(pos, world_up, yaw, pitch, fov, (float)width, (float)height, near_clip, far_clip);
    M4 iv = cam.invView(), ip = cam.invProj();
    std::memcpy(inv_view16, &iv, 64);
    std::memcpy(inv_proj16, &ip, 64);
}
// make_uchar3(colour * 255.f) (kernel.cu:214) as the CUDA compiler builds it: cvt.rzi.u32.f32 followed by an 8-bit store
// (checked with nvcc 12.9: F2I.U32.TRUNC + STG.U8).  The conversion saturates at 0 and 2^32 - 1 and maps NaN to 0; the
// store keeps the low byte, so a component above 255 WRAPS modulo 256 (|normal| > 1 under a scale below 1).
static inline unsigned char u8_as_cuda(float v) {
    if (!(v > 0.0f)) return 0;                 // NaN, zero, negatives
    if (v >= 4294967296.0f) return 0xff;       // saturates to 0xffffffff
    return (unsigned char)((uint32_t)v & 0xffu);
}
// As-coded frame (kernel.cu:186-216, memset kernel.cu:340). mode as in orc_intersect: 0 = the reference's object-space
// path; 1 = the GPU contract: the hit is found on the baked world-space triangles, the colour is still the normal the
// reference shows, inverse(transpose(M)) * the OBJECT-space unit normal (kernel.cu:101,117), not re-normalised.
void orc_render_as_coded(void* s, int mode, const float pos[3], const float world_up[3], float yaw, float pitch,
                         float fov, int width, int height, float near_clip, float far_clip, unsigned char* rgb,
                         int nthreads) {
    OScene* os = (OScene*)s;
    Camera cam = make_camera(pos, world_up, yaw, pitch, fov, (float)width, (float)height, near_clip, far_clip);
    M4 invVP = mulMM(cam.invView(), cam.invProj());
    std::memset(rgb, 0, (size_t)width * height * 3);
    if (mode == 1) os->get_flat();
    parallel_for((size_t)height, nthreads, [=](size_t a, size_t b) {
        for (size_t y = a; y < b; ++y)
            for (int x = 0; x < width; ++x) {
                Ray ray = gen_ray(cam, invVP, (float)x, (float)y);
                V3 n; bool hit;
                if (mode == 0) { Hit h = intersectRays(ray, os->sc); hit = h.hit; n = h.normal; }
                else {
                    FlatHit h = intersectFlat(ray, os->sc, os->flat); hit = h.hit; n = h.normal;
                    if (hit && h.tri >= 0) {
                        const Geom& g = os->sc.geoms[h.geom];
                        const Tri& t = g.tris[h.tri];
                        V3 no = normalize(cross(sub(t.v1, t.v0), sub(t.v2, t.v0)));
                        n = xyz(mulMV(inverse(transpose(g.model)), v4(no, 0.f)));
                    }
                }
                if (hit) {
                    unsigned char* px = rgb + 3 * (y * width + x);
                    px[0] = u8_as_cuda(std::fabs(n.x) * 255.f); px[1] = u8_as_cuda(std::fabs(n.y) * 255.f); px[2] = u8_as_cuda(std::fabs(n.z) * 255.f);
                }
            }
    });
}

// Primary-ray feature buffers on the baked scene; kind as FF_INTEGRATOR_AOV_*: 3 shading normal, 4 uv, 5 barycentrics,
// 6 albedo.  rgba = W*H*4 floats, alpha 1 on a hit, everything 0 on a miss.
void orc_render_aov(void* s, int kind, const float pos[3], const float world_up[3], float yaw, float pitch, float fov, int width,
                    int height, float near_clip, float far_clip, float* rgba, int nthreads) {
    OScene* os = (OScene*)s;
    Camera cam = make_camera(pos, world_up, yaw, pitch, fov, (float)width, (float)height, near_clip, far_clip);
    M4 invVP = mulMM(cam.invView(), cam.invProj());
    std::memset(rgba, 0, sizeof(float) * 4 * (size_t)width * height);
    os->get_flat();
    parallel_for((size_t)height, nthreads, [=](size_t a, size_t b) {
        for (size_t y = a; y < b; ++y)
            for (int x = 0; x < width; ++x) {
                Ray ray = gen_ray(cam, invVP, (float)x, (float)y);
                FlatHit h = intersectFlat(ray, os->sc, os->flat);
                if (!h.hit) continue;
                float* px = rgba + 4 * (y * width + x);
                SurfAttr sa = surface_attributes(os->sc, os->flat, h.flat, ray);
                px[3] = 1.0f;
                if (kind == 3) {
                    V3 n = normalize(h.normal);
                    if (dot(n, ray.d) > 0.0f) n = neg(n);
                    V3 ns = n;
                    if (sa.smooth) { ns = sa.ns; if (dot(ns, n) < 0.0f) ns = neg(ns); }
                    px[0] = ns.x; px[1] = ns.y; px[2] = ns.z;
                } else if (kind == 4) { px[0] = sa.tu; px[1] = sa.tv; }
                else if (kind == 5) { if (h.flat >= 0) { px[0] = (1.0f - sa.u) - sa.v; px[1] = sa.u; px[2] = sa.v; } }
                else { const Material& m = material_of(os->sc, h.geom, h.tri); px[0] = m.albedo.x; px[1] = m.albedo.y; px[2] = m.albedo.z; }
            }
    });
}
// closest hit on the baked scene + the barycentrics of the hit (0 on misses and analytic primitives)
void orc_intersect_uv(void* s, const float* rays6, long n, int* out_geom, int* out_tri, float* out_u, float* out_v, int nthreads) {
    OScene* os = (OScene*)s;
    os->get_flat();
    parallel_for((size_t)n, nthreads, [=](size_t a, size_t b) {
        for (size_t i = a; i < b; ++i) {
            Ray ray{v3(rays6[6 * i], rays6[6 * i + 1], rays6[6 * i + 2]), v3(rays6[6 * i + 3], rays6[6 * i + 4], rays6[6 * i + 5])};
            FlatHit h = intersectFlat(ray, os->sc, os->flat);
            SurfAttr sa = surface_attributes(os->sc, os->flat, h.hit ? h.flat : -1, ray);
            out_geom[i] = h.hit ? h.geom : -1; out_tri[i] = h.hit ? h.tri : -1;
            out_u[i] = sa.u; out_v[i] = sa.v;
        }
    });
}

void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    U4 r = philox4x32_10(U4{ctr[0], ctr[1], ctr[2], ctr[3]}, key[0], key[1]);
    out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
}
void orc_sincos2pi(const float* u, long n, float* s, float* c) { for (long i = 0; i < n; ++i) sincos2pi(u[i], s[i], c[i]); }
void orc_cosine_sample_hemisphere(float u1, float u2, float* out3) {
    V3 v = cosineSampleHemisphere(u1, u2);
    out3[0] = v.x; out3[1] = v.y; out3[2] = v.z;
}
void orc_onb(const float n[3], float* t3, float* b3) {
    V3 t, b; onb(v3(n[0], n[1], n[2]), t, b);
    t3[0] = t.x; t3[1] = t.y; t3[2] = t.z; b3[0] = b.x; b3[1] = b.y; b3[2] = b.z;
}

// Synthetic soup: triangle i has centroid c ~ U[0,1)^3 and vertices c + s*U[-1,1)^3,
// Philox key {0xF1EEF1E5, N}, counter {i, vertex(0=centroid,1..3), 0, 0}.
void orc_make_soup(long n, float s, float* xyz9) {
    uint32_t k0 = 0xF1EEF1E5u, k1 = (uint32_t)n;
    for (long i = 0; i < n; ++i) {
        U4 c = philox4x32_10(U4{(uint32_t)i, 0u, 0u, 0u}, k0, k1);
        float cx = u01(c.x), cy = u01(c.y), cz = u01(c.z);
        for (int v = 0; v < 3; ++v) {
            U4 r = philox4x32_10(U4{(uint32_t)i, (uint32_t)(v + 1), 0u, 0u}, k0, k1);
            xyz9[9 * i + 3 * v + 0] = cx + s * (u01(r.x) * 2.0f - 1.0f);
            xyz9[9 * i + 3 * v + 1] = cy + s * (u01(r.y) * 2.0f - 1.0f);
            xyz9[9 * i + 3 * v + 2] = cz + s * (u01(r.z) * 2.0f - 1.0f);
        }
    }
}
// Batch A rays (uniform-incoherent): origin ~ U[0,1)^3, direction uniform on the sphere.
// Philox key {0xF1EEF1E5, 1}, counter {i, 0|1, batch_salt, 0}.
void orc_make_rays_uniform(long first, long n, uint32_t salt, float* rays6) {
    for (long k = 0; k < n; ++k) {
        uint32_t i = (uint32_t)(first + k);
        U4 a = philox4x32_10(U4{i, 0u, salt, 0u}, 0xF1EEF1E5u, 1u);
        U4 b = philox4x32_10(U4{i, 1u, salt, 0u}, 0xF1EEF1E5u, 1u);
        float z = 1.0f - 2.0f * u01(b.x);
        float r = std::sqrt(std::fmax(0.0f, 1.0f - z * z));
        float sn, cs;
        sincos2pi(u01(b.y), sn, cs);
        float* o = rays6 + 6 * k;
        o[0] = u01(a.x); o[1] = u01(a.y); o[2] = u01(a.z);
        o[3] = r * cs; o[4] = r * sn; o[5] = z;
    }
}

// As-intended path tracer; out_rgba = W*H*4 floats, mean radiance over spp (alpha = 1).
// mat table comes from the scene.  Returns the number of closest-hit rays traced.
long orc_render(void* s, const float pos[3], const float world_up[3], float yaw, float pitch, float fov, int width,
                int height, float near_clip, float far_clip, int spp, int max_bounces, uint32_t seed_lo,
                uint32_t seed_hi, int first_sample, int aa, const float background[3], float ray_eps, int nee,
                float* out_rgba, long* out_shadow_rays, int nthreads) {
    OScene* os = (OScene*)s;
    os->get_flat();
    Camera cam = make_camera(pos, world_up, yaw, pitch, fov, (float)width, (float)height, near_clip, far_clip);
    M4 invVP = mulMM(cam.invView(), cam.invProj());
    RenderParams rp{spp, max_bounces, seed_lo, seed_hi, first_sample, aa, {background[0], background[1], background[2]}, ray_eps, nee};
    std::vector<long> counts((size_t)std::max(1, nthreads) + 1, 0), scounts((size_t)std::max(1, nthreads) + 1, 0);
    long* cptr = counts.data();
    long* sptr = scounts.data();
    int nt = std::max(1, nthreads);
    size_t chunk = ((size_t)height + nt - 1) / nt;
    parallel_for((size_t)height, nthreads, [=](size_t a, size_t b) {
        long local = 0, slocal = 0;
        for (size_t y = a; y < b; ++y)
            for (int x = 0; x < width; ++x) {
                V3 sum{0, 0, 0};
                for (int k = 0; k < spp; ++k) {  // samples accumulated in index order
                    V3 L = trace_path(*os, cam, invVP, x, (int)y, (uint32_t)(first_sample + k), rp, &local, &slocal);
                    sum = add(sum, L);
                }
                float inv = 1.0f / (float)spp;
                float* px = out_rgba + 4 * (y * width + x);
                px[0] = sum.x * inv; px[1] = sum.y * inv; px[2] = sum.z * inv; px[3] = 1.0f;
            }
        cptr[a / chunk] = local;
        sptr[a / chunk] = slocal;
    });
    long total = 0, stotal = 0;
    for (long c : counts) total += c;
    for (long c : scounts) stotal += c;
    if (out_shadow_rays) *out_shadow_rays = stotal;
    return total;
}

// The same tracer over the pixel rectangle [x0, x0 + cw) x [y0, y0 + ch) of the width x height frame (paths are keyed by
// the pixel's index in the FULL frame, so the crop equals those pixels of orc_render bit for bit): lets a tile of a
// full-size benchmark frame be checked in seconds.  out_rgba = cw * ch * 4 floats, row-major inside the crop.
long orc_render_crop(void* s, const float pos[3], const float world_up[3], float yaw, float pitch, float fov, int width,
                     int height, float near_clip, float far_clip, int spp, int max_bounces, uint32_t seed_lo,
                     uint32_t seed_hi, int first_sample, int aa, const float background[3], float ray_eps, int nee,
                     int x0, int y0, int cw, int ch, float* out_rgba, int nthreads) {
    OScene* os = (OScene*)s;
    os->get_flat();
    Camera cam = make_camera(pos, world_up, yaw, pitch, fov, (float)width, (float)height, near_clip, far_clip);
    M4 invVP = mulMM(cam.invView(), cam.invProj());
    RenderParams rp{spp, max_bounces, seed_lo, seed_hi, first_sample, aa, {background[0], background[1], background[2]}, ray_eps, nee};
    std::vector<long> counts((size_t)cw * ch, 0);
    long* cptr = counts.data();
    parallel_for((size_t)cw * ch, nthreads, [=](size_t a, size_t b) {
        for (size_t i = a; i < b; ++i) {
            const int x = x0 + (int)(i % (size_t)cw), y = y0 + (int)(i / (size_t)cw);
            long local = 0, slocal = 0;
            V3 sum{0, 0, 0};
            for (int k = 0; k < spp; ++k) sum = add(sum, trace_path(*os, cam, invVP, x, y, (uint32_t)(first_sample + k), rp, &local, &slocal));
            float inv = 1.0f / (float)spp;
            float* px = out_rgba + 4 * i;
            px[0] = sum.x * inv; px[1] = sum.y * inv; px[2] = sum.z * inv; px[3] = 1.0f;
            cptr[i] = local;
        }
    });
    long total = 0;
    for (long c : counts) total += c;
    return total;
}

}  // extern "C"
