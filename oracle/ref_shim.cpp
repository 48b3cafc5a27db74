// TEST INFRASTRUCTURE ONLY -- never linked into, imported by, or called from the product path.
//
// Host shell around the reference's own routines.  The routine text itself is NOT in this
// repository: oracle/extract_ref.py cuts it out of /root/reference at build time into
// oracle/_ref/*.inc (git-ignored) and this file #includes those fragments with `__device__`
// defined away, against the reference's vendored glm 0.9.9.7.  So the arithmetic executed here
// *is* the reference's arithmetic (kernel.cu:8-176,178-184,198-205; utilities.h:41-66,141-234,
// 257-427).  Compile: g++ -std=c++17 -O2 -ffp-contract=off -fno-fast-math (see oracle/Makefile).
//
// What this file adds (my text, not the reference's): a C ABI, a std::thread batch driver, a
// guarded OBJ gather (utilities.h:823-824 reads index -1 when an OBJ has no vn/vt), and the u8
// conversion of kernel.cu:214.
#define __device__
#define __host__
#define __global__
#include <cmath>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <string>
#include <thread>
#include <vector>
#include <chrono>

#include "vec3.hpp"
#include "glm.hpp"
#include "gtc/matrix_transform.hpp"

#ifndef CUDART_PI_F
#define CUDART_PI_F 3.141592654f  // value of math_constants.h's CUDART_PI_F (utilities.h:13,49)
#endif
#ifndef INFINITY
#define INFINITY (__builtin_inff())
#endif

struct BXDF;  // utilities.h:77-139 is not extracted: it needs curand/threadIdx and is dead code

#include "_ref/x_sampling.inc"
#include "_ref/x_intersect.inc"
#include "_ref/x_geomtypes.inc"
#include "_ref/x_ray.inc"
#include "_ref/x_camera.inc"
#include "_ref/x_hotpath.inc"
#include "_ref/x_shade.inc"

#define TINYOBJLOADER_IMPLEMENTATION
#include "tiny_obj_loader.h"

namespace {

struct RefScene {
    std::vector<Geometry*> owned;
    std::vector<Geometry> geoms;  // shallow copies, like main() (kernel.cu:253-258)
};

Ray ref_gen_ray(Camera camera, int x, int y) {
    Ray ray;
#include "_ref/x_raygen.inc"
    return ray;
}

void fill_camera(Camera& camera, const float pos[3], const float world_up[3], float yaw, float pitch,
                 float fov, float width, float height, float near_clip, float far_clip) {
    // same field assignments as kernel.cu:311-322
    camera.m_position = glm::vec3(pos[0], pos[1], pos[2]);
    camera.m_forward = glm::vec3(0.f, 0.f, -1.f);
    camera.m_worldUp = glm::vec3(world_up[0], world_up[1], world_up[2]);
    camera.m_fov = fov;
    camera.m_screenHeight = height;
    camera.m_screenWidth = width;
    camera.m_nearClip = near_clip;
    camera.m_farClip = far_clip;
    camera.m_pitch = pitch;
    camera.m_yaw = yaw;
    camera.UpdateBasisAxis();
}

template <class F>
void parallel_for(size_t n, int nthreads, F f) {
    if (nthreads <= 1 || n < 2) { f(0, n); return; }
    std::vector<std::thread> th;
    size_t chunk = (n + nthreads - 1) / nthreads;
    for (int t = 0; t < nthreads; ++t) {
        size_t a = (size_t)t * chunk, b = std::min(n, a + chunk);
        if (a >= b) break;
        th.emplace_back([=] { f(a, b); });
    }
    for (auto& t : th) t.join();
}

}  // namespace

extern "C" {

int ref_hardware_threads() { return (int)std::thread::hardware_concurrency(); }

void* ref_scene_new() { return new RefScene(); }

void ref_scene_free(void* s) {
    RefScene* sc = (RefScene*)s;
    for (Geometry* g : sc->owned) { delete[] g->m_triangles; delete g; }
    delete sc;
}

// = Geometry(TRIANGLEMESH, pos, rotDeg, scale, tris) (utilities.h:176) + push_back (kernel.cu:254)
int ref_scene_add_mesh(void* s, const float* xyz9, long n_tris, const float pos[3], const float rot[3],
                       const float scale[3]) {
    RefScene* sc = (RefScene*)s;
    std::vector<Triangle> tris((size_t)n_tris);
    for (long i = 0; i < n_tris; ++i) {
        const float* p = xyz9 + 9 * i;
        tris[i] = Triangle(glm::vec3(p[0], p[1], p[2]), glm::vec3(p[3], p[4], p[5]), glm::vec3(p[6], p[7], p[8]),
                           glm::vec2(0), glm::vec2(0), glm::vec2(0), glm::vec3(0), glm::vec3(0), glm::vec3(0));
    }
    Geometry* g = new Geometry(GeometryType::TRIANGLEMESH, glm::vec3(pos[0], pos[1], pos[2]),
                               glm::vec3(rot[0], rot[1], rot[2]), glm::vec3(scale[0], scale[1], scale[2]), tris);
    if (n_tris == 0) { g->m_numberOfTriangles = 0; g->m_triangles = nullptr; }
    sc->owned.push_back(g);
    sc->geoms.push_back(*g);
    return (int)sc->geoms.size() - 1;
}

int ref_scene_add_plane(void* s, const float pos[3], const float rot[3], const float scale[3]) {
    RefScene* sc = (RefScene*)s;
    Geometry* g = new Geometry(GeometryType::PLANE, glm::vec3(pos[0], pos[1], pos[2]),
                               glm::vec3(rot[0], rot[1], rot[2]), glm::vec3(scale[0], scale[1], scale[2]));
    g->m_numberOfTriangles = 0;
    sc->owned.push_back(g);
    sc->geoms.push_back(*g);
    return (int)sc->geoms.size() - 1;
}

void ref_geometry_matrices(void* s, int gi, float* model16, float* inv16) {
    RefScene* sc = (RefScene*)s;
    std::memcpy(model16, &sc->geoms[gi].m_modelMatrix[0][0], 64);   // column-major
    std::memcpy(inv16, &sc->geoms[gi].m_inverseModelMatrix[0][0], 64);
}

// Batch driver around intersectRays (kernel.cu:127-176).  tri is reported as -1 when the winning
// geometry is a plane (the reference leaves a stale mesh index there, kernel.cu:161-163).
void ref_intersect(void* s, const float* rays6, long n, unsigned char* out_hit, int* out_geom, int* out_tri,
                   float* out_t, float* out_point3, float* out_normal3, int nthreads) {
    RefScene* sc = (RefScene*)s;
    parallel_for((size_t)n, nthreads, [=](size_t a, size_t b) {
        for (size_t i = a; i < b; ++i) {
            Ray ray(glm::vec3(rays6[6 * i], rays6[6 * i + 1], rays6[6 * i + 2]),
                    glm::vec3(rays6[6 * i + 3], rays6[6 * i + 4], rays6[6 * i + 5]));
            Intersect it = intersectRays(ray, sc->geoms.data(), (unsigned)sc->geoms.size());
            bool hit = it.m_hit;
            if (out_hit) out_hit[i] = hit ? 1 : 0;
            int g = hit ? it.geometryIndex : -1;
            int t = -1;
            if (hit && sc->geoms[g].m_geometryType == GeometryType::TRIANGLEMESH) t = it.triangleIndex;
            if (out_geom) out_geom[i] = g;
            if (out_tri) out_tri[i] = t;
            if (out_t) out_t[i] = hit ? it.m_t : INFINITY;
            for (int k = 0; k < 3; ++k) {
                if (out_point3) out_point3[3 * i + k] = hit ? it.m_intersectionPoint[k] : 0.f;
                if (out_normal3) out_normal3[3 * i + k] = hit ? it.m_normal[k] : 0.f;
            }
        }
    });
}

// Single ray/triangle test, object space (kernel.cu:35-108). Returns hit flag.
int ref_intersect_triangle(const float v9[9], const float o[3], const float d[3], float* t, float* p3, float* n3) {
    Triangle tri(glm::vec3(v9[0], v9[1], v9[2]), glm::vec3(v9[3], v9[4], v9[5]), glm::vec3(v9[6], v9[7], v9[8]),
                 glm::vec2(0), glm::vec2(0), glm::vec2(0), glm::vec3(0), glm::vec3(0), glm::vec3(0));
    Ray ray(glm::vec3(o[0], o[1], o[2]), glm::vec3(d[0], d[1], d[2]));
    Intersect it;
    bool h = intersectTriangle(tri, ray, it);
    if (h) {
        *t = it.m_t;
        for (int k = 0; k < 3; ++k) { p3[k] = it.m_intersectionPoint[k]; n3[k] = it.m_normal[k]; }
    }
    return h ? 1 : 0;
}

// Camera rays for every pixel, row-major, as coded (kernel.cu:186-205, integer pixel corner).
void ref_camera_rays(const float pos[3], const float world_up[3], float yaw, float pitch, float fov, int width,
                     int height, float near_clip, float far_clip, float* rays6, float* basis9) {
    Camera camera;
    fill_camera(camera, pos, world_up, yaw, pitch, fov, (float)width, (float)height, near_clip, far_clip);
    if (basis9) {
        for (int k = 0; k < 3; ++k) {
            basis9[k] = camera.m_forward[k]; basis9[3 + k] = camera.m_right[k]; basis9[6 + k] = camera.m_up[k];
        }
    }
    if (!rays6) return;
    for (int y = 0; y < height; ++y)
        for (int x = 0; x < width; ++x) {
            Ray r = ref_gen_ray(camera, x, y);
            float* o = rays6 + 6 * ((size_t)y * width + x);
            for (int k = 0; k < 3; ++k) { o[k] = r.m_origin[k]; o[3 + k] = r.m_direction[k]; }
        }
}

void ref_camera_inverse_matrices(const float pos[3], const float world_up[3], float yaw, float pitch, float fov,
                                 int width, int height, float near_clip, float far_clip, float* inv_view16,
                                 float* inv_proj16) {
    Camera camera;
    fill_camera(camera, pos, world_up, yaw, pitch, fov, (float)width, (float)height, near_clip, far_clip);
    glm::mat4 iv = camera.GetInverseViewMatrix(), ip = camera.GetInverseProjectionMatrix();
    std::memcpy(inv_view16, &iv[0][0], 64);
    std::memcpy(inv_proj16, &ip[0][0], 64);
}

// As-coded frame (kernel.cu:186-216 + memset kernel.cu:340): rgb8 W*H*3, row 0 first.
void ref_render_as_coded(void* s, const float pos[3], const float world_up[3], float yaw, float pitch, float fov,
                         int width, int height, float near_clip, float far_clip, unsigned char* rgb, int nthreads) {
    RefScene* sc = (RefScene*)s;
    Camera camera;
    fill_camera(camera, pos, world_up, yaw, pitch, fov, (float)width, (float)height, near_clip, far_clip);
    std::memset(rgb, 0, (size_t)width * height * 3);
    parallel_for((size_t)height, nthreads, [=](size_t a, size_t b) {
        for (size_t y = a; y < b; ++y)
            for (int x = 0; x < width; ++x) {
                Ray ray = ref_gen_ray(camera, x, (int)y);
                Intersect it = intersectRays(ray, sc->geoms.data(), (unsigned)sc->geoms.size());
                if (it.m_hit) {
                    glm::vec3 dir;
                    glm::vec3 c = shade(ray, it, dir, sc->geoms.data());
                    unsigned char* px = rgb + 3 * (y * width + x);
                    // make_uchar3(c*255.f) (kernel.cu:214) as nvcc builds it: cvt.rzi.u32.f32 + an 8-bit store, i.e. the
                    // value saturates to [0, 2^32 - 1] (NaN -> 0) and the low byte is kept (wraps above 255)
                    for (int k = 0; k < 3; ++k) {
                        float v = c[k] * 255.f;
                        px[k] = !(v > 0.0f) ? 0 : v >= 4294967296.0f ? 0xff : (unsigned char)((uint32_t)v & 0xffu);
                    }
                }
            }
    });
}

void ref_cosine_sample_hemisphere(float u1, float u2, float* out3) {
    glm::vec3 v = CosineSampleHemisphere(u1, u2);
    out3[0] = v.x; out3[1] = v.y; out3[2] = v.z;
}

// LoadMesh (utilities.h:781-840) position gather via the vendored tinyobjloader, with the
// missing-vn/vt read guarded.  Returns triangle count, or -1 on load failure.  If xyz9 is null
// only counts.
long ref_load_obj(const char* path, float* xyz9, long cap_tris) {
    tinyobj::attrib_t attrib;
    std::vector<tinyobj::shape_t> shapes;
    std::vector<tinyobj::material_t> materials;
    std::string warn, err;
    bool ret = tinyobj::LoadObj(&attrib, &shapes, &materials, &warn, &err, path);
    if (!ret) return -1;
    long n = 0;
    for (size_t s = 0; s < shapes.size(); s++) {
        size_t index_offset = 0;
        for (size_t f = 0; f < shapes[s].mesh.num_face_vertices.size(); f++) {
            int fv = shapes[s].mesh.num_face_vertices[f];
            if (xyz9 && n < cap_tris) {
                for (int v = 0; v < 3; v++) {
                    tinyobj::index_t idx = shapes[s].mesh.indices[index_offset + v];
                    for (int k = 0; k < 3; ++k) xyz9[9 * n + 3 * v + k] = attrib.vertices[3 * idx.vertex_index + k];
                }
            }
            index_offset += fv;
            ++n;
        }
    }
    return n;
}

// LoadMesh (utilities.h:806-837) with every attribute it gathers: per-corner position, normal, uv and the per-face
// material id it reads and drops (utilities.h:837); normal / uv reads guarded for files without `vn` / `vt`
// (index -1 -> zeros).  Any output may be null.
long ref_load_obj_attrs(const char* path, float* xyz9, float* n9, float* uv6, int* material_id, long cap_tris) {
    tinyobj::attrib_t attrib;
    std::vector<tinyobj::shape_t> shapes;
    std::vector<tinyobj::material_t> materials;
    std::string warn, err;
    bool ret = tinyobj::LoadObj(&attrib, &shapes, &materials, &warn, &err, path);
    if (!ret) return -1;
    long n = 0;
    for (size_t s = 0; s < shapes.size(); s++) {
        size_t index_offset = 0;
        for (size_t f = 0; f < shapes[s].mesh.num_face_vertices.size(); f++) {
            int fv = shapes[s].mesh.num_face_vertices[f];
            if (n < cap_tris) {
                for (int v = 0; v < 3; v++) {
                    tinyobj::index_t idx = shapes[s].mesh.indices[index_offset + v];
                    for (int k = 0; k < 3; ++k) {
                        if (xyz9) xyz9[9 * n + 3 * v + k] = attrib.vertices[3 * idx.vertex_index + k];
                        if (n9) n9[9 * n + 3 * v + k] = idx.normal_index >= 0 ? attrib.normals[3 * idx.normal_index + k] : 0.f;
                    }
                    for (int k = 0; k < 2; ++k)
                        if (uv6) uv6[6 * n + 2 * v + k] = idx.texcoord_index >= 0 ? attrib.texcoords[2 * idx.texcoord_index + k] : 0.f;
                }
                if (material_id) material_id[n] = shapes[s].mesh.material_ids[f];
            }
            index_offset += fv;
            ++n;
        }
    }
    return n;
}

// Camera::ProcessKeyboard / ProcessMouseMovement / resetCamera (utilities.h:343-427) driven through the reference's own
// struct.  state = {pos xyz, yaw, pitch}; ops = n x {kind, a, b}: kind 0 = ProcessKeyboard((int)a); kind 1 =
// SetClickPosition(a, b) (absolute deltas) then ProcessMouseMovement(true).  Outputs the final state and basis.
void ref_camera_walk(const float world_up[3], float* state5, const float* ops3, long n_ops, float* basis9) {
    Camera cam;
    cam.m_position = glm::vec3(state5[0], state5[1], state5[2]);
    cam.m_worldUp = glm::vec3(world_up[0], world_up[1], world_up[2]);
    cam.m_yaw = state5[3]; cam.m_pitch = state5[4];
    cam.UpdateBasisAxis();
    for (long i = 0; i < n_ops; ++i) {
        const float* op = ops3 + 3 * i;
        if ((int)op[0] == 0) cam.ProcessKeyboard((int)op[1]);
        else { cam.SetClickPosition(op[1], op[2]); cam.ProcessMouseMovement(true); }
    }
    state5[0] = cam.m_position.x; state5[1] = cam.m_position.y; state5[2] = cam.m_position.z; state5[3] = cam.m_yaw; state5[4] = cam.m_pitch;
    const glm::vec3 b[3] = {cam.m_forward, cam.m_right, cam.m_up};
    for (int k = 0; k < 3; ++k) { basis9[3 * k] = b[k].x; basis9[3 * k + 1] = b[k].y; basis9[3 * k + 2] = b[k].z; }
}

// Timed brute-force batch for the CPU baseline: returns best-of-`reps` seconds.
double ref_time_intersect(void* s, const float* rays6, long n, int nthreads, int reps, int* out_geom, int* out_tri,
                          float* out_t) {
    double best = 1e30;
    for (int r = 0; r < reps; ++r) {
        auto t0 = std::chrono::steady_clock::now();
        ref_intersect(s, rays6, n, nullptr, out_geom, out_tri, out_t, nullptr, nullptr, nthreads);
        auto t1 = std::chrono::steady_clock::now();
        double dt = std::chrono::duration<double>(t1 - t0).count();
        if (dt < best) best = dt;
    }
    return best;
}

}  // extern "C"
