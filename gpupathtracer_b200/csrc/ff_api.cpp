// ff_api.cpp -- the C ABI of include/firefly.h: context, scene building / commit, trace and render entry points.
// Host counterpart of the reference's main() set-up code (kernel.cu:223-329) and per-frame launch (kernel.cu:331-344).
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <limits>
#include <string>
#include <vector>

#include "ff_host.h"

namespace ff {
static thread_local std::string g_error;
void set_error(const std::string& msg) { g_error = msg; }
const std::string& get_error() { return g_error; }
bool load_obj_positions(const char* path, std::vector<float>& xyz9);
int write_image_float(const char* path, const float* rgba, int w, int h);
int write_image_u8(const char* path, const unsigned char* rgb, int w, int h);
}  // namespace ff

using namespace ff;

static void free_scene_device(ff_scene* s) {
    if (s->nodes_tex) cudaDestroyTextureObject(s->nodes_tex);
    if (s->tris_tex) cudaDestroyTextureObject(s->tris_tex);
    s->nodes_tex = s->tris_tex = 0;
    cudaFree(s->d_flat); cudaFree(s->d_tris); cudaFree(s->d_nodes); cudaFree(s->d_ord_geom); cudaFree(s->d_ord_tri);
    cudaFree(s->d_materials); cudaFree(s->d_planes); cudaFree(s->d_bvh2_children); cudaFree(s->d_bvh2_lo);
    cudaFree(s->d_bvh2_hi); cudaFree(s->d_leaf_slot); cudaFree(s->d_lights); s->d_lights = nullptr;
    cudaFree(s->d_as_coded_normals); s->d_as_coded_normals = nullptr;
    cudaFree(s->d_bounds); s->d_bounds = nullptr;
    cudaFree(s->d_geom_attr); cudaFree(s->d_attr_normals); cudaFree(s->d_attr_uvs); cudaFree(s->d_attr_facemat);
    s->d_geom_attr = nullptr; s->d_attr_normals = nullptr; s->d_attr_uvs = nullptr; s->d_attr_facemat = nullptr;
    s->d_flat = s->d_tris = s->d_nodes = nullptr; s->d_ord_geom = s->d_ord_tri = nullptr; s->d_materials = nullptr;
    s->d_planes = nullptr; s->d_bvh2_children = nullptr; s->d_bvh2_lo = s->d_bvh2_hi = nullptr; s->d_leaf_slot = nullptr;
    s->committed = false;
}
static void free_render_scratch(ff_scene* s) {
    for (int k = 0; k < 2; ++k) { cudaFree(s->wb.ray_o[k]); cudaFree(s->wb.ray_d[k]); cudaFree(s->wb.beta[k]); s->wb.ray_o[k] = s->wb.ray_d[k] = s->wb.beta[k] = nullptr; }
    cudaFree(s->wb.hits); cudaFree(s->wb.radiance); cudaFree(s->wb.counts); cudaFree(s->wb.sh_o); cudaFree(s->wb.sh_d); cudaFree(s->wb.sh_c);
    s->wb.hits = nullptr; s->wb.radiance = nullptr; s->wb.counts = nullptr; s->wb.sh_o = s->wb.sh_d = s->wb.sh_c = nullptr;
    s->wb_capacity = 0; s->wb_counts = 0; s->wb_shadow = false;
    cudaFree(s->d_accum); cudaFree(s->d_out); s->d_accum = s->d_out = nullptr; s->fb_pixels = 0;
    cudaFree(s->d_pixel_map); s->d_pixel_map = nullptr; s->pixel_map_count = 0;
    cudaFree(s->d_u8); s->d_u8 = nullptr; s->u8_bytes = 0;
}

template <class T>
static cudaError_t upload(T** dptr, const std::vector<T>& h, cudaStream_t st) {
    *dptr = nullptr;
    if (h.empty()) return cudaSuccess;
    cudaError_t e = cudaMalloc((void**)dptr, sizeof(T) * h.size());
    if (e != cudaSuccess) return e;
    return cudaMemcpyAsync(*dptr, h.data(), sizeof(T) * h.size(), cudaMemcpyHostToDevice, st);
}

extern "C" {

const char* ff_last_error(void) { return g_error.c_str(); }
int ff_version(void) { return FF_VERSION; }
int ff_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int ff_context_create(ff_context** out, int device_id) {
    FF_REQUIRE(out, "ff_context_create: out is NULL");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        set_error(std::string("ff_context_create: no CUDA device (") + (e != cudaSuccess ? cudaGetErrorString(e) : "count 0") +
                  "); this library has no CPU fallback");
        return FF_ERR_CUDA;
    }
    FF_REQUIRE(device_id >= 0 && device_id < n, "ff_context_create: device id out of range");
    FF_CUDA(cudaSetDevice(device_id));
    ff_context* c = new ff_context();
    c->device = device_id;
    FF_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    for (int k = 0; k < 4; ++k) FF_CUDA(cudaEventCreate(&c->ev[k]));
    for (int k = 0; k < 2; ++k) FF_CUDA(cudaEventCreate(&c->timer_ev[k]));
    *out = c;
    return FF_OK;
}
static void release_scene(ff_scene* scene);
void ff_context_destroy(ff_context* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (ff_scene* s : ctx->scenes) release_scene(s);  // a scene destroyed later only frees its host object
    ctx->scenes.clear();
    for (auto e : ctx->trace_events) cudaEventDestroy(e);
    for (int k = 0; k < 4; ++k) cudaEventDestroy(ctx->ev[k]);
    for (int k = 0; k < 2; ++k) cudaEventDestroy(ctx->timer_ev[k]);
    cudaFree(ctx->flush_buf);
    cudaFree(ctx->resolve_buf);
    if (ctx->pipe.ready) {
        TracePipe& tp = ctx->pipe;
        for (int k = 0; k < FF_PIPE_SLOTS; ++k) { cudaFree(tp.block[k]); cudaEventDestroy(tp.in_done[k]); cudaEventDestroy(tp.comp_done[k]); cudaEventDestroy(tp.out_done[k]); }
        for (int k = 0; k < 5; ++k) cudaEventDestroy(tp.span[k]);
        cudaFree(tp.work);
        cudaFree(tp.sblock); cudaFree(tp.d_sync); cudaFreeHost(tp.h_flags); cudaFreeHost(tp.h_marks);
        if (tp.s_go) cudaEventDestroy(tp.s_go);
        cudaStreamDestroy(tp.s_in); cudaStreamDestroy(tp.s_out); cudaStreamDestroy(tp.s_comp);
    }
    ctx->arena.release();
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}
int ff_context_device(const ff_context* ctx) { return ctx ? ctx->device : -1; }
int ff_context_synchronize(ff_context* ctx) {
    FF_REQUIRE(ctx, "ff_context_synchronize: NULL context");
    FF_CUDA(cudaStreamSynchronize(ctx->stream));
    return FF_OK;
}

int ff_context_release_scratch(ff_context* ctx) {
    FF_REQUIRE(ctx, "ff_context_release_scratch: NULL context");
    FF_CUDA(cudaSetDevice(ctx->device));
    FF_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->arena.release();
    return FF_OK;
}
int ff_timer_begin(ff_context* ctx) {
    FF_REQUIRE(ctx, "ff_timer_begin: NULL context");
    FF_CUDA(cudaSetDevice(ctx->device));
    FF_CUDA(cudaEventRecord(ctx->timer_ev[0], ctx->stream));
    ctx->timer_open = true;
    return FF_OK;
}
int ff_timer_end(ff_context* ctx, double* out_ms) {
    FF_REQUIRE(ctx && out_ms, "ff_timer_end: NULL argument");
    FF_REQUIRE(ctx->timer_open, "ff_timer_end: no ff_timer_begin on this context");
    FF_CUDA(cudaSetDevice(ctx->device));
    FF_CUDA(cudaEventRecord(ctx->timer_ev[1], ctx->stream));
    FF_CUDA(cudaEventSynchronize(ctx->timer_ev[1]));
    float ms = 0.f;
    FF_CUDA(cudaEventElapsedTime(&ms, ctx->timer_ev[0], ctx->timer_ev[1]));
    ctx->timer_open = false;
    *out_ms = (double)ms;
    return FF_OK;
}

// ------------------------------------------------------------------------------------------ scene building
int ff_scene_create(ff_context* ctx, ff_scene** out) {
    FF_REQUIRE(ctx && out, "ff_scene_create: NULL argument");
    ff_scene* s = new ff_scene();
    s->ctx = ctx;
    memset(&s->dev, 0, sizeof(s->dev));
    memset(&s->info, 0, sizeof(s->info));
    memset(&s->wb, 0, sizeof(s->wb));
    ctx->scenes.push_back(s);
    *out = s;
    return FF_OK;
}
static void release_scene(ff_scene* scene) {  // device side only; the context must still be alive
    free_scene_device(scene);
    free_render_scratch(scene);
    cudaFree(scene->d_work_counter); cudaFree(scene->d_ray_counter); cudaFree(scene->d_trace_counters);
    scene->d_work_counter = nullptr; scene->d_ray_counter = nullptr; scene->d_trace_counters = nullptr;
    scene->ctx = nullptr;
}
void ff_scene_destroy(ff_scene* scene) {
    if (!scene) return;
    if (scene->ctx) {
        ff_context* ctx = scene->ctx;
        cudaSetDevice(ctx->device);
        cudaStreamSynchronize(ctx->stream);
        ctx->scenes.erase(std::remove(ctx->scenes.begin(), ctx->scenes.end(), scene), ctx->scenes.end());
        release_scene(scene);
    }
    delete scene;
}
static ff_material default_material() {
    ff_material m;
    m.type = FF_MAT_DIFFUSE;
    m.albedo[0] = m.albedo[1] = m.albedo[2] = 0.8f;
    m.emissive[0] = m.emissive[1] = m.emissive[2] = 0.f;
    m.intensity = 0.f;
    m.specular[0] = m.specular[1] = m.specular[2] = 1.f;
    m.ior = 1.5f;
    m.transmittance[0] = m.transmittance[1] = m.transmittance[2] = 1.f;
    return m;
}
static int add_geom(ff_scene* scene, int type, std::vector<float>&& xyz9, const float pos[3], const float rot[3], const float scl[3],
                    const ff_material* material, int* out_index) {
    static const float zero[3] = {0, 0, 0}, one[3] = {1, 1, 1};
    HostGeom g;
    g.type = type;
    ffh::model_from_trs(pos ? pos : zero, rot ? rot : zero, scl ? scl : one, g.model, g.inv);
    g.mat = material ? *material : default_material();
    g.xyz9 = std::move(xyz9);
    scene->geoms.push_back(std::move(g));
    scene->committed = false;
    if (out_index) *out_index = (int)scene->geoms.size() - 1;
    return FF_OK;
}
long ff_load_obj(const char* obj_path, float* xyz9, size_t capacity_tris) {
    if (!obj_path) { set_error("ff_load_obj: NULL path"); return FF_ERR_INVALID; }
    std::vector<float> v;
    if (!load_obj_positions(obj_path, v)) return FF_ERR_IO;
    const size_t n = v.size() / 9;
    if (xyz9) memcpy(xyz9, v.data(), sizeof(float) * 9 * std::min(n, capacity_tris));
    return (long)n;
}
int ff_scene_add_obj(ff_scene* scene, const char* obj_path, const float pos[3], const float rot_deg[3], const float scale[3],
                     const ff_material* material, int flip_winding, int* out_geometry_index) {
    FF_REQUIRE(scene && obj_path, "ff_scene_add_obj: NULL argument");
    std::vector<float> v;
    if (!load_obj_positions(obj_path, v)) return FF_ERR_IO;
    if (flip_winding)
        for (size_t t = 0; t + 8 < v.size(); t += 9)
            for (int a = 0; a < 3; ++a) std::swap(v[t + 3 + a], v[t + 6 + a]);
    return add_geom(scene, 2, std::move(v), pos, rot_deg, scale, material, out_geometry_index);
}
int ff_scene_add_triangles(ff_scene* scene, const float* xyz9, size_t n_tris, const float pos[3], const float rot_deg[3],
                           const float scale[3], const ff_material* material, int* out_geometry_index) {
    FF_REQUIRE(scene && (xyz9 || n_tris == 0), "ff_scene_add_triangles: NULL argument");
    std::vector<float> v(xyz9, xyz9 + 9 * n_tris);
    return add_geom(scene, 2, std::move(v), pos, rot_deg, scale, material, out_geometry_index);
}
int ff_scene_add_plane(ff_scene* scene, const float pos[3], const float rot_deg[3], const float scale[3], const ff_material* material,
                       int* out_geometry_index) {
    FF_REQUIRE(scene, "ff_scene_add_plane: NULL scene");
    return add_geom(scene, 1, std::vector<float>(), pos, rot_deg, scale, material, out_geometry_index);
}

int ff_scene_add_sphere(ff_scene* scene, float radius, const float pos[3], const float rot_deg[3], const float scale[3],
                        const ff_material* material, int* out_geometry_index) {
    FF_REQUIRE(scene, "ff_scene_add_sphere: NULL scene");
    FF_REQUIRE(radius > 0.f, "ff_scene_add_sphere: radius must be positive");
    int idx = -1;
    const int rc = add_geom(scene, 0, std::vector<float>(), pos, rot_deg, scale, material, &idx);
    if (rc == FF_OK) {
        scene->geoms[(size_t)idx].radius = radius;
        if (out_geometry_index) *out_geometry_index = idx;
    }
    return rc;
}

static int mesh_geom(ff_scene* scene, int gi, const char* who, HostGeom** out) {
    if (!scene) { set_error(std::string(who) + ": NULL scene"); return FF_ERR_INVALID; }
    if (gi < 0 || (size_t)gi >= scene->geoms.size() || scene->geoms[(size_t)gi].type != 2) {
        set_error(std::string(who) + ": geometry index does not name a triangle mesh"); return FF_ERR_INVALID;
    }
    *out = &scene->geoms[(size_t)gi];
    return FF_OK;
}
int ff_scene_set_normals(ff_scene* scene, int geometry_index, const float* normals9) {
    HostGeom* g;
    int rc = mesh_geom(scene, geometry_index, "ff_scene_set_normals", &g);
    if (rc) return rc;
    if (normals9) g->n9.assign(normals9, normals9 + g->xyz9.size()); else g->n9.clear();
    scene->committed = false;
    return FF_OK;
}
int ff_scene_set_uvs(ff_scene* scene, int geometry_index, const float* uvs6) {
    HostGeom* g;
    int rc = mesh_geom(scene, geometry_index, "ff_scene_set_uvs", &g);
    if (rc) return rc;
    if (uvs6) g->uv6.assign(uvs6, uvs6 + g->xyz9.size() / 9 * 6); else g->uv6.clear();
    scene->committed = false;
    return FF_OK;
}
int ff_scene_set_face_materials(ff_scene* scene, int geometry_index, const int* face_material, const ff_material* materials, int n_materials) {
    HostGeom* g;
    int rc = mesh_geom(scene, geometry_index, "ff_scene_set_face_materials", &g);
    if (rc) return rc;
    scene->committed = false;
    if (!face_material) { g->face_mat.clear(); g->face_materials.clear(); return FF_OK; }
    FF_REQUIRE(n_materials >= 0 && (materials || n_materials == 0), "ff_scene_set_face_materials: NULL materials");
    const size_t nt = g->xyz9.size() / 9;
    for (size_t t = 0; t < nt; ++t)
        FF_REQUIRE(face_material[t] >= -1 && face_material[t] < n_materials, "ff_scene_set_face_materials: material index out of range");
    for (int k = 0; k < n_materials; ++k)
        FF_REQUIRE(materials[k].type >= FF_MAT_EMITTER && materials[k].type <= FF_MAT_GLASS, "ff_scene_set_face_materials: unknown material type");
    g->face_mat.assign(face_material, face_material + nt);
    g->face_materials.assign(materials, materials + n_materials);
    return FF_OK;
}
static std::string dir_of(const std::string& path) {
    const size_t k = path.find_last_of('/');
    return k == std::string::npos ? std::string() : path.substr(0, k + 1);
}
int ff_scene_add_obj_ex(ff_scene* scene, const char* obj_path, const float pos[3], const float rot_deg[3], const float scale[3],
                        const ff_material* material, int flags, int* out_geometry_index) {
    FF_REQUIRE(scene && obj_path, "ff_scene_add_obj_ex: NULL argument");
    FF_REQUIRE((flags & ~15) == 0, "ff_scene_add_obj_ex: unknown flag");
    ObjMesh m;
    if (!load_obj_mesh(obj_path, m, 0)) return FF_ERR_IO;
    const size_t nt = m.xyz9.size() / 9;
    if (flags & FF_OBJ_FLIP_WINDING)
        for (size_t t = 0; t < nt; ++t) {
            for (int a = 0; a < 3; ++a) { std::swap(m.xyz9[9 * t + 3 + a], m.xyz9[9 * t + 6 + a]); std::swap(m.n9[9 * t + 3 + a], m.n9[9 * t + 6 + a]); }
            for (int a = 0; a < 2; ++a) std::swap(m.uv6[6 * t + 2 + a], m.uv6[6 * t + 4 + a]);
        }
    std::vector<ff_material> mats;
    std::vector<int> face_mat;
    if ((flags & FF_OBJ_MATERIALS) && !m.material_names.empty() && !m.mtllib.empty()) {
        std::vector<std::string> lib_names;
        std::vector<ff_material> lib_mats;
        if (!load_mtl((dir_of(obj_path) + m.mtllib).c_str(), lib_names, lib_mats)) return FF_ERR_IO;
        std::vector<int> remap(m.material_names.size(), -1);
        for (size_t k = 0; k < m.material_names.size(); ++k) {
            auto it = std::find(lib_names.begin(), lib_names.end(), m.material_names[k]);
            if (it != lib_names.end()) { remap[k] = (int)mats.size(); mats.push_back(lib_mats[(size_t)(it - lib_names.begin())]); }
        }
        face_mat.resize(nt);
        for (size_t t = 0; t < nt; ++t) face_mat[t] = m.face_material[t] >= 0 ? remap[(size_t)m.face_material[t]] : -1;
    }
    int idx = -1;
    int rc = add_geom(scene, 2, std::move(m.xyz9), pos, rot_deg, scale, material, &idx);
    if (rc) return rc;
    HostGeom& g = scene->geoms[(size_t)idx];
    if ((flags & FF_OBJ_NORMALS) && m.has_normals) g.n9 = std::move(m.n9);
    if ((flags & FF_OBJ_UVS) && m.has_uvs) g.uv6 = std::move(m.uv6);
    if (!mats.empty()) { g.face_mat = std::move(face_mat); g.face_materials = std::move(mats); }
    if (out_geometry_index) *out_geometry_index = idx;
    return FF_OK;
}
long ff_load_obj_ex(const char* obj_path, float* xyz9, float* normals9, float* uvs6, int* face_material, size_t capacity_tris,
                    int threads, int* out_has_normals, int* out_has_uvs, int* out_n_materials, double* out_seconds) {
    if (!obj_path) { set_error("ff_load_obj_ex: NULL path"); return FF_ERR_INVALID; }
    ObjMesh m;
    const auto t0 = std::chrono::steady_clock::now();
    if (!load_obj_mesh(obj_path, m, threads)) return FF_ERR_IO;
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    const size_t n = m.xyz9.size() / 9, k = std::min(n, capacity_tris);
    if (xyz9) memcpy(xyz9, m.xyz9.data(), sizeof(float) * 9 * k);
    if (normals9) memcpy(normals9, m.n9.data(), sizeof(float) * 9 * k);
    if (uvs6) memcpy(uvs6, m.uv6.data(), sizeof(float) * 6 * k);
    if (face_material) memcpy(face_material, m.face_material.data(), sizeof(int) * k);
    if (out_has_normals) *out_has_normals = m.has_normals;
    if (out_has_uvs) *out_has_uvs = m.has_uvs;
    if (out_n_materials) *out_n_materials = (int)m.material_names.size();
    if (out_seconds) *out_seconds = sec;
    return (long)n;
}

int ff_scene_commit(ff_scene* scene, int builder, double* out_build_ms) {
    FF_REQUIRE(scene, "ff_scene_commit: NULL scene");
    FF_REQUIRE(scene->ctx, "ff_scene_commit: the scene's context was destroyed");
    FF_REQUIRE(builder == FF_BUILDER_LBVH || builder == FF_BUILDER_SAH, "ff_scene_commit: unknown builder");
    ff_context* ctx = scene->ctx;
    FF_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    FF_CUDA(cudaStreamSynchronize(st));
    free_scene_device(scene);

    // flat numbering: ord = position in the reference's scan over (geometry, triangle), planes take one slot
    size_t n_tris = 0, n_planes = 0;
    for (auto& g : scene->geoms) { if (g.type == 2) n_tris += g.xyz9.size() / 9; else ++n_planes; }
    FF_REQUIRE(n_tris + n_planes < 0x3fffffffu, "ff_scene_commit: too many primitives");
    std::vector<int> ord_geom, ord_tri;
    ord_geom.reserve(n_tris + n_planes); ord_tri.reserve(n_tris + n_planes);
    std::vector<uint32_t> geom_of(n_tris), ord_of(n_tris);
    std::vector<M4> models(scene->geoms.size());
    size_t n_rows = scene->geoms.size();
    for (auto& g : scene->geoms) n_rows += g.face_materials.size();
    std::vector<float4> materials(4 * n_rows);
    auto put_material = [&](size_t row, const ff_material& m) {
        float type_bits; int ty = m.type; memcpy(&type_bits, &ty, 4);
        materials[4 * row] = make_float4(type_bits, m.albedo[0], m.albedo[1], m.albedo[2]);
        materials[4 * row + 1] = make_float4(m.emissive[0], m.emissive[1], m.emissive[2], m.intensity);
        materials[4 * row + 2] = make_float4(m.specular[0], m.specular[1], m.specular[2], m.ior > 0.f ? m.ior : 1.5f);
        materials[4 * row + 3] = make_float4(m.transmittance[0], m.transmittance[1], m.transmittance[2], 0.f);
    };
    std::vector<int4> geom_attr(scene->geoms.size(), make_int4(0, -1, -1, -1));
    std::vector<float4> attr_normals;
    std::vector<float2> attr_uvs;
    std::vector<uint32_t> attr_facemat;
    size_t next_row = scene->geoms.size();
    bool has_attr = false;
    std::vector<DevPlane> planes;
    std::vector<uint32_t> lights;  // flat indices of emitter-mesh triangles (next-event estimation samples these)
    float maxabs = 0.f;
    size_t flat = 0;
    for (size_t gi = 0; gi < scene->geoms.size(); ++gi) {
        const HostGeom& g = scene->geoms[gi];
        models[gi] = g.model;
        put_material(gi, g.mat);
        if (g.type == 2) {
            const size_t nt = g.xyz9.size() / 9;
            int4& ga = geom_attr[gi];
            ga.x = (int)flat;
            const size_t first_row = next_row;
            for (const ff_material& fm : g.face_materials) put_material(next_row++, fm);
            FF_REQUIRE(g.n9.empty() || g.n9.size() == g.xyz9.size(), "ff_scene_commit: normals do not match the triangle count");
            FF_REQUIRE(g.uv6.empty() || g.uv6.size() == nt * 6, "ff_scene_commit: uvs do not match the triangle count");
            FF_REQUIRE(g.face_mat.empty() || g.face_mat.size() == nt, "ff_scene_commit: face materials do not match the triangle count");
            if (!g.n9.empty() && nt) {
                // world-space corner normals: inverse(transpose(M)) * (n, 0), left un-normalised like kernel.cu:117
                has_attr = true;
                ga.y = (int)attr_normals.size();
                const M4 nm = ffh::inverted(ffh::transposed(g.model));
                for (size_t c = 0; c < 3 * nt; ++c) {
                    const ffm::V4 nw = ffh::matvec(nm, ffh::v4(g.n9[3 * c], g.n9[3 * c + 1], g.n9[3 * c + 2], 0));
                    attr_normals.push_back(make_float4(nw.x, nw.y, nw.z, 0.f));
                }
            }
            if (!g.uv6.empty() && nt) {
                has_attr = true;
                ga.z = (int)attr_uvs.size();
                for (size_t c = 0; c < 3 * nt; ++c) attr_uvs.push_back(make_float2(g.uv6[2 * c], g.uv6[2 * c + 1]));
            }
            if (!g.face_mat.empty() && nt) {
                has_attr = true;
                ga.w = (int)attr_facemat.size();
                for (size_t t = 0; t < nt; ++t) attr_facemat.push_back(g.face_mat[t] < 0 ? (uint32_t)gi : (uint32_t)(first_row + (size_t)g.face_mat[t]));
            }
            for (size_t t = 0; t < nt; ++t, ++flat) {
                geom_of[flat] = (uint32_t)gi; ord_of[flat] = (uint32_t)ord_geom.size();
                const int mt = g.face_mat.empty() || g.face_mat[t] < 0 ? g.mat.type : g.face_materials[(size_t)g.face_mat[t]].type;
                if (mt == FF_MAT_EMITTER) lights.push_back((uint32_t)flat);
                ord_geom.push_back((int)gi); ord_tri.push_back((int)t);
            }
        } else {
            DevPlane p;
            p.model = g.model; p.inv = g.inv;
            const ffm::V4 nw = ffh::matvec(ffh::inverted(ffh::transposed(g.model)), ffh::v4(0, 0, 1, 0));
            p.normal_world[0] = nw.x; p.normal_world[1] = nw.y; p.normal_world[2] = nw.z; p.normal_world[3] = 0;
            p.nrm = ffh::inverted(ffh::transposed(g.model));
            p.ord = (int)ord_geom.size(); p.geom = (int)gi; p.kind = g.type; p.radius = g.radius;
            planes.push_back(p);
            ord_geom.push_back((int)gi); ord_tri.push_back(-1);
        }
    }

    FF_CUDA(upload(&scene->d_ord_geom, ord_geom, st));
    FF_CUDA(upload(&scene->d_ord_tri, ord_tri, st));
    FF_CUDA(upload(&scene->d_materials, materials, st));
    FF_CUDA(upload(&scene->d_planes, planes, st));
    FF_CUDA(upload(&scene->d_lights, lights, st));
    FF_REQUIRE(attr_normals.size() < 0x7fffffffu && attr_uvs.size() < 0x7fffffffu, "ff_scene_commit: too many attribute corners");
    FF_CUDA(upload(&scene->d_geom_attr, geom_attr, st));
    FF_CUDA(upload(&scene->d_attr_normals, attr_normals, st));
    FF_CUDA(upload(&scene->d_attr_uvs, attr_uvs, st));
    FF_CUDA(upload(&scene->d_attr_facemat, attr_facemat, st));

    BuildOut bo; bo.n_nodes = 0; bo.depth = 0; bo.sah_cost = 0; bo.morton_bits = 0;
    float blo[3] = {0, 0, 0}, bhi[3] = {0, 0, 0};
    double build_ms = 0;
    uint32_t launches = 0;
    if (n_tris) {
        // stage object-space triangles, bake on the device (same single-rounding op order as the reference's M * v)
        DevTmp t_obj, t_models, t_geom_of, t_ord_of;  // released on every exit path
        FF_CUDA(t_obj.alloc(sizeof(float) * 9 * n_tris));
        FF_CUDA(t_models.alloc(sizeof(M4) * models.size()));
        FF_CUDA(t_geom_of.alloc(sizeof(uint32_t) * n_tris));
        FF_CUDA(t_ord_of.alloc(sizeof(uint32_t) * n_tris));
        float* d_obj = t_obj.as<float>();
        size_t off = 0;
        for (auto& g : scene->geoms) {
            if (g.type != 2 || g.xyz9.empty()) continue;
            FF_CUDA(cudaMemcpyAsync(d_obj + off, g.xyz9.data(), sizeof(float) * g.xyz9.size(), cudaMemcpyHostToDevice, st));
            off += g.xyz9.size();
        }
        FF_CUDA(cudaMemcpyAsync(t_models.p, models.data(), sizeof(M4) * models.size(), cudaMemcpyHostToDevice, st));
        FF_CUDA(cudaMemcpyAsync(t_geom_of.p, geom_of.data(), sizeof(uint32_t) * n_tris, cudaMemcpyHostToDevice, st));
        FF_CUDA(cudaMemcpyAsync(t_ord_of.p, ord_of.data(), sizeof(uint32_t) * n_tris, cudaMemcpyHostToDevice, st));
        FF_CUDA(cudaMalloc((void**)&scene->d_flat, sizeof(float4) * 3 * n_tris));
        FF_CUDA(cudaMalloc((void**)&scene->d_tris, sizeof(float4) * 3 * n_tris));
        const uint32_t node_cap = (uint32_t)(n_tris / 2 + 64);
        FF_CUDA(cudaMalloc((void**)&scene->d_nodes, sizeof(float4) * FF_NODE_F4 * (size_t)node_cap));
        FF_CUDA(cudaMalloc((void**)&scene->d_bvh2_children, sizeof(int2) * std::max<size_t>(1, n_tris)));
        FF_CUDA(cudaMalloc((void**)&scene->d_bvh2_lo, sizeof(float4) * 2 * n_tris));
        FF_CUDA(cudaMalloc((void**)&scene->d_bvh2_hi, sizeof(float4) * 2 * n_tris));
        FF_CUDA(cudaMalloc((void**)&scene->d_leaf_slot, sizeof(uint32_t) * n_tris));
        if (!scene->d_bounds) FF_CUDA(cudaMalloc((void**)&scene->d_bounds, 12 * sizeof(int)));
        k_bake_triangles(d_obj, t_models.as<M4>(), t_geom_of.as<uint32_t>(), t_ord_of.as<uint32_t>(), (uint32_t)n_tris, scene->d_flat, scene->d_bounds, st);
        FF_CUDA(cudaStreamSynchronize(st));
        t_obj.alloc(0); t_models.alloc(0); t_geom_of.alloc(0); t_ord_of.alloc(0);  // give the staging back before the build
        // the conservative pad needs the scene's largest |coordinate|: take it from the host copies' model-space
        // bounds pushed through |M| (cheap upper bound), refined by the device bounds after the build
        for (auto& g : scene->geoms) {
            if (g.type != 2) continue;
            float lo[3] = {1e30f, 1e30f, 1e30f}, hi[3] = {-1e30f, -1e30f, -1e30f};
            for (size_t k = 0; k < g.xyz9.size(); ++k) { const int a = (int)(k % 3); lo[a] = std::min(lo[a], g.xyz9[k]); hi[a] = std::max(hi[a], g.xyz9[k]); }
            for (int r = 0; r < 3; ++r) {
                float ext = std::fabs(ffh::at(g.model, 3, r));
                for (int c = 0; c < 3; ++c) ext += std::fabs(ffh::at(g.model, c, r)) * std::max(std::fabs(lo[c]), std::fabs(hi[c]));
                maxabs = std::max(maxabs, ext);
            }
        }
        const float pad = maxabs * (1.0f / 65536.0f);
        scene->dev.pad = pad;
        FF_CUDA(cudaEventRecord(ctx->ev[0], st));
        cudaError_t be = bvh_build(scene->d_flat, (uint32_t)n_tris, builder, pad, scene->d_tris, scene->d_nodes, node_cap,
                                   scene->d_bvh2_children, scene->d_bvh2_lo, scene->d_bvh2_hi, scene->d_leaf_slot, blo, bhi, &bo,
                                   &launches, ctx->arena, st, scene->d_bounds);
        if (be != cudaSuccess) { set_error(std::string("bvh_build: ") + cudaGetErrorString(be)); cudaGetLastError(); return FF_ERR_CUDA; }
        FF_CUDA(cudaEventRecord(ctx->ev[1], st));
        FF_CUDA(cudaEventSynchronize(ctx->ev[1]));
        float ms = 0;
        FF_CUDA(cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]));
        build_ms = ms;
        FF_REQUIRE(2 * bo.depth <= FF_MAX_WIDE_DEPTH, "ff_scene_commit: hierarchy deeper than the traversal stack (postponed triangle groups share it)");
    }
    FF_CUDA(cudaStreamSynchronize(st));

    DevScene& d = scene->dev;
    d.tris = scene->d_tris; d.nodes = scene->d_nodes; d.ord_geom = scene->d_ord_geom; d.ord_tri = scene->d_ord_tri;
    d.materials = scene->d_materials; d.planes = scene->d_planes; d.flat = scene->d_flat; d.lights = scene->d_lights;
    d.geom_attr = scene->d_geom_attr; d.attr_normals = scene->d_attr_normals; d.attr_uvs = scene->d_attr_uvs;
    d.attr_facemat = scene->d_attr_facemat; d.has_attr = has_attr ? 1u : 0u;
    d.n_lights = (uint32_t)lights.size(); d.bvh2_children = scene->d_bvh2_children;
    d.bvh2_lo = scene->d_bvh2_lo; d.bvh2_hi = scene->d_bvh2_hi; d.bvh2_leaf_slot = scene->d_leaf_slot;
    d.n_tris = (uint32_t)n_tris; d.n_nodes = bo.n_nodes; d.n_planes = (uint32_t)n_planes; d.n_geoms = (uint32_t)scene->geoms.size();
    d.wide_depth = bo.depth;
    d.prmt_magic = 0x3F800000u;
    d.nodes_tex = d.tris_tex = 0;
#if FF_NODE_TEX || FF_TRI_TEX
    {   // linear float4 texture views for the texture-path kernel variants
        auto make_tex = [](const void* p, size_t bytes, cudaTextureObject_t* out) -> cudaError_t {
            cudaResourceDesc rd; memset(&rd, 0, sizeof(rd));
            rd.resType = cudaResourceTypeLinear; rd.res.linear.devPtr = const_cast<void*>(p);
            rd.res.linear.desc = cudaCreateChannelDesc<float4>(); rd.res.linear.sizeInBytes = bytes;
            cudaTextureDesc td; memset(&td, 0, sizeof(td)); td.readMode = cudaReadModeElementType;
            return cudaCreateTextureObject(out, &rd, &td, nullptr);
        };
        if (bo.n_nodes) FF_CUDA(make_tex(scene->d_nodes, sizeof(float4) * FF_NODE_F4 * (size_t)bo.n_nodes, &scene->nodes_tex));
        if (n_tris) FF_CUDA(make_tex(scene->d_tris, sizeof(float4) * 3 * n_tris, &scene->tris_tex));
        d.nodes_tex = scene->nodes_tex; d.tris_tex = scene->tris_tex;
    }
#endif
    d.postpone_div = (uint32_t)tuning().postpone_div;
    // refill threshold of the persistent kernels: short rays (small hierarchies: C2 visits 5 nodes per ray) finish so often
    // that refilling at 28 live lanes spends 10 % of the issue slots in the fetch path at 11 active lanes; measured optimum
    // 16-20 lanes on C2 (+4 %), 28 on the 1M-triangle soup (profiles/r01g_tune_refill.txt).  ff_debug_set_tuning overrides.
    d.refill_lanes = tuning().refill_lanes > 0 ? (uint32_t)tuning().refill_lanes : (bo.n_nodes <= 4096u ? 20u : 28u);
    for (int k = 0; k < 3; ++k) { d.bounds_lo[k] = blo[k]; d.bounds_hi[k] = bhi[k]; }

    ff_scene_info& in = scene->info;
    memset(&in, 0, sizeof(in));
    in.n_triangles = n_tris; in.n_geometries = scene->geoms.size(); in.n_planes = n_planes; in.n_nodes = bo.n_nodes;
    in.node_bytes = 16ull * FF_NODE_F4 * bo.n_nodes; in.triangle_bytes = 48ull * n_tris; in.depth = bo.depth; in.builder = (uint32_t)builder;
    for (int k = 0; k < 3; ++k) { in.bounds_lo[k] = blo[k]; in.bounds_hi[k] = bhi[k]; }
    in.build_ms = build_ms; in.sah_cost = bo.sah_cost; in.morton_bits = bo.morton_bits;
    scene->committed = true;
    if (out_build_ms) *out_build_ms = build_ms;
    return FF_OK;
}
int ff_scene_get_info(const ff_scene* scene, ff_scene_info* out) {
    FF_REQUIRE(scene && out, "ff_scene_get_info: NULL argument");
    FF_REQUIRE(scene->committed, "ff_scene_get_info: scene not committed");
    *out = scene->info;
    return FF_OK;
}
int ff_scene_get_flat(const ff_scene* scene, float* xyz9, int* geom, int* tri, size_t capacity_tris) {
    FF_REQUIRE(scene && scene->committed, "ff_scene_get_flat: scene not committed");
    const size_t n = std::min<size_t>(scene->dev.n_tris, capacity_tris);
    if (n == 0) return FF_OK;
    FF_CUDA(cudaSetDevice(scene->ctx->device));
    std::vector<float4> h(3 * n);
    FF_CUDA(cudaMemcpy(h.data(), scene->d_flat, sizeof(float4) * 3 * n, cudaMemcpyDeviceToHost));
    std::vector<int> og(scene->info.n_triangles + scene->info.n_planes), ot(og.size());
    FF_CUDA(cudaMemcpy(og.data(), scene->d_ord_geom, sizeof(int) * og.size(), cudaMemcpyDeviceToHost));
    FF_CUDA(cudaMemcpy(ot.data(), scene->d_ord_tri, sizeof(int) * ot.size(), cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < n; ++i) {
        if (xyz9)
            for (int v = 0; v < 3; ++v) { xyz9[9 * i + 3 * v] = h[3 * i + v].x; xyz9[9 * i + 3 * v + 1] = h[3 * i + v].y; xyz9[9 * i + 3 * v + 2] = h[3 * i + v].z; }
        int ord; memcpy(&ord, &h[3 * i].w, 4);
        if (geom) geom[i] = og[ord];
        if (tri) tri[i] = ot[ord];
    }
    return FF_OK;
}
int ff_scene_get_matrices(const ff_scene* scene, int gi, float model16[16], float inverse16[16]) {
    FF_REQUIRE(scene && gi >= 0 && gi < (int)scene->geoms.size(), "ff_scene_get_matrices: bad geometry index");
    if (model16) memcpy(model16, &scene->geoms[gi].model, 64);
    if (inverse16) memcpy(inverse16, &scene->geoms[gi].inv, 64);
    return FF_OK;
}

}  // extern "C"

// ================================================================================================ trace
namespace {
struct ScopedTimer {  // CUDA-event bracket on the context's stream
    ff_context* c; int a, b;
    ScopedTimer(ff_context* c_, int a_, int b_) : c(c_), a(a_), b(b_) { cudaEventRecord(c->ev[a], c->stream); }
    double stop() {
        cudaEventRecord(c->ev[b], c->stream);
        cudaEventSynchronize(c->ev[b]);
        float ms = 0;
        cudaEventElapsedTime(&ms, c->ev[a], c->ev[b]);
        return ms;
    }
};
int check_scene(ff_scene* scene, const char* who) {
    if (!scene) { set_error(std::string(who) + ": NULL scene"); return FF_ERR_INVALID; }
    if (!scene->ctx) { set_error(std::string(who) + ": the scene's context was destroyed"); return FF_ERR_INVALID; }
    if (!scene->committed) { set_error(std::string(who) + ": scene not committed"); return FF_ERR_INVALID; }
    cudaError_t e = cudaSetDevice(scene->ctx->device);
    if (e != cudaSuccess) { set_error(cudaGetErrorString(e)); return FF_ERR_CUDA; }
    return FF_OK;
}
int ensure_counters(ff_scene* s) {
    if (!s->d_trace_counters) {
        FF_CUDA(cudaMalloc((void**)&s->d_trace_counters, sizeof(TraceCounters)));
        FF_CUDA(cudaMalloc((void**)&s->d_work_counter, sizeof(uint32_t)));
        FF_CUDA(cudaMalloc((void**)&s->d_ray_counter, 2 * sizeof(unsigned long long)));  // closest-hit rays, shadow rays
    }
    return FF_OK;
}
bool valid_mode(int mode) { const int m = mode & 0xff; if (mode & ~(0xff | FF_TRACE_COUNT | FF_TRACE_STATIC)) return false; return m == FF_TRACE_BVH8 || m == FF_TRACE_BRUTE || m == FF_TRACE_BVH2; }
}  // namespace

extern "C" {

int ff_trace_closest_device(ff_scene* scene, const void* d_rays, size_t n, int mode, void* d_hits, ff_stats* stats) {
    int rc = check_scene(scene, "ff_trace_closest_device");
    if (rc) return rc;
    FF_REQUIRE(valid_mode(mode), "ff_trace_closest_device: unknown mode");
    FF_REQUIRE(n < 0xffffffffull, "ff_trace_closest_device: batch too large");
    FF_REQUIRE((d_rays && d_hits) || n == 0, "ff_trace_closest_device: NULL buffer");
    ff_context* c = scene->ctx;
    rc = ensure_counters(scene);
    if (rc) return rc;
    const bool count = (mode & FF_TRACE_COUNT) != 0;
    if (count) FF_CUDA(cudaMemsetAsync(scene->d_trace_counters, 0, sizeof(TraceCounters), c->stream));
    const float4* ro = (const float4*)d_rays;
    FF_CUDA(cudaMemsetAsync(scene->d_work_counter, 0, sizeof(uint32_t), c->stream));
    ScopedTimer t(c, 0, 1);
    k_trace_closest(scene->dev, ro, ro + n, nullptr, (uint32_t)n, mode, (HitRec*)d_hits, scene->d_trace_counters, scene->d_work_counter, c->stream);
    const double ms = t.stop();
    FF_CUDA(cudaGetLastError());
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        stats->gpu_ms = ms; stats->trace_ms = ms; stats->rays = n; stats->kernel_launches = n ? 1 : 0;
        if (count) {
            TraceCounters h;
            FF_CUDA(cudaMemcpy(&h, scene->d_trace_counters, sizeof(h), cudaMemcpyDeviceToHost));
            stats->nodes_visited = h.nodes; stats->tris_tested = h.tris;
        }
    }
    return FF_OK;
}

int ff_hits_unpack(ff_scene* scene, const void* d_hits, size_t n, int* out_geom, int* out_tri, float* out_t) {
    int rc = check_scene(scene, "ff_hits_unpack");
    if (rc) return rc;
    if (n == 0) return FF_OK;
    cudaStream_t st = scene->ctx->stream;
    DevTmp tg, tt, tk;
    FF_CUDA(tg.alloc(sizeof(int) * n)); FF_CUDA(tt.alloc(sizeof(int) * n)); FF_CUDA(tk.alloc(sizeof(float) * n));
    int *dg = tg.as<int>(), *dt = tt.as<int>(); float* dk = tk.as<float>();
    k_unpack_hits(scene->dev, (const HitRec*)d_hits, n, dg, dt, dk, st);
    if (out_geom) FF_CUDA(cudaMemcpyAsync(out_geom, dg, sizeof(int) * n, cudaMemcpyDeviceToHost, st));
    if (out_tri) FF_CUDA(cudaMemcpyAsync(out_tri, dt, sizeof(int) * n, cudaMemcpyDeviceToHost, st));
    if (out_t) FF_CUDA(cudaMemcpyAsync(out_t, dk, sizeof(float) * n, cudaMemcpyDeviceToHost, st));
    FF_CUDA(cudaStreamSynchronize(st));
    return FF_OK;
}

int ff_trace_closest(ff_scene* scene, const float* rays6, size_t n, int mode, int* out_geom, int* out_tri, float* out_t, ff_stats* stats) {
    return ff_trace_closest_uv(scene, rays6, n, mode, out_geom, out_tri, out_t, nullptr, nullptr, stats);
}
// Layout of one pipeline slot (bytes per ray): rays6 24 | packed o,d 32 | hit record 16 | geom 4 | tri 4 | t 4 | u,v 8.
static const size_t kPipeBytesPerRay = 24 + 32 + 16 + 4 + 4 + 4 + 8;
static int ensure_pipe(ff_context* c, size_t chunk) {
    TracePipe& tp = c->pipe;
    if (!tp.ready) {
        FF_CUDA(cudaStreamCreateWithFlags(&tp.s_in, cudaStreamNonBlocking));
        FF_CUDA(cudaStreamCreateWithFlags(&tp.s_out, cudaStreamNonBlocking));
        FF_CUDA(cudaStreamCreateWithFlags(&tp.s_comp, cudaStreamNonBlocking));
        for (int k = 0; k < FF_PIPE_SLOTS; ++k) {
            FF_CUDA(cudaEventCreateWithFlags(&tp.in_done[k], cudaEventDisableTiming));
            FF_CUDA(cudaEventCreateWithFlags(&tp.comp_done[k], cudaEventDisableTiming));
            FF_CUDA(cudaEventCreateWithFlags(&tp.out_done[k], cudaEventDisableTiming));
        }
        for (int k = 0; k < 5; ++k) FF_CUDA(cudaEventCreate(&tp.span[k]));
        FF_CUDA(cudaMalloc((void**)&tp.work, sizeof(uint32_t) * FF_PIPE_SLOTS));
        FF_CUDA(cudaMalloc((void**)&tp.d_sync, sizeof(StreamSync)));
        FF_CUDA(cudaHostAlloc((void**)&tp.h_flags, sizeof(uint32_t) * (FF_STREAM_MAX_CHUNKS + 1), cudaHostAllocPortable | cudaHostAllocMapped));
        FF_CUDA(cudaHostAlloc((void**)&tp.h_marks, sizeof(uint32_t) * (FF_STREAM_MAX_CHUNKS + 8), cudaHostAllocPortable));
        FF_CUDA(cudaEventCreateWithFlags(&tp.s_go, cudaEventDisableTiming));
        tp.ready = true;
    }
    chunk = (chunk + 3) & ~(size_t)3;  // keeps every sub-array of a slot 16-byte aligned
    if (tp.chunk < chunk) {
        for (int k = 0; k < FF_PIPE_SLOTS; ++k) { cudaFree(tp.block[k]); tp.block[k] = nullptr; }
        tp.chunk = 0;
        for (int k = 0; k < FF_PIPE_SLOTS; ++k) FF_CUDA(cudaMalloc((void**)&tp.block[k], kPipeBytesPerRay * chunk + 1024));
        tp.chunk = chunk;
    }
    return FF_OK;
}
// Streaming path of ff_trace_closest: the whole batch (in super-batches of <= kStreamSuper rays) is traced by ONE persistent
// launch that starts as soon as the first chunk has landed, draws rays as the copy-in stream delivers them and hands finished
// chunks to the copy-out stream -- so the batch pays for one kernel tail instead of one per chunk (the chunked pipeline below
// lost ~1.2 ms of an 8.3 M-ray batch to eight tails).  Kernel side: ff_trace.cu "Host batches".
static const size_t kStreamSuper = (size_t)1 << 24;
static const size_t kStreamBytesPerRay = 24 + 4 + 4 + 4;
static const int kStreamFellBack = 1;   // (positive: not an ff_status) the streaming launch timed out, nothing was delivered
static int trace_closest_stream(ff_scene* scene, const float* rays6, size_t n, int* out_geom, int* out_tri, float* out_t, ff_stats* stats) {
    ff_context* c = scene->ctx;
    int rc = ensure_pipe(c, 0);
    if (rc) return rc;
    TracePipe& tp = c->pipe;
    cudaStream_t st = c->stream;
    static const uint32_t shift_pref = [] {
        const char* e = getenv("FF_STREAM_CHUNK_SHIFT"); const long v = e ? atol(e) : 0; return (uint32_t)(v >= 12 && v <= 24 ? v : 19); }();
    const size_t super = std::min(n, kStreamSuper);
    if (tp.scap < super) {
        FF_CUDA(cudaStreamSynchronize(st));
        cudaFree(tp.sblock); tp.sblock = nullptr; tp.scap = 0;
        const size_t cap = (super + 63) & ~(size_t)63;   // keeps the four sub-arrays 256-byte aligned
        FF_CUDA(cudaMalloc((void**)&tp.sblock, kStreamBytesPerRay * cap));
        tp.scap = cap;
    }
    float* d6 = (float*)tp.sblock;
    int* dg = (int*)(tp.sblock + 24 * tp.scap);
    int* dt = dg + tp.scap;
    float* dk = (float*)(dt + tp.scap);
    uint32_t* d_flags = nullptr;
    FF_CUDA(cudaHostGetDevicePointer((void**)&d_flags, tp.h_flags, 0));
    while (c->trace_events.size() < 2) { cudaEvent_t e; FF_CUDA(cudaEventCreate(&e)); c->trace_events.push_back(e); }
    FF_CUDA(cudaEventRecord(tp.span[0], st));
    double trace_ms = 0, h2d_ms = 0, d2h_ms = 0;
    uint32_t launches = 0;
    for (size_t first = 0; first < n; first += kStreamSuper) {
        const size_t m = std::min(kStreamSuper, n - first);
        uint32_t shift = shift_pref;
        while (((m + ((size_t)1 << shift) - 1) >> shift) > FF_STREAM_MAX_CHUNKS) ++shift;
        const size_t chunk = (size_t)1 << shift, n_chunks = (m + chunk - 1) >> shift;
        for (size_t k = 0; k < n_chunks; ++k) tp.h_marks[k] = (uint32_t)std::min(m, (k + 1) << shift);
        // completion (copy-out) chunks may be finer than the copy-in chunks: the only copy-out nothing overlaps is the last one
        static const uint32_t dshift_pref = [] { const char* e = getenv("FF_STREAM_DONE_SHIFT"); const long v = e ? atol(e) : 0; return (uint32_t)(v >= 12 && v <= 24 ? v : 0); }();
        uint32_t dshift = dshift_pref ? std::min(dshift_pref, shift) : shift;
        while (((m + ((size_t)1 << dshift) - 1) >> dshift) > FF_STREAM_MAX_CHUNKS) ++dshift;
        const size_t dchunk = (size_t)1 << dshift, d_chunks = (m + dchunk - 1) >> dshift;
        for (size_t k = 0; k < d_chunks; ++k) tp.h_flags[k] = 0u;
        tp.h_flags[FF_STREAM_MAX_CHUNKS] = 0u;   // raised by a lane that timed out
        // zero the shared words, then let the copy streams go (a super-batch reuses the staging of the one before it)
        FF_CUDA(cudaMemsetAsync(tp.d_sync, 0, sizeof(StreamSync), st));
        FF_CUDA(cudaMemsetAsync(tp.work, 0, sizeof(uint32_t), st));
        FF_CUDA(cudaEventRecord(tp.s_go, st));
        FF_CUDA(cudaStreamWaitEvent(tp.s_in, tp.s_go, 0));
        FF_CUDA(cudaStreamWaitEvent(tp.s_out, tp.s_go, 0));
        if (first == 0) FF_CUDA(cudaEventRecord(tp.span[2], tp.s_out));
        auto copy_in = [&](size_t k) -> cudaError_t {
            const size_t a = k << shift, cnt = std::min(chunk, m - a);
            cudaError_t e = cudaMemcpyAsync(d6 + 6 * a, rays6 + 6 * (first + a), sizeof(float) * 6 * cnt, cudaMemcpyHostToDevice, tp.s_in);
            if (e != cudaSuccess) return e;
            return cudaMemcpyAsync(&tp.d_sync->avail, tp.h_marks + k, sizeof(uint32_t), cudaMemcpyHostToDevice, tp.s_in);
        };
        static const bool preload = [] { const char* e = getenv("FF_STREAM_PRELOAD"); return e && atoi(e) != 0; }();
        if (preload) {   // probe: every ray lands before the launch, so the launch's time is the kernel's own cost without copies in
            for (size_t k = 0; k < n_chunks; ++k) FF_CUDA(copy_in(k));
            FF_CUDA(cudaStreamSynchronize(tp.s_in));
        }
        // the launch goes first: its latency (and the first warps' spin on `avail`) overlaps the host's queueing of the copies
        FF_CUDA(cudaEventRecord(c->trace_events[0], st));
        k_trace_closest_stream(scene->dev, d6, (uint32_t)m, dshift, tp.d_sync, d_flags, dg, dt, dk, tp.work, st);
        FF_CUDA(cudaEventRecord(c->trace_events[1], st));
        if (!preload) {
            {   // the first chunk lands in growing pieces (32 Ki, 32 Ki, 64 Ki, ... rays), so the kernel's first warps start after ~20 us of copy
                const size_t c0 = std::min(chunk, m);
                size_t a = 0, piece = (size_t)1 << 15;
                for (int j = 0; a < c0; ++j) {
                    const size_t cnt = (j == 7 || a + piece > c0) ? c0 - a : piece;
                    FF_CUDA(cudaMemcpyAsync(d6 + 6 * a, rays6 + 6 * (first + a), sizeof(float) * 6 * cnt, cudaMemcpyHostToDevice, tp.s_in));
                    a += cnt;
                    tp.h_marks[FF_STREAM_MAX_CHUNKS + j] = (uint32_t)a;
                    FF_CUDA(cudaMemcpyAsync(&tp.d_sync->avail, tp.h_marks + FF_STREAM_MAX_CHUNKS + j, sizeof(uint32_t), cudaMemcpyHostToDevice, tp.s_in));
                    if (j > 0) piece <<= 1;
                }
            }
            for (size_t k = 1; k < n_chunks; ++k) FF_CUDA(copy_in(k));
        }
        if (first + m == n) FF_CUDA(cudaEventRecord(tp.span[1], tp.s_in));
        // hand finished chunks to the copy-out stream in order; the kernel cannot outlive its own time-out, so neither can this loop
        size_t next = 0;
        bool kernel_over = false;
        while (next < d_chunks) {
            if (*(volatile uint32_t*)(tp.h_flags + next) == 0u) {
                if (kernel_over) break;   // the kernel ended without completing the chunk: it timed out (err is read below)
                kernel_over = cudaEventQuery(c->trace_events[1]) != cudaErrorNotReady;
                if (!kernel_over) (void)cudaGetLastError();   // cudaErrorNotReady is recorded as the thread's last error: clear it
                continue;                 // (one more look at the flag after the kernel is known to be over)
            }
            const size_t a = next << dshift, cnt = std::min(dchunk, m - a);
            if (out_geom) FF_CUDA(cudaMemcpyAsync(out_geom + first + a, dg + a, sizeof(int) * cnt, cudaMemcpyDeviceToHost, tp.s_out));
            if (out_tri) FF_CUDA(cudaMemcpyAsync(out_tri + first + a, dt + a, sizeof(int) * cnt, cudaMemcpyDeviceToHost, tp.s_out));
            if (out_t) FF_CUDA(cudaMemcpyAsync(out_t + first + a, dk + a, sizeof(float) * cnt, cudaMemcpyDeviceToHost, tp.s_out));
            ++next;
        }
        FF_CUDA(cudaEventRecord(tp.span[3], tp.s_out));
        FF_CUDA(cudaStreamWaitEvent(st, tp.span[3], 0));   // the context's stream (and an open ff_timer) ends after the last copy
        FF_CUDA(cudaStreamWaitEvent(st, tp.span[1], 0));
        FF_CUDA(cudaEventRecord(tp.span[4], st));
        FF_CUDA(cudaEventSynchronize(tp.span[4]));
        FF_CUDA(cudaStreamSynchronize(tp.s_in));
        FF_CUDA(cudaGetLastError());
        if (tp.h_flags[FF_STREAM_MAX_CHUNKS] || next < d_chunks) {
            // The kernel gave up waiting for rays: something serialises kernels against copies (a profiler that replays
            // kernels, CUDA_LAUNCH_BLOCKING).  The caller redoes the batch through the chunked pipeline; after two such
            // calls the context stops trying.
            ++tp.stream_failures;
            return kStreamFellBack;
        }
        float ms = 0;
        cudaEventElapsedTime(&ms, c->trace_events[0], c->trace_events[1]); trace_ms += ms;
        ++launches;
    }
    if (stats) {
        float ms = 0;
        cudaEventElapsedTime(&ms, tp.span[0], tp.span[4]); stats->gpu_ms = ms;
        cudaEventElapsedTime(&ms, tp.span[0], tp.span[1]); h2d_ms = ms;
        cudaEventElapsedTime(&ms, tp.span[2], tp.span[3]); d2h_ms = ms;
        stats->h2d_ms = h2d_ms; stats->d2h_ms = d2h_ms; stats->trace_ms = trace_ms; stats->rays = n;
        stats->passes = launches; stats->kernel_launches = launches;
    }
    return FF_OK;
}

int ff_trace_closest_uv(ff_scene* scene, const float* rays6, size_t n, int mode, int* out_geom, int* out_tri, float* out_t,
                        float* out_u, float* out_v, ff_stats* stats) {
    int rc = check_scene(scene, "ff_trace_closest");
    if (rc) return rc;
    FF_REQUIRE(valid_mode(mode), "ff_trace_closest: unknown mode");
    FF_REQUIRE(rays6 || n == 0, "ff_trace_closest: NULL rays");
    if (stats) memset(stats, 0, sizeof(*stats));
    if (n == 0) return FF_OK;
    FF_REQUIRE(n < 0xffffffffull, "ff_trace_closest: batch too large");
    ff_context* c = scene->ctx;
    if ((rc = ensure_counters(scene))) return rc;
    static const bool stream_on = [] { const char* e = getenv("FF_TRACE_STREAM"); return !e || atoi(e) != 0; }();
    if (stream_on && c->pipe.stream_failures < 2 && (mode & 0xff) == FF_TRACE_BVH8 && !(mode & (FF_TRACE_COUNT | FF_TRACE_STATIC)) && !out_u && !out_v) {
        rc = trace_closest_stream(scene, rays6, n, out_geom, out_tri, out_t, stats);
        if (rc != kStreamFellBack) return rc;
        if (stats) memset(stats, 0, sizeof(*stats));
    }
    // ~1 Mi rays per chunk: large enough that the persistent kernel's tail is a few per cent of a chunk, small enough that
    // the first copy-in and the last copy-out (the only transfers nothing overlaps) are a small part of the call
    static const size_t chunk_pref = [] { const char* e = getenv("FF_TRACE_CHUNK"); const long v = e ? atol(e) : 0; return v > 0 ? (size_t)v : (size_t)1 << 20; }();
    const size_t chunk = std::min(n, chunk_pref);
    const size_t n_chunks = (n + chunk - 1) / chunk;
    if ((rc = ensure_pipe(c, chunk))) return rc;
    TracePipe& tp = c->pipe;
    cudaStream_t st = c->stream;
    const bool count = (mode & FF_TRACE_COUNT) != 0;
    const bool want_uv = out_u || out_v;
    if (count) FF_CUDA(cudaMemsetAsync(scene->d_trace_counters, 0, sizeof(TraceCounters), st));
    while (c->trace_events.size() < 2 * n_chunks) { cudaEvent_t e; FF_CUDA(cudaEventCreate(&e)); c->trace_events.push_back(e); }
    // everything below is ordered after what is already queued on the context's stream (so an ff_timer bracket sees it)
    FF_CUDA(cudaEventRecord(tp.span[0], st));
    FF_CUDA(cudaStreamWaitEvent(tp.s_in, tp.span[0], 0));
    FF_CUDA(cudaStreamWaitEvent(tp.s_out, tp.span[0], 0));
    FF_CUDA(cudaStreamWaitEvent(tp.s_comp, tp.span[0], 0));
    for (size_t k = 0; k < n_chunks; ++k) {
        // Chunks alternate between two compute streams: a persistent trace launch ends in a tail where a few warps finish
        // their last rays on an otherwise idle machine; with the next chunk's launch already queued on the other stream its
        // CTAs move in as the tail's retire, instead of waiting for the whole launch (a chunked batch cost 28 % more traversal
        // time than one launch when the chunks were serialised on one stream, profiles/r02a_bench.json).
        cudaStream_t cs = (k & 1) ? tp.s_comp : st;
        const int slot = (int)(k % FF_PIPE_SLOTS);
        const size_t first = k * chunk, m = std::min(chunk, n - first);
        char* b = tp.block[slot];
        float* d6 = (float*)b;
        float4* dr = (float4*)(b + 24 * tp.chunk);
        HitRec* dh = (HitRec*)(b + 56 * tp.chunk);
        int* dg = (int*)(b + 72 * tp.chunk);
        int* dt = dg + tp.chunk;
        float* dk = (float*)(dt + tp.chunk);
        float* duv = dk + tp.chunk;
        // copy in (the slot's previous chunk must have been packed: its compute stage is done)
        if (k >= FF_PIPE_SLOTS) FF_CUDA(cudaStreamWaitEvent(tp.s_in, tp.comp_done[slot], 0));
        FF_CUDA(cudaMemcpyAsync(d6, rays6 + 6 * first, sizeof(float) * 6 * m, cudaMemcpyHostToDevice, tp.s_in));
        FF_CUDA(cudaEventRecord(tp.in_done[slot], tp.s_in));
        if (k + 1 == n_chunks) FF_CUDA(cudaEventRecord(tp.span[1], tp.s_in));
        // pack + trace + unpack (the slot's previous outputs must have left the device)
        FF_CUDA(cudaStreamWaitEvent(cs, tp.in_done[slot], 0));
        if (k >= FF_PIPE_SLOTS) FF_CUDA(cudaStreamWaitEvent(cs, tp.out_done[slot], 0));
        k_pack_rays(d6, m, dr, dr + m, cs);
        FF_CUDA(cudaMemsetAsync(tp.work + slot, 0, sizeof(uint32_t), cs));
        FF_CUDA(cudaEventRecord(c->trace_events[2 * k], cs));
        k_trace_closest(scene->dev, dr, dr + m, nullptr, (uint32_t)m, mode, dh, scene->d_trace_counters, tp.work + slot, cs);
        FF_CUDA(cudaEventRecord(c->trace_events[2 * k + 1], cs));
        k_unpack_hits(scene->dev, dh, m, dg, dt, dk, cs);
        if (want_uv) k_hit_barycentrics(scene->dev, dr, dr + m, dh, m, duv, duv + m, cs);
        FF_CUDA(cudaEventRecord(tp.comp_done[slot], cs));
        // copy out
        FF_CUDA(cudaStreamWaitEvent(tp.s_out, tp.comp_done[slot], 0));
        if (k == 0) FF_CUDA(cudaEventRecord(tp.span[2], tp.s_out));
        if (out_geom) FF_CUDA(cudaMemcpyAsync(out_geom + first, dg, sizeof(int) * m, cudaMemcpyDeviceToHost, tp.s_out));
        if (out_tri) FF_CUDA(cudaMemcpyAsync(out_tri + first, dt, sizeof(int) * m, cudaMemcpyDeviceToHost, tp.s_out));
        if (out_t) FF_CUDA(cudaMemcpyAsync(out_t + first, dk, sizeof(float) * m, cudaMemcpyDeviceToHost, tp.s_out));
        if (out_u) FF_CUDA(cudaMemcpyAsync(out_u + first, duv, sizeof(float) * m, cudaMemcpyDeviceToHost, tp.s_out));
        if (out_v) FF_CUDA(cudaMemcpyAsync(out_v + first, duv + m, sizeof(float) * m, cudaMemcpyDeviceToHost, tp.s_out));
        FF_CUDA(cudaEventRecord(tp.out_done[slot], tp.s_out));
    }
    FF_CUDA(cudaEventRecord(tp.span[3], tp.s_out));
    FF_CUDA(cudaStreamWaitEvent(st, tp.span[3], 0));   // the context's stream (and an open ff_timer) ends after the last copy
    FF_CUDA(cudaEventRecord(tp.span[4], st));
    FF_CUDA(cudaEventSynchronize(tp.span[4]));
    FF_CUDA(cudaGetLastError());
    if (stats) {
        float ms = 0;
        cudaEventElapsedTime(&ms, tp.span[0], tp.span[4]); stats->gpu_ms = ms;
        cudaEventElapsedTime(&ms, tp.span[0], tp.span[1]); stats->h2d_ms = ms;   // span of the copy-in stream's activity
        cudaEventElapsedTime(&ms, tp.span[2], tp.span[3]); stats->d2h_ms = ms;   // span of the copy-out stream's activity
        // chunks overlap on the two compute streams: the traversal time is the span from the first launch to the end of the last
        cudaEventElapsedTime(&ms, c->trace_events[0], c->trace_events[2 * (n_chunks - 1) + 1]);
        if (n_chunks > 1) { float m2 = 0; cudaEventElapsedTime(&m2, c->trace_events[0], c->trace_events[2 * (n_chunks - 2) + 1]); ms = std::max(ms, m2); }
        stats->trace_ms = ms; stats->rays = n; stats->passes = (uint32_t)n_chunks;
        stats->kernel_launches = (uint32_t)(n_chunks * (want_uv ? 4 : 3));
        if (count) {
            TraceCounters h;
            FF_CUDA(cudaMemcpy(&h, scene->d_trace_counters, sizeof(h), cudaMemcpyDeviceToHost));
            stats->nodes_visited = h.nodes; stats->tris_tested = h.tris;
        }
    }
    return FF_OK;
}

int ff_trace_any(ff_scene* scene, const float* rays6, const float* tmax, size_t n, int mode, unsigned char* out_occluded, ff_stats* stats) {
    int rc = check_scene(scene, "ff_trace_any");
    if (rc) return rc;
    FF_REQUIRE(valid_mode(mode), "ff_trace_any: unknown mode");
    FF_REQUIRE((rays6 && out_occluded) || n == 0, "ff_trace_any: NULL buffer");
    if (stats) memset(stats, 0, sizeof(*stats));
    if (n == 0) return FF_OK;
    ff_context* c = scene->ctx;
    cudaStream_t st = c->stream;
    FF_REQUIRE(n < 0xffffffffull, "ff_trace_any: batch too large");
    DevTmp t6, tr, ttm, tocc;  // released on every exit path
    FF_CUDA(t6.alloc(sizeof(float) * 6 * n));
    FF_CUDA(tr.alloc(sizeof(float4) * 2 * n));
    FF_CUDA(tocc.alloc(n));
    float* d6 = t6.as<float>(); float4* dr = tr.as<float4>(); unsigned char* docc = tocc.as<unsigned char>();
    if (tmax) { FF_CUDA(ttm.alloc(sizeof(float) * n)); FF_CUDA(cudaMemcpyAsync(ttm.p, tmax, sizeof(float) * n, cudaMemcpyHostToDevice, st)); }
    float* dtm = ttm.as<float>();
    FF_CUDA(cudaMemcpyAsync(d6, rays6, sizeof(float) * 6 * n, cudaMemcpyHostToDevice, st));
    k_pack_rays(d6, n, dr, dr + n, st);
    if ((rc = ensure_counters(scene))) return rc;
    FF_CUDA(cudaMemsetAsync(scene->d_work_counter, 0, sizeof(uint32_t), st));
    ScopedTimer t(c, 0, 1);
    k_trace_any(scene->dev, dr, dr + n, dtm, (uint32_t)n, mode, docc, scene->d_work_counter, st);
    const double ms = t.stop();
    FF_CUDA(cudaGetLastError());
    FF_CUDA(cudaMemcpy(out_occluded, docc, n, cudaMemcpyDeviceToHost));
    if (stats) { stats->gpu_ms = ms; stats->trace_ms = ms; stats->shadow_rays = n; stats->kernel_launches = 2; }
    return FF_OK;
}

// ================================================================================================ camera
static CamDev make_cam(const ff_camera* cam) {
    CamDev d;
    V3 f, r, u;
    ffh::camera_basis(cam->world_up, cam->yaw_deg, cam->pitch_deg, f, r, u);
    const V3 pos = ffm::v3(cam->pos[0], cam->pos[1], cam->pos[2]);
    const M4 iv = ffh::inverted(ffh::look_at_rh(pos, ffm::add(pos, f), u));
    const M4 ip = ffh::inverted(ffh::perspective_fov_rh_no(ffh::to_radians(cam->fov_deg), (float)cam->width, (float)cam->height,
                                                           cam->near_clip, cam->far_clip));
    d.inv_vp = ffh::matmul(iv, ip);
    d.pos[0] = cam->pos[0]; d.pos[1] = cam->pos[1]; d.pos[2] = cam->pos[2];
    d.width = (float)cam->width; d.height = (float)cam->height; d.far_clip = cam->far_clip;
    d.w = cam->width; d.h = cam->height;
    return d;
}
static int check_cam(const ff_camera* cam, const char* who) {
    if (!cam) { set_error(std::string(who) + ": NULL camera"); return FF_ERR_INVALID; }
    if (cam->width <= 0 || cam->height <= 0 || (uint64_t)cam->width * (uint64_t)cam->height > 0x7fffffffull) {
        set_error(std::string(who) + ": bad image size"); return FF_ERR_INVALID;
    }
    return FF_OK;
}
int ff_camera_basis(const ff_camera* cam, float forward[3], float right[3], float up[3]) {
    FF_REQUIRE(cam, "ff_camera_basis: NULL camera");
    V3 f, r, u;
    ffh::camera_basis(cam->world_up, cam->yaw_deg, cam->pitch_deg, f, r, u);
    if (forward) { forward[0] = f.x; forward[1] = f.y; forward[2] = f.z; }
    if (right) { right[0] = r.x; right[1] = r.y; right[2] = r.z; }
    if (up) { up[0] = u.x; up[1] = u.y; up[2] = u.z; }
    return FF_OK;
}
int ff_camera_inverse_matrices(const ff_camera* cam, float inv_view16[16], float inv_proj16[16]) {
    FF_REQUIRE(cam, "ff_camera_inverse_matrices: NULL camera");
    V3 f, r, u;
    ffh::camera_basis(cam->world_up, cam->yaw_deg, cam->pitch_deg, f, r, u);
    const V3 pos = ffm::v3(cam->pos[0], cam->pos[1], cam->pos[2]);
    const M4 iv = ffh::inverted(ffh::look_at_rh(pos, ffm::add(pos, f), u));
    const M4 ip = ffh::inverted(ffh::perspective_fov_rh_no(ffh::to_radians(cam->fov_deg), (float)cam->width, (float)cam->height,
                                                           cam->near_clip, cam->far_clip));
    if (inv_view16) memcpy(inv_view16, &iv, 64);
    if (inv_proj16) memcpy(inv_proj16, &ip, 64);
    return FF_OK;
}
int ff_camera_rays(ff_context* ctx, const ff_camera* cam, float* rays6_host, void* d_rays) {
    FF_REQUIRE(ctx, "ff_camera_rays: NULL context");
    int rc = check_cam(cam, "ff_camera_rays");
    if (rc) return rc;
    FF_CUDA(cudaSetDevice(ctx->device));
    const size_t n = (size_t)cam->width * cam->height;
    DevTmp tr, t6;
    float4* dr = (float4*)d_rays;
    if (!dr) { FF_CUDA(tr.alloc(sizeof(float4) * 2 * n)); dr = tr.as<float4>(); }
    k_camera_rays(make_cam(cam), dr, dr + n, ctx->stream);
    if (rays6_host) {
        FF_CUDA(t6.alloc(sizeof(float) * 6 * n));
        k_unpack_rays(dr, dr + n, n, t6.as<float>(), ctx->stream);
        FF_CUDA(cudaMemcpyAsync(rays6_host, t6.p, sizeof(float) * 6 * n, cudaMemcpyDeviceToHost, ctx->stream));
    }
    FF_CUDA(cudaStreamSynchronize(ctx->stream));
    FF_CUDA(cudaGetLastError());
    return FF_OK;
}

}  // extern "C"

// ================================================================================================ render
namespace {
#ifndef FF_PASS_PATHS
#define FF_PASS_PATHS 64       // default paths in flight per wavefront pass, in Mi (9.6 GB of state on a 180 GB part)
#define FF_PASS_PATHS_MAX 128  // upper bound for an explicit samples_per_pass, in Mi (state is ~150 B per path)
#endif
int ensure_wave(ff_scene* s, size_t capacity, int n_counts, bool want_shadow) {
    if (s->wb_capacity < capacity) {
        for (int k = 0; k < 2; ++k) { cudaFree(s->wb.ray_o[k]); cudaFree(s->wb.ray_d[k]); cudaFree(s->wb.beta[k]); s->wb.ray_o[k] = s->wb.ray_d[k] = s->wb.beta[k] = nullptr; }
        cudaFree(s->wb.hits); cudaFree(s->wb.radiance); cudaFree(s->wb.sh_o); cudaFree(s->wb.sh_d); cudaFree(s->wb.sh_c);
        s->wb.hits = nullptr; s->wb.radiance = nullptr; s->wb.sh_o = s->wb.sh_d = s->wb.sh_c = nullptr; s->wb_capacity = 0; s->wb_shadow = false;
        for (int k = 0; k < 2; ++k) {
            FF_CUDA(cudaMalloc((void**)&s->wb.ray_o[k], sizeof(float4) * capacity));
            FF_CUDA(cudaMalloc((void**)&s->wb.ray_d[k], sizeof(float4) * capacity));
            FF_CUDA(cudaMalloc((void**)&s->wb.beta[k], sizeof(float4) * capacity));
        }
        FF_CUDA(cudaMalloc((void**)&s->wb.hits, sizeof(HitLite) * capacity));
        FF_CUDA(cudaMalloc((void**)&s->wb.radiance, sizeof(float4) * capacity));
        s->wb_capacity = capacity;
        s->wb.capacity = (uint32_t)capacity;
    }
    if (want_shadow && !s->wb_shadow) {  // shadow queue of the next-event-estimation stage, only when asked for
        FF_CUDA(cudaMalloc((void**)&s->wb.sh_o, sizeof(float4) * s->wb_capacity));
        FF_CUDA(cudaMalloc((void**)&s->wb.sh_d, sizeof(float4) * s->wb_capacity));
        FF_CUDA(cudaMalloc((void**)&s->wb.sh_c, sizeof(float4) * s->wb_capacity));
        s->wb_shadow = true;
    }
    if (s->wb_counts < n_counts) {
        cudaFree(s->wb.counts); s->wb.counts = nullptr;
        FF_CUDA(cudaMalloc((void**)&s->wb.counts, sizeof(uint32_t) * 4 * n_counts));
        s->wb_counts = n_counts;
    }
    return FF_OK;
}
int ensure_fb(ff_scene* s, size_t npix) {
    if (s->fb_pixels < npix) {
        cudaFree(s->d_accum); cudaFree(s->d_out); s->d_accum = s->d_out = nullptr; s->fb_pixels = 0;
        FF_CUDA(cudaMalloc((void**)&s->d_accum, sizeof(float4) * npix));
        FF_CUDA(cudaMalloc((void**)&s->d_out, sizeof(float4) * npix));
        s->fb_pixels = npix;
    }
    return FF_OK;
}
int ensure_u8(ff_scene* s, size_t bytes) {
    if (s->u8_bytes < bytes) {
        cudaFree(s->d_u8); s->d_u8 = nullptr; s->u8_bytes = 0;
        FF_CUDA(cudaMalloc((void**)&s->d_u8, bytes));
        s->u8_bytes = bytes;
    }
    return FF_OK;
}
// The normal the reference's frame shows for a mesh hit is inverse(transpose(M)) * the OBJECT-space unit normal
// (kernel.cu:101,117), which the baked world-space triangles cannot give back (a scale changes its length).  The first
// as-coded frame after a commit builds it per flat triangle from the host copy of the object-space vertices (16 B per
// triangle, only for scenes that are ever rendered as-coded).
int ensure_as_coded_normals(ff_scene* scene) {
    if (scene->d_as_coded_normals || scene->dev.n_tris == 0) return FF_OK;
    cudaStream_t st = scene->ctx->stream;
    const size_t n_tris = scene->dev.n_tris;
    std::vector<M4> nrm(scene->geoms.size());
    std::vector<uint32_t> geom_of(n_tris);
    DevTmp t_obj, t_nrm, t_geom;
    FF_CUDA(t_obj.alloc(sizeof(float) * 9 * n_tris));
    size_t off = 0, flat = 0;
    for (size_t gi = 0; gi < scene->geoms.size(); ++gi) {
        const HostGeom& g = scene->geoms[gi];
        nrm[gi] = ffh::inverted(ffh::transposed(g.model));
        if (g.type != 2 || g.xyz9.empty()) continue;
        FF_REQUIRE(off + g.xyz9.size() <= 9 * n_tris, "as-coded frame: geometry was edited after the commit");
        FF_CUDA(cudaMemcpyAsync(t_obj.as<float>() + off, g.xyz9.data(), sizeof(float) * g.xyz9.size(), cudaMemcpyHostToDevice, st));
        off += g.xyz9.size();
        for (size_t t = 0; t < g.xyz9.size() / 9; ++t) geom_of[flat++] = (uint32_t)gi;
    }
    FF_REQUIRE(flat == n_tris, "as-coded frame: geometry was edited after the commit");
    FF_CUDA(t_nrm.alloc(sizeof(M4) * nrm.size()));
    FF_CUDA(t_geom.alloc(sizeof(uint32_t) * n_tris));
    FF_CUDA(cudaMemcpyAsync(t_nrm.p, nrm.data(), sizeof(M4) * nrm.size(), cudaMemcpyHostToDevice, st));
    FF_CUDA(cudaMemcpyAsync(t_geom.p, geom_of.data(), sizeof(uint32_t) * n_tris, cudaMemcpyHostToDevice, st));
    float4* out = nullptr;
    FF_CUDA(cudaMalloc((void**)&out, sizeof(float4) * n_tris));
    k_as_coded_normals(t_obj.as<float>(), t_nrm.as<M4>(), t_geom.as<uint32_t>(), (uint32_t)n_tris, out, st);
    cudaError_t e = cudaStreamSynchronize(st);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) { cudaFree(out); set_error(std::string("as-coded normals: ") + cudaGetErrorString(e)); return FF_ERR_CUDA; }
    scene->d_as_coded_normals = out;
    return FF_OK;
}
int check_params(const ff_render_params* p, const char* who) {
    if (!p) { set_error(std::string(who) + ": NULL params"); return FF_ERR_INVALID; }
    if (p->spp <= 0 || p->max_bounces < 0 || p->max_bounces > 1024 || p->first_sample < 0) {
        set_error(std::string(who) + ": bad spp / max_bounces / first_sample"); return FF_ERR_INVALID;
    }
    if ((p->integrator & 0xff) != FF_INTEGRATOR_WAVEFRONT && (p->integrator & 0xff) != FF_INTEGRATOR_MEGAKERNEL) {
        set_error(std::string(who) + ": integrator must be wavefront or megakernel here"); return FF_ERR_INVALID;
    }
    if (p->nee != 0 && p->nee != 1) { set_error(std::string(who) + ": nee must be 0 or 1"); return FF_ERR_INVALID; }
    if (p->tile_world > 1 && (p->tile_rank < 0 || p->tile_rank >= p->tile_world || p->tile_size < 0 || p->tile_size > 4096)) {
        set_error(std::string(who) + ": bad tile_rank / tile_world / tile_size"); return FF_ERR_INVALID;
    }
    return FF_OK;
}
// The pixels a rank owns under the interleaved-tile split, ascending inside each tile, tiles in row-major order.
void tile_owned_pixels(int width, int height, int tile, int rank, int world, std::vector<uint32_t>& out) {
    out.clear();
    if (tile <= 0) tile = 32;
    const int tx = (width + tile - 1) / tile, ty = (height + tile - 1) / tile;
    for (long t = rank; t < (long)tx * ty; t += world) {
        const int x0 = (int)(t % tx) * tile, y0 = (int)(t / tx) * tile;
        for (int y = y0; y < std::min(height, y0 + tile); ++y)
            for (int x = x0; x < std::min(width, x0 + tile); ++x) out.push_back((uint32_t)y * (uint32_t)width + (uint32_t)x);
    }
}
// Adds params->spp samples (as sums, in sample-index order) into d_accum.
int accumulate_samples(ff_scene* scene, const ff_camera* cam, const ff_render_params* p, float4* d_accum, ff_stats* stats) {
    ff_context* c = scene->ctx;
    cudaStream_t st = c->stream;
    size_t npix = (size_t)cam->width * cam->height;
    const uint32_t* d_map = nullptr;
    if (p->tile_world > 1) {  // this call renders the rank's tiles only
        const int key[5] = {cam->width, cam->height, p->tile_size, p->tile_rank, p->tile_world};
        if (!scene->d_pixel_map || memcmp(key, scene->pixel_map_key, sizeof(key)) != 0) {
            std::vector<uint32_t> own;
            tile_owned_pixels(cam->width, cam->height, p->tile_size, p->tile_rank, p->tile_world, own);
            cudaFree(scene->d_pixel_map); scene->d_pixel_map = nullptr; scene->pixel_map_count = 0;
            FF_CUDA(cudaMalloc((void**)&scene->d_pixel_map, sizeof(uint32_t) * std::max<size_t>(1, own.size())));
            FF_CUDA(cudaMemcpy(scene->d_pixel_map, own.data(), sizeof(uint32_t) * own.size(), cudaMemcpyHostToDevice));
            scene->pixel_map_count = own.size();
            memcpy(scene->pixel_map_key, key, sizeof(key));
        }
        npix = scene->pixel_map_count;
        d_map = scene->d_pixel_map;
        if (npix == 0) return FF_OK;  // more ranks than tiles: nothing to render here
    }
    // samples per pass: ~64M paths in flight by default.  Late bounces of a pass carry few rays (C3: 8 bounces, most paths
    // escape early), and a persistent kernel over a short queue is all tail: C3 (64 spp) 28.5 ms at 8M paths per pass, 23.7 at
    // 16M, 21.3 at 32M, 19.8 at 64M; C2 47.1 / 46.4 / 46.0 ms at 16M / 32M / 64M (profiles/r01j_pass_size.txt).  State is
    // ~150 B per path (9.6 GB at the default); an explicit samples_per_pass may go up to 128M paths.
    size_t spb = p->samples_per_pass > 0 ? (size_t)p->samples_per_pass : std::max<size_t>(1, ((size_t)FF_PASS_PATHS << 20) / npix);
    spb = std::min<size_t>(spb, (size_t)p->spp);
    while (spb > 1 && spb * npix > ((size_t)FF_PASS_PATHS_MAX << 20)) --spb;
    FF_REQUIRE(spb * npix < 0xffffffffull, "render: too many paths per pass");
    int rc = ensure_wave(scene, spb * npix, p->max_bounces + 2, p->nee != 0);
    while (rc != FF_OK && p->samples_per_pass <= 0 && spb > 1) {
        // the DEFAULT pass size did not fit beside what else lives on the device: give the partial allocation back and halve it
        // (an explicit samples_per_pass is the caller's decision and fails as it did)
        cudaGetLastError();
        scene->wb_capacity = 0;
        spb = (spb + 1) / 2;
        rc = ensure_wave(scene, spb * npix, p->max_bounces + 2, p->nee != 0);
    }
    if (rc) return rc;
    rc = ensure_counters(scene);
    if (rc) return rc;
    const CamDev cd = make_cam(cam);
    RenderDev rd;
    rd.seed_lo = (uint32_t)(p->seed & 0xffffffffull); rd.seed_hi = (uint32_t)(p->seed >> 32);
    rd.max_bounces = p->max_bounces; rd.aa = p->aa; rd.nee = p->nee; rd.npix = (uint32_t)npix; rd.pixel_map = d_map;
    rd.background[0] = p->background[0]; rd.background[1] = p->background[1]; rd.background[2] = p->background[2];
    rd.ray_eps = p->ray_eps;
    const int n_pass = (int)((p->spp + spb - 1) / spb);
    const bool wave = (p->integrator & 0xff) == FF_INTEGRATOR_WAVEFRONT;
    const bool count = wave && (p->integrator & FF_TRACE_COUNT) != 0;  // instrumented extend kernels
    if (count) FF_CUDA(cudaMemsetAsync(scene->d_trace_counters, 0, sizeof(TraceCounters), st));
    // one event pair per extend stage, so trace_ms can be reported separately from the bookkeeping
    const size_t need_events = wave ? (size_t)(p->nee ? 4 : 2) * n_pass * (p->max_bounces + 1) : 0;
    const bool time_trace = stats && need_events <= 8192;
    if (time_trace) while (c->trace_events.size() < need_events) { cudaEvent_t e; FF_CUDA(cudaEventCreate(&e)); c->trace_events.push_back(e); }
    int n_ev = 0;
    uint32_t launches = 0;
    uint64_t rays = 0, shadow_rays = 0;
    const size_t nc = (size_t)p->max_bounces + 2;  // counts block: 4 x nc (paths | work | shadow rays | shadow work)
    std::vector<uint32_t> h_counts((size_t)n_pass * 4 * nc);
    FF_CUDA(cudaMemsetAsync(scene->d_ray_counter, 0, 2 * sizeof(unsigned long long), st));
    ScopedTimer t(c, 0, 1);
    for (int pass = 0; pass < n_pass; ++pass) {
        const size_t s0 = (size_t)pass * spb;
        rd.first_sample = (uint32_t)(p->first_sample + s0);
        rd.samples = (uint32_t)std::min<size_t>(spb, p->spp - s0);
        if (wave) {
            k_wavefront_pass(scene->dev, cd, rd, scene->wb, FF_TRACE_BVH8 | (count ? FF_TRACE_COUNT : 0), scene->d_trace_counters, &launches, time_trace ? c->trace_events.data() : nullptr, &n_ev, st);
            FF_CUDA(cudaMemcpyAsync(h_counts.data() + (size_t)pass * 4 * nc, scene->wb.counts, sizeof(uint32_t) * 4 * nc,
                                    cudaMemcpyDeviceToHost, st));
        } else {
            k_megakernel_pass(scene->dev, cd, rd, scene->wb.radiance, scene->d_work_counter, scene->d_ray_counter, &launches, st);
        }
        k_accumulate_pass(scene->wb.radiance, (uint32_t)npix, rd.samples, d_map, d_accum, st);
        ++launches;
    }
    const double ms = t.stop();
    FF_CUDA(cudaGetLastError());
    if (wave) {
        for (int pass = 0; pass < n_pass; ++pass)
            for (int b = 0; b <= p->max_bounces; ++b) {
                rays += h_counts[(size_t)pass * 4 * nc + b];
                if (p->nee && scene->dev.n_lights && b < p->max_bounces) shadow_rays += h_counts[(size_t)pass * 4 * nc + 2 * nc + b];
            }
    } else {
        unsigned long long r[2] = {0, 0};
        FF_CUDA(cudaMemcpy(r, scene->d_ray_counter, sizeof(r), cudaMemcpyDeviceToHost));
        rays = r[0]; shadow_rays = r[1];
    }
    if (stats && count) {
        TraceCounters h;
        FF_CUDA(cudaMemcpy(&h, scene->d_trace_counters, sizeof(h), cudaMemcpyDeviceToHost));
        stats->nodes_visited += h.nodes; stats->tris_tested += h.tris;
    }
    if (stats) {
        stats->gpu_ms += ms; stats->rays += rays; stats->shadow_rays += shadow_rays; stats->paths += (uint64_t)p->spp * npix; stats->kernel_launches += launches; stats->passes += (uint32_t)n_pass;
        if (time_trace) {
            double tr = 0;
            for (int k = 0; k + 1 < n_ev; k += 2) { float m = 0; cudaEventElapsedTime(&m, c->trace_events[k], c->trace_events[k + 1]); tr += m; }
            stats->trace_ms += tr;
        }
    }
    return FF_OK;
}
}  // namespace

namespace ff {
int host_check_render(ff_scene* scene, const ff_camera* cam, const ff_render_params* params, const char* who) {
    int rc = check_scene(scene, who);
    if (rc) return rc;
    if ((rc = check_cam(cam, who))) return rc;
    return check_params(params, who);
}
int host_accumulate_samples(ff_scene* scene, const ff_camera* cam, const ff_render_params* params, float4* d_accum, ff_stats* stats) {
    return accumulate_samples(scene, cam, params, d_accum, stats);
}
}  // namespace ff

extern "C" {

int ff_render_accumulate(ff_scene* scene, const ff_camera* cam, const ff_render_params* params, void* d_accum_rgba, ff_stats* stats) {
    int rc = check_scene(scene, "ff_render_accumulate");
    if (rc) return rc;
    if ((rc = check_cam(cam, "ff_render_accumulate"))) return rc;
    if ((rc = check_params(params, "ff_render_accumulate"))) return rc;
    FF_REQUIRE(d_accum_rgba, "ff_render_accumulate: NULL accumulator");
    if (stats) memset(stats, 0, sizeof(*stats));
    return accumulate_samples(scene, cam, params, (float4*)d_accum_rgba, stats);
}

int ff_resolve(ff_context* ctx, const void* d_accum_rgba, int width, int height, int total_spp, void* d_out_rgba, float* out_rgba_host) {
    const void* one[1] = {d_accum_rgba};
    return ff_reduce_resolve_peers(ctx, one, 1, width, height, total_spp, d_out_rgba, out_rgba_host);
}
int ff_reduce_resolve_peers(ff_context* ctx, const void* const* d_accum_peers, int n_peers, int width, int height, int total_spp,
                            void* d_out_rgba, float* out_rgba_host) {
    FF_REQUIRE(ctx && d_accum_peers && n_peers >= 1 && n_peers <= 8, "ff_reduce_resolve_peers: bad argument");
    FF_REQUIRE(width > 0 && height > 0 && total_spp > 0, "ff_reduce_resolve_peers: bad size");
    FF_REQUIRE(d_out_rgba || out_rgba_host, "ff_reduce_resolve_peers: no output");
    FF_CUDA(cudaSetDevice(ctx->device));
    const size_t npix = (size_t)width * height;
    float4* d_out = (float4*)d_out_rgba;
    if (!d_out) {
        if (ctx->resolve_pixels < npix) {
            cudaFree(ctx->resolve_buf); ctx->resolve_buf = nullptr; ctx->resolve_pixels = 0;
            FF_CUDA(cudaMalloc((void**)&ctx->resolve_buf, sizeof(float4) * npix));
            ctx->resolve_pixels = npix;
        }
        d_out = ctx->resolve_buf;
    }
    const float inv = 1.0f / (float)total_spp;
    k_resolve((const float4* const*)d_accum_peers, n_peers, (uint32_t)npix, inv, d_out, ctx->stream);
    if (out_rgba_host) FF_CUDA(cudaMemcpyAsync(out_rgba_host, d_out, sizeof(float4) * npix, cudaMemcpyDeviceToHost, ctx->stream));
    FF_CUDA(cudaStreamSynchronize(ctx->stream));
    FF_CUDA(cudaGetLastError());
    return FF_OK;
}

int ff_render(ff_scene* scene, const ff_camera* cam, const ff_render_params* params, float* out_rgba_host, void* out_rgba_device,
              ff_stats* stats) {
    int rc = check_scene(scene, "ff_render");
    if (rc) return rc;
    if ((rc = check_cam(cam, "ff_render"))) return rc;
    FF_REQUIRE(params, "ff_render: NULL params");
    if (stats) memset(stats, 0, sizeof(*stats));
    ff_context* c = scene->ctx;
    const size_t npix = (size_t)cam->width * cam->height;
    if ((rc = ensure_fb(scene, npix))) return rc;
    if ((params->integrator & 0xff) == FF_INTEGRATOR_AS_CODED) {
        // 8-bit as-coded frame widened to float RGBA (value / 255) on the device, for callers that want one code path
        if ((rc = ensure_u8(scene, npix * 3))) return rc;
        if ((rc = ff_render_u8(scene, cam, params, scene->d_u8, 1, stats))) return rc;
        float4* d_out = out_rgba_device ? (float4*)out_rgba_device : scene->d_out;
        k_u8_to_rgba(scene->d_u8, (uint32_t)npix, d_out, c->stream);
        if (out_rgba_host) FF_CUDA(cudaMemcpyAsync(out_rgba_host, d_out, sizeof(float4) * npix, cudaMemcpyDeviceToHost, c->stream));
        FF_CUDA(cudaStreamSynchronize(c->stream));
        FF_CUDA(cudaGetLastError());
        if (stats) stats->kernel_launches += 1;
        return FF_OK;
    }
    const int integ = params->integrator & 0xff;
    if (integ >= FF_INTEGRATOR_AOV_NORMAL && integ <= FF_INTEGRATOR_AOV_ALBEDO) {
        float4* d_out = out_rgba_device ? (float4*)out_rgba_device : scene->d_out;
        ScopedTimer t(c, 0, 1);
        k_aov(scene->dev, make_cam(cam), integ, d_out, c->stream);
        if (out_rgba_host) FF_CUDA(cudaMemcpyAsync(out_rgba_host, d_out, sizeof(float4) * npix, cudaMemcpyDeviceToHost, c->stream));
        const double ms = t.stop();
        FF_CUDA(cudaGetLastError());
        if (stats) { stats->gpu_ms = ms; stats->trace_ms = ms; stats->rays = npix; stats->paths = npix; stats->kernel_launches = 1; }
        return FF_OK;
    }
    if ((rc = check_params(params, "ff_render"))) return rc;
    FF_CUDA(cudaMemsetAsync(scene->d_accum, 0, sizeof(float4) * npix, c->stream));
    if ((rc = accumulate_samples(scene, cam, params, scene->d_accum, stats))) return rc;
    float4* d_out = out_rgba_device ? (float4*)out_rgba_device : scene->d_out;
    const int total = params->total_spp > 0 ? params->total_spp : params->spp;
    const float4* one[1] = {scene->d_accum};
    ScopedTimer t(c, 2, 3);
    k_resolve(one, 1, (uint32_t)npix, 1.0f / (float)total, d_out, c->stream);
    if (out_rgba_host) FF_CUDA(cudaMemcpyAsync(out_rgba_host, d_out, sizeof(float4) * npix, cudaMemcpyDeviceToHost, c->stream));
    const double ms = t.stop();
    FF_CUDA(cudaGetLastError());
    if (stats) { stats->gpu_ms += ms; stats->d2h_ms = out_rgba_host ? ms : 0; stats->kernel_launches += 1; }
    return FF_OK;
}

int ff_render_u8(ff_scene* scene, const ff_camera* cam, const ff_render_params* params, unsigned char* out_rgb, int out_is_device,
                 ff_stats* stats) {
    int rc = check_scene(scene, "ff_render_u8");
    if (rc) return rc;
    if ((rc = check_cam(cam, "ff_render_u8"))) return rc;
    FF_REQUIRE(params && out_rgb, "ff_render_u8: NULL argument");
    ff_context* c = scene->ctx;
    const size_t npix = (size_t)cam->width * cam->height;
    unsigned char* d_rgb = out_is_device ? out_rgb : nullptr;
    if (!d_rgb) {  // host output: staging owned by the scene, grown on demand (no allocation per frame)
        if ((rc = ensure_u8(scene, npix * 3))) return rc;
        d_rgb = scene->d_u8;
    }
    if ((params->integrator & 0xff) == FF_INTEGRATOR_AS_CODED) {
        if (stats) memset(stats, 0, sizeof(*stats));
        if ((rc = ensure_as_coded_normals(scene))) return rc;
        ScopedTimer t(c, 0, 1);
        k_as_coded(scene->dev, make_cam(cam), scene->d_as_coded_normals, d_rgb, c->stream);
        const double ms = t.stop();
        FF_CUDA(cudaGetLastError());
        if (stats) { stats->gpu_ms = ms; stats->trace_ms = ms; stats->rays = npix; stats->paths = npix; stats->kernel_launches = 1; }
    } else {
        if ((rc = ensure_fb(scene, npix))) return rc;
        if ((rc = ff_render(scene, cam, params, nullptr, scene->d_out, stats))) return rc;
        k_rgba_to_u8(scene->d_out, (uint32_t)npix, d_rgb, c->stream);
        FF_CUDA(cudaStreamSynchronize(c->stream));
        if (stats) stats->kernel_launches += 1;
    }
    if (!out_is_device) FF_CUDA(cudaMemcpy(out_rgb, d_rgb, npix * 3, cudaMemcpyDeviceToHost));
    return FF_OK;
}

// ================================================================================================ workloads, memory, output
int ff_make_soup(ff_context* ctx, size_t n_tris, float s, float* xyz9_host) {
    FF_REQUIRE(ctx && (xyz9_host || n_tris == 0), "ff_make_soup: NULL argument");
    if (!n_tris) return FF_OK;
    FF_CUDA(cudaSetDevice(ctx->device));
    DevTmp td;
    FF_CUDA(td.alloc(sizeof(float) * 9 * n_tris));
    k_make_soup(n_tris, s, td.as<float>(), ctx->stream);
    FF_CUDA(cudaMemcpyAsync(xyz9_host, td.p, sizeof(float) * 9 * n_tris, cudaMemcpyDeviceToHost, ctx->stream));
    FF_CUDA(cudaStreamSynchronize(ctx->stream));
    return FF_OK;
}
int ff_make_rays_uniform(ff_context* ctx, uint64_t first, size_t n, uint32_t salt, float* rays6_host, void* d_rays) {
    FF_REQUIRE(ctx, "ff_make_rays_uniform: NULL context");
    if (!n) return FF_OK;
    FF_CUDA(cudaSetDevice(ctx->device));
    DevTmp tr, t6;
    float4* dr = (float4*)d_rays;
    if (!dr) { FF_CUDA(tr.alloc(sizeof(float4) * 2 * n)); dr = tr.as<float4>(); }
    k_make_rays_uniform(first, n, salt, dr, dr + n, ctx->stream);
    if (rays6_host) {
        FF_CUDA(t6.alloc(sizeof(float) * 6 * n));
        k_unpack_rays(dr, dr + n, n, t6.as<float>(), ctx->stream);
        FF_CUDA(cudaMemcpyAsync(rays6_host, t6.p, sizeof(float) * 6 * n, cudaMemcpyDeviceToHost, ctx->stream));
    }
    FF_CUDA(cudaStreamSynchronize(ctx->stream));
    FF_CUDA(cudaGetLastError());
    return FF_OK;
}
int ff_make_rays_bounce(ff_scene* scene, const void* d_rays_in, const void* d_hits_in, size_t n, uint64_t seed, void* d_rays_out, size_t* out_n) {
    int rc = check_scene(scene, "ff_make_rays_bounce");
    if (rc) return rc;
    FF_REQUIRE(d_rays_in && d_hits_in && d_rays_out && out_n && n < 0xffffffffull, "ff_make_rays_bounce: bad argument");
    if ((rc = ensure_counters(scene))) return rc;
    cudaStream_t st = scene->ctx->stream;
    FF_CUDA(cudaMemsetAsync(scene->d_work_counter, 0, sizeof(uint32_t), st));
    const float4* ri = (const float4*)d_rays_in;
    // survivors are written compacted: n_out origins at [0, n_out), directions start at offset n (the input count)
    float4* ro = (float4*)d_rays_out;
    k_make_rays_bounce(scene->dev, ri, ri + n, (const HitRec*)d_hits_in, (uint32_t)n, (uint32_t)seed, (uint32_t)(seed >> 32), ro, ro + n,
                       scene->d_work_counter, st);
    uint32_t cnt = 0;
    FF_CUDA(cudaMemcpyAsync(&cnt, scene->d_work_counter, sizeof(cnt), cudaMemcpyDeviceToHost, st));
    FF_CUDA(cudaStreamSynchronize(st));
    // make the output a dense [cnt origins][cnt directions] batch
    if (cnt && cnt != n) FF_CUDA(cudaMemcpy(ro + cnt, ro + n, sizeof(float4) * cnt, cudaMemcpyDeviceToDevice));
    *out_n = cnt;
    return FF_OK;
}

int ff_device_malloc(ff_context* ctx, size_t bytes, void** out) {
    FF_REQUIRE(ctx && out, "ff_device_malloc: NULL argument");
    FF_CUDA(cudaSetDevice(ctx->device));
    FF_CUDA(cudaMalloc(out, bytes ? bytes : 16));
    return FF_OK;
}
int ff_device_free(ff_context* ctx, void* p) {
    FF_REQUIRE(ctx, "ff_device_free: NULL context");
    FF_CUDA(cudaSetDevice(ctx->device));
    FF_CUDA(cudaFree(p));
    return FF_OK;
}
int ff_device_memset(ff_context* ctx, void* p, int value, size_t bytes) {
    FF_REQUIRE(ctx && p, "ff_device_memset: NULL argument");
    FF_CUDA(cudaSetDevice(ctx->device));
    FF_CUDA(cudaMemsetAsync(p, value, bytes, ctx->stream));
    FF_CUDA(cudaStreamSynchronize(ctx->stream));
    return FF_OK;
}
int ff_memcpy_h2d(ff_context* ctx, void* dst, const void* src, size_t bytes) {
    FF_REQUIRE(ctx && dst && src, "ff_memcpy_h2d: NULL argument");
    FF_CUDA(cudaSetDevice(ctx->device));
    FF_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
    FF_CUDA(cudaStreamSynchronize(ctx->stream));
    return FF_OK;
}
int ff_memcpy_d2h(ff_context* ctx, void* dst, const void* src, size_t bytes) {
    FF_REQUIRE(ctx && dst && src, "ff_memcpy_d2h: NULL argument");
    FF_CUDA(cudaSetDevice(ctx->device));
    FF_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    FF_CUDA(cudaStreamSynchronize(ctx->stream));
    return FF_OK;
}
int ff_host_alloc_pinned(size_t bytes, void** out) {
    FF_REQUIRE(out, "ff_host_alloc_pinned: NULL argument");
    FF_CUDA(cudaMallocHost(out, bytes ? bytes : 16));
    return FF_OK;
}
int ff_host_free_pinned(void* p) {
    FF_CUDA(cudaFreeHost(p));
    return FF_OK;
}
int ff_ipc_get_handle(ff_context* ctx, void* d_ptr, unsigned char handle64[64]) {
    FF_REQUIRE(ctx && d_ptr && handle64, "ff_ipc_get_handle: NULL argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    FF_CUDA(cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h;
    FF_CUDA(cudaIpcGetMemHandle(&h, d_ptr));
    memcpy(handle64, &h, 64);
    return FF_OK;
}
int ff_ipc_open_handle(ff_context* ctx, const unsigned char handle64[64], void** out_d_ptr) {
    FF_REQUIRE(ctx && handle64 && out_d_ptr, "ff_ipc_open_handle: NULL argument");
    FF_CUDA(cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    FF_CUDA(cudaIpcOpenMemHandle(out_d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return FF_OK;
}
int ff_ipc_close_handle(ff_context* ctx, void* d_ptr) {
    FF_REQUIRE(ctx && d_ptr, "ff_ipc_close_handle: NULL argument");
    FF_CUDA(cudaSetDevice(ctx->device));
    FF_CUDA(cudaIpcCloseMemHandle(d_ptr));
    return FF_OK;
}
int ff_flush_l2(ff_context* ctx) {
    FF_REQUIRE(ctx, "ff_flush_l2: NULL context");
    FF_CUDA(cudaSetDevice(ctx->device));
    if (!ctx->flush_buf) {
        ctx->flush_n = (256u << 20) / sizeof(float4);  // 256 MiB > 126 MB L2
        FF_CUDA(cudaMalloc((void**)&ctx->flush_buf, ctx->flush_n * sizeof(float4)));
    }
    k_flush(ctx->flush_buf, ctx->flush_n, ctx->stream);
    FF_CUDA(cudaStreamSynchronize(ctx->stream));
    return FF_OK;
}

long ff_tile_owned_pixels(int width, int height, int tile_size, int tile_rank, int tile_world, uint32_t* out, size_t capacity) {
    if (width <= 0 || height <= 0 || tile_size < 0 || tile_world < 1 || tile_rank < 0 || tile_rank >= tile_world) {
        set_error("ff_tile_owned_pixels: bad argument"); return FF_ERR_INVALID;
    }
    std::vector<uint32_t> own;
    tile_owned_pixels(width, height, tile_size, tile_rank, tile_world, own);
    if (out) {
        if (capacity < own.size()) { set_error("ff_tile_owned_pixels: capacity too small"); return FF_ERR_INVALID; }
        memcpy(out, own.data(), sizeof(uint32_t) * own.size());
    }
    return (long)own.size();
}

// ---------------------------------------------------------------- camera controls (utilities.h:343-427)
int ff_camera_process_keyboard(ff_camera* cam, int direction) {
    FF_REQUIRE(cam, "ff_camera_process_keyboard: NULL camera");
    FF_REQUIRE(direction >= 0 && direction <= FF_CAM_RESET, "ff_camera_process_keyboard: direction must be 0..10");
    const float velocity = 0.2f;  // m_cameraMouseSensitivity (utilities.h:288,345)
    V3 f, r, u;
    ffh::camera_basis(cam->world_up, cam->yaw_deg, cam->pitch_deg, f, r, u);
    auto move = [&](const V3& axis, bool plus) {   // m_position +/-= axis * velocity: one rounded multiply, one rounded add
        const V3 step = ffm::mul(axis, velocity);
        const V3 p = ffm::v3(cam->pos[0], cam->pos[1], cam->pos[2]);
        const V3 q = plus ? ffm::add(p, step) : ffm::sub(p, step);
        cam->pos[0] = q.x; cam->pos[1] = q.y; cam->pos[2] = q.z;
    };
    switch (direction) {
        case FF_CAM_FORWARD: move(f, true); break;
        case FF_CAM_BACKWARD: move(f, false); break;
        case FF_CAM_LEFT: move(r, false); break;
        case FF_CAM_RIGHT: move(r, true); break;
        case FF_CAM_UP: move(u, true); break;
        case FF_CAM_DOWN: move(u, false); break;
        case FF_CAM_YAW_LEFT: cam->yaw_deg = ffm::fsub(cam->yaw_deg, 0.5f); break;
        case FF_CAM_YAW_RIGHT: cam->yaw_deg = ffm::fadd(cam->yaw_deg, 0.5f); break;
        case FF_CAM_PITCH_UP: cam->pitch_deg = ffm::fadd(cam->pitch_deg, 0.5f); break;
        case FF_CAM_PITCH_DOWN: cam->pitch_deg = ffm::fsub(cam->pitch_deg, 0.5f); break;
        default:  // resetCamera (utilities.h:420-426)
            cam->pos[0] = 0.f; cam->pos[1] = 0.f; cam->pos[2] = 15.f; cam->pitch_deg = 0.f; cam->yaw_deg = -90.f;
            break;
    }
    return FF_OK;
}
int ff_camera_process_mouse(ff_camera* cam, float x_delta, float y_delta, int constrain_pitch) {
    FF_REQUIRE(cam, "ff_camera_process_mouse: NULL camera");
    const float sens = 0.2f;  // m_cameraMouseSensitivity
    cam->yaw_deg = ffm::fadd(cam->yaw_deg, ffm::fmul(x_delta, sens));
    cam->pitch_deg = ffm::fadd(cam->pitch_deg, ffm::fmul(y_delta, sens));
    if (constrain_pitch) {
        if (cam->pitch_deg > 89.0f) cam->pitch_deg = 89.0f;
        if (cam->pitch_deg < -89.0f) cam->pitch_deg = -89.0f;
    }
    return FF_OK;
}

// ---------------------------------------------------------------- progressive accumulation (viewer loop state, kernel.cu:331-359)
struct ff_progressive {
    ff_scene* scene = nullptr;
    float4* d_accum = nullptr;
    float4* d_out = nullptr;
    unsigned char* d_u8 = nullptr;
    size_t pixels = 0;
    int accumulated = 0;
    bool valid = false;
    ff_camera cam;
    ff_render_params params;
};
int ff_progressive_create(ff_scene* scene, ff_progressive** out) {
    FF_REQUIRE(scene && out, "ff_progressive_create: NULL argument");
    *out = new ff_progressive();
    (*out)->scene = scene;
    return FF_OK;
}
void ff_progressive_destroy(ff_progressive* p) {
    if (!p) return;
    if (p->scene && p->scene->ctx) cudaSetDevice(p->scene->ctx->device);
    cudaFree(p->d_accum); cudaFree(p->d_out); cudaFree(p->d_u8);
    delete p;
}
int ff_progressive_reset(ff_progressive* p) {
    FF_REQUIRE(p, "ff_progressive_reset: NULL session");
    p->valid = false; p->accumulated = 0;
    return FF_OK;
}
int ff_progressive_frame(ff_progressive* p, const ff_camera* cam, const ff_render_params* params, float* out_rgba_host,
                         void* out_rgba_device, unsigned char* out_rgb_u8, int u8_is_device, int* out_total_spp, ff_stats* stats) {
    FF_REQUIRE(p, "ff_progressive_frame: NULL session");
    int rc = check_scene(p->scene, "ff_progressive_frame");
    if (rc) return rc;
    if ((rc = check_cam(cam, "ff_progressive_frame"))) return rc;
    if ((rc = check_params(params, "ff_progressive_frame"))) return rc;
    if (stats) memset(stats, 0, sizeof(*stats));
    ff_context* c = p->scene->ctx;
    FF_CUDA(cudaSetDevice(c->device));
    const size_t npix = (size_t)cam->width * cam->height;
    if (p->pixels != npix) {
        cudaFree(p->d_accum); cudaFree(p->d_out); cudaFree(p->d_u8);
        p->d_accum = p->d_out = nullptr; p->d_u8 = nullptr; p->pixels = 0; p->valid = false;
        FF_CUDA(cudaMalloc((void**)&p->d_accum, sizeof(float4) * npix));
        FF_CUDA(cudaMalloc((void**)&p->d_out, sizeof(float4) * npix));
        FF_CUDA(cudaMalloc((void**)&p->d_u8, npix * 3));
        p->pixels = npix;
    }
    auto same_image = [&]() {
        const ff_render_params& a = p->params;
        return memcmp(&p->cam, cam, sizeof(ff_camera)) == 0 && a.max_bounces == params->max_bounces && a.seed == params->seed &&
               (a.integrator & 0xff) == (params->integrator & 0xff) && a.nee == params->nee && a.aa == params->aa &&
               memcmp(a.background, params->background, sizeof(a.background)) == 0 && a.ray_eps == params->ray_eps &&
               // the tile split decides WHICH pixels this rank accumulates: a changed split must not add to stale pixels
               (a.tile_world > 1 ? a.tile_world : 1) == (params->tile_world > 1 ? params->tile_world : 1) &&
               (params->tile_world <= 1 || (a.tile_rank == params->tile_rank && (a.tile_size > 0 ? a.tile_size : 32) == (params->tile_size > 0 ? params->tile_size : 32)));
    };
    if (!p->valid || !same_image() || (long long)p->accumulated + params->spp > 0x7fffffffll) {
        FF_CUDA(cudaMemsetAsync(p->d_accum, 0, sizeof(float4) * npix, c->stream));
        p->accumulated = 0;
        p->cam = *cam; p->params = *params; p->valid = true;
    }
    ff_render_params q = *params;
    q.first_sample = p->accumulated;   // the Philox stream continues where the previous frame stopped
    if ((rc = accumulate_samples(p->scene, cam, &q, p->d_accum, stats))) { p->valid = false; return rc; }
    p->accumulated += params->spp;
    float4* d_out = out_rgba_device ? (float4*)out_rgba_device : p->d_out;
    const float4* one[1] = {p->d_accum};
    ScopedTimer t(c, 2, 3);
    k_resolve(one, 1, (uint32_t)npix, 1.0f / (float)p->accumulated, d_out, c->stream);
    if (out_rgb_u8) {
        unsigned char* d8 = u8_is_device ? out_rgb_u8 : p->d_u8;
        k_rgba_to_u8(d_out, (uint32_t)npix, d8, c->stream);
        if (!u8_is_device) FF_CUDA(cudaMemcpyAsync(out_rgb_u8, d8, npix * 3, cudaMemcpyDeviceToHost, c->stream));
    }
    if (out_rgba_host) FF_CUDA(cudaMemcpyAsync(out_rgba_host, d_out, sizeof(float4) * npix, cudaMemcpyDeviceToHost, c->stream));
    const double ms = t.stop();
    FF_CUDA(cudaGetLastError());
    if (stats) { stats->gpu_ms += ms; stats->d2h_ms = (out_rgba_host || (out_rgb_u8 && !u8_is_device)) ? ms : 0; stats->kernel_launches += out_rgb_u8 ? 2 : 1; }
    if (out_total_spp) *out_total_spp = p->accumulated;
    return FF_OK;
}

int ff_debug_set_tuning(int postpone_div, int refill_lanes, int blocks_per_sm) {
    if (postpone_div > 0) tuning().postpone_div = postpone_div;
    if (refill_lanes > 0) tuning().refill_lanes = refill_lanes;
    if (blocks_per_sm > 0) tuning().blocks_per_sm = blocks_per_sm;
    return FF_OK;
}
int ff_debug_node_stats(const ff_scene* scene, unsigned long long child_hist[9], unsigned long long leaf_tri_hist[4]) {
    FF_REQUIRE(scene && scene->committed && child_hist && leaf_tri_hist, "ff_debug_node_stats: committed scene and outputs required");
    for (int k = 0; k < 9; ++k) child_hist[k] = 0;
    for (int k = 0; k < 4; ++k) leaf_tri_hist[k] = 0;
    const size_t nn = scene->dev.n_nodes;
    if (nn == 0) return FF_OK;
    std::vector<float4> h(nn * FF_NODE_F4);
    FF_CUDA(cudaMemcpy(h.data(), scene->d_nodes, sizeof(float4) * h.size(), cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < nn; ++i) {
        uint32_t meta[2];
        memcpy(&meta[0], &h[i * FF_NODE_F4 + 1].z, 4);
        memcpy(&meta[1], &h[i * FF_NODE_F4 + 1].w, 4);
        int nch = 0;
        for (int s = 0; s < 8; ++s) {
            const uint32_t m = (meta[s >> 2] >> (8 * (s & 3))) & 0xffu;
            if (m == 0) continue;
            ++nch;
            if ((m & 0xe0u) == 0x20u && (m & 0x1fu) >= 24u) continue;  // inner child
            const uint32_t unary = m >> 5;
            leaf_tri_hist[unary == 1 ? 1 : unary == 3 ? 2 : 3]++;
        }
        child_hist[nch]++;
    }
    return FF_OK;
}
int ff_write_image(const char* path, const float* rgba, int width, int height) { return write_image_float(path, rgba, width, height); }
int ff_write_image_u8(const char* path, const unsigned char* rgb, int width, int height) { return write_image_u8(path, rgb, width, height); }
void ff_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    const ffm::U4 r = ffm::philox4x32_10(ffm::U4{ctr[0], ctr[1], ctr[2], ctr[3]}, key[0], key[1]);
    out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
}

}  // extern "C"
