// ff_host.h -- host-side objects behind the opaque handles of include/firefly.h, shared by the translation units that
// implement the C ABI (ff_api.cpp: context / scene / trace / render; ff_group.cpp: the multi-GPU group).
#pragma once
#include <string>
#include <vector>

#include "ff_hostmath.h"
#include "ff_internal.h"

namespace ff {
void set_error(const std::string& msg);
const std::string& get_error();
}  // namespace ff

using ffm::M4;
using ffm::V3;

#define FF_CUDA(call)                                                                               \
    do {                                                                                            \
        cudaError_t e_ = (call);                                                                    \
        if (e_ != cudaSuccess) {                                                                    \
            ff::set_error(std::string(#call) + ": " + cudaGetErrorString(e_));                          \
            return FF_ERR_CUDA;                                                                     \
        }                                                                                           \
    } while (0)
#define FF_REQUIRE(cond, msg)                                                                       \
    do {                                                                                            \
        if (!(cond)) { ff::set_error(msg); return FF_ERR_INVALID; }                                     \
    } while (0)

// Staging of the host-batch trace pipeline (ff_trace_closest / _uv): FF_PIPE_SLOTS chunks in flight, one device block per
// slot, copy-in and copy-out streams beside the context's own.
#define FF_PIPE_SLOTS 3
struct TracePipe {
    size_t chunk = 0;   // rays per slot
    char* block[FF_PIPE_SLOTS] = {nullptr, nullptr, nullptr};
    uint32_t* work = nullptr;  // one work counter per slot
    cudaStream_t s_in = nullptr, s_out = nullptr, s_comp = nullptr;  // copy-in, copy-out, second compute stream
    cudaEvent_t in_done[FF_PIPE_SLOTS], comp_done[FF_PIPE_SLOTS], out_done[FF_PIPE_SLOTS], span[5];
    bool ready = false;
    // streaming path (one persistent launch per batch, see ff_trace.cu "Host batches"): whole-batch staging
    char* sblock = nullptr;            // rays6 (24 B) | geom | tri | t (4 B each) per ray
    size_t scap = 0;                   // rays the block holds
    ff::StreamSync* d_sync = nullptr;
    uint32_t* h_flags = nullptr;       // page-locked: chunk c is complete on the device (written by the kernel)
    uint32_t* h_marks = nullptr;       // page-locked: values the copy-in stream writes into StreamSync::avail
    cudaEvent_t s_go = nullptr;
    int stream_failures = 0;           // launches that timed out waiting for rays (kernels serialised against copies); 2 = stop trying
};

struct ff_context {
    int device = 0;
    TracePipe pipe;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    std::vector<cudaEvent_t> trace_events;
    float4* flush_buf = nullptr;
    size_t flush_n = 0;
    cudaEvent_t timer_ev[2] = {nullptr, nullptr};  // ff_timer_begin / ff_timer_end
    bool timer_open = false;
    float4* resolve_buf = nullptr;  // device staging of ff_resolve / ff_reduce_resolve_peers when the caller gives no device output
    size_t resolve_pixels = 0;
    ff::Arena arena;                    // builder scratch, reused by every commit on this context
    std::vector<ff_scene*> scenes;  // live scenes: orphaned (device memory released) if the context goes first
};

struct HostGeom {
    int type;  // 0 sphere, 1 plane, 2 triangle mesh (GeometryType, utilities.h:141-146)
    float radius = 1.0f;
    M4 model, inv;
    ff_material mat;
    std::vector<float> xyz9;  // object space
    // optional shading attributes, empty when absent
    std::vector<float> n9, uv6;              // per-corner object-space normals / uvs
    std::vector<int> face_mat;               // per triangle: index into face_materials, -1 = mat
    std::vector<ff_material> face_materials;
};

struct ff_scene {
    ff_context* ctx = nullptr;
    std::vector<HostGeom> geoms;
    bool committed = false;
    // device
    float4* d_flat = nullptr;  // baked triangles in flat order
    int* d_bounds = nullptr;   // scene / centroid bounds reduced by the bake kernel (12 ordered ints), read by the build
    float4* d_tris = nullptr;  // baked triangles in leaf order
    float4* d_nodes = nullptr;
    int* d_ord_geom = nullptr;
    int* d_ord_tri = nullptr;
    float4* d_materials = nullptr;
    ff::DevPlane* d_planes = nullptr;
    uint32_t* d_lights = nullptr;
    int4* d_geom_attr = nullptr;
    float4* d_attr_normals = nullptr;
    float2* d_attr_uvs = nullptr;
    uint32_t* d_attr_facemat = nullptr;
    int2* d_bvh2_children = nullptr;
    float4* d_bvh2_lo = nullptr;
    float4* d_bvh2_hi = nullptr;
    uint32_t* d_leaf_slot = nullptr;
    float4* d_as_coded_normals = nullptr;  // per flat triangle, built by the first as-coded frame after a commit (ensure_as_coded_normals)
    cudaTextureObject_t nodes_tex = 0, tris_tex = 0;  // kernel variants FF_NODE_TEX / FF_TRI_TEX only
    ff::DevScene dev;
    ff_scene_info info;
    // render scratch (grown on demand)
    ff::WaveBuffers wb;
    size_t wb_capacity = 0;
    int wb_counts = 0;
    bool wb_shadow = false;
    float4* d_accum = nullptr;
    float4* d_out = nullptr;
    size_t fb_pixels = 0;
    unsigned char* d_u8 = nullptr;  // device staging of ff_render_u8 when the caller's buffer is host memory
    size_t u8_bytes = 0;
    // tile split: owned-pixel list of the last (width, height, tile, rank, world) rendered
    uint32_t* d_pixel_map = nullptr;
    size_t pixel_map_count = 0;
    int pixel_map_key[5] = {0, 0, 0, 0, 0};
    uint32_t* d_work_counter = nullptr;
    unsigned long long* d_ray_counter = nullptr;
    ff::TraceCounters* d_trace_counters = nullptr;
};

// Device temporary that is released on every exit path (FF_CUDA / FF_REQUIRE return early).
struct DevTmp {
    void* p = nullptr;
    DevTmp() = default;
    DevTmp(const DevTmp&) = delete;
    DevTmp& operator=(const DevTmp&) = delete;
    ~DevTmp() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t bytes) { if (p) { cudaFree(p); p = nullptr; } return bytes ? cudaMalloc(&p, bytes) : cudaSuccess; }
    template <class T> T* as() const { return (T*)p; }
};


namespace ff {
// Argument checks of a render call (scene committed and alive, image size, parameter ranges); sets the device.
int host_check_render(ff_scene* scene, const ff_camera* cam, const ff_render_params* params, const char* who);
// Adds params->spp samples (as sums, in sample-index order) into d_accum on the scene's stream; synchronous.
int host_accumulate_samples(ff_scene* scene, const ff_camera* cam, const ff_render_params* params, float4* d_accum, ff_stats* stats);
}  // namespace ff
