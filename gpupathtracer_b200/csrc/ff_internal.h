// ff_internal.h -- structures shared by the host layer (ff_host.cpp) and the CUDA translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <string>
#include <vector>

#include "../../include/firefly.h"
#include "ff_math.h"

#ifndef FF_SM_STACK
#define FF_SM_STACK 8      // traversal stack entries kept in shared memory per thread
#endif
#ifndef FF_TRI_QUEUE
#define FF_TRI_QUEUE 8     // ring of pending triangle groups per thread in shared memory (power of two)
#endif
#define FF_LOCAL_STACK (64 - FF_SM_STACK)  // spill entries in local memory
#ifndef FF_LEAF_MAX
#define FF_LEAF_MAX 3      // most triangles one leaf child of a wide node may hold (collapse DP, ff_bvh.cu)
#endif
#ifndef FF_TRI_EVERY
#define FF_TRI_EVERY 1     // >1 (probe): a lane that still has node work runs its triangle phase only every k-th iteration
#endif
#define FF_MAX_WIDE_DEPTH (FF_SM_STACK + FF_LOCAL_STACK)
// Binary hierarchy links (both builders, Karras 2012 numbering: an internal node's index lies inside its own leaf range):
// parent[] entries carry FF_PARENT_CROSS when the PARENT's leaf range is not contained in one block of FF_TREELET consecutive
// leaves of the tree's order -- where refit_treelet_kernel (shared memory) hands a subtree over to refit_cross_kernel.
#ifndef FF_TREELET
#define FF_TREELET 256   // leaves per block (r02z: 20 M triangles 6.46 / 6.10 / 7.34 ms with 512 / 256 / 1024)
#endif
#define FF_PARENT_CROSS 0x80000000u
#define FF_PARENT_ROOT 0xffffffffu
#define FF_TRACE_BLOCK 128
#define FF_REFILL_LANES 28            // persistent trace: refill idle lanes when fewer than this many still traverse
#define FF_PERSISTENT_BLOCKS_PER_SM 12
#define FF_TRACE_SELFSKIP 0x400        // OR into a trace mode (internal): rays carry the slot of the triangle they start on in ray_d.w
#define FF_TRACE_STATIC 0x200          // OR into a trace mode: one-thread-per-ray kernel instead of the persistent one
#define FF_POSTPONE_DIV 1000000  // postpone triangle tests when fewer than live/DIV lanes have any (off by default: dynamic fetch does better)

// Kernel-variant switches (tools/probes): FF_NODE_256 = 96-byte nodes fetched with three LDG.E.256; FF_NODE_TEX / FF_TRI_TEX =
// nodes / triangles fetched through the texture path (tex1Dfetch) instead of the LSU.
#ifndef FF_NODE_MIN
#define FF_NODE_MIN 0   // >0: hold a node visit back until this many lanes of the warp want one (see bvh8_step)
#endif
#ifndef FF_TRACE_MINB
#define FF_TRACE_MINB 7   // minimum resident CTAs per SM the persistent trace kernels are compiled for: 72-register cap (they use 71); 8 (64 registers) is 1 % slower on C2 / C3, 9 and more lose 2-50 % (profiles/r01i_variants_minb.txt, r01n_variants_minb7.txt)
#endif
#ifndef FF_SELF_SKIP
#define FF_SELF_SKIP 1   // integrator rays never test the triangle they start on (make_ray_ctx in ff_trace.cuh)
#endif
#ifndef FF_TRI_MIN
#define FF_TRI_MIN 0    // >0: hold triangle tests back until this many lanes of the warp have one pending
#endif
#ifndef FF_NODE_256
#define FF_NODE_256 0
#endif
#ifndef FF_NODE_TEX
#define FF_NODE_TEX 0
#endif
#ifndef FF_TRI_TEX
#define FF_TRI_TEX 0
#endif
#define FF_NODE_F4 (FF_NODE_256 ? 6 : 5)  // float4s per node
#define FF_NODE_EXP_BIAS 15  // node exponents are stored as exponent(scale) + 15 (byte_unit in ff_trace.cuh)

namespace ff {

// ---------------------------------------------------------------- device-side scene (HBM layout)
// All arrays are 16-byte aligned SoA / packed float4; see DESIGN.md section 3.
//   tris      3 float4 per triangle in BVH-leaf order ("slot" order):
//             {v0.xyz, ord}, {v1.xyz, geometry index}, {v2.xyz, flat index}   (integers as raw bits)
//             ord = position in the reference's (geometry, triangle) scan, planes included, i.e. the
//             tie-break order of intersectRays (kernel.cu:133,145).
//   nodes     5 float4 (80 B) per 8-wide compressed node, see ff_bvh.cu for the field layout.
//   ord_geom / ord_tri   ord -> (geometryIndex, triangleIndex) in the reference's numbering.
//   materials 4 float4 per row; rows [0, n_geoms) are the geometries' own, per-face materials follow: {type, albedo.rgb}, {emissive.rgb, intensity}, {specular.rgb, ior}, {transmittance.rgb, 0}
// Analytic primitive of the reference's Geometry: GeometryType::PLANE (unit square, intersectPlane kernel.cu:8-32) or
// GeometryType::SPHERE (utilities.h:143,227; the reference leaves its branch empty, kernel.cu:166-169).
struct DevPlane {
    ffm::M4 model, inv;   // Geometry::m_modelMatrix / m_invModelMatrix (utilities.h:188-189)
    ffm::M4 nrm;          // inverse(transpose(model)): object normal -> world normal, not normalised (kernel.cu:117)
    float normal_world[4];  // nrm * (0,0,1,0) for a plane
    int ord, geom, kind;  // kind: 1 plane, 0 sphere (GeometryType values)
    float radius;
};
struct DevScene {
    const float4* tris;
    const float4* nodes;
    const int* ord_geom;
    const int* ord_tri;
    const float4* materials;
    const DevPlane* planes;
    const float4* flat;          // the same 48-B records in flat (geometry, triangle) order; light sampling reads these
    const uint32_t* lights;      // flat indices of the triangles of emitter meshes, in flat order (next-event estimation)
    // shading attributes, all optional and cold: read once per shading event, never by traversal.
    //   geom_attr[g] = {first flat index of g, base into attr_normals (float4 units) or -1, base into attr_uvs (float2
    //   units) or -1, base into attr_facemat or -1}; corner k of triangle t of g lives at base + 3 t + k, t = flat - first.
    const int4* geom_attr;
    const float4* attr_normals;  // world-space corner normals inverse(transpose(M)) * n, not normalised (Triangle::m_n0..2, utilities.h:168-170)
    const float2* attr_uvs;      // corner uvs (Triangle::m_uv0..2, utilities.h:163-166)
    const uint32_t* attr_facemat;  // per triangle: row of `materials` (per-face OBJ material, utilities.h:837)
    uint32_t has_attr;           // 0: no geometry of the scene carries any of the above (shading skips the lookup)
    // binary LBVH kept for the BVH2 comparison kernel
    const int2* bvh2_children;   // internal node i: {left, right}; id >= n_internal is leaf (id - n_internal = slot2)
    const float4* bvh2_lo;       // per node id (internal then leaves): box min, w unused
    const float4* bvh2_hi;
    const uint32_t* bvh2_leaf_slot;  // sorted position -> slot in tris
    cudaTextureObject_t nodes_tex, tris_tex;   // linear float4 views of nodes / tris (kernel variants FF_NODE_TEX / FF_TRI_TEX)
    uint32_t n_tris, n_nodes, n_planes, n_geoms, n_lights;
    uint32_t wide_depth;
    uint32_t prmt_magic;         // 0x3F800000 (1.0f as float bits), see byte_unit in ff_trace.cuh
    float pad;                   // absolute inflation applied to every leaf box (conservative culling)
    uint32_t postpone_div;       // triangle groups are postponed when fewer than live/postpone_div lanes have any
    uint32_t refill_lanes;       // persistent kernel: refill idle lanes when fewer than this many still traverse
    float bounds_lo[3], bounds_hi[3];
};

struct HitRec {  // 16 B
    float key;      // reference closest-hit key: world distance (kernel.cu:113-115); +inf on a miss
    float t;        // t along the normalised direction (object-space t for planes)
    int slot;       // triangle slot in DevScene::tris; 0x40000000 | plane index for a plane; -1 on a miss
    int ord;        // tie-break order / index into ord_geom, ord_tri; -1 on a miss
};

struct CamDev {
    ffm::M4 inv_vp;      // inverse(view) * inverse(proj)   (kernel.cu:203)
    float pos[3];
    float width, height, far_clip;
    int w, h;
};

struct RenderDev {
    uint32_t seed_lo, seed_hi;
    int max_bounces, aa, nee;
    uint32_t first_sample;   // global sample index of local sample 0 of this pass
    uint32_t samples;        // samples in this pass
    uint32_t npix;           // pixels THIS call renders (the owned pixels under a tile split, else width * height)
    const uint32_t* pixel_map;  // tile split: rendered pixel ordinal -> pixel id of the frame; nullptr = identity
    float background[3];
    float ray_eps;
};

struct TraceCounters { unsigned long long nodes, tris; };

// What the wavefront's extend stage leaves for its shade stage: 8 bytes instead of the 16-byte API record (the closest-hit
// key and the scan-order id matter to callers of ff_trace_closest*, not to shading).
struct HitLite { float t; int slot; };
// Bounce 0 of a wavefront pass: every camera ray starts at the camera position and path p sits at queue index p, so the
// ray-generation kernel writes 16 bytes per path (the direction) instead of 64 (origin + path id, direction, throughput),
// and the extend and shade stages of bounce 0 read those 16 bytes and take the rest from the kernel parameters.
struct WaveGen { float ox, oy, oz; };

// ---------------------------------------------------------------- kernel launch wrappers
// (implemented in the .cu files; all asynchronous on `st`)
void k_bake_triangles(const float* d_xyz9_obj, const ffm::M4* d_model_per_geom, const uint32_t* d_geom_of_flat,
                      const uint32_t* d_ord_of_flat, uint32_t n, float4* d_flat_tris, int* d_bounds, cudaStream_t st);

// Per flat triangle: inverse(transpose(M)) * (normalize(cross(e1, e2)), 0) from OBJECT-space vertices: the normal the
// reference's frame shows (kernel.cu:101,117,183).  d_nrm_per_geom = inverse(transpose(model)) of every geometry.
void k_as_coded_normals(const float* d_xyz9_obj, const ffm::M4* d_nrm_per_geom, const uint32_t* d_geom_of_flat, uint32_t n,
                        float4* d_out, cudaStream_t st);

// Grow-only device scratch of a context, carved by a bump pointer: the builders take all their temporaries from it, so
// a rebuild (the interactive case, and every build after the first) allocates nothing -- cudaMalloc/cudaFree of a dozen
// arrays cost 10-30 ms per build, an order of magnitude more than the build kernels themselves.
struct Arena {
    char* base = nullptr;
    size_t cap = 0, off = 0;
    cudaError_t reserve(size_t bytes) {
        off = 0;
        if (bytes <= cap) return cudaSuccess;
        if (base) cudaFree(base);
        base = nullptr; cap = 0;
        cudaError_t e = cudaMalloc((void**)&base, bytes);
        if (e == cudaSuccess) cap = bytes;
        return e;
    }
    template <class T>
    T* take(size_t count) {
        off = (off + 255) & ~(size_t)255;
        T* p = (T*)(base + off);
        off += sizeof(T) * count;
        return off <= cap ? p : nullptr;
    }
    static size_t padded(size_t bytes) { return ((bytes + 255) & ~(size_t)255) + 256; }
    void release() { if (base) cudaFree(base); base = nullptr; cap = off = 0; }
};

struct BuildOut {
    uint32_t n_nodes, depth, morton_bits;
    double sah_cost;
};
// Builds into caller-allocated arrays; returns via *out after synchronising `st`.
cudaError_t bvh_build(const float4* d_flat_tris, uint32_t n, int builder, float pad, float4* d_tris_out,
                      float4* d_nodes_out, uint32_t node_capacity, int2* d_bvh2_children, float4* d_bvh2_lo,
                      float4* d_bvh2_hi, uint32_t* d_bvh2_leaf_slot, float bounds_lo[3], float bounds_hi[3],
                      BuildOut* out, uint32_t* launches, Arena& arena, cudaStream_t st, const int* d_bounds_ready = nullptr);
// d_bounds_ready (optional): 12 ordered ints already reduced over d_flat_tris by k_bake_triangles; the build then skips its bounds pass.

void k_trace_closest(const DevScene& sc, const float4* d_ray_o, const float4* d_ray_d, const uint32_t* d_count,
                     uint32_t n_max, int mode, HitRec* d_hits, TraceCounters* d_counters, uint32_t* d_work_counter, cudaStream_t st);
// Extend stage of the wavefront: as k_trace_closest, with 8-byte hit records, rays optionally generated from the path id
// (gen != nullptr: bounce 0; d_ray_o / d_ray_d are not read) and the self-skip slot read from ray_d.w (FF_TRACE_SELFSKIP).
void k_trace_extend(const DevScene& sc, const WaveGen* gen, const float4* d_ray_o, const float4* d_ray_d, const uint32_t* d_count,
                    uint32_t n_max, int mode, HitLite* d_hits, TraceCounters* d_counters, uint32_t* d_work_counter, cudaStream_t st);
void k_trace_any(const DevScene& sc, const float4* d_ray_o, const float4* d_ray_d, const float* d_tmax, uint32_t n, int mode,
                 unsigned char* d_occluded, uint32_t* d_work_counter, cudaStream_t st);
// Shadow stage of the wavefront: queue entries {o.xyz, path id}, {d.xyz, tmax}, {contribution.rgb}; unoccluded entries
// add their contribution to radiance[path].  *d_work_counter must be zero.
void k_trace_shadow(const DevScene& sc, const float4* d_ray_o, const float4* d_ray_d, const float4* d_contrib,
                    const uint32_t* d_count, uint32_t n_max, float4* d_radiance, uint32_t* d_work_counter, bool self_skip, cudaStream_t st);

void k_camera_rays(const CamDev& cam, float4* d_ray_o, float4* d_ray_d, cudaStream_t st);
void k_pack_rays(const float* d_rays6, size_t n, float4* d_ray_o, float4* d_ray_d, cudaStream_t st);
void k_unpack_rays(const float4* d_ray_o, const float4* d_ray_d, size_t n, float* d_rays6, cudaStream_t st);
// Host-batch streaming trace (ff_trace.cu, "Host batches"): device words shared by the kernel and the copy streams.
#define FF_STREAM_MAX_CHUNKS 256
struct StreamSync {       // zeroed before every launch
    uint32_t avail;       // rays landed so far (advanced by 4-byte copies on the copy-in stream); alone in its 128-byte line
    uint32_t pad0[31];
    uint32_t err;         // a lane timed out waiting for rays
    uint32_t pad1[31];
    uint32_t done[FF_STREAM_MAX_CHUNKS];   // finished rays per chunk
};
void k_trace_closest_stream(const DevScene& sc, const float* d_rays6, uint32_t n, uint32_t chunk_shift, StreamSync* d_sync, uint32_t* host_flags,
                            int* d_geom, int* d_tri, float* d_t, uint32_t* d_work_counter, cudaStream_t st);
void k_unpack_hits(const DevScene& sc, const HitRec* d_hits, size_t n, int* d_geom, int* d_tri, float* d_t,
                   cudaStream_t st);
void k_make_soup(size_t n, float s, float* d_xyz9, cudaStream_t st);
void k_make_rays_uniform(uint64_t first, size_t n, uint32_t salt, float4* d_ray_o, float4* d_ray_d, cudaStream_t st);
void k_make_rays_bounce(const DevScene& sc, const float4* d_ray_o, const float4* d_ray_d, const HitRec* d_hits, uint32_t n,
                        uint32_t seed_lo, uint32_t seed_hi, float4* d_out_o, float4* d_out_d, uint32_t* d_count, cudaStream_t st);

// wavefront / megakernel integrators.  Path state is SoA, one float4 array per field, double-buffered
// across bounces (queue compaction writes the survivors densely into the other buffer).
struct WaveBuffers {
    float4* ray_o[2];      // {o.xyz, path id}
    float4* ray_d[2];      // {d.xyz, 0}
    float4* beta[2];       // {throughput.rgb, 0}
    HitLite* hits;
    float4* radiance;      // per path of the pass: L.rgb, indexed by path id = sample_in_pass * npix + pixel
    float4* sh_o;          // shadow queue (next-event estimation): {origin.xyz, path id}
    float4* sh_d;          //   {unnormalised direction.xyz, tmax}
    float4* sh_c;          //   {radiance to add when unoccluded, 0}
    uint32_t* counts;      // 4 x [max_bounces + 2]: paths entering bounce b | extend work counters | shadow rays of bounce b | shadow work counters
    uint32_t capacity;     // paths per pass
};
void k_wavefront_pass(const DevScene& sc, const CamDev& cam, const RenderDev& rp, const WaveBuffers& wb, int trace_mode,
                      TraceCounters* d_counters, uint32_t* launches, cudaEvent_t* trace_events, int* n_trace_events,
                      cudaStream_t st);
void k_megakernel_pass(const DevScene& sc, const CamDev& cam, const RenderDev& rp, float4* d_radiance,
                       uint32_t* d_work_counter, unsigned long long* d_ray_counter, uint32_t* launches, cudaStream_t st);
void k_accumulate_pass(const float4* d_radiance, uint32_t npix, uint32_t samples, const uint32_t* d_pixel_map, float4* d_accum, cudaStream_t st);
void k_resolve(const float4* const* accums, int n_accums, uint32_t npix, float inv_spp, float4* d_out, cudaStream_t st);
// FF_REDUCE_PEER: this GPU's slice [first, first + count) of the frame, summed over `world` accumulators in rank order
// (pointers valid on this device: own memory or NVLink peers), resolved and stored into the root's frame.
void k_reduce_scatter_resolve(const float4* const* accums, int world, uint32_t first, uint32_t count, float inv_spp, float4* d_out_root, cudaStream_t st);
void k_as_coded(const DevScene& sc, const CamDev& cam, const float4* d_as_coded_normals, unsigned char* d_rgb, cudaStream_t st);
// Primary-ray feature buffers (FF_INTEGRATOR_AOV_*): shading normal, uv, barycentrics, albedo of the first hit.
void k_aov(const DevScene& sc, const CamDev& cam, int kind, float4* d_rgba, cudaStream_t st);
// Barycentrics of closest hits (u, v as in kernel.cu:80-81); misses and analytic primitives give 0.
void k_hit_barycentrics(const DevScene& sc, const float4* d_ray_o, const float4* d_ray_d, const HitRec* d_hits, size_t n,
                        float* d_u, float* d_v, cudaStream_t st);
void k_rgba_to_u8(const float4* d_rgba, uint32_t npix, unsigned char* d_rgb, cudaStream_t st);
void k_u8_to_rgba(const unsigned char* d_rgb, uint32_t npix, float4* d_rgba, cudaStream_t st);  // value / 255, alpha 1
void k_flush(float4* d_buf, size_t n_float4, cudaStream_t st);

// OBJ ingest (ff_obj.cpp): per-corner attributes in file order, 9 / 9 / 6 floats per triangle; face_material indexes
// material_names (-1: no usemtl in force).
struct ObjMesh {
    std::vector<float> xyz9, n9, uv6;
    std::vector<int> face_material;
    std::vector<std::string> material_names;
    std::string mtllib;
    bool has_normals = false, has_uvs = false;
    size_t bytes = 0;
};
bool load_obj_mesh(const char* path, ObjMesh& out, int threads);
bool load_obj_positions(const char* path, std::vector<float>& xyz9);
bool load_mtl(const char* path, std::vector<std::string>& names, std::vector<ff_material>& mats);
int obj_default_threads();

struct Tuning { int postpone_div, refill_lanes, blocks_per_sm; };
Tuning& tuning();
// Multiprocessors of the CURRENT device (cudaDevAttrMultiProcessorCount, cached per device): every grid is sized as a
// multiple of it (148 on B200).
int sm_count();

}  // namespace ff
