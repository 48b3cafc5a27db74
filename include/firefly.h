/* firefly.h -- C ABI of the B200-native path-tracing core (libfirefly_b200.so).
 *
 * Drop-in boundary for the one hot path of FireflyRenderEngine/GPUPathTracer: scene in (OBJ or
 * triangle arrays + TRS + material), camera in, render() / trace() out.  Plain pointers and sizes
 * only; no C++, glm or torch types.  The reference has no FFI of its own (the path is called inline
 * from main()), so each entry point cites the seam in the reference it replaces; paths are relative
 * to the reference tree, kernel.cu = PathTracer/FireflyEngine/kernel.cu, utilities.h likewise.
 *
 * Conventions: every int-returning function returns FF_OK (0) or a negative ff_status; the message
 * is available from ff_last_error() (thread-local).  Nothing here ever calls exit() (the reference's
 * cudaCheckErrors macro, utilities.h:27-37, does).  The library owns handles and all device memory
 * it allocates; the caller owns every buffer it passes in.  A context is bound to one CUDA device and
 * one internal stream and is not thread-safe; calls are synchronous on return unless stated.
 * There is no CPU fallback: without a CUDA device ff_context_create fails with FF_ERR_CUDA.
 */
#ifndef FIREFLY_H
#define FIREFLY_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FF_VERSION 100

typedef enum ff_status {
    FF_OK = 0,
    FF_ERR_INVALID = -1,     /* bad argument / handle state */
    FF_ERR_IO = -2,          /* file could not be read / written / parsed */
    FF_ERR_CUDA = -3,        /* CUDA runtime error (message has the cudaError string) */
    FF_ERR_NOMEM = -4,
    FF_ERR_UNSUPPORTED = -5
} ff_status;

typedef struct ff_context ff_context; /* replaces PathTracerState, kernel.h:12-20 */
typedef struct ff_scene ff_scene;     /* replaces Geometry[] + per-geometry Triangle[]/BXDF, kernel.cu:268-298 */

/* Camera, mirrors the fields main() sets (kernel.cu:311-322; struct at utilities.h:269-285).
 * forward/right/up are derived from yaw/pitch exactly like Camera::UpdateBasisAxis (utilities.h:407-418). */
typedef struct ff_camera {
    float pos[3];
    float world_up[3];
    float yaw_deg, pitch_deg, fov_deg;
    int width, height;
    float near_clip, far_clip;
} ff_camera;

/* Material, mirrors BXDF (utilities.h:77-88); type follows BXDFTyp (utilities.h:68-75). */
enum { FF_MAT_EMITTER = 0, FF_MAT_DIFFUSE = 1, FF_MAT_MIRROR = 2, FF_MAT_GLASS = 3 };
typedef struct ff_material {
    int type;
    float albedo[3];
    float emissive[3];
    float intensity;
    float specular[3];       /* m_specularColor: mirror reflectance, and the reflected lobe of glass */
    float ior;               /* m_refractiveIndex (glass); 0 is read as 1.5 */
    float transmittance[3];  /* m_transmittanceColor: the refracted lobe of glass */
} ff_material;

enum { FF_INTEGRATOR_WAVEFRONT = 0, FF_INTEGRATOR_MEGAKERNEL = 1, FF_INTEGRATOR_AS_CODED = 2,
       /* primary-ray feature buffers through ff_render (one ray per pixel, spp / bounces ignored): */
       FF_INTEGRATOR_AOV_NORMAL = 3, /* face-forwarded world shading normal (interpolated if the geometry has normals) */
       FF_INTEGRATOR_AOV_UV = 4,     /* interpolated Triangle::m_uv (utilities.h:163-166) */
       FF_INTEGRATOR_AOV_BARY = 5,   /* (1-u-v, u, v), the barycentrics of kernel.cu:80-81 */
       FF_INTEGRATOR_AOV_ALBEDO = 6  /* albedo of the hit's (per-face) material */ };
enum { FF_BUILDER_LBVH = 0, FF_BUILDER_SAH = 1 };
enum { FF_TRACE_BVH8 = 0, FF_TRACE_BRUTE = 1, FF_TRACE_BVH2 = 2 };

typedef struct ff_render_params {
    int spp;                 /* samples per pixel rendered by THIS call */
    int max_bounces;         /* path segments = max_bounces + 1 */
    uint64_t seed;           /* Philox key {lo, hi}; the reference's constant is 1234 (utilities.h:118) */
    int integrator;          /* FF_INTEGRATOR_*, optionally OR FF_TRACE_COUNT (instrumented extend kernels) */
    int nee;                 /* next-event estimation with shadow rays (0 = plain loop) */
    int first_sample;        /* global index of the first sample (multi-GPU sample split, resume) */
    int aa;                  /* 0: integer pixel corner like kernel.cu:200-201; 1: Philox jitter */
    float background[3];     /* radiance of a miss; the reference leaves misses black (kernel.cu:340) */
    float ray_eps;           /* offset of spawned rays along the normal */
    int samples_per_pass;    /* samples in flight per wavefront pass; 0 = about 64 Mi paths (~150 B of device state each), at most 128 Mi */
    int total_spp;           /* divisor of the final mean; 0 = spp (single GPU) */
    /* interleaved screen tiles (multi-GPU split for low-spp / interactive frames): with tile_world > 1 this call
     * renders only the pixels of the tile_size x tile_size tiles t (row-major) with t % tile_world == tile_rank and leaves
     * every other pixel of the accumulator untouched, so the SUM of the ranks' accumulators is the full frame and is
     * bit-identical to a single-GPU render whatever tile_world is.  tile_size 0 = 32.  tile_world <= 1: the whole image. */
    int tile_size, tile_rank, tile_world;
} ff_render_params;

/* Per-call measurements, all taken with CUDA events on the context's stream. */
typedef struct ff_stats {
    double gpu_ms;           /* device time of the whole call's kernels */
    double trace_ms;         /* of which closest-hit / any-hit traversal kernels */
    double h2d_ms, d2h_ms;   /* host<->device copies made by the call (0 for *_device entry points) */
    uint64_t rays;           /* closest-hit queries executed */
    uint64_t shadow_rays;    /* any-hit queries executed */
    uint64_t paths;          /* complete samples */
    uint64_t nodes_visited;  /* instrumented runs only (ff_trace flag FF_TRACE_COUNT), else 0 */
    uint64_t tris_tested;    /* idem */
    uint32_t kernel_launches;
    uint32_t passes;         /* render calls: wavefront / megakernel passes actually run (spp / the samples-per-pass the library chose,
                              * which is smaller than the default when the path state did not fit); trace calls: pipeline chunks */
} ff_stats;

typedef struct ff_scene_info {
    uint64_t n_triangles, n_geometries, n_planes;
    uint64_t n_nodes;        /* 8-wide compressed nodes (80 B each) */
    uint64_t node_bytes, triangle_bytes;
    uint32_t depth;          /* depth of the wide hierarchy */
    uint32_t builder;
    float bounds_lo[3], bounds_hi[3];
    double build_ms;         /* device time of the last commit (sort + hierarchy + refit + collapse) */
    double sah_cost;         /* SAH cost of the wide hierarchy, for comparing builders */
    uint32_t morton_bits;    /* width of the Morton keys the last commit sorted: 30, or 63 when the 30-bit grid left more than
                              * 1/16 of a large scene's triangles sharing a cell with their neighbour; 0 without triangles */
    uint32_t reserved;
} ff_scene_info;

const char* ff_last_error(void);
int ff_version(void);
int ff_device_count(void);

/* -- context: device + stream.  Replaces the implicit default device/stream of main(). */
int ff_context_create(ff_context** out, int device_id);
void ff_context_destroy(ff_context* ctx);
int ff_context_device(const ff_context* ctx);
int ff_context_synchronize(ff_context* ctx);
/* The builders keep their temporaries (~45 B/triangle LBVH, ~120 B/triangle SAH) in a grow-only scratch owned by
 * the context so that rebuilds allocate nothing; this returns it to the driver. */
int ff_context_release_scratch(ff_context* ctx);
/* CUDA-event bracket on the context's stream: device-clock time of everything issued between the two calls
 * (kernels, copies and any gap while another stream or library, e.g. an NCCL reduce, is working).  The
 * reference has no timing of its own; this is what bench.py reports. */
int ff_timer_begin(ff_context* ctx);
int ff_timer_end(ff_context* ctx, double* out_ms);

/* -- scene building.  LoadMesh (utilities.h:781-840) + Geometry ctor (utilities.h:176-213) + the
 *    m_bxdf assignment (kernel.cu:246-250).  model = T * Rx * Ry * Rz * S, rotations in degrees.
 *    Geometry indices are assigned in call order, like the std::vector in kernel.cu:253-258. */
int ff_scene_create(ff_context* ctx, ff_scene** out);
void ff_scene_destroy(ff_scene* scene);
int ff_scene_add_obj(ff_scene* scene, const char* obj_path, const float pos[3], const float rot_deg[3],
                     const float scale[3], const ff_material* material, int flip_winding,
                     int* out_geometry_index);
int ff_scene_add_triangles(ff_scene* scene, const float* xyz9, size_t n_tris, const float pos[3],
                           const float rot_deg[3], const float scale[3], const ff_material* material,
                           int* out_geometry_index);
/* GeometryType::PLANE (utilities.h:141-146, intersectPlane kernel.cu:8-32): unit square at object z=0. */
int ff_scene_add_plane(ff_scene* scene, const float pos[3], const float rot_deg[3], const float scale[3],
                       const ff_material* material, int* out_geometry_index);
/* GeometryType::SPHERE (utilities.h:143, m_sphereRadius utilities.h:227; the reference leaves its intersection branch
 * empty, kernel.cu:166-169): analytic sphere of `radius` at the object-space origin, front side only like every
 * triangle (a ray that starts inside sees nothing). */
int ff_scene_add_sphere(ff_scene* scene, float radius, const float pos[3], const float rot_deg[3], const float scale[3],
                        const ff_material* material, int* out_geometry_index);
/* -- shading attributes the reference's loader carries and its tracer ignores: per-corner normals and
 *    uvs (Triangle::m_n0..2 / m_uv0..2, utilities.h:163-170) and per-face materials (material_ids, read and dropped at
 *    utilities.h:837).  All are optional, per triangle-mesh geometry, set between add and commit, and never touched by
 *    traversal: hit ids and t are unaffected.
 *    normals9: n_tris x 3 corners x xyz, object space.  A geometry with normals is shaded with the interpolated normal
 *    (BSDF frame, cosines); ray offsets and sidedness keep the geometric normal; triangles whose interpolated normal is
 *    zero fall back to the geometric one.  uvs6: n_tris x 3 corners x uv.  face_material: n_tris indices into
 *    `materials` (-1 = the geometry's own material).  NULL clears the attribute. */
int ff_scene_set_normals(ff_scene* scene, int geometry_index, const float* normals9);
int ff_scene_set_uvs(ff_scene* scene, int geometry_index, const float* uvs6);
int ff_scene_set_face_materials(ff_scene* scene, int geometry_index, const int* face_material, const ff_material* materials,
                                int n_materials);
/* ff_scene_add_obj with the attributes of the file.  flags: */
enum { FF_OBJ_FLIP_WINDING = 1, /* swap v1/v2 (and their normals / uvs) of every face */
       FF_OBJ_NORMALS = 2,      /* keep `vn`: smooth shading */
       FF_OBJ_UVS = 4,          /* keep `vt` */
       FF_OBJ_MATERIALS = 8     /* per-face `usemtl`, materials from the file's `mtllib` (searched next to the OBJ); a name
                                 * the library does not define, or a missing library, keeps the geometry's material */ };
int ff_scene_add_obj_ex(ff_scene* scene, const char* obj_path, const float pos[3], const float rot_deg[3],
                        const float scale[3], const ff_material* material, int flags, int* out_geometry_index);
/* The LoadMesh half on its own, all attributes, parsed on `threads` host threads (0 = all).  Any output may be NULL;
 * capacity_tris bounds what is written.  Returns the triangle count or a negative ff_status.  out_seconds: wall time
 * of read + parse (the ingest figure of SURVEY 8f f4). */
long ff_load_obj_ex(const char* obj_path, float* xyz9, float* normals9, float* uvs6, int* face_material,
                    size_t capacity_tris, int threads, int* out_has_normals, int* out_has_uvs, int* out_n_materials,
                    double* out_seconds);
/* OBJ -> xyz9 triangle list without a scene (the LoadMesh half on its own).  Call with xyz9 = NULL
 * to get the count. Returns the triangle count or a negative ff_status. */
long ff_load_obj(const char* obj_path, float* xyz9, size_t capacity_tris);

/* Flatten to world-space SoA, upload, build the BVH on the device.  Replaces kernel.cu:268-298. */
int ff_scene_commit(ff_scene* scene, int builder, double* out_build_ms);
int ff_scene_get_info(const ff_scene* scene, ff_scene_info* out);
/* World-space (baked) triangles in flat order = (geometry, triangle) order, and their source ids. */
int ff_scene_get_flat(const ff_scene* scene, float* xyz9, int* geom, int* tri, size_t capacity_tris);
int ff_scene_get_matrices(const ff_scene* scene, int geometry_index, float model16[16], float inverse16[16]);

/* -- closest hit on caller-supplied ray batches: the parity / benchmark entry.  Replaces
 *    intersectRays (kernel.cu:127-176).  rays6 = n x {ox,oy,oz,dx,dy,dz}; directions need not be
 *    normalised (the routine normalises, kernel.cu:138).  Outputs use the reference's numbering:
 *    geometryIndex, triangleIndex (-1,-1 on a miss; triangleIndex -1 for a plane), t = the
 *    reference's closest-hit key (world-space distance, kernel.cu:113-121), +inf on a miss. */
#define FF_TRACE_COUNT 0x100 /* OR into `mode`: instrumented kernel, fills nodes_visited / tris_tested */
int ff_trace_closest(ff_scene* scene, const float* rays6, size_t n, int mode, int* out_geom, int* out_tri,
                     float* out_t, ff_stats* stats);
/*    Host batches are traced by ONE persistent launch that starts before the rays have arrived: the copy-in stream delivers
 *    them in chunks of 512 Ki rays and advances a device word after each, lanes that draw a ray beyond it wait; results leave
 *    in the reference's numbering, and a chunk whose rays are all done is handed to the copy-out stream while the launch is
 *    still running (device staging: 36 bytes per ray, at most 16 Mi rays at a time, owned by the context and reused).  With
 *    page-locked host buffers (ff_host_alloc_pinned) both PCIe directions overlap the traversal; pageable buffers work and are
 *    simply slower.  FF_TRACE_COUNT, the other modes and ff_trace_closest_uv use a chunked three-stage pipeline (copy in |
 *    pack + trace + unpack | copy out, ~1 Mi rays per chunk), and so does a context on which the streaming launch has twice
 *    timed out waiting for rays (a tool that serialises kernels against copies: a replaying profiler, CUDA_LAUNCH_BLOCKING=1;
 *    the batch is then redone, nothing is lost).  FF_TRACE_STREAM=0 in the environment selects the chunked pipeline outright.
 *    stats: gpu_ms = device time from the first copy to the last, trace_ms = traversal launches only. */
/* Same with device-resident buffers (no copies): d_rays = 2n float4, SoA: n origins {o.xyz, 0} followed by
 * n directions {d.xyz, 0}; d_hits = n 16-byte records {key t (f32), t along the ray (f32), triangle slot
 * (i32), scan-order id (i32)} -- see ff_hits_unpack. */
int ff_trace_closest_device(ff_scene* scene, const void* d_rays, size_t n, int mode, void* d_hits,
                            ff_stats* stats);
/* Any-hit (occlusion) query: out_occluded[i] = 1 when a front-facing triangle is hit with key < tmax[i]. */
int ff_trace_any(ff_scene* scene, const float* rays6, const float* tmax, size_t n, int mode,
                 unsigned char* out_occluded, ff_stats* stats);
/* Unpack device hit records into the reference's numbering (host arrays, may be NULL). */
/* ff_trace_closest that also returns the barycentrics of the hit (u, v weight v1, v2; kernel.cu:80-81, which the
 * reference computes and drops): what a caller needs to interpolate its own per-corner data.  0 on misses / planes. */
int ff_trace_closest_uv(ff_scene* scene, const float* rays6, size_t n, int mode, int* out_geom, int* out_tri,
                        float* out_t, float* out_u, float* out_v, ff_stats* stats);
int ff_hits_unpack(ff_scene* scene, const void* d_hits, size_t n, int* out_geom, int* out_tri, float* out_t);
/* ff_hits_unpack without allocation.  No counterpart in the reference: added for
 * callers that may not call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same
 * arguments, same outputs bit for bit; the three device temporaries ff_hits_unpack allocates per call live instead in a
 * static device array of FF_HITS_UNPACK_STATIC_MAX records (12 bytes each = 12 MiB, held by every GPU that loads the
 * library until the process exits).  Calls no cudaMalloc / cudaFree / allocating routine.  n above the maximum is
 * FF_ERR_INVALID (split the batch, or use ff_hits_unpack).  Every call on a device shares the array: calls that overlap
 * on one GPU (two host threads, two contexts of one device) race -- serialise them.  Synchronous like ff_hits_unpack. */
#define FF_HITS_UNPACK_STATIC_MAX (1u << 20)
int ff_hits_unpack_static(ff_scene* scene, const void* d_hits, size_t n, int* out_geom, int* out_tri, float* out_t);

/* -- camera rays exactly as generateRays builds them (kernel.cu:198-205), host or device output. */
int ff_camera_rays(ff_context* ctx, const ff_camera* cam, float* rays6_host, void* d_rays);
int ff_camera_basis(const ff_camera* cam, float forward[3], float right[3], float up[3]);
int ff_camera_inverse_matrices(const ff_camera* cam, float inv_view16[16], float inv_proj16[16]);

/* -- render.  Replaces launchPathTrace<<<>>>(pbo, geometries, camera, n) (kernel.cu:218-221,342)
 *    with the bounce loop the reference intends (readme.md:39-46, BXDF utilities.h:90-139).
 *    ff_render: mean radiance, RGBA fp32, row 0 = top (kernel orientation), to a host buffer and/or a
 *    caller-owned device buffer.  ff_render_u8: byte-compatible with the uchar3 PBO the viewer maps
 *    (kernel.cu:337-344, 349-351): W*H*3, tightly packed. */
int ff_render(ff_scene* scene, const ff_camera* cam, const ff_render_params* params, float* out_rgba_host,
              void* out_rgba_device, ff_stats* stats);
int ff_render_u8(ff_scene* scene, const ff_camera* cam, const ff_render_params* params,
                 unsigned char* out_rgb, int out_is_device, ff_stats* stats);
/*    With FF_INTEGRATOR_AS_CODED the frame is the reference's own (kernel.cu:178-216), quirks included: the colour is
 *    |inverse(transpose(M)) * n| for the OBJECT-space unit normal n of the hit triangle, not re-normalised (a mesh scaled by s
 *    is 1/s as bright, kernel.cu:117), and make_uchar3 converts with cvt.rzi.u32.f32 + an 8-bit store, so a component above
 *    255 wraps modulo 256 (kernel.cu:214).  Byte-identical to the reference's frame wherever the same primitive is hit
 *    (tests/golden/ref_main.npz, ref_scaled.npz).  With the path-tracing integrators the radiance is clamped to [0, 1]. */
/* Multi-GPU building blocks: add this call's samples (as SUMS, not means) into a caller-owned
 * device accumulator of W*H float4 (zero it first with ff_device_memset), then after the
 * framebuffer reduce, divide by the total sample count. */
int ff_render_accumulate(ff_scene* scene, const ff_camera* cam, const ff_render_params* params,
                         void* d_accum_rgba, ff_stats* stats);
int ff_resolve(ff_context* ctx, const void* d_accum_rgba, int width, int height, int total_spp,
               void* d_out_rgba, float* out_rgba_host);
/* Fused peer reduce + resolve: sums n_peers accumulators (device pointers valid on this device, e.g.
 * CUDA-IPC mapped peers over NVLink) in rank order and divides by total_spp in one kernel. */
int ff_reduce_resolve_peers(ff_context* ctx, const void* const* d_accum_peers, int n_peers, int width,
                            int height, int total_spp, void* d_out_rgba, float* out_rgba_host);

/* -- multi-GPU group (BASELINE.json north_star).  The reference is single-GPU: its `multi_gpu` flag is dead
 *    (utilities.h:484-487).  The image is partitioned over the GPUs of one node by disjoint sample ranges or by interleaved
 *    screen tiles, the scene + BVH are replicated (commit the same scene on every context), and the per-GPU accumulators are
 *    combined by ONE framebuffer reduce per frame over NVLink, inside the library.  Philox is keyed by the GLOBAL sample
 *    index, so the set of paths does not depend on the number of GPUs.
 *      ff_group_create       one process drives n devices: one context (ff_group_context) and one host thread per device
 *      ff_group_create_rank  one process per GPU (torchrun / MPI style): rank 0 makes an id with ff_group_unique_id and hands
 *                            its FF_GROUP_ID_BYTES to the other ranks by any means; every rank passes its own context.
 *                            Collective, like every ff_group_render of such a group.  world 1 needs no id (may be NULL).
 *    NCCL (libnccl.so.2) is loaded at run time when a group of more than one GPU is formed; FF_ERR_UNSUPPORTED without it. */
typedef struct ff_group ff_group;
#define FF_GROUP_ID_BYTES 128
enum { FF_SPLIT_SAMPLES = 0,  /* rank k renders samples [k*spp/G, (k+1)*spp/G) of every pixel (ff_group_sample_range) */
       FF_SPLIT_TILES = 1 };  /* rank k renders all samples of the 32x32 tiles t with t % G == k (ff_render_params.tile_*):
                               * bit-identical to one GPU for every G */
enum { FF_REDUCE_NCCL = 0,    /* ncclReduce(SUM) of the accumulators to the root + resolve kernel */
       FF_REDUCE_PEER = 1 };  /* one fused kernel per GPU: reduce-scatter + resolve + gather over NVLink peer memory (peer
                               * access in one process, CUDA IPC between processes); summation in rank order, so the frame is
                               * bit-reproducible for a given G.  FF_ERR_UNSUPPORTED when the GPUs have no peer path */
int ff_group_create(const int* device_ids, int n, ff_group** out);
int ff_group_unique_id(unsigned char id[FF_GROUP_ID_BYTES]);
int ff_group_create_rank(ff_context* ctx, const unsigned char id[FF_GROUP_ID_BYTES], int rank, int world, ff_group** out);
void ff_group_destroy(ff_group* group);   /* destroys the contexts of ff_group_create (destroy their scenes first); never the one passed to _create_rank */
int ff_group_world(const ff_group* group);                    /* GPUs in the group */
int ff_group_local_count(const ff_group* group);              /* contexts this process drives: n, or 1 */
int ff_group_rank(const ff_group* group, int local_index);    /* global rank of a local context; rank 0 is the root */
ff_context* ff_group_context(ff_group* group, int local_index);
int ff_group_nccl_version(void);                              /* e.g. 22703, 0 when NCCL cannot be loaded */
int ff_group_sample_range(int rank, int world, int spp, int* out_first, int* out_count);
/* One frame.  scenes[i] = the committed scene on the group's i-th local context.  params->spp is the frame's TOTAL sample
 * count (the mean divides by params->total_spp when that is set: accumulating frames), first_sample the global index of its
 * first sample.  The resolved frame (mean radiance, RGBA fp32, as ff_render) arrives on the root: out_rgba_host and / or
 * out_rgba_device (memory of the root's device); other ranks of a one-process-per-GPU group pass NULL.  stats = one ff_stats
 * per local context (may be NULL); *out_reduce_ms = device time of reduce + resolve, max over the local contexts. */
int ff_group_render(ff_group* group, ff_scene* const* scenes, const ff_camera* cam, const ff_render_params* params, int split,
                    int reduce, float* out_rgba_host, void* out_rgba_device, ff_stats* stats, double* out_reduce_ms);

/* -- synthetic workloads of BASELINE.json configs 4/5, generated on the device with the same
 *    Philox streams as the oracle (bit-identical to the host generators). */
int ff_make_soup(ff_context* ctx, size_t n_tris, float s, float* xyz9_host);
int ff_make_rays_uniform(ff_context* ctx, uint64_t first, size_t n, uint32_t salt, float* rays6_host,
                         void* d_rays);
/* Diffuse-bounce batch: cosine-hemisphere rays leaving the given hits (batch B of config 4). */
int ff_make_rays_bounce(ff_scene* scene, const void* d_rays_in, const void* d_hits_in, size_t n,
                        uint64_t seed, void* d_rays_out, size_t* out_n);

/* -- device memory helpers so callers can keep batches resident without another CUDA binding. */
int ff_device_malloc(ff_context* ctx, size_t bytes, void** out);
int ff_device_free(ff_context* ctx, void* p);
int ff_device_memset(ff_context* ctx, void* p, int value, size_t bytes);
int ff_memcpy_h2d(ff_context* ctx, void* dst_device, const void* src_host, size_t bytes);
int ff_memcpy_d2h(ff_context* ctx, void* dst_host, const void* src_device, size_t bytes);
int ff_host_alloc_pinned(size_t bytes, void** out);
int ff_host_free_pinned(void* p);
int ff_ipc_get_handle(ff_context* ctx, void* d_ptr, unsigned char handle64[64]);
int ff_ipc_open_handle(ff_context* ctx, const unsigned char handle64[64], void** out_d_ptr);
int ff_ipc_close_handle(ff_context* ctx, void* d_ptr);
/* Writes > L2-size scratch so the next timed launch starts cold. */
int ff_flush_l2(ff_context* ctx);

/* -- scene description file: replaces the scene and camera literals of main() (kernel.cu:227-259, 311-322; its
 *    "TODO: Load scene from file", kernel.cu:261).  The format is documented at the top of csrc/ff_scenefile.cpp and
 *    in INTEGRATION.md: camera / material / mesh / plane / sphere / triangles / soup / render directives, one per line.
 *    ff_scene_load_file creates the scene (NOT committed: call ff_scene_commit(scene, desc.builder, ..)) and returns
 *    the camera and render parameters the file asks for.  Relative mesh paths resolve against asset_dir, or the
 *    scene file's own directory when asset_dir is NULL.  ff_scene_file_validate only parses (no device needed). */
typedef struct ff_scene_file {
    ff_camera camera;
    ff_render_params params;
    int builder;
    int n_geometries;
} ff_scene_file;
int ff_scene_load_file(ff_context* ctx, const char* path, const char* asset_dir, ff_scene** out_scene, ff_scene_file* out_desc);
int ff_scene_file_validate(const char* path, ff_scene_file* out_desc);

/* -- the pixels rank `tile_rank` of `tile_world` owns under the interleaved-tile split above, ascending (row-major inside
 *    each tile, tiles in row-major order); returns the count (out may be NULL to query it), negative on bad arguments. */
long ff_tile_owned_pixels(int width, int height, int tile_size, int tile_rank, int tile_world, uint32_t* out, size_t capacity);

/* -- camera controls: Camera::ProcessKeyboard / ProcessMouseMovement / resetCamera (utilities.h:343-427) applied to an
 *    ff_camera, so a front end without the GLFW callbacks (utilities.h:858-893) moves the camera exactly like the viewer.
 *    direction = the reference's codes; a step is m_cameraMouseSensitivity = 0.2 (utilities.h:345: the reference steps by
 *    the mouse sensitivity, not by m_cameraMovementSpeed), a turn is 0.5 degrees. */
enum { FF_CAM_FORWARD = 0, FF_CAM_BACKWARD = 1, FF_CAM_LEFT = 2, FF_CAM_RIGHT = 3, FF_CAM_UP = 4, FF_CAM_DOWN = 5,
       FF_CAM_YAW_LEFT = 6, FF_CAM_YAW_RIGHT = 7, FF_CAM_PITCH_UP = 8, FF_CAM_PITCH_DOWN = 9, FF_CAM_RESET = 10 };
int ff_camera_process_keyboard(ff_camera* cam, int direction);
/*    SetClickPosition(x_delta, y_delta) followed by ProcessMouseMovement(constrain_pitch) (utilities.h:385-404). */
int ff_camera_process_mouse(ff_camera* cam, float x_delta, float y_delta, int constrain_pitch);

/* -- progressive accumulation: the state the viewer loop (kernel.cu:331-359) would carry from frame to frame once the
 *    tracer accumulates.  Each ff_progressive_frame adds params->spp samples (global sample indices continue from the
 *    frames before, so k frames of s samples are bit-identical to one ff_render of k*s samples) and resolves the mean;
 *    the accumulation restarts by itself when the camera, the image size or a parameter that changes the image
 *    (max_bounces, seed, integrator, nee, aa, background, ray_eps) differs from the previous frame, and after
 *    ff_progressive_reset (call it after re-committing the scene).  Outputs as ff_render (float RGBA, host and/or
 *    device) plus an optional 8-bit frame with the PBO layout of ff_render_u8 (kernel.cu:214). */
typedef struct ff_progressive ff_progressive;
int ff_progressive_create(ff_scene* scene, ff_progressive** out);
void ff_progressive_destroy(ff_progressive* prog);
int ff_progressive_reset(ff_progressive* prog);
int ff_progressive_frame(ff_progressive* prog, const ff_camera* cam, const ff_render_params* params, float* out_rgba_host,
                         void* out_rgba_device, unsigned char* out_rgb_u8, int u8_is_device, int* out_total_spp, ff_stats* stats);

/* -- headless output.  Replaces saveToPPM (utilities.h:842-856).  Format by extension:
 *    .ppm (binary P6), .png (8-bit sRGB), .exr (fp32 scanline, uncompressed), .pfm. */
int ff_write_image(const char* path, const float* rgba, int width, int height);
int ff_write_image_u8(const char* path, const unsigned char* rgb, int width, int height);

/* -- development knobs of the traversal scheduler (take effect at the next ff_scene_commit); <= 0 keeps a value. */
int ff_debug_set_tuning(int postpone_div, int refill_lanes, int blocks_per_sm);
/* -- development statistics of the committed 8-wide hierarchy: child_hist[k] = nodes with k children (k = 0..8),
 *    leaf_tri_hist[k] = leaf children holding k triangles (k = 1..3). */
int ff_debug_node_stats(const ff_scene* scene, unsigned long long child_hist[9], unsigned long long leaf_tri_hist[4]);

/* -- Philox4x32-10 on the host (the generator all kernels use), for cross-checks. */
void ff_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);

#ifdef __cplusplus
}
#endif
#endif /* FIREFLY_H */
