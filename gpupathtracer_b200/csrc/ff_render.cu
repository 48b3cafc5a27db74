// ff_render.cu -- the two integrators (wavefront stages; persistent megakernel), the as-coded normal shader,
// sample accumulation and the resolve / peer-reduce kernels.
//
// Replaces launchPathTrace / generateRays / shade (reference kernel.cu:178-221) and implements the bounce loop
// the reference intends but stops short of (readme.md:39-46; BXDF::bsdf utilities.h:90-129), exactly as it is
// written down in DESIGN.md section 5 and restated on the CPU in oracle/firefly_oracle.cpp (trace_path).
// Every arithmetic step uses ffm:: single-rounding operations, and samples are summed per pixel in sample-index
// order, so a single-GPU render is bit-identical to the CPU restatement.
#include "ff_trace.cuh"

namespace ff {

using namespace ffm;

static inline int grid_for(size_t n, int block, int max_blocks) {
    size_t g = (n + block - 1) / block;
    if (g < 1) g = 1;
    if (g > (size_t)max_blocks) g = max_blocks;
    return (int)g;
}

// Surface data of a hit: world point, geometric normal as the reference returns it, geometry index.
__device__ __forceinline__ void hit_surface(const DevScene& sc, const HitRec& h, V3 o, V3 d_raw, V3& point, V3& normal, int& geom) {
    if (h.slot & 0x40000000) {
        const DevPlane& pl = sc.planes[h.slot & 0x3fffffff];
        float key, tl;
        V3 lp;
        plane_hit(pl, o, d_raw, key, tl, point, lp);
        if (pl.kind == 1) normal = v3(pl.normal_world[0], pl.normal_world[1], pl.normal_world[2]);
        else { const V4 nw = mul_mv(pl.nrm, lp.x, lp.y, lp.z, 0.0f); normal = v3(nw.x, nw.y, nw.z); }
        geom = pl.geom;
    } else {
        const float4 a = __ldg(sc.tris + 3 * (size_t)h.slot), b = __ldg(sc.tris + 3 * (size_t)h.slot + 1),
                     c = __ldg(sc.tris + 3 * (size_t)h.slot + 2);
        const V3 dn = normalize(d_raw);
        point = add(o, mul(dn, h.t));                                    // kernel.cu:99
        normal = normalize(cross(sub(ld3(b), ld3(a)), sub(ld3(c), ld3(a))));  // kernel.cu:101
        geom = __float_as_int(b.w);
    }
}

// Shading attributes of a triangle hit: the per-corner normals and uvs LoadMesh keeps and the tracer never
// reads (Triangle::m_n0..2, m_uv0..2, utilities.h:163-170), interpolated with the barycentrics intersectTriangle computes
// and drops (kernel.cu:80-81), and the per-face material id LoadMesh reads and drops (utilities.h:837).  CPU twin:
// surface_attributes in oracle/firefly_oracle.cpp, same operations in the same order.
struct SurfAttr {
    int mat;        // row of sc.materials
    bool smooth;    // ns holds a usable interpolated normal
    V3 ns;          // normalised interpolated world normal (not yet face-forwarded)
    float u, v;     // barycentrics (weights of v1, v2)
    float tu, tv;   // interpolated uv
};
__device__ __forceinline__ int material_row(const DevScene& sc, int geom, int flat) {
    if (!sc.has_attr) return geom;
    const int4 ga = __ldg(sc.geom_attr + geom);
    return ga.w >= 0 ? (int)__ldg(sc.attr_facemat + ga.w + (flat - ga.x)) : geom;
}
__device__ __forceinline__ void surface_attributes(const DevScene& sc, const HitRec& h, V3 o, V3 d_raw, int geom, bool want_bary, SurfAttr& a) {
    a.mat = geom; a.smooth = false; a.ns = v3(0.f, 0.f, 0.f); a.u = a.v = a.tu = a.tv = 0.f;
    if (h.slot < 0 || (h.slot & 0x40000000)) return;
    int4 ga = make_int4(0, -1, -1, -1);
    if (sc.has_attr) ga = __ldg(sc.geom_attr + geom);
    if (!want_bary && ga.y < 0 && ga.z < 0 && ga.w < 0) return;
    const float4 a4 = __ldg(sc.tris + 3 * (size_t)h.slot), b4 = __ldg(sc.tris + 3 * (size_t)h.slot + 1),
                 c4 = __ldg(sc.tris + 3 * (size_t)h.slot + 2);
    const int tri = __float_as_int(c4.w) - ga.x;
    if (ga.w >= 0) a.mat = (int)__ldg(sc.attr_facemat + ga.w + tri);
    if (!want_bary && ga.y < 0 && ga.z < 0) return;
    triangle_barycentrics(ld3(a4), ld3(b4), ld3(c4), o, normalize(d_raw), a.u, a.v);
    const float w0 = fsub(fsub(1.0f, a.u), a.v);
    if (ga.y >= 0) {
        const float4 n0 = __ldg(sc.attr_normals + ga.y + 3 * (size_t)tri), n1 = __ldg(sc.attr_normals + ga.y + 3 * (size_t)tri + 1),
                     n2 = __ldg(sc.attr_normals + ga.y + 3 * (size_t)tri + 2);
        const V3 ns = add(add(mul(ld3(n0), w0), mul(ld3(n1), a.u)), mul(ld3(n2), a.v));
        if (dot(ns, ns) > 0.0f) { a.smooth = true; a.ns = normalize(ns); }
    }
    if (ga.z >= 0) {
        const float2 t0 = __ldg(sc.attr_uvs + ga.z + 3 * (size_t)tri), t1 = __ldg(sc.attr_uvs + ga.z + 3 * (size_t)tri + 1),
                     t2 = __ldg(sc.attr_uvs + ga.z + 3 * (size_t)tri + 2);
        a.tu = fadd(fadd(fmul(t0.x, w0), fmul(t1.x, a.u)), fmul(t2.x, a.v));
        a.tv = fadd(fadd(fmul(t0.y, w0), fmul(t1.y, a.u)), fmul(t2.y, a.v));
    }
}

// Next-event estimation, one light sample per diffuse shading event (CPU twin: sample_light in
// oracle/firefly_oracle.cpp, same operations in the same order).  Lights are the triangles of emitter meshes; one is
// picked uniformly, a point uniformly on it; it radiates from its front side only (the side closest-hit can see).
struct ShadowReq {
    bool valid;
    V3 o, d;       // shadow ray: origin, unnormalised direction to the light sample
    float tmax;    // occluded <=> any front-facing hit with key < tmax
    V3 contrib;    // radiance added to the path when unoccluded
};
__device__ __forceinline__ void sample_light(const DevScene& sc, const RenderDev& rp, uint32_t pixel, uint32_t sample, int bounce,
                                             V3 so, V3 n, V3 ng, bool smooth, V3 beta, ShadowReq& sr) {
    sr.valid = false;
    const uint32_t nl = sc.n_lights;
    if (nl == 0) return;
    const U4 r = philox4x32_10(U4{pixel, sample, (uint32_t)bounce, 2u}, rp.seed_lo, rp.seed_hi);
    uint32_t k = (uint32_t)fmul(u01(r.x), (float)nl);
    if (k > nl - 1) k = nl - 1;
    const uint32_t fi = __ldg(sc.lights + k);
    const float4 a4 = __ldg(sc.flat + 3 * (size_t)fi), b4 = __ldg(sc.flat + 3 * (size_t)fi + 1), c4 = __ldg(sc.flat + 3 * (size_t)fi + 2);
    const V3 v0 = ld3(a4), v1 = ld3(b4), v2 = ld3(c4);
    const float su = fsqrt(u01(r.y));
    const float b0 = fsub(1.0f, su), b1 = fmul(u01(r.z), su);
    const float b2 = fsub(fsub(1.0f, b0), b1);
    const V3 y = add(add(mul(v0, b0), mul(v1, b1)), mul(v2, b2));
    const V3 N = cross(sub(v1, v0), sub(v2, v0));
    const float len = length(N);
    const V3 sd = sub(y, so);
    const float dist2 = dot(sd, sd);
    const float dist = fsqrt(dist2);
    const float cosx = fdiv(dot(n, sd), dist);
    const float cosl = fdiv(fdiv(-dot(N, sd), dist), len);
    if (!(cosx > 0.0f && cosl > 0.0f && dist2 > 0.0f)) return;
    if (smooth && !(dot(ng, sd) > 0.0f)) return;  // below the real surface: the shadow ray would leave through culled back faces
    const float w = fmul(fmul(fdiv(fmul(cosx, cosl), dist2), fmul(0.5f, len)), fmul((float)nl, 0.318309886183790672f));
    const int mat = material_row(sc, __float_as_int(b4.w), (int)fi);
    const float4 m1 = __ldg(sc.materials + 4 * mat + 1);
    sr.contrib = mul(mul3(beta, mul(v3(m1.x, m1.y, m1.z), m1.w)), w);
    sr.o = so; sr.d = sd;
    sr.tmax = fmul(dist, 0.999f);
    sr.valid = true;
}

// The self-skip of make_ray_ctx (ff_trace.cuh) rests on the spawned ray starting a clear distance off the plane of the triangle
// it leaves.  For an origin at height h over that plane the reference's test computes t = (e2 . (tvec x e1)) / det, whose numerator
// is -h |N| exactly and carries a rounding error of at most ~6 * 2^-24 * |tvec| |e1| |e2| (|tvec| <= the longer edge: the origin
// lies over the triangle); the origin itself is known to ~2^-23 of the largest coordinate.  The skip is used only when ray_eps
// clears both by a factor of ten:  ray_eps > pad/8 (pad = 2^-16 of the largest coordinate) and
// ray_eps |N| > 4e-6 |e1| |e2| max(|e1|, |e2|) for the triangle in question -- a sliver, or a triangle as large as C3's 100-unit
// floor under the default offset of 1e-4, keeps being tested like any other.
__device__ __forceinline__ bool self_skip_enabled(const DevScene& sc, const RenderDev& rp) {
#ifdef FF_SELF_SKIP_UNGUARDED   // probe only: how far below the guard the skip stays invisible (profiles/r01j_selfskip_unguarded.txt)
    return true;
#else
    return FF_SELF_SKIP && rp.ray_eps > 0.125f * sc.pad;
#endif
}
__device__ __forceinline__ bool self_skip_safe(const DevScene& sc, const RenderDev& rp, int slot) {
    if (!self_skip_enabled(sc, rp) || slot < 0 || (slot & 0x40000000)) return false;
#ifdef FF_SELF_SKIP_UNGUARDED
    return true;
#else
    const float4 a = __ldg(sc.tris + 3 * (size_t)slot), b = __ldg(sc.tris + 3 * (size_t)slot + 1), c = __ldg(sc.tris + 3 * (size_t)slot + 2);
    const V3 e1 = sub(ld3(b), ld3(a)), e2 = sub(ld3(c), ld3(a)), N = cross(e1, e2);
    const float l1 = dot(e1, e1), l2 = dot(e2, e2), nn = dot(N, N);
    return rp.ray_eps * rp.ray_eps * nn > 1.6e-11f * (l1 * l2) * fmaxf(l1, l2);
#endif
}

// One shading event of the written-down algorithm.  Returns true when the path continues (o, d, beta updated, and with
// next-event estimation `sr` describes the shadow ray of this event); otherwise L has received the path's terminal
// contribution.
template <bool NEE>
__device__ __forceinline__ bool shade_event(const DevScene& sc, const RenderDev& rp, const HitRec& h, uint32_t pixel, uint32_t sample,
                                            int bounce, V3& o, V3& d, V3& beta, V3& L, bool& specular, ShadowReq& sr) {
    sr.valid = false;
    if (h.slot < 0) {
        L = add(L, mul3(beta, v3(rp.background[0], rp.background[1], rp.background[2])));
        return false;
    }
    V3 P, nrm;
    int geom;
    hit_surface(sc, h, o, d, P, nrm, geom);
    SurfAttr sa;
    surface_attributes(sc, h, o, d, geom, false, sa);
    const int mat = sa.mat;
    const float4 m0 = __ldg(sc.materials + 4 * mat), m1 = __ldg(sc.materials + 4 * mat + 1);
    const int type = __float_as_int(m0.x);
    if (type == FF_MAT_EMITTER) {
        // with NEE the direct light of mesh emitters arrives through the shadow rays; a hit found by sampling the BSDF
        // counts only for the camera ray and after a specular event (which has no light sample)
        const bool sampled = NEE && rp.nee && !(h.slot & 0x40000000) && sc.n_lights != 0;
        if (!sampled || bounce == 0 || specular) L = add(L, mul3(beta, mul(v3(m1.x, m1.y, m1.z), m1.w)));
        return false;
    }
    if (bounce == rp.max_bounces) return false;
    V3 n = normalize(nrm);
    if (dot(n, d) > 0.0f) n = neg(n);
    // shading normal: the interpolated corner normal when the geometry carries normals, turned to the side of n; every
    // ray offset and every "which side" decision stays with the geometric normal n
    V3 ns = n;
    if (sa.smooth) { ns = sa.ns; if (dot(ns, n) < 0.0f) ns = neg(ns); }
    const U4 r = philox4x32_10(U4{pixel, sample, (uint32_t)bounce, 0u}, rp.seed_lo, rp.seed_hi);
    if (type == FF_MAT_MIRROR || type == FF_MAT_GLASS) {
        // BXDFTyp MIRROR / GLASS (utilities.h:72-73, fields utilities.h:84-88; no code in the reference).  CPU twin:
        // trace_path in oracle/firefly_oracle.cpp, same operations in the same order.  Only front faces are ever hit,
        // so glass is one refracting interface air -> ior; Schlick's Fresnel picks reflection with probability F.
        const float4 m2 = __ldg(sc.materials + 4 * mat + 2);
        const V3 dn = normalize(d);
        float cosi = -dot(dn, ns);
        if (sa.smooth && !(cosi > 0.0f)) { ns = n; cosi = -dot(dn, n); }  // seen from behind the shading hemisphere
        bool reflect = true;
        if (type == FF_MAT_GLASS) {
            const float ior = m2.w;
            const float eta = fdiv(1.0f, ior);
            float r0 = fdiv(fsub(1.0f, ior), fadd(1.0f, ior));
            r0 = fmul(r0, r0);
            const float c1 = fsub(1.0f, cosi);
            const float F = fadd(r0, fmul(fsub(1.0f, r0), fmul(fmul(fmul(c1, c1), fmul(c1, c1)), c1)));
            const float sin2t = fmul(fmul(eta, eta), fsub(1.0f, fmul(cosi, cosi)));
            if (!(sin2t > 1.0f) && !(u01(r.z) < F)) {
                reflect = false;
                const float4 m3 = __ldg(sc.materials + 4 * mat + 3);
                d = add(mul(dn, eta), mul(ns, fsub(fmul(eta, cosi), fsqrt(fsub(1.0f, sin2t)))));
                beta = mul3(beta, v3(m3.x, m3.y, m3.z));
                o = sub(P, mul(n, rp.ray_eps));
            }
        }
        if (reflect) {
            d = add(dn, mul(ns, fmul(2.0f, cosi)));
            if (sa.smooth) { const float c = dot(d, n); if (!(c > 0.0f)) d = sub(d, mul(n, fmul(2.0f, c))); }
            beta = mul3(beta, v3(m2.x, m2.y, m2.z));
            o = add(P, mul(n, rp.ray_eps));
        }
        specular = true;
        return true;
    }
    specular = false;
    beta = mul3(beta, v3(m0.y, m0.z, m0.w));
    const V3 w = cosine_sample_hemisphere(u01(r.x), u01(r.y));
    V3 t, bt;
    onb(ns, t, bt);
    d = add(add(mul(t, w.x), mul(bt, w.y)), mul(ns, w.z));
    // a direction sampled about the shading normal can point below the real surface, where it would leave a closed
    // mesh through culled back faces: mirror it back above the geometric plane
    if (sa.smooth) { const float c = dot(d, n); if (!(c > 0.0f)) d = sub(d, mul(n, fmul(2.0f, c))); }
    o = add(P, mul(n, rp.ray_eps));
    if (NEE && rp.nee) sample_light(sc, rp, pixel, sample, bounce, o, ns, n, sa.smooth, beta, sr);
    return true;
}

// ------------------------------------------------------------------ wavefront stages
// Start of a pass: the queue counters, the camera-ray DIRECTION of every path (16 B; the origin is the camera position and
// path p sits at queue index p, so bounce 0 needs neither ray_o nor beta: 64 B per path written and 2 x 32 B read back
// before this change, C2: 4.2 GB + 4.2 GB per frame), and -- next-event estimation only -- radiance[p] = 0 for the shadow
// stages to add to.
__global__ void wf_raygen_kernel(CamDev cam, RenderDev rp, float4* __restrict__ ray_d, float4* __restrict__ radiance, uint32_t* __restrict__ counts, int n_counts) {
    const uint32_t n = rp.samples * rp.npix;
    if (blockIdx.x == 0)  // paths per bounce | extend work counters | shadow rays per bounce | shadow work counters
        for (int k = threadIdx.x; k < 4 * n_counts; k += blockDim.x) counts[k] = k == 0 ? n : 0u;
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
        const uint32_t s = p / rp.npix, pixel = frame_pixel(rp, p - s * rp.npix);
        V3 o, d;
        primary_ray(cam, rp, pixel, rp.first_sample + s, o, d);
        ray_d[p] = make_float4(d.x, d.y, d.z, 0.f);
        // radiance[p] is written exactly once, when the path ends -- only the shadow stage of next-event estimation adds to it before that
        if (rp.nee) radiance[p] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
}

// shade + BSDF sampling + queue compaction: survivors are appended to the next bounce's queue (ballot + popc inside a
// warp, a shared-memory scan over the CTA's warps, one atomicAdd per CTA and iteration), so the next extend stage reads a
// dense, coalesced ray array.
#ifndef FF_SHADE_BLOCK
#define FF_SHADE_BLOCK 128   // 64 / 128 / 256 within 1 %, 512 -2 % (profiles/r01i_variants_shade.txt)
#endif
#ifndef FF_SHADE_PREFETCH
#define FF_SHADE_PREFETCH 1  // +0.5 .. 1 %
#endif
static_assert(FF_SHADE_BLOCK >= 64 && FF_SHADE_BLOCK <= 512 && (FF_SHADE_BLOCK & (FF_SHADE_BLOCK - 1)) == 0, "two warp groups of warp 0 scan the CTA's warps");
// Ray-coherence regrouping (north star, subsystem 3): the survivors a CTA appends to the next bounce's queue in one iteration
// are ordered by the OCTANT of their new direction (match_any inside each warp, one 32-lane scan over the 8 octants x 4 warps
// of the CTA), so that the extend stage, which hands consecutive queue entries to the lanes of a warp, gets rays that visit
// the children of a node in the same front-to-back order.  No extra pass and no extra traffic: only the position inside the
// CTA's block of the queue changes.  Results do not change (radiance is indexed by path id).
// Measured on BOUNCE rays (profiles/r02k_variants_regroup.txt): C2 frame 49.03 -> 49.20 ms, C3 5.72 -> 5.73 ms, batch B of
// the 1 M soup 1883 -> 1903 Mrays/s: no gain -- these frames are bound by instruction issue with lanes idling in the
// node / triangle phase imbalance, not by the order in which lanes walk a node's children -- so the stage is compiled
// out by default (round 1 had measured +1.5 % on the uniform batch with a full host-side sort).
#ifndef FF_SHADE_REGROUP
#define FF_SHADE_REGROUP 0
#endif
static_assert(!FF_SHADE_REGROUP || FF_SHADE_BLOCK == 128, "the octant scan is one warp wide: 8 octants x 4 warps");
template <bool NEE, bool FIRST>
__global__ void __launch_bounds__(FF_SHADE_BLOCK) wf_shade_kernel(DevScene sc, CamDev cam, RenderDev rp, int bounce, const uint32_t* __restrict__ count_in,
                                                        uint32_t cap, const HitLite* __restrict__ hits,
                                                        const float4* __restrict__ ray_o_in, const float4* __restrict__ ray_d_in,
                                                        const float4* __restrict__ beta_in, float4* __restrict__ ray_o_out,
                                                        float4* __restrict__ ray_d_out, float4* __restrict__ beta_out,
                                                        float4* __restrict__ radiance, uint32_t* __restrict__ count_out,
                                                        float4* __restrict__ sh_o, float4* __restrict__ sh_d, float4* __restrict__ sh_c,
                                                        uint32_t* __restrict__ shadow_count) {
    // Queue compaction is aggregated over the CTA: one atomicAdd per CTA and iteration (per queue) instead of one per warp.  With a
    // warp-level atomic the 430 k same-address atomics of a 14 M-path launch were the kernel's top stall (53 % of the warp
    // samples sat on the shuffle that broadcasts the returned base, profiles/r01i_shade_c2_before_details.txt).
    __shared__ uint32_t s_cnt[2][FF_SHADE_BLOCK / 32], s_off[2][FF_SHADE_BLOCK / 32];
#if FF_SHADE_REGROUP
    __shared__ uint32_t s_oct[8][FF_SHADE_BLOCK / 32];   // counts, then queue offsets, of (octant, warp)
#endif
    const uint32_t n = min(*count_in, cap);
    const uint32_t total = ((n + FF_SHADE_BLOCK - 1u) / FF_SHADE_BLOCK) * FF_SHADE_BLOCK;  // uniform trip count per CTA
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int NW = FF_SHADE_BLOCK / 32;
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < total; k += gridDim.x * blockDim.x) {
        bool alive = false;
        V3 o, d, beta;
        uint32_t path = 0;
        ShadowReq sr;
        sr.valid = false;
        bool specular = false;  // the path's previous event was a mirror / glass bounce (carried in beta.w)
        uint32_t from = 0;      // slot + 1 of the triangle the continued ray starts on (0: a plane, a sphere, or no offset)
#if FF_SHADE_PREFETCH
        {   // pull the next iteration's four streams towards L2 while this one is shaded (one lane per 128-byte line)
            const uint32_t k2 = k + gridDim.x * blockDim.x;
            if ((lane & 7) == 0 && k2 < n) {
#if FF_SHADE_PREFETCH == 2
#define FF_PF "prefetch.global.L1 [%0];"
#else
#define FF_PF "prefetch.global.L2 [%0];"
#endif
                asm volatile(FF_PF ::"l"(ray_d_in + k2));
                if (!FIRST) {
                    asm volatile(FF_PF ::"l"(ray_o_in + k2));
                    asm volatile(FF_PF ::"l"(beta_in + k2));
                }
                if ((lane & 15) == 0) asm volatile(FF_PF ::"l"(hits + k2));
            }
        }
#endif
        if (k < n) {
            const HitLite hl = hits[k];
            HitRec h;
            h.key = 0.f; h.t = hl.t; h.slot = hl.slot; h.ord = 0;   // shading reads t and slot only
            uint32_t s, pixel;
            if (FIRST) {   // camera segment: path k is (sample, pixel) = (k / npix, k % npix), starts at the camera, beta = 1
                path = k;
                s = path / rp.npix; pixel = frame_pixel(rp, path - s * rp.npix);
                o = v3(cam.pos[0], cam.pos[1], cam.pos[2]);
                d = ld3(ray_d_in[k]);
                beta = v3(1.f, 1.f, 1.f);
            } else {
                const float4 o4 = ray_o_in[k], d4 = ray_d_in[k], b4 = beta_in[k];
                path = __float_as_uint(o4.w);
                s = path / rp.npix; pixel = frame_pixel(rp, path - s * rp.npix);
                o = ld3(o4); d = ld3(d4); beta = ld3(b4);
                specular = b4.w != 0.0f;
            }
            V3 L = v3(0.f, 0.f, 0.f);
            alive = shade_event<NEE>(sc, rp, h, pixel, rp.first_sample + s, bounce, o, d, beta, L, specular, sr);
            if (alive && self_skip_safe(sc, rp, h.slot)) from = (uint32_t)h.slot + 1u;
            if (!alive) {
                if (NEE) {  // the shadow stages of earlier bounces have already added to this path
                    const float4 a = radiance[path];
                    L = add(ld3(a), L);
                }
                radiance[path] = make_float4(L.x, L.y, L.z, 0.f);
            }
        }
        const unsigned m = __ballot_sync(0xffffffffu, alive);
        const unsigned ms = NEE ? __ballot_sync(0xffffffffu, sr.valid) : 0u;
#if FF_SHADE_REGROUP
        const uint32_t oct = alive ? ((d.x < 0.f ? 4u : 0u) | (d.y < 0.f ? 2u : 0u) | (d.z < 0.f ? 1u : 0u)) : 8u;
        const unsigned peers = __match_any_sync(0xffffffffu, oct);   // the lanes of this warp whose ray continues into the same octant
        if (lane < 8) s_oct[lane][warp] = 0u;
        __syncwarp();
        if (alive && lane == __ffs(peers) - 1) s_oct[oct][warp] = (uint32_t)__popc(peers);
#endif
        if (lane == 0) {
            s_cnt[0][warp] = (uint32_t)__popc(m);
            if (NEE) s_cnt[1][warp] = (uint32_t)__popc(ms);
        }
        __syncthreads();
        if (warp == 0) {  // lanes [0, NW): the next bounce's queue; lanes [NW, 2 NW): the shadow queue
            const int g = lane / NW, j = lane % NW;
#if FF_SHADE_REGROUP
            const bool scan_g = NEE && g == 1;   // the next bounce's queue is laid out by the octant scan below
#else
            const bool scan_g = g == 0 || (NEE && g == 1);
#endif
            const uint32_t c = scan_g ? s_cnt[g][j] : 0u;
            uint32_t incl = c;
#pragma unroll
            for (int off = 1; off < NW; off <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, incl, off, NW);
                if (j >= off) incl += t;
            }
            const uint32_t sum = __shfl_sync(0xffffffffu, incl, NW - 1, NW);
            uint32_t base = 0;
            if (j == 0 && sum != 0u && scan_g) base = atomicAdd(g == 0 ? count_out : shadow_count, sum);
            base = __shfl_sync(0xffffffffu, base, 0, NW);
            if (scan_g) s_off[g][j] = base + incl - c;
#if FF_SHADE_REGROUP
            {   // lane = octant * NW + warp: the CTA's block of the queue is octant-major
                const uint32_t co = s_oct[lane / NW][lane % NW];
                uint32_t io = co;
#pragma unroll
                for (int off = 1; off < 32; off <<= 1) {
                    const uint32_t t = __shfl_up_sync(0xffffffffu, io, off);
                    if (lane >= off) io += t;
                }
                const uint32_t total = __shfl_sync(0xffffffffu, io, 31);
                uint32_t bo = 0;
                if (lane == 0 && total != 0u) bo = atomicAdd(count_out, total);
                bo = __shfl_sync(0xffffffffu, bo, 0);
                s_oct[lane / NW][lane % NW] = bo + io - co;
            }
#endif
        }
        __syncthreads();
        if (alive) {
#if FF_SHADE_REGROUP
            const uint32_t q = s_oct[oct][warp] + __popc(peers & ((1u << lane) - 1u));
#else
            const uint32_t q = s_off[0][warp] + __popc(m & ((1u << lane) - 1u));
#endif
            ray_o_out[q] = make_float4(o.x, o.y, o.z, __uint_as_float(path));
            ray_d_out[q] = make_float4(d.x, d.y, d.z, __uint_as_float(from));
            beta_out[q] = make_float4(beta.x, beta.y, beta.z, specular ? 1.0f : 0.0f);
        }
        if (NEE && sr.valid) {  // shadow-ray queue of this bounce, compacted the same way
            const uint32_t q = s_off[1][warp] + __popc(ms & ((1u << lane) - 1u));
            sh_o[q] = make_float4(sr.o.x, sr.o.y, sr.o.z, __uint_as_float(path));
            sh_d[q] = make_float4(sr.d.x, sr.d.y, sr.d.z, sr.tmax);
            sh_c[q] = make_float4(sr.contrib.x, sr.contrib.y, sr.contrib.z, __uint_as_float(from));
        }
    }
}


void k_wavefront_pass(const DevScene& sc, const CamDev& cam, const RenderDev& rp, const WaveBuffers& wb, int trace_mode,
                      TraceCounters* d_counters, uint32_t* launches, cudaEvent_t* trace_events, int* n_trace_events,
                      cudaStream_t st) {
    const uint32_t n = rp.samples * rp.npix;
    const int n_counts = rp.max_bounces + 2;
    wf_raygen_kernel<<<grid_for(n, 256, sm_count() * 16), 256, 0, st>>>(cam, rp, wb.ray_d[0], wb.radiance, wb.counts, n_counts);
    ++*launches;
    const WaveGen gen = {cam.pos[0], cam.pos[1], cam.pos[2]};
    const bool nee = rp.nee && sc.n_lights != 0;
    const int shade_grid = grid_for(n, FF_SHADE_BLOCK, sm_count() * 16 * 256 / FF_SHADE_BLOCK);
    for (int b = 0; b <= rp.max_bounces; ++b) {
        const int in = b & 1, out = in ^ 1;
        if (trace_events) cudaEventRecord(trace_events[(*n_trace_events)++], st);
        k_trace_extend(sc, b == 0 ? &gen : nullptr, wb.ray_o[in], wb.ray_d[in], wb.counts + b, n,
                       trace_mode | ((FF_SELF_SKIP && rp.ray_eps > 0.125f * sc.pad) ? FF_TRACE_SELFSKIP : 0), wb.hits, d_counters, wb.counts + n_counts + b, st);
        if (trace_events) cudaEventRecord(trace_events[(*n_trace_events)++], st);
#define FF_SHADE_ARGS sc, cam, rp, b, wb.counts + b, n, wb.hits, wb.ray_o[in], wb.ray_d[in], wb.beta[in], wb.ray_o[out], wb.ray_d[out], wb.beta[out], \
                      wb.radiance, wb.counts + b + 1, wb.sh_o, wb.sh_d, wb.sh_c, wb.counts + 2 * n_counts + b
        if (nee) {
            if (b == 0) wf_shade_kernel<true, true><<<shade_grid, FF_SHADE_BLOCK, 0, st>>>(FF_SHADE_ARGS);
            else wf_shade_kernel<true, false><<<shade_grid, FF_SHADE_BLOCK, 0, st>>>(FF_SHADE_ARGS);
        } else {
            if (b == 0) wf_shade_kernel<false, true><<<shade_grid, FF_SHADE_BLOCK, 0, st>>>(FF_SHADE_ARGS);
            else wf_shade_kernel<false, false><<<shade_grid, FF_SHADE_BLOCK, 0, st>>>(FF_SHADE_ARGS);
        }
#undef FF_SHADE_ARGS
        *launches += 2;
        if (nee && b < rp.max_bounces) {  // shadow stage: any-hit over this bounce's light samples
            if (trace_events) cudaEventRecord(trace_events[(*n_trace_events)++], st);
            k_trace_shadow(sc, wb.sh_o, wb.sh_d, wb.sh_c, wb.counts + 2 * n_counts + b, n, wb.radiance, wb.counts + 3 * n_counts + b,
                           FF_SELF_SKIP && rp.ray_eps > 0.125f * sc.pad, st);
            if (trace_events) cudaEventRecord(trace_events[(*n_trace_events)++], st);
            ++*launches;
        }
    }
}

// ------------------------------------------------------------------ persistent megakernel
// One thread owns one path from camera to termination (the shape of the reference's launchPathTrace, kernel.cu:186-221,
// with the bounce loop it intends).  Warps are persistent: a lane whose path has terminated pulls the next path id from
// a global counter as soon as a few lanes are idle, so a warp never waits for its longest path.
// NEE is a template parameter so that the plain loop does not carry the shadow phase's registers.
template <bool NEE>
__global__ void __launch_bounds__(FF_TRACE_BLOCK) megakernel(DevScene sc, CamDev cam, RenderDev rp, float4* __restrict__ radiance,
                                                              uint32_t* work_counter, unsigned long long* ray_counter) {
    __shared__ uint2 sm_stack[(FF_SM_STACK + FF_TRI_QUEUE) * FF_TRACE_BLOCK];
    const uint32_t n = rp.samples * rp.npix;
    const int lane = threadIdx.x & 31;
    uint2* my = sm_stack + threadIdx.x;
    bool has = false, exhausted = false;
    bool specular = false;  // the previous event was a mirror / glass bounce
    bool shadow = false;  // the traversal in flight is this bounce's shadow ray; (o, d) already hold the next segment
    uint32_t path = 0, pixel = 0, sample = 0;
    int bounce = 0;
    V3 o = v3(0, 0, 0), d = v3(0, 0, 1), beta = v3(1, 1, 1), L = v3(0, 0, 0), pending = v3(0, 0, 0);
    RayCtx r;
    Trav tv;
    uint2 lstack[FF_LOCAL_STACK];
    Best best;
    unsigned rays = 0, shadow_rays = 0, cn = 0, ct = 0;
    // (re)starts the closest-hit traversal of the segment (o, d)
    // (no self-skip here: carrying the origin triangle's slot costs the megakernel 16 / 32 registers and, with next-event
    // estimation, 17 % of its speed -- profiles/r01j_configs.jsonl)
    auto start_segment = [&]() {
        best_init(best, INFINITY);
        test_planes<false>(sc, o, d, best);
        r = make_ray_ctx(o, normalize(d));
        tv.start(best.key);
        if (sc.n_nodes == 0) tv.ngroup.y = 0u;
        shadow = false;
    };
    for (;;) {
        const unsigned need = __ballot_sync(0xffffffffu, !has);
        if (need && !exhausted) {
            uint32_t base = 0;
            const int leader = __ffs(need) - 1;
            if (lane == leader) base = atomicAdd(work_counter, (uint32_t)__popc(need));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (base + (uint32_t)__popc(need) >= n) exhausted = true;
            if (!has) {
                const uint32_t p = base + __popc(need & ((1u << lane) - 1u));
                if (p < n) {
                    path = p;
                    const uint32_t s = p / rp.npix;
                    pixel = frame_pixel(rp, p - s * rp.npix);
                    sample = rp.first_sample + s;
                    beta = v3(1.f, 1.f, 1.f); L = v3(0.f, 0.f, 0.f);
                    bounce = 0;
                    specular = false;
                    primary_ray(cam, rp, pixel, sample, o, d);
                    start_segment();
                    has = true;
                }
            }
        }
        if (!__any_sync(0xffffffffu, has)) break;
        if (has) {
            for (;;) {
                // A shadow ray runs through the same closest-hit step (no second copy of the traversal in the loop, no
                // divergence between lanes in different phases): its search interval starts at tmax and the first accepted
                // hit ends it, which is the any-hit answer "some front-facing hit with key < tmax".
                const bool more = bvh8_step<false, false>(sc, r, tv, best, my, FF_TRACE_BLOCK, lstack, cn, ct);
                if (NEE && shadow) {
                    if (!more || best.slot >= 0) {
                        if (best.slot < 0) L = add(L, pending);
                        start_segment();
                    }
                } else if (!more) {
                    ++rays;
                    HitRec h;
                    h.key = best.key; h.t = best.t; h.slot = best.slot; h.ord = best.slot < 0 ? -1 : best.ord;
                    ShadowReq sr;
                    if (shade_event<NEE>(sc, rp, h, pixel, sample, bounce, o, d, beta, L, specular, sr)) {
                        ++bounce;
                        if (NEE && sr.valid) {
                            ++shadow_rays;
                            pending = sr.contrib;
                            best_init(best, sr.tmax);
                            best.ord = -1;  // strict '<' against tmax: nothing ties with the limit
                            test_planes<false>(sc, sr.o, sr.d, best);
                            if (best.slot >= 0 || sc.n_nodes == 0) {  // a plane already blocks it / nothing else to test
                                if (best.slot < 0) L = add(L, pending);
                                start_segment();
                            } else {
                                r = make_ray_ctx(sr.o, normalize(sr.d));
                                tv.start(best.key);
                                shadow = true;
                            }
                        } else {
                            start_segment();
                        }
                    } else {
                        radiance[path] = make_float4(L.x, L.y, L.z, 0.f);
                        has = false;
                        break;
                    }
                }
                if (!exhausted && __popc(__activemask()) < sc.refill_lanes) break;
            }
        }
        __syncwarp();
    }
    for (int off = 16; off; off >>= 1) {
        rays += __shfl_xor_sync(0xffffffffu, rays, off);
        shadow_rays += __shfl_xor_sync(0xffffffffu, shadow_rays, off);
    }
    if (lane == 0) {
        atomicAdd(ray_counter, (unsigned long long)rays);
        if (shadow_rays) atomicAdd(ray_counter + 1, (unsigned long long)shadow_rays);
    }
}
void k_megakernel_pass(const DevScene& sc, const CamDev& cam, const RenderDev& rp, float4* d_radiance, uint32_t* d_work_counter,
                       unsigned long long* d_ray_counter, uint32_t* launches, cudaStream_t st) {
    cudaMemsetAsync(d_work_counter, 0, sizeof(uint32_t), st);
    const uint32_t n = rp.samples * rp.npix;
    const int g = grid_for(n, FF_TRACE_BLOCK, sm_count() * tuning().blocks_per_sm);
    if (rp.nee && sc.n_lights != 0) megakernel<true><<<g, FF_TRACE_BLOCK, 0, st>>>(sc, cam, rp, d_radiance, d_work_counter, d_ray_counter);
    else megakernel<false><<<g, FF_TRACE_BLOCK, 0, st>>>(sc, cam, rp, d_radiance, d_work_counter, d_ray_counter);
    ++*launches;
}

// ------------------------------------------------------------------ accumulate / resolve
// accum[pixel] = (((accum + L[0]) + L[1]) + ...): samples in index order, like the oracle's per-pixel loop.
__global__ void accumulate_kernel(const float4* __restrict__ radiance, uint32_t npix, uint32_t samples,
                                  const uint32_t* __restrict__ pixel_map, float4* __restrict__ accum) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += gridDim.x * blockDim.x) {
        const uint32_t px = pixel_map ? pixel_map[i] : i;
        float4 a = accum[px];
        for (uint32_t s = 0; s < samples; ++s) {
            const float4 l = radiance[(size_t)s * npix + i];
            a.x = fadd(a.x, l.x); a.y = fadd(a.y, l.y); a.z = fadd(a.z, l.z);
        }
        accum[px] = a;
    }
}
void k_accumulate_pass(const float4* d_radiance, uint32_t npix, uint32_t samples, const uint32_t* d_pixel_map, float4* d_accum, cudaStream_t st) {
    accumulate_kernel<<<grid_for(npix, 256, sm_count() * 16), 256, 0, st>>>(d_radiance, npix, samples, d_pixel_map, d_accum);
}
// out = (sum over accumulators in rank order) * (1 / total_spp); alpha = 1.  With n_accums > 1 the pointers may be
// peer-GPU memory (CUDA IPC over NVLink): the framebuffer reduce and the resolve are one kernel.
struct PeerPtrs { const float4* p[8]; };
__global__ void resolve_kernel(PeerPtrs peers, int n_accums, uint32_t npix, float inv_spp, float4* __restrict__ out) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += gridDim.x * blockDim.x) {
        float4 a = peers.p[0][i];
        for (int r = 1; r < n_accums; ++r) {
            const float4 b = peers.p[r][i];
            a.x = fadd(a.x, b.x); a.y = fadd(a.y, b.y); a.z = fadd(a.z, b.z);
        }
        out[i] = make_float4(fmul(a.x, inv_spp), fmul(a.y, inv_spp), fmul(a.z, inv_spp), 1.0f);
    }
}
void k_resolve(const float4* const* accums, int n_accums, uint32_t npix, float inv_spp, float4* d_out, cudaStream_t st) {
    PeerPtrs pp;
    for (int i = 0; i < 8; ++i) pp.p[i] = i < n_accums ? accums[i] : nullptr;
    resolve_kernel<<<grid_for(npix, 256, sm_count() * 16), 256, 0, st>>>(pp, n_accums, npix, inv_spp, d_out);
}

// Fused framebuffer reduce-scatter + resolve + gather over peer memory (FF_REDUCE_PEER, ff_group.cpp): the GPU that runs this
// sums pixels [first, first + count) of EVERY rank's accumulator -- its own from HBM, the peers' over NVLink -- in rank order,
// scales by 1/spp and stores the finished pixels into the root's frame (a peer store unless this GPU is the root).  All GPUs
// run it at once on their own slice, so each NVLink ingress carries (world-1)/world of one accumulator.  Peer loads bypass L1
// (ld.global.cg): the lines belong to another GPU's memory and were rewritten since the last frame.  Four pixels per thread
// and iteration keep 4 x world independent 16-byte loads in flight per thread, which is what hides the NVLink latency.
#define FF_RS_UNROLL 4
__global__ void __launch_bounds__(256) reduce_scatter_resolve_kernel(PeerPtrs peers, int world, uint32_t first, uint32_t count, float inv_spp,
                                                                      float4* __restrict__ out) {
    const uint32_t stride = gridDim.x * blockDim.x * FF_RS_UNROLL;
    for (uint32_t base = blockIdx.x * blockDim.x * FF_RS_UNROLL + threadIdx.x; base < count; base += stride) {
        float4 a[FF_RS_UNROLL];
#pragma unroll
        for (int u = 0; u < FF_RS_UNROLL; ++u) {
            const uint32_t i = base + u * blockDim.x;
            a[u] = i < count ? __ldcg(peers.p[0] + first + i) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        for (int r = 1; r < world; ++r) {
            float4 b[FF_RS_UNROLL];
#pragma unroll
            for (int u = 0; u < FF_RS_UNROLL; ++u) {
                const uint32_t i = base + u * blockDim.x;
                b[u] = i < count ? __ldcg(peers.p[r] + first + i) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int u = 0; u < FF_RS_UNROLL; ++u) { a[u].x = fadd(a[u].x, b[u].x); a[u].y = fadd(a[u].y, b[u].y); a[u].z = fadd(a[u].z, b[u].z); }
        }
#pragma unroll
        for (int u = 0; u < FF_RS_UNROLL; ++u) {
            const uint32_t i = base + u * blockDim.x;
            if (i < count) out[first + i] = make_float4(fmul(a[u].x, inv_spp), fmul(a[u].y, inv_spp), fmul(a[u].z, inv_spp), 1.0f);
        }
    }
}
void k_reduce_scatter_resolve(const float4* const* accums, int world, uint32_t first, uint32_t count, float inv_spp, float4* d_out_root, cudaStream_t st) {
    if (!count) return;
    PeerPtrs pp;
    for (int i = 0; i < 8; ++i) pp.p[i] = i < world ? accums[i] : nullptr;
    reduce_scatter_resolve_kernel<<<grid_for((count + FF_RS_UNROLL - 1) / FF_RS_UNROLL, 256, sm_count() * 8), 256, 0, st>>>(pp, world, first, count, inv_spp, d_out_root);
}

// ------------------------------------------------------------------ primary-ray feature buffers
// One ray per pixel through the integer pixel corner (kernel.cu:200-201), first hit only.  kind: 3 shading normal
// (interpolated when the geometry has normals, else geometric; face-forwarded, world space), 4 uv, 5 barycentrics
// (1-u-v, u, v), 6 albedo of the hit's material row.  Alpha = 1 on a hit, 0 on a miss (rgb 0).
__global__ void __launch_bounds__(FF_TRACE_BLOCK) aov_kernel(DevScene sc, CamDev cam, int kind, float4* __restrict__ out) {
    __shared__ uint2 sm_stack[(FF_SM_STACK + FF_TRI_QUEUE) * FF_TRACE_BLOCK];
    const uint32_t npix = (uint32_t)cam.w * cam.h;
    unsigned cn = 0, ct = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += gridDim.x * blockDim.x) {
        V3 o, d;
        camera_ray_dev(cam, (float)(i % cam.w), (float)(i / cam.w), o, d);
        const V3 dn = normalize(d);
        Best best;
        best_init(best, INFINITY);
        test_planes<false>(sc, o, d, best);
        trace_bvh8<false, false>(sc, o, dn, best, sm_stack + threadIdx.x, FF_TRACE_BLOCK, cn, ct);
        float4 px = make_float4(0.f, 0.f, 0.f, 0.f);
        if (best.slot >= 0) {
            HitRec h;
            h.key = best.key; h.t = best.t; h.slot = best.slot; h.ord = best.ord;
            V3 P, nrm;
            int geom;
            hit_surface(sc, h, o, d, P, nrm, geom);
            SurfAttr sa;
            surface_attributes(sc, h, o, d, geom, true, sa);
            if (kind == FF_INTEGRATOR_AOV_NORMAL) {
                V3 n = normalize(nrm);
                if (dot(n, d) > 0.0f) n = neg(n);
                V3 ns = n;
                if (sa.smooth) { ns = sa.ns; if (dot(ns, n) < 0.0f) ns = neg(ns); }
                px = make_float4(ns.x, ns.y, ns.z, 1.f);
            } else if (kind == FF_INTEGRATOR_AOV_UV) {
                px = make_float4(sa.tu, sa.tv, 0.f, 1.f);
            } else if (kind == FF_INTEGRATOR_AOV_BARY) {
                px = (h.slot & 0x40000000) ? make_float4(0.f, 0.f, 0.f, 1.f) : make_float4(fsub(fsub(1.0f, sa.u), sa.v), sa.u, sa.v, 1.f);
            } else {
                const float4 m0 = __ldg(sc.materials + 4 * sa.mat);
                px = make_float4(m0.y, m0.z, m0.w, 1.f);
            }
        }
        out[i] = px;
    }
}
void k_aov(const DevScene& sc, const CamDev& cam, int kind, float4* d_rgba, cudaStream_t st) {
    const size_t n = (size_t)cam.w * cam.h;
    aov_kernel<<<grid_for(n, FF_TRACE_BLOCK, sm_count() * 16), FF_TRACE_BLOCK, 0, st>>>(sc, cam, kind, d_rgba);
}
__global__ void hit_barycentrics_kernel(DevScene sc, const float4* __restrict__ ray_o, const float4* __restrict__ ray_d,
                                        const HitRec* __restrict__ hits, size_t n, float* __restrict__ out_u, float* __restrict__ out_v) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const HitRec h = hits[i];
        float u = 0.f, v = 0.f;
        if (h.slot >= 0 && !(h.slot & 0x40000000)) {
            const float4 a4 = __ldg(sc.tris + 3 * (size_t)h.slot), b4 = __ldg(sc.tris + 3 * (size_t)h.slot + 1),
                         c4 = __ldg(sc.tris + 3 * (size_t)h.slot + 2);
            triangle_barycentrics(ld3(a4), ld3(b4), ld3(c4), ld3(ray_o[i]), normalize(ld3(ray_d[i])), u, v);
        }
        out_u[i] = u; out_v[i] = v;
    }
}
void k_hit_barycentrics(const DevScene& sc, const float4* d_ray_o, const float4* d_ray_d, const HitRec* d_hits, size_t n,
                        float* d_u, float* d_v, cudaStream_t st) {
    hit_barycentrics_kernel<<<grid_for(n, 256, sm_count() * 16), 256, 0, st>>>(sc, d_ray_o, d_ray_d, d_hits, n, d_u, d_v);
}

// ------------------------------------------------------------------ as-coded frame (kernel.cu:186-216)
// colour = abs(normal) -> uchar3 (kernel.cu:183,214); misses keep the cleared pixel (kernel.cu:340).  `normal` is what
// setIntersection leaves in the Intersect (kernel.cu:117): inverse(transpose(M)) * object-space unit normal, NOT
// re-normalised -- `nref` holds it per flat triangle (k_as_coded_normals).  make_uchar3(float...) compiles to
// cvt.rzi.u32.f32 followed by an 8-bit store, so a component above 255 wraps modulo 256 (|normal| > 1 under a scale < 1);
// F2I saturates at 2^32 - 1 and turns NaN into 0.  CPU twin: u8_as_cuda in oracle/firefly_oracle.cpp.
__device__ __forceinline__ unsigned char as_coded_u8(float c) { return (unsigned char)(__float2uint_rz(c) & 0xffu); }
__global__ void __launch_bounds__(FF_TRACE_BLOCK) as_coded_kernel(DevScene sc, CamDev cam, const float4* __restrict__ nref, unsigned char* __restrict__ rgb) {
    __shared__ uint2 sm_stack[(FF_SM_STACK + FF_TRI_QUEUE) * FF_TRACE_BLOCK];
    const uint32_t npix = (uint32_t)cam.w * cam.h;
    unsigned cn = 0, ct = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += gridDim.x * blockDim.x) {
        V3 o, d;
        camera_ray_dev(cam, (float)(i % cam.w), (float)(i / cam.w), o, d);
        const V3 dn = normalize(d);
        Best best;
        best_init(best, INFINITY);
        test_planes<false>(sc, o, d, best);
        trace_bvh8<false, false>(sc, o, dn, best, sm_stack + threadIdx.x, FF_TRACE_BLOCK, cn, ct);
        unsigned char px[3] = {0, 0, 0};
        if (best.slot >= 0) {
            V3 n;
            if (best.slot & 0x40000000) {
                HitRec h;
                h.key = best.key; h.t = best.t; h.slot = best.slot; h.ord = best.ord;
                V3 P;
                int geom;
                hit_surface(sc, h, o, d, P, n, geom);
            } else {
                const int flat = __float_as_int(__ldg(sc.tris + 3 * (size_t)best.slot + 2).w);
                n = ld3(__ldg(nref + flat));
            }
            px[0] = as_coded_u8(fmul(fabsf(n.x), 255.f)); px[1] = as_coded_u8(fmul(fabsf(n.y), 255.f)); px[2] = as_coded_u8(fmul(fabsf(n.z), 255.f));
        }
        rgb[3 * (size_t)i] = px[0]; rgb[3 * (size_t)i + 1] = px[1]; rgb[3 * (size_t)i + 2] = px[2];
    }
}
void k_as_coded(const DevScene& sc, const CamDev& cam, const float4* d_as_coded_normals, unsigned char* d_rgb, cudaStream_t st) {
    const size_t n = (size_t)cam.w * cam.h;
    as_coded_kernel<<<grid_for(n, FF_TRACE_BLOCK, sm_count() * 16), FF_TRACE_BLOCK, 0, st>>>(sc, cam, d_as_coded_normals, d_rgb);
}
// linear radiance -> 8-bit: clamp to [0,1], x255, truncate (the kernel.cu:214 conversion)
__global__ void rgba_to_u8_kernel(const float4* __restrict__ rgba, uint32_t npix, unsigned char* __restrict__ rgb) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += gridDim.x * blockDim.x) {
        const float4 v = rgba[i];
        rgb[3 * (size_t)i] = (unsigned char)(fminf(fmaxf(v.x, 0.f), 1.f) * 255.f);
        rgb[3 * (size_t)i + 1] = (unsigned char)(fminf(fmaxf(v.y, 0.f), 1.f) * 255.f);
        rgb[3 * (size_t)i + 2] = (unsigned char)(fminf(fmaxf(v.z, 0.f), 1.f) * 255.f);
    }
}
__global__ void u8_to_rgba_kernel(const unsigned char* __restrict__ rgb, uint32_t npix, float4* __restrict__ rgba) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += gridDim.x * blockDim.x)
        rgba[i] = make_float4(fdiv((float)rgb[3 * (size_t)i], 255.0f), fdiv((float)rgb[3 * (size_t)i + 1], 255.0f), fdiv((float)rgb[3 * (size_t)i + 2], 255.0f), 1.0f);
}
void k_u8_to_rgba(const unsigned char* d_rgb, uint32_t npix, float4* d_rgba, cudaStream_t st) {
    u8_to_rgba_kernel<<<grid_for(npix, 256, sm_count() * 16), 256, 0, st>>>(d_rgb, npix, d_rgba);
}
void k_rgba_to_u8(const float4* d_rgba, uint32_t npix, unsigned char* d_rgb, cudaStream_t st) {
    rgba_to_u8_kernel<<<grid_for(npix, 256, sm_count() * 16), 256, 0, st>>>(d_rgba, npix, d_rgb);
}

}  // namespace ff
