// ff_trace.cuh -- device-side closest-hit / any-hit traversal shared by the trace kernels, the wavefront
// extend stage and the persistent megakernel.
//
// Replaces intersectRays (reference kernel.cu:127-176): the O(rays x triangles) scan becomes a
// traversal of an 8-wide compressed BVH; the leaf test IS the reference's test (ffm::intersect_triangle,
// kernel.cu:35-108) and the accept rule reproduces its closest-hit key and its scan-order tie-break:
//     key = distance(o + d*t, o)                       (setIntersection, kernel.cu:113-115)
//     accept  <=>  key < best || (key == best && ord < best_ord)   (strict '<' in scan order, kernel.cu:115,133,145)
// Box culling is conservative: leaf boxes are inflated by DevScene::pad at build time and a node is
// skipped only when its entry distance exceeds best*(1+2^-15).  DESIGN.md section 4.
#pragma once
#include "ff_internal.h"

namespace ff {

using ffm::V3;


struct Best {
    float key, t;
    int slot, ord;
};
__device__ __forceinline__ void best_init(Best& b, float tmax) {
    b.key = tmax; b.t = 0.f; b.slot = -1; b.ord = 0x7fffffff;
}
__device__ __forceinline__ bool best_accepts(const Best& b, float key, int ord) {
    return key < b.key || (key == b.key && ord < b.ord);
}

__device__ __forceinline__ V3 ld3(const float4& v) { return ffm::v3(v.x, v.y, v.z); }

// One triangle slot against the ray: the reference's test + key + tie-break.
template <bool ANY>
__device__ __forceinline__ bool test_slot(const DevScene& sc, int slot, V3 o, V3 dn, Best& best) {
#if FF_TRI_TEX
    const float4 a = tex1Dfetch<float4>(sc.tris_tex, 3 * slot);
    const float4 b = tex1Dfetch<float4>(sc.tris_tex, 3 * slot + 1);
    const float4 c = tex1Dfetch<float4>(sc.tris_tex, 3 * slot + 2);
#else
    const float4* __restrict__ tris = sc.tris;
    const float4 a = __ldg(tris + 3 * (size_t)slot);
    const float4 b = __ldg(tris + 3 * (size_t)slot + 1);
    const float4 c = __ldg(tris + 3 * (size_t)slot + 2);
#endif
    float t;
    if (!ffm::intersect_triangle(ld3(a), ld3(b), ld3(c), o, dn, t)) return false;
    V3 p;
    float key = ffm::hit_key(o, dn, t, p);
    int ord = __float_as_int(a.w);
    if (ANY) {
        if (key < best.key) { best.slot = slot; best.ord = ord; best.t = t; return true; }
        return false;
    }
    if (best_accepts(best, key, ord)) {
        best.key = key; best.t = t; best.slot = slot; best.ord = ord;
        return true;
    }
    return false;
}

// glm mat4*vec4 (type_mat4x4.inl:536-575)
__device__ __forceinline__ ffm::V4 mul_mv(const ffm::M4& m, float x, float y, float z, float w) {
    using namespace ffm;
    V4 r;
    r.x = fadd(fadd(fmul(m.c[0].x, x), fmul(m.c[1].x, y)), fadd(fmul(m.c[2].x, z), fmul(m.c[3].x, w)));
    r.y = fadd(fadd(fmul(m.c[0].y, x), fmul(m.c[1].y, y)), fadd(fmul(m.c[2].y, z), fmul(m.c[3].y, w)));
    r.z = fadd(fadd(fmul(m.c[0].z, x), fmul(m.c[1].z, y)), fadd(fmul(m.c[2].z, z), fmul(m.c[3].z, w)));
    r.w = fadd(fadd(fmul(m.c[0].w, x), fmul(m.c[1].w, y)), fadd(fmul(m.c[2].w, z), fmul(m.c[3].w, w)));
    return r;
}

// intersectPlane (kernel.cu:8-32) / the sphere routine (CPU twin: intersectSphere in oracle/firefly_oracle.cpp) +
// setIntersection (kernel.cu:110-125) for one analytic primitive.  `d_raw` is the ray direction as given (the reference
// normalises it in object space, kernel.cu:138).  `local_point` is the object-space hit point (the sphere's normal).
__device__ __forceinline__ bool plane_hit(const DevPlane& pl, V3 o, V3 d_raw, float& key, float& t_local, V3& world_point, V3& local_point) {
    using namespace ffm;
    V4 o4 = mul_mv(pl.inv, o.x, o.y, o.z, 1.0f);
    V4 d4 = mul_mv(pl.inv, d_raw.x, d_raw.y, d_raw.z, 0.0f);
    float dd = fadd(fadd(fmul(d4.x, d4.x), fmul(d4.y, d4.y)), fadd(fmul(d4.z, d4.z), fmul(d4.w, d4.w)));
    float inv = inversesqrt(dd);
    V3 od = v3(fmul(d4.x, inv), fmul(d4.y, inv), fmul(d4.z, inv));
    V3 oo = v3(o4.x, o4.y, o4.z);
    float t;
    V3 P;
    if (pl.kind == 1) {
        V3 n = v3(0.f, 0.f, 1.f);
        float denom = dot(n, od);
        if (!((double)fabsf(denom) > 1e-7)) return false;
        t = fdiv(dot(neg(oo), n), denom);
        P = add(oo, mul(od, t));
        if (!(P.x >= -0.5f && P.x <= 0.5f && P.y >= -0.5f && P.y <= 0.5f)) return false;
        if (!(t > 0.0f)) return false;
    } else {
        float b = dot(oo, od);
        float c = fsub(dot(oo, oo), fmul(pl.radius, pl.radius));
        float disc = fsub(fmul(b, b), c);
        if (!(disc > 0.0f)) return false;
        t = fsub(-b, fsqrt(disc));
        if (!(t > 0.0f)) return false;
        P = add(oo, mul(od, t));
    }
    V4 w = mul_mv(pl.model, P.x, P.y, P.z, 1.0f);
    world_point = v3(w.x, w.y, w.z);
    local_point = P;
    key = distance(world_point, o);
    t_local = t;
    return true;
}

template <bool ANY>
__device__ __forceinline__ void test_planes(const DevScene& sc, V3 o, V3 d_raw, Best& best) {
    for (uint32_t i = 0; i < sc.n_planes; ++i) {
        float key, tl;
        V3 wp, lp;
        if (plane_hit(sc.planes[i], o, d_raw, key, tl, wp, lp)) {
            int ord = sc.planes[i].ord;
            if (ANY ? (key < best.key) : best_accepts(best, key, ord)) {
                if (!ANY) best.key = key;
                best.t = tl; best.slot = 0x40000000 | (int)i; best.ord = ord;
                if (ANY) return;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Camera rays (kernel.cu:198-205), shared by the ray-batch kernels, the integrators and the wavefront's extend stage.
__device__ __forceinline__ void camera_ray_dev(const CamDev& cam, float px, float py, V3& o, V3& d) {
    using namespace ffm;
    o = v3(cam.pos[0], cam.pos[1], cam.pos[2]);
    const float Px = fsub(fmul(fdiv(px, cam.width), 2.f), 1.f);
    const float Py = fsub(1.f, fmul(fdiv(py, cam.height), 2.f));
    const V4 look = mul_mv(cam.inv_vp, fmul(Px, cam.far_clip), fmul(Py, cam.far_clip), fmul(1.f, cam.far_clip),
                           fmul(1.f, cam.far_clip));
    d = normalize(sub(v3(look.x, look.y, look.z), o));
}
// primary ray of path (pixel, sample): integer pixel corner (kernel.cu:200-201) or Philox-jittered
// Pixel of the frame behind ordinal q of this call's pixels (identity unless the call renders a tile split).
__device__ __forceinline__ uint32_t frame_pixel(const RenderDev& rp, uint32_t q) { return rp.pixel_map ? __ldg(rp.pixel_map + q) : q; }

__device__ __forceinline__ void primary_ray(const CamDev& cam, const RenderDev& rp, uint32_t pixel, uint32_t sample, V3& o, V3& d) {
    using namespace ffm;
    float px = (float)(pixel % (uint32_t)cam.w), py = (float)(pixel / (uint32_t)cam.w);
    if (rp.aa) {
        const U4 j = philox4x32_10(U4{pixel, sample, 0u, 1u}, rp.seed_lo, rp.seed_hi);
        px = fadd(px, u01(j.z));
        py = fadd(py, u01(j.w));
    }
    camera_ray_dev(cam, px, py, o, d);
}


// Reciprocal direction for the slab tests only (never for the leaf test): MUFU.RCP, 1 ulp.  The culling margins
// (FF_SLACK, DevScene::pad) are seven orders of magnitude wider, and the IEEE divide cost ~25 instructions per ray at the
// low lane occupancy of the refill path.
__device__ __forceinline__ float safe_rcp(float d) {
    const float tiny = 1e-20f;
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(fabsf(d) > tiny ? d : copysignf(tiny, d)));
    return r;
}

// ------------------------------------------------------------------------------------------------
// 8-wide compressed BVH traversal (node layout in ff_bvh.cu / DESIGN.md section 3).
struct RayCtx {
    V3 o, dn;
    float idx, idy, idz;   // reciprocal direction
    uint32_t octinv;       // 7 - octant of the direction: traversal order of the child slots
    uint32_t skip;         // triangle slot the ray starts on (never tested by the SKIP traversals), 0xffffffff = none
};
// A ray the integrators spawn at P + n*eps (or P - n*eps under glass) cannot hit the triangle it starts on: it leaves the
// triangle's plane (reflection, diffuse bounce: d.n > 0, t < 0) or looks at its back face (refraction: culled).  The
// integrators pass that triangle's slot in ray_d.w (as slot + 1; 0 = none, which is what every API ray carries) and the
// SKIP instantiations of the traversal drop it from the leaf mask of the node visit that finds it instead of running a test whose answer is known (C2: 6.5 ->
// 5.7 triangle tests per ray).  The API entry points (ff_trace_closest*) run the plain instantiation: every triangle tested.
__device__ __forceinline__ uint32_t skip_from_bits(float w) { return __float_as_uint(w) - 1u; }
__device__ __forceinline__ RayCtx make_ray_ctx(V3 o, V3 dn, uint32_t skip = 0xffffffffu) {
    RayCtx r;
    r.skip = skip;
    r.o = o; r.dn = dn;
    r.idx = safe_rcp(dn.x); r.idy = safe_rcp(dn.y); r.idz = safe_rcp(dn.z);
    r.octinv = (dn.x < 0.f ? 0u : 4u) | (dn.y < 0.f ? 0u : 2u) | (dn.z < 0.f ? 0u : 1u);
    return r;
}
#define FF_SLACK 1.000030517578125f  // 1 + 2^-15: a node is culled only beyond best * FF_SLACK

// Byte j of q as a float WITHOUT the int->float converter (I2F runs on the quarter-rate XU pipe, and 48 of them per
// node saturated it: profiles/r01b) and without a subtraction: PRMT drops the byte into mantissa bits 8..15 of 1.0f, so
// the result is u = 1 + byte * 2^-15 exactly, and the plane distance is ONE FMA,
//     t = (p + byte*s - o) * id  =  u * A + B,   A = 2^15 * s * id,   B = (p - o) * id - A.
// B carries one extra rounding of relative size 2^-24 at the magnitude of A = 128 node extents, i.e. <= 2^-16 of the
// node extent in world units; the builder pushes the quantisation frame out by 2^-13 of the extent (collapse_kernel), so
// culling stays conservative.  The node stores the exponent of 2^15 * s.  `magic` = 0x3F800000 read from the kernel
// parameters (DevScene::prmt_magic): as a compile-time constant it would occupy PRMT's only immediate slot and push the
// byte selector into a register, which cost ~40 MOVs per node visit.
template <int J>
__device__ __forceinline__ float byte_unit(uint32_t q, uint32_t magic) {
    return __uint_as_float(__byte_perm(q, magic, 0x7604u + (J << 4)));
}
template <int J>
__device__ __forceinline__ float plane_t(uint32_t q, uint32_t magic, float a, float b) {
    return fmaf(byte_unit<J>(q, magic), a, b);
}

// Fetches one node (5 x LDG.128), tests its 8 quantised child boxes against the ray and returns the hit mask:
// bits 24..31 = inner children in traversal order, bits 0..23 = triangles of the leaf children.
__device__ __forceinline__ uint32_t bvh8_child_hits(const DevScene& sc, uint32_t node_index, const RayCtx& r, float tlimit,
                                                    uint32_t magic, uint32_t& child_base, uint32_t& tri_base, uint32_t& imask) {
#if FF_NODE_256
    // 96-byte node, three 32-byte loads (LDG.E.256): {n0, n1}, {qlo x,y,z + 8 spare bytes}, {qhi x,y,z + 8 spare bytes}
    const char* np = (const char*)sc.nodes + (size_t)node_index * 96;
    float4 n0, n1, n2, n3, n4;
    float s0, s1, s2, s3;
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=f"(n0.x), "=f"(n0.y), "=f"(n0.z), "=f"(n0.w), "=f"(n1.x), "=f"(n1.y), "=f"(n1.z), "=f"(n1.w) : "l"(np));
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=f"(n2.x), "=f"(n2.y), "=f"(n2.z), "=f"(n2.w), "=f"(n3.x), "=f"(n3.y), "=f"(s0), "=f"(s1) : "l"(np + 32));
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=f"(n3.z), "=f"(n3.w), "=f"(n4.x), "=f"(n4.y), "=f"(n4.z), "=f"(n4.w), "=f"(s2), "=f"(s3) : "l"(np + 64));
#elif FF_NODE_TEX
    const int ni = (int)node_index * 5;
    const float4 n0 = tex1Dfetch<float4>(sc.nodes_tex, ni), n1 = tex1Dfetch<float4>(sc.nodes_tex, ni + 1), n2 = tex1Dfetch<float4>(sc.nodes_tex, ni + 2),
                 n3 = tex1Dfetch<float4>(sc.nodes_tex, ni + 3), n4 = tex1Dfetch<float4>(sc.nodes_tex, ni + 4);
#else
    const float4* __restrict__ nodes = sc.nodes;
    const size_t ni = (size_t)node_index * FF_NODE_F4;
    const float4 n0 = __ldg(nodes + ni), n1 = __ldg(nodes + ni + 1), n2 = __ldg(nodes + ni + 2), n3 = __ldg(nodes + ni + 3),
                 n4 = __ldg(nodes + ni + 4);
#endif
    const uint32_t eb = __float_as_uint(n0.w);
    const float ax = __uint_as_float((eb & 0xffu) << 23) * r.idx;   // 2^15 * scale * 1/d (exponent stored pre-biased)
    const float ay = __uint_as_float(((eb >> 8) & 0xffu) << 23) * r.idy;
    const float az = __uint_as_float(((eb >> 16) & 0xffu) << 23) * r.idz;
    const float ox = fmaf(n0.x - r.o.x, r.idx, -ax), oy = fmaf(n0.y - r.o.y, r.idy, -ay), oz = fmaf(n0.z - r.o.z, r.idz, -az);
    child_base = __float_as_uint(n1.x);
    tri_base = __float_as_uint(n1.y);
    imask = eb >> 24;
    const uint32_t octinv4 = r.octinv * 0x01010101u;
    const bool nx = r.dn.x < 0.f, ny = r.dn.y < 0.f, nz = r.dn.z < 0.f;
    uint32_t hitmask = 0;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const uint32_t meta4 = __float_as_uint(h ? n1.w : n1.z);
        const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
        const uint32_t inner_mask4 = (is_inner4 >> 4) * 0xffu;
        const uint32_t bit_index4 = (meta4 ^ (octinv4 & inner_mask4)) & 0x1f1f1f1fu;
        const uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;
        const uint32_t qlox = __float_as_uint(h ? n2.y : n2.x), qloy = __float_as_uint(h ? n2.w : n2.z);
        const uint32_t qloz = __float_as_uint(h ? n3.y : n3.x), qhix = __float_as_uint(h ? n3.w : n3.z);
        const uint32_t qhiy = __float_as_uint(h ? n4.y : n4.x), qhiz = __float_as_uint(h ? n4.w : n4.z);
        const uint32_t nearx = nx ? qhix : qlox, farx = nx ? qlox : qhix;
        const uint32_t neary = ny ? qhiy : qloy, fary = ny ? qloy : qhiy;
        const uint32_t nearz = nz ? qhiz : qloz, farz = nz ? qloz : qhiz;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float tnx = (j == 0 ? plane_t<0>(nearx, magic, ax, ox) : j == 1 ? plane_t<1>(nearx, magic, ax, ox) : j == 2 ? plane_t<2>(nearx, magic, ax, ox) : plane_t<3>(nearx, magic, ax, ox));
            const float tny = (j == 0 ? plane_t<0>(neary, magic, ay, oy) : j == 1 ? plane_t<1>(neary, magic, ay, oy) : j == 2 ? plane_t<2>(neary, magic, ay, oy) : plane_t<3>(neary, magic, ay, oy));
            const float tnz = (j == 0 ? plane_t<0>(nearz, magic, az, oz) : j == 1 ? plane_t<1>(nearz, magic, az, oz) : j == 2 ? plane_t<2>(nearz, magic, az, oz) : plane_t<3>(nearz, magic, az, oz));
            const float tfx = (j == 0 ? plane_t<0>(farx, magic, ax, ox) : j == 1 ? plane_t<1>(farx, magic, ax, ox) : j == 2 ? plane_t<2>(farx, magic, ax, ox) : plane_t<3>(farx, magic, ax, ox));
            const float tfy = (j == 0 ? plane_t<0>(fary, magic, ay, oy) : j == 1 ? plane_t<1>(fary, magic, ay, oy) : j == 2 ? plane_t<2>(fary, magic, ay, oy) : plane_t<3>(fary, magic, ay, oy));
            const float tfz = (j == 0 ? plane_t<0>(farz, magic, az, oz) : j == 1 ? plane_t<1>(farz, magic, az, oz) : j == 2 ? plane_t<2>(farz, magic, az, oz) : plane_t<3>(farz, magic, az, oz));
            const float tn = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, 0.0f));
            const float tf = fminf(fminf(tfx, tfy), fminf(tfz, tlimit));
            if (tn <= tf) {
                const uint32_t cb = (child_bits4 >> (8 * j)) & 0xffu;
                const uint32_t bi = (bit_index4 >> (8 * j)) & 0xffu;
                hitmask |= cb << bi;
            }
        }
    }
    return hitmask;
}

// Traversal state of one ray.  `sm` is this thread's column of a [FF_SM_STACK + FF_TRI_QUEUE][stride] shared array:
// rows [0, FF_SM_STACK) are the top of the traversal stack, the rest is a ring of pending triangle groups.
struct Trav {
    uint2 ngroup, tgroup;   // node group being expanded; triangle group being tested
    int sp;
    int qhead, qcount;      // ring of triangle groups waiting behind `tgroup`
    float tlimit;
    // The spill part of the stack is NOT a member: a dynamically indexed array inside the struct would pin every field of
    // the struct to local memory (the compiler then shadows each register update with an STL: ~40 wasted store sectors
    // per ray in profiles/r01c).  Callers own `uint2 lstack[FF_LOCAL_STACK]` and pass it to push/pop.
    __device__ __forceinline__ void start(float best_key) {
        ngroup = make_uint2(0u, 0x80000000u);
        tgroup = make_uint2(0u, 0u);
        sp = 0; qhead = 0; qcount = 0;
        tlimit = best_key * FF_SLACK;
    }
    __device__ __forceinline__ void push(uint2 v, uint2* sm, int stride, uint2* lstack) {
        if (sp < FF_SM_STACK) sm[sp * stride] = v;
        else lstack[sp - FF_SM_STACK] = v;
        ++sp;
    }
    __device__ __forceinline__ uint2 pop(uint2* sm, int stride, const uint2* lstack) {
        --sp;
        return (sp < FF_SM_STACK) ? sm[sp * stride] : lstack[sp - FF_SM_STACK];
    }
    __device__ __forceinline__ void enqueue(uint2 v, uint2* sm, int stride) {
        sm[(FF_SM_STACK + ((qhead + qcount) & (FF_TRI_QUEUE - 1))) * stride] = v;
        ++qcount;
    }
    __device__ __forceinline__ uint2 dequeue(uint2* sm, int stride) {
        const uint2 v = sm[(FF_SM_STACK + qhead) * stride];
        qhead = (qhead + 1) & (FF_TRI_QUEUE - 1);
        --qcount;
        return v;
    }
};

template <bool ANY, bool COUNT>
__device__ __forceinline__ bool bvh8_test_one(const DevScene& sc, const RayCtx& r, Trav& tv, Best& best, unsigned& n_tris) {
    const uint32_t ti = 31u - __clz(tv.tgroup.y);
    tv.tgroup.y &= ~(1u << ti);
    if (COUNT) ++n_tris;
    if (test_slot<ANY>(sc, (int)(tv.tgroup.x + ti), r.o, r.dn, best)) {
        if (ANY) return true;
        tv.tlimit = best.key * FF_SLACK;
    }
    return false;
}

// One traversal iteration = at most ONE node visit followed by at most ONE triangle test.  The triangles a node visit
// finds are queued (in traversal order) instead of being tested in an inner loop: with incoherent rays the lanes of a
// warp find very different numbers of triangles per node, and the inner loop ran at ~14 % lane occupancy for ~70 % of the
// kernel's issue slots (profiles/r01b).  Decoupled, both halves of the iteration keep most lanes busy.  The accept rule
// is order independent, so the result is unchanged.  Returns false when the ray is finished.
template <bool ANY, bool COUNT, bool SKIP = false>
__device__ __forceinline__ bool bvh8_step(const DevScene& sc, const RayCtx& r, Trav& tv, Best& best, uint2* sm, int stride,
                                          uint2* lstack, unsigned& n_nodes, unsigned& n_tris, bool do_node = true, bool do_tri = true) {
    if (do_node && (tv.ngroup.y & 0xff000000u)) {
        const uint32_t hits_imask = tv.ngroup.y;
        const uint32_t bit = 31u - __clz(hits_imask);
        const uint32_t base = tv.ngroup.x;
        tv.ngroup.y &= ~(1u << bit);
        if (tv.ngroup.y & 0xff000000u) tv.push(tv.ngroup, sm, stride, lstack);
        const uint32_t slot_index = (bit - 24u) ^ r.octinv;
        const uint32_t rel = __popc(hits_imask & ~(0xffffffffu << slot_index) & 0xffu);
        uint32_t imask, tri_base;
        const uint32_t hitmask = bvh8_child_hits(sc, base + rel, r, tv.tlimit, sc.prmt_magic, tv.ngroup.x, tri_base, imask);
        if (COUNT) ++n_nodes;
        tv.ngroup.y = (hitmask & 0xff000000u) | imask;
        uint32_t tmask = hitmask & 0x00ffffffu;
        if (SKIP) { const uint32_t rel = r.skip - tri_base; if (rel < 24u) tmask &= ~(1u << rel); }
        if (tmask) {
            if (tv.tgroup.y == 0u) tv.tgroup = make_uint2(tri_base, tmask);
            else {
                while (tv.qcount == FF_TRI_QUEUE) {  // ring full (rare): drain the group in hand
                    while (tv.tgroup.y != 0u)
                        if (bvh8_test_one<ANY, COUNT>(sc, r, tv, best, n_tris)) return false;
                    tv.tgroup = tv.dequeue(sm, stride);
                }
                tv.enqueue(make_uint2(tri_base, tmask), sm, stride);
            }
        }
    }
    // (do_tri false = hold the triangle phase back this iteration, unless the lane has nothing else to do)
    if ((do_tri || (tv.ngroup.y & 0xff000000u) == 0u) && tv.tgroup.y != 0u) {
        if (bvh8_test_one<ANY, COUNT>(sc, r, tv, best, n_tris)) return false;
        if (tv.tgroup.y == 0u && tv.qcount > 0) tv.tgroup = tv.dequeue(sm, stride);
    }
    if ((tv.ngroup.y & 0xff000000u) == 0u) {
        if (tv.sp > 0) tv.ngroup = tv.pop(sm, stride, lstack);
        else if (tv.tgroup.y == 0u) return false;
    }
    return true;
}

template <bool ANY, bool COUNT, bool SKIP = false>
__device__ __forceinline__ void trace_bvh8(const DevScene& sc, V3 o, V3 dn, Best& best, uint2* sm_stack, int sm_stride,
                                           unsigned& n_nodes, unsigned& n_tris, uint32_t skip = 0xffffffffu) {
    if (sc.n_nodes == 0) return;
    const RayCtx r = make_ray_ctx(o, dn, skip);
    Trav tv;
    uint2 lstack[FF_LOCAL_STACK];
    tv.start(best.key);
    while (bvh8_step<ANY, COUNT, SKIP>(sc, r, tv, best, sm_stack, sm_stride, lstack, n_nodes, n_tris)) {}
}

// Binary LBVH traversal (comparison / bridge variant).
template <bool ANY, bool COUNT>
__device__ __forceinline__ void trace_bvh2(const DevScene& sc, V3 o, V3 dn, Best& best, unsigned& n_nodes, unsigned& n_tris) {
    if (sc.n_tris == 0) return;
    const uint32_t n_internal = sc.n_tris - 1;
    const float idx_ = safe_rcp(dn.x), idy_ = safe_rcp(dn.y), idz_ = safe_rcp(dn.z);
    const float slack = FF_SLACK;
    uint32_t stack[64];
    int sp = 0;
    uint32_t node = 0;
    if (n_internal == 0) {
        if (COUNT) ++n_tris;
        test_slot<ANY>(sc, (int)sc.bvh2_leaf_slot[0], o, dn, best);
        return;
    }
    for (;;) {
        if (node >= n_internal) {
            if (COUNT) ++n_tris;
            if (test_slot<ANY>(sc, (int)sc.bvh2_leaf_slot[node - n_internal], o, dn, best) && ANY) return;
        } else {
            if (COUNT) ++n_nodes;
            const int2 ch = sc.bvh2_children[node];
            float tn[2];
            bool hit[2];
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const uint32_t c = (uint32_t)(k ? ch.y : ch.x);
                const float4 lo = __ldg(sc.bvh2_lo + c), hi = __ldg(sc.bvh2_hi + c);
                const float x0 = (lo.x - o.x) * idx_, x1 = (hi.x - o.x) * idx_;
                const float y0 = (lo.y - o.y) * idy_, y1 = (hi.y - o.y) * idy_;
                const float z0 = (lo.z - o.z) * idz_, z1 = (hi.z - o.z) * idz_;
                const float a = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), 0.0f));
                const float b = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), best.key * slack));
                // widen by a relative 2^-20 so rounding in the slab arithmetic can never cull
                hit[k] = a * 0.99999905f <= b * 1.00000095f;
                tn[k] = a;
            }
            if (hit[0] && hit[1]) {
                const bool first0 = tn[0] <= tn[1];
                stack[sp++] = (uint32_t)(first0 ? ch.y : ch.x);
                node = (uint32_t)(first0 ? ch.x : ch.y);
                continue;
            }
            if (hit[0]) { node = (uint32_t)ch.x; continue; }
            if (hit[1]) { node = (uint32_t)ch.y; continue; }
        }
        if (sp == 0) break;
        node = stack[--sp];
    }
}

}  // namespace ff
