// ff_bvh.cu -- GPU BVH builder: bake -> bounds -> Morton codes -> LSD radix sort -> Karras hierarchy ->
// bottom-up refit -> collapse to the 8-wide compressed layout with triangles re-ordered into leaf order.
//
// Replaces the reference's "no acceleration structure" scan (kernel.cu:133-155: every ray loops over every
// geometry and every triangle) and its pointer-laden upload (kernel.cu:268-298).
//
// 8-wide node, 80 B = 5 x float4 (one LDG.128 each):
//   n0 = { p.x, p.y, p.z, ex | ey<<8 | ez<<16 | imask<<24 }     p = frame origin, e* = biased IEEE exponents of 2^15 * scale
//   n1 = { child_base, tri_base, meta[0..3], meta[4..7] }        u32 each
//   n2 = { qlo_x[0..3], qlo_x[4..7], qlo_y[0..3], qlo_y[4..7] }  8-bit quantised child planes
//   n3 = { qlo_z[0..3], qlo_z[4..7], qhi_x[0..3], qhi_x[4..7] }
//   n4 = { qhi_y[0..3], qhi_y[4..7], qhi_z[0..3], qhi_z[4..7] }
//   meta: empty 0; inner child in slot s: 0x20 | (24 + s); leaf: (unary triangle count 1/3/7)<<5 | offset from tri_base
//   child box (decoded) = p + q * 2^(e-127), always a superset of the exact child box.
#include <cooperative_groups.h>
#include <cooperative_groups/reduce.h>
#include <cooperative_groups/scan.h>
#include <cfloat>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include "ff_internal.h"

namespace cg = cooperative_groups;
namespace ff {

using namespace ffm;

static inline int grid_for(size_t n, int block, int max_blocks) {
    size_t g = (n + block - 1) / block;
    if (g < 1) g = 1;
    if (g > (size_t)max_blocks) g = max_blocks;
    return (int)g;
}

// ------------------------------------------------------------------ bake: object space -> world space SoA
// v_w = vec3(M * vec4(v, 1)) in glm's operation order (type_mat4x4.inl:536-575), so the baked vertices are
// bit-identical to the oracle's flatten().
__device__ __forceinline__ V3 xform_point(const M4& m, float x, float y, float z) {
    V3 r;
    r.x = fadd(fadd(fmul(m.c[0].x, x), fmul(m.c[1].x, y)), fadd(fmul(m.c[2].x, z), fmul(m.c[3].x, 1.0f)));
    r.y = fadd(fadd(fmul(m.c[0].y, x), fmul(m.c[1].y, y)), fadd(fmul(m.c[2].y, z), fmul(m.c[3].y, 1.0f)));
    r.z = fadd(fadd(fmul(m.c[0].z, x), fmul(m.c[1].z, y)), fadd(fmul(m.c[2].z, z), fmul(m.c[3].z, 1.0f)));
    return r;
}
__global__ void bake_kernel(const float* __restrict__ xyz9, const M4* __restrict__ models, const uint32_t* __restrict__ geom_of,
                            const uint32_t* __restrict__ ord_of, uint32_t n, float4* __restrict__ out) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t g = geom_of[i];
        const M4 m = models[g];
        const float* p = xyz9 + 9 * (size_t)i;
        const V3 a = xform_point(m, p[0], p[1], p[2]), b = xform_point(m, p[3], p[4], p[5]), c = xform_point(m, p[6], p[7], p[8]);
        out[3 * (size_t)i] = make_float4(a.x, a.y, a.z, __int_as_float((int)ord_of[i]));
        out[3 * (size_t)i + 1] = make_float4(b.x, b.y, b.z, __int_as_float((int)g));
        out[3 * (size_t)i + 2] = make_float4(c.x, c.y, c.z, __int_as_float((int)i));
    }
}
void k_bake_triangles(const float* d_xyz9_obj, const M4* d_model_per_geom, const uint32_t* d_geom_of_flat,
                      const uint32_t* d_ord_of_flat, uint32_t n, float4* d_flat_tris, cudaStream_t st) {
    if (n) bake_kernel<<<grid_for(n, 256, sm_count() * 16), 256, 0, st>>>(d_xyz9_obj, d_model_per_geom, d_geom_of_flat, d_ord_of_flat, n, d_flat_tris);
}

// As-coded shading normal of every triangle, flat order: inverse(transpose(M)) * vec4(normalize(cross(e1, e2)), 0) with
// OBJECT-space edges -- what intersectTriangle leaves in m_normal (kernel.cu:101) pushed through setIntersection
// (kernel.cu:117), un-normalised, so a mesh scaled by s comes out 1/s as bright, exactly as in the reference's frame.
__global__ void as_coded_normals_kernel(const float* __restrict__ xyz9, const M4* __restrict__ nrm_of_geom, const uint32_t* __restrict__ geom_of,
                                        uint32_t n, float4* __restrict__ out) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const M4 m = nrm_of_geom[geom_of[i]];
        const float* p = xyz9 + 9 * (size_t)i;
        const V3 v0 = v3(p[0], p[1], p[2]), v1 = v3(p[3], p[4], p[5]), v2 = v3(p[6], p[7], p[8]);
        const V3 no = normalize(cross(sub(v1, v0), sub(v2, v0)));
        float4 w;  // glm mat4 * vec4(no, 0) (type_mat4x4.inl:536-575)
        w.x = fadd(fadd(fmul(m.c[0].x, no.x), fmul(m.c[1].x, no.y)), fadd(fmul(m.c[2].x, no.z), fmul(m.c[3].x, 0.0f)));
        w.y = fadd(fadd(fmul(m.c[0].y, no.x), fmul(m.c[1].y, no.y)), fadd(fmul(m.c[2].y, no.z), fmul(m.c[3].y, 0.0f)));
        w.z = fadd(fadd(fmul(m.c[0].z, no.x), fmul(m.c[1].z, no.y)), fadd(fmul(m.c[2].z, no.z), fmul(m.c[3].z, 0.0f)));
        w.w = 0.f;
        out[i] = w;
    }
}
void k_as_coded_normals(const float* d_xyz9_obj, const M4* d_nrm_per_geom, const uint32_t* d_geom_of_flat, uint32_t n,
                        float4* d_out, cudaStream_t st) {
    if (n) as_coded_normals_kernel<<<grid_for(n, 256, sm_count() * 16), 256, 0, st>>>(d_xyz9_obj, d_nrm_per_geom, d_geom_of_flat, n, d_out);
}

// ------------------------------------------------------------------ bounds
__device__ __forceinline__ int float_to_ordered(float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__host__ __device__ __forceinline__ float ordered_to_float(int i) {
    int b = i >= 0 ? i : i ^ 0x7fffffff;
#if defined(__CUDA_ARCH__)
    return __int_as_float(b);
#else
    float f; memcpy(&f, &b, 4); return f;
#endif
}
// bounds[0..2] = scene min, [3..5] = scene max, [6..8] = centroid min, [9..11] = centroid max (ordered ints)
__global__ void bounds_init_kernel(int* bounds) {
    if (threadIdx.x < 12) bounds[threadIdx.x] = ((threadIdx.x / 3) & 1) ? float_to_ordered(-FLT_MAX) : float_to_ordered(FLT_MAX);
}
__global__ void bounds_kernel(const float4* __restrict__ tris, uint32_t n, int* bounds) {
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    float clo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, chi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 a = tris[3 * (size_t)i], b = tris[3 * (size_t)i + 1], c = tris[3 * (size_t)i + 2];
        const float av[3] = {a.x, a.y, a.z}, bv[3] = {b.x, b.y, b.z}, cv[3] = {c.x, c.y, c.z};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float l = fminf(fminf(av[k], bv[k]), cv[k]), h = fmaxf(fmaxf(av[k], bv[k]), cv[k]);
            const float ce = 0.5f * (l + h);
            lo[k] = fminf(lo[k], l); hi[k] = fmaxf(hi[k], h);
            clo[k] = fminf(clo[k], ce); chi[k] = fmaxf(chi[k], ce);
        }
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        for (int off = 16; off; off >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], off));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], off));
            clo[k] = fminf(clo[k], __shfl_xor_sync(0xffffffffu, clo[k], off));
            chi[k] = fmaxf(chi[k], __shfl_xor_sync(0xffffffffu, chi[k], off));
        }
        if ((threadIdx.x & 31) == 0) {
            atomicMin(bounds + k, float_to_ordered(lo[k]));
            atomicMax(bounds + 3 + k, float_to_ordered(hi[k]));
            atomicMin(bounds + 6 + k, float_to_ordered(clo[k]));
            atomicMax(bounds + 9 + k, float_to_ordered(chi[k]));
        }
    }
}

// ------------------------------------------------------------------ Morton codes (30 bit, centroid in centroid bounds)
__device__ __forceinline__ uint32_t expand10(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
__global__ void morton_kernel(const float4* __restrict__ tris, uint32_t n, const int* __restrict__ bounds, uint32_t* keys, uint32_t* vals) {
    float clo[3], inv[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        clo[k] = ordered_to_float(bounds[6 + k]);
        const float ext = ordered_to_float(bounds[9 + k]) - clo[k];
        inv[k] = ext > 0.f ? 1024.0f / ext : 0.f;
    }
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 a = tris[3 * (size_t)i], b = tris[3 * (size_t)i + 1], c = tris[3 * (size_t)i + 2];
        const float cx = 0.5f * (fminf(fminf(a.x, b.x), c.x) + fmaxf(fmaxf(a.x, b.x), c.x));
        const float cy = 0.5f * (fminf(fminf(a.y, b.y), c.y) + fmaxf(fmaxf(a.y, b.y), c.y));
        const float cz = 0.5f * (fminf(fminf(a.z, b.z), c.z) + fmaxf(fmaxf(a.z, b.z), c.z));
        const uint32_t x = min(1023u, (uint32_t)max(0.f, (cx - clo[0]) * inv[0]));
        const uint32_t y = min(1023u, (uint32_t)max(0.f, (cy - clo[1]) * inv[1]));
        const uint32_t z = min(1023u, (uint32_t)max(0.f, (cz - clo[2]) * inv[2]));
        keys[i] = (expand10(x) << 2) | (expand10(y) << 1) | expand10(z);
        vals[i] = i;
    }
}

// ------------------------------------------------------------------ LSD radix sort, 8-bit digits, key/value u32
// Three kernels per pass: per-tile digit histogram -> exclusive scan over (digit, tile) -> stable scatter.
#define RS_THREADS 256
#define RS_ITEMS 16
#define RS_TILE (RS_THREADS * RS_ITEMS)
__global__ void __launch_bounds__(RS_THREADS) rs_hist_kernel(const uint32_t* __restrict__ keys, uint32_t n, int shift, uint32_t n_tiles,
                                                              uint32_t* __restrict__ hist /*[256][n_tiles]*/) {
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t base = blockIdx.x * RS_TILE;
    for (int r = 0; r < RS_ITEMS; ++r) {
        const uint32_t i = base + r * RS_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&h[(keys[i] >> shift) & 0xffu], 1u);
    }
    __syncthreads();
    hist[(size_t)threadIdx.x * n_tiles + blockIdx.x] = h[threadIdx.x];
}
// Exclusive scan of the (digit, tile) counters in two steps: block d scans row d (the tiles of digit d) in place and
// leaves the row total; rs_scan_totals_kernel scans the 256 totals; the scatter adds the two.  (One block over all
// 256 * n_tiles counters was 0.4 ms per pass at 20 M keys: 1.6 of the sort's 3.5 ms.)
__global__ void __launch_bounds__(1024) rs_scan_kernel(uint32_t* hist, uint32_t m, uint32_t* __restrict__ totals) {
    uint32_t* data = hist + (size_t)blockIdx.x * m;
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const uint32_t per = 4;
    for (uint32_t base = 0; base < m; base += 1024 * per) {
        uint32_t v[per], sum = 0;
        const uint32_t i0 = base + threadIdx.x * per;
#pragma unroll
        for (uint32_t k = 0; k < per; ++k) { v[k] = (i0 + k < m) ? data[i0 + k] : 0u; sum += v[k]; }
        uint32_t incl = sum;
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, off);
            if ((threadIdx.x & 31) >= off) incl += t;
        }
        if ((threadIdx.x & 31) == 31) warp_sums[threadIdx.x >> 5] = incl;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint32_t w = warp_sums[threadIdx.x], wi = w;
            for (int off = 1; off < 32; off <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, wi, off);
                if (threadIdx.x >= off) wi += t;
            }
            warp_sums[threadIdx.x] = wi - w;
        }
        __syncthreads();
        uint32_t excl = carry + warp_sums[threadIdx.x >> 5] + incl - sum;
#pragma unroll
        for (uint32_t k = 0; k < per; ++k) { if (i0 + k < m) data[i0 + k] = excl; excl += v[k]; }
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl;
        __syncthreads();
    }
    if (threadIdx.x == 0) totals[blockIdx.x] = carry;
}
__global__ void __launch_bounds__(256) rs_scan_totals_kernel(uint32_t* totals) {
    __shared__ uint32_t s[256];
    const uint32_t v = totals[threadIdx.x];
    s[threadIdx.x] = v;
    __syncthreads();
    for (int off = 1; off < 256; off <<= 1) {
        const uint32_t t = threadIdx.x >= (unsigned)off ? s[threadIdx.x - off] : 0u;
        __syncthreads();
        s[threadIdx.x] += t;
        __syncthreads();
    }
    totals[threadIdx.x] = s[threadIdx.x] - v;
}
__global__ void __launch_bounds__(RS_THREADS) rs_scatter_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                                                                 uint32_t n, int shift, uint32_t n_tiles, const uint32_t* __restrict__ hist,
                                                                 uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out) {
    // warp w owns the contiguous chunk [w*512, (w+1)*512) of the tile, processed 32 keys per round, so ranks are stable
    __shared__ uint32_t warp_hist[RS_THREADS / 32][256];
    __shared__ uint32_t digit_base[256];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int k = threadIdx.x; k < (RS_THREADS / 32) * 256; k += RS_THREADS) (&warp_hist[0][0])[k] = 0;
    __syncthreads();
    const uint32_t base = blockIdx.x * RS_TILE + warp * (32 * RS_ITEMS);
    uint32_t key[RS_ITEMS], val[RS_ITEMS], rank[RS_ITEMS];
#pragma unroll
    for (int r = 0; r < RS_ITEMS; ++r) {
        const uint32_t i = base + r * 32 + lane;
        const bool live = i < n;
        key[r] = live ? keys_in[i] : 0xffffffffu;
        val[r] = live ? vals_in[i] : 0u;
        const uint32_t d = (key[r] >> shift) & 0xffu;
        const unsigned peers = __match_any_sync(0xffffffffu, live ? d : 0x100u + 0u);
        const int leader = __ffs(peers) - 1;
        uint32_t start = 0;
        if (live && lane == leader) { start = warp_hist[warp][d]; warp_hist[warp][d] = start + __popc(peers); }
        start = __shfl_sync(0xffffffffu, start, leader);
        rank[r] = start + __popc(peers & ((1u << lane) - 1u));
        __syncwarp();
    }
    __syncthreads();
    {   // per digit: exclusive prefix over warps, plus the tile's global base
        const uint32_t d = threadIdx.x;
        uint32_t run = hist[(size_t)d * n_tiles + blockIdx.x] + hist[(size_t)256 * n_tiles + d];  // row prefix + prefix of the row totals
        digit_base[d] = run;
        uint32_t acc = 0;
        for (int w = 0; w < RS_THREADS / 32; ++w) { const uint32_t c = warp_hist[w][d]; warp_hist[w][d] = acc; acc += c; }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < RS_ITEMS; ++r) {
        const uint32_t i = base + r * 32 + lane;
        if (i < n) {
            const uint32_t d = (key[r] >> shift) & 0xffu;
            const uint32_t pos = digit_base[d] + warp_hist[warp][d] + rank[r];
            keys_out[pos] = key[r];
            vals_out[pos] = val[r];
        }
    }
}
// Sorts (keys, vals) by `bits` low bits of the key; result ends in keys_a/vals_a (bits/8 must be even) .
static void radix_sort_pairs(uint32_t* keys_a, uint32_t* vals_a, uint32_t* keys_b, uint32_t* vals_b, uint32_t n, uint32_t* hist,
                             uint32_t* launches, cudaStream_t st) {
    const uint32_t n_tiles = (n + RS_TILE - 1) / RS_TILE;
    for (int pass = 0; pass < 4; ++pass) {
        const int shift = 8 * pass;
        const uint32_t* ki = (pass & 1) ? keys_b : keys_a;
        const uint32_t* vi = (pass & 1) ? vals_b : vals_a;
        uint32_t* ko = (pass & 1) ? keys_a : keys_b;
        uint32_t* vo = (pass & 1) ? vals_a : vals_b;
        rs_hist_kernel<<<n_tiles, RS_THREADS, 0, st>>>(ki, n, shift, n_tiles, hist);
        rs_scan_kernel<<<256, 1024, 0, st>>>(hist, n_tiles, hist + (size_t)256 * n_tiles);
        rs_scan_totals_kernel<<<1, 256, 0, st>>>(hist + (size_t)256 * n_tiles);
        rs_scatter_kernel<<<n_tiles, RS_THREADS, 0, st>>>(ki, vi, n, shift, n_tiles, hist, ko, vo);
        *launches += 4;
    }
}

// ------------------------------------------------------------------ Karras 2012 hierarchy
__device__ __forceinline__ int delta(const uint32_t* __restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const uint32_t a = keys[i], b = keys[j];
    if (a != b) return __clz(a ^ b);
    return 32 + __clz((uint32_t)i ^ (uint32_t)j);
}
// node ids: internal i in [0, n-1), leaf k -> (n-1) + k
__global__ void hierarchy_kernel(const uint32_t* __restrict__ keys, int n, int2* __restrict__ children, uint32_t* __restrict__ parent) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n - 1; i += gridDim.x * blockDim.x) {
        const int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
        const int dmin = delta(keys, n, i, i - d);
        int lmax = 2;
        while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
        int l = 0;
        for (int t = lmax >> 1; t >= 1; t >>= 1)
            if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
        const int j = i + l * d;
        const int dnode = delta(keys, n, i, j);
        int s = 0, t = l;
        do {
            t = (t + 1) >> 1;
            if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
        } while (t > 1);
        const int gamma = i + s * d + min(d, 0);
        const int lo = min(i, j), hi = max(i, j);
        const int left = (lo == gamma) ? (n - 1 + gamma) : gamma;
        const int right = (hi == gamma + 1) ? (n - 1 + gamma + 1) : gamma + 1;
        children[i] = make_int2(left, right);
        parent[left] = (uint32_t)i;
        parent[right] = (uint32_t)i;
        if (i == 0) parent[0] = 0xffffffffu;
    }
}

#define FF_COUNT_MASK 0x3fffffff
#ifndef FF_REFIT_FENCE
#define FF_REFIT_FENCE 1   // 2 = the two __threadfence() per level of r01h
#endif
#define FF_SLOT_COST 0.25f   // cost of one hit child (mask assembly, queueing), in triangle tests
#define FF_NODE_COST 1.5f    // cost of one wide-node visit (fetch + 8 box tests) in triangle tests (~240 vs ~150 instructions)
__device__ __forceinline__ float box_area_lohi(const float4& lo, const float4& hi) {
    const float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
    return 2.0f * (dx * dy + dy * dz + dz * dx);
}
// ------------------------------------------------------------------ bottom-up refit + collapse DP
// lo.w = triangle count (int bits), hi.w = first sorted position (int bits)
//
// The same pass runs the dynamic programme that decides how the binary tree is cut into 8-wide nodes (after Ylitie,
// Karras & Laine 2017, section 3.1, with this kernel's costs).  For a binary node v and a budget of j parent slots,
//     F(v, j) = cheapest way to hang the subtree of v under a wide node using at most j of its child slots
//     F(v, 1) = area(v) * (slot + count(v))                 count(v) <= 3: one leaf child
//             = area(v) * (slot + node) + min_k F(l, k) + F(r, 8 - k)       otherwise: one inner child, a wide node of its own
//     F(v, j) = min( F(v, 1), min_k F(l, k) + F(r, j - k) )                 j = 2..7
// in units of one triangle test.  dp[2v] = {F1..F4}, dp[2v+1] = {F5, F6, F7, decisions}: decisions bits 3(j-2).. = the
// left child's share k for budget j (0 = take a single slot), bits 21.. = k of the node's own 8 slots.  The greedy
// "open the largest child" collapse this replaces left 40 % of C2's nodes with two children (profiles/r01g_node_stats).
__global__ void refit_kernel(const float4* __restrict__ flat_tris, const uint32_t* __restrict__ sorted_idx, int n, float pad,
                             const int2* __restrict__ children, const uint32_t* __restrict__ parent, uint32_t* __restrict__ flags,
                             float4* __restrict__ box_lo, float4* __restrict__ box_hi, float4* __restrict__ dp, float slot_cost,
                             float node_cost) {
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
        const uint32_t src = sorted_idx[k];
        const float4 a = flat_tris[3 * (size_t)src], b = flat_tris[3 * (size_t)src + 1], c = flat_tris[3 * (size_t)src + 2];
        float4 lo = make_float4(fminf(fminf(a.x, b.x), c.x) - pad, fminf(fminf(a.y, b.y), c.y) - pad, fminf(fminf(a.z, b.z), c.z) - pad, __int_as_float(1));
        float4 hi = make_float4(fmaxf(fmaxf(a.x, b.x), c.x) + pad, fmaxf(fmaxf(a.y, b.y), c.y) + pad, fmaxf(fmaxf(a.z, b.z), c.z) + pad, __int_as_float(k));
        uint32_t node = (uint32_t)(n - 1 + k);
        box_lo[node] = lo;
        box_hi[node] = hi;
        if (n == 1) return;   // (a single triangle has no F table: F(leaf, j) = area * (slot + 1) is formed by its parent)
        for (;;) {
            const uint32_t p = parent[node];
            if (p == 0xffffffffu) break;
            // acq_rel: RELEASE so this thread's box and table are visible before its arrival is counted; ACQUIRE so the second
            // arrival's loads of its sibling's box and table are ordered after the atomic by the PTX memory model itself (a
            // compiler barrier plus a control dependency is not an ordering there).  One atomic instead of the fence.sc + L1
            // invalidate that __threadfence() cost twice per level (39 % of the kernel's stall samples,
            // profiles/r01i_refit_before_stalls.txt); the loads below additionally bypass L1 (ld.global.cg).
#if FF_REFIT_FENCE == 2
            __threadfence();
            if (atomicAdd(&flags[p], 1u) == 0u) break;  // first arrival: the sibling's thread finishes this node
            __threadfence();
#else
            uint32_t arrived;
            asm volatile("atom.global.add.acq_rel.gpu.u32 %0, [%1], 1;" : "=r"(arrived) : "l"(__cvta_generic_to_global(flags + p)) : "memory");
            if (arrived == 0u) break;  // first arrival: the sibling's thread finishes this node
#endif
            const int2 ch = children[p];
#if FF_REFIT_FENCE == 2
            const volatile float4* vlo = box_lo;
            const volatile float4* vhi = box_hi;
            float4 l0, l1, h0, h1;
            l0.x = vlo[ch.x].x; l0.y = vlo[ch.x].y; l0.z = vlo[ch.x].z; l0.w = vlo[ch.x].w;
            l1.x = vlo[ch.y].x; l1.y = vlo[ch.y].y; l1.z = vlo[ch.y].z; l1.w = vlo[ch.y].w;
            h0.x = vhi[ch.x].x; h0.y = vhi[ch.x].y; h0.z = vhi[ch.x].z; h0.w = vhi[ch.x].w;
            h1.x = vhi[ch.y].x; h1.y = vhi[ch.y].y; h1.z = vhi[ch.y].z; h1.w = vhi[ch.y].w;
#else
            // ld.global.cg: 128-bit loads served by L2 (never a stale L1 line), ordered after the atomic by its "memory" clobber
            const float4 l0 = __ldcg(box_lo + ch.x), l1 = __ldcg(box_lo + ch.y), h0 = __ldcg(box_hi + ch.x), h1 = __ldcg(box_hi + ch.y);
#endif
            const int cnt = (__float_as_int(l0.w) & FF_COUNT_MASK) + (__float_as_int(l1.w) & FF_COUNT_MASK);
            lo = make_float4(fminf(l0.x, l1.x), fminf(l0.y, l1.y), fminf(l0.z, l1.z), 0.f);
            hi = make_float4(fmaxf(h0.x, h1.x), fmaxf(h0.y, h1.y), fmaxf(h0.z, h1.z), h0.w);
            // F(left, 1..7), F(right, 1..7)
            float fl[8], fr[8];
            if ((uint32_t)ch.x < (uint32_t)(n - 1)) {
                const float4 d0 = __ldcg(dp + 2 * (size_t)ch.x), d1 = __ldcg(dp + 2 * (size_t)ch.x + 1);
                fl[1] = d0.x; fl[2] = d0.y; fl[3] = d0.z; fl[4] = d0.w; fl[5] = d1.x; fl[6] = d1.y; fl[7] = d1.z;
            } else {  // a triangle: one leaf child whatever the budget (only internal nodes have a table)
                const float f1 = box_area_lohi(l0, h0) * (slot_cost + 1.0f);
#pragma unroll
                for (int q = 1; q <= 7; ++q) fl[q] = f1;
            }
            if ((uint32_t)ch.y < (uint32_t)(n - 1)) {
                const float4 d0 = __ldcg(dp + 2 * (size_t)ch.y), d1 = __ldcg(dp + 2 * (size_t)ch.y + 1);
                fr[1] = d0.x; fr[2] = d0.y; fr[3] = d0.z; fr[4] = d0.w; fr[5] = d1.x; fr[6] = d1.y; fr[7] = d1.z;
            } else {
                const float f1 = box_area_lohi(l1, h1) * (slot_cost + 1.0f);
#pragma unroll
                for (int q = 1; q <= 7; ++q) fr[q] = f1;
            }
            const float area = box_area_lohi(lo, hi);
            float f[8];
            uint32_t dec = 0;
            {
                float g = FLT_MAX;
                uint32_t k8 = 1;
#pragma unroll
                for (int q = 1; q <= 7; ++q) { const float v = fl[q] + fr[8 - q]; if (v < g) { g = v; k8 = (uint32_t)q; } }
                dec |= k8 << 21;
                f[1] = (cnt <= 3) ? area * (slot_cost + (float)cnt) : area * (slot_cost + node_cost) + g;
            }
#pragma unroll
            for (int j = 2; j <= 7; ++j) {
                float g = f[1];
                uint32_t kj = 0;
#pragma unroll
                for (int q = 1; q < j; ++q) { const float v = fl[q] + fr[j - q]; if (v < g) { g = v; kj = (uint32_t)q; } }
                f[j] = g;
                dec |= kj << (3 * (j - 2));
            }
            dp[2 * (size_t)p] = make_float4(f[1], f[2], f[3], f[4]);
            dp[2 * (size_t)p + 1] = make_float4(f[5], f[6], f[7], __uint_as_float(dec));
            lo.w = __int_as_float(cnt);
            box_lo[p] = lo;
            box_hi[p] = hi;
            node = p;
        }
    }
}

// ------------------------------------------------------------------ collapse to 8-wide compressed nodes
struct CollapseCounters {
    uint32_t node_count;   // wide nodes allocated so far
    uint32_t tri_count;    // triangle slots allocated so far
    uint32_t level_begin, level_end;  // wide nodes [begin, end) form the level being collapsed (indices are allocated in level order)
    uint32_t levels;       // non-empty levels processed so far
    uint32_t overflow;     // a node allocation ran past the capacity of the node arrays
    double sah;            // sum over wide nodes of area(node) * (1 + sum_leaf area_ratio...) -- see below
};
__device__ __forceinline__ float box_area(const float4& lo, const float4& hi) {
    const float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
    return 2.0f * (dx * dy + dy * dz + dz * dx);
}
__device__ __forceinline__ uint32_t exponent_for(float extent) {
    // smallest biased exponent eb with 255 * 2^(eb-127) >= extent
    if (!(extent > 1e-30f)) return 1u;
    int ex;
    frexpf(extent / 255.0f, &ex);  // extent/255 = m * 2^ex, m in [0.5, 1)  =>  2^ex >= extent/255
    int eb = ex + 127;
    if (eb < 1) eb = 1;
    if (eb > 254 - FF_NODE_EXP_BIAS) eb = 254 - FF_NODE_EXP_BIAS;
    float sc = __uint_as_float((uint32_t)eb << 23);
    while (sc * 255.0f < extent && eb < 254 - FF_NODE_EXP_BIAS) { ++eb; sc = __uint_as_float((uint32_t)eb << 23); }
    return (uint32_t)eb;
}
__device__ __forceinline__ uint32_t quant_lo(float c, float p, float sc) {
    float q = floorf((c - p) / sc);
    q = fminf(fmaxf(q, 0.f), 255.f);
    while (q > 0.f && p + q * sc > c) q -= 1.f;
    return (uint32_t)q;
}
__device__ __forceinline__ uint32_t quant_hi(float c, float p, float sc) {
    float q = ceilf((c - p) / sc);
    q = fminf(fmaxf(q, 0.f), 255.f);
    while (q < 255.f && p + q * sc < c) q += 1.f;
    return (uint32_t)q;
}

// One thread per wide node.  node_src[w] = the binary node wide node w is built from.  Wide-node indices are handed out
// by one counter, so the nodes of a level are a contiguous index range and the levels need no host round trip: the kernel
// reads its range from the counters block, collapse_advance_kernel moves the range to the nodes the level allocated.
__device__ __forceinline__ void collapse_one(uint32_t bnode, uint32_t wnode, uint32_t* __restrict__ node_src, uint32_t node_capacity, int n,
                                const int2* __restrict__ children, const float4* __restrict__ box_lo, const float4* __restrict__ box_hi,
                                const float4* __restrict__ flat_tris, const uint32_t* __restrict__ sorted_idx,
                                float4* __restrict__ nodes, float4* __restrict__ tris_out, uint32_t* __restrict__ leaf_slot,
                                CollapseCounters* ctr, float root_area, const float4* __restrict__ dp) {
    const uint32_t n_internal = (uint32_t)(n - 1);

    uint32_t ch[8];
    int nch = 0;
    float4 plo = box_lo[bnode], phi = box_hi[bnode];
    if (bnode >= n_internal || (__float_as_int(plo.w) & FF_COUNT_MASK) <= 3) {
        ch[nch++] = bnode;  // tiny scene: the whole tree is one leaf child
    } else {
        // follow the decisions of the refit DP: (binary node, slot budget) pairs, at most 8 slots in flight
        uint32_t st_node[8];
        uint32_t st_bud[8];
        int sp = 0;
        {
            const uint32_t k8 = (__float_as_uint(dp[2 * (size_t)bnode + 1].w) >> 21) & 7u;
            st_node[sp] = (uint32_t)children[bnode].y; st_bud[sp++] = 8u - k8;
            st_node[sp] = (uint32_t)children[bnode].x; st_bud[sp++] = k8;
        }
        while (sp > 0 && nch < 8) {
            const uint32_t m = st_node[--sp], j = st_bud[sp];
            uint32_t d = 0;
            if (m < n_internal && j >= 2u) d = (__float_as_uint(dp[2 * (size_t)m + 1].w) >> (3u * (j - 2u))) & 7u;
            if (d == 0u || d >= j || sp + 2 > 8) { ch[nch++] = m; continue; }   // one slot: a leaf (<= 3 triangles) or an inner child
            const int2 c2 = children[m];
            st_node[sp] = (uint32_t)c2.y; st_bud[sp++] = j - d;
            st_node[sp] = (uint32_t)c2.x; st_bud[sp++] = d;
        }
    }

    // slot assignment: greedy maximisation of dot(child centre - node centre, octant direction of the slot)
    const float pcx = 0.5f * (plo.x + phi.x), pcy = 0.5f * (plo.y + phi.y), pcz = 0.5f * (plo.z + phi.z);
    // (the 8 x 8 cost table used to live in local memory: 512 LDL per node, 80 % of this kernel's stall samples at the
    // 2.4 M-node level of a 20 M-triangle build; the costs are now recomputed from 24 registers, same expression, same order)
    float cdx[8], cdy[8], cdz[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        cdx[k] = cdy[k] = cdz[k] = 0.f;
        if (k < nch) {
            const float4 l = box_lo[ch[k]], h = box_hi[ch[k]];
            cdx[k] = 0.5f * (l.x + h.x) - pcx; cdy[k] = 0.5f * (l.y + h.y) - pcy; cdz[k] = 0.5f * (l.z + h.z) - pcz;
        }
    }
    int slot_child[8];
    for (int s = 0; s < 8; ++s) slot_child[s] = -1;
    uint32_t assigned = 0, used = 0;
    for (int it = 0; it < nch; ++it) {
        float bestc = -FLT_MAX;
        int bk = -1, bs = -1;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            if (k >= nch || (assigned & (1u << k))) continue;
#pragma unroll
            for (int s = 0; s < 8; ++s) {
                if (used & (1u << s)) continue;
                const float c = ((s & 4) ? cdx[k] : -cdx[k]) + ((s & 2) ? cdy[k] : -cdy[k]) + ((s & 1) ? cdz[k] : -cdz[k]);
                if (c > bestc) { bestc = c; bk = k; bs = s; }
            }
        }
        if (bk < 0) {  // NaN costs: fall back to first free pairing
            for (int k = 0; k < nch && bk < 0; ++k) if (!(assigned & (1u << k))) bk = k;
            for (int s = 0; s < 8 && bs < 0; ++s) if (!(used & (1u << s))) bs = s;
        }
        slot_child[bs] = bk;
        used |= 1u << bs;
        assigned |= 1u << bk;
    }

    // classify, allocate
    uint32_t imask = 0, n_inner = 0, n_leaf_tris = 0;
    for (int s = 0; s < 8; ++s) {
        const int k = slot_child[s];
        if (k < 0) continue;
        const uint32_t c = ch[k];
        const int cnt = __float_as_int(box_lo[c].w) & FF_COUNT_MASK;
        if (c < n_internal && cnt > 3) { imask |= 1u << s; ++n_inner; }
        else n_leaf_tris += (uint32_t)cnt;
    }
    // Allocation is aggregated over the lanes that arrive here together (one atomicAdd per counter and warp instead of one
    // per wide node): the three counters of a 20 M-triangle build took 8.4 M same-address atomics, which the L2 serialises.
    uint32_t child_base = 0, tri_base = 0;
    bool fits = true;
    {
        cg::coalesced_group g = cg::coalesced_threads();
        const uint32_t pre_n = cg::exclusive_scan(g, n_inner), pre_t = cg::exclusive_scan(g, n_leaf_tris);
        const unsigned last = g.size() - 1;
        const uint32_t tot_n = g.shfl(pre_n + n_inner, last), tot_t = g.shfl(pre_t + n_leaf_tris, last);
        uint32_t bn = 0, bt = 0;
        if (g.thread_rank() == 0) {
            if (tot_n) bn = atomicAdd(&ctr->node_count, tot_n);
            if (tot_t) bt = atomicAdd(&ctr->tri_count, tot_t);
        }
        bn = g.shfl(bn, 0); bt = g.shfl(bt, 0);
        if (n_inner) child_base = bn + pre_n;
        if (n_leaf_tris) tri_base = bt + pre_t;
    }
    if (n_inner) {
        fits = child_base + n_inner <= node_capacity;
        if (!fits) atomicExch(&ctr->overflow, 1u);   // reported by bvh_build; the children are not queued, nothing is written out of bounds
    }

    // Quantisation frame: the node box pushed out by 2^-13 of its extent per side.  The traversal evaluates a plane as
    // u*A + B with A = 2^15 * scale / d (byte_unit in ff_trace.cuh); the rounding of B moves a plane by at most 2^-16 of
    // the extent, and floor/ceil against the wider frame keeps every child box inside its quantised box by 8x that.
    {
        const float gx = (phi.x - plo.x) * 0x1p-13f, gy = (phi.y - plo.y) * 0x1p-13f, gz = (phi.z - plo.z) * 0x1p-13f;
        plo.x -= gx; plo.y -= gy; plo.z -= gz;
        phi.x += gx; phi.y += gy; phi.z += gz;
    }
    const uint32_t ebx = exponent_for(phi.x - plo.x), eby = exponent_for(phi.y - plo.y), ebz = exponent_for(phi.z - plo.z);
    const float sx = __uint_as_float(ebx << 23), sy = __uint_as_float(eby << 23), sz = __uint_as_float(ebz << 23);
    uint32_t meta[2] = {0, 0}, qlx[2] = {0, 0}, qly[2] = {0, 0}, qlz[2] = {0, 0}, qhx[2] = {0, 0}, qhy[2] = {0, 0}, qhz[2] = {0, 0};
    uint32_t inner_rank = 0, tri_off = 0;
    double sah = 0.0;
    for (int s = 0; s < 8; ++s) {
        const int k = slot_child[s];
        if (k < 0) continue;
        const uint32_t c = ch[k];
        const float4 l = box_lo[c], h = box_hi[c];
        const int cnt = __float_as_int(l.w) & FF_COUNT_MASK;
        uint32_t m;
        if (imask & (1u << s)) {
            m = 0x20u | (24u + (uint32_t)s);
            if (fits) node_src[child_base + inner_rank] = c;
            ++inner_rank;
        } else {
            const uint32_t first = (uint32_t)__float_as_int(h.w);
            m = (((1u << cnt) - 1u) << 5) | tri_off;
            // only the destination is recorded here; gather_tris_kernel moves the 48-byte records afterwards, one thread
            // per triangle (the copies used to be 72 dependent-latency loads and stores in this one thread)
            for (int t = 0; t < cnt; ++t) leaf_slot[first + t] = tri_base + tri_off + (uint32_t)t;
            tri_off += (uint32_t)cnt;
            sah += (double)box_area(l, h) * cnt;
        }
        const int w = s >> 2, sh = 8 * (s & 3);
        meta[w] |= m << sh;
        const float gx = (h.x - l.x) * 0x1p-13f + (phi.x - plo.x) * 0x1p-15f, gy = (h.y - l.y) * 0x1p-13f + (phi.y - plo.y) * 0x1p-15f,
                    gz = (h.z - l.z) * 0x1p-13f + (phi.z - plo.z) * 0x1p-15f;
        qlx[w] |= quant_lo(l.x - gx, plo.x, sx) << sh; qly[w] |= quant_lo(l.y - gy, plo.y, sy) << sh; qlz[w] |= quant_lo(l.z - gz, plo.z, sz) << sh;
        qhx[w] |= quant_hi(h.x + gx, plo.x, sx) << sh; qhy[w] |= quant_hi(h.y + gy, plo.y, sy) << sh; qhz[w] |= quant_hi(h.z + gz, plo.z, sz) << sh;
    }
    float4* out = nodes + FF_NODE_F4 * (size_t)wnode;
    const uint32_t xb = FF_NODE_EXP_BIAS;  // the node stores the exponents of 2^15 * scale
    out[0] = make_float4(plo.x, plo.y, plo.z, __uint_as_float((ebx + xb) | ((eby + xb) << 8) | ((ebz + xb) << 16) | (imask << 24)));
    out[1] = make_float4(__uint_as_float(child_base), __uint_as_float(tri_base), __uint_as_float(meta[0]), __uint_as_float(meta[1]));
#if FF_NODE_256
    out[2] = make_float4(__uint_as_float(qlx[0]), __uint_as_float(qlx[1]), __uint_as_float(qly[0]), __uint_as_float(qly[1]));
    out[3] = make_float4(__uint_as_float(qlz[0]), __uint_as_float(qlz[1]), 0.f, 0.f);
    out[4] = make_float4(__uint_as_float(qhx[0]), __uint_as_float(qhx[1]), __uint_as_float(qhy[0]), __uint_as_float(qhy[1]));
    out[5] = make_float4(__uint_as_float(qhz[0]), __uint_as_float(qhz[1]), 0.f, 0.f);
#else
    out[2] = make_float4(__uint_as_float(qlx[0]), __uint_as_float(qlx[1]), __uint_as_float(qly[0]), __uint_as_float(qly[1]));
    out[3] = make_float4(__uint_as_float(qlz[0]), __uint_as_float(qlz[1]), __uint_as_float(qhx[0]), __uint_as_float(qhx[1]));
    out[4] = make_float4(__uint_as_float(qhy[0]), __uint_as_float(qhy[1]), __uint_as_float(qhz[0]), __uint_as_float(qhz[1]));
#endif
    // SAH cost of the wide tree: C_node * area(node) + C_tri * sum(leaf area * count), normalised by the root area
    if (root_area > 0.f) {
        cg::coalesced_group g = cg::coalesced_threads();
        const double sum = cg::reduce(g, ((double)box_area(plo, phi) * 1.0 + sah * 1.0) / (double)root_area, cg::plus<double>());
        if (g.thread_rank() == 0) atomicAdd(&ctr->sah, sum);
    }
}
__global__ void __launch_bounds__(128) collapse_kernel(uint32_t* __restrict__ node_src, uint32_t node_capacity, int n,
                                const int2* __restrict__ children, const float4* __restrict__ box_lo, const float4* __restrict__ box_hi,
                                const float4* __restrict__ flat_tris, const uint32_t* __restrict__ sorted_idx,
                                float4* __restrict__ nodes, float4* __restrict__ tris_out, uint32_t* __restrict__ leaf_slot,
                                CollapseCounters* ctr, float root_area, const float4* __restrict__ dp) {
    const uint32_t begin = ctr->level_begin, end = ctr->level_end;   // constant during the level
    for (uint32_t w = begin + blockIdx.x * blockDim.x + threadIdx.x; w < end; w += gridDim.x * blockDim.x)
        collapse_one(node_src[w], w, node_src, node_capacity, n, children, box_lo, box_hi, flat_tris, sorted_idx, nodes, tris_out, leaf_slot,
                     ctr, root_area, dp);
}
// Triangles into BVH-leaf order: sorted position p -> slot leaf_slot[p] (written by the collapse).
__global__ void gather_tris_kernel(const float4* __restrict__ flat_tris, const uint32_t* __restrict__ sorted_idx,
                                   const uint32_t* __restrict__ leaf_slot, uint32_t n, float4* __restrict__ tris_out) {
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
        const size_t src = sorted_idx[p], slot = leaf_slot[p];
        const float4 a = __ldg(flat_tris + 3 * src), b = __ldg(flat_tris + 3 * src + 1), c = __ldg(flat_tris + 3 * src + 2);
        tris_out[3 * slot] = a; tris_out[3 * slot + 1] = b; tris_out[3 * slot + 2] = c;
    }
}
__global__ void collapse_advance_kernel(CollapseCounters* ctr, uint32_t node_capacity) {
    const uint32_t end = ctr->level_end, allocated = min(ctr->node_count, node_capacity);
    if (end > ctr->level_begin) ctr->levels += 1;
    ctr->level_begin = end;
    ctr->level_end = ctr->overflow ? end : allocated;
}

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return e_; } while (0)

cudaError_t sah_build_hierarchy(const float4* d_flat_tris, const uint32_t* d_sorted_idx, uint32_t n, const float centre_lo[3],
                                const float centre_hi[3], int2* d_children, uint32_t* d_parent, uint32_t* d_order_out,
                                uint32_t* launches, uint32_t* out_levels, Arena& arena, cudaStream_t st);  // ff_sah.cu
size_t sah_scratch_bytes(uint32_t n);

cudaError_t bvh_build(const float4* d_flat_tris, uint32_t n, int builder, float pad, float4* d_tris_out, float4* d_nodes_out,
                      uint32_t node_capacity, int2* d_bvh2_children, float4* d_bvh2_lo, float4* d_bvh2_hi,
                      uint32_t* d_bvh2_leaf_slot, float bounds_lo[3], float bounds_hi[3], BuildOut* out, uint32_t* launches,
                      Arena& arena, cudaStream_t st) {
    out->n_nodes = 0; out->depth = 0; out->sah_cost = 0;
    for (int k = 0; k < 3; ++k) { bounds_lo[k] = 0; bounds_hi[k] = 0; }
    if (n == 0) return cudaSuccess;
    // FF_BUILD_TRACE=1: per-stage wall times on stderr (synchronises after every stage; development aid)
    const bool trace = getenv("FF_BUILD_TRACE") != nullptr;
    auto t_last = std::chrono::steady_clock::now();
    auto mark = [&](const char* what) {
        if (!trace) return;
        cudaStreamSynchronize(st);
        const auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "[ff build] %-28s %9.3f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
        t_last = now;
    };
    const uint32_t n_tiles = (n + RS_TILE - 1) / RS_TILE;
    int* d_bounds = nullptr;
    uint32_t *d_keys_a = nullptr, *d_keys_b = nullptr, *d_vals_a = nullptr, *d_vals_b = nullptr, *d_hist = nullptr;
    uint32_t *d_parent = nullptr, *d_flags = nullptr;
    float4* d_dp = nullptr;
    uint32_t* d_src = nullptr;
    CollapseCounters* d_ctr = nullptr;
    {
        size_t need = Arena::padded(12 * sizeof(int)) + 4 * Arena::padded(sizeof(uint32_t) * n) + Arena::padded(sizeof(uint32_t) * 256 * ((size_t)n_tiles + 1)) +
                      Arena::padded(sizeof(uint32_t) * 2 * (size_t)n) + Arena::padded(sizeof(uint32_t) * n) + Arena::padded(sizeof(float4) * 2 * (size_t)n) +
                      Arena::padded(sizeof(uint32_t) * (size_t)node_capacity) + Arena::padded(sizeof(CollapseCounters));
        if (builder == FF_BUILDER_SAH) need += sah_scratch_bytes(n);
        CK(arena.reserve(need));
    }
    d_bounds = arena.take<int>(12);
    d_keys_a = arena.take<uint32_t>(n); d_keys_b = arena.take<uint32_t>(n);
    d_vals_a = arena.take<uint32_t>(n); d_vals_b = arena.take<uint32_t>(n);
    d_hist = arena.take<uint32_t>(256 * ((size_t)n_tiles + 1));   // [256][n_tiles] counters + 256 row totals
    d_parent = arena.take<uint32_t>(2 * (size_t)n);
    d_flags = arena.take<uint32_t>(n);
    d_dp = arena.take<float4>(2 * (size_t)n);   // F table + decisions of the n - 1 internal binary nodes
    d_src = arena.take<uint32_t>(node_capacity);
    d_ctr = arena.take<CollapseCounters>(1);
    if (!d_ctr) return cudaErrorMemoryAllocation;
    CK(cudaMemsetAsync(d_flags, 0, sizeof(uint32_t) * n, st));
    mark("reserve scratch");

    bounds_init_kernel<<<1, 32, 0, st>>>(d_bounds);
    bounds_kernel<<<grid_for(n, 256, sm_count() * 8), 256, 0, st>>>(d_flat_tris, n, d_bounds);
    morton_kernel<<<grid_for(n, 256, sm_count() * 16), 256, 0, st>>>(d_flat_tris, n, d_bounds, d_keys_a, d_vals_a);
    *launches += 3;
    mark("bounds + morton");
    radix_sort_pairs(d_keys_a, d_vals_a, d_keys_b, d_vals_b, n, d_hist, launches, st);
    mark("radix sort");
    int h_bounds[12];
    CK(cudaMemcpyAsync(h_bounds, d_bounds, sizeof(h_bounds), cudaMemcpyDeviceToHost, st));
    uint32_t* d_order = d_vals_a;  // position in the binary tree's leaf order -> flat triangle index
    if (n > 1 && builder == FF_BUILDER_SAH) {
        // top-down binned SAH over the Morton-sorted triangles (ff_sah.cu); leaves the order in d_vals_b
        CK(cudaStreamSynchronize(st));
        float clo[3], chi[3];
        for (int k = 0; k < 3; ++k) { clo[k] = ordered_to_float(h_bounds[6 + k]); chi[k] = ordered_to_float(h_bounds[9 + k]); }
        CK(sah_build_hierarchy(d_flat_tris, d_vals_a, n, clo, chi, d_bvh2_children, d_parent, d_vals_b, launches, nullptr, arena, st));
        d_order = d_vals_b;
        mark("binned-SAH hierarchy");
    } else if (n > 1) {
        hierarchy_kernel<<<grid_for(n - 1, 256, sm_count() * 16), 256, 0, st>>>(d_keys_a, (int)n, d_bvh2_children, d_parent);
        ++*launches;
        mark("karras hierarchy");
    }
    const char* sc_env = getenv("FF_SLOT_COST");  // development knob: leaf-or-open trade-off of the refit DP
    const float slot_cost = sc_env ? (float)atof(sc_env) : FF_SLOT_COST;
    const char* nc_env = getenv("FF_NODE_COST");
    const float node_cost = nc_env ? (float)atof(nc_env) : FF_NODE_COST;
    refit_kernel<<<grid_for(n, 256, sm_count() * 16), 256, 0, st>>>(d_flat_tris, d_order, (int)n, pad, d_bvh2_children, d_parent, d_flags,
                                                            d_bvh2_lo, d_bvh2_hi, d_dp, slot_cost, node_cost);
    ++*launches;
    mark("refit + collapse DP");

    float4 root_lo, root_hi;
    const uint32_t root = (n == 1) ? 0u : 0u;  // internal node 0, or the single leaf (id n-1+0 = 0)
    CK(cudaMemcpyAsync(&root_lo, d_bvh2_lo + root, sizeof(float4), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(&root_hi, d_bvh2_hi + root, sizeof(float4), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    for (int k = 0; k < 3; ++k) { bounds_lo[k] = ordered_to_float(h_bounds[k]); bounds_hi[k] = ordered_to_float(h_bounds[3 + k]); }
    const float dx = root_hi.x - root_lo.x, dy = root_hi.y - root_lo.y, dz = root_hi.z - root_lo.z;
    const float root_area = 2.0f * (dx * dy + dy * dz + dz * dx);

    CollapseCounters h_ctr;
    memset(&h_ctr, 0, sizeof(h_ctr));
    h_ctr.node_count = 1; h_ctr.level_begin = 0; h_ctr.level_end = 1;
    CK(cudaMemcpyAsync(d_ctr, &h_ctr, sizeof(h_ctr), cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(d_src, &root, sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    // Levels in flights of 12 without host round trips (a hierarchy of 20 M triangles is 10 levels deep): fixed grid,
    // grid-stride over the level's index range read from the device; an exhausted hierarchy costs empty launches.
    for (int flight = 0; flight < 8; ++flight) {
        for (int l = 0; l < 12; ++l) {
            collapse_kernel<<<sm_count() * 8, 128, 0, st>>>(d_src, node_capacity, (int)n, d_bvh2_children, d_bvh2_lo, d_bvh2_hi, d_flat_tris, d_order,
                                                    d_nodes_out, d_tris_out, d_bvh2_leaf_slot, d_ctr, root_area, d_dp);
            collapse_advance_kernel<<<1, 1, 0, st>>>(d_ctr, node_capacity);
            *launches += 2;
        }
        CK(cudaMemcpyAsync(&h_ctr, d_ctr, sizeof(h_ctr), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        if (h_ctr.overflow) { CK(cudaErrorMemoryAllocation); }
        if (h_ctr.level_begin == h_ctr.level_end) break;
    }
    if (h_ctr.level_begin != h_ctr.level_end) return cudaErrorUnknown;  // deeper than 96 levels: not a hierarchy this builder makes
    if (h_ctr.tri_count == n) {
        gather_tris_kernel<<<grid_for(n, 256, sm_count() * 16), 256, 0, st>>>(d_flat_tris, d_order, d_bvh2_leaf_slot, n, d_tris_out);
        ++*launches;
    }
    const uint32_t depth = h_ctr.levels;
    mark("collapse to 8-wide");
    out->n_nodes = h_ctr.node_count;
    out->depth = depth;
    out->sah_cost = h_ctr.sah;
    return h_ctr.tri_count == n ? cudaSuccess : cudaErrorUnknown;
}

}  // namespace ff
