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
#include <algorithm>
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
// Adds one triangle to a thread's running bounds (scene box and box of the leaf-box centres).
__device__ __forceinline__ void bounds_add(const float av[3], const float bv[3], const float cv[3], float lo[3], float hi[3], float clo[3], float chi[3]) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float l = fminf(fminf(av[k], bv[k]), cv[k]), h = fmaxf(fmaxf(av[k], bv[k]), cv[k]);
        const float ce = 0.5f * (l + h);
        lo[k] = fminf(lo[k], l); hi[k] = fmaxf(hi[k], h);
        clo[k] = fminf(clo[k], ce); chi[k] = fmaxf(chi[k], ce);
    }
}
// Every thread of the CTA calls this once with its running bounds (blockDim.x <= 256).
__device__ __forceinline__ void bounds_block_reduce(float lo[3], float hi[3], float clo[3], float chi[3], int* bounds) {
    // warp shuffle -> shared memory -> warp 0: 12 atomics per CTA.  (One set per WARP was 375 k same-address atomics at 1 M
    // triangles, which the L2 serialises: 80 of the build's 930 us, profiles/r02a_launches_bench_summary.json.)
    __shared__ float s_red[12][8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        for (int off = 16; off; off >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], off));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], off));
            clo[k] = fminf(clo[k], __shfl_xor_sync(0xffffffffu, clo[k], off));
            chi[k] = fmaxf(chi[k], __shfl_xor_sync(0xffffffffu, chi[k], off));
        }
        if (lane == 0) { s_red[k][warp] = lo[k]; s_red[3 + k][warp] = hi[k]; s_red[6 + k][warp] = clo[k]; s_red[9 + k][warp] = chi[k]; }
    }
    __syncthreads();
    if (threadIdx.x < 12) {
        const int q = threadIdx.x;
        const bool is_max = ((q / 3) & 1) != 0;
        float v = s_red[q][0];
        for (int w = 1; w < (int)(blockDim.x >> 5); ++w) v = is_max ? fmaxf(v, s_red[q][w]) : fminf(v, s_red[q][w]);
        if (is_max) atomicMax(bounds + q, float_to_ordered(v)); else atomicMin(bounds + q, float_to_ordered(v));
    }
}
__global__ void bounds_kernel(const float4* __restrict__ tris, uint32_t n, int* bounds) {
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    float clo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, chi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 a = tris[3 * (size_t)i], b = tris[3 * (size_t)i + 1], c = tris[3 * (size_t)i + 2];
        const float av[3] = {a.x, a.y, a.z}, bv[3] = {b.x, b.y, b.z}, cv[3] = {c.x, c.y, c.z};
        bounds_add(av, bv, cv, lo, hi, clo, chi);
    }
    bounds_block_reduce(lo, hi, clo, chi, bounds);
}

// `bounds` (optional): the scene / centroid bounds of bounds_kernel, reduced here while the baked vertices are in registers,
// so the build that follows does not have to read all triangles once more just for them (146 us at 20 M triangles).
__global__ void bake_kernel(const float* __restrict__ xyz9, const M4* __restrict__ models, const uint32_t* __restrict__ geom_of,
                            const uint32_t* __restrict__ ord_of, uint32_t n, float4* __restrict__ out, int* bounds) {
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    float clo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, chi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t g = geom_of[i];
        const M4 m = models[g];
        const float* p = xyz9 + 9 * (size_t)i;
        const V3 a = xform_point(m, p[0], p[1], p[2]), b = xform_point(m, p[3], p[4], p[5]), c = xform_point(m, p[6], p[7], p[8]);
        out[3 * (size_t)i] = make_float4(a.x, a.y, a.z, __int_as_float((int)ord_of[i]));
        out[3 * (size_t)i + 1] = make_float4(b.x, b.y, b.z, __int_as_float((int)g));
        out[3 * (size_t)i + 2] = make_float4(c.x, c.y, c.z, __int_as_float((int)i));
        const float av[3] = {a.x, a.y, a.z}, bv[3] = {b.x, b.y, b.z}, cv[3] = {c.x, c.y, c.z};
        bounds_add(av, bv, cv, lo, hi, clo, chi);
    }
    if (bounds) bounds_block_reduce(lo, hi, clo, chi, bounds);
}
void k_bake_triangles(const float* d_xyz9_obj, const M4* d_model_per_geom, const uint32_t* d_geom_of_flat,
                      const uint32_t* d_ord_of_flat, uint32_t n, float4* d_flat_tris, int* d_bounds, cudaStream_t st) {
    if (d_bounds) bounds_init_kernel<<<1, 32, 0, st>>>(d_bounds);
    if (n) bake_kernel<<<grid_for(n, 256, sm_count() * 16), 256, 0, st>>>(d_xyz9_obj, d_model_per_geom, d_geom_of_flat, d_ord_of_flat, n, d_flat_tris, d_bounds);
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

// ------------------------------------------------------------------ Morton codes of the box centres inside the centroid bounds
// 30 bits (10 per axis, u32) for small scenes, 63 bits (21 per axis, u64) above 2^20 triangles: on a 1024^3 grid a large or
// clustered scene puts many triangles into one cell, and inside a cell the hierarchy can only fall back to the input order.
__device__ __forceinline__ uint32_t expand10(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
__device__ __forceinline__ uint64_t expand21(uint32_t a) {
    uint64_t x = a & 0x1fffffu;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}
template <class KeyT> struct Morton;
template <> struct Morton<uint32_t> {
    static constexpr float CELLS = 1024.0f;
    static constexpr uint32_t MAXQ = 1023u;
    static __device__ __forceinline__ uint32_t code(uint32_t x, uint32_t y, uint32_t z) { return (expand10(x) << 2) | (expand10(y) << 1) | expand10(z); }
};
template <> struct Morton<uint64_t> {
    static constexpr float CELLS = 2097152.0f;
    static constexpr uint32_t MAXQ = 2097151u;
    static __device__ __forceinline__ uint64_t code(uint32_t x, uint32_t y, uint32_t z) { return (expand21(x) << 2) | (expand21(y) << 1) | expand21(z); }
};
template <class KeyT>
__global__ void morton_kernel(const float4* __restrict__ tris, uint32_t n, const int* __restrict__ bounds, KeyT* keys, uint32_t* vals) {
    float clo[3], inv[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        clo[k] = ordered_to_float(bounds[6 + k]);
        const float ext = ordered_to_float(bounds[9 + k]) - clo[k];
        inv[k] = ext > 0.f ? Morton<KeyT>::CELLS / ext : 0.f;
    }
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 a = tris[3 * (size_t)i], b = tris[3 * (size_t)i + 1], c = tris[3 * (size_t)i + 2];
        const float cx = 0.5f * (fminf(fminf(a.x, b.x), c.x) + fmaxf(fmaxf(a.x, b.x), c.x));
        const float cy = 0.5f * (fminf(fminf(a.y, b.y), c.y) + fmaxf(fmaxf(a.y, b.y), c.y));
        const float cz = 0.5f * (fminf(fminf(a.z, b.z), c.z) + fmaxf(fmaxf(a.z, b.z), c.z));
        const uint32_t x = min(Morton<KeyT>::MAXQ, (uint32_t)max(0.f, (cx - clo[0]) * inv[0]));
        const uint32_t y = min(Morton<KeyT>::MAXQ, (uint32_t)max(0.f, (cy - clo[1]) * inv[1]));
        const uint32_t z = min(Morton<KeyT>::MAXQ, (uint32_t)max(0.f, (cz - clo[2]) * inv[2]));
        keys[i] = Morton<KeyT>::code(x, y, z);
        vals[i] = i;
    }
}

// how many sorted keys repeat their predecessor: the measure of whether the 30-bit grid resolves the scene
__global__ void count_duplicate_keys_kernel(const uint32_t* __restrict__ keys, uint32_t n, uint32_t* __restrict__ out) {
    uint32_t c = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x + 1u; i < n; i += gridDim.x * blockDim.x) c += keys[i] == keys[i - 1] ? 1u : 0u;
    for (int off = 16; off; off >>= 1) c += __shfl_xor_sync(0xffffffffu, c, off);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}

// ------------------------------------------------------------------ LSD radix sort, 8-bit digits, ONE sweep per pass
// Keys are 32-bit (30-bit Morton codes, 4 passes) or 64-bit (63-bit codes, 8 passes); values are the 32-bit triangle indices.
//   os_hist_kernel   one read of the keys: the digit histograms of ALL passes (shared-memory atomics), added to global counters
//   os_scan_kernel   exclusive scan of each pass's 256 counters
//   os_pass_kernel   one launch per pass: a CTA takes the next tile (4096 keys) from a counter, ranks its keys by digit
//                    (match_any inside each warp, prefix over the warps), publishes the tile's 256 digit counts, obtains the
//                    counts of all earlier tiles by decoupled look-back (Merrill & Garland 2016: a tile publishes its local
//                    counts, then its inclusive prefix, one flagged 32-bit word per digit; a successor sums local counts
//                    backwards until it meets an inclusive one), reorders keys and values through shared memory and writes
//                    them in digit runs.  One read and one write of the pairs per pass; the three-kernel pass it replaces
//                    (histogram, two scans, scatter) read the keys twice and ran at 1.07 TB/s.
#define OS_THREADS 256
#define OS_ITEMS 16
#define OS_TILE (OS_THREADS * OS_ITEMS)
#define OS_FLAG_LOCAL 0x40000000u
#define OS_FLAG_PREFIX 0x80000000u
#define OS_VALUE_MASK 0x3fffffffu
template <class KeyT> struct KeyTraits;
template <> struct KeyTraits<uint32_t> { static constexpr int PASSES = 4; };
template <> struct KeyTraits<uint64_t> { static constexpr int PASSES = 8; };

template <class KeyT>
__global__ void __launch_bounds__(256) os_hist_kernel(const KeyT* __restrict__ keys, uint32_t n, uint32_t* __restrict__ ghist /*[PASSES][256]*/) {
    constexpr int P = KeyTraits<KeyT>::PASSES;
    __shared__ uint32_t h[P][256];
    for (int k = threadIdx.x; k < P * 256; k += blockDim.x) (&h[0][0])[k] = 0u;
    __syncthreads();
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const KeyT k = keys[i];
#pragma unroll
        for (int p = 0; p < P; ++p) atomicAdd(&h[p][(uint32_t)(k >> (8 * p)) & 0xffu], 1u);
    }
    __syncthreads();
    for (int k = threadIdx.x; k < P * 256; k += blockDim.x) {
        const uint32_t v = (&h[0][0])[k];
        if (v) atomicAdd(ghist + k, v);
    }
}
__global__ void __launch_bounds__(256) os_scan_kernel(uint32_t* ghist) {   // block p: exclusive scan of pass p's 256 counters
    __shared__ uint32_t s[256];
    uint32_t* g = ghist + 256 * blockIdx.x;
    const uint32_t v = g[threadIdx.x];
    s[threadIdx.x] = v;
    __syncthreads();
    for (int off = 1; off < 256; off <<= 1) {
        const uint32_t t = threadIdx.x >= (unsigned)off ? s[threadIdx.x - off] : 0u;
        __syncthreads();
        s[threadIdx.x] += t;
        __syncthreads();
    }
    g[threadIdx.x] = s[threadIdx.x] - v;
}
template <class KeyT>
__global__ void __launch_bounds__(OS_THREADS) os_pass_kernel(const KeyT* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                                                              KeyT* __restrict__ keys_out, uint32_t* __restrict__ vals_out, uint32_t n, int shift,
                                                              const uint32_t* __restrict__ gbase /*[256], exclusive*/, uint32_t* desc /*[tiles][256], zero*/,
                                                              uint32_t* tile_counter /*zero*/) {
    __shared__ uint32_t s_tile;
    __shared__ uint32_t warp_hist[OS_THREADS / 32][256];
    __shared__ uint32_t digit_off[256];    // first local position of the digit inside the reordered tile
    __shared__ uint32_t digit_dst[256];    // global position of that first element, minus digit_off
    __shared__ uint32_t warp_sums[OS_THREADS / 32];
    __shared__ KeyT s_buf[OS_TILE];        // keys, then (its first half, as u32) the values
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // tiles in launch order, not in blockIdx order: a tile only ever waits for tiles that already hold an SM
    if (threadIdx.x == 0) s_tile = atomicAdd(tile_counter, 1u);
    for (int k = threadIdx.x; k < (OS_THREADS / 32) * 256; k += OS_THREADS) (&warp_hist[0][0])[k] = 0u;
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t tile_base = tile * OS_TILE;
    const uint32_t tile_count = min((uint32_t)OS_TILE, n - tile_base);
    // warp w owns the contiguous chunk [w*512, (w+1)*512) of the tile, processed 32 keys per round, so ranks are stable
    const uint32_t base = tile_base + warp * (32 * OS_ITEMS);
    KeyT key[OS_ITEMS];
    uint32_t val[OS_ITEMS], rank[OS_ITEMS];
#pragma unroll
    for (int r = 0; r < OS_ITEMS; ++r) {
        const uint32_t i = base + r * 32 + lane;
        const bool live = i < n;
        key[r] = live ? keys_in[i] : (KeyT)0;
        val[r] = live ? vals_in[i] : 0u;
    }
#pragma unroll
    for (int r = 0; r < OS_ITEMS; ++r) {
        const bool live = base + r * 32 + lane < n;
        const uint32_t d = (uint32_t)(key[r] >> shift) & 0xffu;
        // lanes with the same digit, by eight ballots (one per digit bit) instead of MATCH.ANY: the match unit serialises on the
        // ~30 distinct digits of a warp, and the BREV that consumes its result was 48 % of this kernel's stall samples
        // (short scoreboard, profiles/r02w_build20m_details.txt)
        unsigned peers = __ballot_sync(0xffffffffu, live);
#pragma unroll
        for (int b = 0; b < 8; ++b) {
            const bool bit = (d >> b) & 1u;
            const unsigned m = __ballot_sync(0xffffffffu, bit);
            peers &= bit ? m : ~m;
        }
        if (!live) peers = 1u << lane;
        const int leader = __ffs(peers) - 1;
        uint32_t start = 0;
        if (live && lane == leader) { start = warp_hist[warp][d]; warp_hist[warp][d] = start + __popc(peers); }
        start = __shfl_sync(0xffffffffu, start, leader);
        rank[r] = start + __popc(peers & ((1u << lane) - 1u));
        __syncwarp();
    }
    __syncthreads();
    {   // thread d owns digit d: counts over the warps, the tile's count, its position inside the tile, the look-back
        const uint32_t d = threadIdx.x;
        uint32_t count = 0;
#pragma unroll
        for (int w = 0; w < OS_THREADS / 32; ++w) { const uint32_t c = warp_hist[w][d]; warp_hist[w][d] = count; count += c; }
        uint32_t* my = desc + (size_t)tile * 256 + d;
        if (tile == 0) atomicExch(my, count | OS_FLAG_PREFIX);
        else atomicExch(my, count | OS_FLAG_LOCAL);
        // exclusive scan of `count` over the 256 digits: where each digit starts in the reordered tile
        uint32_t incl = count;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, off);
            if (lane >= off) incl += t;
        }
        if (lane == 31) warp_sums[warp] = incl;
        __syncthreads();
        uint32_t wbase = 0;
#pragma unroll
        for (int w = 0; w < OS_THREADS / 32; ++w) wbase += w < warp ? warp_sums[w] : 0u;
        const uint32_t off_d = wbase + incl - count;
        digit_off[d] = off_d;
        uint32_t prefix = 0;
        if (tile != 0) {
            // four predecessors per round trip (the loads are independent; a tile of the first wave walks back over hundreds)
            for (int t = (int)tile - 1; t >= 0;) {
                uint32_t v[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) v[j] = t - j >= 0 ? *(volatile uint32_t*)(desc + (size_t)(t - j) * 256 + d) : OS_FLAG_PREFIX;
                int used = 0;
                bool done = false;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    if (done || used != j) continue;
                    if ((v[j] & (OS_FLAG_LOCAL | OS_FLAG_PREFIX)) == 0u) continue;   // not published yet: retry from here
                    prefix += v[j] & OS_VALUE_MASK;
                    ++used;
                    if (v[j] & OS_FLAG_PREFIX) done = true;
                }
                if (done) break;
                if (used == 0) __nanosleep(20);
                t -= used;
            }
            atomicExch(my, (prefix + count) | OS_FLAG_PREFIX);
        }
        digit_dst[d] = __ldg(gbase + d) + prefix - off_d;
    }
    __syncthreads();
    // reorder through shared memory so that a digit's keys leave as one contiguous run
    uint32_t lp[OS_ITEMS];
#pragma unroll
    for (int r = 0; r < OS_ITEMS; ++r) {
        const uint32_t d = (uint32_t)(key[r] >> shift) & 0xffu;
        lp[r] = digit_off[d] + warp_hist[warp][d] + rank[r];
        if (base + r * 32 + lane < n) s_buf[lp[r]] = key[r];
    }
    __syncthreads();
    uint32_t dst[OS_ITEMS];
#pragma unroll
    for (int j = 0; j < OS_ITEMS; ++j) {
        const uint32_t idx = j * OS_THREADS + threadIdx.x;
        dst[j] = 0xffffffffu;
        if (idx < tile_count) {
            const KeyT k = s_buf[idx];
            dst[j] = digit_dst[(uint32_t)(k >> shift) & 0xffu] + idx;
            keys_out[dst[j]] = k;
        }
    }
    __syncthreads();
    uint32_t* s_val = reinterpret_cast<uint32_t*>(s_buf);
#pragma unroll
    for (int r = 0; r < OS_ITEMS; ++r)
        if (base + r * 32 + lane < n) s_val[lp[r]] = val[r];
    __syncthreads();
#pragma unroll
    for (int j = 0; j < OS_ITEMS; ++j)
        if (dst[j] != 0xffffffffu) vals_out[dst[j]] = s_val[j * OS_THREADS + threadIdx.x];
}
// Scratch of one sort: [PASSES][256] digit bases | [PASSES] tile counters | [PASSES][tiles][256] tile descriptors, zeroed by one memset.
template <class KeyT>
static size_t sort_scratch_words(uint32_t n) {
    const size_t tiles = (n + OS_TILE - 1) / OS_TILE;
    return (size_t)KeyTraits<KeyT>::PASSES * (256 + 64 + tiles * 256);
}
// Sorts (keys, vals) by the whole key; the result ends in keys_a / vals_a (the number of passes is even).
template <class KeyT>
static cudaError_t radix_sort_pairs(KeyT* keys_a, uint32_t* vals_a, KeyT* keys_b, uint32_t* vals_b, uint32_t n, uint32_t* scratch,
                                    uint32_t* launches, cudaStream_t st) {
    constexpr int P = KeyTraits<KeyT>::PASSES;
    const uint32_t tiles = (n + OS_TILE - 1) / OS_TILE;
    uint32_t* ghist = scratch;
    uint32_t* counters = scratch + (size_t)P * 256;
    uint32_t* desc = counters + (size_t)P * 64;
    cudaError_t e = cudaMemsetAsync(scratch, 0, sizeof(uint32_t) * sort_scratch_words<KeyT>(n), st);
    if (e != cudaSuccess) return e;
    os_hist_kernel<KeyT><<<std::min<uint32_t>((n + 255) / 256, (uint32_t)sm_count() * 8), 256, 0, st>>>(keys_a, n, ghist);
    os_scan_kernel<<<P, 256, 0, st>>>(ghist);
    *launches += 2;
    for (int pass = 0; pass < P; ++pass) {
        const KeyT* ki = (pass & 1) ? keys_b : keys_a;
        const uint32_t* vi = (pass & 1) ? vals_b : vals_a;
        KeyT* ko = (pass & 1) ? keys_a : keys_b;
        uint32_t* vo = (pass & 1) ? vals_a : vals_b;
        os_pass_kernel<KeyT><<<tiles, OS_THREADS, 0, st>>>(ki, vi, ko, vo, n, 8 * pass, ghist + 256 * pass, desc + (size_t)pass * tiles * 256, counters + 64 * pass);
        ++*launches;
    }
    return cudaGetLastError();
}

// ------------------------------------------------------------------ Karras 2012 hierarchy
__device__ __forceinline__ int key_clz(uint32_t x) { return __clz(x); }
__device__ __forceinline__ int key_clz(uint64_t x) { return __clzll((long long)x); }
// length of the common prefix of keys i and j; equal keys are told apart by their position (longer than any key prefix)
template <class KeyT>
__device__ __forceinline__ int delta(const KeyT* __restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const KeyT a = keys[i], b = keys[j];
    if (a != b) return key_clz(a ^ b);
    return 8 * (int)sizeof(KeyT) + __clz((uint32_t)i ^ (uint32_t)j);
}
// node ids: internal i in [0, n-1), leaf k -> (n-1) + k
// parent[] entries carry FF_PARENT_CROSS in bit 31 when the parent's leaf range [lo, hi] is not contained in one block of
// FF_TREELET consecutive sorted leaves: that is where the shared-memory phase of refit_treelet_kernel hands a subtree over to
// the global hand-off protocol.  (An internal node's index always lies inside its own range, so a contained node also has
// its index inside the block.)
// (FF_TREELET, FF_PARENT_CROSS, FF_PARENT_ROOT: ff_internal.h -- the binned-SAH builder tags its links the same way.)
template <class KeyT>
__global__ void hierarchy_kernel(const KeyT* __restrict__ keys, int n, int2* __restrict__ children, uint32_t* __restrict__ parent) {
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
        const uint32_t tag = (uint32_t)i | ((lo / FF_TREELET) != (hi / FF_TREELET) ? FF_PARENT_CROSS : 0u);
        parent[left] = tag;
        parent[right] = tag;
        if (i == 0) parent[0] = FF_PARENT_ROOT;
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
// One step of the bottom-up pass: the box, triangle count and F table of a binary node from those of its two children.
// fl / fr = F(left, 1..7), F(right, 1..7) (index 0 unused); a child that is a single triangle has the constant table
// area * (slot + 1) (leaf_table).  Results: lo (w = count), hi (w = first sorted position, the left child's), f[1..7], dec.
__device__ __forceinline__ void leaf_table(const float4& l, const float4& h, float slot_cost, float f[8]) {
    const float f1 = box_area_lohi(l, h) * (slot_cost + 1.0f);
#pragma unroll
    for (int q = 1; q <= 7; ++q) f[q] = f1;
}
__device__ __forceinline__ void dp_combine(const float4& l0, const float4& h0, const float fl[8], const float4& l1, const float4& h1, const float fr[8],
                                           float slot_cost, float node_cost, float4& lo, float4& hi, float f[8], uint32_t& dec) {
    const int cnt = (__float_as_int(l0.w) & FF_COUNT_MASK) + (__float_as_int(l1.w) & FF_COUNT_MASK);
    lo = make_float4(fminf(l0.x, l1.x), fminf(l0.y, l1.y), fminf(l0.z, l1.z), __int_as_float(cnt));
    hi = make_float4(fmaxf(h0.x, h1.x), fmaxf(h0.y, h1.y), fmaxf(h0.z, h1.z), h0.w);
    const float area = box_area_lohi(lo, hi);
    dec = 0;
    {
        float g = FLT_MAX;
        uint32_t k8 = 1;
#pragma unroll
        for (int q = 1; q <= 7; ++q) { const float v = fl[q] + fr[8 - q]; if (v < g) { g = v; k8 = (uint32_t)q; } }
        dec |= k8 << 21;
        f[1] = (cnt <= FF_LEAF_MAX) ? area * (slot_cost + (float)cnt) : area * (slot_cost + node_cost) + g;
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
}
__device__ __forceinline__ void unpack_table(const float4& d0, const float4& d1, float f[8]) {
    f[1] = d0.x; f[2] = d0.y; f[3] = d0.z; f[4] = d0.w; f[5] = d1.x; f[6] = d1.y; f[7] = d1.z;
}
// Climbs from `node` (whose box / table are already in global memory) towards the root with the global hand-off: the first
// thread to arrive at a parent stops, the second finishes it.
// acq_rel: RELEASE so this thread's box and table are visible before its arrival is counted; ACQUIRE so the second arrival's
// loads of its sibling's box and table are ordered after the atomic by the PTX memory model itself (a compiler barrier plus
// a control dependency is not an ordering there).  One atomic instead of the fence.sc + L1 invalidate that __threadfence()
// cost twice per level (39 % of the stall samples, profiles/r01i_refit_before_stalls.txt); the loads bypass L1 (ld.global.cg).
__device__ __forceinline__ void refit_climb_global(uint32_t node, int n, const int2* __restrict__ children, const uint32_t* __restrict__ parent,
                                                   uint32_t* __restrict__ flags, float4* __restrict__ box_lo, float4* __restrict__ box_hi,
                                                   float4* __restrict__ dp, float slot_cost, float node_cost) {
    for (;;) {
        const uint32_t pp = parent[node];
        if (pp == FF_PARENT_ROOT) break;
        const uint32_t p = pp & ~FF_PARENT_CROSS;
        uint32_t arrived;
        asm volatile("atom.global.add.acq_rel.gpu.u32 %0, [%1], 1;" : "=r"(arrived) : "l"(__cvta_generic_to_global(flags + p)) : "memory");
        if (arrived == 0u) break;  // first arrival: the sibling's thread finishes this node
        const int2 ch = children[p];
        const float4 l0 = __ldcg(box_lo + ch.x), l1 = __ldcg(box_lo + ch.y), h0 = __ldcg(box_hi + ch.x), h1 = __ldcg(box_hi + ch.y);
        float fl[8], fr[8];
        if ((uint32_t)ch.x < (uint32_t)(n - 1)) unpack_table(__ldcg(dp + 2 * (size_t)ch.x), __ldcg(dp + 2 * (size_t)ch.x + 1), fl);
        else leaf_table(l0, h0, slot_cost, fl);  // a triangle: one leaf child whatever the budget (only internal nodes have a table)
        if ((uint32_t)ch.y < (uint32_t)(n - 1)) unpack_table(__ldcg(dp + 2 * (size_t)ch.y), __ldcg(dp + 2 * (size_t)ch.y + 1), fr);
        else leaf_table(l1, h1, slot_cost, fr);
        float4 lo, hi;
        float f[8];
        uint32_t dec;
        dp_combine(l0, h0, fl, l1, h1, fr, slot_cost, node_cost, lo, hi, f, dec);
        dp[2 * (size_t)p] = make_float4(f[1], f[2], f[3], f[4]);
        dp[2 * (size_t)p + 1] = make_float4(f[5], f[6], f[7], __uint_as_float(dec));
        box_lo[p] = lo;
        box_hi[p] = hi;
        node = p;
    }
}
__device__ __forceinline__ void leaf_box(const float4* __restrict__ flat_tris, uint32_t src, int k, float pad, float4& lo, float4& hi) {
    const float4 a = flat_tris[3 * (size_t)src], b = flat_tris[3 * (size_t)src + 1], c = flat_tris[3 * (size_t)src + 2];
    lo = make_float4(fminf(fminf(a.x, b.x), c.x) - pad, fminf(fminf(a.y, b.y), c.y) - pad, fminf(fminf(a.z, b.z), c.z) - pad, __int_as_float(1));
    hi = make_float4(fmaxf(fmaxf(a.x, b.x), c.x) + pad, fmaxf(fmaxf(a.y, b.y), c.y) + pad, fmaxf(fmaxf(a.z, b.z), c.z) + pad, __int_as_float(k));
}
// Any binary hierarchy (the binned-SAH builder's): one thread per leaf, every level through the global hand-off.
__global__ void refit_kernel(const float4* __restrict__ flat_tris, const uint32_t* __restrict__ sorted_idx, int n, float pad,
                             const int2* __restrict__ children, const uint32_t* __restrict__ parent, uint32_t* __restrict__ flags,
                             float4* __restrict__ box_lo, float4* __restrict__ box_hi, float4* __restrict__ dp, float slot_cost,
                             float node_cost) {
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
        float4 lo, hi;
        leaf_box(flat_tris, sorted_idx[k], k, pad, lo, hi);
        const uint32_t node = (uint32_t)(n - 1 + k);
        box_lo[node] = lo;
        box_hi[node] = hi;
        if (n == 1) return;   // (a single triangle has no F table: F(leaf, j) = area * (slot + 1) is formed by its parent)
        refit_climb_global(node, n, children, parent, flags, box_lo, box_hi, dp, slot_cost, node_cost);
    }
}

// The same pass for the LBVH hierarchy, in treelets.  In Karras's numbering a node's index lies inside its own leaf range,
// so the subtrees over a block of FF_TREELET consecutive sorted leaves -- all but the ~log levels above them -- live in a
// contiguous index range: one CTA resolves them in SHARED memory (hand-off by shared-memory atomics, children read back
// from shared memory), writes the block's boxes and tables out with coalesced stores, and only the few nodes whose parent's
// range crosses the block boundary (FF_PARENT_CROSS) continue through the global hand-off above.  At 20 M triangles the
// all-global version was a 25-level chain of global atomic -> L2 loads -> stores per node, 6.5 of the build's 12 ms
// (71 % long-scoreboard, 21 % membar stalls: profiles/r01i_refit_20m_stalls.txt).
#define FF_TREELET_THREADS 256
__global__ void __launch_bounds__(FF_TREELET_THREADS) refit_treelet_kernel(const float4* __restrict__ flat_tris, const uint32_t* __restrict__ sorted_idx,
                                                                            int n, float pad, const int2* __restrict__ children,
                                                                            const uint32_t* __restrict__ parent, uint32_t* __restrict__ flags,
                                                                            float4* __restrict__ box_lo, float4* __restrict__ box_hi,
                                                                            float4* __restrict__ dp, float slot_cost, float node_cost,
                                                                            uint32_t* __restrict__ starters, uint32_t* __restrict__ starter_count) {
    constexpr int S = FF_TREELET, T = FF_TREELET_THREADS;
    extern __shared__ float4 sm4[];
    float4* s_lo = sm4;              // [0, S): internal nodes of the block (index - base), [S, 2S): its leaves
    float4* s_hi = sm4 + 2 * S;
    float4* s_dp = sm4 + 4 * S;      // 2 per internal node
    uint32_t* s_flag = reinterpret_cast<uint32_t*>(sm4 + 6 * S);
    uint32_t* s_parent = s_flag + S;                              // [0, S): parents of the internal nodes, [S, 2S): of the leaves
    int2* s_children = reinterpret_cast<int2*>(s_parent + 2 * S); // children of the internal nodes
    uint32_t (*s_q)[FF_TREELET] = reinterpret_cast<uint32_t (*)[FF_TREELET]>(s_children + S);   // two queues of finished-children nodes
    uint32_t* s_qn = reinterpret_cast<uint32_t*>(s_q + 2);   // [0], [1]: queue lengths; [2]: starters of this block; [3]: their place in the global list
    uint32_t* s_start = s_qn + 4;                               // nodes finished here whose parent crosses the block boundary (global ids)
    const int base = blockIdx.x * S;
    const int m = min(S, n - base);
    const uint32_t n_internal = (uint32_t)(n - 1);
    // The links of the block, staged with coalesced loads: the climb below is then a chain of shared-memory accesses only
    // (with parent[] / children[] read from global memory inside the chain, the ~9 levels of a block cost two dependent L2
    // round trips each and a CTA spent most of its life waiting: 3.1 ms at 20 M triangles instead of 6.5, still latency-bound).
    for (int i = threadIdx.x; i < S; i += T) {
        s_flag[i] = 0u;
        const uint32_t g = (uint32_t)(base + i);
        s_parent[i] = g < n_internal ? parent[g] : FF_PARENT_ROOT;
        s_children[i] = g < n_internal ? children[g] : make_int2(0, 0);
        s_parent[S + i] = i < m ? parent[n_internal + g] : FF_PARENT_ROOT;
    }
    __syncthreads();
    // Level-synchronous inside the block: a node whose second child has just been finished is appended to the NEXT round's
    // queue instead of being computed by the arriving thread, and each round hands its queue to consecutive threads.  (With
    // the classic climb -- the second arrival computes the parent at once -- half of the lanes drop out per level while the
    // warp keeps executing the ~250 instructions of the DP: 7 of 32 lanes active on average, and the kernel was bound by
    // instruction issue, profiles/r02g_build1m_full details.)
    if (threadIdx.x < 4) s_qn[threadIdx.x] = 0u;
    __syncthreads();
    auto arrive = [&](int local, int buf) {   // the node at `local` is finished: count it at its parent
        const uint32_t pp = s_parent[local];
        if (pp == FF_PARENT_ROOT) return;
        if (pp & FF_PARENT_CROSS) {   // handed over to refit_cross_kernel, which starts its climbs from the list this block appends to
            s_start[atomicAdd(&s_qn[2], 1u)] = local < S ? (uint32_t)(base + local) : n_internal + (uint32_t)(base + local - S);
            return;
        }
        const int pl = (int)pp - base;
        if (atomicAdd(&s_flag[pl], 1u) == 1u) s_q[buf][atomicAdd(&s_qn[buf], 1u)] = (uint32_t)pl;
    };
    for (int t = threadIdx.x; t < m; t += T) {
        const int k = base + t;
        float4 lo, hi;
        leaf_box(flat_tris, sorted_idx[k], k, pad, lo, hi);
        s_lo[S + t] = lo;
        s_hi[S + t] = hi;
        if (n > 1) arrive(S + t, 0);
    }
    __syncthreads();
    for (int cur = 0;; cur ^= 1) {
        const uint32_t count = s_qn[cur];
        if (count == 0u) break;
        __syncthreads();                       // everybody has read the count before the other queue is reset
        if (threadIdx.x == 0) s_qn[cur ^ 1] = 0u;
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < count; i += T) {
            const int pl = (int)s_q[cur][i];
            const int2 ch = s_children[pl];
            const bool leaf0 = (uint32_t)ch.x >= n_internal, leaf1 = (uint32_t)ch.y >= n_internal;
            const int c0 = leaf0 ? S + (ch.x - (int)n_internal - base) : ch.x - base, c1 = leaf1 ? S + (ch.y - (int)n_internal - base) : ch.y - base;
            const float4 l0 = s_lo[c0], h0 = s_hi[c0], l1 = s_lo[c1], h1 = s_hi[c1];
            float fl[8], fr[8];
            if (leaf0) leaf_table(l0, h0, slot_cost, fl); else unpack_table(s_dp[2 * c0], s_dp[2 * c0 + 1], fl);
            if (leaf1) leaf_table(l1, h1, slot_cost, fr); else unpack_table(s_dp[2 * c1], s_dp[2 * c1 + 1], fr);
            float4 lo, hi;
            float f[8];
            uint32_t dec;
            dp_combine(l0, h0, fl, l1, h1, fr, slot_cost, node_cost, lo, hi, f, dec);
            s_lo[pl] = lo;
            s_hi[pl] = hi;
            s_dp[2 * pl] = make_float4(f[1], f[2], f[3], f[4]);
            s_dp[2 * pl + 1] = make_float4(f[5], f[6], f[7], __uint_as_float(dec));
            arrive(pl, cur ^ 1);
        }
        __syncthreads();                       // this round's boxes and tables, and the next round's queue, are complete
    }
    __syncthreads();
    // the block's leaves and finished internal nodes to global memory, coalesced
    for (int t = threadIdx.x; t < m; t += T) {
        box_lo[n_internal + (uint32_t)(base + t)] = s_lo[S + t];
        box_hi[n_internal + (uint32_t)(base + t)] = s_hi[S + t];
    }
    for (int i = threadIdx.x; i < S; i += T)
        if (s_flag[i] == 2u) {
            const size_t g = (size_t)base + i;
            box_lo[g] = s_lo[i];
            box_hi[g] = s_hi[i];
            dp[2 * g] = s_dp[2 * i];
            dp[2 * g + 1] = s_dp[2 * i + 1];
        }
    // this block's starters join the global list (one atomic per block); the kernel boundary publishes them and the boxes above
    if (threadIdx.x == 0) s_qn[3] = atomicAdd(starter_count, s_qn[2]);
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < s_qn[2]; i += T) starters[s_qn[3] + i] = s_start[i];
}
// The levels above the treelets: every node (or leaf) that was finished inside a block but whose PARENT's range crosses a block
// boundary starts a climb through the global hand-off.  A separate launch, not a tail of the treelet kernel: a climb is a
// chain of up to ~15 global atomic -> load -> store round trips, and a treelet CTA that waited for its climbs kept its 58 KB
// of shared memory -- and with it a third of an SM -- for tens of microseconds (3.1 ms at 20 M triangles).
__global__ void refit_cross_kernel(int n, const int2* __restrict__ children, const uint32_t* __restrict__ parent, uint32_t* __restrict__ flags,
                                   float4* __restrict__ box_lo, float4* __restrict__ box_hi, float4* __restrict__ dp, float slot_cost, float node_cost,
                                   const uint32_t* __restrict__ starters, const uint32_t* __restrict__ starter_count) {
    // One thread per starter, all resident at once: the launch lasts as long as the longest chain of hand-offs.  (Finding the
    // starters by scanning parent[] of all 2n - 1 nodes took 1.3 ms at 20 M triangles: a starter late in a thread's scan began
    // its chain late, and everything above it waited -- profiles/r02w_build20m_details.txt.)
    const uint32_t count = *starter_count;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x)
        refit_climb_global(starters[i], n, children, parent, flags, box_lo, box_hi, dp, slot_cost, node_cost);
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

// ONE THREAD per wide node: the throughput version, used for the large levels of the hierarchy (collapse_persistent_kernel).
// node_src[w] = the binary node wide node w is built from.  Wide-node indices are handed out by one counter, so the nodes of
// a level are a contiguous index range and the levels need no host round trip.
__device__ __forceinline__ void collapse_one(uint32_t bnode, uint32_t wnode, uint32_t* __restrict__ node_src, uint32_t node_capacity, int n,
                                const int2* __restrict__ children, const float4* __restrict__ box_lo, const float4* __restrict__ box_hi,
                                const float4* __restrict__ flat_tris, const uint32_t* __restrict__ sorted_idx,
                                float4* __restrict__ nodes, float4* __restrict__ tris_out, uint32_t* __restrict__ leaf_slot,
                                CollapseCounters* ctr, float root_area, const float4* __restrict__ dp, uint32_t (*e_node)[8][128], uint32_t (*e_bud)[8][128]) {
    const uint32_t n_internal = (uint32_t)(n - 1);

    uint32_t ch[8];
    int nch = 0;
    float4 plo = box_lo[bnode], phi = box_hi[bnode];
    if (bnode >= n_internal || (__float_as_int(plo.w) & FF_COUNT_MASK) <= FF_LEAF_MAX) {
        ch[nch++] = bnode;  // tiny scene: the whole tree is one leaf child
    } else {
        // Follow the decisions of the refit DP: a list of (binary node, slot budget) pairs in left-to-right order, budgets
        // summing to 8.  Every round expands ALL entries that the DP splits, with the loads of the round (decision word and
        // children of each entry) issued together: 3-4 memory round trips per wide node.  (The depth-first walk this replaces
        // did one entry at a time, two dependent loads each: up to ~30 round trips, and the collapse of 1 M triangles spent
        // 0.42 ms in nine levels of that whether it ran as 24 launches or as one persistent kernel.)
        // e_node / e_bud: this thread's column of two [8][128] shared arrays per buffer (dynamic indexing without local memory).
        const int t = threadIdx.x;
        int cur = 0, c = 2;
        {
            const uint32_t k8 = (__float_as_uint(__ldg(&dp[2 * (size_t)bnode + 1].w)) >> 21) & 7u;
            const int2 c2 = __ldg(children + bnode);
            e_node[0][0][t] = (uint32_t)c2.x; e_bud[0][0][t] = k8;
            e_node[0][1][t] = (uint32_t)c2.y; e_bud[0][1][t] = 8u - k8;
        }
        for (;;) {
            uint32_t dec[8];
            int2 kid[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                dec[i] = 0u; kid[i] = make_int2(0, 0);
                if (i < c) {
                    const uint32_t m = e_node[cur][i][t], j = e_bud[cur][i][t];
                    if (m < n_internal && j >= 2u) { dec[i] = __float_as_uint(__ldg(&dp[2 * (size_t)m + 1].w)); kid[i] = __ldg(children + m); }
                }
            }
            int nc = 0;
            bool changed = false;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (i < c) {
                    const uint32_t m = e_node[cur][i][t], j = e_bud[cur][i][t];
                    const uint32_t d = (m < n_internal && j >= 2u) ? (dec[i] >> (3u * (j - 2u))) & 7u : 0u;
                    if (d == 0u || d >= j || nc + 2 + (c - 1 - i) > 8) {   // one slot: a leaf (<= FF_LEAF_MAX triangles) or an inner child; budget 1 = settled
                        e_node[cur ^ 1][nc][t] = m; e_bud[cur ^ 1][nc][t] = 1u; ++nc;
                    } else {
                        e_node[cur ^ 1][nc][t] = (uint32_t)kid[i].x; e_bud[cur ^ 1][nc][t] = d; ++nc;
                        e_node[cur ^ 1][nc][t] = (uint32_t)kid[i].y; e_bud[cur ^ 1][nc][t] = j - d; ++nc;
                        changed = true;
                    }
                }
            }
            cur ^= 1; c = nc;
            if (!changed) break;
        }
        for (int i = 0; i < c; ++i) ch[i] = e_node[cur][i][t];
        nch = c;
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
        if (c < n_internal && cnt > FF_LEAF_MAX) { imask |= 1u << s; ++n_inner; }
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
// EIGHT LANES per wide node (a cg tile), lane k owning child k through the whole job: the latency version, used for the small
// levels at the top of the hierarchy.
//   One thread per node is ~6000 dependent instructions (512 slot-cost evaluations of the greedy assignment, 48 quantisations
//   with a division each) executed by a single lane: ~30 us per node, and since a level cannot start before its parent level
//   has allocated its nodes, nine levels cost 270 us before any throughput mattered -- the grid barrier was 29 of the 47 stall
//   cycles per issued instruction (profiles/r02g_build1m_full).  Spread over 8 lanes the critical path of a node is ~1/6; on
//   a level with millions of nodes the eight-fold thread count costs more than the shorter path saves (20 M triangles: 3.7 ms
//   against 2.9), hence the two versions.
// The tree is the same as the one-thread version built: same children in the same order, the same greedy assignment with the
// same tie-breaks (lowest child, then lowest slot), the same quantisation arithmetic.
typedef cg::thread_block_tile<8> Tile8;
__device__ __forceinline__ uint32_t tile_or(const Tile8& t, uint32_t v) {
    v |= t.shfl_xor(v, 1); v |= t.shfl_xor(v, 2); v |= t.shfl_xor(v, 4);
    return v;
}
__device__ __forceinline__ uint32_t tile_add(const Tile8& t, uint32_t v) {
    v += t.shfl_xor(v, 1); v += t.shfl_xor(v, 2); v += t.shfl_xor(v, 4);
    return v;
}
__device__ __forceinline__ void collapse_group(const Tile8& tile, uint32_t bnode, uint32_t wnode, uint32_t* __restrict__ node_src, uint32_t node_capacity, int n,
                                const int2* __restrict__ children, const float4* __restrict__ box_lo, const float4* __restrict__ box_hi,
                                float4* __restrict__ nodes, uint32_t* __restrict__ leaf_slot, CollapseCounters* ctr, float root_area,
                                const float4* __restrict__ dp, uint2* e_list /* this tile's 8 entries in shared memory */) {
    const uint32_t n_internal = (uint32_t)(n - 1);
    const int L = (int)tile.thread_rank();
    float4 plo = __ldg(box_lo + bnode), phi = __ldg(box_hi + bnode);

    // ---- the children: follow the decisions of the refit DP.  A list of (binary node, slot budget) pairs in left-to-right order,
    // budgets summing to 8, lane i holding entry i; every round expands ALL entries the DP splits (their loads in flight
    // together) and the list is re-packed through a prefix sum: 3-4 memory round trips per wide node.
    uint32_t my_node = 0xffffffffu, my_bud = 0u;
    int c;
    if (bnode >= n_internal || (__float_as_int(plo.w) & FF_COUNT_MASK) <= FF_LEAF_MAX) {
        c = 1;   // tiny scene: the whole tree is one leaf child
        if (L == 0) { my_node = bnode; my_bud = 1u; }
    } else {
        const uint32_t k8 = (__float_as_uint(__ldg(&dp[2 * (size_t)bnode + 1].w)) >> 21) & 7u;
        const int2 c2 = __ldg(children + bnode);
        c = 2;
        if (L == 0) { my_node = (uint32_t)c2.x; my_bud = k8; }
        if (L == 1) { my_node = (uint32_t)c2.y; my_bud = 8u - k8; }
        for (;;) {
            uint32_t d = 0u;
            int2 kid = make_int2(0, 0);
            if (L < c && my_node < n_internal && my_bud >= 2u) {
                d = (__float_as_uint(__ldg(&dp[2 * (size_t)my_node + 1].w)) >> (3u * (my_bud - 2u))) & 7u;
                kid = __ldg(children + my_node);
            }
            const bool split = d != 0u && d < my_bud;
            const uint32_t splits = tile.ballot(split);
            if (splits == 0u || c + __popc(splits) > 8) break;   // (budgets sum to 8, so the list cannot outgrow the tile)
            const int pos = L + __popc(splits & ((1u << L) - 1u));   // entries before mine after the re-pack
            tile.sync();
            if (L < c) {
                if (split) { e_list[pos] = make_uint2((uint32_t)kid.x, d); e_list[pos + 1] = make_uint2((uint32_t)kid.y, my_bud - d); }
                else e_list[pos] = make_uint2(my_node, 1u);   // one slot: a leaf (<= FF_LEAF_MAX triangles) or an inner child; settled
            }
            c += __popc(splits);
            tile.sync();
            if (L < c) { const uint2 e = e_list[L]; my_node = e.x; my_bud = e.y; }
        }
    }
    const int nch = c;
    const bool have = L < nch;
    const uint32_t ch = my_node;

    // ---- my child's box, and its centre relative to the node's
    float4 cl = make_float4(0.f, 0.f, 0.f, 0.f), chh = make_float4(0.f, 0.f, 0.f, 0.f);
    if (have) { cl = __ldg(box_lo + ch); chh = __ldg(box_hi + ch); }
    const int cnt = __float_as_int(cl.w) & FF_COUNT_MASK;
    const float pcx = 0.5f * (plo.x + phi.x), pcy = 0.5f * (plo.y + phi.y), pcz = 0.5f * (plo.z + phi.z);
    const float cdx = have ? 0.5f * (cl.x + chh.x) - pcx : 0.f, cdy = have ? 0.5f * (cl.y + chh.y) - pcy : 0.f, cdz = have ? 0.5f * (cl.z + chh.z) - pcz : 0.f;

    // ---- slot assignment: greedy maximisation of dot(child centre - node centre, octant direction of the slot); the pair
    // (child, slot) with the largest value is fixed, ties go to the lowest child and then the lowest slot; repeat.
    int my_slot = -1;
    uint32_t used = 0u;
    for (int it = 0; it < nch; ++it) {
        float bestc = -FLT_MAX;
        int bs = -1;
        if (have && my_slot < 0) {
#pragma unroll
            for (int sl = 0; sl < 8; ++sl) {
                if (used & (1u << sl)) continue;
                const float v = ((sl & 4) ? cdx : -cdx) + ((sl & 2) ? cdy : -cdy) + ((sl & 1) ? cdz : -cdz);
                if (v > bestc) { bestc = v; bs = sl; }
            }
        }
        // arg max over the lanes: larger value wins, equal values go to the lower lane
        float wc = bestc;
        int wk = bs >= 0 ? L : 8, ws = bs;
#pragma unroll
        for (int off = 1; off < 8; off <<= 1) {
            const float oc = tile.shfl_xor(wc, off);
            const int ok = tile.shfl_xor(wk, off), os = tile.shfl_xor(ws, off);
            if (ok < 8 && (wk >= 8 || oc > wc || (oc == wc && ok < wk))) { wc = oc; wk = ok; ws = os; }
        }
        if (wk >= 8) {   // NaN costs everywhere: first unassigned child, first free slot
            const uint32_t un = tile.ballot(have && my_slot < 0);
            wk = __ffs(un) - 1;
            ws = __ffs(~used & 0xffu) - 1;
        }
        if (L == wk) my_slot = ws;
        used |= 1u << ws;
    }

    // ---- classify, allocate
    const bool inner = have && ch < n_internal && cnt > FF_LEAF_MAX;
    const bool leaf = have && !inner;
    const uint32_t imask = tile_or(tile, inner ? 1u << my_slot : 0u);
    const uint32_t n_inner = __popc(imask);
    const uint32_t n_leaf_tris = tile_add(tile, leaf ? (uint32_t)cnt : 0u);
    // One atomicAdd per counter and WARP (a scan over the first lanes of the tiles that arrive together): per wide node the
    // three counters of a 20 M-triangle build took 8.4 M same-address atomics, which the L2 serialises.
    uint32_t child_base = 0, tri_base = 0;
    if (L == 0) {
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
    child_base = tile.shfl(child_base, 0);
    tri_base = tile.shfl(tri_base, 0);
    bool fits = true;
    if (n_inner) {
        fits = child_base + n_inner <= node_capacity;
        if (!fits && L == 0) atomicExch(&ctr->overflow, 1u);   // reported by bvh_build; the children are not queued, nothing is written out of bounds
    }
    // children keep slot order in both index spaces: inner children as wide nodes child_base + rank, leaf triangles as
    // slots tri_base + offset
    uint32_t tri_off = 0;
    {
        const uint32_t mine = have ? ((uint32_t)my_slot | (leaf ? (uint32_t)cnt << 4 : 0u)) : 0xfu;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const uint32_t o = tile.shfl(mine, j);
            if ((o & 0xfu) < (uint32_t)(my_slot < 0 ? 0 : my_slot)) tri_off += o >> 4;
        }
    }
    if (inner && fits) node_src[child_base + __popc(imask & ((1u << my_slot) - 1u))] = ch;

    // ---- quantisation frame: the node box pushed out by 2^-13 of its extent per side.  The traversal evaluates a plane as
    // u*A + B with A = 2^15 * scale / d (byte_unit in ff_trace.cuh); the rounding of B moves a plane by at most 2^-16 of
    // the extent, and floor/ceil against the wider frame keeps every child box inside its quantised box by 8x that.
    {
        const float gx = (phi.x - plo.x) * 0x1p-13f, gy = (phi.y - plo.y) * 0x1p-13f, gz = (phi.z - plo.z) * 0x1p-13f;
        plo.x -= gx; plo.y -= gy; plo.z -= gz;
        phi.x += gx; phi.y += gy; phi.z += gz;
    }
    const uint32_t ebx = exponent_for(phi.x - plo.x), eby = exponent_for(phi.y - plo.y), ebz = exponent_for(phi.z - plo.z);
    const float sx = __uint_as_float(ebx << 23), sy = __uint_as_float(eby << 23), sz = __uint_as_float(ebz << 23);
    uint32_t b_meta = 0, b_qlx = 0, b_qly = 0, b_qlz = 0, b_qhx = 0, b_qhy = 0, b_qhz = 0;
    double sah = 0.0;
    if (have) {
        if (inner) b_meta = 0x20u | (24u + (uint32_t)my_slot);
        else {
            const uint32_t first = (uint32_t)__float_as_int(chh.w);
            b_meta = (((1u << cnt) - 1u) << 5) | tri_off;
            // only the destination is recorded here; gather_tris_kernel moves the 48-byte records afterwards, one thread per triangle
            for (int t = 0; t < cnt; ++t) leaf_slot[first + t] = tri_base + tri_off + (uint32_t)t;
            sah = (double)box_area(cl, chh) * cnt;
        }
        const float gx = (chh.x - cl.x) * 0x1p-13f + (phi.x - plo.x) * 0x1p-15f, gy = (chh.y - cl.y) * 0x1p-13f + (phi.y - plo.y) * 0x1p-15f,
                    gz = (chh.z - cl.z) * 0x1p-13f + (phi.z - plo.z) * 0x1p-15f;
        b_qlx = quant_lo(cl.x - gx, plo.x, sx); b_qly = quant_lo(cl.y - gy, plo.y, sy); b_qlz = quant_lo(cl.z - gz, plo.z, sz);
        b_qhx = quant_hi(chh.x + gx, plo.x, sx); b_qhy = quant_hi(chh.y + gy, plo.y, sy); b_qhz = quant_hi(chh.z + gz, plo.z, sz);
    }
    // bytes of the 8 lanes -> the packed words of the node: word w of an array holds slots 4w .. 4w+3
    const int sh = have ? 8 * (my_slot & 3) : 0;
    const bool w0 = have && my_slot < 4, w1 = have && my_slot >= 4;
#define FF_PACK(b, lo_word, hi_word) const uint32_t lo_word = tile_or(tile, w0 ? (b) << sh : 0u), hi_word = tile_or(tile, w1 ? (b) << sh : 0u)
    FF_PACK(b_meta, meta0, meta1);
    FF_PACK(b_qlx, qlx0, qlx1); FF_PACK(b_qly, qly0, qly1); FF_PACK(b_qlz, qlz0, qlz1);
    FF_PACK(b_qhx, qhx0, qhx1); FF_PACK(b_qhy, qhy0, qhy1); FF_PACK(b_qhz, qhz0, qhz1);
#undef FF_PACK
    float4* out = nodes + FF_NODE_F4 * (size_t)wnode;
    const uint32_t xb = FF_NODE_EXP_BIAS;  // the node stores the exponents of 2^15 * scale
    if (L == 0) out[0] = make_float4(plo.x, plo.y, plo.z, __uint_as_float((ebx + xb) | ((eby + xb) << 8) | ((ebz + xb) << 16) | (imask << 24)));
    if (L == 1) out[1] = make_float4(__uint_as_float(child_base), __uint_as_float(tri_base), __uint_as_float(meta0), __uint_as_float(meta1));
#if FF_NODE_256
    if (L == 2) out[2] = make_float4(__uint_as_float(qlx0), __uint_as_float(qlx1), __uint_as_float(qly0), __uint_as_float(qly1));
    if (L == 3) out[3] = make_float4(__uint_as_float(qlz0), __uint_as_float(qlz1), 0.f, 0.f);
    if (L == 4) out[4] = make_float4(__uint_as_float(qhx0), __uint_as_float(qhx1), __uint_as_float(qhy0), __uint_as_float(qhy1));
    if (L == 5) out[5] = make_float4(__uint_as_float(qhz0), __uint_as_float(qhz1), 0.f, 0.f);
#else
    if (L == 2) out[2] = make_float4(__uint_as_float(qlx0), __uint_as_float(qlx1), __uint_as_float(qly0), __uint_as_float(qly1));
    if (L == 3) out[3] = make_float4(__uint_as_float(qlz0), __uint_as_float(qlz1), __uint_as_float(qhx0), __uint_as_float(qhx1));
    if (L == 4) out[4] = make_float4(__uint_as_float(qhy0), __uint_as_float(qhy1), __uint_as_float(qhz0), __uint_as_float(qhz1));
#endif
    // SAH cost of the wide tree: C_node * area(node) + C_tri * sum(leaf area * count), normalised by the root area
    if (root_area > 0.f) {
        double tot = sah;
        tot += tile.shfl_xor(tot, 1); tot += tile.shfl_xor(tot, 2); tot += tile.shfl_xor(tot, 4);
        if (L == 0) {
            cg::coalesced_group g = cg::coalesced_threads();
            const double sum = cg::reduce(g, ((double)box_area(plo, phi) + tot) / (double)root_area, cg::plus<double>());
            if (g.thread_rank() == 0) atomicAdd(&ctr->sah, sum);
        }
    }
}
// Triangles into BVH-leaf order: sorted position p -> slot leaf_slot[p] (written by the collapse).
__global__ void gather_tris_kernel(const float4* __restrict__ flat_tris, const uint32_t* __restrict__ sorted_idx,
                                   const uint32_t* __restrict__ leaf_slot, uint32_t n, float4* __restrict__ tris_out) {
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
        const size_t src = sorted_idx[p], slot = leaf_slot[p];
        if (slot >= n) continue;   // (only after a node-capacity overflow, which bvh_build reports: the slot was never assigned)
        const float4 a = __ldg(flat_tris + 3 * src), b = __ldg(flat_tris + 3 * src + 1), c = __ldg(flat_tris + 3 * src + 2);
        tris_out[3 * slot] = a; tris_out[3 * slot + 1] = b; tris_out[3 * slot + 2] = c;
    }
}
// All levels in ONE launch: a persistent grid (every CTA resident: cooperative launch) walks the levels with a grid barrier in
// between.  The last CTA to arrive at the barrier moves the level range to the nodes the level allocated before it releases the others, so one barrier per level is enough.  The
// per-level launches cost 24 kernel launches and two host round trips per build whatever the depth -- 0.4 of the 0.93 ms of a
// 1 M-triangle build (profiles/r02a_launches_bench_summary.json).
struct GridBarrier { uint32_t arrived; uint32_t generation; };
#ifndef FF_COLLAPSE_TILE_LEVEL
#define FF_COLLAPSE_TILE_LEVEL 16384u   // levels of at most this many wide nodes are collapsed by 8-lane tiles, larger ones by one thread per node
#endif
#ifndef FF_COLLAPSE_MINB
#define FF_COLLAPSE_MINB 6   // resident CTAs per SM the collapse is compiled for: 85-register cap (uncapped: 94 registers, 5 CTAs).  20 M triangles:
                             // 6.00 / 5.80 / 5.95 ms with no cap / 6 / 8; 1 M unchanged (profiles/r04h_build_variants_minb.txt)
#endif
__global__ void __launch_bounds__(128, FF_COLLAPSE_MINB) collapse_persistent_kernel(uint32_t* __restrict__ node_src, uint32_t node_capacity, int n,
                                const int2* __restrict__ children, const float4* __restrict__ box_lo, const float4* __restrict__ box_hi,
                                const float4* __restrict__ flat_tris, const uint32_t* __restrict__ sorted_idx,
                                float4* __restrict__ nodes, float4* __restrict__ tris_out, uint32_t* __restrict__ leaf_slot,
                                CollapseCounters* ctr, const float4* __restrict__ root_lo, const float4* __restrict__ root_hi,
                                const float4* __restrict__ dp, GridBarrier* bar) {
    const float4 rl = __ldg(root_lo), rh = __ldg(root_hi);
    const float root_area = box_area(rl, rh);
    __shared__ uint32_t s_range[2];
    __shared__ uint32_t e_node[2][8][128], e_bud[2][8][128];
    uint2 (*e_lists)[8] = reinterpret_cast<uint2 (*)[8]>(&e_node[0][0][0]);   // the tile version's lists alias the thread version's
    const Tile8 tile = cg::tiled_partition<8>(cg::this_thread_block());
    const uint32_t tile_in_block = threadIdx.x >> 3, tiles_per_block = blockDim.x >> 3;
    for (uint32_t gen = 0;; ++gen) {
        if (threadIdx.x == 0) {
            s_range[0] = *(volatile uint32_t*)&ctr->level_begin;
            s_range[1] = *(volatile uint32_t*)&ctr->level_end;
        }
        __syncthreads();
        const uint32_t begin = s_range[0], end = s_range[1];
#if FF_COLLAPSE_TRACE   // development aid (-DFF_COLLAPSE_TRACE=1): when each level started on block 0, printed at the end.  1 M triangles:
        // levels of 1 / 8 / 64 / 512 / 4096 nodes take 11 / 14 / 15 / 16 / 18 us each (the latency of one node: barrier, DP descent,
        // allocation), 32 768 nodes 56 us, 108 871 nodes 101 us (profiles/r04b_collapse_levels.txt).  Lighter barrier fences: no change.
        __shared__ unsigned long long s_t[40];
        __shared__ uint32_t s_cnt[40];
        if (blockIdx.x == 0 && threadIdx.x == 0 && gen < 40) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); s_t[gen] = t; s_cnt[gen] = end - begin; }
        if (begin == end && blockIdx.x == 0 && threadIdx.x == 0)
            for (uint32_t g = 0; g + 1 <= gen && g < 39; ++g) printf("[collapse] level %u: %u nodes, %.1f us\n", g, s_cnt[g], (double)(s_t[g + 1] - s_t[g]) * 1e-3);
#endif
        if (begin == end) return;   // uniform over the grid: every CTA reads the range the barrier published
        if (end - begin <= FF_COLLAPSE_TILE_LEVEL) {   // a small level: latency matters, eight lanes per node
            for (uint32_t w = begin + blockIdx.x * tiles_per_block + tile_in_block; w < end; w += gridDim.x * tiles_per_block)
                collapse_group(tile, __ldcg(node_src + w), w, node_src, node_capacity, n, children, box_lo, box_hi, nodes, leaf_slot, ctr, root_area, dp,
                               e_lists[tile_in_block]);
        } else {
            for (uint32_t w = begin + blockIdx.x * blockDim.x + threadIdx.x; w < end; w += gridDim.x * blockDim.x)
                collapse_one(__ldcg(node_src + w), w, node_src, node_capacity, n, children, box_lo, box_hi, flat_tris, sorted_idx, nodes, tris_out, leaf_slot,
                             ctr, root_area, dp, e_node, e_bud);
        }
        // ---- grid barrier; the last arrival advances the level
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            const uint32_t ticket = atomicAdd(&bar->arrived, 1u);
            if (ticket == (gen + 1u) * gridDim.x - 1u) {
                const uint32_t allocated = min(*(volatile uint32_t*)&ctr->node_count, node_capacity);
                if (end > begin) ctr->levels += 1;
                ctr->level_begin = end;
                ctr->level_end = *(volatile uint32_t*)&ctr->overflow ? end : allocated;
                __threadfence();
                atomicExch(&bar->generation, gen + 1u);
            } else {
                while (*(volatile uint32_t*)&bar->generation <= gen) __nanosleep(40);
            }
            __threadfence();
        }
        __syncthreads();
    }
}

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return e_; } while (0)

cudaError_t sah_build_hierarchy(const float4* d_flat_tris, const uint32_t* d_sorted_idx, uint32_t n, const float centre_lo[3],
                                const float centre_hi[3], int2* d_children, uint32_t* d_parent, uint32_t* d_order_out,
                                uint32_t* launches, uint32_t* out_levels, Arena& arena, cudaStream_t st);  // ff_sah.cu
size_t sah_scratch_bytes(uint32_t n);

cudaError_t bvh_build(const float4* d_flat_tris, uint32_t n, int builder, float pad, float4* d_tris_out, float4* d_nodes_out,
                      uint32_t node_capacity, int2* d_bvh2_children, float4* d_bvh2_lo, float4* d_bvh2_hi,
                      uint32_t* d_bvh2_leaf_slot, float bounds_lo[3], float bounds_hi[3], BuildOut* out, uint32_t* launches,
                      Arena& arena, cudaStream_t st, const int* d_bounds_ready) {
    out->n_nodes = 0; out->depth = 0; out->sah_cost = 0; out->morton_bits = 0;
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
    // Morton keys: 30 bits (u32, 4 sort passes) unless that grid does not resolve the scene.  Above 2^20 triangles the build
    // sorts with 30-bit keys first, counts the sorted keys that repeat their predecessor and, if more than 1/16 of them do,
    // redoes codes + sort with 63-bit keys (u64, 8 passes): inside a cell the hierarchy can only fall back to the input
    // order.  A uniform 20 M-triangle soup has 2 % duplicates on the 1024^3 grid and gains nothing from the wide keys
    // (35.9108 vs 35.9107 nodes per ray, 9.3 vs 7.7 ms build: profiles/r02f, r02g); a scene whose triangles sit in a small
    // part of its bounding box does.  FF_MORTON_BITS=30 / 63 forces one or the other (development knob).
    const char* mb_env = getenv("FF_MORTON_BITS");
    const bool may_widen = mb_env ? atoi(mb_env) > 32 : n > (1u << 20);
    bool wide_keys = mb_env ? atoi(mb_env) > 32 : false;
    const size_t key_bytes = may_widen ? sizeof(uint64_t) : sizeof(uint32_t);
    const size_t sort_words = may_widen ? sort_scratch_words<uint64_t>(n) : sort_scratch_words<uint32_t>(n);
    int* d_bounds = nullptr;
    void *d_keys_a = nullptr, *d_keys_b = nullptr;
    uint32_t *d_vals_a = nullptr, *d_vals_b = nullptr, *d_hist = nullptr;
    uint32_t *d_parent = nullptr, *d_flags = nullptr;
    float4* d_dp = nullptr;
    uint32_t* d_src = nullptr;
    CollapseCounters* d_ctr = nullptr;
    {
        size_t need = Arena::padded(12 * sizeof(int)) + 2 * Arena::padded(sizeof(uint32_t) * n) + 2 * Arena::padded(key_bytes * n) + Arena::padded(sizeof(uint32_t) * sort_words) +
                      Arena::padded(sizeof(uint32_t) * 2 * (size_t)n) + Arena::padded(sizeof(uint32_t) * n) + Arena::padded(sizeof(float4) * 2 * (size_t)n) +
                      Arena::padded(sizeof(uint32_t) * (size_t)node_capacity) + Arena::padded(sizeof(CollapseCounters)) + Arena::padded(sizeof(GridBarrier));
        if (builder == FF_BUILDER_SAH) need += sah_scratch_bytes(n);
        CK(arena.reserve(need));
    }
    d_bounds = arena.take<int>(12);
    d_keys_a = arena.take<char>(key_bytes * n); d_keys_b = arena.take<char>(key_bytes * n);
    d_vals_a = arena.take<uint32_t>(n); d_vals_b = arena.take<uint32_t>(n);
    d_hist = arena.take<uint32_t>(sort_words);   // digit bases, tile counters, tile descriptors of the sort
    d_parent = arena.take<uint32_t>(2 * (size_t)n);
    d_flags = arena.take<uint32_t>(n);
    d_dp = arena.take<float4>(2 * (size_t)n);   // F table + decisions of the n - 1 internal binary nodes
    d_src = arena.take<uint32_t>(node_capacity);
    d_ctr = arena.take<CollapseCounters>(1);
    GridBarrier* d_bar = arena.take<GridBarrier>(1);
    if (!d_ctr || !d_bar) return cudaErrorMemoryAllocation;
    CK(cudaMemsetAsync(d_flags, 0, sizeof(uint32_t) * n, st));
    mark("reserve scratch");

    if (d_bounds_ready) {
        d_bounds = const_cast<int*>(d_bounds_ready);   // reduced by the bake kernel (k_bake_triangles)
    } else {
        bounds_init_kernel<<<1, 32, 0, st>>>(d_bounds);
        bounds_kernel<<<grid_for(n, 256, sm_count() * 8), 256, 0, st>>>(d_flat_tris, n, d_bounds);
        *launches += 2;
    }
    mark("bounds");
    for (int attempt = 0; attempt < 2; ++attempt) {
        if (wide_keys) {
            morton_kernel<uint64_t><<<grid_for(n, 256, sm_count() * 16), 256, 0, st>>>(d_flat_tris, n, d_bounds, (uint64_t*)d_keys_a, d_vals_a);
            CK(radix_sort_pairs<uint64_t>((uint64_t*)d_keys_a, d_vals_a, (uint64_t*)d_keys_b, d_vals_b, n, d_hist, launches, st));
        } else {
            morton_kernel<uint32_t><<<grid_for(n, 256, sm_count() * 16), 256, 0, st>>>(d_flat_tris, n, d_bounds, (uint32_t*)d_keys_a, d_vals_a);
            CK(radix_sort_pairs<uint32_t>((uint32_t*)d_keys_a, d_vals_a, (uint32_t*)d_keys_b, d_vals_b, n, d_hist, launches, st));
        }
        ++*launches;
        if (wide_keys || !may_widen || mb_env) break;
        uint32_t dups = 0;
        CK(cudaMemsetAsync(d_hist, 0, sizeof(uint32_t), st));
        count_duplicate_keys_kernel<<<grid_for(n, 256, sm_count() * 8), 256, 0, st>>>((const uint32_t*)d_keys_a, n, d_hist);
        CK(cudaMemcpyAsync(&dups, d_hist, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        ++*launches;
        if ((uint64_t)dups * 16u <= n) break;
        wide_keys = true;
    }
    out->morton_bits = wide_keys ? 63u : 30u;
    mark("morton codes + radix sort");
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
        if (wide_keys) hierarchy_kernel<uint64_t><<<grid_for(n - 1, 256, sm_count() * 16), 256, 0, st>>>((const uint64_t*)d_keys_a, (int)n, d_bvh2_children, d_parent);
        else hierarchy_kernel<uint32_t><<<grid_for(n - 1, 256, sm_count() * 16), 256, 0, st>>>((const uint32_t*)d_keys_a, (int)n, d_bvh2_children, d_parent);
        ++*launches;
        mark("karras hierarchy");
    }
    const char* sc_env = getenv("FF_SLOT_COST");  // development knob: leaf-or-open trade-off of the refit DP
    const float slot_cost = sc_env ? (float)atof(sc_env) : FF_SLOT_COST;
    const char* nc_env = getenv("FF_NODE_COST");
    const float node_cost = nc_env ? (float)atof(nc_env) : FF_NODE_COST;
    if (!getenv("FF_REFIT_GLOBAL")) {   // both builders number their nodes like Karras 2012 and tag the links that cross a block
        const size_t smem = sizeof(float4) * 6 * FF_TREELET + sizeof(uint32_t) * 3 * FF_TREELET + sizeof(int2) * FF_TREELET + sizeof(uint32_t) * (2 * FF_TREELET + 4) + sizeof(uint32_t) * 2 * FF_TREELET;   // + s_start: up to 2S - 1 starters
        static bool attr_set[64] = {false};
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev >= 0 && dev < 64 && !attr_set[dev]) {
            CK(cudaFuncSetAttribute(refit_treelet_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            attr_set[dev] = true;
        }
        // the sort's second buffers are free by now: starters (at most one per node finished in a block) and their count
        uint32_t* d_starters = d_order == d_vals_b ? d_vals_a : d_vals_b;   // (the SAH hierarchy leaves its order in the second buffer)
        uint32_t* d_starter_count = (uint32_t*)d_keys_b;
        CK(cudaMemsetAsync(d_starter_count, 0, sizeof(uint32_t), st));
        const uint32_t blocks = (n + FF_TREELET - 1) / FF_TREELET;
        refit_treelet_kernel<<<blocks, FF_TREELET_THREADS, smem, st>>>(d_flat_tris, d_order, (int)n, pad, d_bvh2_children, d_parent, d_flags, d_bvh2_lo,
                                                                       d_bvh2_hi, d_dp, slot_cost, node_cost, d_starters, d_starter_count);
        if (n > 1) {
            // ~ 2 log2(FF_TREELET) starters per block on a balanced hierarchy; the grid-stride loop covers a degenerate one
            refit_cross_kernel<<<grid_for((size_t)blocks * 24, 128, sm_count() * 64), 128, 0, st>>>((int)n, d_bvh2_children, d_parent, d_flags, d_bvh2_lo,
                                                                                                 d_bvh2_hi, d_dp, slot_cost, node_cost, d_starters, d_starter_count);
            ++*launches;
        }
    } else {
        refit_kernel<<<grid_for(n, 256, sm_count() * 16), 256, 0, st>>>(d_flat_tris, d_order, (int)n, pad, d_bvh2_children, d_parent, d_flags,
                                                                d_bvh2_lo, d_bvh2_hi, d_dp, slot_cost, node_cost);
    }
    ++*launches;
    mark("refit + collapse DP");

    const uint32_t root = 0u;  // internal node 0, or the single leaf (id n-1+0 = 0)
    CollapseCounters h_ctr;
    memset(&h_ctr, 0, sizeof(h_ctr));
    h_ctr.node_count = 1; h_ctr.level_begin = 0; h_ctr.level_end = 1;
    CK(cudaMemcpyAsync(d_ctr, &h_ctr, sizeof(h_ctr), cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(d_src, &root, sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    CK(cudaMemsetAsync(d_bar, 0, sizeof(GridBarrier), st));
    {
        // one launch for all levels; grid = what is co-resident on this device (the barrier needs every CTA running)
        static int blocks_per_sm[64] = {0};
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev < 0 || dev >= 64) dev = 0;
        if (blocks_per_sm[dev] == 0) {
            int b = 0;
            CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, collapse_persistent_kernel, 128, 0));
            blocks_per_sm[dev] = b > 0 ? b : 1;
        }
        const int grid = sm_count() * std::min(blocks_per_sm[dev], 8);
        int n_arg = (int)n;
        const float4* rl = d_bvh2_lo + root;
        const float4* rh = d_bvh2_hi + root;
        const int2* ch_arg = d_bvh2_children;
        const float4 *lo_arg = d_bvh2_lo, *hi_arg = d_bvh2_hi, *flat_arg = d_flat_tris, *dp_arg = d_dp;
        const uint32_t* ord_arg = d_order;
        void* args[] = {&d_src, &node_capacity, &n_arg, &ch_arg, &lo_arg, &hi_arg, &flat_arg, &ord_arg, &d_nodes_out, &d_tris_out, &d_bvh2_leaf_slot,
                        &d_ctr, &rl, &rh, &dp_arg, &d_bar};
        CK(cudaLaunchCooperativeKernel((const void*)collapse_persistent_kernel, dim3(grid), dim3(128), args, 0, st));
        ++*launches;
    }
    // the triangle gather needs leaf_slot[] complete, nothing from the host: queue it behind the collapse, then read the counters
    gather_tris_kernel<<<grid_for(n, 256, sm_count() * 16), 256, 0, st>>>(d_flat_tris, d_order, d_bvh2_leaf_slot, n, d_tris_out);
    ++*launches;
    CK(cudaMemcpyAsync(&h_ctr, d_ctr, sizeof(h_ctr), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    for (int k = 0; k < 3; ++k) { bounds_lo[k] = ordered_to_float(h_bounds[k]); bounds_hi[k] = ordered_to_float(h_bounds[3 + k]); }
    if (h_ctr.overflow) { CK(cudaErrorMemoryAllocation); }
    if (h_ctr.level_begin != h_ctr.level_end) return cudaErrorUnknown;
    const uint32_t depth = h_ctr.levels;
    mark("collapse to 8-wide");
    out->n_nodes = h_ctr.node_count;
    out->depth = depth;
    out->sah_cost = h_ctr.sah;
    return h_ctr.tri_count == n ? cudaSuccess : cudaErrorUnknown;
}

}  // namespace ff
