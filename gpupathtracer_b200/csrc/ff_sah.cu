// ff_sah.cu -- GPU binned-SAH builder (FF_BUILDER_SAH): a top-down binary hierarchy over the Morton-sorted triangles,
// built level by level on the device, in the same arrays the LBVH path produces (children / parent / sorted order),
// so the refit + leaf-or-open DP + collapse to the 8-wide compressed layout in ff_bvh.cu run unchanged after it.
//
// The reference has no acceleration structure at all (kernel.cu:133-155 scans every triangle for every ray); the north
// star asks for "a GPU LBVH/binned-SAH builder".  LBVH (ff_bvh.cu) is the fast builder, this is the quality builder.
//
// Layout of the work:
//   * a "segment" is a node's contiguous range [b, e) of the current triangle order.  Each level one kernel runs one
//     CTA per segment of more than 32 triangles: 16 bins x 3 axes in shared memory (centroid binning, shared atomics),
//     45 candidate planes evaluated by 45 threads, then a stable two-way partition of the segment into the other
//     order buffer (ballot/popc ranks), which also reduces the two children's centroid bounds for the next level.
//   * a segment of <= 32 triangles is finished by ONE WARP in registers: lane k holds triangle k's box; every lane
//     evaluates the sweep-SAH candidate "everything not after my centroid goes left" for the three axes (an exact
//     sweep without sorting), the warp takes the minimum, partitions by shuffle, and repeats on the sub-ranges from a
//     small stack until every leaf is one triangle.
//   * node numbering follows Karras 2012 (what hierarchy_kernel emits): the children of a node split after position
//     g are internal nodes g (left: its LAST position) and g+1 (right: its FIRST position), a one-triangle child at
//     position p is leaf n-1+p, the root is 0.  The numbering depends only on positions, so the tree is deterministic
//     although segments are appended to the next level's list with atomics.
//   * splits always go down to single triangles; whether <= 3 of them become one leaf child is decided afterwards by
//     the expected-cost DP in refit_kernel, exactly as for the LBVH tree.
#include <cfloat>
#include "ff_internal.h"

namespace ff {

#define SAH_BINS 16
#define SAH_SMALL 32          // segments up to this size are finished by one warp
#define SAH_BLOCK 256
#define SAH_MAX_LEVELS 96     // beyond this depth segments are halved by position (never reached on sane input)

struct SahSeg {
    uint32_t b, e, id, level;
    float clo[3], chi[3];     // bounds of the triangle-box centres (x2: lo+hi, never halved) of the segment
    uint32_t pad0, pad1;
};

static inline int sah_grid_for(size_t n, int block, int max_blocks) {
    size_t g = (n + block - 1) / block;
    if (g < 1) g = 1;
    if (g > (size_t)max_blocks) g = max_blocks;
    return (int)g;
}

// Triangle boxes in the Morton-sorted order k (so a spatially compact segment gathers from a compact address range).
__global__ void sah_boxes_kernel(const float4* __restrict__ flat_tris, const uint32_t* __restrict__ sorted_idx, uint32_t n,
                                 float4* __restrict__ box_lo, float4* __restrict__ box_hi, uint32_t* __restrict__ order) {
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
        const uint32_t src = sorted_idx[k];
        const float4 a = flat_tris[3 * (size_t)src], b = flat_tris[3 * (size_t)src + 1], c = flat_tris[3 * (size_t)src + 2];
        box_lo[k] = make_float4(fminf(fminf(a.x, b.x), c.x), fminf(fminf(a.y, b.y), c.y), fminf(fminf(a.z, b.z), c.z), 0.f);
        box_hi[k] = make_float4(fmaxf(fmaxf(a.x, b.x), c.x), fmaxf(fmaxf(a.y, b.y), c.y), fmaxf(fmaxf(a.z, b.z), c.z), 0.f);
        order[k] = k;
    }
}

__device__ __forceinline__ int sah_f2o(float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__device__ __forceinline__ float sah_o2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }
__device__ __forceinline__ float sah_half_area(float lx, float ly, float lz, float hx, float hy, float hz) {
    const float dx = hx - lx, dy = hy - ly, dz = hz - lz;
    return dx * dy + dy * dz + dz * dx;
}

// Emits the two children of node `id` = segment [b, e) split at m (absolute positions) and returns their node ids.
__device__ __forceinline__ void sah_link(uint32_t n, uint32_t id, uint32_t b, uint32_t m, uint32_t e, int2* children, uint32_t* parent,
                                         uint32_t& lid, uint32_t& rid) {
    lid = (m - b == 1u) ? (n - 1u + b) : (m - 1u);
    rid = (e - m == 1u) ? (n - 1u + m) : m;
    children[id] = make_int2((int)lid, (int)rid);
    const uint32_t tag = id | ((b / FF_TREELET) != ((e - 1u) / FF_TREELET) ? FF_PARENT_CROSS : 0u);   // ff_internal.h: this node's range crosses a refit block
    parent[lid] = tag;
    parent[rid] = tag;
}

// ------------------------------------------------------------------ one CTA per segment of more than SAH_SMALL triangles
#define SAH_BIG 8192           // segments above this are processed by many CTAs (see the big-segment kernels below)
#define SAH_CHUNK 2048
struct SahLists {
    SahSeg* next;          // segments of the next level with SAH_SMALL < size <= SAH_BIG (one CTA each)
    SahSeg* small;         // segments handed to the warp kernel
    SahSeg* big;           // segments of the next level with more than SAH_BIG triangles
    uint32_t* counters;    // [0] entries in next, [1] entries in small, [2] deepest level seen, [3] entries in big, [4] chunks of big
};
// Child [cb, ce) of a split goes to the list that fits its size; a single triangle is final.
__device__ __forceinline__ void sah_append_child(const SahLists& lists, uint32_t cb, uint32_t ce, uint32_t cid, uint32_t level,
                                                 const int* cbounds /*[6] ordered*/, const uint32_t* idx_out, uint32_t* final_order) {
    const uint32_t cc = ce - cb;
    if (cc == 1u) { final_order[cb] = idx_out[cb]; return; }
    SahSeg c;
    c.b = cb; c.e = ce; c.id = cid; c.level = level; c.pad0 = c.pad1 = 0;
    for (int k = 0; k < 3; ++k) { c.clo[k] = sah_o2f(cbounds[k]); c.chi[k] = sah_o2f(cbounds[3 + k]); }
    if (cc <= SAH_SMALL) lists.small[atomicAdd(lists.counters + 1, 1u)] = c;
    else if (cc <= SAH_BIG) lists.next[atomicAdd(lists.counters + 0, 1u)] = c;
    else {
        c.pad1 = (cc + SAH_CHUNK - 1) / SAH_CHUNK;
        c.pad0 = atomicAdd(lists.counters + 4, c.pad1);
        lists.big[atomicAdd(lists.counters + 3, 1u)] = c;
    }
    atomicMax(lists.counters + 2, level);
}

__device__ __forceinline__ int sah_bin_of(float c2, float lo2, float scale) {
    const int b = (int)((c2 - lo2) * scale);
    return min(max(b, 0), SAH_BINS - 1);
}

// One warp adds (up to) 32 triangle boxes to a shared [3][SAH_BINS][7] bin table.  Morton-sorted neighbours mostly share a bin,
// and 32 lanes doing 7 shared atomics each on the same words serialise (the binning passes were 25 of the 47 ms of a
// 20 M-triangle build at 0.56 TB/s, profiles/r03a_launches_sah_20000000.csv).  When the whole warp agrees on the bin of an axis
// (one match.all) the boxes are reduced by a shuffle butterfly and lane 0 does the 7 atomics; otherwise every lane does its
// own.  min / max / + commute, so the table -- and the tree -- is the same.  (Grouping the lanes of each bin with ballots and
// reducing with redux.sync over the group masks was measured first: 47 -> 77 ms, redux over partial masks is slow.)
__device__ __forceinline__ void sah_bin_warp(int (*bins)[SAH_BINS][7], bool live, const int bin_of[3], const int ol[3], const int oh[3]) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const int b = bin_of[a];
        int same = 0;
        __match_all_sync(0xffffffffu, live ? b : -1 - lane, &same);
        if (same) {
            int v[6] = {ol[0], ol[1], ol[2], oh[0], oh[1], oh[2]};
#pragma unroll
            for (int off = 16; off; off >>= 1) {
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    v[k] = min(v[k], __shfl_xor_sync(0xffffffffu, v[k], off));
                    v[3 + k] = max(v[3 + k], __shfl_xor_sync(0xffffffffu, v[3 + k], off));
                }
            }
            if (lane == 0) {
                int* bin = bins[a][b];
                atomicMin(bin + 0, v[0]); atomicMin(bin + 1, v[1]); atomicMin(bin + 2, v[2]);
                atomicMax(bin + 3, v[3]); atomicMax(bin + 4, v[4]); atomicMax(bin + 5, v[5]);
                atomicAdd(bin + 6, 32);
            }
        } else if (live) {
            int* bin = bins[a][b];
            atomicMin(bin + 0, ol[0]); atomicMin(bin + 1, ol[1]); atomicMin(bin + 2, ol[2]);
            atomicMax(bin + 3, oh[0]); atomicMax(bin + 4, oh[1]); atomicMax(bin + 5, oh[2]);
            atomicAdd(bin + 6, 1);
        }
    }
}

// Candidate plane t (axis t / 15, left = bins [0, t % 15]) of a [3][SAH_BINS][7] bin table: SAH cost and left count.
__device__ __forceinline__ void sah_eval_plane(const int* bins, int t, bool forced, float& cost_out, uint32_t& nl_out) {
    const int a = t / (SAH_BINS - 1), plane = t % (SAH_BINS - 1);
    float l[6] = {FLT_MAX, FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX}, r[6] = {FLT_MAX, FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX};
    uint32_t nl = 0, nr = 0;
    for (int k = 0; k < SAH_BINS; ++k) {
        const int* bin = bins + (a * SAH_BINS + k) * 7;
        if (bin[6] == 0) continue;
        float* d = k <= plane ? l : r;
        d[0] = fminf(d[0], sah_o2f(bin[0])); d[1] = fminf(d[1], sah_o2f(bin[1])); d[2] = fminf(d[2], sah_o2f(bin[2]));
        d[3] = fmaxf(d[3], sah_o2f(bin[3])); d[4] = fmaxf(d[4], sah_o2f(bin[4])); d[5] = fmaxf(d[5], sah_o2f(bin[5]));
        if (k <= plane) nl += (uint32_t)bin[6]; else nr += (uint32_t)bin[6];
    }
    float cost = FLT_MAX;
    if (nl && nr && !forced)
        cost = sah_half_area(l[0], l[1], l[2], l[3], l[4], l[5]) * (float)nl + sah_half_area(r[0], r[1], r[2], r[3], r[4], r[5]) * (float)nr;
    if (!(cost < FLT_MAX)) cost = FLT_MAX;  // NaN boxes never win
    cost_out = cost;
    nl_out = nl;
}

// BLOCK threads per segment: 256 while the segments are thousands of triangles, 128 / 64 for the last levels, where a CTA's life
// is a fixed chain of barriers and the segments per SM are what counts (the three launches over 40..150-triangle segments were
// 7.9 of the 47 ms of a 20 M-triangle build, profiles/r03a).  The partition is stable and the bins commute: same tree for any BLOCK.
template <int BLOCK>
__global__ void __launch_bounds__(BLOCK) sah_level_kernel(const SahSeg* __restrict__ segs, const uint32_t* __restrict__ idx_in,
                                                               uint32_t* __restrict__ idx_out, const float4* __restrict__ box_lo,
                                                               const float4* __restrict__ box_hi, SahLists lists, int2* children,
                                                               uint32_t* parent, uint32_t* __restrict__ final_order, uint32_t n) {
    __shared__ int s_bins[3][SAH_BINS][7];      // lo.xyz, hi.xyz (ordered ints), count
    __shared__ float s_cost[3 * (SAH_BINS - 1)];
    __shared__ uint32_t s_nleft[3 * (SAH_BINS - 1)];
    __shared__ int s_axis, s_plane;
    __shared__ uint32_t s_nl;
    __shared__ uint32_t s_warp_l[BLOCK / 32], s_warp_r[BLOCK / 32];
    __shared__ uint32_t s_base_l, s_base_r;
    __shared__ int s_cb[2][6];                  // children's centre bounds (ordered ints): [side][lo.xyz, hi.xyz]

    const SahSeg s = segs[blockIdx.x];
    const uint32_t b = s.b, e = s.e, cnt = e - b;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    for (int k = tid; k < 3 * SAH_BINS * 7; k += BLOCK) {
        const int f = k % 7;
        (&s_bins[0][0][0])[k] = f < 3 ? sah_f2o(FLT_MAX) : (f < 6 ? sah_f2o(-FLT_MAX) : 0);
    }
    if (tid < 12) (&s_cb[0][0])[tid] = (tid % 6) < 3 ? sah_f2o(FLT_MAX) : sah_f2o(-FLT_MAX);
    float scale[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float ext = s.chi[a] - s.clo[a];
        scale[a] = ext > 0.f ? (float)SAH_BINS / ext : 0.f;
    }
    __syncthreads();

    // pass A: bin the triangle boxes by centre
    const bool forced = s.level >= SAH_MAX_LEVELS;
    if (!forced) {
        for (uint32_t base = b; base < e; base += BLOCK) {   // warp-uniform trip count: sah_bin_warp votes
            const uint32_t i = base + tid;
            const bool live = i < e;
            float4 lo = make_float4(0.f, 0.f, 0.f, 0.f), hi = lo;
            if (live) { const uint32_t p = idx_in[i]; lo = __ldg(box_lo + p); hi = __ldg(box_hi + p); }
            const float c2[3] = {lo.x + hi.x, lo.y + hi.y, lo.z + hi.z};
            const int ol[3] = {sah_f2o(lo.x), sah_f2o(lo.y), sah_f2o(lo.z)}, oh[3] = {sah_f2o(hi.x), sah_f2o(hi.y), sah_f2o(hi.z)};
            const int bo[3] = {sah_bin_of(c2[0], s.clo[0], scale[0]), sah_bin_of(c2[1], s.clo[1], scale[1]), sah_bin_of(c2[2], s.clo[2], scale[2])};
            sah_bin_warp(s_bins, live, bo, ol, oh);
        }
    }
    __syncthreads();

    // 45 candidate planes: thread t -> axis t / 15, left = bins [0, plane]
    if (tid < 3 * (SAH_BINS - 1)) sah_eval_plane(&s_bins[0][0][0], tid, forced, s_cost[tid], s_nleft[tid]);
    __syncthreads();
    if (tid == 0) {
        int best = -1;
        float bc = FLT_MAX;
        for (int k = 0; k < 3 * (SAH_BINS - 1); ++k)
            if (s_cost[k] < bc) { bc = s_cost[k]; best = k; }   // first minimum: deterministic
        if (best >= 0) { s_axis = best / (SAH_BINS - 1); s_plane = best % (SAH_BINS - 1); s_nl = s_nleft[best]; }
        else { s_axis = -1; s_plane = 0; s_nl = cnt / 2; }      // no plane separates the centres: halve by position
        s_base_l = b;
        s_base_r = b + s_nl;
    }
    __syncthreads();
    const int axis = s_axis, plane = s_plane;
    const uint32_t nl = s_nl;

    // pass B: stable partition into idx_out, and the centre bounds of the two sides
    float cl[6] = {FLT_MAX, FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX}, cr[6] = {FLT_MAX, FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (uint32_t base = b; base < e; base += BLOCK) {
        const uint32_t i = base + tid;
        const bool live = i < e;
        uint32_t p = 0;
        bool left = false;
        if (live) {
            p = idx_in[i];
            const float4 lo = __ldg(box_lo + p), hi = __ldg(box_hi + p);
            const float c2[3] = {lo.x + hi.x, lo.y + hi.y, lo.z + hi.z};
            left = axis < 0 ? (i - b) < nl : sah_bin_of(c2[axis], s.clo[axis], scale[axis]) <= plane;
            float* d = left ? cl : cr;
            d[0] = fminf(d[0], c2[0]); d[1] = fminf(d[1], c2[1]); d[2] = fminf(d[2], c2[2]);
            d[3] = fmaxf(d[3], c2[0]); d[4] = fmaxf(d[4], c2[1]); d[5] = fmaxf(d[5], c2[2]);
        }
        const unsigned ml = __ballot_sync(0xffffffffu, live && left), mr = __ballot_sync(0xffffffffu, live && !left);
        if (lane == 0) { s_warp_l[warp] = __popc(ml); s_warp_r[warp] = __popc(mr); }
        __syncthreads();
        uint32_t off_l = 0, off_r = 0, tot_l = 0, tot_r = 0;
#pragma unroll
        for (int w = 0; w < BLOCK / 32; ++w) {
            if (w < warp) { off_l += s_warp_l[w]; off_r += s_warp_r[w]; }
            tot_l += s_warp_l[w]; tot_r += s_warp_r[w];
        }
        if (live) {
            const unsigned lt = (1u << lane) - 1u;
            const uint32_t dst = left ? s_base_l + off_l + __popc(ml & lt) : s_base_r + off_r + __popc(mr & lt);
            idx_out[dst] = p;
        }
        __syncthreads();
        if (tid == 0) { s_base_l += tot_l; s_base_r += tot_r; }
        __syncthreads();
    }
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        float vl = cl[k], vr = cr[k];
        for (int off = 16; off; off >>= 1) {
            const float ol = __shfl_xor_sync(0xffffffffu, vl, off), orr = __shfl_xor_sync(0xffffffffu, vr, off);
            vl = k < 3 ? fminf(vl, ol) : fmaxf(vl, ol);
            vr = k < 3 ? fminf(vr, orr) : fmaxf(vr, orr);
        }
        if (lane == 0) {
            if (k < 3) { atomicMin(&s_cb[0][k], sah_f2o(vl)); atomicMin(&s_cb[1][k], sah_f2o(vr)); }
            else { atomicMax(&s_cb[0][k], sah_f2o(vl)); atomicMax(&s_cb[1][k], sah_f2o(vr)); }
        }
    }
    __syncthreads();

    if (tid < 2) {  // thread 0 emits the left child, thread 1 the right one
        const uint32_t m = b + nl;
        uint32_t lid, rid;
        if (tid == 0) sah_link(n, s.id, b, m, e, children, parent, lid, rid);
        else { lid = (m - b == 1u) ? (n - 1u + b) : (m - 1u); rid = (e - m == 1u) ? (n - 1u + m) : m; }
        sah_append_child(lists, tid == 0 ? b : m, tid == 0 ? m : e, tid == 0 ? lid : rid, s.level + 1, s_cb[tid], idx_out, final_order);
    }
}

// ------------------------------------------------------------------ segments of more than SAH_BIG triangles: many CTAs each
// One CTA per segment would leave the top levels (one segment of N triangles, then two of N/2, ...) on a handful of
// SMs.  A big segment is cut into chunks of SAH_CHUNK triangles (SahSeg::pad0 = first chunk, pad1 = chunk count) and a
// level takes five launches over all big segments of the level:
//   prepare  clears the segment's global bin table, maps its chunks to it
//   bin      one CTA per chunk: shared-memory bins, flushed to the segment's global table with atomics, and the
//            chunk's own per-axis bin counts kept (they give the chunk's left count for ANY plane, so no third pass)
//   split    one CTA per segment: the 45 planes, then an exclusive scan of the chunks' left counts -> every chunk knows
//            where its left and right triangles go (stable, deterministic)
//   scatter  one CTA per chunk: stable partition into the other order buffer; children's centre bounds by atomics
//   emit     one thread per segment: links the node, appends the children to the big / CTA / warp lists
struct SahBigSplit {
    int axis, plane;
    uint32_t nl, pad;
    int cb[2][6];          // children's centre bounds (ordered ints)
};
struct SahBig {
    const SahSeg* segs;        // big segments of this level
    uint32_t n_segs;
    uint32_t* chunk_seg;       // chunk -> index into segs
    int* gbins;                // [n_segs][3][SAH_BINS][7]
    uint32_t* chunk_hist;      // [chunks][3][SAH_BINS]
    uint2* chunk_off;          // [chunks] {lefts before this chunk, rights before this chunk} within the segment
    SahBigSplit* split;        // [n_segs]
};

__global__ void __launch_bounds__(SAH_BLOCK) sah_big_prepare_kernel(SahBig B) {
    const SahSeg s = B.segs[blockIdx.x];
    int* g = B.gbins + (size_t)blockIdx.x * 3 * SAH_BINS * 7;
    for (int k = threadIdx.x; k < 3 * SAH_BINS * 7; k += SAH_BLOCK) {
        const int f = k % 7;
        g[k] = f < 3 ? sah_f2o(FLT_MAX) : (f < 6 ? sah_f2o(-FLT_MAX) : 0);
    }
    for (uint32_t c = threadIdx.x; c < s.pad1; c += SAH_BLOCK) B.chunk_seg[s.pad0 + c] = blockIdx.x;
    if (threadIdx.x < 12) (&B.split[blockIdx.x].cb[0][0])[threadIdx.x] = (threadIdx.x % 6) < 3 ? sah_f2o(FLT_MAX) : sah_f2o(-FLT_MAX);
}

__global__ void __launch_bounds__(SAH_BLOCK) sah_big_bin_kernel(SahBig B, const uint32_t* __restrict__ idx_in,
                                                                 const float4* __restrict__ box_lo, const float4* __restrict__ box_hi) {
    __shared__ int s_bins[3][SAH_BINS][7];
    const uint32_t si = B.chunk_seg[blockIdx.x];
    const SahSeg s = B.segs[si];
    const uint32_t cb = s.b + (blockIdx.x - s.pad0) * SAH_CHUNK, ce = min(cb + SAH_CHUNK, s.e);
    const int tid = threadIdx.x;
    for (int k = tid; k < 3 * SAH_BINS * 7; k += SAH_BLOCK) {
        const int f = k % 7;
        (&s_bins[0][0][0])[k] = f < 3 ? sah_f2o(FLT_MAX) : (f < 6 ? sah_f2o(-FLT_MAX) : 0);
    }
    float scale[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float ext = s.chi[a] - s.clo[a];
        scale[a] = ext > 0.f ? (float)SAH_BINS / ext : 0.f;
    }
    __syncthreads();
    for (uint32_t base = cb; base < ce; base += SAH_BLOCK) {   // warp-uniform trip count: sah_bin_warp votes
        const uint32_t i = base + tid;
        const bool live = i < ce;
        float4 lo = make_float4(0.f, 0.f, 0.f, 0.f), hi = lo;
        if (live) { const uint32_t p = idx_in[i]; lo = __ldg(box_lo + p); hi = __ldg(box_hi + p); }
        const float c2[3] = {lo.x + hi.x, lo.y + hi.y, lo.z + hi.z};
        const int ol[3] = {sah_f2o(lo.x), sah_f2o(lo.y), sah_f2o(lo.z)}, oh[3] = {sah_f2o(hi.x), sah_f2o(hi.y), sah_f2o(hi.z)};
        const int bo[3] = {sah_bin_of(c2[0], s.clo[0], scale[0]), sah_bin_of(c2[1], s.clo[1], scale[1]), sah_bin_of(c2[2], s.clo[2], scale[2])};
        sah_bin_warp(s_bins, live, bo, ol, oh);
    }
    __syncthreads();
    int* g = B.gbins + (size_t)si * 3 * SAH_BINS * 7;
    for (int k = tid; k < 3 * SAH_BINS; k += SAH_BLOCK) {
        const int* bin = &s_bins[0][0][0] + k * 7;
        B.chunk_hist[(size_t)blockIdx.x * 3 * SAH_BINS + k] = (uint32_t)bin[6];
        if (bin[6]) {
            int* d = g + k * 7;
            atomicMin(d + 0, bin[0]); atomicMin(d + 1, bin[1]); atomicMin(d + 2, bin[2]);
            atomicMax(d + 3, bin[3]); atomicMax(d + 4, bin[4]); atomicMax(d + 5, bin[5]);
            atomicAdd(d + 6, bin[6]);
        }
    }
}

__global__ void __launch_bounds__(SAH_BLOCK) sah_big_split_kernel(SahBig B) {
    __shared__ float s_cost[3 * (SAH_BINS - 1)];
    __shared__ uint32_t s_nleft[3 * (SAH_BINS - 1)];
    __shared__ int s_axis, s_plane;
    __shared__ uint32_t s_nl, s_carry;
    __shared__ uint32_t s_warp[SAH_BLOCK / 32];
    const SahSeg s = B.segs[blockIdx.x];
    const uint32_t cnt = s.e - s.b;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool forced = s.level >= SAH_MAX_LEVELS;
    if (tid < 3 * (SAH_BINS - 1)) sah_eval_plane(B.gbins + (size_t)blockIdx.x * 3 * SAH_BINS * 7, tid, forced, s_cost[tid], s_nleft[tid]);
    __syncthreads();
    if (tid == 0) {
        int best = -1;
        float bc = FLT_MAX;
        for (int k = 0; k < 3 * (SAH_BINS - 1); ++k)
            if (s_cost[k] < bc) { bc = s_cost[k]; best = k; }
        if (best >= 0) { s_axis = best / (SAH_BINS - 1); s_plane = best % (SAH_BINS - 1); s_nl = s_nleft[best]; }
        else { s_axis = -1; s_plane = 0; s_nl = cnt / 2; }
        s_carry = 0;
        SahBigSplit& sp = B.split[blockIdx.x];
        sp.axis = s_axis; sp.plane = s_plane; sp.nl = s_nl; sp.pad = 0;
    }
    __syncthreads();
    const int axis = s_axis, plane = s_plane;
    const uint32_t nl = s_nl;
    // exclusive scan of the chunks' left counts, SAH_BLOCK chunks per round
    for (uint32_t c0 = 0; c0 < s.pad1; c0 += SAH_BLOCK) {
        const uint32_t c = c0 + tid;
        uint32_t mine = 0;
        if (c < s.pad1) {
            const uint32_t first = c * SAH_CHUNK, size = min((uint32_t)SAH_CHUNK, cnt - first);
            if (axis < 0) mine = first >= nl ? 0u : min(size, nl - first);
            else {
                const uint32_t* h = B.chunk_hist + (size_t)(s.pad0 + c) * 3 * SAH_BINS + axis * SAH_BINS;
                for (int k = 0; k <= plane; ++k) mine += h[k];
            }
        }
        uint32_t incl = mine;
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, off);
            if (lane >= off) incl += t;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        uint32_t before = s_carry;
        for (int w = 0; w < warp; ++w) before += s_warp[w];
        const uint32_t left_before = before + incl - mine;
        if (c < s.pad1) B.chunk_off[s.pad0 + c] = make_uint2(left_before, c * SAH_CHUNK - left_before);
        __syncthreads();
        if (tid == SAH_BLOCK - 1) s_carry = left_before + mine;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(SAH_BLOCK) sah_big_scatter_kernel(SahBig B, const uint32_t* __restrict__ idx_in, uint32_t* __restrict__ idx_out,
                                                                     const float4* __restrict__ box_lo, const float4* __restrict__ box_hi) {
    __shared__ uint32_t s_warp_l[SAH_BLOCK / 32], s_warp_r[SAH_BLOCK / 32];
    __shared__ uint32_t s_base_l, s_base_r;
    __shared__ int s_cb[12];   // the chunk's contribution to the children's centre bounds (ordered ints)
    const uint32_t si = B.chunk_seg[blockIdx.x];
    const SahSeg s = B.segs[si];
    const SahBigSplit sp = B.split[si];
    const uint32_t cb = s.b + (blockIdx.x - s.pad0) * SAH_CHUNK, ce = min(cb + SAH_CHUNK, s.e);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int axis = sp.axis, plane = sp.plane;
    const uint32_t nl = sp.nl;
    const float ext = axis >= 0 ? s.chi[axis] - s.clo[axis] : 0.f;
    const float scale = ext > 0.f ? (float)SAH_BINS / ext : 0.f;
    if (tid == 0) {
        const uint2 off = B.chunk_off[blockIdx.x];
        s_base_l = s.b + off.x;
        s_base_r = s.b + nl + off.y;
    }
    if (tid < 12) s_cb[tid] = (tid % 6) < 3 ? sah_f2o(FLT_MAX) : sah_f2o(-FLT_MAX);
    __syncthreads();
    float cl[6] = {FLT_MAX, FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX}, cr[6] = {FLT_MAX, FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (uint32_t base = cb; base < ce; base += SAH_BLOCK) {
        const uint32_t i = base + tid;
        const bool live = i < ce;
        uint32_t p = 0;
        bool left = false;
        if (live) {
            p = idx_in[i];
            const float4 lo = __ldg(box_lo + p), hi = __ldg(box_hi + p);
            const float c2[3] = {lo.x + hi.x, lo.y + hi.y, lo.z + hi.z};
            left = axis < 0 ? (i - s.b) < nl : sah_bin_of(axis == 0 ? c2[0] : (axis == 1 ? c2[1] : c2[2]), s.clo[axis], scale) <= plane;
            float* d = left ? cl : cr;
            d[0] = fminf(d[0], c2[0]); d[1] = fminf(d[1], c2[1]); d[2] = fminf(d[2], c2[2]);
            d[3] = fmaxf(d[3], c2[0]); d[4] = fmaxf(d[4], c2[1]); d[5] = fmaxf(d[5], c2[2]);
        }
        const unsigned ml = __ballot_sync(0xffffffffu, live && left), mr = __ballot_sync(0xffffffffu, live && !left);
        if (lane == 0) { s_warp_l[warp] = __popc(ml); s_warp_r[warp] = __popc(mr); }
        __syncthreads();
        uint32_t off_l = 0, off_r = 0, tot_l = 0, tot_r = 0;
#pragma unroll
        for (int w = 0; w < SAH_BLOCK / 32; ++w) {
            if (w < warp) { off_l += s_warp_l[w]; off_r += s_warp_r[w]; }
            tot_l += s_warp_l[w]; tot_r += s_warp_r[w];
        }
        if (live) {
            const unsigned lt = (1u << lane) - 1u;
            idx_out[left ? s_base_l + off_l + __popc(ml & lt) : s_base_r + off_r + __popc(mr & lt)] = p;
        }
        __syncthreads();
        if (tid == 0) { s_base_l += tot_l; s_base_r += tot_r; }
        __syncthreads();
    }
    int* gcb = &B.split[si].cb[0][0];
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        float vl = cl[k], vr = cr[k];
        for (int off = 16; off; off >>= 1) {
            const float ol = __shfl_xor_sync(0xffffffffu, vl, off), orr = __shfl_xor_sync(0xffffffffu, vr, off);
            vl = k < 3 ? fminf(vl, ol) : fmaxf(vl, ol);
            vr = k < 3 ? fminf(vr, orr) : fmaxf(vr, orr);
        }
        if (lane == 0) {   // per CTA first: at the top levels every chunk of the scene lands on the same 12 words (937 k global atomics at 20 M)
            if (k < 3) { atomicMin(s_cb + k, sah_f2o(vl)); atomicMin(s_cb + 6 + k, sah_f2o(vr)); }
            else { atomicMax(s_cb + k, sah_f2o(vl)); atomicMax(s_cb + 6 + k, sah_f2o(vr)); }
        }
    }
    __syncthreads();
    if (tid < 12) {
        if ((tid % 6) < 3) atomicMin(gcb + tid, s_cb[tid]); else atomicMax(gcb + tid, s_cb[tid]);
    }
}

__global__ void sah_big_emit_kernel(SahBig B, SahLists lists, const uint32_t* __restrict__ idx_out, int2* children, uint32_t* parent,
                                    uint32_t* __restrict__ final_order, uint32_t n) {
    const uint32_t si = blockIdx.x * blockDim.x + threadIdx.x;
    if (si >= B.n_segs) return;
    const SahSeg s = B.segs[si];
    const SahBigSplit sp = B.split[si];
    const uint32_t m = s.b + sp.nl;
    uint32_t lid, rid;
    sah_link(n, s.id, s.b, m, s.e, children, parent, lid, rid);
    sah_append_child(lists, s.b, m, lid, s.level + 1, sp.cb[0], idx_out, final_order);
    sah_append_child(lists, m, s.e, rid, s.level + 1, sp.cb[1], idx_out, final_order);
}

// ------------------------------------------------------------------ one warp per segment of <= SAH_SMALL triangles
// Lane k holds the triangle at position b + k.  Exact sweep SAH without sorting: lane j's candidate sends every
// triangle whose (centre, lane) is not after (centre_j, j) to the left, which enumerates the prefixes of the order
// sorted by centre.  The cheapest (cost, axis, lane) wins; the sub-range is partitioned stably through shared memory.
// Level-synchronous: every lane carries the sub-range [rb, re) it currently belongs to, and ALL sub-ranges of a level are
// evaluated and partitioned together (a lane meets the lanes of its own range through shuffles with a per-lane source).  The
// first version took the sub-ranges one at a time from a stack -- 31 serial steps and 160 sweep iterations for a 32-triangle
// segment instead of 5 and 62 -- and was 14.9 of the 47 ms of a 20 M-triangle build (profiles/r03a).  Same candidates, same
// tie-breaks, same tree.
#define SAH_WARPS 4
__global__ void __launch_bounds__(SAH_WARPS * 32) sah_small_kernel(const SahSeg* __restrict__ segs, uint32_t n_segs,
                                                                   const uint32_t* __restrict__ idx_in, const float4* __restrict__ box_lo,
                                                                   const float4* __restrict__ box_hi, int2* children, uint32_t* parent,
                                                                   uint32_t* __restrict__ final_order, uint32_t n) {
    __shared__ float s_box[SAH_WARPS][6][32];
    __shared__ uint32_t s_prim[SAH_WARPS][32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t si = blockIdx.x * SAH_WARPS + warp;
    if (si >= n_segs) return;
    const SahSeg s = segs[si];
    const uint32_t cnt = s.e - s.b;
    uint32_t prim = 0;
    float lo[3] = {0.f, 0.f, 0.f}, hi[3] = {0.f, 0.f, 0.f};
    if ((uint32_t)lane < cnt) {
        prim = idx_in[s.b + lane];
        const float4 l = __ldg(box_lo + prim), h = __ldg(box_hi + prim);
        lo[0] = l.x; lo[1] = l.y; lo[2] = l.z; hi[0] = h.x; hi[1] = h.y; hi[2] = h.z;
    }
    uint32_t rb = 0, re = cnt, id = s.id;                     // the lane's current sub-range (lanes) and its node
    bool active = (uint32_t)lane < cnt && cnt >= 2u;          // ... which still has to be split
    const unsigned lt = (1u << lane) - 1u;
    while (__any_sync(0xffffffffu, active)) {
        const uint32_t rc = re - rb;
        const uint32_t maxlen = __reduce_max_sync(0xffffffffu, active ? rc : 0u);
        const float cj[3] = {lo[0] + hi[0], lo[1] + hi[1], lo[2] + hi[2]};
        float L[3][6], R[3][6];
        uint32_t nl[3] = {0, 0, 0};
#pragma unroll
        for (int a = 0; a < 3; ++a)
#pragma unroll
            for (int k = 0; k < 6; ++k) { L[a][k] = k < 3 ? FLT_MAX : -FLT_MAX; R[a][k] = k < 3 ? FLT_MAX : -FLT_MAX; }
        for (uint32_t t = 0; t < maxlen; ++t) {
            const bool valid = active && t < rc;
            const uint32_t k = valid ? rb + t : (uint32_t)lane;   // the t-th lane of my range
            float kl[3], kh[3];
#pragma unroll
            for (int a = 0; a < 3; ++a) { kl[a] = __shfl_sync(0xffffffffu, lo[a], (int)k); kh[a] = __shfl_sync(0xffffffffu, hi[a], (int)k); }
            if (valid) {
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    const float ck = kl[a] + kh[a];
                    const bool left = ck < cj[a] || (ck == cj[a] && k <= (uint32_t)lane);
                    if (left) {
                        ++nl[a];
#pragma unroll
                        for (int q = 0; q < 3; ++q) { L[a][q] = fminf(L[a][q], kl[q]); L[a][3 + q] = fmaxf(L[a][3 + q], kh[q]); }
                    } else {
#pragma unroll
                        for (int q = 0; q < 3; ++q) { R[a][q] = fminf(R[a][q], kl[q]); R[a][3 + q] = fmaxf(R[a][3 + q], kh[q]); }
                    }
                }
            }
        }
        float best = FLT_MAX;
        uint32_t code = 0xffffffffu;  // axis * 32 + lane of the winning candidate
        if (active) {
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                if (nl[a] == 0 || nl[a] >= rc) continue;
                const float c = sah_half_area(L[a][0], L[a][1], L[a][2], L[a][3], L[a][4], L[a][5]) * (float)nl[a] +
                                sah_half_area(R[a][0], R[a][1], R[a][2], R[a][3], R[a][4], R[a][5]) * (float)(rc - nl[a]);
                if (c < best) { best = c; code = (uint32_t)a * 32u + (uint32_t)lane; }
            }
        }
        // the range's cheapest candidate (ties: smallest code), by a sweep over the range like the one above
        float rbest = FLT_MAX;
        uint32_t rcode = 0xffffffffu;
        for (uint32_t t = 0; t < maxlen; ++t) {
            const bool valid = active && t < rc;
            const int k = valid ? (int)(rb + t) : lane;
            const float ob = __shfl_sync(0xffffffffu, best, k);
            const uint32_t oc = __shfl_sync(0xffffffffu, code, k);
            if (valid && (ob < rbest || (ob == rbest && oc < rcode))) { rbest = ob; rcode = oc; }
        }
        bool left;
        {
            const int a = rcode != 0xffffffffu ? (int)(rcode >> 5) : 0, j = rcode != 0xffffffffu ? (int)(rcode & 31u) : lane;
            const float mine = a == 0 ? cj[0] : (a == 1 ? cj[1] : cj[2]);
            const float cjw = __shfl_sync(0xffffffffu, mine, j);
            if (rcode != 0xffffffffu) left = mine < cjw || (mine == cjw && lane <= j);
            else left = (uint32_t)lane < rb + rc / 2;   // NaN boxes: halve by position
        }
        const unsigned ml = __ballot_sync(0xffffffffu, active && left), mr = __ballot_sync(0xffffffffu, active && !left);
        const unsigned range = (re >= 32u ? 0xffffffffu : (1u << re) - 1u) & ~((1u << rb) - 1u);
        const uint32_t nleft = __popc(ml & range);
        if (active) {
            const uint32_t dst = left ? rb + __popc(ml & range & lt) : rb + nleft + __popc(mr & range & lt);
#pragma unroll
            for (int q = 0; q < 3; ++q) { s_box[warp][q][dst] = lo[q]; s_box[warp][3 + q][dst] = hi[q]; }
            s_prim[warp][dst] = prim;
        }
        __syncwarp();
        if (active) {
#pragma unroll
            for (int q = 0; q < 3; ++q) { lo[q] = s_box[warp][q][lane]; hi[q] = s_box[warp][3 + q][lane]; }
            prim = s_prim[warp][lane];
            const uint32_t rm = rb + nleft;
            uint32_t lid, rid;
            if ((uint32_t)lane == rb) sah_link(n, id, s.b + rb, s.b + rm, s.b + re, children, parent, lid, rid);   // one lane per range writes the links
            lid = (rm - rb == 1u) ? (n - 1u + s.b + rb) : (s.b + rm - 1u);   // (the ids sah_link assigns, for every lane of the range)
            rid = (re - rm == 1u) ? (n - 1u + s.b + rm) : (s.b + rm);
            if ((uint32_t)lane < rm) { re = rm; id = lid; } else { rb = rm; id = rid; }
            active = re - rb >= 2u;
        }
        __syncwarp();
    }
    if ((uint32_t)lane < cnt) final_order[s.b + lane] = prim;
}

__global__ void sah_compose_kernel(const uint32_t* __restrict__ order_k, const uint32_t* __restrict__ sorted_idx, uint32_t n,
                                   uint32_t* __restrict__ out) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) out[i] = sorted_idx[order_k[i]];
}

#define SCK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return e_; } while (0)

static inline size_t sah_seg_cap(uint32_t n) { return (size_t)n / (SAH_SMALL + 1) + 2; }
static inline size_t sah_small_cap(uint32_t n) { return (size_t)n / 2 + 2; }
static inline size_t sah_big_cap(uint32_t n) { return (size_t)n / (SAH_BIG + 1) + 2; }
static inline size_t sah_chunk_cap(uint32_t n) { return (size_t)n / SAH_CHUNK + sah_big_cap(n) + 2; }
// Bytes sah_build_hierarchy takes from the arena (on top of what bvh_build has taken).
size_t sah_scratch_bytes(uint32_t n) {
    return 2 * Arena::padded(sizeof(float4) * n) + 3 * Arena::padded(sizeof(uint32_t) * n) + Arena::padded(sizeof(uint32_t) * 8) +
           2 * Arena::padded(sizeof(SahSeg) * sah_seg_cap(n)) + Arena::padded(sizeof(SahSeg) * sah_small_cap(n)) +
           2 * Arena::padded(sizeof(SahSeg) * sah_big_cap(n)) + Arena::padded(sizeof(uint32_t) * sah_chunk_cap(n)) +
           Arena::padded(sizeof(int) * 3 * SAH_BINS * 7 * sah_big_cap(n)) + Arena::padded(sizeof(uint32_t) * 3 * SAH_BINS * sah_chunk_cap(n)) +
           Arena::padded(sizeof(uint2) * sah_chunk_cap(n)) + Arena::padded(sizeof(SahBigSplit) * sah_big_cap(n));
}

// Builds the binary hierarchy over n >= 2 triangles.  d_sorted_idx: Morton order (position -> flat index).  Writes
// children[0 .. n-1), parent[0 .. 2n-1) and d_order_out (position -> flat index in the SAH tree's leaf order).
cudaError_t sah_build_hierarchy(const float4* d_flat_tris, const uint32_t* d_sorted_idx, uint32_t n, const float centre_lo[3],
                                const float centre_hi[3], int2* d_children, uint32_t* d_parent, uint32_t* d_order_out,
                                uint32_t* launches, uint32_t* out_levels, Arena& arena, cudaStream_t st) {
    const size_t seg_cap = sah_seg_cap(n), small_cap = sah_small_cap(n), big_cap = sah_big_cap(n), chunk_cap = sah_chunk_cap(n);
    float4 *d_lo = arena.take<float4>(n), *d_hi = arena.take<float4>(n);
    uint32_t *d_idx0 = arena.take<uint32_t>(n), *d_idx1 = arena.take<uint32_t>(n), *d_order_k = arena.take<uint32_t>(n);
    uint32_t* d_counters = arena.take<uint32_t>(8);
    SahSeg *d_seg0 = arena.take<SahSeg>(seg_cap), *d_seg1 = arena.take<SahSeg>(seg_cap), *d_small = arena.take<SahSeg>(small_cap);
    SahSeg *d_big0 = arena.take<SahSeg>(big_cap), *d_big1 = arena.take<SahSeg>(big_cap);
    SahBig B;
    B.segs = nullptr; B.n_segs = 0;
    B.chunk_seg = arena.take<uint32_t>(chunk_cap);
    B.gbins = arena.take<int>(3 * SAH_BINS * 7 * big_cap);
    B.chunk_hist = arena.take<uint32_t>(3 * SAH_BINS * chunk_cap);
    B.chunk_off = arena.take<uint2>(chunk_cap);
    B.split = arena.take<SahBigSplit>(big_cap);
    if (!B.split) return cudaErrorMemoryAllocation;
    sah_boxes_kernel<<<sah_grid_for(n, 256, sm_count() * 16), 256, 0, st>>>(d_flat_tris, d_sorted_idx, n, d_lo, d_hi, d_idx0);
    ++*launches;

    SahSeg root;
    memset(&root, 0, sizeof(root));
    root.b = 0; root.e = n; root.id = 0; root.level = 0;
    for (int k = 0; k < 3; ++k) { root.clo[k] = 2.0f * centre_lo[k]; root.chi[k] = 2.0f * centre_hi[k]; }
    const uint32_t none = 0xffffffffu;
    SCK(cudaMemcpyAsync(d_parent, &none, sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    SCK(cudaMemsetAsync(d_counters, 0, sizeof(uint32_t) * 8, st));
    uint32_t n_mid = 0, n_big = 0, n_chunks = 0, levels = 0;
    SahSeg *seg_in = d_seg0, *seg_out = d_seg1, *big_in = d_big0, *big_out = d_big1;
    uint32_t *idx_in = d_idx0, *idx_out = d_idx1;
    if (n <= SAH_SMALL) {
        SCK(cudaMemcpyAsync(d_small, &root, sizeof(root), cudaMemcpyHostToDevice, st));
        sah_small_kernel<<<1, SAH_WARPS * 32, 0, st>>>(d_small, 1, idx_in, d_lo, d_hi, d_children, d_parent, d_order_k, n);
        ++*launches;
    } else if (n <= SAH_BIG) {
        SCK(cudaMemcpyAsync(seg_in, &root, sizeof(root), cudaMemcpyHostToDevice, st));
        n_mid = 1;
    } else {
        root.pad0 = 0; root.pad1 = (n + SAH_CHUNK - 1) / SAH_CHUNK;
        SCK(cudaMemcpyAsync(big_in, &root, sizeof(root), cudaMemcpyHostToDevice, st));
        n_big = 1; n_chunks = root.pad1;
    }
    while (n_mid > 0 || n_big > 0) {
        ++levels;
        SahLists lists;
        lists.next = seg_out; lists.small = d_small; lists.big = big_out; lists.counters = d_counters;
        if (n_big) {
            B.segs = big_in; B.n_segs = n_big;
            sah_big_prepare_kernel<<<n_big, SAH_BLOCK, 0, st>>>(B);
            sah_big_bin_kernel<<<n_chunks, SAH_BLOCK, 0, st>>>(B, idx_in, d_lo, d_hi);
            sah_big_split_kernel<<<n_big, SAH_BLOCK, 0, st>>>(B);
            sah_big_scatter_kernel<<<n_chunks, SAH_BLOCK, 0, st>>>(B, idx_in, idx_out, d_lo, d_hi);
            sah_big_emit_kernel<<<(n_big + 127) / 128, 128, 0, st>>>(B, lists, idx_out, d_children, d_parent, d_order_k, n);
            *launches += 5;
        }
        if (n_mid) {
            const uint32_t avg = n / n_mid;   // upper bound of the level's mean segment size
            if (avg <= 96u) sah_level_kernel<64><<<n_mid, 64, 0, st>>>(seg_in, idx_in, idx_out, d_lo, d_hi, lists, d_children, d_parent, d_order_k, n);
            else if (avg <= 320u) sah_level_kernel<128><<<n_mid, 128, 0, st>>>(seg_in, idx_in, idx_out, d_lo, d_hi, lists, d_children, d_parent, d_order_k, n);
            else sah_level_kernel<SAH_BLOCK><<<n_mid, SAH_BLOCK, 0, st>>>(seg_in, idx_in, idx_out, d_lo, d_hi, lists, d_children, d_parent, d_order_k, n);
            ++*launches;
        }
        uint32_t h[8];
        SCK(cudaMemcpyAsync(h, d_counters, sizeof(h), cudaMemcpyDeviceToHost, st));
        SCK(cudaStreamSynchronize(st));
        if (h[0] > seg_cap || h[1] > small_cap || h[3] > big_cap || h[4] > chunk_cap) return cudaErrorMemoryAllocation;
        if (h[1]) {
            sah_small_kernel<<<(h[1] + SAH_WARPS - 1) / SAH_WARPS, SAH_WARPS * 32, 0, st>>>(d_small, h[1], idx_out, d_lo, d_hi, d_children,
                                                                                          d_parent, d_order_k, n);
            ++*launches;
        }
        SCK(cudaMemsetAsync(d_counters, 0, sizeof(uint32_t) * 8, st));
        n_mid = h[0]; n_big = h[3]; n_chunks = h[4];
        SahSeg* ts = seg_in; seg_in = seg_out; seg_out = ts;
        ts = big_in; big_in = big_out; big_out = ts;
        uint32_t* ti = idx_in; idx_in = idx_out; idx_out = ti;
    }
    sah_compose_kernel<<<sah_grid_for(n, 256, sm_count() * 16), 256, 0, st>>>(d_order_k, d_sorted_idx, n, d_order_out);
    ++*launches;
    SCK(cudaStreamSynchronize(st));
    SCK(cudaGetLastError());
    if (out_levels) *out_levels = levels;
    return cudaSuccess;
}

}  // namespace ff
