// ff_trace.cu -- ray-batch kernels: closest hit (8-wide BVH, binary BVH, brute force), any hit, camera-ray
// generation, synthetic workload generators and the pack/unpack helpers of the C ABI.
// Reference seams: intersectRays kernel.cu:127-176 (trace), generateRays kernel.cu:186-205 (camera rays).
#include "ff_trace.cuh"

namespace ff {

using namespace ffm;

Tuning& tuning() {
    static Tuning t = {FF_POSTPONE_DIV, 0 /* refill_lanes: chosen per scene at commit */, FF_PERSISTENT_BLOCKS_PER_SM};
    return t;
}

int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

static inline int grid_for(size_t n, int block, int max_blocks) {
    size_t g = (n + block - 1) / block;
    if (g < 1) g = 1;
    if (g > (size_t)max_blocks) g = max_blocks;
    return (int)g;
}

#ifndef FF_STREAM_HINT
#define FF_STREAM_HINT 1   // rays are read and hits written with the cache-streaming (evict-first) hint, so that they do not push the scene out of L2:
                           // C4 batch A 1526 -> 1540 Mrays/s, everything else unchanged (profiles/r04q_variants_hint.txt)
#endif
__device__ __forceinline__ void store_hit(HitRec* hits, size_t i, const Best& b) {
    int4 v;
    if (b.slot < 0) { v.x = __float_as_int(INFINITY); v.y = 0; v.z = -1; v.w = -1; }
    else { v.x = __float_as_int(b.key); v.y = __float_as_int(b.t); v.z = b.slot; v.w = b.ord; }
#if FF_STREAM_HINT
    __stcs(reinterpret_cast<int4*>(hits) + i, v);
#else
    reinterpret_cast<int4*>(hits)[i] = v;
#endif
}

// Where a closest-hit kernel takes its rays from and what it leaves behind.  The API entry points read ray arrays and write
// the 16-byte record; the wavefront's extend stage writes 8-byte records and, on bounce 0, reads only the 16-byte direction
// (the origin of every camera ray is the camera position).  Generating the camera ray INSIDE this kernel was tried and lost:
// the fetch path runs at ~11 active lanes, and ~100 instructions of matrix product and IEEE normalise there cost more
// (+1.1 ms per C2 frame) than the 32 bytes per path they saved (profiles/r02c_bench.json).
// `skipbits` = slot + 1 of the triangle the ray starts on (0: none), see make_ray_ctx.
struct FetchArrays {
    static constexpr bool kStreams = false;
    const float4* __restrict__ ray_o;
    const float4* __restrict__ ray_d;
    __device__ __forceinline__ void operator()(uint32_t i, V3& o, V3& draw, uint32_t& skipbits) const {
#if FF_STREAM_HINT
        const float4 d4 = __ldcs(ray_d + i);
        o = ld3(__ldcs(ray_o + i)); draw = ld3(d4); skipbits = __float_as_uint(d4.w);
#else
        const float4 d4 = __ldg(ray_d + i);
        o = ld3(__ldg(ray_o + i)); draw = ld3(d4); skipbits = __float_as_uint(d4.w);
#endif
    }
};
struct FetchPrimary {   // bounce 0 of the wavefront: origin = the camera position, direction from the ray-generation kernel, no self-skip
    static constexpr bool kStreams = false;
    const float4* __restrict__ ray_d;
    float ox, oy, oz;
    __device__ __forceinline__ void operator()(uint32_t i, V3& o, V3& draw, uint32_t& skipbits) const {
#if FF_STREAM_HINT
        o = ffm::v3(ox, oy, oz); draw = ld3(__ldcs(ray_d + i)); skipbits = 0u;
#else
        o = ffm::v3(ox, oy, oz); draw = ld3(__ldg(ray_d + i)); skipbits = 0u;
#endif
    }
};
struct StoreRec {
    static constexpr bool kCompletion = false;
    HitRec* __restrict__ hits;
    __device__ __forceinline__ void operator()(size_t i, const Best& b) const { store_hit(hits, i, b); }
};
struct StoreLite {
    static constexpr bool kCompletion = false;
    HitLite* __restrict__ hits;
    __device__ __forceinline__ void operator()(size_t i, const Best& b) const {
#if FF_STREAM_HINT
        __stcs(reinterpret_cast<int2*>(hits) + i, b.slot < 0 ? make_int2(0, -1) : make_int2(__float_as_int(b.t), b.slot));
#else
        reinterpret_cast<int2*>(hits)[i] = b.slot < 0 ? make_int2(0, -1) : make_int2(__float_as_int(b.t), b.slot);
#endif
    }
};

// Host batches (ff_trace_closest): ONE persistent launch traces the batch while the copy engines are still delivering it and
// already taking results away.  The host queues the rays as chunks of 2^chunk_shift on a copy stream, each followed by a
// 4-byte copy that advances `avail` (rays landed so far); a warp whose refill reaches beyond `avail` waits for it.  Results go out
// in the reference's numbering (no hit-record pass, no unpack pass), written when the warp next refills; each warp counts
// the rays it finished and reports them to the chunk's counter when it moves on to another chunk, and the warp that completes
// a chunk raises its flag in page-locked host memory, on which the host queues the chunk's device-to-host copies.  Nothing here can wait for ever: a lane gives up after kStreamTimeoutNs, sets `err`,
// and every later wait returns at once (the host then redoes the batch through the chunked pipeline).
static constexpr unsigned long long kStreamTimeoutNs = 250000000ull;   // 0.25 s: rays are normally a few ms away at most; a spurious time-out only costs the fallback
// Acquires are rationed: on this machine an acquire (ld.acquire, fence.acq_rel) invalidates the SM's whole L1 (CCTL.IVALL).  The
// first version of this path did one per lane and chunk for `avail` and one fence per refill for the results, and traced the C4
// batch in 6.7 ms against 5.4 ms device-resident: the node / triangle hit rate went with the L1.  Now ONE lane per warp acquires
// `avail`, and only when the warp's refill reaches past what it already knows has landed (about once per warp and chunk: a few
// hundred per SM and batch); results are published with a RELEASE on the counting atomic, which invalidates nothing.
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t atom_add_release_gpu(uint32_t* p, uint32_t v) {
    uint32_t old;
    asm volatile("atom.add.release.gpu.global.u32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
    return old;
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
struct FetchRays6Stream {
    static constexpr bool kStreams = true;
    const float* __restrict__ rays6;
    StreamSync* sync;
    uint32_t* host_err;     // page-locked host word: a lane timed out (the host reads it after the launch, without a copy)
    // Warp-converged.  Returns a count of landed rays that is >= `want` (lane 0 spins on `avail` with acquire loads; the
    // __syncwarp orders every lane's ray loads after them), or 0xffffffff once a time-out has been raised.
    __device__ __forceinline__ uint32_t wait_warp(uint32_t want, int lane) const {
        uint32_t known = 0u;
        if (lane == 0) {
            const unsigned long long t0 = global_ns();
            for (;;) {
                known = ld_acquire_sys(&sync->avail);
                if (known >= want) break;
                if (*(volatile uint32_t*)&sync->err) { known = 0xffffffffu; break; }
                if (global_ns() - t0 > kStreamTimeoutNs) { atomicExch(&sync->err, 1u); *(volatile uint32_t*)host_err = 1u; known = 0xffffffffu; break; }
                __nanosleep(256);
            }
        }
        known = __shfl_sync(0xffffffffu, known, 0);
        __syncwarp();
        return known;
    }
    __device__ __forceinline__ void operator()(uint32_t i, V3& o, V3& draw, uint32_t& skipbits) const {
        // written by the copy engine while this kernel runs: read through L2 (.cg), never through the read-only path
        const float2* p = reinterpret_cast<const float2*>(rays6 + 6 * (size_t)i);
        const float2 a = __ldcg(p), b = __ldcg(p + 1), c = __ldcg(p + 2);
        o = ffm::v3(a.x, a.y, b.x); draw = ffm::v3(b.y, c.x, c.y); skipbits = 0u;
    }
};
struct StoreIdsStream {
    static constexpr bool kCompletion = true;
    const int* __restrict__ ord_geom;
    const int* __restrict__ ord_tri;
    int* __restrict__ geom;
    int* __restrict__ tri;
    float* __restrict__ t;
    StreamSync* sync;
    uint32_t* host_flags;   // page-locked host memory, one word per chunk
    uint32_t chunk_shift, n;
    // Nothing is written when a ray finishes: that happens in the divergent traversal loop, where the two dependent loads of
    // the reference numbering would stall the whole warp once per ray.  The lane keeps its result in `best` (untouched until it
    // draws its next ray) and writes it in retire(), at the warp's next refill, together with the other lanes that finished.
    __device__ __forceinline__ void operator()(size_t, const Best&) const {}
    // Warp-converged.  `pending` = index of the ray this lane finished since the last call (0xffffffff: none), `best` its result
    // (== unpack_hits_kernel on store_hit's record).  Finished rays are counted per warp -- acc_chunk is the newest chunk seen,
    // acc_count holds its count (low half) and that of the chunk before it (high half) -- and reported to the chunk's counter
    // by one release atomic, which is what publishes the warp's result stores, when the warp has moved two chunks on, and
    // when it exits: about one report per warp and chunk (a report costs a MEMBAR; with one open chunk per warp the late rays
    // of a chunk made the warp flip between two chunks and report several times).
    __device__ __forceinline__ void retire(uint32_t& pending, const Best& b, int lane, uint32_t& acc_chunk, uint32_t& acc_count) const {
        unsigned pend = __ballot_sync(0xffffffffu, pending != 0xffffffffu);
        if (pend == 0u) return;
        if (pending != 0xffffffffu) {
            const bool hit = b.slot >= 0 && b.ord >= 0;
            geom[pending] = hit ? __ldg(ord_geom + b.ord) : -1;
            tri[pending] = hit ? __ldg(ord_tri + b.ord) : -1;
            t[pending] = b.slot < 0 ? INFINITY : b.key;
        }
        while (pend) {
            const int leader = __ffs(pend) - 1;
            const uint32_t mine = pending >> chunk_shift;
            const uint32_t c = __shfl_sync(0xffffffffu, mine, leader);
            const unsigned same = __ballot_sync(0xffffffffu, pending != 0xffffffffu && mine == c);
            // two chunks are open per warp (late rays of chunk c finish among the first of c + 1): the counts of the older one
            // ride in the top half of acc_count and are reported when a third chunk shows up
            const uint32_t k = (uint32_t)__popc(same);
            if (c == acc_chunk) acc_count += k;
            else if (c + 1u == acc_chunk) acc_count += k << 16;
            else {
                if (c == acc_chunk + 1u) { flush(lane, acc_chunk - 1u, acc_count >> 16); acc_count = (acc_count & 0xffffu) << 16; }
                else { flush(lane, acc_chunk - 1u, acc_count >> 16); flush(lane, acc_chunk, acc_count & 0xffffu); acc_count = 0u; }
                acc_chunk = c;
                acc_count += k;
            }
            if ((acc_count & 0xffffu) > 0xff00u || (acc_count >> 16) > 0xff00u) {   // (cannot happen with 2^16+ rays per chunk and warp; keeps the halves apart)
                flush(lane, acc_chunk - 1u, acc_count >> 16); flush(lane, acc_chunk, acc_count & 0xffffu); acc_count = 0u;
            }
            pend &= ~same;
        }
        pending = 0xffffffffu;
    }
    __device__ __forceinline__ void flush(int lane, uint32_t c, uint32_t cnt) const {
        if (cnt == 0u) return;
        __syncwarp();   // orders every lane's result stores before lane 0's release below (which is cumulative over them)
        if (lane == 0) {
            const uint32_t old = atom_add_release_gpu(&sync->done[c], cnt);
            if (old + cnt == min(1u << chunk_shift, n - (c << chunk_shift))) {
                __threadfence_system();
                *(volatile uint32_t*)(host_flags + c) = 1u;
            }
        }
    }
};

// ------------------------------------------------------------------ closest hit, one thread per ray
template <int MODE, bool COUNT, bool SKIP, class Fetch, class Store>
__global__ void __launch_bounds__(FF_TRACE_BLOCK) trace_closest_kernel(DevScene sc, Fetch fetch, Store store, const uint32_t* __restrict__ count_ptr,
                                                                        uint32_t n_max, TraceCounters* ctr) {
    __shared__ uint2 sm_stack[(FF_SM_STACK + FF_TRI_QUEUE) * FF_TRACE_BLOCK];
    const uint32_t n = count_ptr ? min(*count_ptr, n_max) : n_max;
    unsigned cn = 0, ct = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        V3 o, draw;
        uint32_t skipbits;
        fetch(i, o, draw, skipbits);
        const V3 dn = normalize(draw);  // kernel.cu:138
        Best best;
        best_init(best, INFINITY);
        test_planes<false>(sc, o, draw, best);
        if (MODE == FF_TRACE_BVH8) trace_bvh8<false, COUNT, SKIP>(sc, o, dn, best, sm_stack + threadIdx.x, FF_TRACE_BLOCK, cn, ct, SKIP ? skipbits - 1u : 0xffffffffu);
        else trace_bvh2<false, COUNT>(sc, o, dn, best, cn, ct);
        store(i, best);
    }
    if (COUNT) {
        atomicAdd(&ctr->nodes, (unsigned long long)cn);
        atomicAdd(&ctr->tris, (unsigned long long)ct);
    }
}

// ------------------------------------------------------------------ closest hit, persistent warps + dynamic ray fetch
// Warps pull rays from a global counter; when fewer than FF_REFILL_LANES lanes of a warp are still traversing, the
// idle lanes fetch new rays instead of waiting for the slowest one (Aila & Laine 2009).  Results do not depend on
// the schedule (the accept rule is order independent).
// MINB = resident CTAs per SM the instantiation is compiled for: FF_TRACE_MINB (7: 72 registers) for scenes that live in L2,
// FF_TRACE_MINB_BIG (8: 64 registers) for scenes far larger than L2, where four more warps per SM hide DRAM latency: the 20 M-triangle
// C5 frame renders 3 % faster with 8 (326.0 -> 315.5 ms at 8 spp), the 1 M-triangle soup does not care, C3 loses 2.5 %
// (profiles/r04n_c5_minb.txt, r02c_variants.txt).
#define FF_TRACE_MINB_BIG 8
#define FF_TRACE_BIG_SCENE (1u << 21)   // triangles: ~130 MB of nodes + triangles, more than the 126 MB L2
template <bool SKIP, class Fetch, class Store, int MINB = FF_TRACE_MINB>
__global__ void __launch_bounds__(FF_TRACE_BLOCK, MINB) trace_closest_persistent_kernel(DevScene sc, Fetch fetch, Store store,
                                                                                   const uint32_t* __restrict__ count_ptr, uint32_t n_max,
                                                                                   uint32_t* work_counter) {
    __shared__ uint2 sm_stack[(FF_SM_STACK + FF_TRI_QUEUE) * FF_TRACE_BLOCK];
    const uint32_t n = count_ptr ? min(*count_ptr, n_max) : n_max;
    const int lane = threadIdx.x & 31;
    uint2* my_stack = sm_stack + threadIdx.x;
    bool has = false, exhausted = false;
    uint32_t ray_index = 0;
    RayCtx r;
    Trav tv;
    uint2 lstack[FF_LOCAL_STACK];
    Best best;
    unsigned cn = 0, ct = 0;
    uint32_t pending = 0xffffffffu;   // Store::kCompletion: the ray this lane finished and has not reported yet
    uint32_t avail_known = 0u;        // Fetch::kStreams: rays known to have landed (warp-uniform)
    uint32_t acc_chunk = 0u, acc_count = 0u;   // Store::kCompletion: finished rays of chunk acc_chunk this warp has not reported yet
    for (;;) {
        const unsigned need = __ballot_sync(0xffffffffu, !has);
        if constexpr (Store::kCompletion) { if (need) store.retire(pending, best, lane, acc_chunk, acc_count); }
        if (need && !exhausted) {
            uint32_t base = 0;
            const int leader = __ffs(need) - 1;
            if (lane == leader) base = atomicAdd(work_counter, (uint32_t)__popc(need));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (base + (uint32_t)__popc(need) >= n) exhausted = true;
            if constexpr (Fetch::kStreams) {   // host batches: the rays of this refill may still be on their way
                const uint32_t want = min(n, base + (uint32_t)__popc(need));
                if (want > avail_known) avail_known = fetch.wait_warp(want, lane);
            }
            if (!has) {
                const uint32_t i = base + __popc(need & ((1u << lane) - 1u));
                if (i < n) {
                    V3 o, draw;
                    uint32_t skipbits;
                    fetch(i, o, draw, skipbits);
                    const V3 dn = normalize(draw);  // kernel.cu:138
                    best_init(best, INFINITY);
                    test_planes<false>(sc, o, draw, best);
                    r = make_ray_ctx(o, dn, SKIP ? skipbits - 1u : 0xffffffffu);
                    tv.start(best.key);
                    ray_index = i;
                    has = sc.n_nodes != 0;
                    if (!has) { store(i, best); pending = i; }
                }
            }
        }
        if (!__any_sync(0xffffffffu, has)) { if (exhausted) break; continue; }  // (no hierarchy / occluded by an analytic primitive: rays finish in the fetch path)
#if FF_NODE_MIN || FF_TRI_MIN
        // Experiment: every lane runs the loop (idle ones predicated off), so the warp votes with a full mask: a node visit
        // is held back while fewer than FF_NODE_MIN lanes want one and some lane still has triangles to test.
        for (;;) {
            const unsigned nb = __ballot_sync(0xffffffffu, has && (tv.ngroup.y & 0xff000000u) != 0u);
            const unsigned tb = __ballot_sync(0xffffffffu, has && tv.tgroup.y != 0u);
            bool do_node = (__popc(nb) >= FF_NODE_MIN) || tb == 0u;
            bool do_tri = (__popc(tb) >= FF_TRI_MIN) || nb == 0u;
            if (!do_node && !do_tri) do_node = do_tri = true;
            if (has && !bvh8_step<false, false, SKIP>(sc, r, tv, best, my_stack, FF_TRACE_BLOCK, lstack, cn, ct, do_node, do_tri)) {
                store(ray_index, best); pending = ray_index;
                has = false;
            }
            const unsigned hb = __ballot_sync(0xffffffffu, has);
            if (hb == 0u || (!exhausted && __popc(hb) < sc.refill_lanes)) break;
        }
#else
        if (has) {
#if FF_TRI_EVERY > 1
            unsigned it = 0;
#endif
            for (;;) {
#if FF_TRI_EVERY > 1
                const bool do_tri = (++it % FF_TRI_EVERY) == 0;
                if (!bvh8_step<false, false, SKIP>(sc, r, tv, best, my_stack, FF_TRACE_BLOCK, lstack, cn, ct, true, do_tri)) {
#else
                if (!bvh8_step<false, false, SKIP>(sc, r, tv, best, my_stack, FF_TRACE_BLOCK, lstack, cn, ct)) {
#endif
                    store(ray_index, best); pending = ray_index;
                    has = false;
                    break;
                }
                if (!exhausted && __popc(__activemask()) < sc.refill_lanes) break;
            }
        }
#endif
        __syncwarp();
    }
    // rays that finished in the fetch path of the last round (no hierarchy) have not been reported yet; the break is warp-uniform
    if constexpr (Store::kCompletion) { store.retire(pending, best, lane, acc_chunk, acc_count); store.flush(lane, acc_chunk - 1u, acc_count >> 16); store.flush(lane, acc_chunk, acc_count & 0xffffu); }
}

// Brute force: the reference's own algorithm (every ray against every triangle), tiles of triangles staged
// in shared memory.  Same leaf test and accept rule, so it is order independent.
#define FF_BRUTE_TILE 256
template <bool COUNT>
__global__ void __launch_bounds__(FF_BRUTE_TILE) trace_brute_kernel(DevScene sc, const float4* __restrict__ ray_o,
                                                                     const float4* __restrict__ ray_d,
                                                                     const uint32_t* __restrict__ count_ptr, uint32_t n_max,
                                                                     HitRec* __restrict__ hits, TraceCounters* ctr) {
    __shared__ float4 sv[3 * FF_BRUTE_TILE];
    const uint32_t n = count_ptr ? min(*count_ptr, n_max) : n_max;
    const uint32_t rounds = (n + gridDim.x * blockDim.x - 1) / (gridDim.x * blockDim.x);
    unsigned long long ct = 0;
    for (uint32_t r = 0; r < rounds; ++r) {
        const uint32_t i = (r * gridDim.x + blockIdx.x) * blockDim.x + threadIdx.x;
        const bool live = i < n;
        V3 o = v3(0, 0, 0), draw = v3(0, 0, 1);
        if (live) { o = ld3(__ldg(ray_o + i)); draw = ld3(__ldg(ray_d + i)); }
        const V3 dn = normalize(draw);
        Best best;
        best_init(best, INFINITY);
        if (live) test_planes<false>(sc, o, draw, best);
        for (uint32_t base = 0; base < sc.n_tris; base += FF_BRUTE_TILE) {
            const uint32_t m = min((uint32_t)FF_BRUTE_TILE, sc.n_tris - base);
            __syncthreads();
            for (uint32_t k = threadIdx.x; k < 3 * m; k += blockDim.x) sv[k] = __ldg(sc.tris + 3 * (size_t)base + k);
            __syncthreads();
            if (live) {
                for (uint32_t k = 0; k < m; ++k) {
                    float t;
                    const float4 a = sv[3 * k];
                    if (intersect_triangle(ld3(a), ld3(sv[3 * k + 1]), ld3(sv[3 * k + 2]), o, dn, t)) {
                        V3 p;
                        const float key = hit_key(o, dn, t, p);
                        const int ord = __float_as_int(a.w);
                        if (best_accepts(best, key, ord)) { best.key = key; best.t = t; best.slot = (int)(base + k); best.ord = ord; }
                    }
                }
                if (COUNT) ct += m;
            }
        }
        if (live) store_hit(hits, i, best);
    }
    if (COUNT) atomicAdd(&ctr->tris, ct);
}

template <class Fetch, class Store>
static void launch_closest(const DevScene& sc, const Fetch& fetch, const Store& store, const uint32_t* d_count, uint32_t n_max, int mode,
                           TraceCounters* d_counters, uint32_t* d_work_counter, cudaStream_t st) {
    const bool count = (mode & FF_TRACE_COUNT) != 0;
    const int m = mode & 0xff;
    const bool skip = (mode & FF_TRACE_SELFSKIP) != 0;
    if (m == FF_TRACE_BVH8 && !count && d_work_counter && !(mode & FF_TRACE_STATIC)) {
        // persistent: enough warps to fill the machine, each pulling rays until the batch is drained.
        // *d_work_counter must be zero (the caller clears it on the same stream).
        const int g = grid_for(n_max, FF_TRACE_BLOCK, sm_count() * tuning().blocks_per_sm);
        if (sc.n_tris > FF_TRACE_BIG_SCENE) {
            if (skip) trace_closest_persistent_kernel<true, Fetch, Store, FF_TRACE_MINB_BIG><<<g, FF_TRACE_BLOCK, 0, st>>>(sc, fetch, store, d_count, n_max, d_work_counter);
            else trace_closest_persistent_kernel<false, Fetch, Store, FF_TRACE_MINB_BIG><<<g, FF_TRACE_BLOCK, 0, st>>>(sc, fetch, store, d_count, n_max, d_work_counter);
        } else {
            if (skip) trace_closest_persistent_kernel<true><<<g, FF_TRACE_BLOCK, 0, st>>>(sc, fetch, store, d_count, n_max, d_work_counter);
            else trace_closest_persistent_kernel<false><<<g, FF_TRACE_BLOCK, 0, st>>>(sc, fetch, store, d_count, n_max, d_work_counter);
        }
        return;
    }
    const int g = grid_for(n_max, FF_TRACE_BLOCK, sm_count() * 64);
    if (m == FF_TRACE_BVH2) {
        if (count) trace_closest_kernel<FF_TRACE_BVH2, true, false><<<g, FF_TRACE_BLOCK, 0, st>>>(sc, fetch, store, d_count, n_max, d_counters);
        else trace_closest_kernel<FF_TRACE_BVH2, false, false><<<g, FF_TRACE_BLOCK, 0, st>>>(sc, fetch, store, d_count, n_max, d_counters);
    } else if (skip) {
        if (count) trace_closest_kernel<FF_TRACE_BVH8, true, true><<<g, FF_TRACE_BLOCK, 0, st>>>(sc, fetch, store, d_count, n_max, d_counters);
        else trace_closest_kernel<FF_TRACE_BVH8, false, true><<<g, FF_TRACE_BLOCK, 0, st>>>(sc, fetch, store, d_count, n_max, d_counters);
    } else {
        if (count) trace_closest_kernel<FF_TRACE_BVH8, true, false><<<g, FF_TRACE_BLOCK, 0, st>>>(sc, fetch, store, d_count, n_max, d_counters);
        else trace_closest_kernel<FF_TRACE_BVH8, false, false><<<g, FF_TRACE_BLOCK, 0, st>>>(sc, fetch, store, d_count, n_max, d_counters);
    }
}
void k_trace_closest(const DevScene& sc, const float4* d_ray_o, const float4* d_ray_d, const uint32_t* d_count,
                     uint32_t n_max, int mode, HitRec* d_hits, TraceCounters* d_counters, uint32_t* d_work_counter, cudaStream_t st) {
    if (n_max == 0) return;
    if ((mode & 0xff) == FF_TRACE_BRUTE) {
        const int g = grid_for(n_max, FF_BRUTE_TILE, sm_count() * 8);
        if (mode & FF_TRACE_COUNT) trace_brute_kernel<true><<<g, FF_BRUTE_TILE, 0, st>>>(sc, d_ray_o, d_ray_d, d_count, n_max, d_hits, d_counters);
        else trace_brute_kernel<false><<<g, FF_BRUTE_TILE, 0, st>>>(sc, d_ray_o, d_ray_d, d_count, n_max, d_hits, d_counters);
        return;
    }
    launch_closest(sc, FetchArrays{d_ray_o, d_ray_d}, StoreRec{d_hits}, d_count, n_max, mode, d_counters, d_work_counter, st);
}
void k_trace_closest_stream(const DevScene& sc, const float* d_rays6, uint32_t n, uint32_t chunk_shift, StreamSync* d_sync, uint32_t* host_flags,
                            int* d_geom, int* d_tri, float* d_t, uint32_t* d_work_counter, cudaStream_t st) {
    if (n == 0) return;
    const FetchRays6Stream fetch{d_rays6, d_sync, host_flags + FF_STREAM_MAX_CHUNKS};
    const StoreIdsStream store{sc.ord_geom, sc.ord_tri, d_geom, d_tri, d_t, d_sync, host_flags, chunk_shift, n};
    const int g = grid_for(n, FF_TRACE_BLOCK, sm_count() * tuning().blocks_per_sm);
    trace_closest_persistent_kernel<false><<<g, FF_TRACE_BLOCK, 0, st>>>(sc, fetch, store, nullptr, n, d_work_counter);
}
void k_trace_extend(const DevScene& sc, const WaveGen* gen, const float4* d_ray_o, const float4* d_ray_d, const uint32_t* d_count,
                    uint32_t n_max, int mode, HitLite* d_hits, TraceCounters* d_counters, uint32_t* d_work_counter, cudaStream_t st) {
    if (n_max == 0) return;
    if (gen) launch_closest(sc, FetchPrimary{d_ray_d, gen->ox, gen->oy, gen->oz}, StoreLite{d_hits}, d_count, n_max, mode & ~FF_TRACE_SELFSKIP, d_counters, d_work_counter, st);
    else launch_closest(sc, FetchArrays{d_ray_o, d_ray_d}, StoreLite{d_hits}, d_count, n_max, mode, d_counters, d_work_counter, st);
}

// ------------------------------------------------------------------ any hit (occlusion)
template <int MODE>
__global__ void __launch_bounds__(FF_TRACE_BLOCK) trace_any_kernel(DevScene sc, const float4* __restrict__ ray_o,
                                                                    const float4* __restrict__ ray_d,
                                                                    const float* __restrict__ tmax, uint32_t n,
                                                                    unsigned char* __restrict__ occluded) {
    __shared__ uint2 sm_stack[(FF_SM_STACK + FF_TRI_QUEUE) * FF_TRACE_BLOCK];
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const V3 o = ld3(__ldg(ray_o + i)), draw = ld3(__ldg(ray_d + i));
        const V3 dn = normalize(draw);
        Best best;
        best_init(best, tmax ? tmax[i] : INFINITY);
        unsigned cn = 0, ct = 0;
        test_planes<true>(sc, o, draw, best);
        if (best.slot < 0) {
            if (MODE == FF_TRACE_BRUTE) {
                for (uint32_t k = 0; k < sc.n_tris && best.slot < 0; ++k) test_slot<true>(sc, (int)k, o, dn, best);
            } else if (MODE == FF_TRACE_BVH2) trace_bvh2<true, false>(sc, o, dn, best, cn, ct);
            else trace_bvh8<true, false>(sc, o, dn, best, sm_stack + threadIdx.x, FF_TRACE_BLOCK, cn, ct);
        }
        occluded[i] = best.slot >= 0 ? 1 : 0;
    }
}
// Any-hit with persistent warps + dynamic fetch (same scheduling as the closest-hit kernel).  Serves ff_trace_any
// (tmax from an array, result = occluded flags) and the wavefront's shadow stage (tmax in ray_d.w, path id in ray_o.w;
// an unoccluded entry adds its contribution to radiance[path] -- one shadow ray per path per bounce, stages are
// stream-ordered, so a plain read-modify-write is race free and the sum order matches the CPU restatement).
template <bool SKIP>
__global__ void __launch_bounds__(FF_TRACE_BLOCK, FF_TRACE_MINB) trace_any_persistent_kernel(DevScene sc, const float4* __restrict__ ray_o,
                                                                               const float4* __restrict__ ray_d,
                                                                               const float* __restrict__ tmax_array, int tmax_in_dw,
                                                                               const uint32_t* __restrict__ count_ptr, uint32_t n_max,
                                                                               unsigned char* __restrict__ occluded,
                                                                               const float4* __restrict__ contrib,
                                                                               float4* __restrict__ radiance, uint32_t* work_counter) {
    __shared__ uint2 sm_stack[(FF_SM_STACK + FF_TRI_QUEUE) * FF_TRACE_BLOCK];
    const uint32_t n = count_ptr ? min(*count_ptr, n_max) : n_max;
    const int lane = threadIdx.x & 31;
    uint2* my_stack = sm_stack + threadIdx.x;
    bool has = false, exhausted = false;
    uint32_t ray_index = 0, path = 0;
    RayCtx r;
    Trav tv;
    uint2 lstack[FF_LOCAL_STACK];
    Best best;
    unsigned cn = 0, ct = 0;
    auto finish = [&]() {
        const bool occ = best.slot >= 0;
        if (occluded) occluded[ray_index] = occ ? 1 : 0;
        if (radiance && !occ) {
            const float4 c = __ldg(contrib + ray_index);
            float4 a = radiance[path];
            a.x = fadd(a.x, c.x); a.y = fadd(a.y, c.y); a.z = fadd(a.z, c.z);
            radiance[path] = a;
        }
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
                const uint32_t i = base + __popc(need & ((1u << lane) - 1u));
                if (i < n) {
                    const float4 o4 = __ldg(ray_o + i), d4 = __ldg(ray_d + i);
                    const V3 o = ld3(o4), draw = ld3(d4);
                    const V3 dn = normalize(draw);
                    best_init(best, tmax_in_dw ? d4.w : (tmax_array ? tmax_array[i] : INFINITY));
                    test_planes<true>(sc, o, draw, best);
                    // shadow stage: the slot of the triangle the light sample was taken from rides in contrib.w (make_ray_ctx)
                    r = make_ray_ctx(o, dn, SKIP ? skip_from_bits(__ldg(&contrib[i].w)) : 0xffffffffu);
                    tv.start(best.key);
                    ray_index = i;
                    path = __float_as_uint(o4.w);
                    has = sc.n_nodes != 0 && best.slot < 0;
                    if (!has) finish();
                }
            }
        }
        if (!__any_sync(0xffffffffu, has)) { if (exhausted) break; continue; }  // (no hierarchy / occluded by an analytic primitive: rays finish in the fetch path)
        if (has) {
            for (;;) {
                if (!bvh8_step<true, false, SKIP>(sc, r, tv, best, my_stack, FF_TRACE_BLOCK, lstack, cn, ct)) {
                    finish();
                    has = false;
                    break;
                }
                if (!exhausted && __popc(__activemask()) < sc.refill_lanes) break;
            }
        }
        __syncwarp();
    }
}

void k_trace_any(const DevScene& sc, const float4* d_ray_o, const float4* d_ray_d, const float* d_tmax, uint32_t n, int mode,
                 unsigned char* d_occluded, uint32_t* d_work_counter, cudaStream_t st) {
    if (n == 0) return;
    const int m = mode & 0xff;
    if (m == FF_TRACE_BVH8 && d_work_counter && !(mode & FF_TRACE_STATIC)) {
        const int g = grid_for(n, FF_TRACE_BLOCK, sm_count() * tuning().blocks_per_sm);
        trace_any_persistent_kernel<false><<<g, FF_TRACE_BLOCK, 0, st>>>(sc, d_ray_o, d_ray_d, d_tmax, 0, nullptr, n, d_occluded, nullptr, nullptr, d_work_counter);
        return;
    }
    const int g = grid_for(n, FF_TRACE_BLOCK, sm_count() * 64);
    if (m == FF_TRACE_BRUTE) trace_any_kernel<FF_TRACE_BRUTE><<<g, FF_TRACE_BLOCK, 0, st>>>(sc, d_ray_o, d_ray_d, d_tmax, n, d_occluded);
    else if (m == FF_TRACE_BVH2) trace_any_kernel<FF_TRACE_BVH2><<<g, FF_TRACE_BLOCK, 0, st>>>(sc, d_ray_o, d_ray_d, d_tmax, n, d_occluded);
    else trace_any_kernel<FF_TRACE_BVH8><<<g, FF_TRACE_BLOCK, 0, st>>>(sc, d_ray_o, d_ray_d, d_tmax, n, d_occluded);
}
void k_trace_shadow(const DevScene& sc, const float4* d_ray_o, const float4* d_ray_d, const float4* d_contrib,
                    const uint32_t* d_count, uint32_t n_max, float4* d_radiance, uint32_t* d_work_counter, bool self_skip, cudaStream_t st) {
    if (n_max == 0) return;
    const int g = grid_for(n_max, FF_TRACE_BLOCK, sm_count() * tuning().blocks_per_sm);
    if (self_skip) trace_any_persistent_kernel<true><<<g, FF_TRACE_BLOCK, 0, st>>>(sc, d_ray_o, d_ray_d, nullptr, 1, d_count, n_max, nullptr, d_contrib, d_radiance, d_work_counter);
    else trace_any_persistent_kernel<false><<<g, FF_TRACE_BLOCK, 0, st>>>(sc, d_ray_o, d_ray_d, nullptr, 1, d_count, n_max, nullptr, d_contrib, d_radiance, d_work_counter);
}

// ------------------------------------------------------------------ camera rays (kernel.cu:198-205)
__device__ __forceinline__ void camera_ray(const CamDev& cam, float px, float py, V3& o, V3& d) {
    o = v3(cam.pos[0], cam.pos[1], cam.pos[2]);
    const float Px = fsub(fmul(fdiv(px, cam.width), 2.f), 1.f);
    const float Py = fsub(1.f, fmul(fdiv(py, cam.height), 2.f));
    const V4 look = mul_mv(cam.inv_vp, fmul(Px, cam.far_clip), fmul(Py, cam.far_clip), fmul(1.f, cam.far_clip),
                           fmul(1.f, cam.far_clip));
    d = normalize(sub(v3(look.x, look.y, look.z), o));
}
__global__ void camera_rays_kernel(CamDev cam, float4* ray_o, float4* ray_d) {
    const uint32_t npix = (uint32_t)cam.w * cam.h;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += gridDim.x * blockDim.x) {
        const uint32_t x = i % cam.w, y = i / cam.w;
        V3 o, d;
        camera_ray(cam, (float)x, (float)y, o, d);
        ray_o[i] = make_float4(o.x, o.y, o.z, 0.f);
        ray_d[i] = make_float4(d.x, d.y, d.z, 0.f);
    }
}
void k_camera_rays(const CamDev& cam, float4* d_ray_o, float4* d_ray_d, cudaStream_t st) {
    const size_t n = (size_t)cam.w * cam.h;
    camera_rays_kernel<<<grid_for(n, 256, sm_count() * 16), 256, 0, st>>>(cam, d_ray_o, d_ray_d);
}

// ------------------------------------------------------------------ pack / unpack
__global__ void pack_rays_kernel(const float* __restrict__ r6, size_t n, float4* ray_o, float4* ray_d) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        ray_o[i] = make_float4(r6[6 * i], r6[6 * i + 1], r6[6 * i + 2], 0.f);
        ray_d[i] = make_float4(r6[6 * i + 3], r6[6 * i + 4], r6[6 * i + 5], 0.f);
    }
}
void k_pack_rays(const float* d_rays6, size_t n, float4* d_ray_o, float4* d_ray_d, cudaStream_t st) {
    if (n) pack_rays_kernel<<<grid_for(n, 256, sm_count() * 16), 256, 0, st>>>(d_rays6, n, d_ray_o, d_ray_d);
}
__global__ void unpack_rays_kernel(const float4* __restrict__ ray_o, const float4* __restrict__ ray_d, size_t n, float* r6) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float4 o = ray_o[i], d = ray_d[i];
        r6[6 * i] = o.x; r6[6 * i + 1] = o.y; r6[6 * i + 2] = o.z;
        r6[6 * i + 3] = d.x; r6[6 * i + 4] = d.y; r6[6 * i + 5] = d.z;
    }
}
void k_unpack_rays(const float4* d_ray_o, const float4* d_ray_d, size_t n, float* d_rays6, cudaStream_t st) {
    if (n) unpack_rays_kernel<<<grid_for(n, 256, sm_count() * 16), 256, 0, st>>>(d_ray_o, d_ray_d, n, d_rays6);
}
__global__ void unpack_hits_kernel(DevScene sc, const HitRec* __restrict__ hits, size_t n, int* geom, int* tri, float* t) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const HitRec h = hits[i];
        if (geom) geom[i] = h.ord >= 0 ? sc.ord_geom[h.ord] : -1;
        if (tri) tri[i] = h.ord >= 0 ? sc.ord_tri[h.ord] : -1;
        if (t) t[i] = h.key;
    }
}
void k_unpack_hits(const DevScene& sc, const HitRec* d_hits, size_t n, int* d_geom, int* d_tri, float* d_t, cudaStream_t st) {
    if (n) unpack_hits_kernel<<<grid_for(n, 256, sm_count() * 16), 256, 0, st>>>(sc, d_hits, n, d_geom, d_tri, d_t);
}

// ------------------------------------------------------------------ synthetic workloads
// Triangle i: centroid c ~ U[0,1)^3, vertices c + s*U[-1,1)^3; Philox key {0xF1EEF1E5, N}, counter {i, vertex, 0, 0}.
__global__ void make_soup_kernel(size_t n, float s, float* xyz9) {
    const uint32_t k0 = 0xF1EEF1E5u, k1 = (uint32_t)n;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        U4 c = philox4x32_10(U4{(uint32_t)i, 0u, 0u, 0u}, k0, k1);
        const float cx = u01(c.x), cy = u01(c.y), cz = u01(c.z);
        for (uint32_t v = 0; v < 3; ++v) {
            U4 r = philox4x32_10(U4{(uint32_t)i, v + 1u, 0u, 0u}, k0, k1);
            xyz9[9 * i + 3 * v + 0] = fadd(cx, fmul(s, fsub(fmul(u01(r.x), 2.0f), 1.0f)));
            xyz9[9 * i + 3 * v + 1] = fadd(cy, fmul(s, fsub(fmul(u01(r.y), 2.0f), 1.0f)));
            xyz9[9 * i + 3 * v + 2] = fadd(cz, fmul(s, fsub(fmul(u01(r.z), 2.0f), 1.0f)));
        }
    }
}
void k_make_soup(size_t n, float s, float* d_xyz9, cudaStream_t st) {
    if (n) make_soup_kernel<<<grid_for(n, 256, sm_count() * 16), 256, 0, st>>>(n, s, d_xyz9);
}
// Batch A: origin ~ U[0,1)^3, direction uniform on the sphere; key {0xF1EEF1E5, 1}, counter {i, 0|1, salt, 0}.
__global__ void make_rays_uniform_kernel(uint64_t first, size_t n, uint32_t salt, float4* ray_o, float4* ray_d) {
    for (size_t k = blockIdx.x * (size_t)blockDim.x + threadIdx.x; k < n; k += (size_t)gridDim.x * blockDim.x) {
        const uint32_t i = (uint32_t)(first + k);
        U4 a = philox4x32_10(U4{i, 0u, salt, 0u}, 0xF1EEF1E5u, 1u);
        U4 b = philox4x32_10(U4{i, 1u, salt, 0u}, 0xF1EEF1E5u, 1u);
        const float z = fsub(1.0f, fmul(2.0f, u01(b.x)));
        const float r = fsqrt(fmaxf(0.0f, fsub(1.0f, fmul(z, z))));
        float sn, cs;
        sincos2pi(u01(b.y), sn, cs);
        ray_o[k] = make_float4(u01(a.x), u01(a.y), u01(a.z), 0.f);
        ray_d[k] = make_float4(fmul(r, cs), fmul(r, sn), z, 0.f);
    }
}
void k_make_rays_uniform(uint64_t first, size_t n, uint32_t salt, float4* d_ray_o, float4* d_ray_d, cudaStream_t st) {
    if (n) make_rays_uniform_kernel<<<grid_for(n, 256, sm_count() * 16), 256, 0, st>>>(first, n, salt, d_ray_o, d_ray_d);
}

// Batch B: one cosine-hemisphere bounce off every hit of an input batch (compacted, warp-aggregated append).
__global__ void make_rays_bounce_kernel(DevScene sc, const float4* __restrict__ ray_o, const float4* __restrict__ ray_d,
                                        const HitRec* __restrict__ hits, uint32_t n, uint32_t seed_lo, uint32_t seed_hi,
                                        float4* out_o, float4* out_d, uint32_t* count) {
    const uint32_t total = ((n + 31) / 32) * 32;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        bool alive = false;
        V3 no, nd;
        if (i < n) {
            const HitRec h = hits[i];
            if (h.slot >= 0 && h.slot < 0x40000000) {
                const V3 o = ld3(ray_o[i]), draw = ld3(ray_d[i]);
                const V3 dn = normalize(draw);
                const V3 P = add(o, mul(dn, h.t));
                const V3 a = ld3(sc.tris[3 * (size_t)h.slot]), b = ld3(sc.tris[3 * (size_t)h.slot + 1]), c = ld3(sc.tris[3 * (size_t)h.slot + 2]);
                V3 nrm = normalize(normalize(cross(sub(b, a), sub(c, a))));
                if (dot(nrm, draw) > 0.0f) nrm = neg(nrm);
                U4 r = philox4x32_10(U4{i, 0u, 0u, 0u}, seed_lo, seed_hi);
                const V3 w = cosine_sample_hemisphere(u01(r.x), u01(r.y));
                V3 t, bt;
                onb(nrm, t, bt);
                nd = add(add(mul(t, w.x), mul(bt, w.y)), mul(nrm, w.z));
                no = add(P, mul(nrm, 1e-4f));
                alive = true;
            }
        }
        const unsigned m = __ballot_sync(0xffffffffu, alive);
        if (m) {
            const int lane = threadIdx.x & 31;
            uint32_t base = 0;
            if (lane == __ffs(m) - 1) base = atomicAdd(count, (uint32_t)__popc(m));
            base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
            if (alive) {
                const uint32_t k = base + __popc(m & ((1u << lane) - 1u));
                out_o[k] = make_float4(no.x, no.y, no.z, 0.f);
                out_d[k] = make_float4(nd.x, nd.y, nd.z, 0.f);
            }
        }
    }
}
void k_make_rays_bounce(const DevScene& sc, const float4* d_ray_o, const float4* d_ray_d, const HitRec* d_hits, uint32_t n,
                        uint32_t seed_lo, uint32_t seed_hi, float4* d_out_o, float4* d_out_d, uint32_t* d_count, cudaStream_t st) {
    if (n) make_rays_bounce_kernel<<<grid_for(n, 256, sm_count() * 16), 256, 0, st>>>(sc, d_ray_o, d_ray_d, d_hits, n, seed_lo, seed_hi,
                                                                              d_out_o, d_out_d, d_count);
}

__global__ void flush_kernel(float4* buf, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        buf[i] = make_float4((float)i, 0.f, 0.f, 0.f);
}
void k_flush(float4* d_buf, size_t n_float4, cudaStream_t st) {
    if (n_float4) flush_kernel<<<sm_count() * 8, 256, 0, st>>>(d_buf, n_float4);
}

}  // namespace ff
