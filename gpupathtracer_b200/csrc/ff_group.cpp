// ff_group.cpp -- multi-GPU rendering behind the C ABI (include/firefly.h, "multi-GPU group"): the image is partitioned
// across the GPUs of one node by disjoint sample ranges or interleaved screen tiles, the scene + BVH are replicated, and
// the per-GPU accumulators are combined by ONE framebuffer reduce per frame over NVLink (BASELINE.json north_star;
// SURVEY.md 8e).  The reference is single-GPU: a dead `multi_gpu` flag is all it has (utilities.h:484-487).
//
// Two ways to form a group, one render entry for both:
//   ff_group_create        one process drives n devices (one ff_context + one host thread per device, ncclCommInitAll)
//   ff_group_create_rank   one process per GPU (torchrun / MPI style): rank 0 makes an id (ff_group_unique_id), the other
//                          ranks get the 128 bytes by any means, everybody calls ncclCommInitRank inside
// Two reduce paths:
//   FF_REDUCE_NCCL   ncclReduce(SUM) of the RGBA fp32 accumulators to the root, then the resolve kernel
//   FF_REDUCE_PEER   ONE kernel per GPU does reduce-scatter + resolve + gather over peer memory: GPU r sums pixel slice r of
//                    every accumulator (its own and the peers', read over NVLink: cudaDeviceEnablePeerAccess in a single
//                    process, CUDA IPC handles exchanged through an ncclAllGather between processes), in rank order,
//                    scales by 1/spp and stores the finished pixels straight into the root's frame.  Every GPU's NVLink
//                    ingress carries (world-1)/world of one accumulator instead of the root's carrying world-1 of them,
//                    and the summation order is fixed, so the frame is bit-reproducible for a given world size.
// NCCL is loaded with dlopen("libnccl.so.2") the first time a group of more than one GPU is formed (in a process that has
// already loaded PyTorch's NCCL the loader hands back that copy); a single-GPU user never needs it.
#include <dlfcn.h>

#include <algorithm>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "ff_host.h"

using namespace ff;

namespace {

// ---- the handful of NCCL entry points used, declared here so that the build needs no NCCL header (ABI of nccl.h 2.x)
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { ncclSuccess = 0 };
enum { ncclChar = 0, ncclFloat = 7 };
enum { ncclSum = 0 };
struct NcclApi {
    void* lib = nullptr;
    bool tried = false;
    std::string error;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Reduce)(const void*, void*, size_t, int, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    int (*GetVersion)(int*) = nullptr;
    bool load() {
        if (tried) return lib != nullptr;
        tried = true;
        const char* names[] = {getenv("FF_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
        for (const char* nm : names) {
            if (!nm || !*nm) continue;
            lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
            if (lib) break;
            error = dlerror();
        }
        if (!lib) { error = "cannot load NCCL (libnccl.so.2): " + error; return false; }
        auto sym = [&](const char* s) { void* p = dlsym(lib, s); if (!p) error = std::string("NCCL symbol missing: ") + s; return p; };
        *(void**)&GetUniqueId = sym("ncclGetUniqueId");
        *(void**)&CommInitRank = sym("ncclCommInitRank");
        *(void**)&CommInitAll = sym("ncclCommInitAll");
        *(void**)&CommDestroy = sym("ncclCommDestroy");
        *(void**)&Reduce = sym("ncclReduce");
        *(void**)&AllReduce = sym("ncclAllReduce");
        *(void**)&AllGather = sym("ncclAllGather");
        *(void**)&GroupStart = sym("ncclGroupStart");
        *(void**)&GroupEnd = sym("ncclGroupEnd");
        *(void**)&GetErrorString = sym("ncclGetErrorString");
        *(void**)&GetVersion = sym("ncclGetVersion");
        if (!GetUniqueId || !CommInitRank || !CommInitAll || !CommDestroy || !Reduce || !AllReduce || !AllGather || !GroupStart || !GroupEnd ||
            !GetErrorString) {
            dlclose(lib); lib = nullptr;
            return false;
        }
        return true;
    }
};
NcclApi& nccl() { static NcclApi api; return api; }

#define FF_NCCL(call)                                                                                          \
    do {                                                                                                       \
        ncclResult_t r_ = (call);                                                                              \
        if (r_ != ncclSuccess) {                                                                               \
            ff::set_error(std::string(#call) + ": " + nccl().GetErrorString(r_));                              \
            return FF_ERR_CUDA;                                                                                \
        }                                                                                                      \
    } while (0)

}  // namespace

struct ff_group {
    int world = 1;           // GPUs in the group
    int n_local = 1;         // contexts this process drives
    int first_rank = 0;      // global rank of local context 0 (ranks of a single-process group are its device order)
    bool single_process = true;
    bool owns_contexts = false;
    std::vector<ff_context*> ctx;
    std::vector<ncclComm_t> comm;
    // per-frame buffers, (re)allocated when the image size changes
    size_t npix = 0;
    std::vector<float4*> accum;        // one accumulator per local context (cudaMalloc base pointers: IPC handles name whole allocations)
    float4* out_root = nullptr;        // resolved frame on the root's device (only where the root is local)
    std::vector<float*> dummy;         // 1 float per local context: the payload of the stream-ordered barrier
    // peer tables of the fused reduce: per local context, every rank's accumulator and the root's frame as pointers valid on that device
    bool peer_ready = false;
    std::vector<std::vector<const float4*>> peer_accum;
    std::vector<float4*> peer_out;
    std::vector<std::pair<int, void*>> ipc_opened;  // (local index, mapped pointer)
    std::vector<cudaEvent_t> ev_ready, ev_done, ev_t0, ev_t1;
    bool root_local() const { return first_rank == 0; }
};

namespace {

int set_device(ff_context* c) {
    FF_CUDA(cudaSetDevice(c->device));
    return FF_OK;
}

void close_peers(ff_group* g) {
    for (auto& kv : g->ipc_opened) {
        cudaSetDevice(g->ctx[kv.first]->device);
        cudaIpcCloseMemHandle(kv.second);
    }
    g->ipc_opened.clear();
    g->peer_accum.clear();
    g->peer_out.clear();
    g->peer_ready = false;
}

void free_buffers(ff_group* g) {
    close_peers(g);
    for (int i = 0; i < (int)g->accum.size(); ++i) {
        cudaSetDevice(g->ctx[i]->device);
        cudaStreamSynchronize(g->ctx[i]->stream);
        cudaFree(g->accum[i]);
        if (i == 0 && g->root_local()) cudaFree(g->out_root);
    }
    g->accum.clear();
    g->out_root = nullptr;
    g->npix = 0;
}

int ensure_buffers(ff_group* g, size_t npix) {
    if (g->npix == npix && !g->accum.empty()) return FF_OK;
    free_buffers(g);
    g->accum.assign(g->n_local, nullptr);
    for (int i = 0; i < g->n_local; ++i) {
        int rc = set_device(g->ctx[i]);
        if (rc) return rc;
        FF_CUDA(cudaMalloc((void**)&g->accum[i], sizeof(float4) * npix));
        if (i == 0 && g->root_local()) FF_CUDA(cudaMalloc((void**)&g->out_root, sizeof(float4) * npix));
    }
    g->npix = npix;
    return FF_OK;
}

// Stream-ordered barrier over the group's GPUs: work queued after it on any member's stream starts only when the work
// queued before it on EVERY member's stream has finished.  Single process: events across the devices' streams; one
// process per GPU: a 4-byte ncclAllReduce (the only transport the ranks share).
int group_barrier(ff_group* g, std::vector<cudaEvent_t>& ev) {
    if (g->world == 1) return FF_OK;
    if (g->single_process) {
        for (int i = 0; i < g->n_local; ++i) {
            int rc = set_device(g->ctx[i]);
            if (rc) return rc;
            FF_CUDA(cudaEventRecord(ev[i], g->ctx[i]->stream));
        }
        for (int j = 0; j < g->n_local; ++j) {
            int rc = set_device(g->ctx[j]);
            if (rc) return rc;
            for (int i = 0; i < g->n_local; ++i)
                if (i != j) FF_CUDA(cudaStreamWaitEvent(g->ctx[j]->stream, ev[i], 0));
        }
        return FF_OK;
    }
    int rc = set_device(g->ctx[0]);
    if (rc) return rc;
    FF_NCCL(nccl().AllReduce(g->dummy[0], g->dummy[0], 1, ncclFloat, ncclSum, g->comm[0], g->ctx[0]->stream));
    return FF_OK;
}

// Builds, for every local context, the table of all ranks' accumulators (and the root's frame) as pointers that device can
// dereference.  Collective in the one-process-per-GPU mode (the IPC handles travel through an ncclAllGather).
int ensure_peer_tables(ff_group* g) {
    if (g->peer_ready) return FF_OK;
    g->peer_accum.assign(g->n_local, std::vector<const float4*>(g->world, nullptr));
    g->peer_out.assign(g->n_local, nullptr);
    if (g->single_process) {
        for (int i = 0; i < g->n_local; ++i) {
            int rc = set_device(g->ctx[i]);
            if (rc) return rc;
            for (int j = 0; j < g->n_local; ++j) {
                if (i == j) continue;
                int can = 0;
                FF_CUDA(cudaDeviceCanAccessPeer(&can, g->ctx[i]->device, g->ctx[j]->device));
                if (!can) {
                    ff::set_error("ff_group_render: device " + std::to_string(g->ctx[i]->device) + " cannot access device " +
                                  std::to_string(g->ctx[j]->device) + " (no peer path); use FF_REDUCE_NCCL");
                    close_peers(g);
                    return FF_ERR_UNSUPPORTED;
                }
                cudaError_t e = cudaDeviceEnablePeerAccess(g->ctx[j]->device, 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
                else if (e != cudaSuccess) { ff::set_error(std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e)); close_peers(g); return FF_ERR_CUDA; }
            }
            for (int j = 0; j < g->n_local; ++j) g->peer_accum[i][j] = g->accum[j];
            g->peer_out[i] = g->out_root;
        }
        g->peer_ready = true;
        return FF_OK;
    }
    // one process per GPU: gather {accumulator handle, root-frame handle} of every rank
    ff_context* c = g->ctx[0];
    int rc = set_device(c);
    if (rc) return rc;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    unsigned char mine[128];
    memset(mine, 0, sizeof(mine));
    cudaIpcMemHandle_t h;
    FF_CUDA(cudaIpcGetMemHandle(&h, g->accum[0]));
    memcpy(mine, &h, 64);
    if (g->root_local()) { FF_CUDA(cudaIpcGetMemHandle(&h, g->out_root)); memcpy(mine + 64, &h, 64); }
    DevTmp d_send, d_recv;
    FF_CUDA(d_send.alloc(128));
    FF_CUDA(d_recv.alloc(128 * (size_t)g->world));
    FF_CUDA(cudaMemcpyAsync(d_send.p, mine, 128, cudaMemcpyHostToDevice, c->stream));
    FF_NCCL(nccl().AllGather(d_send.p, d_recv.p, 128, ncclChar, g->comm[0], c->stream));
    std::vector<unsigned char> all(128 * (size_t)g->world);
    FF_CUDA(cudaMemcpyAsync(all.data(), d_recv.p, all.size(), cudaMemcpyDeviceToHost, c->stream));
    FF_CUDA(cudaStreamSynchronize(c->stream));
    for (int r = 0; r < g->world; ++r) {
        if (r == g->first_rank) { g->peer_accum[0][r] = g->accum[0]; continue; }
        memcpy(&h, all.data() + 128 * (size_t)r, 64);
        void* p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) { ff::set_error(std::string("cudaIpcOpenMemHandle (accumulator of rank ") + std::to_string(r) + "): " + cudaGetErrorString(e)); cudaGetLastError(); close_peers(g); return FF_ERR_CUDA; }
        g->ipc_opened.emplace_back(0, p);
        g->peer_accum[0][r] = (const float4*)p;
    }
    if (g->root_local()) g->peer_out[0] = g->out_root;
    else {
        memcpy(&h, all.data() + 64, 64);
        void* p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) { ff::set_error(std::string("cudaIpcOpenMemHandle (root frame): ") + cudaGetErrorString(e)); cudaGetLastError(); close_peers(g); return FF_ERR_CUDA; }
        g->ipc_opened.emplace_back(0, p);
        g->peer_out[0] = (float4*)p;
    }
    g->peer_ready = true;
    return FF_OK;
}

int make_events(ff_group* g) {
    g->ev_ready.resize(g->n_local); g->ev_done.resize(g->n_local); g->ev_t0.resize(g->n_local); g->ev_t1.resize(g->n_local);
    g->dummy.assign(g->n_local, nullptr);
    for (int i = 0; i < g->n_local; ++i) {
        int rc = set_device(g->ctx[i]);
        if (rc) return rc;
        FF_CUDA(cudaEventCreateWithFlags(&g->ev_ready[i], cudaEventDisableTiming));
        FF_CUDA(cudaEventCreateWithFlags(&g->ev_done[i], cudaEventDisableTiming));
        FF_CUDA(cudaEventCreate(&g->ev_t0[i]));
        FF_CUDA(cudaEventCreate(&g->ev_t1[i]));
        FF_CUDA(cudaMalloc((void**)&g->dummy[i], 256));
        FF_CUDA(cudaMemset(g->dummy[i], 0, 256));
    }
    return FF_OK;
}

}  // namespace

extern "C" {

int ff_group_sample_range(int rank, int world, int spp, int* out_first, int* out_count) {
    FF_REQUIRE(world > 0 && rank >= 0 && rank < world && spp >= 0 && out_first && out_count, "ff_group_sample_range: bad argument");
    const long long a = (long long)rank * spp / world, b = (long long)(rank + 1) * spp / world;
    *out_first = (int)a;
    *out_count = (int)(b - a);
    return FF_OK;
}

int ff_group_unique_id(unsigned char id[FF_GROUP_ID_BYTES]) {
    FF_REQUIRE(id, "ff_group_unique_id: NULL argument");
    if (!nccl().load()) { ff::set_error(nccl().error); return FF_ERR_UNSUPPORTED; }
    static_assert(FF_GROUP_ID_BYTES == sizeof(ncclUniqueId), "NCCL unique id size");
    ncclUniqueId u;
    FF_NCCL(nccl().GetUniqueId(&u));
    memcpy(id, &u, sizeof(u));
    return FF_OK;
}

int ff_group_create(const int* device_ids, int n, ff_group** out) {
    FF_REQUIRE(out, "ff_group_create: out is NULL");
    *out = nullptr;
    FF_REQUIRE(device_ids && n >= 1 && n <= 8, "ff_group_create: 1 to 8 devices");
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < i; ++j) FF_REQUIRE(device_ids[i] != device_ids[j], "ff_group_create: a device is listed twice");
    if (n > 1 && !nccl().load()) { ff::set_error(nccl().error); return FF_ERR_UNSUPPORTED; }
    ff_group* g = new ff_group();
    g->world = n; g->n_local = n; g->first_rank = 0; g->single_process = true; g->owns_contexts = true;
    for (int i = 0; i < n; ++i) {
        ff_context* c = nullptr;
        int rc = ff_context_create(&c, device_ids[i]);
        if (rc) { ff_group_destroy(g); return rc; }
        g->ctx.push_back(c);
    }
    g->comm.assign(n, nullptr);
    if (n > 1) {
        ncclResult_t r = nccl().CommInitAll(g->comm.data(), n, device_ids);
        if (r != ncclSuccess) { ff::set_error(std::string("ncclCommInitAll: ") + nccl().GetErrorString(r)); g->comm.assign(n, nullptr); ff_group_destroy(g); return FF_ERR_CUDA; }
    }
    int rc = make_events(g);
    if (rc) { ff_group_destroy(g); return rc; }
    *out = g;
    return FF_OK;
}

int ff_group_create_rank(ff_context* ctx, const unsigned char id[FF_GROUP_ID_BYTES], int rank, int world, ff_group** out) {
    FF_REQUIRE(out, "ff_group_create_rank: out is NULL");
    *out = nullptr;
    FF_REQUIRE(ctx && world >= 1 && world <= 8 && rank >= 0 && rank < world, "ff_group_create_rank: bad context / rank / world (1 to 8 GPUs)");
    FF_REQUIRE(world == 1 || id, "ff_group_create_rank: a group of more than one GPU needs the id of ff_group_unique_id");
    if (world > 1 && !nccl().load()) { ff::set_error(nccl().error); return FF_ERR_UNSUPPORTED; }
    FF_CUDA(cudaSetDevice(ctx->device));
    ff_group* g = new ff_group();
    g->world = world; g->n_local = 1; g->first_rank = rank; g->single_process = (world == 1); g->owns_contexts = false;
    g->ctx.push_back(ctx);
    g->comm.assign(1, nullptr);
    if (world > 1) {
        ncclUniqueId u;
        memcpy(&u, id, sizeof(u));
        ncclResult_t r = nccl().CommInitRank(&g->comm[0], world, u, rank);
        if (r != ncclSuccess) { ff::set_error(std::string("ncclCommInitRank: ") + nccl().GetErrorString(r)); g->comm[0] = nullptr; ff_group_destroy(g); return FF_ERR_CUDA; }
    }
    int rc = make_events(g);
    if (rc) { ff_group_destroy(g); return rc; }
    *out = g;
    return FF_OK;
}

void ff_group_destroy(ff_group* g) {
    if (!g) return;
    free_buffers(g);
    for (int i = 0; i < (int)g->ctx.size(); ++i) {
        cudaSetDevice(g->ctx[i]->device);
        cudaStreamSynchronize(g->ctx[i]->stream);
        if (i < (int)g->ev_ready.size()) { cudaEventDestroy(g->ev_ready[i]); cudaEventDestroy(g->ev_done[i]); cudaEventDestroy(g->ev_t0[i]); cudaEventDestroy(g->ev_t1[i]); }
        if (i < (int)g->dummy.size()) cudaFree(g->dummy[i]);
        if (i < (int)g->comm.size() && g->comm[i]) nccl().CommDestroy(g->comm[i]);
    }
    if (g->owns_contexts)
        for (ff_context* c : g->ctx) ff_context_destroy(c);
    delete g;
}

int ff_group_world(const ff_group* g) { return g ? g->world : 0; }
int ff_group_local_count(const ff_group* g) { return g ? g->n_local : 0; }
int ff_group_rank(const ff_group* g, int local_index) { return (g && local_index >= 0 && local_index < g->n_local) ? g->first_rank + local_index : -1; }
ff_context* ff_group_context(ff_group* g, int local_index) { return (g && local_index >= 0 && local_index < g->n_local) ? g->ctx[local_index] : nullptr; }
int ff_group_nccl_version(void) {
    int v = 0;
    if (nccl().load() && nccl().GetVersion) nccl().GetVersion(&v);
    return v;
}

int ff_group_render(ff_group* g, ff_scene* const* scenes, const ff_camera* cam, const ff_render_params* params, int split, int reduce,
                    float* out_rgba_host, void* out_rgba_device, ff_stats* stats, double* out_reduce_ms) {
    FF_REQUIRE(g && scenes && cam && params, "ff_group_render: NULL argument");
    FF_REQUIRE(split == FF_SPLIT_SAMPLES || split == FF_SPLIT_TILES, "ff_group_render: unknown split");
    FF_REQUIRE(reduce == FF_REDUCE_NCCL || reduce == FF_REDUCE_PEER, "ff_group_render: unknown reduce");
    FF_REQUIRE(params->tile_world <= 1, "ff_group_render: the group sets the tile split itself");
    const int nl = g->n_local;
    for (int i = 0; i < nl; ++i) {
        FF_REQUIRE(scenes[i] && scenes[i]->ctx == g->ctx[i], "ff_group_render: scenes[i] must live on the group's i-th context");
        int rc = host_check_render(scenes[i], cam, params, "ff_group_render");
        if (rc) return rc;
    }
    if (g->root_local()) FF_REQUIRE(out_rgba_host || out_rgba_device, "ff_group_render: the root needs an output buffer");
    const size_t npix = (size_t)cam->width * cam->height;
    int rc = ensure_buffers(g, npix);
    if (rc) return rc;
    const int total_spp = params->total_spp > 0 ? params->total_spp : params->spp;
    if (stats) memset(stats, 0, sizeof(ff_stats) * nl);
    if (out_reduce_ms) *out_reduce_ms = 0.0;

    // ---- every GPU renders its share into its own accumulator (sums, not means)
    std::vector<int> rcs(nl, FF_OK);
    std::vector<std::string> errs(nl);
    std::vector<ff_stats> st(nl);
    auto work = [&](int i) {
        ff_context* c = g->ctx[i];
        ff_render_params p = *params;
        const int rank = g->first_rank + i;
        if (split == FF_SPLIT_SAMPLES) {
            int first = 0, count = 0;
            ff_group_sample_range(rank, g->world, params->spp, &first, &count);
            p.first_sample = params->first_sample + first;
            p.spp = count;
        } else if (g->world > 1) {
            p.tile_rank = rank; p.tile_world = g->world;
        }
        memset(&st[i], 0, sizeof(ff_stats));
        cudaError_t e = cudaSetDevice(c->device);
        if (e == cudaSuccess) e = cudaMemsetAsync(g->accum[i], 0, sizeof(float4) * npix, c->stream);
        if (e != cudaSuccess) { rcs[i] = FF_ERR_CUDA; errs[i] = cudaGetErrorString(e); return; }
        if (p.spp > 0) {
            rcs[i] = host_accumulate_samples(scenes[i], cam, &p, g->accum[i], &st[i]);
            if (rcs[i]) errs[i] = ff::get_error();
        }
    };
    if (nl == 1) work(0);
    else {
        std::vector<std::thread> th;
        for (int i = 0; i < nl; ++i) th.emplace_back(work, i);
        for (auto& t : th) t.join();
    }
    for (int i = 0; i < nl; ++i)
        if (rcs[i]) { ff::set_error("ff_group_render (rank " + std::to_string(g->first_rank + i) + "): " + errs[i]); return rcs[i]; }
    if (stats) memcpy(stats, st.data(), sizeof(ff_stats) * nl);

    // ---- one framebuffer reduce + resolve
    const float inv = 1.0f / (float)total_spp;
    ff_context* root = g->root_local() ? g->ctx[0] : nullptr;
    for (int i = 0; i < nl; ++i) {
        if ((rc = set_device(g->ctx[i]))) return rc;
        FF_CUDA(cudaEventRecord(g->ev_t0[i], g->ctx[i]->stream));
    }
    float4* d_frame = nullptr;  // where the resolved frame ends up on the root's device
    if (g->world == 1) {
        if ((rc = set_device(root))) return rc;
        d_frame = out_rgba_device ? (float4*)out_rgba_device : g->out_root;
        const float4* one[1] = {g->accum[0]};
        k_resolve(one, 1, (uint32_t)npix, inv, d_frame, root->stream);
    } else if (reduce == FF_REDUCE_NCCL) {
        FF_NCCL(nccl().GroupStart());
        for (int i = 0; i < nl; ++i) {
            ncclResult_t r = nccl().Reduce(g->accum[i], g->accum[i], npix * 4, ncclFloat, ncclSum, 0, g->comm[i], g->ctx[i]->stream);
            if (r != ncclSuccess) { nccl().GroupEnd(); ff::set_error(std::string("ncclReduce: ") + nccl().GetErrorString(r)); return FF_ERR_CUDA; }
        }
        FF_NCCL(nccl().GroupEnd());
        if (root) {
            if ((rc = set_device(root))) return rc;
            d_frame = out_rgba_device ? (float4*)out_rgba_device : g->out_root;
            const float4* one[1] = {g->accum[0]};
            k_resolve(one, 1, (uint32_t)npix, inv, d_frame, root->stream);
        }
    } else {
        if ((rc = ensure_peer_tables(g))) return rc;
        if ((rc = group_barrier(g, g->ev_ready))) return rc;          // every accumulator is complete
        for (int i = 0; i < nl; ++i) {
            if ((rc = set_device(g->ctx[i]))) return rc;
            const int rank = g->first_rank + i;
            const size_t a = npix * (size_t)rank / g->world, b = npix * (size_t)(rank + 1) / g->world;
            k_reduce_scatter_resolve(g->peer_accum[i].data(), g->world, (uint32_t)a, (uint32_t)(b - a), inv, g->peer_out[i], g->ctx[i]->stream);
        }
        if ((rc = group_barrier(g, g->ev_done))) return rc;           // every slice of the root's frame is written
        if (root) {
            if ((rc = set_device(root))) return rc;
            d_frame = g->out_root;
            if (out_rgba_device) FF_CUDA(cudaMemcpyAsync(out_rgba_device, g->out_root, sizeof(float4) * npix, cudaMemcpyDeviceToDevice, root->stream));
        }
    }
    for (int i = 0; i < nl; ++i) {
        if ((rc = set_device(g->ctx[i]))) return rc;
        FF_CUDA(cudaEventRecord(g->ev_t1[i], g->ctx[i]->stream));
    }
    if (root && out_rgba_host) {
        if ((rc = set_device(root))) return rc;
        FF_CUDA(cudaMemcpyAsync(out_rgba_host, d_frame, sizeof(float4) * npix, cudaMemcpyDeviceToHost, root->stream));
    }
    double reduce_ms = 0;
    for (int i = 0; i < nl; ++i) {
        if ((rc = set_device(g->ctx[i]))) return rc;
        FF_CUDA(cudaStreamSynchronize(g->ctx[i]->stream));
        FF_CUDA(cudaGetLastError());
        float ms = 0;
        FF_CUDA(cudaEventElapsedTime(&ms, g->ev_t0[i], g->ev_t1[i]));
        reduce_ms = std::max(reduce_ms, (double)ms);
        if (stats) { stats[i].gpu_ms += ms; stats[i].kernel_launches += 1; }
    }
    if (out_reduce_ms) *out_reduce_ms = reduce_ms;
    return FF_OK;
}

}  // extern "C"
