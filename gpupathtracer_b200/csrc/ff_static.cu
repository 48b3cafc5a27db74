// ff_static.cu -- ff_hits_unpack_static: ff_hits_unpack (ff_api.cpp) for callers that may not allocate.
//
// No counterpart in the reference.  Some callers cannot call cudaMalloc (code
// under CUDA graph capture, tight serving loops, harnesses that forbid it).  ff_hits_unpack turns device hit records into
// the reference's numbering through three device temporaries (geometryIndex, triangleIndex, t; 12 bytes per record) that it
// cudaMallocs and frees on every call.  This variant keeps those temporaries in ONE file-scope `static __device__` array of
// FF_HITS_UNPACK_STATIC_MAX records (12 MiB): every GPU that loads the library holds it until the process exits, and the
// entry point calls no cudaMalloc / cudaFree / allocating library routine.
//
// The array's NAME is a host-side stand-in in host code, not a device address: the address the kernel gets comes from
// cudaGetSymbolAddress, asked once per device with that device current (below).  Every call on a device shares the array
// and finds the previous call's contents in it; nothing in it has to start at zero (the kernel writes every element the
// copies read, for every output the caller asked for), so nothing is cleared.  Calls that overlap on one GPU (two host
// threads, two contexts on the same device) race on the array: serialise them.
#include <mutex>

#include "ff_host.h"

namespace {

constexpr size_t kMaxRecords = FF_HITS_UNPACK_STATIC_MAX;
// [0, kMax) geometryIndex | [kMax, 2 kMax) triangleIndex | [2 kMax, 3 kMax) t (float bits)
static __device__ __align__(256) uint32_t g_unpack_scratch[3 * kMaxRecords];

constexpr int kMaxDevices = 64;
std::mutex g_lock;
uint32_t* g_scratch_addr[kMaxDevices] = {nullptr};

// Device address of the scratch array on the CURRENT device (each device has its own copy of a __device__ variable).
cudaError_t scratch_on_current_device(int device, uint32_t** out) {
    std::lock_guard<std::mutex> hold(g_lock);
    if (!g_scratch_addr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_unpack_scratch);
        if (e != cudaSuccess) return e;
        g_scratch_addr[device] = (uint32_t*)p;
    }
    *out = g_scratch_addr[device];
    return cudaSuccess;
}

// Same arithmetic as unpack_hits_kernel (ff_trace.cu): ord -> the reference's (geometryIndex, triangleIndex), key -> t.
__global__ void unpack_hits_static_kernel(const int* __restrict__ ord_geom, const int* __restrict__ ord_tri,
                                          const ff::HitRec* __restrict__ hits, uint32_t n, uint32_t* scratch,
                                          bool want_geom, bool want_tri, bool want_t) {
    int* geom = (int*)scratch;
    int* tri = (int*)(scratch + kMaxRecords);
    float* t = (float*)(scratch + 2 * kMaxRecords);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const ff::HitRec h = hits[i];
        if (want_geom) geom[i] = h.ord >= 0 ? ord_geom[h.ord] : -1;
        if (want_tri) tri[i] = h.ord >= 0 ? ord_tri[h.ord] : -1;
        if (want_t) t[i] = h.key;
    }
}

}  // namespace

extern "C" int ff_hits_unpack_static(ff_scene* scene, const void* d_hits, size_t n, int* out_geom, int* out_tri, float* out_t) {
    FF_REQUIRE(scene, "ff_hits_unpack_static: NULL scene");
    FF_REQUIRE(scene->ctx, "ff_hits_unpack_static: the scene's context was destroyed");
    FF_REQUIRE(scene->committed, "ff_hits_unpack_static: scene not committed");
    if (n > kMaxRecords) {
        ff::set_error("ff_hits_unpack_static: " + std::to_string(n) + " records exceed the static scratch (FF_HITS_UNPACK_STATIC_MAX = " +
                      std::to_string(kMaxRecords) + "); split the batch or use ff_hits_unpack");
        return FF_ERR_INVALID;
    }
    FF_REQUIRE(d_hits || n == 0, "ff_hits_unpack_static: NULL buffer");
    const int device = scene->ctx->device;
    FF_REQUIRE(device >= 0 && device < kMaxDevices, "ff_hits_unpack_static: device index out of range");
    FF_CUDA(cudaSetDevice(device));
    if (n == 0) return FF_OK;
    uint32_t* scratch = nullptr;
    FF_CUDA(scratch_on_current_device(device, &scratch));
    cudaStream_t st = scene->ctx->stream;
    const int block = 256;
    int grid = (int)((n + block - 1) / block);
    const int cap = ff::sm_count() * 16;
    if (grid > cap) grid = cap;
    unpack_hits_static_kernel<<<grid, block, 0, st>>>(scene->dev.ord_geom, scene->dev.ord_tri, (const ff::HitRec*)d_hits, (uint32_t)n, scratch,
                                                      out_geom != nullptr, out_tri != nullptr, out_t != nullptr);
    FF_CUDA(cudaGetLastError());
    if (out_geom) FF_CUDA(cudaMemcpyAsync(out_geom, scratch, sizeof(int) * n, cudaMemcpyDeviceToHost, st));
    if (out_tri) FF_CUDA(cudaMemcpyAsync(out_tri, scratch + kMaxRecords, sizeof(int) * n, cudaMemcpyDeviceToHost, st));
    if (out_t) FF_CUDA(cudaMemcpyAsync(out_t, scratch + 2 * kMaxRecords, sizeof(float) * n, cudaMemcpyDeviceToHost, st));
    FF_CUDA(cudaStreamSynchronize(st));
    return FF_OK;
}
