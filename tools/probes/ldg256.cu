// Probe: divergent gathers of BVH-node-sized records from an L2-resident array on sm_100a --
// 5 x LDG.128 from 80-byte records vs 3 x LDG.256 from 96-byte (32-byte aligned) records, and
// 3 x LDG.128 from 48-byte triangles vs 2 x LDG.256 from 64-byte triangles.  Prints ms per variant.
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
__device__ __forceinline__ uint32_t hash(uint32_t x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }
struct F8 { float v[8]; };
__device__ __forceinline__ F8 ld256(const void* p) {
    F8 r;
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7]) : "l"(p));
    return r;
}
template <int REC, int MODE>  // MODE 0: LDG.128 x (REC/16); MODE 1: LDG.256 x (REC/32)
__global__ void gather(const char* __restrict__ base, uint32_t n_rec, int iters, float* out) {
    uint32_t idx = hash(blockIdx.x * blockDim.x + threadIdx.x);
    float acc = 0.f;
    for (int it = 0; it < iters; ++it) {
        const char* p = base + (size_t)(idx % n_rec) * REC;
        if (MODE == 0) {
#pragma unroll
            for (int k = 0; k < REC / 16; ++k) { float4 v = __ldg((const float4*)(p + 16 * k)); acc += v.x + v.y + v.z + v.w; }
        } else {
#pragma unroll
            for (int k = 0; k < REC / 32; ++k) { F8 v = ld256(p + 32 * k); for (int j = 0; j < 8; ++j) acc += v.v[j]; }
        }
        idx = hash(idx + __float_as_uint(acc) % 3u);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
template <int REC, int MODE>
float run(const char* d, uint32_t n_rec, float* out) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    gather<REC, MODE><<<148 * 8, 128>>>(d, n_rec, 64, out);
    cudaEventRecord(a);
    gather<REC, MODE><<<148 * 8, 128>>>(d, n_rec, 512, out);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); return ms;
}
int main() {
    const size_t bytes = 64u << 20;
    char* d; float* out;
    cudaMalloc(&d, bytes); cudaMemset(d, 0, bytes); cudaMalloc(&out, 148 * 8 * 128 * 4);
    const double g = 148.0 * 8 * 128 * 512;
    float t;
    t = run<80, 0>(d, 170000, out); printf("node  80B 5xLDG.128: %.3f ms  %.2f Ggather/s\n", t, g / t / 1e6);
    t = run<96, 0>(d, 170000, out); printf("node  96B 6xLDG.128: %.3f ms  %.2f Ggather/s\n", t, g / t / 1e6);
    t = run<96, 1>(d, 170000, out); printf("node  96B 3xLDG.256: %.3f ms  %.2f Ggather/s\n", t, g / t / 1e6);
    t = run<64, 1>(d, 170000, out); printf("node  64B 2xLDG.256: %.3f ms  %.2f Ggather/s\n", t, g / t / 1e6);
    t = run<48, 0>(d, 1000000, out); printf("tri   48B 3xLDG.128: %.3f ms  %.2f Ggather/s\n", t, g / t / 1e6);
    t = run<64, 0>(d, 1000000, out); printf("tri   64B 4xLDG.128: %.3f ms  %.2f Ggather/s\n", t, g / t / 1e6);
    t = run<64, 1>(d, 1000000, out); printf("tri   64B 2xLDG.256: %.3f ms  %.2f Ggather/s\n", t, g / t / 1e6);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
