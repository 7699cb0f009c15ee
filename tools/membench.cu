// Read-bandwidth ceiling probes for the streaming kernels (tuning aid, not product code):
//   ldg   : grid-stride 128-bit loads, xor-reduced (defeats dead-code elimination)
//   tma   : producer-only ring of cp.async.bulk tiles, consumers just release the stage
// Usage: membench <MB> <iters>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include "../semantic-segmentation-server_b200/csrc/ssb_common.cuh"
namespace ssb { char *last_error_buffer() { static char b[512]; return b; } }
using namespace ssb;

template <int UNROLL>
__global__ void __launch_bounds__(256) ldg_kernel(const uint4 *p, size_t n, uint32_t *sink) {
    uint32_t acc = 0;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + (UNROLL - 1) * stride < n; i += UNROLL * stride) {
        uint4 v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u)
            asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(p + i + u * stride));
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
    }
    if (acc == 0x12345678u) *sink = acc;
}

template <int TILE, int STAGES, int NW>
__global__ void __launch_bounds__((NW + 1) * 32) tma_kernel(const uint8_t *p, long long n_tiles, uint32_t *sink, int contiguous) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + STAGES * TILE);
    uint64_t *empty = full + STAGES;
    int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) { for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], NW); } mbar_init_fence(); }
    __syncthreads();
    long long t0 = contiguous ? n_tiles * blockIdx.x / gridDim.x : blockIdx.x;
    long long t1 = contiguous ? n_tiles * (blockIdx.x + 1) / gridDim.x : n_tiles;
    long long step = contiguous ? 1 : gridDim.x;
    if (warp == NW) {
        if (lane == 0) {
            uint64_t pol = l2_policy_evict_first();
            int s = 0; uint32_t ph = 0;
            for (long long t = t0; t < t1; t += step) {
                mbar_wait(&empty[s], ph ^ 1);
                mbar_arrive_expect_tx(&full[s], TILE);
                bulk_copy_g2s(smem + s * TILE, p + t * TILE, TILE, &full[s], pol);
                if (++s == STAGES) { s = 0; ph ^= 1; }
            }
        }
        return;
    }
    int s = 0; uint32_t ph = 0; uint32_t acc = 0;
    for (long long t = t0; t < t1; t += step) {
        mbar_wait(&full[s], ph);
        acc ^= reinterpret_cast<uint32_t *>(smem + s * TILE)[tid];
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
        if (++s == STAGES) { s = 0; ph ^= 1; }
    }
    if (acc == 0x12345678u) *sink = acc;
}

template <typename F> float timeit(F f, int iters) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int i = 0; i < 3; ++i) f(i);
    cudaEventRecord(a);
    for (int i = 0; i < iters; ++i) f(i);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    cudaError_t e = cudaGetLastError(); if (e != cudaSuccess) printf("CUDA error %s\n", cudaGetErrorString(e));
    return ms / iters;
}

template <int TILE, int STAGES, int NW> void run_tma(const uint8_t *d, size_t bytes, int nbuf, uint32_t *sink, int sms, int iters) {
    auto k = tma_kernel<TILE, STAGES, NW>;
    int smem = STAGES * TILE + 2 * STAGES * 8;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    int occ = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k, (NW + 1) * 32, smem);
    long long n_tiles = bytes / TILE;
    for (int cap = 1; cap <= occ && cap <= 8; ++cap) for (int contiguous = 0; contiguous < 2; ++contiguous) {
        int grid = sms * cap;
        float ms = timeit([&](int i) { k<<<grid, (NW + 1) * 32, smem>>>(d + (size_t)(i % nbuf) * bytes, n_tiles, sink, contiguous); }, iters);
        printf("tma tile=%d stages=%d ctas/sm=%d (occ %d) %s: %.2f us  %.1f GB/s\n", TILE, STAGES, cap, occ, contiguous ? "contig " : "strided", ms * 1e3, n_tiles * (double)TILE / ms / 1e6);
    }
}

int main(int argc, char **argv) {
    size_t mb = argc > 1 ? atoi(argv[1]) : 354; int iters = argc > 2 ? atoi(argv[2]) : 40;
    size_t bytes = mb << 20; int nbuf = 4;
    uint8_t *d; cudaMalloc(&d, bytes * nbuf); cudaMemset(d, 1, bytes * nbuf);
    uint32_t *sink; cudaMalloc(&sink, 4);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    for (int mult : {2, 4, 8, 16, 32}) {
        float ms = timeit([&](int i) { ldg_kernel<8><<<sms * mult, 256>>>((const uint4 *)(d + (size_t)(i % nbuf) * bytes), bytes / 16, sink); }, iters);
        printf("ldg unroll8 grid=%dxSM: %.2f us  %.1f GB/s\n", mult, ms * 1e3, bytes / ms / 1e6);
    }
    run_tma<10752, 4, 4>(d, bytes, nbuf, sink, sms, iters);
    run_tma<10752, 8, 4>(d, bytes, nbuf, sink, sms, iters);
    run_tma<21504, 4, 4>(d, bytes, nbuf, sink, sms, iters);
    run_tma<32768, 4, 4>(d, bytes, nbuf, sink, sms, iters);
    run_tma<16384, 6, 4>(d, bytes, nbuf, sink, sms, iters);
    run_tma<8192, 8, 4>(d, bytes, nbuf, sink, sms, iters);
    // copy (the MEASURED_PEAKS method): read+write
    uint8_t *e; cudaMalloc(&e, bytes);
    float ms = timeit([&](int i) { cudaMemcpyAsync(e, d + (size_t)(i % nbuf) * bytes, bytes, cudaMemcpyDeviceToDevice); }, iters);
    printf("cudaMemcpy D2D: %.2f us  %.1f GB/s (read+write)\n", ms * 1e3, 2.0 * bytes / ms / 1e6);
    return 0;
}
