// Experiment: are per-SM streaming speeds stable from launch to launch?  TMA ring (as the product
// kernel: tile 21504 B, 3 stages, 8 consumer warps that read the tile), contiguous spans, 2 CTAs/SM;
// records (smid, cycles) per CTA for several launches.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include "../semantic-segmentation-server_b200/csrc/ssb_common.cuh"
namespace ssb { char *last_error_buffer() { static char b[512]; return b; } }
using namespace ssb;

constexpr int TILE = 21504, STAGES = 3, NW = 8;
__global__ void __launch_bounds__((NW + 1) * 32) k(const uint8_t *p, long long n_tiles, uint32_t *sink, unsigned long long *rec) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + STAGES * TILE);
    uint64_t *empty = full + STAGES;
    int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    long long c0 = clock64();
    if (tid == 0) { for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], NW); } mbar_init_fence(); }
    __syncthreads();
    long long t0 = n_tiles * blockIdx.x / gridDim.x, t1 = n_tiles * (blockIdx.x + 1) / gridDim.x;
    if (warp == NW) {
        if (lane == 0) {
            uint64_t pol = l2_policy_evict_first();
            int s = 0; uint32_t ph = 0;
            for (long long t = t0; t < t1; ++t) {
                mbar_wait(&empty[s], ph ^ 1);
                mbar_arrive_expect_tx(&full[s], TILE);
                bulk_copy_g2s(smem + s * TILE, p + t * TILE, TILE, &full[s], pol);
                if (++s == STAGES) { s = 0; ph ^= 1; }
            }
        }
        return;
    }
    int s = 0; uint32_t ph = 0; uint32_t acc = 0;
    for (long long t = t0; t < t1; ++t) {
        mbar_wait(&full[s], ph);
        const uint32_t *src = reinterpret_cast<const uint32_t *>(smem + s * TILE) + (warp * 32 + lane) * 21;
#pragma unroll
        for (int c = 0; c < 21; ++c) acc ^= src[c];
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
        if (++s == STAGES) { s = 0; ph ^= 1; }
    }
    if (acc == 0x12345678u) *sink = acc;
    if (tid == 0) {
        unsigned smid; asm("mov.u32 %0, %%smid;" : "=r"(smid));
        rec[blockIdx.x] = ((unsigned long long)smid << 48) | (unsigned long long)(clock64() - c0);
    }
}

int main() {
    size_t bytes = 353699136; int nbuf = 4;
    uint8_t *d; cudaMalloc(&d, bytes * nbuf); cudaMemset(d, 1, bytes * nbuf);
    uint32_t *sink; cudaMalloc(&sink, 4);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int grid = sms * 2; int smem = STAGES * TILE + 2 * STAGES * 8;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    unsigned long long *rec; cudaMalloc(&rec, grid * 8 * 8);
    const int L = 6;
    for (int i = 0; i < L; ++i) k<<<grid, (NW + 1) * 32, smem>>>(d + (size_t)(i % nbuf) * bytes, bytes / TILE, sink, rec + i * grid);
    cudaDeviceSynchronize();
    printf("err=%s\n", cudaGetErrorString(cudaGetLastError()));
    std::vector<unsigned long long> h(grid * L); cudaMemcpy(h.data(), rec, grid * L * 8, cudaMemcpyDeviceToHost);
    // per launch: min/avg/max cycles; per SM: mean over launches 2.. of the max of its CTAs
    std::vector<double> smsum(256, 0); std::vector<int> smcnt(256, 0);
    for (int i = 0; i < L; ++i) {
        double mn = 1e18, mx = 0, sum = 0;
        for (int c = 0; c < grid; ++c) { double cyc = (double)(h[i * grid + c] & 0xFFFFFFFFFFFFull); mn = std::min(mn, cyc); mx = std::max(mx, cyc); sum += cyc;
            if (i >= 2) { int sm = h[i * grid + c] >> 48; smsum[sm] += cyc; smcnt[sm]++; } }
        printf("launch %d: min %.0f avg %.0f max %.0f  (max/avg %.3f)\n", i, mn, sum / grid, mx, mx / (sum / grid));
    }
    // stability: correlation of per-SM cycles between launch 2 and launch 5
    std::vector<double> a(256, 0), b(256, 0);
    for (int c = 0; c < grid; ++c) { a[h[2 * grid + c] >> 48] += (double)(h[2 * grid + c] & 0xFFFFFFFFFFFFull); b[h[5 * grid + c] >> 48] += (double)(h[5 * grid + c] & 0xFFFFFFFFFFFFull); }
    double ma = 0, mb = 0; int n = 0; for (int s2 = 0; s2 < 256; ++s2) if (smcnt[s2]) { ma += a[s2]; mb += b[s2]; ++n; }
    ma /= n; mb /= n; double sab = 0, saa = 0, sbb = 0;
    for (int s2 = 0; s2 < 256; ++s2) if (smcnt[s2]) { sab += (a[s2] - ma) * (b[s2] - mb); saa += (a[s2] - ma) * (a[s2] - ma); sbb += (b[s2] - mb) * (b[s2] - mb); }
    printf("per-SM correlation launch2 vs launch5: %.3f over %d SMs\n", sab / sqrt(saa * sbb), n);
    printf("per-SM mean cycles (launches 2..5), by smid:\n");
    for (int s2 = 0; s2 < 256; ++s2) if (smcnt[s2]) printf("%d:%.0f ", s2, smsum[s2] / smcnt[s2]);
    printf("\n");
    return 0;
}
