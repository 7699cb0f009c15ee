// Generic (any C, any alignment) argmax + per-label statistics kernel.  Correct for every shape the
// C-ABI accepts; used when the streaming fast path does not apply (C not compiled in, misaligned
// views, tiny frames).  One CTA handles a chunk of ONE frame: scalar loads, block accumulators in
// shared memory via 64-bit atomics, one global flush per block.  The element of pixel r, channel c
// of a frame sits at r*px_stride + c*ch_stride: (C, 1) for [H,W,C] frames, (1, H*W) for planar
// [C,H,W] frames (argmax_planar.cu).
#pragma once
#include <cuda_fp16.h>

#include "ssb_common.cuh"

namespace ssb {

struct ElemU8 {
    using T = uint8_t;
    using V = uint32_t;
    static __device__ __forceinline__ V load(const T *p) { return *p; }
    static __device__ __forceinline__ bool gt(V a, V b) { return a > b; }
    static __device__ __forceinline__ unsigned long long fixed(const T *p) { return *p; }
};

// value * 2^24 of a finite fp16 bit pattern, exact, as two's complement.
__host__ __device__ inline long long half_bits_to_fixed24(uint32_t h) {
    const long long e = (h >> 10) & 0x1f, m = h & 0x3ff;
    const long long v = e == 0 ? m : ((m | 0x400) << (e - 1));
    return (h & 0x8000u) ? -v : v;
}

struct ElemF16 {
    using T = uint16_t;
    using V = float;
    static __device__ __forceinline__ V load(const T *p) { return __half2float(__ushort_as_half(*p)); }
    static __device__ __forceinline__ bool gt(V a, V b) { return a > b; }
    static __device__ __forceinline__ unsigned long long fixed(const T *p) {
        return static_cast<unsigned long long>(half_bits_to_fixed24(*p));
    }
};

// fp32 scores: Σq adds round-to-nearest-even(value * 2^24) (exact whenever the value is a multiple
// of 2^-24, e.g. every fp16-representable value; |value| < 2^39 keeps it inside an int64).
struct ElemF32 {
    using T = float;
    using V = float;
    static __device__ __forceinline__ V load(const T *p) { return *p; }
    static __device__ __forceinline__ bool gt(V a, V b) { return a > b; }
    static __device__ __forceinline__ unsigned long long fixed(const T *p) {
        return static_cast<unsigned long long>(__float2ll_rn(*p * 16777216.0f));
    }
};

constexpr int kGenericThreads = 256;
constexpr int kGenericMaxC = 255;

template <typename E>
__global__ void __launch_bounds__(kGenericThreads)
argmax_stats_generic_kernel(const typename E::T *logits, uint8_t *labels, unsigned long long *stats, int hw, int w,
                            int C, int roi_h, int roi_w, int chunk_px, long long px_stride, long long ch_stride) {
    __shared__ unsigned long long blk[kGenericMaxC * 4];
    for (int i = threadIdx.x; i < C * 4; i += kGenericThreads) blk[i] = 0ull;
    __syncthreads();
    const long long f = blockIdx.y;
    const int r_begin = blockIdx.x * chunk_px, r_end = min(hw, r_begin + chunk_px);
    // Each thread keeps the sums of its current label in registers and only touches the block
    // accumulators when the label changes: segmentation maps have long same-label runs, and a thread's
    // consecutive pixels (256 apart, usually the same row) mostly share a label.
    int cur = -1;
    unsigned long long cn = 0, cx = 0, cy = 0, cq = 0;
    auto spill = [&]() {
        if (cn) {
            atomicAdd(&blk[cur * 4 + SSB_STAT_N], cn);
            atomicAdd(&blk[cur * 4 + SSB_STAT_SX], cx);
            atomicAdd(&blk[cur * 4 + SSB_STAT_SY], cy);
            atomicAdd(&blk[cur * 4 + SSB_STAT_SQ], cq);
        }
        cn = cx = cy = cq = 0;
    };
    for (int r = r_begin + threadIdx.x; r < r_end; r += kGenericThreads) {
        const typename E::T *q = logits + f * hw * C + r * px_stride;
        int best = 0;
        typename E::V bv = E::load(q);
        for (int c = 1; c < C; ++c) {
            const typename E::V v = E::load(q + c * ch_stride);
            if (E::gt(v, bv)) { bv = v; best = c; }
        }
        if (labels) labels[f * hw + r] = static_cast<uint8_t>(best);
        const int y = r / w, x = r - y * w;
        if (x < roi_w && y < roi_h) {
            if (best != cur) {
                spill();
                cur = best;
            }
            cn += 1;
            cx += static_cast<unsigned long long>(x);
            cy += static_cast<unsigned long long>(y);
            cq += E::fixed(q + best * ch_stride);
        }
    }
    spill();
    __syncthreads();
    for (int i = threadIdx.x; i < C * 4; i += kGenericThreads) {
        const unsigned long long v = blk[i];
        if (v) atomicAdd(&stats[f * C * 4 + i], v);
    }
}

template <typename E>
inline int launch_generic(const typename E::T *logits, int B, int H, int W, int C, int roi_h, int roi_w,
                          uint8_t *labels, unsigned long long *stats, cudaStream_t st, bool planar = false) {
    const int hw = H * W;
    const long long px_stride = planar ? 1 : C, ch_stride = planar ? hw : 1;
    const int chunk_px = 8192;
    dim3 grid((hw + chunk_px - 1) / chunk_px, 1);
    // gridDim.y is limited to 65535: launch in slabs of frames
    for (int b0 = 0; b0 < B; b0 += 65535) {
        grid.y = min(65535, B - b0);
        argmax_stats_generic_kernel<E><<<grid, kGenericThreads, 0, st>>>(
            logits + static_cast<long long>(b0) * hw * C, labels ? labels + static_cast<long long>(b0) * hw : nullptr,
            stats + static_cast<long long>(b0) * C * 4, hw, W, C, roi_h, roi_w, chunk_px, px_stride, ch_stride);
        SSB_CUDA_TRY(cudaGetLastError());
    }
    return SSB_OK;
}

// Shared argument validation for ssb_argmax_stats_{u8,f16} (limits documented in ssb200.h).
inline int check_argmax_args(const void *logits, int B, int H, int W, int C, int roi_h, int roi_w, const void *stats) {
    if (B < 0 || H <= 0 || W <= 0) return fail(SSB_ERR_BAD_SHAPE, "bad shape B=%d H=%d W=%d", B, H, W);
    if (C < 2 || C > kGenericMaxC) return fail(SSB_ERR_LIMIT, "C=%d outside [2,%d]", C, kGenericMaxC);
    if (roi_h < 1 || roi_h > H || roi_w < 1 || roi_w > W)
        return fail(SSB_ERR_BAD_SHAPE, "ROI (%d,%d) outside frame (%d,%d)", roi_h, roi_w, H, W);
    if (static_cast<long long>(H) * W >= (1ll << 31)) return fail(SSB_ERR_LIMIT, "H*W must be < 2^31");
    if (static_cast<long long>(B) * H * W >= (1ll << 40)) return fail(SSB_ERR_LIMIT, "B*H*W must be < 2^40");
    if (B > 0 && (!logits || !stats)) return fail(SSB_ERR_BAD_POINTER, "logits/stats must not be NULL");
    return SSB_OK;
}

}  // namespace ssb
