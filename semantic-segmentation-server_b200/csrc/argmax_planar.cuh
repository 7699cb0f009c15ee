// Planar ([B,C,H,W], "NCHW") variant of the fused argmax + per-label statistics kernel, for class
// scores that come straight out of a convolutional segmentation head on the GPU (a
// logits provider replacing the Edge TPU call of sem_seg_server.py:162; torchvision's DeepLabv3 returns
// fp32 -- fp16 under autocast -- [B,21,H,W]).  Same outputs as ssb_argmax_stats_{u8,f16}: the label map
// of sem_seg_server.py:162-170 and per label n, Σx, Σy, Σ winning score over the ROI.
//
// What the layout changes:
//   * a pixel's C scores are H*W elements apart, so the argmax needs no cross-lane or cross-byte work:
//     a lane walks the C planes of ITS pixels;
//   * consecutive pixels of a plane are contiguous, so a warp's load of 32 consecutive pixels is one
//     fully coalesced request whatever the plane's byte phase (plane starts are not 16-byte aligned:
//     513*513 is odd).  No shared-memory staging: a staging ring fed by one bulk copy per plane and
//     tile (C copies of ~1 KB at arbitrary 16-byte phase, aligned supersets, a plain-load tail for the
//     last bytes of the tensor) was built first and measured slower at every ring shape, bound by
//     issuing C small copies per tile (profiles/experiments/r01_nchw_ring_vs_direct.log);
//   * a tile is NWARPS chunks of 128 (fp32) or 256 (fp16) consecutive pixels of ONE frame, one chunk per
//     warp (the last tile of a frame is partial); a thread issues all its loads of a tile before the
//     first comparison.
// fp32: lane l of warp w owns pixels 128*w + 32*j + l (j = 0..3), one 4-byte load per score.
// fp16: one 2-byte load per score is bound by load instructions in flight, not bytes (3.9 TB/s), so a
//   lane owns the pixel PAIRS 256*w + 64*jj + 2*l + {0,1} (jj = 0..3) and loads a pair as one aligned
//   32-bit word, compared as half2.  A plane whose first element sits at an odd element index has its
//   aligned words one pixel to the left of the pairs; such planes keep their own running maximum over
//   the shifted pairs (2*l-1, 2*l) and the two are merged once per tile with three shuffles (plus a
//   lane-31 scalar for the pixel the shifted words do not reach).  Partial tiles use 2-byte loads.
// Statistics: the lane-private packed histogram of argmax_stream.cuh (16-byte entries: packed counters +
// 64-bit Σq) with this file's pixel <-> lane maps, and the same zero-at-rest scratch protocol
// (per-frame accumulators + pixel counters, no memset launch, completion by pixel count).
#pragma once
#include <cuda_fp16.h>

#include <type_traits>

#include "argmax_stream.cuh"

#ifndef SSB_PLANAR_PREFETCH
#define SSB_PLANAR_PREFETCH 0   // rejected: 124 -> 171 us for 64 fp16 frames (profiles/experiments/r02_nchw_l2_prefetch.log)
#endif

namespace ssb {
namespace {

struct PlanarParams {
    const void *logits;          // [B][C][H*W] elements
    uint8_t *labels;             // may be null
    unsigned long long *stats;   // [B][C][4]
    int B, hw, w, roi_h, roi_w;
    int tiles_per_frame;         // ceil(hw / TILE_PX)
    long long n_tiles;           // B * tiles_per_frame
    int flush_tiles;
    int pairs_ok;                // fp16: tensor base is 4-byte aligned, pair loads allowed
    unsigned long long *acc;     // scratch [B][C][4], zero at rest
    unsigned int *frame_px;      // scratch [B], zero at rest
    FastDiv divw;
};

// Packed lane-private word of this file:  n[0,8)  Σb[8,16)  Σk[24,38)  Σa[38,48)  Σdy[48,64)
// where a pixel's offset inside its warp's chunk is  kA*a + b + kLane*lane.
constexpr int kFieldB = 8;

struct PlaneF32 {
    using Elem = ElemF32;
    static constexpr int kBytes = 4;
    static constexpr int kPx = 4;                                  // pixels per lane and tile
    static constexpr int kDefaultWarps = 8;                        // best of the compiled set on B200 (tools/bench_nchw.py)
    static constexpr int kA = 32, kLane = 1;                       // offset = 32*j + lane
    static __device__ __forceinline__ int field_a(int j) { return j; }
    static __device__ __forceinline__ int field_b(int) { return 0; }
    static __device__ __forceinline__ uint32_t ldg(const float *q) { return __float_as_uint(__ldg(q)); }
    static __device__ __forceinline__ bool gt(uint32_t a, uint32_t b) { return __uint_as_float(a) > __uint_as_float(b); }
    static __device__ __forceinline__ long long fixed(uint32_t a) { return __float2ll_rn(__uint_as_float(a) * 16777216.0f); }
};

struct PlaneF16 {
    using Elem = ElemF16;
    static constexpr int kBytes = 2;
    static constexpr int kPx = 8;                                  // 4 pair words per plane in flight: as many bytes per thread as fp32
    static constexpr int kDefaultWarps = 4;
    static constexpr int kA = 64, kLane = 2;                       // offset = 64*(j>>1) + 2*lane + (j&1)
    static __device__ __forceinline__ int field_a(int j) { return j >> 1; }
    static __device__ __forceinline__ int field_b(int j) { return j & 1; }
    static __device__ __forceinline__ uint32_t ldg(const uint16_t *q) { return __ldg(q); }
    static __device__ __forceinline__ __half as_half(uint32_t a) { return __ushort_as_half(static_cast<unsigned short>(a)); }
    static __device__ __forceinline__ bool gt(uint32_t a, uint32_t b) { return __hgt(as_half(a), as_half(b)); }
    static __device__ __forceinline__ long long fixed(uint32_t a) { return __float2ll_rn(__half2float(as_half(a)) * 16777216.0f); }
};

template <typename P>
__device__ __forceinline__ int pixel_offset(int j, int lane) {
    return P::kA * P::field_a(j) + P::field_b(j) + P::kLane * lane;
}

// Two pixels' read-modify-writes of [packed : u64][Σq : i64] entries; loads first, the second pixel
// forwards the first one's addends when both hit the same entry (see CodecF16::add2).
template <typename P>
__device__ __forceinline__ void planar_add2(uint32_t hist_lane, uint32_t lab0, uint32_t lab1, uint32_t val0, uint32_t val1,
                                            uint32_t lo0, uint32_t lo1, uint32_t hi0, uint32_t hi1, bool c0, bool c1) {
    const uint32_t addr0 = hist_lane + lab0 * (32u * 16u), addr1 = hist_lane + lab1 * (32u * 16u);
    const unsigned long long a0 = c0 ? ((static_cast<unsigned long long>(hi0) << 32) | lo0) : 0ull;
    const unsigned long long a1 = c1 ? ((static_cast<unsigned long long>(hi1) << 32) | lo1) : 0ull;
    const unsigned long long q0 = c0 ? static_cast<unsigned long long>(P::fixed(val0)) : 0ull;
    const unsigned long long q1 = c1 ? static_cast<unsigned long long>(P::fixed(val1)) : 0ull;
    uint32_t x0, x1, x2, x3, y0, y1, y2, y3;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(x0), "=r"(x1), "=r"(x2), "=r"(x3) : "r"(addr0));
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(y0), "=r"(y1), "=r"(y2), "=r"(y3) : "r"(addr1));
    const bool same = addr0 == addr1;
    const unsigned long long va0 = ((static_cast<unsigned long long>(x1) << 32) | x0) + a0;
    const unsigned long long vq0 = ((static_cast<unsigned long long>(x3) << 32) | x2) + q0;
    const unsigned long long va1 = ((static_cast<unsigned long long>(y1) << 32) | y0) + a1 + (same ? a0 : 0ull);
    const unsigned long long vq1 = ((static_cast<unsigned long long>(y3) << 32) | y2) + q1 + (same ? q0 : 0ull);
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr0), "r"(static_cast<uint32_t>(va0)),
                 "r"(static_cast<uint32_t>(va0 >> 32)), "r"(static_cast<uint32_t>(vq0)), "r"(static_cast<uint32_t>(vq0 >> 32)));
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr1), "r"(static_cast<uint32_t>(va1)),
                 "r"(static_cast<uint32_t>(va1 >> 32)), "r"(static_cast<uint32_t>(vq1)), "r"(static_cast<uint32_t>(vq1 >> 32)));
}

struct PlanarTotals {
    long long n, sq, sk, sa, sb, sdy, cw;
};

// Lane `lane` owns label l: sum its row over the 32 lane columns (skewed start: conflict-free).
__device__ __forceinline__ PlanarTotals planar_row_totals(const uint8_t *hist, int l, int lane) {
    const ulonglong2 *row = reinterpret_cast<const ulonglong2 *>(hist) + l * 32;
    unsigned long long nk = 0, sq = 0;
    uint32_t sa = 0, sb = 0, sdy = 0, cw = 0;
#pragma unroll 8
    for (int i = 0; i < 32; ++i) {
        const int c = (lane + i) & 31;
        const ulonglong2 e = row[c];
        nk += e.x & 0x0000003FFF0000FFull;        // n at 0 (headroom to 24), Σk at 24 (headroom to 64)
        sb += (static_cast<uint32_t>(e.x) >> kFieldB) & 0xFFu;
        sa += static_cast<uint32_t>(e.x >> kFieldJ) & 0x3FFu;
        sdy += static_cast<uint32_t>(e.x >> kFieldDy);
        cw += (static_cast<uint32_t>(e.x) & 0xFFu) * c;
        sq += e.y;
    }
    PlanarTotals t;
    t.n = static_cast<long long>(nk & 0xFFFFFFull);
    t.sk = static_cast<long long>(nk >> 24);
    t.sq = static_cast<long long>(sq);
    t.sa = sa;
    t.sb = sb;
    t.sdy = sdy;
    t.cw = cw;
    return t;
}

// ---- per-tile argmax ------------------------------------------------------------------------------
// pc + plane_bytes, opaque to the optimiser: left to itself it keeps a base and a running offset and
// spends four integer instructions per plane on addresses where two suffice.
__device__ __forceinline__ const char *next_plane(const char *pc, long long plane_bytes) {
    unsigned long long a = reinterpret_cast<unsigned long long>(pc);
    asm("add.u64 %0, %0, %1;" : "+l"(a) : "l"(plane_bytes));
    return reinterpret_cast<const char *>(a);
}

// One load per score.  `lgw` points at plane 0 of the warp's chunk; pixels at or beyond
// `n_chunk` (pixels of the chunk inside the frame) are not loaded.
template <typename P, int C, bool FULL>
__device__ __forceinline__ void classify_scalar(const typename P::Elem::T *lgw, long long hw, int lane, int n_chunk,
                                                uint32_t (&best)[P::kPx], uint32_t (&lab)[P::kPx]) {
    // one running BYTE pointer per thread (a 64-bit add per plane); everything else is an immediate
    const char *pc = reinterpret_cast<const char *>(lgw + P::kLane * lane);
    const long long plane_bytes = hw * P::kBytes;
#pragma unroll
    for (int c = 0; c < C; ++c, pc = next_plane(pc, plane_bytes)) {
#pragma unroll
        for (int j = 0; j < P::kPx; ++j) {
            const int off = pixel_offset<P>(j, lane);
            uint32_t v = 0u;
            if (FULL || off < n_chunk)
                v = P::ldg(reinterpret_cast<const typename P::Elem::T *>(pc) + (P::kA * P::field_a(j) + P::field_b(j)));
            if (c == 0) {
                best[j] = v;
                lab[j] = 0u;
            } else {
                const bool g = P::gt(v, best[j]);   // strict: the smallest channel wins ties
                best[j] = g ? v : best[j];
                lab[j] = g ? static_cast<uint32_t>(c) : lab[j];
            }
        }
    }
}

__device__ __forceinline__ uint32_t hgt2_mask_bits(uint32_t a, uint32_t b) {
    return __hgt2_mask(*reinterpret_cast<const __half2 *>(&a), *reinterpret_cast<const __half2 *>(&b));
}

// fp16, full chunk, 4-byte aligned tensor: pairs of pixels as half2 words.  MODE 0: every plane of this
// frame starts at an even element index (H*W even); MODE 1 / 2: H*W is odd and the planes with even /
// odd channel index do.  The other planes ("shifted") are read through the aligned words that hold
// pixels (2l-1, 2l) and merged at the end.
template <int C, int MODE>
__device__ __forceinline__ void classify_pairs(const uint16_t *lgw, long long hw, int lane, uint32_t (&best)[PlaneF16::kPx],
                                               uint32_t (&lab)[PlaneF16::kPx]) {
    constexpr int NJ = PlaneF16::kPx / 2, kChunk = 32 * PlaneF16::kPx;   // pair words per plane, pixels per warp
    uint32_t bA[NJ], lA[NJ], bB[NJ], lB[NJ], bX = 0u, lX = 0u;
    bool haveA = false, haveB = false;
    const char *pc = reinterpret_cast<const char *>(lgw + 2 * lane);   // running byte pointer, see classify_scalar
    const long long plane_bytes = hw * 2;
#pragma unroll
    for (int c = 0; c < C; ++c, pc = next_plane(pc, plane_bytes)) {
        const uint16_t *q = reinterpret_cast<const uint16_t *>(pc);
        const bool aligned = MODE == 0 || ((c & 1) == (MODE == 1 ? 0 : 1));
        const uint32_t cc = static_cast<uint32_t>(c) * 0x00010001u;
        if (aligned) {
#pragma unroll
            for (int jj = 0; jj < NJ; ++jj) {
                const uint32_t v = __ldg(reinterpret_cast<const unsigned int *>(q + 64 * jj));
                if (!haveA) {
                    bA[jj] = v;
                    lA[jj] = cc;
                } else {
                    const uint32_t m = hgt2_mask_bits(v, bA[jj]);
                    bA[jj] = (v & m) | (bA[jj] & ~m);
                    lA[jj] = (cc & m) | (lA[jj] & ~m);
                }
            }
            haveA = true;
        } else {
#pragma unroll
            for (int jj = 0; jj < NJ; ++jj) {
                const uint32_t v = __ldg(reinterpret_cast<const unsigned int *>(q + 64 * jj - 1));
                if (!haveB) {
                    bB[jj] = v;
                    lB[jj] = cc;
                } else {
                    const uint32_t m = hgt2_mask_bits(v, bB[jj]);
                    bB[jj] = (v & m) | (bB[jj] & ~m);
                    lB[jj] = (cc & m) | (lB[jj] & ~m);
                }
            }
            // the last pixel of the chunk is in none of the shifted words: lane 31 tracks it on its own
            uint32_t x = 0u;
            if (lane == 31) x = __ldg(q + (kChunk - 1 - 62));   // q already holds 2*31
            if (!haveB) {
                bX = x;
                lX = static_cast<uint32_t>(c);
            } else {
                const bool g = PlaneF16::gt(x, bX);
                bX = g ? x : bX;
                lX = g ? static_cast<uint32_t>(c) : lX;
            }
            haveB = true;
        }
    }
    if constexpr (MODE == 0) {
#pragma unroll
        for (int j = 0; j < 2 * NJ; ++j) {
            best[j] = (bA[j >> 1] >> (16 * (j & 1))) & 0xFFFFu;
            lab[j] = (lA[j >> 1] >> (16 * (j & 1))) & 0xFFFFu;
        }
    } else {
    // shifted word of lane l holds pixels (2l-1, 2l): pixel 2l is its high half, pixel 2l+1 the low
    // half of lane l+1's word (lane 31: lane 0's word of the next 64-pixel half, or the scalar)
    uint32_t nB[NJ], nL[NJ];
#pragma unroll
    for (int jj = 0; jj < NJ; ++jj) {
        nB[jj] = __shfl_down_sync(0xffffffffu, bB[jj], 1);
        nL[jj] = __shfl_down_sync(0xffffffffu, lB[jj], 1);
        // lane 31's right neighbour: lane 0's word of the next 64-pixel run, or the scalar at the chunk's end
        const uint32_t wb = jj + 1 < NJ ? __shfl_sync(0xffffffffu, bB[jj + 1 < NJ ? jj + 1 : jj], 0) : bX;
        const uint32_t wl = jj + 1 < NJ ? __shfl_sync(0xffffffffu, lB[jj + 1 < NJ ? jj + 1 : jj], 0) : lX;
        if (lane == 31) {
            nB[jj] = wb;
            nL[jj] = wl;
        }
    }
#pragma unroll
    for (int j = 0; j < 2 * NJ; ++j) {
        const int jj = j >> 1, h = j & 1;
        const uint32_t vA = (bA[jj] >> (16 * h)) & 0xFFFFu, cA = (lA[jj] >> (16 * h)) & 0xFFFFu;
        const uint32_t vB = h ? (nB[jj] & 0xFFFFu) : (bB[jj] >> 16), cB = h ? (nL[jj] & 0xFFFFu) : (lB[jj] >> 16);
        const __half ha = PlaneF16::as_half(vA), hb = PlaneF16::as_half(vB);
        const bool takeB = __hgt(hb, ha) || (__heq(hb, ha) && cB < cA);   // equal scores: the smaller channel
        best[j] = takeB ? vB : vA;
        lab[j] = takeB ? cB : cA;
    }
    }
}

// ---- kernel -----------------------------------------------------------------------------------------
template <typename P, int C, int NWARPS, int MINB>
__global__ void __launch_bounds__(NWARPS * 32, MINB) argmax_stats_planar_kernel(const PlanarParams p) {
    using T = typename P::Elem::T;
    constexpr int kThreads = NWARPS * 32, kPx = P::kPx, kChunk = 32 * kPx, kTilePx = NWARPS * kChunk;
    constexpr int kHistWarpBytes = C * 32 * 16, kHistBytes = NWARPS * kHistWarpBytes;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *hist_all = smem;
    unsigned long long *blk = reinterpret_cast<unsigned long long *>(hist_all + kHistBytes);   // [NWARPS][C][4]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    // each CTA owns a contiguous span of tiles: it meets one or two frames, one statistics flush each
    const long long t0 = p.n_tiles * blockIdx.x / gridDim.x;
    const long long t1 = p.n_tiles * (blockIdx.x + 1) / gridDim.x;

    pdl_trigger();
    {
        unsigned long long *z = reinterpret_cast<unsigned long long *>(hist_all);
        for (int i = tid; i < (kHistBytes + NWARPS * C * 4 * 8) / 8; i += kThreads) z[i] = 0ull;
    }
    __syncthreads();
    pdl_wait();

    long long f = t0 / p.tiles_per_frame;                       // frame and tile-of-frame of the current tile
    int i = static_cast<int>(t0 - f * p.tiles_per_frame);

    uint8_t *hist = hist_all + warp * kHistWarpBytes;
    unsigned long long *blk_w = blk + warp * C * 4;
    const uint32_t hist_lane = smem_addr(hist) + lane * 16;
    // the lane-private histograms hold pixels of frame f_acc at r_base + k*TILE + CHUNK*warp + offset(a, b, lane),
    // rows >= y_base
    long long f_acc = f;
    int r_base = i * kTilePx;
    uint32_t y_base = fastdiv(static_cast<uint32_t>(r_base), p.divw);
    int k = 0, covered = 0;
    __shared__ int s_last;

    auto warp_flush = [&]() {
        __syncwarp();
        for (int l = lane; l < C; l += 32) {
            const PlanarTotals t = planar_row_totals(hist, l, lane);
            if (t.n) {
                const long long sr = t.n * r_base + t.sk * kTilePx + static_cast<long long>(kChunk) * warp * t.n + P::kA * t.sa + t.sb + P::kLane * t.cw;
                const long long sy = t.n * y_base + t.sdy;
                unsigned long long *row = blk_w + l * 4;
                row[SSB_STAT_N] += static_cast<unsigned long long>(t.n);
                row[SSB_STAT_SX] += static_cast<unsigned long long>(sr - sy * p.w);
                row[SSB_STAT_SY] += static_cast<unsigned long long>(sy);
                row[SSB_STAT_SQ] += static_cast<unsigned long long>(t.sq);
            }
        }
        __syncwarp();
        ulonglong2 *h = reinterpret_cast<ulonglong2 *>(hist);
#pragma unroll
        for (int l = 0; l < C; ++l) h[l * 32 + lane] = make_ulonglong2(0ull, 0ull);
        __syncwarp();
    };

    // same protocol as argmax_stats_stream_kernel: add the CTA's share of frame f_acc to scratch, report
    // the pixels covered; the CTA that completes the frame publishes it and re-zeroes scratch
    auto frame_flush = [&]() {
        if (covered == 0) return;   // CTA-uniform
        warp_flush();
        __syncthreads();
        for (int e = tid; e < C * 4; e += kThreads) {
            unsigned long long v = 0ull;
#pragma unroll
            for (int wi = 0; wi < NWARPS; ++wi) {
                v += blk[wi * C * 4 + e];
                blk[wi * C * 4 + e] = 0ull;
            }
            if (v) atomicAdd(&p.acc[f_acc * (C * 4) + e], v);
        }
        __threadfence();
        __syncthreads();
        if (tid == 0) {
            const unsigned int before = atomicAdd(&p.frame_px[f_acc], static_cast<unsigned int>(covered));
            s_last = before + static_cast<unsigned int>(covered) == static_cast<unsigned int>(p.hw);
            __threadfence();
        }
        __syncthreads();
        if (s_last) {
            for (int e = tid; e < C * 4; e += kThreads)
                p.stats[f_acc * (C * 4) + e] = atomicExch(&p.acc[f_acc * (C * 4) + e], 0ull);
            if (tid == 0) p.frame_px[f_acc] = 0u;
        }
        covered = 0;
    };

    uint64_t pol_keep;   // label map: keep in L2 for whoever consumes it next
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_keep));

    for (long long t = t0; t < t1; ++t) {
        const int r0 = i * kTilePx;
        if (f != f_acc || k >= p.flush_tiles) {
            // a new frame (hand the finished one over) or a packed field about to overflow: fold the
            // lane-private histograms and restart them at this tile
            if (f != f_acc) frame_flush();
            else warp_flush();
            f_acc = f;
            r_base = r0;
            y_base = fastdiv(static_cast<uint32_t>(r0), p.divw);
            k = 0;
        }
#if SSB_PLANAR_PREFETCH == 2
        // per-thread L2 prefetches of the words this thread will load in the next tile (same frame, full tile)
        if (t + 1 < t1 && i + 1 < p.tiles_per_frame && (i + 2) * kTilePx <= p.hw) {
            const char *q = reinterpret_cast<const char *>(static_cast<const T *>(p.logits) + (f * C * p.hw + r0 + kTilePx + kChunk * warp)) +
                            lane * (P::kBytes == 2 ? 4 : 4);
            const long long plane_bytes = static_cast<long long>(p.hw) * P::kBytes;
#pragma unroll
            for (int c = 0; c < C; ++c, q += plane_bytes) {
#pragma unroll
                for (int jj = 0; jj < 4; ++jj)
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(q + 128 * jj));
            }
        }
#elif SSB_PLANAR_PREFETCH
        // The C plane chunks this warp reads in the NEXT tile are requested from DRAM now (one bulk prefetch into
        // L2 per plane, issued by one lane): a thread holds all its loads of a tile in registers, so it cannot
        // have two tiles in flight itself, and with four warps per scheduler DRAM latency was not covered.
        if (lane == 0 && t + 1 < t1) {
            long long fn = f;
            int in = i + 1;
            if (in == p.tiles_per_frame) { in = 0; ++fn; }
            const int rn = in * kTilePx + kChunk * warp;
            const int n_px = min(kChunk, p.hw - rn);
            if (n_px > 0) {
                const char *q = reinterpret_cast<const char *>(static_cast<const T *>(p.logits) + (fn * C * p.hw + rn));
                const long long plane_bytes = static_cast<long long>(p.hw) * P::kBytes;
#pragma unroll
                for (int c = 0; c < C; ++c, q += plane_bytes) {
                    const uintptr_t a = reinterpret_cast<uintptr_t>(q) & ~static_cast<uintptr_t>(15);
                    const uintptr_t e = (reinterpret_cast<uintptr_t>(q) + static_cast<uintptr_t>(n_px) * P::kBytes) & ~static_cast<uintptr_t>(15);
                    if (e > a) bulk_prefetch_l2(reinterpret_cast<const void *>(a), static_cast<uint32_t>(e - a));
                }
            }
        }
#endif
        const int n_valid = min(kTilePx, p.hw - r0);       // pixels of the tile inside the frame
        const int n_chunk = n_valid - kChunk * warp;        // ... of this warp's chunk (may be <= 0)
        const T *lgw = static_cast<const T *>(p.logits) + (f * C * p.hw + r0 + kChunk * warp);
        uint32_t best[kPx], lab[kPx];
        if (n_valid == kTilePx) {
            if constexpr (P::kBytes == 2) {
                if (p.pairs_ok) {
                    const int mode = (p.hw & 1) ? 1 + static_cast<int>((f * C) & 1) : 0;   // CTA-uniform
                    if (mode == 0) classify_pairs<C, 0>(lgw, p.hw, lane, best, lab);
                    else if (mode == 1) classify_pairs<C, 1>(lgw, p.hw, lane, best, lab);
                    else classify_pairs<C, 2>(lgw, p.hw, lane, best, lab);
                } else {
                    classify_scalar<P, C, true>(lgw, p.hw, lane, n_chunk, best, lab);
                }
            } else {
                classify_scalar<P, C, true>(lgw, p.hw, lane, n_chunk, best, lab);
            }
        } else {
            classify_scalar<P, C, false>(lgw, p.hw, lane, n_chunk, best, lab);
        }

        if (p.labels) {
            uint8_t *out = p.labels + f * p.hw + r0 + kChunk * warp;
#pragma unroll
            for (int j = 0; j < kPx; ++j) {
                const int off = pixel_offset<P>(j, lane);
                if (off < n_chunk)
                    asm volatile("st.global.L2::cache_hint.u8 [%0], %1, %2;" ::"l"(out + off), "r"(lab[j]), "l"(pol_keep) : "memory");
            }
        }

        // statistics.  Tile classes (CTA-uniform): interior = full tile, every pixel inside the ROI (no
        // per-pixel tests); tiles whose rows all lie below the ROI contribute labels only; the rest test
        // every pixel.
        if (n_valid == kTilePx && p.roi_w == p.w &&
            fastdiv(static_cast<uint32_t>(r0 + kTilePx - 1), p.divw) < static_cast<uint32_t>(p.roi_h)) {
            uint32_t add_lo[kPx], add_hi[kPx];
#pragma unroll
            for (int j = 0; j < kPx; ++j) {
                const uint32_t y = fastdiv(static_cast<uint32_t>(r0 + kChunk * warp + pixel_offset<P>(j, lane)), p.divw);
                add_lo[j] = 1u | (static_cast<uint32_t>(P::field_b(j)) << kFieldB) | (static_cast<uint32_t>(k) << kFieldK);
                add_hi[j] = (static_cast<uint32_t>(P::field_a(j)) << (kFieldJ - 32)) | ((y - y_base) << (kFieldDy - 32));
            }
#pragma unroll
            for (int j = 0; j < kPx; j += 2)
                planar_add2<P>(hist_lane, lab[j], lab[j + 1], best[j], best[j + 1], add_lo[j], add_lo[j + 1], add_hi[j], add_hi[j + 1], true, true);
        } else if (fastdiv(static_cast<uint32_t>(r0), p.divw) < static_cast<uint32_t>(p.roi_h)) {
            uint32_t add_lo[kPx], add_hi[kPx];
            bool counted[kPx];
#pragma unroll
            for (int j = 0; j < kPx; ++j) {
                const int off = pixel_offset<P>(j, lane);
                const uint32_t r = static_cast<uint32_t>(r0 + kChunk * warp + off);
                const uint32_t y = fastdiv(r, p.divw), x = r - y * p.w;
                counted[j] = (off < n_chunk) & (x < static_cast<uint32_t>(p.roi_w)) & (y < static_cast<uint32_t>(p.roi_h));
                add_lo[j] = 1u | (static_cast<uint32_t>(P::field_b(j)) << kFieldB) | (static_cast<uint32_t>(k) << kFieldK);
                add_hi[j] = (static_cast<uint32_t>(P::field_a(j)) << (kFieldJ - 32)) | ((y - y_base) << (kFieldDy - 32));
            }
#pragma unroll
            for (int j = 0; j < kPx; j += 2)
                planar_add2<P>(hist_lane, lab[j], lab[j + 1], best[j], best[j + 1], add_lo[j], add_lo[j + 1], add_hi[j], add_hi[j + 1], counted[j],
                               counted[j + 1]);
        }
        covered += n_valid;
        ++k;
        if (++i == p.tiles_per_frame) { i = 0; ++f; }
    }
    frame_flush();
}

// ---- host side ------------------------------------------------------------------------------------

template <typename P, int C, int NWARPS, int MINB>
int launch_planar(PlanarParams p, cudaStream_t st) {
    auto kern = argmax_stats_planar_kernel<P, C, NWARPS, MINB>;
    constexpr int kTilePx = NWARPS * 32 * P::kPx;
    constexpr int kSmem = NWARPS * C * 32 * 16 + NWARPS * C * 4 * 8;
    DeviceInfo di;
    if (int rc = device_info(&di)) return rc;
    if (kSmem > di.max_smem_optin) return fail(SSB_ERR_LIMIT, "config needs %d B shared memory", kSmem);
    static int ctas_per_sm[64] = {0};
    int dev = 0;
    SSB_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(SSB_ERR_LIMIT, "device ordinal %d not supported", dev);
    if (ctas_per_sm[dev] == 0) {
        SSB_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
        int occ = 0;
        SSB_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, NWARPS * 32, kSmem));
        if (occ < 1) return fail(SSB_ERR_LIMIT, "kernel does not fit on an SM");
        ctas_per_sm[dev] = occ;
    }
    p.tiles_per_frame = (p.hw + kTilePx - 1) / kTilePx;
    p.n_tiles = static_cast<long long>(p.B) * p.tiles_per_frame;
    p.flush_tiles = min(pick_flush_tiles(kTilePx, p.w), 252 / P::kPx);   // the 8-bit n field: <= 252 pixels per lane and period
    if (p.flush_tiles < 1) return fail(SSB_ERR_LIMIT, "W=%d too narrow for the packed accumulators", p.w);
    long long grid = static_cast<long long>(di.sm_count) * ctas_per_sm[dev];
    if (const char *env = getenv("SSB_ARGMAX_WAVES")) grid *= atoi(env) > 0 ? atoi(env) : 1;  // tuning knob
    if (grid > p.n_tiles) grid = p.n_tiles;
    SSB_CUDA_TRY(launch_pdl(kern, dim3(static_cast<unsigned>(grid)), dim3(NWARPS * 32), kSmem, st, p));
    return SSB_OK;
}

template <typename P, int C>
int dispatch_planar(const PlanarParams &p, cudaStream_t st, int nw) {
    if (nw == 8) return launch_planar<P, C, 8, 2>(p, st);
    if (nw == 4) return launch_planar<P, C, 4, 4>(p, st);
    return fail(SSB_ERR_LIMIT, "no planar kernel compiled for %d warps per CTA", nw);
}

}  // namespace

// Planar kernels for the class counts other than 19 and 21 (argmax_planar_anyc_*.cu): 2 <= C <= kMaxPlanarC,
// the default CTA shape of each element type.  (PlanarParams lives in an unnamed namespace: the functions are
// declared per translation unit and defined with external linkage on plain argument lists.)
constexpr int kMaxPlanarC = 32;
struct PlanarArgs {              // PlanarParams without the unnamed-namespace type in a cross-TU signature
    const void *logits; uint8_t *labels; unsigned long long *stats;
    int B, hw, w, roi_h, roi_w, pairs_ok;
    unsigned long long *acc; unsigned int *frame_px;
    FastDiv divw;
};
int launch_planar_f32_other_c(int C, const PlanarArgs &a, cudaStream_t st);
int launch_planar_f16_other_c(int C, const PlanarArgs &a, cudaStream_t st);

namespace {
inline PlanarParams to_params(const PlanarArgs &a) {
    PlanarParams p{};
    p.logits = a.logits; p.labels = a.labels; p.stats = a.stats;
    p.B = a.B; p.hw = a.hw; p.w = a.w; p.roi_h = a.roi_h; p.roi_w = a.roi_w; p.pairs_ok = a.pairs_ok;
    p.acc = a.acc; p.frame_px = a.frame_px; p.divw = a.divw;
    return p;
}
}  // namespace

}  // namespace ssb
