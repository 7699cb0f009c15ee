// fp16 codec of the streaming argmax + statistics kernel (argmax_stream.cuh), shared by the translation units
// that instantiate it for different class counts.
//
// A lane owns 4 consecutive pixels = C aligned 64-bit words of the tile (stride C x 8 bytes between
// lanes: conflict-free LDS.64 for odd C).  Per pixel: the maximum with HMNMX2 over half2 words (halves
// that belong to a neighbouring pixel are replaced by -inf), then the FIRST channel equal to it from
// per-word HSET2 equality masks: floating-point `==`, so -0 and +0 tie and the first index wins, as
// np.argmax does.  Inputs must be NaN-free.  Sum q_max is the exact fixed-point sum of value * 2^24, kept
// as a second 64-bit word next to the packed n / k / j / dy word (16-byte entries, LDS.128/STS.128).
#pragma once
#include <cuda_fp16.h>

#include "argmax_stream.cuh"

namespace ssb {
namespace {

__device__ __forceinline__ uint32_t hmax2_bits(uint32_t a, uint32_t b) {
    const __half2 r = __hmax2(*reinterpret_cast<const __half2 *>(&a), *reinterpret_cast<const __half2 *>(&b));
    return *reinterpret_cast<const uint32_t *>(&r);
}
__device__ __forceinline__ uint32_t heq2_mask_bits(uint32_t a, uint32_t b) {
    return __heq2_mask(*reinterpret_cast<const __half2 *>(&a), *reinterpret_cast<const __half2 *>(&b));
}

template <int C>
struct CodecF16 {
    static constexpr int kC = C, kElemBytes = 2, kQuadWords = 2 * C, kEntryBytes = 16;
    using Elem = ElemF16;
    struct Ctx {
        uint32_t ninf2;   // half2(-inf, -inf), kept in a register (see the u8 codec on why via smem)
    };
    static __device__ __forceinline__ void store_table(uint32_t *tab) { tab[0] = 0xFC00FC00u; }
    static __device__ __forceinline__ Ctx load_ctx(const uint32_t *tab) {
        return Ctx{*reinterpret_cast<const volatile uint32_t *>(tab)};
    }

    static __device__ __forceinline__ void load_quad(const uint8_t *stage, int quad, uint32_t (&w)[2 * C]) {
        const uint2 *src = reinterpret_cast<const uint2 *>(stage) + quad * C;
#pragma unroll
        for (int c = 0; c < C; ++c) {
            const uint2 v = src[c];
            w[2 * c] = v.x;
            w[2 * c + 1] = v.y;
        }
    }

    // key = (bits of the maximum << 16) | (255 - label)
    static __device__ __forceinline__ uint32_t pixel_key(const uint32_t (&w)[2 * C], int j, const Ctx &ctx) {
        const int h0 = j * C, h1 = (j + 1) * C;          // halves [h0, h1) of the quad belong to pixel j
        const int wd0 = h0 / 2, wd1 = (h1 - 1) / 2;      // words wd0..wd1
        uint32_t v[C / 2 + 2];
        int nv = 0;
#pragma unroll
        for (int wd = wd0; wd <= wd1; ++wd) {
            const bool lo = 2 * wd >= h0, hi = 2 * wd + 1 < h1;
            // foreign halves -> -inf: PRMT picks bytes 0-3 from the word, 4-7 from half2(-inf,-inf)
            v[nv++] = lo && hi ? w[wd] : __byte_perm(w[wd], ctx.ninf2, lo ? 0x5410u : 0x3254u);
        }
        uint32_t m2 = v[0];
#pragma unroll
        for (int i = 1; i < nv; ++i) m2 = hmax2_bits(m2, v[i]);
        m2 = hmax2_bits(m2, __byte_perm(m2, 0u, 0x1032u));   // both halves = the pixel's maximum
        // equality mask: bit i (low halves) / bit 16+i (high halves) for word wd0+i
        uint32_t acc = 0;
#pragma unroll
        for (int i = 0; i < nv; ++i) {
            const int wd = wd0 + i;
            const bool lo = 2 * wd >= h0, hi = 2 * wd + 1 < h1;
            const uint32_t bits = (lo ? 1u << i : 0u) | (hi ? 1u << (16 + i) : 0u);
            acc |= heq2_mask_bits(v[i], m2) & bits;
        }
        // low half of word wd0+i is channel 2*(wd0+i) - h0, high half is that + 1
        const int off = 2 * wd0 - h0;                     // 0 or -1
        const uint32_t lo16 = acc & 0xffffu, hi16 = acc >> 16;
        const uint32_t cl = lo16 ? 2u * (__ffs(lo16) - 1) + off : 255u;
        const uint32_t ch = hi16 ? 2u * (__ffs(hi16) - 1) + off + 1 : 255u;
        const uint32_t label = min(cl, ch);
        return (m2 << 16) | (255u - label);
    }

    static __device__ __forceinline__ void classify(const uint32_t (&w)[2 * C], const Ctx &ctx, uint32_t (&key)[4]) {
#pragma unroll
        for (int j = 0; j < 4; ++j) key[j] = pixel_key(w, j, ctx);
    }

    // value * 2^24 of a finite fp16 as an exact integer: fp16 -> fp32 and the scaling by a power of two
    // are exact, so one F2I does it (the bit-twiddling form costs ~12 instructions per pixel).
    static __device__ __forceinline__ long long fixed24(uint32_t hbits) {
        const float f = __half2float(__ushort_as_half(static_cast<unsigned short>(hbits)));
        return __float2ll_rn(f * 16777216.0f);
    }

    // 128-bit read-modify-writes of [packed n|Σk|Σj|Σdy : u64][Σ value*2^24 : i64] entries, two pixels
    // at a time: both entries are loaded first and the second pixel adds the first one's addends itself
    // when they share a label (its load is stale; its store, issued later to the same address, wins).
    // Pairs instead of the u8 codec's quads: with two 64-bit words per entry, forwarding among four
    // pixels costs more instructions than the second round of shared-memory latency.
    static __device__ __forceinline__ void add2(uint32_t hist_lane, uint32_t key0, uint32_t key1, uint32_t lo_base,
                                                uint32_t hi0, uint32_t hi1, bool c0, bool c1) {
        const uint32_t addr0 = hist_lane - (key0 & 0xffu) * (32u * kEntryBytes);
        const uint32_t addr1 = hist_lane - (key1 & 0xffu) * (32u * kEntryBytes);
        const unsigned long long a0 = c0 ? ((static_cast<unsigned long long>(hi0) << 32) | lo_base) : 0ull;
        const unsigned long long a1 = c1 ? ((static_cast<unsigned long long>(hi1) << 32) | lo_base) : 0ull;
        const unsigned long long q0 = c0 ? static_cast<unsigned long long>(fixed24(key0 >> 16)) : 0ull;
        const unsigned long long q1 = c1 ? static_cast<unsigned long long>(fixed24(key1 >> 16)) : 0ull;
        uint32_t x0, x1, x2, x3, y0, y1, y2, y3;
        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(x0), "=r"(x1), "=r"(x2), "=r"(x3) : "r"(addr0));
        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(y0), "=r"(y1), "=r"(y2), "=r"(y3) : "r"(addr1));
        const unsigned long long va0 = ((static_cast<unsigned long long>(x1) << 32) | x0) + a0;
        const unsigned long long vq0 = ((static_cast<unsigned long long>(x3) << 32) | x2) + q0;
        const bool same = addr0 == addr1;
        const unsigned long long va1 = ((static_cast<unsigned long long>(y1) << 32) | y0) + a1 + (same ? a0 : 0ull);
        const unsigned long long vq1 = ((static_cast<unsigned long long>(y3) << 32) | y2) + q1 + (same ? q0 : 0ull);
        asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr0), "r"(static_cast<uint32_t>(va0)),
                     "r"(static_cast<uint32_t>(va0 >> 32)), "r"(static_cast<uint32_t>(vq0)),
                     "r"(static_cast<uint32_t>(vq0 >> 32)));
        asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr1), "r"(static_cast<uint32_t>(va1)),
                     "r"(static_cast<uint32_t>(va1 >> 32)), "r"(static_cast<uint32_t>(vq1)),
                     "r"(static_cast<uint32_t>(vq1 >> 32)));
    }

    static __device__ __forceinline__ void add4(uint32_t hist_lane, const uint32_t (&key)[4], uint32_t lo_base,
                                                const uint32_t (&add_hi)[4], const bool (&counted)[4]) {
        add2(hist_lane, key[0], key[1], lo_base, add_hi[0], add_hi[1], counted[0], counted[1]);
        add2(hist_lane, key[2], key[3], lo_base, add_hi[2], add_hi[3], counted[2], counted[3]);
    }

    static __device__ __forceinline__ RowTotals row_totals(const uint8_t *hist, int l, int lane) {
        const ulonglong2 *row = reinterpret_cast<const ulonglong2 *>(hist) + l * 32;
        unsigned long long nk = 0, sq = 0;
        uint32_t sj = 0, sdy = 0, cw = 0;
#pragma unroll 8
        for (int i = 0; i < 32; ++i) {
            const int c = (lane + i) & 31;
            const ulonglong2 e = row[c];
            nk += e.x & 0x0000003FFF0000FFull;        // n at 0 (headroom to 24), Σk at 24 (headroom to 64)
            sj += static_cast<uint32_t>(e.x >> kFieldJ) & 0x3FFu;
            sdy += static_cast<uint32_t>(e.x >> kFieldDy);
            cw += (static_cast<uint32_t>(e.x) & 0xFFu) * c;
            sq += e.y;                                // two's complement: wraps correctly
        }
        RowTotals t;
        t.n = static_cast<long long>(nk & 0xFFFFFFull);
        t.sk = static_cast<long long>(nk >> 24);
        t.sq = static_cast<long long>(sq);
        t.sj = sj;
        t.sdy = sdy;
        t.cw = cw;
        return t;
    }

    static __device__ __forceinline__ void clear_column(uint8_t *hist, int lane) {
        ulonglong2 *h = reinterpret_cast<ulonglong2 *>(hist);
#pragma unroll
        for (int l = 0; l < C; ++l) h[l * 32 + lane] = make_ulonglong2(0ull, 0ull);
    }
};

}  // namespace

// Streaming kernel for the class counts other than 19 and 21 (argmax_stats_f16_anyc_*.cu): 2 <= C <= kMaxStreamCF16.
constexpr int kMaxStreamCF16 = 32;
int launch_stream_f16_other_c(int C, const ArgmaxParams &p, cudaStream_t st);

}  // namespace ssb
