// u8 codec of the streaming argmax + statistics kernel (argmax_stream.cuh), shared by the translation units
// that instantiate it for different class counts.
//
// Argmax on quantised u8 scores: a lane owns 4 consecutive pixels = C aligned 32-bit words of the
// tile.  Scores are packed with their channel into 16-bit keys (value<<8 | 255-channel), two channels per
// PRMT, and reduced with VIMNMX3.U16x2: the maximum key is the maximum value with the smallest channel on
// ties.  The lane stride of C words is free of bank conflicts for odd C with 32-bit loads, for C = 2 mod 4
// with 64-bit loads and for C = 4 mod 8 with 128-bit loads; multiples of 8 keep 2- to 8-way conflicts
// (the tile arrives as one flat bulk copy: it cannot be padded).
#pragma once
#include "argmax_stream.cuh"

namespace ssb {
namespace {

// PRMT selectors are kept in registers (the tag bytes ride in the immediate operand).  They are
// read from shared memory once per thread: a value ptxas can see as a constant gets re-materialised
// (UMOV + IMAD.U32) in front of every one of the ~44 PRMTs of a quad.
struct Selectors {
    uint32_t both[2], only_a[2], only_b[2];  // indexed by byte-lane parity: lanes (0,2) / (1,3)
};
// result bytes: [tag_a, value_a, tag_b, value_b]; nibbles 0-3 pick the data word, 4/5 the two tag
// bytes of the immediate, 6 its zero byte (used for bytes that belong to a neighbouring pixel).
__device__ __forceinline__ void store_selector_table(uint32_t *tab) {
    tab[0] = 0x2504u; tab[1] = 0x3514u; tab[2] = 0x6604u; tab[3] = 0x6614u; tab[4] = 0x2566u; tab[5] = 0x3566u;
}
__device__ __forceinline__ Selectors load_selectors(const uint32_t *tab) {
    Selectors s;
    const volatile uint32_t *t = tab;
    s.both[0] = t[0]; s.both[1] = t[1]; s.only_a[0] = t[2]; s.only_a[1] = t[3]; s.only_b[0] = t[4]; s.only_b[1] = t[5];
    return s;
}

// ---- argmax over the 4 pixels held in C words -------------------------------------------------
template <int C>
__device__ __forceinline__ uint32_t pixel_key(const uint32_t (&w)[C], int j, const Selectors &sel) {
    // Pixel j of the quad occupies bytes [j*C, (j+1)*C) of the C words.  One PRMT builds two 16-bit
    // keys (value<<8 | tag) from byte lanes (0,2) or (1,3) of a word; bytes of a neighbouring pixel
    // become a zero key (valid keys are >= 1 because tag = 255 - channel >= 1).
    const int b0 = j * C, b1 = (j + 1) * C;
    uint32_t ks[C / 2 + 4];
    int nk = 0;
#pragma unroll
    for (int wd = b0 / 4; wd <= (b1 - 1) / 4; ++wd) {
#pragma unroll
        for (int par = 0; par < 2; ++par) {
            const int oa = 4 * wd + par, ob = 4 * wd + par + 2;
            const bool va = oa >= b0 && oa < b1, vb = ob >= b0 && ob < b1;
            if (va || vb) {
                const uint32_t ta = va ? 255u - (oa - b0) : 0u, tb = vb ? 255u - (ob - b0) : 0u;
                const uint32_t sl = va && vb ? sel.both[par] : (va ? sel.only_a[par] : sel.only_b[par]);
                ks[nk++] = __byte_perm(w[wd], ta | (tb << 8), sl);
            }
        }
    }
    uint32_t m = ks[0];
    int i = 1;
#pragma unroll
    for (; i + 1 < nk; i += 2) m = __vimax3_u16x2(m, ks[i], ks[i + 1]);
    if (i < nk) m = __vmaxu2(m, ks[i]);
    return __vmaxu2(m, m >> 16);  // low 16 bits = (q_max << 8) | (255 - label)
}

template <int C>
struct CodecU8 {
    static constexpr int kC = C, kElemBytes = 1, kQuadWords = C, kEntryBytes = 8;
    using Elem = ElemU8;
    struct Ctx {
        Selectors sel;
    };
    static __device__ __forceinline__ void store_table(uint32_t *tab) { store_selector_table(tab); }
    static __device__ __forceinline__ Ctx load_ctx(const uint32_t *tab) { return Ctx{load_selectors(tab)}; }

    static __device__ __forceinline__ void load_quad(const uint8_t *stage, int quad, uint32_t (&w)[C]) {
        if constexpr (C % 4 == 0) {
            const uint4 *src = reinterpret_cast<const uint4 *>(stage) + quad * (C / 4);
#pragma unroll
            for (int c = 0; c < C / 4; ++c) {
                const uint4 v = src[c];
                w[4 * c] = v.x; w[4 * c + 1] = v.y; w[4 * c + 2] = v.z; w[4 * c + 3] = v.w;
            }
        } else if constexpr (C % 2 == 0) {
            const uint2 *src = reinterpret_cast<const uint2 *>(stage) + quad * (C / 2);
#pragma unroll
            for (int c = 0; c < C / 2; ++c) {
                const uint2 v = src[c];
                w[2 * c] = v.x; w[2 * c + 1] = v.y;
            }
        } else {
            const uint32_t *src = reinterpret_cast<const uint32_t *>(stage) + quad * C;
#pragma unroll
            for (int c = 0; c < C; ++c) w[c] = src[c];
        }
    }

    // key = (q_max << 8) | (255 - label)
    static __device__ __forceinline__ void classify(const uint32_t (&w)[C], const Ctx &ctx, uint32_t (&key)[4]) {
#pragma unroll
        for (int j = 0; j < 4; ++j) key[j] = pixel_key<C>(w, j, ctx.sel);
    }

    // 64-bit read-modify-writes of the lane-private entries of a quad's labels, two pixels at a time.
    //   add_lo = 1 | k<<24 (tile-uniform) gets q<<8 from the key; add_hi = j<<6 | dy<<16.
    // Both entries of a pair are loaded first (a dependent load->add->store chain per pixel would
    // expose the shared-memory latency four times per tile); where the two pixels share a label the
    // second one adds the first one's addend itself (its load is stale) and its store, issued later to
    // the same address, wins.
    static __device__ __forceinline__ void add2(uint32_t hist_lane, uint32_t key0, uint32_t key1, uint32_t lo_base,
                                                uint32_t hi0, uint32_t hi1, bool c0, bool c1) {
        const uint32_t addr0 = hist_lane - __byte_perm(key0, 0u, 0x4404u);  // tag * 256
        const uint32_t addr1 = hist_lane - __byte_perm(key1, 0u, 0x4404u);
        const unsigned long long a0 =
            c0 ? ((static_cast<unsigned long long>(hi0) << 32) | ((key0 & 0xff00u) | lo_base)) : 0ull;
        const unsigned long long a1 =
            c1 ? ((static_cast<unsigned long long>(hi1) << 32) | ((key1 & 0xff00u) | lo_base)) : 0ull;
        uint32_t x0, x1, y0, y1;
        asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(x0), "=r"(x1) : "r"(addr0));
        asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(y0), "=r"(y1) : "r"(addr1));
        const unsigned long long v0 = ((static_cast<unsigned long long>(x1) << 32) | x0) + a0;
        const unsigned long long v1 =
            ((static_cast<unsigned long long>(y1) << 32) | y0) + a1 + (addr0 == addr1 ? a0 : 0ull);
        asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(addr0), "r"(static_cast<uint32_t>(v0)),
                     "r"(static_cast<uint32_t>(v0 >> 32)));
        asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(addr1), "r"(static_cast<uint32_t>(v1)),
                     "r"(static_cast<uint32_t>(v1 >> 32)));
    }

    static __device__ __forceinline__ void add4(uint32_t hist_lane, const uint32_t (&key)[4], uint32_t lo_base,
                                                const uint32_t (&add_hi)[4], const bool (&counted)[4]) {
        add2(hist_lane, key[0], key[1], lo_base, add_hi[0], add_hi[1], counted[0], counted[1]);
        add2(hist_lane, key[2], key[3], lo_base, add_hi[2], add_hi[3], counted[2], counted[3]);
    }

    // Lane l owns label l: walk the 32 lane columns of its row (skewed start: conflict-free), adding
    // the packed fields two at a time into wide accumulators.
    static __device__ __forceinline__ RowTotals row_totals(const uint8_t *hist, int l, int lane) {
        const unsigned long long *row = reinterpret_cast<const unsigned long long *>(hist) + l * 32;
        unsigned long long nk = 0, qj = 0;
        uint32_t sdy = 0, cw = 0;
#pragma unroll 8
        for (int i = 0; i < 32; ++i) {
            const int c = (lane + i) & 31;
            const unsigned long long e = row[c];
            nk += e & 0x0000003FFF0000FFull;          // n at 0 (headroom to 24), Σk at 24 (headroom to 64)
            qj += (e >> 8) & 0x000000FFC000FFFFull;   // Σq at 0 (headroom to 30), Σj at 30 (headroom to 64)
            sdy += static_cast<uint32_t>(e >> kFieldDy);
            cw += (static_cast<uint32_t>(e) & 0xFFu) * c;   // Σ n_c * c: lane offsets of the counted pixels
        }
        RowTotals t;
        t.n = static_cast<long long>(nk & 0xFFFFFFull);
        t.sk = static_cast<long long>(nk >> 24);
        t.sq = static_cast<long long>(qj & 0x3FFFFFFFull);
        t.sj = static_cast<long long>(qj >> 30);
        t.sdy = sdy;
        t.cw = cw;
        return t;
    }

    static __device__ __forceinline__ void clear_column(uint8_t *hist, int lane) {
        unsigned long long *h = reinterpret_cast<unsigned long long *>(hist);
#pragma unroll
        for (int l = 0; l < C; ++l) h[l * 32 + lane] = 0ull;
    }
};

}  // namespace

// Streaming kernel for the class counts other than 19 and 21 (argmax_stats_u8_anyc_*.cu): 2 <= C <= kMaxStreamC.
constexpr int kMaxStreamC = 32;
int launch_stream_u8_other_c(int C, const ArgmaxParams &p, cudaStream_t st);

}  // namespace ssb
