// K2+K3: fused per-pixel argmax over C quantised class scores + per-label {n, Σx, Σy, Σq_max}.
//
// Replaces the ArgMax tail of engine.run_inference (sem_seg_server.py:162), the reshape/crop of
// :163-170 and the counting/moment arithmetic of :102-122 restated per label (SURVEY.md §8 a5,a6,a9).
//
// Design (HBM-bound: every logit byte is read exactly once, 1/C of that is written back):
//   * the batch is ONE flat byte stream [B*H*W*C]; frames are not 16-byte aligned (513*513*21 is
//     odd), tiles are.  A tile is TILE_PX = 128*NWARPS pixels (a multiple of 16 => TILE_PX*C bytes
//     is a multiple of 16), moved global->shared by one TMA bulk copy (cp.async.bulk, UBLKCP) into a
//     STAGES-deep ring guarded by full/empty mbarriers; one producer warp, NWARPS consumer warps;
//   * each CTA owns a CONTIGUOUS span of tiles (grid = SMs x CTAs/SM), so it touches one or two
//     frames and flushes its statistics to global memory once per frame it touched;
//   * a lane owns 4 consecutive pixels = C aligned 32-bit words of the tile (stride C words between
//     lanes: conflict-free for odd C).  Argmax packs (value<<8 | 255-channel) into 16-bit keys, two
//     channels per PRMT, and reduces them with VIMNMX3.U16x2: the maximum key is the maximum value
//     with the smallest channel on ties (np.argmax / TFLite ArgMax semantics);
//   * statistics: every lane has a PRIVATE histogram column in shared memory, hist[warp][label][lane]
//     (bank == lane, so no conflicts and no atomics whatever the label distribution); one 64-bit
//     read-modify-write per pixel adds the packed fields
//         n:8 | Σq:16 | Σk:14 | Σj:10 | Σdy:16
//     where k = tiles since the last flush, j = pixel of the quad, dy = row - y_base.  A lane's pixel
//     positions are r = r_base + k*TILE_PX + 4*quad + j, so Σr (hence Σx = Σr - W*Σy) is recovered at
//     flush time from n, Σk, Σj and the lane index: no per-pixel x is needed.  Before a field can
//     overflow (every `flush_tiles` tiles) the warp reduces its 32 columns into the CTA's [C][4] u64
//     block accumulators; on a frame change / at the end the CTA adds those to global memory -- one
//     atomic flush per block and frame;
//   * no separate zeroing launch (a 43 KB memset in front of a 60 us kernel costs ~3 us): the global
//     flush goes to a scratch accumulator that is ZERO AT REST.  Every CTA also adds the number of
//     pixels of the frame it covered to a per-frame counter; the CTA that completes a frame (counter
//     reaches H*W) moves the accumulated block to `stats` with plain stores and re-zeroes scratch.
//     Completion by pixel count is independent of how tiles are assigned to CTAs.
#include "ssb_common.cuh"

namespace ssb {
namespace {

struct ArgmaxParams {
    const uint8_t *logits;
    uint8_t *labels;             // may be null
    unsigned long long *stats;   // [B][C][4]
    long long total_px;          // B*H*W
    long long n_tiles;           // full tiles
    int hw, w, roi_h, roi_w;
    int flush_tiles;             // lane-private histogram flush period, in tiles
    unsigned long long *acc;     // scratch [B][C][4], zero at rest
    unsigned int *frame_px;      // scratch [B]: pixels of the frame already flushed, zero at rest
    FastDiv divw;
};

// packed lane-private entry: n[0,8) Σq[8,24) Σk[24,38) Σj[38,48) Σdy[48,64)
constexpr int kFieldK = 24, kFieldJ = 38, kFieldDy = 48;
constexpr int kMaxFlushTiles = 63;   // k <= 62 -> 6 bits per pixel, 252 pixels per lane and period
constexpr int kMaxRowsPerPeriod = 260;  // 252 * 260 < 2^16

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

// Four 64-bit read-modify-writes of the lane-private histogram entries of a quad's labels.
//   add_lo = 1 | k<<24 (tile-uniform) gets q<<8 from the key; add_hi = j<<6 | dy<<16 (0:0 = not counted).
// A dependent load->add->store chain per pixel costs ~4 x 35 cycles of exposed shared-memory latency
// per warp and tile, so all four entries are loaded first; where two pixels of the quad share a
// label the later one adds the earlier addends itself (its load is stale) and its store, issued
// later to the same address, wins.
__device__ __forceinline__ void hist_add4(uint32_t hist_lane, const uint32_t (&key)[4], uint32_t lo_base,
                                          const uint32_t (&add_hi)[4], const bool (&counted)[4]) {
    uint32_t addr[4];
    unsigned long long a[4], e[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        addr[j] = hist_lane - __byte_perm(key[j], 0u, 0x4404u);  // tag * 256
        const uint32_t add_lo = (key[j] & 0xff00u) | lo_base;
        a[j] = counted[j] ? ((static_cast<unsigned long long>(add_hi[j]) << 32) | add_lo) : 0ull;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        uint32_t e_lo, e_hi;
        asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(e_lo), "=r"(e_hi) : "r"(addr[j]));
        e[j] = (static_cast<unsigned long long>(e_hi) << 32) | e_lo;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        unsigned long long v = e[j] + a[j];
#pragma unroll
        for (int i = 0; i < j; ++i) v += addr[i] == addr[j] ? a[i] : 0ull;
        asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(addr[j]), "r"(static_cast<uint32_t>(v)),
                     "r"(static_cast<uint32_t>(v >> 32)));
    }
}

// ---- warp-private histogram flush ---------------------------------------------------------------
// Lane l (< C) owns label l: it walks the 32 lane columns of its row (skewed start: conflict-free),
// adding the packed fields two at a time into wide accumulators, and folds the totals into this
// warp's [C][4] u64 block with a plain read-modify-write (the block is private to the warp, so no
// shared-memory atomics).  Every lane then clears its own column.
template <int C>
__device__ __forceinline__ void warp_flush(unsigned long long *hist /*[C][32]*/, unsigned long long *blk_w /*[C][4]*/,
                                           int lane, int warp, int tile_px, long long r_base, uint32_t y_base, int w) {
    __syncwarp();
    for (int l = lane; l < C; l += 32) {
        unsigned long long nk = 0, qj = 0;
        uint32_t sdy = 0, cw = 0;
#pragma unroll 8
        for (int i = 0; i < 32; ++i) {
            const int c = (lane + i) & 31;
            const unsigned long long e = hist[l * 32 + c];
            nk += e & 0x0000003FFF0000FFull;          // n at 0 (headroom to 24), Σk at 24 (headroom to 64)
            qj += (e >> 8) & 0x000000FFC000FFFFull;   // Σq at 0 (headroom to 30), Σj at 30 (headroom to 64)
            sdy += static_cast<uint32_t>(e >> kFieldDy);
            cw += (static_cast<uint32_t>(e) & 0xFFu) * c;   // Σ n_c * c: lane offsets of the counted pixels
        }
        const long long n = static_cast<long long>(nk & 0xFFFFFFull), sk = static_cast<long long>(nk >> 24);
        const long long sq = static_cast<long long>(qj & 0x3FFFFFFFull), sj = static_cast<long long>(qj >> 30);
        if (n) {
            // Σr = Σ (r_base + k*TILE + 4*(32*warp + c) + j);  Σy = n*y_base + Σdy;  Σx = Σr - W*Σy
            const long long sr = n * r_base + sk * tile_px + 4 * (32ll * warp * n + cw) + sj;
            const long long sy = n * y_base + sdy;
            blk_w[l * 4 + SSB_STAT_N] += static_cast<unsigned long long>(n);
            blk_w[l * 4 + SSB_STAT_SX] += static_cast<unsigned long long>(sr - sy * w);
            blk_w[l * 4 + SSB_STAT_SY] += static_cast<unsigned long long>(sy);
            blk_w[l * 4 + SSB_STAT_SQ] += static_cast<unsigned long long>(sq);
        }
    }
    __syncwarp();
#pragma unroll
    for (int l = 0; l < C; ++l) hist[l * 32 + lane] = 0ull;
    __syncwarp();
}

// Scalar path for the few pixels past the last full tile (all inside frame f): plain byte loads,
// atomics straight to the scratch accumulator.
template <int C>
__device__ void scalar_pixels_global(const ArgmaxParams &p, long long px_begin, long long px_end, int tid, int nthr) {
    for (long long px = px_begin + tid; px < px_end; px += nthr) {
        const uint8_t *q = p.logits + px * C;
        int best = 0;
        uint32_t bv = q[0];
        for (int c = 1; c < C; ++c) {
            const uint32_t v = q[c];
            if (v > bv) { bv = v; best = c; }
        }
        if (p.labels) p.labels[px] = static_cast<uint8_t>(best);
        const long long f = px / p.hw;
        const uint32_t r = static_cast<uint32_t>(px - f * p.hw), y = r / p.w, x = r - y * p.w;
        if (x < static_cast<uint32_t>(p.roi_w) && y < static_cast<uint32_t>(p.roi_h)) {
            unsigned long long *s = p.acc + (f * C + best) * 4;
            atomicAdd(s + SSB_STAT_N, 1ull);
            atomicAdd(s + SSB_STAT_SX, static_cast<unsigned long long>(x));
            atomicAdd(s + SSB_STAT_SY, static_cast<unsigned long long>(y));
            atomicAdd(s + SSB_STAT_SQ, static_cast<unsigned long long>(bv));
        }
    }
}

template <int C, int NWARPS, int STAGES>
struct Cfg {
    static constexpr int kConsumers = NWARPS * 32;
    static constexpr int kThreads = kConsumers + 32;
    static constexpr int kTilePx = NWARPS * 128;
    static constexpr int kTileBytes = kTilePx * C;
    static constexpr int kHistBytes = NWARPS * C * 32 * 8;
    static constexpr int kBlkBytes = NWARPS * C * 4 * 8;   // one [C][4] u64 block per consumer warp
    static constexpr int kSmem = STAGES * kTileBytes + kHistBytes + kBlkBytes + 2 * STAGES * 8 + 32;
};

template <int C, int NWARPS, int STAGES>
__global__ void __launch_bounds__((NWARPS + 1) * 32) argmax_stats_u8_kernel(const ArgmaxParams p) {
    using K = Cfg<C, NWARPS, STAGES>;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *stage = smem;
    unsigned long long *hist_all = reinterpret_cast<unsigned long long *>(smem + STAGES * K::kTileBytes);
    unsigned long long *blk = hist_all + NWARPS * C * 32;
    uint64_t *full = reinterpret_cast<uint64_t *>(blk + NWARPS * C * 4);
    uint64_t *empty = full + STAGES;
    uint32_t *sel_tab = reinterpret_cast<uint32_t *>(empty + STAGES);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long long t0 = p.n_tiles * blockIdx.x / gridDim.x;
    const long long t1 = p.n_tiles * (blockIdx.x + 1) / gridDim.x;

    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], NWARPS);
        }
        mbar_init_fence();
        store_selector_table(sel_tab);
    }
    for (int i = tid; i < NWARPS * C * 32 + NWARPS * C * 4; i += K::kThreads) hist_all[i] = 0ull;
    __syncthreads();

    if (warp == NWARPS) {
        // ===== producer: one elected lane streams this CTA's span through the ring =====
        if (lane == 0) {
            const uint64_t pol = l2_policy_evict_first();
            int s = 0;
            uint32_t ph = 0;
            for (long long t = t0; t < t1; ++t) {
                mbar_wait(&empty[s], ph ^ 1);
                mbar_arrive_expect_tx(&full[s], K::kTileBytes);
                bulk_copy_g2s(stage + s * K::kTileBytes, p.logits + t * K::kTileBytes, K::kTileBytes, &full[s], pol);
                if (++s == STAGES) { s = 0; ph ^= 1; }
            }
        }
        return;
    }

    // ===== consumers =====
    unsigned long long *hist = hist_all + warp * C * 32;
    const uint32_t hist_lane = smem_addr(hist) + lane * 8 + 255u * 256u;  // entry(tag) = hist_lane - tag*256
    const int quad = warp * 32 + lane;       // quad of the tile this lane owns
    const int i0 = quad * 4;                 // first pixel (tile-relative)
    const Selectors sel = load_selectors(sel_tab);

    // frame bookkeeping (CTA-uniform): the current tile starts at in-frame offset r0 of frame f; the
    // lane-private histograms hold pixels r_base + k*TILE + i0 + j with rows >= y_base.
    const long long px0 = t0 * K::kTilePx;
    long long f = px0 / p.hw;
    int r0 = static_cast<int>(px0 - f * p.hw);
    int r_base = r0;
    uint32_t y_base = fastdiv(static_cast<uint32_t>(r0), p.divw);
    int k = 0;  // tiles since the last lane-private flush

    int covered = 0;  // pixels of frame f this CTA has processed since its last flush for f (ROI or not)
    __shared__ int s_last;

    // All consumer threads: add this CTA's share of frame f to scratch, then report `covered` pixels.
    // The CTA whose report completes the frame publishes it and leaves scratch zero again.
    auto frame_flush = [&]() {
        if (covered == 0) return;  // CTA-uniform: nothing accumulated for f
        warp_flush<C>(hist, blk + warp * C * 4, lane, warp, K::kTilePx, r_base, y_base, p.w);
        named_bar_sync(1, K::kConsumers);
        if (tid < C * 4) {
            unsigned long long v = 0ull;
#pragma unroll
            for (int wi = 0; wi < NWARPS; ++wi) {
                v += blk[wi * C * 4 + tid];
                blk[wi * C * 4 + tid] = 0ull;
            }
            if (v) atomicAdd(&p.acc[f * (C * 4) + tid], v);
            __threadfence();
        }
        named_bar_sync(1, K::kConsumers);
        if (tid == 0) {
            const unsigned int before = atomicAdd(&p.frame_px[f], static_cast<unsigned int>(covered));
            s_last = before + static_cast<unsigned int>(covered) == static_cast<unsigned int>(p.hw);
            __threadfence();
        }
        named_bar_sync(1, K::kConsumers);
        if (s_last) {
            if (tid < C * 4) p.stats[f * (C * 4) + tid] = atomicExch(&p.acc[f * (C * 4) + tid], 0ull);
            if (tid == 0) p.frame_px[f] = 0u;
        }
        covered = 0;
    };

    int s = 0;
    uint32_t ph = 0;
    uint32_t w[C], key[4];
    auto fetch = [&]() {  // wait for the next tile, pull this lane's quad into registers, free the stage
        mbar_wait(&full[s], ph);
        const uint32_t *src = reinterpret_cast<const uint32_t *>(stage + s * K::kTileBytes) + quad * C;
#pragma unroll
        for (int c = 0; c < C; ++c) w[c] = src[c];
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
        if (++s == STAGES) { s = 0; ph ^= 1; }
    };
    uint64_t pol_keep;  // label map: keep in L2 for whoever consumes it next (mask kernel, D2H copy)
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_keep));
    auto classify = [&](long long t) {  // argmax of the fetched quad + label store
#pragma unroll
        for (int j = 0; j < 4; ++j) key[j] = pixel_key<C>(w, j, sel);
        if (p.labels) {
            const uint32_t t01 = __byte_perm(key[0], key[1], 0x0040), t23 = __byte_perm(key[2], key[3], 0x0040);
            const uint32_t tags = __byte_perm(t01, t23, 0x5410);   // label = 255 - tag
            asm volatile("st.global.L2::cache_hint.u32 [%0], %1, %2;" ::"l"(p.labels + t * K::kTilePx + i0),
                         "r"(~tags), "l"(pol_keep)
                         : "memory");
        }
    };

    for (long long t = t0; t < t1; ++t) {
        fetch();
        classify(t);

        // statistics.  Tile classes (CTA-uniform): interior = one frame, every pixel inside
        // the ROI (no per-pixel tests); exterior = one frame, every pixel outside; mixed = the rest.
        const bool one_frame = r0 + K::kTilePx <= p.hw;
        const uint32_t y_first = fastdiv(static_cast<uint32_t>(r0), p.divw);
        const uint32_t y_last = fastdiv(static_cast<uint32_t>(r0 + K::kTilePx - 1), p.divw);
        if (one_frame && p.roi_w == p.w && y_last < static_cast<uint32_t>(p.roi_h)) {
            covered += K::kTilePx;
            const uint32_t rq = static_cast<uint32_t>(r0 + i0);
            const uint32_t yq = fastdiv(rq, p.divw);
            const uint32_t d = p.w - (rq - yq * p.w);           // pixels j >= d sit on the next row
            const uint32_t lo_base = 1u | (static_cast<uint32_t>(k) << kFieldK);
            const uint32_t hi_base = (yq - y_base) << (kFieldDy - 32);
            uint32_t add_hi[4];
#pragma unroll
            for (int j = 0; j < 4; ++j)
                add_hi[j] = hi_base + (j << (kFieldJ - 32)) + (static_cast<uint32_t>(j) >= d ? 1u << (kFieldDy - 32) : 0u);
            const bool all[4] = {true, true, true, true};
            hist_add4(hist_lane, key, lo_base, add_hi, all);
        } else if (one_frame && p.roi_w == p.w && y_first >= static_cast<uint32_t>(p.roi_h)) {
            covered += K::kTilePx;  // padding rows below the ROI: labels only
        } else {
            int lo = 0;  // one segment per frame the tile touches
            while (true) {
                const int hi = min(K::kTilePx, p.hw - r0);
                const uint32_t lo_base = 1u | (static_cast<uint32_t>(k) << kFieldK);
                uint32_t add_hi[4];
                bool counted[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int i = i0 + j;
                    const uint32_t r = static_cast<uint32_t>(r0 + i);
                    const uint32_t y = fastdiv(r, p.divw);
                    const uint32_t x = r - y * p.w;
                    counted[j] = (i >= lo) & (i < hi) & (x < static_cast<uint32_t>(p.roi_w)) &
                                 (y < static_cast<uint32_t>(p.roi_h));
                    add_hi[j] = (j << (kFieldJ - 32)) | ((y - y_base) << (kFieldDy - 32));
                }
                hist_add4(hist_lane, key, lo_base, add_hi, counted);
                covered += hi - lo;
                if (hi == K::kTilePx) break;
                frame_flush();  // frame f ends inside (or exactly before) this tile
                ++f;
                r0 -= p.hw;     // offsets of the rest of this tile in the new frame (k = 0 is this tile)
                r_base = r0;
                y_base = 0;
                k = 0;
                lo = hi;
            }
        }
        r0 += K::kTilePx;
        if (++k >= p.flush_tiles) {
            warp_flush<C>(hist, blk + warp * C * 4, lane, warp, K::kTilePx, r_base, y_base, p.w);
            // rows of all later pixels of this frame are >= the row of the next tile's first pixel
            // (r0 == hw means the next tile opens with a frame flush, which resets the bases)
            r_base = r0;
            y_base = fastdiv(static_cast<uint32_t>(r0), p.divw);
            k = 0;
        }
    }

    // stream tail (< one tile, not a multiple of 16 bytes): the last CTA reads it straight from
    // global memory, frame by frame
    if (blockIdx.x == gridDim.x - 1) {
        long long px = p.n_tiles * K::kTilePx;
        if (px < p.total_px && px == (f + 1) * p.hw) {  // the tiles ended exactly on a frame boundary
            frame_flush();
            ++f;
        }
        while (px < p.total_px) {
            const long long seg_end = min(p.total_px, (f + 1) * static_cast<long long>(p.hw));
            scalar_pixels_global<C>(p, px, seg_end, tid, K::kConsumers);
            covered += static_cast<int>(seg_end - px);
            px = seg_end;
            if (px < p.total_px) {
                frame_flush();
                ++f;
            }
        }
    }
    frame_flush();
}

}  // namespace
}  // namespace ssb

// ---- host side ------------------------------------------------------------------------------------
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "argmax_generic.cuh"

namespace ssb {
namespace {

// Largest lane-private flush period (in tiles) for which no packed field can overflow:
//   n  : 4*ft <= 252  (8 bits; also bounds Σq <= 252*255 < 2^16, Σk <= 252*62 < 2^14, Σj <= 756 < 2^10)
//   Σdy: 252*rows < 2^16 with rows = ft*TILE_PX/W + 2 (rows a flush period can span) <= 260
int pick_flush_tiles(int tile_px, int W) {
    int best = 0;
    for (int ft = 1; ft <= kMaxFlushTiles; ++ft)
        if (static_cast<long long>(ft) * tile_px / W + 2 <= kMaxRowsPerPeriod) best = ft;
    return best;
}

template <int C, int NWARPS, int STAGES>
int launch_fast(const ArgmaxParams &base, cudaStream_t st) {
    using K = Cfg<C, NWARPS, STAGES>;
    auto kern = argmax_stats_u8_kernel<C, NWARPS, STAGES>;
    DeviceInfo di;
    if (int rc = device_info(&di)) return rc;
    if (K::kSmem > di.max_smem_optin) return fail(SSB_ERR_LIMIT, "config needs %d B shared memory", K::kSmem);
    // per-device one-time attribute + occupancy query
    static int ctas_per_sm[64] = {0};
    int dev = 0;
    SSB_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(SSB_ERR_LIMIT, "device ordinal %d not supported", dev);
    if (ctas_per_sm[dev] == 0) {
        SSB_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, K::kSmem));
        int occ = 0;
        SSB_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, K::kThreads, K::kSmem));
        if (occ < 1) return fail(SSB_ERR_LIMIT, "kernel does not fit on an SM");
        ctas_per_sm[dev] = occ;
    }
    ArgmaxParams p = base;
    p.n_tiles = p.total_px / K::kTilePx;
    p.flush_tiles = pick_flush_tiles(K::kTilePx, p.w);
    if (p.flush_tiles < 1) return fail(SSB_ERR_LIMIT, "W=%d too narrow for the packed accumulators", p.w);
    long long grid = static_cast<long long>(di.sm_count) * ctas_per_sm[dev];
    if (const char *env = getenv("SSB_ARGMAX_WAVES")) grid *= atoi(env) > 0 ? atoi(env) : 1;  // tuning knob
    if (grid > p.n_tiles) grid = p.n_tiles > 0 ? p.n_tiles : 1;
    kern<<<static_cast<unsigned>(grid), K::kThreads, K::kSmem, st>>>(p);
    SSB_CUDA_TRY(cudaGetLastError());
    return SSB_OK;
}

template <int C>
int dispatch_cfg(const ArgmaxParams &p, cudaStream_t st, int nw, int stg) {
    if (nw == 4 && stg == 3) return launch_fast<C, 4, 3>(p, st);
    if (nw == 4 && stg == 4) return launch_fast<C, 4, 4>(p, st);
    if (nw == 4 && stg == 6) return launch_fast<C, 4, 6>(p, st);
    if (nw == 8 && stg == 3) return launch_fast<C, 8, 3>(p, st);
    if (nw == 8 && stg == 4) return launch_fast<C, 8, 4>(p, st);
    return fail(SSB_ERR_LIMIT, "no kernel compiled for NWARPS=%d STAGES=%d", nw, stg);
}

}  // namespace
}  // namespace ssb

extern "C" int ssb_argmax_stats_u8(const uint8_t *logits, int B, int H, int W, int C, int roi_h, int roi_w,
                                   uint8_t *labels, uint64_t *stats, void *stream) {
    using namespace ssb;
    if (int rc = check_argmax_args(logits, B, H, W, C, roi_h, roi_w, stats)) return rc;
    if (B == 0) return SSB_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    auto *stats64 = reinterpret_cast<unsigned long long *>(stats);
    // tuning knob (benchmarks only): SSB_ARGMAX_CFG="<consumer warps>,<stages>" or "generic"
    int nw = 8, stg = 3;   // best of the compiled set on B200 (tools/sweep_argmax.py)
    bool force_generic = false;
    if (const char *env = getenv("SSB_ARGMAX_CFG")) {
        if (!strcmp(env, "generic")) force_generic = true;
        else if (sscanf(env, "%d,%d", &nw, &stg) != 2) return fail(SSB_ERR_LIMIT, "bad SSB_ARGMAX_CFG '%s'", env);
    }
    const bool aligned = (reinterpret_cast<uintptr_t>(logits) % 16 == 0) && (reinterpret_cast<uintptr_t>(labels) % 4 == 0);
    const long long total_px = static_cast<long long>(B) * H * W;
    const bool fast = !force_generic && aligned && W >= 4 && (C == 19 || C == 21) && total_px >= 128ll * nw;
    if (!fast) {
        SSB_CUDA_TRY(cudaMemsetAsync(stats64, 0, sizeof(uint64_t) * 4 * C * static_cast<size_t>(B), st));
        return launch_generic<ElemU8>(logits, B, H, W, C, roi_h, roi_w, labels, stats64, st);
    }
    // scratch: [B][C][4] u64 accumulators + [B] u32 per-frame pixel counters, zero at rest
    const size_t acc_bytes = sizeof(uint64_t) * 4 * C * static_cast<size_t>(B);
    WorkspaceLease ws;
    if (int rc = workspace_acquire(acc_bytes + sizeof(unsigned int) * static_cast<size_t>(B), st, &ws)) return rc;

    ArgmaxParams p{};
    p.logits = logits;
    p.labels = labels;
    p.stats = stats64;
    p.total_px = total_px;
    p.acc = static_cast<unsigned long long *>(ws.ptr);
    p.frame_px = reinterpret_cast<unsigned int *>(static_cast<char *>(ws.ptr) + acc_bytes);
    p.hw = H * W;
    p.w = W;
    p.roi_h = roi_h;
    p.roi_w = roi_w;
    p.divw = make_fastdiv(static_cast<uint32_t>(W));
    const int rc = C == 21 ? dispatch_cfg<21>(p, st, nw, stg) : dispatch_cfg<19>(p, st, nw, stg);
    const int rc2 = workspace_release(ws, st);
    return rc ? rc : rc2;
}

// ---- entry point that never allocates ------------------------------------------------------------
// Same computation as ssb_argmax_stats_u8, but the streaming kernel's scratch (the zero-at-rest
// [B][C][4] accumulators and [B] pixel counters) lives in a fixed array inside the loaded module
// instead of coming from the cudaMalloc-backed workspace pool.  See include/ssb200.h for who needs
// this and what it costs.
namespace ssb {
namespace {

constexpr int kNoAllocMaxFrames = SSB_NOALLOC_MAX_FRAMES;
constexpr int kNoAllocMaxClasses = 21;  // the classes the streaming kernel is compiled for: 19 and 21
constexpr size_t kNoAllocAccBytes = sizeof(uint64_t) * 4 * kNoAllocMaxClasses * kNoAllocMaxFrames;
constexpr size_t kNoAllocBytes = kNoAllocAccBytes + sizeof(unsigned int) * kNoAllocMaxFrames;

// Resident on every GPU that loads the library until the process exits (~173 KB).
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

// In host code `g_noalloc_scratch` is only a handle for the symbol, not something a kernel can
// dereference: the device address has to be asked for, per device, with that device current.
int noalloc_scratch_address(unsigned char **out) {
    static unsigned char *cached[64] = {nullptr};
    static std::mutex mu;
    int dev = 0;
    SSB_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(SSB_ERR_LIMIT, "device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (!cached[dev]) {
        void *ptr = nullptr;
        SSB_CUDA_TRY(cudaGetSymbolAddress(&ptr, g_noalloc_scratch));
        cached[dev] = static_cast<unsigned char *>(ptr);
    }
    *out = cached[dev];
    return SSB_OK;
}

}  // namespace
}  // namespace ssb

extern "C" int ssb_argmax_stats_u8_noalloc(const uint8_t *logits, int B, int H, int W, int C, int roi_h, int roi_w,
                                           uint8_t *labels, uint64_t *stats, void *stream) {
    using namespace ssb;
    if (int rc = check_argmax_args(logits, B, H, W, C, roi_h, roi_w, stats)) return rc;
    if (B > kNoAllocMaxFrames)
        return fail(SSB_ERR_LIMIT, "ssb_argmax_stats_u8_noalloc takes at most %d frames per call (got %d); split the batch",
                    kNoAllocMaxFrames, B);
    if (B == 0) return SSB_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    auto *stats64 = reinterpret_cast<unsigned long long *>(stats);
    const int nw = 8, stg = 3;
    const bool aligned = (reinterpret_cast<uintptr_t>(logits) % 16 == 0) && (reinterpret_cast<uintptr_t>(labels) % 4 == 0);
    const long long total_px = static_cast<long long>(B) * H * W;
    if (!(aligned && W >= 4 && (C == 19 || C == 21) && total_px >= 128ll * nw)) {
        // shapes the streaming kernel does not cover use the generic kernel, which needs no scratch
        SSB_CUDA_TRY(cudaMemsetAsync(stats64, 0, sizeof(uint64_t) * 4 * C * static_cast<size_t>(B), st));
        return launch_generic<ElemU8>(logits, B, H, W, C, roi_h, roi_w, labels, stats64, st);
    }
    unsigned char *scratch = nullptr;
    if (int rc = noalloc_scratch_address(&scratch)) return rc;
    const size_t acc_bytes = sizeof(uint64_t) * 4 * C * static_cast<size_t>(B);
    const size_t used = acc_bytes + sizeof(unsigned int) * static_cast<size_t>(B);
    // The array is shared by every call and still holds whatever the previous call left (zeros if
    // it ran to completion, anything if it was cut short): start from a clean slate.
    SSB_CUDA_TRY(cudaMemsetAsync(scratch, 0, used, st));

    ArgmaxParams p{};
    p.logits = logits;
    p.labels = labels;
    p.stats = stats64;
    p.total_px = total_px;
    p.acc = reinterpret_cast<unsigned long long *>(scratch);
    p.frame_px = reinterpret_cast<unsigned int *>(scratch + acc_bytes);
    p.hw = H * W;
    p.w = W;
    p.roi_h = roi_h;
    p.roi_w = roi_w;
    p.divw = make_fastdiv(static_cast<uint32_t>(W));
    return C == 21 ? dispatch_cfg<21>(p, st, nw, stg) : dispatch_cfg<19>(p, st, nw, stg);
}
