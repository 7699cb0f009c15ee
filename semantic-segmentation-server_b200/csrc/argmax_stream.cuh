// K2+K3 streaming kernel: fused per-pixel argmax over C class scores + per-label {n, Σx, Σy, Σq_max},
// templated on a Codec that knows the element type (u8 quantised scores: argmax_stats_u8.cu, IEEE
// fp16: argmax_stats_f16.cu).
//
// Replaces the ArgMax tail of engine.run_inference (sem_seg_server.py:162), the reshape/crop of
// :163-170 and the counting/moment arithmetic of :102-122 restated per label (SURVEY.md §8 a5,a6,a9).
//
// Design (HBM-bound: every logit byte is read exactly once, one label byte per pixel written back):
//   * the batch is ONE flat byte stream [B*H*W*C]; frames are not 16-byte aligned (513*513*21 is
//     odd), tiles are.  A tile is TILE_PX = 128*NWARPS pixels (a multiple of 16 => its byte size is
//     a multiple of 16), moved global->shared by one TMA bulk copy (cp.async.bulk, UBLKCP) into a
//     STAGES-deep ring guarded by full/empty mbarriers; one producer warp, NWARPS consumer warps;
//   * each CTA owns a CONTIGUOUS span of tiles (grid = SMs x CTAs/SM), so it touches one or two
//     frames and flushes its statistics to global memory once per frame it touched;
//   * a lane owns 4 consecutive pixels of the tile; the codec loads them conflict-free, finds each
//     pixel's maximum with the smallest channel on ties (np.argmax / TFLite ArgMax semantics) and
//     returns a 32-bit key per pixel whose low byte is 255 - label;
//   * statistics: every lane has a PRIVATE histogram column in shared memory, hist[warp][label][lane]
//     (bank == lane, so no conflicts and no atomics whatever the label distribution); one
//     read-modify-write per pixel adds the packed 64-bit fields
//         n:8 | Σq:16 (u8 codec only) | Σk:14 | Σj:10 | Σdy:16
//     where k = tiles since the last flush, j = pixel of the quad, dy = row - y_base (the fp16 codec
//     keeps Σq as a second, 64-bit fixed-point word).  A lane's pixel positions are
//     r = r_base + k*TILE_PX + 4*quad + j, so Σr (hence Σx = Σr - W*Σy) is recovered at flush time
//     from n, Σk, Σj and the lane index: no per-pixel x is needed.  Before a field can overflow
//     (every `flush_tiles` tiles) the warp folds its 32 columns into its own [C][4] u64 block;
//   * no separate zeroing launch (a 43 KB memset in front of a 60 us kernel costs ~3 us): CTAs add
//     their per-frame block to a scratch accumulator that is ZERO AT REST and add the number of
//     pixels of the frame they covered to a per-frame counter; the CTA that completes a frame
//     (counter reaches H*W) moves the accumulated block to `stats` with plain stores and re-zeroes
//     the scratch.  Completion by pixel count is independent of how tiles are assigned to CTAs.
#pragma once
#include <cstdlib>
#include <cstring>

#include "argmax_generic.cuh"
#include "ssb_common.cuh"

namespace ssb {

struct ArgmaxParams {
    const uint8_t *logits;       // element type is the codec's; byte pointer for the flat stream
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

// packed lane-private word: n[0,8) Σq[8,24) Σk[24,38) Σj[38,48) Σdy[48,64)
constexpr int kFieldK = 24, kFieldJ = 38, kFieldDy = 48;
constexpr int kMaxFlushTiles = 63;      // k <= 62 -> 6 bits per pixel, 252 pixels per lane and period
constexpr int kMaxRowsPerPeriod = 260;  // 252 * 260 < 2^16

// Sums the packed words of one label row over the 32 lane columns (skewed start: conflict-free).
struct RowTotals {
    long long n, sq, sk, sj, sdy, cw;
};

// Folds one label's totals into the warp's [C][4] block.  Σr = Σ (r_base + k*TILE + 4*(32*warp + c) + j);
// Σy = n*y_base + Σdy;  Σx = Σr - W*Σy.
__device__ __forceinline__ void fold_row(unsigned long long *blk_row, const RowTotals &t, int warp, int tile_px,
                                         long long r_base, uint32_t y_base, int w) {
    if (!t.n) return;
    const long long sr = t.n * r_base + t.sk * tile_px + 4 * (32ll * warp * t.n + t.cw) + t.sj;
    const long long sy = t.n * y_base + t.sdy;
    blk_row[SSB_STAT_N] += static_cast<unsigned long long>(t.n);
    blk_row[SSB_STAT_SX] += static_cast<unsigned long long>(sr - sy * w);
    blk_row[SSB_STAT_SY] += static_cast<unsigned long long>(sy);
    blk_row[SSB_STAT_SQ] += static_cast<unsigned long long>(t.sq);
}

template <typename Codec, int NWARPS, int STAGES>
struct StreamCfg {
    static constexpr int C = Codec::kC;
    static constexpr int kConsumers = NWARPS * 32;
    static constexpr int kThreads = kConsumers + 32;
    static constexpr int kTilePx = NWARPS * 128;
    static constexpr int kTileBytes = kTilePx * C * Codec::kElemBytes;
    static constexpr int kHistWarpBytes = C * 32 * Codec::kEntryBytes;
    static constexpr int kHistBytes = NWARPS * kHistWarpBytes;
    static constexpr int kBlkBytes = NWARPS * C * 4 * 8;   // one [C][4] u64 block per consumer warp
    static constexpr int kSmem = STAGES * kTileBytes + kHistBytes + kBlkBytes + 2 * STAGES * 8 + 32;
};

// Scalar path for the few pixels past the last full tile (all inside frame f): plain loads, atomics
// straight to the scratch accumulator.
template <typename Codec>
__device__ void scalar_pixels_global(const ArgmaxParams &p, long long px_begin, long long px_end, int tid, int nthr) {
    using E = typename Codec::Elem;
    constexpr int C = Codec::kC;
    for (long long px = px_begin + tid; px < px_end; px += nthr) {
        const typename E::T *q = reinterpret_cast<const typename E::T *>(p.logits) + px * C;
        int best = 0;
        typename E::V bv = E::load(q);
        for (int c = 1; c < C; ++c) {
            const typename E::V v = E::load(q + c);
            if (E::gt(v, bv)) { bv = v; best = c; }
        }
        if (p.labels) p.labels[px] = static_cast<uint8_t>(best);
        const long long f = px / p.hw;
        const uint32_t r = static_cast<uint32_t>(px - f * p.hw), y = r / p.w, x = r - y * p.w;
        if (x < static_cast<uint32_t>(p.roi_w) && y < static_cast<uint32_t>(p.roi_h)) {
            unsigned long long *s = p.acc + (f * C + best) * 4;
            atomicAdd(s + SSB_STAT_N, 1ull);
            atomicAdd(s + SSB_STAT_SX, static_cast<unsigned long long>(x));
            atomicAdd(s + SSB_STAT_SY, static_cast<unsigned long long>(y));
            atomicAdd(s + SSB_STAT_SQ, E::fixed(q + best));
        }
    }
}

template <typename Codec, int NWARPS, int STAGES>
__global__ void __launch_bounds__((NWARPS + 1) * 32) argmax_stats_stream_kernel(const ArgmaxParams p) {
    using K = StreamCfg<Codec, NWARPS, STAGES>;
    constexpr int C = Codec::kC;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *stage = smem;
    uint8_t *hist_all = smem + STAGES * K::kTileBytes;
    unsigned long long *blk = reinterpret_cast<unsigned long long *>(hist_all + K::kHistBytes);
    uint64_t *full = reinterpret_cast<uint64_t *>(blk + NWARPS * C * 4);
    uint64_t *empty = full + STAGES;
    uint32_t *codec_tab = reinterpret_cast<uint32_t *>(empty + STAGES);   // 32 bytes of codec constants

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long long t0 = p.n_tiles * blockIdx.x / gridDim.x;
    const long long t1 = p.n_tiles * (blockIdx.x + 1) / gridDim.x;

    // Let the next kernel in the stream (typically the next batch's launch of this same kernel) place
    // its CTAs on SMs as ours retire and run its prologue early.
    pdl_trigger();

    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], NWARPS);
        }
        mbar_init_fence();
        Codec::store_table(codec_tab);
    }
    {
        unsigned long long *z = reinterpret_cast<unsigned long long *>(hist_all);
        for (int i = tid; i < (K::kHistBytes + K::kBlkBytes) / 8; i += K::kThreads) z[i] = 0ull;
    }
    __syncthreads();
    // Everything above touched shared memory only.  From here on global memory is read and written:
    // wait until the kernels this launch depends on have completed and their writes are visible
    // (a no-op when the launch was not programmatic).
    pdl_wait();

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
    uint8_t *hist = hist_all + warp * K::kHistWarpBytes;
    unsigned long long *blk_w = blk + warp * C * 4;
    // entry(tag) = hist_lane - tag * (32 * entry bytes), tag = 255 - label
    const uint32_t hist_lane = smem_addr(hist) + lane * Codec::kEntryBytes + 255u * 32u * Codec::kEntryBytes;
    const int quad = warp * 32 + lane;       // quad of the tile this lane owns
    const int i0 = quad * 4;                 // first pixel (tile-relative)
    const typename Codec::Ctx ctx = Codec::load_ctx(codec_tab);

    // frame bookkeeping (CTA-uniform): the current tile starts at in-frame offset r0 of frame f; the
    // lane-private histograms hold pixels r_base + k*TILE + i0 + j with rows >= y_base.
    const long long px0 = t0 * K::kTilePx;
    long long f = px0 / p.hw;
    int r0 = static_cast<int>(px0 - f * p.hw);
    int r_base = r0;
    uint32_t y_base = fastdiv(static_cast<uint32_t>(r0), p.divw);
    int k = 0;        // tiles since the last lane-private flush
    int covered = 0;  // pixels of frame f this CTA has processed since its last flush for f (ROI or not)
    __shared__ int s_last;

    auto warp_flush = [&]() {
        __syncwarp();
        for (int l = lane; l < C; l += 32) fold_row(blk_w + l * 4, Codec::row_totals(hist, l, lane), warp, K::kTilePx, r_base, y_base, p.w);
        __syncwarp();
        Codec::clear_column(hist, lane);
        __syncwarp();
    };

    // a [C][4] block has more entries than a 2-warp CTA has consumer threads: 1 round for NWARPS >= 4
    constexpr int kEntryIters = (C * 4 + K::kConsumers - 1) / K::kConsumers;
    // All consumer threads: add this CTA's share of frame f to scratch, then report `covered` pixels.
    // The CTA whose report completes the frame publishes it and leaves scratch zero again.
    auto frame_flush = [&]() {
        if (covered == 0) return;  // CTA-uniform: nothing accumulated for f
        warp_flush();
        named_bar_sync(1, K::kConsumers);
#pragma unroll
        for (int it = 0; it < kEntryIters; ++it) {
            const int e = tid + it * K::kConsumers;
            if (e < C * 4) {
                unsigned long long v = 0ull;
#pragma unroll
                for (int wi = 0; wi < NWARPS; ++wi) {
                    v += blk[wi * C * 4 + e];
                    blk[wi * C * 4 + e] = 0ull;
                }
                if (v) atomicAdd(&p.acc[f * (C * 4) + e], v);
            }
        }
        // every consumer thread: its adds to scratch (these, and those of the scalar stream tail) are
        // ordered device-wide before the pixel count that may let another CTA collect the frame
        __threadfence();
        named_bar_sync(1, K::kConsumers);
        if (tid == 0) {
            const unsigned int before = atomicAdd(&p.frame_px[f], static_cast<unsigned int>(covered));
            s_last = before + static_cast<unsigned int>(covered) == static_cast<unsigned int>(p.hw);
            __threadfence();
        }
        named_bar_sync(1, K::kConsumers);
        if (s_last) {
#pragma unroll
            for (int it = 0; it < kEntryIters; ++it) {
                const int e = tid + it * K::kConsumers;
                if (e < C * 4) p.stats[f * (C * 4) + e] = atomicExch(&p.acc[f * (C * 4) + e], 0ull);
            }
            if (tid == 0) p.frame_px[f] = 0u;
        }
        covered = 0;
    };

    uint64_t pol_keep;  // label map: keep in L2 for whoever consumes it next (mask kernel, D2H copy)
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_keep));

    int s = 0;
    uint32_t ph = 0;
    for (long long t = t0; t < t1; ++t) {
        // wait for the tile, pull this lane's quad into registers, free the stage
        mbar_wait(&full[s], ph);
        uint32_t w[Codec::kQuadWords];
        Codec::load_quad(stage + s * K::kTileBytes, quad, w);
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
        if (++s == STAGES) { s = 0; ph ^= 1; }

        uint32_t key[4];  // low byte = 255 - label; the rest belongs to the codec (winning score)
        Codec::classify(w, ctx, key);
        if (p.labels) {
            const uint32_t t01 = __byte_perm(key[0], key[1], 0x0040), t23 = __byte_perm(key[2], key[3], 0x0040);
            const uint32_t tags = __byte_perm(t01, t23, 0x5410);
            asm volatile("st.global.L2::cache_hint.u32 [%0], %1, %2;" ::"l"(p.labels + t * K::kTilePx + i0),
                         "r"(~tags), "l"(pol_keep)
                         : "memory");
        }

        // statistics.  Tile classes (CTA-uniform): interior = one frame, every pixel inside the ROI
        // (no per-pixel tests); exterior = one frame, every row below the ROI; border = one frame,
        // per-pixel ROI tests; the rest = tiles that straddle two frames.
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
            Codec::add4(hist_lane, key, lo_base, add_hi, all);
        } else if (one_frame && y_first >= static_cast<uint32_t>(p.roi_h)) {
            covered += K::kTilePx;  // padding rows below the ROI: labels only
        } else if (one_frame) {
            // the ROI border runs through the tile (or roi_w < W): per-pixel tests, coordinates stepped
            // from the quad's first pixel (W >= 4: at most one row wrap inside a quad)
            covered += K::kTilePx;
            const uint32_t rq = static_cast<uint32_t>(r0 + i0);
            const uint32_t yq = fastdiv(rq, p.divw);
            const uint32_t xq = rq - yq * p.w;
            const uint32_t lo_base = 1u | (static_cast<uint32_t>(k) << kFieldK);
            uint32_t add_hi[4];
            bool counted[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const bool wrap = xq + j >= static_cast<uint32_t>(p.w);
                const uint32_t x = xq + j - (wrap ? p.w : 0), y = yq + (wrap ? 1u : 0u);
                counted[j] = (x < static_cast<uint32_t>(p.roi_w)) & (y < static_cast<uint32_t>(p.roi_h));
                add_hi[j] = (j << (kFieldJ - 32)) | ((y - y_base) << (kFieldDy - 32));
            }
            Codec::add4(hist_lane, key, lo_base, add_hi, counted);
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
                Codec::add4(hist_lane, key, lo_base, add_hi, counted);
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
            warp_flush();
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
            scalar_pixels_global<Codec>(p, px, seg_end, tid, K::kConsumers);
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

// ---- host side ------------------------------------------------------------------------------------

// Largest lane-private flush period (in tiles) for which no packed field can overflow:
//   n  : 4*ft <= 252  (8 bits; also bounds Σq <= 252*255 < 2^16, Σk <= 252*62 < 2^14, Σj <= 756 < 2^10)
//   Σdy: 252*rows < 2^16 with rows = ft*TILE_PX/W + 2 (rows a flush period can span) <= 260
inline int pick_flush_tiles(int tile_px, int W) {
    int best = 0;
    for (int ft = 1; ft <= kMaxFlushTiles; ++ft)
        if (static_cast<long long>(ft) * tile_px / W + 2 <= kMaxRowsPerPeriod) best = ft;
    return best;
}

template <typename Codec, int NWARPS, int STAGES>
int launch_stream(const ArgmaxParams &base, cudaStream_t st) {
    using K = StreamCfg<Codec, NWARPS, STAGES>;
    auto kern = argmax_stats_stream_kernel<Codec, NWARPS, STAGES>;
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
    SSB_CUDA_TRY(launch_pdl(kern, dim3(static_cast<unsigned>(grid)), dim3(K::kThreads), K::kSmem, st, p));
    return SSB_OK;
}

// Streaming kernels need 16-byte aligned logits, 4-byte aligned labels, W >= 4 and at least one tile.
inline bool stream_path_ok(const void *logits, const void *labels, int W, long long total_px, int nw) {
    return reinterpret_cast<uintptr_t>(logits) % 16 == 0 && reinterpret_cast<uintptr_t>(labels) % 4 == 0 && W >= 4 &&
           total_px >= 128ll * nw;
}

inline ArgmaxParams make_params(const void *logits, int B, int H, int W, int C, int roi_h, int roi_w, uint8_t *labels,
                                unsigned long long *stats, void *scratch) {
    ArgmaxParams p{};
    const size_t acc_bytes = sizeof(uint64_t) * 4 * C * static_cast<size_t>(B);
    p.logits = static_cast<const uint8_t *>(logits);
    p.labels = labels;
    p.stats = stats;
    p.total_px = static_cast<long long>(B) * H * W;
    p.acc = static_cast<unsigned long long *>(scratch);
    p.frame_px = reinterpret_cast<unsigned int *>(static_cast<char *>(scratch) + acc_bytes);
    p.hw = H * W;
    p.w = W;
    p.roi_h = roi_h;
    p.roi_w = roi_w;
    p.divw = make_fastdiv(static_cast<uint32_t>(W));
    return p;
}
inline size_t scratch_bytes(int B, int C) {
    return sizeof(uint64_t) * 4 * C * static_cast<size_t>(B) + sizeof(unsigned int) * static_cast<size_t>(B);
}

}  // namespace ssb
