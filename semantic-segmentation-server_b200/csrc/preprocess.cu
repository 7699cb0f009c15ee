// K1: camera frame -> model input.  BGR->RGB swap + Pillow-NEAREST keep-aspect resize + zero pad,
// fused into one gather (replaces sem_seg_server.py:150-161: cv2.cvtColor, Image.fromarray,
// edgetpu resampling_with_original_ratio(NEAREST) = Image.resize + ImageOps.expand, np.asarray().flatten()).
//
// One WARP per (frame, output row), two independent warps per CTA, no block-wide barriers:
//   1. the sampled source row (Wc*3 bytes, 5760 B at 1080p) is copied into the warp's shared memory
//      by ONE bulk async copy (cp.async.bulk, completion on the warp's own mbarrier; plain loads for
//      a ragged head / tail) -- every 32-byte sector of a sampled row is touched by the 11.2-byte
//      sample stride anyway, so reading the row densely costs no extra DRAM traffic;
//   2. each lane gathers quads of pixels (source offsets from the host-built Pillow index table, one
//      128-bit load per quad, issued while the row copy is still in flight: the kernel is bound by
//      the chain table load -> row copy -> gather -> store of a warp, not by bytes), swaps B and R
//      while packing 12 bytes into three words, and stores them word-wise into a staging row;
//   3. the output row (W*3 bytes, 1539 B: its 16-byte phase changes from row to row) leaves as
//      128-bit global stores; the staging row is re-aligned to that phase on the way out with funnel
//      shifts (byte stores only for the ragged head and tail).
// Rows that are pure padding (y >= new_h, ImageOps.expand) skip 1-2 and store zeros.  Index tables
// are built on the host (Pillow accumulates `xo += scale` in double; see the host module's
// nearest_index_table).  An earlier version used one CTA per row with byte-wise staging and was
// issue-bound (65 % issue slots at 30 % DRAM utilisation).
#include <cstring>

#include "ssb_common.cuh"

namespace ssb {
namespace {

#ifndef SSB_PRE_TRIGGER
#define SSB_PRE_TRIGGER 0   // tuning: 1 = let the next kernel's CTAs be placed from the top of every CTA, 2 = once its row is read
#endif
#ifndef SSB_PRE_WARPS
#define SSB_PRE_WARPS 2   // warps per CTA: 1 -> 35.0 us, 2 -> 29.0, 4 -> 29.8, 8 -> 32.5 (64 frames of 1080p)
#endif
constexpr int kPreWarps = SSB_PRE_WARPS, kPreThreads = kPreWarps * 32;
constexpr int kPrePrefetch = 5;   // quads per lane whose table entries are prefetched: 5 * 32 * 4 >= 513 columns

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr(smem_dst)), "l"(gmem_src) : "memory");
}
#ifndef SSB_PRE_BULK
#define SSB_PRE_BULK 1      // 1: the 16-byte aligned middle of a row is ONE bulk async copy per warp; 0: 16-byte cp.async per lane
#endif
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

template <bool kStore32>
__global__ void __launch_bounds__(kPreThreads)
preprocess_nearest_kernel(const uint8_t *__restrict__ frames, long long frame_stride, int Wc,
                          const int *__restrict__ xtab, const int *__restrict__ ytab, int new_w, int new_h,
                          uint8_t *__restrict__ out, int H, int W, int rows_total, int src_bytes, int dst_bytes) {
    extern __shared__ __align__(16) uint8_t smem[];
#if SSB_PRE_TRIGGER == 1
    pdl_trigger();
#endif
    pdl_wait();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row = blockIdx.x * kPreWarps + warp;          // flat (frame, output row)
    if (row >= rows_total) return;
    const int b = row / H, y = row - b * H;
    uint8_t *src = smem + warp * (src_bytes + dst_bytes);    // [<= 15 B phase][Wc*3]
    uint32_t *dst = reinterpret_cast<uint32_t *>(src + src_bytes);   // [W*3 rounded up to words + spare]

    const int out_bytes = W * 3;
    uint8_t *grow = out + static_cast<long long>(row) * out_bytes;
    const int dphase = static_cast<int>(reinterpret_cast<uintptr_t>(grow) & 15);
    const int ohead = min(out_bytes, (16 - dphase) & 15);
    const int obody16 = (out_bytes - ohead) >> 4;
    uint4 *o4 = reinterpret_cast<uint4 *>(grow + ohead);

    if (y >= new_h) {  // padding row: zeros straight to global memory
        for (int i = lane; i < ohead; i += 32) grow[i] = 0;
        for (int i = lane; i < obody16; i += 32) o4[i] = make_uint4(0u, 0u, 0u, 0u);
        for (int i = ohead + obody16 * 16 + lane; i < out_bytes; i += 32) grow[i] = 0;
        return;
    }

    // 1. source row -> shared memory (same 16-byte phase as in global memory)
    const int row_bytes = Wc * 3;
    const uint8_t *srow = frames + b * frame_stride + static_cast<long long>(ytab[y]) * row_bytes;
    const int sphase = static_cast<int>(reinterpret_cast<uintptr_t>(srow) & 15);
    const int head = min(row_bytes, (16 - sphase) & 15);
    const int body16 = (row_bytes - head) >> 4;
    uint8_t *sp = src + sphase;
#if SSB_PRE_BULK
    // the 16-byte aligned middle of the row as ONE bulk async copy (UBLKCP) issued by lane 0, completion on the warp's
    // own mbarrier (a dozen 16-byte cp.async per lane took 16 % of the kernel's instructions)
    __shared__ uint64_t row_bar[kPreWarps];
    if (lane == 0) {
        mbar_init(&row_bar[warp], 1);
        mbar_init_fence();
        if (body16 > 0) {
            mbar_arrive_expect_tx(&row_bar[warp], static_cast<uint32_t>(body16) * 16u);
            bulk_copy_g2s_plain(sp + head, srow + head, static_cast<uint32_t>(body16) * 16u, &row_bar[warp]);
        }
    }
    __syncwarp();
    for (int i = lane; i < head; i += 32) sp[i] = srow[i];
#else
    for (int i = lane; i < head; i += 32) sp[i] = srow[i];
    for (int i = lane; i < body16; i += 32) cp_async16(sp + head + i * 16, srow + head + i * 16);
#endif
    for (int i = head + body16 * 16 + lane; i < row_bytes; i += 32) sp[i] = srow[i];
    // the index-table entries of this lane's first quads are fetched while the row is in flight
    const int quads = (W + 3) >> 2;
    int4 pre[kPrePrefetch];
#pragma unroll
    for (int k = 0; k < kPrePrefetch; ++k) {
        const int x0 = (lane + 32 * k) * 4;
        pre[k] = x0 + 3 < new_w ? *reinterpret_cast<const int4 *>(xtab + x0) : make_int4(0, 0, 0, 0);   // 16-byte aligned table
    }
#if SSB_PRE_BULK
    if (body16 > 0) mbar_wait(&row_bar[warp], 0);
#else
    cp_async_wait_all();
#endif
    __syncwarp();
#if SSB_PRE_TRIGGER == 2
    pdl_trigger();   // this warp's reads are done
#endif

    // 2. gather: quads of output pixels -> three packed words each
    auto do_quad = [&](int q, const int4 &t) {
        const int x0 = q * 4;
        int xs[4];
        if (x0 + 3 < new_w) {
            xs[0] = t.x; xs[1] = t.y; xs[2] = t.z; xs[3] = t.w;
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) xs[j] = x0 + j < new_w ? xtab[x0 + j] : -1;
        }
        uint32_t px[4];   // R | G<<8 | B<<16 of each pixel (source order is B,G,R)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (xs[j] >= 0) {
                // the pixel's 3 bytes sit in one or two aligned words: two 32-bit loads, a funnel shift and a
                // byte permute (three byte loads per pixel took half again as many shared-memory wavefronts
                // at the 11.2-byte sample stride)
                const uint32_t o = sphase + static_cast<uint32_t>(xs[j]) * 3u;
                const uint32_t *wp = reinterpret_cast<const uint32_t *>(src) + (o >> 2);
                const uint32_t bgr = __funnelshift_r(wp[0], wp[1], (o & 3u) * 8u);
                px[j] = __byte_perm(bgr, 0u, 0x4012);
            } else {
                px[j] = 0u;   // right of the resized image: ImageOps.expand padding
            }
        }
        // 12 bytes R0 G0 B0 R1 | G1 B1 R2 G2 | B2 R3 G3 B3
        dst[3 * q + 0] = px[0] | (px[1] << 24);
        dst[3 * q + 1] = (px[1] >> 8) | (px[2] << 16);
        dst[3 * q + 2] = (px[2] >> 16) | (px[3] << 8);
    };
#pragma unroll
    for (int k = 0; k < kPrePrefetch; ++k)
        if (lane + 32 * k < quads) do_quad(lane + 32 * k, pre[k]);
    for (int q = lane + 32 * kPrePrefetch; q < quads; q += 32) {   // model inputs wider than 640 columns
        const int x0 = q * 4;
        do_quad(q, x0 + 3 < new_w ? *reinterpret_cast<const int4 *>(xtab + x0) : make_int4(0, 0, 0, 0));
    }
    __syncwarp();

    const uint8_t *dbytes = reinterpret_cast<const uint8_t *>(dst);
    if (kStore32) {
        // 3. staging row -> global as 32-bit words: lane l moves word l of every group of 32, so both the
        //    shared-memory reads (two per word, for the re-alignment by 0-3 bytes) and the global stores
        //    (128 contiguous bytes per instruction) are conflict-free; byte stores for the ragged ends
        const int h4 = min(out_bytes, static_cast<int>((4 - (reinterpret_cast<uintptr_t>(grow) & 3)) & 3));
        for (int i = lane; i < h4; i += 32) grow[i] = dbytes[i];
        const int nwords = (out_bytes - h4) >> 2;
        const int sh = h4 * 8;
        uint32_t *gw = reinterpret_cast<uint32_t *>(grow + h4);
        for (int i = lane; i < nwords; i += 32) gw[i] = __funnelshift_r(dst[i], dst[i + 1], sh);
        for (int i = h4 + nwords * 4 + lane; i < out_bytes; i += 32) grow[i] = dbytes[i];
        return;
    }
    // 3. staging row -> global: byte head, 128-bit body re-aligned by (ohead mod 4) bytes, byte tail
    for (int i = lane; i < ohead; i += 32) grow[i] = dbytes[i];
    const int wsh = (ohead & 3) * 8;          // bit offset of the body inside the staging words
    const uint32_t *dw = dst + (ohead >> 2);  // first staging word that holds body bytes
    for (int i = lane; i < obody16; i += 32) {
        const uint32_t *w = dw + 4 * i;
        const uint32_t w0 = w[0], w1 = w[1], w2 = w[2], w3 = w[3], w4 = w[4];
        uint4 v;
        v.x = __funnelshift_r(w0, w1, wsh);
        v.y = __funnelshift_r(w1, w2, wsh);
        v.z = __funnelshift_r(w2, w3, wsh);
        v.w = __funnelshift_r(w3, w4, wsh);
        o4[i] = v;
    }
    for (int i = ohead + obody16 * 16 + lane; i < out_bytes; i += 32) grow[i] = dbytes[i];
}

}  // namespace
}  // namespace ssb

extern "C" int ssb_preprocess_nearest_u8(const uint8_t *frames_bgr, int B, int Hc, int Wc, int64_t frame_stride,
                                         const int32_t *xtab, const int32_t *ytab, int new_w, int new_h,
                                         uint8_t *out_rgb, int H, int W, void *stream) {
    using namespace ssb;
    if (B < 0 || Hc <= 0 || Wc <= 0 || H <= 0 || W <= 0)
        return fail(SSB_ERR_BAD_SHAPE, "bad shape B=%d Hc=%d Wc=%d H=%d W=%d", B, Hc, Wc, H, W);
    if (new_w < 1 || new_w > W || new_h < 1 || new_h > H)
        return fail(SSB_ERR_BAD_SHAPE, "resized size (%d,%d) outside model input (%d,%d)", new_w, new_h, W, H);
    if (frame_stride < static_cast<int64_t>(Hc) * Wc * 3)
        return fail(SSB_ERR_BAD_SHAPE, "frame_stride %lld < Hc*Wc*3", static_cast<long long>(frame_stride));
    if (B == 0) return SSB_OK;
    if (!frames_bgr || !xtab || !ytab || !out_rgb) return fail(SSB_ERR_BAD_POINTER, "NULL buffer");
    if (reinterpret_cast<uintptr_t>(xtab) % 16 != 0) return fail(SSB_ERR_BAD_POINTER, "xtab must be 16-byte aligned");
    const long long rows_total = static_cast<long long>(B) * H;
    if (rows_total >= (1ll << 31)) return fail(SSB_ERR_LIMIT, "B*H must be < 2^31");
    const int src_bytes = ((Wc * 3 + 15 + 15) / 16) * 16;
    const int dst_bytes = (((W + 3) / 4 * 12 + 16 + 15) / 16) * 16;   // whole quads + spare words for the funnel
    const int smem = kPreWarps * (src_bytes + dst_bytes);
    DeviceInfo di;
    if (int rc = device_info(&di)) return rc;
    if (smem > di.max_smem_optin) return fail(SSB_ERR_LIMIT, "row of %d px needs %d B shared memory", Wc, smem);
    const char *env = getenv("SSB_PRE_STORE");   // tuning knob: "128" = 128-bit stores from 4-way conflicting reads
    const bool store32 = !(env && !strcmp(env, "128"));
    auto kern = store32 ? preprocess_nearest_kernel<true> : preprocess_nearest_kernel<false>;
    if (smem > 48 * 1024) SSB_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const unsigned grid = static_cast<unsigned>((rows_total + kPreWarps - 1) / kPreWarps);
    SSB_CUDA_TRY(launch_pdl(kern, dim3(grid), dim3(kPreThreads), smem,
                            static_cast<cudaStream_t>(stream), frames_bgr, static_cast<long long>(frame_stride), Wc, xtab,
                            ytab, new_w, new_h, out_rgb, H, W, static_cast<int>(rows_total), src_bytes, dst_bytes));
    return SSB_OK;
}
