// K1: camera frame -> model input.  BGR->RGB swap + Pillow-NEAREST keep-aspect resize + zero pad,
// fused into one gather (replaces sem_seg_server.py:150-161: cv2.cvtColor, Image.fromarray,
// edgetpu resampling_with_original_ratio(NEAREST) = Image.resize + ImageOps.expand, np.asarray().flatten()).
//
// One CTA per (frame, output row).  A sampled source row (Wc*3 bytes, 5760 B at 1080p) is staged in
// shared memory with coalesced 128-bit loads -- every 32-byte sector of a sampled row is touched by
// the 11.2-byte sample stride anyway, so reading the row densely costs no extra DRAM traffic and
// makes the read side fully vectorised.  The output row (W*3 bytes, 1539 B: never 16-byte aligned
// from row to row) is assembled in shared memory at the same 16-byte phase as its global address and
// written back with 128-bit stores (byte stores only for the ragged head and tail).  Rows that are
// pure padding (y >= new_h) skip the read entirely.  Index tables are built on the host
// (Pillow accumulates `xo += scale` in double; see semseg host module nearest_index_table).
#include "ssb_common.cuh"

namespace ssb {
namespace {

constexpr int kPreThreads = 128;

__global__ void __launch_bounds__(kPreThreads)
preprocess_nearest_kernel(const uint8_t *__restrict__ frames, long long frame_stride, int Wc,
                          const int *__restrict__ xtab, const int *__restrict__ ytab, int new_w, int new_h,
                          uint8_t *__restrict__ out, int H, int W, int src_smem_bytes) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint8_t *src = smem;                       // [<= 15 B phase][Wc*3], 16-byte phase of the global row
    uint8_t *dst = smem + src_smem_bytes;      // [<= 15 B phase][W*3]
    const int y = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
    const int out_bytes = W * 3;
    uint8_t *grow = out + (static_cast<long long>(b) * H + y) * out_bytes;
    const int dphase = static_cast<int>(reinterpret_cast<uintptr_t>(grow) & 15);

    if (y < new_h) {
        const int row_bytes = Wc * 3;
        const uint8_t *srow = frames + b * frame_stride + static_cast<long long>(ytab[y]) * row_bytes;
        const int sphase = static_cast<int>(reinterpret_cast<uintptr_t>(srow) & 15);
        // head (to the next 16-byte boundary), 128-bit body, tail
        const int head = min(row_bytes, (16 - sphase) & 15);
        const int body16 = (row_bytes - head) >> 4;
        for (int i = tid; i < head; i += kPreThreads) src[sphase + i] = srow[i];
        const uint4 *g4 = reinterpret_cast<const uint4 *>(srow + head);
        uint4 *s4 = reinterpret_cast<uint4 *>(src + sphase + head);
        for (int i = tid; i < body16; i += kPreThreads) {
            uint4 v;
            asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                         : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                         : "l"(g4 + i));
            s4[i] = v;
        }
        for (int i = head + body16 * 16 + tid; i < row_bytes; i += kPreThreads) src[sphase + i] = srow[i];
        __syncthreads();
        const uint8_t *sp = src + sphase;
        for (int x = tid; x < W; x += kPreThreads) {
            uint8_t r = 0, g = 0, bl = 0;
            if (x < new_w) {
                const uint8_t *px = sp + xtab[x] * 3;  // source is B,G,R
                bl = px[0]; g = px[1]; r = px[2];
            }
            uint8_t *o = dst + dphase + x * 3;
            o[0] = r; o[1] = g; o[2] = bl;
        }
    }
    const int ohead = min(out_bytes, (16 - dphase) & 15);
    const int obody16 = (out_bytes - ohead) >> 4;
    uint4 *o4 = reinterpret_cast<uint4 *>(grow + ohead);
    if (y >= new_h) {  // padding row (ImageOps.expand): zeros straight to global memory
        for (int i = tid; i < ohead; i += kPreThreads) grow[i] = 0;
        for (int i = tid; i < obody16; i += kPreThreads) o4[i] = make_uint4(0u, 0u, 0u, 0u);
        for (int i = ohead + obody16 * 16 + tid; i < out_bytes; i += kPreThreads) grow[i] = 0;
        return;
    }
    __syncthreads();
    for (int i = tid; i < ohead; i += kPreThreads) grow[i] = dst[dphase + i];
    const uint4 *d4 = reinterpret_cast<const uint4 *>(dst + dphase + ohead);
    for (int i = tid; i < obody16; i += kPreThreads) o4[i] = d4[i];
    for (int i = ohead + obody16 * 16 + tid; i < out_bytes; i += kPreThreads) grow[i] = dst[dphase + i];
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
    if (H > 65535 || B > 65535) return fail(SSB_ERR_LIMIT, "H and B must be <= 65535");
    const int src_smem = ((Wc * 3 + 15 + 15) / 16) * 16;
    const int dst_smem = ((W * 3 + 15 + 15) / 16) * 16;
    const int smem = src_smem + dst_smem;
    DeviceInfo di;
    if (int rc = device_info(&di)) return rc;
    if (smem > di.max_smem_optin) return fail(SSB_ERR_LIMIT, "row of %d px needs %d B shared memory", Wc, smem);
    if (smem > 48 * 1024)
        SSB_CUDA_TRY(cudaFuncSetAttribute(preprocess_nearest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    dim3 grid(H, B);
    preprocess_nearest_kernel<<<grid, kPreThreads, smem, static_cast<cudaStream_t>(stream)>>>(
        frames_bgr, frame_stride, Wc, xtab, ytab, new_w, new_h, out_rgb, H, W, src_smem);
    SSB_CUDA_TRY(cudaGetLastError());
    return SSB_OK;
}
