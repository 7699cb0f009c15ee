// Foreground mask of the reference's contour stage, as exact integer arithmetic on the label map:
// colour = colormap[label] (sem_seg_server.py:173) -> cv2.blur 3x3 with BORDER_REFLECT_101 (:79)
// -> cv2.cvtColor(BGR2GRAY) applied to the RGB image (:82; channel 0 gets the blue weight) ->
// threshold 127 (:85).  blur rounds s/9 half-up: b = (2s+9)/18; gray = (b0*1868 + b1*9617 +
// b2*4899 + 8192) >> 14.  The image the reference blurs is the CROPPED roi_h x roi_w map, so the
// reflection happens at the ROI border, not the frame border.
#include "ssb_common.cuh"

namespace ssb {
namespace {

__device__ __forceinline__ int reflect101(int i, int n) {
    if (n == 1) return 0;
    if (i < 0) i = -i;
    if (i >= n) i = 2 * (n - 1) - i;
    return i;
}

// Each thread produces 4 horizontally adjacent mask pixels: 6 column sums (3 rows each) of packed
// colours are shared by the 4 box sums, so a pixel costs 4.5 label loads and 4.5 LUT reads instead
// of 9 + 9.  The colour LUT packs channel 0/1 into 16-bit fields of one word (box sums <= 9*255 fit)
// and channel 2 into a second word, so a box sum is 2 integer adds per neighbour.
constexpr int kMaskPxPerThread = 4, kMaskTileW = 32 * kMaskPxPerThread, kMaskTileH = 8;
constexpr int kMaskTilesPerCta = 8;   // vertical tiles per CTA: the LUT set-up (global loads + barrier) is paid once per 64 rows

__global__ void __launch_bounds__(256)
fg_mask_kernel(const uint8_t *__restrict__ labels, int H, int W, int roi_h, int roi_w,
               const uint8_t *__restrict__ colormap, uint8_t *__restrict__ mask) {
    __shared__ uint2 cmap[256];  // .x = c0 | c1<<16, .y = c2
    pdl_trigger();
    pdl_wait();
    for (int i = threadIdx.x; i < 256; i += blockDim.x)
        cmap[i] = make_uint2(colormap[3 * i] | (colormap[3 * i + 1] << 16), colormap[3 * i + 2]);
    __syncthreads();
    const int x0 = blockIdx.x * kMaskTileW + (threadIdx.x & 31) * kMaskPxPerThread;
    if (x0 >= W) return;
    for (int ty = 0; ty < kMaskTilesPerCta; ++ty) {
    const int y = (blockIdx.y * kMaskTilesPerCta + ty) * kMaskTileH + (threadIdx.x >> 5);
    if (y >= H) return;
    uint8_t *out = mask + static_cast<long long>(blockIdx.z) * H * W + static_cast<long long>(y) * W + x0;
    const int n_out = min(kMaskPxPerThread, W - x0);
    if (y >= roi_h || x0 >= roi_w) {
        for (int j = 0; j < n_out; ++j) out[j] = 0;
        continue;
    }
    const uint8_t *frame = labels + static_cast<long long>(blockIdx.z) * H * W;
    const uint8_t *rows[3] = {frame + static_cast<long long>(reflect101(y - 1, roi_h)) * W,
                              frame + static_cast<long long>(y) * W,
                              frame + static_cast<long long>(reflect101(y + 1, roi_h)) * W};
    uint32_t col01[kMaskPxPerThread + 2], col2[kMaskPxPerThread + 2];
#pragma unroll
    for (int i = 0; i < kMaskPxPerThread + 2; ++i) {
        // columns right of the ROI are never used by a counted pixel; clamp keeps the load in range
        const int xc = reflect101(min(x0 - 1 + i, roi_w), roi_w);
        uint32_t a = 0, c = 0;
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const uint2 e = cmap[rows[r][xc]];
            a += e.x;
            c += e.y;
        }
        col01[i] = a;
        col2[i] = c;
    }
#pragma unroll
    for (int j = 0; j < kMaskPxPerThread; ++j) {
        if (j >= n_out) break;
        uint8_t m = 0;
        if (x0 + j < roi_w) {
            const uint32_t s01 = col01[j] + col01[j + 1] + col01[j + 2], s2 = col2[j] + col2[j + 1] + col2[j + 2];
            const uint32_t b0 = (2 * (s01 & 0xffffu) + 9) / 18, b1 = (2 * (s01 >> 16) + 9) / 18, b2 = (2 * s2 + 9) / 18;
            m = ((b0 * 1868 + b1 * 9617 + b2 * 4899 + 8192) >> 14) > 127 ? 255 : 0;
        }
        out[j] = m;
    }
    }
}

}  // namespace
}  // namespace ssb

extern "C" int ssb_fg_mask_u8(const uint8_t *labels, int B, int H, int W, int roi_h, int roi_w,
                              const uint8_t *colormap, uint8_t *mask, void *stream) {
    using namespace ssb;
    if (B < 0 || H <= 0 || W <= 0) return fail(SSB_ERR_BAD_SHAPE, "bad shape B=%d H=%d W=%d", B, H, W);
    if (roi_h < 1 || roi_h > H || roi_w < 1 || roi_w > W)
        return fail(SSB_ERR_BAD_SHAPE, "ROI (%d,%d) outside frame (%d,%d)", roi_h, roi_w, H, W);
    if (B == 0) return SSB_OK;
    if (!labels || !colormap || !mask) return fail(SSB_ERR_BAD_POINTER, "NULL buffer");
    if (B > 65535 || (H + 7) / 8 > 65535) return fail(SSB_ERR_LIMIT, "B and H/8 must be <= 65535");
    dim3 grid((W + kMaskTileW - 1) / kMaskTileW, (H + kMaskTileH * kMaskTilesPerCta - 1) / (kMaskTileH * kMaskTilesPerCta), B);
    SSB_CUDA_TRY(launch_pdl(fg_mask_kernel, grid, dim3(256), 0, static_cast<cudaStream_t>(stream), labels, H, W, roi_h,
                            roi_w, colormap, mask));
    return SSB_OK;
}
