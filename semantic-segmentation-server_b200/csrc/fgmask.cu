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

__global__ void __launch_bounds__(256)
fg_mask_kernel(const uint8_t *__restrict__ labels, int H, int W, int roi_h, int roi_w,
               const uint8_t *__restrict__ colormap, uint8_t *__restrict__ mask) {
    __shared__ uint32_t cmap[256];  // packed c0 | c1<<8 | c2<<16
    for (int i = threadIdx.x; i < 256; i += blockDim.x)
        cmap[i] = colormap[3 * i] | (colormap[3 * i + 1] << 8) | (colormap[3 * i + 2] << 16);
    __syncthreads();
    const int x = blockIdx.x * 32 + (threadIdx.x & 31);
    const int y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= W || y >= H) return;
    const long long fbase = static_cast<long long>(blockIdx.z) * H * W;
    uint8_t m = 0;
    if (x < roi_w && y < roi_h) {
        int s0 = 0, s1 = 0, s2 = 0;
#pragma unroll
        for (int dy = -1; dy <= 1; ++dy) {
            const uint8_t *row = labels + fbase + static_cast<long long>(reflect101(y + dy, roi_h)) * W;
#pragma unroll
            for (int dx = -1; dx <= 1; ++dx) {
                const uint32_t c = cmap[row[reflect101(x + dx, roi_w)]];
                s0 += c & 0xff; s1 += (c >> 8) & 0xff; s2 += (c >> 16) & 0xff;
            }
        }
        const int b0 = (2 * s0 + 9) / 18, b1 = (2 * s1 + 9) / 18, b2 = (2 * s2 + 9) / 18;
        const int gray = (b0 * 1868 + b1 * 9617 + b2 * 4899 + 8192) >> 14;
        m = gray > 127 ? 255 : 0;
    }
    mask[fbase + static_cast<long long>(y) * W + x] = m;
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
    dim3 grid((W + 31) / 32, (H + 7) / 8, B);
    fg_mask_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(labels, H, W, roi_h, roi_w, colormap, mask);
    SSB_CUDA_TRY(cudaGetLastError());
    return SSB_OK;
}
