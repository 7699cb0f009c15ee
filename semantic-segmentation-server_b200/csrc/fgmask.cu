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

// A warp owns a strip of 128 columns (4 per lane) and kMaskRowsPerWarp consecutive rows and walks down it
// with rolling sums: every row's labels are loaded and looked up ONCE (6 per lane: its 4 columns and one
// on either side), added into 4 horizontal 3-sums, and the box sum of an output row is the sum of three
// rows' horizontal sums.
//
// Colours are summed as one packed word per pixel when every channel of the label's colour is a multiple
// of 64 (true for the first 64 labels of the PASCAL colormap: values 0/64/128/192): code = c0/64 |
// c1/64 << 8 | c2/64 << 16, so a box sum is nine adds of one word (fields <= 27) and the per-channel
// rounded mean times its luma weight comes from three 28-entry tables (28 words sit in 28 different
// banks: no conflicts whatever the values).  Labels with any other colour carry a flag in bit 24; a box
// sum with a flag set is recomputed exactly from the colormap bytes (never taken for 21 VOC classes).
constexpr int kMaskPxPerThread = 4, kMaskTileW = 32 * kMaskPxPerThread;
constexpr int kMaskWarps = 8, kMaskRowsPerWarp = 8, kMaskTileH = kMaskWarps * kMaskRowsPerWarp;
constexpr uint32_t kMaskFlag = 1u << 24;
constexpr int kMaskThreshold = 128 << 14;   // gray > 127  <=>  weighted sum + 8192 >= 128 << 14

// One pixel by the plain formula (flagged box sums only).
__device__ __noinline__ uint8_t mask_pixel_exact(const uint8_t *frame, int W, int roi_h, int roi_w, int x, int y,
                                                 const uint8_t *__restrict__ colormap) {
    int s[3] = {0, 0, 0};
    for (int dy = -1; dy <= 1; ++dy)
        for (int dx = -1; dx <= 1; ++dx) {
            const int l = frame[static_cast<long long>(reflect101(y + dy, roi_h)) * W + reflect101(x + dx, roi_w)];
            for (int c = 0; c < 3; ++c) s[c] += colormap[3 * l + c];
        }
    const int b0 = (2 * s[0] + 9) / 18, b1 = (2 * s[1] + 9) / 18, b2 = (2 * s[2] + 9) / 18;
    return ((b0 * 1868 + b1 * 9617 + b2 * 4899 + 8192) >> 14) > 127 ? 255 : 0;
}

__global__ void __launch_bounds__(kMaskWarps * 32)
fg_mask_kernel(const uint8_t *__restrict__ labels, int H, int W, int roi_h, int roi_w,
               const uint8_t *__restrict__ colormap, uint8_t *__restrict__ mask) {
    __shared__ uint32_t code[256];      // packed colour / 64 per label, or kMaskFlag
    // luma weight x rounded mean, by channel sum / 64; 512-byte aligned so that a table address is
    // base | offset (one LOP3) instead of base + offset
    __shared__ __align__(512) uint32_t lum[3][32];
    pdl_wait();
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
        const uint32_t c0 = colormap[3 * i], c1 = colormap[3 * i + 1], c2 = colormap[3 * i + 2];
        code[i] = ((c0 | c1 | c2) & 63u) ? kMaskFlag : (c0 >> 6) | (c1 >> 6) << 8 | (c2 >> 6) << 16;
    }
    if (threadIdx.x < 96) {
        const int c = threadIdx.x >> 5, k = threadIdx.x & 31;
        const uint32_t w = c == 0 ? 1868u : (c == 1 ? 9617u : 4899u);
        lum[c][k] = w * ((2u * 64u * k + 9u) / 18u);
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int x0 = blockIdx.x * kMaskTileW + lane * kMaskPxPerThread;
    const int y0 = blockIdx.y * kMaskTileH + warp * kMaskRowsPerWarp;
    if (x0 >= W || y0 >= H) return;
    const int n_out = min(kMaskPxPerThread, W - x0);
    const uint8_t *frame = labels + static_cast<long long>(blockIdx.z) * H * W;
    uint8_t *out = mask + static_cast<long long>(blockIdx.z) * H * W + x0;
    const int y_end = min(y0 + kMaskRowsPerWarp, H);
    if (x0 >= roi_w || y0 >= roi_h) {
        for (int y = y0; y < y_end; ++y)
            for (int j = 0; j < n_out; ++j) out[static_cast<long long>(y) * W + j] = 0;
        return;
    }
    // Lanes whose six taps all lie inside the ROI read them through one pointer with immediate offsets;
    // the others (first / last columns) go through reflected column indices.  Taps right of the ROI are
    // never used by a counted pixel; the clamp keeps the load in range.
    const bool simple = x0 >= 1 && x0 + kMaskPxPerThread < roi_w;
    int xc[kMaskPxPerThread + 2];
#pragma unroll
    for (int i = 0; i < kMaskPxPerThread + 2; ++i) xc[i] = reflect101(min(x0 - 1 + i, roi_w), roi_w);
    const uint32_t code_base = smem_addr(code);
    auto lut = [&](uint32_t label) {
        uint32_t v;
        asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(code_base + label * 4u));
        return v;
    };
    // A row's six labels are loaded one iteration before they are looked up, so the global loads of row
    // y+2 fly while row y's pixels are computed.
    auto load_labels = [&](const uint8_t *row, uint32_t (&l)[kMaskPxPerThread + 2]) {
        if (simple) {
            const uint8_t *q = row + (x0 - 1);
#pragma unroll
            for (int i = 0; i < kMaskPxPerThread + 2; ++i) l[i] = q[i];
        } else {
#pragma unroll
            for (int i = 0; i < kMaskPxPerThread + 2; ++i) l[i] = row[xc[i]];
        }
    };
    // horizontal 3-sums of one row
    auto row_sums = [&](const uint32_t (&l)[kMaskPxPerThread + 2], uint32_t (&h)[kMaskPxPerThread]) {
        uint32_t c[kMaskPxPerThread + 2];
#pragma unroll
        for (int i = 0; i < kMaskPxPerThread + 2; ++i) c[i] = lut(l[i]);
#pragma unroll
        for (int j = 0; j < kMaskPxPerThread; ++j) h[j] = c[j] + c[j + 1] + c[j + 2];
    };
    auto row_ptr = [&](int y) { return frame + static_cast<long long>(reflect101(y, roi_h)) * W; };
    const uint32_t lum_base = smem_addr(lum);
    uint32_t prev[kMaskPxPerThread], cur[kMaskPxPerThread], next[kMaskPxPerThread];
    uint32_t lab_a[kMaskPxPerThread + 2], lab_b[kMaskPxPerThread + 2];
    load_labels(row_ptr(y0 - 1), lab_a);
    const uint8_t *rp = row_ptr(y0);
    load_labels(rp, lab_b);
    row_sums(lab_a, prev);
    row_sums(lab_b, cur);
    rp = y0 + 1 < roi_h ? rp + W : row_ptr(y0 + 1);
    load_labels(rp, lab_a);                     // row y0+1, looked up in the first iteration
    uint8_t *orow = out + static_cast<long long>(y0) * W;
    for (int y = y0; y < y_end; ++y, orow += W) {
        if (y >= roi_h) {                       // below the ROI: zeros
            for (int j = 0; j < n_out; ++j) orow[j] = 0;
            continue;
        }
        row_sums(lab_a, next);
        if (y + 1 < y_end && y + 1 < roi_h) {   // labels of row y+2 for the next iteration
            rp = y + 2 < roi_h ? rp + W : row_ptr(y + 2);   // reflection only at the ROI's last row
            load_labels(rp, lab_a);
        }
        uint32_t sum[kMaskPxPerThread], m[kMaskPxPerThread];
#pragma unroll
        for (int j = 0; j < kMaskPxPerThread; ++j) sum[j] = prev[j] + cur[j] + next[j];
        if ((sum[0] | sum[1] | sum[2] | sum[3]) >> 24) {
#pragma unroll 1
            for (int j = 0; j < kMaskPxPerThread; ++j)
                m[j] = x0 + j < roi_w ? mask_pixel_exact(frame, W, roi_h, roi_w, x0 + j, y, colormap) : 0;
        } else {
#pragma unroll
            for (int j = 0; j < kMaskPxPerThread; ++j) {
                uint32_t g0, g1, g2;
                // channel sums are <= 27: offsets stay inside one 128-byte table
                asm("ld.shared.u32 %0, [%1];" : "=r"(g0) : "r"(lum_base | ((sum[j] << 2) & 0x7cu)));
                asm("ld.shared.u32 %0, [%1+128];" : "=r"(g1) : "r"(lum_base | ((sum[j] >> 6) & 0x7cu)));
                asm("ld.shared.u32 %0, [%1+256];" : "=r"(g2) : "r"(lum_base | ((sum[j] >> 14) & 0x7cu)));
                m[j] = g0 + g1 + g2 >= static_cast<uint32_t>(kMaskThreshold - 8192) ? 255u : 0u;
            }
        }
        if (simple) {
#pragma unroll
            for (int j = 0; j < kMaskPxPerThread; ++j) orow[j] = static_cast<uint8_t>(m[j]);
        } else {
#pragma unroll
            for (int j = 0; j < kMaskPxPerThread; ++j)
                if (j < n_out) orow[j] = x0 + j < roi_w ? static_cast<uint8_t>(m[j]) : 0;
        }
#pragma unroll
        for (int j = 0; j < kMaskPxPerThread; ++j) {
            prev[j] = cur[j];
            cur[j] = next[j];
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
    if (B > 65535 || (H + kMaskTileH - 1) / kMaskTileH > 65535) return fail(SSB_ERR_LIMIT, "B and H/64 must be <= 65535");
    dim3 grid((W + kMaskTileW - 1) / kMaskTileW, (H + kMaskTileH - 1) / kMaskTileH, B);
    SSB_CUDA_TRY(launch_pdl(fg_mask_kernel, grid, dim3(kMaskWarps * 32), 0, static_cast<cudaStream_t>(stream), labels, H, W, roi_h,
                            roi_w, colormap, mask));
    return SSB_OK;
}
