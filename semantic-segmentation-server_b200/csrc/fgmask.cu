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
// with rolling sums: every row's labels are loaded and looked up ONCE (4 per lane; the columns left and
// right of a lane's four come from its neighbours by shuffle), and the box sum of an output row is the sum
// of three rows' horizontal sums.
//
// Colours are summed as one 15-bit code per pixel when every channel of the label's colour is a multiple
// of 64 (true for the first 64 labels of the PASCAL colormap: values 0/64/128/192): code = c0/64 |
// c1/64 << 5 | c2/64 << 10, so a box sum keeps three 5-bit fields (<= 27), and TWO pixels share a 32-bit
// word (16 bits each): a lane's four pixels are two words, their horizontal 3-sums three funnel shifts and
// two 3-input adds.  The box sum is its own table index: t01[sum & 1023] holds the luma-weighted rounded
// means of channels 0 and 1, t2[sum >> 10] what channel 2 still has to add to reach the threshold.
// Labels with any other colour carry a flag; a warp that meets one recomputes its rows exactly from the
// colormap bytes afterwards (never taken for 21 VOC classes).
constexpr int kMaskPxPerThread = 4, kMaskStripW = 32 * kMaskPxPerThread;
constexpr int kMaskWarps = 4, kMaskRowsPerWarp = 18, kMaskTileH = kMaskWarps * kMaskRowsPerWarp;
constexpr int kMaskTailMax = 8;             // a last strip of at most this many columns is done a thread per row
constexpr uint32_t kMaskFlag = 1u << 31;
constexpr int kMaskNeed = (128 << 14) - 8192;   // gray > 127  <=>  weighted sum >= (128 << 14) - 8192

struct MaskArgs {
    const uint8_t *labels;
    const uint8_t *colormap;
    uint8_t *mask;
    int H, W, roi_h, roi_w;
    int n_strips;        // 128-column strips (blockIdx.x < n_strips); blockIdx.x == n_strips: the narrow last columns
    int tail_x;          // first column of the narrow last columns (W: none)
    int roi_row_tiles;   // blockIdx.y < roi_row_tiles: rows of the ROI; == roi_row_tiles: the rows below it
};

// One pixel by the plain formula (flagged box sums, the narrow last columns).
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

// Zeros over a byte range, 16 bytes per store where the address allows.
__device__ __forceinline__ void zero_bytes(uint8_t *p, long long n, int tid, int n_threads) {
    const long long head = min(n, static_cast<long long>((16 - (reinterpret_cast<uintptr_t>(p) & 15)) & 15));
    const long long body = (n - head) / 16;
    for (long long i = tid; i < head; i += n_threads) p[i] = 0;
    uint4 *q = reinterpret_cast<uint4 *>(p + head);
    for (long long i = tid; i < body; i += n_threads) q[i] = make_uint4(0u, 0u, 0u, 0u);
    for (long long i = head + body * 16 + tid; i < n; i += n_threads) p[i] = 0;
}

__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}

struct MaskTables {
    uint32_t code, t01, t2;             // shared-memory addresses; t01 is 4096-byte aligned, t2 128-byte aligned
};

// The rows [y0, y_end) of one warp's strip.  kInterior: all 128 columns of the strip are ROI columns (plain
// loads and stores); otherwise columns are clamped to the ROI for loading and zeros are stored right of it.
// Returns true if a label whose colour does not fit the packed code came up: the caller redoes the rows exactly.
template <bool kInterior>
__device__ __forceinline__ bool mask_rows(const uint8_t *frame, uint8_t *mframe, int W, int roi_h, int roi_w, int xs, int x_end,
                                          int y0, int y_end, int lane, const MaskTables tb) {
    const int x0 = xs + lane * kMaskPxPerThread;
    // The column left of lane 0 and right of lane 31 belongs to another strip: one more label for those two
    // lanes.  At the ROI's first and last column the missing neighbour is the reflected one, i.e. the pixel's
    // other neighbour.
    const bool need_left = lane == 0 && xs > 0, need_right = lane == 31 && xs + kMaskStripW < roi_w;
    const bool need_edge = need_left || need_right;
    const int x_edge = need_left ? x0 - 1 : need_right ? x0 + kMaskPxPerThread : 0;
    const bool first_col = lane == 0 && xs == 0;
    // the ROI's last column, relative to the strip: which halves of the lane that holds it take the reflected column
    const int x_last = roi_w - 1 - xs;
    const int last_j = (x_last < kMaskStripW && lane == (x_last >> 2)) ? (x_last & 3) : -1;
    const uint32_t fix01 = last_j == 0 ? 0xffff0000u : 0u;                               // column 1 := column -1
    const uint32_t fix23 = last_j == 1 ? 0x0000ffffu : last_j == 2 ? 0xffff0000u : 0u;   // column 2 := 0  /  column 3 := 1
    const uint32_t fix4 = last_j == 3 ? 0xffffffffu : 0u;                                // column 4 := column 2
    int xc[kMaskPxPerThread];
#pragma unroll
    for (int j = 0; j < kMaskPxPerThread; ++j) xc[j] = min(x0 + j, roi_w - 1);

    auto load_labels = [&](const uint8_t *row, uint32_t (&l)[kMaskPxPerThread], uint32_t &le) {
        if (kInterior) {
            const uint8_t *q = row + x0;
#pragma unroll
            for (int j = 0; j < kMaskPxPerThread; ++j) l[j] = q[j];
        } else {
#pragma unroll
            for (int j = 0; j < kMaskPxPerThread; ++j) l[j] = row[xc[j]];
        }
        if (need_edge) le = row[x_edge];
    };
    // horizontal 3-sums of one row: h01 = (columns 0, 1 of the lane's four), h23 = (columns 2, 3), 16 bits each;
    // returns the OR of the codes looked up (bit 31: a label whose colour does not fit the packed code)
    auto row_sums = [&](const uint32_t (&l)[kMaskPxPerThread], uint32_t le, uint32_t &h01, uint32_t &h23) -> uint32_t {
        const uint32_t c0 = lds_u32(tb.code + l[0] * 4u), c1 = lds_u32(tb.code + l[1] * 4u);
        const uint32_t c2 = lds_u32(tb.code + l[2] * 4u), c3 = lds_u32(tb.code + l[3] * 4u);
        uint32_t flags = c0 | c1 | c2 | c3;
        const uint32_t p01 = __byte_perm(c0, c1, 0x5410);
        uint32_t p23 = __byte_perm(c2, c3, 0x5410);
        uint32_t left = __shfl_up_sync(0xffffffffu, p23, 1);      // high half: the column left of my four
        uint32_t right = __shfl_down_sync(0xffffffffu, p01, 1);   // low half: the column right of my four
        if (need_edge) {
            const uint32_t ce = lds_u32(tb.code + le * 4u);
            flags |= ce;
            left = need_left ? ce << 16 : left;
            right = need_left ? right : ce & 0xffffu;
        }
        left = first_col ? p01 : left;                            // column -1 is column 1
        right = (right & ~fix4) | (p23 & fix4);                   // column roi_w is column roi_w - 2 (no-ops in every other lane)
        const uint32_t q01 = (p01 & ~fix01) | (left & fix01);
        p23 = (p23 & ~fix23) | (p01 & fix23);
        const uint32_t pm0 = __funnelshift_r(left, q01, 16);      // columns (-1, 0)
        const uint32_t p12 = __funnelshift_r(q01, p23, 16);       // columns (1, 2)
        const uint32_t p34 = __funnelshift_r(p23, right, 16);     // columns (3, 4)
        h01 = pm0 + q01 + p12;
        h23 = p12 + p23 + p34;
        return flags;
    };
    // the box sum is its own table index: bits 0..9 channels 0 and 1, bits 10..14 channel 2 (sh: 0 / 16 for the two halves)
    auto lit = [&](uint32_t s, int sh) -> uint32_t {
        const uint32_t v01 = lds_u32(tb.t01 | ((sh ? s >> 14 : s << 2) & 0xffcu));
        const uint32_t need = lds_u32(tb.t2 | ((s >> (8 + sh)) & 0x7cu));
        return static_cast<int>(v01) >= static_cast<int>(need) ? 255u : 0u;
    };
    // rows above / below the ROI are the reflected ones: row -1 is row 1, row roi_h is row roi_h - 2
    const long long step = roi_h > 1 ? W : 0;
    const uint8_t *rp = frame + static_cast<long long>(y0) * W;
    uint32_t hp01, hp23, hc01, hc23, hn01, hn23;
    uint32_t lab[kMaskPxPerThread], lab_e = 0;
    load_labels(y0 > 0 ? rp - step : rp + step, lab, lab_e);
    uint32_t flags = row_sums(lab, lab_e, hp01, hp23);
    load_labels(rp, lab, lab_e);
    flags |= row_sums(lab, lab_e, hc01, hc23);
    rp += y0 + 1 < roi_h ? step : -step;
    load_labels(rp, lab, lab_e);                                 // row y0 + 1, looked up in the first iteration
    uint8_t *orow = mframe + static_cast<long long>(y0) * W + x0;
#pragma unroll 3
    for (int y = y0; y < y_end; ++y, orow += W) {
        flags |= row_sums(lab, lab_e, hn01, hn23);
        if (y + 1 < y_end) {                                     // labels of row y + 2: they fly during this iteration
            rp += y + 2 < roi_h ? step : -step;
            load_labels(rp, lab, lab_e);
        }
        const uint32_t s01 = hp01 + hc01 + hn01, s23 = hp23 + hc23 + hn23;
        const uint32_t m0 = lit(s01, 0), m1 = lit(s01, 16), m2 = lit(s23, 0), m3 = lit(s23, 16);
        if (kInterior) {
            orow[0] = static_cast<uint8_t>(m0);
            orow[1] = static_cast<uint8_t>(m1);
            orow[2] = static_cast<uint8_t>(m2);
            orow[3] = static_cast<uint8_t>(m3);
        } else {
            if (x0 < x_end) orow[0] = x0 < roi_w ? static_cast<uint8_t>(m0) : 0;
            if (x0 + 1 < x_end) orow[1] = x0 + 1 < roi_w ? static_cast<uint8_t>(m1) : 0;
            if (x0 + 2 < x_end) orow[2] = x0 + 2 < roi_w ? static_cast<uint8_t>(m2) : 0;
            if (x0 + 3 < x_end) orow[3] = x0 + 3 < roi_w ? static_cast<uint8_t>(m3) : 0;
        }
        hp01 = hc01; hp23 = hc23; hc01 = hn01; hc23 = hn23;
    }
    return __any_sync(0xffffffffu, flags >> 31);
}

__global__ void __launch_bounds__(kMaskWarps * 32) fg_mask_kernel(const MaskArgs a) {
    // t01[s1 * 32 + s0]: luma weight x rounded mean of channels 0 and 1 (s = channel sum / 64); t2[s2]: what
    // channels 0 and 1 must reach given channel 2's sum.  A table address is base | offset (one LOP3), so t01
    // sits on a 4096-byte and t2 on a 128-byte boundary inside arrays of twice the size.
    __shared__ uint32_t t01_raw[2048];
    __shared__ uint32_t t2_raw[64];
    __shared__ uint32_t code[256];                   // packed colour / 64 per label, or kMaskFlag
    __shared__ uint32_t g01[2][32];
    pdl_wait();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int W = a.W, roi_h = a.roi_h, roi_w = a.roi_w;
    const long long frame_off = static_cast<long long>(blockIdx.z) * a.H * W;
    const uint8_t *frame = a.labels + frame_off;
    uint8_t *mframe = a.mask + frame_off;
    MaskTables tb;
    tb.code = smem_addr(code);
    tb.t01 = (smem_addr(t01_raw) + 4095u) & ~4095u;
    tb.t2 = (smem_addr(t2_raw) + 127u) & ~127u;
    uint32_t *t01 = t01_raw + ((tb.t01 - smem_addr(t01_raw)) >> 2), *t2 = t2_raw + ((tb.t2 - smem_addr(t2_raw)) >> 2);

    if (static_cast<int>(blockIdx.y) == a.roi_row_tiles) {      // the rows below the ROI are one byte range: zeros
        if (blockIdx.x == 0)
            zero_bytes(mframe + static_cast<long long>(roi_h) * W, static_cast<long long>(a.H - roi_h) * W, tid, kMaskWarps * 32);
        return;
    }
    const int y_tile = blockIdx.y * kMaskTileH;
    if (static_cast<int>(blockIdx.x) == a.n_strips) {           // the narrow last columns, a thread per row
        const int y = y_tile + tid;
        if (tid < kMaskTileH && y < roi_h)
            for (int x = a.tail_x; x < W; ++x)
                mframe[static_cast<long long>(y) * W + x] =
                    x < roi_w ? mask_pixel_exact(frame, W, roi_h, roi_w, x, y, a.colormap) : 0;
        return;
    }
    const int xs = blockIdx.x * kMaskStripW;
    const int x_end = min(xs + kMaskStripW, a.tail_x);          // tail_x == W when there are no narrow last columns
    if (xs >= roi_w) {                                          // a strip right of the ROI: zeros
        const int rows = min(kMaskTileH, roi_h - y_tile), cols = x_end - xs;
        for (int i = tid; i < rows * cols; i += kMaskWarps * 32) {
            const int r = i / cols;
            mframe[static_cast<long long>(y_tile + r) * W + xs + (i - r * cols)] = 0;
        }
        return;
    }

    for (int i = tid; i < 256; i += kMaskWarps * 32) {
        const uint32_t c0 = a.colormap[3 * i], c1 = a.colormap[3 * i + 1], c2 = a.colormap[3 * i + 2];
        code[i] = ((c0 | c1 | c2) & 63u) ? kMaskFlag : (c0 >> 6) | (c1 >> 6) << 5 | (c2 >> 6) << 10;
    }
    if (tid < 96) {
        const int c = tid >> 5, k = tid & 31;
        const uint32_t mean = (2u * 64u * k + 9u) / 18u;
        if (c == 0) g01[0][k] = 1868u * mean;
        else if (c == 1) g01[1][k] = 9617u * mean;
        else t2[k] = static_cast<uint32_t>(kMaskNeed) - 4899u * mean;
    }
    __syncthreads();
    for (int i = tid; i < 1024; i += kMaskWarps * 32) t01[i] = g01[0][i & 31] + g01[1][i >> 5];
    __syncthreads();

    const int y0 = y_tile + warp * kMaskRowsPerWarp;
    if (y0 >= roi_h) return;
    const int y_end = min(y0 + kMaskRowsPerWarp, roi_h);
    bool redo = roi_w < 2;                                      // a one-column ROI has no neighbours to shuffle
    if (!redo) {
        if (xs + kMaskStripW <= roi_w) redo = mask_rows<true>(frame, mframe, W, roi_h, roi_w, xs, x_end, y0, y_end, lane, tb);
        else redo = mask_rows<false>(frame, mframe, W, roi_h, roi_w, xs, x_end, y0, y_end, lane, tb);
    }
    if (redo) {                                                 // by the plain formula (never for the 21 VOC classes)
        const int x0 = xs + lane * kMaskPxPerThread;
        for (int y = y0; y < y_end; ++y)
            for (int j = 0; j < kMaskPxPerThread; ++j) {
                const int x = x0 + j;
                if (x < x_end)
                    mframe[static_cast<long long>(y) * W + x] =
                        x < roi_w ? mask_pixel_exact(frame, W, roi_h, roi_w, x, y, a.colormap) : 0;
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
    if (B > 65535 || (H + kMaskTileH - 1) / kMaskTileH > 65534) return fail(SSB_ERR_LIMIT, "B and H/72 must be <= 65534");
    MaskArgs a{};
    a.labels = labels; a.colormap = colormap; a.mask = mask;
    a.H = H; a.W = W; a.roi_h = roi_h; a.roi_w = roi_w;
    const int rem = W % kMaskStripW;
    a.tail_x = (rem > 0 && rem <= kMaskTailMax) ? W - rem : W;
    a.n_strips = (a.tail_x + kMaskStripW - 1) / kMaskStripW;
    a.roi_row_tiles = (roi_h + kMaskTileH - 1) / kMaskTileH;
    dim3 grid(a.n_strips + (a.tail_x < W ? 1 : 0), a.roi_row_tiles + (roi_h < H ? 1 : 0), B);
    SSB_CUDA_TRY(launch_pdl(fg_mask_kernel, grid, dim3(kMaskWarps * 32), 0, static_cast<cudaStream_t>(stream), a));
    return SSB_OK;
}
