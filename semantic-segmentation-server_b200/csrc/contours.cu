// K6: contour integrals of the regions of K5's table, on the GPU.
//
// The reference takes cv2.contourArea and cv2.moments of every contour cv2.findContours returns
// (sem_seg_server.py:88, :93, :115-122).  Those are closed-polygon sums over the border OpenCV follows,
//     a00 = sum (x' y - x y'),  a10 = sum (x' y - x y') (x' + x),  a01 = sum (x' y - x y') (y' + y)
// over consecutive border points (x', y') -> (x, y), with area = |a00| / 2, m00 = a00 / 2, m10 = a10 / 6, m01 = a01 / 6:
// sums of products of integer coordinates, exact in 64-bit integers, the same whichever collinear points
// CHAIN_APPROX_SIMPLE drops and wherever on the border the list starts.  The border itself is the one Suzuki & Abe's
// border following walks, and WHICH pixels it walks does not depend on the marks that algorithm leaves (they only
// decide where the raster scan discovers a border) -- so a thread can follow the border of its region on its own:
//   * a foreground region: its outer border, from the region's first pixel (raster order), where OpenCV starts it;
//   * a hole (background region that does not touch the ROI border): the hole border, from the foreground pixel left
//     of the hole's first pixel (a pixel of that border; OpenCV may start it elsewhere on the same cycle).
// A warp per table row; the walk is serial, rows and frames are parallel.
// The host restatement with marks (csrc/hostpath.cu, ssb_contours_host) is the reference these sums are tested
// against, and the fallback where the ORDER of sibling hole contours matters (see hotpath.candidates_from_moments).
#include <cstdlib>

#include "ssb_common.cuh"

namespace ssb {
namespace {

constexpr long long kMomentsUnknown = static_cast<long long>(0x8000000000000000ull);   // a10 of a row whose border was not walked to the end
constexpr int kDefaultMaxSteps = 32768;

struct MomentParams {
    const uint8_t *mask;      // [B][H][W]
    int B, H, W, roi_h, roi_w, max_regions;
    const int *count, *first, *flags;
    long long *moments;       // [B][max_regions][3]
    int max_steps;            // a border longer than this is left to the host (row marked "unknown")
};

// A warp per table row: eight lanes look at the eight neighbours of the current border pixel at once (one load
// latency per step instead of up to eight in a row), a ballot gives the neighbourhood as a byte, and the next
// direction is the first set bit counter-clockwise from the one after the direction the walk came from.
__global__ void __launch_bounds__(128) contour_moments_kernel(const MomentParams p) {
    pdl_wait();
    const int f = blockIdx.y, lane = threadIdx.x & 31, row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= p.max_regions || row >= p.count[f]) return;                 // warp-uniform
    const long long tab = static_cast<long long>(f) * p.max_regions + row;
    const int fl = p.flags[tab] & 3;
    long long a00 = 0, a10 = 0, a01 = 0;
    bool unknown = false;
    if (fl != 2) {                                       // not the outside
        const bool is_hole = fl == 0;
        const int fp = p.first[tab];
        const int y0 = fp / p.roi_w, x0 = fp - y0 * p.roi_w - (is_hole ? 1 : 0);
        const uint8_t *m = p.mask + static_cast<long long>(f) * p.H * p.W;
        const int dxs[8] = {1, 1, 0, -1, -1, -1, 0, 1}, dys[8] = {0, -1, -1, -1, 0, 1, 1, 1};
        const int my_dx = dxs[lane & 7], my_dy = dys[lane & 7];
        // bit d of the result: the neighbour of (x, y) in direction d is foreground (inside the ROI)
        auto neighbours = [&](int x, int y) -> unsigned int {
            const int nx = x + my_dx, ny = y + my_dy;
            const bool on = lane < 8 && nx >= 0 && ny >= 0 && nx < p.roi_w && ny < p.roi_h && m[static_cast<long long>(ny) * p.W + nx] != 0;
            return __ballot_sync(0xffffffffu, on) & 0xffu;
        };
        auto step_x = [&](int d) { return ((0x01u | 0x02u | 0x80u) >> d & 1u) ? 1 : ((0x08u | 0x10u | 0x20u) >> d & 1u) ? -1 : 0; };
        auto step_y = [&](int d) { return ((0x02u | 0x04u | 0x08u) >> d & 1u) ? -1 : ((0x20u | 0x40u | 0x80u) >> d & 1u) ? 1 : 0; };
        unsigned int nb = x0 >= 0 ? neighbours(x0, y0) : 0u;
        if (nb) {                                        // (an isolated pixel is a border of one point: all sums zero)
            // the first neighbour clockwise from the start direction (east for a hole border, west for an outer one)
            const int s_end = is_hole ? 0 : 4;
            int s = s_end;
            do {
                s = (s - 1) & 7;
            } while (!((nb >> s) & 1u));
            const int x1 = x0 + step_x(s), y1 = y0 + step_y(s);
            int x = x0, y = y0, px = x0, py = y0, steps = 0;
            bool first = true;
            for (;;) {
                // first foreground neighbour counter-clockwise from direction s + 1
                const int from = (s + 1) & 7;
                const unsigned int rot = ((nb >> from) | (nb << (8 - from))) & 0xffu;
                s = (from + __ffs(rot) - 1) & 7;
                const int nx = x + step_x(s), ny = y + step_y(s);
                if (!first) {
                    const long long d = static_cast<long long>(px) * y - static_cast<long long>(x) * py;
                    a00 += d; a10 += d * (px + x); a01 += d * (py + y);
                }
                first = false;
                px = x; py = y;
                if (nx == x0 && ny == y0 && x == x1 && y == y1) break;
                if (++steps > p.max_steps) { unknown = true; break; }       // bounded GPU time: the host walks this one
                x = nx; y = ny;
                s = (s + 4) & 7;
                nb = neighbours(x, y);
            }
            const long long d = static_cast<long long>(px) * y0 - static_cast<long long>(x0) * py;   // back to the first point
            a00 += d; a10 += d * (px + x0); a01 += d * (py + y0);
        }
    }
    if (unknown) { a00 = 0; a10 = kMomentsUnknown; a01 = 0; }
    if (lane == 0) {
        p.moments[tab * 3 + 0] = a00;
        p.moments[tab * 3 + 1] = a10;
        p.moments[tab * 3 + 2] = a01;
    }
}

}  // namespace
}  // namespace ssb

extern "C" int ssb_contour_moments_u8(const uint8_t *mask, int B, int H, int W, int roi_h, int roi_w, int max_regions,
                                      const int32_t *count, const int32_t *first, const int32_t *flags, int64_t *moments,
                                      void *stream) {
    using namespace ssb;
    if (B < 0 || H <= 0 || W <= 0 || roi_h < 1 || roi_h > H || roi_w < 1 || roi_w > W || max_regions < 1)
        return fail(SSB_ERR_BAD_SHAPE, "bad shape B=%d H=%d W=%d ROI (%d,%d) max_regions=%d", B, H, W, roi_h, roi_w, max_regions);
    if (B == 0) return SSB_OK;
    if (B > 65535) return fail(SSB_ERR_LIMIT, "B must be <= 65535");
    if (!mask || !count || !first || !flags || !moments) return fail(SSB_ERR_BAD_POINTER, "NULL buffer");
    int max_steps = kDefaultMaxSteps;
    if (const char *env = getenv("SSB_MOMENTS_MAX_STEPS")) max_steps = atoi(env) > 0 ? atoi(env) : max_steps;   // test knob
    MomentParams p{mask, B, H, W, roi_h, roi_w, max_regions, count, first, flags, reinterpret_cast<long long *>(moments), max_steps};
    const dim3 grid((max_regions + 3) / 4, B), block(128);   // a warp per row
    SSB_CUDA_TRY(launch_pdl(contour_moments_kernel, grid, block, 0, static_cast<cudaStream_t>(stream), p));
    return SSB_OK;
}
