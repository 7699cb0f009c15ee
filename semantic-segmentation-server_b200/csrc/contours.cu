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
#include <cstring>
#include <mutex>

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


// ---- the same walk on bit rows in shared memory: a CTA per frame, a THREAD per table row -------------------
// The warp-per-row walk above spends ~240 cycles per step in one dependent chain (address, byte load, ballot,
// rotate, find-first-set, step).  Here the ROI's mask first becomes bit rows in shared memory (a thread per
// 32-pixel word: nine aligned 32-bit loads, the top bit of every non-zero byte, four bits gathered by one
// multiply; pixel x of row y is bit (x + 32) of bit row (y + 1): one zero guard word left and right, one zero
// guard row above and below, so no bounds tests in the walk); then a thread follows the border of one table row:
// the 3 x 3 neighbourhood is three funnel shifts, its direction byte one table look-up.  32 borders per warp run
// in lock step, and a frame's longest border is walked at ~2x the speed.
constexpr int kBitsThreads = 1024;
#ifndef SSB_MOMENTS_RUNS
#define SSB_MOMENTS_RUNS 1   // 1: straight horizontal and vertical stretches of a border are walked in one step
#endif

__global__ void __launch_bounds__(kBitsThreads) contour_moments_bits_kernel(const MomentParams p) {
    extern __shared__ __align__(16) unsigned int bits_smem[];
    pdl_wait();
    const int f = blockIdx.x, tid = threadIdx.x;
    const int S = (p.roi_w + 31) >> 5, SW = S + 2;
    unsigned int *bits = bits_smem;                                   // [(roi_h + 2)][SW]
    // the transposed plane: pixel y of column x is bit (y + 32) of bit column (x + 1), guards as above
    const int T = (p.roi_h + 31) >> 5, TW = T + 2;
    unsigned int *bits_t = bits + static_cast<size_t>(p.roi_h + 2) * SW + 4;   // [(roi_w + 2)][TW], behind 4 spare zero words
    uint8_t *nbtab = reinterpret_cast<uint8_t *>(bits_t + static_cast<size_t>(p.roi_w + 2) * TW + 4);   // [512]
    const int n_rows = min(p.count[f], p.max_regions);
    if (n_rows <= 0) return;                                         // CTA-uniform
    const uint8_t *m = p.mask + static_cast<long long>(f) * p.H * p.W;
    // guard rows and guard words
    for (int i = tid; i < SW + 4; i += kBitsThreads) { if (i < SW) bits[i] = 0u; bits[(p.roi_h + 1) * SW + i] = 0u; }
    for (int y = tid; y < p.roi_h; y += kBitsThreads) { bits[(y + 1) * SW] = 0u; bits[(y + 1) * SW + S + 1] = 0u; }
    // direction byte of a 3 x 3 neighbourhood given as (up | mid << 3 | down << 6), 3 bits each = columns x-1, x, x+1:
    // bit d = the neighbour in direction d (0 = E, 1 = NE, 2 = N, 3 = NW, 4 = W, 5 = SW, 6 = S, 7 = SE; y grows downwards)
    for (int i = tid; i < 512; i += kBitsThreads) {
        const unsigned int up = i & 7, mid = (i >> 3) & 7, dn = i >> 6;
        nbtab[i] = static_cast<uint8_t>(((mid >> 2) & 1u) | ((up >> 2) & 1u) << 1 | ((up >> 1) & 1u) << 2 | (up & 1u) << 3 |
                                        (mid & 1u) << 4 | (dn & 1u) << 5 | ((dn >> 1) & 1u) << 6 | ((dn >> 2) & 1u) << 7);
    }
    for (int t = tid; t < p.roi_h * S; t += kBitsThreads) {
        const int y = t / S, seg = t - y * S;
        const int n_px = min(32, p.roi_w - seg * 32);
        const uint8_t *src = m + static_cast<long long>(y) * p.W + seg * 32;
        const unsigned int off = static_cast<unsigned int>(reinterpret_cast<uintptr_t>(src) & 3u);
        const unsigned int *w32 = reinterpret_cast<const unsigned int *>(src - off);
        const int n_words = (static_cast<int>(off) + n_px + 3) >> 2;   // at least one byte of every word loaded is a live pixel
        unsigned int v[9];
#pragma unroll
        for (int j = 0; j < 9; ++j) v[j] = j < n_words ? w32[j] : 0u;
        unsigned int word = 0u;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const unsigned int px4 = __funnelshift_r(v[j], v[j + 1], 8 * off);
            const unsigned int top = (((px4 & 0x7f7f7f7fu) + 0x7f7f7f7fu) | px4) & 0x80808080u;
            word |= (((top >> 7) * 0x01020408u) >> 24) << (4 * j);
        }
        bits[(y + 1) * SW + seg + 1] = n_px == 32 ? word : word & ((1u << n_px) - 1u);
    }
    __syncthreads();
    // bit columns: a warp transposes a block of 32 rows x 32 columns in registers, lane j holding row j's word before
    // and column j's word after (guards and ragged edges stay zero); 32 ballots per block took ~4 us more per launch
    for (int i = tid; i < (p.roi_w + 2) * TW + 4; i += kBitsThreads) bits_t[i] = 0u;
    __syncthreads();
    if (SSB_MOMENTS_RUNS) {
        const int lane = tid & 31;
        for (int blk = tid >> 5; blk < T * S; blk += kBitsThreads / 32) {
            const int k = blk / S, w = blk - k * S;
            const int y = k * 32 + lane;
            const unsigned int word = y < p.roi_h ? bits[(y + 1) * SW + w + 1] : 0u;
            // five butterfly stages: stage j swaps the off-diagonal j x j sub-blocks with lane ^ j
            unsigned int mine = word;
#pragma unroll
            for (int st = 0; st < 5; ++st) {
                const int j = 16 >> st;
                const unsigned int m = st == 0 ? 0x0000ffffu : st == 1 ? 0x00ff00ffu : st == 2 ? 0x0f0f0f0fu
                                       : st == 3 ? 0x33333333u : 0x55555555u;
                const unsigned int t = __shfl_xor_sync(0xffffffffu, mine, j);
                mine = (lane & j) ? ((t >> j) & m) | (mine & ~m) : (mine & m) | ((t << j) & ~m);
            }
            const int x = w * 32 + lane;
            if (x < p.roi_w) bits_t[(x + 1) * TW + k + 1] = mine;
        }
    }
    __syncthreads();

    // direction byte of pixel (x, y): bits x-1 .. x+1 of the three bit rows
    auto neighbours = [&](int x, int y) -> unsigned int {
        const unsigned int *r = bits + y * SW + ((x + 31) >> 5);      // bit row y - 1, the word that holds bit (x - 1 + 32)
        const int sh = (x + 31) & 31;
        const unsigned int up = __funnelshift_r(r[0], r[1], sh) & 7u;
        const unsigned int mid = __funnelshift_r(r[SW], r[SW + 1], sh) & 7u;
        const unsigned int dn = __funnelshift_r(r[2 * SW], r[2 * SW + 1], sh) & 7u;
        return nbtab[up | (mid & 5u) << 3 | dn << 6];                 // the centre pixel itself is not a neighbour
    };
    auto step_x = [](int d) { return ((0x01u | 0x02u | 0x80u) >> d & 1u) ? 1 : ((0x08u | 0x10u | 0x20u) >> d & 1u) ? -1 : 0; };
    auto step_y = [](int d) { return ((0x02u | 0x04u | 0x08u) >> d & 1u) ? -1 : ((0x20u | 0x40u | 0x80u) >> d & 1u) ? 1 : 0; };
    for (int row = tid; row < n_rows; row += kBitsThreads) {
        const long long tab = static_cast<long long>(f) * p.max_regions + row;
        const int fl = p.flags[tab] & 3;
        long long a00 = 0, a10 = 0, a01 = 0;
        bool unknown = false;
        if (fl != 2) {                                       // not the outside
            const bool is_hole = fl == 0;
            const int fp = p.first[tab];
            const int y0 = fp / p.roi_w, x0 = fp - y0 * p.roi_w - (is_hole ? 1 : 0);
            unsigned int nb = x0 >= 0 ? neighbours(x0, y0) : 0u;
            if (nb) {                                        // (an isolated pixel is a border of one point: all sums zero)
                const int s_end = is_hole ? 0 : 4;
                int s = s_end;
                do {
                    s = (s - 1) & 7;
                } while (!((nb >> s) & 1u));
                const int x1 = x0 + step_x(s), y1 = y0 + step_y(s);
                int x = x0, y = y0, px = x0, py = y0, steps = 0, travel = -1;
                bool first = true;
                for (;;) {
                    const int from = (s + 1) & 7;
                    const unsigned int rot = ((nb >> from) | (nb << (8 - from))) & 0xffu;
                    s = (from + __ffs(rot) - 1) & 7;
                    const int nx = x + step_x(s), ny = y + step_y(s);
                    if (!first) {
                        const long long d = static_cast<long long>(px * y - x * py);     // |coordinates| < 2^15: fits 32 bits
                        a00 += d; a10 += d * (px + x); a01 += d * (py + y);
                    }
                    first = false;
                    px = x; py = y;
                    if (nx == x0 && ny == y0 && x == x1 && y == y1) break;
                    if (++steps > p.max_steps) { unknown = true; break; }
                    x = nx; y = ny;
                    // A straight horizontal or vertical stretch in one go.  The walk goes on eastwards from a pixel q
                    // iff SW, S, SE of q are background and E is foreground (westwards: NE, N, NW and W; southwards:
                    // NW, W, SW and S; northwards: SE, E, NE and N): for the pixels ahead these are bit tests on windows
                    // of two lines of a bit plane -- two bit rows, or two bit columns of the transposed plane -- and
                    // the length of the stretch is a count of trailing ones.  The points in between are collinear, so
                    // the sums see one long edge (same integers); the stretch stops at (x1, y1) at the latest, where the
                    // walk may end.
                    if (SSB_MOMENTS_RUNS && s == travel && !(s & 1)) {
                        const bool horizontal = !(s & 2), forward = s == 0 || s == 6;   // east and south count upwards
                        const int pos = horizontal ? x : y, line = horizontal ? y : x;
                        const int side = (s == 0 || s == 2) ? line + 1 : line - 1;      // the line that must stay background
                        const int LW = horizontal ? SW : TW;
                        const unsigned int *plane = horizontal ? bits : bits_t;
                        const unsigned int *lm = plane + (line + 1) * LW, *ls = plane + (side + 1) * LW;
                        // forward: windows start at pos + 1 (this line) and pos - 1 (the side line);
                        // backward: they end at pos - 1 and pos + 1 and are read in reverse
                        const int pm = (forward ? pos + 1 : pos - 32) + 32, ps = (forward ? pos - 1 : pos - 30) + 32;
                        unsigned int mw = __funnelshift_r(lm[pm >> 5], lm[(pm >> 5) + 1], pm & 31);
                        unsigned int sw = __funnelshift_r(ls[ps >> 5], ls[(ps >> 5) + 1], ps & 31);
                        if (!forward) { mw = __brev(mw); sw = __brev(sw); }
                        const unsigned int go = mw & ~(sw | sw >> 1 | sw >> 2) & 0x3fffffffu;
                        int n = __ffs(~go) - 1;
                        const int end1 = horizontal ? x1 : y1;                           // (x1, y1), if it lies on this line
                        if ((horizontal ? y == y1 : x == x1)) {
                            if (forward) { if (pos <= end1 && end1 <= pos + n) n = end1 - pos; }
                            else if (end1 <= pos && pos - n <= end1) n = pos - end1;
                        }
                        const int moved = forward ? n : -n;
                        if (horizontal) x += moved; else y += moved;
                        steps += n;
                    }
                    travel = s;
                    s = (s + 4) & 7;
                    nb = neighbours(x, y);
                }
                const long long d = static_cast<long long>(px * y0 - x0 * py);
                a00 += d; a10 += d * (px + x0); a01 += d * (py + y0);
            }
        }
        if (unknown) { a00 = 0; a10 = kMomentsUnknown; a01 = 0; }
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
    // bit rows in shared memory and a thread per row where the ROI fits (and its coordinates 15 bits); else a warp per row
    const size_t smem = sizeof(unsigned int) * (static_cast<size_t>(roi_h + 2) * ((roi_w + 31) / 32 + 2) + 4 +
                                                static_cast<size_t>(roi_w + 2) * ((roi_h + 31) / 32 + 2) + 4) + 512;
    const char *path = getenv("SSB_MOMENTS_PATH");            // tuning / test knob: "warp" forces the warp-per-row kernel
    const bool want_bits = !(path && !strcmp(path, "warp")) && roi_h < 32768 && roi_w < 32768;
    DeviceInfo di;
    if (int rc = device_info(&di)) return rc;
    if (want_bits && smem <= static_cast<size_t>(di.max_smem_optin)) {
        static std::mutex mu;
        static size_t smem_set[64] = {0};
        int dev = 0;
        SSB_CUDA_TRY(cudaGetDevice(&dev));
        if (dev < 0 || dev >= 64) return fail(SSB_ERR_LIMIT, "device ordinal %d not supported", dev);
        {
            std::lock_guard<std::mutex> lock(mu);
            if (smem_set[dev] < smem) {
                SSB_CUDA_TRY(cudaFuncSetAttribute(contour_moments_bits_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                  static_cast<int>(smem)));
                smem_set[dev] = smem;
            }
        }
        SSB_CUDA_TRY(launch_pdl(contour_moments_bits_kernel, dim3(B), dim3(kBitsThreads), smem, static_cast<cudaStream_t>(stream), p));
        return SSB_OK;
    }
    const dim3 grid((max_regions + 3) / 4, B), block(128);   // a warp per row
    SSB_CUDA_TRY(launch_pdl(contour_moments_kernel, grid, block, 0, static_cast<cudaStream_t>(stream), p));
    return SSB_OK;
}
