// Connected regions of the foreground mask and their label histograms, so that the reference's
// per-contour majority label and purity (sem_seg_server.py:98-111: filled contour mask -> labels under
// it -> bincount -> argmax, count/len) come from a table instead of a polygon fill per contour
// (stage 2).  Polygon area and moments (:93, :115) stay with OpenCV on the host.
//
// cv2.findContours(RETR_TREE) follows 8-connected foreground components and 4-connected background
// components (the image border counts as background).  The pixels cv2.drawContours fills for the OUTER
// contour of a foreground component F are exactly those that cannot be reached from the border through
// 4-connected steps that avoid F: F itself, its holes, and everything nested inside them.  With every
// region's own label histogram and its enclosing region ("parent"), the host adds up a contour's
// histogram by folding children into parents (hotpath.objects_from_regions).  Regions are named by
// their first pixel in raster order -- the point cv2 starts an outer contour at -- and the parent of a
// region is the region of the pixel left of that first pixel.
//
// Labelling: union-find over the ROI with atomicMin links (larger root under smaller, so a region's root
// IS its first pixel): init, merge (west / north-west / north / north-east for foreground, west / north
// for background), flatten, then slot assignment for roots and warp-aggregated histogram atomics.
#include "ssb_common.cuh"

namespace ssb {
namespace {

constexpr int kBlobThreads = 256;

struct BlobParams {
    const uint8_t *mask;     // [B][H][W], 0 / 255
    const uint8_t *labels;   // [B][H][W]
    int B, H, W, roi_h, roi_w, C, max_regions;
    int *root;               // scratch [B][roi_h*roi_w]: union-find parent, then the region's first pixel
    int *slot;               // scratch [B][roi_h*roi_w]: table row of a root pixel (valid at roots only)
    int *count;              // [B] regions found (may exceed max_regions: the frame's table is incomplete then)
    int *first;              // [B][max_regions] first pixel (y*roi_w + x)
    int *parent;             // [B][max_regions] first pixel of the enclosing region, -1 = none (first pixel in column 0)
    int *flags;              // [B][max_regions] bit 0: foreground region, bit 1: touches the ROI border
    unsigned int *hist;      // [B][max_regions][C]
};

__device__ __forceinline__ int find_root(const int *par, int a) {
    int p = par[a];
    while (p != a) {
        a = p;
        p = par[a];
    }
    return a;
}

// Links the trees of a and b: the larger root goes under the smaller one.
__device__ __forceinline__ void unite(int *par, int a, int b) {
    while (true) {
        a = find_root(par, a);
        b = find_root(par, b);
        if (a == b) return;
        if (a < b) {
            const int t = a;
            a = b;
            b = t;
        }
        const int old = atomicMin(&par[a], b);   // a was a root (par[a] == a) unless somebody linked it meanwhile
        if (old == a) return;
        a = old;                                  // retry from where a points now
    }
}

// A warp walks 32-pixel segments of rows.  Inside a segment every pixel starts out pointing at the first
// pixel of its horizontal run (ballot of "differs from the pixel on the left", no atomics); what is left
// for the merge pass are the links a run structure does not imply: a segment's first pixel to the run
// that continues from the previous segment, and vertical / diagonal links where the pixel on the left
// has not already made them.
__device__ __forceinline__ bool segment_of(const BlobParams &p, long long seg, int lane, int *f, int *y, int *x) {
    const int segs_per_row = (p.roi_w + 31) >> 5;
    const long long row = seg / segs_per_row;
    *f = static_cast<int>(row / p.roi_h);
    *y = static_cast<int>(row - static_cast<long long>(*f) * p.roi_h);
    *x = static_cast<int>(seg - row * segs_per_row) * 32 + lane;
    return *x < p.roi_w;
}

__global__ void __launch_bounds__(kBlobThreads) blob_init_kernel(const BlobParams p) {
    pdl_trigger();
    pdl_wait();
    const int lane = threadIdx.x & 31, n = p.roi_h * p.roi_w;
    const long long n_seg = static_cast<long long>(p.B) * p.roi_h * ((p.roi_w + 31) >> 5);
    const long long warp0 = (blockIdx.x * static_cast<long long>(kBlobThreads) + threadIdx.x) >> 5;
    const long long n_warps = static_cast<long long>(gridDim.x) * (kBlobThreads >> 5);
    for (long long seg = warp0; seg < n_seg; seg += n_warps) {
        int f, y, x;
        const bool live = segment_of(p, seg, lane, &f, &y, &x);
        const uint8_t *m = p.mask + static_cast<long long>(f) * p.H * p.W + static_cast<long long>(y) * p.W;
        const bool fg = live && m[x] != 0;
        // bit l: lane l starts a run inside this segment (lane 0 always does: its link to the left is made later)
        const bool starts = live && (lane == 0 || (m[x - 1] != 0) != fg);
        const unsigned int bits = __ballot_sync(0xffffffffu, starts);
        if (live) {
            const int start_lane = 31 - __clz(bits & (0xffffffffu >> (31 - lane)));
            p.root[static_cast<long long>(f) * n + y * p.roi_w + x] = y * p.roi_w + x - (lane - start_lane);
        }
    }
}

__global__ void __launch_bounds__(kBlobThreads) blob_merge_kernel(const BlobParams p) {
    pdl_trigger();
    pdl_wait();
    const int lane = threadIdx.x & 31, n = p.roi_h * p.roi_w;
    const long long n_seg = static_cast<long long>(p.B) * p.roi_h * ((p.roi_w + 31) >> 5);
    const long long warp0 = (blockIdx.x * static_cast<long long>(kBlobThreads) + threadIdx.x) >> 5;
    const long long n_warps = static_cast<long long>(gridDim.x) * (kBlobThreads >> 5);
    for (long long seg = warp0; seg < n_seg; seg += n_warps) {
        int f, y, x;
        if (!segment_of(p, seg, lane, &f, &y, &x)) continue;
        const uint8_t *m = p.mask + static_cast<long long>(f) * p.H * p.W + static_cast<long long>(y) * p.W;
        int *par = p.root + static_cast<long long>(f) * n;
        const int r = y * p.roi_w + x;
        const bool fg = m[x] != 0;
        const bool w_same = x > 0 && (m[x - 1] != 0) == fg;
        if (lane == 0 && w_same) unite(par, r, r - 1);              // the run continues from the previous segment
        if (y == 0) continue;
        const uint8_t *up = m - p.W;
        const bool n_same = (up[x] != 0) == fg;
        const bool nw_fg = x > 0 && up[x - 1] != 0, ne_fg = x + 1 < p.roi_w && up[x + 1] != 0;
        if (n_same) {
            // implied by the pixel on the left when it, and the pixel above it, are of this kind too
            const bool nw_same = x > 0 && (up[x - 1] != 0) == fg;
            if (!(w_same && nw_same)) unite(par, r, r - p.roi_w);
        } else if (fg) {
            // diagonals connect foreground only, and only matter when the pixel above is background;
            // the west / east neighbour makes the link itself when it is foreground (it sits below that pixel)
            if (nw_fg && !w_same) unite(par, r, r - p.roi_w - 1);
            const bool e_fg = x + 1 < p.roi_w && m[x + 1] != 0;
            if (ne_fg && !e_fg) unite(par, r, r - p.roi_w + 1);
        }
    }
}

__global__ void __launch_bounds__(kBlobThreads) blob_flatten_kernel(const BlobParams p) {
    pdl_trigger();
    pdl_wait();
    const int n = p.roi_h * p.roi_w;
    const long long total = static_cast<long long>(p.B) * n;
    for (long long i = blockIdx.x * static_cast<long long>(kBlobThreads) + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * kBlobThreads) {
        const int f = static_cast<int>(i / n), r = static_cast<int>(i - static_cast<long long>(f) * n);
        int *par = p.root + static_cast<long long>(f) * n;
        const int root = find_root(par, r);
        // Writing the root shortens other threads' walks and never breaks them: par[r] still points at
        // an ancestor.  A root keeps par[root] == root.
        par[r] = root;
        if (root == r) {
            const int s = atomicAdd(&p.count[f], 1);
            p.slot[i] = s;
            if (s < p.max_regions) {
                const int y = r / p.roi_w, x = r - y * p.roi_w;
                const bool fg = p.mask[static_cast<long long>(f) * p.H * p.W + y * p.W + x] != 0;
                p.first[static_cast<long long>(f) * p.max_regions + s] = r;
                p.flags[static_cast<long long>(f) * p.max_regions + s] = fg ? 1 : 0;
            }
        }
    }
}

// After the flatten pass every par[r] is a root or one step from it (a thread may have read an ancestor
// before ITS par[] was rewritten), so one more hop settles it.
__global__ void __launch_bounds__(kBlobThreads) blob_tables_kernel(const BlobParams p) {
    pdl_trigger();
    pdl_wait();
    const int n = p.roi_h * p.roi_w;
    const long long total = static_cast<long long>(p.B) * n;
    const int lane = threadIdx.x & 31;
    // whole warps stay in the loop (shuffles below): the bound is rounded up to a multiple of 32
    const long long total32 = (total + 31) / 32 * 32;
    for (long long i = blockIdx.x * static_cast<long long>(kBlobThreads) + threadIdx.x; i < total32;
         i += static_cast<long long>(gridDim.x) * kBlobThreads) {
        const bool live = i < total;
        int f = 0, r = 0, root = 0, s = 0x7fffffff, lab = 0;
        bool fg = false, edge = false;
        if (live) {
            f = static_cast<int>(i / n);
            r = static_cast<int>(i - static_cast<long long>(f) * n);
            const int *par = p.root + static_cast<long long>(f) * n;
            root = find_root(par, r);
            s = p.slot[static_cast<long long>(f) * n + root];
            const int y = r / p.roi_w, x = r - y * p.roi_w;
            const long long px = static_cast<long long>(f) * p.H * p.W + y * p.W + x;
            lab = p.labels[px];
            fg = p.mask[px] != 0;
            edge = x == 0 || y == 0 || x == p.roi_w - 1 || y == p.roi_h - 1;
            if (r == root && s < p.max_regions) {
                // the region that encloses this one owns the pixel left of its first pixel
                int par_first = -1;
                if (x > 0) par_first = find_root(par, r - 1);
                p.parent[static_cast<long long>(f) * p.max_regions + s] = par_first;
            }
        }
        const bool counted = live && s < p.max_regions && lab < p.C;
        // one atomic per distinct (frame, region, label) in the warp
        const unsigned long long key = counted ? ((static_cast<unsigned long long>(f) << 40) | (static_cast<unsigned long long>(s) << 8) | lab)
                                               : ~0ull - lane;
        const unsigned int peers = __match_any_sync(0xffffffffu, key);
        if (counted && lane == __ffs(peers) - 1)
            atomicAdd(&p.hist[(static_cast<long long>(f) * p.max_regions + s) * p.C + lab], __popc(peers));
        if (counted && edge && !fg) atomicOr(&p.flags[static_cast<long long>(f) * p.max_regions + s], 2);
    }
}

}  // namespace
}  // namespace ssb

extern "C" size_t ssb_blob_regions_scratch_bytes(int B, int roi_h, int roi_w) {
    if (B < 0 || roi_h < 0 || roi_w < 0) return 0;
    return sizeof(int) * 2 * static_cast<size_t>(B) * roi_h * roi_w;
}

extern "C" int ssb_blob_regions_u8(const uint8_t *mask, const uint8_t *labels, int B, int H, int W, int roi_h, int roi_w,
                                   int C, int max_regions, void *scratch, int32_t *count, int32_t *first,
                                   int32_t *parent, int32_t *flags, uint32_t *hist, void *stream) {
    using namespace ssb;
    if (B < 0 || H <= 0 || W <= 0) return fail(SSB_ERR_BAD_SHAPE, "bad shape B=%d H=%d W=%d", B, H, W);
    if (roi_h < 1 || roi_h > H || roi_w < 1 || roi_w > W)
        return fail(SSB_ERR_BAD_SHAPE, "ROI (%d,%d) outside frame (%d,%d)", roi_h, roi_w, H, W);
    if (C < 1 || C > 256 || max_regions < 1) return fail(SSB_ERR_LIMIT, "C=%d outside [1,256] or max_regions=%d < 1", C, max_regions);
    if (static_cast<long long>(roi_h) * roi_w >= (1ll << 31)) return fail(SSB_ERR_LIMIT, "roi_h*roi_w must be < 2^31");
    if (max_regions >= (1 << 24)) return fail(SSB_ERR_LIMIT, "max_regions must be < 2^24");
    if (B == 0) return SSB_OK;
    if (!mask || !labels || !scratch || !count || !first || !parent || !flags || !hist)
        return fail(SSB_ERR_BAD_POINTER, "NULL buffer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t n = static_cast<size_t>(B) * roi_h * roi_w;
    BlobParams p{};
    p.mask = mask;
    p.labels = labels;
    p.B = B; p.H = H; p.W = W; p.roi_h = roi_h; p.roi_w = roi_w; p.C = C; p.max_regions = max_regions;
    p.root = static_cast<int *>(scratch);
    p.slot = p.root + n;
    p.count = count; p.first = first; p.parent = parent; p.flags = flags; p.hist = hist;
    SSB_CUDA_TRY(cudaMemsetAsync(count, 0, sizeof(int) * B, st));
    SSB_CUDA_TRY(cudaMemsetAsync(hist, 0, sizeof(uint32_t) * static_cast<size_t>(B) * max_regions * C, st));
    DeviceInfo di;
    if (int rc = device_info(&di)) return rc;
    long long blocks = (static_cast<long long>(n) + kBlobThreads - 1) / kBlobThreads;
    const long long cap = static_cast<long long>(di.sm_count) * 16;
    const dim3 grid(static_cast<unsigned>(blocks < cap ? blocks : cap));
    SSB_CUDA_TRY(launch_pdl(blob_init_kernel, grid, dim3(kBlobThreads), 0, st, p));
    SSB_CUDA_TRY(launch_pdl(blob_merge_kernel, grid, dim3(kBlobThreads), 0, st, p));
    SSB_CUDA_TRY(launch_pdl(blob_flatten_kernel, grid, dim3(kBlobThreads), 0, st, p));
    SSB_CUDA_TRY(launch_pdl(blob_tables_kernel, grid, dim3(kBlobThreads), 0, st, p));
    return SSB_OK;
}
