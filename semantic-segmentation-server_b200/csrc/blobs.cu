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
// region is the region of the pixel left of that first pixel.  A HOLE contour runs through the pixels of
// the enclosing region that are 4-adjacent to the hole: it encloses the hole, what is nested in it, and
// that ring, whose label histogram is the `ring` row of the hole.
//
// Labelling: union-find with atomicMin links (larger root under smaller, so a region's root IS its first
// pixel): run-based initial links, merge (west / north-west / north / north-east for foreground, west /
// north for background), flatten, then table rows for roots and warp-aggregated histogram atomics.
//
// Two implementations with the same tables:
//   * blob_cluster_kernel (ROIs whose eighth fits 16-bit pixel indices: everything up to ~720x720): ONE
//     launch, one thread-block cluster of 8 CTAs per frame, a CTA per horizontal strip.  The mask lives in
//     shared memory as bit rows (all neighbour tests are word operations, uniform over the warp that owns
//     a 32-pixel segment), the union-find of a strip in shared memory with 16-bit ids.  Global memory only
//     sees what crosses strips: one parent word and one row number per LOCAL ROOT (sparse use of the two
//     scratch planes), the ids of each strip's first and last row, and the tables.  Phases are separated
//     by cluster barriers.  DRAM traffic: mask and label map once.
//   * blob_{init,merge,flatten,tables}_kernel: the whole union-find in global memory, any ROI size.
#include <cstring>
#include <mutex>

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
    unsigned int *ring;      // [B][max_regions][C] or null: for a hole, labels of the enclosing region's pixels that touch it
};

__device__ __forceinline__ int find_root(const int *par, int a) {
    int p = par[a];
    while (p != a) {
        a = p;
        p = par[a];
    }
    return a;
}

// The same walk, halving the path as it goes (every visited node is re-pointed at its grandparent).  The
// plain stores race with other threads' walks and links, harmlessly: a node only ever gets re-pointed at
// one of its own ancestors, which keeps it in its tree and keeps parent < child.
__device__ __forceinline__ int find_root_halving(int *par, int a) {
    while (true) {
        const int p = par[a];
        if (p == a) return a;
        const int gp = par[p];
        if (gp == p) return p;
        par[a] = gp;
        a = gp;
    }
}

// Links the trees of a and b: the larger root goes under the smaller one.
__device__ __forceinline__ void unite(int *par, int a, int b) {
    while (true) {
        a = find_root_halving(par, a);
        b = find_root_halving(par, b);
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

// Thread mapping of all four kernels: block (32, 8) = 8 warps, warp w of block (bx, by, bz) owns the
// 32-pixel segment bx of rows 64*by + 8*w .. +7 of frame bz (no index divisions anywhere; one pixel per
// thread made the kernels bound by the rate at which an SM can start CTAs).
//
// Inside a segment every pixel starts out pointing at the first pixel of its horizontal run (ballot of
// "differs from the pixel on the left", no atomics); what is left for the merge pass are the links a run
// structure does not imply: a segment's first pixel to the run that continues from the previous segment,
// and vertical / diagonal links where the pixel on the left has not already made them.
struct PixelId {
    int f, y, x, r;
    bool live;
};
constexpr int kBlobRowsPerWarp = 8;
__device__ __forceinline__ PixelId pixel_of(const BlobParams &p, int k) {
    PixelId q;
    q.f = blockIdx.z;
    q.y = (blockIdx.y * 8 + threadIdx.y) * kBlobRowsPerWarp + k;
    q.x = blockIdx.x * 32 + threadIdx.x;
    q.live = q.y < p.roi_h && q.x < p.roi_w;
    q.r = q.y * p.roi_w + q.x;
    return q;
}

__global__ void __launch_bounds__(kBlobThreads) blob_init_kernel(const BlobParams p) {
    pdl_wait();
    for (int k = 0; k < kBlobRowsPerWarp; ++k) {
    const PixelId q = pixel_of(p, k);
    const int lane = threadIdx.x, n = p.roi_h * p.roi_w;
    const uint8_t *m = p.mask + static_cast<long long>(q.f) * p.H * p.W + static_cast<long long>(q.y) * p.W;
    const bool fg = q.live && m[q.x] != 0;
    // bit l: lane l starts a run inside this segment (lane 0 always does: its link to the left is made later)
    const bool starts = q.live && (lane == 0 || (m[q.x - 1] != 0) != fg);
    const unsigned int bits = __ballot_sync(0xffffffffu, starts);
    if (q.live) {
        const int start_lane = 31 - __clz(bits & (0xffffffffu >> (31 - lane)));
        p.root[static_cast<long long>(q.f) * n + q.r] = q.r - (lane - start_lane);
    }
    }
}

__global__ void __launch_bounds__(kBlobThreads) blob_merge_kernel(const BlobParams p) {
    pdl_wait();
    for (int k = 0; k < kBlobRowsPerWarp; ++k) {
    const PixelId q = pixel_of(p, k);
    if (!q.live) continue;
    const int lane = threadIdx.x, n = p.roi_h * p.roi_w, x = q.x, r = q.r;
    const uint8_t *m = p.mask + static_cast<long long>(q.f) * p.H * p.W + static_cast<long long>(q.y) * p.W;
    int *par = p.root + static_cast<long long>(q.f) * n;
    const bool fg = m[x] != 0;
    const bool w_same = x > 0 && (m[x - 1] != 0) == fg;
    if (lane == 0 && w_same) unite(par, r, r - 1);              // the run continues from the previous segment
    if (q.y == 0) continue;
    const uint8_t *up = m - p.W;
    const bool n_same = (up[x] != 0) == fg;
    if (n_same) {
        // implied by the pixel on the left when it, and the pixel above it, are of this kind too
        const bool nw_same = x > 0 && (up[x - 1] != 0) == fg;
        if (!(w_same && nw_same)) unite(par, r, r - p.roi_w);
    } else if (fg) {
        // diagonals connect foreground only, and only matter when the pixel above is background;
        // the west / east neighbour makes the link itself when it is foreground (it sits below that pixel)
        const bool nw_fg = x > 0 && up[x - 1] != 0, ne_fg = x + 1 < p.roi_w && up[x + 1] != 0;
        if (nw_fg && !w_same) unite(par, r, r - p.roi_w - 1);
        const bool e_fg = x + 1 < p.roi_w && m[x + 1] != 0;
        if (ne_fg && !e_fg) unite(par, r, r - p.roi_w + 1);
    }
    }
}

__global__ void __launch_bounds__(kBlobThreads) blob_flatten_kernel(const BlobParams p) {
    pdl_wait();
    for (int k = 0; k < kBlobRowsPerWarp; ++k) {
    const PixelId q = pixel_of(p, k);
    if (!q.live) continue;
    const int n = p.roi_h * p.roi_w, r = q.r;
    int *par = p.root + static_cast<long long>(q.f) * n;
    const int root = find_root_halving(par, r);
    // Writing the root shortens other threads' walks and never breaks them: par[r] still points at an
    // ancestor.  A root keeps par[root] == root.
    par[r] = root;
    if (root == r) {
        const int s = atomicAdd(&p.count[q.f], 1);
        p.slot[static_cast<long long>(q.f) * n + r] = s;
        if (s < p.max_regions) {
            const bool fg = p.mask[static_cast<long long>(q.f) * p.H * p.W + static_cast<long long>(q.y) * p.W + q.x] != 0;
            p.first[static_cast<long long>(q.f) * p.max_regions + s] = r;
            p.flags[static_cast<long long>(q.f) * p.max_regions + s] = fg ? 1 : 0;
        }
    }
    }
}

// After the flatten pass every par[r] is a root or one step from it (a thread may have read an ancestor
// before ITS par[] was rewritten), so one more hop settles it.
__global__ void __launch_bounds__(kBlobThreads) blob_tables_kernel(const BlobParams p) {
    pdl_wait();
    for (int k = 0; k < kBlobRowsPerWarp; ++k) {   // warp-uniform: the shuffles below see all 32 lanes
    const PixelId q = pixel_of(p, k);
    const int n = p.roi_h * p.roi_w, lane = threadIdx.x, f = q.f;
    int s = 0x7fffffff, lab = 0;
    bool fg = false, edge = false;
    if (q.live) {
        const int *par = p.root + static_cast<long long>(f) * n;
        const int root = find_root(par, q.r);
        s = p.slot[static_cast<long long>(f) * n + root];
        const long long px = static_cast<long long>(f) * p.H * p.W + static_cast<long long>(q.y) * p.W + q.x;
        lab = p.labels[px];
        fg = p.mask[px] != 0;
        edge = q.x == 0 || q.y == 0 || q.x == p.roi_w - 1 || q.y == p.roi_h - 1;
        if (q.r == root && s < p.max_regions) {
            // the region that encloses this one owns the pixel left of its first pixel
            p.parent[static_cast<long long>(f) * p.max_regions + s] = q.x > 0 ? find_root(par, q.r - 1) : -1;
        }
    }
    const bool counted = q.live && s < p.max_regions && lab < p.C;
    // one atomic per distinct (region, label) in the warp (a warp is 32 consecutive pixels of one row)
    const unsigned int key = counted ? (static_cast<unsigned int>(s) << 8) | static_cast<unsigned int>(lab) : 0xffffffffu - lane;
    const unsigned int peers = __match_any_sync(0xffffffffu, key);
    if (counted && lane == __ffs(peers) - 1)
        atomicAdd(&p.hist[(static_cast<long long>(f) * p.max_regions + s) * p.C + lab], __popc(peers));
    if (counted && edge && !fg) atomicOr(&p.flags[static_cast<long long>(f) * p.max_regions + s], 2);
    // Hole contours run through the pixels of the ENCLOSING foreground region that are 4-adjacent to the
    // hole (not through islands inside the hole): such a pixel adds its label to the hole's ring row.
    if (p.ring && q.live && fg && lab < p.C) {
        const int *par = p.root + static_cast<long long>(f) * n;
        const uint8_t *m = p.mask + static_cast<long long>(f) * p.H * p.W + static_cast<long long>(q.y) * p.W;
        const int mine = find_root(par, q.r);
        int seen[4], n_seen = 0;
        const int dr[4] = {-1, 1, -p.roi_w, p.roi_w};
        const bool ok[4] = {q.x > 0 && m[q.x - 1] == 0, q.x + 1 < p.roi_w && m[q.x + 1] == 0,
                            q.y > 0 && m[q.x - p.W] == 0, q.y + 1 < p.roi_h && m[q.x + p.W] == 0};
#pragma unroll
        for (int d = 0; d < 4; ++d) {
            if (!ok[d]) continue;
            const int g = find_root(par, q.r + dr[d]);          // the hole's first pixel
            bool dup = false;
            for (int e = 0; e < n_seen; ++e) dup |= seen[e] == g;
            if (dup) continue;
            seen[n_seen++] = g;
            if (g % p.roi_w == 0) continue;                     // column 0: touches the border, not a hole
            if (find_root(par, g - 1) != mine) continue;        // this pixel belongs to an island inside the hole
            const int sg = p.slot[static_cast<long long>(f) * n + g];
            if (sg < p.max_regions) atomicAdd(&p.ring[(static_cast<long long>(f) * p.max_regions + sg) * p.C + lab], 1u);
        }
    }
    }
}


#include "blobs_cluster.cuh"   // blob_cluster_kernel: one thread-block cluster per frame

}  // namespace
}  // namespace ssb

extern "C" size_t ssb_blob_regions_scratch_bytes(int B, int roi_h, int roi_w) {
    if (B < 0 || roi_h < 0 || roi_w < 0) return 0;
    // two int32 planes (union-find parents, table rows) + the ids of the first / last row of every strip
    return sizeof(int) * static_cast<size_t>(B) * (2 * static_cast<size_t>(roi_h) * roi_w + ssb::kStrips * 2 * static_cast<size_t>(roi_w));
}

extern "C" int ssb_blob_regions_u8(const uint8_t *mask, const uint8_t *labels, int B, int H, int W, int roi_h, int roi_w,
                                   int C, int max_regions, void *scratch, int32_t *count, int32_t *first,
                                   int32_t *parent, int32_t *flags, uint32_t *hist, uint32_t *ring, void *stream) {
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
    p.count = count; p.first = first; p.parent = parent; p.flags = flags; p.hist = hist; p.ring = ring;
    if (B > 65535) return fail(SSB_ERR_LIMIT, "B must be <= 65535");
    // one cluster per frame when a strip (an eighth of the ROI's rows) fits 16-bit pixel indices and shared memory
    {
        StripParams sp{};
        sp.b = p;
        sp.rows = (roi_h + kStrips - 1) / kStrips;
        sp.S = (roi_w + 31) / 32;
        sp.divS = make_fastdiv(static_cast<uint32_t>(sp.S > 1 ? sp.S : 2));
        sp.divW = make_fastdiv(static_cast<uint32_t>(roi_w > 1 ? roi_w : 2));
        sp.edge = p.slot + n;
        sp.L = strip_layout(sp.rows, sp.S, roi_w, C);
        const size_t smem = static_cast<size_t>(sp.L.total);
        DeviceInfo di;
        if (int rc = device_info(&di)) return rc;
        const char *force = getenv("SSB_BLOB_PATH");   // tuning / test knob: "global" = the global-memory kernels
        const bool want_global = force && !strcmp(force, "global");
        if (!want_global && static_cast<long long>(sp.rows) * roi_w <= 65535 && max_regions <= 0xFFFF &&
            smem <= static_cast<size_t>(di.max_smem_optin)) {
            static std::mutex mu;
            static size_t smem_set[64] = {0};
            int dev = 0;
            SSB_CUDA_TRY(cudaGetDevice(&dev));
            if (dev < 0 || dev >= 64) return fail(SSB_ERR_LIMIT, "device ordinal %d not supported", dev);
            {
                std::lock_guard<std::mutex> lock(mu);
                if (smem_set[dev] < smem) {
                    SSB_CUDA_TRY(cudaFuncSetAttribute(blob_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                      static_cast<int>(smem)));
                    // all of the SM's L1 / shared-memory array as shared memory: with the default split only three
                    // CTAs fit an SM and 48 clusters the GPU -- a batch of 64 frames then runs as two waves
                    SSB_CUDA_TRY(cudaFuncSetAttribute(blob_cluster_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                                      cudaSharedmemCarveoutMaxShared));
                    smem_set[dev] = smem;
                }
            }
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3(kStrips, B);
            cfg.blockDim = dim3(kStripThreads);
            cfg.dynamicSmemBytes = smem;
            cfg.stream = st;
            cudaLaunchAttribute attr[2];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = kStrips;
            attr[0].val.clusterDim.y = 1;
            attr[0].val.clusterDim.z = 1;
            attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[1].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = attr;
            cfg.numAttrs = 2;
            if (getenv("SSB_BLOB_DEBUG")) {   // tuning aid: how many clusters the GPU holds at once
                int n_clusters = 0;
                cudaOccupancyMaxActiveClusters(&n_clusters, blob_cluster_kernel, &cfg);
                fprintf(stderr, "blob_cluster_kernel: %d threads, %zu B shared memory, %d clusters of %d active at once\n",
                        kStripThreads, smem, n_clusters, kStrips);
            }
            if (cudaLaunchKernelEx(&cfg, blob_cluster_kernel, sp) == cudaSuccess) return SSB_OK;
            cudaGetLastError();   // a device / context that cannot place the cluster: the global-memory kernels below
        }
    }
    SSB_CUDA_TRY(cudaMemsetAsync(count, 0, sizeof(int) * B, st));
    SSB_CUDA_TRY(cudaMemsetAsync(hist, 0, sizeof(uint32_t) * static_cast<size_t>(B) * max_regions * C, st));
    if (ring) SSB_CUDA_TRY(cudaMemsetAsync(ring, 0, sizeof(uint32_t) * static_cast<size_t>(B) * max_regions * C, st));
    constexpr int kRowsPerCta = 8 * kBlobRowsPerWarp;
    if ((roi_h + kRowsPerCta - 1) / kRowsPerCta > 65535) return fail(SSB_ERR_LIMIT, "roi_h/64 must be <= 65535");
    const dim3 grid((roi_w + 31) / 32, (roi_h + kRowsPerCta - 1) / kRowsPerCta, B), block(32, 8);
    SSB_CUDA_TRY(launch_pdl(blob_init_kernel, grid, block, 0, st, p));
    SSB_CUDA_TRY(launch_pdl(blob_merge_kernel, grid, block, 0, st, p));
    SSB_CUDA_TRY(launch_pdl(blob_flatten_kernel, grid, block, 0, st, p));
    SSB_CUDA_TRY(launch_pdl(blob_tables_kernel, grid, block, 0, st, p));
    return SSB_OK;
}
