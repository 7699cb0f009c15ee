// Host-side helpers of the producer loop (sem_seg_server.py:144-195 for a batch of frames), native because
// the per-frame work they replace was Python under the GIL:
//   ssb_frames_to_device   cap.read() frames (separate host arrays) -> one device batch: frames that already
//                          sit in page-locked memory go by DMA straight from where they are, the others are
//                          gathered into the caller's pinned staging buffer by a pool of host threads and sent
//                          from there, frame by frame, so gathering and DMA overlap;
//   ssb_fold_regions_host  region tables of a batch (ssb_blob_regions_u8, copied to the host) -> per contour
//                          look-up tables: label histogram of everything an outer contour / a hole contour
//                          encloses, keyed by the region's first pixel (what sem_seg_server.py:98-107 gets from
//                          a polygon fill per contour).
// No kernels here; compiled with the rest of the library because it uses the CUDA runtime.
#include <algorithm>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif
#include <atomic>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

#include "ssb_common.cuh"

namespace ssb {
namespace {

// A small pool of host threads, created on first use and kept.  run(n, threads, fn) calls fn(i) for every
// i in [0, n) from up to `threads` threads (the caller's included) and returns when all are done.
class HostPool {
  public:
    void run(int n, int threads, const std::function<void(int)> &fn) {
        if (n <= 0) return;
        threads = std::max(1, std::min(threads, n));
        if (threads == 1) {
            for (int i = 0; i < n; ++i) fn(i);
            return;
        }
        std::lock_guard<std::mutex> one_job(job_mu_);
        {
            std::unique_lock<std::mutex> lk(mu_);
            while (static_cast<int>(workers_.size()) < threads - 1) {
                const int id = static_cast<int>(workers_.size());
                workers_.emplace_back([this, id] { worker(id); });
                workers_.back().detach();
            }
            fn_ = &fn;
            n_ = n;
            next_.store(0, std::memory_order_relaxed);
            helpers_ = threads - 1;
            running_ = helpers_;
            ++generation_;
        }
        work_.notify_all();
        drain();
        std::unique_lock<std::mutex> lk(mu_);
        done_.wait(lk, [this] { return running_ == 0; });
        fn_ = nullptr;
    }

  private:
    void drain() {
        for (;;) {
            const int i = next_.fetch_add(1, std::memory_order_relaxed);
            if (i >= n_) return;
            (*fn_)(i);
        }
    }
    void worker(int id) {
        unsigned seen = 0;
        std::unique_lock<std::mutex> lk(mu_);
        for (;;) {
            work_.wait(lk, [&] { return generation_ != seen; });
            seen = generation_;
            if (id >= helpers_) continue;      // not invited to this job
            lk.unlock();
            drain();
            lk.lock();
            if (--running_ == 0) done_.notify_one();
        }
    }
    std::mutex job_mu_, mu_;
    std::condition_variable work_, done_;
    std::vector<std::thread> workers_;
    const std::function<void(int)> *fn_ = nullptr;
    std::atomic<int> next_{0};
    int n_ = 0, helpers_ = 0, running_ = 0;
    unsigned generation_ = 0;
};

// Two pools: a pool runs one job at a time, and the feeder's frame gather (milliseconds per batch with pageable frames)
// must not hold up the consumer's table fold / contour stage of the previous batch.
// Copy for the frame gather: the destination (page-locked staging) is written once and read next by the DMA engine,
// never by a CPU -- streaming (non-temporal) stores skip the read-for-ownership of every destination line, a third of
// the memory traffic of a plain memcpy at these sizes (6 MB per frame is below the size at which glibc switches to
// streaming stores by itself).
void copy_streaming(uint8_t *dst, const uint8_t *src, size_t n) {
#if defined(__SSE2__)
    const size_t head = (16 - (reinterpret_cast<uintptr_t>(dst) & 15)) & 15;
    if (n < 4096 || head >= n) { std::memcpy(dst, src, n); return; }
    std::memcpy(dst, src, head);
    dst += head; src += head; n -= head;
    const size_t blocks = n / 64;
    for (size_t i = 0; i < blocks; ++i) {
        const __m128i a = _mm_loadu_si128(reinterpret_cast<const __m128i *>(src) + 4 * i + 0);
        const __m128i b = _mm_loadu_si128(reinterpret_cast<const __m128i *>(src) + 4 * i + 1);
        const __m128i c = _mm_loadu_si128(reinterpret_cast<const __m128i *>(src) + 4 * i + 2);
        const __m128i d = _mm_loadu_si128(reinterpret_cast<const __m128i *>(src) + 4 * i + 3);
        _mm_stream_si128(reinterpret_cast<__m128i *>(dst) + 4 * i + 0, a);
        _mm_stream_si128(reinterpret_cast<__m128i *>(dst) + 4 * i + 1, b);
        _mm_stream_si128(reinterpret_cast<__m128i *>(dst) + 4 * i + 2, c);
        _mm_stream_si128(reinterpret_cast<__m128i *>(dst) + 4 * i + 3, d);
    }
    _mm_sfence();                                   // the stores are globally visible before the DMA is enqueued
    std::memcpy(dst + blocks * 64, src + blocks * 64, n - blocks * 64);
#else
    std::memcpy(dst, src, n);
#endif
}

HostPool &pool() {
    static HostPool *p = new HostPool();   // never destroyed: its detached workers may outlive static destruction
    return *p;
}
HostPool &upload_pool() {
    static HostPool *p = new HostPool();
    return *p;
}

}  // namespace
}  // namespace ssb

extern "C" int ssb_frames_to_device(const uint8_t *const *frames_host, int n, size_t frame_bytes, uint8_t *staging_host,
                                    uint8_t *dst_device, int threads, void *stream) {
    using namespace ssb;
    if (n < 0) return fail(SSB_ERR_BAD_SHAPE, "n=%d", n);
    if (n == 0 || frame_bytes == 0) return SSB_OK;
    if (!frames_host || !dst_device) return fail(SSB_ERR_BAD_POINTER, "NULL buffer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // which frames can be sent from where they are?
    std::vector<char> direct(n, 0);
    int n_staged = 0;
    for (int i = 0; i < n; ++i) {
        if (!frames_host[i]) return fail(SSB_ERR_BAD_POINTER, "frame %d is NULL", i);
        cudaPointerAttributes attr{};
        const cudaError_t e = cudaPointerGetAttributes(&attr, frames_host[i]);
        if (e != cudaSuccess) cudaGetLastError();   // older runtimes report unregistered host memory as an error
        direct[i] = e == cudaSuccess && attr.type == cudaMemoryTypeHost;
        n_staged += !direct[i];
    }
    if (n_staged && !staging_host)
        return fail(SSB_ERR_BAD_POINTER, "%d frames are in pageable memory and no pinned staging buffer was given", n_staged);
    std::atomic<int> failed{0};
    const char *env = getenv("SSB_GATHER_COPY");          // tuning knob: "memcpy" = plain memcpy for the gather
    const bool streaming = !(env && !strcmp(env, "memcpy"));
    auto send = [&](int i) {
        const uint8_t *src = frames_host[i];
        if (!direct[i]) {
            if (streaming) copy_streaming(staging_host + frame_bytes * i, src, frame_bytes);
            else std::memcpy(staging_host + frame_bytes * i, src, frame_bytes);
            src = staging_host + frame_bytes * i;
        }
        if (cudaMemcpyAsync(dst_device + frame_bytes * i, src, frame_bytes, cudaMemcpyHostToDevice, st) != cudaSuccess)
            failed.store(1);
    };
    int dev = 0;
    SSB_CUDA_TRY(cudaGetDevice(&dev));
    // direct frames first (no host work; frames that follow each other in memory go as one copy), then the
    // gathered ones as their copies finish
    for (int i = 0; i < n;) {
        if (!direct[i]) { ++i; continue; }
        int j = i + 1;
        while (j < n && direct[j] && frames_host[j] == frames_host[j - 1] + frame_bytes) ++j;
        if (cudaMemcpyAsync(dst_device + frame_bytes * i, frames_host[i], frame_bytes * (j - i), cudaMemcpyHostToDevice, st) !=
            cudaSuccess)
            failed.store(1);
        i = j;
    }
    if (n_staged) {
        std::vector<int> todo;
        for (int i = 0; i < n; ++i)
            if (!direct[i]) todo.push_back(i);
        upload_pool().run(static_cast<int>(todo.size()), threads, [&](int j) {
            cudaSetDevice(dev);            // pool threads start on device 0
            send(todo[j]);
        });
    }
    if (failed.load()) return fail(SSB_ERR_CUDA, "cudaMemcpyAsync failed: %s", cudaGetErrorString(cudaGetLastError()));
    return SSB_OK;
}

extern "C" int ssb_fold_regions_host(int B, int C, int max_regions, const int32_t *count, const int32_t *first,
                                     const int32_t *parent, const int32_t *flags, const uint32_t *hist,
                                     const uint32_t *ring, int64_t *enclosed, int32_t *order, uint8_t *complete,
                                     int threads) {
    using namespace ssb;
    if (B < 0 || C < 1 || max_regions < 1) return fail(SSB_ERR_BAD_SHAPE, "B=%d C=%d max_regions=%d", B, C, max_regions);
    if (B == 0) return SSB_OK;
    if (!count || !first || !parent || !flags || !hist || !enclosed || !order || !complete)
        return fail(SSB_ERR_BAD_POINTER, "NULL buffer");
    const size_t R = static_cast<size_t>(max_regions);
    pool().run(B, threads, [&](int b) {
        const int n = count[b];
        complete[b] = n >= 0 && n <= max_regions;
        if (!complete[b]) return;
        const int32_t *f = first + R * b, *par = parent + R * b, *fl = flags + R * b;
        const uint32_t *h = hist + R * C * b, *rg = ring ? ring + R * C * b : nullptr;
        int64_t *enc = enclosed + R * C * b;
        int32_t *ord = order + R * b;
        for (int i = 0; i < n; ++i) ord[i] = i;
        std::sort(ord, ord + n, [&](int32_t a, int32_t c) { return f[a] < f[c]; });   // rows by first pixel
        for (size_t e = 0; e < static_cast<size_t>(n) * C; ++e) enc[e] = h[e];
        auto outside = [&](int i) { return (fl[i] & 3) == 2; };    // background touching the ROI border
        // A region's children all have larger first pixels than it: one pass in descending order of first
        // pixel folds every subtree into its root.
        for (int k = n - 1; k >= 0; --k) {
            const int i = ord[k];
            if (outside(i) || par[i] < 0) continue;               // the outside, or a region in column 0: nothing encloses it
            const int32_t *pos = std::lower_bound(ord, ord + n, par[i], [&](int32_t a, int32_t key) { return f[a] < key; });
            if (pos == ord + n || f[*pos] != par[i]) continue;     // cannot happen with a complete table
            const int j = *pos;
            if (outside(j)) continue;                              // enclosed by the outside only
            for (int c = 0; c < C; ++c) enc[static_cast<size_t>(j) * C + c] += enc[static_cast<size_t>(i) * C + c];
        }
        // a hole contour also encloses the ring of the surrounding region's pixels it runs through
        if (rg)
            for (int i = 0; i < n; ++i)
                if (!(fl[i] & 1) && !outside(i))
                    for (int c = 0; c < C; ++c) enc[static_cast<size_t>(i) * C + c] += rg[static_cast<size_t>(i) * C + c];
    });
    return SSB_OK;
}

// ---- contour stage of one frame: border following with marks (Suzuki & Abe), as cv2.findContours(RETR_TREE) ----
// The reference calls cv2.findContours / contourArea / moments per frame (sem_seg_server.py:88, :93, :115-122).
// cv2.findContours does not run in parallel on threads of a Python process; this restatement does (a frame per
// pool thread).  It reproduces, on the ROI crop of a mask:
//   * the borders OpenCV follows (8-connected foreground, outer and hole borders), in OpenCV's OUTPUT ORDER
//     (pre-order of the border tree, siblings in reverse order of discovery) -- discovery depends on the marks a
//     traced border leaves (a pixel whose right-hand neighbour was examined and found empty is marked negative
//     and cannot start a hole border any more), so the marks are kept exactly;
//   * contourArea and the moments m00, m10, m01 of every border: sums of products of integer coordinates, exact in
//     64-bit integers, and independent of which collinear points CHAIN_APPROX_SIMPLE drops and of where on the
//     border the point list starts; the final doubles are formed with OpenCV's own constants (0.5, 1/6).
// Equality with OpenCV is a test (tests/test_hostpath.py: contour order, hole flags, start points of outer
// borders, areas and centroids over > 10 000 random masks), not an assumption.
namespace ssb {
namespace {

struct BorderInfo {
    int parent, last_child, prev_sibling;   // tree over border numbers; children chained newest first
    int start_x, start_y;                   // where the raster scan met the border
    int key;                                // first pixel (y*w + x) of the region (outer) / of the hole (hole border), -1 unknown
    bool is_hole;
    long long a00, a10, a01;
};

struct ContourScratch {
    std::vector<int> img;
    std::vector<BorderInfo> borders;
    std::vector<int> stack;
};

// Follows the border the raster scan met at pixel (x0, y0) of the ROI (unpadded coordinates; `img` is the
// padded image, pixel (x, y) at img[(y + 1) * pitch + x + 1]) and numbers it `nbd`.
void follow_border(int *img, int pitch, int x0, int y0, bool is_hole, int nbd, int w, BorderInfo &b) {
    static const int dx[8] = {1, 1, 0, -1, -1, -1, 0, 1}, dy[8] = {0, -1, -1, -1, 0, 1, 1, 1};
    int delta[16];
    for (int s = 0; s < 16; ++s) delta[s] = dy[s & 7] * pitch + dx[s & 7];
    int *i0 = img + (y0 + 1) * pitch + x0 + 1;
    int s_end = is_hole ? 0 : 4, s = s_end;
    int *i1;
    do {
        s = (s - 1) & 7;
        i1 = i0 + delta[s];
    } while (*i1 == 0 && s != s_end);
    long long a00 = 0, a10 = 0, a01 = 0;
    int y_min = y0, best_x = 1 << 30;      // hole borders: top row of the border, leftmost point of it with background below
    if (*i1 == 0) {                         // an isolated pixel: a border of one point, all sums zero
        *i0 = -nbd;
    } else {
        int *i3 = i0;
        int x = x0, y = y0, prev_x = x0, prev_y = y0;
        bool first = true;
        y_min = 1 << 30;
        for (;;) {
            s_end = s;
            int *i4;
            for (;;) {
                i4 = i3 + delta[++s];
                if (*i4 != 0) break;
            }
            s &= 7;
            if (static_cast<unsigned>(s - 1) < static_cast<unsigned>(s_end)) *i3 = -nbd;   // right-hand neighbour examined and empty
            else if (*i3 == 1) *i3 = nbd;
            // (x, y) is a point of the closed polygon
            if (!first) {
                const long long d = static_cast<long long>(prev_x) * y - static_cast<long long>(x) * prev_y;
                a00 += d; a10 += d * (prev_x + x); a01 += d * (prev_y + y);
            }
            first = false;
            if (is_hole) {
                if (y < y_min) { y_min = y; best_x = 1 << 30; }
                if (y == y_min && x < best_x && i3[pitch] == 0) best_x = x;
            }
            prev_x = x; prev_y = y;
            if (i4 == i0 && i3 == i1) break;
            i3 = i4;
            x += dx[s]; y += dy[s];
            s = (s + 4) & 7;
        }
        const long long d = static_cast<long long>(prev_x) * y0 - static_cast<long long>(x0) * prev_y;   // back to the first point
        a00 += d; a10 += d * (prev_x + x0); a01 += d * (prev_y + y0);
    }
    b.a00 = a00; b.a10 = a10; b.a01 = a01;
    b.key = !is_hole ? y0 * w + x0 : (best_x < (1 << 30) ? (y_min + 1) * w + best_x : -1);
}

// All borders of one ROI (h x w, row pitch `ld` bytes), in OpenCV's output order, into sc.borders / `out_order`.
void frame_borders(const uint8_t *mask, int h, int w, int ld, ContourScratch &sc, std::vector<int> &out_order) {
    const int pitch = w + 2;
    sc.img.assign(static_cast<size_t>(h + 2) * pitch, 0);
    for (int y = 0; y < h; ++y) {
        const uint8_t *m = mask + static_cast<size_t>(y) * ld;
        int *row = sc.img.data() + static_cast<size_t>(y + 1) * pitch + 1;
        for (int x = 0; x < w; ++x) row[x] = m[x] != 0;
    }
    sc.borders.clear();
    sc.borders.push_back(BorderInfo{});          // 0: unused
    BorderInfo frame{};                          // 1: the frame, a hole border, root of the tree
    frame.parent = 0; frame.last_child = 0; frame.prev_sibling = 0; frame.is_hole = true;
    sc.borders.push_back(frame);
    int *img = sc.img.data();
    for (int y = 0; y < h; ++y) {
        int lnbd = 1;
        int *row = img + static_cast<size_t>(y + 1) * pitch + 1;
        for (int x = 0; x < w; ++x) {
            int f = row[x];
            if (f == 0) continue;
            int start = -1;                      // 0: outer border starts here, 1: hole border
            if (f == 1 && row[x - 1] == 0) {
                start = 0;
            } else if (f >= 1 && row[x + 1] == 0) {
                start = 1;
                if (f > 1) lnbd = f;
            }
            if (start >= 0) {
                const int nbd = static_cast<int>(sc.borders.size());
                const BorderInfo &prev = sc.borders[lnbd];
                BorderInfo b{};
                b.is_hole = start == 1;
                // Suzuki & Abe, table 1: same kind as the last border met on this row -> its parent; else that border
                b.parent = b.is_hole == prev.is_hole ? prev.parent : lnbd;
                b.start_x = x; b.start_y = y;
                b.last_child = 0;
                b.prev_sibling = sc.borders[b.parent].last_child;
                sc.borders.push_back(b);
                sc.borders[b.parent].last_child = nbd;
                follow_border(img, pitch, x, y, start == 1, nbd, w, sc.borders[nbd]);
                f = row[x];
            }
            if (f != 1) lnbd = f < 0 ? -f : f;
        }
    }
    // pre-order of the tree, children newest first
    out_order.clear();
    sc.stack.clear();
    for (int c = sc.borders[1].last_child; c; c = sc.borders[c].prev_sibling) sc.stack.push_back(c);
    // the stack pops the LAST pushed first: push siblings oldest-first so that the newest comes out first
    std::reverse(sc.stack.begin(), sc.stack.end());
    while (!sc.stack.empty()) {
        const int n = sc.stack.back();
        sc.stack.pop_back();
        out_order.push_back(n);
        const size_t at = sc.stack.size();
        for (int c = sc.borders[n].last_child; c; c = sc.borders[c].prev_sibling) sc.stack.push_back(c);
        std::reverse(sc.stack.begin() + at, sc.stack.end());
    }
}

}  // namespace
}  // namespace ssb

extern "C" int ssb_contours_host(const uint8_t *masks_host, int B, int H, int W, int roi_h, int roi_w, double min_area,
                                 int max_candidates, int32_t *n_candidates, int32_t *cand_i32, double *cand_area,
                                 int32_t *n_contours, int threads) {
    using namespace ssb;
    if (B < 0 || H <= 0 || W <= 0 || roi_h < 1 || roi_h > H || roi_w < 1 || roi_w > W || max_candidates < 1)
        return fail(SSB_ERR_BAD_SHAPE, "bad shape B=%d H=%d W=%d ROI (%d,%d) max_candidates=%d", B, H, W, roi_h, roi_w, max_candidates);
    if (B == 0) return SSB_OK;
    if (!masks_host || !n_candidates || !cand_i32 || !cand_area) return fail(SSB_ERR_BAD_POINTER, "NULL buffer");
    pool().run(B, threads, [&](int b) {
        thread_local ContourScratch sc;
        thread_local std::vector<int> order;
        frame_borders(masks_host + static_cast<size_t>(b) * H * W, roi_h, roi_w, W, sc, order);
        if (n_contours) n_contours[b] = static_cast<int32_t>(order.size());
        int n = 0;
        for (int nbd : order) {
            const BorderInfo &c = sc.borders[nbd];
            // cv2.contourArea: |sum of cross products| / 2;  cv2.moments: m00 = a00/2, m10 = a10/6, m01 = a01/6 with the
            // sign of a00 folded into the constants -- the same doubles, from exact integer sums
            double a00 = static_cast<double>(c.a00);
            const double area = (a00 < 0 ? -a00 : a00) * 0.5;
            if (area < min_area) continue;                               // sem_seg_server.py:94
            if (c.a00 == 0) continue;                                    // m00 == 0, sem_seg_server.py:117
            const double db1_2 = c.a00 > 0 ? 0.5 : -0.5, db1_6 = c.a00 > 0 ? 0.16666666666666666666666666666667 : -0.16666666666666666666666666666667;
            const double m00 = a00 * db1_2, m10 = static_cast<double>(c.a10) * db1_6, m01 = static_cast<double>(c.a01) * db1_6;
            if (n < max_candidates) {
                int32_t *o = cand_i32 + (static_cast<size_t>(b) * max_candidates + n) * 4;
                o[0] = c.is_hole; o[1] = c.key;
                o[2] = static_cast<int32_t>(m10 / m00);                  // int(M['m10'] / M['m00']), sem_seg_server.py:121
                o[3] = static_cast<int32_t>(m01 / m00);
                cand_area[static_cast<size_t>(b) * max_candidates + n] = area;
            }
            ++n;
        }
        n_candidates[b] = n;                                             // > max_candidates: the list is cut short
    });
    return SSB_OK;
}

// Contour candidates of a batch from the region tables and the GPU's contour integrals (ssb_contour_moments_u8),
// without touching a mask: the list ssb_contours_host returns, in the same order.
//   * area / centroid: from the exact sums, with OpenCV's constants;
//   * order: OpenCV lists contours in pre-order of the border tree, siblings in reverse order of discovery.  The tree is
//     the enclosure tree of the table (outside background regions are transparent: what they hold is top level).  An
//     outer border is discovered at its region's first pixel.  A hole border is discovered at the pixel left of the
//     hole's first pixel UNLESS an earlier border left that pixel marked -- which can only change the order among the
//     holes of one region, and only matters if two of them are candidates: such a frame is reported with
//     n_candidates = -1 and goes through ssb_contours_host (so does a frame whose table is incomplete).
extern "C" int ssb_candidates_from_moments_host(int B, int max_regions, int roi_w, const int32_t *count, const int32_t *first,
                                                const int32_t *parent, const int32_t *flags, const int64_t *moments,
                                                double min_area, int max_candidates, int32_t *n_candidates,
                                                int32_t *cand_i32, double *cand_area, int threads) {
    using namespace ssb;
    if (B < 0 || max_regions < 1 || roi_w < 1 || max_candidates < 1) return fail(SSB_ERR_BAD_SHAPE, "bad shape");
    if (B == 0) return SSB_OK;
    if (!count || !first || !parent || !flags || !moments || !n_candidates || !cand_i32 || !cand_area)
        return fail(SSB_ERR_BAD_POINTER, "NULL buffer");
    const size_t R = static_cast<size_t>(max_regions);
    pool().run(B, threads, [&](int b) {
        const int n = count[b];
        if (n < 0 || n > max_regions) { n_candidates[b] = -1; return; }
        const int32_t *f = first + R * b, *par = parent + R * b, *fl = flags + R * b;
        const int64_t *mom = moments + R * 3 * b;
        thread_local std::vector<int> ord, node_parent, kids, kid_begin, stack;
        thread_local std::vector<char> cand;
        ord.resize(n);
        for (int i = 0; i < n; ++i) ord[i] = i;
        std::sort(ord.begin(), ord.end(), [&](int a, int c) { return f[a] < f[c]; });
        auto row_of = [&](int key) -> int {
            auto pos = std::lower_bound(ord.begin(), ord.end(), key, [&](int a, int k) { return f[a] < k; });
            return pos != ord.end() && f[*pos] == key ? *pos : -1;
        };
        auto outside = [&](int i) { return (fl[i] & 3) == 2; };
        auto is_hole = [&](int i) { return (fl[i] & 3) == 0; };
        auto area_of = [&](int i) { const double a = static_cast<double>(mom[3 * i]); return (a < 0 ? -a : a) * 0.5; };
        auto disc_key = [&](int i) { return is_hole(i) ? f[i] - 1 : f[i]; };     // where the raster scan meets the border
        for (int i = 0; i < n; ++i)                       // a border the GPU gave up on (longer than its step limit)
            if (mom[3 * i] == 0 && mom[3 * i + 1] == static_cast<int64_t>(0x8000000000000000ull)) { n_candidates[b] = -1; return; }
        // candidates (sem_seg_server.py:94, :117)
        cand.assign(n, 0);
        int n_cand = 0;
        for (int i = 0; i < n; ++i)
            if (!outside(i) && mom[3 * i] != 0 && area_of(i) >= min_area) { cand[i] = 1; ++n_cand; }
        if (n_cand == 0) { n_candidates[b] = 0; return; }
        // the border tree: parent node = enclosing region unless that is the outside (-1: top level)
        node_parent.assign(n, -1);
        for (int i = 0; i < n; ++i) {
            if (outside(i) || par[i] < 0) continue;
            const int j = row_of(par[i]);
            node_parent[i] = (j < 0 || outside(j)) ? -1 : j;
        }
        if (n_cand > 1) {      // two candidate holes of one region: their order depends on marks this path does not have
            for (int i = 0; i < n; ++i) {
                if (!cand[i] || !is_hole(i)) continue;
                for (int k = i + 1; k < n; ++k)
                    if (cand[k] && is_hole(k) && node_parent[k] == node_parent[i]) { n_candidates[b] = -1; return; }
            }
        }
        // children of every node (and of the top level, index n), newest (largest discovery key) first
        kids.clear();
        for (int i = 0; i < n; ++i)
            if (!outside(i)) kids.push_back(i);
        auto pkey = [&](int i) { return node_parent[i] < 0 ? n : node_parent[i]; };
        std::sort(kids.begin(), kids.end(), [&](int a, int c) {
            return pkey(a) != pkey(c) ? pkey(a) < pkey(c) : disc_key(a) > disc_key(c);
        });
        kid_begin.assign(n + 2, 0);
        for (int i : kids) ++kid_begin[pkey(i) + 1];
        for (int i = 0; i <= n; ++i) kid_begin[i + 1] += kid_begin[i];
        // pre-order
        int out_n = 0;
        stack.clear();
        for (int k = kid_begin[n + 1] - 1; k >= kid_begin[n]; --k) stack.push_back(kids[k]);
        while (!stack.empty()) {
            const int i = stack.back();
            stack.pop_back();
            if (cand[i]) {
                if (out_n < max_candidates) {
                    const long long a00i = mom[3 * i];
                    const double a00 = static_cast<double>(a00i);
                    const double db1_2 = a00i > 0 ? 0.5 : -0.5;
                    const double db1_6 = a00i > 0 ? 0.16666666666666666666666666666667 : -0.16666666666666666666666666666667;
                    const double m00 = a00 * db1_2, m10 = static_cast<double>(mom[3 * i + 1]) * db1_6,
                                 m01 = static_cast<double>(mom[3 * i + 2]) * db1_6;
                    int32_t *o = cand_i32 + (static_cast<size_t>(b) * max_candidates + out_n) * 4;
                    o[0] = is_hole(i); o[1] = f[i];
                    o[2] = static_cast<int32_t>(m10 / m00);
                    o[3] = static_cast<int32_t>(m01 / m00);
                    cand_area[static_cast<size_t>(b) * max_candidates + out_n] = area_of(i);
                }
                ++out_n;
            }
            for (int k = kid_begin[i + 1] - 1; k >= kid_begin[i]; --k) stack.push_back(kids[k]);
        }
        n_candidates[b] = out_n;
    });
    return SSB_OK;
}
