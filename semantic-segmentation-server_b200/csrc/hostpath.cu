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
#include <atomic>
#include <condition_variable>
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

HostPool &pool() {
    static HostPool *p = new HostPool();   // never destroyed: its detached workers may outlive static destruction
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
    auto send = [&](int i) {
        const uint8_t *src = frames_host[i];
        if (!direct[i]) {
            std::memcpy(staging_host + frame_bytes * i, src, frame_bytes);
            src = staging_host + frame_bytes * i;
        }
        if (cudaMemcpyAsync(dst_device + frame_bytes * i, src, frame_bytes, cudaMemcpyHostToDevice, st) != cudaSuccess)
            failed.store(1);
    };
    int dev = 0;
    SSB_CUDA_TRY(cudaGetDevice(&dev));
    // direct frames first (no host work), then the gathered ones as their copies finish
    for (int i = 0; i < n; ++i)
        if (direct[i]) send(i);
    if (n_staged) {
        std::vector<int> todo;
        for (int i = 0; i < n; ++i)
            if (!direct[i]) todo.push_back(i);
        pool().run(static_cast<int>(todo.size()), threads, [&](int j) {
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
