// Library-wide pieces of the C-ABI: version, thread-local status text, cached device attributes.
#include <condition_variable>
#include <mutex>

#include "ssb_common.cuh"

namespace ssb {

char *last_error_buffer() {
    static thread_local char buf[kErrLen] = {0};
    return buf;
}

int device_info(DeviceInfo *out) {
    static DeviceInfo cache[64];
    static bool have[64] = {false};
    static std::mutex mu;
    int dev = 0;
    SSB_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(SSB_ERR_LIMIT, "device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (!have[dev]) {
        DeviceInfo di;
        SSB_CUDA_TRY(cudaDeviceGetAttribute(&di.sm_count, cudaDevAttrMultiProcessorCount, dev));
        SSB_CUDA_TRY(cudaDeviceGetAttribute(&di.max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
        cache[dev] = di;
        have[dev] = true;
    }
    *out = cache[dev];
    return SSB_OK;
}

namespace {
constexpr int kMaxDevices = 64, kSlots = 8, kCaptureSlots = 16;
struct Slot {
    void *ptr = nullptr;
    size_t bytes = 0;
    cudaEvent_t done = nullptr;
    cudaStream_t last_stream = nullptr;
    bool used = false;     // `done` has been recorded at least once
    bool leased = false;   // handed out by workspace_acquire, not yet returned by workspace_release
    bool dirty = false;    // the last launch on it failed: contents unknown, re-zero before the next use
};
// Scratch for calls recorded into a CUDA graph.  Such a call bakes its scratch pointer into the graph and
// may be replayed at any later time, so it gets a block of its own for good (never shared with eager
// calls or with another recorded call).  Nothing can be allocated while a stream is being captured:
// the blocks that are still free are (re)sized by EAGER calls, to the largest request seen so far.
struct CaptureSlot {
    void *ptr = nullptr;
    size_t bytes = 0;
    bool owned = false;
};
struct Pool {
    std::mutex mu;
    std::condition_variable freed;   // a leased slot came back
    Slot slots[kSlots];
    unsigned next = 0;
    CaptureSlot capture[kCaptureSlots];
    size_t capture_bytes = 0;
};
Pool g_pools[kMaxDevices];

// A fresh block is cleared on the CALLER's stream and the clear is waited for: the block may be used next
// by any stream (the non-blocking ones of ssb_pipeline and of torch do not order against the legacy
// default stream, so a cudaMemset there could still be in flight when the first kernel reads the block).
// Only runs when the pool grows.
int alloc_zeroed(void **ptr, size_t bytes, cudaStream_t stream) {
    SSB_CUDA_TRY(cudaMalloc(ptr, bytes));
    SSB_CUDA_TRY(cudaMemsetAsync(*ptr, 0, bytes, stream));
    SSB_CUDA_TRY(cudaStreamSynchronize(stream));
    return SSB_OK;
}
}  // namespace

int workspace_acquire(size_t bytes, cudaStream_t stream, WorkspaceLease *out) {
    int dev = 0;
    SSB_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return fail(SSB_ERR_LIMIT, "device ordinal %d not supported", dev);
    Pool &pool = g_pools[dev];
    std::unique_lock<std::mutex> lock(pool.mu);
    cudaStreamCaptureStatus capturing = cudaStreamCaptureStatusNone;
    SSB_CUDA_TRY(cudaStreamIsCapturing(stream, &capturing));
    if (capturing != cudaStreamCaptureStatusNone) {
        for (int i = 0; i < kCaptureSlots; ++i) {
            CaptureSlot &c = pool.capture[i];
            if (!c.owned && c.bytes >= bytes) {
                c.owned = true;   // for the life of the process: the graph may be replayed at any time
                out->ptr = c.ptr;
                out->dev = dev;
                out->slot = -1;
                return SSB_OK;
            }
        }
        return fail(SSB_ERR_LIMIT,
                    "no scratch block of %zu bytes for a call recorded into a CUDA graph: run the same call once outside "
                    "stream capture first (at most %d recorded calls per device)", bytes, kCaptureSlots);
    }
    const size_t want = (bytes + 65535) / 65536 * 65536;
    if (pool.capture_bytes < want) {   // keep the free capture blocks as large as the largest eager request
        for (int i = 0; i < kCaptureSlots; ++i) {
            CaptureSlot &c = pool.capture[i];
            if (c.owned) continue;
            if (c.ptr) SSB_CUDA_TRY(cudaFree(c.ptr));
            c.ptr = nullptr;
            c.bytes = 0;
            if (int rc = alloc_zeroed(&c.ptr, want, stream)) return rc;
            c.bytes = want;
        }
        pool.capture_bytes = want;
    }
    // A slot is the caller's alone from here until workspace_release: leased slots are never handed out,
    // resized or freed.  Preference: the slot this stream used last (no cross-stream wait), else the next
    // free one in ring order; with every slot leased (more than kSlots threads between acquire and
    // release) the caller waits for one to come back -- a lease only spans a kernel launch.
    int idx = -1;
    while (idx < 0) {
        for (int i = 0; i < kSlots && idx < 0; ++i)
            if (!pool.slots[i].leased && pool.slots[i].used && pool.slots[i].last_stream == stream) idx = i;
        for (int i = 0; i < kSlots && idx < 0; ++i) {
            const int j = static_cast<int>((pool.next + i) % kSlots);
            if (!pool.slots[j].leased) {
                idx = j;
                pool.next = j + 1;
            }
        }
        if (idx < 0) pool.freed.wait(lock);
    }
    Slot &s = pool.slots[idx];
    if (!s.done) SSB_CUDA_TRY(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
    if (s.bytes < bytes) {
        if (s.used) SSB_CUDA_TRY(cudaEventSynchronize(s.done));  // nobody may still be using the old block
        if (s.ptr) SSB_CUDA_TRY(cudaFree(s.ptr));
        s.ptr = nullptr;
        s.bytes = 0;
        if (int rc = alloc_zeroed(&s.ptr, want, stream)) return rc;
        s.bytes = want;
        s.used = false;
        s.dirty = false;
    }
    if (s.used && s.last_stream != stream) SSB_CUDA_TRY(cudaStreamWaitEvent(stream, s.done, 0));
    if (s.dirty) {   // a launch on this block failed: it may not be zero any more
        SSB_CUDA_TRY(cudaMemsetAsync(s.ptr, 0, s.bytes, stream));
        s.dirty = false;
    }
    s.leased = true;
    out->ptr = s.ptr;
    out->dev = dev;
    out->slot = idx;
    return SSB_OK;
}

int workspace_release(const WorkspaceLease &lease, cudaStream_t stream, bool launch_ok) {
    if (lease.slot < 0) return SSB_OK;   // block owned by a recorded call: ordered by the graph itself
    Pool &pool = g_pools[lease.dev];
    int rc = SSB_OK;
    {
        std::lock_guard<std::mutex> lock(pool.mu);
        Slot &s = pool.slots[lease.slot];
        const cudaError_t e = cudaEventRecord(s.done, stream);
        if (e != cudaSuccess) rc = fail(SSB_ERR_CUDA, "cudaEventRecord failed: %s", cudaGetErrorString(e));
        s.last_stream = stream;
        s.used = true;
        s.leased = false;
        if (!launch_ok || e != cudaSuccess) s.dirty = true;
    }
    pool.freed.notify_one();
    return rc;
}

}  // namespace ssb

extern "C" int ssb_version(void) { return SSB_VERSION; }

extern "C" const char *ssb_last_error(void) { return ssb::last_error_buffer(); }
