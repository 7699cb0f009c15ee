// Library-wide pieces of the C-ABI: version, thread-local status text, cached device attributes.
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
    int dev = 0;
    SSB_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(SSB_ERR_LIMIT, "device ordinal %d not supported", dev);
    if (!have[dev]) {
        DeviceInfo di;
        SSB_CUDA_TRY(cudaDeviceGetAttribute(&di.sm_count, cudaDevAttrMultiProcessorCount, dev));
        SSB_CUDA_TRY(cudaDeviceGetAttribute(&di.max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
        cache[dev] = di;  // benign race: every thread writes the same values
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
    bool used = false;
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
    Slot slots[kSlots];
    unsigned next = 0;
    CaptureSlot capture[kCaptureSlots];
    size_t capture_bytes = 0;
};
Pool g_pools[kMaxDevices];

int alloc_zeroed(void **ptr, size_t bytes) {
    SSB_CUDA_TRY(cudaMalloc(ptr, bytes));
    SSB_CUDA_TRY(cudaMemset(*ptr, 0, bytes));
    return SSB_OK;
}
}  // namespace

int workspace_acquire(size_t bytes, cudaStream_t stream, WorkspaceLease *out) {
    int dev = 0;
    SSB_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return fail(SSB_ERR_LIMIT, "device ordinal %d not supported", dev);
    Pool &pool = g_pools[dev];
    std::lock_guard<std::mutex> lock(pool.mu);
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
            if (int rc = alloc_zeroed(&c.ptr, want)) return rc;
            c.bytes = want;
        }
        pool.capture_bytes = want;
    }
    const int idx = static_cast<int>(pool.next++ % kSlots);
    Slot &s = pool.slots[idx];
    if (!s.done) SSB_CUDA_TRY(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
    if (s.bytes < bytes) {
        if (s.used) SSB_CUDA_TRY(cudaEventSynchronize(s.done));  // nobody may still be using the old block
        if (s.ptr) SSB_CUDA_TRY(cudaFree(s.ptr));
        s.ptr = nullptr;
        s.bytes = 0;
        if (int rc = alloc_zeroed(&s.ptr, want)) return rc;
        s.bytes = want;
        s.used = false;
    }
    if (s.used && s.last_stream != stream) SSB_CUDA_TRY(cudaStreamWaitEvent(stream, s.done, 0));
    out->ptr = s.ptr;
    out->dev = dev;
    out->slot = idx;
    return SSB_OK;
}

int workspace_release(const WorkspaceLease &lease, cudaStream_t stream) {
    if (lease.slot < 0) return SSB_OK;   // block owned by a recorded call: ordered by the graph itself
    Pool &pool = g_pools[lease.dev];
    std::lock_guard<std::mutex> lock(pool.mu);
    Slot &s = pool.slots[lease.slot];
    SSB_CUDA_TRY(cudaEventRecord(s.done, stream));
    s.last_stream = stream;
    s.used = true;
    return SSB_OK;
}

}  // namespace ssb

extern "C" int ssb_version(void) { return SSB_VERSION; }

extern "C" const char *ssb_last_error(void) { return ssb::last_error_buffer(); }
