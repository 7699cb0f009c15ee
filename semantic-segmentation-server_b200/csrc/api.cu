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
constexpr int kMaxDevices = 64, kSlots = 8;
struct Slot {
    void *ptr = nullptr;
    size_t bytes = 0;
    cudaEvent_t done = nullptr;
    cudaStream_t last_stream = nullptr;
    bool used = false;
};
struct Pool {
    std::mutex mu;
    Slot slots[kSlots];
    unsigned next = 0;
};
Pool g_pools[kMaxDevices];
}  // namespace

int workspace_acquire(size_t bytes, cudaStream_t stream, WorkspaceLease *out) {
    int dev = 0;
    SSB_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return fail(SSB_ERR_LIMIT, "device ordinal %d not supported", dev);
    Pool &pool = g_pools[dev];
    std::lock_guard<std::mutex> lock(pool.mu);
    const int idx = static_cast<int>(pool.next++ % kSlots);
    Slot &s = pool.slots[idx];
    if (!s.done) SSB_CUDA_TRY(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
    if (s.bytes < bytes) {
        if (s.used) SSB_CUDA_TRY(cudaEventSynchronize(s.done));  // nobody may still be using the old block
        if (s.ptr) SSB_CUDA_TRY(cudaFree(s.ptr));
        s.ptr = nullptr;
        s.bytes = 0;
        const size_t want = (bytes + 65535) / 65536 * 65536;
        SSB_CUDA_TRY(cudaMalloc(&s.ptr, want));
        SSB_CUDA_TRY(cudaMemset(s.ptr, 0, want));
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
