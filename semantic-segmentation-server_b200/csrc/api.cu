// Library-wide pieces of the C-ABI: version, thread-local status text, cached device attributes.
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

}  // namespace ssb

extern "C" int ssb_version(void) { return SSB_VERSION; }

extern "C" const char *ssb_last_error(void) { return ssb::last_error_buffer(); }
