// Shared host/device helpers for libssb200: status reporting, launch checks, and the sm_100a
// PTX wrappers (mbarrier + bulk async copy) the streaming kernels use.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>

#include "ssb200.h"

namespace ssb {

// ---- status -------------------------------------------------------------------------------
char *last_error_buffer();  // thread-local, defined in api.cu
constexpr int kErrLen = 512;

inline int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(last_error_buffer(), kErrLen, fmt, ap);
    va_end(ap);
    return code;
}

#define SSB_CUDA_TRY(expr)                                                                   \
    do {                                                                                     \
        cudaError_t e__ = (expr);                                                            \
        if (e__ != cudaSuccess)                                                              \
            return ::ssb::fail(SSB_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e__)); \
    } while (0)

struct DeviceInfo {
    int sm_count;
    int max_smem_optin;
};
// Cached per device (the only global state the library keeps); SSB_ERR_CUDA if no device.
int device_info(DeviceInfo *out);

// ---- scratch workspaces ------------------------------------------------------------------------
// Device scratch that is ZERO AT REST (kernels leave it zeroed).  A small per-device pool, grown on
// demand with cudaMalloc (first use, or a larger request than seen before).  A slot belongs to one
// caller between acquire and release (other callers take another slot, or wait when all are out), and a
// slot that was last used on another stream is ordered behind that use with an event, so concurrent
// streams never share live scratch.  `launch_ok = false` on release (the kernel could not be launched)
// marks the block for re-zeroing before its next use.  A call made while its stream is being captured into a CUDA graph gets a block of its
// own that no other call ever uses (slot == -1 in the lease); those blocks are sized by earlier eager
// calls, so the same call must have run once outside capture.
struct WorkspaceLease {
    void *ptr;
    int dev, slot;
};
int workspace_acquire(size_t bytes, cudaStream_t stream, WorkspaceLease *out);
int workspace_release(const WorkspaceLease &lease, cudaStream_t stream, bool launch_ok = true);

// ---- programmatic dependent launch ---------------------------------------------------------------
// Every kernel of the library waits for the kernels it depends on (`pdl_wait`) only after the part of its
// prologue that needs no global memory, which hides the ~1.5 us between two dependent launches;
// semantics are those of ordinary stream order.  Only the SINGLE-WAVE kernels (the argmax kernels: one
// persistent CTA set that fills the GPU) also call `pdl_trigger` at their top, so that the next kernel's
// CTAs are placed as theirs retire.  Kernels with many waves of small CTAs (pre-process, mask, regions)
// must not: a dependent grid becomes schedulable while most of their CTAs have not started yet, and its
// CTAs then sit in `pdl_wait` on SM resources the remaining waves need (measured: 256-frame pre-process
// followed by the argmax kernel, 410 us with the early trigger against 330 us without).
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    const char *no_pdl = getenv("SSB_NO_PDL");   // tuning knob: "1" = plain stream-ordered launches
    attr[0].val.programmaticStreamSerializationAllowed = (no_pdl && no_pdl[0] == '1') ? 0 : 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

// ---- device-side PTX wrappers ----------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ uint32_t smem_addr(const void *p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}

// Makes freshly initialised barriers visible to the async (TMA) proxy.
__device__ __forceinline__ void mbar_init_fence() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes)
                 : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t"
        "}" ::"r"(smem_addr(bar)),
        "r"(parity)
        : "memory");
}

// L2 eviction policy for data that is read exactly once.
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}

// 1-D TMA bulk copy global -> shared (SASS: UBLKCP).  dst/src 16-byte aligned, bytes % 16 == 0.
__device__ __forceinline__ void bulk_copy_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar,
                                              uint64_t policy) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::
            "r"(smem_addr(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_addr(bar)), "l"(policy)
        : "memory");
}

// The same copy without a cache hint (data that other kernels read again).
__device__ __forceinline__ void bulk_copy_g2s_plain(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
                     "r"(smem_addr(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_addr(bar))
                 : "memory");
}

// Asks L2 for [p, p + bytes) ahead of the loads that will want it (16-byte aligned address and size).
__device__ __forceinline__ void bulk_prefetch_l2(const void *src_gmem, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src_gmem), "r"(bytes) : "memory");
}

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

__device__ __forceinline__ void st_global_cs_u32(void *p, uint32_t v) {
    asm volatile("st.global.cs.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
#endif  // __CUDACC__

// ---- exact division by a runtime constant (row width) for 31-bit dividends ------------------
// y = r / d for 0 <= r < 2^31, 2 <= d < 2^31:  m = ceil(2^(31+L) / d), L = ceil(log2 d);
// y = umulhi(r, m) >> (L - 1).   (Round-up method; m < 2^32 because 2^(L-1) < d.)
struct FastDiv {
    uint32_t mul;
    uint32_t shift;
    uint32_t d;
};
inline FastDiv make_fastdiv(uint32_t d) {
    FastDiv f;
    uint32_t L = 0;
    while ((1ull << L) < d) ++L;
    f.mul = static_cast<uint32_t>(((1ull << (31 + L)) + d - 1) / d);
    f.shift = L - 1;
    f.d = d;
    return f;
}
#ifdef __CUDACC__
__device__ __forceinline__ uint32_t fastdiv(uint32_t r, const FastDiv &f) { return __umulhi(r, f.mul) >> f.shift; }
#endif

}  // namespace ssb
