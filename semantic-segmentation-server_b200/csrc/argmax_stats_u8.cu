// u8 codec + C-ABI entry points of the fused argmax + per-label statistics kernel
// (argmax_stream.cuh has the design notes and the kernel itself).
//
// Argmax on quantised u8 scores: a lane owns 4 consecutive pixels = C aligned 32-bit words of the
// tile (stride C words between lanes: conflict-free for odd C).  Scores are packed with their
// channel into 16-bit keys (value<<8 | 255-channel), two channels per PRMT, and reduced with
// VIMNMX3.U16x2: the maximum key is the maximum value with the smallest channel on ties.
#include <mutex>


#include "codec_u8.cuh"

namespace ssb {
namespace {

template <int C>
int dispatch_cfg(const ArgmaxParams &p, cudaStream_t st, int nw, int stg) {
    if (nw == 4 && stg == 2) return launch_stream<CodecU8<C>, 4, 2>(p, st);
    if (nw == 4 && stg == 3) return launch_stream<CodecU8<C>, 4, 3>(p, st);
    if (nw == 8 && stg == 2) return launch_stream<CodecU8<C>, 8, 2>(p, st);
    if (nw == 4 && stg == 4) return launch_stream<CodecU8<C>, 4, 4>(p, st);
    if (nw == 8 && stg == 3) return launch_stream<CodecU8<C>, 8, 3>(p, st);
    return fail(SSB_ERR_LIMIT, "no kernel compiled for NWARPS=%d STAGES=%d", nw, stg);
}

// Best of the compiled set on B200 (tools/sweep_argmax.py): 4 consumer warps, 2 stages = 46 KB per CTA,
// 4 CTAs per SM, 86 KB in flight per SM.  Deeper rings or more bytes in flight are slower.
constexpr int kDefaultWarps = 4, kDefaultStages = 2;

}  // namespace
}  // namespace ssb

extern "C" int ssb_argmax_stats_u8(const uint8_t *logits, int B, int H, int W, int C, int roi_h, int roi_w,
                                   uint8_t *labels, uint64_t *stats, void *stream) {
    using namespace ssb;
    if (int rc = check_argmax_args(logits, B, H, W, C, roi_h, roi_w, stats)) return rc;
    if (B == 0) return SSB_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    auto *stats64 = reinterpret_cast<unsigned long long *>(stats);
    // tuning knob (benchmarks only): SSB_ARGMAX_CFG="<consumer warps>,<stages>" or "generic"
    int nw = kDefaultWarps, stg = kDefaultStages;
    bool force_generic = false;
    if (const char *env = getenv("SSB_ARGMAX_CFG")) {
        if (!strcmp(env, "generic")) force_generic = true;
        else if (sscanf(env, "%d,%d", &nw, &stg) != 2) return fail(SSB_ERR_LIMIT, "bad SSB_ARGMAX_CFG '%s'", env);
    }
    const long long total_px = static_cast<long long>(B) * H * W;
    const bool tuned = C == 19 || C == 21;     // the class counts of the two model families: every compiled ring shape
    if (!tuned) { nw = kDefaultWarps; stg = kDefaultStages; }
    const bool fast = !force_generic && C >= 2 && C <= kMaxStreamC && stream_path_ok(logits, labels, W, total_px, nw);
    if (!fast) {
        SSB_CUDA_TRY(cudaMemsetAsync(stats64, 0, sizeof(uint64_t) * 4 * C * static_cast<size_t>(B), st));
        return launch_generic<ElemU8>(logits, B, H, W, C, roi_h, roi_w, labels, stats64, st);
    }
    // scratch: [B][C][4] u64 accumulators + [B] u32 per-frame pixel counters, zero at rest
    WorkspaceLease ws;
    if (int rc = workspace_acquire(scratch_bytes(B, C), st, &ws)) return rc;
    const ArgmaxParams p = make_params(logits, B, H, W, C, roi_h, roi_w, labels, stats64, ws.ptr);
    const int rc = C == 21 ? dispatch_cfg<21>(p, st, nw, stg) : C == 19 ? dispatch_cfg<19>(p, st, nw, stg)
                                                                          : launch_stream_u8_other_c(C, p, st);
    const int rc2 = workspace_release(ws, st, rc == SSB_OK);
    return rc ? rc : rc2;
}

// ---- entry point that never allocates ------------------------------------------------------------
// Same computation as ssb_argmax_stats_u8, but the streaming kernel's scratch (the zero-at-rest
// [B][C][4] accumulators and [B] pixel counters) lives in a fixed array inside the loaded module
// instead of coming from the cudaMalloc-backed workspace pool.  See include/ssb200.h for who needs
// this and what it costs.
namespace ssb {
namespace {

constexpr int kNoAllocMaxFrames = SSB_NOALLOC_MAX_FRAMES;
constexpr int kNoAllocMaxClasses = 21;  // the classes the streaming kernel is compiled for: 19 and 21
constexpr size_t kNoAllocAccBytes = sizeof(uint64_t) * 4 * kNoAllocMaxClasses * kNoAllocMaxFrames;
constexpr size_t kNoAllocBytes = kNoAllocAccBytes + sizeof(unsigned int) * kNoAllocMaxFrames;

// Resident on every GPU that loads the library until the process exits (~173 KB).
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

// In host code `g_noalloc_scratch` is only a handle for the symbol, not something a kernel can
// dereference: the device address has to be asked for, per device, with that device current.
int noalloc_scratch_address(unsigned char **out) {
    static unsigned char *cached[64] = {nullptr};
    static std::mutex mu;
    int dev = 0;
    SSB_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(SSB_ERR_LIMIT, "device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (!cached[dev]) {
        void *ptr = nullptr;
        SSB_CUDA_TRY(cudaGetSymbolAddress(&ptr, g_noalloc_scratch));
        cached[dev] = static_cast<unsigned char *>(ptr);
    }
    *out = cached[dev];
    return SSB_OK;
}

}  // namespace
}  // namespace ssb

extern "C" int ssb_argmax_stats_u8_noalloc(const uint8_t *logits, int B, int H, int W, int C, int roi_h, int roi_w,
                                           uint8_t *labels, uint64_t *stats, void *stream) {
    using namespace ssb;
    if (int rc = check_argmax_args(logits, B, H, W, C, roi_h, roi_w, stats)) return rc;
    if (B > kNoAllocMaxFrames)
        return fail(SSB_ERR_LIMIT, "ssb_argmax_stats_u8_noalloc takes at most %d frames per call (got %d); split the batch",
                    kNoAllocMaxFrames, B);
    if (B == 0) return SSB_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    auto *stats64 = reinterpret_cast<unsigned long long *>(stats);
    const int nw = kDefaultWarps, stg = kDefaultStages;
    const long long total_px = static_cast<long long>(B) * H * W;
    if (!((C == 19 || C == 21) && stream_path_ok(logits, labels, W, total_px, nw))) {
        // shapes the streaming kernel does not cover use the generic kernel, which needs no scratch
        SSB_CUDA_TRY(cudaMemsetAsync(stats64, 0, sizeof(uint64_t) * 4 * C * static_cast<size_t>(B), st));
        return launch_generic<ElemU8>(logits, B, H, W, C, roi_h, roi_w, labels, stats64, st);
    }
    unsigned char *scratch = nullptr;
    if (int rc = noalloc_scratch_address(&scratch)) return rc;
    // The array is shared by every call and still holds whatever the previous call left (zeros if
    // it ran to completion, anything if it was cut short): start from a clean slate.
    SSB_CUDA_TRY(cudaMemsetAsync(scratch, 0, scratch_bytes(B, C), st));
    const ArgmaxParams p = make_params(logits, B, H, W, C, roi_h, roi_w, labels, stats64, scratch);
    return C == 21 ? dispatch_cfg<21>(p, st, nw, stg) : dispatch_cfg<19>(p, st, nw, stg);
}

// Where the fixed array lives on the current device and how large it is (no allocation, no launch).
extern "C" int ssb_noalloc_scratch_info(void **device_ptr, size_t *bytes) {
    using namespace ssb;
    unsigned char *scratch = nullptr;
    if (int rc = noalloc_scratch_address(&scratch)) return rc;
    if (device_ptr) *device_ptr = scratch;
    if (bytes) *bytes = kNoAllocBytes;
    return SSB_OK;
}
