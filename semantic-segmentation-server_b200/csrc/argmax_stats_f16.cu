// fp16 entry point of the fused argmax + per-label statistics (Cityscapes-shaped config:
// 1024x2048x19 fp16).  Same contract as the u8 path (ssb200.h); SQ is the exact fixed-point sum
// value * 2^24 of the winning logits.
#include "argmax_generic.cuh"

extern "C" int ssb_argmax_stats_f16(const uint16_t *logits, int B, int H, int W, int C, int roi_h, int roi_w,
                                    uint8_t *labels, uint64_t *stats, void *stream) {
    using namespace ssb;
    if (int rc = check_argmax_args(logits, B, H, W, C, roi_h, roi_w, stats)) return rc;
    if (B == 0) return SSB_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    auto *stats64 = reinterpret_cast<unsigned long long *>(stats);
    SSB_CUDA_TRY(cudaMemsetAsync(stats64, 0, sizeof(uint64_t) * 4 * C * static_cast<size_t>(B), st));
    return launch_generic<ElemF16>(logits, B, H, W, C, roi_h, roi_w, labels, stats64, st);
}
