// C-ABI entry points of the planar ([B,C,H,W]) argmax + per-label statistics kernels (argmax_planar.cuh has the
// design notes and the kernels).
#include "argmax_planar.cuh"

namespace ssb {
namespace {

template <typename P>
int argmax_stats_planar(const void *logits, int B, int H, int W, int C, int roi_h, int roi_w, uint8_t *labels,
                        uint64_t *stats, void *stream) {
    using E = typename P::Elem;
    if (int rc = check_argmax_args(logits, B, H, W, C, roi_h, roi_w, stats)) return rc;
    if (reinterpret_cast<uintptr_t>(logits) % P::kBytes != 0) return fail(SSB_ERR_BAD_POINTER, "logits not aligned to its element size");
    if (B == 0) return SSB_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    auto *stats64 = reinterpret_cast<unsigned long long *>(stats);
    int nw = P::kDefaultWarps;
    bool force_generic = false;
    if (const char *env = getenv("SSB_ARGMAX_NCHW_CFG")) {   // tuning knob (benchmarks only): warps per CTA, or "generic"
        if (!strcmp(env, "generic")) force_generic = true;
        else if (sscanf(env, "%d", &nw) != 1) return fail(SSB_ERR_LIMIT, "bad SSB_ARGMAX_NCHW_CFG '%s'", env);
    }
    // the unrolled kernel exists for the class counts of the two model families; everything else
    // (and frames too narrow for the packed counters) takes the generic kernel
    const bool tuned = C == 19 || C == 21;     // the class counts of the two model families: both compiled CTA shapes
    if (!tuned) nw = P::kDefaultWarps;
    const bool fast = !force_generic && C >= 2 && C <= kMaxPlanarC && W >= 2 && pick_flush_tiles(nw * 32 * P::kPx, W) >= 1;
    if (!fast) {
        SSB_CUDA_TRY(cudaMemsetAsync(stats64, 0, sizeof(uint64_t) * 4 * C * static_cast<size_t>(B), st));
        return launch_generic<E>(static_cast<const typename E::T *>(logits), B, H, W, C, roi_h, roi_w, labels, stats64, st,
                                 /*planar=*/true);
    }
    WorkspaceLease ws;
    if (int rc = workspace_acquire(scratch_bytes(B, C), st, &ws)) return rc;
    PlanarParams p{};
    p.logits = logits;
    p.labels = labels;
    p.stats = stats64;
    p.B = B;
    p.hw = H * W;
    p.w = W;
    p.roi_h = roi_h;
    p.roi_w = roi_w;
    p.pairs_ok = reinterpret_cast<uintptr_t>(logits) % 4 == 0;
    p.acc = static_cast<unsigned long long *>(ws.ptr);
    p.frame_px = reinterpret_cast<unsigned int *>(static_cast<char *>(ws.ptr) + sizeof(uint64_t) * 4 * C * static_cast<size_t>(B));
    p.divw = make_fastdiv(static_cast<uint32_t>(W));
    int rc;
    if (C == 21) rc = dispatch_planar<P, 21>(p, st, nw);
    else if (C == 19) rc = dispatch_planar<P, 19>(p, st, nw);
    else {
        PlanarArgs a{p.logits, p.labels, p.stats, p.B, p.hw, p.w, p.roi_h, p.roi_w, p.pairs_ok, p.acc, p.frame_px, p.divw};
        rc = P::kBytes == 4 ? launch_planar_f32_other_c(C, a, st) : launch_planar_f16_other_c(C, a, st);
    }
    const int rc2 = workspace_release(ws, st, rc == SSB_OK);
    return rc ? rc : rc2;
}

}  // namespace
}  // namespace ssb

extern "C" int ssb_argmax_stats_nchw_f32(const float *logits, int B, int H, int W, int C, int roi_h, int roi_w,
                                         uint8_t *labels, uint64_t *stats, void *stream) {
    return ssb::argmax_stats_planar<ssb::PlaneF32>(logits, B, H, W, C, roi_h, roi_w, labels, stats, stream);
}

extern "C" int ssb_argmax_stats_nchw_f16(const uint16_t *logits, int B, int H, int W, int C, int roi_h, int roi_w,
                                         uint8_t *labels, uint64_t *stats, void *stream) {
    return ssb::argmax_stats_planar<ssb::PlaneF16>(logits, B, H, W, C, roi_h, roi_w, labels, stats, stream);
}
