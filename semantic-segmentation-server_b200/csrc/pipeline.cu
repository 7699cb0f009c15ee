// Host-buffer pipeline: the reference holds frames and model outputs in HOST memory
// (cap.read() / engine.run_inference return numpy arrays, sem_seg_server.py:146,162).  This wraps
// the device-pointer entry points for such callers: chunks of frames go host -> device on one of
// `n_slots` streams, are reduced there, and only the label map (optional) and the 32*C bytes of
// statistics per frame come back.  Copies of chunk i+1 overlap the kernel of chunk i.
#include <new>

#include "ssb_common.cuh"

struct ssb_pipeline {
    int device, H, W, C, elem_bytes, max_frames, n_slots;
    size_t frame_in_bytes, frame_px;
    struct Slot {
        cudaStream_t stream;
        void *d_logits;
        uint8_t *d_labels;
        uint64_t *d_stats;
    } *slots;
};

extern "C" int ssb_pipeline_destroy(ssb_pipeline *p) {
    using namespace ssb;
    if (!p) return SSB_OK;
    int prev = 0;
    cudaGetDevice(&prev);
    cudaSetDevice(p->device);
    if (p->slots) {
        for (int i = 0; i < p->n_slots; ++i) {
            if (p->slots[i].stream) cudaStreamSynchronize(p->slots[i].stream);
            cudaFree(p->slots[i].d_logits);
            cudaFree(p->slots[i].d_labels);
            cudaFree(p->slots[i].d_stats);
            if (p->slots[i].stream) cudaStreamDestroy(p->slots[i].stream);
        }
        delete[] p->slots;
    }
    delete p;
    cudaSetDevice(prev);
    return SSB_OK;
}

extern "C" int ssb_pipeline_create(ssb_pipeline **out, int H, int W, int C, int elem_bytes, int max_frames_per_chunk,
                                   int n_slots) {
    using namespace ssb;
    if (!out) return fail(SSB_ERR_BAD_POINTER, "out is NULL");
    *out = nullptr;
    if (H <= 0 || W <= 0 || C < 2 || C > 255 || (elem_bytes != 1 && elem_bytes != 2) || max_frames_per_chunk < 1 ||
        n_slots < 1 || n_slots > 16)
        return fail(SSB_ERR_BAD_SHAPE, "bad pipeline geometry");
    ssb_pipeline *p = new (std::nothrow) ssb_pipeline();
    if (!p) return fail(SSB_ERR_LIMIT, "out of host memory");
    p->H = H; p->W = W; p->C = C; p->elem_bytes = elem_bytes;
    p->max_frames = max_frames_per_chunk; p->n_slots = n_slots;
    p->frame_px = static_cast<size_t>(H) * W;
    p->frame_in_bytes = p->frame_px * C * elem_bytes;
    p->slots = new (std::nothrow) ssb_pipeline::Slot[n_slots]();
    cudaError_t e = cudaGetDevice(&p->device);
    for (int i = 0; e == cudaSuccess && p->slots && i < n_slots; ++i) {
        auto &s = p->slots[i];
        e = cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaMalloc(&s.d_logits, p->frame_in_bytes * max_frames_per_chunk);
        if (e == cudaSuccess) e = cudaMalloc(&s.d_labels, p->frame_px * max_frames_per_chunk);
        if (e == cudaSuccess) e = cudaMalloc(&s.d_stats, sizeof(uint64_t) * 4 * C * max_frames_per_chunk);
    }
    if (e != cudaSuccess || !p->slots) {
        ssb_pipeline_destroy(p);
        return fail(SSB_ERR_CUDA, "pipeline allocation failed: %s", cudaGetErrorString(e));
    }
    *out = p;
    return SSB_OK;
}

extern "C" int ssb_pipeline_argmax_stats_host(ssb_pipeline *p, const void *logits_host, int B, int roi_h, int roi_w,
                                              uint8_t *labels_host, uint64_t *stats_host) {
    using namespace ssb;
    if (!p) return fail(SSB_ERR_BAD_POINTER, "pipeline is NULL");
    if (B < 0) return fail(SSB_ERR_BAD_SHAPE, "B=%d", B);
    if (B == 0) return SSB_OK;
    if (!logits_host || !stats_host) return fail(SSB_ERR_BAD_POINTER, "logits_host/stats_host must not be NULL");
    int dev = 0;
    SSB_CUDA_TRY(cudaGetDevice(&dev));
    if (dev != p->device) return fail(SSB_ERR_BAD_SHAPE, "pipeline belongs to device %d, current is %d", p->device, dev);
    const size_t stats_frame = sizeof(uint64_t) * 4 * p->C;
    int chunk = 0;
    for (int b0 = 0; b0 < B; b0 += p->max_frames, ++chunk) {
        const int nb = B - b0 < p->max_frames ? B - b0 : p->max_frames;
        auto &s = p->slots[chunk % p->n_slots];
        // stream order makes the slot's buffers safe to reuse: the previous chunk's D2H copies are
        // ahead of this H2D copy on the same stream
        const char *src = static_cast<const char *>(logits_host) + p->frame_in_bytes * b0;
        SSB_CUDA_TRY(cudaMemcpyAsync(s.d_logits, src, p->frame_in_bytes * nb, cudaMemcpyHostToDevice, s.stream));
        int rc;
        if (p->elem_bytes == 1)
            rc = ssb_argmax_stats_u8(static_cast<const uint8_t *>(s.d_logits), nb, p->H, p->W, p->C, roi_h, roi_w,
                                     labels_host ? s.d_labels : nullptr, s.d_stats, s.stream);
        else
            rc = ssb_argmax_stats_f16(static_cast<const uint16_t *>(s.d_logits), nb, p->H, p->W, p->C, roi_h, roi_w,
                                      labels_host ? s.d_labels : nullptr, s.d_stats, s.stream);
        if (rc) return rc;
        SSB_CUDA_TRY(cudaMemcpyAsync(reinterpret_cast<char *>(stats_host) + stats_frame * b0, s.d_stats,
                                     stats_frame * nb, cudaMemcpyDeviceToHost, s.stream));
        if (labels_host)
            SSB_CUDA_TRY(cudaMemcpyAsync(labels_host + p->frame_px * b0, s.d_labels, p->frame_px * nb,
                                         cudaMemcpyDeviceToHost, s.stream));
    }
    for (int i = 0; i < p->n_slots; ++i) SSB_CUDA_TRY(cudaStreamSynchronize(p->slots[i].stream));
    return SSB_OK;
}
