// Host-buffer pipeline: the reference holds frames and model outputs in HOST memory
// (cap.read() / engine.run_inference return numpy arrays, sem_seg_server.py:146,162).  This wraps
// the device-pointer entry points for such callers: chunks of frames go host -> device on one of
// `n_slots` streams, are reduced there, and only the label map (optional) and the 32*C bytes of
// statistics per frame come back.  Copies of chunk i+1 overlap the kernel of chunk i.
#include <new>

#include "ssb_common.cuh"

struct ssb_pipeline {
    int device, H, W, C, elem_bytes, max_frames, n_slots;
    int cam_h, cam_w;            // > 0: ingest pipeline (camera frames -> model input), else logits pipeline
    size_t frame_in_bytes, frame_px;
    int32_t *d_xtab, *d_ytab;    // ingest only: device copies of the index tables of the last call
    cudaEvent_t tables;          // ingest only: "tables are on the device", recorded on slot 0's stream per call
    struct Slot {
        cudaStream_t stream;
        void *d_logits;          // logits chunk, or camera frames chunk (ingest)
        uint8_t *d_labels;       // label maps, or model inputs (ingest)
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
    cudaFree(p->d_xtab);
    cudaFree(p->d_ytab);
    if (p->tables) cudaEventDestroy(p->tables);
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
    int rc = SSB_OK, chunk = 0;
    auto cuda_ok = [&](cudaError_t e, const char *what) {
        if (e != cudaSuccess) rc = fail(SSB_ERR_CUDA, "%s failed: %s", what, cudaGetErrorString(e));
        return e == cudaSuccess;
    };
    for (int b0 = 0; b0 < B && rc == SSB_OK; b0 += p->max_frames, ++chunk) {
        const int nb = B - b0 < p->max_frames ? B - b0 : p->max_frames;
        auto &s = p->slots[chunk % p->n_slots];
        // stream order makes the slot's buffers safe to reuse: the previous chunk's D2H copies are
        // ahead of this H2D copy on the same stream
        const char *src = static_cast<const char *>(logits_host) + p->frame_in_bytes * b0;
        if (!cuda_ok(cudaMemcpyAsync(s.d_logits, src, p->frame_in_bytes * nb, cudaMemcpyHostToDevice, s.stream), "H2D copy"))
            break;
        if (p->elem_bytes == 1)
            rc = ssb_argmax_stats_u8(static_cast<const uint8_t *>(s.d_logits), nb, p->H, p->W, p->C, roi_h, roi_w,
                                     labels_host ? s.d_labels : nullptr, s.d_stats, s.stream);
        else
            rc = ssb_argmax_stats_f16(static_cast<const uint16_t *>(s.d_logits), nb, p->H, p->W, p->C, roi_h, roi_w,
                                      labels_host ? s.d_labels : nullptr, s.d_stats, s.stream);
        if (rc) break;
        if (!cuda_ok(cudaMemcpyAsync(reinterpret_cast<char *>(stats_host) + stats_frame * b0, s.d_stats, stats_frame * nb,
                                     cudaMemcpyDeviceToHost, s.stream), "D2H copy of the statistics"))
            break;
        if (labels_host)
            cuda_ok(cudaMemcpyAsync(labels_host + p->frame_px * b0, s.d_labels, p->frame_px * nb, cudaMemcpyDeviceToHost,
                                    s.stream), "D2H copy of the label maps");
    }
    // Also on failure: copies of earlier chunks into the caller's buffers may still be in flight, and the
    // caller is free to release those buffers as soon as this returns.
    for (int i = 0; i < p->n_slots; ++i) {
        const cudaError_t e = cudaStreamSynchronize(p->slots[i].stream);
        if (e != cudaSuccess && rc == SSB_OK) rc = fail(SSB_ERR_CUDA, "stream sync failed: %s", cudaGetErrorString(e));
    }
    return rc;
}

// ---- ingest: camera frames held on the host -> model inputs (row f3 of SURVEY.md section 8) ----------
extern "C" int ssb_pipeline_create_ingest(ssb_pipeline **out, int Hc, int Wc, int H, int W, int max_frames_per_chunk,
                                          int n_slots) {
    using namespace ssb;
    if (!out) return fail(SSB_ERR_BAD_POINTER, "out is NULL");
    *out = nullptr;
    if (Hc <= 0 || Wc <= 0 || H <= 0 || W <= 0 || max_frames_per_chunk < 1 || n_slots < 1 || n_slots > 16)
        return fail(SSB_ERR_BAD_SHAPE, "bad ingest geometry");
    ssb_pipeline *p = new (std::nothrow) ssb_pipeline();
    if (!p) return fail(SSB_ERR_LIMIT, "out of host memory");
    p->H = H; p->W = W; p->C = 3; p->elem_bytes = 1; p->cam_h = Hc; p->cam_w = Wc;
    p->max_frames = max_frames_per_chunk; p->n_slots = n_slots;
    p->frame_px = static_cast<size_t>(H) * W * 3;                 // bytes of one model input
    p->frame_in_bytes = static_cast<size_t>(Hc) * Wc * 3;         // bytes of one camera frame
    p->slots = new (std::nothrow) ssb_pipeline::Slot[n_slots]();
    cudaError_t e = cudaGetDevice(&p->device);
    if (e == cudaSuccess) e = cudaMalloc(&p->d_xtab, sizeof(int32_t) * W);
    if (e == cudaSuccess) e = cudaMalloc(&p->d_ytab, sizeof(int32_t) * H);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->tables, cudaEventDisableTiming);
    for (int i = 0; e == cudaSuccess && p->slots && i < n_slots; ++i) {
        auto &s = p->slots[i];
        e = cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaMalloc(&s.d_logits, p->frame_in_bytes * max_frames_per_chunk);
        if (e == cudaSuccess) e = cudaMalloc(&s.d_labels, p->frame_px * max_frames_per_chunk);
    }
    if (e != cudaSuccess || !p->slots) {
        ssb_pipeline_destroy(p);
        return fail(SSB_ERR_CUDA, "ingest pipeline allocation failed: %s", cudaGetErrorString(e));
    }
    *out = p;
    return SSB_OK;
}

extern "C" int ssb_pipeline_preprocess_host(ssb_pipeline *p, const uint8_t *frames_bgr_host, int B,
                                            const int32_t *xtab_host, const int32_t *ytab_host, int new_w, int new_h,
                                            uint8_t *out_rgb_host) {
    using namespace ssb;
    if (!p || p->cam_h <= 0) return fail(SSB_ERR_BAD_POINTER, "not an ingest pipeline");
    if (B < 0) return fail(SSB_ERR_BAD_SHAPE, "B=%d", B);
    if (new_w < 1 || new_w > p->W || new_h < 1 || new_h > p->H)
        return fail(SSB_ERR_BAD_SHAPE, "resized size (%d,%d) outside model input (%d,%d)", new_w, new_h, p->W, p->H);
    if (B == 0) return SSB_OK;
    if (!frames_bgr_host || !xtab_host || !ytab_host || !out_rgb_host) return fail(SSB_ERR_BAD_POINTER, "NULL buffer");
    int dev = 0;
    SSB_CUDA_TRY(cudaGetDevice(&dev));
    if (dev != p->device) return fail(SSB_ERR_BAD_SHAPE, "pipeline belongs to device %d, current is %d", p->device, dev);
    // tables first, on slot 0's stream; the other slots wait for them through an event.  Nothing below
    // returns before the final synchronisation of every slot.
    int rc = SSB_OK, chunk = 0;
    auto cuda_ok = [&](cudaError_t e, const char *what) {
        if (e != cudaSuccess && rc == SSB_OK) rc = fail(SSB_ERR_CUDA, "%s failed: %s", what, cudaGetErrorString(e));
        return e == cudaSuccess;
    };
    cudaStream_t s0 = p->slots[0].stream;
    cuda_ok(cudaMemcpyAsync(p->d_xtab, xtab_host, sizeof(int32_t) * new_w, cudaMemcpyHostToDevice, s0), "table copy") &&
        cuda_ok(cudaMemcpyAsync(p->d_ytab, ytab_host, sizeof(int32_t) * new_h, cudaMemcpyHostToDevice, s0), "table copy") &&
        cuda_ok(cudaEventRecord(p->tables, s0), "cudaEventRecord");
    for (int i = 1; i < p->n_slots && rc == SSB_OK; ++i)
        cuda_ok(cudaStreamWaitEvent(p->slots[i].stream, p->tables, 0), "cudaStreamWaitEvent");
    for (int b0 = 0; b0 < B && rc == SSB_OK; b0 += p->max_frames, ++chunk) {
        const int nb = B - b0 < p->max_frames ? B - b0 : p->max_frames;
        auto &s = p->slots[chunk % p->n_slots];
        cudaError_t e = cudaMemcpyAsync(s.d_logits, frames_bgr_host + p->frame_in_bytes * b0, p->frame_in_bytes * nb,
                                        cudaMemcpyHostToDevice, s.stream);
        if (e != cudaSuccess) { rc = fail(SSB_ERR_CUDA, "H2D copy failed: %s", cudaGetErrorString(e)); break; }
        rc = ssb_preprocess_nearest_u8(static_cast<const uint8_t *>(s.d_logits), nb, p->cam_h, p->cam_w,
                                       static_cast<int64_t>(p->frame_in_bytes), p->d_xtab, p->d_ytab, new_w, new_h,
                                       s.d_labels, p->H, p->W, s.stream);
        if (rc) break;
        e = cudaMemcpyAsync(out_rgb_host + p->frame_px * b0, s.d_labels, p->frame_px * nb, cudaMemcpyDeviceToHost, s.stream);
        if (e != cudaSuccess) rc = fail(SSB_ERR_CUDA, "D2H copy failed: %s", cudaGetErrorString(e));
    }
    for (int i = 0; i < p->n_slots; ++i) {
        const cudaError_t e = cudaStreamSynchronize(p->slots[i].stream);
        if (e != cudaSuccess && rc == SSB_OK) rc = fail(SSB_ERR_CUDA, "stream sync failed: %s", cudaGetErrorString(e));
    }
    return rc;
}
