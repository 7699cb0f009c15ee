/*
 * ssb200 -- C-ABI of the B200-native segmentation hot path.
 *
 * Drop-in boundary for the per-frame data-parallel path of goruck/semantic-segmentation-server
 * (reference: /root/reference/sem_seg_server.py).  The reference has no plugin/FFI interface of its
 * own for this path; the seam is the body of `recognize_and_segment` (sem_seg_server.py:135-213).
 * Each entry point below names the reference lines it replaces.  A Python maintainer binds these
 * with ctypes; nothing in the signatures is a torch or C++ type.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the parameter name ends in `_host`;
 *   - the caller owns every input/output buffer and `stats` needs no initialisation.  The only device
 *     memory the library owns is a small zero-at-rest scratch pool per device (a few hundred KB,
 *     cudaMalloc'ed on first use or when a larger batch than ever before arrives; that call clears the
 *     new block on `stream` and waits for the clear) and the ssb_pipeline objects the caller creates
 *     explicitly.  The argmax kernels leave the scratch zeroed; a block whose kernel could not be
 *     launched is cleared again before its next use.  A kernel that dies while running (a sticky CUDA
 *     error: illegal address, ECC, reset) takes the whole CUDA context with it, so there is nothing to
 *     recover in that case;
 *   - any number of host threads may call at the same time (each on its own stream for overlap); a
 *     scratch block belongs to one call from before its launch until the launch is enqueued, and with
 *     more than 8 calls in that window on one device the ninth waits for one of them;
 *   - CUDA graphs: every device-pointer entry point except ssb_argmax_stats_u8_noalloc may be called
 *     while `stream` is being captured, after the same call (same or larger batch) has run once
 *     outside capture.  A recorded argmax call keeps a scratch block of its own for the life of the
 *     process (at most 16 recorded calls per device; SSB_ERR_LIMIT otherwise);
 *   - all work is enqueued on `stream` (a cudaStream_t passed as void*, 0 = legacy default stream)
 *     of the CURRENT device; calls are asynchronous with respect to the host and never
 *     synchronise the device; re-entrant across (device, stream) pairs;
 *   - return value: SSB_OK (0) or a negative SSB_ERR_* code; a thread-local message is available
 *     from ssb_last_error().  Nothing throws or aborts across the ABI.
 *   - there is NO CPU fallback: without a CUDA device every compute entry point fails with
 *     SSB_ERR_CUDA.
 */
#ifndef SSB200_H_
#define SSB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SSB_VERSION 100 /* 0.1.0 */

#if defined(__GNUC__)
#define SSB_API __attribute__((visibility("default")))
#else
#define SSB_API
#endif

#define SSB_OK 0
#define SSB_ERR_BAD_SHAPE (-1)  /* non-positive / inconsistent dimensions, ROI outside the frame */
#define SSB_ERR_BAD_POINTER (-2) /* NULL where a buffer is required */
#define SSB_ERR_LIMIT (-3)      /* documented size limit exceeded */
#define SSB_ERR_CUDA (-4)       /* a CUDA runtime call failed (message has the CUDA error string) */

/* Statistics block per frame: stats[b][c][SSB_STAT_*], unsigned 64-bit.
 *   N  = pixels of label c inside the ROI          (len(mask_img), sem_seg_server.py:111)
 *   SX = sum of column indices x of those pixels    (m10 of sem_seg_server.py:115-121, as a pixel sum)
 *   SY = sum of row indices y                       (m01 of sem_seg_server.py:115-122)
 *   SQ = sum of the winning logit: raw u8 q for the u8 entry point; value * 2^24 as a
 *        two's-complement int64 for the fp16 entry point (exact, order-independent). */
#define SSB_STAT_N 0
#define SSB_STAT_SX 1
#define SSB_STAT_SY 2
#define SSB_STAT_SQ 3
#define SSB_F16_FIXED_SHIFT 24

SSB_API int ssb_version(void);

/* Message for the last non-zero status returned to the calling thread ("" if none). */
SSB_API const char *ssb_last_error(void);

/* Replaces sem_seg_server.py:150-161 (cvtColor BGR2RGB -> Image.fromarray -> edgetpu
 * resampling_with_original_ratio(NEAREST) -> np.asarray().flatten()) for a batch of frames.
 *
 *   frames_bgr   [B][Hc][Wc][3] u8, frame b at frames_bgr + b*frame_stride (bytes, >= Hc*Wc*3)
 *   xtab, ytab   int32[new_w], int32[new_h]: source column / row of each destination column / row
 *                (Pillow's accumulated-double NEAREST tables, built on the host by the caller --
 *                 semseg host module `nearest_index_table`; values in [0,Wc) / [0,Hc)); xtab must be
 *                 16-byte aligned (any device allocation is)
 *   out_rgb      [B][H][W][3] u8, dense; pixels with x >= new_w or y >= new_h are zero
 *                (ImageOps.expand padding).  new_w <= W, new_h <= H.
 * The stretch branch (sem_seg_server.py:157-159) is the same call with new_w = W, new_h = H. */
SSB_API int ssb_preprocess_nearest_u8(const uint8_t *frames_bgr, int B, int Hc, int Wc, int64_t frame_stride,
                              const int32_t *xtab, const int32_t *ytab, int new_w, int new_h,
                              uint8_t *out_rgb, int H, int W, void *stream);

/* Replaces the ArgMax tail of engine.run_inference (sem_seg_server.py:162), the reshape / crop /
 * astype of :163-170 and the counting + moment arithmetic of :102-122, per label, in one pass.
 *
 *   logits       [B][H][W][C] u8 quantised class scores (dequantisation s*(q-zp), s>0, is monotone,
 *                so argmax on q is exact)
 *   roi_h, roi_w crop of :164-167: statistics cover rows < roi_h and columns < roi_w only;
 *                1 <= roi_h <= H, 1 <= roi_w <= W
 *   labels       [B][H][W] u8 label map (first index of the maximum, np.argmax semantics), written
 *                for the FULL frame; may be NULL to skip it
 *   stats        [B][C][4] u64, zeroed by the call, then accumulated
 * Limits: 2 <= C <= 255 (the streaming fast path is compiled for every C from 2 to 32), H*W < 2^31,
 *         B*H*W < 2^40.  The fast path needs `logits` 16-byte aligned and `labels` 4-byte aligned
 *         (any torch allocation; a frame offset that is a multiple of 16 frames keeps both); other
 *         shapes/alignments take a slower generic kernel with identical results. */
SSB_API int ssb_argmax_stats_u8(const uint8_t *logits, int B, int H, int W, int C, int roi_h, int roi_w,
                        uint8_t *labels, uint64_t *stats, void *stream);

/* Same contract for IEEE fp16 logits (Cityscapes-shaped config; streaming fast path for 2 <= C <= 32).  `logits` is [B][H][W][C] of
 * 16-bit halves passed as uint16_t bit patterns.  Comparison is the floating-point `>` (so
 * -0 == +0 and the first index wins); inputs must be NaN-free (a NaN never wins here, numpy
 * lets the first NaN win).  SQ accumulates value * 2^24 exactly; +-inf winners are not supported
 * in SQ (labels and N/SX/SY stay exact). */
SSB_API int ssb_argmax_stats_f16(const uint16_t *logits, int B, int H, int W, int C, int roi_h, int roi_w,
                         uint8_t *labels, uint64_t *stats, void *stream);

/* Planar class scores, [B][C][H][W] ("NCHW"): what a convolutional segmentation head leaves on the
 * GPU when it stands in for the Edge TPU call of sem_seg_server.py:162 (torchvision's
 * DeepLabv3 returns fp32, fp16 under autocast).  Outputs, ROI, limits and tie rule are those of
 * ssb_argmax_stats_u8; inputs must be NaN-free.
 *   _f32: SQ accumulates round-to-nearest-even(value * 2^24) as a two's-complement int64: exact for
 *         values that are multiples of 2^-24 (all fp16-representable ones), otherwise within 2^-25 per
 *         pixel of the real sum; |value| < 2^39.
 *   _f16: 16-bit halves as uint16_t bit patterns; SQ = sum of value * 2^24, exact (finite winners).
 * `logits` needs element alignment only (fp16 loads pixel pairs when it is 4-byte aligned); `labels`
 * may have any alignment.  Class counts from 2 to 32 run an unrolled kernel, larger ones a generic one. */
SSB_API int ssb_argmax_stats_nchw_f32(const float *logits, int B, int H, int W, int C, int roi_h, int roi_w,
                              uint8_t *labels, uint64_t *stats, void *stream);
SSB_API int ssb_argmax_stats_nchw_f16(const uint16_t *logits, int B, int H, int W, int C, int roi_h, int roi_w,
                              uint8_t *labels, uint64_t *stats, void *stream);

/* No counterpart in the reference: an extra way to call the u8 argmax + statistics pass for hosts
 * that must not trigger a device allocation.
 *
 * ssb_argmax_stats_u8 borrows a small scratch block from a pool that the library grows with
 * cudaMalloc when it first sees a device (or a bigger batch than before).  Some hosts of this code
 * cannot tolerate that: a serving process that has already given all of HBM to its own allocator
 * (the PyTorch caching allocator, RMM, a fixed MPS/MIG memory budget) can see a raw cudaMalloc fail
 * or fragment its pool, and a frame loop with a latency target cannot afford cudaMalloc's
 * device-wide synchronisation stalling its other streams in the middle of a run.  For them this
 * entry point keeps the scratch in an array that is part of the library image itself: it calls no
 * cudaMalloc, cudaFree or other allocating routine, ever.
 *
 * Same arguments, results and limits as ssb_argmax_stats_u8, plus:
 *   - at most SSB_NOALLOC_MAX_FRAMES frames per call (more is refused with SSB_ERR_LIMIT -- split
 *     the batch); the array is sized for that many frames of 21 classes, about 173 KB, and every GPU
 *     that loads the library carries it until the process exits;
 *   - the first call on each device looks the array's device address up (cudaGetSymbolAddress with
 *     that device current) and remembers it; make that first call before starting any stream
 *     capture;
 *   - all calls on a device share the one array.  Each call clears the part it uses first, so a
 *     call that was cut short cannot poison the next one, but two calls whose kernels overlap on
 *     the GPU (different streams, no ordering between them) race on it: order them yourself. */
#define SSB_NOALLOC_MAX_FRAMES 256
SSB_API int ssb_argmax_stats_u8_noalloc(const uint8_t *logits, int B, int H, int W, int C, int roi_h, int roi_w,
                                        uint8_t *labels, uint64_t *stats, void *stream);
/* No counterpart in the reference either: device address and size of that fixed array on the current
 * device, for a host that wants to look at it or clear it itself (and for the test that fills it with
 * garbage to show that ssb_argmax_stats_u8_noalloc does not depend on its contents).  Allocates and
 * launches nothing; the first call on a device does the same address look-up as above. */
SSB_API int ssb_noalloc_scratch_info(void **device_ptr, size_t *bytes);

/* Replaces sem_seg_server.py:173 (label_to_color_image) + :79-85 (blur 3x3, BGR2GRAY, threshold
 * 127) -- the foreground mask handed to cv2.findContours -- with the exact integer formula.
 *
 *   labels       [B][H][W] u8 label map (as written by ssb_argmax_stats_*)
 *   colormap     [256][3] u8 (PASCAL VOC colormap, sem_seg_server.py:36-50)
 *   mask         [B][H][W] u8: 255 where foreground inside the ROI, 0 elsewhere (also outside the
 *                ROI).  Border handling is BORDER_REFLECT_101 on the cropped roi_h x roi_w image. */
SSB_API int ssb_fg_mask_u8(const uint8_t *labels, int B, int H, int W, int roi_h, int roi_w,
                   const uint8_t *colormap, uint8_t *mask, void *stream);

/* Replaces the per-contour fill of sem_seg_server.py:98-111 (filled contour mask -> label pixels under
 * it -> bincount -> majority label and its share) with one table per frame: the connected regions of
 * the mask inside the ROI -- 8-connected foreground, 4-connected background, as cv2.findContours sees
 * them -- each with its first pixel in raster order, the region that encloses it, and the histogram
 * of the label map over its own pixels.  The pixels OpenCV fills for the outer contour of a
 * foreground region are that region plus everything enclosed by it, so the host sums the table along
 * the enclosure tree (Python: ssb200.objects_from_regions); polygon area and moments (:93, :115)
 * stay with OpenCV.
 *
 *   mask, labels [B][H][W] u8 (ssb_fg_mask_u8 / ssb_argmax_stats_* outputs); label values >= C are not counted.
 *                Both are read as aligned 32-bit words: the words that hold a frame's first and last ROI byte are
 *                read whole (up to 3 bytes either side, never across a 4-byte boundary of its own -- inside any
 *                CUDA allocation); the same holds for `mask` in ssb_contour_moments_u8
 *   scratch      ssb_blob_regions_scratch_bytes(B, roi_h, roi_w) bytes of device memory, contents ignored
 *   count        [B] regions found; a frame with count > max_regions has an incomplete table (use the
 *                per-contour fill for that frame)
 *   first        [B][max_regions] first pixel of the region, y*roi_w + x (row order is arbitrary)
 *   parent       [B][max_regions] first pixel of the enclosing region = the region of the pixel left
 *                of `first`; -1 when `first` is in column 0
 *   flags        [B][max_regions] bit 0: foreground region; bit 1: background region touching the ROI
 *                border (not a hole: it is the outside)
 *   (rows at or beyond count[b] are not written: the tables need no clearing and hold garbage there)
 *   hist         [B][max_regions][C] u32 label histogram of the region's own pixels
 *   ring         [B][max_regions][C] u32 or NULL: for a hole (background region not touching the border), the
 *                label histogram of the enclosing foreground region's pixels that are 4-adjacent to it --
 *                the pixels OpenCV's hole contour runs through.  What drawContours fills for a hole contour
 *                is the hole, everything nested in it, and this ring. */
SSB_API size_t ssb_blob_regions_scratch_bytes(int B, int roi_h, int roi_w);
SSB_API int ssb_blob_regions_u8(const uint8_t *mask, const uint8_t *labels, int B, int H, int W, int roi_h, int roi_w,
                        int C, int max_regions, void *scratch, int32_t *count, int32_t *first, int32_t *parent,
                        int32_t *flags, uint32_t *hist, uint32_t *ring, void *stream);

/* Replaces cv2.contourArea (sem_seg_server.py:93) and cv2.moments (:115) for the contours of the regions in a
 * table written by ssb_blob_regions_u8: for every row, the closed-polygon sums over the border OpenCV follows
 * (outer border of a foreground region; hole border of a background region that does not touch the ROI border),
 *   moments [B][max_regions][3] i64 = a00, a10, a01:  contourArea = |a00| / 2,  m00 = a00 / 2,  m10 = a10 / 6,
 *   m01 = a01 / 6 (sign of a00 folded in, as OpenCV does) -- exact integers, so the doubles formed from them equal
 *   OpenCV's.  Rows of the outside background and rows at or beyond count[b] (or beyond max_regions) are zero /
 *   not written.  `mask` is the array the table was made from; same ROI.  A thread per row walks the border on bit
 *   planes of the frame's mask held in shared memory (a warp per row on the mask bytes where they do not fit); a border
 *   longer than 32 768 steps is given up (bounded GPU time) and its row marked a00 = 0, a10 = INT64_MIN:
 *   ssb_candidates_from_moments_host then reports the frame with n_candidates = -1. */
SSB_API int ssb_contour_moments_u8(const uint8_t *mask, int B, int H, int W, int roi_h, int roi_w, int max_regions,
                                   const int32_t *count, const int32_t *first, const int32_t *flags, int64_t *moments,
                                   void *stream);

/* Host-buffer convenience for callers that hold frames/logits in host memory (the reference's
 * situation: cap.read() and engine.run_inference both return host arrays).
 *
 * An ssb_pipeline owns device buffers and CUDA streams (no host memory) for up to
 * `max_frames_per_chunk` frames per chunk and slot on the current device, created once and reused; the
 * _host call splits the batch into chunks, overlaps host->device copies with the kernels and
 * device->host copies of the results, and returns when `stats_host` (and `labels_host`) are complete
 * -- also on failure: it never returns while a copy into or out of the caller's buffers is in flight.
 * Pass PINNED host memory (cudaHostAlloc / cudaHostRegister / torch pin_memory): it is copied by DMA at
 * link speed and the chunks overlap.  Pageable pointers are accepted and give the same results, but the
 * driver then stages every copy through its own bounce buffer and cudaMemcpyAsync blocks the calling
 * thread until the staging is done, so the overlap between chunks is lost (several times slower). */
typedef struct ssb_pipeline ssb_pipeline;

SSB_API int ssb_pipeline_create(ssb_pipeline **out, int H, int W, int C, int elem_bytes /*1=u8, 2=f16*/,
                        int max_frames_per_chunk, int n_slots);
SSB_API int ssb_pipeline_destroy(ssb_pipeline *p);

/* logits_host [B][H][W][C]; labels_host [B][H][W] or NULL; stats_host [B][C][4] u64.
 * Replaces sem_seg_server.py:162-181 for a batch of frames held on the host. */
SSB_API int ssb_pipeline_argmax_stats_host(ssb_pipeline *p, const void *logits_host, int B, int roi_h, int roi_w,
                                   uint8_t *labels_host, uint64_t *stats_host);

/* Ingest pipeline for camera frames held on the host (cap.read(), sem_seg_server.py:144-152): frames go
 * host -> device in chunks on `n_slots` streams, run through ssb_preprocess_nearest_u8 there, and the
 * model inputs [B][H][W][3] come back to `out_rgb_host`; copies of chunk i+1 overlap the kernel of
 * chunk i.  Replaces sem_seg_server.py:150-161 for a batch of host frames.  xtab/ytab are HOST tables
 * (as for ssb_preprocess_nearest_u8); destroy with ssb_pipeline_destroy. */
SSB_API int ssb_pipeline_create_ingest(ssb_pipeline **out, int Hc, int Wc, int H, int W, int max_frames_per_chunk,
                                       int n_slots);
SSB_API int ssb_pipeline_preprocess_host(ssb_pipeline *p, const uint8_t *frames_bgr_host, int B,
                                         const int32_t *xtab_host, const int32_t *ytab_host, int new_w, int new_h,
                                         uint8_t *out_rgb_host);

/* Host-side helpers of a producer loop that feeds the entry points above (sem_seg_server.py:144-195 for a
 * batch of frames).  Both run on host threads the library keeps in a small pool; both take HOST pointers
 * except `dst_device`.
 *
 * ssb_frames_to_device: the frames cap.read() returned (sem_seg_server.py:146; n separate host arrays of
 * frame_bytes each) -> one device batch [n][frame_bytes] on `stream`.  A frame that already sits in
 * page-locked memory is sent by DMA from where it is; the others are copied by up to `threads` host threads
 * into `staging_host` (page-locked, n * frame_bytes, may be NULL if no frame needs it) and sent from there
 * frame by frame, so gathering and DMA overlap.  Returns when every copy is enqueued: do not reuse
 * `staging_host` (or the frames) before `stream` has passed this point. */
SSB_API int ssb_frames_to_device(const uint8_t *const *frames_host, int n, size_t frame_bytes, uint8_t *staging_host,
                                 uint8_t *dst_device, int threads, void *stream);

/* ssb_fold_regions_host: the region tables of B frames (outputs of ssb_blob_regions_u8 copied to the host, same
 * shapes) -> what each contour encloses, i.e. what sem_seg_server.py:98-107 counts under a filled polygon.
 *   enclosed  [B][max_regions][C] i64: for a foreground region, the label histogram of the region, its holes and
 *             everything nested in them (its OUTER contour); for a hole, that of the hole, everything nested in
 *             it and its ring (its HOLE contour; without `ring`, NULL, the ring is left out); rows of the outside
 *             background are not meaningful
 *   order     [B][max_regions] i32: the first count[b] entries are the table rows sorted by `first`
 *             (look a contour up by the first pixel of its region with a binary search)
 *   complete  [B] u8: 0 where count[b] > max_regions (table incomplete: use the per-contour fill for that frame) */
SSB_API int ssb_fold_regions_host(int B, int C, int max_regions, const int32_t *count, const int32_t *first,
                                  const int32_t *parent, const int32_t *flags, const uint32_t *hist, const uint32_t *ring,
                                  int64_t *enclosed, int32_t *order, uint8_t *complete, int threads);

/* ssb_contours_host: the reference's contour stage (sem_seg_server.py:88 cv2.findContours(RETR_TREE,
 * CHAIN_APPROX_SIMPLE), :93-94 cv2.contourArea + area threshold, :115-122 cv2.moments + integer centroid) for the
 * masks of B frames held on the host, a frame per pool thread.  Border following with marks (Suzuki & Abe) on the
 * roi_h x roi_w crop of every mask; areas and moments are exact integer sums turned into doubles with OpenCV's
 * constants, so every number equals what OpenCV returns, and the contours come in OpenCV's order.
 * (Equality with OpenCV holds while the moment sums stay below 2^53, where OpenCV's own double accumulation is
 * exact: ROIs up to about 2000 x 2000.)
 *   masks_host    [B][H][W] u8, 0 = background (ssb_fg_mask_u8 output copied to the host)
 *   n_candidates  [B] contours with area >= min_area and m00 != 0; more than max_candidates: the lists below are cut
 *   cand_i32      [B][max_candidates][4] i32: is_hole, key, cx, cy -- key = first pixel (y*roi_w + x) of the region
 *                 whose outer border this is / of the hole the border goes around (-1: not identified): the row of
 *                 the region table that holds what the contour encloses; cx, cy = int(m10/m00), int(m01/m00)
 *   cand_area     [B][max_candidates] f64 contourArea
 *   n_contours    [B] all contours found (may be NULL) */
SSB_API int ssb_contours_host(const uint8_t *masks_host, int B, int H, int W, int roi_h, int roi_w, double min_area,
                              int max_candidates, int32_t *n_candidates, int32_t *cand_i32, double *cand_area,
                              int32_t *n_contours, int threads);

/* ssb_candidates_from_moments_host: the list ssb_contours_host returns (same layout, same order), made from the
 * region tables and the contour integrals of ssb_contour_moments_u8 (all copied to the host) instead of from the
 * masks.  A frame whose table is incomplete, or in which one region has two candidate holes (their order depends on
 * marks only the border following with marks has), gets n_candidates = -1: run ssb_contours_host on it. */
SSB_API int ssb_candidates_from_moments_host(int B, int max_regions, int roi_w, const int32_t *count, const int32_t *first,
                                             const int32_t *parent, const int32_t *flags, const int64_t *moments,
                                             double min_area, int max_candidates, int32_t *n_candidates, int32_t *cand_i32,
                                             double *cand_area, int threads);

#ifdef __cplusplus
}
#endif
#endif /* SSB200_H_ */
