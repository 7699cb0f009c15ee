"""Host side of the hot path: mirrors the functions of the reference's `sem_seg_server.py` that sit
on the per-frame path, with the heavy lifting done by libssb200 (CUDA, sm_100a) through the C-ABI.

Names, argument meaning and error behaviour follow the reference where a counterpart exists:
    load_labels                     sem_seg_server.py:30-34
    create_pascal_label_colormap    sem_seg_server.py:36-50
    SegmentContour, Centroid        sem_seg_server.py:21-25
    preprocess_frames               sem_seg_server.py:150-161   (K1, batched, on device)
    argmax_stats                    sem_seg_server.py:162-170 + 102-122 (K2+K3, batched, on device)
    foreground_mask                 sem_seg_server.py:173 + 79-85 (device)
    objects_from_stats              sem_seg_server.py:92-133 + 183-192 restated per label
    objects_from_contours           sem_seg_server.py:77-133 + 183-192, reference-literal (cv2 on the
                                    GPU-made mask; one polygon fill per contour, as the reference does)
    blob_regions, fold_region_histograms, objects_from_regions
                                    the same objects with :98-111 read from the GPU's region tables
                                    (contours, polygon area and moments stay with cv2 on the CPU)
PyTorch is used for device memory and streams only.
"""
import collections
import re

import numpy as np
import torch

from . import _cabi

SegmentContour = collections.namedtuple("SegmentContour", ["label", "score", "area", "centroid"])


class Centroid(collections.namedtuple("Centroid", ["cx", "cy"])):
    __slots__ = ()


def load_labels(path):
    """{int: name} from '<int> <name>' lines (sem_seg_server.py:30-34)."""
    pat = re.compile(r"\s*(\d+)(.+)")
    with open(path, "r", encoding="utf-8") as f:
        pairs = (pat.match(line).groups() for line in f.readlines())
        return {int(num): text.strip() for num, text in pairs}


def create_pascal_label_colormap():
    """256x3 PASCAL VOC colormap (sem_seg_server.py:36-50), vectorised over labels."""
    idx = np.arange(256, dtype=np.int64)
    cmap = np.zeros((256, 3), dtype=np.int64)
    for shift in range(7, -1, -1):
        for ch in range(3):
            cmap[:, ch] |= ((idx >> ch) & 1) << shift
        idx = idx >> 3
    return cmap


def foreground_label_set(num_classes, colormap=None):
    """Labels whose uniform colour passes the reference's threshold (sem_seg_server.py:82-85):
    BGR2GRAY fixed-point luma of the RGB colour, > 127.  {7, 15} for the 21 VOC classes."""
    cm = create_pascal_label_colormap() if colormap is None else colormap
    gray = (cm[:, 0] * 1868 + cm[:, 1] * 9617 + cm[:, 2] * 4899 + 8192) >> 14
    return {int(l) for l in range(num_classes) if gray[l] > 127}


# ---- geometry (host scalars; sem_seg_server.py:139-141,155,164) ---------------------------------
def nearest_index_table(n_in, n_out):
    """Pillow NEAREST source indices: position accumulated in double from scale/2, truncated."""
    if n_in <= 0 or n_out <= 0:
        raise ValueError("empty axis")
    scale = n_in / n_out
    tab = np.empty(n_out, dtype=np.int32)
    pos = scale / 2.0
    for i in range(n_out):
        tab[i] = int(pos)
        pos += scale
    return tab


class FrameGeometry:
    """Everything the kernels need to know about one camera size / model input size pair."""

    def __init__(self, cam_w, cam_h, width, height, keep_aspect_ratio=True):
        self.cam_w, self.cam_h, self.width, self.height = cam_w, cam_h, width, height
        if keep_aspect_ratio:
            # edgetpu resampling_with_original_ratio, call site sem_seg_server.py:155-156
            r = min(width / cam_w, height / cam_h)
            self.new_w, self.new_h = int(cam_w * r), int(cam_h * r)
            if self.new_w <= 0 or self.new_h <= 0:
                raise ValueError("frame collapses to zero pixels")
            self.ratio = (self.new_w / width, self.new_h / height)
        else:
            self.new_w, self.new_h, self.ratio = width, height, (1.0, 1.0)   # :157-159
        # crop applied to the model output, by the reference's formula (:164)
        self.roi_w, self.roi_h = int(width * self.ratio[0]), int(height * self.ratio[1])
        self.xtab = nearest_index_table(cam_w, self.new_w)
        self.ytab = nearest_index_table(cam_h, self.new_h)
        self._dev_tabs = {}

    def device_tables(self, device):
        key = str(device)
        if key not in self._dev_tabs:
            self._dev_tabs[key] = (torch.from_numpy(self.xtab).to(device), torch.from_numpy(self.ytab).to(device))
        return self._dev_tabs[key]


def _stream_ptr(stream):
    s = torch.cuda.current_stream() if stream is None else stream
    return s.cuda_stream


def _require_cuda(t, name):
    if not (isinstance(t, torch.Tensor) and t.is_cuda and t.is_contiguous()):
        raise ValueError("%s must be a contiguous CUDA tensor (there is no CPU path)" % name)


# ---- K1 -------------------------------------------------------------------------------------------
def preprocess_frames(frames_bgr, geom, out=None, stream=None):
    """[B,Hc,Wc,3] u8 BGR camera frames (device) -> [B,H,W,3] u8 RGB model input (device).

    Replaces sem_seg_server.py:150-161 for a batch; `geom` is a FrameGeometry.  Returns `out`."""
    _require_cuda(frames_bgr, "frames_bgr")
    if frames_bgr.dtype != torch.uint8 or frames_bgr.dim() != 4 or frames_bgr.shape[3] != 3:
        raise ValueError("frames_bgr must be [B,Hc,Wc,3] uint8")
    b, hc, wc, _ = frames_bgr.shape
    if (wc, hc) != (geom.cam_w, geom.cam_h):
        raise ValueError("frames are %dx%d, geometry is for %dx%d" % (wc, hc, geom.cam_w, geom.cam_h))
    if out is None:
        out = torch.empty((b, geom.height, geom.width, 3), dtype=torch.uint8, device=frames_bgr.device)
    _require_cuda(out, "out")
    xt, yt = geom.device_tables(frames_bgr.device)
    with torch.cuda.device(frames_bgr.device):
        _cabi.check(_cabi.lib.ssb_preprocess_nearest_u8(
            frames_bgr.data_ptr(), b, hc, wc, hc * wc * 3, xt.data_ptr(), yt.data_ptr(), geom.new_w, geom.new_h,
            out.data_ptr(), geom.height, geom.width, _stream_ptr(stream)))
    return out


# ---- K2 + K3 ----------------------------------------------------------------------------------------
def argmax_stats(logits, roi_h=None, roi_w=None, labels=None, stats=None, want_labels=True, stream=None):
    """[B,H,W,C] u8 or fp16 class scores (device) -> (labels [B,H,W] u8 | None, stats [B,C,4] int64).

    labels = first index of the per-pixel maximum (the ArgMax the tflite graph ends with,
    sem_seg_server.py:162-163); stats[b,c] = (n, sum x, sum y, sum winning logit) over the crop
    [:roi_h,:roi_w] of sem_seg_server.py:164-167.  The stats tensor is int64 holding the library's u64
    bit patterns (torch has no uint64 arithmetic); view as numpy uint64 on the host."""
    _require_cuda(logits, "logits")
    if logits.dim() != 4 or logits.dtype not in (torch.uint8, torch.float16):
        raise ValueError("logits must be [B,H,W,C] uint8 or float16")
    b, h, w, c = logits.shape
    roi_h = h if roi_h is None else roi_h
    roi_w = w if roi_w is None else roi_w
    dev = logits.device
    if want_labels and labels is None:
        labels = torch.empty((b, h, w), dtype=torch.uint8, device=dev)
    if stats is None:
        stats = torch.empty((b, c, 4), dtype=torch.int64, device=dev)
    _require_cuda(stats, "stats")
    fn = _cabi.lib.ssb_argmax_stats_u8 if logits.dtype == torch.uint8 else _cabi.lib.ssb_argmax_stats_f16
    with torch.cuda.device(dev):
        _cabi.check(fn(logits.data_ptr(), b, h, w, c, roi_h, roi_w,
                       labels.data_ptr() if want_labels else None, stats.data_ptr(), _stream_ptr(stream)))
    return (labels if want_labels else None), stats


def argmax_stats_nchw(logits, roi_h=None, roi_w=None, labels=None, stats=None, want_labels=True, stream=None):
    """[B,C,H,W] fp32 or fp16 class scores (device, contiguous) -> (labels [B,H,W] u8 | None, stats
    [B,C,4] int64): argmax_stats for the planar layout a convolutional head leaves on the GPU when it
    replaces engine.run_inference (sem_seg_server.py:162).  The fourth statistic is the sum of
    round(score * 2**24) (fixed_shift=_cabi.F16_FIXED_SHIFT in objects_from_stats)."""
    _require_cuda(logits, "logits")
    if logits.dim() != 4 or logits.dtype not in (torch.float32, torch.float16) or not logits.is_contiguous():
        raise ValueError("logits must be a contiguous [B,C,H,W] float32 or float16 tensor")
    b, c, h, w = logits.shape
    roi_h = h if roi_h is None else roi_h
    roi_w = w if roi_w is None else roi_w
    dev = logits.device
    if want_labels and labels is None:
        labels = torch.empty((b, h, w), dtype=torch.uint8, device=dev)
    if stats is None:
        stats = torch.empty((b, c, 4), dtype=torch.int64, device=dev)
    _require_cuda(stats, "stats")
    fn = _cabi.lib.ssb_argmax_stats_nchw_f32 if logits.dtype == torch.float32 else _cabi.lib.ssb_argmax_stats_nchw_f16
    with torch.cuda.device(dev):
        _cabi.check(fn(logits.data_ptr(), b, h, w, c, roi_h, roi_w,
                       labels.data_ptr() if want_labels else None, stats.data_ptr(), _stream_ptr(stream)))
    return (labels if want_labels else None), stats


_COLORMAPS = {}


def _pascal_colormap_on(device):
    """The PASCAL colormap as a [256,3] u8 device tensor, uploaded once per device."""
    key = str(device)
    if key not in _COLORMAPS:
        _COLORMAPS[key] = torch.from_numpy(create_pascal_label_colormap().astype(np.uint8)).to(device)
    return _COLORMAPS[key]


def foreground_mask(labels, roi_h, roi_w, colormap_dev=None, out=None, stream=None):
    """[B,H,W] u8 label maps (device) -> [B,H,W] u8 mask, 255 where the reference's
    blur/gray/threshold stage (sem_seg_server.py:173,79-85) is foreground inside the ROI."""
    _require_cuda(labels, "labels")
    b, h, w = labels.shape
    if colormap_dev is None:
        colormap_dev = _pascal_colormap_on(labels.device)
    if out is None:
        out = torch.empty_like(labels)
    with torch.cuda.device(labels.device):
        _cabi.check(_cabi.lib.ssb_fg_mask_u8(labels.data_ptr(), b, h, w, roi_h, roi_w, colormap_dev.data_ptr(),
                                             out.data_ptr(), _stream_ptr(stream)))
    return out


# ---- host finalisation (floats from exact integers; sem_seg_server.py:111,121-122,187-190) ----------
def objects_from_stats(stats_frame, width, height, labels, min_area_ratio, scale=1.0, zero_point=0,
                       fixed_shift=0, only_foreground=True):
    """One frame's [C,4] integer block -> list of SegmentContour, in the order the reference's
    producer would push them (one per label, ascending label index).

    area = n (pixels); centroid = (int(sx/n), int(sy/n)) as :121-122; dropped when n is below
    min_area_ratio*H*W (:94,:141) or, with only_foreground, when the label's colour would not survive
    the reference's threshold (:82-85).  score = purity of the label's own pixel set = 1.0 (:111);
    the dequantised mean winning logit is returned as a second value in `score_conf`."""
    st = np.asarray(stats_frame).view(np.int64).reshape(-1, 4)
    min_area = min_area_ratio * height * width
    fg = foreground_label_set(st.shape[0]) if only_foreground else None
    out, conf = [], []
    for lab in range(st.shape[0]):
        n, sx, sy, sq = (int(v) for v in st[lab])
        if n == 0 or n < min_area or (fg is not None and lab not in fg):
            continue
        out.append(SegmentContour(label=labels.get(lab, 0), score=1.0, area=float(n),
                                  centroid=Centroid(cx=int(sx / n), cy=int(sy / n))))
        conf.append(scale * ((sq / float(1 << fixed_shift)) / n - zero_point))
    return out, conf


def objects_from_contours(mask, segment_img, min_area, labels):
    """Reference-literal objects of one frame from the GPU-made foreground mask and label map
    (both already cropped to the ROI, host numpy): cv2.findContours on the mask, then per contour
    the polygon area, majority label, purity and moment centroid of sem_seg_server.py:92-133."""
    import cv2
    contours, _ = cv2.findContours(np.ascontiguousarray(mask), cv2.RETR_TREE, cv2.CHAIN_APPROX_SIMPLE)
    objs = []
    for i in range(len(contours)):
        area = cv2.contourArea(contours[i])
        if area < min_area:
            continue
        # the reference fills a full-frame temporary per contour (sem_seg_server.py:98-103); the same
        # pixels come out of a fill restricted to the contour's bounding box
        bx, by, bw, bh = cv2.boundingRect(contours[i])
        filled = np.zeros((bh, bw), dtype=np.uint8)
        cv2.drawContours(filled, contours, i, color=255, thickness=-1, offset=(-bx, -by))
        inside = segment_img[by:by + bh, bx:bx + bw][filled == 255]
        counts = np.bincount(inside)
        idx = int(np.argmax(counts))
        score = int(counts[idx]) / len(inside)
        m = cv2.moments(contours[i])
        if not m["m00"]:
            continue
        objs.append(SegmentContour(label=labels.get(idx, 0), score=score, area=area,
                                   centroid=Centroid(cx=int(m["m10"] / m["m00"]), cy=int(m["m01"] / m["m00"]))))
    return objs


class HostStaging:
    """Pinned host buffers for the results a producer loop reads back every batch (masks, region tables,
    statistics).  `fetch` issues all copies on the current stream, waits once, and returns numpy views that
    stay valid until the next fetch of the same names.  A pageable `.cpu()` per tensor moves these few MB
    at 1-2 GB/s and synchronises every time; pinned, they take well under a millisecond."""

    def __init__(self):
        self._buf = {}

    def _pinned(self, name, t):
        b = self._buf.get(name)
        if b is None or b.dtype != t.dtype or b.numel() < t.numel():
            b = torch.empty(t.numel(), dtype=t.dtype).pin_memory()
            self._buf[name] = b
        return b[:t.numel()].view(t.shape)

    def fetch(self, tensors):
        out = {}
        for name, t in tensors.items():
            _require_cuda(t, name)
            dst = self._pinned(name, t)
            dst.copy_(t, non_blocking=True)
            out[name] = dst
        torch.cuda.current_stream().synchronize()
        return {k: v.numpy() for k, v in out.items()}


class RegionTables(dict):
    """The tensors blob_regions returns; `.scratch` is the device block the kernels worked in (kept so that
    `out=` can reuse it)."""
    scratch = None


def blob_regions(masks, label_maps, roi_h, roi_w, num_classes=21, max_regions=256, stream=None, out=None):
    """[B,H,W] u8 masks + label maps (device) -> region tables of every frame's ROI, as a dict of device
    tensors: count [B], first / parent / flags [B,R] int32, hist and ring [B,R,num_classes] int32 (C-ABI:
    ssb_blob_regions_u8; rows at or beyond count[b] are not written).  Copy them to the host and hand one
    frame's rows to fold_region_histograms.  `out`: the dict an earlier call with the same shapes returned,
    to reuse its tensors (and the scratch block it carries) instead of allocating."""
    _require_cuda(masks, "masks")
    _require_cuda(label_maps, "label_maps")
    b, h, w = masks.shape
    dev = masks.device
    n_scratch = _cabi.lib.ssb_blob_regions_scratch_bytes(b, roi_h, roi_w)
    t = out
    if (t is None or t["count"].shape[0] != b or t["first"].shape[1] != max_regions or t["hist"].shape[2] != num_classes
            or t.scratch.numel() < n_scratch or t["count"].device != dev):
        t = RegionTables(
            {"count": torch.empty((b,), dtype=torch.int32, device=dev),
             "first": torch.empty((b, max_regions), dtype=torch.int32, device=dev),
             "parent": torch.empty((b, max_regions), dtype=torch.int32, device=dev),
             "flags": torch.empty((b, max_regions), dtype=torch.int32, device=dev),
             "hist": torch.empty((b, max_regions, num_classes), dtype=torch.int32, device=dev),
             "ring": torch.empty((b, max_regions, num_classes), dtype=torch.int32, device=dev)})
        t.scratch = torch.empty(n_scratch, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        _cabi.check(_cabi.lib.ssb_blob_regions_u8(
            masks.data_ptr(), label_maps.data_ptr(), b, h, w, roi_h, roi_w, num_classes, max_regions, t.scratch.data_ptr(),
            t["count"].data_ptr(), t["first"].data_ptr(), t["parent"].data_ptr(), t["flags"].data_ptr(),
            t["hist"].data_ptr(), t["ring"].data_ptr(), _stream_ptr(stream)))
    return t


class EnclosedHistograms(dict):
    """{first pixel of a foreground region: label histogram of everything its outer contour encloses};
    `.holes` is the same for hole contours, keyed by the first pixel of the hole."""
    holes = {}


def fold_region_histograms(count, first, parent, flags, hist, ring=None):
    """One frame's region table (host numpy) -> EnclosedHistograms, or None if the table is incomplete
    (count > rows).  An outer contour encloses its region, the region's holes and whatever is nested in
    them; a hole contour encloses the hole, what is nested in it, and the ring of the surrounding region's
    pixels it runs through (`ring`; without it `.holes` stays empty).  A region's children all have larger
    first pixels than it, so one pass in descending order of first pixel folds every subtree."""
    n = int(count)
    if n > first.shape[0]:
        return None
    first, parent, flags = first[:n], parent[:n], flags[:n]
    enc = hist[:n].astype(np.int64)
    row_of = {int(p): i for i, p in enumerate(first)}
    outside = (flags & 3) == 2                         # background touching the ROI border
    for i in np.argsort(-first):
        if outside[i] or parent[i] < 0:
            continue                                   # the outside, or a region in column 0: nothing encloses it
        j = row_of[int(parent[i])]
        if outside[j]:
            continue                                   # enclosed by the outside only
        enc[j] += enc[i]
    out = EnclosedHistograms((int(first[i]), enc[i]) for i in range(n) if flags[i] & 1)
    out.holes = {}
    if ring is not None:
        out.holes = {int(first[i]): enc[i] + ring[i] for i in range(n) if not (flags[i] & 1) and not outside[i]}
    return out


def fold_regions_batch(tables_host, threads=8, scratch=None):
    """Region tables of a batch (host numpy arrays: the dict blob_regions returned, copied to the host) ->
    one EnclosedHistograms (or None: table incomplete) per frame.  Same results as fold_region_histograms
    frame by frame; the folding runs in the library on `threads` host threads (C-ABI: ssb_fold_regions_host).
    `scratch`: a dict kept by the caller between calls (work arrays are reused)."""
    count, first, parent, flags = (np.ascontiguousarray(tables_host[k]) for k in ("count", "first", "parent", "flags"))
    hist = np.ascontiguousarray(tables_host["hist"])
    ring = tables_host.get("ring")
    ring = None if ring is None else np.ascontiguousarray(ring)
    b, r, c = hist.shape
    scratch = {} if scratch is None else scratch
    if scratch.get("shape") != (b, r, c):
        scratch.update(shape=(b, r, c), enc=np.empty((b, r, c), np.int64), order=np.empty((b, r), np.int32),
                       complete=np.empty((b,), np.uint8))
    enc, order, complete = scratch["enc"], scratch["order"], scratch["complete"]
    _cabi.check(_cabi.lib.ssb_fold_regions_host(
        b, c, r, count.ctypes.data, first.ctypes.data, parent.ctypes.data, flags.ctypes.data, hist.ctypes.data,
        None if ring is None else ring.ctypes.data, enc.ctypes.data, order.ctypes.data, complete.ctypes.data, threads))
    out = []
    for i in range(b):
        if not complete[i]:
            out.append(None)
            continue
        rows = order[i, :int(count[i])]
        keys = first[i][rows].tolist()
        kind = (flags[i][rows] & 3).tolist()          # 1 / 3: foreground; 0: hole; 2: the outside
        rows = rows.tolist()
        enc_i = enc[i]
        e = EnclosedHistograms((k, enc_i[j].copy()) for k, j, f in zip(keys, rows, kind) if f & 1)
        e.holes = {k: enc_i[j].copy() for k, j, f in zip(keys, rows, kind) if f == 0} if ring is not None else {}
        out.append(e)
    return out


def contour_candidates_batch(masks_host, roi_h, roi_w, min_area, threads=8, max_candidates=64, scratch=None):
    """The reference's contour stage (sem_seg_server.py:88, :93-94, :115-122) for a batch of masks on the host
    (numpy [B,H,W] u8), natively and a frame per thread (C-ABI: ssb_contours_host): per frame the list of
    (is_hole, key, area, cx, cy) of the contours that pass the area threshold, in OpenCV's contour order --
    what contour_worker.contour_candidates computes with cv2.  A frame with more than `max_candidates`
    candidates comes back as None."""
    m = masks_host
    if not (isinstance(m, np.ndarray) and m.dtype == np.uint8 and m.ndim == 3 and m.flags.c_contiguous):
        raise ValueError("masks_host must be a C-contiguous [B,H,W] uint8 array")
    b, h, w = m.shape
    scratch = {} if scratch is None else scratch
    if scratch.get("cand_shape") != (b, max_candidates):
        scratch.update(cand_shape=(b, max_candidates), n=np.empty((b,), np.int32), ci=np.empty((b, max_candidates, 4), np.int32),
                       ca=np.empty((b, max_candidates), np.float64))
    n, ci, ca = scratch["n"], scratch["ci"], scratch["ca"]
    _cabi.check(_cabi.lib.ssb_contours_host(m.ctypes.data, b, h, w, roi_h, roi_w, float(min_area), max_candidates,
                                            n.ctypes.data, ci.ctypes.data, ca.ctypes.data, None, threads))
    out = []
    for i in range(b):
        k = int(n[i])
        if k > max_candidates:
            out.append(None)
        elif k == 0:
            out.append([])
        else:
            rows, areas = ci[i, :k].tolist(), ca[i, :k].tolist()
            out.append([(r[0], r[1] if r[1] >= 0 else None, a, r[2], r[3]) for r, a in zip(rows, areas)])
    return out


def contour_moments(masks, tables, roi_h, roi_w, stream=None):
    """Contour integrals (contourArea / moments as exact integers, C-ABI: ssb_contour_moments_u8) of every region in
    `tables` (what blob_regions returned for these masks): adds tables["moments"], [B,R,3] int64, and returns it."""
    _require_cuda(masks, "masks")
    b, h, w = masks.shape
    r = tables["first"].shape[1]
    mom = tables.get("moments")
    if mom is None or mom.shape != (b, r, 3) or mom.device != masks.device:
        mom = torch.empty((b, r, 3), dtype=torch.int64, device=masks.device)
        tables["moments"] = mom
    with torch.cuda.device(masks.device):
        _cabi.check(_cabi.lib.ssb_contour_moments_u8(masks.data_ptr(), b, h, w, roi_h, roi_w, r, tables["count"].data_ptr(),
                                                     tables["first"].data_ptr(), tables["flags"].data_ptr(), mom.data_ptr(),
                                                     _stream_ptr(stream)))
    return mom


def candidates_from_moments_batch(tables_host, roi_w, min_area, threads=8, max_candidates=64, scratch=None):
    """contour_candidates_batch without the masks: the same per-frame candidate lists, in OpenCV's contour order, from
    the region tables and their contour integrals (host numpy arrays; C-ABI: ssb_candidates_from_moments_host).
    None for a frame that needs the masks after all (incomplete table, two candidate holes in one region, more than
    `max_candidates` candidates)."""
    count, first, parent, flags = (np.ascontiguousarray(tables_host[k]) for k in ("count", "first", "parent", "flags"))
    mom = np.ascontiguousarray(tables_host["moments"])
    b, r = first.shape
    scratch = {} if scratch is None else scratch
    if scratch.get("cand_shape") != (b, max_candidates):
        scratch.update(cand_shape=(b, max_candidates), n=np.empty((b,), np.int32), ci=np.empty((b, max_candidates, 4), np.int32),
                       ca=np.empty((b, max_candidates), np.float64))
    n, ci, ca = scratch["n"], scratch["ci"], scratch["ca"]
    _cabi.check(_cabi.lib.ssb_candidates_from_moments_host(b, r, roi_w, count.ctypes.data, first.ctypes.data, parent.ctypes.data,
                                                           flags.ctypes.data, mom.ctypes.data, float(min_area), max_candidates,
                                                           n.ctypes.data, ci.ctypes.data, ca.ctypes.data, threads))
    out = []
    for i in range(b):
        k = int(n[i])
        if k < 0 or k > max_candidates:
            out.append(None)
        elif k == 0:
            out.append([])
        else:
            rows, areas = ci[i, :k].tolist(), ca[i, :k].tolist()
            out.append([(rw[0], rw[1], a, rw[2], rw[3]) for rw, a in zip(rows, areas)])
    return out


def frames_to_device(frames, dst, staging=None, threads=8, stream=None):
    """The frames cap.read() returned (a list of equally shaped C-contiguous uint8 arrays) -> dst[:len(frames)]
    (device tensor [B,Hc,Wc,3]) on `stream` (C-ABI: ssb_frames_to_device).  Frames in page-locked memory are
    sent from where they are, the others through `staging` (pinned tensor of dst's shape).  The copies are
    asynchronous: keep `frames` and `staging` untouched until the stream has passed this point."""
    import ctypes
    n = len(frames)
    if n == 0:
        return dst
    _require_cuda(dst, "dst")
    shape = tuple(dst.shape[1:])
    ptrs = (ctypes.c_void_p * n)()
    for i, f in enumerate(frames):
        if not (isinstance(f, np.ndarray) and f.dtype == np.uint8 and f.shape == shape and f.flags.c_contiguous):
            raise ValueError("frame %d is not a C-contiguous uint8 array of shape %s" % (i, shape))
        ptrs[i] = f.ctypes.data
    frame_bytes = int(np.prod(shape))
    if staging is not None and (not staging.is_pinned() or staging.numel() < n * frame_bytes):
        raise ValueError("staging must be a pinned tensor of at least %d bytes" % (n * frame_bytes))
    with torch.cuda.device(dst.device):
        _cabi.check(_cabi.lib.ssb_frames_to_device(ptrs, n, frame_bytes, None if staging is None else staging.data_ptr(),
                                                   dst.data_ptr(), threads, _stream_ptr(stream)))
    return dst


from .contour_worker import hole_first_pixel  # noqa: E402,F401  (shared with the worker processes; no torch there)


def objects_from_regions(mask, enclosed, min_area, labels, label_map=None):
    """Reference-literal objects of one frame (sem_seg_server.py:92-133) with the per-contour label
    histogram taken from the GPU's region table: `enclosed` is fold_region_histograms(...) of the frame.
    Contours, polygon area and moments come from OpenCV as in the reference.  Frames without a complete
    table (and hole contours when the table has no ring rows) use the polygon fill on `label_map`
    (callable or array; only touched in those cases)."""
    import cv2
    mask = np.ascontiguousarray(mask)
    contours, hierarchy = cv2.findContours(mask, cv2.RETR_TREE, cv2.CHAIN_APPROX_SIMPLE)
    objs, w = [], mask.shape[1]
    segment_img = None
    for i in range(len(contours)):
        area = cv2.contourArea(contours[i])
        if area < min_area:
            continue
        counts = None
        if enclosed is not None:
            depth, j = 0, hierarchy[0][i][3]
            while j >= 0:                              # outer contours sit at even depth of the RETR_TREE hierarchy
                depth, j = depth + 1, hierarchy[0][j][3]
            if depth % 2 == 0:
                x, y = contours[i][0][0]               # cv2 starts an outer contour at the region's first pixel
                counts = enclosed.get(int(y) * w + int(x))
            elif enclosed.holes:
                counts = enclosed.holes.get(hole_first_pixel(contours[i], mask))
        if counts is None:
            if segment_img is None:
                segment_img = label_map() if callable(label_map) else label_map
            bx, by, bw, bh = cv2.boundingRect(contours[i])
            filled = np.zeros((bh, bw), dtype=np.uint8)
            cv2.drawContours(filled, contours, i, color=255, thickness=-1, offset=(-bx, -by))
            counts = np.bincount(segment_img[by:by + bh, bx:bx + bw][filled == 255])
        idx = int(np.argmax(counts))
        score = int(counts[idx]) / int(counts.sum())
        m = cv2.moments(contours[i])
        if not m["m00"]:
            continue
        objs.append(SegmentContour(label=labels.get(idx, 0), score=score, area=area,
                                   centroid=Centroid(cx=int(m["m10"] / m["m00"]), cy=int(m["m01"] / m["m00"]))))
    return objs


def normalise(obj, width, height):
    """(label, score, area, cx, cy) with the reply normalisation of sem_seg_server.py:184-192."""
    return (obj.label, obj.score, min(1.0, obj.area / (width * height)),
            min(1.0, obj.centroid.cx / width), min(1.0, obj.centroid.cy / height))


# ---- host-buffer pipeline ---------------------------------------------------------------------------
class HostPipeline:
    """Owns an ssb_pipeline (device buffers + streams) for logits that live in host memory."""

    def __init__(self, height, width, num_classes, dtype=np.uint8, frames_per_chunk=16, n_slots=3, device=0):
        import ctypes
        self.h, self.w, self.c = height, width, num_classes
        self.dtype = np.dtype(dtype)
        self.device = device
        self._p = ctypes.c_void_p()
        with torch.cuda.device(device):
            _cabi.check(_cabi.lib.ssb_pipeline_create(ctypes.byref(self._p), height, width, num_classes,
                                                      self.dtype.itemsize, frames_per_chunk, n_slots))

    def argmax_stats(self, logits_host, roi_h=None, roi_w=None, labels_out=None, stats_out=None,
                     want_labels=True):
        """logits_host: numpy [B,H,W,C] (C-contiguous; pinned memory is fastest) -> (labels|None, stats u64)."""
        a = logits_host
        if not (isinstance(a, np.ndarray) and a.flags.c_contiguous and a.dtype == self.dtype
                and a.shape[1:] == (self.h, self.w, self.c)):
            raise ValueError("logits_host must be a C-contiguous [B,%d,%d,%d] %s array"
                             % (self.h, self.w, self.c, self.dtype))
        b = a.shape[0]
        if want_labels and labels_out is None:
            labels_out = np.empty((b, self.h, self.w), np.uint8)
        if stats_out is None:
            stats_out = np.empty((b, self.c, 4), np.uint64)
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib.ssb_pipeline_argmax_stats_host(
                self._p, a.ctypes.data, b, self.h if roi_h is None else roi_h, self.w if roi_w is None else roi_w,
                labels_out.ctypes.data if want_labels else None, stats_out.ctypes.data))
        return (labels_out if want_labels else None), stats_out

    def close(self):
        if self._p:
            _cabi.lib.ssb_pipeline_destroy(self._p)
            self._p = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class IngestPipeline:
    """Owns an ingest ssb_pipeline: camera frames in host memory -> model inputs in host memory,
    through K1 on the GPU with copies and kernels overlapped (sem_seg_server.py:144-161 for a batch)."""

    def __init__(self, geom, frames_per_chunk=16, n_slots=3, device=0):
        import ctypes
        self.geom, self.device = geom, device
        self._p = ctypes.c_void_p()
        with torch.cuda.device(device):
            _cabi.check(_cabi.lib.ssb_pipeline_create_ingest(ctypes.byref(self._p), geom.cam_h, geom.cam_w, geom.height,
                                                             geom.width, frames_per_chunk, n_slots))

    def preprocess(self, frames_bgr_host, out=None):
        a, g = frames_bgr_host, self.geom
        if not (isinstance(a, np.ndarray) and a.flags.c_contiguous and a.dtype == np.uint8
                and a.shape[1:] == (g.cam_h, g.cam_w, 3)):
            raise ValueError("frames must be a C-contiguous [B,%d,%d,3] uint8 array" % (g.cam_h, g.cam_w))
        if out is None:
            out = np.empty((a.shape[0], g.height, g.width, 3), np.uint8)
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib.ssb_pipeline_preprocess_host(self._p, a.ctypes.data, a.shape[0], g.xtab.ctypes.data,
                                                               g.ytab.ctypes.data, g.new_w, g.new_h, out.ctypes.data))
        return out

    def close(self):
        if self._p:
            _cabi.lib.ssb_pipeline_destroy(self._p)
            self._p = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
