"""CPU oracle for the segmentation hot path -- numpy restatement.

TEST INFRASTRUCTURE, NOT PRODUCT.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / `--impl reference` legs may import this module, and only as the checker or
the timed CPU baseline.  The product package never imports it and has no CPU fallback.

Parity status (see DESIGN.md "Oracle"):
  * PINNED against the reference's own functions run in the authoring container
    (oracle/make_golden.py -> tests/golden/*.npz, regenerated through oracle/refshim.py):
      - frame pre-process   (sem_seg_server.py:150-161, keep-aspect NEAREST + zero pad)
      - decode / ROI crop   (sem_seg_server.py:163-170)
      - colormap            (sem_seg_server.py:36-50)
      - foreground mask     (sem_seg_server.py:77-85, blur 3x3 -> BGR2GRAY -> >127)
      - reference-literal objects and reply normalisation (sem_seg_server.py:92-133, 183-195)
      - reply padding / LIFO order (sem_seg_server.py:225-234)
  * PARITY UNPINNED (no counterpart inside /root/reference, which has no tests and whose
    ArgMax lives inside the absent tflite model): per-pixel argmax and the per-label
    {n, sum x, sum y, sum q_max} reduction.  They restate BASELINE.json's north-star with
    numpy semantics (np.argmax: first index of the maximum; np.bincount sums) and are
    anchored on the SURVEY.md §4 known-answer digests KAT-1 / KAT-3.
  * edgetpu==2.14.0 `image_processing.resampling_with_original_ratio` (requirements.txt:6)
    is third-party and absent; its published algorithm is restated in
    `keep_aspect_geometry` / `preprocess_frame_pil` and pinned through the reference's call
    site (sem_seg_server.py:155-156) with Pillow's NEAREST resize.
"""
import math

import numpy as np

STAT_N, STAT_SX, STAT_SY, STAT_SQ = 0, 1, 2, 3
F16_FIXED_SHIFT = 24  # every finite fp16 times 2**24 is an integer


# --------------------------------------------------------------------------------------
# pre-process: sem_seg_server.py:150-161
# --------------------------------------------------------------------------------------
def nearest_index_table(n_in, n_out):
    """Source index for each destination index of Pillow's NEAREST resize.

    Pillow walks the destination axis accumulating `xo += scale` in double precision,
    starting from scale/2, and truncates (the closed form
    floor((x+0.5)*scale) differs at x=85 and x=256 for 1920->513).  Reached from
    sem_seg_server.py:155-156 via img.resize(new_size, Image.NEAREST).
    """
    if n_in <= 0 or n_out <= 0:
        raise ValueError("empty axis")
    scale = n_in / n_out
    pos = scale / 2.0
    tab = np.empty(n_out, dtype=np.int32)
    for i in range(n_out):
        tab[i] = int(pos)
        pos += scale
    return tab


def keep_aspect_geometry(cam_w, cam_h, width, height):
    """(new_w, new_h, ratio) as edgetpu 2.14 resampling_with_original_ratio computes them.

    Call site sem_seg_server.py:155-156.  r = min(W/Wc, H/Hc); new = (int(Wc*r), int(Hc*r));
    the image is resized to `new` and zero-padded right/bottom; ratio = (new_w/W, new_h/H).
    """
    r = min(width / cam_w, height / cam_h)
    new_w, new_h = int(cam_w * r), int(cam_h * r)
    if new_w <= 0 or new_h <= 0:
        raise ValueError("frame collapses to zero pixels")
    return new_w, new_h, (new_w / width, new_h / height)


def stretch_geometry(width, height):
    """sem_seg_server.py:157-159 (unreachable from the CLI): full-size NEAREST, ratio (1,1)."""
    return width, height, (1.0, 1.0)


def preprocess_frame(frame_bgr, width, height, keep_aspect_ratio=True):
    """numpy gather restatement of sem_seg_server.py:150-161.  Returns (rgb[H,W,3] u8, ratio)."""
    cam_h, cam_w = frame_bgr.shape[:2]
    if keep_aspect_ratio:
        new_w, new_h, ratio = keep_aspect_geometry(cam_w, cam_h, width, height)
    else:
        new_w, new_h, ratio = stretch_geometry(width, height)
    xt = nearest_index_table(cam_w, new_w)
    yt = nearest_index_table(cam_h, new_h)
    out = np.zeros((height, width, 3), dtype=np.uint8)
    out[:new_h, :new_w] = frame_bgr[yt][:, xt][:, :, ::-1]
    return out, ratio


def preprocess_frame_pil(frame_bgr, width, height, keep_aspect_ratio=True):
    """Same result through the libraries the reference calls (cv2 + Pillow), used to pin
    `preprocess_frame` and as the timed reference-literal CPU pre-process."""
    import cv2
    from PIL import Image, ImageOps
    rgb = cv2.cvtColor(frame_bgr, cv2.COLOR_BGR2RGB)            # :151
    img = Image.fromarray(rgb)                                  # :152
    if keep_aspect_ratio:
        new_w, new_h, ratio = keep_aspect_geometry(img.size[0], img.size[1], width, height)
        img = img.resize((new_w, new_h), Image.NEAREST)
        img = ImageOps.expand(img, (0, 0, width - new_w, height - new_h))
    else:
        img = img.resize((width, height), Image.NEAREST)        # Pillow 6.0.0 default filter
        ratio = (1.0, 1.0)
    return np.asarray(img), ratio


def decode_roi(width, height, ratio):
    """sem_seg_server.py:164 -- the crop applied to the model output.  Returns (roi_w, roi_h)."""
    return int(width * ratio[0]), int(height * ratio[1])


# --------------------------------------------------------------------------------------
# argmax + per-label statistics (north-star restatement)
# --------------------------------------------------------------------------------------
def argmax_labels(logits):
    """First index of the per-pixel maximum over the last axis (np.argmax; TFLite ArgMax)."""
    return np.argmax(logits, axis=-1).astype(np.uint8)


def winning_logit_fixed(logits, labels):
    """Integer value of the winning logit: raw q for u8, value*2**24 for fp16 (exact), the same
    rounded half-to-even for fp32 (the planar entry points' contract, include/ssb200.h)."""
    win = np.take_along_axis(logits, labels[..., None].astype(np.int64), axis=-1)[..., 0]
    if logits.dtype == np.uint8:
        return win.astype(np.int64)
    if logits.dtype == np.float16:
        return np.ldexp(win.astype(np.float64), F16_FIXED_SHIFT).astype(np.int64)
    if logits.dtype == np.float32:
        return np.rint(np.ldexp(win.astype(np.float64), F16_FIXED_SHIFT)).astype(np.int64)
    raise TypeError(logits.dtype)


def label_stats(labels, win_fixed, num_classes, roi_h, roi_w):
    """[C,4] u64 = per label (n, sum x, sum y, sum winning logit) over [:roi_h,:roi_w].

    x is the column index, y the row index.  This is the arithmetic of
    sem_seg_server.py:102-122 (count, majority, centroid sums) restated per label.
    The fourth column is a two's-complement int64 (negative for fp16 logits).
    """
    lab = labels[:roi_h, :roi_w].astype(np.int64).ravel()
    ys, xs = np.mgrid[0:roi_h, 0:roi_w]
    out = np.zeros((num_classes, 4), dtype=np.int64)
    out[:, STAT_N] = np.bincount(lab, minlength=num_classes)[:num_classes]
    for col, w in ((STAT_SX, xs), (STAT_SY, ys), (STAT_SQ, win_fixed[:roi_h, :roi_w])):
        acc = np.zeros(num_classes, dtype=np.int64)
        np.add.at(acc, lab, w.ravel().astype(np.int64))
        out[:, col] = acc
    return out.view(np.uint64)


def argmax_stats(logits, roi_h=None, roi_w=None):
    """Batched oracle for ssb_argmax_stats_{u8,f16}: logits [B,H,W,C] -> (labels [B,H,W] u8,
    stats [B,C,4] u64)."""
    if logits.ndim != 4:
        raise ValueError("expect [B,H,W,C]")
    b, h, w, c = logits.shape
    roi_h = h if roi_h is None else roi_h
    roi_w = w if roi_w is None else roi_w
    labels = np.empty((b, h, w), dtype=np.uint8)
    stats = np.zeros((b, c, 4), dtype=np.uint64)
    for i in range(b):
        labels[i] = argmax_labels(logits[i])
        stats[i] = label_stats(labels[i], winning_logit_fixed(logits[i], labels[i]), c, roi_h, roi_w)
    return labels, stats


def argmax_stats_nchw(logits, roi_h=None, roi_w=None):
    """Oracle for ssb_argmax_stats_nchw_{f32,f16}: planar [B,C,H,W] scores, same outputs."""
    return argmax_stats(np.ascontiguousarray(np.transpose(logits, (0, 2, 3, 1))), roi_h, roi_w)


# --------------------------------------------------------------------------------------
# colormap / foreground mask: sem_seg_server.py:36-90
# --------------------------------------------------------------------------------------
def pascal_colormap():
    """256x3 PASCAL VOC colormap by bit interleaving (sem_seg_server.py:36-50)."""
    cmap = np.zeros((256, 3), dtype=np.int64)
    for label in range(256):
        v = label
        for shift in range(7, -1, -1):
            for ch in range(3):
                cmap[label, ch] |= ((v >> ch) & 1) << shift
            v >>= 3
    return cmap


def gray_of_uniform_label(cmap=None):
    """Grey level cv2.cvtColor(BGR2GRAY) gives a uniform region of each label's colour.

    The RGB visualisation is fed to BGR2GRAY (sem_seg_server.py:82), so colormap channel 0
    takes the blue weight: gray = (c0*1868 + c1*9617 + c2*4899 + 8192) >> 14.
    """
    cmap = pascal_colormap() if cmap is None else cmap
    return ((cmap[:, 0] * 1868 + cmap[:, 1] * 9617 + cmap[:, 2] * 4899 + 8192) >> 14).astype(np.int64)


def foreground_labels(num_classes=21):
    """Labels whose uniform colour survives threshold(127) -- {7, 15} for VOC."""
    g = gray_of_uniform_label()
    return [int(l) for l in range(num_classes) if g[l] > 127]


def _reflect101(i, n):
    if n == 1:
        return 0
    while i < 0 or i >= n:
        i = -i if i < 0 else 2 * (n - 1) - i
    return i


def foreground_mask(label_map):
    """Integer restatement of sem_seg_server.py:173 + :79-85 on a [h,w] label map.

    colour = colormap[label]; 3x3 box sum with BORDER_REFLECT_101; b = (2*s + 9) // 18
    (round-half-up of s/9); gray as in `gray_of_uniform_label`; mask = 255 where gray > 127.
    """
    h, w = label_map.shape
    cmap = pascal_colormap()
    col = cmap[label_map.astype(np.int64)]                       # [h,w,3] int64
    yi = np.array([_reflect101(i, h) for i in range(-1, h + 1)])
    xi = np.array([_reflect101(i, w) for i in range(-1, w + 1)])
    pad = col[yi][:, xi]
    s = np.zeros_like(col)
    for dy in range(3):
        for dx in range(3):
            s += pad[dy:dy + h, dx:dx + w]
    b = (2 * s + 9) // 18
    gray = (b[..., 0] * 1868 + b[..., 1] * 9617 + b[..., 2] * 4899 + 8192) >> 14
    return np.where(gray > 127, 255, 0).astype(np.uint8)


# --------------------------------------------------------------------------------------
# result records and normalisation: sem_seg_server.py:21-25, 92-133, 183-195
# --------------------------------------------------------------------------------------
def normalise_object(label_name, score, area_px, cx_px, cy_px, width, height):
    """sem_seg_server.py:184-192: area and centroid relative to the full model input."""
    return {
        "label": label_name,
        "score": float(score),
        "area": min(1.0, area_px / (width * height)),
        "cx": min(1.0, cx_px / width),
        "cy": min(1.0, cy_px / height),
    }


def per_label_objects(stats, width, height, label_names, min_area_ratio=0.05,
                      scale=1.0, zero_point=0, only_foreground=True, fixed_shift=0):
    """Per-label objects from a [C,4] stats block, in the order the reference's producer
    would leave them in its LIFO deque (appendleft per object => last label first).

    n is the area in pixels; centroid = int(sum/n) as sem_seg_server.py:121-122 do with
    moments; min-area test as :94 against min_area_ratio*H*W (:141); purity score (:111)
    of a label's own pixel set is 1.0; `score_conf` = scale*(mean winning logit - zero_point).
    """
    st = np.asarray(stats).view(np.int64)
    min_area = min_area_ratio * height * width
    fg = set(foreground_labels(st.shape[0])) if only_foreground else None
    objs = []
    for lab in range(st.shape[0]):
        n, sx, sy, sq = (int(v) for v in st[lab])
        if n == 0 or n < min_area:
            continue
        if fg is not None and lab not in fg:
            continue
        cx_px, cy_px = int(sx / n), int(sy / n)
        name = label_names.get(lab, 0)
        o = normalise_object(name, 1.0, n, cx_px, cy_px, width, height)
        o["label_index"] = lab
        o["score_conf"] = scale * (math.ldexp(sq, -fixed_shift) / n - zero_point)
        objs.insert(0, o)
    return objs


def reference_literal_objects(segment_img, width, height, label_names, min_area_ratio=0.05):
    """Objects exactly as sem_seg_server.py:173-195 produce them from the cropped u8 label
    map: per contour of the thresholded colour image, polygon area/moments from OpenCV.
    Returned newest-first (deque order after appendleft)."""
    import cv2
    min_area = min_area_ratio * height * width
    vis = pascal_colormap()[segment_img.astype(np.int64)].astype(np.uint8)   # :173
    blurred = cv2.blur(vis, (3, 3), 0)                                        # :79
    gray = cv2.cvtColor(blurred, cv2.COLOR_BGR2GRAY)                          # :82
    _, thresh = cv2.threshold(gray, 127, 255, 0)                              # :85
    contours, _ = cv2.findContours(thresh, cv2.RETR_TREE, cv2.CHAIN_APPROX_SIMPLE)  # :88
    objs = []
    for i in range(len(contours)):
        area = cv2.contourArea(contours[i])                                   # :93
        if area < min_area:
            continue
        cimg = np.zeros_like(segment_img)
        cv2.drawContours(cimg, contours, i, color=255, thickness=-1)          # :99
        inside = segment_img[cimg == 255]                                     # :102-103
        idx = int(np.argmax(np.bincount(inside)))                             # :106-107
        score = np.count_nonzero(inside == idx) / len(inside)                 # :111
        m = cv2.moments(contours[i])                                          # :115
        if not m["m00"]:
            continue
        cx_px, cy_px = int(m["m10"] / m["m00"]), int(m["m01"] / m["m00"])     # :121-122
        o = normalise_object(label_names.get(idx, 0), score, area, cx_px, cy_px, width, height)
        o["label_index"] = idx
        objs.insert(0, o)
    return objs


def load_labels(path):
    """sem_seg_server.py:30-34: '<int> <name>' per line -> {int: name}."""
    import re
    pat = re.compile(r"\s*(\d+)(.+)")
    out = {}
    with open(path, "r", encoding="utf-8") as f:
        for line in f:
            num, text = pat.match(line).groups()
            out[int(num)] = text.strip()
    return out


# --------------------------------------------------------------------------------------
# seeded parity inputs (SURVEY.md §8d "Value distributions & seeds")
# --------------------------------------------------------------------------------------
def blobby_logits_u8(seed, h, w, c, cell=16, bg_frac=0.7):
    """U1: low-res label field nearest-upsampled; winner = 200+noise, others noise < 180."""
    rng = np.random.default_rng(seed)
    gh, gw = -(-h // cell), -(-w // cell)
    field = rng.integers(1, c, (gh, gw))
    field[rng.random((gh, gw)) < bg_frac] = 0
    lab = np.repeat(np.repeat(field, cell, 0), cell, 1)[:h, :w]
    q = rng.integers(0, 180, (h, w, c), dtype=np.uint8)
    np.put_along_axis(q, lab[..., None], rng.integers(200, 256, (h, w, 1), dtype=np.uint8), -1)
    return q


def blobby_label_map(seed, h, w, n_blobs=4, labels=(7, 15, 12, 19)):
    """Rectangles/ellipses of a few VOC labels on background, for the object-level fixtures."""
    rng = np.random.default_rng(seed)
    m = np.zeros((h, w), dtype=np.uint8)
    yy, xx = np.mgrid[0:h, 0:w]
    for _ in range(n_blobs):
        lab = labels[int(rng.integers(0, len(labels)))]
        cy, cx = rng.integers(h // 8, h - h // 8), rng.integers(w // 8, w - w // 8)
        ry, rx = rng.integers(max(2, h // 10), max(3, h // 3)), rng.integers(max(2, w // 10), max(3, w // 3))
        if rng.random() < 0.5:
            m[max(0, cy - ry):cy + ry, max(0, cx - rx):cx + rx] = lab
        else:
            m[((yy - cy) / ry) ** 2 + ((xx - cx) / rx) ** 2 <= 1.0] = lab
    return m
