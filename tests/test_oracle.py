"""Pins the oracle (oracle/segpath.py + oracle/segpath_c.c) against the fixtures the reference's
own code produced (tests/golden/, made by oracle/make_golden.py) and the SURVEY.md §4 KATs."""
import hashlib

import numpy as np
import pytest

from oracle import cport, segpath as sp


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()[:16]


def test_colormap_and_labels_match_reference(golden):
    g = golden("labels_colormap.npz")
    assert (sp.pascal_colormap() == g["colormap"]).all()           # sem_seg_server.py:36-50
    names = dict(zip(g["label_ids"].tolist(), g["label_names"].tolist()))
    assert names[15] == "person" and names[255] == "ignore" and len(names) == 22
    assert sp.foreground_labels() == [7, 15]                        # KAT-5
    assert sp.gray_of_uniform_label()[[7, 15, 19, 14, 255]].tolist() == [128, 135, 127, 121, 214]


def test_index_tables_kat2():
    xt = sp.nearest_index_table(1920, 513)
    assert xt[0:3].tolist() == [1, 5, 9]
    assert xt[84:87].tolist() == [316, 319, 323]      # closed form gives a different x=85
    assert xt[255:258].tolist() == [956, 959, 963]
    assert sp.nearest_index_table(1080, 288)[0:3].tolist() == [1, 5, 9]
    assert (sp.nearest_index_table(7, 7) == np.arange(7)).all()


def _pre_cases(g):
    return sorted(k[:-5] for k in g.files if k.endswith("_geom"))


def test_preprocess_matches_reference_loop(golden):
    """sem_seg_server.py:150-161 driven with a fake camera/engine -> engine input tensor."""
    g = golden("preprocess.npz")
    for name in _pre_cases(g):
        cw, ch, W, H, seed, keep = (int(v) for v in g[name + "_geom"])
        frame = np.random.default_rng(seed).integers(0, 256, (ch, cw, 3), dtype=np.uint8)
        out, ratio = sp.preprocess_frame(frame, W, H, bool(keep))
        assert digest(out) == str(g[name + "_digest"]), name
        out_pil, ratio_pil = sp.preprocess_frame_pil(frame, W, H, bool(keep))
        assert (out_pil == out).all() and ratio_pil == ratio, name
        if name + "_out" in g.files:
            assert (out == g[name + "_out"]).all(), name
        else:
            assert (out[0] == g[name + "_row0"]).all()
            assert (out[:, [85, 256]] == g[name + "_col85_256"]).all()
        # C port, same tables
        if keep:
            nw, nh, _ = sp.keep_aspect_geometry(cw, ch, W, H)
        else:
            nw, nh = W, H
        c_out = cport.preprocess(frame[None], sp.nearest_index_table(cw, nw),
                                 sp.nearest_index_table(ch, nh), H, W)
        assert (c_out[0] == out).all(), name


def test_stretch_branch_is_nearest():
    """sem_seg_server.py:157-159 under the pinned Pillow 6.0.0 default filter (NEAREST);
    parity unpinned by fixtures (see oracle/make_golden.py)."""
    frame = np.random.default_rng(16).integers(0, 256, (48, 64, 3), dtype=np.uint8)
    out, ratio = sp.preprocess_frame(frame, 33, 33, keep_aspect_ratio=False)
    out_pil, _ = sp.preprocess_frame_pil(frame, 33, 33, keep_aspect_ratio=False)
    assert (out == out_pil).all() and ratio == (1.0, 1.0)


def test_kat2_digest():
    f = np.random.default_rng(1).integers(0, 256, (1080, 1920, 3), dtype=np.uint8)
    assert digest(f) == "ab57ba0358ea770e"
    out, ratio = sp.preprocess_frame(f, 513, 513)
    assert digest(out) == "b8c5ed5577aafd67"
    assert not out[288:].any()
    assert sp.decode_roi(513, 513, ratio) == (513, 288)


def test_roi_and_fgmask_match_reference(golden):
    """sem_seg_server.py:164-167 crop (mask shape) and :173,:79-85 foreground mask."""
    g = golden("objects.npz")
    cm = sp.pascal_colormap().astype(np.uint8)
    for name in sorted(k[:-10] for k in g.files if k.endswith("_label_map")):
        lm, fm = g[name + "_label_map"], g[name + "_fgmask"]
        if name + "_geom" in g.files:
            cw, ch, W, H = (int(v) for v in g[name + "_geom"][:4])
            _, _, ratio = sp.keep_aspect_geometry(cw, ch, W, H)
            roi_w, roi_h = sp.decode_roi(W, H, ratio)
            assert (roi_h, roi_w) == fm.shape, name
        roi_h, roi_w = fm.shape
        assert (sp.foreground_mask(lm[:roi_h, :roi_w]) == fm).all(), name
        c_mask = cport.fg_mask(lm[None], roi_h, roi_w, cm)[0]
        assert (c_mask[:roi_h, :roi_w] == fm).all(), name
        assert not c_mask[roi_h:].any() and not c_mask[:, roi_w:].any()


def test_reference_literal_objects_match_reference(golden):
    """sem_seg_server.py:92-133 + 183-195: deque contents after one frame."""
    g = golden("objects.npz")
    lab = golden("labels_colormap.npz")
    names = dict(zip(lab["label_ids"].tolist(), lab["label_names"].tolist()))
    n_objs = 0
    for name in sorted(k[:-5] for k in g.files if k.endswith("_geom")):
        cw, ch, W, H = (int(v) for v in g[name + "_geom"][:4])
        _, _, ratio = sp.keep_aspect_geometry(cw, ch, W, H)
        roi_w, roi_h = sp.decode_roi(W, H, ratio)
        seg = g[name + "_label_map"][:roi_h, :roi_w]
        objs = sp.reference_literal_objects(seg, W, H, names, float(g[name + "_min_area_ratio"]))
        assert [o["label"] for o in objs] == g[name + "_obj_names"].tolist(), name
        vals = np.array([[o["score"], o["area"], o["cx"], o["cy"]] for o in objs],
                        dtype=np.float32).reshape(-1, 4)
        assert (vals == g[name + "_obj_vals"]).all(), name   # bit-exact after FP32 rounding
        n_objs += len(objs)
    assert n_objs >= 8


def test_kat1_argmax_stats_u8():
    q = np.random.default_rng(0).integers(0, 256, (513, 513, 21), dtype=np.uint8)
    assert digest(q) == "83509ced2bf5ff0f"
    labels, stats = sp.argmax_stats(q[None])
    assert digest(labels[0]) == "06ccd4046908a08f"
    st = stats[0].astype(np.int64)
    assert st[0].tolist() == [13026, 3345589, 3334508, 3175759]
    assert st[7].tolist() == [12640, 3214057, 3245255, 3080501]
    assert st[15].tolist() == [12322, 3135500, 3166364, 3003540]
    assert st[20].tolist() == [12140, 3102486, 3113402, 2958649]
    srt = np.sort(q, axis=-1)
    assert int((srt[..., -1] == srt[..., -2]).sum()) == 10649      # tied maxima -> first index
    c_labels, c_stats = cport.argmax_stats(q[None], threads=2)
    assert (c_labels == labels).all() and (c_stats == stats).all()


def test_kat3_argmax_f16():
    h = np.random.default_rng(2).standard_normal((1024, 2048, 19), dtype=np.float32).astype(np.float16)
    assert digest(h) == "b11c61b711a3e572"
    c_labels, c_stats = cport.argmax_stats(h[None], threads=4)
    assert digest(c_labels[0]) == "7e7612187b8000af"
    assert int(c_stats[0, 0, 0]) == 109977 and int(c_stats[0, 18, 0]) == 110044
    # numpy restatement on a strip (full frame is slow in numpy)
    labels, stats = sp.argmax_stats(h[None, :64], 64, 2048)
    c2_labels, c2_stats = cport.argmax_stats(h[None, :64])
    assert (labels == c2_labels).all() and (stats == c2_stats).all()


def test_f16_special_values_c_vs_numpy():
    rng = np.random.default_rng(7)
    h = rng.standard_normal((2, 40, 50, 19), dtype=np.float32).astype(np.float16)
    u = h.view(np.uint16)
    u[0, 0, 0, :] = [0x8000, 0x0000] + [0xfc00] * 17      # -0 vs +0 tie -> index 0
    u[0, 0, 1, :] = [0x0001, 0x0002, 0x8001] + [0x8400] * 16   # subnormals
    u[0, 1, :, :5] = rng.integers(0, 0x400, (50, 5))
    h[1, 2] *= np.float16(1e-3)
    h[1, 3] *= np.float16(1000)
    labels, stats = sp.argmax_stats(h)
    c_labels, c_stats = cport.argmax_stats(h)
    assert (labels == c_labels).all() and (stats == c_stats).all()
    assert labels[0, 0, 0] == 0 and labels[0, 0, 1] == 1


@pytest.mark.parametrize("roi", [(1, 1), (37, 1), (1, 41), (20, 30), (37, 41)])
def test_roi_edges_c_vs_numpy(roi):
    q = np.random.default_rng(9).integers(0, 256, (3, 37, 41, 21), dtype=np.uint8)
    labels, stats = sp.argmax_stats(q, *roi)
    c_labels, c_stats = cport.argmax_stats(q, *roi, threads=3)
    assert (labels == c_labels).all() and (stats == c_stats).all()
    assert int(stats[..., 0].sum()) == 3 * roi[0] * roi[1]


def test_adversarial_u8():
    h, w, c = 33, 47, 21
    a0 = np.full((1, h, w, c), 77, np.uint8)                    # all tied -> label 0
    l, s = cport.argmax_stats(a0)
    assert not l.any() and int(s[0, 0, 0]) == h * w
    a3 = np.zeros((1, h, w, c), np.uint8); a3[..., c - 1] = 1   # max only in last channel
    l, s = cport.argmax_stats(a3)
    assert (l == c - 1).all() and int(s[0, c - 1, 3]) == h * w
    a2 = np.zeros((1, h, w, c), np.uint8)
    xs = np.arange(w) % c
    a2[0, :, np.arange(w), xs] = 255                            # label = x mod C
    l, s = cport.argmax_stats(a2)
    assert (l[0] == xs[None, :]).all()
    l2, s2 = sp.argmax_stats(a2)
    assert (s == s2).all()


def test_per_label_objects_formulas():
    """Reply normalisation sem_seg_server.py:187-190 and centroid truncation :121-122."""
    stats = np.zeros((21, 4), np.uint64)
    stats[15] = [40000, 40000 * 199 + 7, 40000 * 120 + 39999, 40000 * 250]
    stats[7] = [20000, 20000 * 10, 20000 * 500, 20000 * 200]
    stats[3] = [90000, 1, 1, 1]          # not a foreground colour -> filtered
    stats[12] = [10, 1, 1, 1]            # below min area
    names = {i: "l%d" % i for i in range(21)}
    objs = sp.per_label_objects(stats, 513, 513, names, 0.05, scale=0.5, zero_point=128)
    assert [o["label_index"] for o in objs] == [15, 7]          # LIFO: last appended first
    o = objs[0]
    assert o["area"] == 40000 / (513 * 513) and o["cx"] == 199 / 513 and o["cy"] == 120 / 513
    assert o["score"] == 1.0 and abs(o["score_conf"] - 0.5 * (250 - 128)) < 1e-12
    allobjs = sp.per_label_objects(stats, 513, 513, names, 0.0, only_foreground=False)
    assert [o["label_index"] for o in allobjs] == [15, 12, 7, 3]


def test_packed_box_sum_indexes_the_threshold_table():
    """The arithmetic of the mask kernel (csrc/fgmask.cu) restated in numpy against the oracle: a colour whose channels
    are multiples of 64 becomes one 15-bit code (c/64 in three 5-bit fields), two pixels share a 32-bit word (16 bits
    each), the 3 x 3 box sum of the words never carries from one field (or pixel) into the next, and the sum itself
    indexes the tables -- t01[sum & 1023] >= t2[sum >> 10]  <=>  gray > 127."""
    rng = np.random.default_rng(12)
    cm = sp.pascal_colormap().astype(np.uint8)
    assert not (cm[:64] & 63).any()                       # what the packed path relies on for the first 64 labels
    code = (cm[:, 0].astype(np.uint32) >> 6) | (cm[:, 1].astype(np.uint32) >> 6) << 5 | (cm[:, 2].astype(np.uint32) >> 6) << 10
    mean = (2 * 64 * np.arange(32) + 9) // 18
    t01 = (1868 * mean)[None, :] + (9617 * mean)[:, None]           # [s1][s0]
    t2 = (128 << 14) - 8192 - 4899 * mean
    for shape, roi in (((3, 40, 70), (40, 70)), ((2, 33, 64), (20, 63)), ((1, 9, 5), (9, 5)), ((2, 30, 31), (30, 2))):
        for kind in range(2):
            if kind == 0:
                lm = rng.integers(0, 64, shape).astype(np.uint8)
            else:
                small = rng.choice([0, 7, 15, 12, 63], size=(shape[0], shape[1] // 4 + 1, shape[2] // 4 + 1))
                lm = np.kron(small, np.ones((1, 4, 4), int))[:, :shape[1], :shape[2]].astype(np.uint8)
            want = cport.fg_mask(lm, roi[0], roi[1], cm)
            for b in range(shape[0]):
                if min(roi) < 2:
                    continue                                         # one-column / one-row ROIs take the plain formula
                cpad = np.pad(code[lm[b, :roi[0], :roi[1]]], 1, mode="reflect")      # BORDER_REFLECT_101 at the ROI border
                box = sum(cpad[dy:dy + roi[0], dx:dx + roi[1]] for dy in range(3) for dx in range(3))
                assert ((box & 31) <= 27).all() and ((box >> 5 & 31) <= 27).all() and (box >> 15 == 0).all()   # no carries
                # two pixels per 32-bit word, as the kernel holds them: nine packed words summed give the two box sums
                c2 = np.pad(cpad, ((0, 0), (0, cpad.shape[1] % 2)))
                words = c2[:, 0::2] | c2[:, 1::2] << 16
                assert words.dtype == np.uint32
                wsum = words[0:-2] + words[1:-1] + words[2:]                        # vertical 3-sums of packed words
                v = cpad[0:-2] + cpad[1:-1] + cpad[2:]
                v = np.pad(v, ((0, 0), (0, v.shape[1] % 2)))
                assert ((wsum & 0xffff) == v[:, 0::2]).all() and ((wsum >> 16) == v[:, 1::2]).all()
                lit = t01[(box >> 5) & 31, box & 31] >= t2[box >> 10]
                assert (np.where(lit, 255, 0) == want[b, :roi[0], :roi[1]]).all(), (shape, roi, kind, b)


def test_byte_to_bit_gather_used_by_the_region_and_contour_kernels():
    """The word arithmetic csrc/blobs_cluster.cuh and csrc/contours.cu use to turn four mask bytes into four bits and to
    count a word's pixels by label, restated on numpy uint32: top bit of every non-zero byte, then one multiply by
    0x01020408 gathers the four top bits into bits 24..27 (no two partial products meet in those bits)."""
    rng = np.random.default_rng(4)
    px4 = rng.integers(0, 2 ** 32, 20000, dtype=np.uint64).astype(np.uint32)
    px4[rng.random(px4.size) < 0.5] &= np.uint32(0x00ff00ff)          # plenty of zero bytes
    px4[rng.random(px4.size) < 0.3] &= np.uint32(0xff0000ff)
    px4[:4] = (0, 0xffffffff, 0x01000000, 0x00000080)
    top = (((px4 & np.uint32(0x7f7f7f7f)) + np.uint32(0x7f7f7f7f)) | px4) & np.uint32(0x80808080)
    bits4 = ((top >> np.uint32(7)) * np.uint32(0x01020408)) >> np.uint32(24)      # uint32 arithmetic wraps like the GPU's
    want = sum((((px4 >> np.uint32(8 * k)) & np.uint32(0xff)) != 0).astype(np.uint32) << np.uint32(k) for k in range(4))
    assert (bits4 == want).all()
    # pixels of a 32-pixel word that carry a given label: XOR with the replicated label, then the same gather, inverted
    words = rng.integers(0, 3, (500, 32)).astype(np.uint8)            # three labels, long enough runs of each
    for lab in range(3):
        other = np.zeros(500, np.uint32)
        for j in range(8):
            w = words[:, 4 * j:4 * j + 4].copy().view(np.uint32)[:, 0]                   # little endian: pixel 4j in byte 0
            x4 = w ^ np.uint32(lab * 0x01010101)
            t = (((x4 & np.uint32(0x7f7f7f7f)) + np.uint32(0x7f7f7f7f)) | x4) & np.uint32(0x80808080)
            other |= (((t >> np.uint32(7)) * np.uint32(0x01020408)) >> np.uint32(24)) << np.uint32(4 * j)
        same = ~other
        want = sum((words[:, i] == lab).astype(np.uint32) << np.uint32(i) for i in range(32))
        assert (same == want).all()
