"""CPU-side checks of the C-ABI boundary: the library loads, exports every symbol include/ssb200.h
declares, validates arguments before touching CUDA, and fails loudly (no CPU fallback) without a GPU."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def ssb():
    import ssb200
    return ssb200


def test_header_symbols_all_exported(ssb):
    hdr = open(os.path.join(ROOT, "include", "ssb200.h")).read()
    declared = set(re.findall(r"SSB_API\s+[\w\s\*]+?\b(ssb_\w+)\s*\(", hdr))
    assert len(declared) >= 9
    assert declared == set(ssb._cabi.SIGNATURES), declared ^ set(ssb._cabi.SIGNATURES)
    raw = ctypes.CDLL(ssb._cabi.LIB_PATH)
    for name in declared:
        assert hasattr(raw, name), name
    assert ssb._cabi.lib.ssb_version() == 100


def test_no_torch_or_cxx_types_in_abi(ssb):
    hdr = open(os.path.join(ROOT, "include", "ssb200.h")).read()
    code = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)       # signatures only, comments stripped
    assert "torch" not in code and "std::" not in code and "at::" not in code and "#include <cuda" not in code


def test_argument_validation_before_cuda(ssb):
    lib = ssb._cabi.lib
    one = ctypes.c_void_p(16)   # never dereferenced: validation fails first
    assert lib.ssb_argmax_stats_u8(one, 1, 0, 4, 21, 1, 1, None, one, None) == ssb._cabi.SSB_ERR_BAD_SHAPE
    assert lib.ssb_argmax_stats_u8(one, 1, 4, 4, 21, 5, 4, None, one, None) == ssb._cabi.SSB_ERR_BAD_SHAPE
    assert b"ROI" in lib.ssb_last_error()
    assert lib.ssb_argmax_stats_u8(one, 1, 4, 4, 1, 4, 4, None, one, None) == ssb._cabi.SSB_ERR_LIMIT
    assert lib.ssb_argmax_stats_u8(one, 1, 4, 4, 256, 4, 4, None, one, None) == ssb._cabi.SSB_ERR_LIMIT
    assert lib.ssb_argmax_stats_u8(None, 1, 4, 4, 21, 4, 4, None, one, None) == ssb._cabi.SSB_ERR_BAD_POINTER
    assert lib.ssb_argmax_stats_f16(one, 1, 65536, 65536, 19, 4, 4, None, one, None) == ssb._cabi.SSB_ERR_LIMIT
    assert lib.ssb_argmax_stats_u8(one, 0, 4, 4, 21, 4, 4, None, one, None) == 0          # empty batch
    for fn in (lib.ssb_argmax_stats_nchw_f32, lib.ssb_argmax_stats_nchw_f16):
        assert fn(one, 1, 4, 4, 21, 4, 5, None, one, None) == ssb._cabi.SSB_ERR_BAD_SHAPE
        assert fn(one, 1, 4, 4, 300, 4, 4, None, one, None) == ssb._cabi.SSB_ERR_LIMIT
        assert fn(None, 1, 4, 4, 21, 4, 4, None, one, None) == ssb._cabi.SSB_ERR_BAD_POINTER
        assert fn(ctypes.c_void_p(17), 1, 4, 4, 21, 4, 4, None, one, None) == ssb._cabi.SSB_ERR_BAD_POINTER   # element alignment
        assert fn(one, 0, 4, 4, 21, 4, 4, None, one, None) == 0
    assert lib.ssb_preprocess_nearest_u8(one, 1, 4, 4, 48, one, one, 5, 4, one, 4, 4, None) == \
        ssb._cabi.SSB_ERR_BAD_SHAPE
    assert lib.ssb_preprocess_nearest_u8(one, 1, 4, 4, 47, one, one, 4, 4, one, 4, 4, None) == \
        ssb._cabi.SSB_ERR_BAD_SHAPE
    assert lib.ssb_fg_mask_u8(one, 1, 4, 4, 0, 4, one, one, None) == ssb._cabi.SSB_ERR_BAD_SHAPE
    assert lib.ssb_pipeline_create(None, 4, 4, 21, 1, 1, 1) == ssb._cabi.SSB_ERR_BAD_POINTER


def test_no_cpu_fallback(ssb):
    import torch
    with pytest.raises(ValueError, match="no CPU path"):
        ssb.argmax_stats(torch.zeros(1, 4, 4, 21, dtype=torch.uint8))
    with pytest.raises(ValueError, match="no CPU path"):
        ssb.preprocess_frames(torch.zeros(1, 4, 4, 3, dtype=torch.uint8), ssb.FrameGeometry(4, 4, 4, 4))
    if not torch.cuda.is_available():
        # a raw call with plausible arguments must report a CUDA failure, not compute anything
        buf = ctypes.create_string_buffer(4 * 4 * 21 + 64)
        st = ctypes.create_string_buffer(21 * 32)
        rc = ssb._cabi.lib.ssb_argmax_stats_u8(ctypes.addressof(buf), 1, 4, 4, 21, 4, 4, None,
                                               ctypes.addressof(st), None)
        assert rc == ssb._cabi.SSB_ERR_CUDA


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "semantic-segmentation-server_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f
                assert "oracle/" not in src and "liboracle" not in src, f


def test_host_geometry_matches_oracle(ssb):
    """Host-side tables/geometry of the product vs the oracle restatement (sem_seg_server.py:155,164)."""
    from oracle import segpath as sp
    for cw, ch, W, H in [(1920, 1080, 513, 513), (640, 480, 513, 513), (480, 640, 513, 513), (100, 37, 33, 33),
                         (3840, 2160, 513, 513), (513, 513, 513, 513), (31, 97, 24, 40)]:
        g = ssb.FrameGeometry(cw, ch, W, H)
        nw, nh, ratio = sp.keep_aspect_geometry(cw, ch, W, H)
        assert (g.new_w, g.new_h, g.ratio) == (nw, nh, ratio)
        assert (g.roi_w, g.roi_h) == sp.decode_roi(W, H, ratio)
        assert (g.xtab == sp.nearest_index_table(cw, nw)).all() and (g.ytab == sp.nearest_index_table(ch, nh)).all()
    assert (ssb.create_pascal_label_colormap() == sp.pascal_colormap()).all()
    assert ssb.foreground_label_set(21) == set(sp.foreground_labels(21))


def test_objects_from_stats_matches_oracle(ssb):
    import numpy as np
    from oracle import segpath as sp
    rng = np.random.default_rng(3)
    names = {i: "l%d" % i for i in range(21)}
    for _ in range(20):
        st = np.zeros((21, 4), np.uint64)
        for lab in rng.choice(21, 6, replace=False):
            n = int(rng.integers(1, 200000))
            st[lab] = [n, n * int(rng.integers(0, 513)) + int(rng.integers(0, n)),
                       n * int(rng.integers(0, 513)) + int(rng.integers(0, n)), n * int(rng.integers(0, 256))]
        for only_fg in (True, False):
            objs, conf = ssb.objects_from_stats(st, 513, 513, names, 0.05, scale=0.25, zero_point=3,
                                                only_foreground=only_fg)
            want = sp.per_label_objects(st, 513, 513, names, 0.05, scale=0.25, zero_point=3,
                                        only_foreground=only_fg)
            got = [ssb.normalise(o, 513, 513) for o in objs][::-1]     # oracle lists newest-first
            assert [g[0] for g in got] == [w["label"] for w in want]
            for g, c, w in zip(got, conf[::-1], want):
                assert g[1:] == (w["score"], w["area"], w["cx"], w["cy"])
                assert abs(c - w["score_conf"]) <= 1e-6 * max(1.0, abs(w["score_conf"]))


def test_objects_from_contours_matches_reference_literal_oracle(ssb, golden):
    """Host tail of the reference-literal path (sem_seg_server.py:88-133) on masks from the oracle:
    identical objects to the fixtures the reference produced, and to the oracle on random scenes."""
    import numpy as np
    from oracle import segpath as sp
    g = golden("objects.npz")
    lab = golden("labels_colormap.npz")
    names = dict(zip(lab["label_ids"].tolist(), lab["label_names"].tolist()))
    for name in sorted(k[:-5] for k in g.files if k.endswith("_geom")):
        W, H = (int(v) for v in g[name + "_geom"][2:4])
        fm = g[name + "_fgmask"]
        seg = np.ascontiguousarray(g[name + "_label_map"][:fm.shape[0], :fm.shape[1]])
        objs = ssb.objects_from_contours(fm, seg, float(g[name + "_min_area_ratio"]) * H * W, names)
        got = [ssb.normalise(o, W, H) for o in objs][::-1]
        assert [o[0] for o in got] == g[name + "_obj_names"].tolist()
        assert (np.array([o[1:] for o in got], dtype=np.float32).reshape(-1, 4) == g[name + "_obj_vals"]).all()
    for seed in range(12):
        lm = sp.blobby_label_map(seed, 288, 513, n_blobs=8)
        a = [ssb.normalise(o, 513, 513) for o in ssb.objects_from_contours(sp.foreground_mask(lm), lm, 50.0, names)][::-1]
        b = sp.reference_literal_objects(lm, 513, 513, names, 50.0 / (513 * 513))
        assert [x[0] for x in a] == [o["label"] for o in b]
        assert [x[1:] for x in a] == [(o["score"], o["area"], o["cx"], o["cy"]) for o in b]
