"""ctypes loader for oracle/segpath_c.c (oracle/_build/liboracle.so).

TEST INFRASTRUCTURE: the checker and the timed CPU baseline, never the product path.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle.so")
_lib = None


def build(force=False):
    src = os.path.join(_HERE, "segpath_c.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["sh", os.path.join(_HERE, "build_oracle.sh")])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_SO)
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def argmax_stats(logits, roi_h=None, roi_w=None, threads=1, want_labels=True, out=None):
    """logits [B,H,W,C] u8 or f16 (C-contiguous) -> (labels [B,H,W] u8 | None, stats [B,C,4] u64).
    `out=(labels, stats)` reuses the caller's arrays (a timed loop should not fault fresh pages in)."""
    logits = np.ascontiguousarray(logits)
    b, h, w, c = logits.shape
    roi_h = h if roi_h is None else roi_h
    roi_w = w if roi_w is None else roi_w
    if out is not None:
        labels, stats = out
        assert stats.shape == (b, c, 4) and stats.dtype == np.uint64 and stats.flags.c_contiguous
        assert labels is None or (labels.shape == (b, h, w) and labels.dtype == np.uint8 and labels.flags.c_contiguous)
        want_labels = labels is not None
    else:
        labels = np.empty((b, h, w), np.uint8) if want_labels else None
        stats = np.zeros((b, c, 4), np.uint64)
    fn = {np.dtype(np.uint8): "oracle_argmax_stats_u8",
          np.dtype(np.float16): "oracle_argmax_stats_f16"}[logits.dtype]
    rc = getattr(lib(), fn)(_p(logits), b, h, w, c, roi_h, roi_w,
                            _p(labels) if want_labels else None, _p(stats), threads)
    if rc != 0:
        raise RuntimeError("%s failed" % fn)
    return labels, stats


def interleaved_copy(a, threads):
    """A copy of `a` whose pages were first touched by the pooled threads, 64 KB chunks handed out dynamically: spread
    over the memory nodes the threads run on (what `numactl --interleave` does for a bandwidth-bound benchmark)."""
    a = np.ascontiguousarray(a)
    out = np.empty_like(a)
    f = lib().oracle_interleaved_copy
    f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int]
    if f(_p(out), _p(a), a.nbytes, threads) != 0:
        raise RuntimeError("oracle_interleaved_copy failed")
    return out


def preprocess(frames_bgr, xtab, ytab, height, width, threads=1):
    """frames [B,Hc,Wc,3] u8 + NEAREST tables -> [B,H,W,3] u8 RGB, zero padded."""
    frames_bgr = np.ascontiguousarray(frames_bgr)
    b, hc, wc, _ = frames_bgr.shape
    xtab = np.ascontiguousarray(xtab, np.int32)
    ytab = np.ascontiguousarray(ytab, np.int32)
    out = np.empty((b, height, width, 3), np.uint8)
    f = lib().oracle_preprocess_nearest_u8
    f.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int64,
                  ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                  ctypes.c_int, ctypes.c_int, ctypes.c_int]
    rc = f(_p(frames_bgr), b, hc, wc, hc * wc * 3, _p(xtab), _p(ytab), len(xtab), len(ytab),
           _p(out), height, width, threads)
    if rc != 0:
        raise RuntimeError("oracle_preprocess_nearest_u8 failed")
    return out


def fg_mask(labels, roi_h, roi_w, colormap_u8):
    """labels [B,H,W] u8 -> mask [B,H,W] u8 (255 inside the ROI where foreground, else 0)."""
    labels = np.ascontiguousarray(labels)
    b, h, w = labels.shape
    cm = np.ascontiguousarray(colormap_u8, np.uint8)
    mask = np.empty((b, h, w), np.uint8)
    rc = lib().oracle_fg_mask_u8(_p(labels), b, h, w, roi_h, roi_w, _p(cm), _p(mask))
    if rc != 0:
        raise RuntimeError("oracle_fg_mask_u8 failed")
    return mask
