"""ctypes binding of libssb200.so (include/ssb200.h).  This is the only way the host code reaches
the kernels; there is no other implementation behind these names.  If the library is not built
the import fails loudly -- run `python semantic-segmentation-server_b200/build.py`."""
import ctypes
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SSB200_LIB") or os.path.join(_PKG, "lib", "libssb200.so")  # override: A/B tuning only

SSB_OK = 0
SSB_ERR_BAD_SHAPE, SSB_ERR_BAD_POINTER, SSB_ERR_LIMIT, SSB_ERR_CUDA = -1, -2, -3, -4
STAT_N, STAT_SX, STAT_SY, STAT_SQ = 0, 1, 2, 3
F16_FIXED_SHIFT = 24

_c_int, _c_void_p, _c_i64 = ctypes.c_int, ctypes.c_void_p, ctypes.c_int64

# name -> (restype, argtypes): every symbol include/ssb200.h declares
SIGNATURES = {
    "ssb_version": (_c_int, []),
    "ssb_last_error": (ctypes.c_char_p, []),
    "ssb_preprocess_nearest_u8": (_c_int, [_c_void_p, _c_int, _c_int, _c_int, _c_i64, _c_void_p, _c_void_p,
                                           _c_int, _c_int, _c_void_p, _c_int, _c_int, _c_void_p]),
    "ssb_argmax_stats_u8": (_c_int, [_c_void_p, _c_int, _c_int, _c_int, _c_int, _c_int, _c_int,
                                     _c_void_p, _c_void_p, _c_void_p]),
    "ssb_argmax_stats_u8_noalloc": (_c_int, [_c_void_p, _c_int, _c_int, _c_int, _c_int, _c_int, _c_int,
                                             _c_void_p, _c_void_p, _c_void_p]),
    "ssb_noalloc_scratch_info": (_c_int, [ctypes.POINTER(_c_void_p), ctypes.POINTER(ctypes.c_size_t)]),
    "ssb_argmax_stats_f16": (_c_int, [_c_void_p, _c_int, _c_int, _c_int, _c_int, _c_int, _c_int,
                                      _c_void_p, _c_void_p, _c_void_p]),
    "ssb_argmax_stats_nchw_f32": (_c_int, [_c_void_p, _c_int, _c_int, _c_int, _c_int, _c_int, _c_int,
                                           _c_void_p, _c_void_p, _c_void_p]),
    "ssb_argmax_stats_nchw_f16": (_c_int, [_c_void_p, _c_int, _c_int, _c_int, _c_int, _c_int, _c_int,
                                           _c_void_p, _c_void_p, _c_void_p]),
    "ssb_fg_mask_u8": (_c_int, [_c_void_p, _c_int, _c_int, _c_int, _c_int, _c_int, _c_void_p, _c_void_p,
                                _c_void_p]),
    "ssb_blob_regions_scratch_bytes": (ctypes.c_size_t, [_c_int, _c_int, _c_int]),
    "ssb_blob_regions_u8": (_c_int, [_c_void_p, _c_void_p, _c_int, _c_int, _c_int, _c_int, _c_int, _c_int, _c_int,
                                     _c_void_p, _c_void_p, _c_void_p, _c_void_p, _c_void_p, _c_void_p, _c_void_p, _c_void_p]),
    "ssb_frames_to_device": (_c_int, [_c_void_p, _c_int, ctypes.c_size_t, _c_void_p, _c_void_p, _c_int, _c_void_p]),
    "ssb_fold_regions_host": (_c_int, [_c_int, _c_int, _c_int, _c_void_p, _c_void_p, _c_void_p, _c_void_p, _c_void_p,
                                       _c_void_p, _c_void_p, _c_void_p, _c_void_p, _c_int]),
    "ssb_contours_host": (_c_int, [_c_void_p, _c_int, _c_int, _c_int, _c_int, _c_int, ctypes.c_double, _c_int, _c_void_p,
                                   _c_void_p, _c_void_p, _c_void_p, _c_int]),
    "ssb_contour_moments_u8": (_c_int, [_c_void_p, _c_int, _c_int, _c_int, _c_int, _c_int, _c_int, _c_void_p, _c_void_p,
                                        _c_void_p, _c_void_p, _c_void_p]),
    "ssb_candidates_from_moments_host": (_c_int, [_c_int, _c_int, _c_int, _c_void_p, _c_void_p, _c_void_p, _c_void_p, _c_void_p,
                                                  ctypes.c_double, _c_int, _c_void_p, _c_void_p, _c_void_p, _c_int]),
    "ssb_pipeline_create": (_c_int, [ctypes.POINTER(_c_void_p), _c_int, _c_int, _c_int, _c_int, _c_int, _c_int]),
    "ssb_pipeline_destroy": (_c_int, [_c_void_p]),
    "ssb_pipeline_create_ingest": (_c_int, [ctypes.POINTER(_c_void_p), _c_int, _c_int, _c_int, _c_int, _c_int, _c_int]),
    "ssb_pipeline_preprocess_host": (_c_int, [_c_void_p, _c_void_p, _c_int, _c_void_p, _c_void_p, _c_int, _c_int,
                                              _c_void_p]),
    "ssb_pipeline_argmax_stats_host": (_c_int, [_c_void_p, _c_void_p, _c_int, _c_int, _c_int, _c_void_p,
                                                _c_void_p]),
}


class SsbError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("ssb200 error %d: %s" % (code, message))
        self.code = code


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError("libssb200.so is not built (expected %s); run "
                          "`python semantic-segmentation-server_b200/build.py` -- there is no CPU fallback"
                          % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)  # CDLL calls release the GIL
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    return lib


lib = _load()


def check(rc):
    if rc != SSB_OK:
        raise SsbError(rc, lib.ssb_last_error().decode("utf-8", "replace"))
