"""Import the UNMODIFIED reference module (/root/reference/sem_seg_server.py) in this container.

TEST INFRASTRUCTURE ONLY.  Used by oracle/make_golden.py (fixture generation) and by the
CPU tests that re-check the oracle against the live reference when /root/reference is
mounted.  Nothing on the product path, nothing under `-m gpu`, and neither smoke() nor
bench.py may import this file: /root/reference does not exist on the GPU box.

Three shims are needed:
  1. `edgetpu` is not installed       -> stub modules (sem_seg_server.py:18-19 import them).
  2. the generated *_pb2.py predates protobuf 4 -> force the pure-python protobuf backend
     (PROTOCOL_BUFFERS_PYTHON_IMPLEMENTATION=python) before google.protobuf is imported.
  3. sem_seg_server.py:88 unpacks three values from cv2.findContours (OpenCV 3 API);
     OpenCV 4 returns two -> adapter on the module's private `cv2` name only.
Also sem_seg_server.py:16 does `import concurrent` and relies on grpc having imported
`concurrent.futures`; we import it explicitly.
"""
import importlib
import os
import sys
import types

REFERENCE_DIR = os.environ.get("SSB_REFERENCE_DIR", "/root/reference")


def available():
    return os.path.isfile(os.path.join(REFERENCE_DIR, "sem_seg_server.py"))


class _Cv2Adapter(types.ModuleType):
    """Forwards everything to the real cv2, but gives findContours the OpenCV-3 arity."""

    def __init__(self, real):
        super().__init__("cv2")
        self.__dict__["_real"] = real

    def __getattr__(self, name):
        return getattr(self.__dict__["_real"], name)

    def findContours(self, image, mode, method, *a, **k):
        contours, hierarchy = self.__dict__["_real"].findContours(image, mode, method, *a, **k)
        return image, list(contours), hierarchy


def _resampling_with_original_ratio(img, required_size, sample):
    """Restatement of edgetpu==2.14.0 image_processing.resampling_with_original_ratio
    (third-party, requirements.txt:6, NOT under /root/reference; algorithm recalled from
    the published package): scale by the smaller of the two axis ratios, truncate the new
    size to int, resize with `sample`, pad right/bottom with zeros to the required size,
    and return the fraction of the target each axis covers."""
    from PIL import ImageOps
    src_w, src_h = img.size
    r = min(required_size[0] / src_w, required_size[1] / src_h)
    new_w, new_h = int(src_w * r), int(src_h * r)
    resized = img.resize((new_w, new_h), sample)
    padded = ImageOps.expand(resized, (0, 0, required_size[0] - new_w, required_size[1] - new_h))
    return padded, (new_w / required_size[0], new_h / required_size[1])


_cached = None


def load():
    """Returns the imported reference module `sem_seg_server` (cached)."""
    global _cached
    if _cached is not None:
        return _cached
    if not available():
        raise RuntimeError("reference not mounted at %s" % REFERENCE_DIR)
    if "google.protobuf" in sys.modules and \
            os.environ.get("PROTOCOL_BUFFERS_PYTHON_IMPLEMENTATION") != "python":
        raise RuntimeError("refshim.load() must run before google.protobuf is imported "
                           "(or with PROTOCOL_BUFFERS_PYTHON_IMPLEMENTATION=python set)")
    os.environ["PROTOCOL_BUFFERS_PYTHON_IMPLEMENTATION"] = "python"
    import concurrent.futures  # noqa: F401  (sem_seg_server.py:16,272)
    for name in ("edgetpu", "edgetpu.basic", "edgetpu.basic.basic_engine",
                 "edgetpu.utils", "edgetpu.utils.image_processing"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["edgetpu.basic.basic_engine"].BasicEngine = object
    sys.modules["edgetpu.utils"].image_processing = sys.modules["edgetpu.utils.image_processing"]
    sys.modules["edgetpu.utils.image_processing"].resampling_with_original_ratio = \
        _resampling_with_original_ratio
    if REFERENCE_DIR not in sys.path:
        sys.path.insert(0, REFERENCE_DIR)
    mod = importlib.import_module("sem_seg_server")
    import cv2 as real_cv2
    mod.cv2 = _Cv2Adapter(real_cv2)
    _cached = mod
    return mod
