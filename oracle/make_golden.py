"""Generate tests/golden/*.npz by running the UNMODIFIED reference (through oracle/refshim.py)
on seeded inputs.  Run in the authoring container only (needs /root/reference):

    PROTOCOL_BUFFERS_PYTHON_IMPLEMENTATION=python python -m oracle.make_golden

TEST INFRASTRUCTURE.  The fixtures travel to the GPU box; this script and the reference
do not.  Every array written here is an output of reference code (or an input to it); the
oracle (oracle/segpath.py) and the CUDA path are both checked against them.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(os.path.dirname(HERE), "tests", "golden")


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()[:16]


class FakeCapture:
    """Stands in for cv2.VideoCapture (sem_seg_server.py:144-148): yields the given frames."""
    frames = []

    def __init__(self, idx):
        self._it = iter(FakeCapture.frames)

    def isOpened(self):
        return True

    def read(self):
        try:
            return True, next(self._it)
        except StopIteration:
            return False, None

    def release(self):
        pass


class FakeEngine:
    """Duck-types BasicEngine for sem_seg_server.py:139 and :162.  Records every input tensor
    and answers with the queued label maps as flat float32 (what edgetpu returns)."""

    def __init__(self, height, width, label_maps):
        self.shape = (1, height, width, 3)
        self.inputs = []
        self._maps = list(label_maps)

    def get_input_tensor_shape(self):
        return self.shape

    def run_inference(self, input_tensor):
        assert input_tensor.dtype == np.uint8 and input_tensor.size == np.prod(self.shape)
        self.inputs.append(np.array(input_tensor, copy=True))
        return 1.0, self._maps.pop(0).astype(np.float32).ravel()


def run_reference_loop(ref, frames, height, width, label_maps, labels, min_area_ratio=0.05,
                       keep_aspect_ratio=True):
    """Drive recognize_and_segment (sem_seg_server.py:135-213) once over `frames`.
    Returns (engine inputs, thresholded masks seen by findContours, deque contents)."""
    FakeCapture.frames = frames
    engine = FakeEngine(height, width, label_maps)
    seen_masks = []
    adapter = ref.cv2
    real_find = type(adapter).findContours

    def spy(self, image, mode, method, *a, **k):
        seen_masks.append(np.array(image, copy=True))
        return real_find(self, image, mode, method, *a, **k)

    type(adapter).findContours = spy
    saved = adapter.__dict__.get("VideoCapture")
    adapter.__dict__["VideoCapture"] = FakeCapture
    try:
        ref.recognize_and_segment(keep_aspect_ratio, 0, engine, labels, min_area_ratio)
    finally:
        type(adapter).findContours = real_find
        if saved is None:
            del adapter.__dict__["VideoCapture"]
        else:
            adapter.__dict__["VideoCapture"] = saved
    objs = list(ref.segmented_objects)
    ref.segmented_objects.clear()
    return engine.inputs, seen_masks, objs


def objs_to_arrays(objs):
    names = np.array([o.label for o in objs], dtype="U32")
    vals = np.array([[o.score, o.area, o.centroid.cx, o.centroid.cy] for o in objs],
                    dtype=np.float32).reshape(-1, 4)
    return names, vals


def main():
    from oracle import refshim, segpath as sp
    ref = refshim.load()
    os.makedirs(GOLDEN, exist_ok=True)
    labels = ref.load_labels(os.path.join(refshim.REFERENCE_DIR, "models",
                                          "pascal_voc_segmentation_labels.txt"))

    # ---- labels + colormap (sem_seg_server.py:30-50) ---------------------------------
    np.savez_compressed(
        os.path.join(GOLDEN, "labels_colormap.npz"),
        label_ids=np.array(sorted(labels), dtype=np.int32),
        label_names=np.array([labels[k] for k in sorted(labels)], dtype="U32"),
        colormap=ref.create_pascal_label_colormap().astype(np.int64))

    # ---- pre-process (sem_seg_server.py:150-161) -------------------------------------
    # small cases keep the whole model input; large ones keep a digest + the rows that
    # the survey's index-table trap (x=85, x=256) touches.  The stretch branch (:157-159) is
    # NOT recorded: it calls resize() with Pillow's default filter, which is NEAREST in the
    # pinned Pillow 6.0.0 (requirements.txt:23) but BICUBIC in this container's Pillow 12,
    # so a run here would pin the wrong behaviour; the oracle implements the pinned one.
    pre = {}
    cases = [  # (name, cam_w, cam_h, W, H, seed, keep_aspect)
        ("s0", 100, 37, 33, 33, 10, True), ("s1", 64, 48, 33, 33, 11, True),
        ("s2", 48, 64, 33, 33, 12, True), ("s3", 33, 33, 33, 33, 13, True),
        ("s4", 77, 51, 40, 24, 14, True), ("s5", 31, 97, 24, 40, 15, True),
        ("voc1080", 1920, 1080, 513, 513, 1, True), ("voc720", 1280, 720, 513, 513, 2, True),
        ("voc480", 640, 480, 513, 513, 3, True), ("voc4k", 3840, 2160, 513, 513, 4, True),
    ]
    for name, cw, ch, W, H, seed, keep in cases:
        frame = np.random.default_rng(seed).integers(0, 256, (ch, cw, 3), dtype=np.uint8)
        zero_map = np.zeros((H, W), np.uint8)
        inputs, _, _ = run_reference_loop(ref, [frame], H, W, [zero_map], labels,
                                          keep_aspect_ratio=keep)
        out = inputs[0].reshape(H, W, 3)
        pre[name + "_geom"] = np.array([cw, ch, W, H, seed, int(keep)], dtype=np.int64)
        pre[name + "_digest"] = np.array(digest(out))
        if out.size <= 4096:
            pre[name + "_out"] = out
        else:
            pre[name + "_row0"] = out[0]
            pre[name + "_col85_256"] = out[:, [85, 256]]
    np.savez_compressed(os.path.join(GOLDEN, "preprocess.npz"), **pre)

    # ---- decode/ROI + foreground mask + objects (sem_seg_server.py:163-195) ----------
    obj = {}
    ocases = [  # (name, cam_w, cam_h, W, H, seed, n_blobs, min_area_ratio)
        ("o0", 64, 48, 65, 65, 20, 3, 0.02), ("o1", 48, 64, 65, 65, 21, 4, 0.02),
        ("o2", 65, 65, 65, 65, 22, 5, 0.01), ("o3", 100, 37, 97, 97, 23, 6, 0.005),
        ("v0", 1920, 1080, 513, 513, 24, 4, 0.05), ("v1", 640, 480, 513, 513, 25, 6, 0.05),
        ("v2", 513, 513, 513, 513, 26, 8, 0.05), ("v3", 1280, 720, 513, 513, 27, 5, 0.01),
    ]
    for name, cw, ch, W, H, seed, nb, mar in ocases:
        frame = np.zeros((ch, cw, 3), np.uint8)
        lm = sp.blobby_label_map(seed, H, W, n_blobs=nb)
        _, masks, objs = run_reference_loop(ref, [frame], H, W, [lm], labels, min_area_ratio=mar)
        names, vals = objs_to_arrays(objs)
        obj[name + "_geom"] = np.array([cw, ch, W, H, seed, nb], dtype=np.int64)
        obj[name + "_min_area_ratio"] = np.array(mar)
        obj[name + "_label_map"] = lm
        obj[name + "_fgmask"] = masks[0]          # shape = cropped ROI, pins :164-167 too
        obj[name + "_obj_names"] = names
        obj[name + "_obj_vals"] = vals
    # random-noise label maps stress the 3x3 blur/threshold formula
    for k in range(3):
        lm = np.random.default_rng(40 + k).integers(0, 21, (57, 91)).astype(np.uint8)
        vis = ref.label_to_color_image(lm.astype(int)).astype(np.uint8)
        seen = []
        real_find = type(ref.cv2).findContours

        def spy(self, image, mode, method, *a, **kw):
            seen.append(np.array(image, copy=True))
            return real_find(self, image, mode, method, *a, **kw)
        type(ref.cv2).findContours = spy
        try:
            ref.get_segment_contours(vis)
        finally:
            type(ref.cv2).findContours = real_find
        obj["n%d_label_map" % k] = lm
        obj["n%d_fgmask" % k] = seen[0]
    np.savez_compressed(os.path.join(GOLDEN, "objects.npz"), **obj)

    # ---- servicer / wire format (sem_seg_server.py:216-234, sem_seg_server.proto) ----
    pb2 = sys.modules["sem_seg_server_pb2"]
    wire = {}
    ref.segmented_objects.clear()
    for lab, score, area, cx, cy in (("person", 1.0, .114, .29, .27), ("car", .93, .103, .74, .48)):
        ref.segmented_objects.appendleft(pb2.SegmentedObject(
            label=lab, score=score, area=area,
            centroid=pb2.SegmentedObject.Centroid(cx=cx, cy=cy)))
    servicer = ref.SemanticSegmentationServicer(3, (1920, 1080))
    wire["reply_two_objects"] = np.frombuffer(
        servicer.GetSegmentedObjects(pb2.Empty(), None).SerializeToString(), dtype=np.uint8)
    wire["reply_drained"] = np.frombuffer(
        servicer.GetSegmentedObjects(pb2.Empty(), None).SerializeToString(), dtype=np.uint8)
    wire["camera_resolution"] = np.frombuffer(
        servicer.GetCameraResolution(pb2.Empty(), None).SerializeToString(), dtype=np.uint8)
    wire["descriptor"] = np.frombuffer(pb2.DESCRIPTOR.serialized_pb, dtype=np.uint8)
    np.savez_compressed(os.path.join(GOLDEN, "wire.npz"), **wire)

    for f in sorted(os.listdir(GOLDEN)):
        print(f, os.path.getsize(os.path.join(GOLDEN, f)))


if __name__ == "__main__":
    main()
