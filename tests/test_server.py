"""Drop-in boundary kept unchanged: wire types, servicer semantics, flags."""
import concurrent.futures
import importlib

import numpy as np
import pytest


@pytest.fixture(scope="module")
def srv():
    import ssb200  # noqa: F401  (loads the package)
    return importlib.import_module("semantic-segmentation-server_b200.server")


@pytest.fixture(scope="module")
def proto():
    import ssb200  # noqa: F401
    return importlib.import_module("semantic-segmentation-server_b200.proto")


def test_descriptor_is_byte_identical_to_reference(proto, golden):
    """sem_seg_server.proto / sem_seg_server_pb2.py:21: same serialized file descriptor."""
    assert proto.FILE_DESCRIPTOR_PROTO.SerializeToString() == bytes(golden("wire.npz")["descriptor"])


def _push(srv, proto, label, score, area, cx, cy):
    srv.segmented_objects.appendleft(proto.SegmentedObject(
        label=label, score=score, area=area, centroid=proto.SegmentedObject.Centroid(cx=cx, cy=cy)))


def test_servicer_lifo_and_padding_bytes(srv, proto, golden):
    """sem_seg_server.py:225-234 (KAT-4): newest first, padded with empty objects."""
    w = golden("wire.npz")
    srv.segmented_objects.clear()
    _push(srv, proto, "person", 1.0, .114, .29, .27)
    _push(srv, proto, "car", .93, .103, .74, .48)
    servicer = srv.SemanticSegmentationServicer(3, (1920, 1080))
    assert servicer.GetSegmentedObjects(proto.Empty(), None).SerializeToString() == bytes(w["reply_two_objects"])
    assert servicer.GetSegmentedObjects(proto.Empty(), None).SerializeToString() == bytes(w["reply_drained"])
    assert servicer.GetCameraResolution(proto.Empty(), None).SerializeToString() == bytes(w["camera_resolution"])


def test_grpc_round_trip(srv, proto):
    import grpc
    srv.segmented_objects.clear()
    _push(srv, proto, "person", 1.0, .114, .29, .27)
    server = grpc.server(concurrent.futures.ThreadPoolExecutor(max_workers=2))
    srv.add_SemanticSegmentationServicer_to_server(srv.SemanticSegmentationServicer(2, (640, 480)), server)
    port = server.add_insecure_port("127.0.0.1:0")
    server.start()
    try:
        with grpc.insecure_channel("127.0.0.1:%d" % port) as channel:
            stub = srv.SemanticSegmentationStub(channel)
            res = stub.GetCameraResolution(proto.Empty(), timeout=10)
            assert (res.width, res.height) == (640, 480)
            objs = stub.GetSegmentedObjects(proto.Empty(), timeout=10)
            assert [o.label for o in objs.data] == ["person", ""]
            assert abs(objs.data[0].area - 0.114) < 1e-7 and objs.data[1].score == 0.0
    finally:
        server.stop(None)


def test_cli_flags_match_reference(srv):
    """sem_seg_server.py:240-260: same flags and defaults; --keep_aspect_ratio is always true."""
    a = srv.build_arg_parser().parse_args([])
    assert (a.camera_idx, a.num_detections, a.min_area_ratio, a.keep_aspect_ratio) == (1, 3, 0.05, True)
    assert a.labels.endswith("pascal_voc_segmentation_labels.txt")
    assert srv.build_arg_parser().parse_args(["--keep_aspect_ratio"]).keep_aspect_ratio is True


def test_load_labels_matches_reference(golden, tmp_path):
    import ssb200
    g = golden("labels_colormap.npz")
    path = tmp_path / "labels.txt"
    path.write_text("\n".join("%d %s" % (i, n) for i, n in zip(g["label_ids"], g["label_names"])))
    got = ssb200.load_labels(str(path))
    assert got == dict(zip(g["label_ids"].tolist(), g["label_names"].tolist()))


# ---- the producer loop on the GPU --------------------------------------------------------------------
class _MapEngine:
    """Engine whose class scores make the argmax equal to queued label maps (like the fake engine the
    golden fixtures were made with: oracle/make_golden.py)."""

    def __init__(self, h, w, maps):
        self.h, self.w, self.maps, self.seen = h, w, list(maps), []

    def get_input_tensor_shape(self):
        return (1, self.h, self.w, 3)

    def run_logits(self, x):
        import torch
        self.seen.append(x.cpu().numpy())
        out = []
        for _ in range(x.shape[0]):
            lm = self.maps.pop(0)
            q = np.random.default_rng(1).integers(0, 100, (self.h, self.w, 21), dtype=np.uint8)
            np.put_along_axis(q, lm[..., None].astype(np.int64), 200, -1)
            out.append(q)
        return torch.from_numpy(np.stack(out)).cuda(), 1.0, 0


@pytest.mark.gpu
@pytest.mark.parametrize("contour_stage", [None, "host", "gpu-short-walks", 2, 0])
def test_recognize_and_segment_reproduces_reference_replies(srv, golden, contour_stage, monkeypatch):
    """Whole producer loop (camera frame -> K1 -> model stand-in -> K2/K3 -> mask -> contours ->
    normalised protos in the deque) against what the reference's own loop left in ITS deque -- with every
    implementation of the contour stage: contour integrals on the GPU (default), the library's border following on
    host threads, the GPU stage with a walk limit that sends every frame to the host, cv2 in worker processes, cv2
    in this process."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    if contour_stage == "gpu-short-walks":                 # every real border exceeds the limit: all frames fall back to the host
        monkeypatch.setenv("SSB_MOMENTS_MAX_STEPS", "25")
        contour_stage = "gpu"
    g = golden("objects.npz")
    lab = golden("labels_colormap.npz")
    names = dict(zip(lab["label_ids"].tolist(), lab["label_names"].tolist()))
    checked = 0
    for name in sorted(k[:-5] for k in g.files if k.endswith("_geom")):
        cw, ch, W, H = (int(v) for v in g[name + "_geom"][:4])
        frames = [np.zeros((ch, cw, 3), np.uint8)] * 3                     # a batch of 3 identical frames
        engine = _MapEngine(H, W, [g[name + "_label_map"]] * 3)
        srv.recognize_and_segment(True, srv.ArrayCamera(frames), engine, names,
                                  float(g[name + "_min_area_ratio"]), batch_frames=2, mode="contours",
                                  contour_workers=contour_stage)
        got = list(srv.segmented_objects)
        want_names, want_vals = g[name + "_obj_names"].tolist(), g[name + "_obj_vals"]
        assert len(got) == 3 * len(want_names), name
        for rep in range(3):                                               # deque: newest frame first
            part = got[rep * len(want_names):(rep + 1) * len(want_names)]
            assert [o.label for o in part] == want_names, name
            vals = np.array([[o.score, o.area, o.centroid.cx, o.centroid.cy] for o in part],
                            dtype=np.float32).reshape(-1, 4)
            assert (vals == want_vals).all(), name
            checked += len(part)
    assert checked >= 24
    srv.segmented_objects.clear()


@pytest.mark.gpu
def test_recognize_and_segment_labels_mode(srv, golden):
    import torch
    from oracle import cport, segpath as sp
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    lab = golden("labels_colormap.npz")
    names = dict(zip(lab["label_ids"].tolist(), lab["label_names"].tolist()))
    frames = [np.random.default_rng(i).integers(0, 256, (480, 640, 3), dtype=np.uint8) for i in range(5)]
    engine = srv.SyntheticEngine()
    srv.recognize_and_segment(True, srv.ArrayCamera(frames), engine, names, 0.0, batch_frames=4, mode="labels")
    got = list(srv.segmented_objects)
    srv.segmented_objects.clear()
    # oracle: same pipeline on the CPU
    want = []
    for f in frames:
        rgb, ratio = sp.preprocess_frame(f, 513, 513)
        roi_w, roi_h = sp.decode_roi(513, 513, ratio)
        q, scale, zp = engine.run_logits(torch.from_numpy(rgb[None]).cuda())
        _, st = cport.argmax_stats(q.cpu().numpy(), roi_h, roi_w)
        want = sp.per_label_objects(st[0], 513, 513, names, 0.0, scale, zp) + want
    assert [o.label for o in got] == [w["label"] for w in want] and len(got) >= 5
    for o, w in zip(got, want):
        for a, b in ((o.area, w["area"]), (o.centroid.cx, w["cx"]), (o.centroid.cy, w["cy"]), (o.score, w["score"])):
            assert abs(a - b) <= 1e-6 * max(abs(b), 1e-30)


@pytest.mark.gpu
@pytest.mark.parametrize("half", [False, True])
def test_recognize_and_segment_with_a_network_as_logits_provider(srv, golden, half):
    """Row f4: a DeepLabv3 network on the GPU hands planar [B,21,H,W] scores to the planar argmax +
    statistics kernel; replies are checked against the oracle run on the very scores the network produced."""
    import torch
    from oracle import segpath as sp
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    pytest.importorskip("torchvision")
    lab = golden("labels_colormap.npz")
    names = dict(zip(lab["label_ids"].tolist(), lab["label_names"].tolist()))

    class Recording(srv.TorchvisionEngine):
        def run_logits(self, x):
            out = super().run_logits(x)
            self.seen.append(out[0].cpu().numpy())
            return out

    engine = Recording(half=half)
    engine.seen = []
    frames = [np.random.default_rng(10 + i).integers(0, 256, (480, 640, 3), dtype=np.uint8) for i in range(3)]
    srv.recognize_and_segment(True, srv.ArrayCamera(frames), engine, names, 0.0, batch_frames=2, mode="labels")
    got = list(srv.segmented_objects)
    srv.segmented_objects.clear()
    assert [s.shape for s in engine.seen] == [(2, 21, 513, 513), (1, 21, 513, 513)]
    assert engine.seen[0].dtype == (np.float16 if half else np.float32)
    want, n_px = [], 0
    for scores in engine.seen:
        _, st = sp.argmax_stats_nchw(scores, 384, 513)                      # 640x480 -> 513x384 inside 513x513
        n_px += int(st[:, :, 0].sum())
        for i in range(scores.shape[0]):
            want = sp.per_label_objects(st[i], 513, 513, names, 0.0, 1.0, 0, fixed_shift=24) + want
    assert n_px == 3 * 384 * 513
    assert [o.label for o in got] == [w["label"] for w in want]
    for o, w in zip(got, want):
        for a, b in ((o.area, w["area"]), (o.centroid.cx, w["cx"]), (o.centroid.cy, w["cy"]), (o.score, w["score"])):
            assert abs(a - b) <= 1e-6 * max(abs(b), 1e-30)


def test_network_logits_provider_shapes_on_cpu(srv):
    """The provider itself needs no GPU: planar scores of the model-input size."""
    import torch
    pytest.importorskip("torchvision")
    e = srv.TorchvisionEngine(height=65, width=65, device="cpu")
    out, scale, zp = e.run_logits(torch.zeros((2, 65, 65, 3), dtype=torch.uint8))
    assert out.shape == (2, 21, 65, 65) and out.dtype == torch.float32 and out.is_contiguous()
    assert (scale, zp) == (1.0, 0) and e.layout == "nchw" and e.get_input_tensor_shape() == (1, 65, 65, 3)
