"""BASELINE.json configs at their full per-GPU sizes: the oracle cannot chew through these in
seconds, so they are checked through size-independent invariants on EVERY frame (each ROI pixel
counted once; closed-form sums of all x and all y; sum of winning scores against an independent
torch reduction; labels point at the first maximum) plus a full oracle comparison on a few frames."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from oracle import cport, segpath as sp  # noqa: E402  (checker only)


@pytest.fixture(scope="module")
def ssb():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ssb200
    return ssb200


def check_invariants(lg, labels, stats, roi_h, roi_w, frames_vs_oracle, fixed_shift=0):
    b = lg.shape[0]
    st = stats.cpu().numpy().view(np.uint64).astype(np.int64)
    assert (st[:, :, 0].sum(1) == roi_h * roi_w).all()
    assert (st[:, :, 1].sum(1) == roi_h * (roi_w * (roi_w - 1) // 2)).all()
    assert (st[:, :, 2].sum(1) == roi_w * (roi_h * (roi_h - 1) // 2)).all()
    for s in range(0, b, 32):                      # chunked to bound temporary memory
        part = lg[s:s + 32]
        mx = part.amax(dim=-1)
        first = (part == mx.unsqueeze(-1)).to(torch.uint8).argmax(dim=-1)
        assert bool((first == labels[s:s + 32].long()).all())
        if fixed_shift == 0:
            want = mx[:, :roi_h, :roi_w].sum(dim=(1, 2), dtype=torch.int64)
        else:
            want = (mx[:, :roi_h, :roi_w].double() * float(1 << fixed_shift)).round().long().sum(dim=(1, 2))
        assert (st[s:s + 32, :, 3].sum(1) == want.cpu().numpy()).all()
    for f in frames_vs_oracle:
        want = cport.argmax_stats(lg[f:f + 1].cpu().numpy(), roi_h, roi_w, threads=8)
        assert (labels[f].cpu().numpy() == want[0][0]).all() and (st[f].view(np.uint64) == want[1][0]).all()


def test_config3_preprocess_plus_postprocess_batch256(ssb):
    """1920x1080 camera frames -> 513x513 + post-process, batch 256 on one GPU."""
    b = 256
    g = torch.Generator(device="cuda"); g.manual_seed(3)
    frames = torch.randint(0, 256, (b, 1080, 1920, 3), dtype=torch.uint8, device="cuda", generator=g)
    geom = ssb.FrameGeometry(1920, 1080, 513, 513)
    rgb = ssb.preprocess_frames(frames, geom)
    assert (geom.roi_h, geom.roi_w) == (288, 513)
    assert not bool(rgb[:, 288:].any())                                           # ImageOps.expand padding
    yt = torch.from_numpy(geom.ytab.astype(np.int64)).cuda()
    xt = torch.from_numpy(geom.xtab.astype(np.int64)).cuda()
    for s in range(0, b, 64):                                                     # torch gather as a 2nd witness
        want = frames[s:s + 64][:, yt][:, :, xt].flip(-1)
        assert torch.equal(rgb[s:s + 64, :288], want)
    for f in (0, 100, 255):                                                       # oracle = Pillow-pinned restatement
        assert (rgb[f].cpu().numpy() == sp.preprocess_frame(frames[f].cpu().numpy(), 513, 513)[0]).all()
    del frames
    lg = torch.randint(0, 256, (b, 513, 513, 21), dtype=torch.uint8, device="cuda", generator=g)
    labels, stats = ssb.argmax_stats(lg, geom.roi_h, geom.roi_w)
    check_invariants(lg, labels, stats, geom.roi_h, geom.roi_w, (0, 255))


def test_config4_shard_of_4096_frames(ssb):
    """Batch 4096 of 513x513x21 frame-sharded over 8 GPUs: one rank's 512-frame shard (2.8 GB), chosen
    by the scheduler's own partition, processed in one launch."""
    sched = ssb.FrameScheduler(rank=5, world=8)
    s, e = sched.local_range(4096)
    assert (s, e) == (2560, 3072)
    g = torch.Generator(device="cuda"); g.manual_seed(4)
    lg = torch.randint(0, 256, (e - s, 513, 513, 21), dtype=torch.uint8, device="cuda", generator=g)
    labels, stats = sched.argmax_stats_local(lg, 513, 513)
    check_invariants(lg, labels, stats, 513, 513, (0, 511))


def test_config5_cityscapes_fp16_shard(ssb):
    """1024x2048x19 fp16, batch 512 over 8 GPUs: one rank's 64 frames (5.1 GB)."""
    b = 64
    g = torch.Generator(device="cuda"); g.manual_seed(5)
    lg = torch.randn((b, 1024, 2048, 19), dtype=torch.float16, device="cuda", generator=g)
    labels, stats = ssb.argmax_stats(lg)
    check_invariants(lg, labels, stats, 1024, 2048, (0, 63), fixed_shift=24)


def test_config1_single_frame_latency_path(ssb, golden):
    """Config 1: one frame end to end (identity-size camera -> K1 -> synthetic scores -> K2/K3 -> reply)."""
    import importlib
    srv = importlib.import_module("semantic-segmentation-server_b200.server")
    lab = golden("labels_colormap.npz")
    names = dict(zip(lab["label_ids"].tolist(), lab["label_names"].tolist()))
    frame = np.random.default_rng(0).integers(0, 256, (513, 513, 3), dtype=np.uint8)
    engine = srv.SyntheticEngine()
    srv.recognize_and_segment(True, srv.ArrayCamera([frame]), engine, names, 0.0, batch_frames=1, mode="labels")
    got = list(srv.segmented_objects)
    srv.segmented_objects.clear()
    rgb, ratio = sp.preprocess_frame(frame, 513, 513)
    assert ratio == (1.0, 1.0) and (rgb == frame[:, :, ::-1]).all()
    q, scale, zp = engine.run_logits(torch.from_numpy(rgb[None].copy()).cuda())
    _, st = cport.argmax_stats(q.cpu().numpy())
    want = sp.per_label_objects(st[0], 513, 513, names, 0.0, scale, zp)
    assert [o.label for o in got] == [w["label"] for w in want] and len(got) >= 1
    for o, w in zip(got, want):
        assert abs(o.area - w["area"]) <= 1e-6 * w["area"] and abs(o.centroid.cx - w["cx"]) <= 1e-6 * max(w["cx"], 1e-30)
