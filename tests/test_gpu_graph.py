"""The device-pointer entry points recorded into a CUDA graph (torch.cuda.graph captures the stream the
calls are made on): replays must give the bytes an eager call gives, also with eager calls in between."""
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


def test_whole_frame_path_recorded_and_replayed(ssb):
    """camera frames -> K1, class scores -> K2+K3 -> K4, one graph, three replays on fresh inputs."""
    b, cam_h, cam_w = 3, 480, 640
    geom = ssb.FrameGeometry(cam_w, cam_h, 513, 513)
    rng = np.random.default_rng(77)
    frames = torch.zeros((b, cam_h, cam_w, 3), dtype=torch.uint8, device="cuda")
    q = torch.zeros((b, 513, 513, 21), dtype=torch.uint8, device="cuda")
    rgb = torch.empty((b, 513, 513, 3), dtype=torch.uint8, device="cuda")
    labels = torch.empty((b, 513, 513), dtype=torch.uint8, device="cuda")
    stats = torch.empty((b, 21, 4), dtype=torch.int64, device="cuda")
    mask = torch.empty_like(labels)
    cm = torch.from_numpy(ssb.create_pascal_label_colormap().astype(np.uint8)).cuda()

    def run():
        ssb.preprocess_frames(frames, geom, out=rgb)
        ssb.argmax_stats(q, geom.roi_h, geom.roi_w, labels=labels, stats=stats)
        ssb.foreground_mask(labels, geom.roi_h, geom.roi_w, cm, out=mask)

    run()                                   # once outside capture: sizes the scratch a recorded call will own
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        run()
    for rep in range(3):
        f_np = rng.integers(0, 256, (b, cam_h, cam_w, 3), dtype=np.uint8)
        q_np = rng.integers(0, 256, (b, 513, 513, 21), dtype=np.uint8)
        frames.copy_(torch.from_numpy(f_np))
        q.copy_(torch.from_numpy(q_np))
        stats.fill_(-1)
        graph.replay()
        if rep == 1:                        # eager traffic on the same device between replays
            ssb.argmax_stats(torch.from_numpy(rng.integers(0, 256, (5, 513, 513, 21), dtype=np.uint8)).cuda())
        torch.cuda.synchronize()
        want_labels, want_stats = cport.argmax_stats(q_np, geom.roi_h, geom.roi_w, threads=4)
        assert (labels.cpu().numpy() == want_labels).all()
        assert (stats.cpu().numpy().view(np.uint64) == want_stats).all()
        for i in range(b):
            assert (rgb[i].cpu().numpy() == sp.preprocess_frame(f_np[i], 513, 513)[0]).all()
        cmap = sp.pascal_colormap().astype(np.uint8)
        assert (mask.cpu().numpy() == cport.fg_mask(want_labels, geom.roi_h, geom.roi_w, cmap)).all()


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16])
def test_planar_entry_recorded_and_replayed(ssb, dtype):
    x = torch.zeros((2, 21, 513, 513), dtype=dtype, device="cuda")
    labels = torch.empty((2, 513, 513), dtype=torch.uint8, device="cuda")
    stats = torch.empty((2, 21, 4), dtype=torch.int64, device="cuda")
    ssb.argmax_stats_nchw(x, 288, 513, labels=labels, stats=stats)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        ssb.argmax_stats_nchw(x, 288, 513, labels=labels, stats=stats)
    for seed in (1, 2):
        x_np = np.random.default_rng(seed).standard_normal((2, 21, 513, 513)).astype(np.float32)
        x.copy_(torch.from_numpy(x_np).to(dtype))
        graph.replay()
        torch.cuda.synchronize()
        want_labels, want_stats = sp.argmax_stats_nchw(x.cpu().numpy(), 288, 513)
        assert (labels.cpu().numpy() == want_labels).all()
        assert (stats.cpu().numpy().view(np.uint64) == want_stats).all()


@pytest.mark.filterwarnings("ignore:The CUDA Graph is empty")
def test_recording_a_larger_batch_than_ever_run_is_refused(ssb):
    big = torch.zeros((700, 16, 16, 21), dtype=torch.uint8, device="cuda")     # scratch of 700 frames: more than any test ran eagerly
    graph = torch.cuda.CUDAGraph()
    with pytest.raises(ssb._cabi.SsbError, match="outside"):
        with torch.cuda.graph(graph):
            ssb.argmax_stats(big)
    ssb.argmax_stats(big)                   # the library is still usable afterwards
    torch.cuda.synchronize()
