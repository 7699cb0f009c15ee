"""GPU parity tests: the CUDA path (through the C-ABI in libssb200.so) against the oracle and the
golden fixtures produced by the reference's own code.  Bit-exact for label maps and integer
statistics; floats derived from them within 1e-6 relative (BASELINE.json north_star)."""
import ctypes
import hashlib

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from oracle import cport, segpath as sp  # noqa: E402  (checker only)

REL_TOL = 1e-6   # north_star: scores and normalised centroids within 1e-6 relative


@pytest.fixture(scope="module")
def ssb():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ssb200
    return ssb200


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()[:16]


def run_gpu(ssb, logits_np, roi_h=None, roi_w=None, want_labels=True):
    lg = torch.from_numpy(logits_np).cuda()
    # poisoned output: torch.empty hands back the block of the previous call, which would hide an entry
    # the kernel forgot to write
    stats = torch.full((lg.shape[0], lg.shape[3], 4), -0x5A5A5A5A, dtype=torch.int64, device="cuda")
    labels, stats = ssb.argmax_stats(lg, roi_h, roi_w, stats=stats, want_labels=want_labels)
    torch.cuda.synchronize()
    return (labels.cpu().numpy() if want_labels else None), stats.cpu().numpy().view(np.uint64)


def check_vs_oracle(ssb, logits_np, roi_h=None, roi_w=None):
    labels, stats = run_gpu(ssb, logits_np, roi_h, roi_w)
    o_labels, o_stats = cport.argmax_stats(logits_np, roi_h, roi_w, threads=8)
    assert (labels == o_labels).all(), "label map differs at %d pixels" % int((labels != o_labels).sum())
    assert (stats == o_stats).all(), "stats differ:\n%s\n%s" % (stats[stats != o_stats][:8], o_stats[stats != o_stats][:8])
    return labels, stats


# ---- K2 + K3, u8 --------------------------------------------------------------------------------
def test_kat1_u8(ssb):
    q = np.random.default_rng(0).integers(0, 256, (1, 513, 513, 21), dtype=np.uint8)
    labels, stats = check_vs_oracle(ssb, q)
    assert digest(labels[0]) == "06ccd4046908a08f"
    assert stats[0, 0].tolist() == [13026, 3345589, 3334508, 3175759]
    assert stats[0, 20].tolist() == [12140, 3102486, 3113402, 2958649]
    n_labels, n_stats = sp.argmax_stats(q)                      # numpy restatement agrees too
    assert (labels == n_labels).all() and (stats == n_stats).all()


@pytest.mark.parametrize("batch", [2, 3, 5, 17])
def test_u8_batches_straddle_tiles(ssb, batch):
    """513*513 is odd: frames are misaligned to tiles and quads, the stream has a ragged tail."""
    q = np.random.default_rng(100 + batch).integers(0, 256, (batch, 513, 513, 21), dtype=np.uint8)
    check_vs_oracle(ssb, q)


@pytest.mark.parametrize("roi", [(1, 1), (513, 1), (1, 513), (288, 513), (384, 513), (513, 384), (200, 300)])
def test_u8_roi(ssb, roi):
    q = np.random.default_rng(7).integers(0, 256, (3, 513, 513, 21), dtype=np.uint8)
    labels, stats = check_vs_oracle(ssb, q, *roi)
    assert (stats[:, :, 0].sum(axis=1) == roi[0] * roi[1]).all()


def test_u8_blobby_long_runs(ssb):
    q = np.stack([sp.blobby_logits_u8(3 + i, 513, 513, 21) for i in range(4)])
    check_vs_oracle(ssb, q, 288, 513)


def test_u8_adversarial(ssb):
    h, w, c = 513, 513, 21
    a0 = np.full((2, h, w, c), 77, np.uint8)                      # every pixel tied -> label 0
    labels, stats = check_vs_oracle(ssb, a0)
    assert not labels.any()
    a1 = np.zeros((2, h, w, c), np.uint8); a1[..., 13] = 255      # one label everywhere
    check_vs_oracle(ssb, a1)
    a2 = np.zeros((1, h, w, c), np.uint8)                         # label = x mod C
    a2[0, :, np.arange(w), np.arange(w) % c] = 200
    check_vs_oracle(ssb, a2)
    a3 = np.zeros((1, h, w, c), np.uint8); a3[..., c - 1] = 1     # maximum only in the last channel
    labels, _ = check_vs_oracle(ssb, a3)
    assert (labels == c - 1).all()
    a4 = np.full((1, h, w, c), 255, np.uint8)                     # saturated, all tied
    check_vs_oracle(ssb, a4)


@pytest.mark.parametrize("shape", [(1, 7, 9, 5), (3, 16, 16, 21), (2, 33, 47, 21), (2, 64, 128, 19),
                                   (1, 1, 2, 21), (1, 3, 1, 21), (4, 40, 24, 2), (1, 300, 301, 19),
                                   (9, 31, 33, 21), (1, 1024, 2048, 19)])
def test_u8_shapes(ssb, shape):
    q = np.random.default_rng(sum(shape)).integers(0, 256, shape, dtype=np.uint8)
    check_vs_oracle(ssb, q)
    check_vs_oracle(ssb, q, max(1, shape[1] // 2), max(1, shape[2] - 1))


@pytest.mark.parametrize("c", list(range(2, 34)))
def test_u8_every_class_count_on_the_streaming_path(ssb, c, monkeypatch):
    """The streaming kernel is compiled for every class count from 2 to 32 (32-, 64- or 128-bit shared-memory
    loads by the parity of C); 33 takes the generic kernel.  Uniform scores, a narrow range with many ties, a
    crop that cuts rows and columns, frames that do not end on tile boundaries -- and the same bytes from the
    generic kernel."""
    rng = np.random.default_rng(1000 + c)
    b, h, w = int(rng.integers(2, 6)), int(rng.integers(40, 90)), int(rng.integers(37, 140))
    q = rng.integers(0, 256, (b, h, w, c), dtype=np.uint8)
    labels, stats = check_vs_oracle(ssb, q)
    check_vs_oracle(ssb, q, h - 7, w - 5)
    ties = rng.integers(0, 3, (b, h, w, c), dtype=np.uint8)
    check_vs_oracle(ssb, ties, h, w - 1)
    last = np.zeros((2, 64, 64, c), np.uint8)
    last[..., c - 1] = 9                                # the maximum only in the last channel
    check_vs_oracle(ssb, last)
    monkeypatch.setenv("SSB_ARGMAX_CFG", "generic")
    g_labels, g_stats = run_gpu(ssb, q)
    assert (g_labels == labels).all() and (g_stats == stats).all()


@pytest.mark.parametrize("shape", [(200, 8, 8, 21), (64, 3, 4, 21), (7, 4000, 4, 19), (5, 1, 3000, 21),
                                   (1000, 5, 5, 19), (3, 2, 1025, 21)])
def test_u8_extreme_geometries_on_the_streaming_path(ssb, shape):
    """Frames much smaller than a tile (many frame changes inside one tile), very narrow and very wide
    rows: the flat-stream kernel's frame bookkeeping and packed row counters."""
    q = np.random.default_rng(sum(shape)).integers(0, 256, shape, dtype=np.uint8)
    check_vs_oracle(ssb, q)
    check_vs_oracle(ssb, q, max(1, shape[1] - 1), max(1, shape[2] // 2))


def test_u8_narrow_value_range_many_ties(ssb):
    q = np.random.default_rng(5).integers(0, 3, (2, 513, 513, 21), dtype=np.uint8)
    check_vs_oracle(ssb, q)


def test_u8_kernel_configs_agree(ssb, monkeypatch):
    """Every compiled pipeline configuration and the generic kernel give identical bytes."""
    q = np.random.default_rng(11).integers(0, 256, (5, 513, 513, 21), dtype=np.uint8)
    want = cport.argmax_stats(q, 288, 513, threads=8)
    for cfg in ["4,2", "4,3", "4,4", "8,2", "8,3", "generic"]:
        monkeypatch.setenv("SSB_ARGMAX_CFG", cfg)
        labels, stats = run_gpu(ssb, q, 288, 513)
        assert (labels == want[0]).all() and (stats == want[1]).all(), cfg
    monkeypatch.setenv("SSB_ARGMAX_CFG", "7,7")
    with pytest.raises(ssb._cabi.SsbError):
        run_gpu(ssb, q)


def test_u8_misaligned_views_take_generic_path(ssb):
    """A view starting at frame 1 is not 16-byte aligned (frame = 5 526 549 B): same results."""
    q = np.random.default_rng(12).integers(0, 256, (3, 513, 513, 21), dtype=np.uint8)
    big = torch.from_numpy(q).cuda()
    view = big[1:]
    assert view.data_ptr() % 16 != 0
    labels, stats = ssb.argmax_stats(view)
    want = cport.argmax_stats(q[1:], threads=4)
    assert (labels.cpu().numpy() == want[0]).all()
    assert (stats.cpu().numpy().view(np.uint64) == want[1]).all()


def test_u8_labels_optional_and_stats_rezeroed(ssb):
    q = np.random.default_rng(13).integers(0, 256, (2, 513, 513, 21), dtype=np.uint8)
    lg = torch.from_numpy(q).cuda()
    stats = torch.full((2, 21, 4), 0x5555555555555555, dtype=torch.int64, device="cuda")   # garbage in
    _, s1 = ssb.argmax_stats(lg, stats=stats, want_labels=False)
    _, s2 = ssb.argmax_stats(lg, stats=stats, want_labels=False)        # second call: no accumulation
    want = cport.argmax_stats(q, threads=4, want_labels=False)[1]
    assert (s2.cpu().numpy().view(np.uint64) == want).all()


def test_u8_non_default_stream(ssb):
    q = np.random.default_rng(14).integers(0, 256, (2, 513, 513, 21), dtype=np.uint8)
    want = cport.argmax_stats(q, threads=4)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        lg = torch.from_numpy(q).cuda()
        labels, stats = ssb.argmax_stats(lg, stream=s)
    s.synchronize()
    assert (labels.cpu().numpy() == want[0]).all() and (stats.cpu().numpy().view(np.uint64) == want[1]).all()


def test_u8_config2_full_size_properties(ssb):
    """BASELINE config 2 (batch 64 of 513x513x21): size-independent invariants on every frame,
    full oracle comparison on a few frames."""
    b, h, w, c = 64, 513, 513, 21
    g = torch.Generator(device="cuda"); g.manual_seed(1234)
    lg = torch.randint(0, 256, (b, h, w, c), dtype=torch.uint8, device="cuda", generator=g)
    roi_h, roi_w = 288, 513
    labels, stats = ssb.argmax_stats(lg, roi_h, roi_w)
    st = stats.cpu().numpy().view(np.uint64).astype(np.int64)
    assert (st[:, :, 0].sum(1) == roi_h * roi_w).all()                             # every ROI pixel once
    assert (st[:, :, 1].sum(1) == roi_h * (roi_w * (roi_w - 1) // 2)).all()        # sum of all x
    assert (st[:, :, 2].sum(1) == roi_w * (roi_h * (roi_h - 1) // 2)).all()        # sum of all y
    mx = lg.amax(dim=-1)                                                            # torch as a 2nd witness
    assert (st[:, :, 3].sum(1) == mx[:, :roi_h, :roi_w].sum(dim=(1, 2), dtype=torch.int64).cpu().numpy()).all()
    win = torch.gather(lg, 3, labels.long().unsqueeze(-1)).squeeze(-1)
    assert bool((win == mx).all())                                                  # label points at a maximum
    first = (lg == mx.unsqueeze(-1)).to(torch.uint8).argmax(dim=-1)
    assert bool((first == labels.long()).all())                                     # ... the FIRST one
    for f in (0, 31, 63):
        want = cport.argmax_stats(lg[f:f + 1].cpu().numpy(), roi_h, roi_w, threads=4)
        assert (labels[f].cpu().numpy() == want[0][0]).all() and (st[f].view(np.uint64) == want[1][0]).all()


def test_u8_hypothesis_shapes(ssb):
    from hypothesis import given, settings, strategies as st_

    @settings(max_examples=40, deadline=None)
    @given(b=st_.integers(1, 4), h=st_.integers(1, 70), w=st_.integers(1, 70), c=st_.sampled_from([2, 3, 19, 21, 32]),
           seed=st_.integers(0, 2 ** 31), lowrange=st_.booleans(), data=st_.data())
    def prop(b, h, w, c, seed, lowrange, data):
        roi_h = data.draw(st_.integers(1, h)); roi_w = data.draw(st_.integers(1, w))
        q = np.random.default_rng(seed).integers(0, 4 if lowrange else 256, (b, h, w, c), dtype=np.uint8)
        check_vs_oracle(ssb, q, roi_h, roi_w)
    prop()


def test_u8_hypothesis_voc_width(ssb):
    """Fast path (C=21/19, wide frames) with random batch / height / ROI / value range."""
    from hypothesis import given, settings, strategies as st_

    @settings(max_examples=15, deadline=None)
    @given(b=st_.integers(1, 6), h=st_.integers(20, 200), w=st_.sampled_from([513, 512, 257, 640]),
           c=st_.sampled_from([19, 21]), seed=st_.integers(0, 2 ** 31), data=st_.data())
    def prop(b, h, w, c, seed, data):
        roi_h = data.draw(st_.integers(1, h)); roi_w = data.draw(st_.sampled_from([w, w, max(1, w // 2)]))
        q = np.random.default_rng(seed).integers(0, 256, (b, h, w, c), dtype=np.uint8)
        check_vs_oracle(ssb, q, roi_h, roi_w)
    prop()


# ---- fp16 -------------------------------------------------------------------------------------------
def test_kat3_f16(ssb):
    h = np.random.default_rng(2).standard_normal((1, 1024, 2048, 19), dtype=np.float32).astype(np.float16)
    labels, stats = check_vs_oracle(ssb, h)
    assert digest(labels[0]) == "7e7612187b8000af"
    assert int(stats[0, 0, 0]) == 109977 and int(stats[0, 18, 0]) == 110044


def test_f16_special_values(ssb):
    rng = np.random.default_rng(7)
    h = rng.standard_normal((2, 40, 50, 19), dtype=np.float32).astype(np.float16)
    u = h.view(np.uint16)
    u[0, 0, 0, :] = [0x8000, 0x0000] + [0xfc00] * 17          # -0 vs +0 tie -> first index
    u[0, 0, 1, :] = [0x0001, 0x0002, 0x8001] + [0x8400] * 16  # subnormals
    u[0, 1, :, :5] = rng.integers(0, 0x400, (50, 5))
    h[1, 2] *= np.float16(1e-3)
    h[1, 3] *= np.float16(1000)
    labels, stats = check_vs_oracle(ssb, h, 33, 44)
    n_labels, n_stats = sp.argmax_stats(h, 33, 44)
    assert (labels == n_labels).all() and (stats == n_stats).all()


@pytest.mark.parametrize("batch", [2, 3])
def test_f16_streaming_batches_and_roi(ssb, batch):
    """513x513 fp16 frames: odd pixel count, so frames straddle tiles and quads; ROI variants."""
    for c in (19, 21):
        h = (np.random.default_rng(50 + batch + c).standard_normal((batch, 513, 513, c), dtype=np.float32) * 4
             ).astype(np.float16)
        check_vs_oracle(ssb, h)
        check_vs_oracle(ssb, h, 288, 513)
        check_vs_oracle(ssb, h, 513, 384)


def test_f16_ties_and_infinities(ssb):
    rng = np.random.default_rng(8)
    h = rng.integers(-2, 3, (2, 200, 300, 19)).astype(np.float16)        # few distinct values: many ties
    h[0, :50][h[0, :50] == 0] = np.float16(-0.0)                         # mixed -0 / +0 maxima
    h[1, 10:20, :, 3] = np.float16(np.inf)
    h[1, 30:40, :, :] = np.float16(-np.inf)                              # all -inf -> label 0
    labels, stats = run_gpu(ssb, h)
    o_labels, o_stats = cport.argmax_stats(h, threads=4)
    assert (labels == o_labels).all()
    assert (stats[..., :3] == o_stats[..., :3]).all()                    # n, Σx, Σy exact everywhere
    assert (stats[0] == o_stats[0]).all()                                # Σq exact where winners are finite
    finite = np.ones(19, bool); finite[[0, 3]] = False                   # frame 1: labels 0 / 3 win with -inf / +inf
    assert (stats[1, finite, 3] == o_stats[1, finite, 3]).all()          # (infinite winners: Σq unspecified, ssb200.h)
    assert (labels[1, 10:20] == 3).all() and not labels[1, 30:40].any()
    n_labels = sp.argmax_labels(h[0])
    assert (labels[0] == n_labels).all()                                 # numpy: -0 == +0, first index wins


def test_f16_kernel_configs_agree(ssb, monkeypatch):
    h = (np.random.default_rng(51).standard_normal((2, 300, 400, 19), dtype=np.float32)).astype(np.float16)
    want = cport.argmax_stats(h, 250, 400, threads=8)
    for cfg in ["4,2", "2,2", "2,3", "4,3", "4,4", "8,2", "generic"]:
        monkeypatch.setenv("SSB_ARGMAX_F16_CFG", cfg)
        labels, stats = run_gpu(ssb, h, 250, 400)
        assert (labels == want[0]).all() and (stats == want[1]).all(), cfg


@pytest.mark.parametrize("shape", [(2, 64, 128, 19), (1, 33, 47, 21), (3, 17, 5, 4)])
def test_f16_shapes(ssb, shape):
    h = (np.random.default_rng(sum(shape)).standard_normal(shape, dtype=np.float32) * 8).astype(np.float16)
    check_vs_oracle(ssb, h)
    check_vs_oracle(ssb, h, max(1, shape[1] // 3), max(1, shape[2] // 2))


# ---- K1 pre-process ---------------------------------------------------------------------------------
@pytest.mark.parametrize("c", list(range(2, 34)))
def test_f16_every_class_count_on_the_streaming_path(ssb, c, monkeypatch):
    """fp16 scores, every class count the streaming kernel is compiled for (2..32; 33: generic kernel): normal
    scores, ties with both zeros, a crop; and the same bytes from the generic kernel."""
    rng = np.random.default_rng(2000 + c)
    b, h, w = int(rng.integers(2, 5)), int(rng.integers(40, 80)), int(rng.integers(37, 120))
    x = (rng.standard_normal((b, h, w, c)) * 4).astype(np.float16)
    labels, stats = check_vs_oracle(ssb, x)
    check_vs_oracle(ssb, x, h - 7, w - 5)
    ties = rng.choice(np.array([-1.0, -0.0, 0.0, 1.0], np.float16), size=(b, h, w, c))
    check_vs_oracle(ssb, ties, h, w - 1)
    monkeypatch.setenv("SSB_ARGMAX_F16_CFG", "generic")
    g_labels, g_stats = run_gpu(ssb, x)
    assert (g_labels == labels).all() and (g_stats == stats).all()


@pytest.mark.parametrize("shape", [(200, 8, 8, 19), (64, 3, 4, 21), (7, 4000, 4, 19), (5, 1, 3000, 21),
                                   (600, 5, 5, 19), (3, 2, 1025, 21)])
def test_f16_extreme_geometries_on_the_streaming_path(ssb, shape):
    """The fp16 codec through the same frame bookkeeping: frames much smaller than a tile, very narrow
    and very wide rows, ROI cut both ways."""
    h = np.random.default_rng(sum(shape) + 1).standard_normal(shape, dtype=np.float32).astype(np.float16)
    check_vs_oracle(ssb, h)
    check_vs_oracle(ssb, h, max(1, shape[1] - 1), max(1, shape[2] // 2))


def gpu_preprocess(ssb, frames_np, W, H, keep=True):
    b, ch, cw, _ = frames_np.shape
    geom = ssb.FrameGeometry(cw, ch, W, H, keep)
    out = ssb.preprocess_frames(torch.from_numpy(frames_np).cuda(), geom)
    torch.cuda.synchronize()
    return out.cpu().numpy(), geom


def test_preprocess_matches_reference_fixtures(ssb, golden):
    g = golden("preprocess.npz")
    for name in sorted(k[:-5] for k in g.files if k.endswith("_geom")):
        cw, ch, W, H, seed, keep = (int(v) for v in g[name + "_geom"])
        frame = np.random.default_rng(seed).integers(0, 256, (ch, cw, 3), dtype=np.uint8)
        out, geom = gpu_preprocess(ssb, frame[None], W, H, bool(keep))
        assert digest(out[0]) == str(g[name + "_digest"]), name
        if name + "_out" in g.files:
            assert (out[0] == g[name + "_out"]).all(), name


def test_preprocess_kat2_and_batch(ssb):
    frames = np.random.default_rng(1).integers(0, 256, (1, 1080, 1920, 3), dtype=np.uint8)
    out, geom = gpu_preprocess(ssb, frames, 513, 513)
    assert digest(out[0]) == "b8c5ed5577aafd67"
    assert (geom.roi_h, geom.roi_w) == (288, 513) and not out[0, 288:].any()
    more = np.random.default_rng(21).integers(0, 256, (5, 480, 640, 3), dtype=np.uint8)
    out, geom = gpu_preprocess(ssb, more, 513, 513)
    for i in range(5):
        assert (out[i] == sp.preprocess_frame(more[i], 513, 513)[0]).all()


@pytest.mark.parametrize("cam", [(101, 37), (37, 101), (5, 5), (641, 479), (1, 9), (1000, 3)])
def test_preprocess_ragged_sizes(ssb, cam):
    """Row lengths that are not multiples of 16 bytes: head/tail byte paths on both sides."""
    cw, ch = cam
    frames = np.random.default_rng(cw * 7 + ch).integers(0, 256, (3, ch, cw, 3), dtype=np.uint8)
    for (W, H) in [(33, 33), (40, 24), (513, 513)]:
        try:
            want = [sp.preprocess_frame(f, W, H)[0] for f in frames]
        except ValueError:
            with pytest.raises(ValueError):
                ssb.FrameGeometry(cw, ch, W, H)
            continue
        out, _ = gpu_preprocess(ssb, frames, W, H)
        assert (out == np.stack(want)).all()
    out, _ = gpu_preprocess(ssb, frames, 33, 33, keep=False)                     # stretch branch
    assert (out == np.stack([sp.preprocess_frame(f, 33, 33, False)[0] for f in frames])).all()


def test_preprocess_hypothesis_sizes(ssb):
    from hypothesis import given, settings, strategies as st_

    @settings(max_examples=40, deadline=None)
    @given(cw=st_.integers(1, 300), ch=st_.integers(1, 200), W=st_.integers(1, 150), H=st_.integers(1, 150),
           b=st_.integers(1, 3), keep=st_.booleans(), seed=st_.integers(0, 2 ** 31))
    def prop(cw, ch, W, H, b, keep, seed):
        frames = np.random.default_rng(seed).integers(0, 256, (b, ch, cw, 3), dtype=np.uint8)
        try:
            want = np.stack([sp.preprocess_frame(f, W, H, keep)[0] for f in frames])
        except ValueError:                                   # frame collapses to zero pixels
            with pytest.raises(ValueError):
                ssb.FrameGeometry(cw, ch, W, H, keep)
            return
        out, _ = gpu_preprocess(ssb, frames, W, H, keep)
        assert (out == want).all()
    prop()


def test_preprocess_strided_frames_and_views(ssb):
    """Frames embedded in a larger buffer (frame_stride > Hc*Wc*3) and output written into a view."""
    lib = ssb._cabi.lib
    rng = np.random.default_rng(44)
    b, ch, cw, W, H = 4, 37, 53, 40, 30
    frames = rng.integers(0, 256, (b, ch, cw, 3), dtype=np.uint8)
    stride = ch * cw * 3 + 37                                # odd padding between frames
    buf = torch.zeros(b * stride + 5, dtype=torch.uint8, device="cuda")
    for i in range(b):
        buf[3 + i * stride: 3 + i * stride + ch * cw * 3] = torch.from_numpy(frames[i].ravel()).cuda()
    geom = ssb.FrameGeometry(cw, ch, W, H)
    xt, yt = geom.device_tables(torch.device("cuda", 0))
    out = torch.full((b * H * W * 3 + 7,), 0xAB, dtype=torch.uint8, device="cuda")
    rc = lib.ssb_preprocess_nearest_u8(buf.data_ptr() + 3, b, ch, cw, stride, xt.data_ptr(), yt.data_ptr(), geom.new_w,
                                       geom.new_h, out.data_ptr() + 7, H, W, None)
    assert rc == 0, lib.ssb_last_error()
    torch.cuda.synchronize()
    got = out[7:].cpu().numpy().reshape(b, H, W, 3)
    assert (got == np.stack([sp.preprocess_frame(f, W, H)[0] for f in frames])).all()
    assert (out[:7].cpu().numpy() == 0xAB).all()             # nothing written in front of the view


# ---- foreground mask + reference-literal objects (f1) --------------------------------------------------
def test_fg_mask_matches_reference_fixtures(ssb, golden):
    g = golden("objects.npz")
    for name in sorted(k[:-10] for k in g.files if k.endswith("_label_map")):
        lm, fm = g[name + "_label_map"], g[name + "_fgmask"]
        roi_h, roi_w = fm.shape
        mask = ssb.foreground_mask(torch.from_numpy(lm[None]).cuda(), roi_h, roi_w).cpu().numpy()[0]
        assert (mask[:roi_h, :roi_w] == fm).all(), name
        assert not mask[roi_h:].any() and not mask[:, roi_w:].any()


@pytest.mark.parametrize("shape,roi", [((3, 513, 513), (288, 513)), ((2, 513, 513), (513, 513)), ((2, 70, 131), (69, 130)),
                                       ((1, 9, 5), (9, 5)), ((2, 64, 128), (1, 128)), ((1, 200, 300), (200, 1)),
                                       ((1, 130, 257), (100, 200))])
def test_fg_mask_random_maps_and_rois(ssb, shape, roi):
    """Noise (every 3x3 window mixes labels) and blocky maps, strips that straddle the ROI's bottom and right
    edges, single-row and single-column ROIs."""
    rng = np.random.default_rng(shape[1] * 7 + roi[0])
    cm = sp.pascal_colormap().astype(np.uint8)
    for kind in range(2):
        if kind == 0:
            lm = rng.integers(0, 21, shape).astype(np.uint8)
        else:
            small = rng.choice([0, 7, 15, 12], size=(shape[0], shape[1] // 6 + 1, shape[2] // 6 + 1))
            lm = np.kron(small, np.ones((1, 6, 6), int))[:, :shape[1], :shape[2]].astype(np.uint8)
        mask = ssb.foreground_mask(torch.from_numpy(lm).cuda(), roi[0], roi[1]).cpu().numpy()
        assert (mask == cport.fg_mask(lm, roi[0], roi[1], cm)).all()


@pytest.mark.parametrize("shape,roi", [((2, 80, 513), (80, 513)), ((1, 75, 520), (75, 519)), ((1, 75, 521), (75, 521)),
                                       ((2, 150, 260), (149, 258)), ((1, 30, 130), (30, 129)), ((1, 30, 131), (30, 2)),
                                       ((1, 77, 256), (77, 253)), ((1, 77, 256), (76, 254)), ((1, 77, 256), (77, 255)),
                                       ((1, 77, 256), (2, 256)), ((1, 40, 7), (40, 7)), ((1, 145, 384), (145, 129)),
                                       ((1, 74, 300), (73, 128)), ((3, 513, 513), (288, 512))])
def test_fg_mask_strip_and_roi_borders(ssb, shape, roi):
    """Frame widths with and without narrow last columns (W mod 128 <= 8), ROIs whose last column falls on each of a
    lane's four pixels and on strip borders, two-column and two-row ROIs; the output buffer starts as garbage (zeros
    outside the ROI are the kernel's job)."""
    rng = np.random.default_rng(shape[2] * 5 + roi[1])
    cm = sp.pascal_colormap().astype(np.uint8)
    for kind in range(2):
        if kind == 0:
            lm = rng.choice([0, 7, 15, 12, 20], size=shape).astype(np.uint8)
        else:
            small = rng.choice([0, 7, 15, 12], size=(shape[0], shape[1] // 5 + 1, shape[2] // 5 + 1))
            lm = np.kron(small, np.ones((1, 5, 5), int))[:, :shape[1], :shape[2]].astype(np.uint8)
        out = torch.full(shape, 0xAB, dtype=torch.uint8, device="cuda")
        ssb.foreground_mask(torch.from_numpy(lm).cuda(), roi[0], roi[1], out=out)
        assert (out.cpu().numpy() == cport.fg_mask(lm, roi[0], roi[1], cm)).all()


def test_fg_mask_colours_outside_the_packed_range(ssb):
    """Labels >= 64 have colours that are not multiples of 64 (and a caller may pass any colormap): those
    box sums are recomputed from the colormap bytes."""
    rng = np.random.default_rng(8)
    cm = sp.pascal_colormap().astype(np.uint8)
    lm = rng.integers(0, 256, (2, 120, 140)).astype(np.uint8)
    lm[1, 30:90, 20:120] = 15                        # a region where only the rim meets such labels
    mask = ssb.foreground_mask(torch.from_numpy(lm).cuda(), 110, 133).cpu().numpy()
    assert (mask == cport.fg_mask(lm, 110, 133, cm)).all()
    odd = rng.integers(0, 256, (256, 3)).astype(np.uint8)      # an arbitrary colormap
    lm2 = rng.integers(0, 21, (1, 64, 64)).astype(np.uint8)
    mask2 = ssb.foreground_mask(torch.from_numpy(lm2).cuda(), 64, 64, torch.from_numpy(odd).cuda()).cpu().numpy()
    assert (mask2 == cport.fg_mask(lm2, 64, 64, odd)).all()


def test_reference_literal_objects_from_gpu_mask(ssb, golden):
    """sem_seg_server.py:173-195 reproduced from GPU products: label map (K2) -> mask (GPU) ->
    contours/moments (cv2 on host) -> normalised reply values; bit-exact in FP32."""
    g = golden("objects.npz")
    lab = golden("labels_colormap.npz")
    names = dict(zip(lab["label_ids"].tolist(), lab["label_names"].tolist()))
    for name in sorted(k[:-5] for k in g.files if k.endswith("_geom")):
        cw, ch, W, H = (int(v) for v in g[name + "_geom"][:4])
        geom = ssb.FrameGeometry(cw, ch, W, H)
        lm = g[name + "_label_map"]
        # logits whose argmax is the fixture's label map
        q = np.random.default_rng(1).integers(0, 100, (1, H, W, 21), dtype=np.uint8)
        np.put_along_axis(q[0], lm[..., None].astype(np.int64), 200, -1)
        labels, _ = ssb.argmax_stats(torch.from_numpy(q).cuda(), geom.roi_h, geom.roi_w)
        assert (labels.cpu().numpy()[0] == lm).all()
        mask = ssb.foreground_mask(labels, geom.roi_h, geom.roi_w).cpu().numpy()[0]
        seg = labels.cpu().numpy()[0][:geom.roi_h, :geom.roi_w]
        objs = ssb.objects_from_contours(mask[:geom.roi_h, :geom.roi_w], seg,
                                         float(g[name + "_min_area_ratio"]) * H * W, names)
        got = [ssb.normalise(o, W, H) for o in objs][::-1]                         # deque is newest-first
        assert [o[0] for o in got] == g[name + "_obj_names"].tolist(), name
        vals = np.array([o[1:] for o in got], dtype=np.float32).reshape(-1, 4)
        assert (vals == g[name + "_obj_vals"]).all(), name


def test_per_label_objects_floats(ssb):
    """Scores / normalised centroids from GPU integers vs the oracle's floats (1e-6 relative)."""
    q = np.stack([sp.blobby_logits_u8(30 + i, 513, 513, 21, cell=64, bg_frac=0.5) for i in range(2)])
    _, stats = run_gpu(ssb, q, 288, 513)
    names = {i: "l%d" % i for i in range(21)}
    for f in range(2):
        objs, conf = ssb.objects_from_stats(stats[f], 513, 513, names, 0.01, scale=0.1, zero_point=128,
                                            only_foreground=False)
        want = sp.per_label_objects(cport.argmax_stats(q[f:f + 1], 288, 513)[1][0], 513, 513, names, 0.01,
                                    scale=0.1, zero_point=128, only_foreground=False)[::-1]
        assert len(objs) == len(want) and len(objs) >= 2
        for o, cf, w in zip(objs, conf, want):
            got = ssb.normalise(o, 513, 513)
            for a, b in zip(got[1:] + (cf,), (w["score"], w["area"], w["cx"], w["cy"], w["score_conf"])):
                assert abs(a - b) <= REL_TOL * max(abs(b), 1e-30)


# ---- host-buffer pipeline (C-ABI with HOST pointers) ---------------------------------------------------
def test_host_pipeline_matches_device_path(ssb):
    q = np.random.default_rng(40).integers(0, 256, (11, 513, 513, 21), dtype=np.uint8)
    want = cport.argmax_stats(q, 288, 513, threads=8)
    pipe = ssb.HostPipeline(513, 513, 21, np.uint8, frames_per_chunk=4, n_slots=2)
    labels, stats = pipe.argmax_stats(q, 288, 513)
    assert (labels == want[0]).all() and (stats == want[1]).all()
    _, stats2 = pipe.argmax_stats(q[:3], want_labels=False)                       # reuse, other ROI
    assert (stats2 == cport.argmax_stats(q[:3], threads=4)[1]).all()
    pipe.close()
    h = np.random.default_rng(41).standard_normal((3, 64, 128, 19), dtype=np.float32).astype(np.float16)
    pipe = ssb.HostPipeline(64, 128, 19, np.float16, frames_per_chunk=2, n_slots=2)
    labels, stats = pipe.argmax_stats(h)
    want = cport.argmax_stats(h)
    assert (labels == want[0]).all() and (stats == want[1]).all()


def test_ingest_pipeline_host_frames(ssb):
    """Row f3: camera frames in host memory -> model inputs, chunked over streams."""
    frames = np.random.default_rng(43).integers(0, 256, (11, 480, 640, 3), dtype=np.uint8)
    geom = ssb.FrameGeometry(640, 480, 513, 513)
    pipe = ssb.IngestPipeline(geom, frames_per_chunk=4, n_slots=2)
    out = pipe.preprocess(frames)
    for i in range(11):
        assert (out[i] == sp.preprocess_frame(frames[i], 513, 513)[0]).all()
    out2 = pipe.preprocess(frames[:3])                    # reuse
    assert (out2 == out[:3]).all()
    pipe.close()
    with pytest.raises(ValueError):
        ssb.IngestPipeline(geom).preprocess(frames[:, :100])


def test_raw_cabi_call_with_ctypes(ssb):
    """The call a non-torch host would make: raw device pointers from cudaMalloc via the runtime."""
    q = np.random.default_rng(42).integers(0, 256, (2, 65, 65, 21), dtype=np.uint8)
    lg = torch.from_numpy(q).cuda()
    labels = torch.empty((2, 65, 65), dtype=torch.uint8, device="cuda")
    stats = torch.empty((2, 21, 4), dtype=torch.int64, device="cuda")
    rc = ssb._cabi.lib.ssb_argmax_stats_u8(ctypes.c_void_p(lg.data_ptr()), 2, 65, 65, 21, 65, 65,
                                           ctypes.c_void_p(labels.data_ptr()), ctypes.c_void_p(stats.data_ptr()),
                                           ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    torch.cuda.synchronize()
    want = cport.argmax_stats(q)
    assert (labels.cpu().numpy() == want[0]).all() and (stats.cpu().numpy().view(np.uint64) == want[1]).all()
