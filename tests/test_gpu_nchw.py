"""GPU parity tests of the planar ([B,C,H,W]) argmax + statistics entry points against the oracle
(row f4 of the scope table: class scores left on the GPU by a convolutional head).  Bit-exact label
maps and integer statistics; the fp32 score sum is the sum of round-half-even(value * 2**24)."""
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


def run_gpu(ssb, planar_np, roi_h=None, roi_w=None, want_labels=True):
    lg = torch.from_numpy(planar_np).cuda()
    stats = torch.full((lg.shape[0], lg.shape[1], 4), -0x5A5A5A5A, dtype=torch.int64, device="cuda")   # poisoned
    labels, stats = ssb.argmax_stats_nchw(lg, roi_h, roi_w, stats=stats, want_labels=want_labels)
    torch.cuda.synchronize()
    return (labels.cpu().numpy() if want_labels else None), stats.cpu().numpy().view(np.uint64)


def oracle(planar_np, roi_h=None, roi_w=None):
    if planar_np.dtype == np.float16:   # the C port is much faster than numpy on full-size frames
        return cport.argmax_stats(np.ascontiguousarray(planar_np.transpose(0, 2, 3, 1)), roi_h, roi_w, threads=8)
    return sp.argmax_stats_nchw(planar_np, roi_h, roi_w)


def check(ssb, planar_np, roi_h=None, roi_w=None):
    labels, stats = run_gpu(ssb, planar_np, roi_h, roi_w)
    o_labels, o_stats = oracle(planar_np, roi_h, roi_w)
    assert (labels == o_labels).all(), "label map differs at %d pixels" % int((labels != o_labels).sum())
    assert (stats == o_stats).all(), "stats differ:\n%s\n%s" % (stats[stats != o_stats][:8], o_stats[stats != o_stats][:8])
    return labels, stats


def normal(seed, shape, dtype, scale=1.0):
    return (np.random.default_rng(seed).standard_normal(shape) * scale).astype(dtype)


@pytest.mark.parametrize("dtype", [np.float32, np.float16])
@pytest.mark.parametrize("batch", [1, 2, 3])
def test_voc_frames(ssb, dtype, batch):
    """513*513 is odd: plane starts have every 16-byte phase and the tensor ends off a 16-byte
    boundary (the last pixels of the last frame go through the plain-load tail)."""
    x = normal(40 + batch, (batch, 21, 513, 513), dtype)
    labels, stats = check(ssb, x)
    assert int(stats[:, :, 0].sum()) == batch * 513 * 513


@pytest.mark.parametrize("dtype", [np.float32, np.float16])
@pytest.mark.parametrize("roi", [(1, 1), (513, 1), (1, 513), (288, 513), (513, 384), (200, 300)])
def test_roi(ssb, dtype, roi):
    x = normal(7, (2, 21, 513, 513), dtype)
    _, stats = check(ssb, x, *roi)
    assert int(stats[:, :, 0].sum()) == 2 * roi[0] * roi[1]


@pytest.mark.parametrize("dtype", [np.float32, np.float16])
def test_ties_first_index_wins(ssb, dtype):
    x = np.random.default_rng(3).integers(-2, 3, (2, 21, 513, 513)).astype(dtype)   # ~20 % of the scores tie for the maximum
    x[0, :, :100] = 0                      # whole rows of all-equal scores, incl. -0 vs +0
    x[0, 1::2, :100] = -0.0
    labels, _ = check(ssb, x)
    assert (labels[0, :100] == 0).all()


@pytest.mark.parametrize("dtype", [np.float32, np.float16])
def test_adversarial_label_layouts(ssb, dtype):
    h, w, c = 513, 513, 21
    x = np.zeros((4, c, h, w), dtype)
    x[0, 20] = 1                                              # one label everywhere: maximum contention
    x[1] = (np.arange(c)[:, None, None] == (np.arange(w) % c)[None, None, :])   # label = x mod C
    x[2, :, :, :] = -1
    x[2, 20] = -0.5                                           # negative scores, maximum in the last plane
    yy, xx = np.mgrid[0:h, 0:w]
    x[3] = (np.arange(c)[:, None, None] == ((yy // 37 + xx // 53) % c)[None]) * 3.0   # blocks of labels
    check(ssb, x.astype(dtype), 288, 513)


def test_f32_score_sum_rounds_half_even(ssb):
    """Values that are not multiples of 2**-24: each winner is rounded on its own, then summed."""
    x = normal(9, (1, 21, 513, 513), np.float32, scale=1e-4)
    x[0, :, 0, :8] = -1.0
    x[0, 3, 0, :8] = np.array([0.5, 1.5, 2.5, 3.5, -0.5, 4.5, 1.25, 2.75], np.float32) * np.float32(2.0 ** -24)   # exact halves
    check(ssb, x)


@pytest.mark.parametrize("dtype", [np.float32, np.float16])
@pytest.mark.parametrize("shape", [(1, 5, 7, 9), (3, 21, 16, 16), (2, 21, 33, 47), (2, 19, 64, 128), (1, 19, 1024, 2048),
                                   (7, 21, 4000, 4), (5, 21, 1, 3000), (64, 21, 3, 4), (200, 21, 8, 8), (2, 21, 4, 4),
                                   (1, 21, 2, 9), (3, 2, 40, 50), (1, 255, 9, 11)])
def test_shapes(ssb, dtype, shape):
    """Unrolled kernel (C = 19/21) and the generic kernel (everything else); odd and even H*W."""
    check(ssb, normal(sum(shape), shape, dtype), max(1, shape[2] - 1), max(1, shape[3] - 2))


@pytest.mark.parametrize("dtype", [np.float32, np.float16])
def test_kernel_configs_agree(ssb, dtype, monkeypatch):
    x = normal(12, (5, 21, 513, 513), dtype)
    want = oracle(x, 288, 513)
    for cfg in ["8", "4", "generic"]:
        monkeypatch.setenv("SSB_ARGMAX_NCHW_CFG", cfg)
        labels, stats = run_gpu(ssb, x, 288, 513)
        assert (labels == want[0]).all() and (stats == want[1]).all(), cfg
    monkeypatch.setenv("SSB_ARGMAX_NCHW_CFG", "7")
    with pytest.raises(ssb._cabi.SsbError):
        run_gpu(ssb, x)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16])
def test_tensor_starting_one_element_into_an_allocation(ssb, dtype):
    """Only element alignment is required (an fp16 tensor at 2 mod 4 bytes cannot use pair loads)."""
    n = 2 * 21 * 65 * 33
    buf = torch.randn(n + 1, device="cuda").to(dtype)
    lg = buf[1:].view(2, 21, 65, 33)
    labels, stats = ssb.argmax_stats_nchw(lg, 60, 33)
    aligned = lg.clone()
    labels2, stats2 = ssb.argmax_stats_nchw(aligned, 60, 33)
    torch.cuda.synchronize()
    assert torch.equal(labels, labels2) and torch.equal(stats, stats2)
    o_labels, o_stats = oracle(lg.cpu().numpy(), 60, 33)
    assert (labels.cpu().numpy() == o_labels).all() and (stats.cpu().numpy().view(np.uint64) == o_stats).all()


def test_labels_optional_stats_rezeroed_and_repeatable(ssb):
    x = torch.from_numpy(normal(21, (3, 21, 513, 513), np.float32)).cuda()
    stats = torch.full((3, 21, 4), -1, dtype=torch.int64, device="cuda")
    _, s1 = ssb.argmax_stats_nchw(x, stats=stats, want_labels=False)
    first = s1.clone()
    for _ in range(3):                       # scratch is zero at rest: repeated calls give the same bytes
        labels, s2 = ssb.argmax_stats_nchw(x, stats=stats)
        assert torch.equal(first, s2)
    o_labels, o_stats = oracle(x.cpu().numpy())
    assert (labels.cpu().numpy() == o_labels).all() and (first.cpu().numpy().view(np.uint64) == o_stats).all()


def test_agrees_with_nhwc_entry_on_transposed_fp16(ssb):
    """The planar and the interleaved fp16 kernels are two independent implementations."""
    x = torch.from_numpy(normal(33, (4, 19, 256, 512), np.float16)).cuda()
    l1, s1 = ssb.argmax_stats_nchw(x, 200, 500)
    l2, s2 = ssb.argmax_stats(x.permute(0, 2, 3, 1).contiguous(), 200, 500)
    torch.cuda.synchronize()
    assert torch.equal(l1, l2) and torch.equal(s1, s2)


def test_full_batch_agrees_with_interleaved_kernel_and_invariants(ssb):
    """Config-2 sized batch (64 x 21 x 513 x 513 fp16, 707 MB): the planar kernel against the interleaved
    one on the permuted tensor, plus size-independent properties of the statistics."""
    torch.manual_seed(5)
    x = torch.randn((64, 21, 513, 513), device="cuda", dtype=torch.float16)
    labels, stats = ssb.argmax_stats_nchw(x, 288, 513)
    l2, s2 = ssb.argmax_stats(x.permute(0, 2, 3, 1).contiguous(), 288, 513)
    torch.cuda.synchronize()
    assert torch.equal(labels, l2) and torch.equal(stats, s2)
    n = stats[:, :, 0]
    assert bool((n.sum(dim=1) == 288 * 513).all())
    assert int(stats[:, :, 1].sum()) == 64 * 288 * (513 * 512 // 2)           # every column index, 288 times per frame
    assert int(stats[:, :, 2].sum()) == 64 * 513 * (288 * 287 // 2)
    want_labels, want_stats = oracle(x[:2].cpu().numpy(), 288, 513)            # and the oracle on two of the frames
    assert (labels[:2].cpu().numpy() == want_labels).all()
    assert (stats[:2].cpu().numpy().view(np.uint64) == want_stats).all()


def test_hypothesis_shapes(ssb):
    """Random batch / frame size / ROI / class count / dtype / value range (ties when the range is narrow):
    frames smaller than a tile (several frames per CTA), odd and even H*W (shifted-pair modes), the
    generic kernel for other class counts."""
    from hypothesis import given, settings, strategies as st_

    @settings(max_examples=60, deadline=None)
    @given(b=st_.integers(1, 5), h=st_.integers(1, 90), w=st_.integers(1, 150), c=st_.sampled_from([2, 19, 21, 21, 33]),
           half=st_.booleans(), seed=st_.integers(0, 2 ** 31), narrow=st_.booleans(), data=st_.data())
    def prop(b, h, w, c, half, seed, narrow, data):
        roi_h = data.draw(st_.integers(1, h)); roi_w = data.draw(st_.integers(1, w))
        rng = np.random.default_rng(seed)
        x = rng.integers(-2, 3, (b, c, h, w)).astype(np.float32) if narrow else rng.standard_normal((b, c, h, w)).astype(np.float32)
        check(ssb, x.astype(np.float16) if half else x, roi_h, roi_w)
    prop()


def test_argument_errors(ssb):
    x = torch.zeros((1, 21, 8, 8), device="cuda")
    with pytest.raises(ssb._cabi.SsbError):
        ssb.argmax_stats_nchw(x, 9, 8)
    with pytest.raises(ValueError):
        ssb.argmax_stats_nchw(x.permute(0, 2, 3, 1))          # not contiguous
    with pytest.raises(ValueError):
        ssb.argmax_stats_nchw(x.to(torch.float64))


@pytest.mark.parametrize("dtype", [np.float32, np.float16])
@pytest.mark.parametrize("c", list(range(2, 34)))
def test_every_class_count(ssb, dtype, c, monkeypatch):
    """The unrolled planar kernels are compiled for every class count from 2 to 32 (33: the generic kernel).  Odd
    H*W so that planes alternate between even and odd element phases (the fp16 pair loads), a crop that cuts rows and
    columns, many ties; and the same bytes from the generic kernel."""
    rng = np.random.default_rng(500 + c)
    b, h, w = int(rng.integers(2, 5)), int(rng.integers(35, 80)) | 1, int(rng.integers(41, 120)) | 1
    x = (rng.standard_normal((b, c, h, w)) * 3).astype(dtype)
    labels, stats = check(ssb, x)
    check(ssb, x, h - 6, w - 3)
    ties = rng.integers(-1, 2, (b, c, h, w)).astype(dtype)
    check(ssb, ties, h, w - 1)
    monkeypatch.setenv("SSB_ARGMAX_NCHW_CFG", "generic")
    g_labels, g_stats = run_gpu(ssb, x)
    assert (g_labels == labels).all() and (g_stats == stats).all()
