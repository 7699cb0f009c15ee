"""ssb_argmax_stats_u8_noalloc: the call for hosts that cannot let the library allocate device
memory (their own allocator already owns HBM, or a cudaMalloc stall in the frame loop is not
acceptable).  Its scratch is a fixed array inside the library, shared by all calls, so the things
to prove are: it gives the same bytes as the regular entry point, and a second call straight after
the first is not disturbed by what the first left behind.

Collected last on purpose (file name): a run that stops at the first failure has then already run
every other test.  Works in a fresh process and after the rest of the suite alike -- it depends on
nothing the other tests set up.
"""
import ctypes

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ssb():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ssb200
    return ssb200


def _call(fn, logits, roi_h, roi_w):
    b, h, w, c = logits.shape
    labels = torch.full((b, h, w), 0xEE, dtype=torch.uint8, device="cuda")
    stats = torch.full((b, c, 4), -1, dtype=torch.int64, device="cuda")       # garbage in: nothing to pre-zero
    rc = fn(ctypes.c_void_p(logits.data_ptr()), b, h, w, c, roi_h, roi_w, ctypes.c_void_p(labels.data_ptr()),
            ctypes.c_void_p(stats.data_ptr()), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    return rc, labels.cpu().numpy(), stats.cpu().numpy()


def test_noalloc_matches_regular_entry_point_twice_in_a_row(ssb):
    lib = ssb._cabi.lib
    rng = np.random.default_rng(77)
    # two back-to-back calls with different batch sizes, ROIs and contents: the second one starts
    # from whatever the first left in the shared array
    cases = [(rng.integers(0, 256, (5, 513, 513, 21), dtype=np.uint8), 288, 513),
             (rng.integers(0, 4, (3, 513, 513, 21), dtype=np.uint8), 513, 384)]
    got = []
    for q, roi_h, roi_w in cases:                      # the two no-alloc calls, nothing in between
        got.append(_call(lib.ssb_argmax_stats_u8_noalloc, torch.from_numpy(q).cuda(), roi_h, roi_w))
    for (q, roi_h, roi_w), (rc, labels, stats) in zip(cases, got):
        assert rc == 0, lib.ssb_last_error()
        rc_ref, want_labels, want_stats = _call(lib.ssb_argmax_stats_u8, torch.from_numpy(q).cuda(), roi_h, roi_w)
        assert rc_ref == 0
        # integer outputs: the tolerance the rest of the suite uses is zero
        assert (labels == want_labels).all()
        assert (stats == want_stats).all()
        assert int(stats[:, :, 0].sum()) == q.shape[0] * roi_h * roi_w


def test_noalloc_other_shapes_and_limit(ssb):
    lib = ssb._cabi.lib
    for shape in [(2, 64, 128, 19), (3, 17, 23, 5)]:   # compiled class count / generic kernel
        q = np.random.default_rng(sum(shape)).integers(0, 256, shape, dtype=np.uint8)
        rc, labels, stats = _call(lib.ssb_argmax_stats_u8_noalloc, torch.from_numpy(q).cuda(), shape[1], shape[2])
        rc_ref, want_labels, want_stats = _call(lib.ssb_argmax_stats_u8, torch.from_numpy(q).cuda(), shape[1], shape[2])
        assert rc == 0 and rc_ref == 0
        assert (labels == want_labels).all() and (stats == want_stats).all()
    # one frame too many: refused with a message, nothing launched
    too_many = ssb._cabi.lib.ssb_argmax_stats_u8_noalloc(ctypes.c_void_p(16), 257, 8, 8, 21, 8, 8, None,
                                                        ctypes.c_void_p(16), None)
    assert too_many == ssb._cabi.SSB_ERR_LIMIT
    assert b"at most 256 frames" in lib.ssb_last_error()
