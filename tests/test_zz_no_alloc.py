"""ssb_argmax_stats_u8_noalloc: the call for hosts that cannot let the library allocate device
memory (their own allocator already owns HBM, or a cudaMalloc stall in the frame loop is not
acceptable).  Its scratch is a fixed array inside the library, shared by all calls, so the things
to prove are: it gives the bytes the oracle gives (and therefore those of the regular entry
point), a second call straight after the first is not disturbed by what the first left behind --
nor by garbage anybody else left in the array -- and the amount of free device memory does not
move across a call.

Collected last on purpose (file name): a run that stops at the first failure has then already run
every other test.  Works in a fresh process and after the rest of the suite alike -- it depends on
nothing the other tests set up.
"""
import ctypes

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from oracle import cport  # noqa: E402  (checker only)


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
    return rc, labels.cpu().numpy(), stats.cpu().numpy().view(np.uint64)


class _DeviceBytes:
    """A raw device address as something torch can wrap without copying."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def _scratch_view(lib):
    ptr, n = ctypes.c_void_p(), ctypes.c_size_t()
    assert lib.ssb_noalloc_scratch_info(ctypes.byref(ptr), ctypes.byref(n)) == 0, lib.ssb_last_error()
    assert ptr.value and ptr.value % 256 == 0 and n.value >= 256 * (21 * 32 + 4)
    return torch.as_tensor(_DeviceBytes(ptr.value, n.value), device="cuda")


def test_noalloc_matches_oracle_twice_in_a_row(ssb):
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
        want_labels, want_stats = cport.argmax_stats(q, roi_h, roi_w, threads=4)     # the oracle
        # integer outputs: the tolerance the rest of the suite uses is zero
        assert (labels == want_labels).all()
        assert (stats == want_stats).all()
        assert int(stats[:, :, 0].sum()) == q.shape[0] * roi_h * roi_w
        rc_ref, ref_labels, ref_stats = _call(lib.ssb_argmax_stats_u8, torch.from_numpy(q).cuda(), roi_h, roi_w)
        assert rc_ref == 0 and (labels == ref_labels).all() and (stats == ref_stats).all()


def test_noalloc_ignores_what_the_array_held(ssb):
    """The array keeps the last call's contents and is shared: whatever is in it -- here, every byte
    set to 0xA5, as a call that was cut short or a foreign writer might leave it -- must not reach the
    results."""
    lib = ssb._cabi.lib
    q = np.random.default_rng(5).integers(0, 256, (7, 513, 513, 21), dtype=np.uint8)
    want_labels, want_stats = cport.argmax_stats(q, 400, 500, threads=4)
    dq = torch.from_numpy(q).cuda()
    view = _scratch_view(lib)
    for _ in range(2):
        view.fill_(0xA5)
        torch.cuda.synchronize()
        rc, labels, stats = _call(lib.ssb_argmax_stats_u8_noalloc, dq, 400, 500)
        assert rc == 0, lib.ssb_last_error()
        assert (labels == want_labels).all() and (stats == want_stats).all()
    # a call that ran to completion leaves the part it used zeroed
    used = 7 * 21 * 32 + 7 * 4
    assert int(view[:used].count_nonzero()) == 0


def test_noalloc_leaves_free_device_memory_alone(ssb):
    """The property the entry point exists for.  The first call on a device may make the driver load
    the kernel's code (lazy module loading), so the measurement brackets the calls after it."""
    lib = ssb._cabi.lib
    q = torch.from_numpy(np.random.default_rng(6).integers(0, 256, (16, 513, 513, 21), dtype=np.uint8)).cuda()
    labels = torch.empty((16, 513, 513), dtype=torch.uint8, device="cuda")
    stats = torch.empty((16, 21, 4), dtype=torch.int64, device="cuda")
    args = (ctypes.c_void_p(q.data_ptr()), 16, 513, 513, 21, 513, 513, ctypes.c_void_p(labels.data_ptr()),
            ctypes.c_void_p(stats.data_ptr()), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert lib.ssb_argmax_stats_u8_noalloc(*args) == 0
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        assert lib.ssb_argmax_stats_u8_noalloc(*args) == 0
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after == free_before
    want_labels, want_stats = cport.argmax_stats(q.cpu().numpy(), 513, 513, threads=4)
    assert (labels.cpu().numpy() == want_labels).all()
    assert (stats.cpu().numpy().view(np.uint64) == want_stats).all()


def test_noalloc_other_shapes_and_limit(ssb):
    lib = ssb._cabi.lib
    for shape in [(2, 64, 128, 19), (3, 17, 23, 5)]:   # compiled class count / generic kernel
        q = np.random.default_rng(sum(shape)).integers(0, 256, shape, dtype=np.uint8)
        rc, labels, stats = _call(lib.ssb_argmax_stats_u8_noalloc, torch.from_numpy(q).cuda(), shape[1], shape[2])
        want_labels, want_stats = cport.argmax_stats(q, shape[1], shape[2], threads=2)
        assert rc == 0
        assert (labels == want_labels).all() and (stats == want_stats).all()
    # one frame too many: refused with a message, nothing launched
    too_many = ssb._cabi.lib.ssb_argmax_stats_u8_noalloc(ctypes.c_void_p(16), 257, 8, 8, 21, 8, 8, None,
                                                        ctypes.c_void_p(16), None)
    assert too_many == ssb._cabi.SSB_ERR_LIMIT
    assert b"at most 256 frames" in lib.ssb_last_error()
