"""The C-ABI is documented as re-entrant across (device, stream) pairs: host threads with their own CUDA
streams call the scratch-using entry points at the same time; every result must be the bytes a lone call
gives (the zero-at-rest scratch slots are handed out per call and ordered across streams with events)."""
import threading

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


def test_threads_on_their_own_streams(ssb):
    rng = np.random.default_rng(2024)
    n_threads, rounds = 4, 12
    # per thread: a u8 batch of its own size (scratch slots of different sizes get reused across streams),
    # an fp16 batch and a planar fp32 batch
    work = []
    for t in range(n_threads):
        q = rng.integers(0, 256, (3 + 2 * t, 513, 513, 21), dtype=np.uint8)
        h = rng.standard_normal((2, 160, 200, 19)).astype(np.float16)
        x = rng.standard_normal((2, 21, 130, 257)).astype(np.float32)
        work.append({"q": q, "h": h, "x": x,
                     "want_q": cport.argmax_stats(q, 288, 513, threads=4),
                     "want_h": cport.argmax_stats(h, 150, 199, threads=2),
                     "want_x": sp.argmax_stats_nchw(x, 100, 257)})
    errors = []
    start = threading.Barrier(n_threads)

    def run(t):
        try:
            w = work[t]
            stream = torch.cuda.Stream()
            with torch.cuda.stream(stream):
                q, h, x = (torch.from_numpy(w[k]).cuda() for k in ("q", "h", "x"))
                stream.synchronize()
                start.wait()
                for r in range(rounds):
                    lq, sq = ssb.argmax_stats(q, 288, 513, stream=stream)
                    lh, sh = ssb.argmax_stats(h, 150, 199, stream=stream)
                    lx, sx = ssb.argmax_stats_nchw(x, 100, 257, stream=stream)
                    if r % 4 == 3 or r == rounds - 1:
                        stream.synchronize()
                        for (lab, st), want in (((lq, sq), w["want_q"]), ((lh, sh), w["want_h"]), ((lx, sx), w["want_x"])):
                            if not ((lab.cpu().numpy() == want[0]).all() and (st.cpu().numpy().view(np.uint64) == want[1]).all()):
                                raise AssertionError("thread %d round %d: result differs from the oracle" % (t, r))
        except Exception as e:   # surfaced in the main thread
            errors.append(e)
            try:
                start.abort()
            except Exception:
                pass

    threads = [threading.Thread(target=run, args=(t,)) for t in range(n_threads)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    torch.cuda.synchronize()
    assert not errors, errors[0]


def test_more_threads_than_scratch_slots(ssb):
    """The scratch pool has 8 slots per device.  Twelve threads on twelve streams keep more calls than
    that between 'slot taken' and 'kernel enqueued': a slot must never be handed to two of them (the
    ninth caller waits for a slot to come back), so every result stays the lone call's."""
    rng = np.random.default_rng(99)
    n_threads, rounds = 12, 40
    work = []
    for t in range(n_threads):
        q = rng.integers(0, 256, (1 + t % 3, 120, 131, 21), dtype=np.uint8)
        work.append((q, cport.argmax_stats(q, 100, 131, threads=2)))
    errors = []
    start = threading.Barrier(n_threads)

    def run(t):
        try:
            q_np, want = work[t]
            stream = torch.cuda.Stream()
            with torch.cuda.stream(stream):
                q = torch.from_numpy(q_np).cuda()
                stream.synchronize()
                start.wait()
                for r in range(rounds):
                    lab, st = ssb.argmax_stats(q, 100, 131, stream=stream)
                    if r % 8 == 7:
                        stream.synchronize()
                        if not ((lab.cpu().numpy() == want[0]).all() and (st.cpu().numpy().view(np.uint64) == want[1]).all()):
                            raise AssertionError("thread %d round %d: result differs from the oracle" % (t, r))
        except Exception as e:
            errors.append(e)
            try:
                start.abort()
            except Exception:
                pass

    threads = [threading.Thread(target=run, args=(t,)) for t in range(n_threads)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    torch.cuda.synchronize()
    assert not errors, errors[0]
