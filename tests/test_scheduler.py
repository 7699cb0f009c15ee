"""Frame sharding: host logic on CPU (incl. a world_size-2 gloo run), kernels on one GPU with the
ranks emulated one after the other (N-rank result == 1-rank result, byte for byte)."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_partition_frames_properties():
    import ssb200
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=200, deadline=None)
    @given(n=st.integers(0, 5000), parts=st.integers(1, 16))
    def prop(n, parts):
        r = ssb200.partition_frames(n, parts)
        assert len(r) == parts and r[0][0] == 0 and r[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
        sizes = [e - s for s, e in r]
        assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
    prop()
    assert ssb200.partition_frames(4096, 8)[3] == (1536, 2048)          # config 4: 512 frames per GPU
    with pytest.raises(ValueError):
        ssb200.partition_frames(4, 0)


def test_aligned_chunks_keep_16_frame_alignment():
    import ssb200
    for start, stop, chunk in [(0, 70, 32), (0, 15, 16), (0, 64, 10), (16, 100, 48)]:
        c = ssb200.aligned_chunks(start, stop, chunk)
        assert c[0][0] == start and c[-1][1] == stop
        assert all(a[1] == b[0] for a, b in zip(c, c[1:]))
        assert all((s - start) % 16 == 0 for s, _ in c)
        # 16 VOC frames are a multiple of 16 bytes although one frame is not
        assert (16 * 513 * 513 * 21) % 16 == 0 and (513 * 513 * 21) % 16 != 0


_WORKER = r'''
import os, sys
sys.path.insert(0, %(root)r)
import numpy as np, torch.distributed as dist
import ssb200
from oracle import cport                      # test only: stands in for the GPU kernels on this CPU box
dist.init_process_group("gloo")
sched = ssb200.FrameScheduler()
n = 11
logits = np.random.default_rng(5).integers(0, 256, (n, 33, 47, 21), dtype=np.uint8)
s, e = sched.local_range(n)
_, local = cport.argmax_stats(logits[s:e], 20, 47)
full = sched.gather_stats(local, n)
if sched.rank == 0:
    _, want = cport.argmax_stats(logits, 20, 47)
    assert full.shape == want.shape and (full == want).all()
    print("GATHER_OK", full.shape)
else:
    assert full is None
dist.destroy_process_group()
'''


def test_two_rank_gloo_gather(tmp_path):
    """world_size 2 over gloo on CPU: each rank handles its contiguous frame range, rank 0 gets the
    statistics of the whole batch in frame order, identical to a single-rank run."""
    script = tmp_path / "worker.py"
    script.write_text(_WORKER % {"root": ROOT})
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29611", str(script)],
                         capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "GATHER_OK (11, 21, 4)" in out.stdout


@pytest.mark.gpu
def test_emulated_ranks_equal_single_pass():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ssb200
    n = 37
    q = np.random.default_rng(9).integers(0, 256, (n, 513, 513, 21), dtype=np.uint8)
    full = torch.from_numpy(q).cuda()
    labels1, stats1 = ssb200.argmax_stats(full, 288, 513)
    for world in (2, 4, 8):
        parts_l, parts_s = [], []
        for rank in range(world):
            sched = ssb200.FrameScheduler(rank=rank, world=world)
            s, e = sched.local_range(n)
            local = torch.from_numpy(q[s:e]).cuda()                     # each rank owns its own buffer
            l, st = sched.argmax_stats_local(local, 288, 513)
            parts_l.append(l); parts_s.append(st)
        assert torch.equal(torch.cat(parts_l), labels1) and torch.equal(torch.cat(parts_s), stats1)


@pytest.mark.gpu
def test_frames_sharded_over_the_visible_gpus_equal_one_gpu():
    """N-GPU output byte-identical to 1-GPU output.  Uses every visible device (one host
    thread per GPU, as one process per GPU would); skipped on a single-GPU box."""
    import threading
    import torch
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs at least two CUDA devices")
    import ssb200
    world, n = torch.cuda.device_count(), 41
    q = np.random.default_rng(12).integers(0, 256, (n, 513, 513, 21), dtype=np.uint8)
    labels1, stats1 = ssb200.argmax_stats(torch.from_numpy(q).cuda(0), 288, 513)
    parts, errors = [None] * world, []

    def run(rank):
        try:
            with torch.cuda.device(rank):
                sched = ssb200.FrameScheduler(rank=rank, world=world, device=rank)
                s, e = sched.local_range(n)
                local = torch.from_numpy(q[s:e]).cuda(rank)
                lab, st = sched.argmax_stats_local(local, 288, 513)
                torch.cuda.synchronize(rank)
                parts[rank] = (lab.cpu(), st.cpu())
        except Exception as exc:
            errors.append(exc)
    threads = [threading.Thread(target=run, args=(r,)) for r in range(world)]
    [t.start() for t in threads]
    [t.join() for t in threads]
    assert not errors, errors[0]
    assert torch.equal(torch.cat([p[0] for p in parts]), labels1.cpu())
    assert torch.equal(torch.cat([p[1] for p in parts]), stats1.cpu())


@pytest.mark.gpu
def test_chunked_streams_equal_single_launch():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ssb200
    q = torch.from_numpy(np.random.default_rng(10).integers(0, 256, (50, 513, 513, 21), dtype=np.uint8)).cuda()
    labels1, stats1 = ssb200.argmax_stats(q, 288, 513)
    sched = ssb200.FrameScheduler(rank=0, world=1)
    streams = [torch.cuda.Stream() for _ in range(3)]
    labels2, stats2 = sched.argmax_stats_local(q, 288, 513, chunk_frames=16, streams=streams)
    torch.cuda.synchronize()
    assert torch.equal(labels1, labels2) and torch.equal(stats1, stats2)


@pytest.mark.gpu
def test_concurrent_streams_share_the_scratch_pool_safely():
    """Back-to-back calls of varying size on four streams: every call borrows a zero-at-rest scratch
    slot; slots reused across streams are ordered with events.  All results must be exact."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ssb200
    from oracle import cport
    rng = np.random.default_rng(0)
    streams = [torch.cuda.Stream() for _ in range(4)]
    pending = []
    for it in range(24):
        b = int(rng.integers(1, 24))
        roi = (int(rng.integers(1, 514)), 513 if it % 3 else int(rng.integers(1, 514)))
        q = rng.integers(0, 256, (b, 513, 513, 21), dtype=np.uint8)
        s = streams[it % 4]
        with torch.cuda.stream(s):
            lg = torch.from_numpy(q).cuda(non_blocking=True)
            labels, stats = ssb200.argmax_stats(lg, roi[0], roi[1], stream=s)
        pending.append((q, roi, labels, stats, lg))
    torch.cuda.synchronize()
    for q, roi, labels, stats, _ in pending:
        want_l, want_s = cport.argmax_stats(q, roi[0], roi[1], threads=8)
        assert (labels.cpu().numpy() == want_l).all()
        assert (stats.cpu().numpy().view(np.uint64) == want_s).all()
