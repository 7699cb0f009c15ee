"""Frame-batching scheduler: frames are independent (sem_seg_server.py:145-195 keeps no cross-frame
state), so a batch is cut into contiguous frame ranges -- across the GPUs of one box (one process per
GPU, `torch.distributed` for the plumbing) and, inside a process, into chunks that flow through CUDA
streams.  There is no collective on the data path: each GPU writes its own [b,C,4] statistics and the
host concatenates 32*C bytes per frame.
"""
import os

import numpy as np
import torch

from . import hotpath


def partition_frames(n_frames, parts):
    """Contiguous, balanced ranges: part i gets [start_i, stop_i); the first n % parts get one extra."""
    if parts < 1:
        raise ValueError("parts must be >= 1")
    base, extra = divmod(n_frames, parts)
    out, start = [], 0
    for i in range(parts):
        stop = start + base + (1 if i < extra else 0)
        out.append((start, stop))
        start = stop
    return out


def aligned_chunks(start, stop, chunk, align=16):
    """Cuts [start, stop) into chunks whose starts are multiples of `align` frames relative to
    `start`, so that views into one big device buffer keep the streaming kernels' 16-byte alignment
    (16 frames of any odd-sized frame are a multiple of 16 bytes)."""
    chunk = max(align, chunk // align * align)
    return [(s, min(stop, s + chunk)) for s in range(start, stop, chunk)]


def bind_process_to_device_numa(device_index):
    """Pins this process to the CPU cores closest to GPU `device_index` (NVML's ideal CPU affinity), so
    that pinned host buffers allocated afterwards live on the GPU's NUMA node.  One process per GPU:
    call it before allocating pinned memory.  Returns the core count, or 0 if NVML is unavailable."""
    try:
        import pynvml
        pynvml.nvmlInit()
        handle = pynvml.nvmlDeviceGetHandleByIndex(device_index)
        n_words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(handle, n_words)
        cores = {64 * w + b for w, word in enumerate(mask) for b in range(64) if (word >> b) & 1}
        cores &= os.sched_getaffinity(0)
        if cores:
            os.sched_setaffinity(0, cores)
        return len(cores)
    except Exception:
        return 0


class FrameScheduler:
    """One instance per process (= per GPU).  `world`/`rank` default to the torchrun environment."""

    def __init__(self, rank=None, world=None, device=None):
        self.world = int(os.environ.get("WORLD_SIZE", "1")) if world is None else world
        self.rank = int(os.environ.get("RANK", "0")) if rank is None else rank
        self.device = device

    def local_range(self, n_frames):
        return partition_frames(n_frames, self.world)[self.rank]

    # -- device-resident inputs ----------------------------------------------------------------
    def argmax_stats_local(self, logits_local, roi_h, roi_w, want_labels=True, chunk_frames=None, streams=None):
        """Runs K2+K3 on this rank's frames (a device tensor holding only the local range).  With
        `chunk_frames` the range is cut into aligned chunks issued round-robin on `streams`."""
        b = logits_local.shape[0]
        c = logits_local.shape[3]
        labels = torch.empty(logits_local.shape[:3], dtype=torch.uint8, device=logits_local.device) if want_labels else None
        stats = torch.empty((b, c, 4), dtype=torch.int64, device=logits_local.device)
        if not chunk_frames or not streams:
            hotpath.argmax_stats(logits_local, roi_h, roi_w, labels=labels, stats=stats, want_labels=want_labels)
            return labels, stats
        cur = torch.cuda.current_stream(logits_local.device)
        for i, (s, e) in enumerate(aligned_chunks(0, b, chunk_frames)):
            st = streams[i % len(streams)]
            st.wait_stream(cur)
            hotpath.argmax_stats(logits_local[s:e], roi_h, roi_w, labels=labels[s:e] if want_labels else None,
                                 stats=stats[s:e], want_labels=want_labels, stream=st)
        for st in streams:
            cur.wait_stream(st)
        return labels, stats

    # -- gathering the per-frame statistics on rank 0 (host side, tiny) -----------------------------
    def gather_stats(self, stats_local, n_frames):
        """[b_local,C,4] (device or host) -> on rank 0 the full [n_frames,C,4] uint64 numpy array in
        frame order, None elsewhere.  Uses the default process group (gloo or nccl); with one
        process it is a plain host copy."""
        local = stats_local.detach().cpu().numpy() if isinstance(stats_local, torch.Tensor) else np.asarray(stats_local)
        local = np.ascontiguousarray(local).view(np.uint64)
        if self.world == 1:
            return local
        import torch.distributed as dist
        parts = [None] * self.world if self.rank == 0 else None
        dist.gather_object(local, parts, dst=0)
        if self.rank != 0:
            return None
        ranges = partition_frames(n_frames, self.world)
        for p, (s, e) in zip(parts, ranges):
            if p.shape[0] != e - s:
                raise RuntimeError("rank returned %d frames for range [%d,%d)" % (p.shape[0], s, e))
        return np.concatenate(parts, axis=0)
