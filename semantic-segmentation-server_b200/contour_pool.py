"""Parent side of the contour worker processes (contour_worker.py): a shared-memory file that is also
page-locked for the GPU, so the masks of a batch land in it by DMA and the workers read them in place."""
import mmap
import os
import subprocess
import sys

import numpy as np
import torch

from . import contour_worker as cw
from . import hotpath


class ContourWorkers:
    """`n_workers` processes running the reference's OpenCV contour stage on masks in `n_slots` shared slots
    of `slot_bytes` each.  submit() hands a batch to them, collect() returns one candidate list per frame."""

    def __init__(self, n_workers, slot_bytes, n_slots=3):
        self.slot_bytes = (slot_bytes + 4095) // 4096 * 4096
        self.n_slots = n_slots
        total = self.slot_bytes * n_slots
        path = "/dev/shm/ssb200_masks_%d_%x" % (os.getpid(), id(self))
        fd = os.open(path, os.O_CREAT | os.O_EXCL | os.O_RDWR, 0o600)
        try:
            os.ftruncate(fd, total)
            self._map = mmap.mmap(fd, total)
        finally:
            os.close(fd)
        self.array = np.frombuffer(self._map, dtype=np.uint8)
        self.array[::4096] = 0                                    # touch the pages before locking them
        self._registered = False
        if torch.cuda.is_available():
            rc = torch.cuda.cudart().cudaHostRegister(self.array.ctypes.data, total, 0)
            if int(rc) != 0:
                raise RuntimeError("cudaHostRegister of the mask buffer failed: %s" % rc)
            self._registered = True
        self._procs, self._pending = [], []
        script = os.path.abspath(cw.__file__)
        try:
            for _ in range(max(1, n_workers)):
                self._procs.append(subprocess.Popen([sys.executable, script, path, str(total)], stdin=subprocess.PIPE,
                                                    stdout=subprocess.PIPE))
            for p in self._procs:
                if cw.recv_msg(p.stdout) != "ready":
                    raise RuntimeError("contour worker did not start")
        finally:
            os.unlink(path)                                        # the mappings keep the memory alive

    def slot(self, slot, shape):
        """uint8 tensor view of `shape` at the start of a slot (page-locked: copies into it are DMA)."""
        n = int(np.prod(shape))
        if n > self.slot_bytes:
            raise ValueError("batch of %d mask bytes does not fit a slot of %d" % (n, self.slot_bytes))
        off = slot * self.slot_bytes
        return torch.from_numpy(self.array[off:off + n]).view(shape)

    def submit(self, slot, n_frames, h, w, roi_h, roi_w, min_area):
        ranges = [(n_frames * i // len(self._procs), n_frames * (i + 1) // len(self._procs)) for i in range(len(self._procs))]
        self._pending = []
        for p, (i0, i1) in zip(self._procs, ranges):
            if i1 > i0:
                cw.send_msg(p.stdin, (slot * self.slot_bytes, h, w, roi_h, roi_w, i0, i1, float(min_area)))
                self._pending.append(p)

    def collect(self):
        out = []
        for p in self._pending:
            res = cw.recv_msg(p.stdout)
            if res is None:
                raise RuntimeError("contour worker died")
            out.extend(res)
        self._pending = []
        return out

    def close(self):
        for p in self._procs:
            try:
                p.stdin.close()
                p.wait(timeout=5)
            except Exception:
                p.kill()
        self._procs = []
        if self._registered:
            torch.cuda.cudart().cudaHostUnregister(self.array.ctypes.data)
            self._registered = False

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def objects_from_candidates(candidates, enclosed, labels):
    """One frame's reference-literal objects from the workers' contour candidates and the frame's folded region
    table; None if the table cannot answer for one of them (incomplete table, unidentified hole): the caller
    then runs hotpath.objects_from_regions on the frame."""
    if enclosed is None:
        return None if candidates else []
    objs = []
    for is_hole, key, area, cx, cy in candidates:
        counts = (enclosed.holes if is_hole else enclosed).get(key) if key is not None else None
        if counts is None:
            return None
        idx = int(np.argmax(counts))
        objs.append(hotpath.SegmentContour(label=labels.get(idx, 0), score=int(counts[idx]) / int(counts.sum()), area=area,
                                           centroid=hotpath.Centroid(cx=cx, cy=cy)))
    return objs
