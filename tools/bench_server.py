"""Frames per second through the drop-in producer loop (server.recognize_and_segment) with a model stand-in
that costs nothing (a fixed batch of class scores), so the figure is the path's own: frame gathering, H2D,
K1, K2+K3, and for "contours" K4 + K5 + the OpenCV tail on the host.

    python tools/bench_server.py [frames] [batch_frames]
"""
import importlib
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ssb200  # noqa: E402

srv = importlib.import_module(ssb200.__name__ + ".server")


class FixedScores:
    """Blobby label maps turned into u8 class scores once; run_logits hands out views of them."""

    def __init__(self, batch, h=513, w=513, c=21):
        g = torch.Generator(device="cuda")
        g.manual_seed(1)
        field = torch.randint(1, c, (batch, 9, 9), device="cuda", generator=g)
        field[torch.rand((batch, 9, 9), device="cuda", generator=g) < 0.5] = 0
        field[field > 0] = torch.where(torch.rand(int((field > 0).sum()), device="cuda", generator=g) < 0.5, 15, 7)
        lab = field.repeat_interleave(64, 1).repeat_interleave(64, 2)[:, :h, :w]
        q = torch.randint(0, 150, (batch, h, w, c), dtype=torch.uint8, device="cuda", generator=g)
        q.scatter_(3, lab.unsqueeze(-1), 255)
        self.q, self.h, self.w = q, h, w

    def get_input_tensor_shape(self):
        return (1, self.h, self.w, 3)

    def run_logits(self, x):
        return self.q[:x.shape[0]], 1.0 / 16, 128


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    batch = int(sys.argv[2]) if len(sys.argv) > 2 else 32
    rng = np.random.default_rng(0)
    frames = [rng.integers(0, 256, (1080, 1920, 3), dtype=np.uint8) for _ in range(8)]
    engine = FixedScores(batch)
    names = {i: "label%d" % i for i in range(21)}
    for mode in ("labels", "contours"):
        for rep in range(2):                          # first pass warms allocations and thread pools
            cam = srv.ArrayCamera([frames[i % 8] for i in range(n)])
            t0 = time.perf_counter()
            srv.recognize_and_segment(True, cam, engine, names, 0.05, batch_frames=batch, mode=mode)
            dt = time.perf_counter() - t0
        print("%s: %d frames of 1920x1080 in %.1f ms = %.0f frames/s, %d objects in the deque" %
              (mode, n, dt * 1e3, n / dt, len(srv.segmented_objects)))
        srv.segmented_objects.clear()


if __name__ == "__main__":
    main()
