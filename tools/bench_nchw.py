"""Times ssb_argmax_stats_nchw_{f32,f16} over the compiled CTA shapes (env SSB_ARGMAX_NCHW_CFG).

    python tools/bench_nchw.py [f32|f16] [frames] [warps-per-CTA ...]
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ssb200  # noqa: E402


def main():
    dt = sys.argv[1] if len(sys.argv) > 1 else "f32"
    frames = int(sys.argv[2]) if len(sys.argv) > 2 else (16 if dt == "f32" else 32)
    cfgs = sys.argv[3:] or ["8", "4"]
    dtype = torch.float32 if dt == "f32" else torch.float16
    h = w = 513
    c = 21
    bufs = [torch.randn((frames, c, h, w), device="cuda", dtype=dtype) for _ in range(3)]
    labels = torch.empty((frames, h, w), dtype=torch.uint8, device="cuda")
    stats = torch.empty((frames, c, 4), dtype=torch.int64, device="cuda")
    alg = frames * (c * h * w * bufs[0].element_size() + h * w + c * 32)
    for cfg in cfgs:
        os.environ["SSB_ARGMAX_NCHW_CFG"] = cfg
        for i in range(6):
            ssb200.argmax_stats_nchw(bufs[i % 3], labels=labels, stats=stats)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 60
        e0.record()
        for i in range(n):
            ssb200.argmax_stats_nchw(bufs[i % 3], labels=labels, stats=stats)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / n
        print("%s frames=%d warps/CTA=%s  %.1f us/launch  %.0f GB/s  %.0f frames/s" % (dt, frames, cfg, us, alg / us / 1e3, frames / us * 1e6),
              flush=True)


if __name__ == "__main__":
    main()
