"""Tuning sweep for the fused argmax+stats kernel (run on the GPU box):
times every compiled (consumer warps, stages) configuration x grid multiplier with CUDA events."""
import itertools
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ssb200  # noqa: E402

F16 = os.environ.get("SWEEP_DTYPE", "u8") == "f16"
H, W, C, B = (1024, 2048, 19, 8) if F16 else (513, 513, 21, 64)
ROI = (H, W) if F16 else (288, 513)
ALGO = (H * W * C * (2 if F16 else 1) + H * W + C * 32) * B
CFG_ENV = "SSB_ARGMAX_F16_CFG" if F16 else "SSB_ARGMAX_CFG"


def time_cfg(batches, labels, stats, iters=40):
    s = torch.cuda.current_stream()
    for i in range(4):
        ssb200.argmax_stats(batches[i % len(batches)], ROI[0], ROI[1], labels=labels, stats=stats)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(s)
    for i in range(iters):
        ssb200.argmax_stats(batches[i % len(batches)], ROI[0], ROI[1], labels=labels, stats=stats)
    e1.record(s)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3   # us


def main():
    g = torch.Generator(device="cuda"); g.manual_seed(0)
    if F16:
        batches = [torch.randn((B, H, W, C), dtype=torch.float16, device="cuda", generator=g) for _ in range(4)]
    else:
        batches = [torch.randint(0, 256, (B, H, W, C), dtype=torch.uint8, device="cuda", generator=g) for _ in range(4)]
    labels = torch.empty((B, H, W), dtype=torch.uint8, device="cuda")
    stats = torch.empty((B, C, 4), dtype=torch.int64, device="cuda")
    cfgs = os.environ.get("SWEEP_CFGS", "4,3 4,4 8,2" if F16 else "4,3 4,4 8,3").split()
    waves = os.environ.get("SWEEP_WAVES", "1 2 3 4 6").split()
    extra = json.loads(os.environ.get("SWEEP_EXTRA", "{}"))   # {"ENV": ["a","b"]}
    keys = sorted(extra)
    out = []
    for cfg, wv in itertools.product(cfgs, waves):
        for combo in itertools.product(*[extra[k] for k in keys]):
            os.environ[CFG_ENV] = cfg
            os.environ["SSB_ARGMAX_WAVES"] = wv
            for k, v in zip(keys, combo):
                os.environ[k] = v
            try:
                us = min(time_cfg(batches, labels, stats) for _ in range(3))
            except Exception as e:   # config not compiled / does not fit
                print(cfg, wv, combo, "ERR", e)
                continue
            row = {"cfg": cfg, "waves": int(wv), "extra": dict(zip(keys, combo)), "us": round(us, 2),
                   "GBps": round(ALGO / us / 1e3, 1)}
            out.append(row)
            print(json.dumps(row), flush=True)
    best = min(out, key=lambda r: r["us"])
    print("BEST", json.dumps(best))


if __name__ == "__main__":
    main()
