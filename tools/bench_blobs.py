"""Device time of the mask + region-table kernels (K4, K5) on 64 blobby 513x513 label maps with the 16:9
camera crop, for the cluster kernel and (SSB_BLOB_PATH=global) the global-memory kernels.

    python tools/bench_blobs.py [max_regions]
"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ssb200  # noqa: E402

H = W = 513
ROI_H, ROI_W, C = 288, 513, 21


def time_us(fn, iters=30):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


def main():
    rows = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
    b = 64
    g = torch.Generator(device="cuda"); g.manual_seed(7)
    field = torch.randint(1, C, (b, 33, 33), device="cuda", generator=g)
    field[torch.rand((b, 33, 33), device="cuda", generator=g) < 0.7] = 0
    lab = field.repeat_interleave(16, 1).repeat_interleave(16, 2)[:, :H, :W].to(torch.uint8).contiguous()
    noise = torch.randint(0, C, (b, H, W), dtype=torch.uint8, device="cuda", generator=g)
    # what a segmentation network gives: a few large objects (96-pixel cells, half of them person / car)
    big = torch.randint(1, C, (b, 7, 7), device="cuda", generator=g)
    big[torch.rand((b, 7, 7), device="cuda", generator=g) < 0.5] = 0
    big[torch.rand((b, 7, 7), device="cuda", generator=g) < 0.25] = 15
    large = big.repeat_interleave(96, 1).repeat_interleave(96, 2)[:, 40:H + 40, 50:W + 50].to(torch.uint8).contiguous()
    cases = os.environ.get("BLOBS_CASES", "blobby,noise,large").split(",")
    for name, lm in (("blobby", lab), ("noise", noise), ("large", large)):
        if name not in cases:
            continue
        mask = ssb200.foreground_mask(lm, ROI_H, ROI_W)
        res = {}
        for path in os.environ.get("BLOBS_PATHS", "cluster,global").split(","):
            os.environ["SSB_BLOB_PATH"] = path
            t = ssb200.blob_regions(mask, lm, ROI_H, ROI_W, C, rows)
            us = time_us(lambda: ssb200.blob_regions(mask, lm, ROI_H, ROI_W, C, rows, out=t))
            t = ssb200.blob_regions(mask, lm, ROI_H, ROI_W, C, rows, out=t)
            torch.cuda.synchronize()
            res[path] = {k: v.cpu().numpy() for k, v in t.items()}
            print("%-7s %-8s %8.1f us per %d frames, %.1f regions per frame" % (name, path, us, b, res[path]["count"].mean()))
        if os.environ.get("BLOBS_NO_CHECK"):
            continue
        a, c = res["cluster"], res["global"]
        same = (a["count"] == c["count"]).all()
        for f in range(b):
            n = min(int(a["count"][f]), rows)
            if a["count"][f] > rows:
                continue
            oa, oc = np.argsort(a["first"][f, :n]), np.argsort(c["first"][f, :n])
            for k in ("first", "parent", "flags", "hist"):
                same = same and (a[k][f, :n][oa] == c[k][f, :n][oc]).all()
            holes = (a["flags"][f, :n][oa] & 3) == 0
            same = same and (a["ring"][f, :n][oa][holes] == c["ring"][f, :n][oc][holes]).all()
        print("%-7s tables of the two paths agree: %s" % (name, bool(same)))
    os.environ.pop("SSB_BLOB_PATH", None)


if __name__ == "__main__":
    main()
