"""Device time of the foreground-mask kernel (K4) on 64 513x513 label maps with the 16:9 camera crop, checked against
the oracle.   python tools/bench_fgmask.py        (SSB200_LIB=... for another build of the library)"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ssb200  # noqa: E402
from oracle import cport  # noqa: E402
from oracle import segpath as sp  # noqa: E402

H = W = 513
C = 21


def time_us(fn, iters=50):
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
    b = 64
    g = torch.Generator(device="cuda"); g.manual_seed(7)
    field = torch.randint(1, C, (b, 33, 33), device="cuda", generator=g)
    field[torch.rand((b, 33, 33), device="cuda", generator=g) < 0.7] = 0
    blobby = field.repeat_interleave(16, 1).repeat_interleave(16, 2)[:, :H, :W].to(torch.uint8).contiguous()
    noise = torch.randint(0, C, (b, H, W), dtype=torch.uint8, device="cuda", generator=g)
    cm = sp.pascal_colormap().astype(np.uint8)
    for name, lm in (("blobby", blobby), ("noise", noise)):
        for roi in ((288, 513), (513, 513)):
            out = torch.full_like(lm, 0xAB)
            us = time_us(lambda: ssb200.foreground_mask(lm, roi[0], roi[1], out=out))
            ok = (out[:4].cpu().numpy() == cport.fg_mask(lm[:4].cpu().numpy(), roi[0], roi[1], cm)).all()
            print("%-7s ROI %3dx%3d  %7.1f us per %d frames  %.0f GB/s  oracle: %s" %
                  (name, roi[0], roi[1], us, b, 2 * b * H * W / us / 1e3, bool(ok)), flush=True)


if __name__ == "__main__":
    main()
