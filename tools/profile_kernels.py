"""Launches every kernel of the library a few times at its benchmark shape (for `ncu --set full`; the
numbers a profiler prints are never bench values).

    python tools/profile_kernels.py            # plain run first
    ncu --set full --clock-control none --import-source on -o out -f python tools/profile_kernels.py
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ssb200  # noqa: E402

REPS = 2


def main():
    g = torch.Generator(device="cuda")
    g.manual_seed(3)
    H = W = 513
    # K1: 128 x 1080p -> 513x513
    geom = ssb200.FrameGeometry(1920, 1080, H, W)
    cam = torch.randint(0, 256, (128, 1080, 1920, 3), dtype=torch.uint8, device="cuda", generator=g)
    rgb = torch.empty((128, H, W, 3), dtype=torch.uint8, device="cuda")
    for _ in range(REPS):
        ssb200.preprocess_frames(cam, geom, out=rgb)
    del cam
    # K2+K3 u8 (headline), full-frame ROI and the camera crop
    q = torch.randint(0, 256, (64, H, W, 21), dtype=torch.uint8, device="cuda", generator=g)
    for _ in range(REPS):
        labels, _ = ssb200.argmax_stats(q)
    ssb200.argmax_stats(q, geom.roi_h, geom.roi_w)
    # K4 on the label maps just produced
    cm = torch.from_numpy(ssb200.create_pascal_label_colormap().astype(np.uint8)).cuda()
    for _ in range(REPS):
        ssb200.foreground_mask(labels, geom.roi_h, geom.roi_w, cm)
    # K5 on blobby maps (regions of 16x16 cells) and their masks
    field = torch.randint(1, 21, (64, 33, 33), device="cuda", generator=g)
    field[torch.rand((64, 33, 33), device="cuda", generator=g) < 0.7] = 0
    blobby = field.repeat_interleave(16, 1).repeat_interleave(16, 2)[:, :H, :W].to(torch.uint8).contiguous()
    bmask = ssb200.foreground_mask(blobby, geom.roi_h, geom.roi_w, cm)
    for _ in range(REPS):
        tabs = ssb200.blob_regions(bmask, blobby, geom.roi_h, geom.roi_w, 21, 1024)
    # K6 on those tables
    for _ in range(REPS):
        ssb200.contour_moments(bmask, tabs, geom.roi_h, geom.roi_w)
    del q
    # K2+K3 fp16, Cityscapes shape
    h16 = torch.randn((8, 1024, 2048, 19), dtype=torch.float16, device="cuda", generator=g)
    for _ in range(REPS):
        ssb200.argmax_stats(h16)
    del h16
    # planar fp32 / fp16
    for dt, nb in ((torch.float32, 32), (torch.float16, 64)):
        x = torch.randn((nb, 21, H, W), dtype=dt, device="cuda", generator=g)
        for _ in range(REPS):
            ssb200.argmax_stats_nchw(x)
        del x
    torch.cuda.synchronize()
    print("profile_kernels ok")


if __name__ == "__main__":
    main()
