"""Times the reference-literal tail (row f1) for a batch of blobby label maps: mask kernel, region-table
kernel, device->host copies, and the host part per frame with and without the region tables.

    python tools/bench_contours.py [frames]
"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ssb200  # noqa: E402


def blobby(seed, h, w):
    rng = np.random.default_rng(seed)
    lab = np.zeros((h, w), np.uint8)
    yy, xx = np.mgrid[0:h, 0:w]
    for k in range(5):
        cy, cx, ry, rx = rng.integers(30, h - 30), rng.integers(40, w - 40), rng.integers(20, 110), rng.integers(30, 150)
        lab[((yy - cy) / ry) ** 2 + ((xx - cx) / rx) ** 2 < 1] = (7, 15, 12, 15, 7)[k]
    cy, cx = rng.integers(60, h - 60), rng.integers(80, w - 80)
    lab[cy - 10:cy + 10, cx - 14:cx + 14] = 3          # a patch of another class: usually a hole in a blob
    return lab


def timed(fn, n):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


def main():
    b = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    H = W = 513
    roi_h, roi_w = 288, 513
    labs = np.zeros((b, H, W), np.uint8)
    for i in range(b):
        labs[i, :roi_h] = blobby(i, roi_h, roi_w)
    lm = torch.from_numpy(labs).cuda()
    names = {i: "l%d" % i for i in range(21)}
    min_area = 0.05 * H * W
    masks = ssb200.foreground_mask(lm, roi_h, roi_w)
    us_mask = timed(lambda: ssb200.foreground_mask(lm, roi_h, roi_w, out=masks), 20)
    us_regions = timed(lambda: ssb200.blob_regions(masks, lm, roi_h, roi_w, 21, 256), 20)
    t = ssb200.blob_regions(masks, lm, roi_h, roi_w, 21, 256)
    torch.cuda.synchronize()
    staging = ssb200.HostStaging()
    for rep in range(2):                                   # second round: buffers exist, paths are warm
        t0 = time.perf_counter()
        th = staging.fetch(dict(t, mask=masks))
        t_copy_new = time.perf_counter() - t0
        t0 = time.perf_counter()
        masks_h = masks[:, :roi_h, :roi_w].cpu().numpy()
        labs_h = lm[:, :roi_h, :roi_w].cpu().numpy()
        t_copy_old = time.perf_counter() - t0
    masks_h = th["mask"][:, :roi_h, :roi_w]
    for i in range(2):                                     # warm both host paths before timing them
        ssb200.objects_from_contours(masks_h[i], labs_h[i], min_area, names)
        ssb200.objects_from_regions(masks_h[i], None, min_area, names, label_map=labs_h[i])
    t0 = time.perf_counter()
    new = [ssb200.objects_from_regions(masks_h[i], ssb200.fold_region_histograms(th["count"][i], th["first"][i], th["parent"][i],
                                                                                  th["flags"][i], th["hist"][i], th["ring"][i]),
                                       min_area, names, label_map=labs_h[i]) for i in range(b)]
    t_new = time.perf_counter() - t0
    t0 = time.perf_counter()
    old = [ssb200.objects_from_contours(masks_h[i], labs_h[i], min_area, names) for i in range(b)]
    t_old = time.perf_counter() - t0
    assert new == old and sum(len(o) for o in new) > 0
    print("frames %d, objects %d, regions/frame avg %.1f max %d" % (b, sum(len(o) for o in new), th["count"].mean(), th["count"].max()))
    print("GPU: fg_mask %.1f us, blob_regions %.1f us (%.2f us/frame)" % (us_mask, us_regions, us_regions / b))
    print("D2H: mask+tables through pinned staging %.2f ms, mask+labels with .cpu() %.2f ms" % (t_copy_new * 1e3, t_copy_old * 1e3))
    print("host tail per frame, 1 thread: region tables %.3f ms, polygon fill %.3f ms" % (t_new / b * 1e3, t_old / b * 1e3))


if __name__ == "__main__":
    main()
