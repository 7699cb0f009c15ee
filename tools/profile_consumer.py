"""Where the host side of the contours path spends its time, per batch of 32 blobby 1080p frames (tuning aid):
read-back, table fold, OpenCV contour tail (serial and on the thread pool), reply objects."""
import importlib
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import ssb200  # noqa: E402

srv = importlib.import_module(ssb200.__name__ + ".server")
proto = importlib.import_module(ssb200.__name__ + ".proto")


def main():
    b = 32
    eng = bench.FixedScores(torch, b)
    geom = ssb200.FrameGeometry(1920, 1080, 513, 513)
    lab, stats = ssb200.argmax_stats(eng.q, geom.roi_h, geom.roi_w)
    mask = ssb200.foreground_mask(lab, geom.roi_h, geom.roi_w)
    tabs = ssb200.blob_regions(mask, lab, geom.roi_h, geom.roi_w, 21, 256)
    staging = ssb200.HostStaging()
    names = {i: "label%d" % i for i in range(21)}

    def timed(fn, n=20):
        fn()
        t0 = time.perf_counter()
        for _ in range(n):
            r = fn()
        return (time.perf_counter() - t0) / n * 1e3, r

    ms, th = timed(lambda: staging.fetch(dict(tabs, mask=mask)))
    print("read-back of masks + tables: %.3f ms per batch" % ms)
    scratch = {}
    ms, enclosed = timed(lambda: ssb200.fold_regions_batch(th, srv.HOST_THREADS, scratch))
    print("fold (native + python wrapper): %.3f ms" % ms)
    ms, _ = timed(lambda: [ssb200.fold_region_histograms(th["count"][i], th["first"][i], th["parent"][i], th["flags"][i],
                                                         th["hist"][i], th["ring"][i]) for i in range(b)], 5)
    print("fold (python, frame by frame): %.3f ms" % ms)
    masks_h = th["mask"][:, :geom.roi_h, :geom.roi_w]
    min_area = 0.05 * 513 * 513

    def tail(i):
        return ssb200.objects_from_regions(masks_h[i], enclosed[i], min_area, names, label_map=None)
    ms, objs = timed(lambda: [tail(i) for i in range(b)], 5)
    print("contour tails, serial: %.3f ms (%.1f us per frame), %d objects" % (ms, ms / b * 1e3, sum(len(o) for o in objs)))
    import cv2
    ms, _ = timed(lambda: [cv2.findContours(np.ascontiguousarray(masks_h[i]), cv2.RETR_TREE, cv2.CHAIN_APPROX_SIMPLE) for i in range(b)], 5)
    print("   of which ascontiguousarray + cv2.findContours: %.3f ms" % ms)
    pool = srv._contour_pool()
    ms, objs = timed(lambda: list(pool.map(tail, range(b))), 10)
    print("contour tails on the pool: %.3f ms" % ms)

    def protos():
        out = []
        for fo in objs:
            for o in fo:
                lab_, score, area, cx, cy = ssb200.normalise(o, 513, 513)
                out.append(proto.SegmentedObject(label=lab_, score=score, area=area,
                                                 centroid=proto.SegmentedObject.Centroid(cx=cx, cy=cy)))
        return out
    ms, _ = timed(protos)
    print("reply objects: %.3f ms" % ms)


if __name__ == "__main__":
    main()
