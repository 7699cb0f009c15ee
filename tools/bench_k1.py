"""Times K1 (pre-process) and the whole config-3 chain on the GPU box (tuning aid)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ssb200  # noqa: E402


def timeit(fn, iters=30):
    for _ in range(3):
        fn(0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


def main():
    B = int(os.environ.get("K1_B", "128"))
    g = torch.Generator(device="cuda"); g.manual_seed(0)
    frames = [torch.randint(0, 256, (B, 1080, 1920, 3), dtype=torch.uint8, device="cuda", generator=g) for _ in range(2)]
    geom = ssb200.FrameGeometry(1920, 1080, 513, 513)
    out = torch.empty((B, 513, 513, 3), dtype=torch.uint8, device="cuda")
    us = min(timeit(lambda i: ssb200.preprocess_frames(frames[i % 2], geom, out=out)) for _ in range(3))
    sampled = (geom.new_h * geom.new_w * 3 + 513 * 513 * 3) * B
    sector = (geom.new_h * 1920 * 3 + 513 * 513 * 3) * B
    print(json.dumps({"kernel": "preprocess 1080p->513", "B": B, "us": round(us, 1), "frames_per_s": round(B / us * 1e6),
                      "GBps_sampled": round(sampled / us / 1e3, 1), "GBps_sector_effective": round(sector / us / 1e3, 1)}))
    for cam in [(640, 480), (1280, 720), (3840, 2160)]:
        fr = torch.randint(0, 256, (32, cam[1], cam[0], 3), dtype=torch.uint8, device="cuda", generator=g)
        gm = ssb200.FrameGeometry(cam[0], cam[1], 513, 513)
        o = torch.empty((32, 513, 513, 3), dtype=torch.uint8, device="cuda")
        us = min(timeit(lambda i: ssb200.preprocess_frames(fr, gm, out=o)) for _ in range(2))
        sector = (gm.new_h * cam[0] * 3 + 513 * 513 * 3) * 32
        print(json.dumps({"kernel": "preprocess %dx%d->513" % cam, "B": 32, "us": round(us, 1),
                          "GBps_sector_effective": round(sector / us / 1e3, 1)}))
    # mask kernel
    labels = torch.randint(0, 21, (B, 513, 513), dtype=torch.uint8, device="cuda", generator=g)
    mask = torch.empty_like(labels)
    cm = torch.from_numpy(ssb200.create_pascal_label_colormap().astype("uint8")).cuda()
    us = min(timeit(lambda i: ssb200.foreground_mask(labels, 288, 513, cm, out=mask)) for _ in range(2))
    print(json.dumps({"kernel": "fg_mask", "B": B, "us": round(us, 1), "GBps": round(2 * 513 * 513 * B / us / 1e3, 1)}))


if __name__ == "__main__":
    main()
