"""Device time of the contour-integral kernel (K6) on 64 blobby / noisy 513x513 label maps, ROI 288x513."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ssb200  # noqa: E402


def main():
    b, H, W, rh, rw = 64, 513, 513, 288, 513
    g = torch.Generator(device="cuda"); g.manual_seed(7)
    field = torch.randint(1, 21, (b, 9, 9), device="cuda", generator=g)
    field[torch.rand((b, 9, 9), device="cuda", generator=g) < 0.5] = 0
    field[field > 0] = torch.where(torch.rand(int((field > 0).sum()), device="cuda", generator=g) < 0.5, 15, 7)
    big = field.repeat_interleave(64, 1).repeat_interleave(64, 2)[:, :H, :W].to(torch.uint8).contiguous()
    noise = torch.randint(0, 21, (b, H, W), dtype=torch.uint8, device="cuda", generator=g)
    cells16 = torch.randint(1, 21, (b, 33, 33), device="cuda", generator=g)
    cells16[torch.rand((b, 33, 33), device="cuda", generator=g) < 0.7] = 0
    blobby = cells16.repeat_interleave(16, 1).repeat_interleave(16, 2)[:, :H, :W].to(torch.uint8).contiguous()
    one = torch.zeros((1, H, W), dtype=torch.uint8, device="cuda")
    one[0, 20:270, 30:480] = 15
    one[0, 60:230, 90:420] = 0
    for name, lab in (("64-px cells", big), ("16-px cells", blobby), ("noise", noise), ("1 frame, ring", one)):
        b = lab.shape[0]
        mask = ssb200.foreground_mask(lab, rh, rw)
        t = ssb200.blob_regions(mask, lab, rh, rw, 21, 1024)
        for _ in range(3):
            ssb200.contour_moments(mask, t, rh, rw)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            ssb200.contour_moments(mask, t, rh, rw)
        e1.record()
        torch.cuda.synchronize()
        print("%-12s contour integrals %8.1f us per %d frames, %.1f regions per frame" %
              (name, e0.elapsed_time(e1) / 10 * 1e3, b, t["count"].float().mean().item()))


if __name__ == "__main__":
    main()
