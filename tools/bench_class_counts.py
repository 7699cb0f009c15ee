"""Rate of ssb_argmax_stats_u8 against the class count C (batch sized to ~350 MB of scores for every C):
streaming kernel for 2 <= C <= 32, generic kernel beyond and with SSB_ARGMAX_CFG=generic.

    python tools/bench_class_counts.py [C ...]
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ssb200  # noqa: E402

H = W = 513


def time_us(fn, n=30):
    for i in range(5):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


def main():
    cs = [int(a) for a in sys.argv[1:]] or list(range(2, 34)) + [40, 64]
    g = torch.Generator(device="cuda"); g.manual_seed(0)
    for c in cs:
        b = max(16, (350_000_000 // (H * W * c)) // 16 * 16)
        q = [torch.randint(0, 256, (b, H, W, c), dtype=torch.uint8, device="cuda", generator=g) for _ in range(3)]
        lab = torch.empty((b, H, W), dtype=torch.uint8, device="cuda")
        st = torch.empty((b, c, 4), dtype=torch.int64, device="cuda")
        us = time_us(lambda i: ssb200.argmax_stats(q[i % 3], labels=lab, stats=st))
        gbs = b * (H * W * c + H * W + c * 32) / us / 1e3
        os.environ["SSB_ARGMAX_CFG"] = "generic"
        us_g = time_us(lambda i: ssb200.argmax_stats(q[i % 3], labels=lab, stats=st), 5)
        os.environ.pop("SSB_ARGMAX_CFG")
        print("C=%2d  batch %4d  %7.1f us  %6.0f GB/s   (generic kernel: %6.0f GB/s)" %
              (c, b, us, gbs, b * (H * W * c + H * W + c * 32) / us_g / 1e3), flush=True)
        del q


def planar():
    """The same sweep for the planar entry points (fp32 and fp16, [B,C,513,513])."""
    cs = [int(a) for a in sys.argv[2:]] or [2, 3, 4, 8, 12, 16, 19, 20, 21, 24, 28, 32, 33]
    g = torch.Generator(device="cuda"); g.manual_seed(0)
    for dt in (torch.float32, torch.float16):
        for c in cs:
            es = 4 if dt == torch.float32 else 2
            b = max(8, (700_000_000 // (H * W * c * es)) // 8 * 8)
            q = [torch.randn((b, c, H, W), dtype=dt, device="cuda", generator=g) for _ in range(2)]
            lab = torch.empty((b, H, W), dtype=torch.uint8, device="cuda")
            st = torch.empty((b, c, 4), dtype=torch.int64, device="cuda")
            us = time_us(lambda i: ssb200.argmax_stats_nchw(q[i % 2], labels=lab, stats=st), 20)
            gbs = b * (H * W * c * es + H * W + c * 32) / us / 1e3
            os.environ["SSB_ARGMAX_NCHW_CFG"] = "generic"
            us_g = time_us(lambda i: ssb200.argmax_stats_nchw(q[i % 2], labels=lab, stats=st), 3)
            os.environ.pop("SSB_ARGMAX_NCHW_CFG")
            print("planar %s C=%2d  batch %4d  %7.1f us  %6.0f GB/s   (generic kernel: %6.0f GB/s)" %
                  ("fp32" if es == 4 else "fp16", c, b, us, gbs, b * (H * W * c * es + H * W + c * 32) / us_g / 1e3), flush=True)
            del q


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "planar":
        planar()
    else:
        main()
