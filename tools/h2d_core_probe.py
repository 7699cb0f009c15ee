"""Does it matter WHICH CORE first touches a page-locked buffer?  (On a VM that shows one NUMA node, the pages of a
buffer still land on the physical node of the vCPU that touches them first.)

Phase 1, one GPU at a time: for a sample of cores c, a 256 MiB buffer is first touched by a thread bound to c, page-locked
(cudaHostRegister) and copied to the GPU alone: GB/s per (gpu, core).
Phase 2, all GPUs at once: every rank's buffer touched (a) by whatever core the process runs on, (b) by the best core
found for its GPU in phase 1, (c) by many threads (pages spread): per-rank GB/s and the aggregate, fixed time.

    python tools/h2d_core_probe.py [--gpus N] > profiles/rNN_h2d_core_probe.json
"""
import argparse
import json
import mmap
import multiprocessing as mp
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
MIB = 256


def touched_buffer(nbytes, cores=None, spread_threads=0):
    """Anonymous mapping whose pages are first touched by a thread bound to `cores` (or by `spread_threads` threads,
    64 KB chunks interleaved), as a numpy uint8 array."""
    import numpy as np
    m = mmap.mmap(-1, nbytes, flags=mmap.MAP_PRIVATE | mmap.MAP_ANONYMOUS)
    a = np.frombuffer(m, dtype=np.uint8)

    def touch(lo, step):
        if cores:
            os.sched_setaffinity(0, cores)
        for off in range(lo << 16, nbytes, step << 16):
            a[off:off + 65536:4096] = 1
    if spread_threads:
        ts = [threading.Thread(target=touch, args=(i, spread_threads)) for i in range(spread_threads)]
    else:
        ts = [threading.Thread(target=touch, args=(0, 1))]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    return m, a


def copy_rate(torch, a, seconds, barrier=None):
    rc = torch.cuda.cudart().cudaHostRegister(a.ctypes.data, a.nbytes, 0)
    assert int(rc) == 0, rc
    src = torch.from_numpy(a)
    dst = torch.empty(a.nbytes, dtype=torch.uint8, device="cuda")
    dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    if barrier is not None:
        barrier.wait(timeout=120)
    n, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < seconds:
        dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        n += 1
    el = time.perf_counter() - t0
    torch.cuda.cudart().cudaHostUnregister(a.ctypes.data)
    return n * a.nbytes / el / 1e9


def phase1(gpu, cores, q):
    try:
        import torch
        torch.cuda.set_device(gpu)
        out = {}
        for c in cores:
            m, a = touched_buffer(MIB << 20, {c})
            out[c] = copy_rate(torch, a, 0.25)
            del a
        q.put((gpu, out))
    except Exception as e:
        q.put((gpu, {"error": repr(e)}))


def phase2(gpu, mode, best, barrier, q):
    try:
        import torch
        torch.cuda.set_device(gpu)
        if mode == "best":
            m, a = touched_buffer(MIB << 20, set(best))
        elif mode == "spread":
            m, a = touched_buffer(MIB << 20, None, 16)
        else:
            m, a = touched_buffer(MIB << 20)
        q.put((gpu, copy_rate(torch, a, 1.0, barrier)))
    except Exception as e:
        q.put((gpu, repr(e)))
        try:
            barrier.abort()
        except Exception:
            pass


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=0)
    args = ap.parse_args()
    import torch
    n_gpu = args.gpus or torch.cuda.device_count()
    allowed = sorted(os.sched_getaffinity(0))
    sample = allowed[::max(1, len(allowed) // 8)][:8]
    ctx = mp.get_context("spawn")
    doc = {"cores_allowed": len(allowed), "cores_sampled": sample, "per_gpu_alone_GBps": {}, "concurrent": []}
    for g in range(n_gpu):
        q = ctx.Queue()
        p = ctx.Process(target=phase1, args=(g, sample, q))
        p.start()
        gpu, res = q.get(timeout=300)
        p.join(timeout=30)
        doc["per_gpu_alone_GBps"][str(gpu)] = res
        sys.stderr.write("gpu %d alone: %s\n" % (gpu, {k: round(v, 1) for k, v in res.items() if isinstance(v, float)}))
    best = {}
    for g in range(n_gpu):
        res = {k: v for k, v in doc["per_gpu_alone_GBps"][str(g)].items() if isinstance(v, float)}
        top = max(res.values()) if res else 0
        best[g] = [c for c, v in res.items() if v >= 0.97 * top] or allowed
    doc["best_cores"] = {str(g): best[g] for g in best}
    for mode in ("default", "best", "spread"):
        barrier, q = ctx.Barrier(n_gpu), ctx.Queue()
        ps = [ctx.Process(target=phase2, args=(g, mode, best[g], barrier, q)) for g in range(n_gpu)]
        for p in ps:
            p.start()
        res = dict(q.get(timeout=300) for _ in range(n_gpu))
        for p in ps:
            p.join(timeout=30)
        rates = [res.get(g) for g in range(n_gpu)]
        doc["concurrent"].append({"mode": mode, "per_rank_GBps": rates,
                                  "aggregate_GBps": sum(r for r in rates if isinstance(r, float))})
        sys.stderr.write("%d GPUs, %-7s: %s\n" % (n_gpu, mode, [round(r, 1) if isinstance(r, float) else r for r in rates]))
    print(json.dumps(doc, indent=1))


if __name__ == "__main__":
    main()
