"""Host->device copy rate of N GPUs pulling at the same time from pinned host memory, by where that
memory sits:

    none        no CPU binding, pages first-touched wherever the process happens to run
    nvml        the process is bound to NVML's ideal CPUs for its GPU first (what bench.py did in round 1)
    sysfs       CPUs and pages bound to the NUMA node sysfs reports for the GPU's PCIe function
    interleave  pages interleaved over every node the process may allocate on

    python tools/h2d_probe.py [--gpus 1,2,4,8] [--mib 1024] [--seconds 1.0] > profiles/rNN_h2d_probe.json

One process per GPU (spawned here), all copies start behind a barrier; per rank GB/s = bytes / own time,
aggregate = all bytes / slowest rank's time.  Prints one JSON document.
"""
import argparse
import json
import multiprocessing as mp
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def worker(rank, world, mode, mib, seconds, barrier, out_q):
    try:
        import importlib

        import torch
        import ssb200
        hostmem = importlib.import_module(ssb200.__name__ + ".hostmem")
        torch.cuda.set_device(rank)
        info = {"rank": rank, "gpu_node_sysfs": hostmem.gpu_numa_node(rank), "bdf": hostmem.gpu_pci_bdf(rank)}
        nbytes = mib << 20
        if mode == "nvml":
            info["nvml_cores"] = ssb200.bind_process_to_device_numa(rank)
            buf = hostmem.PlacedPinned(nbytes, "default")
        elif mode == "sysfs":
            node = info["gpu_node_sysfs"]
            if node is not None and node in hostmem.allowed_nodes():
                cpus = hostmem.node_cpus(node) & os.sched_getaffinity(0)
                if cpus:
                    os.sched_setaffinity(0, cpus)
                buf = hostmem.PlacedPinned(nbytes, "bind", [node])
            else:
                buf = hostmem.PlacedPinned(nbytes, "default")
        elif mode == "interleave":
            buf = hostmem.PlacedPinned(nbytes, "interleave", hostmem.allowed_nodes())
        else:
            buf = hostmem.PlacedPinned(nbytes, "default")
        info["policy_applied"] = buf.policy_applied
        info["pages_on_nodes"] = buf.nodes_of_pages()
        info["cpus"] = len(os.sched_getaffinity(0))
        src = torch.from_numpy(buf.array)
        dst = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
        stream = torch.cuda.Stream()
        with torch.cuda.stream(stream):
            for _ in range(2):
                dst.copy_(src, non_blocking=True)
            stream.synchronize()
            barrier.wait(timeout=120)
            n, t0 = 0, time.perf_counter()
            while time.perf_counter() - t0 < seconds:
                dst.copy_(src, non_blocking=True)
                stream.synchronize()
                n += 1
            el = time.perf_counter() - t0
        info.update(copies=n, seconds=el, GBps=n * nbytes / el / 1e9)
        barrier.wait(timeout=120)
        del src
        buf.close()
        out_q.put(info)
    except Exception as e:      # never leave the parent waiting
        out_q.put({"rank": rank, "error": repr(e)})
        try:
            barrier.abort()
        except Exception:
            pass


def run(world, mode, mib, seconds):
    ctx = mp.get_context("spawn")
    barrier, q = ctx.Barrier(world), ctx.Queue()
    procs = [ctx.Process(target=worker, args=(r, world, mode, mib, seconds, barrier, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = []
    for _ in range(world):
        try:
            res.append(q.get(timeout=300))
        except Exception:
            res.append({"error": "timeout"})
    for p in procs:
        p.join(timeout=30)
        if p.is_alive():
            p.terminate()
    res.sort(key=lambda r: r.get("rank", 99))
    good = [r for r in res if "GBps" in r]
    agg = sum(r["copies"] for r in good) * (mib << 20) / max(r["seconds"] for r in good) / 1e9 if good else None
    return {"gpus": world, "mode": mode, "aggregate_GBps": agg, "per_rank": res}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", default="1,2,4,8")
    ap.add_argument("--modes", default="none,nvml,sysfs,interleave")
    ap.add_argument("--mib", type=int, default=1024)
    ap.add_argument("--seconds", type=float, default=1.0)
    args = ap.parse_args()
    import torch
    have = torch.cuda.device_count()
    doc = {"host": {"cpus_allowed": len(os.sched_getaffinity(0)), "cpu_count": os.cpu_count(), "gpus_visible": have},
           "runs": []}
    try:
        with open("/proc/self/status") as f:
            doc["host"]["status"] = {l.split(":")[0]: l.split(":", 1)[1].strip() for l in f
                                     if l.startswith(("Mems_allowed_list", "Cpus_allowed_list"))}
        doc["host"]["nodes"] = sorted(int(d[4:]) for d in os.listdir("/sys/devices/system/node") if d.startswith("node") and d[4:].isdigit())
    except OSError:
        pass
    for n in [int(x) for x in args.gpus.split(",")]:
        if n > have:
            continue
        for mode in args.modes.split(","):
            doc["runs"].append(run(n, mode, args.mib, args.seconds))
            sys.stderr.write("%d GPUs %-10s %s GB/s\n" % (n, mode, doc["runs"][-1]["aggregate_GBps"]))
    print(json.dumps(doc, indent=1))


if __name__ == "__main__":
    main()
