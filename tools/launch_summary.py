"""Summarises an ncu launch list (--metrics gpu__time_duration.sum --csv) per kernel and refreshes
profiles/traffic_per_launch.json from a --set full summary written by tools/ncu_summary.py.

    python tools/launch_summary.py <launches.csv> <out.csv> "<command that was profiled>"
    python tools/launch_summary.py --traffic profiles/r01_argmax_u8_ncu_full_summary.csv
"""
import collections
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def launches(src, out, cmd):
    rows = [r for r in csv.reader(l for l in open(src) if l.startswith('"'))]
    hdr, rows = rows[0], rows[1:]
    kn, val = hdr.index("Kernel Name"), hdr.index("Metric Value")
    tot = collections.OrderedDict()
    for r in rows:
        t = tot.setdefault(r[kn], [0, 0.0])
        t[0] += 1
        t[1] += float(r[val].replace(",", "")) / 1e3
    total = sum(t[1] for t in tot.values())
    with open(out, "w") as f:
        f.write("# ncu launch list summary: %s\n" % cmd)
        f.write("# (gpu__time_duration.sum, --clock-control none; per-launch times are cold-cache and serialised)\n")
        f.write("kernel,launches,total_us,avg_us,share\n")
        for k, (n, us) in tot.items():
            f.write('"%s",%d,%.1f,%.1f,%.3f\n' % (k, n, us, us / n, us / total))
    print(open(out).read())


def traffic(summary):
    rd = wr = None
    for r in csv.reader(open(summary)):
        if r and r[0] == "dram__bytes_read.sum":
            rd = [float(v) for v in r[2:]]
            unit = r[1]
        if r and r[0] == "dram__bytes_write.sum":
            wr = [float(v) for v in r[2:]]
    scale = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}[unit]
    per = sum(a + b for a, b in zip(rd, wr)) / len(rd) * scale
    out = {"argmax_stats_u8_b64_bytes": per,
           "source": "%s (dram__bytes_read.sum + dram__bytes_write.sum, mean of %d launches)" % (os.path.relpath(summary, ROOT), len(rd)),
           "algorithmic_bytes": 5790390 * 64}
    json.dump(out, open(os.path.join(ROOT, "profiles", "traffic_per_launch.json"), "w"), indent=1)
    print(out)


if __name__ == "__main__":
    if sys.argv[1] == "--traffic":
        traffic(sys.argv[2])
    else:
        launches(sys.argv[1], sys.argv[2], sys.argv[3])
