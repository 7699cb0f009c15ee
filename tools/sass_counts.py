"""Opcode histogram per kernel of the built library (cuobjdump -sass), for the instructions the design notes
name: bulk async copies (UBLKCP), mbarrier operations (SYNCS), packed integer / half maxima (VIMNMX3, HMNMX2),
Ampere-style async copies (LDGSTS), matches / votes / shuffles, cluster barriers (UCGABAR), atomics.

    python tools/sass_counts.py > profiles/rNN_sass_counts.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "semantic-segmentation-server_b200", "lib", "libssb200.so")
WATCH = ["UBLKCP", "UBLKPF", "SYNCS", "VIMNMX3", "VIMNMX", "HMNMX2", "PRMT", "LDGSTS", "LDS", "STS", "LDG", "STG", "ATOMS", "ATOMG", "RED",
         "MATCH", "VOTE", "SHFL", "UCGABAR_ARV", "UCGABAR_WAIT", "BAR", "MEMBAR", "ACQBULK", "UTMALDG", "HMMA", "IMMA", "UTCHMMA", "FFMA", "DFMA"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels, name = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"ssb::\(anonymous namespace\)::", "", name)
            kernels[name] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)((?:\.[A-Z0-9_]+)*)", line)
        if m and name:
            kernels[name][m.group(1)] += 1
            kernels[name]["_total"] += 1
            if m.group(1) in ("VIMNMX3", "VIMNMX", "HMNMX2", "MATCH", "LDS", "STS"):
                kernels[name][m.group(1) + m.group(2)] += 1
    print("# cuobjdump -sass %s: instructions per kernel (static counts), sm_100a" % os.path.relpath(LIB, ROOT))
    print("# tensor-core opcodes (HMMA / IMMA / UTCHMMA) and tensor-map TMA (UTMALDG) are expected to be absent: nothing here is a")
    print("# contraction and the streams are flat byte ranges (1-D bulk copies = UBLKCP)")
    for name, cnt in kernels.items():
        short = name if len(name) < 150 else name[:147] + "..."
        print("\n%s\n  total %d" % (short, cnt["_total"]))
        row = ["%s %d" % (op, cnt[op]) for op in WATCH if cnt[op]]
        print("  " + ", ".join(row))
        detail = ["%s %d" % (k, v) for k, v in sorted(cnt.items()) if "." in k and k.split(".")[0] in ("VIMNMX3", "HMNMX2", "MATCH")]
        if detail:
            print("  " + ", ".join(detail))


if __name__ == "__main__":
    main()
