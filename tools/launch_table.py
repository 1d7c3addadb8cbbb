#!/usr/bin/env python
"""Per-kernel totals of one step from an ncu launch list (gpu__time_duration.sum csv)."""
import collections
import csv
import re
import sys


def load(path):
    lines = open(path).read().splitlines()
    i = [k for k, l in enumerate(lines) if l.startswith('"ID"')][0]
    rows = list(csv.DictReader(lines[i:]))
    out = []
    for r in rows:
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        n = r["Kernel Name"]
        m = re.search(r"<(?:clb::)?(\w+)", n)
        short = m.group(1) if (m and "clb_kernel" in n) else n.split("(")[0].replace("void ", "").replace("clb::", "")
        scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r["Metric Unit"], 1e-6)
        out.append((short, float(r["Metric Value"].replace(",", "")) * scale, r["Grid Size"], r["Block Size"]))
    return out


def steps_of(launches):
    steps, cur = [], []
    for s in launches:
        if s[0] == "PrepKernel" and cur and any(x[0] != "PrepKernel" for x in cur):
            steps.append(cur)
            cur = []
        cur.append(s)
    steps.append(cur)
    return steps


if __name__ == "__main__":
    st = steps_of(load(sys.argv[1]))
    k = int(sys.argv[2]) if len(sys.argv) > 2 else max(range(len(st)), key=lambda q: sum(x[1] for x in st[q]))
    agg = collections.OrderedDict()
    for name, ms, g, b in st[k]:
        key = name + (" " + b if name == "FrontierKernel" else "")
        agg.setdefault(key, [0, 0.0])
        agg[key][0] += 1
        agg[key][1] += ms
    tot = sum(v[1] for v in agg.values())
    print("step %d of %d: %d launches, %.3f ms" % (k, len(st), len(st[k]), tot))
    for key, v in agg.items():
        print("%-34s x%-3d %9.3f ms %5.1f%%" % (key, v[0], v[1], 100 * v[1] / tot))
