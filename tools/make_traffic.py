#!/usr/bin/env python
"""profiles/r2_traffic.json from an ncu launch list that carries gpu__time_duration.sum, dram__bytes_read.sum
and dram__bytes_write.sum per launch (tools/gpu_profile.sh): DRAM bytes and device time of one step per stage.
usage: make_traffic.py launches.csv reads_per_step ref_bp [step index] > profiles/r2_traffic.json"""
import collections
import csv
import json
import re
import sys

STAGE = [("prep", r"PrepKernel"), ("partition", r"SeedKernel|GreedyKernel|EmitKernel"),
         ("search", r"FrontierKernel"), ("verify", r"LocateKernel|DedupKernel|ScreenKernel|VerifyKernel"),
         ("fm_expand", r"Fm\w+Kernel"), ("text_filter", r"Tx\w+Kernel|CigarJobKernel")]


def main():
    path, reads, ref_bp = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    lines = open(path).read().splitlines()
    i = [k for k, l in enumerate(lines) if l.startswith('"ID"')][0]
    rows = list(csv.DictReader(lines[i:]))
    launches = collections.OrderedDict()
    for r in rows:
        d = launches.setdefault(r["ID"], {"name": r["Kernel Name"]})
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1.0)
        d[r["Metric Name"]] = v * scale
    seq = list(launches.values())
    steps, cur = [], []
    for l in seq:
        if "PrepKernel" in l["name"] and cur and any("PrepKernel" not in x["name"] for x in cur):
            steps.append(cur)
            cur = []
        cur.append(l)
    steps.append(cur)
    # default: the full step with the longest device time (attempts that a list overflow voided are short)
    dur = [sum(x.get("gpu__time_duration.sum", 0.0) for x in st_) for st_ in steps]
    k = int(sys.argv[4]) if len(sys.argv) > 4 else max(range(len(steps)), key=lambda q: dur[q])
    st = steps[k]
    out = {"reads_per_step": reads, "ref_bp": ref_bp, "source": path, "step": k, "launches": len(st),
           "stage_dram_bytes": {}, "stage_ms_under_ncu": {}, "kernel_dram_bytes": {}}
    stage_of_sort = None
    for l in st:
        name = l["name"]
        stage = None
        for s, pat in STAGE:
            if re.search(pat, name):
                stage = s
        if stage is None:  # scans and radix sort passes belong to the stage whose kernels surround them
            stage = stage_of_sort or "fm_expand"
        if stage in ("fm_expand", "text_filter"):
            stage_of_sort = stage
        b = l.get("dram__bytes_read.sum", 0.0) + l.get("dram__bytes_write.sum", 0.0)
        out["stage_dram_bytes"][stage] = out["stage_dram_bytes"].get(stage, 0.0) + b
        out["stage_ms_under_ncu"][stage] = out["stage_ms_under_ncu"].get(stage, 0.0) + l.get("gpu__time_duration.sum", 0.0)
        m = re.search(r"<(?:clb::)?(\w+)", name)
        short = m.group(1) if (m and "clb_kernel" in name) else name.split("(")[0].replace("void ", "").replace("clb::", "")
        if "FrontierKernel" in name:
            short += "_cta" if "clb_kernel_cta" in name else "_warp"
        out["kernel_dram_bytes"][short] = out["kernel_dram_bytes"].get(short, 0.0) + b
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
