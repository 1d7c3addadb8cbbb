#!/usr/bin/env python3
"""Source-line hot spots of one kernel from an ncu capture that was taken with --import-source on.

ncu's CSV export of the source page lists SASS instructions with their executed counts and warp-state samples but
without their source lines; nvdisasm -gi of the same build lists the same instructions with their (inlined) source
lines.  The two are joined instruction by instruction.

  ncu -i X.ncu-rep --page source --csv --print-source sass --launch-skip K --launch-count 1 > k.csv
  cuobjdump -xelf all columba_b200/libcolumba_b200.so && nvdisasm -gi clb_api.sm_100a.cubin > lib.dis
  tools/ncu_hotspots.py lib.dis <mangled kernel name substring> k.csv columba_b200/csrc/clb_frontier.h [top]
"""
import collections
import csv
import re
import sys


def main():
    dis, kname, srccsv, srcfile = sys.argv[1:5]
    top = int(sys.argv[5]) if len(sys.argv) > 5 else 40
    lines = open(dis).read().split("\n")
    start = [i for i, l in enumerate(lines) if l.startswith(".text.") and kname in l][0]
    end = start + 1
    while end < len(lines) and not lines[end].startswith(".text."):
        end += 1
    fileline = re.compile(r'//## File "([^"]+)", line (\d+)')
    fresh, chain, ins = True, [], []
    for l in lines[start:end]:
        m = fileline.search(l)
        if m:
            if fresh:
                chain, fresh = [], False
            chain.append((m.group(1).split("/")[-1], int(m.group(2))))
        elif re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
            fresh = True
            ins.append(list(chain))
    rows = list(csv.reader(open(srccsv)))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
    hdr, data = rows[hi], rows[hi + 1:]
    stop = [i for i, r in enumerate(data) if not (r and r[0].startswith("0x"))]
    if stop:  # ncu prints the table once per view
        data = data[:stop[0]]
    if len(ins) != len(data):
        sys.exit("instruction counts differ (%d in the listing, %d in the capture): not the same build" % (len(ins), len(data)))
    ci, ti, si = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
    stall = {h: i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h}
    src = open(srcfile).read().split("\n")
    sname = srcfile.split("/")[-1]
    inner, outer = collections.defaultdict(lambda: [0.0, 0.0, 0.0]), collections.defaultdict(lambda: [0.0, 0.0, 0.0])
    stalls = collections.Counter()
    tot = tots = 0.0
    for ch, r in zip(ins, data):
        n, t, s = float(r[ci] or 0), float(r[ti] or 0), float(r[si] or 0)
        mine = [c for c in ch if c[0] == sname]
        ki = mine[0] if mine else (ch[0] if ch else ("?", 0))
        ko = mine[-1] if mine else ki
        for d, k in ((inner, ki), (outer, ko)):
            d[k][0] += n
            d[k][1] += t
            d[k][2] += s
        tot += n
        tots += s
        for h, i in stall.items():
            try:
                stalls[h] += float(r[i] or 0)
            except ValueError:
                pass
    print("%d instructions; executed: %.4g warp instructions, %.4g thread instructions (%.1f lanes); %d samples"
          % (len(ins), tot, sum(v[1] for v in inner.values()), sum(v[1] for v in inner.values()) / max(tot, 1), tots))
    print("stall reasons, % of samples:", {k: round(v / tots * 100, 1) for k, v in stalls.most_common(8)})
    for title, d in (("outermost line in %s (the phase)" % sname, outer), ("innermost line in %s" % sname, inner)):
        print("-- by", title)
        print("  instr%  samples%  lanes  line")
        for k, v in sorted(d.items(), key=lambda kv: -kv[1][2])[:top]:
            txt = src[k[1] - 1].strip()[:90] if k[0] == sname and 0 < k[1] <= len(src) else ""
            print("  %5.1f   %5.1f    %4.1f   %s:%d  %s" % (v[0] / tot * 100, v[2] / tots * 100, v[1] / max(v[0], 1), k[0], k[1], txt))


if __name__ == "__main__":
    main()
