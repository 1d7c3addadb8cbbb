#!/usr/bin/env python
"""Per-source-line summary of an ncu report: joins the SASS page of one kernel
with the line table of the matching function in the built library.
usage: ncu_lines.py report.ncu-rep KernelFunctor [top]"""
import collections, csv, io, os, re, subprocess, sys, tempfile
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(root, "columba_b200", "libcolumba_b200.so")
tmp = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", lib], cwd=tmp, stdout=subprocess.DEVNULL)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
sass = subprocess.run(["nvdisasm", "-g", "-c", cubin], cwd=tmp, capture_output=True, text=True).stdout.split("\n")
start = [i for i, l in enumerate(sass) if l.startswith(".text.") and kern in l][0]
cur, instrs = None, []
for l in sass[start + 1:]:
    if l.startswith(".text.") or l.startswith(".section"):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        instrs.append((m.group(2), cur))
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name-base", "demangled", "-k",
                      "regex:" + kern], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr, data = rows[h], [r for r in rows[h + 1:] if r and r[0].startswith("0x")]
iI, iT, iS = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
n = min(len(data), len(instrs))
data = data[:n]
print("sass instrs: report %d, library %d" % (len(data), len(instrs)))
agg = collections.defaultdict(lambda: [0, 0, 0])
for k in range(n):
    a = agg[instrs[k][1]]
    a[0] += int(data[k][iI]); a[1] += int(data[k][iT]); a[2] += int(data[k][iS])
tot = sum(a[0] for a in agg.values()); ts = sum(a[2] for a in agg.values())
print("warp instructions %d, thread/warp-inst %.1f, samples %d" % (tot, sum(a[1] for a in agg.values()) / tot, ts))
for k, a in sorted(agg.items(), key=lambda x: -x[1][2])[:top]:
    print("%-28s inst %5.1f%%  thr/inst %4.1f  samples %5.1f%%" % (k, 100 * a[0] / tot, a[1] / max(a[0], 1), 100 * a[2] / ts))
