#!/usr/bin/env python
"""SAM -> canonical occurrence tuples (read, strand flag, seq, pos, dist, CIGAR), sorted."""
import sys

def sam_tuples(path, flag_mask=16):
    out = []
    for line in open(path):
        if line.startswith("@"):
            continue
        f = line.rstrip("\n").split("\t")
        flag = int(f[1])
        if flag & 4:
            continue
        nm = [x for x in f[11:] if x.startswith("NM:i:")]
        dist = int(nm[0][5:]) if nm else -1
        out.append((f[0], flag & flag_mask, f[2], int(f[3]), dist, f[5]))
    out.sort()
    return out

def tsv_tuples(path):
    out = []
    for line in open(path):
        f = line.rstrip("\n").split("\t")
        out.append((f[0], int(f[1]), f[2], int(f[3]), int(f[4]), f[5]))
    out.sort()
    return out

if __name__ == "__main__":
    a = sam_tuples(sys.argv[1]) if sys.argv[1].endswith(".sam") else tsv_tuples(sys.argv[1])
    b = sam_tuples(sys.argv[2]) if sys.argv[2].endswith(".sam") else tsv_tuples(sys.argv[2])
    sa, sb = set(a), set(b)
    print("A=%d B=%d common=%d onlyA=%d onlyB=%d" % (len(a), len(b), len(sa & sb), len(sa - sb), len(sb - sa)))
    for t in sorted(sa - sb)[:10]:
        print("onlyA", t)
    for t in sorted(sb - sa)[:10]:
        print("onlyB", t)
    sys.exit(0 if sa == sb and len(a) == len(b) else 1)
