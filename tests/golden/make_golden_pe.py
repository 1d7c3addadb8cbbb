#!/usr/bin/env python
"""Paired-end golden fixtures, produced by RUNNING THE UNMODIFIED REFERENCE
(oracle/_ref/columba_dev -f R1 -F R2 -nI ...) on pairs drawn from the committed
golden references.  Build container only.

  python tests/golden/make_golden_pe.py

Pairs: FR fragments of length ~N(400, 50) in [200, 700] with 0..3 edits per
mate; besides those, pairs whose mates lie on different sequences or too far
apart (discordant), pairs with one random mate (half mapped), pairs of two
random reads (unmapped), pairs in FF orientation, and -- on the repeat-rich
reference -- pairs inside the high-copy family (many pairs per read pair).
Writes pe_<set>_1.fq.gz / _2.fq.gz, expected_pe/<case>.tsv.gz (every SAM record:
qname flag rname pos mapq cigar rnext pnext tlen NM) + .json (pair counters),
pe_cases.json.
"""
import gzip
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import gen_data  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref")
COMP = bytes.maketrans(b"ACGT", b"TGCA")

CASES = {
    # name: (set, reference, mode args, parameters)
    "pe_all_k2": ("g1", ["-a", "all", "-e", "2"], dict(mode="all", value=2, ori="fr", max_frag=700, min_frag=200, disc=0, x=0)),
    "pe_all_k3_disc": ("g1", ["-a", "all", "-e", "3", "-D"], dict(mode="all", value=3, ori="fr", max_frag=700, min_frag=200, disc=1, x=0)),
    "pe_all_k2_ff": ("g1", ["-a", "all", "-e", "2", "-O", "ff"], dict(mode="all", value=2, ori="ff", max_frag=700, min_frag=200, disc=0, x=0)),
    "pe_all_k2_rf_disc": ("g1", ["-a", "all", "-e", "2", "-O", "rf", "-D"], dict(mode="all", value=2, ori="rf", max_frag=700, min_frag=200, disc=1, x=0)),
    "pe_best_I95": ("g1", ["-a", "best", "-I", "95"], dict(mode="best", value=95, ori="fr", max_frag=700, min_frag=200, disc=0, x=0)),
    "pe_best_I95_disc": ("g1", ["-a", "best", "-I", "95", "-D"], dict(mode="best", value=95, ori="fr", max_frag=700, min_frag=200, disc=1, x=0)),
    "pe_best_I97_x1": ("g1", ["-a", "best", "-I", "97", "-x", "1"], dict(mode="best", value=97, ori="fr", max_frag=700, min_frag=200, disc=0, x=1)),
    "pe_best_I95_ff_disc": ("g1", ["-a", "best", "-I", "95", "-O", "ff", "-D"], dict(mode="best", value=95, ori="ff", max_frag=700, min_frag=200, disc=1, x=0)),
    # parameter inference on (the reference's default): all pairs go through the single-end pre-pass
    "pe_infer_all_k2": ("g1", ["-a", "all", "-e", "2"], dict(mode="all", value=2, ori="fr", max_frag=500, min_frag=0, disc=0, x=0, infer=1)),
    "pe_infer_best_I95": ("g1", ["-a", "best", "-I", "95"], dict(mode="best", value=95, ori="fr", max_frag=500, min_frag=0, disc=0, x=0, infer=1)),
    "pe_infer_best_I95_disc": ("g1", ["-a", "best", "-I", "95", "-D"], dict(mode="best", value=95, ori="fr", max_frag=500, min_frag=0, disc=1, x=0, infer=1)),
    "pe_g2_infer_all_k2_disc": ("g2", ["-a", "all", "-e", "2", "-D"], dict(mode="all", value=2, ori="fr", max_frag=500, min_frag=0, disc=1, x=0, infer=1)),
    "pe_g2_all_k2": ("g2", ["-a", "all", "-e", "2"], dict(mode="all", value=2, ori="fr", max_frag=700, min_frag=200, disc=0, x=0)),
    "pe_g2_best_I95_disc": ("g2", ["-a", "best", "-I", "95", "-D"], dict(mode="best", value=95, ori="fr", max_frag=700, min_frag=200, disc=1, x=0)),
}
LOGPAT = {"pairs": r"Total no\. matches : (\d+)", "mapped": r"Mapped pairs: (\d+)",
          "discordant": r"Discordantly mapped pairs: (\d+)", "half": r"No\. unpaired reads that did match: (\d+)",
          "unpaired": r"Total read pairs both mapped but unpaired: (\d+)"}


def revcomp(b):
    return b.translate(COMP)[::-1]


def mutate(rng, seq, e):
    s = bytearray(seq)
    for _ in range(e):
        p = int(rng.integers(0, len(s)))
        kind = rng.random()
        if kind < 0.7:
            s[p] = b"ACGT"[(b"ACGT".index(bytes([s[p]])) + 1 + int(rng.integers(0, 3))) % 4] if bytes([s[p]]) in b"ACGT" else s[p]
        elif kind < 0.85:
            s.insert(p, b"ACGT"[int(rng.integers(0, 4))])
        else:
            del s[p]
    return bytes(s)


def make_pairs(seqs, n, seed, read_len=150):
    rng = np.random.default_rng(seed)
    out1, out2 = [], []
    cat_lens = [len(s) for s in seqs]

    def frag(sid, p, ln):
        return seqs[sid][p:p + ln].tobytes()

    for i in range(n):
        kind = rng.random()
        sid = int(rng.integers(0, len(seqs)))
        ln = int(np.clip(rng.normal(400, 50), 200, 700))
        p = int(rng.integers(0, cat_lens[sid] - 800))
        f = frag(sid, p, ln)
        m1, m2 = f[:read_len], revcomp(f[-read_len:])  # FR
        tag = "fr"
        if kind < 0.06:        # mates on different sequences / far apart
            sid2 = int(rng.integers(0, len(seqs)))
            p2 = int(rng.integers(0, cat_lens[sid2] - 800))
            m2 = revcomp(frag(sid2, p2, read_len))
            tag = "far"
        elif kind < 0.10:      # one mate random
            m2 = gen_data.ACGT[rng.integers(0, 4, read_len)].tobytes()
            tag = "half"
        elif kind < 0.12:      # both random
            m1 = gen_data.ACGT[rng.integers(0, 4, read_len)].tobytes()
            m2 = gen_data.ACGT[rng.integers(0, 4, read_len)].tobytes()
            tag = "none"
        elif kind < 0.18:      # FF
            m2 = f[-read_len:]
            tag = "ff"
        elif kind < 0.22:      # RF
            m1, m2 = revcomp(f[:read_len]), f[-read_len:]
            tag = "rf"
        if rng.random() < 0.5:  # which mate is read 1
            m1, m2 = m2, m1
        e1, e2 = int(rng.integers(0, 4)), int(rng.integers(0, 4))
        out1.append(("p%d_%s/1" % (i, tag), mutate(rng, m1, e1)))
        out2.append(("p%d_%s/2" % (i, tag), mutate(rng, m2, e2)))
    return out1, out2


def read_fasta(path):
    seqs, cur = [], []
    for line in open(path):
        if line.startswith(">"):
            if cur:
                seqs.append(np.frombuffer("".join(cur).encode(), dtype=np.uint8))
            cur = []
        else:
            cur.append(line.strip())
    seqs.append(np.frombuffer("".join(cur).encode(), dtype=np.uint8))
    return seqs


def sam_records(path):
    rows = []
    for line in open(path):
        if line.startswith("@"):
            continue
        f = line.rstrip("\n").split("\t")
        nm = [x for x in f[11:] if x.startswith("NM:i:")]
        rows.append((f[0], int(f[1]), f[2], int(f[3]), int(f[4]), f[5], f[6], int(f[7]), int(f[8]),
                     int(nm[0][5:]) if nm else -1))
    rows.sort()
    return rows


def main():
    tmp = tempfile.mkdtemp(prefix="goldenpe_")
    out = os.path.join(HERE, "expected_pe")
    shutil.rmtree(out, ignore_errors=True)
    os.makedirs(out)
    for g, n, seed in (("g1", 400, 71), ("g2", 150, 72)):
        with gzip.open(os.path.join(HERE, g + ".fa.gz"), "rb") as i, open(os.path.join(tmp, g + ".fa"), "wb") as o:
            o.write(i.read())
        subprocess.check_call([os.path.join(REF, "columba_build"), "-r", os.path.join(tmp, g), "-f",
                               os.path.join(tmp, g + ".fa")], stdout=subprocess.DEVNULL)
        r1, r2 = make_pairs(read_fasta(os.path.join(tmp, g + ".fa")), n, seed)
        for k, rr in ((1, r1), (2, r2)):
            fq = os.path.join(tmp, "pe_%s_%d.fq" % (g, k))
            gen_data.write_fastq(fq, rr)
            with open(fq, "rb") as i, gzip.GzipFile(os.path.join(HERE, "pe_%s_%d.fq.gz" % (g, k)), "wb", mtime=0) as o:
                o.write(i.read())
    meta = {}
    for name, (g, args, params) in CASES.items():
        sam = os.path.join(tmp, name + ".sam")
        cmd = [os.path.join(REF, "columba_dev"), "-r", os.path.join(tmp, g), "-f", os.path.join(tmp, "pe_%s_1.fq" % g),
               "-F", os.path.join(tmp, "pe_%s_2.fq" % g), "-t", "1", "-v", "-o", sam] + args
        if not params.get("infer"):
            cmd += ["-nI", "-X", str(params["max_frag"]), "-N", str(params["min_frag"])]
            if "-O" not in args:
                cmd += ["-O", "fr"]
        log = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
        if log.returncode != 0:
            print(log.stdout[-1500:], log.stderr[-1500:])
            raise SystemExit("reference failed on " + name)
        rows = sam_records(sam)
        with gzip.GzipFile(os.path.join(out, name + ".tsv.gz"), "wb", mtime=0) as o:
            for r in rows:
                o.write(("\t".join(map(str, r)) + "\n").encode())
        ctr = {}
        for key, pat in LOGPAT.items():
            m = re.search(pat, log.stdout)
            ctr[key] = int(m.group(1)) if m else None
        json.dump(ctr, open(os.path.join(out, name + ".json"), "w"), sort_keys=True)
        meta[name] = dict(set=g, **params)
        inf = re.findall(r"(?i)(inferred[^\n]*|orientation[^\n]*|insert size[^\n]*)", log.stdout)
        print(name, len(rows), ctr, inf[-4:] if params.get("infer") else "", flush=True)
    json.dump(meta, open(os.path.join(HERE, "pe_cases.json"), "w"), indent=1, sort_keys=True)
    shutil.rmtree(tmp)


if __name__ == "__main__":
    main()
