#!/usr/bin/env python
"""Generate the golden fixtures under tests/golden/ by RUNNING THE UNMODIFIED
REFERENCE (oracle/_ref/columba_dev, columba_build -- built from /root/reference
by `make -C oracle ref`).  Run in the build container only; the outputs are
committed so the GPU box (which has no /root/reference) can check against them.

  python tests/golden/make_golden.py

Writes:  g1.fa.gz, g1_edit.fq.gz, g1_ham.fq.gz          (inputs)
         schemes/pigeon5/, schemes/multi4/                (custom scheme dirs)
         expected/<case>.tsv.gz + expected/<case>.json    (reference outputs)
         cases.json                                       (case -> parameters)
"""
import gzip
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import gen_data  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref")


def pigeon_searches(k):
    """Pigeonhole scheme for k errors, P=k+1 parts (same shape as the reference's
    search_schemes/pigeon: search i starts at part i, goes right, then left)."""
    P = k + 1
    lines = []
    for i in range(P):
        order = list(range(i, P)) + list(range(i - 1, -1, -1))
        L = [0] * P
        U = [0] + [k] * (P - 1)
        lines.append("{%s} {%s} {%s}" % (",".join(map(str, order)), ",".join(map(str, L)), ",".join(map(str, U))))
    return lines


def write_scheme_dirs():
    sd = os.path.join(HERE, "schemes")
    shutil.rmtree(sd, ignore_errors=True)
    # custom scheme: pigeonhole for k = 1..5
    p = os.path.join(sd, "pigeon5")
    os.makedirs(p)
    open(os.path.join(p, "name.txt"), "w").write("pigeon5\n")
    for k in range(1, 6):
        os.makedirs(os.path.join(p, str(k)))
        open(os.path.join(p, str(k), "searches.txt"), "w").write("\n".join(pigeon_searches(k)) + "\n")
    # a custom scheme with its own partitioning parameters at k = 3
    open(os.path.join(p, "3", "static_partitioning.txt"), "w").write("0.22 0.5 0.78\n")
    open(os.path.join(p, "3", "dynamic_partitioning.txt"), "w").write("0.3 0.7\n3 1 1 2\n")
    # multiple schemes with dynamic selection at k = 2 and 4 (same #parts each)
    m = os.path.join(sd, "multi4")
    os.makedirs(m)
    open(os.path.join(m, "name.txt"), "w").write("multi4\n")
    builtin = os.path.join(ROOT, "columba_b200", "schemes")
    for k in (2, 4):
        os.makedirs(os.path.join(m, str(k)))
        srcs = [open(os.path.join(builtin, "kuch1", str(k), "searches.txt")).read(),
                "\n".join(pigeon_searches(k)) + "\n",
                open(os.path.join(builtin, "minU", str(k), "searches.txt")).read()]
        for x, s in enumerate(srcs, 1):
            open(os.path.join(m, str(k), "scheme%d.txt" % x), "w").write(s)


CASES = {
    # name: (reads, reference-CLI args, product/oracle parameters)
    "edit_k1_columba": ("edit", ["-e", "1"], dict(k=1)),
    "edit_k2_columba": ("edit", ["-e", "2"], dict(k=2)),
    "edit_k3_columba": ("edit", ["-e", "3"], dict(k=3)),
    "edit_k4_columba": ("edit", ["-e", "4"], dict(k=4)),
    "edit_k6_columba": ("edit", ["-e", "6"], dict(k=6)),
    "edit_k4_kuch1": ("edit", ["-e", "4", "-S", "kuch1"], dict(k=4, scheme=("named", "kuch1"))),
    "edit_k3_kuch2": ("edit", ["-e", "3", "-S", "kuch2"], dict(k=3, scheme=("named", "kuch2"), kmer=4)),  # -K is capped at 4 for kuch2/01*0 (alignparameters.cpp:1272-1279)
    "edit_k2_01star0": ("edit", ["-e", "2", "-S", "01*0"], dict(k=2, scheme=("named", "01*0"), kmer=4)),
    "edit_k4_minU": ("edit", ["-e", "4", "-S", "minU"], dict(k=4, scheme=("named", "minU"))),
    "edit_k4_pigeon_custom": ("edit", ["-e", "4", "-c", "@pigeon5", "-nD"], dict(k=4, scheme=("custom", "pigeon5"))),
    "edit_k3_pigeon_custom_dyn": ("edit", ["-e", "3", "-c", "@pigeon5"], dict(k=3, scheme=("custom_dynamic", "pigeon5"))),
    "edit_k4_multi": ("edit", ["-e", "4", "-d", "@multi4"], dict(k=4, scheme=("multi", "multi4"))),
    "edit_k3_uniform": ("edit", ["-e", "3", "-p", "uniform"], dict(k=3, partition="uniform")),
    "edit_k3_static": ("edit", ["-e", "3", "-p", "static"], dict(k=3, partition="static")),
    "edit_k3_pigeon_static": ("edit", ["-e", "3", "-c", "@pigeon5", "-nD", "-p", "static"],
                              dict(k=3, scheme=("custom", "pigeon5"), partition="static")),
    "edit_k3_pigeon_dynparams": ("edit", ["-e", "3", "-c", "@pigeon5", "-nD"], dict(k=3, scheme=("custom", "pigeon5"))),
    "edit_k4_i0": ("edit", ["-e", "4", "-i", "0"], dict(k=4, in_text=0)),
    "edit_k4_i50": ("edit", ["-e", "4", "-i", "50"], dict(k=4, in_text=50)),
    "edit_k4_K8": ("edit", ["-e", "4", "-K", "8"], dict(k=4, kmer=8)),
    "ham_k1_columba": ("ham", ["-m", "hamming", "-e", "1"], dict(k=1, metric="hamming")),
    "ham_k4_columba": ("ham", ["-m", "hamming", "-e", "4"], dict(k=4, metric="hamming")),
    "ham_k5_pigeon_custom": ("ham", ["-m", "hamming", "-e", "5", "-c", "@pigeon5", "-nD"],
                             dict(k=5, metric="hamming", scheme=("custom", "pigeon5"))),
    "ham_k4_multi": ("ham", ["-m", "hamming", "-e", "4", "-d", "@multi4"],
                     dict(k=4, metric="hamming", scheme=("multi", "multi4"))),
    "ham_k3_i0": ("ham", ["-m", "hamming", "-e", "3", "-i", "0"], dict(k=3, metric="hamming", in_text=0)),
}

LOGPAT = {
    "nodes": r"Total no\. Nodes: (\d+)",
    "matches": r"Total no\. matches: (\d+)",
    "reported": r"Total no\. reported matches: (\d+)",
    "mapped": r"Mapped reads: (\d+)",
    "reads": r"Number of reads: (\d+)",
    "intext": r"In text verification procedures (\d+)",
    "aborted": r"Failed in-text verifications procedures: (\d+)",
    "immediate": r"Immediate switch after first part: (\d+)",
    "searches": r"Searches started \(does not include immediate switches\) : (\d+)",
}


def main():
    write_scheme_dirs()
    # ---- inputs -----------------------------------------------------------
    names, seqs = gen_data.gen_reference(300_000, 11, n_seqs=3, profile="dense")
    tmp = tempfile.mkdtemp(prefix="golden_")
    fa = os.path.join(tmp, "g1.fa")
    gen_data.write_fasta(fa, names, seqs)
    edit_reads = gen_data.gen_reads(seqs, 600, 150, 150, "edit", 4, over_e=6, n_frac=0.03)
    # a few reads that straddle sequence boundaries (exercise trimming)
    cat = b"".join(s.tobytes() for s in seqs)
    off = 0
    extra = []
    for i, s in enumerate(seqs[:-1]):
        off += len(s)
        for d in (-148, -140, -75, -10, -3):
            extra.append(("b%d_%d" % (i, -d), cat[off + d: off + d + 150]))
    edit_reads += extra
    ham_reads = gen_data.gen_reads(seqs, 400, 100, 100, "hamming", 5, over_e=7, n_frac=0.02)
    gen_data.write_fastq(os.path.join(tmp, "g1_edit.fq"), edit_reads)
    gen_data.write_fastq(os.path.join(tmp, "g1_ham.fq"), ham_reads)
    for f in ("g1.fa", "g1_edit.fq", "g1_ham.fq"):
        with open(os.path.join(tmp, f), "rb") as i, gzip.GzipFile(os.path.join(HERE, f + ".gz"), "wb", mtime=0) as o:
            o.write(i.read())
    subprocess.check_call([os.path.join(REF, "columba_build"), "-r", os.path.join(tmp, "g1"), "-f", fa],
                          stdout=subprocess.DEVNULL)
    # keep the reference-built index too: the product's index writer is checked byte-for-byte against it
    idx_dir = os.path.join(HERE, "ref_index")
    shutil.rmtree(idx_dir, ignore_errors=True)
    os.makedirs(idx_dir)
    import hashlib
    sums = {}
    for f in sorted(os.listdir(tmp)):
        if f.startswith("g1.") and f != "g1.fa":
            sums[f] = hashlib.sha256(open(os.path.join(tmp, f), "rb").read()).hexdigest()
    json.dump(sums, open(os.path.join(idx_dir, "sha256.json"), "w"), indent=1, sort_keys=True)
    # ---- reference runs ---------------------------------------------------
    exp = os.path.join(HERE, "expected")
    shutil.rmtree(exp, ignore_errors=True)
    os.makedirs(exp)
    meta = {}
    for name, (rd, args, params) in CASES.items():
        a = [x.replace("@", os.path.join(HERE, "schemes") + "/") + ("/" if x.startswith("@") else "") for x in args]
        sam = os.path.join(tmp, name + ".sam")
        cmd = [os.path.join(REF, "columba_dev"), "-r", os.path.join(tmp, "g1"), "-f",
               os.path.join(tmp, "g1_%s.fq" % rd), "-a", "all", "-t", "1", "-v", "-o", sam] + a
        log = subprocess.run(cmd, capture_output=True, text=True)
        if log.returncode != 0:
            print(log.stdout, log.stderr)
            raise SystemExit("reference failed on " + name)
        rows = []
        for line in open(sam):
            if line.startswith("@"):
                continue
            f = line.rstrip("\n").split("\t")
            flag = int(f[1])
            if flag & 4:
                continue
            nm = [x for x in f[11:] if x.startswith("NM:i:")]
            rows.append((f[0], flag & 16, f[2], int(f[3]), int(nm[0][5:]), f[5]))
        rows.sort()
        with gzip.GzipFile(os.path.join(exp, name + ".tsv.gz"), "wb", mtime=0) as o:
            for r in rows:
                o.write(("\t".join(map(str, r)) + "\n").encode())
        ctr = {}
        for key, pat in LOGPAT.items():
            m = re.search(pat, log.stdout)
            ctr[key] = int(m.group(1)) if m else None
        json.dump(ctr, open(os.path.join(exp, name + ".json"), "w"), sort_keys=True)
        meta[name] = dict(reads=rd, **params)
        print(name, len(rows), ctr)
    json.dump(meta, open(os.path.join(HERE, "cases.json"), "w"), indent=1, sort_keys=True)
    shutil.rmtree(tmp)


if __name__ == "__main__":
    main()
