#!/usr/bin/env python
"""Golden cases for the reference's OWN search_schemes/ directories (copied as
data into tests/golden/schemes/ref), mapped through the custom (-c DIR -nD),
custom + dynamic selection (-c DIR) and multiple-scheme (-d DIR) paths of the
unmodified reference (oracle/_ref/columba_dev).  Build container only.

  python tests/golden/make_golden_schemes.py
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
sys.path.insert(0, HERE)
from make_golden import LOGPAT, REF  # noqa: E402

S = "ref/"
CASES = {
    # name: (reads, CLI args ("@" = schemes dir), parameters)
    "ref_kuch1_k2_dyn": ("edit", ["-e", "2", "-c", "@kuch_k+1"], dict(k=2, scheme=("custom_dynamic", S + "kuch_k+1"))),
    "ref_kuch1_k4": ("edit", ["-e", "4", "-c", "@kuch_k+1", "-nD"], dict(k=4, scheme=("custom", S + "kuch_k+1"))),
    "ref_kuch1_adapted_k3": ("edit", ["-e", "3", "-c", "@kuch_k+1_adapted", "-nD"],
                             dict(k=3, scheme=("custom", S + "kuch_k+1_adapted"))),
    "ref_kuch2_k3_dyn": ("edit", ["-e", "3", "-c", "@kuch_k+2"], dict(k=3, scheme=("custom_dynamic", S + "kuch_k+2"))),
    "ref_kuch2_adapted_k4": ("edit", ["-e", "4", "-c", "@kuch_k+2_adapted", "-nD"],
                             dict(k=4, scheme=("custom", S + "kuch_k+2_adapted"))),
    "ref_kianfar_k3": ("edit", ["-e", "3", "-c", "@kianfar", "-nD"], dict(k=3, scheme=("custom", S + "kianfar"))),
    "ref_01star0_k3_dyn": ("edit", ["-e", "3", "-c", "@01star0"], dict(k=3, scheme=("custom_dynamic", S + "01star0"))),
    "ref_01star0_adapted_k5": ("edit", ["-e", "5", "-c", "@01star0_adapted", "-nD"],
                               dict(k=5, scheme=("custom", S + "01star0_adapted"))),
    "ref_manbest_k4": ("edit", ["-e", "4", "-c", "@manbest", "-nD"], dict(k=4, scheme=("custom", S + "manbest"))),
    "ref_pigeon_k4_dyn": ("edit", ["-e", "4", "-c", "@pigeon"], dict(k=4, scheme=("custom_dynamic", S + "pigeon"))),
    "ref_pigeon_adapted_k6": ("edit", ["-e", "6", "-c", "@pigeon_adapted", "-nD"],
                              dict(k=6, scheme=("custom", S + "pigeon_adapted"))),
    "ref_suffix_filter_k3": ("edit", ["-e", "3", "-c", "@suffix_filter", "-nD"],
                             dict(k=3, scheme=("custom", S + "suffix_filter"))),
    "ref_multiple_opt_k4": ("edit", ["-e", "4", "-d", "@multiple_opt"], dict(k=4, scheme=("multi", S + "multiple_opt"))),
    "ref_multiple_opt_k6": ("edit", ["-e", "6", "-d", "@multiple_opt"], dict(k=6, scheme=("multi", S + "multiple_opt"))),
    # BASELINE.json configs[4]: Hamming distance k = 5, pigeon and multiple_opt's individual scheme 1, dynamic partitioning
    "ref_ham_pigeon_k5": ("ham", ["-m", "hamming", "-e", "5", "-c", "@pigeon"],
                          dict(k=5, metric="hamming", scheme=("custom_dynamic", S + "pigeon"))),
    "ref_ham_mopt_scheme1_k5": ("ham", ["-m", "hamming", "-e", "5", "-c", "@multiple_opt/individual_schemes/scheme1"],
                                dict(k=5, metric="hamming",
                                     scheme=("custom_dynamic", S + "multiple_opt/individual_schemes/scheme1"))),
    "ref_ham_multiple_opt_k4": ("ham", ["-m", "hamming", "-e", "4", "-d", "@multiple_opt"],
                                dict(k=4, metric="hamming", scheme=("multi", S + "multiple_opt"))),
}


def main():
    tmp = tempfile.mkdtemp(prefix="goldens_")
    for f in ("g1.fa", "g1_edit.fq", "g1_ham.fq"):
        with gzip.open(os.path.join(HERE, f + ".gz"), "rb") as i, open(os.path.join(tmp, f), "wb") as o:
            o.write(i.read())
    subprocess.check_call([os.path.join(REF, "columba_build"), "-r", os.path.join(tmp, "g1"), "-f",
                           os.path.join(tmp, "g1.fa")], stdout=subprocess.DEVNULL)
    exp = os.path.join(HERE, "expected")
    meta = json.load(open(os.path.join(HERE, "cases.json")))
    sdir = os.path.join(HERE, "schemes", "ref")
    for name, (rd, args, params) in CASES.items():
        a = [os.path.join(sdir, x[1:]) + "/" if x.startswith("@") else x for x in args]
        sam = os.path.join(tmp, name + ".sam")
        cmd = [os.path.join(REF, "columba_dev"), "-r", os.path.join(tmp, "g1"), "-f",
               os.path.join(tmp, "g1_%s.fq" % rd), "-a", "all", "-t", "1", "-v", "-o", sam] + a
        import time
        t0 = time.time()
        try:
            log = subprocess.run(cmd, capture_output=True, text=True, timeout=120)
        except subprocess.TimeoutExpired:
            print(name, "SKIPPED: the reference needs more than 120 s (it hangs after a scheme-loading error)", flush=True)
            continue
        print("  %.1f s" % (time.time() - t0), flush=True)
        if log.returncode != 0:
            print(name, "SKIPPED: the reference rejects it:", (log.stdout + log.stderr).strip().splitlines()[-1][:300], flush=True)
            continue
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
        print(name, len(rows), ctr, flush=True)
    json.dump(meta, open(os.path.join(HERE, "cases.json"), "w"), indent=1, sort_keys=True)
    shutil.rmtree(tmp)


if __name__ == "__main__":
    main()
