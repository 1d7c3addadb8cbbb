#!/usr/bin/env python
"""BEST-mode golden fixtures, produced by RUNNING THE UNMODIFIED REFERENCE
(oracle/_ref/columba_dev -a best) on the committed golden inputs.  Build
container only.  Writes expected_best/<case>.tsv.gz and best_cases.json."""
import gzip
import json
import os
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import sam_tuples  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref")
CASES = {"best_I95": dict(identity=95, x=0), "best_I97": dict(identity=97, x=0), "best_I99": dict(identity=99, x=0),
         "best_I90": dict(identity=90, x=0), "best_I95_x2": dict(identity=95, x=2),
         # repeat-rich reference: many co-optimal alignments per read, on both strands
         "g2_best_I95": dict(identity=95, x=0, fasta="g2", reads="g2"),
         "g2_best_I97_x1": dict(identity=97, x=1, fasta="g2", reads="g2"),
         # most reads have exact occurrences (tandem repeats among them): best stratum 0
         "g3_best_I95": dict(identity=95, x=0, fasta="g2", reads="g3"),
         "g3_best_I95_x1": dict(identity=95, x=1, fasta="g2", reads="g3")}


def main():
    tmp = tempfile.mkdtemp(prefix="goldenbest_")
    for f in ("g1.fa", "g1_edit.fq", "g2.fa", "g2.fq", "g3.fq"):
        with gzip.open(os.path.join(HERE, f + ".gz"), "rb") as i, open(os.path.join(tmp, f), "wb") as o:
            o.write(i.read())
    for g in ("g1", "g2"):
        subprocess.check_call([os.path.join(REF, "columba_build"), "-r", os.path.join(tmp, g), "-f",
                               os.path.join(tmp, g + ".fa")], stdout=subprocess.DEVNULL)
    out = os.path.join(HERE, "expected_best")
    os.makedirs(out, exist_ok=True)
    for name, p in CASES.items():
        sam = os.path.join(tmp, name + ".sam")
        fq = os.path.join(tmp, (p["reads"] if "reads" in p else "g1_edit") + ".fq")
        cmd = [os.path.join(REF, "columba_dev"), "-r", os.path.join(tmp, p.get("fasta", "g1")), "-f", fq,
               "-a", "best", "-I", str(p["identity"]), "-x", str(p["x"]), "-t", "1", "-o", sam]
        log = subprocess.run(cmd, capture_output=True, text=True)
        if log.returncode != 0:
            raise SystemExit(log.stdout + log.stderr)
        # every record: strand and the secondary flag (the first record of a read is its primary alignment)
        rows = sam_tuples.sam_tuples(sam, flag_mask=16 | 256)
        with gzip.GzipFile(os.path.join(out, name + ".tsv.gz"), "wb", mtime=0) as o:
            for r in rows:
                o.write(("\t".join(map(str, r)) + "\n").encode())
        print(name, len(rows))
    json.dump(CASES, open(os.path.join(HERE, "best_cases.json"), "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
