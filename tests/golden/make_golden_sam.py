#!/usr/bin/env python
"""Golden SAM records (single-end, ALL mode) from the UNMODIFIED reference
binary oracle/_ref/columba_dev (DEVELOPER_MODE = stable sorts, built from
/root/reference by oracle/Makefile), on the golden index and reads:
  columba_dev -r g1 -f g1_edit.fq -a all -t 1 -e 4 -o out.sam
Header lines are dropped; the body is stored as it was written.
Run in the build container (the GPU box has no /root/reference)."""
import gzip
import os
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import golden_util as gu  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref")


def main():
    base = os.path.join(gu.WORK, "g1")  # written by the golden_index fixture (byte-identical to columba_build's)
    tmp = tempfile.mkdtemp(prefix="golden_sam_")
    fq = os.path.join(tmp, "g1_edit.fq")
    with gzip.open(os.path.join(HERE, "g1_edit.fq.gz"), "rb") as i, open(fq, "wb") as o:
        o.write(i.read())
    os.makedirs(os.path.join(HERE, "expected_sam"), exist_ok=True)
    for name, args in {"edit_k4_columba": ["-e", "4"], "edit_k2_columba": ["-e", "2"]}.items():
        sam = os.path.join(tmp, name + ".sam")
        cmd = [os.path.join(REF, "columba_dev"), "-r", base, "-f", fq, "-a", "all", "-t", "1", "-o", sam] + args
        log = subprocess.run(cmd, capture_output=True, text=True)
        if log.returncode != 0:
            raise SystemExit(log.stdout + log.stderr)
        body = [line for line in open(sam) if not line.startswith("@")]
        with gzip.GzipFile(os.path.join(HERE, "expected_sam", name + ".sam.gz"), "wb", mtime=0) as o:
            o.write("".join(body).encode())
        print(name, len(body), "records")


if __name__ == "__main__":
    main()
