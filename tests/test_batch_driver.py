"""Batch driver (columba_b200/csrc/host/clb_map_main.cpp: FASTQ in, SAM out, worker threads with one context each):
its SAM body must equal the reference's byte for byte, whatever the batch size and the number of threads."""
import os
import subprocess

import pytest

import columba_b200 as cb
import golden_util as gu
from test_sam_records import CASES, expected_sam

SRC = os.path.join(gu.ROOT, "columba_b200", "csrc", "host", "clb_map_main.cpp")


def fastq():
    import gzip
    fq = os.path.join(gu.WORK, "g1_edit.fq")
    if not os.path.exists(fq):
        with gzip.open(os.path.join(gu.GOLD, "g1_edit.fq.gz"), "rb") as i, open(fq, "wb") as o:
            o.write(i.read())
    return fq


def run(exe, base, k, threads, batch, tmp_path, fmt=3, fq=None):
    sam = str(tmp_path / "out.sam")
    p = subprocess.run([exe, "-r", base, "-f", fq or fastq(), "-o", sam, "-e", str(k), "-t", str(threads), "-b", str(batch),
                        "-F", str(fmt), "-D", cb.SCHEMES_DIR], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr[-500:]
    assert "mapping threads finished mapping in" in p.stderr
    return [l for l in open(sam) if not l.startswith("@")]


@pytest.mark.parametrize("case", sorted(CASES))
def test_batch_driver_kernel_logic(case, golden_index, hostemu_lib, tmp_path):
    exe = os.path.join(gu.WORK, "clb_map_emu")
    libdir = os.path.dirname(hostemu_lib)
    if not os.path.exists(exe) or os.path.getmtime(exe) < max(os.path.getmtime(SRC), os.path.getmtime(hostemu_lib)):
        subprocess.check_call(["g++", "-std=c++14", "-O2", SRC, "-L", libdir, "-lclb_hostemu", "-Wl,-rpath," + libdir,
                               "-lpthread", "-o", exe])
    for threads, batch, fmt in ((1, 1 << 20, 5), (2, 64, 1), (3, 7, 3), (2, 200, 2)):
        assert run(exe, golden_index, CASES[case], threads, batch, tmp_path, fmt) == expected_sam(case)


def test_batch_driver_reader_variants(golden_index, hostemu_lib, tmp_path):
    """The reader cuts records at line ends and the workers take names, reads and qualities out of the raw text:
    CRLF line ends, lower-case bases, a last line without a line end and a trailing incomplete record must not
    change the records (batch sizes that put the cuts at different places of the 16 MB read buffer included)."""
    exe = os.path.join(gu.WORK, "clb_map_emu")
    libdir = os.path.dirname(hostemu_lib)
    if not os.path.exists(exe) or os.path.getmtime(exe) < max(os.path.getmtime(SRC), os.path.getmtime(hostemu_lib)):
        subprocess.check_call(["g++", "-std=c++14", "-O2", SRC, "-L", libdir, "-lclb_hostemu", "-Wl,-rpath," + libdir,
                               "-lpthread", "-o", exe])
    case = sorted(CASES)[0]
    want = expected_sam(case)
    lines = open(fastq(), "rb").read().split(b"\n")
    if lines[-1] == b"":
        lines.pop()
    assert len(lines) % 4 == 0
    variants = {
        "crlf": b"\r\n".join(lines) + b"\r\n",
        "no_final_newline": b"\n".join(lines),
        "lower_case": b"\n".join(l.lower() if i % 4 == 1 else l for i, l in enumerate(lines)) + b"\n",
        "partial_record": b"\n".join(lines) + b"\n@cut_off\nACGT\n",
    }
    for name, data in variants.items():
        fq = str(tmp_path / (name + ".fq"))
        open(fq, "wb").write(data)
        for threads, batch in ((1, 1 << 20), (2, 33)):
            assert run(exe, golden_index, CASES[case], threads, batch, tmp_path, 2, fq) == want, name


@pytest.mark.gpu
@pytest.mark.parametrize("case", sorted(CASES))
def test_batch_driver_gpu(case, golden_index, cuda_lib, tmp_path):
    import __graft_entry__ as ge
    exe = ge.build_mapper()
    for threads, batch in ((1, 1 << 20), (2, 64)):
        assert run(exe, golden_index, CASES[case], threads, batch, tmp_path) == expected_sam(case)
