"""SAM records of single-end ALL mode (flags, MAPQ, primary / secondary,
unmapped records; columba_b200/host.py: sam_single_end) against the SAM body
written by the unmodified reference (tests/golden/expected_sam, made by
tests/golden/make_golden_sam.py)."""
import gzip
import os

import pytest

import columba_b200 as cb
import golden_util as gu
from columba_b200 import host

CASES = {"edit_k4_columba": 4, "edit_k2_columba": 2}


def expected_sam(case):
    with gzip.open(os.path.join(gu.GOLD, "expected_sam", case + ".sam.gz"), "rt") as f:
        return f.readlines()


def read_fastq_with_quals(path):
    names, seqs, quals = [], [], []
    with gzip.open(path, "rt") as f:
        while True:
            h = f.readline()
            if not h:
                break
            names.append(h[1:].split()[0])
            seqs.append(f.readline().strip())
            f.readline()
            quals.append(f.readline().strip())
    return names, seqs, quals


def run_case(case, ix):
    names, seqs, quals = read_fastq_with_quals(os.path.join(gu.GOLD, "g1_edit.fq.gz"))
    m = cb.Mapper(ix)
    reads, off = cb.Mapper.pack([s.upper() for s in seqs])
    res = m.map_batch(reads, off, CASES[case], 0)
    got = host.sam_single_end(res, names, seqs, quals, ix.seq_names)
    m.close()
    assert got == expected_sam(case)


def test_map_q_formula():
    # src/indexhelpers.h:378-388
    assert host.map_q(2, 1, 2) == 60 and host.map_q(3, 1, 2) == 0
    assert host.map_q(0, 2, 0) == 3 and host.map_q(0, 3, 0) == 2 and host.map_q(0, 10, 0) == 0


@pytest.mark.parametrize("case", sorted(CASES))
def test_sam_records_kernel_logic(case, golden_index, hostemu_lib):
    ix = cb.Index(golden_index, 4, 10, 4, lib_path=hostemu_lib)
    run_case(case, ix)
    ix.close()


@pytest.mark.gpu
@pytest.mark.parametrize("case", sorted(CASES))
def test_sam_records_gpu(case, golden_index, cuda_lib):
    ix = cb.Index(golden_index, 4, 10, 4, device=0, lib_path=cuda_lib)
    run_case(case, ix)
    ix.close()
