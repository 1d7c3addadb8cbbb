"""The oracle (CPU restatement, oracle/) against the golden outputs produced by
the UNMODIFIED reference binary (tests/golden/make_golden.py).  CPU only."""
import os

import pytest

import golden_util as gu
import oracle_util as ou

CASES = gu.cases()
ORACLE_CASES = [c for c in sorted(CASES) if c in (
    "edit_k2_columba", "edit_k4_columba", "edit_k4_kuch1", "edit_k3_kuch2", "edit_k4_multi", "edit_k3_static",
    "edit_k4_i0", "edit_k4_pigeon_custom", "ham_k4_columba", "ham_k5_pigeon_custom")]


@pytest.mark.parametrize("case", ORACLE_CASES)
def test_oracle_matches_reference_golden(case, golden_index):
    p = CASES[case]
    names, seqs = gu.read_fastq_gz(os.path.join(gu.GOLD, "g1_%s.fq.gz" % p["reads"]))
    orc = ou.Oracle(golden_index, 4, p.get("kmer", 10), p.get("in_text", 4), gu.scheme_arg(p),
                    p.get("partition", "dynamic"), p.get("metric", "edit"))
    got = orc.tuples(names, seqs, p["k"])
    exp, ctr = gu.expected(case)
    assert got == exp
    c = list(orc.counters)
    assert c[0] == ctr["nodes"]
    assert c[3] == ctr["intext"]
    assert c[4] == ctr["aborted"]
    assert c[6] == ctr["immediate"]
    assert c[7] == ctr["searches"]
    assert c[1] == ctr["reported"]
    orc.close()
