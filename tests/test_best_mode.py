"""BEST mode (host strata loop over the device pass, columba_b200/host.py)
against golden outputs of the unmodified reference run with -a best."""
import gzip
import json
import os

import pytest

import columba_b200 as cb
import golden_util as gu
from columba_b200 import host

CASES = json.load(open(os.path.join(gu.GOLD, "best_cases.json")))


def expected_best(case):
    rows = []
    with gzip.open(os.path.join(gu.GOLD, "expected_best", case + ".tsv.gz"), "rt") as f:
        for line in f:
            a = line.rstrip("\n").split("\t")
            rows.append((a[0], int(a[1]), a[2], int(a[3]), int(a[4]), a[5]))
    return rows


def test_strata_schedule_follows_reference():
    # src/searchstrategy.cpp:672-708: 0; 1, 3, 5, 9, 13 (step 2 below 5, then 4), clipped to the cut-off
    assert host.strata(7) == [0, 1, 3, 5, 7]
    assert host.strata(13) == [0, 1, 3, 5, 9, 13]
    assert host.strata(4) == [0, 1, 3, 4]
    assert host.strata(0) == [0]
    assert host.strata(7, x=1) == [1, 4, 7]
    # the per-read state machine walks the same schedule while nothing is found
    s = host._BestState(13, 0)
    s.after_exact()
    seen = []
    while not s.finished:
        seen.append(s.k)
        s.step(mapped=True)
    assert [0] + seen == host.strata(13)
    assert host.max_ed(95, 150) == 7 and host.max_ed(97, 150) == 4 and host.max_ed(100, 150) == 0


def run_case(case, lib, device):
    """Every record of `columba -a best` (all co-optimal alignments, which of them is primary, strands,
    positions, distances, CIGARs) for every read."""
    p = CASES[case]
    base = gu.index_of(p, lib)
    ix = cb.Index(base, 4, 10, 4, device=device, lib_path=lib)
    m = cb.Mapper(ix)
    names, seqs = gu.read_fastq_gz(os.path.join(gu.GOLD, (p["reads"] if "reads" in p else "g1_edit") + ".fq.gz"))
    occs, found = host.map_best(m, seqs, p["identity"], p["x"])
    got = host.best_records(occs, found, names, ix.seq_names)
    m.close()
    ix.close()
    assert got == expected_best(case)


@pytest.mark.parametrize("case", sorted(CASES))
def test_best_mode_kernel_logic(case, golden_index, hostemu_lib):
    run_case(case, hostemu_lib, 0)


@pytest.mark.gpu
@pytest.mark.parametrize("case", sorted(CASES))
def test_best_mode_gpu(case, golden_index, cuda_lib):
    run_case(case, cuda_lib, 0)


def test_pair_occurrences_window_join():
    mk = lambda b, n=150: host.Occ((0, 0, b, b + n, b, 0, "150M"))
    up = [mk(1000), mk(5000)]
    down = [mk(1200), mk(1600), mk(5300), mk(900)]
    pairs = host.pair_occurrences(up, down, 200, 700)
    assert sorted(pairs) == [(0, 0, 350), (1, 2, 450)]
