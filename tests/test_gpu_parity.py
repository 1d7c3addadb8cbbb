"""Parity tests proper: the CUDA path through the C ABI against the golden
outputs of the unmodified reference and against the oracle."""
import os

import numpy as np
import pytest

import columba_b200 as cb
import golden_util as gu
import oracle_util as ou

pytestmark = pytest.mark.gpu
CASES = gu.cases()


@pytest.fixture(scope="module")
def gpu_indexes(golden_index, cuda_lib):
    cache = {}

    def get(kmer, in_text, p=None):
        base = gu.index_of(p or {}, cuda_lib)
        if (base, kmer, in_text) not in cache:
            cache[(base, kmer, in_text)] = cb.Index(base, 4, kmer, in_text, device=0, lib_path=cuda_lib)
        return cache[(base, kmer, in_text)]
    yield get
    for ix in cache.values():
        ix.close()


@pytest.mark.parametrize("case", sorted(CASES))
def test_gpu_matches_reference_golden(case, gpu_indexes):
    p = CASES[case]
    ix = gpu_indexes(p.get("kmer", 10), p.get("in_text", 4), p)
    m = cb.Mapper(ix, gu.scheme_arg(p), p.get("metric", "edit"), p.get("partition", "dynamic"))
    names, seqs = gu.read_fastq_gz(gu.reads_file(p))
    reads, off = cb.Mapper.pack(seqs)
    res = m.map_batch(reads, off, p["k"])
    exp, ctr = gu.expected(case)
    assert res.tuples(names, ix.seq_names) == exp
    for k, v in gu.CTR_MAP.items():
        assert res.counters[v] == ctr[k], k
    assert res.gpu_launches > 0
    m.close()


@pytest.mark.parametrize("case", [c for c in sorted(CASES) if c.startswith("g2_")])
@pytest.mark.parametrize("group", ["32", "16"])
def test_gpu_heavy_search_launch_and_small_groups(case, group, gpu_indexes, monkeypatch):
    """Small search scratch: a third of the tasks of the repeat-rich set go to the heavy launch (one CTA per task);
    the first launch also with half-warp groups."""
    monkeypatch.setenv("COLUMBA_B200_SEARCH_SLAB", "4096")
    monkeypatch.setenv("COLUMBA_B200_SEARCH_GROUP", group)
    if group == "16":  # tasks that start from a wide range go to the heavy launch without a first attempt
        monkeypatch.setenv("COLUMBA_B200_HEAVY_WIDTH", "64")
    p = CASES[case]
    ix = gpu_indexes(p.get("kmer", 10), p.get("in_text", 4), p)
    m = cb.Mapper(ix, gu.scheme_arg(p), p.get("metric", "edit"), p.get("partition", "dynamic"))
    names, seqs = gu.read_fastq_gz(gu.reads_file(p))
    reads, off = cb.Mapper.pack(seqs)
    res = m.map_batch(reads, off, p["k"])
    exp, ctr = gu.expected(case)
    assert res.stats["HEAVY_TASKS"] > 0
    assert res.tuples(names, ix.seq_names) == exp
    for k, v in gu.CTR_MAP.items():
        assert res.counters[v] == ctr[k], k
    m.close()


def test_gpu_rank_extend_locate_match_oracle(gpu_indexes, golden_index):
    ix = gpu_indexes(10, 4)
    m = cb.Mapper(ix)
    orc = ou.Oracle(golden_index)
    n = ix.text_length
    rng = np.random.default_rng(2)
    pos = np.concatenate([[0, 1, 63, 64, 65, n - 1, n], rng.integers(0, n + 1, 2000)]).astype(np.uint32)
    for rev in (False, True):
        got = m.occ4(pos, rev)
        for i, p in enumerate(pos):
            assert list(got[i]) == orc.occ4(int(p), rev)
    parents = [[0, n, 0, n]]
    for _ in range(500):
        b = int(rng.integers(0, n))
        w = int(rng.integers(1, 3000))
        parents.append([b, min(n, b + w), b, min(n, b + w)])
    for mode in (0, 1, 2):
        got = m.extend(parents, mode)
        for i, par in enumerate(parents):
            exp = orc.extend(par, mode)
            for c in range(4):
                g = list(got[i, c])
                if exp[c][1] <= exp[c][0]:
                    assert g[1] <= g[0]
                else:
                    assert g[:2] == exp[c][:2]
                    if mode != 2:
                        assert g == exp[c]
    rows = rng.integers(0, n, 2000).astype(np.uint32)
    got = m.find_sa(rows)
    for i, r in enumerate(rows):
        assert int(got[i]) == orc.find_sa(int(r))
    m.close()
    orc.close()
