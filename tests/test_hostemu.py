"""CPU-side unit tests (no GPU): the kernels' task functions compiled for the
host (tests/hostemu) against the reference's golden outputs, the host-side index
writer and scheme parser, and the exported ABI of the CUDA library."""
import hashlib
import json
import os
import re

import numpy as np
import pytest

import columba_b200 as cb
import golden_util as gu
import oracle_util as ou

CASES = gu.cases()


def test_index_writer_is_byte_identical_to_reference(golden_index):
    sums = json.load(open(os.path.join(gu.GOLD, "ref_index", "sha256.json")))
    for f, h in sums.items():
        assert hashlib.sha256(open(os.path.join(gu.WORK, f), "rb").read()).hexdigest() == h, f


@pytest.fixture(scope="module")
def emu_indexes(golden_index, hostemu_lib):
    cache = {}

    def get(kmer, in_text, p=None):
        base = gu.index_of(p or {}, hostemu_lib)
        if (base, kmer, in_text) not in cache:
            cache[(base, kmer, in_text)] = cb.Index(base, 4, kmer, in_text, lib_path=hostemu_lib)
        return cache[(base, kmer, in_text)]
    yield get
    for ix in cache.values():
        ix.close()


@pytest.mark.parametrize("case", sorted(CASES))
def test_kernel_logic_matches_reference_golden(case, emu_indexes):
    p = CASES[case]
    ix = emu_indexes(p.get("kmer", 10), p.get("in_text", 4), p)
    m = cb.Mapper(ix, gu.scheme_arg(p), p.get("metric", "edit"), p.get("partition", "dynamic"))
    names, seqs = gu.read_fastq_gz(gu.reads_file(p))
    reads, off = cb.Mapper.pack(seqs)
    res = m.map_batch(reads, off, p["k"])
    exp, ctr = gu.expected(case)
    assert res.tuples(names, ix.seq_names) == exp
    for k, v in gu.CTR_MAP.items():
        assert res.counters[v] == ctr[k], k
    m.close()


@pytest.mark.parametrize("case", [c for c in sorted(CASES) if c.startswith("g2_")])
@pytest.mark.parametrize("group", ["32", "8"])
def test_heavy_search_launch_and_small_groups(case, group, emu_indexes, monkeypatch):
    """Small search scratch: the tasks that outgrow it are run again by the heavy launch (a group of 128 lanes,
    children counted then built) -- same occurrences and counters; the first launch with quarter-warp groups."""
    monkeypatch.setenv("COLUMBA_B200_SEARCH_SLAB", "4096")
    monkeypatch.setenv("COLUMBA_B200_SEARCH_GROUP", group)
    if group == "8":  # tasks that start from a wide range go to the heavy launch without a first attempt
        monkeypatch.setenv("COLUMBA_B200_HEAVY_WIDTH", "64")
    p = CASES[case]
    ix = emu_indexes(p.get("kmer", 10), p.get("in_text", 4), p)
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


def test_rank_extend_locate_match_oracle(emu_indexes, golden_index):
    ix = emu_indexes(10, 4)
    m = cb.Mapper(ix)
    orc = ou.Oracle(golden_index)
    n = ix.text_length
    rng = np.random.default_rng(1)
    pos = np.concatenate([[0, 1, 63, 64, 65, n - 1, n], rng.integers(0, n + 1, 300)]).astype(np.uint32)
    for rev in (False, True):
        got = m.occ4(pos, rev)
        for i, p in enumerate(pos):
            assert list(got[i]) == orc.occ4(int(p), rev)
    parents = []
    for _ in range(200):
        b = int(rng.integers(0, n))
        w = int(rng.integers(1, 2000))
        parents.append([b, min(n, b + w), b, min(n, b + w)])
    parents.append([0, n, 0, n])
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
    rows = rng.integers(0, n, 300).astype(np.uint32)
    got = m.find_sa(rows)
    for i, r in enumerate(rows):
        assert int(got[i]) == orc.find_sa(int(r))
    m.close()
    orc.close()


def test_error_conventions(emu_indexes, hostemu_lib, tmp_path):
    with pytest.raises(cb.ClbError) as e:
        cb.Index(str(tmp_path / "nope"), lib_path=hostemu_lib)
    assert "Cannot open file" in str(e.value)
    ix = emu_indexes(10, 4)
    with pytest.raises(cb.ClbError) as e:
        cb.Mapper(ix, ("custom", str(tmp_path)))
    assert "name.txt" in str(e.value)
    bad = tmp_path / "bad"
    (bad / "1").mkdir(parents=True)
    (bad / "name.txt").write_text("bad\n")
    (bad / "1" / "searches.txt").write_text("{0,2,1} {0,0,0} {0,1,1}\n")
    with pytest.raises(cb.ClbError) as e:
        cb.Mapper(ix, ("custom", str(bad)))
    assert "Connectivity" in str(e.value)
    m = cb.Mapper(ix, ("named", "kuch1"))
    assert m.supports(4) and not m.supports(5)
    reads, off = cb.Mapper.pack(["ACGT" * 30])
    with pytest.raises(cb.ClbError):
        m.map_batch(reads, off, 5)
    # reads not longer than the number of parts (the reference falls back to naive backtracking for them) and
    # empty reads are flagged per read; the other reads of the batch keep their results
    names, seqs = gu.read_fastq_gz(os.path.join(gu.GOLD, "g1_edit.fq.gz"))
    sub = seqs[:40]
    mixed = sub[:10] + ["ACG"] + sub[10:20] + [""] + sub[20:] + ["ACGTA"]
    reads, off = cb.Mapper.pack(mixed)
    res = m.map_batch(reads, off, 4)
    assert res.n_short_reads == 3 and list(np.flatnonzero(res.read_status)) == [10, 21, 42]
    assert res.read_count[10] == 0 and res.read_count[21] == 0 and res.read_count[42] == 0
    ref = m.map_batch(*cb.Mapper.pack(sub), 4)
    keep = [i for i in range(len(mixed)) if i not in (10, 21, 42)]
    assert [int(res.read_count[i]) for i in keep] == [int(c) for c in ref.read_count]
    a = np.concatenate([res.occ[int(res.read_start[i]):int(res.read_start[i]) + int(res.read_count[i])] for i in keep])
    for f in ("seq_id", "pos", "end", "strand", "dist"):
        assert (a[f] == ref.occ[f]).all()
    m.close()


def test_cuda_library_exports_the_whole_abi(cuda_lib):
    hdr = open(os.path.join(gu.ROOT, "include", "columba_b200.h")).read()
    declared = sorted(set(re.findall(r"\b(clb_[a-z0-9_]+)\s*\(", hdr)))
    assert set(declared) == set(cb.api.EXPORTS)
    lib = cb.load_library(cuda_lib)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.clb_abi_version() == 4


def test_product_has_no_cpu_fallback(cuda_lib, golden_index):
    lib = cb.load_library(cuda_lib)
    if lib.clb_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(cb.ClbError) as e:
        cb.Index(golden_index, lib_path=cuda_lib)
    assert e.value.code == -3 and "no CPU fallback" in str(e.value)


def test_list_overflow_reruns_give_the_same_result(emu_indexes, monkeypatch):
    """Lists that start far too small (test hook COLUMBA_B200_MIN_LIST_CAP): the batch is re-run with lists sized
    from the counts of the overflowing run -- the search stage keeps counting after a candidate / index-occurrence
    overflow -- and returns what a context with ample lists returns."""
    ix = emu_indexes(10, 4)
    names, seqs = gu.read_fastq_gz(os.path.join(gu.GOLD, "g1_edit.fq.gz"))
    reads, off = cb.Mapper.pack(seqs[:200])
    for metric, k in (("edit", 4), ("hamming", 3)):
        m = cb.Mapper(ix, ("named", "columba"), metric)
        ref = m.map_batch(reads, off, k)
        m.close()
        monkeypatch.setenv("COLUMBA_B200_MIN_LIST_CAP", "16")
        m = cb.Mapper(ix, ("named", "columba"), metric)
        monkeypatch.delenv("COLUMBA_B200_MIN_LIST_CAP")
        got = m.map_batch(reads, off, k)
        m.close()
        assert got.counters == ref.counters
        assert (got.read_count == ref.read_count).all()
        for f in ("seq_id", "pos", "end", "strand", "dist", "n_cigar"):
            assert (got.occ[f] == ref.occ[f]).all(), f
