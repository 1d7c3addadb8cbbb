"""Parity at BASELINE.json's full size (configs[1]: 4.6 Mbp, 1 M reads of
150 bp, edit distance k = 4, ALL mode) on the GPU:
  * a 40 000-read sample of the batch against the UNMODIFIED reference binary
    (oracle/_ref/columba_dev; the oracle port on a smaller sample when the
    binary did not travel) -- bit-exact tuples;
  * size-independent properties of the whole result: batch-split invariance
    (checksum of per-read results), counter identities, every CIGAR replayed
    against the text reproduces the reported distance and span.
Nothing here reads /root/reference."""
import hashlib
import os
import subprocess
import sys
import tempfile

import numpy as np
import pytest

import columba_b200 as cb
import golden_util as gu
import oracle_util as ou

sys.path.insert(0, os.path.join(gu.ROOT, "tools"))
pytestmark = pytest.mark.gpu

REF_BP, REF_SEED, N_READS, K = 4_600_000, 4600, 1_000_000, 4
REF_DEV = os.path.join(gu.ROOT, "oracle", "_ref", "columba_dev")


@pytest.fixture(scope="module")
def full(cuda_lib):
    import gen_data
    d = os.path.join(tempfile.gettempdir(), "clb_bench_%d_%d" % (REF_BP, REF_SEED))  # shared with bench.py
    os.makedirs(d, exist_ok=True)
    base = os.path.join(d, "ref")
    names, seqs = gen_data.gen_reference(REF_BP, REF_SEED, 1, "ecoli")
    if not os.path.exists(base + ".meta"):
        gen_data.write_fasta(base + ".fa", names, seqs)
        cb.build_index(base + ".fa", base, 4, lib_path=cuda_lib)
    reads = gen_data.gen_reads_fast(seqs, N_READS, 150, 150, "edit", K, over_e=K + 2)
    ix = cb.Index(base, 4, 10, 4, device=0, lib_path=cuda_lib)
    m = cb.Mapper(ix)
    off = (np.arange(N_READS + 1, dtype=np.int64) * 150).astype(np.uint32)
    res = m.map_batch(reads.reshape(-1), off, K)
    yield {"dir": d, "base": base, "text": seqs[0], "reads": reads, "ix": ix, "m": m, "res": res}
    m.close()
    ix.close()


def per_read_digest(res, lo, hi, shift=0):
    """sha256 over the per-read results of reads [lo, hi) (read ids relative)."""
    h = hashlib.sha256()
    for r in range(lo, hi):
        s, c = int(res.read_start[r]), int(res.read_count[r])
        o = res.occ[s:s + c]
        h.update(np.int64(c).tobytes())
        for f in ("seq_id", "pos", "end", "index_begin", "strand", "dist", "n_cigar", "trimmed"):
            h.update(np.ascontiguousarray(o[f]).tobytes())
        for i in range(s, s + c):
            h.update(res.cigar[i, :int(res.occ[i]["n_cigar"])].tobytes())
    return h.hexdigest()


def test_counter_identities(full):
    res = full["res"]
    c = res.counters
    assert c["NUMBER_OF_READS"] == N_READS
    assert c["MAPPED_READS"] == int((res.read_count > 0).sum())
    assert c["TOTAL_UNIQUE_MATCHES"] == int(res.read_count.sum())
    assert c["IN_TEXT_STARTED"] >= c["ABORTED_IN_TEXT_VERIF"]
    # reads with at most k edits must be found (the scheme is lossless): >= 90 % of the batch
    assert c["MAPPED_READS"] >= 0.93 * N_READS
    occ = res.occ
    assert int(occ["dist"].max()) <= K
    # occurrences of a read are in text order
    for r in np.random.default_rng(0).integers(0, N_READS, 20000):
        s, n = int(res.read_start[r]), int(res.read_count[r])
        b = occ["index_begin"][s:s + n]
        assert (np.diff(b.astype(np.int64)) >= 0).all()


def test_cigars_replay_to_reported_distance(full):
    res, text, reads = full["res"], full["text"], full["reads"]
    comp = np.zeros(256, np.uint8)
    for a, b in zip(b"ACGT", b"TGCA"):
        comp[a] = b
    rng = np.random.default_rng(1)
    n_checked = 0
    for r in rng.integers(0, N_READS, 30000):
        s, n = int(res.read_start[r]), int(res.read_count[r])
        for i in range(s, s + n):
            o = res.occ[i]
            rd = reads[r] if not o["strand"] else comp[reads[r][::-1]]
            ti, qi, dist = int(o["index_begin"]), 0, 0
            for run in res.cigar[i, :int(o["n_cigar"])]:
                ln, op = int(run) >> 2, int(run) & 3
                if op == 0:
                    dist += int((text[ti:ti + ln] != rd[qi:qi + ln]).sum())
                    ti += ln
                    qi += ln
                elif op == 1:
                    qi += ln
                    dist += ln
                else:
                    ti += ln
                    dist += ln
            assert qi == 150
            assert ti - int(o["index_begin"]) == int(o["end"]) - int(o["pos"])
            assert dist == int(o["dist"])
            n_checked += 1
    assert n_checked > 25000


def test_batch_split_invariance(full):
    """Reads are independent: mapping a slice of the batch on its own gives the
    same per-read results as the slice of the full batch."""
    res, reads, m = full["res"], full["reads"], full["m"]
    lo, hi = 300_000, 360_000
    off = (np.arange(hi - lo + 1, dtype=np.int64) * 150).astype(np.uint32)
    sub = m.map_batch(reads[lo:hi].reshape(-1), off, K)
    a = per_read_digest(sub, 0, hi - lo)
    b = per_read_digest(res, lo, hi)
    assert a == b
    # and the whole batch is reproducible run to run
    off = (np.arange(N_READS + 1, dtype=np.int64) * 150).astype(np.uint32)
    again = m.map_batch(reads.reshape(-1), off, K)
    assert again.counters == res.counters
    assert (again.read_count == res.read_count).all()
    assert per_read_digest(again, 0, 50_000) == per_read_digest(res, 0, 50_000)


def test_sample_equals_reference(full):
    import gen_data
    import sam_tuples
    res, reads, ix = full["res"], full["reads"], full["ix"]
    have_ref = os.path.exists(REF_DEV)
    step = 25 if have_ref else 500
    idx = np.arange(0, N_READS, step)
    names = ["r%d" % i for i in range(len(idx))]
    got = []
    for j, r in enumerate(idx):
        s, n = int(res.read_start[r]), int(res.read_count[r])
        for i in range(s, s + n):
            o = res.occ[i]
            got.append((names[j], 16 if o["strand"] else 0, ix.seq_names[int(o["seq_id"])], int(o["pos"]) + 1,
                        int(o["dist"]), cb.cigar_string(res.cigar[i, :int(o["n_cigar"])])))
    got.sort()
    if have_ref:
        fq = os.path.join(full["dir"], "sample_fullsize.fq")
        gen_data.write_fastq_matrix(fq, reads[idx])
        sam = os.path.join(full["dir"], "sample_fullsize.sam")
        p = subprocess.run([REF_DEV, "-r", full["base"], "-f", fq, "-a", "all", "-e", str(K), "-t",
                            str(os.cpu_count() or 1), "-o", sam], capture_output=True, text=True)
        assert p.returncode == 0, p.stdout[-400:] + p.stderr[-400:]
        exp = sam_tuples.sam_tuples(sam)
    else:
        orc = ou.Oracle(full["base"])
        exp = orc.tuples(names, [reads[r].tobytes().decode() for r in idx], K)
    assert len(got) == len(exp)
    assert got == exp


def test_hamming_250bp_k5_sample_equals_reference(full, cuda_lib):
    """configs[4] shape on the 4.6 Mbp reference: 250 bp reads, Hamming distance
    k = 5, custom pigeonhole scheme, dynamic partitioning; 200 000 reads on the
    GPU, every 10th against the unmodified reference binary."""
    import gen_data
    import sam_tuples
    if not os.path.exists(REF_DEV):
        pytest.skip("reference binary did not travel")
    n = 200_000
    reads = gen_data.gen_reads_fast([full["text"]], n, 250, 250, "hamming", 5, over_e=7)
    scheme = os.path.join(gu.GOLD, "schemes", "pigeon5")
    m = cb.Mapper(full["ix"], ("custom", scheme), "hamming", "dynamic")
    off = (np.arange(n + 1, dtype=np.int64) * 250).astype(np.uint32)
    res = m.map_batch(reads.reshape(-1), off, 5)
    assert res.counters["NUMBER_OF_READS"] == n
    assert res.counters["MAPPED_READS"] >= 0.9 * n
    idx = np.arange(0, n, 10)
    names = ["r%d" % i for i in range(len(idx))]
    got = []
    for j, r in enumerate(idx):
        s, c = int(res.read_start[r]), int(res.read_count[r])
        for i in range(s, s + c):
            o = res.occ[i]
            got.append((names[j], 16 if o["strand"] else 0, full["ix"].seq_names[int(o["seq_id"])], int(o["pos"]) + 1,
                        int(o["dist"]), cb.cigar_string(res.cigar[i, :int(o["n_cigar"])])))
    got.sort()
    fq = os.path.join(full["dir"], "sample_ham.fq")
    gen_data.write_fastq_matrix(fq, reads[idx])
    sam = os.path.join(full["dir"], "sample_ham.sam")
    p = subprocess.run([REF_DEV, "-r", full["base"], "-f", fq, "-a", "all", "-m", "hamming", "-e", "5", "-c",
                        scheme + "/", "-nD", "-t", str(os.cpu_count() or 1), "-o", sam], capture_output=True, text=True)
    assert p.returncode == 0, p.stdout[-400:] + p.stderr[-400:]
    exp = sam_tuples.sam_tuples(sam)
    m.close()
    assert len(got) == len(exp)
    assert got == exp


def test_best_mode_sample_equals_reference(full):
    """BEST mode (minimal identity 95 %: strata up to k = 7) on 20 000 reads of the
    bench batch against `columba -a best` of the unmodified reference."""
    import gen_data
    import sam_tuples
    from columba_b200 import host
    if not os.path.exists(REF_DEV):
        pytest.skip("reference binary did not travel")
    n = 20_000
    reads = full["reads"][:n]
    seqs = [reads[i].tobytes().decode() for i in range(n)]
    names = ["r%d" % i for i in range(n)]
    m = cb.Mapper(full["ix"])
    occs, found = host.map_best(m, seqs, 95, 0)
    m.close()
    got = host.best_records(occs, found, names, full["ix"].seq_names)
    fq = os.path.join(full["dir"], "sample_best.fq")
    gen_data.write_fastq_matrix(fq, reads)
    sam = os.path.join(full["dir"], "sample_best.sam")
    p = subprocess.run([REF_DEV, "-r", full["base"], "-f", fq, "-a", "best", "-I", "95", "-t", str(os.cpu_count() or 1),
                        "-o", sam], capture_output=True, text=True)
    assert p.returncode == 0, p.stdout[-400:] + p.stderr[-400:]
    # every record of every read: all co-optimal alignments and which of them is the primary one
    exp = sam_tuples.sam_tuples(sam, flag_mask=16 | 256)
    assert len(got) == len(exp)
    assert got == exp
