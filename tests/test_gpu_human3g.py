"""Parity on the metric's own workload (BASELINE.json configs[2]): the synthetic
3.1 Gbp GRCh38-sized reference with the "human" repeat profile, index resident
in HBM (7.8 GB), 150 bp reads, edit distance k = 4, ALL mode.

  * 20 000 reads drawn like the bench batches against the UNMODIFIED
    reference binary (oracle/_ref/columba_dev) on the same index files: every
    (read, strand, sequence, position, distance, CIGAR) tuple and the per-read
    hit counts; the sample holds reads with more than 64, more than 8 192 and
    more than 100 000 occurrences (microsatellites, tandem arrays, SINE-like
    families), i.e. the long sorted runs, wide SA ranges and long filter chains
    of stages 4-6;
  * BEST mode (-a best -I 95) on 5 000 of them: every record incl. which one is
    primary;
  * batch-split invariance of the device path on 100 000 reads.
The index is built by the standalone builder (columba_b200/bin/clb_gpu_build,
about 15 s) into the directory bench.py uses, so one box builds it once.
Nothing here reads /root/reference."""
import os
import subprocess
import sys
import tempfile

import numpy as np
import pytest

import columba_b200 as cb
import golden_util as gu

sys.path.insert(0, os.path.join(gu.ROOT, "tools"))
pytestmark = pytest.mark.gpu

REF_BP, REF_SEED, PROFILE, K = 3_100_000_000, 3100, "human", 4
REF_DEV = os.path.join(gu.ROOT, "oracle", "_ref", "columba_dev")
BUILDER = os.path.join(gu.ROOT, "columba_b200", "bin", "clb_gpu_build")
N_SAMPLE = 20_000


@pytest.fixture(scope="module")
def human(cuda_lib):
    import gen_data
    import torch
    if torch.cuda.get_device_properties(0).total_memory < 60 * 2**30:
        pytest.skip("needs a GPU with the memory of a B200 for the 3.1 Gbp suffix sort")
    d = os.path.join(tempfile.gettempdir(), "clb_bench_%s_%d_%d" % (PROFILE, REF_BP, REF_SEED))  # shared with bench.py
    os.makedirs(d, exist_ok=True)
    base = os.path.join(d, "ref")
    if not os.path.exists(base + ".done"):
        subprocess.check_call([BUILDER, "--synthetic", PROFILE, "--bp", str(REF_BP), "--seed", str(REF_SEED), "--out",
                               base, "--device", "0", "--quiet"])
        open(base + ".done", "w").write("test\n")
    n = int(np.fromfile(base + ".txt.bin", dtype=np.uint32, count=1)[0])
    cat = np.memmap(base + ".txt.bin", dtype=np.uint8, mode="r", offset=4, shape=(n,))
    starts = np.fromfile(base + ".pos", dtype=np.uint32).astype(np.int64)
    reads = gen_data.gen_reads_cat(cat, starts, 100_000, 150, 150, "edit", K, over_e=K + 2)  # drawn like the bench batches
    ix = cb.Index(base, 4, 10, 4, device=0, lib_path=cuda_lib)
    yield {"dir": d, "base": base, "reads": reads, "ix": ix}
    ix.close()


def result_tuples(res, names, ix, lo=0, hi=None):
    out = []
    for r in range(lo, res.n_reads if hi is None else hi):
        s, c = int(res.read_start[r]), int(res.read_count[r])
        for i in range(s, s + c):
            o = res.occ[i]
            out.append((names[r], 16 if o["strand"] else 0, ix.seq_names[int(o["seq_id"])], int(o["pos"]) + 1,
                        int(o["dist"]), cb.cigar_string(res.cigar[i, :int(o["n_cigar"])])))
    out.sort()
    return out


def test_all_mode_sample_equals_reference(human):
    import gen_data
    import sam_tuples
    if not os.path.exists(REF_DEV):
        pytest.skip("reference binary did not travel")
    reads, ix = human["reads"][:N_SAMPLE], human["ix"]
    names = ["r%d" % i for i in range(N_SAMPLE)]
    m = cb.Mapper(ix)
    off = (np.arange(N_SAMPLE + 1, dtype=np.int64) * 150).astype(np.uint32)
    res = m.map_batch(reads.reshape(-1), off, K)
    m.close()
    counts = res.read_count
    # the sample must exercise the heavy paths of the filter stages
    assert (counts > 64).sum() > 50 and (counts > 8192).sum() >= 1, (int((counts > 64).sum()), int(counts.max()))
    got = result_tuples(res, names, ix)
    fq = os.path.join(human["dir"], "sample_h3g.fq")
    gen_data.write_fastq_matrix(fq, reads)
    sam = os.path.join(human["dir"], "sample_h3g.sam")
    p = subprocess.run([REF_DEV, "-r", human["base"], "-f", fq, "-a", "all", "-e", str(K), "-t",
                        str(os.cpu_count() or 1), "-o", sam], capture_output=True, text=True)
    assert p.returncode == 0, p.stdout[-400:] + p.stderr[-400:]
    exp = sam_tuples.sam_tuples(sam)
    os.unlink(sam)
    assert len(got) == len(exp), (len(got), len(exp))
    assert got == exp
    print("3.1 Gbp sample: %d reads, %d occurrences, max %d per read, %d reads > 64, %d reads > 8192" %
          (N_SAMPLE, len(got), int(counts.max()), int((counts > 64).sum()), int((counts > 8192).sum())))


def test_best_mode_sample_equals_reference(human):
    import gen_data
    import sam_tuples
    from columba_b200 import host
    if not os.path.exists(REF_DEV):
        pytest.skip("reference binary did not travel")
    n = 5_000
    reads, ix = human["reads"][:n], human["ix"]
    seqs = [reads[i].tobytes().decode() for i in range(n)]
    names = ["r%d" % i for i in range(n)]
    m = cb.Mapper(ix)
    occs, found = host.map_best(m, seqs, 95, 0)
    m.close()
    got = host.best_records(occs, found, names, ix.seq_names)
    fq = os.path.join(human["dir"], "sample_h3g_best.fq")
    gen_data.write_fastq_matrix(fq, reads)
    sam = os.path.join(human["dir"], "sample_h3g_best.sam")
    p = subprocess.run([REF_DEV, "-r", human["base"], "-f", fq, "-a", "best", "-I", "95", "-t",
                        str(os.cpu_count() or 1), "-o", sam], capture_output=True, text=True)
    assert p.returncode == 0, p.stdout[-400:] + p.stderr[-400:]
    exp = sam_tuples.sam_tuples(sam, flag_mask=16 | 256)
    os.unlink(sam)
    assert len(got) == len(exp), (len(got), len(exp))
    assert got == exp


def test_paired_end_sample_equals_reference(human, cuda_lib):
    """BASELINE.json configs[3] shape on the 3.1 Gbp index: FR pairs of 2 x 150 bp, fragments ~N(400, 50), BEST mode
    with minimal identity 95 (and ALL mode k = 2), through the C++ paired-end layer on the CUDA library against
    `columba_dev -f/-F -nI`: every SAM record (flags, MAPQ, mate fields, TLEN, NM) of every pair."""
    import time

    import gen_data
    from test_paired_end import SRC as PE_SRC
    if not os.path.exists(REF_DEV):
        pytest.skip("reference binary did not travel")
    n = 3_000
    n_text = int(np.fromfile(human["base"] + ".txt.bin", dtype=np.uint32, count=1)[0])
    cat = np.memmap(human["base"] + ".txt.bin", dtype=np.uint8, mode="r", offset=4, shape=(n_text,))
    starts = np.fromfile(human["base"] + ".pos", dtype=np.uint32).astype(np.int64)
    m1, m2 = gen_data.gen_pairs_cat(cat, starts, n, 150, 2150)
    fq1, fq2 = os.path.join(human["dir"], "pe_1.fq"), os.path.join(human["dir"], "pe_2.fq")
    gen_data.write_fastq_matrix(fq1, m1)
    gen_data.write_fastq_matrix(fq2, m2)
    exe = os.path.join(gu.WORK, "pe_cpp_test_cuda")
    os.makedirs(gu.WORK, exist_ok=True)
    libdir = os.path.dirname(cuda_lib)
    subprocess.check_call(["g++", "-std=c++14", "-O2", "-I", os.path.join(gu.ROOT, "include"), PE_SRC, "-L", libdir,
                           "-lcolumba_b200", "-Wl,-rpath," + libdir, "-o", exe])

    def records(path):
        rows = []
        for line in open(path):
            if line.startswith("@"):
                continue
            f = line.rstrip("\n").split("\t")
            nm = [x for x in f[11:] if x.startswith("NM:i:")]
            rows.append((f[0], int(f[1]), f[2], int(f[3]), int(f[4]), f[5], f[6], int(f[7]), int(f[8]),
                         int(nm[0][5:]) if nm else -1))
        return sorted(rows)

    for mode, value, ref_args in (("best", 95, ["-a", "best", "-I", "95"]), ("all", 2, ["-a", "all", "-e", "2"])):
        t0 = time.perf_counter()
        r = subprocess.run([exe, human["base"], fq1, fq2, cb.SCHEMES_DIR, mode, str(value), "fr", "700", "200", "0", "0"],
                           capture_output=True, text=True)
        t_gpu = time.perf_counter() - t0
        assert r.returncode == 0, r.stderr[-500:]
        got = []
        for line in r.stdout.splitlines():
            a = line.split("\t")
            got.append((a[0], int(a[1]), a[2], int(a[3]), int(a[4]), a[5], a[6], int(a[7]), int(a[8]), int(a[9])))
        got.sort()
        sam = os.path.join(human["dir"], "pe_ref.sam")
        t0 = time.perf_counter()
        p = subprocess.run([REF_DEV, "-r", human["base"], "-f", fq1, "-F", fq2, "-nI", "-O", "fr", "-X", "700", "-N", "200",
                            "-t", str(os.cpu_count() or 1), "-o", sam] + ref_args, capture_output=True, text=True)
        t_ref = time.perf_counter() - t0
        assert p.returncode == 0, p.stdout[-400:] + p.stderr[-400:]
        exp = records(sam)
        os.unlink(sam)
        print("paired-end %s: %d pairs, %d records; C++ layer on the GPU %.1f s (incl. index load), reference %.1f s "
              "(incl. index load)" % (mode, n, len(exp), t_gpu, t_ref))
        assert len(got) == len(exp), (mode, len(got), len(exp))
        assert got == exp, mode


def test_batch_split_invariance(human):
    """One batch of 100 000 reads == four batches of 25 000 (device-side list growth, re-runs after list
    overflows and the global sort do not depend on how reads are batched)."""
    reads, ix = human["reads"], human["ix"]
    n = len(reads)
    m = cb.Mapper(ix)
    off = (np.arange(n + 1, dtype=np.int64) * 150).astype(np.uint32)
    whole = m.map_batch(reads.reshape(-1), off, K)
    q = n // 4
    for part in range(4):
        sub = m.map_batch(reads[part * q:(part + 1) * q].reshape(-1), off[:q + 1], K)
        assert (sub.read_count == whole.read_count[part * q:(part + 1) * q]).all()
        s0 = int(whole.read_start[part * q])
        cnt = int(sub.read_count.sum())
        a, b = sub.occ[:cnt], whole.occ[s0:s0 + cnt]
        for f in ("seq_id", "pos", "end", "index_begin", "strand", "dist", "n_cigar", "trimmed"):
            assert (a[f] == b[f]).all(), f
        used = np.arange(sub.cigar.shape[1])[None, :] < a["n_cigar"][:, None]  # slots beyond n_cigar are not written
        assert (np.where(used, sub.cigar[:cnt], 0) == np.where(used, whole.cigar[s0:s0 + cnt], 0)).all()
    m.close()
