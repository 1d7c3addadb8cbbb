"""clb_gpu_build (the standalone index builder of the benchmark) against the
unmodified reference's columba_build: every index file byte-identical; and
the .gidx sidecar against the re-packing loader: identical mapping results."""
import hashlib
import json
import os
import subprocess

import pytest

import golden_util as gu

pytestmark = pytest.mark.gpu

REF_BUILD = os.path.join(gu.ROOT, "oracle", "_ref", "columba_build")
EXTS = [".cct", ".txt.bin", ".pos", ".sna", ".fsid", ".headerSN.bin", ".bwt", ".brt", ".rev.brt", ".sa.4", ".sa.bv.4",
        ".meta"]


def _builder():
    import __graft_entry__ as ge
    if not os.path.exists(ge.BUILDER):
        ge.build_builder()
    return ge.BUILDER


def _sha(fn):
    h = hashlib.sha256()
    with open(fn, "rb") as f:
        for blk in iter(lambda: f.read(1 << 22), b""):
            h.update(blk)
    return h.hexdigest()


def test_golden_fasta_sha256(tmp_path):
    """Same sha256 as the files the reference's columba_build wrote for the golden FASTA."""
    base = str(tmp_path / "g1")
    subprocess.check_call([_builder(), "--fasta", gu.golden_fasta(), "--out", base, "--quiet"])
    sums = json.load(open(os.path.join(gu.GOLD, "ref_index", "sha256.json")))
    for f, h in sums.items():
        ext = f[f.index("."):]
        assert _sha(base + ext) == h, ext


@pytest.mark.skipif(not os.path.exists(REF_BUILD), reason="reference build binary did not travel")
@pytest.mark.parametrize("profile,bp,nseq", [("human", 6_000_000, 5), ("ecoli", 4_600_000, 1)])
def test_synthetic_equals_columba_build(tmp_path, profile, bp, nseq):
    """Synthetic reference -> FASTA -> the unmodified columba_build; all files equal."""
    base = str(tmp_path / "syn")
    subprocess.check_call([_builder(), "--synthetic", profile, "--bp", str(bp), "--seed", "7", "--nseq", str(nseq),
                           "--out", base, "--write-fasta", "--quiet"])
    ref = str(tmp_path / "ref")
    subprocess.check_call([REF_BUILD, "-r", ref, "-f", base + ".fa"], stdout=subprocess.DEVNULL,
                          stderr=subprocess.DEVNULL)
    for ext in EXTS:
        assert _sha(base + ext) == _sha(ref + ext), ext
    # determinism of the generator
    base2 = str(tmp_path / "syn2")
    subprocess.check_call([_builder(), "--synthetic", profile, "--bp", str(bp), "--seed", "7", "--nseq", str(nseq),
                           "--out", base2, "--no-gidx", "--quiet"])
    assert _sha(base + ".txt.bin") == _sha(base2 + ".txt.bin")


def test_gidx_equals_repacked_index(tmp_path, cuda_lib):
    """Mapping through the .gidx sidecar == mapping through the re-packed reference files."""
    import columba_b200 as cb
    base = str(tmp_path / "g1")
    subprocess.check_call([_builder(), "--fasta", gu.golden_fasta(), "--out", base, "--quiet"])
    names, seqs = gu.read_fastq_gz(os.path.join(gu.GOLD, "g1_edit.fq.gz"))
    reads, off = cb.Mapper.pack(seqs)
    out = []
    for no_gidx in (False, True):
        if no_gidx:
            os.environ["CLB_NO_GIDX"] = "1"
        try:
            ix = cb.Index(base, 4, 10, 4, device=0)
            m = cb.Mapper(ix)
            res = m.map_batch(reads, off, 4)
            out.append((res.tuples(names, ix.seq_names), dict(res.counters)))
        finally:
            os.environ.pop("CLB_NO_GIDX", None)
    assert out[0] == out[1]
    exp, _ = gu.expected("edit_k4_columba")
    assert out[0][0] == exp
