#!/usr/bin/env python
"""Seeded synthetic workload generator.

Writes a FASTA reference (iid uniform ACGT + injected repeat families) and a
FASTQ read set (mutated substrings, both strands).  Everything is a pure
function of the arguments, so the GPU arm, the CPU reference arm and the tests
all see the same bytes.  numpy only (PCG64 streams are stable across versions).
"""
import argparse
import os

import numpy as np

ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)
COMP = np.zeros(256, dtype=np.uint8)
for a, b in zip(b"ACGTN", b"TGCAN"):
    COMP[a] = b


def _mutate_copy(rng, block, div):
    """Return a copy of block with substitution divergence div."""
    out = block.copy()
    n = len(out)
    m = rng.random(n) < div
    k = int(m.sum())
    if k:
        # substitute by a different base
        cur = np.searchsorted(ACGT, out[m])
        out[m] = ACGT[(cur + rng.integers(1, 4, k)) % 4]
    return out


def gen_reference(total_bp, seed, n_seqs=1, profile="auto"):
    """Return (names, list of uint8 arrays)."""
    rng = np.random.default_rng(seed)
    g = ACGT[rng.integers(0, 4, total_bp)]
    scale = total_bp / 3.1e9

    def inject(unit_len, copies, dlo, dhi, trunc_min=None):
        copies = int(copies)
        if copies <= 0 or unit_len >= total_bp // 4:
            return
        master = ACGT[rng.integers(0, 4, unit_len)]
        for _ in range(copies):
            blk = master
            if trunc_min is not None and unit_len > trunc_min:
                cut = int(rng.integers(0, unit_len - trunc_min))
                blk = master[cut:]
            d = rng.uniform(dlo, dhi)
            c = _mutate_copy(rng, blk, d)
            p = int(rng.integers(0, total_bp - len(c)))
            g[p:p + len(c)] = c

    if profile == "auto":
        profile = "ecoli" if total_bp < 20_000_000 else "human"
    if profile == "ecoli":
        f = total_bp / 4.6e6
        inject(5000, max(2, round(7 * f)), 0.005, 0.005)
        inject(1300, max(3, round(30 * f)), 0.01, 0.03)
    elif profile == "human":
        # SINE-like, LINE-like, 1 kbp family, segmental dups, microsatellites,
        # 171-bp tandem arrays -- counts scaled by genome size
        inject(300, max(50, 1_000_000 * scale), 0.05, 0.15)
        inject(6000, max(8, 80_000 * scale), 0.02, 0.12, trunc_min=900)
        inject(1000, max(10, 100_000 * scale), 0.01, 0.10)
        nseg = max(2, int(200 * scale))
        for _ in range(nseg):
            ln = min(40_000, total_bp // 50)
            src = int(rng.integers(0, total_bp - ln))
            dst = int(rng.integers(0, total_bp - ln))
            g[dst:dst + ln] = _mutate_copy(rng, g[src:src + ln].copy(),
                                           rng.uniform(0.005, 0.02))
        for _ in range(max(10, int(100_000 * scale))):
            u = int(rng.integers(1, 7))
            tot = int(rng.integers(60, 201))
            unit = ACGT[rng.integers(0, 4, u)]
            blk = np.tile(unit, tot // u + 1)[:tot]
            p = int(rng.integers(0, total_bp - tot))
            g[p:p + tot] = blk
        for _ in range(max(1, int(200 * scale))):
            units = min(2000, max(20, total_bp // (171 * 400)))
            unit = ACGT[rng.integers(0, 4, 171)]
            arr = np.concatenate([_mutate_copy(rng, unit, rng.uniform(0.01, 0.03))
                                  for _ in range(units)])
            p = int(rng.integers(0, total_bp - len(arr)))
            g[p:p + len(arr)] = arr
    elif profile == "dense":
        # repeat-dense test profile (a few big families, 1-10 % divergence):
        # makes most reads start real FM-index searches
        mb = total_bp / 1e6
        inject(300, max(20, 1000 * mb), 0.01, 0.10)
        inject(1000, max(6, 100 * mb), 0.01, 0.10)
        inject(3000, max(3, 20 * mb), 0.005, 0.05)
        for _ in range(max(4, int(30 * mb))):
            u = int(rng.integers(1, 7))
            tot = int(rng.integers(60, 201))
            unit = ACGT[rng.integers(0, 4, u)]
            blk = np.tile(unit, tot // u + 1)[:tot]
            p = int(rng.integers(0, total_bp - tot))
            g[p:p + tot] = blk
    elif profile == "none":
        pass
    else:
        raise ValueError(profile)

    # split into n_seqs sequences with GRCh38-like (decreasing) proportions
    if n_seqs == 1:
        return ["seq1_syn"], [g]
    w = np.array([1.0 / (1 + 0.12 * i) for i in range(n_seqs)])
    cuts = np.concatenate([[0], np.round(np.cumsum(w / w.sum()) * total_bp).astype(np.int64)])
    cuts[-1] = total_bp
    names = ["chr%d_syn" % (i + 1) for i in range(n_seqs)]
    return names, [g[cuts[i]:cuts[i + 1]] for i in range(n_seqs)]


def write_fasta(path, names, seqs, width=80):
    with open(path, "wb") as f:
        for nm, s in zip(names, seqs):
            f.write(b">" + nm.encode() + b" synthetic\n")
            n = len(s)
            full = (n // width) * width
            if full:
                body = np.empty((full // width, width + 1), dtype=np.uint8)
                body[:, :width] = s[:full].reshape(-1, width)
                body[:, width] = 10
                f.write(body.tobytes())
            if n > full:
                f.write(s[full:].tobytes() + b"\n")


def gen_reads(seqs, n_reads, read_len, seed, metric="edit", max_e=4,
              frac_over=0.05, frac_exact=0.05, over_e=6, n_frac=0.0):
    """Return list of (name, bytes).  Reads never cross a sequence end."""
    rng = np.random.default_rng(seed)
    lens = np.array([len(s) for s in seqs], dtype=np.int64)
    ok = lens > read_len + 16
    prob = np.where(ok, lens, 0).astype(np.float64)
    prob /= prob.sum()
    sid = rng.choice(len(seqs), size=n_reads, p=prob)
    u = rng.random(n_reads)
    strand = rng.integers(0, 2, n_reads)
    cls = rng.random(n_reads)
    ne = rng.integers(0, max_e + 1, n_reads)
    ne = np.where(cls < frac_over, over_e, ne)
    ne = np.where(cls > 1.0 - frac_exact, 0, ne)
    out = []
    for r in range(n_reads):
        s = seqs[sid[r]]
        span = read_len + 8
        p = int(u[r] * (len(s) - span))
        frag = bytearray(s[p:p + span].tobytes())
        for _ in range(int(ne[r])):
            q = int(rng.integers(0, read_len))
            t = rng.random() if metric == "edit" else 0.0
            if t < 0.70:
                cur = b"ACGT".index(frag[q]) if frag[q] in b"ACGT" else 0
                frag[q] = b"ACGT"[(cur + int(rng.integers(1, 4))) % 4]
            elif t < 0.85:
                frag.insert(q, b"ACGT"[int(rng.integers(0, 4))])
            else:
                del frag[q]
        rd = np.frombuffer(bytes(frag[:read_len]), dtype=np.uint8)
        if n_frac > 0 and rng.random() < n_frac:
            rd = rd.copy()
            rd[int(rng.integers(0, read_len))] = ord("N")
        if strand[r]:
            rd = COMP[rd[::-1]]
        out.append(("r%d" % r, rd.tobytes()))
    return out


def gen_reads_fast(seqs, n_reads, read_len, seed, metric="edit", max_e=4, frac_over=0.05, frac_exact=0.05,
                   over_e=6):
    """Vectorised read generator for the benchmark-sized sets (same error model
    as gen_reads: edits ~ U{0..max_e} for 90 %, over_e for 5 %, exact for 5 %;
    each edit substitution .70 / insertion .15 / deletion .15; strands 50/50).
    Returns a uint8 matrix n_reads x read_len (ASCII)."""
    lens = np.array([len(s) for s in seqs], dtype=np.int64)
    starts = np.concatenate([[0], np.cumsum(lens)])
    cat = np.concatenate(seqs) if len(seqs) > 1 else seqs[0]
    return gen_reads_cat(cat, starts, n_reads, read_len, seed, metric, max_e, frac_over, frac_exact, over_e)


def gen_reads_cat(cat, starts, n_reads, read_len, seed, metric="edit", max_e=4, frac_over=0.05, frac_exact=0.05,
                  over_e=6):
    """gen_reads_fast on the concatenated text `cat` (uint8 ASCII, e.g. a memory-mapped .txt.bin) with the
    sequence start offsets `starts` (numSeq + 1 entries)."""
    rng = np.random.default_rng(seed)
    starts = np.asarray(starts, dtype=np.int64)
    lens = np.diff(starts)
    span = read_len + 8
    ok = lens > span + 8
    prob = np.where(ok, lens, 0).astype(np.float64)
    prob /= prob.sum()
    sid = rng.choice(len(lens), size=n_reads, p=prob)
    p = (rng.random(n_reads) * (lens[sid] - span)).astype(np.int64) + starts[sid]
    strand = rng.integers(0, 2, n_reads)
    cls = rng.random(n_reads)
    ne = rng.integers(0, max_e + 1, n_reads)
    ne = np.where(cls < frac_over, over_e, ne)
    ne = np.where(cls > 1.0 - frac_exact, 0, ne)
    out = np.empty((n_reads, read_len), dtype=np.uint8)
    J = np.arange(span, dtype=np.int64)[None, :]
    CH = 1 << 18
    for c0 in range(0, n_reads, CH):
        c1 = min(n_reads, c0 + CH)
        m = c1 - c0
        frag = cat[p[c0:c1, None] + J]
        for e in range(int(max(over_e, max_e))):
            act = ne[c0:c1] > e
            q = rng.integers(0, read_len, m)
            t = rng.random(m) if metric == "edit" else np.zeros(m)
            base = ACGT[rng.integers(0, 4, m)]
            off = rng.integers(1, 4, m)
            rows = np.nonzero(act & (t < 0.70))[0]
            if len(rows):
                cur = np.searchsorted(ACGT, frag[rows, q[rows]])
                frag[rows, q[rows]] = ACGT[(cur + off[rows]) % 4]
            rows = np.nonzero(act & (t >= 0.70) & (t < 0.85))[0]
            if len(rows):
                qq = q[rows, None]
                src = np.where(J > qq, J - 1, J)
                sub = np.take_along_axis(frag[rows], src, axis=1)
                sub[np.arange(len(rows)), q[rows]] = base[rows]
                frag[rows] = sub
            rows = np.nonzero(act & (t >= 0.85))[0]
            if len(rows):
                qq = q[rows, None]
                src = np.minimum(np.where(J >= qq, J + 1, J), span - 1)
                frag[rows] = np.take_along_axis(frag[rows], src, axis=1)
        rd = frag[:, :read_len]
        rc = COMP[rd[:, ::-1]]
        out[c0:c1] = np.where(strand[c0:c1, None] == 1, rc, rd)
    return out


def gen_pairs_cat(cat, starts, n_pairs, read_len, seed, max_e=3, frag_mean=400, frag_sd=50, frag_min=200, frag_max=700):
    """FR read pairs (SURVEY.md section 8d, configs[3]): fragments of length ~N(frag_mean, frag_sd) clipped to
    [frag_min, frag_max] drawn from the concatenated text `cat` (never across a sequence end), mate 1 = the first
    read_len characters, mate 2 = the reverse complement of the last read_len characters, fragment strand 50/50,
    0..max_e substitutions per mate.  Returns two uint8 matrices."""
    rng = np.random.default_rng(seed)
    starts = np.asarray(starts, dtype=np.int64)
    lens = np.diff(starts)
    ok = lens > frag_max + 16
    prob = np.where(ok, lens, 0).astype(np.float64)
    prob /= prob.sum()
    sid = rng.choice(len(lens), size=n_pairs, p=prob)
    fl = np.clip(rng.normal(frag_mean, frag_sd, n_pairs), frag_min, frag_max).astype(np.int64)
    p = (rng.random(n_pairs) * (lens[sid] - frag_max - 8)).astype(np.int64) + starts[sid]
    J = np.arange(read_len, dtype=np.int64)[None, :]
    a = np.array(cat[p[:, None] + J])
    b = np.array(cat[(p + fl - read_len)[:, None] + J])
    b = COMP[b[:, ::-1]]
    flip = rng.integers(0, 2, n_pairs) == 1  # fragment from the reverse strand: the mates swap roles
    m1 = np.where(flip[:, None], b, a)
    m2 = np.where(flip[:, None], a, b)
    for m in (m1, m2):
        ne = rng.integers(0, max_e + 1, n_pairs)
        for e in range(max_e):
            rows = np.nonzero(ne > e)[0]
            q = rng.integers(0, read_len, len(rows))
            cur = np.searchsorted(ACGT, m[rows, q])
            m[rows, q] = ACGT[(cur + rng.integers(1, 4, len(rows))) % 4]
    return m1, m2


def write_fastq_matrix(path, reads, first=0):
    """FASTQ of a uint8 read matrix; read i is named r<first+i>."""
    n, L = reads.shape
    qual = b"I" * L
    with open(path, "wb") as f:
        for c0 in range(0, n, 1 << 16):
            blk = reads[c0:c0 + (1 << 16)]
            f.write(b"".join(b"@r%d\n%s\n+\n%s\n" % (first + c0 + i, blk[i].tobytes(), qual)
                             for i in range(len(blk))))


def write_fastq(path, reads):
    with open(path, "wb") as f:
        for nm, s in reads:
            f.write(b"@" + nm.encode() + b"\n" + s + b"\n+\n" + b"I" * len(s) + b"\n")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref-bp", type=int, required=True)
    ap.add_argument("--ref-seed", type=int, default=4600)
    ap.add_argument("--n-seqs", type=int, default=1)
    ap.add_argument("--profile", default="auto")
    ap.add_argument("--reads", type=int, default=0)
    ap.add_argument("--read-len", type=int, default=150)
    ap.add_argument("--read-seed", type=int, default=150)
    ap.add_argument("--metric", default="edit")
    ap.add_argument("--max-e", type=int, default=4)
    ap.add_argument("--n-frac", type=float, default=0.0)
    ap.add_argument("--out", required=True, help="output prefix")
    a = ap.parse_args()
    names, seqs = gen_reference(a.ref_bp, a.ref_seed, a.n_seqs, a.profile)
    os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
    write_fasta(a.out + ".fa", names, seqs)
    if a.reads:
        reads = gen_reads(seqs, a.reads, a.read_len, a.read_seed, a.metric,
                          a.max_e, n_frac=a.n_frac,
                          over_e=(a.max_e + 2))
        write_fastq(a.out + ".fq", reads)


if __name__ == "__main__":
    main()
