"""Host-side mode drivers on top of the device pass (Python binding of the logic
that include/columba_b200.hpp implements for C++ callers).

In the reference BEST mode and paired-end pairing are host loops around the same
per-read mapping call; here they are
loops around `Mapper.map_batch(max_dist, min_dist)`:

  * map_best      -- SearchStrategy::matchApproxBestPlusX / findBestAlignments,
                     src/searchstrategy.cpp:623-744: exact stratum, then the strata
                     1, 3, 5, 9, 13 (or best + x), cut-off from the minimal identity
                     (getMaxED, src/searchstrategy.h:1797-1806).  Batched: every
                     stratum is one device pass over the reads that still need it.
                     As in the reference the two strands of a read are mapped and
                     filtered separately (clb_ctx_set_filter_per_strand), occurrences
                     are kept per strand and distance, clipped alignments move to the
                     level of their new distance (checkAlignments, :537-568) and the
                     result is combineOccVectors (:570-621).
  * pair_occurrences -- SearchStrategy::pairOccurrences, src/searchstrategy.cpp:1281-1343:
                     window join of two position-sorted occurrence lists on
                     [min_frag, max_frag].
"""
import numpy as np

from .api import Mapper, cigar_string

MAX_K = 13  # definitions.h:52 BEST_CUTOFF_COLUMBA = MAX_K


def max_ed(min_identity, seq_len, max_supported=MAX_K):
    """getMaxED, src/searchstrategy.h:1797-1806"""
    if not 50 <= min_identity <= 100:
        raise ValueError("minimal identity must be in [50, 100]")
    return min(MAX_K, max_supported, (seq_len * (100 - min_identity)) // 100)


def strata(cutoff, x=0):
    """The distances at which findBestAlignments maps while nothing has been
    found (src/searchstrategy.cpp:672-708): 0 (only for x == 0), then
    max(x, 1), and from there k -> min(k + x + (2 if k < 5 else 4), cutoff)."""
    out = [0] if x == 0 else []
    k = max(x, 1)
    while k <= cutoff:
        out.append(k)
        if k == cutoff:
            break
        k = min(k + x + (2 if k < 5 else 4), cutoff)
    return out


class Occ(tuple):
    """(strand, seq_id, pos, end, index_begin, dist, cigar)"""
    __slots__ = ()


def _group_occs(res, g):
    """Occurrences of group g of a batch result as (Occ, clipped, distance before clipping)."""
    s, c = int(res.read_start[g]), int(res.read_count[g])
    out = []
    for i in range(s, s + c):
        o = res.occ[i]
        occ = Occ((int(o["strand"]), int(o["seq_id"]), int(o["pos"]), int(o["end"]), int(o["index_begin"]),
                   int(o["dist"]), cigar_string(res.cigar[i, :int(o["n_cigar"])]) if o["n_cigar"] else "*"))
        tr = int(o["trimmed"])
        out.append((occ, tr != 0, tr - 1 if tr else int(o["dist"])))
    return out


def _occs_of(res, r):
    return [t[0] for t in _group_occs(res, r)]


class _BestState:
    """findBestAlignments' per-read state: occurrences per strand and distance, which distances have
    been mapped, and the stratum loop's variables."""

    def __init__(self, cutoff, x):
        self.cutoff, self.x = cutoff, x
        self.lv = [[[] for _ in range(cutoff + 1)] for _ in (0, 1)]   # [strand][distance] -> [(Occ, clipped)]
        self.done_upto = -1       # distances 0..done_upto have been mapped (both strands move in step)
        self.best = cutoff + 1
        self.best_found = False
        self.finished = False
        self.max_ed = cutoff
        self.k = max(x, 1)
        self.prev_k = 0

    def add(self, strand, items):
        for occ, clipped, level in items:
            if level <= self.cutoff:
                self.lv[strand][level].append((occ, clipped))

    def check(self, level):
        """checkAlignments (src/searchstrategy.cpp:537-568) for both strands"""
        for strand in (0, 1):
            keep, moved = [], []
            for occ, clipped in self.lv[strand][level]:
                if clipped:
                    if occ[5] > level:
                        moved.append(occ)
                else:
                    keep.append((occ, False))
                    if level < self.best:
                        self.best = level
            self.lv[strand][level] = keep
            for occ in moved:
                if occ[5] <= self.cutoff:
                    self.lv[strand][occ[5]].append((occ, False))

    def any_upto(self, k):
        return any(self.lv[s][d] for s in (0, 1) for d in range(min(k, self.cutoff) + 1))

    def after_exact(self):
        """x == 0: the exact stratum has been mapped (src/searchstrategy.cpp:637-662)"""
        self.done_upto = 0
        if self.lv[0][0] or self.lv[1][0]:
            self.check(0)
            if self.best == 0:
                self.best_found = True
        self.max_ed = self.x if self.best == 0 else self.cutoff
        if self.k > self.max_ed:
            self.finished = True

    def needs_mapping(self):
        return not self.finished and self.k > self.done_upto

    def step(self, mapped):
        """one iteration of the stratum loop (src/searchstrategy.cpp:677-708); `mapped`: distance k was
        mapped in this iteration (else its occurrences were already there)"""
        k = self.k
        if mapped:
            self.done_upto = k
            update = self.any_upto(k)
        else:
            update = bool(self.lv[0][k] or self.lv[1][k])
        if update:
            level = self.prev_k + 1
            while level <= min(k, self.best + self.x):
                self.check(level)
                level += 1
        if self.best_found:
            self.finished = True
            return
        if update and self.best < self.cutoff + 1:
            self.best_found = True
            if self.x == 0:
                self.finished = True
                return
            self.prev_k, self.k = k, min(self.best + self.x, self.max_ed)
        else:
            if k == self.max_ed:
                self.finished = True
                return
            self.prev_k, self.k = k, min(k + self.x + (2 if k < 5 else 4), self.max_ed)

    def result(self):
        """combineOccVectors (src/searchstrategy.cpp:570-621) -> (occurrences, number of best hits)"""
        if not self.best_found:
            return [], 0
        n_hits = len(self.lv[0][self.best]) + len(self.lv[1][self.best])
        out = []
        for d in range(self.best, min(self.best + self.x, self.cutoff) + 1):
            for strand in (0, 1):
                last = None
                for occ in sorted((o for o, _ in self.lv[strand][d]), key=lambda o: (o[1], o[2])):  # stable
                    if last is not None and (occ[1], occ[2]) == last:
                        continue
                    last = (occ[1], occ[2])
                    out.append(occ)
        return out, n_hits


def map_best(mapper, seqs, min_identity=95, x=0, with_hits=False):
    """BEST mode for a list of reads.  Returns (per-read list of Occ in the reference's output order --
    the first one is the primary alignment --, per-read best distance or None[, per-read nHits])."""
    n = len(seqs)
    supported = 0  # getMaxSupportedDistanceForBestMapping: largest k with all of 1..k supported
    while supported < MAX_K and mapper.supports(supported + 1):
        supported += 1
    st = [_BestState(max_ed(min_identity, len(s), supported), x) for s in seqs]
    mapper.set_filter_per_strand(True)
    try:
        def device_pass(ids, k, min_d, cutoff):
            mapper.set_trim_max(cutoff)  # checkAlignments clips with the cut-off
            reads, off = Mapper.pack([seqs[i] for i in ids])
            res = mapper.map_batch(reads, off, k, min_d, allow_short=True)
            for j, i in enumerate(ids):
                for strand in (0, 1):
                    st[i].add(strand, _group_occs(res, 2 * j + strand))
        if x == 0:
            by_cut = {}
            for i in range(n):
                by_cut.setdefault(st[i].cutoff, []).append(i)
            for cutoff, ids in sorted(by_cut.items()):
                device_pass(ids, 0, 0, cutoff)
            for s in st:
                s.after_exact()
        else:
            for s in st:
                if s.k > s.max_ed:
                    s.finished = True
        while True:
            groups = {}
            for i, s in enumerate(st):
                if s.needs_mapping():
                    groups.setdefault((s.cutoff, s.k, min(s.done_upto + 1, s.k)), []).append(i)
            for (cutoff, k, min_d), ids in sorted(groups.items()):
                device_pass(ids, k, min_d, cutoff)
            progressed = False
            for i, s in enumerate(st):
                if s.finished:
                    continue
                s.step(mapped=s.k > s.done_upto)
                progressed = True
            if not progressed:
                break
    finally:
        mapper.set_trim_max(0)
        mapper.set_filter_per_strand(False)
    occs, found, hits = [], [], []
    for s in st:
        o, h = s.result()
        occs.append(o)
        found.append(s.best if s.best_found else None)
        hits.append(h)
    return (occs, found, hits) if with_hits else (occs, found)


def best_records(occs, found, read_names, seq_names):
    """Every record the reference writes in BEST mode, as (read, flag & (16 | 256), sequence, pos, NM, CIGAR):
    the first occurrence of a read is its primary alignment, the others are secondary (flag 256)."""
    out = []
    for i, o in enumerate(occs):
        if found[i] is None:
            continue
        for j, t in enumerate(o):
            out.append((read_names[i], (16 if t[0] else 0) | (256 if j else 0), seq_names[t[1]], t[2] + 1, t[5], t[6]))
    return sorted(out)


def pair_occurrences(up, down, min_frag, max_frag):
    """pairOccurrences, src/searchstrategy.cpp:1281-1343, for the FR orientation:
    `up` = occurrences of the upstream mate on the forward strand, `down` = of
    the downstream mate on the reverse strand, both lists of Occ of ONE sequence
    id sorted by index_begin.  A pair is valid when the fragment from the begin
    of the upstream mate to the end of the downstream mate has a length in
    [min_frag, max_frag].  Returns (i_up, i_down, fragment length) triples."""
    ub = np.array([o[4] for o in up], dtype=np.int64)
    de = np.array([o[4] + (o[3] - o[2]) for o in down], dtype=np.int64)
    order = np.argsort(de, kind="stable")
    des = de[order]
    out = []
    for i, b in enumerate(ub):
        lo = np.searchsorted(des, b + min_frag, side="left")
        hi = np.searchsorted(des, b + max_frag, side="right")
        for j in range(lo, hi):
            d = int(order[j])
            if down[d][4] >= b:  # the downstream mate does not start before the upstream one
                out.append((i, d, int(des[j] - b)))
    return out


MAX_MAPQ = 60  # definitions.h:49
_COMP = bytes.maketrans(b"ACGTacgt", b"TGCAtgca")


def map_q(distance, n_hits, min_score):
    """TextOcc::getMapQ, src/indexhelpers.h:378-388"""
    if distance != min_score:
        return 0
    if n_hits == 1:
        return MAX_MAPQ
    import math
    return int(round(-10.0 * math.log10(1.0 - 1.0 / n_hits)))


def sam_single_end(res, read_names, seqs, quals, seq_names, unmapped=True):
    """The SAM body the reference writes for a chunk of single-end reads in ALL
    mode, from a batch result: generateOutputSingleEnd
    (src/searchstrategy.cpp:1824-1902, stable sort by distance: the first
    occurrence is the primary one), generateSE_SAM (src/searchstrategy.h:1633-1641),
    TextOcc::generateSAMSingleEnd (src/indexhelpers.cpp:56-87), flags
    getFlagsSE (src/indexhelpers.h:321-331), unmapped records
    createUnmappedRecordSE.  Sequence and quality are printed for the primary
    record only (reverse-complemented / reversed on the reverse strand).
    Returns the lines in read order."""
    out = []
    for r, name in enumerate(read_names):
        occ = _occs_of(res, r)
        qual = quals[r] if quals is not None and quals[r] else "*"
        if not occ:
            if unmapped:
                out.append("%s\t4\t*\t0\t0\t*\t*\t0\t0\t%s\t%s\tPG:Z:Columba\n" % (name, seqs[r], qual))
            continue
        occ = sorted(occ, key=lambda t: t[5])  # stable: text order within one distance
        min_score = occ[0][5]
        n_hits = sum(1 for t in occ if t[5] == min_score)
        for i, (strand, seq_id, pos, end, index_begin, dist, cigar) in enumerate(occ):
            flag = (16 if strand else 0) | (0 if i == 0 else 256)
            if i == 0:
                s = seqs[r].encode().translate(_COMP)[::-1].decode() if strand else seqs[r]
                q = (qual[::-1] if strand and qual != "*" else qual)
            else:
                s, q = "*", "*"
            out.append("%s\t%d\t%s\t%d\t%d\t%s\t*\t0\t0\t%s\t%s\tAS:i:%d\tNM:i:%d\tPG:Z:Columba\n" % (
                name, flag, seq_names[seq_id], pos + 1, map_q(dist, n_hits, min_score), cigar, s, q, dist, dist))
    return out
