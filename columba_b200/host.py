"""Host-side mode drivers on top of the ALL-mode device pass.

In the reference BEST mode and paired-end pairing are host loops around the same
per-read mapping call; here they are
loops around `Mapper.map_batch(max_dist, min_dist)`:

  * map_best      -- SearchStrategy::findBestAlignments, src/searchstrategy.cpp:623-712
                     (strata 0; 1, 3, 5, 9, 13 ... or best + x), cut-off from the
                     minimal identity (getMaxED, src/searchstrategy.h:1797-1806),
                     batched: every stratum is one device pass over the reads
                     that have no alignment yet.
  * pair_occurrences -- SearchStrategy::pairOccurrences, src/searchstrategy.cpp:1281-1343:
                     window join of two position-sorted occurrence lists on
                     [min_frag, max_frag].

Known difference to the reference, documented in DESIGN.md section 7: the
reference maps the two strands of a read in separate calls in BEST mode (so its
redundancy filter sees one strand at a time), the device pass filters both
strands together as the reference's ALL mode does.
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


def _occs_of(res, r):
    s, c = int(res.read_start[r]), int(res.read_count[r])
    out = []
    for i in range(s, s + c):
        o = res.occ[i]
        out.append(Occ((int(o["strand"]), int(o["seq_id"]), int(o["pos"]), int(o["end"]), int(o["index_begin"]),
                        int(o["dist"]), cigar_string(res.cigar[i, :int(o["n_cigar"])]) if o["n_cigar"] else "*")))
    return out


def map_best(mapper, seqs, min_identity=95, x=0):
    """BEST mode for a list of reads.  Returns (per-read list of Occ with
    distance <= best + x, per-read best distance or None)."""
    n = len(seqs)
    found = [None] * n
    occs = [[] for _ in range(n)]
    # reads of one length share the cut-off and therefore the strata
    by_cut = {}
    supported = 0  # getMaxSupportedDistanceForBestMapping: largest k with all of 1..k supported
    while supported < MAX_K and mapper.supports(supported + 1):
        supported += 1
    for i, s in enumerate(seqs):
        by_cut.setdefault(max_ed(min_identity, len(s), supported), []).append(i)
    for cutoff, ids in sorted(by_cut.items()):
        pending = list(ids)
        prev = -1
        sched = strata(cutoff, x)
        # The exact stratum (exactMatchesOutput, src/indexinterface.cpp:947-1014) has no
        # search scheme of its own: it is mapped together with stratum 1 (min_dist 0)
        # and only its distance-0 occurrences are kept when there are any.
        if sched and sched[0] == 0:
            sched = sched[1:] if len(sched) > 1 else [1]
        mapper.set_trim_max(cutoff)  # checkAlignments clips with the cut-off, src/searchstrategy.cpp:537-568
        for k in sched:
            if not pending:
                break
            reads, off = Mapper.pack([seqs[i] for i in pending])
            res = mapper.map_batch(reads, off, k, prev + 1, allow_short=True)
            still = []
            for j, i in enumerate(pending):
                o = [t for t in _occs_of(res, j) if t[5] <= cutoff]
                if o:
                    found[i] = min(t[5] for t in o)
                    occs[i] = o
                else:
                    still.append(i)
            # reads found at this stratum with x > 0: one more pass up to best + x
            if x > 0:
                more = {}
                for j, i in enumerate(pending):
                    if found[i] is not None and occs[i] and min(found[i] + x, cutoff) > k:
                        more.setdefault(min(found[i] + x, cutoff), []).append(i)
                for k2, ids2 in sorted(more.items()):
                    reads2, off2 = Mapper.pack([seqs[i] for i in ids2])
                    res2 = mapper.map_batch(reads2, off2, k2, k + 1, allow_short=True)
                    for j, i in enumerate(ids2):
                        occs[i] += _occs_of(res2, j)
            pending = still
            prev = k
    mapper.set_trim_max(0)
    for i in range(n):
        if found[i] is not None:
            occs[i] = sorted(t for t in occs[i] if t[5] <= found[i] + x)
    return occs, found


def best_tuples(occs, found, read_names, seq_names):
    """The record the reference writes per read in BEST mode (one primary
    alignment): smallest distance, then text order."""
    out = []
    for i, o in enumerate(occs):
        if found[i] is None:
            continue
        prim = min(o, key=lambda t: (t[5], t[4], t[3] - t[2]))
        out.append((read_names[i], 16 if prim[0] else 0, seq_names[prim[1]], prim[2] + 1, prim[5], prim[6]))
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
