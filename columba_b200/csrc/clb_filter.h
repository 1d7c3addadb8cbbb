// columba_b200 -- stages 4-6: per-read dedup of index occurrences, SA-range ->
// text positions, sort + unique + redundancy filter of the text occurrences,
// CIGAR strings for index hits, assignment to a reference sequence (with
// trimming at sequence boundaries).  One task per read.
//
// Replaces (reference): Occurrences::eraseDoublesFM src/indexhelpers.h:2135-2143,
// eraseDoublesAndSortText :2149-2157, FMOcc::operator< :1506-1522 / == :1529-1533,
// TextOcc::operator< :779-798 / == :811-813,
// IndexInterface::getUniqueTextOccurrences src/indexinterface.cpp:1373-1491,
// getTextOccHamming :1331-1371, generateCIGAR src/indexinterface.h:959-997,
// findSeqName src/indexinterface.cpp:799-897, and the assignment loop of
// SearchStrategy::generateOutputSingleEnd src/searchstrategy.cpp:1824-1856.
//
// The reference relies on stable sorts (DEVELOPER_MODE) and insertion order to
// pick one of several equal occurrences; here the insertion order is an
// explicit part of the sort key (strand, emission order), so any sorting
// algorithm yields the reference's survivor.
#pragma once
#include "clb_verify.h"

namespace clb {

template <typename Less> CLB_HD void heapSortIdx(u32* a, u32 n, Less less) {
    if (n < 2)
        return;
    // small segments: insertion sort
    if (n <= 16) {
        for (u32 i = 1; i < n; i++) {
            u32 v = a[i];
            u32 j = i;
            while (j > 0 && less(v, a[j - 1])) {
                a[j] = a[j - 1];
                j--;
            }
            a[j] = v;
        }
        return;
    }
    for (u32 start = n / 2; start-- > 0;) {
        u32 root = start;
        for (;;) {
            u32 child = 2 * root + 1;
            if (child >= n)
                break;
            if (child + 1 < n && less(a[child], a[child + 1]))
                child++;
            if (!less(a[root], a[child]))
                break;
            u32 t = a[root];
            a[root] = a[child];
            a[child] = t;
            root = child;
        }
    }
    for (u32 end = n - 1; end > 0; end--) {
        u32 t = a[0];
        a[0] = a[end];
        a[end] = t;
        u32 root = 0;
        for (;;) {
            u32 child = 2 * root + 1;
            if (child >= end)
                break;
            if (child + 1 < end && less(a[child], a[child + 1]))
                child++;
            if (!less(a[root], a[child]))
                break;
            u32 t2 = a[root];
            a[root] = a[child];
            a[child] = t2;
            root = child;
        }
    }
}

struct FMLess {
    const FMOccRec* d;
    CLB_HD bool operator()(u32 x, u32 y) const {
        const FMOccRec &a = d[x], &b = d[y];
        if (a.sa.b != b.sa.b) return a.sa.b < b.sa.b;
        if (a.dist != b.dist) return a.dist < b.dist;
        u32 wa = a.sa.width(), wb = b.sa.width();
        if (wa != wb) return wa < wb;
        if (a.shift != b.shift) return a.shift < b.shift;
        // insertion order of the reference: forward strand first, then
        // search by search in emission order
        if ((a.t & 1) != (b.t & 1)) return (a.t & 1) < (b.t & 1);
        return a.order < b.order;
    }
};
struct TxLess {
    const TextOccRec* d;
    CLB_HD bool operator()(u32 x, u32 y) const {
        const TextOccRec &a = d[x], &b = d[y];
        if (a.b != b.b) return a.b < b.b;
        if (a.dist != b.dist) return a.dist < b.dist;
        u32 wa = a.e <= a.b ? 0 : a.e - a.b, wb = b.e <= b.b ? 0 : b.e - b.b;
        if (wa != wb) return wa < wb;
        bool ca = a.ncig != 0, cb = b.ncig != 0;
        if (ca != cb) return ca;
        if (a.cls != b.cls) return a.cls < b.cls;
        if (a.cls == 0 && (a.t & 1) != (b.t & 1)) return (a.t & 1) < (b.t & 1);
        if (a.order != b.order) return a.order < b.order;
        return a.order2 < b.order2;
    }
};

struct FilterScratch {
    void* rows;  // rowsCap rows of Row<u128>
    u16* ops;    // 2 x cigW
    u32 rowsCap;
};

// 64-bit fast path of bpFindCigar (bitparallelmatrix.h:460-526): row state in
// registers, text as prefetched 64-bit windows, and per row only the band bits
// the traceback needs (HP and match | !D0, see BandStore) in `band`.
CLB_HD u32 findCigarFast(const IndexView& ix, const Batch& B, const PartSeq& ps, u32 b, u32 refLen, u32 score,
                         u64* band, u16* ops, u32 cap) {
    typedef BPC<u64> K;
    const u32 n = ps.len + 1, Wv = score, Wh = score, sh0 = K::DIAG_R0 - Wv;
    u64 HP = (~0ull) << K::LEFT, HN = ~HP;
    band[0] = bandPack(HP, 0, sh0);
    u64 M0 = matchVec<u64>(ps, 0, 0), M1 = matchVec<u64>(ps, 0, 1), M2 = matchVec<u64>(ps, 0, 2),
        M3 = matchVec<u64>(ps, 0, 3);
    u64 tcur = textWindow64(ix, b) << 2, tnxt = textWindow64(ix, b + K::BLOCK - 1);
    for (u32 i = 1; i <= refLen; i++) {
        const u32 l = i % K::BLOCK;
        if (l == 0) {
            HP >>= K::BLOCK;
            HN >>= K::BLOCK;
            tcur = tnxt;
            tnxt = textWindow64(ix, b + i + K::BLOCK - 1);
            M0 = matchVec<u64>(ps, i / K::BLOCK, 0);
            M1 = matchVec<u64>(ps, i / K::BLOCK, 1);
            M2 = matchVec<u64>(ps, i / K::BLOCK, 2);
            M3 = matchVec<u64>(ps, i / K::BLOCK, 3);
        }
        const u64 M = pickM(M0, M1, M2, M3, (u32)(tcur >> (2 * l)) & 3u);
        const u64 D0 = bandRow(HP, HN, M);
        band[i] = bandPack(HP, M | ~D0, l + sh0);
    }
    u32 i = refLen, j = n - 1, nn = 0;
    while (j > 0 || i > 0) {
        const u64 w = band[i];
        const u32 bidx = j + Wv - i; // wraps when the cell is left of the band
        const bool inBand = bidx <= Wv + Wh;
        u32 op;
        if (j > 0 && inBand && (((u32)w >> bidx) & 1u)) {
            --j;
            op = CIG_I;
        } else if (i > 0 && j > 0 && inBand && (((u32)(w >> 32) >> bidx) & 1u)) {
            --i;
            --j;
            op = CIG_M;
        } else {
            if (i == 0 || (j > 0 && !inBand)) { // the alignment left the band: not an occurrence with `score` errors
                atomOr(B.err, ERR_INTERNAL);
                return 0xFFFFFFFFu;
            }
            --i;
            op = CIG_D;
        }
        nn = cigPush(ops, nn, cap, op);
        if (nn == 0xFFFFFFFFu)
            return nn;
    }
    cigReverse(ops, nn);
    return nn;
}

// generateCIGAR, indexinterface.h:959-997: re-align the read against text[b,e)
CLB_HD u32 cigarForOcc(const IndexView& ix, const Batch& B, u32 t, u32 b, u32 e, u32 dist, FilterScratch& fs,
                       u16* ops, LocalCtr& lc) {
    PartSeq ps = fullSeq(B, t);
    TextRef ref;
    ref.ix = &ix;
    ref.start = b;
    u32 ee = tmin(e, ix.textLength);
    u32 refLen = ee <= b ? 0u : ee - b;
    lc.c[CIGARS_IN_INDEX]++;
    if (refLen + 1 > fs.rowsCap) {
        atomOr(B.err, ERR_ROWS_OVERFLOW);
        return 0;
    }
    u32 n;
    if (BPC<u64>::MAX_ED >= dist)
        n = findCigarFast(ix, B, ps, b, refLen, dist, (u64*)fs.rows, ops, B.cigW);
    else
        n = bpFindCigar<u128>((Row<u128>*)fs.rows, ps, ref, refLen, dist, ops, B.cigW);
    if (n == 0xFFFFFFFFu) {
        atomOr(B.err, ERR_INTERNAL);
        return 0;
    }
    return n;
}

// ---- stage 4: index occurrences -> text occurrences, as global passes ------
// (the reference does this per read: Occurrences::eraseDoublesFM
// src/indexhelpers.h:2135-2143 then getTextPositionsFromSARange per unique
// occurrence, src/indexinterface.cpp:1373-1491).  Here every index occurrence of
// the batch gets the key (group, SA begin), the keys are sorted by the
// backend's radix sort, runs of equal keys are put in FMLess order and made
// unique by their first lane, a prefix sum over the range widths of the
// survivors gives every (occurrence, row) pair its slot in the text-occurrence
// list, and one thread per slot does the locate.
CLB_HD u64 groupKey(u32 g, u32 b) { return ((u64)g << 32) | b; }
CLB_HD u32 keyGroup(u64 k) { return (u32)(k >> 32); }
CLB_HD u32 listSize(const u32* count, u32 cap) { return tmin(*count, cap); }
// consecutive sorted keys of the same group whose begins are at most maxDiff apart
CLB_HD bool keysLinked(u64 prev, u64 cur, u32 maxDiff) {
    return keyGroup(prev) == keyGroup(cur) && (u32)cur - (u32)prev <= maxDiff;
}

CLB_HD void fmKeyTask(const Batch& B, u32 i) {
    const FMOccRec& f = B.fmOccs.data[i];
    B.sortKey[0][i] = groupKey(groupOf(B, f.t), f.sa.b);
    B.sortVal[0][i] = i;
}
// run head: order the run, mark the unique occurrences with their widths
CLB_HD void fmRunTask(const Batch& B, u32 i, LocalCtr& lc) {
    const u32 n = listSize(B.fmOccs.count, B.fmOccs.cap);
    const u64* keys = B.sortKey[B.sortRes];
    u32* vals = B.sortVal[B.sortRes];
    u32* width = B.sortVal[1 - B.sortRes];
    if (i > 0 && keys[i] == keys[i - 1])
        return;
    u32 e = i + 1;
    while (e < n && keys[e] == keys[i])
        e++;
    const FMOccRec* d = B.fmOccs.data;
    FMLess less;
    less.d = d;
    heapSortIdx(vals + i, e - i, less);
    u32 last = i;
    u64 total = 0;
    for (u32 j = i; j < e; j++) {
        if (j > i) {
            const FMOccRec &a = d[vals[last]], &b = d[vals[j]];
            if (a.sa.e == b.sa.e && a.dist == b.dist && a.depth == b.depth && a.shift == b.shift &&
                (a.t & 1) == (b.t & 1)) {
                width[j] = 0;
                continue;
            }
        }
        last = j;
        const u32 w = d[vals[j]].sa.width();
        width[j] = w;
        total += w;
    }
    lc.c[TOTAL_REPORTED_POSITIONS] += total;
    lc.s[ST_TEXT_OCCS] += total;
}
// slots of the expansion in the text-occurrence list (one thread)
CLB_HD void fmReserveTask(const Batch& B) {
    const u32 n = listSize(B.fmOccs.count, B.fmOccs.cap);
    const u32* wOff = (const u32*)B.sortKey[1 - B.sortRes];
    const u32 total = n ? wOff[n] : 0u;
    const u32 base = *B.textOccs.count;
    B.fmSlots[0] = base;
    B.fmSlots[1] = total;
    if (base > B.textOccs.cap || total > B.textOccs.cap - base) {
        atomOr(B.err, ERR_TEXTOCC_OVERFLOW);
        B.fmSlots[1] = 0;
    }
    const u64 sum = (u64)base + total;
    *B.textOccs.count = sum > 0xFFFFFFFFull ? 0xFFFFFFFFu : (u32)sum;
}
// CIGAR of an occurrence with more than CIGAR_THRESHOLD rows: computed once for its first
// row and shared by all of them (getTextPositionsFromSARange, src/indexinterface.cpp:1416-1431)
CLB_HD void fmCigarTask(const IndexView& ix, const Batch& B, u32 j, FilterScratch& fs, LocalCtr& lc) {
    const u32* vals = B.sortVal[B.sortRes];
    const u32* width = B.sortVal[1 - B.sortRes];
    if (B.metric != METRIC_EDIT || B.noCigarFlag || width[j] <= 10)
        return;
    const FMOccRec f = B.fmOccs.data[vals[j]];
    u64 lf = 0;
    const u32 b = findSA(ix, f.sa.b, lf) + f.shift;
    B.fmNcig[j] = (u8)cigarForOcc(ix, B, f.t, b, b + f.depth, f.dist, fs, B.fmCig + (size_t)j * B.cigW, lc);
}
// one (occurrence, row) pair: locate
CLB_HD void fmExpandTask(const IndexView& ix, const Batch& B, u32 s, LocalCtr& lc) {
    const u32 n = listSize(B.fmOccs.count, B.fmOccs.cap);
    const u32* vals = B.sortVal[B.sortRes];
    const u32* wOff = (const u32*)B.sortKey[1 - B.sortRes];
    u32 lo = 0, hi = n; // last j with wOff[j] <= s
    while (hi - lo > 1) {
        const u32 mid = (lo + hi) >> 1;
        if (wOff[mid] <= s)
            lo = mid;
        else
            hi = mid;
    }
    const u32 j = lo, i = s - wOff[j], w = wOff[j + 1] - wOff[j];
    const FMOccRec f = B.fmOccs.data[vals[j]];
    const bool edit = B.metric == METRIC_EDIT;
    const u32 pos = findSA(ix, f.sa.b + i, lc.s[ST_LF_STEPS]);
    const u32 at = B.fmSlots[0] + s;
    TextOccRec o;
    o.t = f.t;
    o.b = edit ? pos + f.shift : pos;
    o.e = o.b + f.depth;
    o.order = wOff[j]; // increases with the occurrence's place in its read's sorted list
    o.order2 = i;
    o.dist = (u8)f.dist;
    o.cls = 1;
    o.pad = 0;
    o.ncig = 0;
    u16* dst = B.textCig + (size_t)at * B.cigW;
    if (!edit) {
        if (!B.noCigarFlag) {
            o.ncig = 1;
            dst[0] = (u16)((f.depth << 2) | CIG_M);
        }
    } else if (!B.noCigarFlag && w > 10) { // CIGAR_THRESHOLD
        const u32 ncig = B.fmNcig[j];
        const u16* src = B.fmCig + (size_t)j * B.cigW;
        o.ncig = (u8)ncig;
        for (u32 q = 0; q < ncig; q++)
            dst[q] = src[q];
    }
    B.textOccs.data[at] = o;
}

struct TrimSink {
    VerifyHit best;
    bool any;
    u16* bestOps;
    CLB_HD void operator()(const VerifyHit& h, const u16* ops) {
        bool better = !any;
        if (any) { // TextOcc::operator<
            u32 wa = h.e - h.b, wb = best.e - best.b;
            if (h.b != best.b) better = h.b < best.b;
            else if (h.dist != best.dist) better = h.dist < best.dist;
            else if (wa != wb) better = wa < wb;
            else better = false;
        }
        if (better) {
            best = h;
            any = true;
            for (u32 i = 0; i < h.ncig && h.ncig != 0xFFFFFFFFu; i++)
                bestOps[i] = ops[i];
        }
    }
};

// ---- stages 5-6: sort, unique, redundancy filter, sequence assignment --------
// (the reference: Occurrences::eraseDoublesAndSortText src/indexhelpers.h:2149-2157,
// the filter loop of getUniqueTextOccurrences src/indexinterface.cpp:1451-1485,
// findSeqName :799-897 in the assignment loop of generateOutputSingleEnd
// src/searchstrategy.cpp:1824-1856; all of it per read.)  Global passes here:
// key (group, begin) per text occurrence, radix sort, runs of equal keys put in
// TxLess order by their first lane.  The filter is a sequential scan whose
// state is the last kept occurrence; an occurrence that begins more than 2k
// after its predecessor in sorted order is kept whatever the state is, so the
// sorted list splits into independent CHAINS at those gaps and at group
// boundaries, and one thread walks one chain.  Survivors get their output slot
// from a prefix sum over the keep flags, so every read's occurrences are
// contiguous and in text order without gaps.
CLB_HD void txKeyTask(const Batch& B, u32 i) {
    const TextOccRec& o = B.textOccs.data[i];
    B.sortKey[0][i] = groupKey(groupOf(B, o.t), o.b);
    B.sortVal[0][i] = i;
}
CLB_HD void txRunTask(const Batch& B, u32 i) {
    const u32 n = listSize(B.textOccs.count, B.textOccs.cap);
    const u64* keys = B.sortKey[B.sortRes];
    u32* vals = B.sortVal[B.sortRes];
    if (i + 1 >= n || keys[i + 1] != keys[i] || (i > 0 && keys[i] == keys[i - 1]))
        return;
    u32 e = i + 2;
    while (e < n && keys[e] == keys[i])
        e++;
    TxLess less;
    less.d = B.textOccs.data;
    heapSortIdx(vals + i, e - i, less);
}
// findSeqName for a kept occurrence at sorted slot `slot`: 1 = inside one sequence; 0 = dropped,
// or it crosses a sequence end and is queued for txTrimTask, which may set the flag again
CLB_HD u32 txAssign(const IndexView& ix, const Batch& B, const TextOccRec& o, u32 slot) {
    if (o.dist < B.minD) // SearchStrategy::mapRead, src/searchstrategy.h:510-515: removed after the filter
        return 0u;
    u32 lo = 0, hi = ix.numSeq + 1; // upper_bound over seqStart[0..numSeq]
    while (lo < hi) {
        const u32 mid = (lo + hi) >> 1;
        if (ix.seqStart[mid] <= o.b)
            lo = mid + 1;
        else
            hi = mid;
    }
    if (o.e <= ix.seqStart[lo])
        return 1u;
    if (B.metric != METRIC_EDIT)
        return 0u;
    B.trimJobs.push(slot, B.err, ERR_OUT_OVERFLOW);
    return 0u;
}
CLB_HD void txChainTask(const IndexView& ix, const Batch& B, u32 i) {
    const u32 n = listSize(B.textOccs.count, B.textOccs.cap);
    const u64* keys = B.sortKey[B.sortRes];
    const u32* vals = B.sortVal[B.sortRes];
    u32* flag = B.sortVal[1 - B.sortRes];
    const TextOccRec* d = B.textOccs.data;
    const bool filter = B.metric == METRIC_EDIT && !B.noFilter;
    const u32 maxDiff = filter ? 2 * B.k : 0u;
    if (i > 0 && keysLinked(keys[i - 1], keys[i], maxDiff))
        return; // same group and within reach of the predecessor: not the head of a chain
    u32 prevBegin = 0xFFFFFFFFu, prevDepth = 0xFFFFFFFFu, prevED = B.k + 1;
    u32 pending = 0xFFFFFFFFu, lastU = 0xFFFFFFFFu;
    for (u32 j = i; j < n; j++) {
        if (j > i && !keysLinked(keys[j - 1], keys[j], maxDiff))
            break;
        const TextOccRec& o = d[vals[j]];
        if (!B.noFilter && lastU != 0xFFFFFFFFu) { // unique
            const TextOccRec& a = d[vals[lastU]];
            if (a.b == o.b && a.e == o.e && a.dist == o.dist) {
                flag[j] = 0;
                continue;
            }
        }
        lastU = j;
        if (!filter) {
            flag[j] = txAssign(ix, B, o, j);
            continue;
        }
        const u32 diff = o.b > prevBegin ? o.b - prevBegin : prevBegin - o.b;
        const u32 width = o.e <= o.b ? 0u : o.e - o.b;
        if (diff == 0) {
            flag[j] = 0;
            continue;
        }
        if (diff <= maxDiff) {
            if (o.dist > prevED || (o.dist == prevED && width >= prevDepth)) {
                flag[j] = 0;
                continue;
            }
            flag[pending] = 0; // the better occurrence replaces the last kept one
        } else if (pending != 0xFFFFFFFFu) {
            flag[pending] = txAssign(ix, B, d[vals[pending]], pending);
        }
        prevBegin = o.b;
        prevED = o.dist;
        prevDepth = width;
        pending = j;
    }
    if (pending != 0xFFFFFFFFu)
        flag[pending] = txAssign(ix, B, d[vals[pending]], pending);
}
// an occurrence that crosses a sequence end: clip and re-verify
// (findSeqName src/indexinterface.cpp:838-897, inTextVerificationOneString src/fmindex/fmindex.cpp:312-342)
CLB_HD void txTrimTask(const IndexView& ix, const Batch& B, u32 job, FilterScratch& fs, LocalCtr& lc) {
    const u32 slot = B.trimJobs.data[job];
    const u32* vals = B.sortVal[B.sortRes];
    u32* flag = B.sortVal[1 - B.sortRes];
    const TextOccRec o = B.textOccs.data[vals[slot]];
    flag[slot] = 0;
    u32 lo = 0, hi = ix.numSeq + 1;
    while (lo < hi) {
        const u32 mid = (lo + hi) >> 1;
        if (ix.seqStart[mid] <= o.b)
            lo = mid + 1;
        else
            hi = mid;
    }
    u32 index = lo - 1;
    if (B.metric == METRIC_EDIT && !B.noCigarFlag && o.ncig == 0)
        lc.c[CIGARS_IN_INDEX]++; // computed by the reference, then superseded by the clipped alignment
    u32 nb, ne;
    const u32 largest = B.trimMax; // ALL mode: k; BEST mode: the cut-off
    if ((ix.seqStart[index + 1] - o.b) <= largest) {
        index++;
        nb = ix.seqStart[index];
        ne = tmin(o.e, ix.seqStart[index + 1]);
    } else if ((o.e - ix.seqStart[index + 1]) <= largest) {
        nb = o.b;
        ne = ix.seqStart[index + 1];
    } else
        return;
    TrimSink sink;
    sink.any = false;
    sink.bestOps = B.trimCig + (size_t)job * B.cigW;
    bool considered = true, ok;
    lc.c[IN_TEXT_STARTED]++;
    if (use64(1, largest))
        ok = verifyWindow<u64>(ix, B, o.t, (Row<u64>*)fs.rows, nb, ne, 1, largest, 0, fs.ops, B.cigW, lc, considered,
                               sink);
    else
        ok = verifyWindow<u128>(ix, B, o.t, (Row<u128>*)fs.rows, nb, ne, 1, largest, 0, fs.ops, B.cigW, lc, considered,
                                sink);
    if (!ok && considered)
        lc.c[ABORTED_IN_TEXT_VERIF]++;
    if (!sink.any)
        return;
    TrimRec tr;
    tr.b = sink.best.b;
    tr.e = sink.best.e;
    tr.seqID = index;
    tr.dist = (u8)sink.best.dist;
    tr.ncig = (u8)(B.noCigarFlag ? 0u : sink.best.ncig);
    tr.pad[0] = o.dist; // distance before clipping
    tr.pad[1] = 0;
    B.trimRecs[job] = tr;
    B.trimOf[slot] = job;
    flag[slot] = 1;
}
// number of survivors (one thread)
CLB_HD void txReserveTask(const Batch& B) {
    const u32 n = listSize(B.textOccs.count, B.textOccs.cap);
    const u32* offs = (const u32*)B.sortKey[1 - B.sortRes];
    const u32 total = n ? offs[n] : 0u;
    *B.out.count = total;
    if (total > B.out.cap)
        atomOr(B.err, ERR_OUT_OVERFLOW);
}
// survivor at sorted slot i -> output record; first slot of a group -> the group's start
CLB_HD void txEmitTask(const IndexView& ix, const Batch& B, u32 i) {
    const u32 n = listSize(B.textOccs.count, B.textOccs.cap);
    const u64* keys = B.sortKey[B.sortRes];
    const u32* vals = B.sortVal[B.sortRes];
    const u32* flag = B.sortVal[1 - B.sortRes];
    const u32* offs = (const u32*)B.sortKey[1 - B.sortRes];
    const u32 g = keyGroup(keys[i]);
    {
        const u32 first = i ? keyGroup(keys[i - 1]) + 1 : 0u;
        for (u32 q = first; q <= g; q++)
            B.outStart[q] = offs[i];
        if (i + 1 == n)
            for (u32 q = g + 1; q <= B.G; q++)
                B.outStart[q] = offs[n];
    }
    if (!flag[i])
        return;
    const u32 at = offs[i];
    if (at >= B.out.cap)
        return;
    const u32 ti = vals[i];
    const TextOccRec o = B.textOccs.data[ti];
    OutOcc out;
    out.read = B.grpShift ? g : (g >> 1);
    out.strand = (u8)(o.t & 1);
    const u16* src = B.textCig + (size_t)ti * B.cigW;
    u32 ncig = o.ncig;
    u32 lo = 0, hi = ix.numSeq + 1;
    while (lo < hi) {
        const u32 mid = (lo + hi) >> 1;
        if (ix.seqStart[mid] <= o.b)
            lo = mid + 1;
        else
            hi = mid;
    }
    const u32 index = lo - 1;
    if (o.e > ix.seqStart[index + 1]) { // kept although it crosses a sequence end: clipped by txTrimTask
        const u32 job = B.trimOf[i];
        const TrimRec tr = B.trimRecs[job];
        out.trimmed = (u8)(1u + tr.pad[0]);
        out.indexBegin = tr.b;
        out.seqID = tr.seqID;
        out.pos = tr.b - ix.seqStart[tr.seqID];
        out.end = tr.e - ix.seqStart[tr.seqID];
        out.dist = tr.dist;
        ncig = tr.ncig;
        src = B.trimCig + (size_t)job * B.cigW;
    } else {
        out.trimmed = 0;
        out.indexBegin = o.b;
        out.seqID = index;
        out.pos = o.b - ix.seqStart[index];
        out.end = o.e - ix.seqStart[index];
        out.dist = o.dist;
        // generateCIGARS for index occurrences (indexinterface.h:939-949): the alignment itself is left to
        // a dense follow-up kernel (cigarJobTask)
        if (B.metric == METRIC_EDIT && !B.noCigarFlag && o.ncig == 0)
            B.cigJobs.push(at, B.err, ERR_OUT_OVERFLOW);
    }
    out.ncig = (u8)ncig;
    B.out.data[at] = out;
    u16* dst = B.outCig + (size_t)at * B.cigW;
    for (u32 q = 0; q < ncig; q++)
        dst[q] = src[q];
}
// hit count and the per-read counters of one group
CLB_HD void txGroupTask(const Batch& B, u32 g, LocalCtr& lc) {
    const u32 n = B.outStart[g + 1] - B.outStart[g];
    B.outCount[g] = n;
    if (B.grpShift == 0) { // per-strand groups (BEST mode strata): MAPPED_READS is the host loop's business
        if (!(g & 1u))
            lc.c[NUMBER_OF_READS]++;
        lc.c[TOTAL_UNIQUE_MATCHES] += n;
        return;
    }
    if (n) {
        lc.c[MAPPED_READS]++;
        lc.c[TOTAL_UNIQUE_MATCHES] += n;
    }
    lc.c[NUMBER_OF_READS]++;
}

// Exact stratum of read r, both strands: IndexInterface::exactMatchesOutput
// src/indexinterface.cpp:947-1014 + verifyInTextExact :918-943.  Backward
// exact match on the forward BWT until the range is no wider than the switch
// point, then the rest of the read is compared with the text.  No redundancy
// filter applies to these occurrences (B.noFilter).  The reference's counter
// arithmetic is reproduced as it stands: in ALL mode both strands append to
// the same vector and `p.size() - tOcc.size()` / `tOcc.size()` use its
// cumulative size (unsigned wrap-around included).
CLB_HD void exactTask(const IndexView& ix, const Batch& B, u32 r, LocalCtr& lc) {
    u32 cumul = 0; // size of the reference's tOcc vector
    for (u32 strand = 0; strand < 2; strand++) {
        const u32 t = 2 * r + strand;
        const u32 len = seqLen(B, t);
        if (B.grpShift == 0)
            cumul = 0; // BEST mode keeps one vector per strand
        if (len == 0)
            continue;
        const u8* codes = seqCodes(B, t);
        RangePair rg = completeRange(ix);
        u32 i = len;
        bool dead = false;
        for (; i-- > 0;) {
            const u32 c = codes[i];
            if (c > 3) {
                dead = true;
                break;
            }
            lc.s[ST_RANK_LOOKUPS] += 2;
            if (!extendOne(ix, DIR_BWD, true, c, rg)) {
                dead = true;
                break;
            }
            lc.c[NODE_COUNTER]++;
            if (rg.sa.width() <= ix.switchPoint)
                break;
        }
        if (dead)
            continue;
        const bool verify = (i != 0xFFFFFFFFu);
        const u32 remaining = verify ? i : 0u;
        const u32 w = rg.sa.width();
        u32 found = 0;
        for (u32 q = 0; q < w; q++) {
            const u32 pos = findSA(ix, rg.sa.b + q, lc.s[ST_LF_STEPS]);
            if (pos < remaining)
                continue;
            const u32 b = pos - remaining;
            bool ok = true;
            for (u32 j = 0; j < remaining && ok; j++)
                ok = textAt(ix, b + j) == codes[j];
            if (!ok)
                continue;
            TextOccRec o;
            o.t = t;
            o.b = b;
            o.e = b + len;
            o.order = q;
            o.order2 = 0;
            o.dist = 0;
            o.cls = 0;
            o.pad = 0;
            o.ncig = B.noCigarFlag ? 0 : 1;
            const u32 at = B.textOccs.push(o, B.err, ERR_TEXTOCC_OVERFLOW);
            if (at != 0xFFFFFFFFu && o.ncig)
                B.textCig[(size_t)at * B.cigW] = (u16)((len << 2) | CIG_M);
            found++;
        }
        cumul += found;
        if (verify) {
            lc.c[IN_TEXT_STARTED] += w;
            lc.c[ABORTED_IN_TEXT_VERIF] += (u64)w - (u64)cumul; // wraps like the reference's size_t arithmetic
        }
        lc.c[TOTAL_REPORTED_POSITIONS] += cumul;
        lc.s[ST_TEXT_OCCS] += found;
    }
}

// CIGAR of one surviving index occurrence (out[j]); one task per job so that
// the alignments run in full warps instead of inside the per-read filter.
CLB_HD void cigarJobTask(const IndexView& ix, const Batch& B, u32 job, FilterScratch& fs, LocalCtr& lc) {
    const u32 j = B.cigJobs.data[job];
    OutOcc& o = B.out.data[j];
    const u32 t = 2 * o.read + o.strand;
    u16* dst = B.outCig + (size_t)j * B.cigW;
    o.ncig = (u8)cigarForOcc(ix, B, t, o.indexBegin, o.indexBegin + (o.end - o.pos), o.dist, fs, dst, lc);
}

} // namespace clb
