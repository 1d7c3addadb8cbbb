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

// stage 4 task: dedup the FM occurrences of read r and expand them to text
// occurrences.
CLB_HD void fmPerReadTask(const IndexView& ix, const Batch& B, u32 r, FilterScratch& fs, LocalCtr& lc) {
    const u32 s0 = B.fmStart[r], s1 = B.fmStart[r + 1];
    if (s1 == s0)
        return;
    u32* idx = B.fmIdx + s0;
    const u32 n = s1 - s0;
    const FMOccRec* d = B.fmOccs.data;
    FMLess less;
    less.d = d;
    heapSortIdx(idx, n, less);
    // unique (adjacent)
    u32 m = 0;
    for (u32 i = 0; i < n; i++) {
        if (m > 0) {
            const FMOccRec &a = d[idx[m - 1]], &b = d[idx[i]];
            if (a.sa.b == b.sa.b && a.sa.e == b.sa.e && a.dist == b.dist && a.depth == b.depth &&
                a.shift == b.shift && (a.t & 1) == (b.t & 1))
                continue;
        }
        idx[m++] = idx[i];
    }
    const u32 hamSize = d[idx[0]].depth;
    for (u32 j = 0; j < m; j++) {
        const FMOccRec f = d[idx[j]];
        const u32 w = f.sa.width();
        lc.c[TOTAL_REPORTED_POSITIONS] += w;
        u32 at = B.textOccs.reserve(w, B.err, ERR_TEXTOCC_OVERFLOW);
        if (at == 0xFFFFFFFFu)
            return;
        const bool edit = B.metric == METRIC_EDIT;
        const bool calc = edit && !B.noCigarFlag && w > 10; // CIGAR_THRESHOLD
        u32 ncig = 0;
        for (u32 i = 0; i < w; i++) {
            u32 pos = findSA(ix, f.sa.b + i, lc.s[ST_LF_STEPS]);
            TextOccRec o;
            o.t = f.t;
            o.b = edit ? pos + f.shift : pos;
            o.e = o.b + (edit ? f.depth : hamSize);
            o.order = j;
            o.order2 = i;
            o.dist = (u8)f.dist;
            o.cls = 1;
            o.pad = 0;
            o.ncig = 0;
            u16* dst = B.textCig + (size_t)(at + i) * B.cigW;
            if (!edit) {
                if (!B.noCigarFlag) {
                    o.ncig = 1;
                    dst[0] = (u16)((hamSize << 2) | CIG_M);
                }
            } else if (calc) {
                if (i == 0) {
                    ncig = cigarForOcc(ix, B, f.t, o.b, o.e, f.dist, fs, fs.ops, lc);
                }
                o.ncig = (u8)ncig;
                for (u32 q = 0; q < ncig; q++)
                    dst[q] = fs.ops[q];
            }
            B.textOccs.data[at + i] = o;
        }
        lc.s[ST_TEXT_OCCS] += w;
    }
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

// stage 6 task: filter the text occurrences of read r and write the output
// `act`: the lanes of the warp that run this function side by side (they meet
// again before the output slots are reserved, so that one atomic serves them all)
CLB_HD void txPerReadTask(const IndexView& ix, const Batch& B, u32 r, FilterScratch& fs, LocalCtr& lc, unsigned act) {
    const u32 s0 = B.txStart[r], s1 = B.txStart[r + 1];
    B.outStart[r] = 0;
    B.outCount[r] = 0;
    u32* idx = B.txIdx + s0;
    const u32 n = s1 - s0;
    TextOccRec* d = B.textOccs.data;
    TxLess less;
    less.d = d;
    if (n > 1 && (n <= B.txPresorted || n > B.txPresortedMax)) // the others: see Backend::sortBigSegments
        heapSortIdx(idx, n, less);
    u32 m = 0;
    for (u32 i = 0; i < n; i++) {
        if (m > 0 && !B.noFilter) {
            const TextOccRec &a = d[idx[m - 1]], &b = d[idx[i]];
            if (a.b == b.b && a.e == b.e && a.dist == b.dist)
                continue;
        }
        idx[m++] = idx[i];
    }
    const bool edit = B.metric == METRIC_EDIT;
    u32 nOut = m;
    if (edit && !B.noFilter) { // redundancy filter, indexinterface.cpp:1451-1485
        const u32 maxDiff = 2 * B.k;
        u32 prevBegin = 0xFFFFFFFFu, prevDepth = 0xFFFFFFFFu, prevED = B.k + 1;
        nOut = 0;
        for (u32 i = 0; i < m; i++) {
            const TextOccRec& o = d[idx[i]];
            u32 diff = o.b > prevBegin ? o.b - prevBegin : prevBegin - o.b;
            if (diff == 0)
                continue;
            u32 width = o.e <= o.b ? 0 : o.e - o.b;
            if (diff <= maxDiff) {
                if (o.dist > prevED || (o.dist == prevED && width >= prevDepth))
                    continue;
                nOut--;
            }
            prevBegin = o.b;
            prevED = o.dist;
            prevDepth = width;
            idx[nOut++] = idx[i];
        }
    }
    warpSync(act);
    u32 at = B.out.reserve(nOut, B.err, ERR_OUT_OVERFLOW);
    if (at == 0xFFFFFFFFu || nOut == 0)
        return;
    u32 w = 0;
    for (u32 i = 0; i < nOut; i++) {
        TextOccRec o = d[idx[i]];
        u16* cig = B.textCig + (size_t)idx[i] * B.cigW;
        // generateCIGARS for index occurrences (indexinterface.h:939-949): the
        // alignment itself is left to a dense follow-up kernel (cigarJobTask)
        const bool pending = edit && !B.noCigarFlag && o.ncig == 0;
        // findSeqName, indexinterface.cpp:799-897
        u32 lo = 0, hi = ix.numSeq + 1; // upper_bound over seqStart[0..numSeq]
        while (lo < hi) {
            u32 mid = (lo + hi) >> 1;
            if (ix.seqStart[mid] <= o.b)
                lo = mid + 1;
            else
                hi = mid;
        }
        u32 index = lo - 1;
        OutOcc out;
        out.read = B.grpShift ? r : (r >> 1);
        out.strand = (u8)(o.t & 1);
        u32 ncig = o.ncig;
        const u16* src = cig;
        if (o.e <= ix.seqStart[index + 1]) {
            if (pending)
                B.cigJobs.push(at + w, B.err, ERR_OUT_OVERFLOW);
            out.trimmed = 0;
            out.indexBegin = o.b;
            out.pos = o.b - ix.seqStart[index];
            out.end = o.e - ix.seqStart[index];
            out.dist = o.dist;
        } else {
            if (!edit)
                continue;
            if (pending)
                lc.c[CIGARS_IN_INDEX]++; // computed by the reference, then superseded by the trimmed alignment
            u32 nb, ne;
            const u32 largest = B.trimMax; // ALL mode: k; BEST mode: the cut-off (indexinterface.cpp:799-897)
            if ((ix.seqStart[index + 1] - o.b) <= largest) {
                index++;
                nb = ix.seqStart[index];
                ne = tmin(o.e, ix.seqStart[index + 1]);
            } else if ((o.e - ix.seqStart[index + 1]) <= largest) {
                nb = o.b;
                ne = ix.seqStart[index + 1];
            } else
                continue;
            // inTextVerificationOneString, fmindex.cpp:312-342
            TrimSink sink;
            sink.any = false;
            sink.bestOps = fs.ops + B.cigW;
            bool considered = true, ok;
            lc.c[IN_TEXT_STARTED]++;
            if (use64(1, largest))
                ok = verifyWindow<u64>(ix, B, o.t, (Row<u64>*)fs.rows, nb, ne, 1, largest, 0, fs.ops, B.cigW, lc,
                                       considered, sink);
            else
                ok = verifyWindow<u128>(ix, B, o.t, (Row<u128>*)fs.rows, nb, ne, 1, largest, 0, fs.ops, B.cigW, lc,
                                        considered, sink);
            if (!ok && considered)
                lc.c[ABORTED_IN_TEXT_VERIF]++;
            if (!sink.any)
                continue;
            out.trimmed = 1;
            out.indexBegin = sink.best.b;
            out.pos = sink.best.b - ix.seqStart[index];
            out.end = sink.best.e - ix.seqStart[index];
            out.dist = (u8)sink.best.dist;
            ncig = B.noCigarFlag ? 0 : sink.best.ncig;
            src = sink.bestOps;
        }
        out.seqID = index;
        out.ncig = (u8)ncig;
        B.out.data[at + w] = out;
        u16* dst = B.outCig + (size_t)(at + w) * B.cigW;
        for (u32 q = 0; q < ncig; q++)
            dst[q] = src[q];
        w++;
    }
    B.outStart[r] = at;
    B.outCount[r] = w;
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
