// columba_b200 -- stage 3: locate + in-text verification.  One task per
// candidate (one SA row of a narrow range).
//
// Replaces (reference): FMIndex::getBeginPositions src/fmindex/fmindex.h:366-382,
// FMIndex::inTextVerification src/fmindex/fmindex.cpp:267-310,
// InTextVerificationTask::doTask src/indexhelpers.cpp:517-574,
// BitParallelED::findClusterCenters src/bitparallelmatrix.h:591-614 + traceBack
// :531-586, FMIndex::inTextVerificationHamming src/fmindex/fmindex.cpp:358-407,
// FMIndex::inTextVerificationOneString :312-342 (used by boundary trimming).
#pragma once
#include "clb_batch.h"
#include "clb_matrix.h"

namespace clb {

// fmindex.h:240-246: 64-bit words suffice iff the zero zone and the band fit
CLB_HD bool use64(u32 nZeros, u32 maxED) { return BPC<u64>::LEFT >= nZeros + maxED && BPC<u64>::MAX_ED >= maxED; }

CLB_HD PartSeq fullSeq(const Batch& B, u32 t) {
    PartSeq ps;
    ps.planes = seqPlanes(B, t, DIR_FWD);
    ps.nw = B.nw;
    ps.base = 0;
    ps.len = seqLen(B, t);
    return ps;
}

struct TextRef {
    const IndexView* ix;
    u32 start;
    CLB_HD u32 operator()(u32 i) const { return textAt(*ix, start + i); }
};

struct VerifyHit {
    u32 b, e, dist, ncig;
};

// Verify the read t against text[start, end) with the matrix initialised for
// nZeros free start rows.  For every cluster centre calls sink(hit, ops).
// Returns false when the verification was aborted.  rows: >= len+1+3*maxED+2.
template <typename W, typename Sink>
CLB_HD bool verifyWindow(const IndexView& ix, const Batch& B, u32 t, Row<W>* rows, u32 start, u32 end, u32 nZeros,
                         u32 maxED, u32 minED, u16* ops, u32 opsCap, LocalCtr& lc, bool& considered, Sink& sink) {
    PartSeq ps = fullSeq(B, t);
    MatGeom g;
    u16 zeros[2 * BPC<u128>::MAX_ED + 2];
    for (u32 i = 0; i < nZeros; i++)
        zeros[i] = 0;
    bpInit<W, u16>(g, rows, ps.len, maxED, zeros, nZeros, 0);
    const u32 size = end <= start ? 0u : end - start;
    considered = true;
    if (!g.inFinalColumn(size)) {
        considered = false; // skipped without counting as aborted (indexhelpers.cpp:528)
        return false;
    }
    TextRef ref;
    ref.ix = &ix;
    ref.start = start;
    u32 i;
    for (i = 0; i < size; ++i) {
        lc.s[ST_VERIFY_ROWS]++;
        if (!bpComputeRow<W>(g, rows, ps, i + 1, ref(i)))
            break;
    }
    const u32 sofc = g.sizeOfFinalColumn();
    if (i <= size - sofc) // unsigned, as in the reference
        return false;
    // findClusterCenters(i, ...)
    const u32 lastRow = i, firstRow = (g.m - 1) - sofc, col = g.n - 1;
    bool any = false;
    for (u32 r = lastRow; r > firstRow; r--) {
        u32 ED = bpAt<W>(rows, r, col);
        if (ED > maxED || ED < minED)
            continue;
        bool above = (r == firstRow) || ED <= bpAt<W>(rows, r - 1, col);
        bool below = (r == lastRow) || ED <= bpAt<W>(rows, r + 1, col);
        if (above && below) {
            any = true;
            VerifyHit h;
            u32 beg = 0, score = maxED + 1;
            u32 n = bpTraceBack<W>(g, rows, ps, ref, r, beg, score, ops, opsCap);
            lc.c[CIGARS_IN_TEXT_VERIFICATION]++;
            h.b = start + beg;
            h.e = start + r;
            h.dist = score;
            h.ncig = n;
            sink(h, ops);
        }
    }
    return any;
}

CLB_HD u32 pushTextOcc(const Batch& B, const TextOccRec& rec, const u16* ops) {
    u32 at = B.textOccs.push(rec, B.err, ERR_TEXTOCC_OVERFLOW);
    if (at != 0xFFFFFFFFu && ops) {
        u16* dst = B.textCig + (size_t)at * B.cigW;
        for (u32 i = 0; i < rec.ncig; i++)
            dst[i] = ops[i];
    }
    return at;
}

struct VerifyScratch {
    void* rows;  // Row<u64> or Row<u128>, rowsCap entries of the larger type
    u16* ops;    // cigW entries
    u32 rowsCap;
};

struct CandSink {
    const Batch& B;
    const Cand& c;
    u32 sub;
    CLB_HD CandSink(const Batch& B_, const Cand& c_) : B(B_), c(c_), sub(0) {}
    CLB_HD void operator()(const VerifyHit& h, const u16* ops) {
        TextOccRec r;
        r.t = c.t;
        r.b = h.b;
        r.e = h.e;
        r.order = c.order;
        r.order2 = sub++;
        r.dist = (u8)h.dist;
        r.cls = 0;
        r.pad = 0;
        if (h.ncig == 0xFFFFFFFFu || h.ncig > B.cigW) {
            atomOr(B.err, ERR_INTERNAL);
            r.ncig = 0;
            pushTextOcc(B, r, 0);
            return;
        }
        r.ncig = B.noCigarFlag ? 0 : (u8)h.ncig;
        pushTextOcc(B, r, ops);
    }
};

// stage 3 task (edit distance)
CLB_HD void verifyEditTask(const IndexView& ix, const Batch& B, u32 ci, VerifyScratch& vs, LocalCtr& lc) {
    const Cand c = B.cands.data[ci];
    u32 pos = findSA(ix, c.row, lc.s[ST_LF_STEPS]);
    u32 sum = pos + c.add;
    u32 start = sum >= c.sub ? sum - c.sub : 0u;
    const u32 maxED = c.maxED, minED = c.minED;
    const u32 nZeros = c.fixedStart ? 1u : 2 * maxED + 1;
    const u32 len = seqLen(B, c.t);
    lc.c[IN_TEXT_STARTED]++;
    CandSink sink(B, c);
    bool considered = true, ok;
    if (use64(nZeros, maxED)) {
        // number of rows: Wv + n with Wv = nZeros - 1 + maxED
        u32 nRows = nZeros - 1 + maxED + len + 1;
        u32 end = tmin(ix.textLength - 1, start + nRows - 1);
        ok = verifyWindow<u64>(ix, B, c.t, (Row<u64>*)vs.rows, start, end, nZeros, maxED, minED, vs.ops, B.cigW, lc,
                               considered, sink);
    } else {
        u32 nRows = nZeros - 1 + maxED + len + 1;
        u32 end = tmin(ix.textLength - 1, start + nRows - 1);
        ok = verifyWindow<u128>(ix, B, c.t, (Row<u128>*)vs.rows, start, end, nZeros, maxED, minED, vs.ops, B.cigW, lc,
                                considered, sink);
    }
    if (!ok && considered)
        lc.c[ABORTED_IN_TEXT_VERIF]++;
    lc.c[TOTAL_REPORTED_POSITIONS] += sink.sub;
    lc.s[ST_TEXT_OCCS] += sink.sub;
}

// stage 3 task (Hamming distance), fmindex.cpp:358-407
CLB_HD void verifyHammingTask(const IndexView& ix, const Batch& B, u32 ci, LocalCtr& lc) {
    const Cand c = B.cands.data[ci];
    u32 Tb = findSA(ix, c.row, lc.s[ST_LF_STEPS]);
    lc.c[IN_TEXT_STARTED]++;
    Tb = (Tb > c.sub) ? Tb - c.sub : 0u;
    const u32 pSize = seqLen(B, c.t);
    u32 Te = Tb + pSize;
    if (Te > ix.textLength)
        return;
    const u8* codes = seqCodes(B, c.t);
    u32 score = 0;
    for (u32 j = 0; j < pSize; j++) {
        u32 p = Tb + j;
        bool mism = (p == ix.textLength - 1) || textAt(ix, p) != codes[j];
        score += mism ? 1u : 0u;
        if (score > c.maxED)
            break;
    }
    if (score <= c.maxED && score >= c.minED) {
        TextOccRec r;
        r.t = c.t;
        r.b = Tb;
        r.e = Te;
        r.order = c.order;
        r.order2 = 0;
        r.dist = (u8)score;
        r.cls = 0;
        r.pad = 0;
        r.ncig = B.noCigarFlag ? 0 : 1;
        u16 op = (u16)((pSize << 2) | CIG_M);
        pushTextOcc(B, r, &op);
        lc.c[TOTAL_REPORTED_POSITIONS]++;
        lc.s[ST_TEXT_OCCS]++;
    }
}

} // namespace clb
