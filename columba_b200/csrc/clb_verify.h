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

// Storage of the 64-bit fast path.  No matrix row goes to HBM: the traceback
// only needs, per cell of the band, "horizontal delta is +1" (HP) and "the
// diagonal predecessor explains the value" (match | !D0), i.e. 2 bits per
// cell of a band that is at most 31 cells wide.  The forward pass keeps these
// for the rows of the current 32-row block in `ring` (shared memory on the
// GPU) and one checkpoint (HP, HN entering the block) per block in `cp`; the
// traceback re-computes an earlier block from its checkpoint into the ring
// when it gets there.  Element i of a worker lives at [i * stride] (stride =
// number of workers that interleave: lanes of a warp / threads of a CTA).
struct BandStore {
    u64* ring;      // 32 slots: low word = HP band, high word = (M | ~D0) band
    u32 ringStride;
    u64* cp;        // 2 x cpCap: HP, HN
    u32 cpStride;
    u32 cpCap;
};
// checkpoints the full verification has room for (VerifyKernel::cpCapOf of the batch's rows capacity)
CLB_HD u32 VerifyKernelCpCap(const Batch& B) { return B.verifyRowsCap / BPC<u64>::BLOCK + 2; }
struct VerifyScratch {
    void* rows;   // Row<u128>[rowsCap], thread private (wide path only; may be null)
    BandStore bs; // fast path
    u16* ops;     // cigW entries
    u32 rowsCap;
    unsigned warpMask; // lanes of the warp that run a task in this round (1 on the host)
    u32 band32;        // forward pass and block re-computation on the 32-bit band where it fits (screenStep32)
};

struct CandSink {
    const Batch& B;
    const Cand& c;
    u32 sub;
    CLB_HD CandSink(const Batch& B_, const Cand& c_, u32 first) : B(B_), c(c_), sub(first) {}
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

// One row of the banded matrix (bitparallelmatrix.h:352-398) on the state in
// registers; returns D0.
CLB_HD u64 bandRow(u64& HP, u64& HN, u64 M) {
    const u64 D0 = (((M & HP) + HP) ^ HP) | M | HN;
    const u64 VP = HN | ~(D0 | HP);
    const u64 VN = D0 & HP;
    HP = (VN << 1) | ~(D0 | (VP << 1));
    HN = D0 & (VP << 1);
    return D0;
}
// match vector of character code y (0..3) out of four registers (a dynamically
// indexed array would live in local memory)
CLB_HD u64 pickM(u64 m0, u64 m1, u64 m2, u64 m3, u32 y) {
    const u64 a = (y & 1u) ? m1 : m0, b = (y & 1u) ? m3 : m2;
    return (y & 2u) ? b : a;
}
CLB_HD u64 bandPack(u64 HP, u64 X, u32 sh) { return (u64)(u32)(HP >> sh) | ((u64)(u32)(X >> sh) << 32); }
CLB_HD u64 bandPack32(u32 HP, u32 X, u32 sh) { return (u64)(HP >> sh) | ((u64)(X >> sh) << 32); }
// one row of the 32-bit band that moves with the diagonal (see screenStep32): the previous row (HP, HN) is
// re-aligned to this row's columns first; returns D0
CLB_HD u32 bandRow32(u32& HP, u32& HN, u32 M) {
    const u32 HPa = (HP >> 1) | 0x80000000u, HNa = HN >> 1;
    const u32 D0 = (((M & HPa) + HPa) ^ HPa) | M | HNa;
    const u32 VP = HNa | ~(D0 | HPa);
    const u32 VN = D0 & HPa;
    HP = (VN << 1) | ~(D0 | (VP << 1));
    HN = D0 & (VP << 1);
    return D0;
}

// 64-bit fast path of verifyWindow: the row state (HP, HN, RAC, score) stays in
// registers, the four match vectors of the current 32-row block are cut out
// of the read's bit planes once per block, the text is read one 16-character
// word at a time, the final-column distances are kept in a small local array
// and the traceback information lives in the BandStore (see there).  Same
// results as verifyWindow<u64> (bitparallelmatrix.h:352-415, :591-614, :531-586).
// `wm` = the lanes of the warp that call this function together: they are
// brought back together after the forward pass (their row counts differ:
// fixed-start and free-start candidates, aborted candidates) and walk the
// traceback phase in lock step.
template <typename Sink>
CLB_HD bool verifyWindowFast(const IndexView& ix, const Batch& B, u32 t, const BandStore& bs, u32 start, u32 end,
                             u32 nZeros, u32 maxED, u32 minED, u16* ops, u32 opsCap, LocalCtr& lc, bool& considered,
                             Sink& sink, unsigned wm, u32 firstCentre, u32 maxCentres, u32& totalCentres, u32 mult,
                             bool band32 = false) {
    typedef BPC<u64> K;
    const PartSeq ps = fullSeq(B, t);
    const u32 n = ps.len + 1;
    const u32 Wv = nZeros - 1 + maxED, Wh = maxED;
    const u32 m = Wv + n;
    const u32 sofc = Wh + Wv + 1;
    const u32 size = end <= start ? 0u : end - start;
    considered = true;
    bool run = true;
    if (!(size >= m - sofc)) {
        considered = false;
        run = false;
    } else if (size / K::BLOCK + 1 > bs.cpCap) {
        atomOr(B.err, ERR_ROWS_OVERFLOW);
        run = false;
    }
    const u32 firstRow = (m - 1) - sofc, col = n - 1;
    const u32 sh0 = K::DIAG_R0 - Wv; // band of row i = bits [i%32 + sh0, i%32 + sh0 + Wv + Wh]
    u8 eds[3 * K::MAX_ED + 4];
    u64 HP = (~0ull) << K::LEFT, HN = ~HP;
    {
        const u32 nn = tmin<u32>(nZeros, K::LEFT + 1);
        for (u32 i = 1; i < nn; ++i)
            HN ^= 1ull << (K::LEFT - i);
    }
    u64 RAC = 1ull << (K::DIAG_R0 + Wh);
    u32 score = 0;
    // distance in the final column of the row held in registers
    auto edAtCol = [&](u32 i) -> u32 {
        const u32 bit = (i % K::BLOCK) + K::DIAG_R0;
        u32 b = (i > col) ? bit - (i - col) + 1 : bit + 1;
        u32 e = (i > col) ? bit + 1 : bit + (col - i) + 1;
        u64 mask = ((1ull << (e - b)) - 1ull) << b;
        int neg = popc64(HN & mask), pos = popc64(HP & mask);
        return score + ((i > col) ? (u32)(neg - pos) : (u32)(pos - neg));
    };
    u32 i = 1;
    // 32-bit band: bit dB = the main diagonal, the band starts at bit mu (screenStep32)
    const bool b32 = band32 && Wv + Wh + 1 + 2 <= 32;
    const u32 mu = b32 ? (32 - (Wv + Wh + 1)) / 2 : 0u, dB = Wv + mu;
    if (run && b32) {
        u32 hp = (~0u) << (dB + 1), hn = (1u << (dB + 2 - nZeros)) - 1u, rac = 1u << (dB + Wh);
        const u32 stop = 1u << mu;
        auto edAtCol32 = [&](u32 r) -> u32 {
            const u32 b = (r > col) ? dB - (r - col) + 1 : dB + 1;
            const u32 e = (r > col) ? dB + 1 : dB + (col - r) + 1;
            const u32 mask = ((1u << (e - b)) - 1u) << b;
            const int neg = popc32(hn & mask), pos = popc32(hp & mask);
            return score + ((r > col) ? (u32)(neg - pos) : (u32)(pos - neg));
        };
        bs.ring[0] = bandPack32(hp, 0, mu);
        bs.cp[0] = hp;
        bs.cp[bs.cpStride] = hn;
        if (firstRow == 0)
            eds[0] = (u8)edAtCol32(0);
        u64 M0 = seqWindow64(ps, 0, -1 - (i32)dB), M1 = seqWindow64(ps, 1, -1 - (i32)dB),
            M2 = seqWindow64(ps, 2, -1 - (i32)dB), M3 = seqWindow64(ps, 3, -1 - (i32)dB);
        u64 tcur = textWindow64(ix, start) << 2, tnxt = textWindow64(ix, start + K::BLOCK - 1);
        bool ok = true;
        for (i = 1; i <= size; ++i) {
            const u32 l = i % K::BLOCK;
            if (l == 0) {
                tcur = tnxt;
                tnxt = textWindow64(ix, start + i + K::BLOCK - 1);
                const i32 w0 = (i32)i - 1 - (i32)dB;
                M0 = seqWindow64(ps, 0, w0);
                M1 = seqWindow64(ps, 1, w0);
                M2 = seqWindow64(ps, 2, w0);
                M3 = seqWindow64(ps, 3, w0);
                const u32 b = i / K::BLOCK; // the row before the block, on its own columns
                bs.cp[(size_t)(2 * b) * bs.cpStride] = hp;
                bs.cp[(size_t)(2 * b + 1) * bs.cpStride] = hn;
            }
            const u32 Y = (u32)(tcur >> (2 * l)) & 3u;
            const u32 M = (u32)(pickM(M0, M1, M2, M3, Y) >> l);
            const u32 D0 = bandRow32(hp, hn, M);
            score += ((D0 >> dB) & 1u) ^ 1u;
            if (!(D0 & rac))
                ok = racWalk(hp, hn, rac, stop);
            if (!ok)
                break;
            bs.ring[(size_t)l * bs.ringStride] = bandPack32(hp, M | ~D0, mu);
            if (i >= firstRow)
                eds[i - firstRow] = (u8)edAtCol32(i);
        }
        if (firstCentre == 0)
            lc.s[ST_VERIFY_ROWS] += (i <= size) ? i : size;
    } else if (run) {
        bs.ring[0] = bandPack(HP, 0, sh0);
        bs.cp[0] = HP;
        bs.cp[bs.cpStride] = HN;
        if (firstRow == 0)
            eds[0] = (u8)edAtCol(0);
        u64 M0 = matchVec<u64>(ps, 0, 0), M1 = matchVec<u64>(ps, 0, 1), M2 = matchVec<u64>(ps, 0, 2),
            M3 = matchVec<u64>(ps, 0, 3);
        // text of the rows of a block as one 64-bit window (row i <-> bits 2 * (i % 32));
        // the next block's window is fetched a block ahead, at the same iteration
        // in every lane
        u64 tcur = textWindow64(ix, start) << 2, tnxt = textWindow64(ix, start + K::BLOCK - 1);
        bool ok = true;
        for (i = 1; i <= size; ++i) {
            const u32 l = i % K::BLOCK;
            RAC <<= 1;
            if (l == 0) {
                tcur = tnxt;
                tnxt = textWindow64(ix, start + i + K::BLOCK - 1);
                HP >>= K::BLOCK;
                HN >>= K::BLOCK;
                RAC >>= K::BLOCK;
                const u32 b = i / K::BLOCK;
                M0 = matchVec<u64>(ps, b, 0);
                M1 = matchVec<u64>(ps, b, 1);
                M2 = matchVec<u64>(ps, b, 2);
                M3 = matchVec<u64>(ps, b, 3);
                bs.cp[(size_t)(2 * b) * bs.cpStride] = HP;
                bs.cp[(size_t)(2 * b + 1) * bs.cpStride] = HN;
            }
            const u32 Y = (u32)(tcur >> (2 * l)) & 3u;
            const u64 M = pickM(M0, M1, M2, M3, Y);
            const u64 D0 = bandRow(HP, HN, M);
            const u32 diagBit = l + K::DIAG_R0;
            score += (D0 >> diagBit) & 1ull ? 0u : 1u;
            if (!(D0 & RAC)) {
                u32 val = 1u;
                while (val > 0) {
                    if (HP & RAC)
                        val--;
                    if (HN & RAC)
                        val++;
                    if (RAC == (1ull << (diagBit - Wv))) {
                        ok = false;
                        break;
                    }
                    RAC >>= 1;
                }
            }
            if (!ok)
                break;
            bs.ring[(size_t)l * bs.ringStride] = bandPack(HP, M | ~D0, l + sh0);
            if (i >= firstRow)
                eds[i - firstRow] = (u8)edAtCol(i);
        }
        if (firstCentre == 0)
            lc.s[ST_VERIFY_ROWS] += (i <= size) ? i : size;
    }
    const u32 lastRow = i - 1; // number of rows computed successfully
    // cluster centres first (findClusterCenters, bitparallelmatrix.h:591-614), then
    // one traceback per centre
    u8 centres[3 * K::MAX_ED + 4];
    u32 nCentres = 0;
    if (run && !(lastRow <= size - sofc)) { // unsigned compare, as in the reference
        for (u32 r = lastRow; r > firstRow; r--) {
            const u32 ED = eds[r - firstRow];
            if (ED > maxED || ED < minED)
                continue;
            const bool above = ED <= eds[r - 1 - firstRow];
            const bool below = (r == lastRow) || ED <= eds[r + 1 - firstRow];
            if (above && below)
                centres[nCentres++] = (u8)(r - firstRow);
        }
    }
    // this pass traces centres [firstCentre, firstCentre + maxCentres): the lanes
    // of a warp almost always have exactly one, and a lane with more would keep
    // the whole warp waiting for its extra rounds
    totalCentres = nCentres;
    const u32 endCentre = tmin(nCentres, firstCentre + maxCentres);
    warpSync(wm);
    u32 ringBlock = lastRow / K::BLOCK; // the block whose rows the ring holds
    // traceBack from (r, n-1), bitparallelmatrix.h:531-586 (precedence I, diagonal,
    // D), one 32-row block per round of the warp
    u32 ci = firstCentre, r = 0, ti = 0, tj = 0, nops = 0, runOp = 4, runLen = 0;
    bool started = false, over = false, bad = false;
    for (;;) {
        const bool active = ci < endCentre;
        if (!warpAny(wm, active))
            break;
        if (!active)
            continue;
        if (!started) {
            r = firstRow + centres[ci];
            ti = r;
            tj = n - 1;
            nops = 0;
            runOp = 4;
            runLen = 0;
            over = false;
            bad = false;
            started = true;
        }
        const u32 cb = ti / K::BLOCK;
        if (cb != ringBlock && b32) {
            // rows of block cb again, from its checkpoint
            u32 hp = (u32)bs.cp[(size_t)(2 * cb) * bs.cpStride], hn = (u32)bs.cp[(size_t)(2 * cb + 1) * bs.cpStride];
            const i32 w0 = (i32)(cb * K::BLOCK) - 1 - (i32)dB;
            const u64 R0 = seqWindow64(ps, 0, w0), R1 = seqWindow64(ps, 1, w0), R2 = seqWindow64(ps, 2, w0),
                      R3 = seqWindow64(ps, 3, w0);
            u32 i0 = cb * K::BLOCK;
            if (cb == 0) {
                bs.ring[0] = bandPack32(hp, 0, mu);
                i0 = 1;
            }
            const u32 i1 = tmin(cb * K::BLOCK + K::BLOCK - 1, lastRow);
            const u64 tw = (cb == 0) ? textWindow64(ix, start) << 2 : textWindow64(ix, start + cb * K::BLOCK - 1);
            for (u32 q = i0; q <= i1; q++) {
                const u32 l = q % K::BLOCK;
                const u32 M = (u32)(pickM(R0, R1, R2, R3, (u32)(tw >> (2 * l)) & 3u) >> l);
                const u32 D0 = bandRow32(hp, hn, M);
                bs.ring[(size_t)l * bs.ringStride] = bandPack32(hp, M | ~D0, mu);
            }
            ringBlock = cb;
        } else if (cb != ringBlock) {
            // rows of block cb again, from its checkpoint
            u64 hp = bs.cp[(size_t)(2 * cb) * bs.cpStride], hn = bs.cp[(size_t)(2 * cb + 1) * bs.cpStride];
            const u64 R0 = matchVec<u64>(ps, cb, 0), R1 = matchVec<u64>(ps, cb, 1), R2 = matchVec<u64>(ps, cb, 2),
                      R3 = matchVec<u64>(ps, cb, 3);
            u32 i0 = cb * K::BLOCK;
            if (cb == 0) {
                bs.ring[0] = bandPack(hp, 0, sh0);
                i0 = 1;
            }
            const u32 i1 = tmin(cb * K::BLOCK + K::BLOCK - 1, lastRow);
            const u64 tw = (cb == 0) ? textWindow64(ix, start) << 2 : textWindow64(ix, start + cb * K::BLOCK - 1);
            for (u32 q = i0; q <= i1; q++) {
                const u32 l = q % K::BLOCK;
                const u64 M = pickM(R0, R1, R2, R3, (u32)(tw >> (2 * l)) & 3u);
                const u64 D0 = bandRow(hp, hn, M);
                bs.ring[(size_t)l * bs.ringStride] = bandPack(hp, M | ~D0, l + sh0);
            }
            ringBlock = cb;
        }
        do {
            const u64 w = bs.ring[(size_t)(ti % K::BLOCK) * bs.ringStride];
            const u32 bidx = tj + Wv - ti; // cell (ti, tj) within the band of row ti
            if (bidx > Wv + Wh) {
                bad = true;
                break;
            }
            const u32 isI = ((u32)w >> bidx) & 1u;
            const u32 isM = (isI ^ 1u) & ((u32)(w >> 32) >> bidx) & 1u;
            const u32 op = isI ? (u32)CIG_I : (isM ? (u32)CIG_M : (u32)CIG_D);
            tj -= (isI | isM);
            ti -= (isI ^ 1u);
            if (op == runOp) {
                runLen++;
            } else {
                if (runLen) { // the run in registers is complete: one store per run
                    if (nops < opsCap)
                        ops[nops++] = (u16)((runLen << 2) | runOp);
                    else
                        over = true;
                }
                runOp = op;
                runLen = 1;
            }
        } while (tj > 0 && ti / K::BLOCK == cb);
        if (tj == 0 || bad) {
            if (bad)
                atomOr(B.err, ERR_INTERNAL);
            if (runLen) {
                if (nops < opsCap)
                    ops[nops++] = (u16)((runLen << 2) | runOp);
                else
                    over = true;
            }
            if (over)
                nops = 0xFFFFFFFFu;
            else
                cigReverse(ops, nops);
            lc.c[CIGARS_IN_TEXT_VERIFICATION] += mult;
            VerifyHit h;
            h.b = start + ti;
            h.e = start + r;
            h.dist = eds[r - firstRow];
            h.ncig = nops;
            sink(h, ops);
            ci++;
            started = false;
        }
    }
    return nCentres > 0;
}

// stage 3a (edit distance): text position of a candidate -> start of its
// verification window (getBeginPositions, fmindex.h:366-382)
CLB_HD void locateTask(const IndexView& ix, const Batch& B, u32 ci, LocalCtr& lc) {
    const Cand c = B.cands.data[ci];
    const u32 pos = findSA(ix, c.row, lc.s[ST_LF_STEPS]);
    const u32 sum = pos + c.add;
    B.candStart[ci] = sum >= c.sub ? sum - c.sub : 0u;
    lc.s[ST_CANDIDATES]++;
}
// stage 3b: the seeds of a read that lie on the same diagonal lead to the SAME
// verification (same sequence, same window, same matrix): the reference runs
// it once per seed, here the first candidate of such a group within an
// aligned block of 32 candidates is verified once and stands for `mult` of
// them in the counters (the duplicates would produce identical occurrences,
// which the per-read filter removes anyway).
CLB_HD void dedupTask(const Batch& B, u32 ci, u32 n, bool enable, unsigned act) {
    const Cand c = B.cands.data[ci];
    const u32 s = B.candStart[ci];
    bool leader = true;
    u32 mult = 1;
    if (enable) {
        const u32 b0 = ci & ~31u, b1 = tmin(b0 + 32u, n);
        for (u32 j = b0; j < b1; j++) {
            if (j == ci)
                continue;
            const Cand o = B.cands.data[j];
            const bool same = o.t == c.t && B.candStart[j] == s && o.fixedStart == c.fixedStart && o.maxED == c.maxED &&
                              o.minED == c.minED;
            if (same && j < ci)
                leader = false;
            mult += (same && j > ci) ? 1u : 0u;
        }
    }
    const u32 at = B.uniq.warpReserve(leader ? 1u : 0u, act, B.err, ERR_CAND_OVERFLOW);
    if (leader && at != 0xFFFFFFFFu) {
        B.uniq.data[at] = ci;
        B.candMult[ci] = mult;
    }
}

// stage 3c task (edit distance).  pass 0: the candidate's first cluster centre
// (all of them on the 128-bit path); candidates with more centres go to the
// `extra` list, whose entries pass 1 verifies again for the remaining centres.
CLB_HD void verifyEditTask(const IndexView& ix, const Batch& B, u32 ci, VerifyScratch& vs, LocalCtr& lc, u32 pass) {
    const Cand c = B.cands.data[ci];
    const u32 start = B.candStart[ci];
    const u32 mult = B.candMult[ci];
    const u32 maxED = c.maxED, minED = c.minED;
    const u32 nZeros = c.fixedStart ? 1u : 2 * maxED + 1;
    const u32 len = seqLen(B, c.t);
    if (pass == 0)
        lc.c[IN_TEXT_STARTED] += mult;
    const u32 first = pass == 0 ? 0u : 1u;
    CandSink sink(B, c, first);
    bool considered = true, ok;
    // number of rows: Wv + n with Wv = nZeros - 1 + maxED (fmindex.cpp:290-299)
    const u32 nRows = nZeros - 1 + maxED + len + 1;
    const u32 end = tmin(ix.textLength - 1, start + nRows - 1);
    const bool fast = use64(nZeros, maxED);
    const unsigned fm = warpBallot(vs.warpMask, fast);
    if (fast) {
        u32 total = 0;
        ok = verifyWindowFast(ix, B, c.t, vs.bs, start, end, nZeros, maxED, minED, vs.ops, B.cigW, lc, considered,
                              sink, fm, first, pass == 0 ? 1u : 0xFFFFu, total, mult, vs.band32 != 0);
        if (pass == 0 && total > 1)
            B.extra.push(ci, B.err, ERR_CAND_OVERFLOW);
    } else { // (mult is 1: no grouping when 128-bit matrices can occur)
        ok = verifyWindow<u128>(ix, B, c.t, (Row<u128>*)vs.rows, start, end, nZeros, maxED, minED, vs.ops, B.cigW, lc,
                                considered, sink);
    }
    if (pass == 0 && !ok && considered)
        lc.c[ABORTED_IN_TEXT_VERIF] += mult;
    lc.c[TOTAL_REPORTED_POSITIONS] += (u64)(sink.sub - first) * mult;
    lc.s[ST_TEXT_OCCS] += sink.sub - first;
}

// Screening pass of the edit-distance verification.  Nine candidates out of ten are
// aborted after a fraction of their rows (the right-most active column leaves the
// band), at a different row in every lane, so a warp that verifies 32 candidates side
// by side idles most of its lanes.  The screening pass runs only the forward
// recurrence -- no band storage, no checkpoints, no traceback -- as a stepped task
// (start / step / finish): a lane whose candidate is decided takes the next one while
// the others keep stepping (clb_kernel_stepped).  A candidate without a cluster centre
// is settled here (counters as the full verification would count them); one with a
// centre goes to the `surv` list, which the full verification (forward pass with band
// storage + traceback) works off as a dense list of equally long tasks.
struct ScreenState {
    u64 HP, HN, RAC, M0, M1, M2, M3, tcur, tnxt;
    const u64* planes;
    u32 ci, start, size, i, score, mult;
    u32 Wv, firstRow, col, sofc, nw, len;
    u8 maxED, minED, run, considered, ok, done;
    u8 b32, dB; // 32-bit band (see screenStep32): in use, bit of the main diagonal
    u8 eds[3 * BPC<u64>::MAX_ED + 4];
};
enum { SCREEN_ROWS = 8 };
CLB_HD u32 screenEdAtCol(const ScreenState& s, u32 i) {
    typedef BPC<u64> K;
    const u32 bit = (i % K::BLOCK) + K::DIAG_R0, col = s.col;
    const u32 b = (i > col) ? bit - (i - col) + 1 : bit + 1;
    const u32 e = (i > col) ? bit + 1 : bit + (col - i) + 1;
    const u64 mask = ((1ull << (e - b)) - 1ull) << b;
    const int neg = popc64(s.HN & mask), pos = popc64(s.HP & mask);
    return s.score + ((i > col) ? (u32)(neg - pos) : (u32)(pos - neg));
}
// The same recurrence on a 32-bit word that moves with the band.  Bit b of row i is the read position
// i - 1 - dB + b, i.e. the diagonal b - dB (bit dB = the main diagonal, whose value `score` follows): the band
// [-Wv, Wh] sits in the middle of the word with mu >= 1 spare columns on its left and R >= 1 on its right, the
// word moves one column per row (the previous row is re-aligned with a shift by one, the column that enters on
// the right gets a horizontal delta of +1), the right-most active column stays on its bit.  What the two spare
// sides compute is not the true matrix, as little as what the 64-bit window computes outside its band: a value
// that enters the band from the left edge is at least maxED + 2 mu, one from the right edge at least
// maxED + 2 R + 2, and every cell of the band whose true value is <= maxED + 1 has its optimal path within the
// diagonals [-Wv - 1, Wh + 1].  The abort decision, the position of the right-most active column and the
// final-column values <= maxED depend on nothing else, so they equal those of the 64-bit window
// (bitparallelmatrix.h:352-415) -- at half the instructions per row.  tests/test_prep_unit.py compares the two
// forms row by row on random and repeat-rich windows.
CLB_HD u32 screenEdAtCol32(const ScreenState& s, u32 HP, u32 HN, u32 i) {
    const u32 bit = s.dB, col = s.col;
    const u32 b = (i > col) ? bit - (i - col) + 1 : bit + 1;
    const u32 e = (i > col) ? bit + 1 : bit + (col - i) + 1;
    const u32 mask = ((1u << (e - b)) - 1u) << b;
    const int neg = popc32(HN & mask), pos = popc32(HP & mask);
    return s.score + ((i > col) ? (u32)(neg - pos) : (u32)(pos - neg));
}
// The rows of one step are gone through in segments that neither cross a block boundary (where the text window
// and the four match-vector windows are fetched) nor the first row of the final-column zone (from where on the
// distance in the final column is kept per row): the loop over a segment then has no such checks -- per row
// the text character, the match vector, the recurrence, the score and the test of the right-most active column.
template <bool EDS>
CLB_HD bool screenRows32(ScreenState& s, u32& HP, u32& HN, u32& RAC, u32& i, u32 segEnd, u32 dB, u32 stop) {
    u32 score = s.score;
    const u64 tcur = s.tcur, M0 = s.M0, M1 = s.M1, M2 = s.M2, M3 = s.M3;
    bool ok = true;
    for (; i <= segEnd; ++i) {
        const u32 l = i % BPC<u64>::BLOCK;
        const u32 Y = (u32)(tcur >> (2 * l)) & 3u;
        const u32 M = (u32)(pickM(M0, M1, M2, M3, Y) >> l);
        const u32 D0 = bandRow32(HP, HN, M);
        score += ((D0 >> dB) & 1u) ^ 1u;
        if (!(D0 & RAC) && !racWalk(HP, HN, RAC, stop)) {
            ok = false;
            break;
        }
        if (EDS) {
            s.score = score;
            s.eds[i - s.firstRow] = (u8)screenEdAtCol32(s, HP, HN, i);
        }
    }
    s.score = score;
    return ok;
}
CLB_HD bool screenStep32(const IndexView& ix, ScreenState& s) {
    typedef BPC<u64> K;
    u32 i = s.i;
    const u32 last = tmin(s.size, i + SCREEN_ROWS - 1);
    u32 HP = (u32)s.HP, HN = (u32)s.HN, RAC = (u32)s.RAC;
    const u32 dB = s.dB, stop = 1u << (dB - s.Wv);
    while (i <= last) {
        if (i % K::BLOCK == 0) {
            PartSeq ps;
            ps.planes = s.planes;
            ps.nw = s.nw;
            ps.base = 0;
            ps.len = s.len;
            s.tcur = s.tnxt;
            s.tnxt = textWindow64(ix, s.start + i + K::BLOCK - 1);
            const i32 w0 = (i32)i - 1 - (i32)dB;
            s.M0 = seqWindow64(ps, 0, w0);
            s.M1 = seqWindow64(ps, 1, w0);
            s.M2 = seqWindow64(ps, 2, w0);
            s.M3 = seqWindow64(ps, 3, w0);
        }
        u32 segEnd = tmin(last, i | (K::BLOCK - 1));
        bool ok;
        if (i >= s.firstRow) {
            ok = screenRows32<true>(s, HP, HN, RAC, i, segEnd, dB, stop);
        } else {
            segEnd = tmin(segEnd, s.firstRow - 1);
            ok = screenRows32<false>(s, HP, HN, RAC, i, segEnd, dB, stop);
        }
        if (!ok) {
            s.ok = 0;
            break;
        }
    }
    s.HP = HP;
    s.HN = HN;
    s.RAC = RAC;
    s.i = i;
    if (!s.ok || i > s.size)
        s.done = 1;
    return s.done != 0;
}
CLB_HD void screenStart(const IndexView& ix, const Batch& B, u32 ci, ScreenState& s, bool screenBand32) {
    typedef BPC<u64> K;
    const Cand c = B.cands.data[ci];
    s.ci = ci;
    s.start = B.candStart[ci];
    s.mult = B.candMult[ci];
    s.maxED = c.maxED;
    s.minED = c.minED;
    const u32 nZeros = c.fixedStart ? 1u : 2 * (u32)c.maxED + 1;
    const u32 len = seqLen(B, c.t);
    s.done = 0;
    s.ok = 1;
    s.run = 1;
    s.considered = 1;
    if (!use64(nZeros, c.maxED)) { // 128-bit matrices: straight to the full verification
        s.done = 2;
        return;
    }
    const u32 nRows = nZeros - 1 + c.maxED + len + 1;
    const u32 end = tmin(ix.textLength - 1, s.start + nRows - 1);
    const PartSeq ps = fullSeq(B, c.t);
    s.planes = ps.planes;
    s.nw = ps.nw;
    s.len = ps.len;
    const u32 n = len + 1;
    s.Wv = nZeros - 1 + c.maxED;
    const u32 Wh = c.maxED, m = s.Wv + n;
    s.sofc = Wh + s.Wv + 1;
    s.size = end <= s.start ? 0u : end - s.start;
    s.firstRow = (m - 1) - s.sofc;
    s.col = n - 1;
    s.i = 1;
    s.score = 0;
    if (!(s.size >= m - s.sofc)) {
        s.considered = 0;
        s.run = 0;
        s.done = 1;
        return;
    }
    if (s.size / K::BLOCK + 1 > VerifyKernelCpCap(B)) { // the full verification would raise ERR_ROWS_OVERFLOW: let it
        s.done = 2;
        return;
    }
    s.b32 = 0;
    if (screenBand32 && s.Wv + Wh + 1 + 2 <= 32) {
        // band + at least one column on either side fit a 32-bit word: screenStep32
        const u32 mu = (32 - (s.Wv + Wh + 1)) / 2;
        s.b32 = 1;
        s.dB = (u8)(s.Wv + mu);
        s.HP = (u64)((~0u) << (s.dB + 1));                       // read positions >= 0: +1 per column
        s.HN = (u64)((1u << (s.dB + 2 - nZeros)) - 1u);          // left of the free start rows: -1 per column
        s.RAC = (u64)(1u << (s.dB + Wh));
        if (s.firstRow == 0)
            s.eds[0] = (u8)screenEdAtCol32(s, (u32)s.HP, (u32)s.HN, 0);
        const i32 w0 = -1 - (i32)s.dB;
        s.M0 = seqWindow64(ps, 0, w0);
        s.M1 = seqWindow64(ps, 1, w0);
        s.M2 = seqWindow64(ps, 2, w0);
        s.M3 = seqWindow64(ps, 3, w0);
        s.tcur = textWindow64(ix, s.start) << 2;
        s.tnxt = textWindow64(ix, s.start + K::BLOCK - 1);
        if (s.size == 0)
            s.done = 1;
        return;
    }
    s.HP = (~0ull) << K::LEFT;
    s.HN = ~s.HP;
    {
        const u32 nn = tmin<u32>(nZeros, K::LEFT + 1);
        for (u32 i = 1; i < nn; ++i)
            s.HN ^= 1ull << (K::LEFT - i);
    }
    s.RAC = 1ull << (K::DIAG_R0 + Wh);
    if (s.firstRow == 0)
        s.eds[0] = (u8)screenEdAtCol(s, 0);
    s.M0 = matchVec<u64>(ps, 0, 0);
    s.M1 = matchVec<u64>(ps, 0, 1);
    s.M2 = matchVec<u64>(ps, 0, 2);
    s.M3 = matchVec<u64>(ps, 0, 3);
    s.tcur = textWindow64(ix, s.start) << 2;
    s.tnxt = textWindow64(ix, s.start + K::BLOCK - 1);
    if (s.size == 0)
        s.done = 1;
}
// up to SCREEN_ROWS rows; true when the candidate is decided
CLB_HD bool screenStep(const IndexView& ix, ScreenState& s) {
    typedef BPC<u64> K;
    if (s.done)
        return true;
    if (s.b32)
        return screenStep32(ix, s);
    PartSeq ps;
    ps.planes = s.planes;
    ps.nw = s.nw;
    ps.base = 0;
    ps.len = s.len;
    u32 i = s.i;
    const u32 last = tmin(s.size, i + SCREEN_ROWS - 1);
    for (; i <= last; ++i) {
        const u32 l = i % K::BLOCK;
        s.RAC <<= 1;
        if (l == 0) {
            s.tcur = s.tnxt;
            s.tnxt = textWindow64(ix, s.start + i + K::BLOCK - 1);
            s.HP >>= K::BLOCK;
            s.HN >>= K::BLOCK;
            s.RAC >>= K::BLOCK;
            const u32 b = i / K::BLOCK;
            s.M0 = matchVec<u64>(ps, b, 0);
            s.M1 = matchVec<u64>(ps, b, 1);
            s.M2 = matchVec<u64>(ps, b, 2);
            s.M3 = matchVec<u64>(ps, b, 3);
        }
        const u32 Y = (u32)(s.tcur >> (2 * l)) & 3u;
        const u64 M = pickM(s.M0, s.M1, s.M2, s.M3, Y);
        const u64 D0 = bandRow(s.HP, s.HN, M);
        const u32 diagBit = l + K::DIAG_R0;
        s.score += (D0 >> diagBit) & 1ull ? 0u : 1u;
        if (!(D0 & s.RAC)) {
            u32 val = 1u;
            while (val > 0) {
                if (s.HP & s.RAC)
                    val--;
                if (s.HN & s.RAC)
                    val++;
                if (s.RAC == (1ull << (diagBit - s.Wv))) {
                    s.ok = 0;
                    break;
                }
                s.RAC >>= 1;
            }
        }
        if (!s.ok)
            break;
        if (i >= s.firstRow)
            s.eds[i - s.firstRow] = (u8)screenEdAtCol(s, i);
    }
    s.i = i;
    if (!s.ok || i > s.size)
        s.done = 1;
    return s.done != 0;
}
CLB_HD void screenFinish(const Batch& B, ScreenState& s, LocalCtr& lc) {
    bool survivor = s.done == 2;
    u32 rows = 0;
    if (!survivor && s.run) {
        rows = (s.i <= s.size) ? s.i : s.size;
        const u32 lastRow = s.i - 1;
        if (!(lastRow <= s.size - s.sofc)) // unsigned compare, as in the reference
            for (u32 r = lastRow; r > s.firstRow && !survivor; r--) {
                const u32 ED = s.eds[r - s.firstRow];
                if (ED > s.maxED || ED < s.minED)
                    continue;
                const bool above = ED <= s.eds[r - 1 - s.firstRow];
                const bool below = (r == lastRow) || ED <= s.eds[r + 1 - s.firstRow];
                survivor = above && below;
            }
    }
    if (survivor) {
        B.surv.push(s.ci, B.err, ERR_CAND_OVERFLOW);
        return;
    }
    lc.c[IN_TEXT_STARTED] += s.mult;
    if (s.considered)
        lc.c[ABORTED_IN_TEXT_VERIF] += s.mult;
    lc.s[ST_VERIFY_ROWS] += rows;
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
    // 32 characters per round: the read's codes (one byte each, 4 = not ACGT) are packed to the text's
    // 2-bit layout eight at a time, XOR with the text window, one population count per round
    const u64* cw = (const u64*)seqCodes(B, c.t); // seqStride is a multiple of 64: whole words are readable
    u32 score = 0;
    for (u32 j0 = 0; j0 < pSize && score <= c.maxED; j0 += 32) {
        const u64 tw = textWindow64(ix, Tb + j0);
        u64 rd = 0, bad = 0; // read in 2-bit layout; both bits set where the read has no ACGT
#pragma unroll
        for (u32 q = 0; q < 4; q++) {
            const u64 w = cw[j0 / 8 + q];
            u64 x = w & 0x0303030303030303ull;
            x = (x | (x >> 6)) & 0x000F000F000F000Full;
            x = (x | (x >> 12)) & 0x000000FF000000FFull;
            x = (x | (x >> 24)) & 0xFFFFull;
            u64 y = (w >> 2) & 0x0101010101010101ull; // code 4
            y |= y << 1;
            y = (y | (y >> 6)) & 0x000F000F000F000Full;
            y = (y | (y >> 12)) & 0x000000FF000000FFull;
            y = (y | (y >> 24)) & 0xFFFFull;
            rd |= x << (16 * q);
            bad |= y << (16 * q);
        }
        const u32 nc = tmin<u32>(32u, pSize - j0);
        const u64 live = nc == 32 ? ~0ull : ((1ull << (2 * nc)) - 1ull);
        u64 d = (tw ^ rd) | bad;
        d = (d | (d >> 1)) & 0x5555555555555555ull & live;
        score += (u32)popc64(d);
        if (Te == ix.textLength && j0 + nc == pSize && !((d >> (2 * (nc - 1))) & 1ull))
            score++; // the sentinel at the end of the text matches nothing (fmindex.cpp:386-388)
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
