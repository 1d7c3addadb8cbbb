// columba_b200 -- HBM layout of the bidirectional FM-index and the rank /
// extend / locate primitives on it.
//
// Replaces (reference, Vanilla flavour):
//   BitvecIntl<4>::rank            src/bitvec.h:356-372
//   BWTRepresentation::occ/cumOcc  src/fmindex/bwtrepr.h:80-107
//   FMIndex::findRangesWithExtraChar{Backward,Forward}, findRangeWithExtraChar
//   BackwardUniDirectional         src/fmindex/fmindex.cpp:137-243
//   FMIndex::findLF / findSA       src/fmindex/fmindex.cpp:47-60
//   SparseSuffixArray::operator[] / get   src/fmindex/suffixArray.h:131-148
//
// Layout (one record = one 32-byte DRAM sector = one rank lookup):
//   BwtRec  : occ counts of A,C,G,T before the 64-position block (4 x u32) and
//             the 64 symbols as two bit planes (lo,hi) of the 2-bit code.
//             The sentinel '$' is stored as code 0 and corrected by position.
//   SsaRec  : 64 marker bits + number of sampled rows before the block.
//   text    : 2 bits per character, 16 characters per u32.
#pragma once
#include "clb_common.h"

namespace clb {

struct alignas(32) BwtRec {
    u32 cnt[4];
    u64 lo, hi;
};
struct alignas(16) SsaRec {
    u64 bits;
    u32 rank;
    u32 pad;
};

struct BwtView {
    const BwtRec* rec;
    u32 dollarPos;
};

struct IndexView {
    BwtView fwd, rev;
    u32 C[5];        // C[c] = number of characters smaller than c ($ACGT order)
    u32 textLength;  // including the final '$'
    const u32* text; // 2-bit packed
    const SsaRec* ssaRec;
    const u32* ssa;
    const RangePair* kmer; // dense 4^K table (indexinterface.cpp:294-335)
    u32 K;
    const u32* seqStart; // numSeq+1 entries (.pos)
    u32 numSeq;
    u32 switchPoint; // in-text verification switch point
    u32 noCigar;
};

CLB_HD BwtRec loadRec(const BwtRec* p) {
#if defined(__CUDA_ARCH__)
    // one 32-byte sector, read through the non-coherent path as ONE 256-bit load
    // (LDG.E.256, sm_100): the lanes of a warp hit 32 different lines, and the
    // L1 tag stage -- one lookup per lane and load instruction -- is what bounds
    // the rank throughput of a warp
    u64 a, b, lo, hi;
    asm("ld.global.nc.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(a), "=l"(b), "=l"(lo), "=l"(hi) : "l"(p));
    BwtRec r;
    r.cnt[0] = (u32)a; r.cnt[1] = (u32)(a >> 32); r.cnt[2] = (u32)b; r.cnt[3] = (u32)(b >> 32);
    r.lo = lo;
    r.hi = hi;
    return r;
#else
    return *p;
#endif
}

// occ of A,C,G,T in BWT[0,p) out of the record r of p's block
CLB_HD void occ4rec(const BwtRec& r, u32 dollarPos, u32 p, u32 o[4]) {
    u32 off = p & 63;
    u64 m = (1ull << off) - 1ull;
    u64 lo = r.lo & m, hi = r.hi & m;
    u32 nT = popc64(hi & lo), nG = popc64(hi) - nT, nC = popc64(lo) - nT;
    u32 nA = off - nT - nG - nC;
    // '$' is stored as an 'A'
    nA -= ((dollarPos >> 6) == (p >> 6) && dollarPos < p) ? 1u : 0u;
    o[0] = r.cnt[0] + nA;
    o[1] = r.cnt[1] + nC;
    o[2] = r.cnt[2] + nG;
    o[3] = r.cnt[3] + nT;
}
// occ of A,C,G,T in BWT[0,p)
CLB_HD void occ4raw(const BwtRec* rec, u32 dollarPos, u32 p, u32 o[4]) {
    const BwtRec r = loadRec(rec + (p >> 6));
    occ4rec(r, dollarPos, p, o);
}
CLB_HD void occ4(const BwtView& v, u32 p, u32 o[4]) { occ4raw(v.rec, v.dollarPos, p, o); }
// occ at both ends of a range [b, e].  Once a range is narrower than a block
// (after the k-mer seed almost always) both ends sit in the same 32-byte
// record, and the second load -- for scattered ranges a wavefront of the L1
// data pipe per lane, the resource that bounds the extension kernels -- is
// skipped.
CLB_HD void occ4pair(const BwtRec* rec, u32 dollarPos, u32 b, u32 e, u32 ob[4], u32 oe[4]) {
    const BwtRec rb = loadRec(rec + (b >> 6));
    BwtRec re = rb;
    if ((b >> 6) != (e >> 6))
        re = loadRec(rec + (e >> 6));
    occ4rec(rb, dollarPos, b, ob);
    occ4rec(re, dollarPos, e, oe);
}
// symbol (0..3) of BWT[p]; caller handles p == dollarPos
CLB_HD u32 bwtSym(const BwtRec& r, u32 off) {
    return (u32)((r.lo >> off) & 1) | ((u32)((r.hi >> off) & 1) << 1);
}

// Extend a range pair by every character at once.  mode: DIR_FWD uses the BWT
// of the reversed text and the rev range as the driving range; DIR_BWD uses
// the forward BWT and the SA range.  Children with an empty range have
// width 0.  Two rank lookups.
CLB_HD void extendAll(const IndexView& ix, int dir, bool uniBack, const RangePair& p, RangePair ch[4]) {
    const BwtView& bw = (dir == DIR_BWD) ? ix.fwd : ix.rev;
    const Range t = (dir == DIR_BWD) ? p.sa : p.rev;
    u32 ob[4], oe[4];
    occ4pair(bw.rec, bw.dollarPos, t.b, t.e, ob, oe);
    u32 s = (dir == DIR_BWD) ? p.rev.b : p.sa.b;
    u32 x = (t.b <= bw.dollarPos && bw.dollarPos < t.e) ? 1u : 0u;
    for (int c = 0; c < 4; c++) {
        Range r1;
        r1.b = ob[c] + ix.C[c + 1];
        r1.e = oe[c] + ix.C[c + 1];
        u32 w = r1.width();
        Range r2;
        r2.b = s + x;
        r2.e = s + x + w;
        x += oe[c] - ob[c];
        if (uniBack) {
            ch[c].sa = r1;
            ch[c].rev.b = 0;
            ch[c].rev.e = 0;
        } else if (dir == DIR_BWD) {
            ch[c].sa = r1;
            ch[c].rev = r2;
        } else {
            ch[c].sa = r2;
            ch[c].rev = r1;
        }
    }
}
// single character (code 0..3) -- indexinterface.cpp:1034-1049 (addChar).
// Written without branches on dir / uniBack / c: the lanes of a warp extend
// different parts in different directions in the same instruction stream.
CLB_HD bool extendOne(const IndexView& ix, int dir, bool uniBack, u32 c, RangePair& r) {
    const bool bwd = (dir == DIR_BWD);
    const BwtRec* rec = bwd ? ix.fwd.rec : ix.rev.rec;
    const u32 dollar = bwd ? ix.fwd.dollarPos : ix.rev.dollarPos;
    const u32 tb = bwd ? r.sa.b : r.rev.b, te = bwd ? r.sa.e : r.rev.e;
    const u32 s = bwd ? r.rev.b : r.sa.b;
    u32 ob[4], oe[4];
    occ4pair(rec, dollar, tb, te, ob, oe);
    u32 x = (tb <= dollar && dollar < te) ? 1u : 0u;
    x += (c > 0) ? oe[0] - ob[0] : 0u;
    x += (c > 1) ? oe[1] - ob[1] : 0u;
    x += (c > 2) ? oe[2] - ob[2] : 0u;
    const u32 obc = (c == 0) ? ob[0] : (c == 1) ? ob[1] : (c == 2) ? ob[2] : ob[3];
    const u32 oec = (c == 0) ? oe[0] : (c == 1) ? oe[1] : (c == 2) ? oe[2] : oe[3];
    const u32 Cc = (c == 0) ? ix.C[1] : (c == 1) ? ix.C[2] : (c == 2) ? ix.C[3] : ix.C[4];
    const u32 b1 = obc + Cc, e1 = oec + Cc;
    const u32 w = e1 <= b1 ? 0u : e1 - b1;
    const u32 b2 = s + x, e2 = s + x + w;
    r.sa.b = bwd ? b1 : b2;
    r.sa.e = bwd ? e1 : e2;
    r.rev.b = uniBack ? 0u : (bwd ? b2 : b1);
    r.rev.e = uniBack ? 0u : (bwd ? e2 : e1);
    return w != 0;
}

CLB_HD RangePair completeRange(const IndexView& ix) { return mkPair(0, ix.textLength, 0, ix.textLength); }
// fmindex.cpp:434-445
CLB_HD RangePair rangeOfSingleChar(const IndexView& ix, u32 code) {
    if (code > 3)
        return emptyPair();
    u32 b = ix.C[code + 1], e = (code < 3) ? ix.C[code + 2] : ix.textLength;
    return mkPair(b, e, b, e);
}

// locate one SA row: LF-walk to a sampled row (fmindex.cpp:47-60)
CLB_HD u32 findSA(const IndexView& ix, u32 i, u64& lfSteps) {
    u32 l = 0;
    for (;;) {
        SsaRec s = ix.ssaRec[i >> 6];
        u32 off = i & 63;
        if ((s.bits >> off) & 1)
            return ix.ssa[s.rank + popc64(s.bits & ((1ull << off) - 1ull))] + l;
        BwtRec r = loadRec(ix.fwd.rec + (i >> 6));
        u32 c = bwtSym(r, off);
        u64 m = (1ull << off) - 1ull;
        u64 plane = ((c & 1) ? r.lo : ~r.lo) & ((c & 2) ? r.hi : ~r.hi) & m;
        u32 occ = r.cnt[c] + popc64(plane);
        if (c == 0 && (ix.fwd.dollarPos >> 6) == (i >> 6) && ix.fwd.dollarPos < i)
            occ--;
        i = ix.C[c + 1] + occ;
        l++;
        lfSteps++;
    }
}

// 32 characters (2 bits each) of the text starting at position p (any
// alignment); the text array is padded so that p may run up to 64 characters
// past the end
CLB_HD u64 textWindow64(const IndexView& ix, u32 p) {
    const u32 wi = p >> 4, sh = (p & 15) * 2;
    const u32 w0 = ix.text[wi], w1 = ix.text[wi + 1], w2 = ix.text[wi + 2];
    const u32 lo = (u32)((((u64)w1 << 32) | w0) >> sh);
    const u32 hi = (u32)((((u64)w2 << 32) | w1) >> sh);
    return ((u64)hi << 32) | lo;
}

CLB_HD u32 textAt(const IndexView& ix, u32 p) { return (ix.text[p >> 4] >> ((p & 15) * 2)) & 3u; }

} // namespace clb
