// columba_b200 -- banded bit-parallel (Hyyro / Myers) edit distance matrix in
// registers + caller-provided row storage.
//
// Replaces BitParallelED<uint64_t / UInt128>, reference
// src/bitparallelmatrix.h:300-750 and src/bitparallelmatrix.cpp:33-123.
// Differences in organisation (results are bit-identical):
//   * the per-part match vectors (setSequence, .cpp:33-74) are not
//     materialised: a read's four "equal to A/C/G/T" bit planes (forward and
//     reversed) are built once per (read,strand) and the 64/128-bit window of
//     a block is cut out of them with a funnel shift when a row is computed;
//   * rows live wherever the caller puts them (per-thread slabs in HBM).
#pragma once
#include "clb_common.h"

namespace clb {

template <typename W> struct BPC {
    static constexpr u32 WORD = sizeof(W) * 8;
    static constexpr u32 BLOCK = WORD / 2;
    static constexpr u32 MAX_ED = (WORD - BLOCK - 2) / 3; // 10 / 20
    static constexpr u32 LEFT = 2 * MAX_ED + 1;           // 21 / 41
    static constexpr u32 DIAG_R0 = 2 * MAX_ED;
};

// the sequence laid along the columns of a matrix: a part of the read in its
// search direction.  planes -> 4 planes x nw words (nw includes one zero pad
// word) of the read in the direction of the part; base = first bit.
struct PartSeq {
    const u64* planes;
    u32 nw;
    u32 base;
    u32 len;
};

CLB_HD u64 planeBits64(const u64* pl, u32 nw, u32 pos) {
    u32 w = pos >> 6, o = pos & 63;
    if (w >= nw)
        return 0;
    u64 v = pl[w] >> o;
    if (o && w + 1 < nw)
        v |= pl[w + 1] << (64 - o);
    return v;
}
// 64 bits of the "sequence == c" string S_c starting at signed position s,
// where S_c[j<0] = 1, S_c[j>=len] = 0
CLB_HD u64 seqWindow64(const PartSeq& ps, u32 c, i32 s) {
    const u64* pl = ps.planes + (size_t)c * ps.nw;
    if (s >= (i32)ps.len)
        return 0;
    u64 v;
    i32 valid; // number of low bits of v that map to positions < len
    if (s < 0) {
        u32 sh = (u32)(-s);
        if (sh >= 64)
            return ~0ull;
        v = (planeBits64(pl, ps.nw, ps.base) << sh) | ((1ull << sh) - 1ull);
        valid = (i32)ps.len + (i32)sh;
    } else {
        v = planeBits64(pl, ps.nw, ps.base + (u32)s);
        valid = (i32)ps.len - s;
    }
    if (valid < 64)
        v &= (1ull << valid) - 1ull;
    return v;
}
template <typename W> CLB_HD W matchVec(const PartSeq& ps, u32 block, u32 c);
template <> CLB_HD u64 matchVec<u64>(const PartSeq& ps, u32 block, u32 c) {
    return seqWindow64(ps, c, (i32)(block * BPC<u64>::BLOCK) - (i32)BPC<u64>::LEFT);
}
template <> CLB_HD u128 matchVec<u128>(const PartSeq& ps, u32 block, u32 c) {
    i32 s = (i32)(block * BPC<u128>::BLOCK) - (i32)BPC<u128>::LEFT;
    return (u128)seqWindow64(ps, c, s) | ((u128)seqWindow64(ps, c, s + 64) << 64);
}

// The right-most active column after a row whose diagonal successor of the old one is not active any more
// (bitparallelmatrix.h:381-397): the reference walks left from RAC, one column per iteration, with the value
// relative to the bound (val = 1 at RAC; a column with a horizontal +1 lowers it, one with -1 raises it), until
// val reaches 0 -- the new RAC is the column left of that one -- or the left edge of the band (`stop`, a single
// bit <= RAC) has been looked at, which aborts the alignment.  Almost always no -1 column comes before the first
// +1 column: then the walk ends at the highest +1 bit at or below RAC, found with one count of leading zeros
// instead of a divergent loop (the loop was 23 % of the screening kernel's instructions at 3 of 32 lanes).
// Returns false when the band is left.
CLB_HD bool racWalk(u32 HP, u32 HN, u32& RAC, u32 stop) {
    const u32 span = (RAC | (RAC - 1u)) & ~(stop - 1u); // columns stop .. RAC
    const u32 hp = HP & span;
    if (!hp)
        return false;
    const u32 p1 = (u32)msb32(hp);
    if (!(HN & span & ~((2u << p1) - 1u))) { // no -1 between the first +1 and RAC
        if ((1u << p1) == stop)
            return false;
        RAC = 1u << (p1 - 1);
        return true;
    }
    u32 val = 1u;
    while (val > 0) {
        if (HP & RAC)
            val--;
        if (HN & RAC)
            val++;
        if (RAC == stop)
            return false;
        RAC >>= 1;
    }
    return true;
}
CLB_HD bool racWalk(u64 HP, u64 HN, u64& RAC, u64 stop) {
    const u64 span = (RAC | (RAC - 1ull)) & ~(stop - 1ull);
    const u64 hp = HP & span;
    if (!hp)
        return false;
    const u32 p1 = (u32)msb64(hp);
    if (!(HN & span & ~((2ull << p1) - 1ull))) {
        if ((1ull << p1) == stop)
            return false;
        RAC = 1ull << (p1 - 1);
        return true;
    }
    u32 val = 1u;
    while (val > 0) {
        if (HP & RAC)
            val--;
        if (HN & RAC)
            val++;
        if (RAC == stop)
            return false;
        RAC >>= 1;
    }
    return true;
}
CLB_HD bool racWalk(u128 HP, u128 HN, u128& RAC, u128 stop) {
    u32 val = 1u;
    while (val > 0) {
        if (HP & RAC)
            val--;
        if (HN & RAC)
            val++;
        if (RAC == stop)
            return false;
        RAC >>= 1;
    }
    return true;
}

template <typename W> struct Row {
    W HP, HN, D0, RAC;
    u32 score;
};

struct MatGeom {
    u32 n;     // columns = len + 1
    u32 m;     // rows
    u32 Wv, Wh;
    u32 maxED;
    CLB_HD u32 sizeOfFinalColumn() const { return Wh + Wv + 1; }
    CLB_HD bool inFinalColumn(u32 i) const { return i >= m - sizeOfFinalColumn(); }
    CLB_HD u32 firstColumn(u32 i) const { return (i <= Wv) ? 0u : i - Wv; }
};

// bitparallelmatrix.cpp:76-123.  initED may be null (nInit == 0).
template <typename W, typename E>
CLB_HD void bpInit(MatGeom& g, Row<W>* bv, u32 seqLen, u32 maxED, const E* initED, u32 nInit, u32 add) {
    typedef BPC<W> K;
    g.n = seqLen + 1;
    g.maxED = maxED;
    u32 last = nInit ? (u32)initED[nInit - 1] + add : 0;
    g.Wv = nInit ? nInit - 1 + maxED - last : maxED;
    g.m = g.Wv + g.n;
    bv[0].score = nInit ? (u32)initED[0] + add : 0;
    g.Wh = maxED - bv[0].score;
    if (g.Wv + g.Wh + 1 > g.m)
        g.m = g.Wv + g.Wh + 1;
    W HP = (~W(0)) << K::LEFT;
    W HN = ~HP;
    u32 nn = tmin<u32>(nInit, K::LEFT + 1);
    for (u32 i = 1; i < nn; ++i) {
        if (initED[i] < initED[i - 1]) {
            HP ^= W(1) << (K::LEFT - i);
            HN ^= W(1) << (K::LEFT - i);
        } else if (initED[i] == initED[i - 1]) {
            HN ^= W(1) << (K::LEFT - i);
        }
    }
    bv[0].HP = HP;
    bv[0].HN = HN;
    bv[0].D0 = 0;
    bv[0].RAC = W(1) << (K::DIAG_R0 + g.Wh);
}

// bitparallelmatrix.h:352-415; Y = character code 0..3 of the row
template <typename W>
CLB_HD bool bpComputeRow(const MatGeom& g, Row<W>* bv, const PartSeq& ps, u32 i, u32 Y) {
    typedef BPC<W> K;
    const u32 b = i / K::BLOCK, l = i % K::BLOCK;
    const W M = matchVec<W>(ps, b, Y);
    W HP = bv[i - 1].HP, HN = bv[i - 1].HN, RAC = bv[i - 1].RAC << 1u;
    if (l == 0) {
        HP >>= K::BLOCK;
        HN >>= K::BLOCK;
        RAC >>= K::BLOCK;
    }
    W D0 = (((M & HP) + HP) ^ HP) | M | HN;
    W VP = HN | ~(D0 | HP);
    W VN = D0 & HP;
    HP = (VN << 1u) | ~(D0 | (VP << 1u));
    HN = (D0 & (VP << 1u));
    const u32 diagBit = l + K::DIAG_R0;
    bv[i].score = bv[i - 1].score + ((D0 & (W(1) << diagBit)) ? 0u : 1u);
    bool ok = true;
    if (!(D0 & RAC))
        ok = racWalk(HP, HN, RAC, W(1) << (diagBit - g.Wv));
    bv[i].HP = HP;
    bv[i].HN = HN;
    bv[i].D0 = D0;
    bv[i].RAC = RAC;
    return ok;
}

// bitparallelmatrix.h:622-639
template <typename W> CLB_HD u32 bpAt(const Row<W>* bv, u32 i, u32 j) {
    typedef BPC<W> K;
    const u32 bit = (i % K::BLOCK) + K::DIAG_R0;
    u32 b = (i > j) ? bit - (i - j) + 1 : bit + 1;
    u32 e = (i > j) ? bit + 1 : bit + (j - i) + 1;
    W mask = ((W(1) << (e - b)) - W(1)) << b;
    int neg = popcw(bv[i].HN & mask), pos = popcw(bv[i].HP & mask);
    u32 score = bv[i].score;
    score += (i > j) ? (u32)(neg - pos) : (u32)(pos - neg);
    return score;
}
// CIGAR ops are emitted back to front as (len << 2) | op with op 0=M 1=I 2=D;
// the caller reverses them.  Returns the number of runs (0xFFFFFFFF on
// overflow of cap).
enum { CIG_M = 0, CIG_I = 1, CIG_D = 2 };
CLB_HD u32 cigPush(u16* ops, u32 n, u32 cap, u32 op) {
    if (n > 0 && (u32)(ops[n - 1] & 3) == op) {
        ops[n - 1] = (u16)(ops[n - 1] + 4);
        return n;
    }
    if (n >= cap)
        return 0xFFFFFFFFu;
    ops[n] = (u16)((1u << 2) | op);
    return n + 1;
}
CLB_HD void cigReverse(u16* ops, u32 n) {
    for (u32 a = 0, b = n; a + 1 < b; a++) {
        b--;
        u16 t = ops[a];
        ops[a] = ops[b];
        ops[b] = t;
    }
}

// bitparallelmatrix.h:531-586.  refAt(i) = code of the i-th reference char.
template <typename W, typename RefFn>
CLB_HD u32 bpTraceBack(const MatGeom& g, const Row<W>* bv, const PartSeq& ps, RefFn refAt, u32 refEnd,
                       u32& refBegin, u32& ED, u16* ops, u32 cap) {
    typedef BPC<W> K;
    u32 i = refEnd, j = g.n - 1;
    ED = bpAt<W>(bv, i, j);
    u32 n = 0;
    while (j > 0) {
        const u32 b = i / K::BLOCK;
        const W bit = W(1) << ((j - b * K::BLOCK) + K::DIAG_R0);
        u32 op;
        if (bv[i].HP & bit) {
            --j;
            op = CIG_I;
        } else if ((i > 0) && ((matchVec<W>(ps, b, refAt(i - 1)) | ~bv[i].D0) & bit)) {
            --i;
            --j;
            op = CIG_M;
        } else {
            --i;
            op = CIG_D;
        }
        n = cigPush(ops, n, cap, op);
        if (n == 0xFFFFFFFFu)
            return n;
    }
    refBegin = i;
    cigReverse(ops, n);
    return n;
}

// bitparallelmatrix.h:460-526: global alignment of the whole sequence against
// ref[0,refLen) with at most `score` errors; rows must hold refLen+1 entries.
template <typename W, typename RefFn>
CLB_HD u32 bpFindCigar(Row<W>* bv, const PartSeq& ps, RefFn refAt, u32 refLen, u32 score, u16* ops, u32 cap) {
    typedef BPC<W> K;
    MatGeom g;
    bpInit<W, u32>(g, bv, ps.len, score, (const u32*)0, 0, 0);
    for (u32 i = 0; i < refLen; i++)
        bpComputeRow<W>(g, bv, ps, i + 1, refAt(i));
    u32 i = refLen, j = g.n - 1;
    u32 n = 0;
    while (j > 0 || i > 0) {
        const u32 b = i / K::BLOCK;
        const W bit = W(1) << ((j - b * K::BLOCK) + K::DIAG_R0);
        u32 op;
        if ((j > 0) && (bv[i].HP & bit)) {
            --j;
            op = CIG_I;
        } else if (i > 0 && j > 0 && ((matchVec<W>(ps, b, refAt(i - 1)) | ~bv[i].D0) & bit)) {
            --i;
            --j;
            op = CIG_M;
        } else {
            --i;
            op = CIG_D;
        }
        n = cigPush(ops, n, cap, op);
        if (n == 0xFFFFFFFFu)
            return n;
    }
    cigReverse(ops, n);
    return n;
}

} // namespace clb
