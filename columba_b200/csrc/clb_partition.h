// columba_b200 -- stage 0/1: sequence preparation, read partitioning, scheme
// selection, task generation.  One task per (read,strand).
//
// Replaces (reference): SearchStrategy::partition src/searchstrategy.cpp:141-156,
// seed :381-419, partitionDynamic :299-379 (+ seedIfNoKmers :255-271,
// extendParts :283-297), partitionUniform :194-209, partitionOptimalStatic /
// setParts :212-238, calculateExactMatchRanges :158-190,
// MultipleSchemes::createSearches src/searchstrategy.h:2505-2537, and the head
// of matchWithSearches src/searchstrategy.cpp:425-476 (immediate switch to
// in-text verification for narrow parts) and of doRecSearch :1181-1254 (which
// searches start at all).
#pragma once
#include "clb_batch.h"

namespace clb {

CLB_HD u32 asciiCode(u8 c) {
    switch (c) {
    case 'A': case 'a': return 0;
    case 'C': case 'c': return 1;
    case 'G': case 'g': return 2;
    case 'T': case 't': return 3;
    default: return 4;
    }
}

// 64 bytes base[x, x + 64) as 16 little-endian words; bytes outside [0, len)
// read as 0.  On the GPU the bytes come in as aligned 32-bit words (17 loads
// instead of 64 byte loads: the load/tag stage, not DRAM, bounds this kernel)
// re-aligned with funnel shifts.
CLB_HD void loadWindow64(const u8* base, i32 x, u32 len, u32 U[16]) {
#if defined(__CUDA_ARCH__)
    const size_t a0 = (size_t)base + (size_t)(ptrdiff_t)x; // address of byte 0 of the window
    const size_t lo = (size_t)base, hi = (size_t)base + len;
    const size_t ab = a0 & ~(size_t)3;
    const u32 sh = (u32)(a0 & 3) * 8;
    u32 V[17];
#pragma unroll
    for (int m = 0; m < 17; m++) {
        const size_t wa = ab + 4 * (size_t)m;
        V[m] = (wa + 4 > lo && wa < hi) ? __ldg((const u32*)wa) : 0u; // only words that hold a byte of the read
    }
#pragma unroll
    for (int m = 0; m < 16; m++) {
        u32 v = __funnelshift_r(V[m], V[m + 1], sh);
        // zero the bytes outside the read
        const i32 p0 = x + 4 * m;
        u32 keep = 0;
        keep |= (p0 + 0 >= 0 && (u32)(p0 + 0) < len) ? 0x000000FFu : 0u;
        keep |= (p0 + 1 >= 0 && (u32)(p0 + 1) < len) ? 0x0000FF00u : 0u;
        keep |= (p0 + 2 >= 0 && (u32)(p0 + 2) < len) ? 0x00FF0000u : 0u;
        keep |= (p0 + 3 >= 0 && (u32)(p0 + 3) < len) ? 0xFF000000u : 0u;
        U[m] = v & keep;
    }
#else
    for (int m = 0; m < 16; m++) {
        u32 v = 0;
        for (int q = 0; q < 4; q++) {
            const i32 p = x + 4 * m + q;
            if (p >= 0 && (u32)p < len)
                v |= (u32)base[p] << (8 * q);
        }
        U[m] = v;
    }
#endif
}
// A/C/G/T (either case) -> 0..3, anything else (and 0) -> 4
CLB_HD u32 asciiCodeFast(u32 c) {
    const u32 u = c & 0xDFu; // upper case
    const bool ok = (u == 'A') || (u == 'C') || (u == 'G') || (u == 'T');
    return ok ? (((c >> 1) ^ (c >> 2)) & 3u) : 4u;
}

// Four characters per 32-bit word (SWAR).  nzBytes: 0x80 in every byte of z
// that is not zero (exact, no carries between bytes).
CLB_HD u32 nzBytes(u32 z) { return (((z & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | z) & 0x80808080u; }
// code2: per byte the 2-bit code of A/C/G/T (either case); valid: 0x80 in every
// byte that holds one of ACGTacgt (same mapping as asciiCodeFast)
CLB_HD void asciiCodes4(u32 w, u32& code2, u32& valid) {
    const u32 u = w & 0xDFDFDFDFu;
    code2 = ((w >> 1) ^ (w >> 2)) & 0x03030303u;
    const u32 bad = nzBytes(u ^ 0x41414141u) & nzBytes(u ^ 0x43434343u) & nzBytes(u ^ 0x47474747u) &
                    nzBytes(u ^ 0x54545454u);
    valid = bad ^ 0x80808080u;
}
// the four byte-LSBs of x as a nibble at bit 4 * slot of a 32-bit word
// (x * 0x00204081 puts bit 8i at 21 + i; the partial products do not overlap)
CLB_HD u32 gatherNibble(u32 x, int slot) {
    const u32 t = x * 0x00204081u;
    const int sh = 4 * slot;
    return (sh <= 21 ? (t >> (21 - sh)) : (t << (sh - 21))) & (0xFu << sh);
}
CLB_HD u32 byteRev32(u32 v) {
#if defined(__CUDA_ARCH__)
    return __byte_perm(v, 0, 0x0123);
#else
    return __builtin_bswap32(v);
#endif
}
CLB_HD u64 bitRev64(u64 v) {
#if defined(__CUDA_ARCH__)
    return ((u64)__brev((u32)v) << 32) | (u64)__brev((u32)(v >> 32));
#else
    u64 r = 0;
    for (int i = 0; i < 64; i++)
        r |= ((v >> i) & 1ull) << (63 - i);
    return r;
#endif
}
// 64 characters (16 words): bit i of lo / hi / ok = code bit 0 / code bit 1 /
// "is ACGT" of character i; cw = codes as bytes (4 = not ACGT), cmp = the same
// for the complemented codes
template <bool complement> CLB_HD void codes64(const u32 U[16], u64& lo, u64& hi, u64& ok, u32 cw[16]) {
    u32 l[2] = {0, 0}, h[2] = {0, 0}, v[2] = {0, 0};
#pragma unroll
    for (int m = 0; m < 16; m++) {
        u32 c2, val;
        asciiCodes4(U[m], c2, val);
        const u32 vb = val >> 7;
        l[m >> 3] |= gatherNibble(c2 & 0x01010101u, m & 7);
        h[m >> 3] |= gatherNibble((c2 >> 1) & 0x01010101u, m & 7);
        v[m >> 3] |= gatherNibble(vb, m & 7);
        const u32 code = complement ? (c2 ^ 0x03030303u) : c2;
        cw[m] = (code & (vb * 3u)) | ((val ^ 0x80808080u) >> 5);
    }
    lo = (u64)l[0] | ((u64)l[1] << 32);
    hi = (u64)h[0] | ((u64)h[1] << 32);
    ok = (u64)v[0] | ((u64)v[1] << 32);
}

// stage 0: codes of both strands of a read (strand 1 = reverse complement,
// nucleotide.h) and their "equals A/C/G/T" bit planes, forward and reversed.
// One task = 64 characters of one read: the planes of the reverse complement
// are the complement-swapped planes of the read, so both strands come out of
// one pass.  Four characters per instruction: codes and validity per 32-bit
// word, code bits gathered into 64-bit masks, planes derived from the masks.
CLB_HD void prepTask(const Batch& B, u32 task, u32 firstRead) {
    const u32 nc = B.nw - 1; // chunks of 64 characters
    const u32 r = firstRead + task / nc, j = task % nc;
    const u32 off = B.readOff[r], len = B.readOff[r + 1] - off;
    const u32 t0 = 2 * r, t1 = 2 * r + 1;
    const u8* rd = B.readsAscii + off;
    u64* c0 = (u64*)(B.codes + (size_t)t0 * B.seqStride);
    u64* c1 = (u64*)(B.codes + (size_t)t1 * B.seqStride);
    const u32 words = B.seqStride / 8;
    u32 F[16], G[16], cf[16], cg[16];
    loadWindow64(rd, (i32)(64 * j), len, F);                   // read[64j + i]
    loadWindow64(rd, (i32)len - 64 - (i32)(64 * j), len, G);   // G byte 63 - i = read[len - 1 - (64j + i)]
    u64 lo, hi, ok, glo, ghi, gok;
    codes64<false>(F, lo, hi, ok, cf);
    codes64<true>(G, glo, ghi, gok, cg);
    glo = bitRev64(glo);
    ghi = bitRev64(ghi);
    gok = bitRev64(gok);
    const u64 P[4] = {ok & ~hi & ~lo, ok & ~hi & lo, ok & hi & ~lo, ok & hi & lo};
    const u64 Q[4] = {gok & ~ghi & ~glo, gok & ~ghi & glo, gok & ghi & ~glo, gok & ghi & glo};
#pragma unroll
    for (int w = 0; w < 8; w++) {
        const u32 wi = 8 * j + (u32)w;
        if (wi < words) {
            c0[wi] = (u64)cf[2 * w] | ((u64)cf[2 * w + 1] << 32);
            // bytes q = 0..7 of the reverse complement = G bytes 63 - 8w - q
            c1[wi] = (u64)byteRev32(cg[15 - 2 * w]) | ((u64)byteRev32(cg[14 - 2 * w]) << 32);
        }
    }
    u64* p0f = B.planes + ((size_t)t0 * 2 + DIR_FWD) * 4 * B.nw;
    u64* p0b = B.planes + ((size_t)t0 * 2 + DIR_BWD) * 4 * B.nw;
    u64* p1f = B.planes + ((size_t)t1 * 2 + DIR_FWD) * 4 * B.nw;
    u64* p1b = B.planes + ((size_t)t1 * 2 + DIR_BWD) * 4 * B.nw;
#pragma unroll
    for (u32 c = 0; c < 4; c++) {
        p0f[c * B.nw + j] = P[c];
        p0b[c * B.nw + j] = Q[c];
        p1f[c * B.nw + j] = Q[3 - c];
        p1b[c * B.nw + j] = P[3 - c];
        if (j == nc - 1) { // zero pad word
            p0f[c * B.nw + nc] = 0;
            p0b[c * B.nw + nc] = 0;
            p1f[c * B.nw + nc] = 0;
            p1b[c * B.nw + nc] = 0;
        }
    }
}

// indexinterface.h:590-595.  The K byte codes at pat[pos, pos + K) come out of
// (at most) three 8-byte words of the code row (rowWords words, 8-byte aligned).
CLB_HD RangePair lookUpKmer(const IndexView& ix, const u8* pat, u32 pos, u32 rowWords) {
    const u64* src = (const u64*)pat;
    const u32 w0 = pos >> 3, last = rowWords - 1;
    const u64 x0 = src[w0], x1 = src[tmin(w0 + 1, last)], x2 = src[tmin(w0 + 2, last)];
    u32 code = 0, bad = 0;
    for (u32 i = 0; i < ix.K; i++) {
        const u32 q = (pos & 7) + i;
        const u64 xw = (q < 8) ? x0 : ((q < 16) ? x1 : x2);
        const u32 c = (u32)(xw >> ((q & 7) * 8)) & 0xFFu;
        bad |= (c > 3) ? 1u : 0u;
        code = (code << 2) | (c & 3u);
    }
    if (bad)
        return emptyPair();
    return ix.kmer[code];
}
CLB_HD bool addChar(const IndexView& ix, int dir, bool uni, u32 c, RangePair& r, LocalCtr& lc) {
    if (c < 4) {
        lc.s[ST_RANK_LOOKUPS] += 2;
        if (extendOne(ix, dir, uni, c, r)) {
            lc.c[NODE_COUNTER]++;
            return true;
        }
    }
    r = emptyPair();
    return false;
}
// indexinterface.cpp:1016-1032; chars pat[b..e) read in direction dir
CLB_HD RangePair matchString(const IndexView& ix, int dir, bool uni, const u8* pat, u32 b, u32 e, int readDir,
                             RangePair r, LocalCtr& lc) {
    u32 n = e <= b ? 0 : e - b;
    for (u32 i = 0; i < n; i++) {
        u32 c = (readDir == DIR_FWD) ? pat[b + i] : pat[e - 1 - i];
        if (!addChar(ix, dir, uni, c, r, lc))
            break;
    }
    return r;
}

// Working state of one partitioning task: per part the exact range (16 B), the
// greedy key (8 B) and begin / end (2 x 2 B).  The workers of a group (the
// threads of a CTA; 1 in the host emulation) interleave element-wise in three
// arrays [P x stride] so that a range is ONE 128-bit and a key or a begin/end
// ONE 64-bit (begin/end: 32-bit) shared-memory access without bank conflicts.  On the GPU the
// region is shared memory (no local-memory traffic through L2); the host
// emulation uses a slab.
struct PartScratch {
    u8* em;  // RangePair of part i at em + i * stride * 16
    u8* kq;  // key at kq + i * stride * 8
    u8* be;  // (begin, end) as two u16 at be + i * stride * 4
    u32 stride;
    CLB_HD void bind(u8* base, u32 nWorkers, u32 slot, u32 P) {
        stride = nWorkers;
        em = base + (size_t)slot * 16;
        kq = base + (size_t)P * nWorkers * 16 + (size_t)slot * 8;
        be = base + (size_t)P * nWorkers * 24 + (size_t)slot * 4;
    }
    CLB_HD u32* emW(u32 i) const {
#if defined(__CUDA_ARCH__)
        __builtin_assume(__isShared(em)); // the CUDA backend always binds shared memory: LDS/STS, 32-bit addresses
#endif
        return (u32*)(em + i * stride * 16u);
    }
    CLB_HD u32* keyW(u32 i) const {
#if defined(__CUDA_ARCH__)
        __builtin_assume(__isShared(kq));
#endif
        return (u32*)(kq + i * stride * 8u);
    }
    CLB_HD u16* beW(u32 i) const {
#if defined(__CUDA_ARCH__)
        __builtin_assume(__isShared(be));
#endif
        return (u16*)(be + i * stride * 4u);
    }
    CLB_HD u16& b(u32 i) const { return beW(i)[0]; }
    CLB_HD u16& e(u32 i) const { return beW(i)[1]; }
    // begin in the low, end in the high half
    CLB_HD u32 be32(u32 i) const { return *(const u32*)beW(i); }
    CLB_HD RangePair emr(u32 i) const {
#if defined(__CUDA_ARCH__)
        const uint4 v = *(const uint4*)emW(i);
        return mkPair(v.x, v.y, v.z, v.w);
#else
        const u32* w = emW(i);
        return mkPair(w[0], w[1], w[2], w[3]);
#endif
    }
    CLB_HD void setEmr(u32 i, const RangePair& r) const {
#if defined(__CUDA_ARCH__)
        *(uint4*)emW(i) = make_uint4(r.sa.b, r.sa.e, r.rev.b, r.rev.e);
#else
        u32* w = emW(i);
        w[0] = r.sa.b; w[1] = r.sa.e; w[2] = r.rev.b; w[3] = r.rev.e;
#endif
    }
    CLB_HD u32 width(u32 i) const {
        const u64 v = *(const u64*)emW(i); // sa.b, sa.e
        const u32 sb = (u32)v, se = (u32)(v >> 32);
        return se <= sb ? 0u : se - sb;
    }
    CLB_HD u64 key(u32 i) const { return *(const u64*)keyW(i); }
    CLB_HD void setKey(u32 i, u64 k) const { *(u64*)keyW(i) = k; }
    static CLB_HD size_t bytesPerWorker(u32 P) { return (size_t)P * 28; }
};

CLB_HD void calcExactRanges(const IndexView& ix, const u8* pat, u32 rowWords, u32 P, const PartScratch& ps,
                            LocalCtr& lc) {
    const u32 ws = ix.K;
    for (u32 i = 0; i < P; ++i) {
        u32 size = ps.e(i) <= ps.b(i) ? 0 : ps.e(i) - ps.b(i);
        u32 start = ps.b(i) + ((size >= ws) ? ws : 0);
        RangePair init = (size >= ws) ? lookUpKmer(ix, pat, ps.b(i), rowWords) : completeRange(ix);
        ps.setEmr(i, matchString(ix, DIR_FWD, false, pat, start, ps.e(i), DIR_FWD, init, lc));
    }
    u32 lb = ps.b(P - 1), le = ps.e(P - 1);
    u32 size = le <= lb ? 0 : le - lb;
    u32 end = (size >= ws) ? le - ws : le;
    RangePair init = (size >= ws) ? lookUpKmer(ix, pat, end, rowWords) : completeRange(ix);
    ps.setEmr(P - 1, matchString(ix, DIR_BWD, true, pat, lb, end, DIR_BWD, init, lc));
}

// State of one (read,strand) partitioning task.  The dynamic strategy is a
// resumable state machine (start / step / finish) so that the lanes of a warp
// can each be at a different task: a sequence that does not occur in the index
// (e.g. the reverse complement of a read from the forward strand) runs out of
// extendable parts after ~20 of its ~100 steps, and its lane takes the next
// task while the other lanes keep stepping.
struct PartState {
    PartScratch ps;
    // the sequence as 2-bit codes, 16 per word, in shared memory (word j at
    // patS[j * patStride]); null, or patN set (the sequence holds a character
    // other than ACGT): read the byte codes from global memory
    u32* patS;
    u32 patStride;
    u32 patN;
    const u8* pat;
    u32 rowWords; // 8-byte words per code row
    u32 t, pSize, j;
    u32 stepping; // 1: greedy steps outstanding
};

// width x weight of part i while it can still grow, else 0
CLB_HD u64 partKey(const DevStrategy& st, const PartScratch& ps, u32 P, u32 i) {
    const bool noL = (i == 0) || ps.b(i) == ps.e(i - 1);
    const bool noR = (i == P - 1) || ps.e(i) == ps.b(i + 1);
    return (noL && noR) ? 0ull : (u64)ps.width(i) * st.weights[i];
}

// seed(), searchstrategy.cpp:381-419 (+ seedIfNoKmers :255-271)
CLB_HD void dynSeed(const IndexView& ix, const DevStrategy& st, u32 P, PartState& S, LocalCtr& lc) {
    const PartScratch& ps = S.ps;
    const u8* pat = S.pat;
    const u32 pSize = S.pSize;
    const bool useK = ((u64)P * ix.K < ((u64)pSize * 2) / 3) && pSize >= st.useKmerCutOff;
    const int wSize = useK ? (int)ix.K : 1;
    ps.b(0) = 0;
    for (u32 i = 1; i + 1 < P; i++)
        ps.b(i) = (u32)(int)dsub(dmul(st.seeds[i - 1], (double)pSize), (double)(wSize / 2));
    for (u32 i = 0; i + 1 < P; i++)
        ps.e(i) = ps.b(i) + wSize;
    ps.b(P - 1) = pSize - wSize;
    ps.e(P - 1) = pSize;
    for (u32 i = 0; i < P; i++)
        ps.setEmr(i, useK ? lookUpKmer(ix, pat, ps.b(i), S.rowWords) : rangeOfSingleChar(ix, pat[ps.b(i)]));
    u32 matched = P * (u32)wSize;
    if (matched == 0) { // seedIfNoKmers, :255-271
        u32 size = (pSize >= 4 * P) ? 4 : 1;
        for (u32 i = 0; i + 1 < P; i++) {
            ps.e(i) = tmin(ps.e(i) + size, pSize);
            RangePair r = ps.emr(i);
            for (u32 j = ps.b(i); j < ps.e(i); j++) {
                addChar(ix, DIR_FWD, false, pat[j], r, lc);
                matched++;
            }
            ps.setEmr(i, r);
        }
    }
    for (u32 i = 0; i < P; i++)
        ps.setKey(i, partKey(st, ps, P, i));
    S.j = matched;
}

// One step of the greedy loop of partitionDynamic (:340-376): only the
// extended part and the neighbour it grew towards change, so a step is one
// argmax over P cached keys, one single-character extension and two key
// updates.  Returns true when the partitioning is complete.
CLB_HD bool dynStep(const IndexView& ix, const DevStrategy& st, u32 P, PartState& S, LocalCtr& lc) {
    const PartScratch& ps = S.ps;
    if (S.j >= S.pSize)
        return true;
    u64 maxW = 0;
    u32 p = 0;
    for (u32 i = 0; i < P; i++) {
        const u64 w = ps.key(i);
        p = (w > maxW) ? i : p;
        maxW = (w > maxW) ? w : maxW;
    }
    if (maxW == 0) { // extendParts, :283-297
        for (u32 i = 0; i < P; i++) {
            if ((i != P - 1) && (ps.e(i) != ps.b(i + 1)))
                ps.e(i) = ps.b(i + 1);
            if ((i != 0) && (ps.b(i) != ps.e(i - 1)))
                ps.b(i) = ps.e(i - 1);
        }
        return true;
    }
    const u32 pl = (p == 0) ? 0 : p - 1, pr = (p == P - 1) ? p : p + 1;
    const u32 beP = ps.be32(p), beL = ps.be32(pl), beR = ps.be32(pr);
    const u32 pb = beP & 0xFFFFu, pe = beP >> 16;
    const u32 el = beL >> 16, br = beR & 0xFFFFu; // end of the left / begin of the right neighbour
    const bool noL = (p == 0) || pb == el;
    const bool noR = (p == P - 1) || pe == br;
    // both neighbours exist when neither side is closed
    const u32 wl = ps.width(pl), wr = ps.width(pr);
    const int dir = noL ? DIR_FWD : (noR ? DIR_BWD : ((wl < wr) ? DIR_BWD : DIR_FWD));
    const bool fwd = (dir == DIR_FWD);
    const u32 cpos = fwd ? pe : pb - 1;
    u32 c;
    if (S.patS && !S.patN) {
#if defined(__CUDA_ARCH__)
        __builtin_assume(__isShared(S.patS));
#endif
        c = (S.patS[(cpos >> 4) * S.patStride] >> ((cpos & 15) * 2)) & 3u;
    } else {
        c = S.pat[cpos];
    }
    const u32 nb = fwd ? pb : pb - 1, ne = fwd ? pe + 1 : pe;
    ps.beW(p)[fwd ? 1 : 0] = (u16)(fwd ? ne : nb);
    RangePair r = ps.emr(p);
    addChar(ix, dir, p == P - 1, c, r, lc);
    ps.setEmr(p, r);
    // keys: the extended part, and the neighbour it grew towards once they touch
    // (a part's key is width x weight until it is closed on both sides)
    const bool cL = (p == 0) || nb == el, cR = (p == P - 1) || ne == br;
    ps.setKey(p, (cL && cR) ? 0ull : (u64)r.width() * st.weights[p]);
    // (no branch on the direction: lanes extend different parts in different directions)
    const u32 q = fwd ? pr : pl;
    const bool touch = fwd ? (p != P - 1 && ne == br) : (p != 0 && nb == el);
    const u32 q2 = fwd ? ((q == P - 1) ? q : q + 1) : ((q == 0) ? q : q - 1); // the neighbour's other neighbour
    const u32 beQ2 = ps.be32(q2);
    // (q != p whenever `touch` holds, so the neighbour's begin / end read above are current)
    const bool closed = fwd ? (q == P - 1 || (beR >> 16) == (beQ2 & 0xFFFFu)) : (q == 0 || (beL & 0xFFFFu) == (beQ2 >> 16));
    if (touch && closed)
        ps.setKey(q, 0ull);
    S.j++;
    return S.j >= S.pSize;
}

CLB_HD void storeParts(const PartStore& st, u32 t, const PartScratch& ps, u32 P) {
    for (u32 i = 0; i < P; i++) {
        st.be[(size_t)i * st.T + t] = (ps.b(i) & 0xFFFFu) | (ps.e(i) << 16);
        st.emr[(size_t)i * st.T + t] = ps.emr(i);
    }
}
CLB_HD void loadParts(const PartStore& st, u32 t, const PartScratch& ps, u32 P) {
    for (u32 i = 0; i < P; i++) {
        const u32 v = st.be[(size_t)i * st.T + t];
        ps.b(i) = v & 0xFFFFu;
        ps.e(i) = v >> 16;
        ps.setEmr(i, st.emr[(size_t)i * st.T + t]);
    }
}

// Stage 1 runs as three kernels: seeding (one thread per (read,strand), full
// warps), the greedy extension steps (stepped kernel with per-lane refill) and
// the emission of candidates and search tasks (one thread per (read,strand)).
// The parts travel between them in PartInfo.
//
// stage 1a: everything up to the greedy steps.
CLB_HD void partStart(const IndexView& ix, const DevStrategy& st, const Batch& B, u32 t, PartState& S, LocalCtr& lc) {
    S.t = t;
    S.pat = seqCodes(B, t);
    S.rowWords = B.seqStride / 8;
    S.pSize = seqLen(B, t);
    S.stepping = 0;
    const u32 pSize = S.pSize;
    const u32 P = st.P;
    PartHdr h;
    h.len = (u16)pSize;
    h.matched = 0;
    h.scheme = 0;
    h.valid = 0;
    h.stepping = 0;
    h.pad = 0;
    if (P >= pSize || P == 1) { // searchstrategy.cpp:141-156: no partitioning possible
        atomOr(B.err, ERR_SHORT_READ);
        B.readStatus[t >> 1] = 1; // both strands write the same value
        B.parts.hdr[t] = h;
        return;
    }
    h.valid = 1;
    const PartScratch& ps = S.ps;
    if (B.partStrategy == PART_DYNAMIC) {
        dynSeed(ix, st, P, S, lc);
        S.stepping = 1;
        h.stepping = 1;
        h.matched = (u16)S.j;
        B.parts.hdr[t] = h;
        storeParts(B.parts, t, ps, P);
        return;
    }
    B.parts.hdr[t] = h;
    if (B.partStrategy == PART_UNIFORM) { // :194-209
        for (u32 i = 0; i < P; i++) {
            ps.b(i) = (u32)dmul(ddiv(dmul((double)(int)i, 1.0), (double)P), (double)pSize);
            ps.e(i) = (u32)dmul(ddiv(dmul((double)(int)(i + 1), 1.0), (double)P), (double)pSize);
        }
        ps.e(P - 1) = pSize;
    } else { // :212-238
        ps.b(0) = 0;
        ps.e(0) = (u32)dmul(st.begins[0], (double)(int)pSize);
        for (u32 i = 0; i + 2 < P; i++) {
            ps.b(i + 1) = (u32)dmul(st.begins[i], (double)(int)pSize);
            ps.e(i + 1) = (u32)dmul(st.begins[i + 1], (double)(int)pSize);
        }
        ps.b(P - 1) = (u32)dmul(st.begins[P - 2], (double)(int)pSize);
        ps.e(P - 1) = pSize;
    }
    calcExactRanges(ix, S.pat, S.rowWords, P, ps, lc);
    storeParts(B.parts, t, ps, P);
}
// stage 1b: the greedy steps of a task seeded by partStart
CLB_HD void greedyStart(const DevStrategy& st, const Batch& B, u32 t, PartState& S) {
    const PartHdr pi = B.parts.hdr[t];
    S.t = t;
    S.stepping = 0;
    if (!pi.valid || !pi.stepping)
        return;
    S.pat = seqCodes(B, t);
    S.pSize = pi.len;
    S.j = pi.matched;
    S.patN = 0;
    if (S.patS) { // 8 byte codes -> 8 x 2 bits (the code rows are 8-byte aligned and padded)
        const u64* src = (const u64*)S.pat;
        const u32 nwords = (S.pSize + 7) / 8;
        u64 bad = 0;
        u32 acc = 0;
        for (u32 j = 0; j < nwords; j++) {
            const u64 v = src[j];
            const u32 rest = S.pSize - 8 * j; // characters of the sequence in this word (>= 1)
            bad |= v & 0x0404040404040404ull & (rest >= 8 ? ~0ull : ((1ull << (8 * rest)) - 1ull));
            u64 x = v & 0x0303030303030303ull;
            x = (x | (x >> 6)) & 0x000F000F000F000Full;
            x = (x | (x >> 12)) & 0x000000FF000000FFull;
            x = (x | (x >> 24)) & 0xFFFFull;
            acc |= (u32)x << (16 * (j & 1));
            if ((j & 1) || j + 1 == nwords) {
                S.patS[(j >> 1) * S.patStride] = acc;
                acc = 0;
            }
        }
        S.patN = bad ? 1u : 0u;
    }
    loadParts(B.parts, t, S.ps, st.P);
    for (u32 i = 0; i < st.P; i++)
        S.ps.setKey(i, partKey(st, S.ps, st.P, i));
    S.stepping = 1;
}
CLB_HD void greedyFinish(const DevStrategy& st, const Batch& B, PartState& S) {
    PartHdr& pi = B.parts.hdr[S.t];
    if (pi.valid && pi.stepping) {
        storeParts(B.parts, S.t, S.ps, st.P);
        pi.stepping = 0;
    }
}
// true when the task has no (more) greedy steps to do
CLB_HD bool partStep(const IndexView& ix, const DevStrategy& st, PartState& S, LocalCtr& lc) {
    if (!S.stepping)
        return true;
    if (dynStep(ix, st, st.P, S, lc))
        S.stepping = 0;
    return !S.stepping;
}

// stage 1c: emits the immediate-switch candidates (order slot 0) and the
// search tasks of the selected scheme.
// `act`: the lanes of the warp that run this function together.
CLB_HD void partFinish(const IndexView& ix, const DevStrategy& st, const Batch& B, u32 t, PartState& S, LocalCtr& lc,
                       unsigned act) {
    const u32 P = st.P;
    PartHdr& pi = B.parts.hdr[t];
    const bool valid = pi.valid != 0;
    const PartScratch& ps = S.ps;
    if (valid)
        loadParts(B.parts, t, ps, P);
    // slots for all immediate-switch candidates of the warp at once
    u32 nCand = 0;
    if (valid)
        for (u32 i = 0; i < P; i++) {
            const u32 width = ps.width(i);
            if (width != 0 && width <= ix.switchPoint)
                nCand += width;
        }
    u32 at = B.cands.warpReserve(nCand, act, B.err, ERR_CAND_OVERFLOW);
    if (!valid)
        return;
    // A) immediate switch, searchstrategy.cpp:464-476 -> fmindex.cpp:245-265 /
    //    :344-356
    u32 emit = 0;
    for (u32 i = 0; i < P; i++) {
        u32 width = ps.width(i);
        if (width != 0 && width <= ix.switchPoint) {
            Cand c;
            c.t = t;
            c.maxED = (u8)B.k;
            c.minED = 0; // mapRead maps with minimal distance 0 and drops what is below minD after the filter
            c.pad = 0;
            c.add = 0;
            if (B.metric == METRIC_EDIT) {
                lc.c[IMMEDIATE_SWITCH]++;
                u32 startDiff = (ps.b(i) == 0) ? 0 : ps.b(i) + B.k;
                c.sub = startDiff;
                c.fixedStart = (ps.b(i) == 0);
            } else {
                c.sub = ps.b(i);
                c.fixedStart = 1;
            }
            for (u32 row = 0; row < width; row++) {
                c.row = ps.emW(i)[0] + row;
                c.order = emit++;
                if (at != 0xFFFFFFFFu)
                    B.cands.data[at++] = c;
            }
        }
    }
    // B) scheme selection, searchstrategy.h:2505-2537
    u32 sel = 0;
    if (st.nSchemes > 1) {
        u32 total = 0;
        for (u32 i = 0; i < P; i++)
            total += ps.width(i);
        if (total > P) {
            u32 minValue = ps.width(st.sc[0].criticalPart);
            for (u32 i = 1; i < st.nSchemes; i++) {
                u32 w = ps.width(st.sc[i].criticalPart);
                if (w < minValue) {
                    minValue = w;
                    sel = i;
                }
            }
        }
    }
    pi.scheme = (u8)sel;
    // which searches start (doRecSearch, searchstrategy.cpp:1181-1254)
    const DevScheme& sc = st.sc[sel];
    for (u32 s = 0; s < sc.nSearches; s++) {
        const DevSearch& se = sc.s[s];
        bool start = se.U[0] > 0 || ps.width(se.order[0]) > ix.switchPoint;
        if (start) {
            SearchTask task;
            task.t = t;
            task.search = s;
            B.tasks.push(task, B.err, ERR_TASK_OVERFLOW);
        }
    }
}

} // namespace clb
