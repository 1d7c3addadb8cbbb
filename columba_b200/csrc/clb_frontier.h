// columba_b200 -- stage 2: search-scheme exploration over the bidirectional
// FM-index as a level-synchronous frontier, one warp per (read, strand,
// search) task.
//
// What it computes is what the reference's depth-first recursion computes
// (SearchStrategy::doRecSearch src/searchstrategy.cpp:1181-1254,
// IndexInterface::recApproxMatchEditEntry src/indexinterface.cpp:1306-1325,
// recApproxMatchEdit :377-527, branchAndBound :529-561, goDeeper :563-669,
// goToInTextVerificationEdit :340-375, extendFMPos :675-697, the cluster of the
// final column src/indexhelpers.h:1677-1838 with getClusterCentra
// src/indexhelpers.cpp:276-382, recApproxMatchHamming
// src/indexinterface.cpp:1211-1304): the same set of index nodes (NODE_COUNTER),
// the same occurrences and in-text candidates, in the same emission order.
// How it is computed is different:
//
//  * A task is a chain of at most P phase groups.  Phase group q holds every
//    instance of phase q of the search (one per hand-over from phase q-1) and
//    is expanded level by level: the frontier of a level is a contiguous run of
//    64-byte records (SA range pair, the bit-parallel matrix row HP/HN/RAC/score
//    of the node, parent link, phase instance), each lane of the warp expands
//    one record (2 rank lookups -> at most 4 children, one matrix row per
//    child), and the surviving children are appended in depth-first order with
//    a warp prefix sum (order-preserving stream compaction).  A node's state
//    depends on its ancestors only, so the set of nodes equals the reference's.
//  * The cluster of final-column values the reference keeps per recursion
//    frame is read off the parent links (at most Wv + Wh + 1 ancestors).
//  * The reference's depth-first ORDER matters twice: "report once" flags
//    (src/indexhelpers.h:1613-1629) give a cluster node to the first leaf in
//    depth-first order that reports it, and the emission order of occurrences
//    breaks ties in the per-read filter.  Both are recovered exactly: the
//    records of a level are in depth-first order, two scans over the stored
//    levels (subtree leaf counts up, offsets down) give every leaf its
//    depth-first rank, report-once becomes an atomic minimum over ranks, and
//    emission indices are prefix sums over the rank-ordered leaves, composed
//    across the phase groups at the end of the task.
//  * Nothing is written to the batch's lists before the task is complete, so a
//    task whose frontier outgrows the warp's scratch is simply handed to a
//    second launch with large scratch (B.heavy).
#pragma once
#include "clb_batch.h"
#include "clb_matrix.h"
#include "clb_partition.h"
#include "clb_simt.h"

namespace clb {

// indexhelpers.h:1353-1535 (FMOcc) as the hand-over between phases
struct LocOcc {
    RangePair r;
    u32 depth, dist, shift;
    u32 valid;
};

template <typename W> struct RowS {
    W HP, HN, RAC;
    u32 score;
};

// One row of the banded bit-parallel matrix from its predecessor (Hyyro's
// recurrences on a sliding window of WORD columns; rows of a block of BLOCK
// rows share the window).  Returns false when no cell of the row is within the
// bound (the right-most active column RAC ran out of the band).
// M = match vector of the row's character within the window of the row's block
template <typename W>
CLB_HD bool rowStepM(const MatGeom& g, const RowS<W>& p, u32 i, W M, RowS<W>& o) {
    typedef BPC<W> K;
    const u32 l = i % K::BLOCK;
    W HP = p.HP, HN = p.HN, RAC = p.RAC << 1u;
    if (l == 0) {
        HP >>= K::BLOCK;
        HN >>= K::BLOCK;
        RAC >>= K::BLOCK;
    }
    const W D0 = (((M & HP) + HP) ^ HP) | M | HN;
    const W VP = HN | ~(D0 | HP);
    const W VN = D0 & HP;
    HP = (VN << 1u) | ~(D0 | (VP << 1u));
    HN = D0 & (VP << 1u);
    const u32 diagBit = l + K::DIAG_R0;
    o.score = p.score + (((D0 >> diagBit) & W(1)) ? 0u : 1u);
    bool ok = true;
    if (!(D0 & RAC))
        ok = racWalk(HP, HN, RAC, W(1) << (diagBit - g.Wv));
    o.HP = HP;
    o.HN = HN;
    o.RAC = RAC;
    return ok;
}
template <typename W>
CLB_HD bool rowStep(const MatGeom& g, const PartSeq& ps, const RowS<W>& p, u32 i, u32 Y, RowS<W>& o) {
    return rowStepM<W>(g, p, i, matchVec<W>(ps, i / BPC<W>::BLOCK, Y), o);
}
// value of cell (i, j) from the row's horizontal deltas
template <typename W> CLB_HD u32 cellAt(const RowS<W>& r, u32 i, u32 j) {
    typedef BPC<W> K;
    const u32 bit = (i % K::BLOCK) + K::DIAG_R0;
    const u32 b = (i > j) ? bit - (i - j) + 1 : bit + 1;
    const u32 e = (i > j) ? bit + 1 : bit + (j - i) + 1;
    const W mask = ((W(1) << (e - b)) - W(1)) << b;
    const int neg = popcw(r.HN & mask), pos = popcw(r.HP & mask);
    return r.score + ((i > j) ? (u32)(neg - pos) : (u32)(pos - neg));
}
template <typename W> CLB_HD bool onlyVerticalLeft(const MatGeom& g, const RowS<W>& r, u32 i) {
    typedef BPC<W> K;
    if (i + K::LEFT < g.n)
        return false;
    const u32 b = i / K::BLOCK, q = i % K::BLOCK;
    const u32 bb = K::DIAG_R0 - g.Wv + q + 1;
    const u32 be = K::DIAG_R0 + g.n - b * K::BLOCK;
    return (((~r.HN >> bb) << bb) << (K::WORD - be)) == W(0);
}

enum { FK_ALIVE = 0, FK_IT = 1, FK_GD = 2, FK_DEAD = 3, FK_HAM_END = 4 };
enum { FF_CHAIN = 1, FF_USE_SM = 2, FF_FROM_CHAIN = 4 };
enum { F_NONE = 0xFFFFFFFFu, F_ED_NONE = 63 };

// one index node: frontier record and tree node in one
template <typename W> struct alignas(16) FNode {
    RangePair r;
    W HP, HN, RAC;
    u32 score;   // edit: score of the row's diagonal cell; Hamming: mismatches so far in this phase
    u32 parent;  // node index within the phase group, F_NONE for the root of a phase instance
    u32 cnt;     // leaves with an action below this node; after the ranking: report-once mark (first rank)
    u32 base;    // depth-first rank of the first such leaf
    u32 pi;      // phase instance within the group
    u16 row;
    u16 meta;    // c (2 bits) | kind FK_* (3) | flags FF_* (3) | ed (6): value in the final column, F_ED_NONE
                 // when the row is not in the final-column zone
    CLB_HD u32 c() const { return meta & 3u; }
    CLB_HD u32 kind() const { return (meta >> 2) & 7u; }
    CLB_HD u32 flags() const { return (meta >> 5) & 7u; }
    CLB_HD u32 ed() const { return (meta >> 8) & 63u; }
    CLB_HD void setMeta(u32 c, u32 kind, u32 flags, u32 ed) { meta = (u16)(c | (kind << 2) | (flags << 5) | (ed << 8)); }
    CLB_HD void setKind(u32 kind) { meta = (u16)((meta & ~(7u << 2)) | (kind << 2)); }
};

// descendants handed from a phase to the next (src/indexinterface.cpp:613-668)
struct FDesc {
    RangePair r;
    u32 c;
    u32 pad;
};

// one instance of a phase: what the reference's recursion frame holds
struct FPI {
    LocOcc sm;
    MatGeom g;
    u32 clSize;
    u32 descOff, nDesc, ieOff, nIe;     // the descendants / initial eds of this direction (byte offsets in the scratch)
    u32 oDescOff, nDescO, oIeOff, nIeO; // those of the other direction
    u32 actBegin;                       // rank of the instance's first leaf action in the group
    u32 total;                          // emissions of the instance and everything it hands over to
    u32 base;                           // emission index of its first emission
    u32 sBegin;
};

// a leaf with an action, in depth-first order within its phase group
struct FAct {
    u32 node;
    u32 nEv;    // emissions of this leaf itself
    u32 spawn;  // phase instance of the next group that this leaf hands over to, or F_NONE
    u32 off;    // emission index of the leaf's first emission / prefix sum scratch
    u64 mask;   // end phase: cluster indices that this leaf reports
    u32 pi;     // the leaf's phase instance and kind: the numbering passes work on the action list alone and
    u32 kind;   // do not chase the node records (random 64-byte reads in a scratch far larger than the L2)
};

struct FGroup {
    u32 idx;      // phase
    u32 nodeOff;  // byte offsets into the scratch
    u32 nNodes;
    u32 piOff, nPI;
    u32 actOff, nActs;
    u32 lvlOff, nLevels;
};

struct FCounters {
    u64 nodes, lookups, rows, fmOccs;
};

template <typename W> struct FrontierCaps {
    static constexpr u32 CL = 3 * BPC<W>::MAX_ED + 2; // Wv + Wh + 1 <= CL
    static constexpr u32 LS = 2 * CL;
};

// The search of one task by one group of GS lanes: a warp (GS = 32), or a whole CTA of GS threads for the
// tasks whose trees are too large for a warp to finish in time.  `scr` is the group's scratch.
template <typename W, int GS> struct Frontier {
    typedef FNode<W> Node;
    static constexpr u32 CL = FrontierCaps<W>::CL, LS = FrontierCaps<W>::LS;

    const IndexView& ix;
    const Batch& B;
    const DevSearch& s;
    const PartInfo& pi;
    u32 t, searchIdx;
    u8* scr;
    u32 cap, top; // bump allocator (bytes)
    u32 err;      // ERR_* bits; ERR_STACK_OVERFLOW = scratch exhausted (retry as a heavy task)
    bool edit;
    FGroup grp[MAXP];
    u32 nGroups;
    bool rankPrefetch; // the ranking passes ask for the next level's records one level ahead
    u32 heavyWidth; // first launch: start ranges wider than this go to the heavy launch at once (0: never)
    LV<FCounters> ctr;

    CLB_HD Frontier(const IndexView& ix_, const Batch& B_, const DevSearch& s_, const PartInfo& pi_, u32 t_, u32 si,
                    u8* scr_, u32 cap_)
        : ix(ix_), B(B_), s(s_), pi(pi_), t(t_), searchIdx(si), scr(scr_), cap(cap_), top(0), err(0),
          edit(B_.metric == METRIC_EDIT), nGroups(0), rankPrefetch(false), heavyWidth(0) {
        CLB_LANES_G(l, GS) {
            ctr[l].nodes = 0;
            ctr[l].lookups = 0;
            ctr[l].rows = 0;
            ctr[l].fmOccs = 0;
        }
    }

    CLB_HD u32 alloc(size_t bytes) {
        const size_t a = ((size_t)top + 15) & ~(size_t)15;
        if (a + bytes > cap) {
            err |= ERR_STACK_OVERFLOW;
            return F_NONE;
        }
        top = (u32)(a + bytes);
        return (u32)a;
    }
    CLB_HD Node* nodesOf(const FGroup& g) const { return (Node*)(scr + g.nodeOff); }
    CLB_HD FPI* pisOf(const FGroup& g) const { return (FPI*)(scr + g.piOff); }
    CLB_HD FAct* actsOf(const FGroup& g) const { return (FAct*)(scr + g.actOff); }
    CLB_HD u32* lvlOf(const FGroup& g) const { return (u32*)(scr + g.lvlOff); }
    CLB_HD u32 partLen(u32 p) const { return pi.end[p] <= pi.begin[p] ? 0u : (u32)(pi.end[p] - pi.begin[p]); }
    CLB_HD PartSeq partSeq(u32 p, u32 dir) const {
        PartSeq ps;
        ps.planes = seqPlanes(B, t, dir);
        ps.nw = B.nw;
        ps.len = partLen(p);
        ps.base = (dir == DIR_FWD) ? pi.begin[p] : (u32)pi.len - pi.end[p];
        return ps;
    }
    CLB_HD u32 partChar(u32 p, u32 dir, u32 j) const {
        const u8* codes = seqCodes(B, t);
        return (dir == DIR_FWD) ? codes[pi.begin[p] + j] : codes[pi.end[p] - 1 - j];
    }
    CLB_HD u32 orderKey(u32 emit) const { return ((searchIdx + 1) << 27) | (emit & 0x7FFFFFFu); }
    CLB_HD u32 maxRows() const { return (u32)pi.len + 4 * BPC<W>::MAX_ED + 8; }

    // ------------------------------------------------------------------ phase instances
    // first column and geometry of a phase instance (recApproxMatchEdit prologue,
    // src/indexinterface.cpp:377-470, BitParallelED::initializeMatrix
    // src/bitparallelmatrix.cpp:76-123); writes the instance's root node
    CLB_HD void initInstance(const FGroup& G, u32 p, const PartSeq& ps) {
        FPI& P = pisOf(G)[p];
        Node& n = nodesOf(G)[p];
        const u32 idx = G.idx;
        n.r = P.sm.r;
        n.parent = F_NONE;
        n.cnt = 0;
        n.base = 0;
        n.pi = p;
        n.row = 0;
        n.setMeta(0, FK_ALIVE, 0, F_ED_NONE);
        n.HP = n.HN = n.RAC = W(0);
        if (!edit) {
            n.score = P.sm.dist;
            P.g.n = ps.len + 1;
            P.g.m = ps.len + 1;
            P.g.Wv = P.g.Wh = 0;
            P.g.maxED = s.U[idx];
            P.clSize = 0;
            return;
        }
        const u32 maxED = s.U[idx];
        const u16* ie = (const u16*)(scr + P.ieOff);
        u16 single[1];
        u32 add = 0, nInit = P.nIe;
        if (nInit == 0) {
            single[0] = (u16)P.sm.dist;
            ie = single;
            nInit = 1;
        } else {
            u32 prevED = ie[0];
            if (s.dsw[idx])
                for (u32 i = 1; i < nInit; i++)
                    prevED = tmin<u32>(prevED, ie[i]);
            add = P.sm.dist - prevED;
        }
        Row<W> row0;
        bpInit<W, u16>(P.g, &row0, ps.len, maxED, ie, nInit, add);
        P.clSize = P.g.sizeOfFinalColumn();
        if (P.clSize > CL || P.g.m > maxRows()) {
            n.setKind(FK_DEAD);
            atomOr(B.err, ERR_CLUSTER_OVERFLOW);
            return;
        }
        n.HP = row0.HP;
        n.HN = row0.HN;
        n.RAC = row0.RAC;
        n.score = row0.score;
        u32 ed0 = F_ED_NONE, fl = 0, kind = FK_ALIVE;
        if (P.g.inFinalColumn(0)) {
            RowS<W> r0;
            r0.HP = n.HP;
            r0.HN = n.HN;
            r0.RAC = n.RAC;
            r0.score = n.score;
            ed0 = tmin<u32>(cellAt<W>(r0, 0, ps.len), 62u);
        }
        if (P.nDesc > 0) {
            const u32 maxRow = P.g.m - 1;
            if (P.nDesc > maxRow) {
                kind = FK_DEAD;
                atomOr(B.err, ERR_INTERNAL);
            } else {
                fl = FF_CHAIN;
            }
        }
        n.setMeta(0, kind, fl, ed0);
    }

    // classification of a freshly computed edit-distance node (branchAndBound
    // + the in-text switch of the reference's node loop)
    CLB_HD u32 classify(const MatGeom& g, const RowS<W>& rs, bool valid, u32 row, u32 width, bool chain, u32 idx,
                        u32& edOut) {
        edOut = F_ED_NONE;
        if (g.inFinalColumn(row)) {
            edOut = cellAt<W>(rs, row, g.n - 1);
            if (edOut > 62)
                edOut = 62;
            if (!valid || onlyVerticalLeft<W>(g, rs, row))
                return FK_GD;
        }
        if (!valid)
            return 0xFFu; // silent
        if (!chain && width <= ix.switchPoint && idx != 0)
            return FK_IT;
        return FK_ALIVE;
    }

    // expand the levels of group G until no node is alive
    // what a lane keeps of the node it expands between the two passes over its children
    struct Expand {
        RowS<W> prs;     // matrix row of the node
        RangePair ch[4]; // ranges of its children (a node on a chain: ch[0] = the next descendant)
        const FPI* P;
        u32 i, row, pi;
        u32 mode;        // 0: nothing to expand, 1: next descendant of a chain, 2: index extension
        u32 chainC;
        u32 sv;          // children that survive, bit q = q-th child in the reference's order
    };
    static CLB_HD RangePair pick4(const RangePair ch[4], u32 c) {
        const RangePair lo = (c & 1u) ? ch[1] : ch[0], hi = (c & 1u) ? ch[3] : ch[2];
        return (c & 2u) ? hi : lo;
    }

    // Levels of group G until no node is alive.  A lane expands one node of the level: two rank lookups give
    // the ranges of its children, one matrix row each decides which of them survive.  The children of a chunk
    // of GS nodes are appended in depth-first order, which needs the number of surviving children of the
    // lanes before (one prefix sum) before anything can be written: the children are therefore gone through
    // twice -- once to count, once to build and store them at their final place -- instead of being parked
    // between the two (four 64-byte records per lane, dynamically indexed, used to be the bulk of the
    // kernel's local-memory traffic).  The loop body is written once for both passes and for the chain /
    // extension / Hamming variants, which keeps the hot code within the instruction cache.
    CLB_HD void expandGroup(FGroup& G, const PartSeq& ps) {
        const u32 idx = G.idx, dir = s.dir[idx], uni = s.isUniBack(idx) ? 1u : 0u, maxED = s.U[idx];
        const u32 part = s.order[idx];
        Node* nodes = nodesOf(G);
        u32* lvl = lvlOf(G);
        const u32 lvlCap = maxRows() + 2;
        u32 b0 = 0, b1 = G.nPI, level = 0;
        lvl[0] = 0;
        lvl[1] = b1;
        // Every node of a level sits on the same row of its instance's matrix (the roots are level 0), all
        // instances of the group lay the same part along the columns, and the match vectors of a character are
        // cut per block of BLOCK rows: the four vectors that the rows of this level's children need are the same
        // for every node of the level and change once per BLOCK levels.  They are cut out of the read's bit
        // planes here, once, instead of once per child and pass.
        W M0 = W(0), M1 = W(0), M2 = W(0), M3 = W(0);
        u32 mBlock = F_NONE;
        while (b1 > b0 && !err) {
            if ((edit || level == 0) && (level + 1) / BPC<W>::BLOCK != mBlock) {
                mBlock = (level + 1) / BPC<W>::BLOCK;
                M0 = matchVec<W>(ps, mBlock, 0);
                M1 = matchVec<W>(ps, mBlock, 1);
                M2 = matchVec<W>(ps, mBlock, 2);
                M3 = matchVec<W>(ps, mBlock, 3);
            }
            u32 outBase = b1; // children are appended to the node array
            for (u32 base = b0; base < b1 && !err; base += GS) {
                LV<u32> nKids, off;
                LV<Expand> st;
                u32 total = 0;
#pragma unroll 1
                for (u32 pass = 0; pass < 2; pass++) {
                    CLB_LANES_G(l, GS) {
                        Expand& S = st[l];
                        if (pass == 0) {
                            nKids[l] = 0;
                            off[l] = 0;
                            S.mode = 0;
                            S.sv = 0;
                            const u32 i = base + (u32)l;
                            if (i >= b1)
                                continue;
                            const Node x = nodes[i];
                            if (x.kind() != FK_ALIVE)
                                continue;
                            const FPI* P = pisOf(G) + x.pi;
                            S.P = P;
                            S.i = i;
                            S.pi = x.pi;
                            S.row = (u32)x.row + 1;
                            S.prs.HP = x.HP;
                            S.prs.HN = x.HN;
                            S.prs.RAC = x.RAC;
                            S.prs.score = x.score;
                            if (x.flags() & FF_CHAIN) {
                                // the next node is the next descendant handed over by the previous phase
                                const FDesc d = ((const FDesc*)(scr + P->descOff))[x.row];
                                S.ch[0] = d.r;
                                S.chainC = d.c;
                                S.mode = 1;
                            } else {
                                // extendFMPos: all four children with two rank lookups
                                const RangePair parentRange = (x.flags() & FF_USE_SM) ? P->sm.r : x.r;
                                extendAll(ix, (int)dir, uni != 0, parentRange, S.ch);
                                ctr[l].lookups += 2;
                                S.mode = 2;
                            }
                        }
                        if (S.mode == 0)
                            continue;
                        const FPI& P = *S.P;
                        const bool chain = S.mode == 1, count = pass == 0;
                        const u32 row = S.row;
                        u32 at = outBase + off[l];
                        const u32 nq = chain ? 1u : 4u;
#pragma unroll 1
                        for (u32 q = 0; q < nq; q++) {
                            if (!count && !((S.sv >> q) & 1u))
                                continue;
                            // children in the order the reference's stack pops them (T, G, C, A)
                            const u32 cc = chain ? S.chainC : 3u - q;
                            const RangePair r = chain ? S.ch[0] : pick4(S.ch, cc);
                            if (!chain) {
                                if (r.empty())
                                    continue;
                                if (count)
                                    ctr[l].nodes++;
                            }
                            RowS<W> rs;
                            rs.HP = rs.HN = rs.RAC = W(0);
                            rs.score = 0;
                            u32 kind, ked = F_ED_NONE, kfl = 0;
                            if (chain || edit) {
                                if ((!chain && row >= P.g.m) || row != level + 1) {
                                    if (count)
                                        atomOr(B.err, ERR_INTERNAL);
                                    continue;
                                }
                                const W Mlo = (cc & 1u) ? M1 : M0, Mhi = (cc & 1u) ? M3 : M2;
                                const bool valid = rowStepM<W>(P.g, S.prs, row, (cc & 2u) ? Mhi : Mlo, rs);
                                if (count)
                                    ctr[l].rows++;
                                kind = classify(P.g, rs, valid, row, chain ? 0xFFFFFFFFu : r.width(), chain, idx, ked);
                                if (kind == 0xFFu)
                                    continue;
                                if (chain) {
                                    kfl = FF_FROM_CHAIN;
                                    if (kind == FK_ALIVE) {
                                        if (row < P.nDesc) {
                                            kfl |= FF_CHAIN;
                                        } else if (P.nDesc == P.g.m - 1) {
                                            kind = FK_DEAD; // the last descendant sits on the last row
                                        } else if (s.dsw[idx]) {
                                            kfl |= FF_USE_SM; // after a switch the children come from the start range
                                        }
                                    }
                                }
                            } else { // recApproxMatchHamming, src/indexinterface.cpp:1211-1304
                                if (r.width() <= ix.switchPoint) {
                                    kind = FK_IT;
                                    rs.score = S.prs.score;
                                } else {
                                    const u32 v = S.prs.score + ((cc != partChar(part, dir, row - 1)) ? 1u : 0u);
                                    rs.score = v;
                                    if (v > maxED)
                                        continue;
                                    if (row == ps.len) {
                                        if (v < s.L[idx])
                                            continue;
                                        kind = FK_HAM_END;
                                    } else {
                                        kind = FK_ALIVE;
                                    }
                                }
                            }
                            if (count) {
                                S.sv |= 1u << q;
                                nKids[l]++;
                                continue;
                            }
                            Node k;
                            k.r = r;
                            k.HP = rs.HP;
                            k.HN = rs.HN;
                            k.RAC = rs.RAC;
                            k.score = rs.score;
                            k.parent = S.i;
                            k.cnt = chain ? (kind == FK_GD ? 1u : 0u) : (kind != FK_ALIVE ? 1u : 0u);
                            k.base = 0;
                            k.pi = S.pi;
                            k.row = (u16)row;
                            k.setMeta(cc & 3u, kind, kfl, ked);
                            // the node array of the group is the last allocation: it grows in place
                            nodes[at] = k;
                            at++;
                        }
                    }
                    if (pass == 0) {
                        total = grpExScanG<GS>(nKids, off);
                        if (!total)
                            break;
                        (void)alloc((size_t)total * sizeof(Node));
                        if (err)
                            break;
                    }
                }
                if (err)
                    break;
                outBase += total;
            }
            if (err)
                break;
            // the children become visible to the lanes that expand them: once per level (the chunks of a level
            // touch disjoint records; a barrier per chunk was a third of the CTA launch's barriers)
            grpSyncG<GS>();
            level++;
            if (level + 1 >= lvlCap) {
                err |= ERR_INTERNAL;
                break;
            }
            b0 = b1;
            b1 = outBase;
            lvl[level + 1] = b1;
        }
        G.nNodes = b1;
        G.nLevels = level + 1; // levels 0 .. level, lvl[level+1] = number of nodes
    }

    // The same expansion for groups of at most a warp, with one lane per (node, child) pair: lane l of a chunk
    // takes the child q = l % 4 (in the reference's order T, G, C, A) of the node base + l / 4.  The light tasks
    // that warps run have narrow levels (ncu: 9.5 of 32 lanes had a node, 5.9 a child to compute, and half of
    // the kernel's issue slots were in use): with a lane per pair a level of up to GS / 4 nodes computes all its
    // children's matrix rows in ONE pass over the loop body instead of four, every lane holds exactly one
    // child, so the record stays in registers between the prefix sum and the store -- no second pass that
    // builds the rows again -- and the prefix sum over the lanes is the depth-first order as before.  The four
    // lanes of a node read the same two rank records (one L1 request) and keep the child of their character.
    CLB_HD void expandGroupPairs(FGroup& G, const PartSeq& ps) {
        const u32 idx = G.idx, dir = s.dir[idx], uni = s.isUniBack(idx) ? 1u : 0u, maxED = s.U[idx];
        const u32 part = s.order[idx];
        constexpr u32 NS = (u32)GS / 4u; // nodes per chunk
        Node* nodes = nodesOf(G);
        u32* lvl = lvlOf(G);
        const u32 lvlCap = maxRows() + 2;
        u32 b0 = 0, b1 = G.nPI, level = 0;
        lvl[0] = 0;
        lvl[1] = b1;
        W M0 = W(0), M1 = W(0), M2 = W(0), M3 = W(0); // match vectors of the level's row (see expandGroup)
        u32 mBlock = F_NONE;
        while (b1 > b0 && !err) {
            if ((edit || level == 0) && (level + 1) / BPC<W>::BLOCK != mBlock) {
                mBlock = (level + 1) / BPC<W>::BLOCK;
                M0 = matchVec<W>(ps, mBlock, 0);
                M1 = matchVec<W>(ps, mBlock, 1);
                M2 = matchVec<W>(ps, mBlock, 2);
                M3 = matchVec<W>(ps, mBlock, 3);
            }
            u32 outBase = b1;
            for (u32 base = b0; base < b1 && !err; base += NS) {
                LV<u32> has, off;
                LV<Node> kid;
                CLB_LANES_G(l, GS) {
                    has[l] = 0;
                    const u32 i = base + ((u32)l >> 2), q = (u32)l & 3u;
                    if (i >= b1)
                        continue;
                    const Node x = nodes[i];
                    if (x.kind() != FK_ALIVE)
                        continue;
                    const FPI& P = pisOf(G)[x.pi];
                    const u32 row = (u32)x.row + 1;
                    const bool chain = (x.flags() & FF_CHAIN) != 0;
                    RangePair r;
                    u32 cc;
                    if (chain) {
                        // the next node is the next descendant handed over by the previous phase
                        if (q != 0)
                            continue;
                        const FDesc d = ((const FDesc*)(scr + P.descOff))[x.row];
                        r = d.r;
                        cc = d.c;
                    } else {
                        // extendFMPos: the two rank lookups serve all four children, this lane computes its own
                        r = (x.flags() & FF_USE_SM) ? P.sm.r : x.r;
                        if (q == 0)
                            ctr[l].lookups += 2;
                        cc = 3u - q;
                        if (!extendOne(ix, (int)dir, uni != 0, cc, r))
                            continue;
                        ctr[l].nodes++;
                    }
                    RowS<W> rs;
                    rs.HP = rs.HN = rs.RAC = W(0);
                    rs.score = 0;
                    u32 kind, ked = F_ED_NONE, kfl = 0;
                    if (chain || edit) {
                        if ((!chain && row >= P.g.m) || row != level + 1) {
                            atomOr(B.err, ERR_INTERNAL);
                            continue;
                        }
                        RowS<W> prs;
                        prs.HP = x.HP;
                        prs.HN = x.HN;
                        prs.RAC = x.RAC;
                        prs.score = x.score;
                        const W Mlo = (cc & 1u) ? M1 : M0, Mhi = (cc & 1u) ? M3 : M2;
                        const bool valid = rowStepM<W>(P.g, prs, row, (cc & 2u) ? Mhi : Mlo, rs);
                        ctr[l].rows++;
                        kind = classify(P.g, rs, valid, row, chain ? 0xFFFFFFFFu : r.width(), chain, idx, ked);
                        if (kind == 0xFFu)
                            continue;
                        if (chain) {
                            kfl = FF_FROM_CHAIN;
                            if (kind == FK_ALIVE) {
                                if (row < P.nDesc) {
                                    kfl |= FF_CHAIN;
                                } else if (P.nDesc == P.g.m - 1) {
                                    kind = FK_DEAD; // the last descendant sits on the last row
                                } else if (s.dsw[idx]) {
                                    kfl |= FF_USE_SM; // after a switch the children come from the start range
                                }
                            }
                        }
                    } else { // recApproxMatchHamming, src/indexinterface.cpp:1211-1304
                        if (r.width() <= ix.switchPoint) {
                            kind = FK_IT;
                            rs.score = x.score;
                        } else {
                            const u32 v = x.score + ((cc != partChar(part, dir, row - 1)) ? 1u : 0u);
                            rs.score = v;
                            if (v > maxED)
                                continue;
                            if (row == ps.len) {
                                if (v < s.L[idx])
                                    continue;
                                kind = FK_HAM_END;
                            } else {
                                kind = FK_ALIVE;
                            }
                        }
                    }
                    Node& k = kid[l];
                    k.r = r;
                    k.HP = rs.HP;
                    k.HN = rs.HN;
                    k.RAC = rs.RAC;
                    k.score = rs.score;
                    k.parent = i;
                    k.cnt = chain ? (kind == FK_GD ? 1u : 0u) : (kind != FK_ALIVE ? 1u : 0u);
                    k.base = 0;
                    k.pi = x.pi;
                    k.row = (u16)row;
                    k.setMeta(cc & 3u, kind, kfl, ked);
                    has[l] = 1;
                }
                const u32 total = grpExScanG<GS>(has, off);
                if (!total)
                    continue;
                (void)alloc((size_t)total * sizeof(Node)); // the node array of the group is the last allocation
                if (err)
                    break;
                CLB_LANES_G(l, GS) {
                    if (has[l])
                        nodes[outBase + off[l]] = kid[l];
                }
                outBase += total;
            }
            if (err)
                break;
            grpSyncG<GS>(); // the children become visible to the lanes that expand them: once per level
            level++;
            if (level + 1 >= lvlCap) {
                err |= ERR_INTERNAL;
                break;
            }
            b0 = b1;
            b1 = outBase;
            lvl[level + 1] = b1;
        }
        G.nNodes = b1;
        G.nLevels = level + 1; // levels 0 .. level, lvl[level+1] = number of nodes
    }

    // depth-first ranks of the leaves with an action: subtree counts up the
    // stored levels, offsets down; fills the group's action list
    CLB_HD void rankLeaves(FGroup& G) {
        Node* nodes = nodesOf(G);
        const u32* lvl = lvlOf(G);
        for (u32 L = G.nLevels; L-- > 1;) {
            const u32 b0 = lvl[L], b1 = lvl[L + 1];
            for (u32 base = b0; base < b1; base += GS) {
                CLB_LANES_G(l, GS) {
                    const u32 i = base + (u32)l;
                    if (i < b1) {
                        // the level above is read next, after the sync: ask for its records now (the scratch of
                        // a heavy task is far larger than the L2, the passes over the levels wait on HBM)
                        if (rankPrefetch && L > 1 && lvl[L - 1] + (i - b0) < b0)
                            prefetchL2(&nodes[lvl[L - 1] + (i - b0)].parent);
                        const u32 c = nodes[i].cnt;
                        if (c)
                            atomAdd<u32>(&nodes[nodes[i].parent].cnt, c);
                    }
                }
            }
            grpSyncG<GS>();
        }
        // roots in instance order
        u32 run = 0;
        FPI* pis = pisOf(G);
        for (u32 base = 0; base < G.nPI; base += GS) {
            LV<u32> c, off;
            CLB_LANES_G(l, GS) {
                const u32 i = base + (u32)l;
                c[l] = i < G.nPI ? nodes[i].cnt : 0u;
            }
            const u32 tot = grpExScanG<GS>(c, off);
            CLB_LANES_G(l, GS) {
                const u32 i = base + (u32)l;
                if (i < G.nPI) {
                    nodes[i].base = run + off[l];
                    pis[i].actBegin = run + off[l];
                }
            }
            run += tot;
        }
        G.nActs = run;
        grpSyncG<GS>();
        G.actOff = alloc((size_t)(run ? run : 1) * sizeof(FAct));
        if (err)
            return;
        FAct* acts = actsOf(G);
        for (u32 L = 1; L < G.nLevels; L++) {
            const u32 b0 = lvl[L], b1 = lvl[L + 1];
            for (u32 base = b0; base < b1; base += GS) {
                CLB_LANES_G(l, GS) {
                    const u32 i = base + (u32)l;
                    if (i < b1) {
                        if (rankPrefetch && L + 1 < G.nLevels && b1 + (i - b0) < lvl[L + 2])
                            prefetchL2(&nodes[b1 + (i - b0)].parent);
                        const u32 p = nodes[i].parent;
                        u32 b = nodes[p].base;
                        for (u32 q = 1; q <= 3 && i >= b0 + q && nodes[i - q].parent == p; q++)
                            b += nodes[i - q].cnt;
                        nodes[i].base = b;
                    }
                }
            }
            grpSyncG<GS>();
        }
        for (u32 base = 0; base < G.nNodes; base += GS) {
            CLB_LANES_G(l, GS) {
                const u32 i = base + (u32)l;
                if (i < G.nNodes) {
                    const Node& n = nodes[i];
                    if (n.kind() != FK_ALIVE && n.kind() != FK_DEAD && n.cnt == 1) {
                        FAct a;
                        a.node = i;
                        a.nEv = n.kind() == FK_IT ? n.r.sa.width() : 0u; // one candidate per row of the range
                        a.spawn = F_NONE;
                        a.off = 0;
                        a.mask = 0;
                        a.pi = n.pi;
                        a.kind = n.kind();
                        acts[n.base] = a;
                    }
                }
            }
        }
        grpSyncG<GS>();
    }

    // the cluster of final-column values above a leaf: entry j belongs to row
    // m - clSize + j of the instance (MatrixMetaInfo, src/indexhelpers.h:1677-1727)
    struct Cluster {
        u16 eds[CL];
        u32 node[CL];
        i32 last;
    };
    CLB_HD void walkCluster(const FGroup& G, const FPI& P, u32 leaf, Cluster& C) const {
        const Node* nodes = nodesOf(G);
        u32 x = leaf;
        C.last = (i32)(P.clSize + nodes[leaf].row) - (i32)P.g.m;
        for (;;) {
            const i32 ci = (i32)(P.clSize + nodes[x].row) - (i32)P.g.m;
            if (ci < 0)
                break;
            C.eds[ci] = (u16)nodes[x].ed();
            C.node[ci] = x;
            if (nodes[x].parent == F_NONE)
                break;
            x = nodes[x].parent;
        }
    }
    static CLB_HD bool isCentre(const Cluster& C, i32 i, u32 maxED) {
        return C.eds[i] <= maxED && (i == 0 || C.eds[i] <= C.eds[i - 1]) && (i == C.last || C.eds[i] <= C.eds[i + 1]);
    }
    // deepest minimum of the cluster (reportDeepestMinimum, src/indexhelpers.h:1770-1798)
    static CLB_HD void deepestMin(const Cluster& C, u32 maxED, u32& minED, i32& hi, i32& deep) {
        minED = maxED + 1;
        hi = deep = -1;
        for (i32 i = 0; i <= C.last; i++) {
            if (C.eds[i] < minED) {
                minED = C.eds[i];
                hi = i;
                deep = i;
            }
            if (C.eds[i] == minED)
                deep = i;
        }
    }
    // first centre at or above the lower bound and the initial eds handed to the
    // next phase (getClusterCentra, src/indexhelpers.cpp:276-382).  Returns the
    // centre's cluster index or -1; ie[0..ni) = eds from the centre downwards
    // after the correction of clusters below the lower bound.
    static CLB_HD i32 centreAndEds(const Cluster& C, u32 lowerBound, u32 maxED, u16* ie, u32& ni) {
        ni = 0;
        for (i32 i = 0; i <= C.last; i++) {
            if (C.eds[i] > maxED || C.eds[i] < lowerBound)
                continue;
            if (!((i == 0 || C.eds[i] <= C.eds[i - 1]) && (i == C.last || C.eds[i] <= C.eds[i + 1])))
                continue;
            for (i32 j = i; j <= C.last; j++)
                ie[ni++] = C.eds[j];
            for (u32 k = 1; k < ni; k++) {
                if (!(ie[k] < lowerBound && ie[k] <= ie[k - 1] && (k == ni - 1 || ie[k] <= ie[k + 1])))
                    continue;
                // a centre below the lower bound: its slopes are straightened
                u32 up = 0, down = ni - 1;
                for (u32 q = k; q-- > 0;)
                    if (ie[q] != ie[q + 1] + 1) {
                        up = q + 1;
                        break;
                    }
                for (u32 q = k + 1; q < ni; q++)
                    if (ie[q] != ie[q - 1] + 1) {
                        down = q - 1;
                        break;
                    }
                if (up != 0 && down != ni - 1) {
                    u32 lo = down, hiq = up;
                    bool fromTop = true;
                    while (lo > hiq) {
                        if (fromTop) {
                            ie[hiq] = (u16)tmin<int>((int)maxED + 1, (int)ie[hiq - 1] + 1);
                            hiq++;
                        } else {
                            ie[lo] = (u16)tmin<int>((int)maxED + 1, (int)ie[lo + 1] + 1);
                            lo--;
                        }
                        fromTop = !fromTop;
                    }
                    if (lo == hiq)
                        ie[lo] = (u16)tmin<int>((int)ie[lo + 1] + 1, (int)ie[lo - 1] + 1);
                } else if (up == 0 && down != ni - 1) {
                    for (u32 q = down; q-- > 0;)
                        ie[q] = (u16)(ie[q + 1] + 1);
                } else if (up != 0 && down == ni - 1) {
                    for (u32 q = up; q < ni; q++)
                        ie[q] = (u16)(ie[q - 1] + 1);
                }
            }
            return i;
        }
        return -1;
    }

    // what a goDeeper leaf does, as far as it can be said without the
    // report-once marks: `rep` = cluster indices it wants to report once
    // (end phase: every centre; other edge phases: the deepest minimum)
    struct GdPlan {
        u64 rep;     // cluster indices reported with the once flag
        i32 centre;  // inner phase: the centre handed over (-1: none)
        i32 hi, deep;
        u32 minED;
    };
    CLB_HD void planGd(const Cluster& C, u32 idx, GdPlan& pl) const {
        const u32 maxED = s.U[idx];
        pl.rep = 0;
        pl.centre = -1;
        pl.hi = pl.deep = -1;
        pl.minED = maxED + 1;
        if (s.isEdge(idx)) {
            if (idx + 1 == s.P) {
                for (i32 i = 0; i <= C.last; i++)
                    if (isCentre(C, i, maxED))
                        pl.rep |= 1ull << i;
            } else {
                deepestMin(C, maxED, pl.minED, pl.hi, pl.deep);
                if (pl.minED <= maxED)
                    pl.rep = 1ull << pl.deep;
            }
        }
    }

    // start difference of the in-text verification of a node
    // (goToInTextVerificationEdit, src/indexinterface.cpp:340-375)
    CLB_HD u32 startDiffOf(const FGroup& G, const FPI& P, const Node& n, bool& fixedStart) const {
        const u32 idx = G.idx;
        const u32 stp = pi.begin[s.lowestBefore(idx)];
        u32 startDiff = stp + s.maxED();
        fixedStart = (stp == 0);
        if (stp == 0)
            return 0;
        if (s.dir[idx] == DIR_BWD) {
            RowS<W> rs;
            rs.HP = n.HP;
            rs.HN = n.HN;
            rs.RAC = n.RAC;
            rs.score = n.score;
            const u32 row = n.row, col = P.g.firstColumn(row);
            startDiff -= col + cellAt<W>(rs, row, col);
        } else if (P.nDescO > 0) {
            const u16* ieO = (const u16*)(scr + P.oIeOff);
            startDiff -= (P.nDescO - P.nIeO + (u32)ieO[P.nIeO - 1]);
        }
        return startDiff;
    }

    // actions of group G: report-once resolution, emission counts, and the
    // phase instances of the next group.  Returns the next group's instance count.
    CLB_HD u32 actGroup(FGroup& G, FGroup& N) {
        const u32 idx = G.idx, nIdx = idx + 1;
        Node* nodes = nodesOf(G);
        FPI* pis = pisOf(G);
        FAct* acts = actsOf(G);
        const u32 nA = G.nActs;
        const bool isEnd = (nIdx == s.P);
        const bool edge = edit && s.isEdge(idx);
        // pass 1: clear the marks of every node some leaf may report once
        if (edge) {
            for (u32 base = 0; base < nA; base += GS) {
                CLB_LANES_G(l, GS) {
                    const u32 a = base + (u32)l;
                    if (a >= nA)
                        continue;
                    const Node& n = nodes[acts[a].node];
                    if (n.kind() != FK_GD)
                        continue;
                    Cluster C;
                    walkCluster(G, pis[n.pi], acts[a].node, C);
                    GdPlan pl;
                    planGd(C, idx, pl);
                    for (u64 m = pl.rep; m; m &= m - 1)
                        nodes[C.node[ctz64(m)]].cnt = F_NONE;
                }
            }
            grpSyncG<GS>();
            for (u32 base = 0; base < nA; base += GS) {
                CLB_LANES_G(l, GS) {
                    const u32 a = base + (u32)l;
                    if (a >= nA)
                        continue;
                    const Node& n = nodes[acts[a].node];
                    if (n.kind() != FK_GD)
                        continue;
                    Cluster C;
                    walkCluster(G, pis[n.pi], acts[a].node, C);
                    GdPlan pl;
                    planGd(C, idx, pl);
                    for (u64 m = pl.rep; m; m &= m - 1)
                        atomMinU32(&nodes[C.node[ctz64(m)]].cnt, a);
                }
            }
            grpSyncG<GS>();
        }
        // pass 2: emissions and hand-overs of every leaf; sizes of the lists handed over
        u32 nSpawn = 0;
        size_t listBytes = 0;
        for (u32 base = 0; base < nA; base += GS) {
            LV<u32> sp, lb;
            CLB_LANES_G(l, GS) {
                sp[l] = 0;
                lb[l] = 0;
                const u32 a = base + (u32)l;
                if (a >= nA)
                    continue;
                FAct& A = acts[a];
                if (A.kind == FK_IT) // nEv was set with the action
                    continue;
                if (A.kind == FK_HAM_END) {
                    if (isEnd)
                        A.nEv = 1;
                    else
                        sp[l] = 1;
                    continue;
                }
                const Node& n = nodes[A.node];
                const FPI& P = pis[n.pi];
                Cluster C;
                walkCluster(G, P, A.node, C);
                const u32 maxED = s.U[idx], lowerBound = s.L[idx];
                if (edge) {
                    GdPlan pl;
                    planGd(C, idx, pl);
                    u64 mine = 0;
                    for (u64 m = pl.rep; m; m &= m - 1) {
                        const i32 i = (i32)ctz64(m);
                        const Node& cn = nodes[C.node[i]];
                        const u32 dist = isEnd ? C.eds[i] : pl.minED;
                        if (cn.cnt == a && !cn.r.empty() && dist >= lowerBound)
                            mine |= 1ull << i;
                    }
                    A.mask = mine;
                    if (isEnd)
                        A.nEv = (u32)popc64(mine);
                    else
                        sp[l] = mine ? 1u : 0u;
                } else {
                    u16 ie[CL + 1];
                    u32 ni;
                    const i32 ci = centreAndEds(C, lowerBound, maxED, ie, ni);
                    if (ci < 0 || nodes[C.node[ci]].r.empty())
                        continue;
                    u32 nd = (u32)(C.last - ci);
                    if (n.flags() & FF_FROM_CHAIN) // a descendant of the previous phase: the rest of them is handed on
                        nd += P.nDesc - n.row;
                    const u32 maxEDNext = s.U[nIdx];
                    while (ni > 0 && ie[ni - 1] > maxEDNext)
                        ni--;
                    if (nd > LS) {
                        atomOr(B.err, ERR_CLUSTER_OVERFLOW);
                        continue;
                    }
                    A.mask = (u64)(u32)ci;
                    sp[l] = 1;
                    lb[l] = (u32)(((size_t)nd * sizeof(FDesc) + (size_t)ni * 2 + 15) & ~(size_t)15);
                }
            }
            LV<u32> so, lo;
            const u32 ts = grpExScanG<GS>(sp, so);
            const u32 tl = grpExScanG<GS>(lb, lo);
            CLB_LANES_G(l, GS) {
                const u32 a = base + (u32)l;
                if (a < nA && sp[l]) {
                    acts[a].spawn = nSpawn + so[l];
                    acts[a].off = (u32)listBytes + lo[l]; // list offset relative to the list block, for pass 3
                }
            }
            nSpawn += ts;
            listBytes += tl;
            grpSyncG<GS>();
        }
        if (!nSpawn)
            return 0;
        // next group: instances + their roots + lists
        N.idx = nIdx;
        N.nPI = nSpawn;
        N.piOff = alloc((size_t)nSpawn * sizeof(FPI));
        const u32 listOff = alloc(listBytes ? listBytes : 16);
        N.lvlOff = alloc((size_t)(maxRows() + 4) * sizeof(u32));
        N.nodeOff = alloc((size_t)nSpawn * sizeof(Node)); // grows in place during the expansion
        if (err)
            return 0;
        FPI* npis = pisOf(N);
        const bool dswN = nIdx < s.P && s.dsw[nIdx] != 0;
        // pass 3: fill the instances
        for (u32 base = 0; base < nA; base += GS) {
            CLB_LANES_G(l, GS) {
                const u32 a = base + (u32)l;
                if (a >= nA || acts[a].spawn == F_NONE)
                    continue;
                const FAct A = acts[a];
                const Node& n = nodes[A.node];
                const FPI& P = pis[n.pi];
                FPI& Q = npis[A.spawn];
                u32 newDescOff = 0, newNd = 0, newIeOff = 0, newNi = 0;
                LocOcc m;
                m.valid = 1;
                m.shift = P.sm.shift;
                if (n.kind() == FK_HAM_END) {
                    m.r = n.r;
                    m.dist = n.score;
                    m.depth = P.sm.depth + (P.g.n - 1);
                    m.shift = 0;
                } else {
                    Cluster C;
                    walkCluster(G, P, A.node, C);
                    if (edge) { // reportDeepestMinimum
                        u32 minED;
                        i32 hi, deep;
                        deepestMin(C, s.U[idx], minED, hi, deep);
                        const Node& cn = nodes[C.node[deep]];
                        m.r = cn.r;
                        m.dist = minED;
                        m.depth = (u32)cn.row + P.sm.depth - (u32)(deep - hi);
                        m.shift = ((s.dir[idx] == DIR_BWD) ? (u32)(deep - hi) : 0u) + P.sm.shift;
                    } else {
                        u16 ie[CL + 1];
                        u32 ni;
                        const i32 ci = centreAndEds(C, s.L[idx], s.U[idx], ie, ni);
                        const Node& cn = nodes[C.node[ci]];
                        m.r = cn.r;
                        m.dist = C.eds[ci];
                        m.depth = (u32)cn.row + P.sm.depth;
                        u32 nd = (u32)(C.last - ci);
                        FDesc* dl = (FDesc*)(scr + listOff + A.off);
                        for (u32 j = 0; j < nd; j++) {
                            const Node& dn = nodes[C.node[ci + 1 + (i32)j]];
                            dl[j].r = dn.r;
                            dl[j].c = dn.c();
                            dl[j].pad = 0;
                        }
                        if (n.flags() & FF_FROM_CHAIN) {
                            const FDesc* rem = (const FDesc*)(scr + P.descOff);
                            for (u32 j = n.row; j < P.nDesc; j++)
                                dl[nd++] = rem[j];
                        }
                        const u32 maxEDNext = s.U[nIdx];
                        while (ni > 0 && ie[ni - 1] > maxEDNext)
                            ni--;
                        u16* il = (u16*)(dl + nd);
                        for (u32 j = 0; j < ni; j++)
                            il[j] = ie[j];
                        newDescOff = listOff + A.off;
                        newNd = nd;
                        newIeOff = newDescOff + nd * (u32)sizeof(FDesc);
                        newNi = ni;
                        if (dswN && nd > 0) {
                            m.r = dl[nd - 1].r;
                            u32 mn = ie[0];
                            for (u32 j = 1; j < ni; j++)
                                mn = tmin<u32>(mn, ie[j]);
                            m.dist = mn;
                        }
                    }
                }
                Q.sm = m;
                // pushFrame(m, nIdx, listP = new lists, listN = this instance's other lists)
                if (dswN) {
                    Q.descOff = P.oDescOff; Q.nDesc = P.nDescO; Q.ieOff = P.oIeOff; Q.nIe = P.nIeO;
                    Q.oDescOff = newDescOff; Q.nDescO = newNd; Q.oIeOff = newIeOff; Q.nIeO = newNi;
                } else {
                    Q.descOff = newDescOff; Q.nDesc = newNd; Q.ieOff = newIeOff; Q.nIe = newNi;
                    Q.oDescOff = P.oDescOff; Q.nDescO = P.nDescO; Q.oIeOff = P.oIeOff; Q.nIeO = P.nIeO;
                }
                Q.total = 0;
                Q.base = 0;
                Q.actBegin = 0;
                Q.sBegin = 0;
                acts[a].off = 0;
            }
        }
        grpSyncG<GS>();
        return nSpawn;
    }

    // roots of a group (after its instances are filled)
    CLB_HD void initGroup(FGroup& G, const PartSeq& ps) {
        for (u32 base = 0; base < G.nPI; base += GS) {
            CLB_LANES_G(l, GS) {
                const u32 p = base + (u32)l;
                if (p < G.nPI)
                    initInstance(G, p, ps);
            }
        }
        grpSyncG<GS>();
    }

    // emission indices over all groups, then the emissions themselves
    CLB_HD void numberAndEmit() {
        // totals, last group first
        for (u32 q = nGroups; q-- > 0;) {
            FGroup& G = grp[q];
            FAct* acts = actsOf(G);
            FPI* pis = pisOf(G);
            FPI* npis = (q + 1 < nGroups) ? pisOf(grp[q + 1]) : (FPI*)0;
            for (u32 base = 0; base < G.nActs; base += GS) {
                CLB_LANES_G(l, GS) {
                    const u32 a = base + (u32)l;
                    if (a < G.nActs) {
                        const u32 w = acts[a].spawn != F_NONE ? npis[acts[a].spawn].total : acts[a].nEv;
                        if (w)
                            atomAdd<u32>(&pis[acts[a].pi].total, w);
                    }
                }
            }
            grpSyncG<GS>();
        }
        // offsets, first group first
        for (u32 q = 0; q < nGroups; q++) {
            FGroup& G = grp[q];
            FAct* acts = actsOf(G);
            FPI* pis = pisOf(G);
            FPI* npis = (q + 1 < nGroups) ? pisOf(grp[q + 1]) : (FPI*)0;
            u32 run = 0;
            for (u32 base = 0; base < G.nActs; base += GS) {
                LV<u32> w, off;
                CLB_LANES_G(l, GS) {
                    const u32 a = base + (u32)l;
                    w[l] = 0;
                    if (a < G.nActs)
                        w[l] = acts[a].spawn != F_NONE ? npis[acts[a].spawn].total : acts[a].nEv;
                }
                const u32 tot = grpExScanG<GS>(w, off);
                CLB_LANES_G(l, GS) {
                    const u32 a = base + (u32)l;
                    if (a < G.nActs)
                        acts[a].off = run + off[l];
                }
                run += tot;
            }
            grpSyncG<GS>();
            for (u32 base = 0; base < G.nPI; base += GS) {
                CLB_LANES_G(l, GS) {
                    const u32 p = base + (u32)l;
                    if (p < G.nPI)
                        pis[p].sBegin = pis[p].actBegin < G.nActs ? acts[pis[p].actBegin].off : 0u;
                }
            }
            grpSyncG<GS>();
            for (u32 base = 0; base < G.nActs; base += GS) {
                CLB_LANES_G(l, GS) {
                    const u32 a = base + (u32)l;
                    if (a < G.nActs) {
                        const FPI& P = pis[acts[a].pi];
                        const u32 o = P.base + (acts[a].off - P.sBegin);
                        acts[a].off = o;
                        if (acts[a].spawn != F_NONE)
                            npis[acts[a].spawn].base = o;
                    }
                }
            }
            grpSyncG<GS>();
        }
        // emissions
        for (u32 q = 0; q < nGroups; q++) {
            FGroup& G = grp[q];
            FAct* acts = actsOf(G);
            Node* nodes = nodesOf(G);
            FPI* pis = pisOf(G);
            const bool isEnd = (G.idx + 1 == s.P);
            for (u32 base = 0; base < G.nActs; base += GS) {
                LV<u32> nc, nf;
                CLB_LANES_G(l, GS) {
                    const u32 a = base + (u32)l;
                    nc[l] = 0;
                    nf[l] = 0;
                    if (a < G.nActs && acts[a].nEv) {
                        if (acts[a].kind == FK_IT)
                            nc[l] = acts[a].nEv;
                        else
                            nf[l] = acts[a].nEv;
                    }
                }
                LV<u32> co, fo;
                const u32 tc = grpExScanG<GS>(nc, co), tf = grpExScanG<GS>(nf, fo);
                if (!tc && !tf)
                    continue;
                LV<u32> cb, fb;
                CLB_LANES_G(l, GS) {
                    cb[l] = 0;
                    fb[l] = 0;
                    if (l == 0) {
                        if (tc)
                            cb[l] = B.cands.reserve(tc, B.err, ERR_CAND_OVERFLOW);
                        if (tf)
                            fb[l] = B.fmOccs.reserve(tf, B.err, ERR_FMOCC_OVERFLOW);
                    }
                }
                const u32 candAt = grpBcastG<GS>(cb, 0), fmAt = grpBcastG<GS>(fb, 0);
                CLB_LANES_G(l, GS) {
                    const u32 a = base + (u32)l;
                    if (a >= G.nActs || !acts[a].nEv)
                        continue;
                    const FAct A = acts[a];
                    const Node& n = nodes[A.node];
                    const FPI& P = pis[n.pi];
                    if (n.kind() == FK_IT) {
                        if (candAt == F_NONE)
                            continue;
                        Cand c;
                        c.t = t;
                        c.maxED = (u8)s.maxED();
                        c.minED = (u8)s.minED();
                        c.pad = 0;
                        if (edit) {
                            bool fixedStart;
                            c.sub = startDiffOf(G, P, n, fixedStart);
                            c.add = P.sm.shift;
                            c.fixedStart = fixedStart ? 1 : 0;
                        } else { // inTextVerificationHamming, src/fmindex/fmindex.cpp:409-428
                            c.add = 0;
                            c.sub = ((G.idx == 0) ? 0u : (u32)pi.begin[s.lowestBefore(G.idx)]) -
                                    ((s.dir[G.idx] == DIR_BWD) ? (u32)n.row : 0u);
                            c.fixedStart = 1;
                        }
                        for (u32 j = 0; j < A.nEv; j++) {
                            c.row = n.r.sa.b + j;
                            c.order = orderKey(A.off + j);
                            B.cands.data[candAt + co[l] + j] = c;
                        }
                    } else {
                        if (fmAt == F_NONE)
                            continue;
                        ctr[l].fmOccs += A.nEv;
                        FMOccRec f;
                        f.t = t;
                        if (n.kind() == FK_HAM_END) {
                            f.sa = n.r.sa;
                            f.depth = P.sm.depth + (P.g.n - 1);
                            f.dist = (u16)n.score;
                            f.shift = 0;
                            f.order = orderKey(A.off);
                            B.fmOccs.data[fmAt + fo[l]] = f;
                            continue;
                        }
                        Cluster C;
                        walkCluster(G, P, A.node, C);
                        u32 j = 0;
                        for (u64 m = A.mask; m; m &= m - 1, j++) {
                            const i32 i = (i32)ctz64(m);
                            const Node& cn = nodes[C.node[i]];
                            f.sa = cn.r.sa;
                            f.depth = (u32)cn.row + P.sm.depth;
                            f.dist = (u16)C.eds[i];
                            f.shift = (u16)P.sm.shift;
                            f.order = orderKey(A.off + j);
                            B.fmOccs.data[fmAt + fo[l] + j] = f;
                        }
                    }
                }
            }
            (void)isEnd;
        }
    }

    // doRecSearch + recApproxMatchEditEntry: the task.  Returns false when the
    // scratch was too small (nothing has been emitted; the task can be re-run).
    CLB_HD bool run(LV<u64>& searchStarted, LV<u64>& immediate) {
        // ---- start of the search: by lane 0, result through the scratch
        struct Start {
            LocOcc sm;
            u32 idx, go; // go: 0 nothing to do, 1 search, 2 immediate switch
            u64 lookups, nodes;
        };
        const u32 startOff = alloc(sizeof(Start));
        if (err)
            return false;
        Start* st = (Start*)(scr + startOff);
        CLB_LANES_G(l, GS) {
            if (l != 0)
                continue;
            LocalCtr lc;
            lc.clear();
            Start S;
            S.go = 0;
            S.idx = 0;
            S.sm.shift = 0;
            S.sm.valid = 1;
            S.sm.dist = 0;
            S.sm.depth = 0;
            S.sm.r = emptyPair();
            if (s.U[0] > 0) {
                S.sm.r = completeRange(ix);
                S.go = 1;
            } else {
                const u32 first = s.order[0];
                RangePair r = pi.emr[first];
                if (r.width() > ix.switchPoint) {
                    u32 pis = 1, exactLength = partLen(first);
                    while (pis < s.P && s.U[pis] == 0) {
                        const u32 p = s.order[pis];
                        r = matchString(ix, s.dir[pis], s.isUniBack(pis), seqCodes(B, t), pi.begin[p], pi.end[p],
                                        s.dir[pis], r, lc);
                        if (r.empty())
                            break;
                        exactLength += partLen(p);
                        pis++;
                    }
                    if (!r.empty() && pis < s.P) {
                        S.sm.r = r;
                        S.sm.depth = exactLength;
                        S.idx = pis;
                        S.go = (r.width() > ix.switchPoint) ? 1u : (edit ? 2u : 1u);
                    }
                }
            }
            S.lookups = lc.s[ST_RANK_LOOKUPS];
            S.nodes = lc.c[NODE_COUNTER];
            *st = S;
        }
        grpSyncG<GS>();
        const Start S = *st;
        CLB_LANES_G(l, GS) {
            if (l == 0) {
                ctr[l].lookups += S.lookups;
                ctr[l].nodes += S.nodes;
            }
        }
        if (S.go == 0)
            return true;
        // A task that starts from a wide range outgrows the first launch's scratch with near certainty: it is handed
        // to the heavy launch here instead of after 16 k nodes of work that the heavy launch then repeats.
        if (S.go == 1 && S.idx > 0 && heavyWidth && S.sm.r.width() > heavyWidth) {
            err |= ERR_STACK_OVERFLOW;
            return false;
        }
        if (S.go == 2) { // verifyExactPartialMatchInText, src/fmindex/fmindex.cpp:245-265
            CLB_LANES_G(l, GS) {
                if (l != 0)
                    continue;
                immediate[l]++;
                const u32 bip = pi.begin[s.lowestBefore(S.idx)];
                const u32 w = S.sm.r.sa.width();
                const u32 at = B.cands.reserve(w, B.err, ERR_CAND_OVERFLOW);
                Cand c;
                c.t = t;
                c.add = 0;
                c.sub = (bip == 0) ? 0 : bip + s.maxED();
                c.fixedStart = (bip == 0);
                c.maxED = (u8)s.maxED();
                c.minED = (u8)s.minED();
                c.pad = 0;
                for (u32 i = 0; i < w && at != F_NONE; i++) {
                    c.row = S.sm.r.sa.b + i;
                    c.order = orderKey(i);
                    B.cands.data[at + i] = c;
                }
            }
            return true;
        }
        CLB_LANES_G(l, GS) {
            if (l == 0 && edit)
                searchStarted[l]++;
        }
        // ---- first group: one instance
        {
            FGroup& G = grp[0];
            G.idx = S.idx;
            G.nPI = 1;
            G.piOff = alloc(sizeof(FPI));
            G.lvlOff = alloc((size_t)(maxRows() + 4) * sizeof(u32));
            G.nodeOff = alloc(sizeof(Node));
            if (err)
                return false;
            FPI P;
            P.sm = S.sm;
            P.descOff = P.nDesc = P.ieOff = P.nIe = 0;
            P.oDescOff = P.nDescO = P.oIeOff = P.nIeO = 0;
            P.actBegin = P.total = P.base = P.sBegin = 0;
            P.clSize = 0;
            P.g.n = P.g.m = P.g.Wv = P.g.Wh = P.g.maxED = 0;
            CLB_LANES_G(l, GS) {
                if (l == 0)
                    pisOf(G)[0] = P;
            }
            grpSyncG<GS>();
            nGroups = 1;
        }
        for (;;) {
            FGroup& G = grp[nGroups - 1];
            const PartSeq ps = partSeq(s.order[G.idx], s.dir[G.idx]);
            initGroup(G, ps);
            if (GS <= 32)
                expandGroupPairs(G, ps);
            else
                expandGroup(G, ps);
            if (err)
                return false;
            rankLeaves(G);
            if (err)
                return false;
            if (!G.nActs)
                break;
            FGroup N;
            const u32 ns = actGroup(G, N);
            if (err)
                return false;
            if (!ns)
                break;
            if (nGroups >= MAXP) {
                err |= ERR_INTERNAL;
                return false;
            }
            grp[nGroups++] = N;
        }
        numberAndEmit();
        return true;
    }
};

} // namespace clb
