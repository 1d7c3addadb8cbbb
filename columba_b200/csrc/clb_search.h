// columba_b200 -- stage 2: search-scheme exploration over the bidirectional
// FM-index.  One task per (read, strand, search); the reference's recursion
// over the phases of a search is an explicit frame stack here, the node stack,
// the matrix rows, the cluster of the final column and the descendant lists
// live in a per-worker slab in HBM.
//
// Replaces (reference): SearchStrategy::doRecSearch src/searchstrategy.cpp:1181-1254,
// IndexInterface::recApproxMatchEditEntry src/indexinterface.cpp:1306-1325,
// recApproxMatchEdit :377-527, branchAndBound :529-561, goDeeper :563-669,
// goToInTextVerificationEdit :340-375, extendFMPos :675-697,
// recApproxMatchHamming :1211-1304, FMIndex::inTextVerificationHamming
// src/fmindex/fmindex.cpp:409-428, Cluster (MatrixMetaInfo)
// src/indexhelpers.h:1677-1838 and getClusterCentra src/indexhelpers.cpp:276-382.
#pragma once
#include "clb_batch.h"
#include "clb_matrix.h"
#include "clb_partition.h"

namespace clb {

// indexhelpers.h:1544-1655 (FMPosExt)
struct NodeRec {
    RangePair r;
    u32 depth;
    u8 c;
    u8 reported;
    u16 pad;
};
struct LocOcc {
    RangePair r;
    u32 depth, dist, shift;
    u32 valid;
};

// capacity of a final-column cluster (Wh + Wv + 1 <= MAX_ED + 2*MAX_ED + 1) and of
// a descendant list, per matrix word type
template <typename W> struct SearchCaps {
    static constexpr u32 CL = 3 * BPC<W>::MAX_ED + 2;
    static constexpr u32 LS = 2 * CL;
};

template <typename W> struct SearchScratch {
    NodeRec* stack;
    u32 stackCap;
    Row<W>* rows;
    u32 rowsCap;
    NodeRec* clNodes; // [nFrames][CL]
    u16* clEds;       // [nFrames][CL]
    NodeRec* lsDesc;  // [nFrames+1][LS]
    u16* lsIe;        // [nFrames+1][LS]
    u32 nFrames;
};
template <typename W> CLB_HD size_t searchScratchBytes(u32 stackCap, u32 rowsCap, u32 nFrames) {
    const u32 CL_CAP = SearchCaps<W>::CL, LS_CAP = SearchCaps<W>::LS;
    size_t b = 0;
    b += (size_t)stackCap * sizeof(NodeRec);
    b += (size_t)rowsCap * sizeof(Row<W>);
    b += (size_t)nFrames * CL_CAP * (sizeof(NodeRec) + sizeof(u16));
    b += (size_t)(nFrames + 1) * LS_CAP * (sizeof(NodeRec) + sizeof(u16));
    return (b + 255) & ~(size_t)63;
}
template <typename W> CLB_HD SearchScratch<W> carveSearchScratch(u8* slab, u32 stackCap, u32 rowsCap, u32 nFrames) {
    const u32 CL_CAP = SearchCaps<W>::CL, LS_CAP = SearchCaps<W>::LS;
    const u32 MAXP = nFrames;
    SearchScratch<W> s;
    s.rows = (Row<W>*)slab;
    slab += (size_t)rowsCap * sizeof(Row<W>);
    s.stack = (NodeRec*)slab;
    slab += (size_t)stackCap * sizeof(NodeRec);
    s.clNodes = (NodeRec*)slab;
    slab += (size_t)MAXP * CL_CAP * sizeof(NodeRec);
    s.lsDesc = (NodeRec*)slab;
    slab += (size_t)(MAXP + 1) * LS_CAP * sizeof(NodeRec);
    s.clEds = (u16*)slab;
    slab += (size_t)MAXP * CL_CAP * sizeof(u16);
    s.lsIe = (u16*)slab;
    s.stackCap = stackCap;
    s.rowsCap = rowsCap;
    s.nFrames = nFrames;
    return s;
}

struct Frame {
    LocOcc sm;
    u32 idx;
    i32 listP, listN;       // list slots passed in (prev-direction / other)
    i32 descSlot, otherSlot;
    u32 maxED, dir, uni, dSwitch;
    MatGeom g;
    u32 rowBase, stackBase;
    PartSeq ps;
    u32 clSize;
    i32 lastCell;
    u32 state, di;
};
enum { FS_DESC = 1, FS_LOOP = 2, FS_POP_AFTER_CHILD = 3 };

template <typename W> struct SearchCtx {
    const IndexView& ix;
    const Batch& B;
    const DevSearch& s;
    const PartInfo& pi;
    SearchScratch<W> sc;
    LocalCtr& lc;
    u32 t, searchIdx;
    u32 sp;      // node stack pointer
    u32 emit;    // emission order counter
    u32 err;
    Frame fr[MAXP];
    u32 nDesc[MAXP + 1], nIe[MAXP + 1];
    int top;     // index of the top frame, -1 if none

    CLB_HD SearchCtx(const IndexView& ix_, const Batch& B_, const DevSearch& s_, const PartInfo& pi_,
                     SearchScratch<W> sc_, LocalCtr& lc_, u32 t_, u32 si)
        : ix(ix_), B(B_), s(s_), pi(pi_), sc(sc_), lc(lc_), t(t_), searchIdx(si), sp(0), emit(0), err(0), top(-1) {}

    CLB_HD u32 orderKey() { return ((searchIdx + 1) << 27) | (emit++ & 0x7FFFFFFu); }
    CLB_HD u32 partBegin(u32 p) const { return pi.begin[p]; }
    CLB_HD u32 partLen(u32 p) const { return pi.end[p] <= pi.begin[p] ? 0u : (u32)(pi.end[p] - pi.begin[p]); }
    CLB_HD PartSeq partSeq(u32 p, u32 dir) const {
        PartSeq ps;
        ps.planes = seqPlanes(B, t, dir);
        ps.nw = B.nw;
        ps.len = partLen(p);
        ps.base = (dir == DIR_FWD) ? pi.begin[p] : (u32)pi.len - pi.end[p];
        return ps;
    }
    // character j of part p read in direction dir
    CLB_HD u32 partChar(u32 p, u32 dir, u32 j) const {
        const u8* codes = seqCodes(B, t);
        return (dir == DIR_FWD) ? codes[pi.begin[p] + j] : codes[pi.end[p] - 1 - j];
    }

    // indexinterface.cpp:675-697
    CLB_HD void extendFMPos(const RangePair& parent, u32 dir, u32 uni, u32 row) {
        RangePair ch[4];
        extendAll(ix, (int)dir, uni != 0, parent, ch);
        lc.s[ST_RANK_LOOKUPS] += 2;
        for (u32 c = 0; c < 4; c++) {
            if (!ch[c].empty()) {
                if (sp >= sc.stackCap) {
                    err |= ERR_STACK_OVERFLOW;
                    return;
                }
                NodeRec n;
                n.r = ch[c];
                n.depth = row + 1;
                n.c = (u8)c;
                n.reported = 0;
                n.pad = 0;
                sc.stack[sp++] = n;
                lc.c[NODE_COUNTER]++;
            }
        }
    }
    CLB_HD void emitCands(const Range& sa, u32 add, u32 sub, bool fixedStart, u32 maxED, u32 minED) {
        u32 w = sa.width();
        if (!w)
            return;
        u32 at = B.cands.reserve(w, B.err, ERR_CAND_OVERFLOW);
        Cand c;
        c.t = t;
        c.add = add;
        c.sub = sub;
        c.fixedStart = fixedStart;
        c.maxED = (u8)maxED;
        c.minED = (u8)minED;
        c.pad = 0;
        for (u32 i = 0; i < w; i++) {
            c.row = sa.b + i;
            c.order = orderKey();
            if (at != 0xFFFFFFFFu)
                B.cands.data[at + i] = c;
        }
    }
    CLB_HD void emitFMOcc(const LocOcc& m) {
        FMOccRec f;
        f.t = t;
        f.sa = m.r.sa;
        f.depth = m.depth;
        f.dist = (u16)m.dist;
        f.shift = (u16)m.shift;
        f.order = orderKey();
        B.fmOccs.push(f, B.err, ERR_FMOCC_OVERFLOW);
        lc.s[ST_FM_OCCS]++;
    }
};

} // namespace clb

#include "clb_search_edit.h"
#include "clb_search_hamming.h"
