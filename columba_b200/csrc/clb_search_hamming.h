// columba_b200 -- Hamming-distance search of one (read,strand,search) task.
// Replaces IndexInterface::recApproxMatchHamming src/indexinterface.cpp:1211-1304
// and FMIndex::inTextVerificationHamming(node,...) src/fmindex/fmindex.cpp:409-428.
#pragma once

namespace clb {

struct HamFrame {
    LocOcc sm;
    u32 idx, part, pSize, dir, uni, maxED, minED;
    u32 vecBase, stackBase;
};

// the mismatch counts along the current path are kept in the `score` field
// of the worker's row slab
template <typename W> CLB_HD void hammingSearchTask(SearchCtx<W>& X) {
    const DevSearch& s = X.s;
    const IndexView& ix = X.ix;
    LocOcc sm;
    sm.shift = 0;
    sm.valid = 1;
    sm.dist = 0;
    u32 idx0;
    if (s.U[0] > 0) {
        sm.r = completeRange(ix);
        sm.depth = 0;
        idx0 = 0;
    } else {
        u32 first = s.order[0];
        RangePair r = X.pi.emr[first];
        if (r.width() <= ix.switchPoint)
            return;
        u32 pis = 1, exactLength = X.partLen(first);
        while (pis < s.P && s.U[pis] == 0) {
            u32 p = s.order[pis];
            r = matchString(ix, s.dir[pis], s.isUniBack(pis), seqCodes(X.B, X.t), X.pi.begin[p], X.pi.end[p],
                            s.dir[pis], r, X.lc);
            if (r.empty())
                return;
            exactLength += X.partLen(p);
            pis++;
        }
        if (pis >= s.P)
            return;
        sm.r = r;
        sm.depth = exactLength;
        idx0 = pis;
    }
    HamFrame fr[MAXP];
    int top = -1;
    Row<W>* vec = X.sc.rows;
    // push the root frame
    auto push = [&](const LocOcc& m, u32 idx) {
        int f = top + 1;
        HamFrame& F = fr[f];
        F.sm = m;
        F.idx = idx;
        F.part = s.order[idx];
        F.pSize = X.partLen(F.part);
        F.dir = s.dir[idx];
        F.uni = s.isUniBack(idx);
        F.maxED = s.U[idx];
        F.minED = s.L[idx];
        F.vecBase = (f == 0) ? 0u : fr[f - 1].vecBase + fr[f - 1].pSize + 1;
        F.stackBase = X.sp;
        if (F.vecBase + F.pSize + 1 > X.sc.rowsCap) {
            X.err |= ERR_ROWS_OVERFLOW;
            return;
        }
        vec[F.vecBase].score = m.dist;
        top = f;
        X.extendFMPos(m.r, F.dir, F.uni, 0);
    };
    push(sm, idx0);
    while (top >= 0 && !X.err) {
        HamFrame& F = fr[top];
        if (X.sp == F.stackBase) {
            top--;
            continue;
        }
        const NodeRec node = X.sc.stack[--X.sp];
        if (node.r.width() <= ix.switchPoint) {
            u32 lengthBefore = ((F.idx == 0) ? 0u : X.partBegin(s.lowestBefore(F.idx))) -
                               ((F.dir == DIR_BWD) ? node.depth : 0u);
            X.emitCands(node.r.sa, 0, lengthBefore, true, s.maxED(), s.minED());
            continue;
        }
        const u32 row = node.depth;
        u32 v = vec[F.vecBase + row - 1].score + (node.c != X.partChar(F.part, F.dir, row - 1) ? 1u : 0u);
        vec[F.vecBase + row].score = v;
        if (v > F.maxED)
            continue;
        if (row == F.pSize) {
            if (v >= F.minED) {
                LocOcc m;
                m.r = node.r;
                m.dist = v;
                m.depth = F.sm.depth + F.pSize;
                m.shift = 0;
                m.valid = 1;
                if (s.isEnd(F.idx))
                    X.emitFMOcc(m);
                else
                    push(m, F.idx + 1);
            }
            continue;
        }
        X.extendFMPos(node.r, F.dir, F.uni, node.depth);
    }
}

} // namespace clb
