// columba_b200 -- edit-distance search of one (read,strand,search) task.
// See clb_search.h for the reference lines this replaces.
#pragma once

namespace clb {

template <typename W> struct EditSearch {
    static constexpr u32 CL_CAP = SearchCaps<W>::CL, LS_CAP = SearchCaps<W>::LS;
    SearchCtx<W>& X;
    CLB_HD EditSearch(SearchCtx<W>& x) : X(x) {}

    CLB_HD NodeRec* clNodes(int f) { return X.sc.clNodes + (size_t)f * CL_CAP; }
    CLB_HD u16* clEds(int f) { return X.sc.clEds + (size_t)f * CL_CAP; }
    CLB_HD NodeRec* lsDesc(int slot) { return X.sc.lsDesc + (size_t)slot * LS_CAP; }
    CLB_HD u16* lsIe(int slot) { return X.sc.lsIe + (size_t)slot * LS_CAP; }
    CLB_HD u32 nDescOf(int slot) { return slot < 0 ? 0u : X.nDesc[slot]; }
    CLB_HD u32 nIeOf(int slot) { return slot < 0 ? 0u : X.nIe[slot]; }
    CLB_HD Row<W>* rowsOf(const Frame& F) { return X.sc.rows + F.rowBase; }

    // push a frame for phase idx (recApproxMatchEdit prologue,
    // indexinterface.cpp:377-470)
    CLB_HD void pushFrame(const LocOcc& sm, u32 idx, i32 listP, i32 listN) {
        int f = X.top + 1;
        if (f >= (int)X.sc.nFrames || f >= MAXP) {
            X.err |= ERR_INTERNAL;
            return;
        }
        Frame& F = X.fr[f];
        const DevSearch& s = X.s;
        F.sm = sm;
        F.idx = idx;
        F.listP = listP;
        F.listN = listN;
        F.maxED = s.U[idx];
        F.dir = s.dir[idx];
        F.uni = s.isUniBack(idx);
        F.dSwitch = s.dsw[idx];
        F.descSlot = F.dSwitch ? listN : listP;
        F.otherSlot = F.dSwitch ? listP : listN;
        const u32 part = s.order[idx];
        F.ps = X.partSeq(part, F.dir);
        F.rowBase = (f == 0) ? 0u : X.fr[f - 1].rowBase + X.fr[f - 1].g.m;
        F.stackBase = X.sp;
        // first column
        const u32 nInit = nIeOf(F.descSlot);
        const u16* ie = nInit ? lsIe(F.descSlot) : (const u16*)0;
        u16 single[1];
        u32 add = 0, n = nInit;
        if (nInit == 0) {
            single[0] = (u16)sm.dist;
            ie = single;
            n = 1;
        } else {
            u32 prevED = ie[0];
            if (F.dSwitch)
                for (u32 i = 1; i < nInit; i++)
                    prevED = tmin<u32>(prevED, ie[i]);
            add = sm.dist - prevED;
        }
        // size check before touching the rows: m <= n - 1 + maxED + len + 1
        if (F.rowBase + n + F.maxED + F.ps.len + 2 + F.maxED > X.sc.rowsCap) {
            X.err |= ERR_ROWS_OVERFLOW;
            return;
        }
        bpInit<W, u16>(F.g, rowsOf(F), F.ps.len, F.maxED, ie, n, add);
        F.clSize = F.g.sizeOfFinalColumn();
        if (F.clSize > CL_CAP || F.rowBase + F.g.m > X.sc.rowsCap) {
            X.err |= ERR_CLUSTER_OVERFLOW;
            return;
        }
        u16* eds = clEds(f);
        NodeRec* nodes = clNodes(f);
        NodeRec z;
        z.r = emptyPair();
        z.depth = 0;
        z.c = 0;
        z.reported = 0;
        z.pad = 0;
        for (u32 i = 0; i < F.clSize; i++) {
            eds[i] = (u16)(F.maxED + 1);
            nodes[i] = z;
        }
        F.lastCell = -1;
        X.top = f;
        if (F.g.inFinalColumn(0)) {
            NodeRec n0 = z;
            n0.r = sm.r;
            setValue(f, 0, n0, bpAt<W>(rowsOf(F), 0, F.ps.len));
        }
        if (nDescOf(F.descSlot) > 0) {
            F.state = FS_DESC;
            F.di = 0;
        } else {
            X.extendFMPos(sm.r, F.dir, F.uni, 0);
            F.state = FS_LOOP;
        }
    }
    CLB_HD void setValue(int f, u32 idx, const NodeRec& n, u32 ed) {
        clEds(f)[idx] = (u16)ed;
        clNodes(f)[idx] = n;
        X.fr[f].lastCell = (i32)idx;
    }
    // FMPosExt::report, indexhelpers.h:1613-1629
    CLB_HD void report(NodeRec& n, LocOcc& occ, u32 startDepth, u32 ed, bool once, u32 shift) {
        if (!n.reported) {
            occ.r = n.r;
            occ.dist = ed;
            occ.depth = n.depth + startDepth;
            occ.shift = shift;
            occ.valid = !n.r.empty();
            if (once)
                n.reported = 1;
        }
    }

    // goToInTextVerificationEdit, indexinterface.cpp:340-375
    CLB_HD void goInText(int f, const NodeRec& node) {
        Frame& F = X.fr[f];
        const DevSearch& s = X.s;
        u32 stp = X.partBegin(s.lowestBefore(F.idx));
        u32 startDiff = stp + s.maxED();
        if (stp == 0) {
            startDiff = 0;
        } else if (F.dir == DIR_BWD) {
            u32 row = node.depth;
            u32 col = F.g.firstColumn(row);
            startDiff -= col + bpAt<W>(rowsOf(F), row, col);
        } else if (nDescOf(F.otherSlot) > 0) {
            u32 nd = nDescOf(F.otherSlot), ni = nIeOf(F.otherSlot);
            startDiff -= (nd - ni + (u32)lsIe(F.otherSlot)[ni - 1]);
        }
        X.emitCands(node.r.sa, F.sm.shift, startDiff, stp == 0, s.maxED(), s.minED());
    }

    // returns true when the branch is finished (indexinterface.cpp:529-561)
    CLB_HD bool branchAndBound(int f, const NodeRec& cur, i32 remSlot, u32 remFrom) {
        Frame& F = X.fr[f];
        const u32 row = cur.depth;
        if (row >= F.g.m) {
            X.err |= ERR_INTERNAL;
            return true;
        }
        Row<W>* rows = rowsOf(F);
        bool valid = bpComputeRow<W>(F.g, rows, F.ps, row, cur.c);
        X.lc.s[ST_SEARCH_ROWS]++;
        if (F.g.inFinalColumn(row)) {
            u32 cIdx = F.clSize + row - F.g.m;
            setValue(f, cIdx, cur, bpAt<W>(rows, row, F.g.n - 1));
            if (!valid || bpOnlyVerticalGapsLeft<W>(F.g, rows, row)) {
                goDeeper(f, remSlot, remFrom);
                return true;
            }
        }
        return !valid;
    }

    // indexinterface.cpp:563-669.  May push one frame (the next phase).
    CLB_HD void goDeeper(int f, i32 remSlot, u32 remFrom) {
        Frame& F = X.fr[f];
        const DevSearch& s = X.s;
        const u32 nIdx = F.idx + 1;
        const bool isEdge = s.isEdge(nIdx - 1);
        const u32 lowerBound = s.L[nIdx - 1];
        u16* eds = clEds(f);
        NodeRec* nodes = clNodes(f);
        const i32 last = F.lastCell;
        const u32 maxED = F.maxED;
        if (isEdge) {
            if (nIdx == s.P) { // reportCentersAtEnd, indexhelpers.h:1738-1756
                for (i32 i = 0; i <= last; i++) {
                    if (eds[i] <= maxED && (i == 0 || eds[i] <= eds[i - 1]) && (i == last || eds[i] <= eds[i + 1])) {
                        LocOcc m;
                        m.valid = 0;
                        report(nodes[i], m, F.sm.depth, eds[i], true, F.sm.shift);
                        if (m.valid && m.dist >= lowerBound)
                            X.emitFMOcc(m);
                    }
                }
            } else { // reportDeepestMinimum, indexhelpers.h:1770-1798
                u32 minED = maxED + 1;
                i32 hi = -1, deep = -1;
                for (i32 i = 0; i <= last; i++) {
                    if (eds[i] < minED) {
                        minED = eds[i];
                        hi = i;
                        deep = i;
                    }
                    if (eds[i] == minED)
                        deep = i;
                }
                LocOcc m;
                m.valid = 0;
                if (minED <= maxED)
                    report(nodes[deep], m, F.sm.depth - (u32)(deep - hi), minED, true,
                           ((F.dir == DIR_BWD) ? (u32)(deep - hi) : 0u) + F.sm.shift);
                if (m.valid && m.dist >= lowerBound)
                    pushFrame(m, nIdx, -1, F.otherSlot);
            }
            return;
        }
        // getClusterCentra, indexhelpers.cpp:276-382
        const int slot = f + 1;
        NodeRec* desc = lsDesc(slot);
        u16* ie = lsIe(slot);
        u32 nd = 0, ni = 0;
        LocOcc m;
        m.valid = 0;
        for (i32 i = 0; i <= last; i++) {
            if (eds[i] > maxED || eds[i] < lowerBound)
                continue;
            bool bp = (i == 0) || eds[i] <= eds[i - 1];
            bool bc = (i == last) || eds[i] <= eds[i + 1];
            if (!(bp && bc))
                continue;
            report(nodes[i], m, F.sm.depth, eds[i], false, F.sm.shift);
            ie[ni++] = eds[i];
            for (i32 j = i + 1; j <= last; j++) {
                desc[nd++] = nodes[j];
                ie[ni++] = eds[j];
            }
            for (u32 k = 1; k < ni; k++) {
                if (ie[k] < lowerBound && ie[k] <= ie[k - 1] && (k == ni - 1 || ie[k] <= ie[k + 1])) {
                    u32 highestPoint = 0, lowestPoint = ni - 1;
                    for (u32 l = k; l-- > 0;)
                        if (ie[l] != ie[l + 1] + 1) {
                            highestPoint = l + 1;
                            break;
                        }
                    for (u32 l = k + 1; l < ni; l++)
                        if (ie[l] != ie[l - 1] + 1) {
                            lowestPoint = l - 1;
                            break;
                        }
                    if (highestPoint != 0 && lowestPoint != ni - 1) {
                        u32 lC = lowestPoint, hC = highestPoint;
                        bool highest = true;
                        while (lC > hC) {
                            if (highest) {
                                ie[hC] = (u16)tmin<int>((int)maxED + 1, (int)ie[hC - 1] + 1);
                                hC++;
                            } else {
                                ie[lC] = (u16)tmin<int>((int)maxED + 1, (int)ie[lC + 1] + 1);
                                lC--;
                            }
                            highest = !highest;
                        }
                        if (lC == hC)
                            ie[lC] = (u16)tmin<int>((int)ie[lC + 1] + 1, (int)ie[lC - 1] + 1);
                    } else if (highestPoint == 0 && lowestPoint != ni - 1) {
                        for (u32 l = lowestPoint; l-- > 0;)
                            ie[l] = (u16)(ie[l + 1] + 1);
                    } else if (highestPoint != 0 && lowestPoint == ni - 1) {
                        for (u32 l = highestPoint; l < ni; l++)
                            ie[l] = (u16)(ie[l - 1] + 1);
                    }
                }
            }
            break;
        }
        if (!m.valid)
            return;
        // append the remaining descendants of the current phase
        if (remSlot >= 0) {
            const NodeRec* rem = lsDesc(remSlot);
            u32 nrem = X.nDesc[remSlot];
            for (u32 i = remFrom; i < nrem; i++) {
                if (nd >= LS_CAP) {
                    X.err |= ERR_CLUSTER_OVERFLOW;
                    return;
                }
                desc[nd++] = rem[i];
            }
        }
        for (u32 i = 0; i < nd; i++)
            desc[i].depth = i + 1;
        const u32 maxEDNext = s.U[nIdx];
        while (ni > 0 && ie[ni - 1] > maxEDNext)
            ni--;
        X.nDesc[slot] = nd;
        X.nIe[slot] = ni;
        if (s.dsw[nIdx] && nd > 0) {
            m.r = desc[nd - 1].r;
            u32 mn = ie[0];
            for (u32 i = 1; i < ni; i++)
                mn = tmin<u32>(mn, ie[i]);
            m.dist = mn;
        }
        pushFrame(m, nIdx, slot, F.otherSlot);
    }

    // run the frames until the root frame is done
    CLB_HD void run() {
        while (X.top >= 0 && !X.err) {
            const int f = X.top;
            Frame& F = X.fr[f];
            if (F.state == FS_POP_AFTER_CHILD) {
                X.sp = F.stackBase;
                X.top = f - 1;
                continue;
            }
            if (F.state == FS_DESC) {
                const u32 maxRow = F.g.m - 1;
                const NodeRec* desc = lsDesc(F.descSlot);
                const u32 nd = X.nDesc[F.descSlot];
                if (F.di < nd && desc[F.di].depth <= maxRow) {
                    u32 i = F.di++;
                    NodeRec cur = desc[i];
                    if (branchAndBound(f, cur, F.descSlot, i + 1)) {
                        // return from this phase (after the child phase, if any)
                        if (X.top > f)
                            F.state = FS_POP_AFTER_CHILD;
                        else {
                            X.sp = F.stackBase;
                            X.top = f - 1;
                        }
                    }
                    continue;
                }
                if (desc[nd - 1].depth == maxRow) {
                    X.sp = F.stackBase;
                    X.top = f - 1;
                    continue;
                }
                X.extendFMPos(F.dSwitch ? F.sm.r : desc[nd - 1].r, F.dir, F.uni, desc[nd - 1].depth);
                F.state = FS_LOOP;
                continue;
            }
            // FS_LOOP
            if (X.sp == F.stackBase) {
                X.top = f - 1;
                continue;
            }
            const NodeRec cur = X.sc.stack[--X.sp];
            if (branchAndBound(f, cur, -1, 0))
                continue;
            if (cur.r.width() <= X.ix.switchPoint && F.idx != 0) {
                goInText(f, cur);
                continue;
            }
            X.extendFMPos(cur.r, F.dir, F.uni, cur.depth);
        }
    }
};

// doRecSearch + recApproxMatchEditEntry for one task
template <typename W> CLB_HD void editSearchTask(SearchCtx<W>& X) {
    const DevSearch& s = X.s;
    const IndexView& ix = X.ix;
    LocOcc sm;
    sm.shift = 0;
    sm.valid = 1;
    u32 idx;
    if (s.U[0] > 0) {
        sm.r = completeRange(ix);
        sm.dist = 0;
        sm.depth = 0;
        idx = 0;
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
        sm.dist = 0;
        sm.depth = exactLength;
        idx = pis;
    }
    if (sm.r.width() > ix.switchPoint) {
        X.lc.c[SEARCH_STARTED]++;
        EditSearch<W> E(X);
        E.pushFrame(sm, idx, -1, -1);
        E.run();
    } else { // verifyExactPartialMatchInText, fmindex.cpp:245-265
        X.lc.c[IMMEDIATE_SWITCH]++;
        u32 bip = X.partBegin(s.lowestBefore(idx));
        u32 startDiff = (bip == 0) ? 0 : bip + s.maxED();
        X.emitCands(sm.r.sa, 0, startDiff, bip == 0, s.maxED(), s.minED());
    }
}

} // namespace clb
