// ORACLE (test infrastructure, NOT product code).
// CPU restatement of the reference's approximate matching engine: partitioning,
// search-scheme DFS with banded edit distance / Hamming, in-text verification,
// locate, filter.  Sequential and recursive like the reference, so it can be
// compared line by line; the product (columba_b200/csrc) is organised
// differently (batched task queues on the GPU).
#pragma once
#include "orc_matrix.h"
#include "orc_scheme.h"
#include <memory>

namespace orc {

// indexhelpers.h:1544-1655
struct Node {
    RangePair r;
    len_t depth = 0;
    char c = 0;
    bool reported = false;
    Node() {}
    Node(char c_, const RangePair& r_, len_t d) : r(r_), depth(d), c(c_) {}
};
// indexhelpers.h:1353-1535
struct FMOcc {
    RangePair r;
    len_t depth = 0, dist = 0, shift = 0;
    uint8_t strand = 0;
    FMOcc() {}
    FMOcc(const RangePair& r_, len_t dist_, len_t depth_, uint8_t strand_ = 0, len_t shift_ = 0)
        : r(r_), depth(depth_), dist(dist_), shift(shift_), strand(strand_) {}
    bool valid() const { return !r.empty(); }
    bool operator<(const FMOcc& o) const { // indexhelpers.h:1506-1522
        if (r.sa.b != o.r.sa.b)
            return r.sa.b < o.r.sa.b;
        if (dist != o.dist)
            return dist < o.dist;
        if (r.width() != o.r.width())
            return r.width() < o.r.width();
        return shift < o.shift;
    }
    bool operator==(const FMOcc& o) const { // indexhelpers.h:1529-1533
        return r.sa == o.r.sa && dist == o.dist && depth == o.depth && shift == o.shift &&
               strand == o.strand;
    }
};
// indexhelpers.h:289-930 (the fields that matter for the occurrence set)
struct TextOcc {
    len_t b = 0, e = 0, dist = 0;
    std::string cigar;
    uint8_t strand = 0;
    len_t seqID = (len_t)-1;
    len_t indexBegin = 0;
    int found = 2; // FOUND=0, FOUND_WITH_TRIMMING=1, NOT_FOUND=2
    TextOcc() {}
    TextOcc(len_t b_, len_t e_, len_t d, const std::string& cg, uint8_t s)
        : b(b_), e(e_), dist(d), cigar(cg), strand(s) {}
    len_t width() const { return e <= b ? 0 : e - b; }
    bool hasCigar() const { return !cigar.empty(); }
    bool operator<(const TextOcc& r) const { // indexhelpers.h:779-798
        if (b != r.b)
            return b < r.b;
        if (dist != r.dist)
            return dist < r.dist;
        if (width() != r.width())
            return width() < r.width();
        return hasCigar() && !r.hasCigar();
    }
    bool operator==(const TextOcc& r) const { return b == r.b && e == r.e && dist == r.dist; }
};
struct Occs {
    std::vector<FMOcc> fm;
    std::vector<TextOcc> txt;
};

// type-erased matrix (IBitParallelED, bitparallelmatrix.h:77-282)
struct IBP {
    virtual ~IBP() {}
    virtual bool sequenceSet() const = 0;
    virtual void reset() = 0;
    virtual void setSequence(const Sub& x) = 0;
    virtual void initializeMatrix(uint32_t maxED, const std::vector<uint32_t>& initED = {}) = 0;
    virtual bool computeRow(uint32_t i, char y) = 0;
    virtual bool inFinalColumn(uint32_t i) const = 0;
    virtual uint32_t at(uint32_t i, uint32_t j) const = 0;
    virtual bool onlyVerticalGapsLeft(uint32_t i) const = 0;
    virtual uint32_t firstColumn(uint32_t i) const = 0;
    virtual uint32_t nCols() const = 0;
    virtual uint32_t nRows() const = 0;
    virtual uint32_t sizeOfFinalColumn() const = 0;
    virtual void findClusterCenters(uint32_t lastRow, std::vector<len_t>& ends, uint32_t maxED,
                                    uint32_t minED) const = 0;
    virtual void traceBack(const Sub& ref, len_t refEnd, len_t& refBegin, len_t& ED,
                           std::string& cigar) const = 0;
    virtual void findCIGAR(const Sub& ref, uint32_t score, std::string& cigar) = 0;
};
template <typename W> struct BPImpl : IBP {
    BPMatrix<W> m;
    bool sequenceSet() const override { return m.sequenceSet(); }
    void reset() override { m.reset(); }
    void setSequence(const Sub& x) override { m.setSequence(x); }
    void initializeMatrix(uint32_t maxED, const std::vector<uint32_t>& initED) override {
        m.initializeMatrix(maxED, initED);
    }
    bool computeRow(uint32_t i, char y) override { return m.computeRow(i, y); }
    bool inFinalColumn(uint32_t i) const override { return m.inFinalColumn(i); }
    uint32_t at(uint32_t i, uint32_t j) const override { return m.at(i, j); }
    bool onlyVerticalGapsLeft(uint32_t i) const override { return m.onlyVerticalGapsLeft(i); }
    uint32_t firstColumn(uint32_t i) const override { return m.firstColumn(i); }
    uint32_t nCols() const override { return m.n; }
    uint32_t nRows() const override { return m.m; }
    uint32_t sizeOfFinalColumn() const override { return m.sizeOfFinalColumn(); }
    void findClusterCenters(uint32_t lastRow, std::vector<len_t>& ends, uint32_t maxED,
                            uint32_t minED) const override {
        m.findClusterCenters(lastRow, ends, maxED, minED);
    }
    void traceBack(const Sub& ref, len_t refEnd, len_t& refBegin, len_t& ED,
                   std::string& cigar) const override {
        m.traceBack(ref, refEnd, refBegin, ED, cigar);
    }
    void findCIGAR(const Sub& ref, uint32_t score, std::string& cigar) override {
        m.findCIGAR(ref, score, cigar);
    }
};

// indexhelpers.h:1677-1838 + indexhelpers.cpp:276-382
struct Cluster {
    std::vector<uint16_t> eds;
    std::vector<Node> nodes;
    len_t lastCell = (len_t)-1;
    uint16_t maxED;
    len_t startDepth, shift;
    Cluster(len_t size, len_t maxED_, len_t startDepth_, len_t shift_)
        : eds(size, maxED_ + 1), nodes(size), maxED(maxED_), startDepth(startDepth_), shift(shift_) {}
    len_t size() const { return (len_t)eds.size(); }
    void setValue(len_t idx, const Node& n, len_t ed) {
        eds[idx] = ed;
        nodes[idx] = n;
        lastCell = idx;
    }
    // Node::report, indexhelpers.h:1613-1629
    static void report(Node& n, FMOcc& occ, len_t startDepth, len_t ed, bool once, len_t shift) {
        if (!n.reported) {
            occ = FMOcc(n.r, ed, n.depth + startDepth, 0, shift);
            if (once)
                n.reported = true;
        }
    }
    std::vector<FMOcc> reportCentersAtEnd() {
        std::vector<FMOcc> c;
        for (len_t i = 0; i <= lastCell; i++) {
            if (eds[i] <= maxED && (i == 0 || eds[i] <= eds[i - 1]) &&
                (i == lastCell || eds[i] <= eds[i + 1])) {
                FMOcc m;
                report(nodes[i], m, startDepth, eds[i], true, shift);
                c.push_back(m);
            }
        }
        return c;
    }
    FMOcc reportDeepestMinimum(Dir dir) {
        uint16_t minED = maxED + 1;
        len_t hi = (len_t)-1, deep = (len_t)-1;
        for (len_t i = 0; i <= lastCell; i++) {
            if (eds[i] < minED) {
                minED = eds[i];
                hi = i;
                deep = i;
            }
            if (eds[i] == minED)
                deep = i;
        }
        FMOcc m;
        if (minED <= maxED)
            report(nodes[deep], m, startDepth - (deep - hi), minED, true,
                   ((dir == BWD) ? (deep - hi) : 0) + shift);
        return m;
    }
    FMOcc getClusterCentra(uint16_t lowerBound, std::vector<Node>& desc, std::vector<uint16_t>& ie) {
        FMOcc m;
        for (len_t i = 0; i <= lastCell; i++) {
            if (eds[i] > maxED || eds[i] < lowerBound)
                continue;
            bool bp = (i == 0) || eds[i] <= eds[i - 1];
            bool bc = (i == lastCell) || eds[i] <= eds[i + 1];
            if (!(bp && bc))
                continue;
            report(nodes[i], m, startDepth, eds[i], false, shift);
            ie.push_back(eds[i]);
            for (len_t j = i + 1; j <= lastCell; j++) {
                desc.push_back(nodes[j]);
                ie.push_back(eds[j]);
            }
            for (len_t k = 1; k < ie.size(); k++) {
                if (ie[k] < lowerBound && ie[k] <= ie[k - 1] &&
                    (k == ie.size() - 1 || ie[k] <= ie[k + 1])) {
                    len_t highestPoint = 0, lowestPoint = (len_t)ie.size() - 1;
                    for (len_t l = k; l-- > 0;)
                        if (ie[l] != ie[l + 1] + 1) {
                            highestPoint = l + 1;
                            break;
                        }
                    for (len_t l = k + 1; l < ie.size(); l++)
                        if (ie[l] != ie[l - 1] + 1) {
                            lowestPoint = l - 1;
                            break;
                        }
                    if (highestPoint != 0 && lowestPoint != ie.size() - 1) {
                        len_t lC = lowestPoint, hC = highestPoint;
                        bool highest = true;
                        while (lC > hC) {
                            if (highest) {
                                ie[hC] = std::min(maxED + 1, ie[hC - 1] + 1);
                                hC++;
                            } else {
                                ie[lC] = std::min(maxED + 1, ie[lC + 1] + 1);
                                lC--;
                            }
                            highest = !highest;
                        }
                        if (lC == hC)
                            ie[lC] = std::min(ie[lC + 1] + 1, ie[lC - 1] + 1);
                    } else if (highestPoint == 0 && lowestPoint != ie.size() - 1) {
                        for (len_t l = lowestPoint; l-- > 0;)
                            ie[l] = ie[l + 1] + 1;
                    } else if (highestPoint != 0 && lowestPoint == ie.size() - 1) {
                        for (len_t l = highestPoint; l < ie.size(); l++)
                            ie[l] = ie[l - 1] + 1;
                    }
                }
            }
            break;
        }
        return m;
    }
};

static inline std::string revComp(const std::string& s) { // nucleotide.h
    std::string r(s.size(), 'N');
    for (size_t i = 0; i < s.size(); i++) {
        char c = s[s.size() - 1 - i], o;
        switch (c) {
        case 'A': o = 'T'; break;
        case 'C': o = 'G'; break;
        case 'G': o = 'C'; break;
        case 'T': o = 'A'; break;
        default: o = 'N';
        }
        r[i] = o;
    }
    return r;
}

struct Part { // a Substring of the read with a direction
    len_t b, e;
    Dir d = FWD;
    len_t size() const { return e <= b ? 0 : e - b; }
};

struct Mapper {
    Index& ix;
    const Strategy& st;
    uint8_t strand = 0;
    // per-call scratch, like the reference's thread_local members
    std::vector<std::vector<Node>> stacks;
    std::vector<std::unique_ptr<BPImpl<uint64_t>>> mats;
    std::vector<std::unique_ptr<BPImpl<u128>>> mats128;
    BPImpl<uint64_t> full64[2];
    BPImpl<u128> full128[2];
    const std::string* curSeq = nullptr; // read in the orientation being mapped

    Mapper(Index& ix_, const Strategy& st_) : ix(ix_), st(st_) {}

    Sub partSub(const Part& p) const { return Sub(curSeq, p.b, p.e, p.d); }

    // ---------------- extend (indexinterface.cpp:675-697) ----------------
    void extendFMPos(const RangePair& parent, std::vector<Node>& stack, Counters& ct, len_t row) {
        for (int i = 1; i < 5; ++i) {
            RangePair ch;
            if (ix.extraChar(i, parent, ch)) {
                stack.emplace_back(I2C[i], ch, row + 1);
                ct.inc(NODE_COUNTER);
            }
        }
    }
    // indexinterface.cpp:1016-1032
    RangePair matchString(const Sub& pat, RangePair r, Counters& ct) {
        for (len_t i = 0; i < pat.size(); i++)
            if (!ix.addChar(pat[i], r, ct))
                break;
        return r;
    }

    // ---------------- partitioning (searchstrategy.cpp:141-419) ----------
    int seed(const std::string& pat, std::vector<Part>& parts, int numParts, int maxScore,
             std::vector<RangePair>& emr) {
        size_t pSize = pat.size();
        bool useK = ((size_t)numParts * ix.wordSize < (pSize * 2) / 3) && pSize >= st.useKmerCutOff;
        int wSize = useK ? (int)ix.wordSize : 1;
        auto sp = st.seedingPositions(numParts, maxScore);
        std::vector<int> seeds;
        seeds.push_back(0);
        for (int i = 1; i < numParts - 1; i++)
            seeds.push_back((int)((sp[i - 1] * pSize) - (wSize / 2)));
        for (int i = 0; i < numParts - 1; i++)
            parts.push_back(Part{(len_t)seeds[i], (len_t)(seeds[i] + wSize)});
        parts.push_back(Part{(len_t)(pSize - wSize), (len_t)pSize});
        emr.resize(numParts);
        for (int i = 0; i < numParts; i++)
            emr[i] = useK ? ix.lookUpKmer(pat.data() + parts[i].b) : ix.rangeOfSingleChar(pat[parts[i].b]);
        return numParts * wSize;
    }
    void partitionDynamic(const std::string& pat, std::vector<Part>& parts, int numParts, int maxScore,
                          std::vector<RangePair>& emr, Counters& ct) {
        int matched = seed(pat, parts, numParts, maxScore, emr);
        len_t pSize = (len_t)pat.size();
        auto weights = st.weights(numParts, maxScore);
        Dir dir = FWD;
        int partToExtend = 0;
        if (matched == 0) { // seedIfNoKmers, searchstrategy.cpp:255-271
            len_t size = (pSize >= 4 * parts.size()) ? 4 : 1;
            for (int i = 0; i < numParts - 1; i++) {
                parts[i].e = std::min(parts[i].e + size, pSize);
                ix.setDirection(FWD, false);
                for (len_t j = parts[i].b; j < parts[i].e; j++) {
                    ix.addChar(pat[j], emr.at(i), ct);
                    matched++;
                }
            }
        }
        for (len_t j = matched; j < pSize; j++) {
            uint64_t maxW = 0;
            for (int i = 0; i < numParts; i++) {
                bool noL = (i == 0) || parts[i].b == parts[i - 1].e;
                bool noR = (i == numParts - 1) || parts[i].e == parts[i + 1].b;
                if (noL && noR)
                    continue;
                if (emr[i].width() * weights[i] > maxW) {
                    maxW = emr[i].width() * weights[i];
                    partToExtend = i;
                    if (noL)
                        dir = FWD;
                    else if (noR)
                        dir = BWD;
                    else
                        dir = (emr[i - 1].width() < emr[i + 1].width()) ? BWD : FWD;
                }
            }
            if (maxW == 0) { // extendParts, searchstrategy.cpp:283-297
                for (len_t i = 0; i < parts.size(); i++) {
                    if ((i != parts.size() - 1) && (parts[i].e != parts[i + 1].b))
                        parts[i].e = parts[i + 1].b;
                    if ((i != 0) && (parts[i].b != parts[i - 1].e))
                        parts[i].b = parts[i - 1].e;
                }
                return;
            }
            char c;
            if (dir == FWD) {
                parts[partToExtend].e++;
                c = pat[parts[partToExtend].e - 1];
            } else {
                parts[partToExtend].b--;
                c = pat[parts[partToExtend].b];
            }
            ix.setDirection(dir, partToExtend == numParts - 1);
            ix.addChar(c, emr.at(partToExtend), ct);
        }
    }
    void calcExactRanges(const std::string& pat, std::vector<Part>& parts, std::vector<RangePair>& emr,
                         Counters& ct) { // searchstrategy.cpp:158-190
        ix.setDirection(FWD, false);
        len_t ws = ix.wordSize;
        for (len_t i = 0; i < parts.size(); ++i) {
            Part& cur = parts[i];
            len_t size = cur.size();
            len_t start = cur.b + ((size >= ws) ? ws : 0);
            RangePair init = (size >= ws) ? ix.lookUpKmer(pat.data() + cur.b) : ix.completeRange();
            emr[i] = matchString(Sub(curSeq, start, cur.e, cur.d), init, ct);
        }
        ix.setDirection(BWD, true);
        Part& last = parts.back();
        last.d = BWD;
        len_t size = last.size();
        len_t end = (size >= ws) ? last.e - ws : last.e;
        RangePair init = (size >= ws) ? ix.lookUpKmer(pat.data() + end) : ix.completeRange();
        emr.back() = matchString(Sub(curSeq, last.b, end, BWD), init, ct);
    }
    void partition(const std::string& pat, std::vector<Part>& parts, int numParts, int maxScore,
                   std::vector<RangePair>& emr, Counters& ct) {
        parts.clear();
        if (numParts >= (int)pat.size() || numParts == 1)
            return;
        if (st.part == DYNAMIC) {
            partitionDynamic(pat, parts, numParts, maxScore, emr, ct);
        } else if (st.part == UNIFORM) { // searchstrategy.cpp:194-209
            for (int i = 0; i < numParts; i++)
                parts.push_back(Part{(len_t)((i * 1.0 / numParts) * pat.size()),
                                     (len_t)(((i + 1) * 1.0 / numParts) * pat.size())});
            parts.back().e = (len_t)pat.size();
            calcExactRanges(pat, parts, emr, ct);
        } else { // searchstrategy.cpp:212-238
            auto begins = st.begins(numParts, maxScore);
            int pSize = (int)pat.size();
            parts.push_back(Part{0, (len_t)(begins[0] * pSize)});
            for (unsigned i = 0; i + 1 < begins.size(); i++)
                parts.push_back(Part{(len_t)(begins[i] * pSize), (len_t)(begins[i + 1] * pSize)});
            parts.push_back(Part{(len_t)(begins.back() * pSize), (len_t)pat.size()});
            calcExactRanges(pat, parts, emr, ct);
        }
    }

    // ---------------- locate (fmindex.h:366-382) -------------------------
    std::vector<len_t> beginPositions(const Range& sa, len_t startDiff, len_t shift = 0) {
        std::vector<len_t> pos(sa.width(), 0);
        for (len_t i = sa.b; i < sa.e; i++) {
            len_t sum = ix.findSA(i) + shift;
            pos[i - sa.b] = sum >= startDiff ? sum - startDiff : 0;
        }
        return pos;
    }

    // ---------------- in-text verification (edit) ------------------------
    static bool use64(len_t nZeros, len_t maxED) { // fmindex.h:240-246
        return BP64::LEFT >= nZeros + maxED && BP64::MAX_ED >= maxED;
    }
    IBP* initFull(len_t nZeros, len_t maxED) { // fmindex.h:247-282
        IBP* mat = use64(nZeros, maxED) ? (IBP*)&full64[strand] : (IBP*)&full128[strand];
        if (!mat->sequenceSet())
            mat->setSequence(Sub(curSeq, 0, (len_t)curSeq->size(), FWD));
        mat->initializeMatrix(maxED, std::vector<uint32_t>(nZeros, 0));
        return mat;
    }
    // indexhelpers.cpp:517-574
    void doTask(IBP* mat, const std::vector<Sub>& refs, len_t maxED, len_t minED, Occs& occ,
                Counters& ct) {
        ct.inc(IN_TEXT_STARTED, refs.size());
        for (const Sub& ref : refs) {
            len_t refBegin = ref.b, i;
            const len_t size = (len_t)ref.size();
            if (!mat->inFinalColumn(size))
                continue;
            for (i = 0; i < size; ++i)
                if (!mat->computeRow(i + 1, ref.fwdAt(i)))
                    break;
            if (i <= size - mat->sizeOfFinalColumn()) {
                ct.inc(ABORTED_IN_TEXT_VERIF);
                continue;
            }
            std::vector<len_t> ends;
            mat->findClusterCenters(i, ends, maxED, minED);
            if (ends.empty()) {
                ct.inc(ABORTED_IN_TEXT_VERIF);
                continue;
            }
            for (len_t refEnd : ends) {
                len_t score = maxED + 1, beg = 0;
                std::string cigar;
                mat->traceBack(ref, refEnd, beg, score, cigar);
                ct.inc(CIGARS_IN_TEXT_VERIFICATION);
                if (ix.noCIGAR)
                    cigar = "";
                occ.txt.emplace_back(refBegin + beg, refBegin + refEnd, score, cigar, strand);
            }
        }
    }
    // fmindex.cpp:267-310
    void inTextVerification(const std::vector<len_t>& startPos, len_t maxED, len_t minED, Occs& occ,
                            Counters& ct, bool fixedStart) {
        len_t nZeros = fixedStart ? 1 : 2 * maxED + 1;
        IBP* mat = initFull(nZeros, maxED);
        len_t nRows = mat->nRows();
        std::vector<Sub> refs;
        for (len_t start : startPos) {
            len_t maxEnd = ix.textLength - 1;
            len_t hEnd = std::min(maxEnd, start + nRows - 1);
            refs.emplace_back(&ix.text, start, hEnd);
        }
        doTask(mat, refs, maxED, minED, occ, ct);
    }
    // fmindex.cpp:245-265
    void verifyExactPartialMatchInText(const FMOcc& sm, len_t beginInPattern, len_t maxED, Occs& occ,
                                       Counters& ct, len_t minED) {
        ct.inc(IMMEDIATE_SWITCH);
        len_t startDiff = (beginInPattern == 0) ? 0 : beginInPattern + maxED;
        auto starts = beginPositions(sm.r.sa, startDiff);
        inTextVerification(starts, maxED, minED, occ, ct, beginInPattern == 0);
    }
    // fmindex.cpp:312-342
    void inTextVerificationOneString(len_t startPos, len_t endPos, len_t maxED, len_t minED, Occs& occ,
                                     Counters& ct) {
        IBP* mat = initFull(1, maxED);
        std::vector<Sub> refs = {Sub(&ix.text, startPos, endPos)};
        doTask(mat, refs, maxED, minED, occ, ct);
    }

    // ---------------- edit distance DFS (indexinterface.cpp:340-669) -----
    IBP* phaseMatrix(len_t maxED, size_t matrixIdx) {
        if (maxED <= BP64::MAX_ED)
            return mats[matrixIdx].get();
        return mats128[matrixIdx].get();
    }
    void goToInTextVerificationEdit(const Node& node, const Search& s, const std::vector<Part>& parts,
                                    Occs& occ, Counters& ct, len_t idx, IBP* bp, const FMOcc& sMatch,
                                    const std::vector<Node>& dOther, const std::vector<uint16_t>& iOther) {
        len_t stp = parts[s.lowestBefore(idx)].b;
        len_t startDiff = stp + s.maxED();
        if (stp == 0) {
            startDiff = 0;
        } else if (ix.dir == BWD) {
            len_t row = node.depth;
            len_t col = bp->firstColumn(row);
            startDiff -= col + bp->at(row, col);
        } else if (!dOther.empty()) {
            startDiff -= (len_t)(dOther.size() - iOther.size() + iOther.back());
        }
        auto pos = beginPositions(node.r.sa, startDiff, sMatch.shift);
        inTextVerification(pos, s.maxED(), s.minED(), occ, ct, stp == 0);
    }
    void recApproxMatchEdit(const Search& s, const FMOcc& startMatch, Occs& occ,
                            const std::vector<Part>& parts, Counters& ct, int idx,
                            const std::vector<Node>& descPrevDir = {},
                            const std::vector<uint16_t>& initPrevDir = {},
                            const std::vector<Node>& descNotPrevDir = {},
                            const std::vector<uint16_t>& initNotPrevDir = {}) {
        const Part& p = parts[s.order[idx]];
        const len_t maxED = s.U[idx];
        const Dir dir = s.directions[idx];
        const bool dSwitch = s.dswitch[idx];
        auto& stack = stacks[idx];
        size_t matrixIdx = s.order[idx] + (dir == BWD) * s.numParts();
        IBP* bp = phaseMatrix(maxED, matrixIdx);
        const auto& initEds = dSwitch ? initNotPrevDir : initPrevDir;
        const auto& descendants = dSwitch ? descNotPrevDir : descPrevDir;
        const auto& initOther = dSwitch ? initPrevDir : initNotPrevDir;
        const auto& descOther = dSwitch ? descPrevDir : descNotPrevDir;
        ix.setDirection(dir, s.isUniBack(idx));
        std::vector<uint32_t> initED;
        if (initEds.empty()) {
            initED = std::vector<uint32_t>(1, startMatch.dist);
        } else {
            uint16_t prevED = dSwitch ? *std::min_element(initEds.begin(), initEds.end()) : initEds[0];
            uint32_t increase = startMatch.dist - prevED;
            initED.resize(initEds.size());
            for (size_t i = 0; i < initED.size(); i++)
                initED[i] = initEds[i] + increase;
        }
        if (!bp->sequenceSet())
            bp->setSequence(Sub(curSeq, p.b, p.e, p.d));
        bp->initializeMatrix(maxED, initED);
        Cluster cluster(bp->sizeOfFinalColumn(), maxED, startMatch.depth, startMatch.shift);
        if (bp->inFinalColumn(0))
            cluster.setValue(0, Node((char)0, startMatch.r, 0), bp->at(0, p.size()));
        if (!descendants.empty()) {
            len_t maxRow = bp->nRows() - 1;
            for (len_t i = 0; i < descendants.size() && descendants[i].depth <= maxRow; i++) {
                std::vector<Node> rem(descendants.begin() + i + 1, descendants.end());
                if (branchAndBound(bp, cluster, descendants[i], s, idx, parts, occ, ct, initOther,
                                   descOther, rem))
                    return;
            }
            if (descendants.back().depth == maxRow)
                return;
            RangePair pair = dSwitch ? startMatch.r : descendants.back().r;
            extendFMPos(pair, stack, ct, descendants.back().depth);
        } else {
            extendFMPos(startMatch.r, stack, ct, 0);
        }
        bool idxZero = idx == 0;
        while (!stack.empty()) {
            const Node cur = stack.back();
            stack.pop_back();
            if (branchAndBound(bp, cluster, cur, s, idx, parts, occ, ct, initOther, descOther, {}))
                continue;
            if (cur.r.width() <= ix.switchPoint && !idxZero) {
                goToInTextVerificationEdit(cur, s, parts, occ, ct, idx, bp, startMatch, descOther,
                                           initOther);
                continue;
            }
            extendFMPos(cur.r, stack, ct, cur.depth);
        }
    }
    bool branchAndBound(IBP* bp, Cluster& cluster, const Node& cur, const Search& s, len_t idx,
                        const std::vector<Part>& parts, Occs& occ, Counters& ct,
                        const std::vector<uint16_t>& initOther, const std::vector<Node>& descOther,
                        const std::vector<Node>& remDesc) {
        const len_t row = cur.depth;
        bool valid = bp->computeRow(row, cur.c);
        if (bp->inFinalColumn(row)) {
            len_t cIdx = cluster.size() + row - bp->nRows();
            cluster.setValue(cIdx, cur, bp->at(row, bp->nCols() - 1));
            if (!valid || bp->onlyVerticalGapsLeft(row)) {
                goDeeper(cluster, idx + 1, s, parts, occ, ct, descOther, initOther, remDesc);
                return true;
            }
        }
        return !valid;
    }
    void goDeeper(Cluster& cluster, len_t nIdx, const Search& s, const std::vector<Part>& parts,
                  Occs& occ, Counters& ct, const std::vector<Node>& descOtherD,
                  const std::vector<uint16_t>& initOtherD, const std::vector<Node>& remDesc) {
        bool isEdge = s.isEdge(nIdx - 1);
        const len_t lowerBound = s.L[nIdx - 1];
        if (isEdge) {
            if (nIdx == parts.size()) {
                auto matches = cluster.reportCentersAtEnd();
                for (auto& m : matches)
                    if (m.valid() && m.dist >= lowerBound) {
                        m.strand = strand;
                        occ.fm.push_back(m);
                    }
            } else {
                FMOcc m = cluster.reportDeepestMinimum(ix.dir);
                if (m.valid() && m.dist >= lowerBound) {
                    Dir od = ix.dir;
                    recApproxMatchEdit(s, m, occ, parts, ct, nIdx, {}, {}, descOtherD, initOtherD);
                    ix.setDirection(od, s.isUniBack(nIdx - 1));
                }
            }
            return;
        }
        std::vector<Node> descendants;
        std::vector<uint16_t> initEds;
        FMOcc newMatch = cluster.getClusterCentra(lowerBound, descendants, initEds);
        if (!newMatch.valid())
            return;
        descendants.insert(descendants.end(), remDesc.begin(), remDesc.end());
        for (len_t i = 0; i < descendants.size(); i++)
            descendants[i].depth = i + 1;
        len_t maxEDNext = s.U[nIdx];
        while (initEds.back() > maxEDNext)
            initEds.pop_back();
        bool switchAfter = s.dswitch[nIdx];
        if (switchAfter) {
            if (!descendants.empty()) {
                newMatch.r = descendants.back().r;
                newMatch.dist = *std::min_element(initEds.begin(), initEds.end());
            }
            Dir od = ix.dir;
            recApproxMatchEdit(s, newMatch, occ, parts, ct, nIdx, descendants, initEds, descOtherD,
                               initOtherD);
            ix.setDirection(od, s.isUniBack(nIdx - 1));
        } else {
            recApproxMatchEdit(s, newMatch, occ, parts, ct, nIdx, descendants, initEds, descOtherD,
                               initOtherD);
        }
    }
    // indexinterface.cpp:1306-1325
    void recApproxMatchEditEntry(const Search& s, const FMOcc& sm, Occs& occ,
                                 const std::vector<Part>& parts, Counters& ct, int idx) {
        if (sm.r.width() > ix.switchPoint) {
            ct.inc(SEARCH_STARTED);
            recApproxMatchEdit(s, sm, occ, parts, ct, idx);
            return;
        }
        verifyExactPartialMatchInText(sm, parts[s.lowestBefore(idx)].b, s.maxED(), occ, ct, s.minED());
    }

    // ---------------- Hamming (indexinterface.cpp:1211-1304, fmindex.cpp:344-428)
    void inTextHamming(const Range& r, len_t maxD, len_t minD, len_t lengthBefore, Occs& occ,
                       Counters& ct) {
        const len_t pSize = (len_t)curSeq->size();
        std::string cigar = ix.noCIGAR ? "*" : std::to_string(pSize) + "M";
        for (len_t i = r.b; i < r.e; i++) {
            len_t Tb = ix.findSA(i);
            ct.inc(IN_TEXT_STARTED);
            Tb = (Tb > lengthBefore) ? Tb - lengthBefore : 0;
            len_t Te = Tb + pSize;
            if (Te > ix.textLength)
                continue;
            len_t score = 0;
            for (len_t j = 0; j < pSize; j++) {
                score += (ix.text[Tb + j] != (*curSeq)[j]);
                if (score > maxD)
                    break;
            }
            if (score <= maxD && score >= minD)
                occ.txt.emplace_back(Tb, Te, score, cigar, strand);
        }
    }
    void recApproxMatchHamming(const Search& s, const FMOcc& sm, Occs& occ,
                               const std::vector<Part>& parts, Counters& ct, int idx) {
        const Part& p = parts[s.order[idx]];
        Sub ps = partSub(p);
        const len_t pSize = p.size();
        const Dir d = s.directions[idx];
        const len_t maxED = s.U[idx], minED = s.L[idx];
        ix.setDirection(d, s.isUniBack(idx));
        std::vector<len_t> vec(p.size() + 1, 0);
        vec[0] = sm.dist;
        auto& stack = stacks[idx];
        extendFMPos(sm.r, stack, ct, 0);
        while (!stack.empty()) {
            const Node node = stack.back();
            stack.pop_back();
            if (node.r.width() <= ix.switchPoint) {
                len_t lengthBefore = ((idx == 0) ? 0 : parts[s.lowestBefore(idx)].b) -
                                     (ix.dir == BWD) * node.depth;
                inTextHamming(node.r.sa, s.maxED(), s.minED(), lengthBefore, occ, ct);
                continue;
            }
            len_t row = node.depth;
            vec[row] = vec[row - 1] + (node.c != ps[row - 1]);
            if (vec[row] > maxED)
                continue;
            if (row == pSize) {
                if (vec[row] >= minED) {
                    FMOcc m(node.r, vec[row], sm.depth + pSize, strand);
                    if (s.isEnd(idx)) {
                        occ.fm.push_back(m);
                    } else {
                        recApproxMatchHamming(s, m, occ, parts, ct, idx + 1);
                        ix.setDirection(s.directions[idx], s.isUniBack(idx));
                    }
                }
                continue;
            }
            extendFMPos(node.r, stack, ct, node.depth);
        }
    }

    // ---------------- search drivers (searchstrategy.cpp:425-493,1181-1254)
    void startIdx(const Search& s, const FMOcc& sm, Occs& occ, const std::vector<Part>& parts,
                  Counters& ct, int idx) {
        if (st.metric == EDIT)
            recApproxMatchEditEntry(s, sm, occ, parts, ct, idx);
        else
            recApproxMatchHamming(s, sm, occ, parts, ct, idx);
    }
    void doRecSearch(const Search& s, std::vector<Part>& parts, Occs& occ,
                     const std::vector<RangePair>& emr, Counters& ct) {
        if (s.U[0] > 0) {
            for (len_t i = 0; i < s.order.size(); i++)
                parts[s.order[i]].d = s.directions[i];
            startIdx(s, FMOcc(ix.completeRange(), 0, 0), occ, parts, ct, 0);
            return;
        }
        int first = s.order[0];
        RangePair startRange = emr[first];
        if (startRange.width() > ix.switchPoint) {
            for (len_t i = 0; i < s.order.size(); i++)
                parts[s.order[i]].d = s.directions[i];
            uint16_t pis = 1;
            len_t exactLength = parts[first].size();
            while (s.U[pis] == 0) {
                ix.setDirection(s.directions[pis], s.isUniBack(pis));
                const Part& part = parts[s.order[pis]];
                startRange = matchString(partSub(part), startRange, ct);
                if (startRange.empty())
                    return;
                exactLength += part.size();
                pis++;
            }
            startIdx(s, FMOcc(startRange, 0, exactLength), occ, parts, ct, pis);
        }
    }
    void matchWithSearches(const std::string& seq, len_t k, Counters& ct, Occs& occ, len_t minD = 0) {
        curSeq = &seq;
        len_t numParts = st.numParts(k);
        std::vector<RangePair> emr(numParts);
        std::vector<Part> parts;
        partition(seq, parts, numParts, k, emr, ct);
        if (parts.empty())
            throw std::runtime_error("oracle: read shorter than the number of parts (naive path not restated)");
        for (uint16_t i = 0; i < numParts; i++) {
            size_t width = emr[i].width();
            if (width != 0 && width <= ix.switchPoint) {
                FMOcc sm(emr[i], 0, parts[i].size());
                if (st.metric == EDIT)
                    verifyExactPartialMatchInText(sm, parts[i].b, k, occ, ct, minD);
                else
                    inTextHamming(sm.r.sa, k, minD, parts[i].b, occ, ct);
            }
        }
        const auto& searches = st.createSearches(k, emr);
        stacks.assign(numParts, {});
        mats.clear();
        mats128.clear();
        for (size_t i = 0; i < 2 * parts.size(); i++) {
            mats.emplace_back(new BPImpl<uint64_t>());
            mats128.emplace_back(new BPImpl<u128>());
        }
        for (const Search& s : searches)
            doRecSearch(s, parts, occ, emr, ct);
    }

    // ---------------- filters (indexinterface.cpp:1331-1491) -------------
    void generateCIGAR(TextOcc& t, Counters& ct, const std::string& read) { // indexinterface.h:959-997
        if (ix.noCIGAR || t.hasCigar())
            return;
        Sub ref(&ix.text, t.b, t.e);
        IBP* mat = (BP64::MAX_ED >= t.dist) ? (IBP*)&full64[t.strand] : (IBP*)&full128[t.strand];
        if (!mat->sequenceSet())
            mat->setSequence(Sub(&read, 0, (len_t)read.size(), FWD));
        std::string cg;
        mat->findCIGAR(ref, t.dist, cg);
        t.cigar = cg;
        ct.inc(CIGARS_IN_INDEX);
    }
    std::vector<TextOcc> uniqueTextOccurrences(Occs& occ, len_t maxED, Counters& ct,
                                               const std::string* seqs[2]) {
        ct.inc(TOTAL_REPORTED_POSITIONS, occ.txt.size());
        std::stable_sort(occ.fm.begin(), occ.fm.end());
        occ.fm.erase(std::unique(occ.fm.begin(), occ.fm.end()), occ.fm.end());
        for (const auto& f : occ.fm) {
            ct.inc(TOTAL_REPORTED_POSITIONS, f.r.sa.width());
            const bool calc = !ix.noCIGAR && (f.r.sa.width() > 10); // CIGAR_THRESHOLD
            std::string cg;
            bool first = true;
            for (len_t i = f.r.sa.b; i < f.r.sa.e; i++) {
                len_t startPos = ix.findSA(i) + f.shift;
                TextOcc t(startPos, startPos + f.depth, f.dist, "", f.strand);
                if (calc) {
                    if (first) {
                        generateCIGAR(t, ct, *seqs[t.strand]);
                        cg = t.cigar;
                        first = false;
                    } else
                        t.cigar = cg;
                }
                occ.txt.push_back(t);
            }
        }
        std::stable_sort(occ.txt.begin(), occ.txt.end());
        occ.txt.erase(std::unique(occ.txt.begin(), occ.txt.end()), occ.txt.end());
        std::vector<TextOcc> out;
        len_t maxDiff = 2 * maxED, prevBegin = (len_t)-1, prevDepth = (len_t)-1, prevED = maxED + 1;
        for (auto& o : occ.txt) {
            len_t diff = o.b > prevBegin ? o.b - prevBegin : prevBegin - o.b;
            if (diff == 0)
                continue;
            if (diff <= maxDiff) {
                if (o.dist > prevED || (o.dist == prevED && o.width() >= prevDepth))
                    continue;
                out.pop_back();
            }
            prevBegin = o.b;
            prevED = o.dist;
            prevDepth = o.width();
            out.push_back(o);
        }
        return out;
    }
    std::vector<TextOcc> textOccHamming(Occs& occ, Counters& ct) {
        ct.inc(TOTAL_REPORTED_POSITIONS, occ.txt.size());
        std::stable_sort(occ.fm.begin(), occ.fm.end());
        occ.fm.erase(std::unique(occ.fm.begin(), occ.fm.end()), occ.fm.end());
        len_t size = occ.fm.empty() ? 0 : occ.fm[0].depth;
        std::string cigar = ix.noCIGAR ? "*" : std::to_string(size) + "M";
        for (const auto& f : occ.fm) {
            ct.inc(TOTAL_REPORTED_POSITIONS, f.r.sa.width());
            for (len_t i = f.r.sa.b; i < f.r.sa.e; i++) {
                len_t p = ix.findSA(i);
                occ.txt.emplace_back(p, p + size, f.dist, cigar, f.strand);
            }
        }
        std::stable_sort(occ.txt.begin(), occ.txt.end());
        occ.txt.erase(std::unique(occ.txt.begin(), occ.txt.end()), occ.txt.end());
        return occ.txt;
    }

    // ---------------- boundary trimming (indexinterface.cpp:799-897) -----
    int findSeqName(TextOcc& t, len_t& seqID, Counters& ct, len_t largestStratum) {
        const len_t begin = t.b, end = t.e;
        auto it = std::upper_bound(ix.startPos.begin(), ix.startPos.end(), begin);
        len_t index = (len_t)std::distance(ix.startPos.begin(), --it);
        if (end <= ix.startPos[index + 1]) {
            t.indexBegin = begin;
            t.b = begin - ix.startPos[index];
            t.e = end - ix.startPos[index];
            seqID = index;
            return 0;
        }
        if (st.metric == HAMMING)
            return 2;
        len_t nb, ne;
        if ((ix.startPos[index + 1] - begin) <= largestStratum) {
            index++;
            nb = ix.startPos[index];
            ne = std::min(end, ix.startPos[index + 1]);
        } else if ((end - ix.startPos[index + 1]) <= largestStratum) {
            nb = begin;
            ne = ix.startPos[index + 1];
        } else
            return 2;
        Occs occ;
        inTextVerificationOneString(nb, ne, largestStratum, 0, occ, ct);
        if (occ.txt.empty())
            return 2;
        t = *std::min_element(occ.txt.begin(), occ.txt.end());
        seqID = index;
        len_t sp = ix.startPos[index];
        t.indexBegin = t.b;
        t.b -= sp;
        t.e -= sp;
        return 1;
    }

    // ---------------- ALL mode (searchstrategy.cpp:495-535) --------------
    // returns the occurrences that would be written to SAM (assigned to a
    // sequence), in concatenated-text order before the distance sort.
    std::vector<TextOcc> matchApproxAll(const std::string& read, len_t maxED, Counters& ct) {
        std::string rc = revComp(read);
        const std::string* seqs[2] = {&read, &rc};
        std::vector<TextOcc> result;
        for (auto& m : full64)
            m.reset();
        for (auto& m : full128)
            m.reset();
        if (maxED == 0)
            throw std::runtime_error("oracle: k=0 exact path not restated yet");
        Occs occ;
        strand = 0;
        matchWithSearches(read, maxED, ct, occ);
        strand = 1;
        matchWithSearches(rc, maxED, ct, occ);
        if (st.metric == EDIT) {
            result = uniqueTextOccurrences(occ, maxED, ct, seqs);
            for (auto& t : result)
                if (!t.hasCigar())
                    generateCIGAR(t, ct, *seqs[t.strand]);
        } else {
            result = textOccHamming(occ, ct);
        }
        // generateOutputSingleEnd (searchstrategy.cpp:1824-1902)
        std::vector<TextOcc> assigned;
        for (auto& t : result) {
            strand = t.strand;
            curSeq = seqs[t.strand];
            len_t seqID;
            int found = findSeqName(t, seqID, ct, maxED);
            if (found != 2) {
                t.seqID = seqID;
                t.found = found;
                assigned.push_back(t);
            }
        }
        return assigned;
    }
};

} // namespace orc
