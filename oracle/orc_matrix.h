// ORACLE (test infrastructure, NOT product code).
// CPU restatement of the reference's banded bit-parallel (Hyyro/Myers) edit
// distance matrix: bitparallelmatrix.h:300-750, bitparallelmatrix.cpp:33-123.
#pragma once
#include "orc_index.h"
#include <array>

namespace orc {

typedef unsigned __int128 u128;

static inline int popc(uint64_t x) { return __builtin_popcountll(x); }
static inline int popc(u128 x) {
    return __builtin_popcountll((uint64_t)(x >> 64)) + __builtin_popcountll((uint64_t)x);
}

// direction-aware view on a string (substring.h)
struct Sub {
    const std::string* t = nullptr;
    len_t b = 0, e = 0;
    Dir d = FWD;
    Sub() {}
    Sub(const std::string* t_, len_t b_, len_t e_, Dir d_ = FWD) : t(t_), b(b_), e(e_), d(d_) {
        if (e > t->size())
            e = (len_t)t->size();
    }
    size_t size() const { return e <= b ? 0 : e - b; }
    char fwdAt(len_t i) const { return (*t)[b + i]; }
    char operator[](len_t i) const { return d == FWD ? (*t)[b + i] : (*t)[e - i - 1]; }
};

static inline int mvIdx(char c) { // bitparallelmatrix.h:85-93
    switch (c) {
    case 'A': return 0;
    case 'C': return 1;
    case 'G': return 2;
    case 'T': return 3;
    case 'N': return 4;
    default: return 5;
    }
}

template <typename W> struct BPMatrix {
    static const uint32_t WORD = sizeof(W) * 8;
    static const uint32_t BLOCK = WORD / 2;
    static const uint32_t MAX_ED = (WORD - BLOCK - 2) / 3; // 10 / 20
    static const uint32_t LEFT = 2 * MAX_ED + 1;           // 21 / 41
    static const uint32_t DIAG_R0 = 2 * MAX_ED;

    struct Row { W HP, HN, D0, RAC; uint64_t score; };
    uint32_t maxED = 0, m = 0, n = 0, Wv = 0, Wh = 0;
    std::vector<Row> bv;
    std::vector<std::array<W, 5>> mv;

    bool sequenceSet() const { return !mv.empty(); }
    void reset() { mv.clear(); }

    // bitparallelmatrix.cpp:33-74
    template <typename S> void setSequence(const S& X) {
        n = (uint32_t)X.size() + 1;
        m = 2 * MAX_ED + n;
        mv.assign((m + BLOCK - 1) / BLOCK, std::array<W, 5>());
        const W init = (W(1) << LEFT) - W(1);
        for (auto& v : mv[0])
            v = init;
        W bit = W(1) << LEFT;
        size_t je = std::min<size_t>(X.size(), WORD - LEFT);
        for (size_t j = 0; j < je; j++) {
            mv[0][mvIdx(X[j])] |= bit;
            bit <<= 1;
        }
        for (size_t b = 1; b < mv.size(); b++) {
            for (size_t i = 0; i < 5; i++)
                mv[b][i] = mv[b - 1][i] >> BLOCK;
            bit = W(1) << (WORD - BLOCK);
            size_t jb = WORD - LEFT + (b - 1) * BLOCK;
            size_t je2 = std::min<size_t>(X.size(), jb + BLOCK);
            for (size_t j = jb; j < je2; j++) {
                mv[b][mvIdx(X[j])] |= bit;
                bit <<= 1;
            }
        }
    }

    // bitparallelmatrix.cpp:76-123
    void initializeMatrix(uint32_t maxED_, const std::vector<uint32_t>& initED = {}) {
        maxED = maxED_;
        Wv = initED.empty() ? maxED : (uint32_t)initED.size() - 1 + maxED - initED.back();
        m = Wv + n;
        bv.resize(m);
        bv[0].score = initED.empty() ? 0 : initED[0];
        Wh = maxED - (uint32_t)bv[0].score;
        if (Wv + Wh + 1 > m) {
            m = Wv + Wh + 1;
            bv.resize(m);
        }
        bv[0].HP = (~W(0)) << LEFT;
        bv[0].HN = ~bv[0].HP;
        const size_t nn = std::min<size_t>(initED.size(), LEFT + 1);
        for (uint32_t i = 1; i < nn; ++i) {
            if (initED[i] < initED[i - 1]) {
                bv[0].HP ^= W(1) << (LEFT - i);
                bv[0].HN ^= W(1) << (LEFT - i);
            } else if (initED[i] == initED[i - 1]) {
                bv[0].HN ^= W(1) << (LEFT - i);
            }
        }
        bv[0].RAC = W(1) << (DIAG_R0 + Wh);
    }

    // bitparallelmatrix.h:352-415
    bool computeRow(uint32_t i, char Y) {
        const uint32_t b = i / BLOCK, l = i % BLOCK;
        W& HP = bv[i].HP;
        W& HN = bv[i].HN;
        W& D0 = bv[i].D0;
        W& RAC = bv[i].RAC;
        const W& M = mv[b][mvIdx(Y)];
        HP = bv[i - 1].HP;
        HN = bv[i - 1].HN;
        RAC = bv[i - 1].RAC << 1u;
        if (i % BLOCK == 0) {
            HP >>= BLOCK;
            HN >>= BLOCK;
            RAC >>= BLOCK;
        }
        D0 = (((M & HP) + HP) ^ HP) | M | HN;
        W VP = HN | ~(D0 | HP);
        W VN = D0 & HP;
        HP = (VN << 1u) | ~(D0 | (VP << 1u));
        HN = (D0 & (VP << 1u));
        const uint32_t diagBit = l + DIAG_R0;
        bv[i].score = bv[i - 1].score + ((D0 & (W(1) << diagBit)) ? 0 : 1);
        if (!(D0 & RAC)) {
            size_t val = 1u;
            while (val > 0) {
                if (HP & RAC)
                    val--;
                if (HN & RAC)
                    val++;
                if (RAC == (W(1) << (diagBit - Wv)))
                    return false;
                RAC >>= 1u;
            }
        }
        return true;
    }

    // bitparallelmatrix.h:622-639
    uint32_t at(uint32_t i, uint32_t j) const {
        const uint32_t bit = (i % BLOCK) + DIAG_R0;
        uint32_t b = (i > j) ? bit - (i - j) + 1 : bit + 1;
        uint32_t e = (i > j) ? bit + 1 : bit + (j - i) + 1;
        W mask = ((W(1) << (e - b)) - W(1)) << b;
        int neg = popc(bv[i].HN & mask), pos = popc(bv[i].HP & mask);
        uint32_t score = (uint32_t)bv[i].score;
        score += (i > j) ? (neg - pos) : (pos - neg);
        return score;
    }
    uint32_t sizeOfFinalColumn() const { return Wh + Wv + 1; }
    bool inFinalColumn(uint32_t i) const { return i >= m - sizeOfFinalColumn(); }
    uint32_t firstColumn(uint32_t i) const { return (i <= Wv) ? 0u : i - Wv; }
    // bitparallelmatrix.h:651-665
    bool onlyVerticalGapsLeft(uint32_t i) const {
        if (i + LEFT < n)
            return false;
        const uint32_t b = i / BLOCK, r = i % BLOCK;
        uint32_t bb = DIAG_R0 - Wv + r + 1;
        uint32_t be = DIAG_R0 + n - b * BLOCK;
        return (((~bv[i].HN >> bb) << bb) << (WORD - be)) == W(0);
    }
    // bitparallelmatrix.h:591-614
    void findClusterCenters(uint32_t lastRow, std::vector<len_t>& refEnds, uint32_t maxED_,
                            uint32_t minED_) const {
        refEnds.clear();
        uint32_t firstRow = (m - 1) - sizeOfFinalColumn();
        uint32_t col = n - 1;
        for (uint32_t i = lastRow; i > (m - 1) - sizeOfFinalColumn(); i--) {
            uint32_t ED = at(i, col);
            if (ED > maxED_ || ED < minED_)
                continue;
            bool above = (i == firstRow) || ED <= at(i - 1, col);
            bool below = (i == lastRow) || ED <= at(i + 1, col);
            if (above && below)
                refEnds.push_back(i);
        }
    }

    static void emitCigar(const std::vector<std::pair<char, uint32_t>>& v, std::string& out) {
        for (auto it = v.crbegin(); it != v.crend(); ++it) {
            out += std::to_string(it->second);
            out += it->first;
        }
    }
    // bitparallelmatrix.h:531-586
    template <typename S>
    void traceBack(const S& ref, len_t refEnd, len_t& refBegin, len_t& ED, std::string& CIGAR) const {
        std::vector<std::pair<char, uint32_t>> v;
        uint32_t i = refEnd, j = n - 1;
        ED = at(i, j);
        char state = 0;
        while (j > 0) {
            const uint32_t b = i / BLOCK;
            const W bit = W(1) << ((j - b * BLOCK) + DIAG_R0);
            char op;
            if (bv[i].HP & bit) {
                --j;
                op = 'I';
            } else if ((i > 0) && ((mv[b][mvIdx(ref.fwdAt(i - 1))] | ~bv[i].D0) & bit)) {
                --i;
                --j;
                op = 'M';
            } else {
                --i;
                op = 'D';
            }
            if (state != op) {
                v.emplace_back(op, 0);
                state = op;
            }
            v.back().second++;
        }
        refBegin = i;
        emitCigar(v, CIGAR);
    }
    // bitparallelmatrix.h:460-526
    template <typename S> void findCIGAR(const S& ref, uint32_t score, std::string& CIGAR) {
        initializeMatrix(score);
        for (unsigned int i = 0; i < ref.size(); i++)
            computeRow(i + 1, ref.fwdAt(i));
        uint32_t i = (uint32_t)ref.size(), j = n - 1;
        char state = 0;
        std::vector<std::pair<char, uint32_t>> v;
        while (j > 0 || i > 0) {
            const uint32_t b = i / BLOCK;
            const W bit = W(1) << ((j - b * BLOCK) + DIAG_R0);
            char op;
            if ((j > 0) && (bv[i].HP & bit)) {
                --j;
                op = 'I';
            } else if (i > 0 && j > 0 && ((mv[b][mvIdx(ref.fwdAt(i - 1))] | ~bv[i].D0) & bit)) {
                --i;
                --j;
                op = 'M';
            } else {
                --i;
                op = 'D';
            }
            if (state != op) {
                v.emplace_back(op, 0);
                state = op;
            }
            v.back().second++;
        }
        emitCigar(v, CIGAR);
    }
};

typedef BPMatrix<uint64_t> BP64;
typedef BPMatrix<u128> BP128;

// common virtual-free handle picking 64 or 128 bit, mirrors IBitParallelED use
struct BPAny {
    BP64 m64;
    BP128 m128;
};

} // namespace orc
