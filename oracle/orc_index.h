// ORACLE (test infrastructure, NOT product code).
// CPU restatement of the reference's bidirectional FM-index (Columba Vanilla,
// 32-bit length_t).  It reads the reference's own index files and keeps the
// reference's in-memory layout, so every function below can be checked
// one-to-one against the cited reference lines.  Only tests/, smoke() and
// bench.py's cpu_baseline leg may use anything under oracle/.
#pragma once
#include <algorithm>
#include <cassert>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace orc {

typedef uint32_t len_t; // reference: definitions.h:69-75 with THIRTY_TWO

enum Dir { FWD = 0, BWD = 1 };

// reference: indexhelpers.h:63-127
struct Range {
    len_t b = 0, e = 0;
    Range() {}
    Range(len_t b_, len_t e_) : b(b_), e(e_) {}
    bool empty() const { return e <= b; }
    len_t width() const { return empty() ? 0 : e - b; }
    bool operator==(const Range& o) const { return b == o.b && e == o.e; }
};
// reference: indexhelpers.h:1117-1243 (equality compares the SA range only)
struct RangePair {
    Range sa, rev;
    RangePair() {}
    RangePair(Range a, Range r) : sa(a), rev(r) {}
    bool empty() const { return sa.empty(); }
    len_t width() const { return sa.width(); }
};

enum Ctr {
    NODE_COUNTER, TOTAL_REPORTED_POSITIONS, CIGARS_IN_INDEX, IN_TEXT_STARTED,
    ABORTED_IN_TEXT_VERIF, CIGARS_IN_TEXT_VERIFICATION, IMMEDIATE_SWITCH,
    SEARCH_STARTED, DROPPED_UNIQUE_MATCHES, NUMBER_OF_READS,
    TOTAL_UNIQUE_MATCHES, MAPPED_READS, TOTAL_UNIQUE_PAIRS, MAPPED_PAIRS,
    DISCORDANTLY_MAPPED_PAIRS, MAPPED_HALF_PAIRS, UNPAIRED_BUT_MAPPED_PAIRS,
    N_COUNTERS
}; // reference: indexhelpers.h:1848-1883
struct Counters {
    uint64_t c[N_COUNTERS];
    Counters() { memset(c, 0, sizeof(c)); }
    void inc(Ctr t, uint64_t a = 1) { c[t] += a; }
};

// reference: bitvec.h:234-478 (BitvecIntl<4>) + bwtrepr.h:34-140
struct BwtRep {
    uint64_t dollarPos = 0, N = 0;
    std::vector<uint64_t> bv, counts;
    static const uint64_t S = 4;
    bool read(const std::string& fn) {
        std::ifstream f(fn, std::ios::binary);
        if (!f)
            return false;
        f.read((char*)&dollarPos, 8);
        f.read((char*)&N, 8);
        bv.resize(S * ((N + 63) / 64));
        counts.resize(2 * S * ((N + 511) / 512));
        f.read((char*)bv.data(), bv.size() * 8);
        f.read((char*)counts.data(), counts.size() * 8);
        return (bool)f;
    }
    // bitvec.h:356-372
    uint64_t rank(uint64_t c, uint64_t p) const {
        uint64_t w = (p / 64) * S + c, b = p % 64;
        uint64_t q = (p / 512) * 2 * S + 2 * c;
        uint64_t rv = counts[q];
        int64_t t = ((p / 64) % 8) - 1;
        rv += counts[q + 1] >> (t + (t >> 60 & 8)) * 9 & 0x1FF;
        return rv + __builtin_popcountll((bv[w] << 1) << (63 - b));
    }
    // bwtrepr.h:80-89
    uint64_t occ(int cIdx, uint64_t k) const {
        if (cIdx == 0)
            return (k <= dollarPos) ? 0 : 1;
        return (cIdx == 1) ? rank(0, k) : rank(cIdx - 1, k) - rank(cIdx - 2, k);
    }
    // bwtrepr.h:98-107
    uint64_t cumOcc(int cIdx, uint64_t k) const {
        if (cIdx == 0)
            return 0;
        return (cIdx == 1) ? ((k <= dollarPos) ? 0 : 1)
                           : rank(cIdx - 2, k) + ((k <= dollarPos) ? 0 : 1);
    }
    // character index (0='$',1..4=ACGT) of BWT[p]; planes are cumulative
    // (bwtrepr.h:56-72): plane j set iff BWT[p] <= char j+1.
    int charAt(uint64_t p) const {
        uint64_t w = (p / 64) * S, b = p % 64;
        int set = 0;
        for (uint64_t c = 0; c < S; c++)
            set += (bv[w + c] >> b) & 1;
        return set == 0 ? 0 : 5 - set;
    }
};

// reference: bitvec.h:99-224 (rank9)
struct Rank9 {
    uint64_t N = 0;
    std::vector<uint64_t> bv, counts;
    bool read(const std::string& fn) {
        std::ifstream f(fn, std::ios::binary);
        if (!f)
            return false;
        f.read((char*)&N, 8);
        bv.resize((N + 63) / 64);
        f.read((char*)bv.data(), bv.size() * 8);
        counts.resize((bv.size() + 7) / 4);
        f.read((char*)counts.data(), counts.size() * 8);
        return (bool)f;
    }
    bool get(uint64_t p) const { return (bv[p / 64] >> (p % 64)) & 1; }
    uint64_t rank(uint64_t p) const {
        uint64_t w = p / 64, b = p % 64, q = (w / 8) * 2;
        uint64_t rv = counts[q];
        int64_t t = (w % 8) - 1;
        rv += counts[q + 1] >> (t + (t >> 60 & 8)) * 9 & 0x1FF;
        return rv + __builtin_popcountll((bv[w] << 1) << (63 - b));
    }
};

static inline int c2i(unsigned char c) { // alphabet.h:52-64 for "$ACGT"
    switch (c) {
    case '$': return 0;
    case 'A': return 1;
    case 'C': return 2;
    case 'G': return 3;
    case 'T': return 4;
    default: return -1;
    }
}
static const char I2C[6] = "$ACGT";

template <typename T> static bool readArrayFile(const std::string& fn, std::vector<T>& v) {
    std::ifstream f(fn, std::ios::binary | std::ios::ate);
    if (!f)
        return false;
    size_t sz = f.tellg();
    if (sz % sizeof(T))
        return false;
    v.resize(sz / sizeof(T));
    f.seekg(0);
    f.read((char*)v.data(), sz);
    return true;
}

struct Index {
    len_t counts[5];   // C array, indexinterface.cpp:137-151
    len_t textLength = 0;
    std::string text;  // fmindex.cpp:77-94
    BwtRep fwd, rev;   // .brt / .rev.brt
    Rank9 ssaBv;       // .sa.bv.<s>
    std::vector<len_t> ssa; // .sa.<s>
    std::vector<len_t> startPos;
    std::vector<std::string> seqNames;
    len_t switchPoint = 4; // alignparameters.h:90-91
    len_t wordSize = 10;
    bool noCIGAR = false;
    std::vector<RangePair> kmerTable; // dense 4^K

    // thread state of the reference (indexinterface.h:62-109)
    Dir dir = BWD;
    bool uniBack = false;

    void load(const std::string& base, int saSparse, len_t K, len_t inTextSwitch) {
        switchPoint = inTextSwitch;
        wordSize = K;
        {
            std::ifstream m(base + ".meta");
            if (m) {
                len_t tag;
                size_t lsz;
                std::string fl;
                m >> tag >> lsz >> fl;
                if (lsz != sizeof(len_t))
                    throw std::runtime_error("index built with different length_t");
                if (fl != "VANILLA")
                    throw std::runtime_error("index flavour is not VANILLA");
            }
        }
        std::vector<len_t> cc;
        if (!readArrayFile(base + ".cct", cc) || cc.size() != 256)
            throw std::runtime_error("Cannot open file: " + base + ".cct");
        len_t cum = 0;
        int na = 0;
        for (int i = 0; i < 256; i++) {
            if (!cc[i])
                continue;
            if (na >= 5 || c2i((unsigned char)i) != na)
                throw std::runtime_error("oracle supports the $ACGT alphabet only");
            counts[na++] = cum;
            cum += cc[i];
        }
        if (na != 5)
            throw std::runtime_error("oracle needs all of $ACGT present");
        textLength = cum;
        {
            std::ifstream f(base + ".txt.bin", std::ios::binary);
            if (!f)
                throw std::runtime_error("Error opening file for reading: " + base + ".txt.bin");
            len_t tl;
            f.read((char*)&tl, sizeof(tl));
            text.resize(tl);
            f.read(&text[0], tl);
        }
        if (!fwd.read(base + ".brt"))
            throw std::runtime_error("Cannot open file: " + base + ".brt");
        if (!rev.read(base + ".rev.brt"))
            throw std::runtime_error("Cannot open file: " + base + ".rev.brt");
        std::string sfx = std::to_string(saSparse);
        if (!ssaBv.read(base + ".sa.bv." + sfx))
            throw std::runtime_error("Cannot open file: " + base + ".sa.bv." + sfx);
        if (!readArrayFile(base + ".sa." + sfx, ssa))
            throw std::runtime_error("Cannot open file: " + base + ".sa." + sfx);
        if (!readArrayFile(base + ".pos", startPos))
            throw std::runtime_error("Cannot open file: " + base + ".pos");
        {
            std::ifstream f(base + ".sna", std::ios::binary);
            if (!f)
                throw std::runtime_error("Cannot open file: " + base + ".sna");
            while (f.good()) { // indexinterface.cpp:180-188 (keeps a trailing empty)
                size_t len = 0;
                f.read((char*)&len, sizeof(len));
                if (!f.good())
                    break;
                std::string s(len, ' ');
                f.read(&s[0], len);
                seqNames.push_back(s);
            }
        }
        populateTable();
    }

    RangePair completeRange() const {
        return RangePair(Range(0, textLength), Range(0, textLength));
    }
    void setDirection(Dir d, bool uni) { dir = d; uniBack = uni; }

    // fmindex.cpp:137-243 behind the dispatch of indexinterface.h:771-779
    bool extraChar(int ci, const RangePair& p, RangePair& child) const {
        if (uniBack) { // fmindex.cpp:226-243
            len_t ob = fwd.occ(ci, p.sa.b), oe = fwd.occ(ci, p.sa.e);
            child = RangePair(Range(ob + counts[ci], oe + counts[ci]), Range());
            return !child.empty();
        }
        if (dir == BWD) { // fmindex.cpp:137-172
            const Range& t = p.sa;
            len_t ob = fwd.occ(ci, t.b), oe = fwd.occ(ci, t.e);
            Range r1(ob + counts[ci], oe + counts[ci]);
            len_t s = p.rev.b;
            len_t x = fwd.cumOcc(ci, t.e) - fwd.cumOcc(ci, t.b);
            child = RangePair(r1, Range(s + x, s + x + r1.width()));
            return !child.empty();
        }
        const Range& t = p.rev; // fmindex.cpp:174-211
        len_t ob = rev.occ(ci, t.b), oe = rev.occ(ci, t.e);
        Range r1(ob + counts[ci], oe + counts[ci]);
        len_t s = p.sa.b;
        len_t x = rev.cumOcc(ci, t.e) - rev.cumOcc(ci, t.b);
        child = RangePair(Range(s + x, s + x + r1.width()), r1);
        return !child.empty();
    }
    // fmindex.cpp:213-224
    bool rangeBackward(int ci, const Range& p, Range& child) const {
        len_t ob = fwd.occ(ci, p.b), oe = fwd.occ(ci, p.e);
        child = Range(ob + counts[ci], oe + counts[ci]);
        return !child.empty();
    }
    // indexinterface.cpp:1034-1049
    bool addChar(char c, RangePair& r, Counters& ct) const {
        int ci = c2i((unsigned char)c);
        if (ci > -1) {
            if (extraChar(ci, r, r)) {
                ct.inc(NODE_COUNTER);
                return true;
            }
        }
        r = RangePair();
        return false;
    }
    // fmindex.cpp:434-445
    RangePair rangeOfSingleChar(char c) const {
        int i = c2i((unsigned char)c);
        if (i < 0)
            return RangePair();
        len_t e = (i < 4) ? counts[i + 1] : textLength;
        return RangePair(Range(counts[i], e), Range(counts[i], e));
    }
    // fmindex.cpp:47-60
    len_t findLF(len_t k) const {
        int pos = fwd.charAt(k);
        return counts[pos] + (len_t)fwd.occ(pos, k);
    }
    len_t findSA(len_t i) const {
        len_t l = 0;
        while (!ssaBv.get(i)) {
            i = findLF(i);
            l++;
        }
        return ssa[ssaBv.rank(i)] + l;
    }

    // indexinterface.cpp:294-335 -- same table content, dense instead of hashed
    void populateTable() {
        kmerTable.assign(wordSize ? (size_t)1 << (2 * wordSize) : 1, RangePair());
        if (wordSize == 0) {
            kmerTable[0] = completeRange();
            return;
        }
        setDirection(FWD, false);
        fill(completeRange(), 0, 0);
    }
    void fill(const RangePair& r, len_t depth, uint32_t code) {
        if (depth == wordSize) {
            kmerTable[code] = r;
            return;
        }
        for (int ci = 1; ci < 5; ci++) {
            RangePair ch;
            if (extraChar(ci, r, ch))
                fill(ch, depth + 1, (code << 2) | (ci - 1));
        }
    }
    // indexinterface.h:590-595; p points at wordSize chars of the read
    RangePair lookUpKmer(const char* p) const {
        uint32_t code = 0;
        for (len_t i = 0; i < wordSize; i++) {
            int ci = c2i((unsigned char)p[i]);
            if (ci < 1)
                return RangePair();
            code = (code << 2) | (ci - 1);
        }
        return kmerTable[code];
    }
};

} // namespace orc
