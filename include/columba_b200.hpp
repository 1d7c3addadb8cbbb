// columba_b200 -- C++ host layer over the C ABI (include/columba_b200.h).
//
// Header only.  Mirrors the part of the reference's C++ API that sits on the
// accelerated path, so that host code written against Columba keeps its shape:
//
//   clb::FMIndex          <- FMIndex(baseFile, inTextSwitch, noCIGAR, sa_sparse, verbose, wordSize)
//                            src/fmindex/fmindex.h:403-417; getSeqNames etc. src/indexinterface.h:900-930
//   clb::SearchStrategy   <- SearchStrategy, src/searchstrategy.h:1724-2081, as built by
//                            Parameters::createStrategy, src/parameters/alignparameters.cpp:1312-1387
//     matchApproxBatch    <- matchApprox (:2021) / matchApproxAllMap src/searchstrategy.cpp:495-535 per chunk
//     matchBestBatch      <- matchApproxBestPlusX / findBestAlignments / checkAlignments / combineOccVectors,
//                            src/searchstrategy.cpp:537-744 (per-read state machine over batched device passes)
//     supportsDistanceScore, calculateNumParts, getMaxED (:1797-1806)
//   clb::TextOcc          <- the fields of TextOcc the SAM writer uses, src/indexhelpers.h:289-930
//   clb::Counters         <- Counters, src/indexhelpers.h:1848-1883 (same enum order)
//
// Errors are std::runtime_error with the library's message (the reference
// throws std::runtime_error with the same texts for index / scheme problems).
// One SearchStrategy per host thread; an FMIndex may be shared.
#ifndef COLUMBA_B200_HPP
#define COLUMBA_B200_HPP
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "columba_b200.h"

namespace clb {

enum Strand { FORWARD_STRAND = 0, REVERSE_C_STRAND = 1 };
enum DistanceMetric { HAMMING = CLB_METRIC_HAMMING, EDIT = CLB_METRIC_EDIT };
enum PartitionStrategy { UNIFORM = CLB_PART_UNIFORM, STATIC = CLB_PART_STATIC, DYNAMIC = CLB_PART_DYNAMIC };

struct Counters {
    uint64_t c[CLB_N_COUNTERS] = {};
    uint64_t get(int type) const { return c[type]; }
    void add(const uint64_t* v) {
        for (int i = 0; i < CLB_N_COUNTERS; i++)
            c[i] += v[i];
    }
};

struct TextOcc {
    uint32_t seqID = 0;      // assigned sequence
    uint32_t begin = 0;      // 0-based begin within the sequence (SAM POS - 1)
    uint32_t end = 0;        // end (exclusive) within the sequence
    uint32_t indexBegin = 0; // begin in the concatenated text
    uint32_t distance = 0;
    Strand strand = FORWARD_STRAND;
    bool trimmed = false;    // FOUND_WITH_TRIMMING: clipped at a sequence boundary
    uint32_t distanceBeforeTrimming = 0; // the distance the occurrence was found with (== distance unless trimmed)
    std::string cigar;       // "*" when CIGARs are switched off
};

// The SAM records the reference writes for one single-end read (ALL mode): generateOutputSingleEnd
// src/searchstrategy.cpp:1824-1902 (stable sort on the distance: the first occurrence is the primary one),
// generateSE_SAM src/searchstrategy.h:1633-1641, TextOcc::generateSAMSingleEnd src/indexhelpers.cpp:56-87, flags
// getFlagsSE src/indexhelpers.h:321-331, MAPQ getMapQ :378-388, unmapped record createUnmappedSAMOccurrenceSE.
// Sequence and quality are printed on the primary record only (reverse-complemented / reversed on the reverse strand).
inline void samSingleEnd(const std::string& name, const std::string& seq, const std::string& qual, std::vector<TextOcc> occs,
                         const std::vector<std::string>& seqNames, std::string& out, bool unmappedRecord = true) {
    const std::string q = qual.empty() ? "*" : qual;
    if (occs.empty()) {
        if (unmappedRecord)
            out += name + "\t4\t*\t0\t0\t*\t*\t0\t0\t" + seq + "\t" + q + "\tPG:Z:Columba\n";
        return;
    }
    std::stable_sort(occs.begin(), occs.end(), [](const TextOcc& a, const TextOcc& b) { return a.distance < b.distance; });
    const uint32_t minScore = occs.front().distance;
    uint32_t nHits = 0;
    for (const auto& t : occs)
        nHits += t.distance == minScore;
    for (size_t i = 0; i < occs.size(); i++) {
        const TextOcc& t = occs[i];
        const bool rc = t.strand == REVERSE_C_STRAND;
        std::string s = "*", ql = "*";
        if (i == 0) {
            s = seq;
            ql = q;
            if (rc) {
                std::reverse(s.begin(), s.end());
                for (auto& c : s)
                    c = c == 'A' ? 'T' : c == 'C' ? 'G' : c == 'G' ? 'C' : c == 'T' ? 'A' : c;
                if (ql != "*")
                    std::reverse(ql.begin(), ql.end());
            }
        }
        int mapq = 0;
        if (t.distance == minScore)
            mapq = nHits == 1 ? 60 : (int)std::lround(-10.0 * std::log10(1.0 - 1.0 / nHits));
        out += name + "\t" + std::to_string((rc ? 16 : 0) | (i ? 256 : 0)) + "\t" + seqNames[t.seqID] + "\t" +
               std::to_string(t.begin + 1) + "\t" + std::to_string(mapq) + "\t" + t.cigar + "\t*\t0\t0\t" + s + "\t" + ql +
               "\tAS:i:" + std::to_string(t.distance) + "\tNM:i:" + std::to_string(t.distance) + "\tPG:Z:Columba\n";
    }
}

class FMIndex {
  public:
    FMIndex(const std::string& baseFile, int inTextSwitch = 4, bool noCIGAR = false, int sa_sparse = 4,
            int kmerSize = 10, int device = 0) {
        if (clb_index_load(baseFile.c_str(), sa_sparse, kmerSize, inTextSwitch, noCIGAR ? 1 : 0, device, &ix))
            throw std::runtime_error(clb_last_error(nullptr));
        for (uint32_t i = 0; i < clb_index_num_seqs(ix); i++)
            names.push_back(clb_index_seq_name(ix, i));
    }
    ~FMIndex() { clb_index_free(ix); }
    FMIndex(const FMIndex&) = delete;
    FMIndex& operator=(const FMIndex&) = delete;
    const std::vector<std::string>& getSeqNames() const { return names; }
    uint32_t getTextLength() const { return clb_index_text_length(ix); }
    clb_index* handle() const { return ix; }

  private:
    clb_index* ix = nullptr;
    std::vector<std::string> names;
};

class SearchStrategy {
  public:
    // kind / arg as clb_ctx_set_strategy: ("named", "columba"), ("custom", dir), ("custom_dynamic", dir), ("multi", dir)
    SearchStrategy(FMIndex& index, const std::string& kind, const std::string& arg, const std::string& schemesDir,
                   PartitionStrategy p = DYNAMIC, DistanceMetric metric = EDIT) {
        if (clb_ctx_create(index.handle(), &ctx))
            throw std::runtime_error(clb_last_error(nullptr));
        try {
            check(clb_ctx_set_strategy(ctx, kind.c_str(), arg.c_str(), schemesDir.c_str()));
            check(clb_ctx_set_options(ctx, (int)metric, (int)p));
        } catch (...) {
            clb_ctx_free(ctx);
            throw;
        }
    }
    ~SearchStrategy() { clb_ctx_free(ctx); }
    SearchStrategy(const SearchStrategy&) = delete;
    SearchStrategy& operator=(const SearchStrategy&) = delete;

    // BEST and paired-end mode map and filter the two strands of a read separately (the results then come in
    // groups 2 * read + strand) and clip alignments at sequence ends up to the cut-off instead of the distance
    void setFilterPerStrand(bool perStrand) { check(clb_ctx_set_filter_per_strand(ctx, perStrand ? 1 : 0)); }
    void setTrimMax(uint32_t largestStratum) { check(clb_ctx_set_trim_max(ctx, (int)largestStratum)); }

    bool supportsDistanceScore(int maxScore) const { return maxScore == 0 || clb_ctx_supports_distance(ctx, maxScore); }
    int calculateNumParts(int maxED) const { return clb_ctx_num_parts(ctx, maxED); }
    // src/searchstrategy.h:1797-1806
    uint32_t getMaxED(uint32_t minIdentity, uint32_t seqSize) const {
        uint32_t supported = 0;
        while (supported < 13 && clb_ctx_supports_distance(ctx, (int)supported + 1))
            supported++;
        return std::min(std::min<uint32_t>(13u, supported), (seqSize * (100 - minIdentity)) / 100);
    }

    // One device pass over a chunk of reads: groups[g] = occurrences of group g in text order, g = read (ALL mode)
    // or 2 * read + strand after setFilterPerStrand(true)
    void mapChunk(const std::vector<std::string>& reads, uint32_t maxED, uint32_t minED, Counters& counters,
                  std::vector<std::vector<TextOcc>>& groups) {
        groups.clear();
        if (reads.empty())
            return;
        std::string cat;
        std::vector<uint32_t> off(1, 0);
        for (const auto& r : reads) {
            cat += r;
            off.push_back((uint32_t)cat.size());
        }
        clb_result r;
        int rc = clb_map_batch(ctx, (const uint8_t*)cat.data(), off.data(), (uint32_t)reads.size(), (int)maxED,
                               (int)minED, &r);
        check(rc); // reads too short for the scheme are skipped (clb_result::read_status), not an error
        groups.assign(r.n_groups, {});
        for (uint32_t g = 0; g < r.n_groups; g++)
            for (uint32_t j = r.read_start[g]; j < r.read_start[g] + r.read_count[g]; j++) {
                const clb_occ& o = r.occ[j];
                TextOcc t;
                t.seqID = o.seq_id;
                t.begin = o.pos;
                t.end = o.end;
                t.indexBegin = o.index_begin;
                t.distance = o.dist;
                t.strand = o.strand ? REVERSE_C_STRAND : FORWARD_STRAND;
                t.trimmed = o.trimmed != 0;
                t.distanceBeforeTrimming = o.trimmed ? (uint32_t)o.trimmed - 1u : (uint32_t)o.dist;
                for (uint32_t c = 0; c < o.n_cigar; c++) {
                    const uint16_t run = r.cigar[(size_t)j * r.cigar_stride + c];
                    t.cigar += std::to_string(run >> 2);
                    t.cigar += "MID"[run & 3];
                }
                if (!o.n_cigar)
                    t.cigar = "*";
                groups[g].push_back(std::move(t));
            }
        counters.add(r.counters);
    }

    // ALL mode for a chunk of reads: result[i] = occurrences of reads[i] in text order
    void matchApproxBatch(const std::vector<std::string>& reads, uint32_t maxED, Counters& counters,
                          std::vector<std::vector<TextOcc>>& result, uint32_t minED = 0) {
        mapChunk(reads, maxED, minED, counters, result);
        result.resize(reads.size());
    }

    // BEST mode (-a best, -x strata after the best one) for a chunk of reads.  result[i] = the occurrences the
    // reference reports for reads[i], in its order: the first one is the primary alignment.  nHits (optional) =
    // number of alignments in the best stratum (the reference's MAPQ input).
    void matchBestBatch(const std::vector<std::string>& reads, uint32_t minIdentity, uint32_t x, Counters& counters,
                        std::vector<std::vector<TextOcc>>& result, std::vector<uint32_t>* nHits = nullptr) {
        const size_t n = reads.size();
        result.assign(n, {});
        if (nHits)
            nHits->assign(n, 0);
        std::vector<BestState> st;
        st.reserve(n);
        for (size_t i = 0; i < n; i++)
            st.emplace_back(getMaxED(minIdentity, (uint32_t)reads[i].size()), x);
        check(clb_ctx_set_filter_per_strand(ctx, 1));
        try {
            // one pass per (cut-off, k, minimal distance) over the reads that need it
            struct Key {
                uint32_t cutoff, k, minD;
                bool operator<(const Key& o) const {
                    return cutoff != o.cutoff ? cutoff < o.cutoff : (k != o.k ? k < o.k : minD < o.minD);
                }
                bool operator==(const Key& o) const { return cutoff == o.cutoff && k == o.k && minD == o.minD; }
            };
            auto devicePass = [&](std::vector<std::pair<Key, size_t>>& want) {
                std::sort(want.begin(), want.end());
                for (size_t a = 0; a < want.size();) {
                    size_t b = a;
                    std::vector<std::string> sub;
                    while (b < want.size() && want[b].first == want[a].first)
                        sub.push_back(reads[want[b++].second]);
                    check(clb_ctx_set_trim_max(ctx, (int)want[a].first.cutoff)); // checkAlignments clips with the cut-off
                    std::vector<std::vector<TextOcc>> groups;
                    mapChunk(sub, want[a].first.k, want[a].first.minD, counters, groups);
                    for (size_t j = a; j < b; j++)
                        for (int strand = 0; strand < 2; strand++)
                            st[want[j].second].add(strand, groups[2 * (j - a) + strand]);
                    a = b;
                }
            };
            std::vector<std::pair<Key, size_t>> want;
            if (x == 0) {
                for (size_t i = 0; i < n; i++)
                    want.push_back({Key{st[i].cutoff, 0, 0}, i});
                devicePass(want);
                for (auto& s : st)
                    s.afterExact();
            } else {
                for (auto& s : st)
                    if (s.k > s.maxED)
                        s.finished = true;
            }
            for (;;) {
                want.clear();
                for (size_t i = 0; i < n; i++)
                    if (!st[i].finished && (int)st[i].k > st[i].doneUpto)
                        want.push_back({Key{st[i].cutoff, st[i].k, std::min((uint32_t)(st[i].doneUpto + 1), st[i].k)}, i});
                devicePass(want);
                bool progressed = false;
                for (auto& s : st)
                    if (!s.finished) {
                        s.step((int)s.k > s.doneUpto);
                        progressed = true;
                    }
                if (!progressed)
                    break;
            }
        } catch (...) {
            clb_ctx_set_trim_max(ctx, 0);
            clb_ctx_set_filter_per_strand(ctx, 0);
            throw;
        }
        check(clb_ctx_set_trim_max(ctx, 0));
        check(clb_ctx_set_filter_per_strand(ctx, 0));
        for (size_t i = 0; i < n; i++) {
            const uint32_t h = st[i].result(result[i]);
            if (nHits)
                (*nHits)[i] = h;
        }
    }

    clb_ctx* handle() const { return ctx; }

  private:
    // findBestAlignments' state of one read (src/searchstrategy.cpp:623-712): occurrences per strand and distance,
    // the distances mapped so far, and the variables of the stratum loop
    struct BestState {
        uint32_t cutoff, x, best, maxED, k, prevK = 0;
        int doneUpto = -1;
        bool bestFound = false, finished = false;
        std::vector<std::vector<TextOcc>> lv[2];
        BestState(uint32_t cutoff_, uint32_t x_) : cutoff(cutoff_), x(x_), best(cutoff_ + 1), maxED(cutoff_), k(std::max(x_, 1u)) {
            lv[0].resize(cutoff + 1);
            lv[1].resize(cutoff + 1);
        }
        void add(int strand, std::vector<TextOcc>& occs) {
            for (auto& t : occs)
                if (t.distanceBeforeTrimming <= cutoff)
                    lv[strand][t.distanceBeforeTrimming].push_back(std::move(t));
        }
        // checkAlignments, src/searchstrategy.cpp:537-568
        void checkLevel(uint32_t level) {
            for (int strand = 0; strand < 2; strand++) {
                std::vector<TextOcc> keep, moved;
                for (auto& t : lv[strand][level]) {
                    if (t.trimmed) {
                        if (t.distance > level)
                            moved.push_back(std::move(t));
                    } else {
                        keep.push_back(std::move(t));
                        best = std::min(best, level);
                    }
                }
                lv[strand][level].swap(keep);
                for (auto& t : moved)
                    if (t.distance <= cutoff) {
                        t.trimmed = false; // removeTrimmingLabel
                        t.distanceBeforeTrimming = t.distance;
                        lv[strand][t.distance].push_back(std::move(t));
                    }
            }
        }
        bool anyUpto(uint32_t d) const {
            for (int strand = 0; strand < 2; strand++)
                for (uint32_t i = 0; i <= std::min(d, cutoff); i++)
                    if (!lv[strand][i].empty())
                        return true;
            return false;
        }
        void afterExact() { // :637-668
            doneUpto = 0;
            if (!lv[0][0].empty() || !lv[1][0].empty()) {
                checkLevel(0);
                if (best == 0)
                    bestFound = true;
            }
            maxED = best == 0 ? x : cutoff;
            if (k > maxED)
                finished = true;
        }
        void step(bool mapped) { // one iteration of the loop at :677-708
            bool update;
            if (mapped) {
                doneUpto = (int)k;
                update = anyUpto(k);
            } else {
                update = !lv[0][k].empty() || !lv[1][k].empty();
            }
            if (update)
                for (uint32_t level = prevK + 1; level <= std::min(k, best + x); level++)
                    checkLevel(level);
            if (bestFound) {
                finished = true;
                return;
            }
            if (update && best < cutoff + 1) {
                bestFound = true;
                if (x == 0) {
                    finished = true;
                    return;
                }
                prevK = k;
                k = std::min(best + x, maxED);
            } else {
                if (k == maxED) {
                    finished = true;
                    return;
                }
                prevK = k;
                k = std::min(k + x + (k < 5 ? 2u : 4u), maxED);
            }
        }
        // combineOccVectors, src/searchstrategy.cpp:570-621; returns nHits
        uint32_t result(std::vector<TextOcc>& out) {
            out.clear();
            if (!bestFound)
                return 0;
            const uint32_t nHits = (uint32_t)(lv[0][best].size() + lv[1][best].size());
            for (uint32_t d = best; d <= std::min(best + x, cutoff); d++)
                for (int strand = 0; strand < 2; strand++) {
                    auto& v = lv[strand][d];
                    std::stable_sort(v.begin(), v.end(), [](const TextOcc& a, const TextOcc& b) {
                        return a.seqID < b.seqID || (a.seqID == b.seqID && a.begin < b.begin);
                    });
                    for (size_t i = 0; i < v.size(); i++)
                        if (i == 0 || v[i].seqID != v[i - 1].seqID || v[i].begin != v[i - 1].begin)
                            out.push_back(v[i]);
                }
            return nHits;
        }
    };
    void check(int rc) const {
        if (rc)
            throw std::runtime_error(clb_last_error(ctx));
    }
    clb_ctx* ctx = nullptr;
};

} // namespace clb
#endif
