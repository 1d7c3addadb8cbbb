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
//     matchBestBatch      <- findBestAlignments src/searchstrategy.cpp:623-712 (strata loop, batched)
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
    bool trimmed = false;    // FOUND_WITH_TRIMMING
    std::string cigar;       // "*" when CIGARs are switched off
};

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

    bool supportsDistanceScore(int maxScore) const { return maxScore == 0 || clb_ctx_supports_distance(ctx, maxScore); }
    int calculateNumParts(int maxED) const { return clb_ctx_num_parts(ctx, maxED); }
    // src/searchstrategy.h:1797-1806
    uint32_t getMaxED(uint32_t minIdentity, uint32_t seqSize) const {
        uint32_t supported = 0;
        while (supported < 13 && clb_ctx_supports_distance(ctx, (int)supported + 1))
            supported++;
        return std::min(std::min<uint32_t>(13u, supported), (seqSize * (100 - minIdentity)) / 100);
    }

    // ALL mode for a chunk of reads: result[i] = occurrences of reads[i] in text order
    void matchApproxBatch(const std::vector<std::string>& reads, uint32_t maxED, Counters& counters,
                          std::vector<std::vector<TextOcc>>& result, uint32_t minED = 0) {
        result.assign(reads.size(), {});
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
        if (rc && !(rc == CLB_E_UNSUPPORTED && r.n_reads == reads.size())) // reads too short for the scheme are skipped
            check(rc);
        for (uint32_t i = 0; i < r.n_reads; i++)
            for (uint32_t j = r.read_start[i]; j < r.read_start[i] + r.read_count[i]; j++) {
                const clb_occ& o = r.occ[j];
                TextOcc t;
                t.seqID = o.seq_id;
                t.begin = o.pos;
                t.end = o.end;
                t.indexBegin = o.index_begin;
                t.distance = o.dist;
                t.strand = o.strand ? REVERSE_C_STRAND : FORWARD_STRAND;
                t.trimmed = o.trimmed != 0;
                for (uint32_t c = 0; c < o.n_cigar; c++) {
                    const uint16_t run = r.cigar[(size_t)j * r.cigar_stride + c];
                    t.cigar += std::to_string(run >> 2);
                    t.cigar += "MID"[run & 3];
                }
                if (!o.n_cigar)
                    t.cigar = "*";
                result[i].push_back(std::move(t));
            }
        counters.add(r.counters);
    }

    // BEST mode (strata 0; 1, 3, 5, 9, 13 or best + x) for a chunk of reads of one
    // length class: result[i] = occurrences with distance <= best + x
    void matchBestBatch(const std::vector<std::string>& reads, uint32_t minIdentity, uint32_t x, Counters& counters,
                        std::vector<std::vector<TextOcc>>& result) {
        result.assign(reads.size(), {});
        std::vector<uint32_t> cut(reads.size());
        uint32_t maxCut = 0;
        for (size_t i = 0; i < reads.size(); i++) {
            cut[i] = getMaxED(minIdentity, (uint32_t)reads[i].size());
            maxCut = std::max(maxCut, cut[i]);
        }
        for (uint32_t cutoff = 0; cutoff <= maxCut; cutoff++) {
            std::vector<size_t> pending;
            for (size_t i = 0; i < reads.size(); i++)
                if (cut[i] == cutoff)
                    pending.push_back(i);
            if (pending.empty())
                continue;
            check(clb_ctx_set_trim_max(ctx, (int)cutoff));
            // the exact stratum is mapped together with stratum 1
            std::vector<uint32_t> sched;
            for (uint32_t k = std::max(x, 1u); k <= std::max(cutoff, 1u);) {
                sched.push_back(k);
                if (k >= cutoff)
                    break;
                k = std::min(k + x + (k < 5 ? 2u : 4u), cutoff);
            }
            int prev = -1;
            for (uint32_t k : sched) {
                if (pending.empty())
                    break;
                std::vector<std::string> sub;
                for (size_t i : pending)
                    sub.push_back(reads[i]);
                std::vector<std::vector<TextOcc>> occ;
                matchApproxBatch(sub, k, counters, occ, (uint32_t)(prev + 1));
                std::vector<size_t> still;
                for (size_t j = 0; j < pending.size(); j++) {
                    auto& o = occ[j];
                    o.erase(std::remove_if(o.begin(), o.end(), [&](const TextOcc& t) { return t.distance > cutoff; }),
                            o.end());
                    if (o.empty()) {
                        still.push_back(pending[j]);
                        continue;
                    }
                    uint32_t best = o[0].distance;
                    for (const auto& t : o)
                        best = std::min(best, t.distance);
                    const uint32_t lim = std::min(best + x, cutoff);
                    if (x > 0 && lim > k) { // the strata up to best + x
                        std::vector<std::vector<TextOcc>> more;
                        matchApproxBatch({reads[pending[j]]}, lim, counters, more, k + 1);
                        for (auto& t : more[0])
                            o.push_back(std::move(t));
                    }
                    o.erase(std::remove_if(o.begin(), o.end(), [&](const TextOcc& t) { return t.distance > lim; }),
                            o.end());
                    result[pending[j]] = std::move(o);
                }
                pending.swap(still);
                prev = (int)k;
            }
        }
        check(clb_ctx_set_trim_max(ctx, 0));
    }

    clb_ctx* handle() const { return ctx; }

  private:
    void check(int rc) const {
        if (rc)
            throw std::runtime_error(clb_last_error(ctx));
    }
    clb_ctx* ctx = nullptr;
};

} // namespace clb
#endif
