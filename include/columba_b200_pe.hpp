// columba_b200 -- paired-end mode of the C++ host layer (header only).
//
// In the reference paired-end mapping is host logic around the per-(mate, strand)
// mapping call SearchStrategy::mapRead (src/searchstrategy.h:490-517): which of
// the four (mate, strand) sequences is mapped, at which distance and in which
// order depends on what the previous mappings returned (the downstream mate is
// only mapped when the upstream one has occurrences, BEST mode raises the total
// distance stratum by stratum, ...), and because the redundancy filter depends on
// the maximal distance of a mapping, the same requests have to be made to get the
// same pairs.  Here every mapping request is served by a batched device pass:
//
//   * the per-pair logic below is a direct restatement of the reference's
//     functions and asks for mappings through PairContext::mapRead;
//   * a mapping that has not been done yet aborts the evaluation of that pair
//     (NeedMapping) after recording the request; the driver (matchApproxPEBatch)
//     collects the requests of all pairs of the chunk, serves them with ONE
//     device pass per (maximal distance, minimal distance) -- per-strand
//     filtering, clb_ctx_set_filter_per_strand -- and evaluates the pairs again;
//     the logic is deterministic given the mappings, so a pair needs as many
//     rounds as it has dependent mapping steps (2 in ALL mode, a handful in BEST
//     mode).
//
// Restated (file:line of the reference):
//   matchApproxPairedEndAll        src/searchstrategy.cpp:746-775
//   processComb / processComb*All  src/searchstrategy.h:753-861
//   pairOccurrences                src/searchstrategy.cpp:1281-1343
//   pairDiscordantly, addDiscPairs, addUnpairedMatches, addOneUnmapped, addBothUnmapped
//                                  src/searchstrategy.cpp:1398-1647, src/searchstrategy.h:1161-1290
//   matchApproxPairedEndBestPlusX  src/searchstrategy.cpp:1091-1179
//   processComb (BEST), processCombFR/FF/RF, mergeOrMovePairs, handleTrimmedOccs, processSeq
//                                  src/searchstrategy.cpp:777-1063
//   pairOccurrencesForBestMapping  src/searchstrategy.cpp:1743-1819
//   pairDiscordantlyBest, findBestMapping, mapStratum, findBestAlignments, checkAlignments, combineOccVectors
//                                  src/searchstrategy.cpp:537-712,1649-1741, src/searchstrategy.h:1354-1361
//   generateSAMPairedEnd (order, primary/secondary, nPairs, best score)   src/searchstrategy.cpp:1904-1976
//   flags / MAPQ / TLEN of the records   src/indexhelpers.h:340-408, src/indexhelpers.cpp:115-262
//   pair counters                  src/fastq.cpp:662-688
//
// Known limit: an occurrence that crosses a sequence boundary comes back from the
// device already clipped (the reference clips it lazily, when a pair is formed,
// and pairs it with its unclipped coordinates first); pairs with such a mate can
// differ from the reference's.
#ifndef COLUMBA_B200_PE_HPP
#define COLUMBA_B200_PE_HPP
#include <cmath>
#include <map>
#include <set>
#include <tuple>

#include "columba_b200.hpp"

namespace clb {

enum Orientation { FR = 0, RF = 1, FF = 2 };

struct ReadPair {
    std::string name1, read1, name2, read2;
};

// one alignment of one mate (or the placeholder of an unmapped mate: valid == false)
struct MateOcc : TextOcc {
    bool first = true; // FIRST_IN_PAIR
    bool valid = true;
    uint8_t unmappedFlags = 0; // flags of the placeholder record (createUnmappedSAMOccurrencePE)
    uint32_t indexEnd() const { return indexBegin + (end - begin); }
};

struct PairedTextOccs {
    MateOcc up, down;
    uint32_t fragSize = 0;
    bool discordant = false;
    uint32_t distance() const { return up.distance + down.distance; }
};

// one SAM record of a pair, fields 1-9 and NM
struct SamRecord {
    std::string qname;
    uint32_t flag = 0;
    std::string rname = "*";
    uint32_t pos = 0;
    int mapq = 0;
    std::string cigar = "*";
    std::string rnext = "*";
    uint32_t pnext = 0;
    int64_t tlen = 0;
    int nm = -1; // -1: no NM tag (unmapped)
};

struct PairResult {
    std::vector<PairedTextOccs> pairs;
    std::vector<MateOcc> unpaired;
    std::vector<uint32_t> unpairedBestCount, unpairedBest; // per unpaired record: hits with the best score, best score
    std::vector<bool> unpairedPrimary;
};

// result of the parameter inference (inferPairedEndParameters, src/parallel.cpp:376-465)
struct PEInference {
    bool inferred = false;       // false: no pair mapped unambiguously, the given options were kept
    Orientation orientation = FR;
    uint32_t maxInsertSize = 500, minInsertSize = 0;
    float meanInsertSize = 0, stddevInsertSize = 0;
    size_t pairsProcessed = 0;   // leading pairs that went through the single-end pre-pass
    size_t unambiguousPairs = 0;
};

struct PEOptions {
    Orientation orientation = FR;
    uint32_t maxFragSize = 500, minFragSize = 0;
    bool discordantAllowed = false;
    bool unmappedSAM = true;
    uint32_t x = 0; // strata after the best one (BEST mode)
};

class PairedEndMapper {
  public:
    explicit PairedEndMapper(SearchStrategy& s) : strategy(s) {}

    // ALL mode (best == false: maxEDOrIdentity = maximal distance) or BEST mode (minimal identity) with the given
    // orientation and fragment bounds (-nI)
    void matchApproxPEBatch(const std::vector<ReadPair>& input, bool best, uint32_t maxEDOrIdentity,
                            const PEOptions& opt, Counters& counters, std::vector<PairResult>& out) {
        runBatch(input, best, maxEDOrIdentity, opt, counters, out, nullptr, nullptr, nullptr);
        addPairCounters(out, counters);
    }

    // The reference's default paired-end flow (inferParametersAndStartWorkers, src/parallel.cpp:880-930): the
    // leading pairs are mapped single-ended in ALL mode (mate 2 only when mate 1 has exactly one occurrence) until
    // 750 pairs are mapped unambiguously or 10 000 reads have been given out (the reference stops at the granularity
    // of its reader's chunks, which depends on timing; here: chunks of `inferChunk` pairs), orientation and
    // insert-size bounds are inferred from them (median / MAD outlier removal, mean +- 6 sd), those pairs are paired
    // from their single-end results (pairSingleEnd, src/searchstrategy.h:1433-1463) and the remaining pairs go
    // through matchApproxPE with the inferred parameters.  All sequences are taken to come from the first reference
    // file (IndexInterface::isInFirstFile, src/indexinterface.h:622-625).
    void matchApproxPEInfer(const std::vector<ReadPair>& input, bool best, uint32_t maxEDOrIdentity, PEOptions opt,
                            Counters& counters, std::vector<PairResult>& out, PEInference& inf, size_t inferChunk = 512) {
        const size_t n = input.size();
        std::vector<std::vector<MateOcc>> se1, se2;
        std::vector<bool> read2done;
        std::vector<uint32_t> fragSizes;
        std::vector<Orientation> orientations;
        size_t given = 0, unambiguous = 0;
        while (given < n && unambiguous < 750 && 2 * given < 10000) {
            const size_t c0 = given, c1 = std::min(n, given + inferChunk);
            std::vector<std::string> r1;
            std::vector<uint32_t> k1;
            for (size_t i = c0; i < c1; i++) {
                r1.push_back(input[i].read1);
                k1.push_back(best ? strategy.getMaxED(maxEDOrIdentity, (uint32_t)input[i].read1.size()) : maxEDOrIdentity);
            }
            std::vector<std::vector<MateOcc>> m1 = mapSingleEndAll(r1, k1, true, counters);
            std::vector<std::string> r2;
            std::vector<uint32_t> k2;
            std::vector<size_t> who;
            for (size_t i = c0; i < c1; i++)
                if (m1[i - c0].size() == 1) { // hasUnambiguousMatchInFirstFile, src/parallel.cpp:229-258
                    who.push_back(i);
                    r2.push_back(input[i].read2);
                    k2.push_back(best ? strategy.getMaxED(maxEDOrIdentity, (uint32_t)input[i].read2.size()) : maxEDOrIdentity);
                }
            std::vector<std::vector<MateOcc>> m2 = mapSingleEndAll(r2, k2, false, counters);
            se1.resize(c1);
            se2.resize(c1);
            read2done.resize(c1, false);
            for (size_t i = c0; i < c1; i++)
                se1[i] = std::move(m1[i - c0]);
            for (size_t j = 0; j < who.size(); j++) {
                const size_t i = who[j];
                read2done[i] = true;
                se2[i] = std::move(m2[j]);
                if (se2[i].size() == 1) { // addFragmentAndOrientation, src/parallel.cpp:327-358
                    const MateOcc &a = se1[i][0], &b = se2[i][0];
                    const bool firstFirst = a.indexBegin < b.indexBegin;
                    fragSizes.push_back(firstFirst ? b.indexEnd() - a.indexBegin : a.indexEnd() - b.indexBegin);
                    const bool aRc = a.strand == REVERSE_C_STRAND, bRc = b.strand == REVERSE_C_STRAND;
                    orientations.push_back(aRc == bRc ? FF : (firstFirst == aRc ? RF : FR));
                    unambiguous++;
                }
            }
            given = c1;
        }
        inf = PEInference();
        inf.orientation = opt.orientation;
        inf.maxInsertSize = opt.maxFragSize;
        inf.minInsertSize = opt.minFragSize;
        inf.pairsProcessed = given;
        inf.unambiguousPairs = unambiguous;
        inferParameters(fragSizes, orientations, inf);
        opt.orientation = inf.orientation;
        opt.maxFragSize = inf.maxInsertSize;
        opt.minFragSize = inf.minInsertSize;
        out.assign(n, PairResult());
        // the pairs of the pre-pass: paired from their single-end results
        {
            std::vector<ReadPair> head(input.begin(), input.begin() + given);
            std::vector<PairResult> res;
            PEOptions o2 = opt;
            if (best)
                o2.x = 0; // pairSingleEndedMatchesBest: x = 0
            runBatch(head, best, maxEDOrIdentity, o2, counters, res, &se1, &se2, &read2done);
            for (size_t i = 0; i < given; i++)
                out[i] = std::move(res[i]);
        }
        if (given < n) {
            std::vector<ReadPair> tail(input.begin() + given, input.end());
            std::vector<PairResult> res;
            runBatch(tail, best, maxEDOrIdentity, opt, counters, res, nullptr, nullptr, nullptr);
            for (size_t i = given; i < n; i++)
                out[i] = std::move(res[i - given]);
        }
        addPairCounters(out, counters);
    }

    // the records the reference writes for one pair, in its order (OutputWriter, src/fastq.cpp:690-710)
    std::vector<SamRecord> samRecords(const ReadPair& rp, const PairResult& r, const std::vector<std::string>& seqNames) const {
        std::vector<SamRecord> out;
        if (!r.pairs.empty()) {
            const uint32_t bestScore = r.pairs.front().distance(); // the pairs are sorted by distance
            uint32_t nPairs = 0;
            for (const auto& p : r.pairs)
                nPairs += p.distance() == bestScore;
            const bool mapped = r.pairs.front().up.valid && r.pairs.front().down.valid;
            const bool mappedHalf = !mapped && (r.pairs.front().up.valid || r.pairs.front().down.valid);
            bool primary = true, firstWrite = true;
            for (const auto& p : r.pairs) {
                out.push_back(pairedRecord(rp, p.up, p.down, p, nPairs, bestScore, primary, seqNames));
                if (!mappedHalf || firstWrite)
                    out.push_back(pairedRecord(rp, p.down, p.up, p, nPairs, bestScore, primary, seqNames));
                firstWrite = false;
                primary = false;
            }
        }
        for (size_t i = 0; i < r.unpaired.size(); i++) {
            const MateOcc& o = r.unpaired[i];
            SamRecord s;
            s.qname = o.first ? rp.name1 : rp.name2;
            if (!o.valid) {
                s.flag = o.unmappedFlags;
            } else { // generateSAMUnpaired, src/indexhelpers.cpp:213-262
                s.flag = 1u + (o.first ? 64u : 128u);
                s.flag |= r.unpairedPrimary[i] ? 0u : 256u;
                s.rname = seqNames[o.seqID];
                s.pos = o.begin + 1;
                s.mapq = mapQ(o.distance, r.unpairedBestCount[i], r.unpairedBest[i]);
                s.cigar = o.cigar;
                s.nm = (int)o.distance;
            }
            out.push_back(s);
        }
        return out;
    }

  private:
    struct NeedMapping {};
    // the replay driver: evaluates every pair, serves the mappings they ask for, evaluates again.  se1 / se2 /
    // read2done: single-end results of the inference pre-pass (pairSingleEnd), else null.
    void runBatch(const std::vector<ReadPair>& input, bool best, uint32_t maxEDOrIdentity, const PEOptions& opt,
                  Counters& counters, std::vector<PairResult>& out, const std::vector<std::vector<MateOcc>>* se1,
                  const std::vector<std::vector<MateOcc>>* se2, const std::vector<bool>* read2done) {
        const size_t n = input.size();
        out.assign(n, PairResult());
        std::vector<PairContext> ctx(n);
        std::vector<bool> done(n, false);
        for (size_t i = 0; i < n; i++) {
            ctx[i].owner = this;
            ctx[i].pair = &input[i];
            if (se1) {
                ctx[i].se[0] = &(*se1)[i];
                ctx[i].se[1] = &(*se2)[i];
                ctx[i].read2done = (*read2done)[i];
            }
        }
        for (int round = 0; round < 200; round++) {
            bool pending = false;
            for (size_t i = 0; i < n; i++) {
                if (done[i])
                    continue;
                try {
                    PairResult r;
                    if (best)
                        evalBest(ctx[i], maxEDOrIdentity, opt, r);
                    else
                        evalAll(ctx[i], maxEDOrIdentity, opt, r);
                    out[i] = std::move(r);
                    done[i] = true;
                    ctx[i].cache.clear();
                } catch (const NeedMapping&) {
                    pending = true;
                }
            }
            if (!pending)
                break;
            serveRequests(ctx, best, maxEDOrIdentity, counters);
        }
        for (size_t i = 0; i < n; i++)
            if (!done[i])
                throw std::runtime_error("paired-end mapping did not converge");
    }
    // pair counters, src/fastq.cpp:662-688
    static void addPairCounters(const std::vector<PairResult>& out, Counters& counters) {
        uint64_t c[CLB_N_COUNTERS] = {};
        for (const auto& r : out) {
            const bool mapped = !r.pairs.empty() && r.pairs.front().up.valid && r.pairs.front().down.valid;
            const bool mappedHalf = !mapped && !r.pairs.empty() && (r.pairs.front().up.valid || r.pairs.front().down.valid);
            c[CLB_CTR_TOTAL_UNIQUE_PAIRS] += mapped ? r.pairs.size() : 0;
            c[CLB_CTR_MAPPED_PAIRS] += mapped;
            c[CLB_CTR_DISCORDANTLY_MAPPED_PAIRS] += mapped && r.pairs.front().discordant;
            c[CLB_CTR_MAPPED_HALF_PAIRS] += mappedHalf;
            c[CLB_CTR_UNPAIRED_BUT_MAPPED_PAIRS] += !r.unpaired.empty();
        }
        counters.add(c);
    }
    // single-end ALL mode (both strands filtered together, matchApproxAllMap src/searchstrategy.cpp:495-535) for the
    // inference pre-pass: reads[i] at distance k[i]
    std::vector<std::vector<MateOcc>> mapSingleEndAll(const std::vector<std::string>& reads, const std::vector<uint32_t>& k,
                                                      bool first, Counters& counters) {
        std::vector<std::vector<MateOcc>> out(reads.size());
        std::map<uint32_t, std::vector<size_t>> byK;
        for (size_t i = 0; i < reads.size(); i++)
            byK[k[i]].push_back(i);
        for (auto& g : byK) {
            std::vector<std::string> sub;
            for (size_t i : g.second)
                sub.push_back(reads[i]);
            std::vector<std::vector<TextOcc>> res;
            strategy.mapChunk(sub, g.first, 0, counters, res);
            for (size_t j = 0; j < g.second.size(); j++)
                for (auto& t : res[j]) {
                    MateOcc m;
                    static_cast<TextOcc&>(m) = t;
                    m.first = first;
                    out[g.second[j]].push_back(std::move(m));
                }
        }
        return out;
    }
    // inferPairedEndParameters, src/parallel.cpp:376-465
    static uint32_t calcMedian(std::vector<uint32_t>& d) {
        std::sort(d.begin(), d.end());
        return d.size() % 2 == 0 ? (d[d.size() / 2 - 1] + d[d.size() / 2]) / 2 : d[d.size() / 2];
    }
    static float average(const std::vector<uint32_t>& v) {
        double s = 0.0;
        for (uint32_t x : v)
            s += x;
        return (float)(s / v.size());
    }
    static float stddev(const std::vector<uint32_t>& v, float mean) {
        float accum = 0.0f;
        for (uint32_t x : v) {
            const float d = (float)x;
            accum += (d - mean) * (d - mean);
        }
        return std::sqrt(accum / (v.size() - 1));
    }
    static void inferParameters(std::vector<uint32_t>& fragSizes, const std::vector<Orientation>& orientations, PEInference& inf) {
        if (fragSizes.empty())
            return;
        inf.inferred = true;
        const int32_t median = (int32_t)calcMedian(fragSizes);
        std::vector<uint32_t> absDev;
        for (uint32_t f : fragSizes)
            absDev.push_back((uint32_t)std::abs((int32_t)f - median));
        const uint32_t mad = calcMedian(absDev);
        std::vector<uint32_t> filtered;
        for (uint32_t f : fragSizes)
            if ((uint32_t)std::abs((int32_t)f - median) < 6u * mad)
                filtered.push_back(f);
        float mean = average(filtered), sd = stddev(filtered, mean);
        if (mean == 0 || sd == 0) {
            mean = average(fragSizes);
            sd = stddev(fragSizes, mean);
        }
        inf.meanInsertSize = mean;
        inf.stddevInsertSize = sd;
        const uint32_t maxDev = (uint32_t)(6 * sd);
        inf.maxInsertSize = (uint32_t)(mean + maxDev);
        inf.minInsertSize = mean > maxDev ? (uint32_t)(mean - maxDev) : 0u;
        size_t rf = 0, fr = 0, ff = 0;
        for (Orientation o : orientations) {
            rf += o == RF;
            fr += o == FR;
            ff += o == FF;
        }
        inf.orientation = (fr >= rf && fr >= ff) ? FR : (rf >= ff ? RF : FF);
    }
    typedef std::tuple<int, int, uint32_t, uint32_t> Key; // mate, strand, maximal distance, minimal distance
    struct BoolAndVector {
        bool first = false;
        std::vector<MateOcc> second;
    };
    typedef std::vector<BoolAndVector> OccVector;
    enum SeqNameFound { FOUND, FOUND_WITH_TRIMMING, NOT_FOUND };

    struct PairContext {
        PairedEndMapper* owner = nullptr;
        const ReadPair* pair = nullptr;
        std::map<Key, std::vector<MateOcc>> cache;
        std::set<Key> wanted;
        const std::vector<MateOcc>* se[2] = {nullptr, nullptr}; // single-end results of the inference pre-pass
        bool read2done = false;
        const std::string& seq(int mate) const { return mate == 0 ? pair->read1 : pair->read2; }
        // SearchStrategy::mapRead, src/searchstrategy.h:490-517
        std::vector<MateOcc> mapRead(int mate, int strand, uint32_t maxD, uint32_t minD = 0) {
            const Key k(mate, strand, maxD, maxD == 0 ? 0u : minD);
            auto it = cache.find(k);
            if (it == cache.end()) {
                wanted.insert(k);
                throw NeedMapping();
            }
            return it->second;
        }
    };

    // ---- device passes ---------------------------------------------------------------------------------------
    void serveRequests(std::vector<PairContext>& ctx, bool best, uint32_t maxEDOrIdentity, Counters& counters) {
        // (maxD, minD, trimMax) -> (pair, mate) list
        std::map<std::tuple<uint32_t, uint32_t, uint32_t>, std::vector<std::pair<size_t, int>>> groups;
        for (size_t i = 0; i < ctx.size(); i++) {
            std::set<std::pair<int, std::pair<uint32_t, uint32_t>>> seen;
            for (const Key& k : ctx[i].wanted) {
                const int mate = std::get<0>(k);
                const uint32_t maxD = std::get<2>(k), minD = std::get<3>(k);
                if (!seen.insert({mate, {maxD, minD}}).second)
                    continue;
                // clipping bound of findSeqName: the maximal distance (ALL mode) or the mate's cut-off (BEST mode)
                const uint32_t trim = best ? strategy.getMaxED(maxEDOrIdentity, (uint32_t)ctx[i].seq(mate).size()) : 0u;
                groups[std::make_tuple(maxD, minD, trim)].push_back({i, mate});
            }
            ctx[i].wanted.clear();
        }
        strategy.setFilterPerStrand(true);
        try {
            for (auto& g : groups) {
                const uint32_t maxD = std::get<0>(g.first), minD = std::get<1>(g.first), trim = std::get<2>(g.first);
                strategy.setTrimMax(trim);
                std::vector<std::string> reads;
                for (auto& pm : g.second)
                    reads.push_back(ctx[pm.first].seq(pm.second));
                std::vector<std::vector<TextOcc>> res;
                strategy.mapChunk(reads, maxD, minD, counters, res);
                for (size_t j = 0; j < g.second.size(); j++)
                    for (int strand = 0; strand < 2; strand++) {
                        std::vector<MateOcc> v;
                        for (auto& t : res[2 * j + strand]) {
                            MateOcc m;
                            static_cast<TextOcc&>(m) = t;
                            m.first = g.second[j].second == 0;
                            v.push_back(std::move(m));
                        }
                        ctx[g.second[j].first].cache[Key(g.second[j].second, strand, maxD, minD)] = std::move(v);
                    }
            }
        } catch (...) {
            strategy.setTrimMax(0);
            strategy.setFilterPerStrand(false);
            throw;
        }
        strategy.setTrimMax(0);
        strategy.setFilterPerStrand(false);
    }

    // ---- shared helpers --------------------------------------------------------------------------------------
    // assignSequence(AndCIGAR): the device has assigned the sequence and computed the CIGAR; what is left of
    // the reference's return value is whether the alignment was clipped.  Occurrences that cannot be assigned
    // never come back from the device (the reference drops them at this point).
    static SeqNameFound assignSequence(MateOcc& o) {
        if (o.trimmed) {
            o.trimmed = false; // a second call on an assigned occurrence finds it inside its sequence
            o.distanceBeforeTrimming = o.distance;
            return FOUND_WITH_TRIMMING;
        }
        return FOUND;
    }
    static MateOcc unmappedOcc(bool first, bool mateMapped = false, Strand mateStrand = FORWARD_STRAND) {
        MateOcc t; // createUnmappedSAMOccurrencePE, src/indexhelpers.cpp:181-211
        t.valid = false;
        t.first = first;
        t.distance = 0;
        t.strand = FORWARD_STRAND;
        t.unmappedFlags = (uint8_t)(1 | 4 | (mateMapped ? 0 : 8) | (mateStrand == REVERSE_C_STRAND ? 32 : 0) | (first ? 64 : 128));
        return t;
    }
    static int mapQ(uint32_t distance, uint32_t hits, uint32_t minScore) { // src/indexhelpers.h:378-388
        if (distance != minScore)
            return 0;
        if (hits == 1)
            return 60;
        return (int)std::round(-10.0 * std::log10(1 - 1.0 / hits));
    }
    SamRecord pairedRecord(const ReadPair& rp, const MateOcc& o, const MateOcc& mate, const PairedTextOccs& p,
                           uint32_t nPairs, uint32_t minScore, bool primary, const std::vector<std::string>& seqNames) const {
        SamRecord s;
        s.qname = o.first ? rp.name1 : rp.name2;
        if (!o.valid) {
            s.flag = o.unmappedFlags;
            return s;
        }
        // getFlagsPE, src/indexhelpers.h:340-370
        uint32_t f = 1;
        f |= (uint32_t)(!p.discordant && mate.valid && o.valid) << 1;
        f |= (uint32_t)(!mate.valid) << 3;
        f |= (uint32_t)(o.strand == REVERSE_C_STRAND) << 4;
        f |= (uint32_t)(mate.strand == REVERSE_C_STRAND) << 5;
        f |= (uint32_t)o.first << 6;
        f |= (uint32_t)mate.first << 7;
        f |= (uint32_t)(!primary) << 8;
        s.flag = f;
        s.rname = seqNames[o.seqID];
        s.pos = o.begin + 1;
        // getMapQPairedEnd, src/indexhelpers.h:397-408
        s.mapq = (o.distance + mate.distance > minScore) ? 0 : (nPairs == 1 ? 60 : (int)std::round(-10.0 * std::log10(1 - 1.0 / nPairs)));
        s.cigar = o.cigar;
        s.rnext = mate.valid ? seqNames[mate.seqID] : "*";
        s.pnext = mate.valid ? mate.begin + 1 : 0;
        const int64_t ins = mate.valid ? (int64_t)p.fragSize : 0;
        s.tlen = (mate.valid && o.begin > mate.begin) ? -ins : ins;
        s.nm = (int)o.distance;
        return s;
    }
    // generateSAMPairedEnd: stable sort on the total distance, src/searchstrategy.cpp:1904-1925
    static void sortPairs(std::vector<PairedTextOccs>& pairs) {
        std::stable_sort(pairs.begin(), pairs.end(),
                         [](const PairedTextOccs& a, const PairedTextOccs& b) { return a.distance() < b.distance(); });
    }
    static std::vector<MateOcc>::iterator findFirstOccAfter(std::vector<MateOcc>& occs, uint32_t position) {
        return std::lower_bound(occs.begin(), occs.end(), position,
                                [](const MateOcc& o, uint32_t pos) { return o.indexBegin < pos; });
    }
    static bool occLess(const MateOcc& a, const MateOcc& b) { // TextOcc::operator<, src/indexhelpers.h:779-798
        if (a.indexBegin != b.indexBegin)
            return a.indexBegin < b.indexBegin;
        if (a.distance != b.distance)
            return a.distance < b.distance;
        return (a.end - a.begin) < (b.end - b.begin);
    }
    static PairedTextOccs makePair(const MateOcc& up, const MateOcc& down) {
        PairedTextOccs p;
        p.up = up;
        p.down = down;
        p.fragSize = down.end - up.begin;
        return p;
    }

    // addUnpairedMatches (one read), src/searchstrategy.cpp:1398-1449
    void addUnpairedMatches(std::vector<MateOcc>& all, bool first, const PEOptions& opt, PairResult& r) {
        std::vector<MateOcc> temp;
        for (auto& o : all) {
            assignSequence(o);
            temp.push_back(o);
        }
        if (temp.empty()) {
            if (opt.unmappedSAM) {
                r.unpaired.push_back(unmappedOcc(first));
                r.unpairedBestCount.push_back(0);
                r.unpairedBest.push_back(0);
                r.unpairedPrimary.push_back(true);
            }
            return;
        }
        std::stable_sort(temp.begin(), temp.end(), [](const MateOcc& a, const MateOcc& b) { return a.distance < b.distance; });
        const uint32_t best = temp.front().distance;
        uint32_t bestCount = 0;
        for (auto& o : temp)
            bestCount += o.distance == best;
        bool firstOcc = true;
        for (auto& o : temp) {
            r.unpaired.push_back(o);
            r.unpairedBestCount.push_back(bestCount);
            r.unpairedBest.push_back(best);
            r.unpairedPrimary.push_back(firstOcc);
            firstOcc = false;
        }
    }
    void addUnpairedMatches4(std::vector<MateOcc>& fw1, std::vector<MateOcc>& rc1, std::vector<MateOcc>& fw2,
                             std::vector<MateOcc>& rc2, const PEOptions& opt, PairResult& r) {
        r.unpaired.clear(); // src/searchstrategy.h:1216-1228
        r.unpairedBestCount.clear();
        r.unpairedBest.clear();
        r.unpairedPrimary.clear();
        std::vector<MateOcc> a1(fw1), a2(fw2);
        a1.insert(a1.end(), rc1.begin(), rc1.end());
        a2.insert(a2.end(), rc2.begin(), rc2.end());
        fw1.clear(), rc1.clear(), fw2.clear(), rc2.clear();
        addUnpairedMatches(a1, true, opt, r);
        addUnpairedMatches(a2, false, opt, r);
    }
    void addBothUnmapped(const PEOptions& opt, std::vector<PairedTextOccs>& pairs) {
        if (!opt.unmappedSAM)
            return;
        PairedTextOccs p;
        p.up = unmappedOcc(true);
        p.down = unmappedOcc(false);
        p.fragSize = 0;
        pairs.push_back(p);
    }
    // addOneUnmapped, src/searchstrategy.cpp:1451-1516
    void addOneUnmapped(std::vector<MateOcc>& matches1, std::vector<MateOcc>& matches2, const PEOptions& opt,
                        std::vector<PairedTextOccs>& pairs) {
        const bool firstMapped = !matches1.empty();
        for (auto& o : firstMapped ? matches1 : matches2) {
            assignSequence(o);
            PairedTextOccs p;
            p.up = o;
            p.down = unmappedOcc(!firstMapped, true, o.strand);
            p.fragSize = 0;
            pairs.push_back(p);
        }
        if (pairs.empty())
            addBothUnmapped(opt, pairs);
    }
    // addDiscPairs, src/searchstrategy.cpp:1518-1580
    void addDiscPairs(std::vector<MateOcc>& fw1, std::vector<MateOcc>& rc1, std::vector<MateOcc>& fw2,
                      std::vector<MateOcc>& rc2, std::vector<PairedTextOccs>& pairs) {
        if ((fw1.empty() && rc1.empty()) || (fw2.empty() && rc2.empty()))
            return;
        auto pairOccs = [&](MateOcc& first, MateOcc& second) {
            assignSequence(first);
            assignSequence(second);
            const bool sameRef = first.seqID == second.seqID;
            const bool firstUpstream = first.begin < second.begin;
            const uint32_t fragSize = sameRef ? (firstUpstream ? second.end - first.begin : first.end - second.begin) : 0;
            PairedTextOccs p;
            p.up = firstUpstream ? first : second;
            p.down = firstUpstream ? second : first;
            p.fragSize = fragSize;
            p.discordant = true;
            pairs.push_back(p);
        };
        for (auto& a : fw1) {
            for (auto& b : fw2)
                pairOccs(a, b);
            for (auto& b : rc2)
                pairOccs(a, b);
        }
        for (auto& a : rc1) {
            for (auto& b : fw2)
                pairOccs(a, b);
            for (auto& b : rc2)
                pairOccs(a, b);
        }
    }

    // ---- ALL mode --------------------------------------------------------------------------------------------
    // pairOccurrences, src/searchstrategy.cpp:1281-1343
    void pairOccurrences(std::vector<MateOcc>& ups, std::vector<MateOcc>& downs, std::vector<PairedTextOccs>& pairs,
                         uint32_t maxFrag, uint32_t minFrag) {
        if (ups.empty() || downs.empty())
            return;
        for (auto& up : ups) {
            const uint32_t upstreamPos = up.indexBegin;
            for (auto it = findFirstOccAfter(downs, upstreamPos); it != downs.end(); ++it) {
                const uint32_t fragSize = it->indexEnd() - upstreamPos;
                if (fragSize <= maxFrag && fragSize >= minFrag) {
                    assignSequence(up);
                    assignSequence(*it);
                    if (up.seqID != it->seqID)
                        continue;
                    pairs.push_back(makePair(up, *it));
                } else if (fragSize > maxFrag) {
                    break;
                }
            }
        }
    }
    void evalAll(PairContext& c, uint32_t maxED, const PEOptions& opt, PairResult& r) {
        BoolAndVector m[2][2]; // [mate][strand]
        if (c.se[0]) { // pairSingleEndedMatchesAll, src/searchstrategy.cpp:1345-1396: split by strand, sort on position
            for (int mate = 0; mate < 2; mate++) {
                for (const MateOcc& o : *c.se[mate])
                    m[mate][o.strand == REVERSE_C_STRAND ? 1 : 0].second.push_back(o);
                for (int strand = 0; strand < 2; strand++) {
                    auto& v = m[mate][strand].second;
                    std::stable_sort(v.begin(), v.end(), occLess);
                    m[mate][strand].first = mate == 0 ? true : c.read2done;
                }
            }
        }
        auto lazy = [&](int mate, int strand) -> BoolAndVector& {
            BoolAndVector& v = m[mate][strand];
            if (!v.first) {
                v.second = c.mapRead(mate, strand, maxED);
                v.first = true;
            }
            return v;
        };
        // processComb, src/searchstrategy.h:753-775
        auto processComb = [&](int uMate, int uStrand, int dMate, int dStrand) {
            BoolAndVector& u = lazy(uMate, uStrand);
            if (u.second.empty())
                return;
            BoolAndVector& d = lazy(dMate, dStrand);
            if (d.second.empty())
                return;
            pairOccurrences(u.second, d.second, r.pairs, opt.maxFragSize, opt.minFragSize);
        };
        switch (opt.orientation) { // processCombFRAll / FFAll / RFAll, src/searchstrategy.h:790-861
        case FR:
            processComb(0, 0, 1, 1);
            processComb(1, 0, 0, 1);
            break;
        case FF:
            processComb(0, 0, 1, 0);
            processComb(1, 1, 0, 1);
            break;
        case RF:
            processComb(0, 1, 1, 0);
            processComb(1, 1, 0, 0);
            break;
        }
        if (r.pairs.empty()) { // pairDiscordantly, src/searchstrategy.cpp:1582-1647
            for (int mate = 0; mate < 2; mate++)
                for (int strand = 0; strand < 2; strand++)
                    lazy(mate, strand);
            auto &fw1 = m[0][0].second, &rc1 = m[0][1].second, &fw2 = m[1][0].second, &rc2 = m[1][1].second;
            const size_t matchesFirst = fw1.size() + rc1.size(), matchesSecond = fw2.size() + rc2.size();
            const bool firstMapped = matchesFirst != 0, secondMapped = matchesSecond != 0;
            bool doneDisc = false;
            if (opt.discordantAllowed && firstMapped && secondMapped) {
                if (matchesFirst * matchesSecond > 10000) {
                    addUnpairedMatches4(fw1, rc1, fw2, rc2, opt, r);
                } else {
                    addDiscPairs(fw1, rc1, fw2, rc2, r.pairs);
                    doneDisc = !r.pairs.empty();
                }
            }
            if (!doneDisc) {
                if (firstMapped && secondMapped) {
                    addUnpairedMatches4(fw1, rc1, fw2, rc2, opt, r);
                } else if (!firstMapped && !secondMapped) {
                    addBothUnmapped(opt, r.pairs);
                } else {
                    std::vector<MateOcc> m1(fw1), m2(fw2);
                    m1.insert(m1.end(), rc1.begin(), rc1.end());
                    m2.insert(m2.end(), rc2.begin(), rc2.end());
                    addOneUnmapped(m1, m2, opt, r.pairs);
                }
            }
        }
        sortPairs(r.pairs);
    }

    // ---- BEST mode -------------------------------------------------------------------------------------------
    // processSeq, src/searchstrategy.cpp:777-811
    bool processSeq(PairContext& c, int mate, int strand, uint32_t maxDist, OccVector& v) {
        if (!v[maxDist].first) {
            uint32_t firstOpen = 0;
            while (firstOpen < v.size() && v[firstOpen].first)
                firstOpen++;
            const uint32_t minD = std::min(firstOpen, maxDist);
            std::vector<MateOcc> occs = c.mapRead(mate, strand, maxDist, minD);
            for (auto& o : occs)
                if (o.distanceBeforeTrimming < v.size())
                    v[o.distanceBeforeTrimming].second.push_back(o);
            for (uint32_t i = minD; i <= maxDist; i++)
                v[i].first = true;
        }
        for (uint32_t i = 0; i <= maxDist; i++)
            if (!v[i].second.empty())
                return true;
        return false;
    }
    // handleTrimmedOccs, src/searchstrategy.cpp:813-832
    static void handleTrimmedOccs(std::vector<uint32_t>& ids, uint32_t oDist, OccVector& v) {
        std::stable_sort(ids.begin(), ids.end());
        for (auto it = ids.rbegin(); it != ids.rend(); ++it) {
            MateOcc occ = v[oDist].second[*it];
            v[oDist].second.erase(v[oDist].second.begin() + *it);
            if (occ.distance >= v.size())
                continue;
            auto& tv = v[occ.distance].second;
            auto point = std::lower_bound(tv.begin(), tv.end(), occ, [](const MateOcc& a, const MateOcc& b) {
                if (a.indexBegin != b.indexBegin) // TextOcc::operator<, src/indexhelpers.h:779-798
                    return a.indexBegin < b.indexBegin;
                if (a.distance != b.distance)
                    return a.distance < b.distance;
                return (a.end - a.begin) < (b.end - b.begin);
            });
            tv.insert(point, occ);
        }
    }
    // pairOccurrencesForBestMapping, src/searchstrategy.cpp:1743-1819
    void pairOccurrencesForBestMapping(std::vector<MateOcc>& ups, std::vector<MateOcc>& downs,
                                       std::vector<PairedTextOccs>& pairs, uint32_t maxFrag, uint32_t minFrag,
                                       std::set<uint32_t>& uTrimmed, std::set<uint32_t>& dTrimmed) {
        if (ups.empty() || downs.empty())
            return;
        std::stable_sort(downs.begin(), downs.end(), [](const MateOcc& a, const MateOcc& b) { return a.indexBegin < b.indexBegin; });
        for (uint32_t i = 0; i < ups.size(); i++) {
            MateOcc& up = ups[i];
            const uint32_t upstreamPos = up.indexBegin;
            for (auto it = findFirstOccAfter(downs, upstreamPos); it != downs.end(); ++it) {
                const uint32_t fragSize = it->indexEnd() - upstreamPos;
                if (fragSize <= maxFrag && fragSize >= minFrag) {
                    const SeqNameFound uFound = assignSequence(up);
                    if (uFound != FOUND) {
                        if (uFound == FOUND_WITH_TRIMMING)
                            uTrimmed.insert(i);
                        break;
                    }
                    const SeqNameFound dFound = assignSequence(*it);
                    if (dFound != FOUND) {
                        if (dFound == FOUND_WITH_TRIMMING)
                            dTrimmed.insert((uint32_t)(it - downs.begin()));
                        continue;
                    }
                    if (up.seqID != it->seqID)
                        continue;
                    pairs.push_back(makePair(up, *it));
                } else if (fragSize > maxFrag) {
                    break;
                }
            }
        }
    }
    // processComb (BEST), src/searchstrategy.cpp:834-915
    void processCombBest(PairContext& c, int uMate, int uStrand, int dMate, int dStrand, OccVector& uVector,
                         OccVector& dVector, std::vector<PairedTextOccs>& pairs, const PEOptions& opt, uint32_t totDist) {
        auto findFirstPosDist = [](const OccVector& v) {
            uint32_t i = 0;
            while (i < v.size() && !(!v[i].second.empty() || !v[i].first))
                i++;
            return i;
        };
        const uint32_t minDDist = findFirstPosDist(dVector), minUDist = findFirstPosDist(uVector);
        // unsigned arithmetic as in the reference (length_t)
        uint32_t maxUp = std::min<uint32_t>(totDist - minDDist, (uint32_t)uVector.size() - 1);
        uint32_t maxDown = std::min<uint32_t>(totDist - minUDist, (uint32_t)dVector.size() - 1);
        auto processRead = [&](int mate, int strand, OccVector& v, const uint32_t& max, uint32_t& maxOther) {
            if (!processSeq(c, mate, strand, max, v))
                return false;
            const uint32_t minD = findFirstPosDist(v);
            maxOther = std::min<uint32_t>(totDist - minD, maxOther);
            return true;
        };
        if (maxUp <= maxDown) {
            if (!(processRead(uMate, uStrand, uVector, maxUp, maxDown) && processRead(dMate, dStrand, dVector, maxDown, maxUp)))
                return;
        } else {
            if (!(processRead(dMate, dStrand, dVector, maxDown, maxUp) && processRead(uMate, uStrand, uVector, maxUp, maxDown)))
                return;
        }
        for (uint32_t dist = minUDist + minDDist; dist <= totDist; dist++) {
            for (uint32_t uDist = minUDist; uDist <= std::min(maxUp, dist); uDist++) {
                const uint32_t dDist = dist - uDist;
                if (dDist > maxDown || dDist < minDDist)
                    continue;
                std::set<uint32_t> uTrimmed, dTrimmed;
                pairOccurrencesForBestMapping(uVector[uDist].second, dVector[dDist].second, pairs, opt.maxFragSize,
                                              opt.minFragSize, uTrimmed, dTrimmed);
                std::vector<uint32_t> uIds(uTrimmed.begin(), uTrimmed.end()), dIds(dTrimmed.begin(), dTrimmed.end());
                handleTrimmedOccs(uIds, uDist, uVector);
                handleTrimmedOccs(dIds, dDist, dVector);
            }
            if (!pairs.empty())
                return;
        }
    }
    static void mergeOrMovePairs(std::vector<PairedTextOccs>& p12, std::vector<PairedTextOccs>& p21,
                                 std::vector<PairedTextOccs>& pairs) { // src/searchstrategy.cpp:917-939
        if (p12.empty() || p21.empty()) {
            pairs = p12.empty() ? p21 : p12;
            return;
        }
        const uint32_t d12 = p12.front().distance(), d21 = p21.front().distance();
        if (d12 <= d21) {
            pairs = p12;
            if (d12 == d21)
                pairs.insert(pairs.end(), p21.begin(), p21.end());
            return;
        }
        pairs = p21;
    }
    // processCombFR / FF / RF, src/searchstrategy.cpp:941-1063.  v[mate][strand]
    void processOriBest(PairContext& c, OccVector (&v)[2][2], std::vector<PairedTextOccs>& pairs, const PEOptions& opt,
                        uint32_t totDist, uint32_t minTotDist) {
        std::vector<PairedTextOccs> pA, pB;
        int a[4], b[4]; // combination A / B: upstream mate, strand, downstream mate, strand
        switch (opt.orientation) {
        case FR: { int aa[4] = {0, 0, 1, 1}, bb[4] = {1, 0, 0, 1}; std::copy(aa, aa + 4, a); std::copy(bb, bb + 4, b); break; }
        case FF: { int aa[4] = {0, 0, 1, 0}, bb[4] = {1, 1, 0, 1}; std::copy(aa, aa + 4, a); std::copy(bb, bb + 4, b); break; }
        default: { int aa[4] = {0, 1, 1, 0}, bb[4] = {1, 1, 0, 0}; std::copy(aa, aa + 4, a); std::copy(bb, bb + 4, b); break; }
        }
        auto anyBelow = [&](OccVector& ov, bool whole) {
            const size_t lim = whole ? ov.size() : std::min<size_t>(minTotDist, ov.size());
            for (size_t i = 0; i < lim; i++)
                if (!ov[i].second.empty())
                    return true;
            return false;
        };
        auto runA = [&](uint32_t td) { processCombBest(c, a[0], a[1], a[2], a[3], v[a[0]][a[1]], v[a[2]][a[3]], pA, opt, td); };
        auto runB = [&](uint32_t td) { processCombBest(c, b[0], b[1], b[2], b[3], v[b[0]][b[1]], v[b[2]][b[3]], pB, opt, td); };
        // FF looks at the whole vectors of combination A (matches1, matches2), FR / RF at the strata below minTotDist
        const bool whole = opt.orientation == FF;
        if (anyBelow(v[a[0]][a[1]], whole) || anyBelow(v[a[2]][a[3]], whole)) {
            runA(totDist);
            runB(pA.empty() ? totDist : pA.front().distance());
        } else {
            runB(totDist);
            runA(pB.empty() ? totDist : pB.front().distance());
        }
        mergeOrMovePairs(pA, pB, pairs);
    }
    // checkAlignments, src/searchstrategy.cpp:537-568
    static void checkAlignments(OccVector& ov, uint32_t& best, uint32_t l) {
        std::vector<MateOcc> trimmedOccs, assigned;
        for (auto& occ : ov[l].second) {
            const SeqNameFound f = occ.trimmed ? FOUND_WITH_TRIMMING : FOUND;
            if (f != FOUND) {
                if (occ.distance > l)
                    trimmedOccs.push_back(occ);
            } else {
                assigned.push_back(occ);
                if (l < best)
                    best = l;
            }
        }
        ov[l].second.swap(assigned);
        for (auto& occ : trimmedOccs) {
            occ.trimmed = false;
            occ.distanceBeforeTrimming = occ.distance;
            if (occ.distance < ov.size())
                ov[occ.distance].second.push_back(occ);
        }
    }
    // findBestAlignments, src/searchstrategy.cpp:623-712
    bool findBestAlignments(PairContext& c, int mate, OccVector& fw, OccVector& rc, uint32_t x, uint32_t& best) {
        const uint32_t cutOff = (uint32_t)fw.size() - 1;
        best = cutOff + 1;
        bool bestFound = false;
        if (x == 0) {
            if (!fw[0].first) {
                fw[0].second = c.mapRead(mate, 0, 0);
                fw[0].first = true;
            }
            if (!rc[0].first) {
                rc[0].second = c.mapRead(mate, 1, 0);
                rc[0].first = true;
            }
            if (!fw[0].second.empty() || !rc[0].second.empty()) {
                checkAlignments(fw, best, 0);
                checkAlignments(rc, best, 0);
                if (best == 0)
                    bestFound = true;
            }
        }
        const uint32_t maxED = best == 0 ? x : cutOff;
        uint32_t prevK = 0;
        auto hasUpdate = [&](int strand, uint32_t k, OccVector& v) {
            if (v[k].first)
                return !v[k].second.empty();
            return processSeq(c, mate, strand, k, v);
        };
        for (uint32_t k = std::max(x, 1u); k <= maxED;) {
            bool update = false;
            update |= hasUpdate(0, k, fw);
            update |= hasUpdate(1, k, rc);
            if (update)
                for (uint32_t l = prevK + 1; l <= std::min(k, best + x); l++) {
                    checkAlignments(fw, best, l);
                    checkAlignments(rc, best, l);
                }
            if (bestFound)
                break;
            if (update && best < cutOff + 1) {
                bestFound = true;
                if (x == 0)
                    break;
                prevK = k;
                k = std::min(best + x, maxED);
            } else {
                if (k == maxED)
                    break;
                const uint32_t step = k < 5 ? 2 : 4;
                prevK = k;
                k = std::min(k + x + step, maxED);
            }
        }
        return bestFound;
    }
    // combineOccVectors, src/searchstrategy.cpp:570-621
    static std::vector<MateOcc> combineOccVectors(OccVector& fw, OccVector& rc, uint32_t best, uint32_t max) {
        std::vector<MateOcc> out;
        for (uint32_t i = best; i <= max; i++)
            for (OccVector* ov : {&fw, &rc}) {
                auto& v = (*ov)[i].second;
                std::stable_sort(v.begin(), v.end(), [](const MateOcc& a, const MateOcc& b) {
                    return a.seqID < b.seqID || (a.seqID == b.seqID && a.begin < b.begin);
                });
                for (size_t j = 0; j < v.size(); j++)
                    if (j == 0 || v[j].seqID != v[j - 1].seqID || v[j].begin != v[j - 1].begin)
                        out.push_back(v[j]);
            }
        return out;
    }
    std::vector<MateOcc> findBestMapping(PairContext& c, int mate, OccVector& fw, OccVector& rc, uint32_t x) {
        uint32_t best;
        if (findBestAlignments(c, mate, fw, rc, x, best))
            return combineOccVectors(fw, rc, best, std::min(best + x, (uint32_t)fw.size() - 1));
        return {};
    }
    // pairDiscordantlyBest, src/searchstrategy.cpp:1664-1741
    void pairDiscordantlyBest(PairContext& c, OccVector (&v)[2][2], const PEOptions& opt, PairResult& r) {
        OccVector &fw1 = v[0][0], &rc1 = v[0][1], &fw2 = v[1][0], &rc2 = v[1][1];
        const uint32_t max1 = (uint32_t)fw1.size() - 1, max2 = (uint32_t)fw2.size() - 1;
        if (opt.discordantAllowed) {
            uint32_t bestStratum = (uint32_t)(fw1.size() + fw2.size() + 1);
            bool bestFound = false;
            auto mapStratum = [&](OccVector& ov, uint32_t d, int mate, int strand) { // src/searchstrategy.h:1354-1361
                if (!ov[d].first) {
                    ov[d].second = c.mapRead(mate, strand, d, d);
                    ov[d].first = true;
                }
            };
            for (uint32_t i = 0; i < fw1.size() + fw2.size(); i++) {
                if (i <= max1) {
                    mapStratum(fw1, i, 0, 0);
                    mapStratum(rc1, i, 0, 1);
                }
                if (i <= max2) {
                    mapStratum(fw2, i, 1, 0);
                    mapStratum(rc2, i, 1, 1);
                }
                const uint32_t min1 = i > max2 ? i - max2 : 0;
                for (uint32_t e1 = min1; e1 <= std::min(i, max1); e1++) {
                    const uint32_t e2 = i - e1;
                    addDiscPairs(fw1[e1].second, rc1[e1].second, fw2[e2].second, rc2[e2].second, r.pairs);
                }
                if (!r.pairs.empty()) {
                    if (!bestFound) {
                        bestStratum = i;
                        bestFound = true;
                    }
                    if (i == bestStratum + opt.x)
                        return;
                }
            }
        }
        std::vector<MateOcc> best1 = findBestMapping(c, 0, fw1, rc1, opt.x);
        std::vector<MateOcc> best2 = findBestMapping(c, 1, fw2, rc2, opt.x);
        if (best1.empty() && best2.empty()) {
            addBothUnmapped(opt, r.pairs);
        } else if (best1.empty() || best2.empty()) {
            addOneUnmapped(best1, best2, opt, r.pairs);
        } else {
            addUnpairedMatches(best1, true, opt, r);
            addUnpairedMatches(best2, false, opt, r);
        }
    }
    // matchApproxPairedEndBestPlusX, src/searchstrategy.cpp:1091-1179
    void evalBest(PairContext& c, uint32_t minIdentity, const PEOptions& opt, PairResult& r) {
        const uint32_t cutOff1 = strategy.getMaxED(minIdentity, (uint32_t)c.pair->read1.size());
        const uint32_t cutOff2 = strategy.getMaxED(minIdentity, (uint32_t)c.pair->read2.size());
        const uint32_t x = opt.x;
        uint32_t best = cutOff1 + cutOff2 + 1;
        OccVector v[2][2];
        v[0][0].resize(cutOff1 + 1);
        v[0][1].resize(cutOff1 + 1);
        v[1][0].resize(cutOff2 + 1);
        v[1][1].resize(cutOff2 + 1);
        if (c.se[0]) { // addSingleEndedForBest, src/searchstrategy.cpp:1065-1089
            for (int mate = 0; mate < 2; mate++) {
                for (const MateOcc& o : *c.se[mate]) {
                    OccVector& ov = v[mate][o.strand == REVERSE_C_STRAND ? 1 : 0];
                    if (o.distanceBeforeTrimming < ov.size())
                        ov[o.distanceBeforeTrimming].second.push_back(o);
                }
                for (int strand = 0; strand < 2; strand++)
                    for (auto& bv : v[mate][strand])
                        bv.first = mate == 0 ? true : c.read2done;
            }
        }
        uint32_t minDistNotExplored = 0;
        if (x == 0) {
            processOriBest(c, v, r.pairs, opt, 0, 0);
            minDistNotExplored = 1;
        }
        bool bestFound = false;
        if (!r.pairs.empty()) {
            best = 0;
            bestFound = true;
        }
        uint32_t maxStratum = best == 0 ? x : cutOff1 + cutOff2;
        for (uint32_t k = std::max(x, 1u); k <= maxStratum;) {
            processOriBest(c, v, r.pairs, opt, k, minDistNotExplored);
            if (!bestFound) {
                if (!r.pairs.empty()) {
                    best = k;
                    bestFound = true;
                    maxStratum = std::min(best + x, cutOff1 + cutOff2);
                    minDistNotExplored = k + 1;
                    if (x == 0)
                        break;
                    k = maxStratum;
                } else {
                    if (k == maxStratum)
                        break;
                    const uint32_t step = k < 6 ? 2 : 4;
                    k = std::min(maxStratum, k + x + step);
                }
            } else {
                break;
            }
        }
        if (r.pairs.empty())
            pairDiscordantlyBest(c, v, opt, r);
        sortPairs(r.pairs);
    }

    SearchStrategy& strategy;
};

} // namespace clb
#endif
