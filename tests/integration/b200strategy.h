// The reference-side binding of INTEGRATION.md, as a file that is compiled against the reference's OWN headers
// (/root/reference/src, with the header shims of oracle/shim) by `make -C oracle dropin`: the batched seam a
// Columba maintainer would add next to SearchStrategy::matchApprox.  Test infrastructure: it exists to prove that
// the seam fits the reference's types (ReadBundle, TextOcc, Range, Counters, SeqNameFound); the product's own
// host layer is include/columba_b200.hpp.
#ifndef B200STRATEGY_H
#define B200STRATEGY_H
#include <stdexcept>
#include <string>
#include <vector>

#include "columba_b200.h"   // this repository: include/columba_b200.h
#include "indexhelpers.h"   // reference: TextOcc, Range, Counters
#include "reads.h"          // reference: ReadBundle

class B200Strategy {
  public:
    // mirrors FMIndex(baseFile, inTextSwitch, noCIGAR, sa_sparse, verbose, wordSize) + Parameters::createStrategy
    B200Strategy(const std::string& base, int sparsenessFactor, int kmerSize, int inTextSwitch, bool noCIGAR,
                 const std::string& kind, const std::string& arg, const std::string& schemesDir, bool editDistance,
                 int partitioning, int device) {
        if (clb_index_load(base.c_str(), sparsenessFactor, kmerSize, inTextSwitch, noCIGAR, device, &ix))
            throw std::runtime_error(clb_last_error(nullptr)); // same texts as the reference
        if (clb_ctx_create(ix, &ctx))
            throw std::runtime_error(clb_last_error(nullptr));
        check(clb_ctx_set_strategy(ctx, kind.c_str(), arg.c_str(), schemesDir.c_str()));
        check(clb_ctx_set_options(ctx, editDistance ? CLB_METRIC_EDIT : CLB_METRIC_HAMMING, partitioning));
        for (uint32_t i = 0; i < clb_index_num_seqs(ix); i++)
            seqNames.push_back(clb_index_seq_name(ix, i));
    }
    ~B200Strategy() {
        clb_ctx_free(ctx);
        clb_index_free(ix);
    }
    const std::vector<std::string>& getSeqNames() const { return seqNames; }

    // the new seam: SearchStrategy::matchApprox (src/searchstrategy.h:2021) for a whole chunk, ALL mode
    void matchApproxBatch(std::vector<ReadBundle>& chunk, length_t maxED, Counters& counters,
                          std::vector<std::vector<TextOcc>>& result) {
        std::string cat;
        std::vector<uint32_t> off(1, 0);
        for (auto& b : chunk) {
            cat += b.getRead();
            off.push_back((uint32_t)cat.size());
        }
        clb_result r;
        check(clb_map_batch(ctx, (const uint8_t*)cat.data(), off.data(), (uint32_t)chunk.size(), (int)maxED, 0, &r));
        result.clear();
        result.resize(chunk.size());
        for (uint32_t i = 0; i < r.n_reads; i++)
            for (uint32_t j = r.read_start[i]; j < r.read_start[i] + r.read_count[i]; j++) {
                const clb_occ& o = r.occ[j];
                std::string cigar; // runs are (len << 2) | op, op 0=M 1=I 2=D
                for (uint32_t c = 0; c < o.n_cigar; c++) {
                    const uint16_t run = r.cigar[(size_t)j * r.cigar_stride + c];
                    cigar += std::to_string(run >> 2);
                    cigar += "MID"[run & 3];
                }
                // after findSeqName a TextOcc's range is relative to its sequence (src/indexinterface.cpp:799-897)
                TextOcc t(Range(o.pos, o.end), o.dist, cigar, o.strand ? REVERSE_C_STRAND : FORWARD_STRAND, FIRST_IN_PAIR);
                t.setIndexBegin(o.index_begin);
                t.setAssignedSequence(o.trimmed ? FOUND_WITH_TRIMMING : FOUND, o.seq_id);
                result[i].emplace_back(std::move(t));
            }
        for (int c = 0; c < CLB_N_COUNTERS; c++) // enum order == Counters::CounterType
            counters.inc((Counters::CounterType)c, r.counters[c]);
    }

  private:
    void check(int rc) {
        if (rc)
            throw std::runtime_error(clb_last_error(ctx));
    }
    clb_index* ix = nullptr;
    clb_ctx* ctx = nullptr;
    std::vector<std::string> seqNames;
};
#endif
