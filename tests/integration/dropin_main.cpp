// Drop-in check (test infrastructure): a driver that maps a FASTQ file in ALL mode through B200Strategy and
// writes the SAM body with the REFERENCE's own record formatting (TextOcc::generateSAMSingleEnd*,
// src/indexhelpers.h:625-680, src/indexhelpers.cpp:56-87), following SearchStrategy::generateOutputSingleEnd
// (src/searchstrategy.cpp:1824-1902, DEVELOPER_MODE branch) and generateSE_SAM (src/searchstrategy.h:1633-1641).
// Its output must equal the record lines of `columba_dev -a all` byte for byte.
//   columba_dropin <index base> <fastq> <schemes dir> <k> <out.sam>
#include <algorithm>
#include <fstream>
#include <iostream>

#include "b200strategy.h"

int main(int argc, char** argv) {
    if (argc < 6) {
        std::cerr << "usage: columba_dropin base fastq schemes k out.sam\n";
        return 2;
    }
    try {
        B200Strategy strategy(argv[1], 4, 10, 4, false, "named", "columba", argv[3], true, 2 /* dynamic */, 0);
        const length_t maxED = (length_t)std::stoul(argv[4]);
        std::vector<ReadBundle> chunk;
        {
            std::ifstream f(argv[2]);
            std::string id, seq, plus, qual;
            while (std::getline(f, id) && std::getline(f, seq) && std::getline(f, plus) && std::getline(f, qual)) {
                Read r(id, seq, qual);
                chunk.emplace_back(r);
            }
        }
        Counters counters;
        counters.resetCounters();
        std::vector<std::vector<TextOcc>> result;
        strategy.matchApproxBatch(chunk, maxED, counters, result);
        std::ofstream out(argv[5]);
        const auto& seqNames = strategy.getSeqNames();
        for (size_t i = 0; i < chunk.size(); i++) {
            auto& occs = result[i];
            if (occs.empty()) {
                out << TextOcc::createUnmappedSAMOccurrenceSE(chunk[i]).getOutputLine();
                continue;
            }
            std::stable_sort(occs.begin(), occs.end(),
                             [](const TextOcc& a, const TextOcc& b) { return a.getDistance() < b.getDistance(); });
            const uint32_t minScore = occs.front().getDistance();
            length_t nHits = 0;
            for (const auto& o : occs)
                nHits += o.getDistance() == minScore;
            occs.front().generateSAMSingleEndFirst(chunk[i], nHits, minScore, seqNames);
            for (size_t j = 1; j < occs.size(); j++)
                occs[j].generateSAMSingleEndNotFirst(chunk[i].getSeqID(), nHits, minScore, seqNames);
            for (const auto& o : occs)
                out << o.getOutputLine();
        }
        std::cerr << "reads " << chunk.size() << " nodes " << counters.get(Counters::NODE_COUNTER) << "\n";
    } catch (std::exception& e) {
        std::cerr << "error: " << e.what() << "\n";
        return 1;
    }
    return 0;
}
