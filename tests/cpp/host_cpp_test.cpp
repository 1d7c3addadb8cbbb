// Test program of the C++ host layer (include/columba_b200.hpp): maps a FASTQ
// file in ALL or BEST mode and prints canonical tuples.  Linked against the
// CUDA library on the GPU box and against the host emulation of the kernels
// (same C ABI) in the CPU tests.
//   host_cpp_test <index base> <fastq> <schemes dir> all <k> | sam <k> | best <min identity> [x]
#include <fstream>
#include <iostream>

#include "columba_b200.hpp"

int main(int argc, char** argv) {
    if (argc < 6) {
        std::cerr << "usage: host_cpp_test base fastq schemes all|best value\n";
        return 2;
    }
    try {
        clb::FMIndex index(argv[1]);
        clb::SearchStrategy strategy(index, "named", "columba", argv[3]);
        std::vector<std::string> names, reads, quals;
        std::ifstream f(argv[2]);
        std::string l1, l2, l3, l4;
        while (std::getline(f, l1) && std::getline(f, l2) && std::getline(f, l3) && std::getline(f, l4)) {
            names.push_back(l1.substr(1, l1.find_first_of(" \t") == std::string::npos ? std::string::npos
                                                                                       : l1.find_first_of(" \t") - 1));
            for (auto& c : l2)
                c = (char)toupper(c);
            reads.push_back(l2);
            quals.push_back(l4);
        }
        clb::Counters counters;
        std::vector<std::vector<clb::TextOcc>> occ;
        const std::string mode = argv[4];
        const uint32_t v = (uint32_t)std::stoul(argv[5]);
        if (mode == "all" || mode == "sam")
            strategy.matchApproxBatch(reads, v, counters, occ);
        else
            strategy.matchBestBatch(reads, v, argc > 6 ? (uint32_t)std::stoul(argv[6]) : 0u, counters, occ);
        const auto& sn = index.getSeqNames();
        if (mode == "sam") { // the SAM body of single-end ALL mode
            std::string body;
            for (size_t i = 0; i < reads.size(); i++)
                clb::samSingleEnd(names[i], reads[i], quals[i], occ[i], sn, body);
            std::cout << body;
            return 0;
        }
        for (size_t i = 0; i < reads.size(); i++)
            for (size_t j = 0; j < occ[i].size(); j++) { // BEST mode: the first record is the primary alignment
                const auto& t = occ[i][j];
                std::cout << names[i] << '\t' << ((t.strand == clb::REVERSE_C_STRAND ? 16 : 0) | (mode == "best" && j ? 256 : 0))
                          << '\t' << sn[t.seqID] << '\t' << t.begin + 1 << '\t' << t.distance << '\t' << t.cigar << '\n';
            }
        std::cerr << "reads " << counters.get(CLB_CTR_NUMBER_OF_READS) << " mapped " << counters.get(CLB_CTR_MAPPED_READS)
                  << "\n";
    } catch (std::exception& e) {
        std::cerr << "error: " << e.what() << "\n";
        return 1;
    }
    return 0;
}
