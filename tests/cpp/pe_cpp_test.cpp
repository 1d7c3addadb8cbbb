// Test program of the paired-end host layer (include/columba_b200_pe.hpp): maps two FASTQ files as pairs
// and prints the fields of every SAM record the reference would write.
//   pe_cpp_test <index base> <fastq 1> <fastq 2> <schemes dir> all <k> | best <min identity>
//               <fr|rf|ff> <max frag> <min frag> <discordant 0|1> <x> [infer]
#include <fstream>
#include <iostream>

#include "columba_b200_pe.hpp"

static void readFastq(const char* path, std::vector<std::string>& names, std::vector<std::string>& reads) {
    std::ifstream f(path);
    std::string l1, l2, l3, l4;
    while (std::getline(f, l1) && std::getline(f, l2) && std::getline(f, l3) && std::getline(f, l4)) {
        const size_t sp = l1.find_first_of(" \t");
        names.push_back(l1.substr(1, sp == std::string::npos ? std::string::npos : sp - 1));
        for (auto& c : l2)
            c = (char)toupper(c);
        reads.push_back(l2);
    }
}

int main(int argc, char** argv) {
    if (argc < 12) {
        std::cerr << "usage: pe_cpp_test base fq1 fq2 schemes all|best value fr|rf|ff maxFrag minFrag discordant x\n";
        return 2;
    }
    try {
        clb::FMIndex index(argv[1]);
        clb::SearchStrategy strategy(index, "named", "columba", argv[4]);
        std::vector<std::string> n1, r1, n2, r2;
        readFastq(argv[2], n1, r1);
        readFastq(argv[3], n2, r2);
        std::vector<clb::ReadPair> pairs(r1.size());
        for (size_t i = 0; i < r1.size(); i++)
            pairs[i] = clb::ReadPair{n1[i], r1[i], n2[i], r2[i]};
        clb::PEOptions opt;
        const std::string ori = argv[7];
        opt.orientation = ori == "fr" ? clb::FR : (ori == "rf" ? clb::RF : clb::FF);
        opt.maxFragSize = (uint32_t)std::stoul(argv[8]);
        opt.minFragSize = (uint32_t)std::stoul(argv[9]);
        opt.discordantAllowed = std::stoi(argv[10]) != 0;
        opt.x = (uint32_t)std::stoul(argv[11]);
        clb::PairedEndMapper pe(strategy);
        clb::Counters counters;
        std::vector<clb::PairResult> res;
        const bool infer = argc > 12 && std::string(argv[12]) == "infer";
        if (infer) {
            clb::PEInference inf;
            pe.matchApproxPEInfer(pairs, std::string(argv[5]) == "best", (uint32_t)std::stoul(argv[6]), opt, counters, res, inf);
            std::cerr << "inferred " << inf.inferred << " orientation " << (inf.orientation == clb::FR ? "FR" : inf.orientation == clb::RF ? "RF" : "FF")
                      << " min " << inf.minInsertSize << " max " << inf.maxInsertSize << " unambiguous " << inf.unambiguousPairs << "\n";
        } else
            pe.matchApproxPEBatch(pairs, std::string(argv[5]) == "best", (uint32_t)std::stoul(argv[6]), opt, counters, res);
        for (size_t i = 0; i < pairs.size(); i++)
            for (const auto& s : pe.samRecords(pairs[i], res[i], index.getSeqNames()))
                std::cout << s.qname << '\t' << s.flag << '\t' << s.rname << '\t' << s.pos << '\t' << s.mapq << '\t' << s.cigar
                          << '\t' << s.rnext << '\t' << s.pnext << '\t' << s.tlen << '\t' << s.nm << '\n';
        std::cerr << "pairs " << counters.get(CLB_CTR_TOTAL_UNIQUE_PAIRS) << " mapped " << counters.get(CLB_CTR_MAPPED_PAIRS)
                  << " discordant " << counters.get(CLB_CTR_DISCORDANTLY_MAPPED_PAIRS) << " half "
                  << counters.get(CLB_CTR_MAPPED_HALF_PAIRS) << " unpaired " << counters.get(CLB_CTR_UNPAIRED_BUT_MAPPED_PAIRS)
                  << "\n";
    } catch (std::exception& e) {
        std::cerr << "error: " << e.what() << "\n";
        return 1;
    }
    return 0;
}
