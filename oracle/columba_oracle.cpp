// ORACLE (test infrastructure, NOT product code).
// Driver + C entry points around the CPU restatement in orc_*.h.  Builds into
// oracle/_build/liboracle.so (ctypes from tests/) and, with -DORC_MAIN, into
// the command line tool oracle/_build/columba_oracle.
//
// Parity status: pinned against the unmodified reference binary built by
// `make -C oracle ref` (oracle/_ref/columba_dev) -- see tests/golden/ and
// tools/make_golden.py.
#include "orc_match.h"
#include <chrono>
#include <iostream>

using namespace orc;

struct OracleCtx {
    Index ix;
    Strategy st;
    std::string err;
};

static void readFastq(const std::string& fn, std::vector<std::string>& names,
                      std::vector<std::string>& seqs) {
    std::ifstream f(fn);
    if (!f)
        throw std::runtime_error("cannot open " + fn);
    std::string l1, l2, l3, l4;
    while (std::getline(f, l1)) {
        if (l1.empty())
            continue;
        bool fq = l1[0] == '@';
        std::getline(f, l2);
        if (fq) {
            std::getline(f, l3);
            std::getline(f, l4);
        }
        std::string nm = l1.substr(1);
        size_t sp = nm.find_first_of(" \t");
        if (sp != std::string::npos)
            nm = nm.substr(0, sp);
        for (auto& c : l2)
            c = (char)toupper(c);
        names.push_back(nm);
        seqs.push_back(l2);
    }
}

extern "C" {

void* orc_open(const char* base, int saSparse, int kmer, int inTextSwitch, const char* schemeKind,
               const char* schemeArg, const char* dataDir, int partition, int metric) {
    OracleCtx* c = new OracleCtx();
    try {
        std::string kind = schemeKind;
        if (kind == "named")
            c->st = Strategy::named(schemeArg, dataDir);
        else if (kind == "custom")
            c->st = Strategy::custom(schemeArg, 50);
        else if (kind == "custom_dynamic")
            c->st = Strategy::dynamicFrom(Strategy::custom(schemeArg, 50));
        else if (kind == "multi")
            c->st = Strategy::multi(schemeArg);
        else
            throw std::runtime_error("unknown scheme kind " + kind);
        c->st.part = (PartitionStrategy)partition;
        c->st.metric = (Metric)metric;
        c->ix.load(base, saSparse, kmer, inTextSwitch);
    } catch (std::exception& e) {
        fprintf(stderr, "oracle: %s\n", e.what());
        delete c;
        return nullptr;
    }
    return c;
}
void orc_close(void* h) { delete (OracleCtx*)h; }

// rank parity helpers: occ of all 4 nucleotides at position p of one BWT
void orc_occ4(void* h, int reverse, uint32_t p, uint32_t* out4) {
    OracleCtx* c = (OracleCtx*)h;
    const BwtRep& b = reverse ? c->ix.rev : c->ix.fwd;
    for (int ci = 1; ci <= 4; ci++)
        out4[ci - 1] = (uint32_t)b.occ(ci, p);
}
// child ranges of one parent for all 4 characters; mode 0=fwd 1=bwd 2=uni-bwd
// out: 4 x (sa.b, sa.e, rev.b, rev.e)
void orc_extend(void* h, int mode, const uint32_t* parent4, uint32_t* out16) {
    OracleCtx* c = (OracleCtx*)h;
    c->ix.setDirection(mode == 0 ? FWD : BWD, mode == 2);
    RangePair p(Range(parent4[0], parent4[1]), Range(parent4[2], parent4[3]));
    for (int ci = 1; ci <= 4; ci++) {
        RangePair ch;
        c->ix.extraChar(ci, p, ch);
        out16[(ci - 1) * 4 + 0] = ch.sa.b;
        out16[(ci - 1) * 4 + 1] = ch.sa.e;
        out16[(ci - 1) * 4 + 2] = ch.rev.b;
        out16[(ci - 1) * 4 + 3] = ch.rev.e;
    }
}
uint32_t orc_find_sa(void* h, uint32_t row) { return ((OracleCtx*)h)->ix.findSA(row); }
uint32_t orc_text_length(void* h) { return ((OracleCtx*)h)->ix.textLength; }
const char* orc_seq_name(void* h, uint32_t i) { return ((OracleCtx*)h)->ix.seqNames[i].c_str(); }

// Map one read (ALL mode).  Output records are appended to a malloc'ed text
// buffer "strand\tseqID\tpos1\tdist\tcigar\tindexBegin\n"; counters[17] are
// accumulated.  Returns number of records or -1.
int orc_map_read(void* h, const char* read, int k, uint64_t* counters, char** outText) {
    OracleCtx* c = (OracleCtx*)h;
    try {
        Mapper mp(c->ix, c->st);
        Counters ct;
        std::string rd(read);
        auto occs = mp.matchApproxAll(rd, k, ct);
        std::string out;
        for (auto& t : occs) {
            out += std::to_string((int)t.strand) + "\t" + std::to_string(t.seqID) + "\t" +
                   std::to_string(t.b + 1) + "\t" + std::to_string(t.dist) + "\t" + t.cigar + "\t" +
                   std::to_string(t.indexBegin) + "\n";
        }
        for (int i = 0; i < N_COUNTERS; i++)
            counters[i] += ct.c[i];
        *outText = (char*)malloc(out.size() + 1);
        memcpy(*outText, out.c_str(), out.size() + 1);
        return (int)occs.size();
    } catch (std::exception& e) {
        c->err = e.what();
        fprintf(stderr, "oracle: %s\n", e.what());
        return -1;
    }
}
void orc_free(char* p) { free(p); }
}

#ifdef ORC_MAIN
int main(int argc, char** argv) {
    std::string base, reads, out = "-", kind = "named", arg = "columba", dataDir = "columba_b200/schemes";
    int k = 4, sparse = 4, kmer = 10, sw = 4, part = DYNAMIC, metric = EDIT;
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        auto nx = [&]() { return std::string(argv[++i]); };
        if (a == "-r") base = nx();
        else if (a == "-f") reads = nx();
        else if (a == "-o") out = nx();
        else if (a == "-e") k = std::stoi(nx());
        else if (a == "-s") sparse = std::stoi(nx());
        else if (a == "-K") kmer = std::stoi(nx());
        else if (a == "-i") sw = std::stoi(nx());
        else if (a == "-S") { kind = "named"; arg = nx(); }
        else if (a == "-c") { kind = "custom"; arg = nx(); }
        else if (a == "-cd") { kind = "custom_dynamic"; arg = nx(); }
        else if (a == "-d") { kind = "multi"; arg = nx(); }
        else if (a == "-D") dataDir = nx();
        else if (a == "-p") { std::string p = nx(); part = p == "uniform" ? UNIFORM : p == "static" ? STATIC : DYNAMIC; }
        else if (a == "-m") { std::string m = nx(); metric = m == "hamming" ? HAMMING : EDIT; }
        else { fprintf(stderr, "unknown arg %s\n", a.c_str()); return 2; }
    }
    OracleCtx* c = (OracleCtx*)orc_open(base.c_str(), sparse, kmer, sw, kind.c_str(), arg.c_str(),
                                        dataDir.c_str(), part, metric);
    if (!c)
        return 1;
    std::vector<std::string> names, seqs;
    readFastq(reads, names, seqs);
    FILE* fo = out == "-" ? stdout : fopen(out.c_str(), "w");
    uint64_t counters[N_COUNTERS] = {0};
    auto t0 = std::chrono::steady_clock::now();
    size_t mapped = 0, total = 0;
    for (size_t r = 0; r < seqs.size(); r++) {
        char* txt = nullptr;
        int n = orc_map_read(c, seqs[r].c_str(), k, counters, &txt);
        if (n < 0)
            return 1;
        mapped += n > 0;
        total += n;
        // one line per record: name strand seqname pos dist cigar
        const char* p = txt;
        while (*p) {
            const char* e = strchr(p, '\n');
            std::string line(p, e);
            std::stringstream ss(line);
            int strand;
            uint32_t sid, pos, dist, ib;
            std::string cg;
            ss >> strand >> sid >> pos >> dist >> cg >> ib;
            fprintf(fo, "%s\t%d\t%s\t%u\t%u\t%s\n", names[r].c_str(), strand ? 16 : 0,
                    c->ix.seqNames[sid].c_str(), pos, dist, cg.c_str());
            p = e + 1;
        }
        orc_free(txt);
    }
    double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    fprintf(stderr, "oracle: %zu reads in %.3fs (%.0f reads/s)\n", seqs.size(), secs, seqs.size() / secs);
    fprintf(stderr, "nodes %lu reported %lu intext_started %lu aborted %lu immediate %lu searches %lu "
                    "cigars_index %lu cigars_text %lu matches %zu mapped %zu\n",
            counters[NODE_COUNTER], counters[TOTAL_REPORTED_POSITIONS], counters[IN_TEXT_STARTED],
            counters[ABORTED_IN_TEXT_VERIF], counters[IMMEDIATE_SWITCH], counters[SEARCH_STARTED],
            counters[CIGARS_IN_INDEX], counters[CIGARS_IN_TEXT_VERIFICATION], total, mapped);
    if (fo != stdout)
        fclose(fo);
    orc_close(c);
    return 0;
}
#endif
