// columba_b200 -- batch driver: FASTQ in, SAM out, single-end ALL mode, on top of
// the C ABI.  What the reference does with a reader thread, worker threads and a writer thread around
// SearchStrategy::matchApprox (src/parallel.cpp:67-118, src/fastq.cpp:283-424,512-733), for a GPU: every worker
// thread owns one context (stream + scratch), takes the next batch of reads from the shared reader (64 K .. 1 M
// reads instead of the reference's 64), maps it with one clb_map_batch call, formats the SAM records of its batch
// and hands the text to the writer in input order.  While one context's batch is on the GPU the other threads
// parse and format, so the copies, the kernels and the host work overlap.
//
// Records follow generateOutputSingleEnd (src/searchstrategy.cpp:1824-1902, stable sort on the distance, the first
// occurrence is primary), generateSE_SAM (src/searchstrategy.h:1633-1641) and TextOcc::generateSAMSingleEnd
// (src/indexhelpers.cpp:56-87); reads are cleaned like Read::cleanUpRecord (src/reads.h:43-58).
//
//   clb_map -r <index base> -f <reads.fq> -o <out.sam> [-e 4] [-t 2] [-b 262144] [-S columba] [-d 0] [-D schemes dir]
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../../include/columba_b200.h"

struct Batch {
    size_t id = 0;
    std::vector<std::string> names, quals;
    std::string seqs; // cleaned reads, concatenated
    std::vector<uint32_t> off;
};

class FastqReader {
  public:
    explicit FastqReader(const char* path) : f(fopen(path, "rb")) { buf.resize(1 << 24); }
    ~FastqReader() {
        if (f)
            fclose(f);
    }
    bool ok() const { return f != nullptr; }
    // next batch of at most n reads; false at the end of the file
    bool next(Batch& b, size_t n) {
        std::lock_guard<std::mutex> lock(m);
        b = Batch();
        b.id = nextId;
        b.off.push_back(0);
        std::string l1, l2, l3, l4;
        while (b.names.size() < n && line(l1) && line(l2) && line(l3) && line(l4)) {
            const size_t sp = l1.find(' ');
            b.names.push_back(l1.substr(1, sp == std::string::npos ? std::string::npos : sp - 1));
            for (char& c : l2) { // uppercase, everything but ACGT becomes N
                c = (char)(c >= 'a' && c <= 'z' ? c - 32 : c);
                if (c != 'A' && c != 'C' && c != 'G' && c != 'T')
                    c = 'N';
            }
            b.seqs += l2;
            b.off.push_back((uint32_t)b.seqs.size());
            b.quals.push_back(l4);
        }
        if (b.names.empty())
            return false;
        nextId++;
        return true;
    }

  private:
    bool line(std::string& out) {
        out.clear();
        for (;;) {
            if (pos == len) {
                len = f ? fread(&buf[0], 1, buf.size(), f) : 0;
                pos = 0;
                if (len == 0)
                    return !out.empty();
            }
            const char* p = (const char*)memchr(&buf[pos], '\n', len - pos);
            if (p) {
                out.append(&buf[pos], p - &buf[pos]);
                pos = (size_t)(p - &buf[0]) + 1;
                if (!out.empty() && out.back() == '\r')
                    out.pop_back();
                return true;
            }
            out.append(&buf[pos], len - pos);
            pos = len;
        }
    }
    FILE* f;
    std::string buf;
    size_t pos = 0, len = 0, nextId = 0;
    std::mutex m;
};

static inline void putU(std::string& s, uint32_t v) {
    char t[12];
    int n = 0;
    do {
        t[n++] = (char)('0' + v % 10);
        v /= 10;
    } while (v);
    while (n)
        s.push_back(t[--n]);
}

// SAM records of one batch
static void formatBatch(const Batch& b, const clb_result& r, const std::vector<std::string>& seqNames, std::string& out) {
    out.clear();
    out.reserve((size_t)r.n_occ * 96 + b.seqs.size() * 2 + b.names.size() * 64);
    std::vector<uint32_t> idx;
    std::string rc, rq;
    for (uint32_t i = 0; i < r.n_reads; i++) {
        const char* seq = b.seqs.data() + b.off[i];
        const uint32_t len = b.off[i + 1] - b.off[i];
        const std::string& qual = b.quals[i];
        const uint32_t s0 = r.read_start[i], cnt = r.read_count[i];
        if (cnt == 0) { // TextOcc::createUnmappedSAMOccurrenceSE
            out += b.names[i];
            out += "\t4\t*\t0\t0\t*\t*\t0\t0\t";
            out.append(seq, len);
            out.push_back('\t');
            out += qual.empty() ? "*" : qual;
            out += "\tPG:Z:Columba\n";
            continue;
        }
        idx.resize(cnt);
        for (uint32_t j = 0; j < cnt; j++)
            idx[j] = s0 + j;
        std::stable_sort(idx.begin(), idx.end(), [&](uint32_t x, uint32_t y) { return r.occ[x].dist < r.occ[y].dist; });
        const uint32_t minScore = r.occ[idx[0]].dist;
        uint32_t nHits = 0;
        while (nHits < cnt && r.occ[idx[nHits]].dist == minScore)
            nHits++;
        const int mapqBest = nHits == 1 ? 60 : (int)lround(-10.0 * log10(1.0 - 1.0 / nHits));
        for (uint32_t j = 0; j < cnt; j++) {
            const clb_occ& o = r.occ[idx[j]];
            out += b.names[i];
            out.push_back('\t');
            putU(out, (o.strand ? 16u : 0u) | (j ? 256u : 0u));
            out.push_back('\t');
            out += seqNames[o.seq_id];
            out.push_back('\t');
            putU(out, o.pos + 1);
            out.push_back('\t');
            putU(out, o.dist == minScore ? (uint32_t)mapqBest : 0u);
            out.push_back('\t');
            if (o.n_cigar == 0)
                out.push_back('*');
            for (uint32_t c = 0; c < o.n_cigar; c++) {
                const uint16_t run = r.cigar[(size_t)idx[j] * r.cigar_stride + c];
                putU(out, run >> 2);
                out.push_back("MID"[run & 3]);
            }
            out += "\t*\t0\t0\t";
            if (j == 0) { // sequence and quality on the primary record only
                if (o.strand) {
                    rc.assign(seq, len);
                    std::reverse(rc.begin(), rc.end());
                    for (char& c : rc)
                        c = c == 'A' ? 'T' : c == 'C' ? 'G' : c == 'G' ? 'C' : c == 'T' ? 'A' : c;
                    out += rc;
                    out.push_back('\t');
                    rq = qual;
                    std::reverse(rq.begin(), rq.end());
                    out += rq.empty() ? "*" : rq;
                } else {
                    out.append(seq, len);
                    out.push_back('\t');
                    out += qual.empty() ? "*" : qual;
                }
            } else {
                out += "*\t*";
            }
            out += "\tAS:i:";
            putU(out, o.dist);
            out += "\tNM:i:";
            putU(out, o.dist);
            out += "\tPG:Z:Columba\n";
        }
    }
}

int main(int argc, char** argv) {
    std::string base, fastq, outPath, scheme = "columba", schemesDir;
    int k = 4, threads = 2, device = 0;
    size_t batchReads = 1 << 18;
    for (int i = 1; i + 1 < argc; i += 2) {
        const std::string a = argv[i], v = argv[i + 1];
        if (a == "-r") base = v;
        else if (a == "-f") fastq = v;
        else if (a == "-o") outPath = v;
        else if (a == "-e") k = atoi(v.c_str());
        else if (a == "-t") threads = std::max(1, atoi(v.c_str()));
        else if (a == "-b") batchReads = (size_t)atol(v.c_str());
        else if (a == "-S") scheme = v;
        else if (a == "-d") device = atoi(v.c_str());
        else if (a == "-D") schemesDir = v;
    }
    if (base.empty() || fastq.empty() || outPath.empty()) {
        fprintf(stderr, "usage: clb_map -r <index base> -f <reads.fq> -o <out.sam> [-e 4] [-t 2] [-b 262144] [-S columba] [-d 0] "
                        "[-D schemes dir]\n");
        return 2;
    }
    const auto t0 = std::chrono::steady_clock::now();
    clb_index* ix = nullptr;
    if (clb_index_load(base.c_str(), 4, 10, 4, 0, device, &ix)) {
        fprintf(stderr, "error: %s\n", clb_last_error(nullptr));
        return 1;
    }
    std::vector<std::string> seqNames;
    for (uint32_t i = 0; i < clb_index_num_seqs(ix); i++)
        seqNames.push_back(clb_index_seq_name(ix, i));
    FastqReader reader(fastq.c_str());
    FILE* out = fopen(outPath.c_str(), "wb");
    if (!reader.ok() || !out) {
        fprintf(stderr, "error: cannot open %s\n", reader.ok() ? outPath.c_str() : fastq.c_str());
        return 1;
    }
    fprintf(out, "@HD\tVN:1.6\tSO:queryname\n@PG\tID:columba_b200\tPN:columba_b200\n");
    for (uint32_t i = 0; i < seqNames.size(); i++)
        fprintf(out, "@SQ\tSN:%s\tLN:%u\n", seqNames[i].c_str(), clb_index_seq_start(ix, i + 1) - clb_index_seq_start(ix, i));
    const auto t1 = std::chrono::steady_clock::now();
    std::mutex wm;
    std::condition_variable wcv;
    size_t nextWrite = 0;
    uint64_t counters[CLB_N_COUNTERS] = {};
    uint64_t totalReads = 0, totalOcc = 0;
    bool failed = false;
    auto worker = [&]() {
        clb_ctx* ctx = nullptr;
        if (clb_ctx_create(ix, &ctx) || clb_ctx_set_strategy(ctx, "named", scheme.c_str(), schemesDir.c_str()) ||
            clb_ctx_set_options(ctx, CLB_METRIC_EDIT, CLB_PART_DYNAMIC)) {
            fprintf(stderr, "error: %s\n", clb_last_error(ctx));
            failed = true;
            return;
        }
        Batch b;
        std::string text;
        while (!failed && reader.next(b, batchReads)) {
            clb_result r;
            if (clb_map_batch(ctx, (const uint8_t*)b.seqs.data(), b.off.data(), (uint32_t)b.names.size(), k, 0, &r)) {
                fprintf(stderr, "error: %s\n", clb_last_error(ctx));
                failed = true;
                wcv.notify_all();
                break;
            }
            formatBatch(b, r, seqNames, text);
            std::unique_lock<std::mutex> lock(wm);
            wcv.wait(lock, [&] { return nextWrite == b.id || failed; });
            fwrite(text.data(), 1, text.size(), out);
            for (int c = 0; c < CLB_N_COUNTERS; c++)
                counters[c] += r.counters[c];
            totalReads += r.n_reads;
            totalOcc += r.n_occ;
            nextWrite++;
            lock.unlock();
            wcv.notify_all();
        }
        clb_ctx_free(ctx);
    };
    std::vector<std::thread> ths;
    for (int t = 0; t < threads; t++)
        ths.emplace_back(worker);
    for (auto& t : ths)
        t.join();
    fclose(out);
    const auto t2 = std::chrono::steady_clock::now();
    const double load = std::chrono::duration<double>(t1 - t0).count(), map = std::chrono::duration<double>(t2 - t1).count();
    fprintf(stderr, "index loaded in %.3fs\n%d mapping threads finished mapping in %.6fs\nreads %llu occurrences %llu mapped %llu "
                    "reads/s %.0f\n",
            load, threads, map, (unsigned long long)totalReads, (unsigned long long)totalOcc,
            (unsigned long long)counters[CLB_CTR_MAPPED_READS], totalReads / map);
    clb_index_free(ix);
    return failed ? 1 : 0;
}
