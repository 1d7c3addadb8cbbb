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
//           [-F formatting threads per worker; default: the idle host threads shared among the workers]
#include <math.h>
#include <stdio.h>
#include <sched.h>
#include <sys/types.h>
#include <unistd.h>
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

// One batch of reads.  The reader only cuts the file into records (four lines each) and copies their bytes; names,
// qualities and the cleaned sequences are taken out of the copy by the worker that owns the batch (parse()), so
// that the part of the input path that all workers share is a memchr scan and a memcpy.
struct Batch {
    size_t id = 0;
    std::string raw;                  // the records' text
    std::vector<uint32_t> lineOff;    // start of every line in raw (4 per record) + end
    std::vector<uint32_t> nameLen;    // name = raw[lineOff[4 i] + 1, + nameLen[i])
    std::string seqs;                 // cleaned reads, concatenated
    std::vector<uint32_t> off;
    size_t size() const { return lineOff.empty() ? 0 : (lineOff.size() - 1) / 4; }
    const char* name(uint32_t i) const { return raw.data() + lineOff[4 * (size_t)i] + 1; }
    // quality line without its line end
    const char* qual(uint32_t i, uint32_t& len) const {
        const uint32_t b = lineOff[4 * (size_t)i + 3];
        uint32_t e = lineOff[4 * (size_t)i + 4];
        while (e > b && (raw[e - 1] == '\n' || raw[e - 1] == '\r'))
            e--;
        len = e - b;
        return raw.data() + b;
    }
    void parse() {
        const size_t n = size();
        nameLen.resize(n);
        off.resize(n + 1);
        off[0] = 0;
        seqs.resize(raw.size()); // upper bound; cut to size below
        char* w = &seqs[0];
        static const struct Clean { // upper case, everything but ACGT becomes N
            char t[256];
            Clean() {
                for (int c = 0; c < 256; c++)
                    t[c] = 'N';
                for (const char* p = "ACGTacgt"; *p; p++)
                    t[(unsigned char)*p] = (char)(*p & ~32);
            }
        } clean;
        for (size_t i = 0; i < n; i++) {
            // name: up to the first blank of the header line (Read::cleanUpRecord, src/reads.h:43-58)
            const uint32_t hb = lineOff[4 * i] + 1;
            uint32_t he = lineOff[4 * i + 1];
            while (he > hb && (raw[he - 1] == '\n' || raw[he - 1] == '\r'))
                he--;
            const void* sp = he > hb ? memchr(raw.data() + hb, ' ', he - hb) : nullptr;
            nameLen[i] = sp ? (uint32_t)((const char*)sp - (raw.data() + hb)) : (he > hb ? he - hb : 0u);
            const uint32_t sb = lineOff[4 * i + 1];
            uint32_t se = lineOff[4 * i + 2];
            while (se > sb && (raw[se - 1] == '\n' || raw[se - 1] == '\r'))
                se--;
            for (uint32_t p = sb; p < se; p++)
                *w++ = clean.t[(unsigned char)raw[p]];
            off[i + 1] = (uint32_t)(w - &seqs[0]);
        }
        seqs.resize((size_t)(w - &seqs[0]));
    }
};

class FastqReader {
  public:
    explicit FastqReader(const char* path) : f(fopen(path, "rb")) { buf.resize(1 << 24); }
    ~FastqReader() {
        if (f)
            fclose(f);
    }
    bool ok() const { return f != nullptr; }
    // next batch of at most n records; false at the end of the file (a trailing partial record is dropped, as the
    // four-lines-or-nothing loop of the line reader did)
    bool next(Batch& b, size_t n) {
        std::lock_guard<std::mutex> lock(m);
        b.id = nextId;
        b.raw.clear();
        b.raw.reserve(rawHint);
        b.lineOff.clear();
        b.lineOff.reserve(4 * n + 2);
        b.lineOff.push_back(0);
        const size_t want = 4 * n;
        while (b.lineOff.size() - 1 < want) {
            if (pos == len) {
                len = f ? fread(&buf[0], 1, buf.size(), f) : 0;
                pos = 0;
                if (len == 0) {
                    if (b.raw.size() > b.lineOff.back()) // last line without a line end
                        b.lineOff.push_back((uint32_t)b.raw.size());
                    break;
                }
            }
            // the lines of this buffer that the batch still wants: one scan, one copy
            const size_t base = b.raw.size(), from = pos;
            size_t at = pos;
            while (b.lineOff.size() - 1 < want) {
                const char* p = (const char*)memchr(&buf[at], '\n', len - at);
                if (!p) {
                    at = len;
                    break;
                }
                at = (size_t)(p - &buf[0]) + 1;
                b.lineOff.push_back((uint32_t)(base + (at - from)));
            }
            b.raw.append(&buf[from], at - from);
            pos = at;
        }
        const size_t lines = b.lineOff.size() - 1;
        if (lines % 4) { // incomplete record at the end of the file
            b.lineOff.resize(lines - lines % 4 + 1);
            b.raw.resize(b.lineOff.back());
        }
        if (b.size() == 0)
            return false;
        rawHint = std::max(rawHint, b.raw.size() + b.raw.size() / 16);
        nextId++;
        return true;
    }

  private:
    FILE* f;
    std::string buf;
    size_t pos = 0, len = 0, nextId = 0, rawHint = 0;
    std::mutex m;
};

// text of a run of reads: a buffer that keeps its memory from batch to batch and is never value-initialised
struct TextBuf {
    char* p = nullptr;
    size_t cap = 0, len = 0;
    TextBuf() = default;
    TextBuf(const TextBuf&) = delete;
    TextBuf& operator=(const TextBuf&) = delete;
    TextBuf(TextBuf&& o) noexcept : p(o.p), cap(o.cap), len(o.len) { o.p = nullptr; o.cap = o.len = 0; }
    ~TextBuf() { free(p); }
    void need(size_t n) {
        if (n > cap) {
            free(p);
            cap = n + n / 8;
            p = (char*)malloc(cap);
            if (!p) {
                fprintf(stderr, "error: out of memory (%zu bytes of SAM text)\n", cap);
                abort();
            }
        }
    }
    size_t size() const { return len; }
    const char* data() const { return p; }
};

// The records are written through a cursor into a buffer of a size known beforehand (std::string::push_back checks
// the capacity for every character: 340 ns per record, a quarter of a GB/s per thread).
static inline char* putU(char* w, uint32_t v) {
    static const char pairs[] = "00010203040506070809101112131415161718192021222324252627282930313233343536373839"
                                "40414243444546474849505152535455565758596061626364656667686970717273747576777879"
                                "8081828384858687888990919293949596979899";
    if (v < 10) {
        *w++ = (char)('0' + v);
        return w;
    }
    char t[12];
    int n = 12;
    while (v >= 100) {
        const uint32_t q = v / 100, d = v - q * 100;
        t[--n] = pairs[2 * d + 1];
        t[--n] = pairs[2 * d];
        v = q;
    }
    if (v >= 10) {
        t[--n] = pairs[2 * v + 1];
        t[--n] = pairs[2 * v];
    } else {
        t[--n] = (char)('0' + v);
    }
    memcpy(w, t + n, (size_t)(12 - n));
    return w + (12 - n);
}
static inline char* putS(char* w, const char* p, size_t n) {
    memcpy(w, p, n);
    return w + n;
}
#define PUT_LIT(w, lit) putS(w, lit, sizeof(lit) - 1)

// SAM records of the reads [first, last) of one batch
static void formatReads(const Batch& b, const clb_result& r, const std::vector<std::string>& seqNames, uint32_t first,
                        uint32_t last, TextBuf& out) {
    out.len = 0;
    if (first >= last)
        return;
    size_t maxSeqName = 1;
    for (const std::string& n : seqNames)
        maxSeqName = std::max(maxSeqName, n.size());
    // upper bound of the text: per record name + sequence name + numbers and fixed text (<= 80) + CIGAR runs
    // (<= 6 characters each); per read once more its sequence and quality
    size_t bound = 0;
    for (uint32_t i = first; i < last; i++) {
        const size_t len = b.off[i + 1] - b.off[i];
        const size_t rec = b.nameLen[i] + maxSeqName + 80 + (size_t)r.cigar_stride * 6;
        bound += (size_t)std::max<uint32_t>(1u, r.read_count[i]) * rec + 2 * len + 2 +
                 (b.lineOff[4 * (size_t)i + 4] - b.lineOff[4 * (size_t)i + 3]);
    }
    out.need(bound);
    char* const begin = out.p;
    char* w = begin;
    std::vector<uint32_t> idx;
    for (uint32_t i = first; i < last; i++) {
        const char* seq = b.seqs.data() + b.off[i];
        const uint32_t len = b.off[i + 1] - b.off[i];
        uint32_t qlen = 0;
        const char* qual = b.qual(i, qlen);
        const char* name = b.name(i);
        const uint32_t nlen = b.nameLen[i];
        const uint32_t s0 = r.read_start[i], cnt = r.read_count[i];
        if (cnt == 0) { // TextOcc::createUnmappedSAMOccurrenceSE
            w = putS(w, name, nlen);
            w = PUT_LIT(w, "\t4\t*\t0\t0\t*\t*\t0\t0\t");
            w = putS(w, seq, len);
            *w++ = '\t';
            if (qlen)
                w = putS(w, qual, qlen);
            else
                *w++ = '*';
            w = PUT_LIT(w, "\tPG:Z:Columba\n");
            continue;
        }
        // stable order on the distance (generateOutputSingleEnd): the distances are a handful of small values, one
        // pass per value keeps the text order within a value
        idx.resize(cnt);
        uint32_t minScore = 255, maxScore = 0;
        for (uint32_t j = 0; j < cnt; j++) {
            const uint32_t d = r.occ[s0 + j].dist;
            minScore = std::min(minScore, d);
            maxScore = std::max(maxScore, d);
        }
        uint32_t filled = 0, nHits = 0;
        for (uint32_t d = minScore; d <= maxScore && filled < cnt; d++) {
            for (uint32_t j = 0; j < cnt; j++)
                if (r.occ[s0 + j].dist == d)
                    idx[filled++] = s0 + j;
            if (d == minScore)
                nHits = filled;
        }
        const int mapqBest = nHits == 1 ? 60 : (int)lround(-10.0 * log10(1.0 - 1.0 / nHits));
        for (uint32_t j = 0; j < cnt; j++) {
            const clb_occ& o = r.occ[idx[j]];
            w = putS(w, name, nlen);
            *w++ = '\t';
            w = putU(w, (o.strand ? 16u : 0u) | (j ? 256u : 0u));
            *w++ = '\t';
            w = putS(w, seqNames[o.seq_id].data(), seqNames[o.seq_id].size());
            *w++ = '\t';
            w = putU(w, o.pos + 1);
            *w++ = '\t';
            w = putU(w, o.dist == minScore ? (uint32_t)mapqBest : 0u);
            *w++ = '\t';
            if (o.n_cigar == 0)
                *w++ = '*';
            for (uint32_t c = 0; c < o.n_cigar; c++) {
                const uint16_t run = r.cigar[(size_t)idx[j] * r.cigar_stride + c];
                w = putU(w, run >> 2);
                *w++ = "MID"[run & 3];
            }
            w = PUT_LIT(w, "\t*\t0\t0\t");
            if (j == 0) { // sequence and quality on the primary record only
                if (o.strand) {
                    for (uint32_t q = len; q-- > 0;) {
                        const char c = seq[q];
                        *w++ = c == 'A' ? 'T' : c == 'C' ? 'G' : c == 'G' ? 'C' : c == 'T' ? 'A' : c;
                    }
                    *w++ = '\t';
                    if (qlen)
                        for (uint32_t q = qlen; q-- > 0;)
                            *w++ = qual[q];
                    else
                        *w++ = '*';
                } else {
                    w = putS(w, seq, len);
                    *w++ = '\t';
                    if (qlen)
                        w = putS(w, qual, qlen);
                    else
                        *w++ = '*';
                }
            } else {
                w = PUT_LIT(w, "*\t*");
            }
            w = PUT_LIT(w, "\tAS:i:");
            w = putU(w, o.dist);
            w = PUT_LIT(w, "\tNM:i:");
            w = putU(w, o.dist);
            w = PUT_LIT(w, "\tPG:Z:Columba\n");
        }
    }
    if ((size_t)(w - begin) > bound) { // cannot happen; never write past the buffer silently
        fprintf(stderr, "internal error: SAM text bound exceeded\n");
        abort();
    }
    out.len = (size_t)(w - begin);
}

// SAM text of one batch in `parts` (in read order): the reads are cut into nFmt runs of about equal output size
// and formatted by nFmt threads -- 37 records per read on a repeat-rich genome make the text 3 GB per 1 M reads,
// and formatting it is what bounds a worker once the mapping itself takes a third of a second
static void formatBatch(const Batch& b, const clb_result& r, const std::vector<std::string>& seqNames, int nFmt,
                        std::vector<TextBuf>& parts) {
    const uint32_t n = r.n_reads;
    nFmt = (int)std::max<uint32_t>(1u, std::min<uint32_t>((uint32_t)nFmt, n / 64 + 1));
    if (parts.size() < (size_t)nFmt)
        parts.resize((size_t)nFmt);
    for (TextBuf& t : parts)
        t.len = 0;
    std::vector<uint32_t> cut((size_t)nFmt + 1, n);
    cut[0] = 0;
    if (nFmt > 1) {
        // weight of a read = its records + its own line
        const double total = (double)r.n_occ + 2.0 * n;
        double run = 0;
        int c = 1;
        for (uint32_t i = 0; i < n && c < nFmt; i++) {
            run += (double)r.read_count[i] + 2.0;
            if (run >= total * c / nFmt)
                cut[(size_t)c++] = i + 1;
        }
    }
    std::vector<std::thread> ths;
    for (int t = 1; t < nFmt; t++)
        ths.emplace_back([&, t] { formatReads(b, r, seqNames, cut[(size_t)t], cut[(size_t)t + 1], parts[(size_t)t]); });
    formatReads(b, r, seqNames, cut[0], cut[1], parts[0]);
    for (auto& t : ths)
        t.join();
}

int main(int argc, char** argv) {
    std::string base, fastq, outPath, scheme = "columba", schemesDir;
    int k = 4, threads = 2, device = 0, fmtThreads = 0;
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
        else if (a == "-F") fmtThreads = atoi(v.c_str());
    }
    if (fmtThreads <= 0) { // the host threads that the workers leave idle, shared among them
        cpu_set_t set;
        int hw = (int)std::thread::hardware_concurrency();
        if (sched_getaffinity(0, sizeof(set), &set) == 0) // the CPUs this process may run on (numactl, taskset)
            hw = CPU_COUNT(&set);
        fmtThreads = std::max(1, std::min(16, hw / threads));
    }
    if (base.empty() || fastq.empty() || outPath.empty()) {
        fprintf(stderr, "usage: clb_map -r <index base> -f <reads.fq> -o <out.sam> [-e 4] [-t 2] [-b 262144] [-S columba] [-d 0] "
                        "[-D schemes dir] [-F formatting threads per worker]\n");
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
    fflush(out);
    const int outFd = fileno(out);
    uint64_t fileOff = (uint64_t)ftell(out); // the records follow the header
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
        std::vector<TextBuf> text;
        while (!failed && reader.next(b, batchReads)) {
            b.parse();
            clb_result r;
            if (clb_map_batch(ctx, (const uint8_t*)b.seqs.data(), b.off.data(), (uint32_t)b.size(), k, 0, &r)) {
                fprintf(stderr, "error: %s\n", clb_last_error(ctx));
                failed = true;
                wcv.notify_all();
                break;
            }
            formatBatch(b, r, seqNames, fmtThreads, text);
            // the batches take their place in the file in input order; the bytes are then written by all workers
            // at once (3 GB of text per 1 M reads: one writer would be the slowest stage of the pipeline)
            size_t bytes = 0;
            for (const TextBuf& part : text)
                bytes += part.size();
            std::unique_lock<std::mutex> lock(wm);
            wcv.wait(lock, [&] { return nextWrite == b.id || failed; });
            uint64_t at = fileOff;
            fileOff += bytes;
            for (int c = 0; c < CLB_N_COUNTERS; c++)
                counters[c] += r.counters[c];
            totalReads += r.n_reads;
            totalOcc += r.n_occ;
            nextWrite++;
            lock.unlock();
            wcv.notify_all();
            std::vector<std::thread> wr;
            bool bad = false;
            for (size_t q = 0; q < text.size(); q++) {
                const uint64_t pAt = at;
                at += text[q].size();
                auto put = [&, q, pAt] {
                    const char* p = text[q].data();
                    size_t left = text[q].size();
                    uint64_t o = pAt;
                    while (left) {
                        const ssize_t w = pwrite(outFd, p, left, (off_t)o);
                        if (w <= 0) {
                            bad = true;
                            return;
                        }
                        p += w;
                        left -= (size_t)w;
                        o += (uint64_t)w;
                    }
                };
                if (q + 1 < text.size())
                    wr.emplace_back(put);
                else
                    put();
            }
            for (auto& t : wr)
                t.join();
            if (bad) {
                fprintf(stderr, "error: writing %s failed\n", outPath.c_str());
                failed = true;
                wcv.notify_all();
                break;
            }
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
                    "reads/s %.0f formatting threads per mapping thread %d\n",
            load, threads, map, (unsigned long long)totalReads, (unsigned long long)totalOcc,
            (unsigned long long)counters[CLB_CTR_MAPPED_READS], totalReads / map, fmtThreads);
    clb_index_free(ix);
    return failed ? 1 : 0;
}
