// columba_b200 -- loads the reference's Vanilla index files (32-bit length_t)
// and re-packs them into the HBM layout of clb_index.h.
//
// Input contract = the file formats written by the reference's columba_build:
//   .meta  src/buildindex.cpp:688-698 (read: src/indexinterface.cpp:77-128)
//   .cct   src/buildindex.cpp:432-437 (read: src/indexinterface.cpp:137-151)
//   .txt.bin  src/buildindex.cpp:591-605 (read: src/fmindex/fmindex.cpp:77-94)
//   .brt / .rev.brt  src/fmindex/bwtrepr.h:113-135 + src/bitvec.h:378-394
//   .sa.bv.<s> / .sa.<s>  src/fmindex/suffixArray.h:229-243, src/bitvec.h:176-195
//   .pos / .sna  src/buildindex.cpp:361-381 (read: src/indexinterface.cpp:161-191)
// Error messages follow the reference's (std::runtime_error texts).
#pragma once
#include "clb_gidx.h"
#include "clb_kernels.h"
#include "clb_pipeline.h"
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <cstdlib>
#include <fstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace clb {

struct HostIndexMeta {
    std::vector<u32> seqStart;
    std::vector<std::string> seqNames;
    u32 textLength = 0;
    u32 C[5];
    u32 K = 0;
    u32 saSparse = 0;
    size_t deviceBytes = 0;
    bool fromGidx = false;
};

template <typename T> static bool readWholeFile(const std::string& fn, std::vector<T>& v) {
    std::ifstream f(fn, std::ios::binary | std::ios::ate);
    if (!f)
        return false;
    size_t sz = (size_t)f.tellg();
    if (sz % sizeof(T))
        return false;
    v.resize(sz / sizeof(T));
    f.seekg(0);
    f.read((char*)v.data(), (std::streamsize)sz);
    return (bool)f;
}

// one direction's BWT: reference planes are cumulative (bwtrepr.h:56-72):
// plane j set iff BWT[p] in {A..char j+1}; '$' sets none.
static void convertBwt(const std::string& fn, u32 textLength, std::vector<BwtRec>& rec, u32& dollarPos) {
    std::ifstream f(fn, std::ios::binary);
    if (!f)
        throw std::runtime_error("Cannot open file: " + fn);
    u64 dp = 0, N = 0;
    f.read((char*)&dp, 8);
    f.read((char*)&N, 8);
    if (N != (u64)textLength + 1)
        throw std::runtime_error("Corrupt BWT bitvector file (size mismatch): " + fn);
    const size_t nBlocks = (size_t)((N + 63) / 64);
    std::vector<u64> bv(nBlocks * 4);
    f.read((char*)bv.data(), (std::streamsize)(bv.size() * 8));
    if (!f)
        throw std::runtime_error("Corrupt BWT bitvector file (truncated): " + fn);
    dollarPos = (u32)dp;
    const size_t nRec = (size_t)textLength / 64 + 1;
    rec.resize(nRec);
    u32 cnt[4] = {0, 0, 0, 0};
    for (size_t q = 0; q < nRec; q++) {
        u64 w0 = 0, w1 = 0, w2 = 0, w3 = 0;
        if (q < nBlocks) {
            w0 = bv[q * 4 + 0];
            w1 = bv[q * 4 + 1];
            w2 = bv[q * 4 + 2];
            w3 = bv[q * 4 + 3];
        }
        BwtRec r;
        for (int c = 0; c < 4; c++)
            r.cnt[c] = cnt[c];
        r.lo = (w1 & ~w0) | (w3 & ~w2);
        r.hi = w3 & ~w1;
        rec[q] = r;
        cnt[0] += (u32)__builtin_popcountll(w0);
        cnt[1] += (u32)__builtin_popcountll(w1 & ~w0);
        cnt[2] += (u32)__builtin_popcountll(w2 & ~w1);
        cnt[3] += (u32)__builtin_popcountll(w3 & ~w2);
    }
}

template <typename BE> struct DeviceIndex {
    BE& be;
    IndexView view;
    HostIndexMeta meta;
    std::vector<void*> allocs;

    DeviceIndex(BE& b) : be(b) { memset(&view, 0, sizeof(view)); }
    ~DeviceIndex() {
        for (void* p : allocs)
            be.free(p);
    }
    template <typename T> const T* upload(const std::vector<T>& v) {
        size_t bytes = v.size() * sizeof(T);
        void* d = be.alloc(bytes ? bytes : 16);
        allocs.push_back(d);
        if (bytes)
            be.h2d(d, v.data(), bytes);
        meta.deviceBytes += bytes;
        return (const T*)d;
    }

    const void* uploadRaw(const void* src, size_t bytes) {
        void* d = be.alloc(bytes ? bytes : 16);
        allocs.push_back(d);
        if (bytes)
            be.h2d(d, src, bytes);
        meta.deviceBytes += bytes;
        return d;
    }
    // the .gidx sidecar (clb_gidx.h): the device layout as written by
    // clb_gpu_build; returns false when there is none for this sparseness
    bool loadGidx(const std::string& base, u32 saSparse) {
        const std::string fn = base + ".gidx";
        if (getenv("CLB_NO_GIDX")) // tests: force the re-packing path
            return false;
        int fd = ::open(fn.c_str(), O_RDONLY);
        if (fd < 0)
            return false;
        struct stat sb;
        if (fstat(fd, &sb) != 0 || (size_t)sb.st_size < sizeof(GidxHeader)) {
            ::close(fd);
            throw std::runtime_error("Corrupt index sidecar: " + fn);
        }
        const size_t fileBytes = (size_t)sb.st_size;
        void* mp = mmap(nullptr, fileBytes, PROT_READ, MAP_PRIVATE, fd, 0);
        ::close(fd);
        if (mp == MAP_FAILED)
            throw std::runtime_error("Cannot map file: " + fn);
        struct Unmap {
            void* p;
            size_t n;
            ~Unmap() { munmap(p, n); }
        } guard{mp, fileBytes};
        madvise(mp, fileBytes, MADV_SEQUENTIAL);
        const u8* m8 = (const u8*)mp;
        const GidxHeader& H = *(const GidxHeader*)mp;
        if (memcmp(H.magic, CLB_GIDX_MAGIC, 8) != 0 || H.version != CLB_GIDX_VERSION || H.fileBytes != fileBytes)
            throw std::runtime_error("Corrupt index sidecar (header): " + fn);
        if (H.saSparse != saSparse)
            return false; // built for another sparseness: fall back to the reference's files
        auto inside = [&](u64 off, u64 bytes) { return off <= fileBytes && bytes <= fileBytes - off; };
        if (!inside(H.offSeqStart, (u64)(H.numSeq + 1) * 4) || !inside(H.offNames, H.namesBytes) ||
            !inside(H.offFwd, H.nRec * sizeof(BwtRec)) || !inside(H.offRev, H.nRec * sizeof(BwtRec)) ||
            !inside(H.offSsaRec, H.nSsaRec * sizeof(SsaRec)) || !inside(H.offSsa, H.nSsa * 4) ||
            !inside(H.offText, H.nTextWords * 4) || H.nRec != (u64)H.textLength / 64 + 1 ||
            H.nTextWords < (u64)H.textLength / 16 + 8)
            throw std::runtime_error("Corrupt index sidecar (sections): " + fn);
        view.textLength = H.textLength;
        for (int a = 0; a < 5; a++)
            view.C[a] = H.C[a];
        meta.textLength = view.textLength;
        memcpy(meta.C, view.C, sizeof(meta.C));
        view.text = (const u32*)uploadRaw(m8 + H.offText, H.nTextWords * 4);
        view.fwd.rec = (const BwtRec*)uploadRaw(m8 + H.offFwd, H.nRec * sizeof(BwtRec));
        view.fwd.dollarPos = H.dollarFwd;
        view.rev.rec = (const BwtRec*)uploadRaw(m8 + H.offRev, H.nRec * sizeof(BwtRec));
        view.rev.dollarPos = H.dollarRev;
        view.ssaRec = (const SsaRec*)uploadRaw(m8 + H.offSsaRec, H.nSsaRec * sizeof(SsaRec));
        view.ssa = (const u32*)uploadRaw(m8 + H.offSsa, H.nSsa * 4);
        meta.saSparse = saSparse;
        meta.seqStart.assign((const u32*)(m8 + H.offSeqStart), (const u32*)(m8 + H.offSeqStart) + H.numSeq + 1);
        view.seqStart = upload(meta.seqStart);
        view.numSeq = H.numSeq;
        const char* nm = (const char*)(m8 + H.offNames);
        for (u64 o = 0; o < H.namesBytes && meta.seqNames.size() < H.numSeq;) {
            std::string s(nm + o);
            o += s.size() + 1;
            meta.seqNames.push_back(s);
        }
        be.sync(); // the copies read the mapping
        meta.fromGidx = true;
        return true;
    }

    void load(const std::string& base, u32 saSparse, u32 K, u32 inTextSwitch, bool noCigar) {
        if (K > 13)
            throw std::runtime_error("k-mer size too large for the dense k-mer table (max 13)");
        if (loadGidx(base, saSparse)) {
            view.switchPoint = inTextSwitch;
            view.noCigar = noCigar;
            buildKmerTable(K);
            return;
        }
        { // .meta
            std::ifstream m(base + ".meta");
            if (m) {
                u32 tag = 0;
                size_t lsz = 0;
                std::string fl;
                m >> tag >> lsz >> fl;
                if (lsz != 4)
                    throw std::runtime_error("The index was built with a compiled version that uses " +
                                             std::to_string(lsz * 8) +
                                             "-bit numbers, while the current programme was compiled using 32-bit numbers. "
                                             "Recompile the programme with the correct THIRTY_TWO flag set or rebuild the index.");
                if (fl != "VANILLA")
                    throw std::runtime_error("The index was built with a different flavor of Columba. \n\tCurrent flavor: "
                                             "VANILLA\n\tIndex flavor: " + fl);
            }
        }
        std::vector<u32> cc;
        if (!readWholeFile(base + ".cct", cc) || cc.size() != 256)
            throw std::runtime_error("Cannot open file: " + base + ".cct");
        {
            static const char order[5] = {'$', 'A', 'C', 'G', 'T'};
            u32 cum = 0;
            for (int i = 0; i < 256; i++) {
                bool inAlphabet = false;
                for (int a = 0; a < 5; a++)
                    inAlphabet |= (i == order[a]);
                if (cc[i] && !inAlphabet)
                    throw std::runtime_error("The index contains characters outside $ACGT; only DNA indexes are supported");
            }
            for (int a = 0; a < 5; a++) {
                view.C[a] = cum;
                cum += cc[(unsigned char)order[a]];
            }
            view.textLength = cum;
            if (cc[(unsigned char)'$'] != 1)
                throw std::runtime_error("The index text must contain exactly one sentinel character");
        }
        meta.textLength = view.textLength;
        memcpy(meta.C, view.C, sizeof(meta.C));
        { // text -> 2 bit
            std::ifstream f(base + ".txt.bin", std::ios::binary);
            if (!f)
                throw std::runtime_error("Error opening file for reading: " + base + ".txt.bin");
            u32 tl = 0;
            f.read((char*)&tl, 4);
            if (tl != view.textLength)
                throw std::runtime_error("Text length does not match the character counts: " + base + ".txt.bin");
            std::vector<u32> packed((size_t)tl / 16 + 8, 0); // padded: textWindow64 reads up to 3 words past a position
            std::vector<char> buf(1 << 20);
            size_t pos = 0;
            while (pos < tl) {
                size_t n = std::min(buf.size(), (size_t)tl - pos);
                f.read(buf.data(), (std::streamsize)n);
                if (!f)
                    throw std::runtime_error("Truncated file: " + base + ".txt.bin");
                for (size_t i = 0; i < n; i++, pos++) {
                    u32 c;
                    switch (buf[i]) {
                    case 'C': c = 1; break;
                    case 'G': c = 2; break;
                    case 'T': c = 3; break;
                    default: c = 0;
                    }
                    packed[pos >> 4] |= c << ((pos & 15) * 2);
                }
            }
            view.text = upload(packed);
        }
        {
            std::vector<BwtRec> rec;
            u32 dp;
            convertBwt(base + ".brt", view.textLength, rec, dp);
            view.fwd.rec = upload(rec);
            view.fwd.dollarPos = dp;
            convertBwt(base + ".rev.brt", view.textLength, rec, dp);
            view.rev.rec = upload(rec);
            view.rev.dollarPos = dp;
        }
        { // sparse suffix array
            std::string sfx = std::to_string(saSparse);
            std::string fn = base + ".sa.bv." + sfx;
            std::ifstream f(fn, std::ios::binary);
            if (!f)
                throw std::runtime_error("Cannot open file: " + fn +
                                         ". Did you set an incorrect suffix array sparseness factor using the -s flag or "
                                         "move your index files?");
            u64 N = 0;
            f.read((char*)&N, 8);
            if (N != view.textLength)
                throw std::runtime_error("Corrupt suffix array bitvector: " + fn);
            std::vector<u64> bits((size_t)((N + 63) / 64));
            f.read((char*)bits.data(), (std::streamsize)(bits.size() * 8));
            if (!f)
                throw std::runtime_error("Truncated file: " + fn);
            std::vector<SsaRec> rec(bits.size() + 1);
            u32 run = 0;
            for (size_t i = 0; i < bits.size(); i++) {
                rec[i].bits = bits[i];
                rec[i].rank = run;
                rec[i].pad = 0;
                run += (u32)__builtin_popcountll(bits[i]);
            }
            rec[bits.size()].bits = 0;
            rec[bits.size()].rank = run;
            rec[bits.size()].pad = 0;
            view.ssaRec = upload(rec);
            std::vector<u32> ssa;
            if (!readWholeFile(base + ".sa." + sfx, ssa))
                throw std::runtime_error("Cannot open file: " + base + ".sa." + sfx);
            if (ssa.size() != run)
                throw std::runtime_error("Sparse suffix array does not match its bitvector: " + base + ".sa." + sfx);
            view.ssa = upload(ssa);
            meta.saSparse = saSparse;
        }
        if (!readWholeFile(base + ".pos", meta.seqStart) || meta.seqStart.size() < 2)
            throw std::runtime_error("Cannot open file: " + base + ".pos");
        view.seqStart = upload(meta.seqStart);
        view.numSeq = (u32)meta.seqStart.size() - 1;
        {
            std::ifstream f(base + ".sna", std::ios::binary);
            if (!f)
                throw std::runtime_error("Cannot open file: " + base + ".sna");
            while (f.good()) {
                size_t len = 0;
                f.read((char*)&len, sizeof(len));
                if (!f.good())
                    break;
                std::string s(len, ' ');
                f.read(&s[0], (std::streamsize)len);
                meta.seqNames.push_back(s);
            }
        }
        view.switchPoint = inTextSwitch;
        view.noCigar = noCigar;
        buildKmerTable(K);
    }

    // indexinterface.cpp:294-335, level-synchronous on the device
    void buildKmerTable(u32 K) {
        view.K = K;
        meta.K = K;
        size_t n = (size_t)1 << (2 * K);
        RangePair* a = (RangePair*)be.alloc(n * sizeof(RangePair));
        RangePair* b = (RangePair*)be.alloc((n / 4 ? n / 4 : 1) * sizeof(RangePair));
        RangePair root = completeRange(view);
        // levels alternate between the two buffers so that level K ends in `a`
        RangePair* cur = (K % 2 == 0) ? a : b;
        be.h2d(cur, &root, sizeof(root));
        for (u32 d = 1; d <= K; d++) {
            RangePair* nxt = (cur == a) ? b : a;
            KmerLevelKernel k{view, cur, nxt};
            be.launch(k, (u32)((size_t)1 << (2 * d)), nullptr, WK_LIGHT);
            cur = nxt;
        }
        be.sync();
        be.free(b);
        allocs.push_back(a);
        meta.deviceBytes += n * sizeof(RangePair);
        view.kmer = a;
    }
};

} // namespace clb
