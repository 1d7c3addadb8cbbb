// columba_b200 -- host-side index construction for synthetic benchmarks and
// tests.  NOT the hot path (index construction stays on the host per the
// project scope); it exists so a benchmark box without the reference can turn
// a generated FASTA into the reference's Vanilla index file set.  The files are
// byte-identical to what the reference's columba_build writes (32-bit
// length_t) for ACGT-only FASTA input: formats per src/buildindex.cpp:341-387,
// :432-437, :591-605, :688-712, :575-584, src/fmindex/bwtrepr.h:56-72,:113-120,
// src/bitvec.h:138-153,:329-349,:378-382, src/fmindex/suffixArray.h:161-166,
// :229-243, src/fmindex/encodedtext.h:95-130,:275-285.
// The suffix sorter is a plain prefix-doubling sorter written for this file.
#pragma once
#include "../../include/columba_b200.h"
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace clb {
namespace build {

// suffix array of s[0,n) over small integer codes (>0), shorter suffix first
static void suffixSort(const std::vector<uint8_t>& s, std::vector<uint32_t>& sa) {
    const size_t n = s.size();
    sa.resize(n);
    if (n == 0)
        return;
    const int H0 = 21; // 3 bits per symbol
    std::vector<std::pair<uint64_t, uint32_t>> keyed(n);
    {
        uint64_t key = 0;
        // rolling key from the back
        for (size_t ii = n; ii-- > 0;) {
            key = (key >> 3) | ((uint64_t)s[ii] << (3 * (H0 - 1)));
            keyed[ii] = std::make_pair(key, (uint32_t)ii);
        }
    }
    std::sort(keyed.begin(), keyed.end());
    std::vector<uint32_t> rank(n + 1), tmp(n + 1);
    rank[n] = 0; // the empty suffix sorts before everything; real ranks are 1-based
    std::vector<uint32_t> groups; // starts of unsorted groups, pairs (start,end)
    {
        size_t i = 0;
        while (i < n) {
            size_t j = i + 1;
            while (j < n && keyed[j].first == keyed[i].first)
                j++;
            for (size_t q = i; q < j; q++) {
                sa[q] = keyed[q].second;
                rank[keyed[q].second] = (uint32_t)i + 1;
            }
            if (j - i > 1) {
                groups.push_back((uint32_t)i);
                groups.push_back((uint32_t)j);
            }
            i = j;
        }
    }
    keyed.clear();
    keyed.shrink_to_fit();
    size_t h = H0;
    std::vector<uint32_t> next;
    std::vector<std::pair<uint32_t, uint32_t>> buf;
    while (!groups.empty()) {
        next.clear();
        // new ranks are written to tmp for the members of unsorted groups only
        for (size_t g = 0; g < groups.size(); g += 2) {
            const uint32_t b = groups[g], e = groups[g + 1];
            buf.resize(e - b);
            for (uint32_t q = b; q < e; q++) {
                uint32_t p = sa[q];
                // members of an unsorted group have at least h symbols: p + h <= n
                buf[q - b] = std::make_pair(rank[p + h], p);
            }
            std::sort(buf.begin(), buf.end());
            uint32_t i = 0;
            const uint32_t m = e - b;
            while (i < m) {
                uint32_t j = i + 1;
                while (j < m && buf[j].first == buf[i].first)
                    j++;
                for (uint32_t q = i; q < j; q++) {
                    sa[b + q] = buf[q].second;
                    tmp[buf[q].second] = b + i + 1;
                }
                if (j - i > 1) {
                    next.push_back(b + i);
                    next.push_back(b + j);
                }
                i = j;
            }
        }
        for (size_t g = 0; g < groups.size(); g += 2)
            for (uint32_t q = groups[g]; q < groups[g + 1]; q++)
                rank[sa[q]] = tmp[sa[q]];
        groups.swap(next);
        h *= 2;
    }
}

static inline int c2i(char c) {
    switch (c) {
    case '$': return 0;
    case 'A': return 1;
    case 'C': return 2;
    case 'G': return 3;
    case 'T': return 4;
    default: return -1;
    }
}

// BWTRepresentation + BitvecIntl<4> file (.brt / .rev.brt)
static void writeBrt(const std::string& fn, const std::string& bwt) {
    const uint64_t N = bwt.size() + 1, S = 4;
    const uint64_t bvSize = S * ((N + 63) / 64), countsSize = 2 * S * ((N + 511) / 512);
    std::vector<uint64_t> bv(bvSize, 0), counts(countsSize, 0);
    uint64_t dollarPos = bwt.size();
    for (size_t i = 0; i < bwt.size(); i++) {
        if (bwt[i] == '$') {
            dollarPos = i;
            continue;
        }
        for (uint64_t c = (uint64_t)c2i(bwt[i]); c < 5; c++)
            bv[(i / 64) * S + (c - 1)] |= 1ull << (i % 64);
    }
    for (uint64_t c = 0; c < S; c++) {
        uint64_t countL1 = 0, countL2 = 0;
        for (uint64_t w = c, q = 2 * c; w < bvSize; w += S) {
            uint64_t numBits = (uint64_t)__builtin_popcountll(bv[w]);
            if (w % (8 * S) == c) {
                countL1 += countL2;
                counts[q] = countL1;
                countL2 = numBits;
                q += 2 * S;
            } else {
                uint64_t L2offs = 9 * ((w / S % 8) - 1);
                counts[q + 1 - 2 * S] |= (countL2 << L2offs);
                countL2 += numBits;
            }
        }
    }
    std::ofstream o(fn, std::ios::binary);
    if (!o)
        throw std::runtime_error("Cannot open file: " + fn);
    o.write((const char*)&dollarPos, 8);
    o.write((const char*)&N, 8);
    o.write((const char*)bv.data(), (std::streamsize)(bvSize * 8));
    o.write((const char*)counts.data(), (std::streamsize)(countsSize * 8));
}

// EncodedText<5> file (.bwt): 3 bits per symbol, most significant bits first
static void writeEncodedBwt(const std::string& fn, const std::string& bwt) {
    const size_t tSize = bwt.size();
    std::vector<uint64_t> enc((tSize * 3) / 64 + 1, 0);
    for (size_t i = 0; i < tSize; i++) {
        uint64_t bits = (uint64_t)c2i(bwt[i]);
        size_t bitPos = i * 3, w = bitPos / 64, b = bitPos % 64;
        if (b <= 61) {
            enc[w] |= bits << (61 - b);
        } else {
            unsigned over = (unsigned)(b + 3 - 64); // bits that go to the next word
            enc[w] |= bits >> over;
            enc[w + 1] |= (bits & ((1ull << over) - 1)) << (64 - over);
        }
    }
    std::ofstream o(fn, std::ios::binary);
    if (!o)
        throw std::runtime_error("Cannot open file: " + fn);
    size_t vs = enc.size();
    o.write((const char*)&tSize, sizeof(size_t));
    o.write((const char*)&vs, sizeof(size_t));
    o.write((const char*)enc.data(), (std::streamsize)(vs * 8));
}

static void buildIndex(const std::string& fasta, const std::string& base, int saSparse) {
    std::ifstream in(fasta);
    if (!in)
        throw std::runtime_error("Cannot open file: " + fasta);
    std::string T, line, sequence;
    std::vector<uint32_t> positions;
    std::vector<std::string> names;
    uint32_t startPosition = 0;
    auto flush = [&]() {
        if (!sequence.empty()) {
            for (char& c : sequence) {
                c = (char)toupper((unsigned char)c);
                if (c2i(c) < 1)
                    throw std::runtime_error("clb_build_index handles ACGT-only FASTA input (use the reference's "
                                             "columba_build for anything else)");
            }
            T += sequence;
            positions.push_back(startPosition);
            startPosition += (uint32_t)sequence.size();
            sequence.clear();
        }
    };
    while (std::getline(in, line)) {
        if (!line.empty() && line.back() == '\r')
            line.pop_back();
        if (line.empty())
            continue;
        if (line[0] == '>') {
            flush();
            std::string desc = line.substr(1);
            std::string name = desc.substr(0, desc.find(' '));
            if (std::find(names.begin(), names.end(), name) != names.end())
                throw std::runtime_error("Error: Sequence name " + name + " in file " + fasta + " is not unique!");
            names.push_back(name);
        } else {
            sequence += line;
        }
    }
    flush();
    if (names.empty() || positions.size() != names.size())
        throw std::runtime_error("No sequences found in " + fasta);
    positions.push_back((uint32_t)T.size());
    T += '$';
    if (T.size() >= 0xFFFFFFF0ull)
        throw std::runtime_error("The size of the reference is too large for 32-bit indexes");
    const uint32_t n = (uint32_t)T.size();
    { // .txt.bin
        std::ofstream o(base + ".txt.bin", std::ios::binary);
        if (!o)
            throw std::runtime_error("Error opening file for writing: " + base + ".txt.bin");
        o.write((const char*)&n, 4);
        o.write(T.data(), n);
    }
    { // .headerSN.bin, .pos, .sna, .fsid
        std::ofstream h(base + ".headerSN.bin", std::ios::binary);
        for (size_t i = 0; i < names.size(); i++)
            h << "@SQ\tSN:" << names[i] << "\tLN:" << positions[i + 1] - positions[i] << "\n";
        std::ofstream p(base + ".pos", std::ios::binary);
        p.write((const char*)positions.data(), (std::streamsize)(positions.size() * 4));
        std::ofstream s(base + ".sna", std::ios::binary);
        for (auto& nm : names) {
            size_t len = nm.size();
            s.write((const char*)&len, sizeof(len));
            s.write(nm.data(), (std::streamsize)len);
        }
        std::ofstream f(base + ".fsid", std::ios::binary);
        uint32_t zero = 0;
        f.write((const char*)&zero, 4);
    }
    { // .cct
        std::vector<uint32_t> cc(256, 0);
        for (char c : T)
            cc[(unsigned char)c]++;
        for (char c : std::string("ACGT"))
            if (!cc[(unsigned char)c])
                throw std::runtime_error("the reference must contain all of A, C, G and T");
        std::ofstream o(base + ".cct", std::ios::binary);
        o.write((const char*)cc.data(), 1024);
    }
    std::vector<uint8_t> codes(n);
    std::vector<uint32_t> sa;
    std::string bwt(n, '$');
    { // forward: SA, sparse SA, BWT
        for (uint32_t i = 0; i < n; i++)
            codes[i] = (uint8_t)(c2i(T[i]) + 1);
        suffixSort(codes, sa);
        const uint64_t N = n;
        std::vector<uint64_t> bits((N + 63) / 64, 0);
        std::vector<uint32_t> sparse;
        sparse.reserve(n / saSparse + 1);
        for (uint32_t i = 0; i < n; i++) {
            if (sa[i] % (uint32_t)saSparse == 0) {
                sparse.push_back(sa[i]);
                bits[i / 64] |= 1ull << (i % 64);
            }
            bwt[i] = sa[i] > 0 ? T[sa[i] - 1] : T.back();
        }
        std::vector<uint64_t> counts((bits.size() + 7) / 4, 0);
        uint64_t countL1 = 0, countL2 = 0;
        for (uint64_t w = 0, q = 0; w < bits.size(); w++) {
            if (w % 8 == 0) {
                countL1 += countL2;
                counts[q] = countL1;
                countL2 = (uint64_t)__builtin_popcountll(bits[w]);
                q += 2;
            } else {
                counts[q - 1] |= (countL2 << (((w % 8) - 1) * 9));
                countL2 += (uint64_t)__builtin_popcountll(bits[w]);
            }
        }
        const std::string sfx = std::to_string(saSparse);
        std::ofstream o(base + ".sa.bv." + sfx, std::ios::binary);
        o.write((const char*)&N, 8);
        o.write((const char*)bits.data(), (std::streamsize)(bits.size() * 8));
        o.write((const char*)counts.data(), (std::streamsize)(counts.size() * 8));
        std::ofstream o2(base + ".sa." + sfx, std::ios::binary);
        o2.write((const char*)sparse.data(), (std::streamsize)(sparse.size() * 4));
        writeEncodedBwt(base + ".bwt", bwt);
        writeBrt(base + ".brt", bwt);
    }
    { // reverse
        for (uint32_t i = 0; i < n; i++)
            codes[i] = (uint8_t)(c2i(T[n - 1 - i]) + 1);
        suffixSort(codes, sa);
        for (uint32_t i = 0; i < n; i++)
            bwt[i] = sa[i] > 0 ? T[n - sa[i]] : T.front();
        writeBrt(base + ".rev.brt", bwt);
    }
    {
        std::ofstream m(base + ".meta");
        m << 21 << "\n" << 4 << "\n" << "VANILLA" << "\n";
    }
}

} // namespace build
} // namespace clb

extern "C" int clb_build_index(const char* fasta, const char* base, int sa_sparse, char* err, size_t err_len) {
    if (!fasta || !base || sa_sparse <= 0 || (sa_sparse & (sa_sparse - 1)) || sa_sparse > 256) {
        if (err && err_len)
            snprintf(err, err_len, "clb_build_index: invalid argument");
        return CLB_E_INVALID;
    }
    try {
        clb::build::buildIndex(fasta, base, sa_sparse);
    } catch (std::exception& e) {
        if (err && err_len)
            snprintf(err, err_len, "%s", e.what());
        return CLB_E_IO;
    }
    return CLB_OK;
}
