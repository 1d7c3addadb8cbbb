// clb_gpu_build -- standalone index builder for benchmark-sized references.
//
// NOT part of libcolumba_b200.so and NOT on the hot path (index construction is
// out of scope per BASELINE.json north_star); it exists so that a benchmark box
// can turn a 3.1 Gbp synthetic reference into (a) the reference's own Vanilla
// index file set, byte-identical with what the unmodified columba_build writes
// (formats: src/buildindex.cpp:341-387,432-437,575-605,688-712,
// src/fmindex/bwtrepr.h:56-72,113-120, src/bitvec.h:138-153,329-349,378-382,
// src/fmindex/suffixArray.h:161-166,229-243, src/fmindex/encodedtext.h:95-130,
// 275-285) and (b) the .gidx sidecar: the HBM layout of clb_index.h as one file,
// so that clb_index_load does not re-pack 12 GB at every start.
//
// Because it is a separate executable, the benchmark's reference arm can call
// it without loading any product kernel.  Suffix sorting = prefix doubling over
// radix-sorted (key, suffix) pairs with CUB (library primitives are fine here:
// this is not the hot path); everything derived from the suffix array (BWT,
// cumulative bit planes + rank9 counts, sparse SA, 3-bit BWT text, 32-byte rank
// records) is computed on the device and streamed to the files.
//
//   clb_gpu_build --out BASE (--fasta F | --synthetic human|ecoli|none --bp N [--seed S] [--nseq K])
//                 [--sparse 4] [--write-fasta] [--no-ref-files] [--no-gidx] [--device D]
#include <cuda_runtime.h>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>

#include <algorithm>
#include <chrono>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "../clb_gidx.h"

typedef uint8_t u8;
typedef uint32_t u32;
typedef uint64_t u64;

#define CK(x)                                                                                         \
    do {                                                                                              \
        cudaError_t e_ = (x);                                                                         \
        if (e_ != cudaSuccess)                                                                        \
            throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(e_) + " at " #x + \
                                     " (" __FILE__ ":" + std::to_string(__LINE__) + ")");            \
    } while (0)

static bool g_verbose = true;
static double nowS() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
static double g_t0 = 0;
static void logf(const char* fmt, ...) {
    if (!g_verbose)
        return;
    va_list ap;
    va_start(ap, fmt);
    fprintf(stderr, "[clb_gpu_build %7.2fs] ", nowS() - g_t0);
    vfprintf(stderr, fmt, ap);
    fprintf(stderr, "\n");
    va_end(ap);
}

template <typename T> struct DBuf {
    T* p = nullptr;
    size_t n = 0;
    DBuf() {}
    explicit DBuf(size_t n_) { alloc(n_); }
    DBuf(const DBuf&) = delete;
    DBuf& operator=(const DBuf&) = delete;
    ~DBuf() { release(); }
    void alloc(size_t n_) {
        release();
        n = n_;
        CK(cudaMalloc(&p, (n ? n : 1) * sizeof(T)));
    }
    void release() {
        if (p)
            cudaFree(p);
        p = nullptr;
        n = 0;
    }
    void zero() { CK(cudaMemset(p, 0, n * sizeof(T))); }
};

static inline u32 gridFor(size_t n, u32 block = 256) {
    size_t b = (n + block - 1) / block;
    if (b > 148u * 64u)
        b = 148u * 64u;
    if (b < 1)
        b = 1;
    return (u32)b;
}

// ------------------------------------------------------------------ hashing
__host__ __device__ inline u64 mix64(u64 x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
struct HostRng { // splitmix64 stream: identical on every machine
    u64 s;
    explicit HostRng(u64 seed) : s(seed) {}
    u64 next() { return mix64(s++); }
    u64 below(u64 n) { return n ? next() % n : 0; }
    double uni() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }
    double uni(double a, double b) { return a + (b - a) * uni(); }
};

// ------------------------------------------------------------------ synthetic text
// codes: 1..4 = A,C,G,T (5 = '$' never generated here)
struct CopyDesc {
    u64 pos;       // first position in the text
    u32 len;
    u32 master;    // offset into the masters buffer (type 0) / seed (type 1)
    u32 thresh;    // substitution probability * 2^32 (type 0)
    u32 type;      // 0: mutated copy of masters[master ...], 1: periodic unit of length `unit`
    u32 unit;
    u32 id;        // global copy id (later copies overwrite earlier ones)
};

__global__ void kBaseText(u8* codes, size_t n, u64 seed) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        codes[i] = (u8)(1 + (mix64(seed * 0x100000001B3ull + i) >> 61 & 3));
}
__device__ inline u32 findDesc(const u64* pre, u32 nd, u64 w) {
    u32 lo = 0, hi = nd; // pre[lo] <= w < pre[lo+1]
    while (hi - lo > 1) {
        u32 mid = (lo + hi) >> 1;
        if (pre[mid] <= w)
            lo = mid;
        else
            hi = mid;
    }
    return lo;
}
__global__ void kOwner(const CopyDesc* d, const u64* pre, u32 nd, u64 total, u32* owner) {
    for (u64 w = (u64)blockIdx.x * blockDim.x + threadIdx.x; w < total; w += (u64)gridDim.x * blockDim.x) {
        const u32 j = findDesc(pre, nd, w);
        atomicMax(owner + d[j].pos + (w - pre[j]), d[j].id + 1);
    }
}
__global__ void kApply(const CopyDesc* d, const u64* pre, u32 nd, u64 total, const u32* owner, const u8* masters,
                       u8* codes, u64 seed) {
    for (u64 w = (u64)blockIdx.x * blockDim.x + threadIdx.x; w < total; w += (u64)gridDim.x * blockDim.x) {
        const u32 j = findDesc(pre, nd, w);
        const CopyDesc c = d[j];
        const u64 off = w - pre[j];
        if (owner[c.pos + off] != c.id + 1)
            continue;
        u32 b;
        if (c.type == 1) {
            b = (u32)(mix64(seed ^ ((u64)c.master << 20) ^ (off % c.unit)) >> 61) & 3;
        } else {
            b = masters[c.master + off] - 1;
            const u64 h = mix64(seed ^ ((u64)(c.id + 1) << 36) ^ off);
            if ((u32)h < c.thresh)
                b = (b + 1 + (u32)((h >> 32) % 3)) & 3;
        }
        codes[c.pos + off] = (u8)(b + 1);
    }
}
__global__ void kGather(const u8* codes, u8* out, u64 src, u32 len) {
    for (u32 i = blockIdx.x * blockDim.x + threadIdx.x; i < len; i += gridDim.x * blockDim.x)
        out[i] = codes[src + i];
}

struct SeqTable {
    std::vector<std::string> names;
    std::vector<u32> start; // numSeq+1 (last = position of '$')
};

static void applyCopies(std::vector<CopyDesc>& ds, u8* dCodes, u32* dOwner, const u8* dMasters, u64 seed) {
    if (ds.empty())
        return;
    std::vector<u64> pre(ds.size() + 1, 0);
    for (size_t i = 0; i < ds.size(); i++)
        pre[i + 1] = pre[i] + ds[i].len;
    DBuf<CopyDesc> dd(ds.size());
    DBuf<u64> dp(pre.size());
    CK(cudaMemcpy(dd.p, ds.data(), ds.size() * sizeof(CopyDesc), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dp.p, pre.data(), pre.size() * 8, cudaMemcpyHostToDevice));
    const u64 total = pre.back();
    kOwner<<<gridFor(total), 256>>>(dd.p, dp.p, (u32)ds.size(), total, dOwner);
    kApply<<<gridFor(total), 256>>>(dd.p, dp.p, (u32)ds.size(), total, dOwner, dMasters, dCodes, seed);
    CK(cudaDeviceSynchronize());
}

// iid uniform ACGT + injected repeat families, counts
// scaled by bp / 3.1e9.  Pure function of (bp, seed, profile).
static void makeSynthetic(const std::string& profile, u64 bp, u64 seed, u32 nSeq, DBuf<u8>& codes, SeqTable& st) {
    const size_t n = bp + 1;
    codes.alloc(n + 64);
    CK(cudaMemset(codes.p, 0, n + 64));
    kBaseText<<<gridFor(bp), 256>>>(codes.p, bp, seed);
    CK(cudaDeviceSynchronize());
    const double scale = (double)bp / 3.1e9;
    HostRng rng(seed * 7919 + 17);
    std::vector<u8> masters;
    auto newMaster = [&](u32 len) {
        u32 off = (u32)masters.size();
        for (u32 i = 0; i < len; i++)
            masters.push_back((u8)(1 + rng.below(4)));
        return off;
    };
    std::vector<CopyDesc> ds;
    u32 nextId = 0;
    auto family = [&](u32 unitLen, double copies, double dlo, double dhi, u32 truncMin) {
        u64 nc = (u64)copies;
        if (nc == 0 || (u64)unitLen >= bp / 4)
            return;
        const u32 m = newMaster(unitLen);
        for (u64 c = 0; c < nc; c++) {
            u32 cut = 0;
            if (truncMin && unitLen > truncMin)
                cut = (u32)rng.below(unitLen - truncMin);
            CopyDesc d;
            d.len = unitLen - cut;
            d.master = m + cut;
            d.thresh = (u32)(rng.uni(dlo, dhi) * 4294967296.0);
            d.pos = rng.below(bp - d.len);
            d.type = 0;
            d.unit = 0;
            d.id = nextId++;
            ds.push_back(d);
        }
    };
    u32 nSeg = 0;
    if (profile == "human") {
        family(300, std::max(50.0, 1000000.0 * scale), 0.05, 0.15, 0);
        family(6000, std::max(8.0, 80000.0 * scale), 0.02, 0.12, 900);
        family(1000, std::max(10.0, 100000.0 * scale), 0.01, 0.10, 0);
        const u64 nMicro = (u64)std::max(10.0, 100000.0 * scale);
        for (u64 i = 0; i < nMicro; i++) {
            CopyDesc d;
            d.unit = 1 + (u32)rng.below(6);
            d.len = 60 + (u32)rng.below(141);
            d.pos = rng.below(bp - d.len);
            d.type = 1;
            d.master = (u32)rng.below(1u << 30);
            d.thresh = 0;
            d.id = nextId++;
            ds.push_back(d);
        }
        const u64 nArr = (u64)std::max(1.0, 200.0 * scale);
        for (u64 a = 0; a < nArr; a++) {
            u64 units = std::min<u64>(2000, std::max<u64>(20, bp / (171 * 400)));
            const u32 m = newMaster(171);
            const u64 p0 = rng.below(bp - units * 171);
            for (u64 q = 0; q < units; q++) {
                CopyDesc d;
                d.len = 171;
                d.master = m;
                d.thresh = (u32)(rng.uni(0.01, 0.03) * 4294967296.0);
                d.pos = p0 + q * 171;
                d.type = 0;
                d.unit = 0;
                d.id = nextId++;
                ds.push_back(d);
            }
        }
        nSeg = (u32)std::max(2.0, 200.0 * scale);
    } else if (profile == "ecoli") {
        const double f = (double)bp / 4.6e6;
        family(5000, std::max(2.0, 7.0 * f + 0.5), 0.005, 0.005, 0);
        family(1300, std::max(3.0, 30.0 * f + 0.5), 0.01, 0.03, 0);
    } else if (profile != "none") {
        throw std::runtime_error("unknown synthetic profile " + profile);
    }
    DBuf<u32> owner(n);
    owner.zero();
    {
        DBuf<u8> dM(masters.size() + 1);
        if (!masters.empty())
            CK(cudaMemcpy(dM.p, masters.data(), masters.size(), cudaMemcpyHostToDevice));
        applyCopies(ds, codes.p, owner.p, dM.p, seed);
    }
    if (nSeg) { // segmental duplications: copies of regions of the text as it is now
        const u32 segLen = (u32)std::min<u64>(40000, bp / 50);
        DBuf<u8> dM((size_t)nSeg * segLen);
        std::vector<CopyDesc> sd;
        for (u32 s = 0; s < nSeg; s++) {
            const u64 src = rng.below(bp - segLen);
            kGather<<<gridFor(segLen), 256>>>(codes.p, dM.p + (size_t)s * segLen, src, segLen);
            CopyDesc d;
            d.len = segLen;
            d.master = s * segLen;
            d.thresh = (u32)(rng.uni(0.005, 0.02) * 4294967296.0);
            d.pos = rng.below(bp - segLen);
            d.type = 0;
            d.unit = 0;
            d.id = nextId++;
            sd.push_back(d);
        }
        CK(cudaDeviceSynchronize());
        applyCopies(sd, codes.p, owner.p, dM.p, seed);
    }
    u8 dollar = 5;
    CK(cudaMemcpy(codes.p + bp, &dollar, 1, cudaMemcpyHostToDevice));
    // sequences with GRCh38-like (decreasing) proportions
    st.names.clear();
    st.start.clear();
    std::vector<double> w(nSeq);
    double ws = 0;
    for (u32 i = 0; i < nSeq; i++)
        ws += (w[i] = 1.0 / (1.0 + 0.12 * i));
    double cum = 0;
    st.start.push_back(0);
    for (u32 i = 0; i < nSeq; i++) {
        cum += w[i] / ws;
        u64 cut = (i + 1 == nSeq) ? bp : (u64)(cum * (double)bp + 0.5);
        if (cut <= st.start.back())
            cut = st.start.back() + 1;
        st.start.push_back((u32)cut);
        st.names.push_back(nSeq == 1 ? "seq1_syn" : "chr" + std::to_string(i + 1) + "_syn");
    }
}

static int c2code(char c) {
    switch (c) {
    case 'A': case 'a': return 1;
    case 'C': case 'c': return 2;
    case 'G': case 'g': return 3;
    case 'T': case 't': return 4;
    default: return 0;
    }
}
static void readFasta(const std::string& fasta, DBuf<u8>& codes, SeqTable& st, size_t& n) {
    std::ifstream in(fasta);
    if (!in)
        throw std::runtime_error("Cannot open file: " + fasta);
    std::vector<u8> T;
    std::string line;
    size_t cur = 0;
    bool open = false;
    auto flush = [&]() {
        if (open && T.size() > cur) {
            st.start.push_back((u32)cur);
            cur = T.size();
        } else if (open) {
            st.names.pop_back();
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
            st.names.push_back(desc.substr(0, desc.find(' ')));
            open = true;
        } else {
            for (char c : line) {
                int v = c2code(c);
                if (!v)
                    throw std::runtime_error("clb_gpu_build handles ACGT-only FASTA input");
                T.push_back((u8)v);
            }
        }
    }
    flush();
    if (st.names.empty())
        throw std::runtime_error("No sequences found in " + fasta);
    st.start.push_back((u32)T.size());
    T.push_back(5);
    n = T.size();
    codes.alloc(n + 64);
    CK(cudaMemset(codes.p, 0, n + 64));
    CK(cudaMemcpy(codes.p, T.data(), n, cudaMemcpyHostToDevice));
}

// ------------------------------------------------------------------ suffix sorting
// Sort codes: '$' smallest: 1, A..T: 2..5, padding past the end: 0 (the empty
// suffix sorts first).  21 symbols x 3 bits per initial key.
__device__ __forceinline__ u32 sortCode(u8 c) { return c == 5 ? 1u : (c ? (u32)c + 1u : 0u); }

__global__ void kInitKeys(const u8* codes, size_t n, u64* key, u32* suf) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        u64 k = 0;
#pragma unroll
        for (int j = 0; j < 21; j++)
            k = (k << 3) | sortCode(codes[i + j]); // codes is padded with zeros
        key[i] = k;
        suf[i] = (u32)i;
    }
}
__global__ void kHeadIdx(const u64* key, size_t m, const u32* slot, u32* head) {
    for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < m; k += (size_t)gridDim.x * blockDim.x) {
        const bool h = (k == 0) || key[k] != key[k - 1];
        head[k] = h ? (slot ? slot[k] : (u32)k) + 1u : 0u; // +1: 0 = "not a head"
    }
}
// after the inclusive max scan head[k] = slot of the run's head + 1 = the new rank
__global__ void kAssign(const u32* suf, const u32* head, const u32* slot, size_t m, u32* rank, u32* sa, u8* active) {
    for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < m; k += (size_t)gridDim.x * blockDim.x) {
        const u32 s = suf[k], mySlot = slot ? slot[k] : (u32)k, r = head[k];
        rank[s] = r;
        sa[mySlot] = s;
        const bool isHead = (r == mySlot + 1);
        const bool nextIsHead = (k + 1 == m) || (head[k + 1] == (slot ? slot[k + 1] : (u32)(k + 1)) + 1);
        active[k] = (isHead && nextIsHead) ? 0 : 1;
    }
}
__global__ void kRoundKeys(const u32* suf, size_t m, const u32* rank, size_t n, u64 h, u64* key) {
    for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < m; k += (size_t)gridDim.x * blockDim.x) {
        const u64 s = suf[k];
        const u32 r2 = (s + h < n) ? rank[s + h] : 0u;
        key[k] = ((u64)rank[s] << 32) | r2;
    }
}
__global__ void kIota(u32* a, size_t m) {
    for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < m; k += (size_t)gridDim.x * blockDim.x)
        a[k] = (u32)k;
}

__global__ void kReverse(const u8* in, size_t n, u8* out) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        out[i] = in[n - 1 - i];
}

struct MaxOp {
    __host__ __device__ u32 operator()(u32 a, u32 b) const { return a > b ? a : b; }
};

// order-preserving compaction of up to 2^32 items in pieces of 2^30
static size_t compactFlagged(const u32* in, const u8* flags, size_t m, u32* out) {
    const size_t CH = (size_t)1 << 30;
    DBuf<u32> dCount(1);
    size_t tb = 0;
    CK(cub::DeviceSelect::Flagged(nullptr, tb, in, flags, out, dCount.p, (int)std::min(m, CH)));
    DBuf<u8> tmp(tb);
    size_t outN = 0;
    for (size_t o = 0; o < m; o += CH) {
        const int cnt = (int)std::min(CH, m - o);
        size_t t2 = tb;
        CK(cub::DeviceSelect::Flagged(tmp.p, t2, in + o, flags + o, out + outN, dCount.p, cnt));
        u32 c = 0;
        CK(cudaMemcpy(&c, dCount.p, 4, cudaMemcpyDeviceToHost));
        outN += c;
    }
    return outN;
}

// suffix array of codes[0,n) (codes padded with >= 32 zero bytes) into sa[n]
static void suffixSort(const u8* codes, size_t n, u32* sa) {
    DBuf<u32> rank(n);
    DBuf<u32> suf, slot;
    size_t m = 0;
    { // initial sort by the first 21 symbols
        DBuf<u64> keyA(n), keyB(n);
        DBuf<u32> sufA(n), sufB(n);
        kInitKeys<<<gridFor(n), 256>>>(codes, n, keyA.p, sufA.p);
        cub::DoubleBuffer<u64> dk(keyA.p, keyB.p);
        cub::DoubleBuffer<u32> dv(sufA.p, sufB.p);
        size_t tb = 0;
        CK(cub::DeviceRadixSort::SortPairs(nullptr, tb, dk, dv, (u64)n, 0, 63));
        DBuf<u8> tmp(tb);
        CK(cub::DeviceRadixSort::SortPairs(tmp.p, tb, dk, dv, (u64)n, 0, 63));
        tmp.release();
        u32* head = dv.Alternate();
        kHeadIdx<<<gridFor(n), 256>>>(dk.Current(), n, nullptr, head);
        (dk.Current() == keyA.p ? keyB : keyA).release();
        size_t sb = 0;
        CK(cub::DeviceScan::InclusiveScan(nullptr, sb, head, head, MaxOp(), (u64)n));
        tmp.alloc(sb);
        CK(cub::DeviceScan::InclusiveScan(tmp.p, sb, head, head, MaxOp(), (u64)n));
        tmp.release();
        (dk.Current() == keyA.p ? keyA : keyB).release();
        DBuf<u8> active(n);
        kAssign<<<gridFor(n), 256>>>(dv.Current(), head, nullptr, n, rank.p, sa, active.p);
        CK(cudaDeviceSynchronize());
        // active suffixes and their slots
        u32* sufSorted = dv.Current();
        // compaction of iota gives the slots; m = number selected
        {
            kIota<<<gridFor(n), 256>>>(head, n); // head buffer is free now: reuse as iota
            DBuf<u32> slotFull;
            // the slot list can be as long as n in the worst case; compact into the free key-sized space
            slotFull.alloc(n);
            m = compactFlagged(head, active.p, n, slotFull.p);
            slot.alloc(m + 1);
            CK(cudaMemcpy(slot.p, slotFull.p, m * 4, cudaMemcpyDeviceToDevice));
            slotFull.release();
            suf.alloc(m + 1);
            DBuf<u32> sufFull(n);
            size_t m2 = compactFlagged(sufSorted, active.p, n, sufFull.p);
            if (m2 != m)
                throw std::runtime_error("internal: compaction mismatch");
            CK(cudaMemcpy(suf.p, sufFull.p, m * 4, cudaMemcpyDeviceToDevice));
        }
        logf("  initial sort done, %zu of %zu suffixes still unsorted", m, n);
    }
    u64 h = 21;
    while (m > 0) {
        DBuf<u64> keyA(m), keyB(m);
        DBuf<u32> sufB(m), head(m);
        DBuf<u8> active(m);
        kRoundKeys<<<gridFor(m), 256>>>(suf.p, m, rank.p, n, h, keyA.p);
        cub::DoubleBuffer<u64> dk(keyA.p, keyB.p);
        cub::DoubleBuffer<u32> dv(suf.p, sufB.p);
        size_t tb = 0;
        CK(cub::DeviceRadixSort::SortPairs(nullptr, tb, dk, dv, (u64)m, 0, 64));
        DBuf<u8> tmp(tb);
        CK(cub::DeviceRadixSort::SortPairs(tmp.p, tb, dk, dv, (u64)m, 0, 64));
        kHeadIdx<<<gridFor(m), 256>>>(dk.Current(), m, slot.p, head.p);
        size_t sb = 0;
        CK(cub::DeviceScan::InclusiveScan(nullptr, sb, head.p, head.p, MaxOp(), (u64)m));
        tmp.alloc(sb);
        CK(cub::DeviceScan::InclusiveScan(tmp.p, sb, head.p, head.p, MaxOp(), (u64)m));
        kAssign<<<gridFor(m), 256>>>(dv.Current(), head.p, slot.p, m, rank.p, sa, active.p);
        CK(cudaDeviceSynchronize());
        keyA.release();
        keyB.release();
        DBuf<u32> suf2(m), slot2(m);
        const size_t mNew = compactFlagged(dv.Current(), active.p, m, suf2.p);
        const size_t mNew2 = compactFlagged(slot.p, active.p, m, slot2.p);
        if (mNew != mNew2)
            throw std::runtime_error("internal: compaction mismatch");
        std::swap(suf.p, suf2.p);
        std::swap(suf.n, suf2.n);
        std::swap(slot.p, slot2.p);
        std::swap(slot.n, slot2.n);
        h *= 2;
        logf("  h = %llu: %zu unsorted", (unsigned long long)h, mNew);
        m = mNew;
        if (h > 2 * (u64)n + 64)
            throw std::runtime_error("internal: prefix doubling did not converge");
    }
}

// ------------------------------------------------------------------ derived structures
// BWT as codes 0..4 ($,A,C,G,T) : bwt[i] = S[(sa[i] + n - 1) % n]; text code 5 -> 0
__global__ void kBwt(const u8* codes, const u32* sa, size_t n, int reversed, u8* bwt, u64* dollarPos) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const size_t s = sa[i];
        const size_t j = s ? s - 1 : n - 1;           // index in the sorted string
        const u8 c = codes[reversed ? n - 1 - j : j]; // sorted string = text or reversed text
        const u8 v = (c == 5) ? 0 : c;
        bwt[i] = v;
        if (v == 0)
            *dollarPos = i;
    }
}
// one thread per 64 positions: reference planes (cumulative) + per-word popcounts,
// gidx record planes + per-block symbol counts
__global__ void kPlanes(const u8* bwt, size_t n, size_t nWords, u64* bv /*4 per word*/, u32* pc /*4 arrays of nWords*/,
                        u64* lo, u64* hi, u32* sym /*4 arrays of nWords*/) {
    for (size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x; w < nWords; w += (size_t)gridDim.x * blockDim.x) {
        u64 p0 = 0, p1 = 0, p2 = 0, p3 = 0, l = 0, h = 0;
        const size_t base = w * 64;
        for (u32 b = 0; b < 64; b++) {
            const size_t i = base + b;
            if (i >= n)
                break;
            const u32 c = bwt[i];
            if (c == 0)
                continue; // '$': no plane bit, stored as A (code 0) in the 2-bit planes
            const u64 bit = 1ull << b;
            if (c <= 1) p0 |= bit;
            if (c <= 2) p1 |= bit;
            if (c <= 3) p2 |= bit;
            p3 |= bit;
            if ((c - 1) & 1) l |= bit;
            if ((c - 1) & 2) h |= bit;
        }
        if (bv) {
            bv[w * 4 + 0] = p0;
            bv[w * 4 + 1] = p1;
            bv[w * 4 + 2] = p2;
            bv[w * 4 + 3] = p3;
            pc[0 * nWords + w] = __popcll(p0);
            pc[1 * nWords + w] = __popcll(p1);
            pc[2 * nWords + w] = __popcll(p2);
            pc[3 * nWords + w] = __popcll(p3);
        }
        lo[w] = l;
        hi[w] = h;
        sym[0 * nWords + w] = __popcll(p0);
        sym[1 * nWords + w] = __popcll(p1 & ~p0);
        sym[2 * nWords + w] = __popcll(p2 & ~p1);
        sym[3 * nWords + w] = __popcll(p3 & ~p2);
    }
}
// rank9 counts of the interleaved bit vector (bitvec.h:329-349): per 512-position
// block and plane c: counts[blk*8 + 2c] = ones before the block, counts[.. + 1] =
// 7 x 9-bit partial sums after words 0..6 of the block (only for words that exist)
__global__ void kIntlCounts(const u32* pcEx /*exclusive scan per plane*/, const u32* pc, size_t nWords, size_t nBlocks,
                            u64* counts) {
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < nBlocks * 4; t += (size_t)gridDim.x * blockDim.x) {
        const size_t blk = t >> 2, c = t & 3, w0 = blk * 8;
        const u32* ex = pcEx + c * nWords;
        const u32* p = pc + c * nWords;
        counts[blk * 8 + 2 * c] = ex[w0];
        u64 l2 = 0, run = p[w0];
        for (u32 j = 1; j < 8 && w0 + j < nWords; j++) {
            l2 |= run << (9 * (j - 1));
            run += p[w0 + j];
        }
        counts[blk * 8 + 2 * c + 1] = l2;
    }
}
__global__ void kRecs(const u32* symEx, const u64* lo, const u64* hi, size_t nWords, size_t nRec, clb::GidxBwtRec* rec,
                      const u32* symTotal) {
    for (size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x; q < nRec; q += (size_t)gridDim.x * blockDim.x) {
        clb::GidxBwtRec r;
        if (q < nWords) {
            for (int c = 0; c < 4; c++)
                r.cnt[c] = symEx[c * nWords + q];
            r.lo = lo[q];
            r.hi = hi[q];
        } else {
            for (int c = 0; c < 4; c++)
                r.cnt[c] = symTotal[c];
            r.lo = r.hi = 0;
        }
        rec[q] = r;
    }
}
// sparse suffix array: marker bits per 64 rows + popcounts
__global__ void kSsaBits(const u32* sa, size_t n, u32 sparseMask, size_t nWords, u64* bits, u32* pc) {
    for (size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x; w < nWords; w += (size_t)gridDim.x * blockDim.x) {
        u64 v = 0;
        for (u32 b = 0; b < 64; b++) {
            const size_t i = w * 64 + b;
            if (i < n && (sa[i] & sparseMask) == 0)
                v |= 1ull << b;
        }
        bits[w] = v;
        pc[w] = __popcll(v);
    }
}
__global__ void kSsaScatter(const u32* sa, size_t n, const u64* bits, const u32* pcEx, size_t nWords, u32* sparse) {
    for (size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x; w < nWords; w += (size_t)gridDim.x * blockDim.x) {
        u64 v = bits[w];
        u32 o = pcEx[w];
        while (v) {
            const u32 b = __ffsll((long long)v) - 1;
            v &= v - 1;
            sparse[o++] = sa[w * 64 + b];
        }
    }
}
// Bitvec rank9 counts (bitvec.h:138-153): per 8 words one absolute count + 7 x 9-bit partial sums
__global__ void kRank9(const u32* pcEx, const u32* pc, size_t nWords, size_t nGroups, u64* counts) {
    for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < nGroups; g += (size_t)gridDim.x * blockDim.x) {
        const size_t w0 = g * 8;
        counts[2 * g] = pcEx[w0];
        u64 l2 = 0, run = pc[w0];
        for (u32 j = 1; j < 8 && w0 + j < nWords; j++) {
            l2 |= run << (9 * (j - 1));
            run += pc[w0 + j];
        }
        counts[2 * g + 1] = l2;
    }
}
__global__ void kSsaRecs(const u64* bits, const u32* pcEx, size_t nWords, u32 total, clb::GidxSsaRec* rec) {
    for (size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x; w <= nWords; w += (size_t)gridDim.x * blockDim.x) {
        clb::GidxSsaRec r;
        r.bits = w < nWords ? bits[w] : 0;
        r.rank = w < nWords ? pcEx[w] : total;
        r.pad = 0;
        rec[w] = r;
    }
}
// EncodedText<5> (.bwt): 3 bits per symbol, most significant bits first
__global__ void kEncode3(const u8* bwt, size_t n, size_t nWords, u64* out) {
    for (size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x; w < nWords; w += (size_t)gridDim.x * blockDim.x) {
        u64 v = 0;
        const u64 bit0 = (u64)w * 64;
        for (u64 i = bit0 / 3; i * 3 < bit0 + 64 && i < n; i++) {
            const u64 c = bwt[i];
            for (u32 q = 0; q < 3; q++) { // bit q of the symbol (0 = most significant) sits at absolute bit 3i+q
                const u64 a = i * 3 + q;
                if (a >= bit0 && a < bit0 + 64 && ((c >> (2 - q)) & 1))
                    v |= 1ull << (63 - (a - bit0));
            }
        }
        out[w] = v;
    }
}
__global__ void kPackText(const u8* codes, size_t n, size_t nWords, u32* out) {
    for (size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x; w < nWords; w += (size_t)gridDim.x * blockDim.x) {
        u32 v = 0;
        for (u32 b = 0; b < 16; b++) {
            const size_t i = w * 16 + b;
            if (i < n) {
                const u32 c = codes[i];
                if (c >= 2 && c <= 4)
                    v |= (c - 1) << (2 * b);
            }
        }
        out[w] = v;
    }
}
__global__ void kAscii(const u8* codes, size_t n, u8* out) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const u8 c = codes[i];
        out[i] = c == 1 ? 'A' : c == 2 ? 'C' : c == 3 ? 'G' : c == 4 ? 'T' : '$';
    }
}
__global__ void kHist(const u8* codes, size_t n, unsigned long long* hist) {
    __shared__ unsigned long long h[8];
    if (threadIdx.x < 8)
        h[threadIdx.x] = 0;
    __syncthreads();
    u32 loc[6] = {0, 0, 0, 0, 0, 0};
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        loc[codes[i] < 6 ? codes[i] : 0]++;
    for (int c = 0; c < 6; c++)
        if (loc[c])
            atomicAdd(&h[c], (unsigned long long)loc[c]);
    __syncthreads();
    if (threadIdx.x < 6 && h[threadIdx.x])
        atomicAdd(&hist[threadIdx.x], h[threadIdx.x]);
}

static void exclusiveSumU32(const u32* in, u32* out, size_t n) {
    size_t tb = 0;
    CK(cub::DeviceScan::ExclusiveSum(nullptr, tb, in, out, (u64)n));
    DBuf<u8> tmp(tb);
    CK(cub::DeviceScan::ExclusiveSum(tmp.p, tb, in, out, (u64)n));
}

// stream a device array into a file through a page-locked bounce buffer
struct FileOut {
    FILE* f = nullptr;
    static u8*& bounce() {
        static u8* b = nullptr;
        return b;
    }
    static const size_t BOUNCE = (size_t)128 << 20;
    explicit FileOut(const std::string& fn) {
        f = fopen(fn.c_str(), "wb");
        if (!f)
            throw std::runtime_error("Error opening file for writing: " + fn);
        if (!bounce())
            CK(cudaHostAlloc(&bounce(), BOUNCE, cudaHostAllocDefault));
    }
    ~FileOut() {
        if (f)
            fclose(f);
    }
    void host(const void* p, size_t bytes) {
        if (bytes && fwrite(p, 1, bytes, f) != bytes)
            throw std::runtime_error("write failed");
    }
    void dev(const void* d, size_t bytes) {
        for (size_t o = 0; o < bytes; o += BOUNCE) {
            const size_t c = std::min(BOUNCE, bytes - o);
            CK(cudaMemcpy(bounce(), (const u8*)d + o, c, cudaMemcpyDeviceToHost));
            host(bounce(), c);
        }
    }
    void pad(size_t align) {
        long pos = ftell(f);
        size_t r = (size_t)pos % align;
        if (r) {
            std::vector<u8> z(align - r, 0);
            host(z.data(), z.size());
        }
    }
    u64 tell() { return (u64)ftell(f); }
};

struct Options {
    std::string out, fasta, synthetic;
    u64 bp = 0, seed = 3100;
    u32 nSeq = 0, sparse = 4;
    bool writeFasta = false, refFiles = true, gidx = true;
    int device = 0;
};

// one direction: SA -> BWT -> files; returns the gidx records on the device
struct DirResult {
    DBuf<clb::GidxBwtRec> rec;
    u64 dollarPos = 0;
    size_t nRec = 0;
};

static void buildDirection(const Options& o, const DBuf<u8>& codes, size_t n, bool reversed, DirResult& R,
                           DBuf<clb::GidxSsaRec>* ssaRec, DBuf<u32>* ssa, size_t* nSsa) {
    DBuf<u8> bwt(n);
    DBuf<u64> dDollar(1);
    {
        DBuf<u32> sa(n);
        {
            DBuf<u8> srt;
            const u8* sortInput = codes.p;
            if (reversed) {
                srt.alloc(n + 64);
                CK(cudaMemset(srt.p, 0, n + 64));
                kReverse<<<gridFor(n), 256>>>(codes.p, n, srt.p);
                sortInput = srt.p;
            }
            logf("suffix sorting the %s text (%zu symbols)", reversed ? "reversed" : "forward", n);
            suffixSort(sortInput, n, sa.p);
        }
        kBwt<<<gridFor(n), 256>>>(codes.p, sa.p, n, reversed ? 1 : 0, bwt.p, dDollar.p);
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(&R.dollarPos, dDollar.p, 8, cudaMemcpyDeviceToHost));
        if (!reversed) { // sparse suffix array
            const size_t nWords = (n + 63) / 64;
            DBuf<u64> bits(nWords);
            DBuf<u32> pc(nWords + 1), pcEx(nWords + 1);
            CK(cudaMemset(pc.p, 0, (nWords + 1) * 4));
            kSsaBits<<<gridFor(nWords), 256>>>(sa.p, n, o.sparse - 1, nWords, bits.p, pc.p);
            exclusiveSumU32(pc.p, pcEx.p, nWords + 1);
            u32 total = 0;
            CK(cudaMemcpy(&total, pcEx.p + nWords, 4, cudaMemcpyDeviceToHost));
            DBuf<u32> sparse(total + 1);
            kSsaScatter<<<gridFor(nWords), 256>>>(sa.p, n, bits.p, pcEx.p, nWords, sparse.p);
            CK(cudaDeviceSynchronize());
            if (o.refFiles) {
                const std::string sfx = std::to_string(o.sparse);
                const size_t nCounts = (nWords + 7) / 4, nGroups = (nWords + 7) / 8;
                DBuf<u64> counts(nCounts + 2);
                counts.zero();
                kRank9<<<gridFor(nGroups), 256>>>(pcEx.p, pc.p, nWords, nGroups, counts.p);
                CK(cudaDeviceSynchronize());
                FileOut f(o.out + ".sa.bv." + sfx);
                u64 N = n;
                f.host(&N, 8);
                f.dev(bits.p, nWords * 8);
                f.dev(counts.p, nCounts * 8);
                FileOut f2(o.out + ".sa." + sfx);
                f2.dev(sparse.p, (size_t)total * 4);
            }
            if (ssaRec) {
                ssaRec->alloc(nWords + 1);
                kSsaRecs<<<gridFor(nWords + 1), 256>>>(bits.p, pcEx.p, nWords, total, ssaRec->p);
                CK(cudaDeviceSynchronize());
                std::swap(ssa->p, sparse.p);
                std::swap(ssa->n, sparse.n);
                *nSsa = total;
            }
        }
    }
    // planes, counts, records
    const u64 N = (u64)n + 1; // the reference's bit vectors have one position more than the text
    const size_t nWords = (size_t)((N + 63) / 64), nBlocks = (size_t)((N + 511) / 512);
    DBuf<u64> lo(nWords), hi(nWords);
    DBuf<u32> sym(4 * nWords), symEx(4 * nWords);
    {
        DBuf<u64> bv;
        DBuf<u32> pc, pcEx;
        if (o.refFiles) {
            bv.alloc(4 * nWords);
            pc.alloc(4 * nWords);
            pcEx.alloc(4 * nWords);
        }
        kPlanes<<<gridFor(nWords), 256>>>(bwt.p, n, nWords, o.refFiles ? bv.p : nullptr, pc.p, lo.p, hi.p, sym.p);
        CK(cudaDeviceSynchronize());
        if (o.refFiles) {
            for (int c = 0; c < 4; c++)
                exclusiveSumU32(pc.p + (size_t)c * nWords, pcEx.p + (size_t)c * nWords, nWords);
            DBuf<u64> counts(8 * nBlocks);
            counts.zero();
            kIntlCounts<<<gridFor(nBlocks * 4), 256>>>(pcEx.p, pc.p, nWords, nBlocks, counts.p);
            CK(cudaDeviceSynchronize());
            FileOut f(o.out + (reversed ? ".rev.brt" : ".brt"));
            f.host(&R.dollarPos, 8);
            f.host(&N, 8);
            f.dev(bv.p, 4 * nWords * 8);
            f.dev(counts.p, 8 * nBlocks * 8);
        }
    }
    if (o.refFiles && !reversed) { // .bwt
        const size_t ew = (n * 3) / 64 + 1;
        DBuf<u64> enc(ew);
        kEncode3<<<gridFor(ew), 256>>>(bwt.p, n, ew, enc.p);
        CK(cudaDeviceSynchronize());
        FileOut f(o.out + ".bwt");
        size_t ts = n, vs = ew;
        f.host(&ts, sizeof(size_t));
        f.host(&vs, sizeof(size_t));
        f.dev(enc.p, ew * 8);
    }
    if (o.gidx) {
        for (int c = 0; c < 4; c++)
            exclusiveSumU32(sym.p + (size_t)c * nWords, symEx.p + (size_t)c * nWords, nWords);
        // totals per symbol for the records past the last word
        std::vector<u32> lastEx(4), lastPc(4);
        DBuf<u32> dTot(4);
        for (int c = 0; c < 4; c++) {
            CK(cudaMemcpy(&lastEx[c], symEx.p + (size_t)c * nWords + nWords - 1, 4, cudaMemcpyDeviceToHost));
            CK(cudaMemcpy(&lastPc[c], sym.p + (size_t)c * nWords + nWords - 1, 4, cudaMemcpyDeviceToHost));
            lastEx[c] += lastPc[c];
        }
        CK(cudaMemcpy(dTot.p, lastEx.data(), 16, cudaMemcpyHostToDevice));
        R.nRec = n / 64 + 1;
        R.rec.alloc(R.nRec);
        kRecs<<<gridFor(R.nRec), 256>>>(symEx.p, lo.p, hi.p, nWords, R.nRec, R.rec.p, dTot.p);
        CK(cudaDeviceSynchronize());
    }
}

static void run(const Options& o) {
    CK(cudaSetDevice(o.device));
    DBuf<u8> codes;
    SeqTable st;
    size_t n = 0;
    if (!o.fasta.empty()) {
        readFasta(o.fasta, codes, st, n);
    } else {
        if (o.bp < 1000 || o.bp >= 0xFFFFFF00ull)
            throw std::runtime_error("--bp out of range for a 32-bit index");
        logf("generating the %s-profile synthetic reference, %llu bp, seed %llu", o.synthetic.c_str(),
             (unsigned long long)o.bp, (unsigned long long)o.seed);
        makeSynthetic(o.synthetic, o.bp, o.seed, o.nSeq ? o.nSeq : (o.synthetic == "human" ? 24u : 1u), codes, st);
        n = o.bp + 1;
    }
    if (n >= 0xFFFFFFF0ull)
        throw std::runtime_error("The size of the reference is too large for 32-bit indexes");
    // character counts
    std::vector<unsigned long long> hist(8, 0);
    {
        DBuf<unsigned long long> dh(8);
        dh.zero();
        kHist<<<gridFor(n), 256>>>(codes.p, n, dh.p);
        CK(cudaMemcpy(hist.data(), dh.p, 64, cudaMemcpyDeviceToHost));
    }
    for (int c = 1; c <= 4; c++)
        if (!hist[c])
            throw std::runtime_error("the reference must contain all of A, C, G and T");
    { // small files + text
        std::vector<u8> ascii;
        if (o.refFiles || o.writeFasta) {
            DBuf<u8> da(n);
            kAscii<<<gridFor(n), 256>>>(codes.p, n, da.p);
            CK(cudaDeviceSynchronize());
            if (o.refFiles) {
                FileOut f(o.out + ".txt.bin");
                u32 n32 = (u32)n;
                f.host(&n32, 4);
                f.dev(da.p, n);
            }
            if (o.writeFasta) {
                ascii.resize(n);
                CK(cudaMemcpy(ascii.data(), da.p, n, cudaMemcpyDeviceToHost));
            }
        }
        if (o.writeFasta) {
            FileOut f(o.out + ".fa");
            std::string buf;
            for (size_t s = 0; s + 1 < st.start.size(); s++) {
                buf = ">" + st.names[s] + " synthetic\n";
                f.host(buf.data(), buf.size());
                buf.clear();
                for (size_t p = st.start[s]; p < st.start[s + 1]; p += 80) {
                    size_t e = std::min<size_t>(p + 80, st.start[s + 1]);
                    buf.append((const char*)ascii.data() + p, e - p);
                    buf.push_back('\n');
                    if (buf.size() > (1u << 22)) {
                        f.host(buf.data(), buf.size());
                        buf.clear();
                    }
                }
                f.host(buf.data(), buf.size());
            }
        }
        if (o.refFiles) {
            std::ofstream h(o.out + ".headerSN.bin", std::ios::binary);
            for (size_t i = 0; i < st.names.size(); i++)
                h << "@SQ\tSN:" << st.names[i] << "\tLN:" << st.start[i + 1] - st.start[i] << "\n";
            std::ofstream p(o.out + ".pos", std::ios::binary);
            p.write((const char*)st.start.data(), (std::streamsize)(st.start.size() * 4));
            std::ofstream s(o.out + ".sna", std::ios::binary);
            for (auto& nm : st.names) {
                size_t len = nm.size();
                s.write((const char*)&len, sizeof(len));
                s.write(nm.data(), (std::streamsize)len);
            }
            std::ofstream fs(o.out + ".fsid", std::ios::binary);
            u32 zero = 0;
            fs.write((const char*)&zero, 4);
            std::vector<u32> cc(256, 0);
            cc[(unsigned char)'$'] = (u32)hist[5];
            cc[(unsigned char)'A'] = (u32)hist[1];
            cc[(unsigned char)'C'] = (u32)hist[2];
            cc[(unsigned char)'G'] = (u32)hist[3];
            cc[(unsigned char)'T'] = (u32)hist[4];
            std::ofstream c(o.out + ".cct", std::ios::binary);
            c.write((const char*)cc.data(), 1024);
            std::ofstream m(o.out + ".meta");
            m << 21 << "\n" << 4 << "\n" << "VANILLA" << "\n";
        }
    }
    DirResult fwd, rev;
    DBuf<clb::GidxSsaRec> ssaRec;
    DBuf<u32> ssa;
    size_t nSsa = 0;
    buildDirection(o, codes, n, false, fwd, o.gidx ? &ssaRec : nullptr, &ssa, &nSsa);
    logf("forward direction done");
    if (!o.gidx) {
        ssa.release();
        ssaRec.release();
    }
    buildDirection(o, codes, n, true, rev, nullptr, nullptr, nullptr);
    logf("reverse direction done");
    if (o.gidx) {
        const size_t tw = n / 16 + 8;
        DBuf<u32> packed(tw);
        kPackText<<<gridFor(tw), 256>>>(codes.p, n, tw, packed.p);
        CK(cudaDeviceSynchronize());
        clb::GidxHeader H;
        memset(&H, 0, sizeof(H));
        memcpy(H.magic, CLB_GIDX_MAGIC, 8);
        H.version = CLB_GIDX_VERSION;
        H.textLength = (u32)n;
        H.saSparse = o.sparse;
        H.C[0] = 0;
        H.C[1] = (u32)hist[5];
        H.C[2] = H.C[1] + (u32)hist[1];
        H.C[3] = H.C[2] + (u32)hist[2];
        H.C[4] = H.C[3] + (u32)hist[3];
        H.dollarFwd = (u32)fwd.dollarPos;
        H.dollarRev = (u32)rev.dollarPos;
        H.numSeq = (u32)st.names.size();
        H.nRec = fwd.nRec;
        H.nSsaRec = ssaRec.n;
        H.nSsa = nSsa;
        H.nTextWords = tw;
        std::string names;
        for (auto& nm : st.names) {
            names += nm;
            names.push_back('\0');
        }
        H.namesBytes = names.size();
        FileOut f(o.out + ".gidx.tmp");
        f.host(&H, sizeof(H)); // rewritten below with the section offsets
        f.pad(4096);
        H.offSeqStart = f.tell();
        f.host(st.start.data(), st.start.size() * 4);
        H.offNames = f.tell();
        f.host(names.data(), names.size());
        f.pad(4096);
        H.offFwd = f.tell();
        f.dev(fwd.rec.p, fwd.nRec * sizeof(clb::GidxBwtRec));
        f.pad(4096);
        H.offRev = f.tell();
        f.dev(rev.rec.p, rev.nRec * sizeof(clb::GidxBwtRec));
        f.pad(4096);
        H.offSsaRec = f.tell();
        f.dev(ssaRec.p, ssaRec.n * sizeof(clb::GidxSsaRec));
        f.pad(4096);
        H.offSsa = f.tell();
        f.dev(ssa.p, nSsa * 4);
        f.pad(4096);
        H.offText = f.tell();
        f.dev(packed.p, tw * 4);
        H.fileBytes = f.tell();
        fseek(f.f, 0, SEEK_SET);
        f.host(&H, sizeof(H));
        fclose(f.f);
        f.f = nullptr;
        if (rename((o.out + ".gidx.tmp").c_str(), (o.out + ".gidx").c_str()) != 0)
            throw std::runtime_error("cannot rename the .gidx file");
    }
    logf("done: %zu symbols, %zu sequences", n, st.names.size());
}

int main(int argc, char** argv) {
    g_t0 = nowS();
    Options o;
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        auto val = [&]() -> std::string {
            if (i + 1 >= argc) {
                fprintf(stderr, "missing value for %s\n", a.c_str());
                exit(2);
            }
            return argv[++i];
        };
        if (a == "--out") o.out = val();
        else if (a == "--fasta") o.fasta = val();
        else if (a == "--synthetic") o.synthetic = val();
        else if (a == "--bp") o.bp = strtoull(val().c_str(), nullptr, 10);
        else if (a == "--seed") o.seed = strtoull(val().c_str(), nullptr, 10);
        else if (a == "--nseq") o.nSeq = (u32)atoi(val().c_str());
        else if (a == "--sparse") o.sparse = (u32)atoi(val().c_str());
        else if (a == "--write-fasta") o.writeFasta = true;
        else if (a == "--no-ref-files") o.refFiles = false;
        else if (a == "--no-gidx") o.gidx = false;
        else if (a == "--device") o.device = atoi(val().c_str());
        else if (a == "--quiet") g_verbose = false;
        else {
            fprintf(stderr, "unknown option %s\n", a.c_str());
            return 2;
        }
    }
    if (o.out.empty() || (o.fasta.empty() == o.synthetic.empty()) || !o.sparse || (o.sparse & (o.sparse - 1)) || o.sparse > 256) {
        fprintf(stderr, "usage: clb_gpu_build --out BASE (--fasta F | --synthetic human|ecoli|none --bp N [--seed S] [--nseq K]) "
                        "[--sparse 4] [--write-fasta] [--no-ref-files] [--no-gidx] [--device D]\n");
        return 2;
    }
    try {
        run(o);
    } catch (std::exception& e) {
        fprintf(stderr, "clb_gpu_build: %s\n", e.what());
        return 1;
    }
    return 0;
}
