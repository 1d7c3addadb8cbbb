// columba_b200 -- common types for the device code.
//
// Every per-task routine of the hot path is written as a CLB_HD function so
// that the same source compiles (a) with nvcc for sm_100a -- the product --
// and (b) with g++ inside tests/hostemu/ where the kernels' task functions are
// run in a plain loop, so the kernel logic can be unit-tested on a machine
// without a GPU.  (b) is a test harness only: the shipped library
// (libcolumba_b200.so) contains no CPU execution path for these routines.
#pragma once
#include <stdint.h>
#include <stddef.h>

#if defined(__CUDACC__)
#define CLB_HD __host__ __device__ __forceinline__
#define CLB_HDN __host__ __device__ __noinline__
#else
#define CLB_HD inline
#define CLB_HDN
#endif

namespace clb {

typedef uint8_t u8;
typedef uint16_t u16;
typedef uint32_t u32;
typedef uint64_t u64;
typedef int32_t i32;
typedef unsigned __int128 u128;

CLB_HD int popc64(u64 x) {
#if defined(__CUDA_ARCH__)
    return __popcll(x);
#else
    return __builtin_popcountll(x);
#endif
}
CLB_HD int popc32(u32 x) {
#if defined(__CUDA_ARCH__)
    return __popc(x);
#else
    return __builtin_popcount(x);
#endif
}
CLB_HD void prefetchL2(const void* p) {
#if defined(__CUDA_ARCH__)
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
#else
    (void)p;
#endif
}
CLB_HD int msb32(u32 x) { // index of the highest set bit, x != 0
#if defined(__CUDA_ARCH__)
    return 31 - __clz((int)x);
#else
    return 31 - __builtin_clz(x);
#endif
}
CLB_HD int msb64(u64 x) { // x != 0
#if defined(__CUDA_ARCH__)
    return 63 - __clzll((long long)x);
#else
    return 63 - __builtin_clzll(x);
#endif
}
CLB_HD int ctz64(u64 x) { // x != 0
#if defined(__CUDA_ARCH__)
    return __ffsll((long long)x) - 1;
#else
    return __builtin_ctzll(x);
#endif
}
CLB_HD int popcw(u64 x) { return popc64(x); }
CLB_HD int popcw(u128 x) { return popc64((u64)x) + popc64((u64)(x >> 64)); }

// atomics: real ones on the device; the host emulation is single threaded
template <typename T> CLB_HD T atomAdd(T* p, T v) {
#if defined(__CUDA_ARCH__)
    return atomicAdd(p, v);
#else
    T o = *p;
    *p = o + v;
    return o;
#endif
}
CLB_HD u64 atomAdd64(u64* p, u64 v) {
#if defined(__CUDA_ARCH__)
    return (u64)atomicAdd((unsigned long long*)p, (unsigned long long)v);
#else
    u64 o = *p;
    *p = o + v;
    return o;
#endif
}
CLB_HD void atomOr(u32* p, u32 v) {
#if defined(__CUDA_ARCH__)
    atomicOr(p, v);
#else
    *p |= v;
#endif
}
CLB_HD void atomMaxU32(u32* p, u32 v) {
#if defined(__CUDA_ARCH__)
    atomicMax(p, v);
#else
    if (v > *p)
        *p = v;
#endif
}

// warp-level votes over the lanes in `mask` (the lanes that run a task side
// by side); the host emulation runs one task at a time
CLB_HD unsigned warpBallot(unsigned mask, bool p) {
#if defined(__CUDA_ARCH__)
    return __ballot_sync(mask, p);
#else
    (void)mask;
    return p ? 1u : 0u;
#endif
}
CLB_HD bool warpAny(unsigned mask, bool p) {
#if defined(__CUDA_ARCH__)
    return __any_sync(mask, p) != 0;
#else
    (void)mask;
    return p;
#endif
}
CLB_HD void warpSync(unsigned mask) {
#if defined(__CUDA_ARCH__)
    __syncwarp(mask);
#else
    (void)mask;
#endif
}

// IEEE double ops without FMA contraction (partition boundaries must truncate
// exactly as the reference's x86-64 code does: searchstrategy.cpp:194-238,396)
CLB_HD double dmul(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __dmul_rn(a, b);
#else
    volatile double r = a * b;
    return r;
#endif
}
CLB_HD double dsub(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __dsub_rn(a, b);
#else
    volatile double r = a - b;
    return r;
#endif
}
CLB_HD double ddiv(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __ddiv_rn(a, b);
#else
    volatile double r = a / b;
    return r;
#endif
}

template <typename T> CLB_HD T tmin(T a, T b) { return a < b ? a : b; }
template <typename T> CLB_HD T tmax(T a, T b) { return a > b ? a : b; }

// reference: indexhelpers.h:63-127 (Range), :1117-1243 (SARangePair)
struct Range {
    u32 b, e;
    CLB_HD bool empty() const { return e <= b; }
    CLB_HD u32 width() const { return e <= b ? 0u : e - b; }
};
struct RangePair {
    Range sa, rev;
    CLB_HD bool empty() const { return sa.empty(); }
    CLB_HD u32 width() const { return sa.width(); }
};
CLB_HD RangePair mkPair(u32 a, u32 b, u32 c, u32 d) {
    RangePair r;
    r.sa.b = a; r.sa.e = b; r.rev.b = c; r.rev.e = d;
    return r;
}
CLB_HD RangePair emptyPair() { return mkPair(0, 0, 0, 0); }

// reference: indexhelpers.h:1848-1883 (Counters, Vanilla build: 17 counters)
enum Ctr {
    NODE_COUNTER, TOTAL_REPORTED_POSITIONS, CIGARS_IN_INDEX, IN_TEXT_STARTED,
    ABORTED_IN_TEXT_VERIF, CIGARS_IN_TEXT_VERIFICATION, IMMEDIATE_SWITCH,
    SEARCH_STARTED, DROPPED_UNIQUE_MATCHES, NUMBER_OF_READS,
    TOTAL_UNIQUE_MATCHES, MAPPED_READS, TOTAL_UNIQUE_PAIRS, MAPPED_PAIRS,
    DISCORDANTLY_MAPPED_PAIRS, MAPPED_HALF_PAIRS, UNPAIRED_BUT_MAPPED_PAIRS,
    N_COUNTERS
};
// extra device statistics kept next to the reference's counters
enum Stat {
    ST_RANK_LOOKUPS,  // 32-byte BWT record reads (the roofline unit)
    ST_LF_STEPS,      // LF steps of locate
    ST_VERIFY_ROWS,   // bit-parallel rows computed by in-text verification
    ST_SEARCH_ROWS,   // bit-parallel rows computed by the index search
    ST_SEARCH_TASKS,  // (read,strand,search) tasks launched
    ST_CANDIDATES,    // in-text verification candidates
    ST_FM_OCCS,       // index occurrences reported by the searches
    ST_TEXT_OCCS,     // text occurrences before the per-read filter
    ST_SEARCH_LOOKUPS, // the part of ST_RANK_LOOKUPS done by the search kernel (rest: partitioning)
    ST_UNIQUE_CANDS,   // candidates left after collapsing equal verification windows (the ones that are verified)
    ST_HEAVY_TASKS,    // search tasks re-run by a whole CTA with large scratch
    N_STATS
};

// error bits raised by kernels (checked by the host after each batch)
enum ErrBit {
    ERR_CAND_OVERFLOW = 1, ERR_FMOCC_OVERFLOW = 2, ERR_TEXTOCC_OVERFLOW = 4,
    ERR_STACK_OVERFLOW = 8, ERR_CLUSTER_OVERFLOW = 16, ERR_TASK_OVERFLOW = 32,
    ERR_SHORT_READ = 64, ERR_ROWS_OVERFLOW = 128, ERR_OUT_OVERFLOW = 256,
    ERR_INTERNAL = 512
};

// list overflows: the attempt is void (the host grows the list and runs the batch again), and the tail of an
// overflowed list holds unwritten records, so the kernels after the one that raised the bit do nothing
enum { ERR_ABORT_MASK = ERR_CAND_OVERFLOW | ERR_FMOCC_OVERFLOW | ERR_TEXTOCC_OVERFLOW | ERR_OUT_OVERFLOW |
                        ERR_TASK_OVERFLOW | ERR_ROWS_OVERFLOW | ERR_STACK_OVERFLOW };

enum { DIR_FWD = 0, DIR_BWD = 1 };
enum { PART_UNIFORM = 0, PART_STATIC = 1, PART_DYNAMIC = 2 };
enum { METRIC_HAMMING = 0, METRIC_EDIT = 1 };

// per-thread counter block, flushed with one atomic per touched counter
struct LocalCtr {
    u64 c[N_COUNTERS];
    u64 s[N_STATS];
    CLB_HD void clear() {
        for (int i = 0; i < N_COUNTERS; i++) c[i] = 0;
        for (int i = 0; i < N_STATS; i++) s[i] = 0;
    }
    CLB_HD void flush(u64* gc, u64* gs) {
        for (int i = 0; i < N_COUNTERS; i++)
            if (c[i]) atomAdd64(gc + i, c[i]);
        for (int i = 0; i < N_STATS; i++)
            if (s[i]) atomAdd64(gs + i, s[i]);
    }
};

} // namespace clb
