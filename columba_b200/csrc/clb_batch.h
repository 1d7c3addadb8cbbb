// columba_b200 -- per-batch buffers in HBM shared by all stages of the path.
//
// One batch = R reads mapped at one maximal distance k.  t = 2*read + strand
// indexes a (read, strand) "sequence" (strand 1 = reverse complement), which is
// the unit the reference maps with one matchWithSearches call
// (src/searchstrategy.cpp:425-493).  Stages append variable-length records to
// global lists with one atomic per record; capacities are set by the host and
// an overflow raises an error bit so the host can re-run with larger lists.
#pragma once
#include "clb_index.h"
#include "clb_scheme.h"

namespace clb {

// per (read,strand) partitioning result: parts + exact ranges + chosen scheme
// (the form the search kernel works with; in HBM it is kept as PartStore)
struct PartInfo {
    u16 begin[MAXP];
    u16 end[MAXP];
    RangePair emr[MAXP];
    u8 scheme; // index into DevStrategy::sc
    u8 valid;  // 0: read too short for partitioning
    u16 len;
};
// header of a (read,strand) in the PartStore
struct PartHdr {
    u16 len;
    u16 matched;  // characters matched by the seeding (dynamic partitioning)
    u8 scheme;
    u8 valid;
    u8 stepping;  // 1: the greedy steps of the dynamic partitioning are still to be done
    u8 pad;
};
// Partitioning results of all T (read,strand) sequences of a batch in HBM,
// structure of arrays over t: the kernels of stage 1 run one thread per t, so
// the lanes of a warp read and write consecutive elements (coalesced) instead
// of one 328-byte record each.
struct PartStore {
    u32* be;        // [part * T + t] : begin | end << 16
    RangePair* emr; // [part * T + t]
    PartHdr* hdr;   // [t]
    u32 T;
    static CLB_HD size_t bytes(u32 T, u32 P) {
        return (size_t)T * P * (sizeof(u32) + sizeof(RangePair)) + (size_t)T * sizeof(PartHdr) + 64;
    }
    CLB_HD PartInfo load(u32 t, u32 P) const {
        PartInfo pi;
        const PartHdr h = hdr[t];
        pi.scheme = h.scheme;
        pi.valid = h.valid;
        pi.len = h.len;
        for (u32 i = 0; i < P; i++) {
            const u32 v = be[(size_t)i * T + t];
            pi.begin[i] = (u16)v;
            pi.end[i] = (u16)(v >> 16);
            pi.emr[i] = emr[(size_t)i * T + t];
        }
        return pi;
    }
};

// one search of the chosen scheme for one (read,strand)
struct SearchTask {
    u32 t;
    u32 search;
};

// candidate for in-text verification: one SA row + how to turn its text
// position into the start of the verification window
struct Cand {
    u32 t;
    u32 row;      // SA row
    u32 add;      // shift added to SA[row] (u32 wrap like the reference's length_t)
    u32 sub;      // startDiff / lengthBefore: start = sum >= sub ? sum - sub : 0
    u32 order;    // reference emission order within t (tie-break of the filter)
    u8 fixedStart;
    u8 maxED, minED;
    u8 pad;
};

// occurrence in the index (an SA range), indexhelpers.h:1353-1535
struct FMOccRec {
    u32 t;
    Range sa;
    u32 depth;
    u16 dist;
    u16 shift;
    u32 order;
};

// occurrence in the text before the per-read filter, indexhelpers.h:289-930
struct TextOccRec {
    u32 t;
    u32 b, e;
    u32 order;  // emission order of the candidate / index of the FM occurrence
    u32 order2; // cluster centre within the candidate / row within the SA range
    u8 dist;
    u8 ncig;    // number of CIGAR runs (0 = none yet)
    u8 cls;     // 0: found by in-text verification, 1: expanded from an FM occurrence
    u8 pad;
};

// final occurrence handed back to the host
struct OutOcc {
    u32 read;
    u32 seqID;
    u32 pos;        // 0-based begin within the sequence
    u32 end;        // end within the sequence
    u32 indexBegin; // begin in the concatenated text
    u8 strand;
    u8 dist;
    u8 ncig;
    u8 trimmed;     // FOUND_WITH_TRIMMING
};

// alignment of an occurrence after clipping at a sequence boundary (stage 6)
struct TrimRec {
    u32 b, e, seqID;
    u8 dist, ncig, pad[2];
};

template <typename T> struct DevList {
    T* data;
    u32* count; // device counter
    u32 cap;
    // slots for the lanes of a warp that happen to arrive together come from one
    // atomic (every list has ONE counter: millions of single atomics on it would
    // serialise in the L2)
    CLB_HD u32 grab(u32 n) const {
#if defined(__CUDA_ARCH__)
        const unsigned m = __match_any_sync(__activemask(), (unsigned long long)count);
        const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
        u32 off = 0, total = 0;
        for (unsigned mm = m; mm; mm &= mm - 1) {
            const int b = __ffs(mm) - 1;
            const u32 v = __shfl_sync(m, n, b);
            total += v;
            off += (b < lane) ? v : 0u;
        }
        u32 base = 0;
        if (lane == leader && total)
            base = atomicAdd(count, total);
        base = __shfl_sync(m, base, leader);
        return base + off;
#else
        u32 o = *count;
        *count = o + n;
        return o;
#endif
    }
    CLB_HD u32 push(const T& v, u32* err, u32 errBit) const {
        u32 i = grab(1u);
        if (i >= cap) {
            atomOr(err, errBit);
            return 0xFFFFFFFFu;
        }
        data[i] = v;
        return i;
    }
    // reserve n slots for every lane in `act` (all of them call this together; the
    // lanes outside act are the top lanes of the warp) with one atomic per warp
    CLB_HD u32 warpReserve(u32 n, unsigned act, u32* err, u32 errBit) const {
#if defined(__CUDA_ARCH__)
        const u32 lane = threadIdx.x & 31;
        u32 incl = n;
        for (u32 o = 1; o < 32; o <<= 1) {
            const u32 v = __shfl_up_sync(act, incl, o);
            if (lane >= o)
                incl += v;
        }
        const int last = 31 - __clz((int)act);
        u32 base = 0;
        if ((int)lane == last && incl)
            base = atomicAdd(count, incl);
        base = __shfl_sync(act, base, last);
        const u32 i = base + incl - n;
        if (i + n > cap || i + n < i) {
            if (n)
                atomOr(err, errBit);
            return 0xFFFFFFFFu;
        }
        return i;
#else
        (void)act;
        return reserve(n, err, errBit);
#endif
    }
    // reserve n consecutive slots; returns first index or 0xFFFFFFFF
    CLB_HD u32 reserve(u32 n, u32* err, u32 errBit) const {
        u32 i = grab(n);
        if (i + n > cap || i + n < i) {
            atomOr(err, errBit);
            return 0xFFFFFFFFu;
        }
        return i;
    }
};

struct Batch {
    // inputs
    const u8* readsAscii;  // concatenated reads
    const u32* readOff;    // R+1 offsets
    u32 R;
    u32 maxLen;            // longest read of the batch
    u32 k;                 // maximal distance
    u32 minD;              // minimal distance (BEST mode strata)
    u32 trimMax;           // largest distance accepted when clipping at a sequence boundary (findSeqName's largestStratum)
    u32 metric;            // METRIC_*
    u32 partStrategy;      // PART_*
    u32 cigW;              // CIGAR slots per occurrence (2k+2)
    u32 noCigarFlag;       // -nC: do not keep CIGAR strings
    u32 grpShift;          // 1: occurrences are grouped and filtered per read (ALL mode, searchstrategy.cpp:495-535);
                           // 0: per (read, strand), as BEST mode maps and filters each strand (:777-811)
    u32 G;                 // number of groups: R (grpShift 1) or 2R
    u32 noFilter;          // exact stratum (indexinterface.cpp:947-1014): no unique, no redundancy filter
    // derived per (read,strand)
    u8* codes;             // 2R x seqStride codes (0..3 ACGT, 4 other)
    u32 seqStride;
    u64* planes;           // 2R x 2 dirs x 4 chars x nw words
    u32 nw;
    PartStore parts;
    // lists
    DevList<SearchTask> tasks;
    DevList<u32> heavy;    // tasks whose frontier did not fit the warp's scratch (second search launch)
    DevList<Cand> cands;
    DevList<u32> extra;    // candidates with more than one cluster centre (second verification pass)
    u32* candStart;        // cands.cap : start of the verification window in the text (edit distance)
    u32* candMult;         // cands.cap : number of candidates a unique candidate stands for
    DevList<u32> uniq;     // candidates with distinct (sequence, window): the ones that are verified
    DevList<u32> surv;     // unique candidates that the screening pass could not settle (full verification)
    u32 verifyRowsCap;     // rows the verification scratch is sized for
    DevList<FMOccRec> fmOccs;
    DevList<TextOccRec> textOccs;
    u16* textCig;          // textOccs.cap x cigW
    // grouping by read (stages 4-6): keys (group << 32 | begin) and list indices, sorted by the backend's
    // radix sort (ping-pong buffers; sortRes = the pair that holds the sorted result).  The other pair is
    // scratch afterwards: its values hold the widths / keep flags, its keys the prefix sums.
    u64* sortKey[2];
    u32* sortVal[2];
    u32 sortRes;
    u32* fmSlots;   // [0] first slot of the expanded index occurrences in textOccs, [1] their number
    u16* fmCig;     // fmOccs.cap x cigW : shared CIGAR of the occurrences with more than CIGAR_THRESHOLD rows
    u8* fmNcig;     // fmOccs.cap
    DevList<u32> trimJobs;      // sorted slots of kept occurrences that cross a sequence end
    TrimRec* trimRecs;          // per job: the clipped alignment
    u16* trimCig;               // per job x cigW
    u32* trimOf;                // per sorted slot: its job
    u8* readStatus; // R : 1 = too short to be partitioned (no search was run for it)
    u32* outCount;  // G+1 : surviving occurrences per group
    u32* outStart;  // G+1 : first output slot of the group (outStart[G] = number of outputs)
    DevList<OutOcc> out;
    u16* outCig;    // out.cap x cigW
    DevList<u32> cigJobs;  // indices into out[] whose CIGAR is still to be computed (index occurrences)
    // statistics
    u64* counters;  // N_COUNTERS
    u64* stats;     // N_STATS
    u32* err;
};

CLB_HD u32 groupOf(const Batch& B, u32 t) { return t >> B.grpShift; }
CLB_HD const u8* seqCodes(const Batch& B, u32 t) { return B.codes + (size_t)t * B.seqStride; }
CLB_HD u32 seqLen(const Batch& B, u32 t) { return B.readOff[(t >> 1) + 1] - B.readOff[t >> 1]; }
// planes of sequence t in direction d (DIR_FWD: as is, DIR_BWD: reversed)
CLB_HD const u64* seqPlanes(const Batch& B, u32 t, u32 d) { return B.planes + ((size_t)t * 2 + d) * 4 * B.nw; }

} // namespace clb
