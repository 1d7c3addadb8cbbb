// columba_b200 -- host-side orchestration of one batch: buffer management and
// the sequence of kernel launches.  Templated on the backend (BE) that owns
// device memory and launches the task functors of clb_kernels.h.
//
// BE must provide:
//   void* alloc(size_t), void free(void*), void zero(void*, size_t),
//   void* hostAlloc(size_t), void hostFree(void*)  (result arrays on the host),
//   void d2hAsync(dst, src, n) (completed by sync()),
//   void sortPairs(u64* k0, u32* v0, u64* k1, u32* v1, u32 nMax, const u32* nDev, u32 keyBits): sorts the first
//   min(nMax, *nDev) (key, value) pairs by the low keyBits bits of the keys; static u32 sortResult(keyBits) tells
//   which of the two buffer pairs holds the result,
//   void copyBegin(), void copySlice(dst, src, n, slice), void copyWait(slice)  (host-to-device copies that
//   overlap the kernels: slice s is usable by work launched after copyWait(s)),
//   void h2d(void* dst, const void* src, size_t), void d2h(void* dst, const void* src, size_t),
//   void sync(), u32 workers(int kind) (number of workers for scratch slabs),
//   u32 fallbackWorkers() (workers that need slabs instead of shared memory; 0 on the GPU),
//   template<class K> void launch(const K&, u32 nMax, const u32* nDev, int kind),
//   void exclusiveScan(const u32* in, u32* out, u32 nMax, const u32* nDev)   (n = min(nMax, *nDev); out[n] = total),
//   void mark(int stage) (timing hook)
#pragma once
#include "clb_kernels.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <chrono>
#include <stdexcept>
#include <string>
#include <vector>

namespace clb {

struct BackendError : public std::runtime_error {
    explicit BackendError(const std::string& m) : std::runtime_error(m) {}
};

enum { WK_LIGHT = 0, WK_SEARCH = 1, WK_VERIFY = 2, WK_FILTER = 3, WK_SEARCH_HEAVY = 4 };
enum { STG_PREP, STG_PARTITION, STG_SEARCH, STG_VERIFY, STG_FM, STG_TEXT, N_STAGES };

template <typename BE> struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    void ensure(BE& be, size_t n) {
        if (n <= bytes)
            return;
        // released first and forgotten before the new allocation: alloc() throws when the device is full
        release(be);
        // head-room so that batches of slightly different size do not re-allocate: a quarter, at most 256 MB
        const size_t want = n + tmin<size_t>(n / 4, (size_t)256 << 20) + 256;
        p = be.alloc(want);
        bytes = want;
    }
    void release(BE& be) {
        if (p)
            be.free(p);
        p = nullptr;
        bytes = 0;
    }
};

struct MapOptions {
    u32 k = 0;
    u32 minD = 0;
    u32 metric = METRIC_EDIT;
    u32 partStrategy = PART_DYNAMIC;
    u32 noCigar = 0;
    u32 trimMax = 0; // largest distance of an alignment clipped at a sequence boundary (0: k); BEST mode: the cut-off
    u32 perStrand = 0; // 1: group and filter the occurrences per (read, strand) instead of per read (BEST mode)
};

// host array that the backend allocates (page-locked on the GPU, so that the
// device-to-host copies of the results run at full PCIe speed)
template <typename BE, typename T> struct HostBuf {
    T* p = nullptr;
    size_t n = 0, cap = 0;
    void resize(BE& be, size_t want) {
        if (want > cap) {
            if (p)
                be.hostFree(p);
            cap = want + want / 4 + 1024;
            p = (T*)be.hostAlloc(cap * sizeof(T));
        }
        n = want;
    }
    void release(BE& be) {
        if (p)
            be.hostFree(p);
        p = nullptr;
        n = cap = 0;
    }
    size_t size() const { return n; }
    T* data() { return p; }
};

template <typename BE> struct MapResult {
    // host copies
    HostBuf<BE, OutOcc> occ;
    HostBuf<BE, u16> cig; // occ.size() x cigW
    HostBuf<BE, u32> outStart, outCount; // per read
    HostBuf<BE, u8> readStatus;          // per read
    u32 nShort = 0;
    u32 cigW = 0;
    u64 counters[N_COUNTERS];
    u64 stats[N_STATS];
    u32 err = 0;
    void release(BE& be) {
        occ.release(be);
        cig.release(be);
        outStart.release(be);
        outCount.release(be);
        readStatus.release(be);
    }
};

template <typename BE> struct Pipeline {
    BE& be;
    IndexView ix;
    DevBuf<BE> bReads, bOff, bCodes, bPlanes, bParts, bTasks, bCands, bCandAux, bExtra, bFm, bTx, bTxCig, bOut, bOutCig, bCigJobs, bPerRead, bSort, bFmCig, bTrim,
        bMisc, bSlabPart, bSlabSearch, bSlabVerify, bSlabVerifyWide, bSlabFilter, bStrategy;
    // list capacities (per read multipliers, grown after an overflow)
    double candPerRead = 24, fmPerRead = 4, txPerRead = 24, outPerRead = 8;
    size_t searchSlabBytes = (size_t)1 << 20, heavySlabBytes = (size_t)16 << 20; // 592 heavy groups: 9.5 GB; x4 on overflow // scratch per warp of the search
    u32 minCap = 1 << 16;
    bool dedupPays = true;   // the last batch's candidate de-duplication removed at least 3 %
    u32 dedupProbe = 0;      // batches since dedup was last tried (0: try it in the next batch)
    bool dedupOn = true;     // de-duplication was on in the batch in flight
    std::string lastError;
    u32 rankPrefetch = 1;    // the ranking passes of the search prefetch one level ahead (COLUMBA_B200_RANK_PREFETCH=0: off; 152.6 -> 151.4 ms)
    u32 heavyWidth = 0;      // search tasks that start from a wider SA range skip the first launch (COLUMBA_B200_HEAVY_WIDTH)
    bool verify32 = true;    // the same for the full verification (COLUMBA_B200_VERIFY32=0)
    bool screen32 = true;    // screening pass on a 32-bit band where it fits (COLUMBA_B200_SCREEN32=0: 64-bit window)

    Pipeline(BE& b) : be(b) {
        // test hook: small search scratch hands most tasks of a small batch to the heavy launch
        if (const char* e = getenv("COLUMBA_B200_SEARCH_SLAB"))
            searchSlabBytes = tmax<size_t>(4096, (size_t)atol(e));
        if (const char* e = getenv("COLUMBA_B200_RANK_PREFETCH"))
            rankPrefetch = atoi(e) != 0 ? 1u : 0u;
        if (const char* e = getenv("COLUMBA_B200_HEAVY_WIDTH"))
            heavyWidth = (u32)atol(e);
        if (const char* e = getenv("COLUMBA_B200_VERIFY32"))
            verify32 = atoi(e) != 0;
        if (const char* e = getenv("COLUMBA_B200_SCREEN32"))
            screen32 = atoi(e) != 0;
        // test hook: small lists make the overflow / re-run path run on small batches
        if (const char* e = getenv("COLUMBA_B200_MIN_LIST_CAP")) {
            minCap = tmax<u32>(16u, (u32)atoi(e));
            candPerRead = fmPerRead = txPerRead = outPerRead = 0;
        }
    }
    ~Pipeline() {
        DevBuf<BE>* all[] = {&bReads, &bOff, &bCodes, &bPlanes, &bParts, &bTasks, &bCands, &bCandAux, &bExtra, &bFm, &bTx, &bTxCig, &bCigJobs, &bOut,
                             &bOutCig, &bPerRead, &bSort, &bFmCig, &bTrim, &bMisc, &bSlabPart, &bSlabSearch, &bSlabVerify, &bSlabVerifyWide, &bSlabFilter,
                             &bStrategy};
        for (auto* d : all)
            d->release(be);
    }

    // misc block layout: counters[N_COUNTERS] stats[N_STATS] (u64) then u32s
    struct Misc {
        u64 counters[N_COUNTERS];
        u64 stats[N_STATS];
        u32 err;
        u32 nTasks, nCands, nFm, nTx, nOut, nExtra, nCigJobs, nUniq, nHeavy, nTrim, nSurv;
        u32 fmSlots[2];
        u32 nT, nR; // 2R and R as device-side task counts
        u32 errSearch; // err without the candidate / index-occurrence overflow bits: what the search stage aborts on
    };

    // longest read of a batch whose offsets are on the device
    u32 maxReadLen(const u32* dOff, u32 R) {
        bMisc.ensure(be, sizeof(Misc));
        be.zero(bMisc.p, sizeof(Misc));
        Misc* M = (Misc*)bMisc.p;
        be.setAbortWord(nullptr);
        MaxLenKernel k{dOff, &M->nT};
        be.launch(k, R, nullptr, WK_LIGHT);
        u32 v = 0;
        be.d2h(&v, &M->nT, sizeof(u32));
        return v;
    }

    void uploadStrategy(const DevStrategy& st) {
        bStrategy.ensure(be, sizeof(DevStrategy));
        be.h2d(bStrategy.p, &st, sizeof(DevStrategy));
    }

    // reads: device pointers when readsOnDevice, else host pointers that are
    // copied in.  Returns 0 or an error bit mask (after growing the lists the
    // caller may simply call again).
    u32 runOnce(const DevStrategy& st, const u8* reads, const u32* off, u32 R, u32 maxLen, size_t totalBytes,
                bool readsOnDevice, const MapOptions& opt, MapResult<BE>* res, bool fetch) {
        be.restartStages();
        Batch B;
        B.R = R;
        B.maxLen = maxLen;
        B.k = opt.k;
        B.minD = opt.minD;
        B.metric = opt.metric;
        B.partStrategy = opt.partStrategy;
        const u32 kk = tmax<u32>(opt.k, opt.trimMax);
        B.trimMax = kk;
        B.cigW = 2 * kk + 2;
        B.noCigarFlag = opt.noCigar;
        B.grpShift = opt.perStrand ? 0u : 1u;
        B.G = opt.perStrand ? 2 * R : R;
        B.noFilter = (opt.k == 0) ? 1u : 0u;
        const u32 G = B.G;
        const u32 T = 2 * R;
        // host reads go up in slices on the copy stream; the preparation kernel of a
        // slice starts as soon as its bytes have arrived (see stage 0 below)
        const u32 nSlices = readsOnDevice ? 1u : tmin<u32>(8u, R / 65536u + 1u);
        u32 sliceEnd[8];
        if (readsOnDevice) {
            B.readsAscii = reads;
            B.readOff = off;
            sliceEnd[0] = R;
        } else {
            bReads.ensure(be, totalBytes + 16);
            bOff.ensure(be, (size_t)(R + 1) * 4);
            be.h2d(bOff.p, off, (size_t)(R + 1) * 4);
            be.copyBegin();
            for (u32 s = 0; s < nSlices; s++) {
                const u32 r0 = (u32)((u64)R * s / nSlices), r1 = (u32)((u64)R * (s + 1) / nSlices);
                sliceEnd[s] = r1;
                be.copySlice((u8*)bReads.p + off[r0], reads + off[r0], (size_t)off[r1] - off[r0], s);
            }
            B.readsAscii = (const u8*)bReads.p;
            B.readOff = (const u32*)bOff.p;
        }
        B.seqStride = (maxLen + 63) & ~63u;
        B.nw = (maxLen + 63) / 64 + 1;
        bCodes.ensure(be, (size_t)T * B.seqStride);
        bPlanes.ensure(be, (size_t)T * 2 * 4 * B.nw * 8);
        bParts.ensure(be, PartStore::bytes(T, st.P));
        B.codes = (u8*)bCodes.p;
        B.planes = (u64*)bPlanes.p;
        B.parts.T = T;
        B.parts.emr = (RangePair*)bParts.p;
        B.parts.be = (u32*)(B.parts.emr + (size_t)T * st.P);
        B.parts.hdr = (PartHdr*)(B.parts.be + (size_t)T * st.P);
        const u32 P = st.P;
        u32 maxSearches = 0;
        for (u32 i = 0; i < st.nSchemes; i++)
            maxSearches = tmax<u32>(maxSearches, st.sc[i].nSearches);
        auto capOf = [&](double perRead) {
            double c = perRead * R;
            if (c < minCap)
                c = minCap;
            if (c > 4.0e9)
                c = 4.0e9;
            return (u32)c;
        };
        const u32 capTasks = T * maxSearches + 16;
        const u32 capCands = capOf(candPerRead), capFm = capOf(fmPerRead), capTx = capOf(txPerRead),
                  capOut = capOf(outPerRead);
        bTasks.ensure(be, (size_t)capTasks * (sizeof(SearchTask) + sizeof(u32)));
        bCands.ensure(be, (size_t)capCands * sizeof(Cand));
        bCandAux.ensure(be, (size_t)capCands * 4 * sizeof(u32));
        const u32 capExtra = capCands / 4 + 1024;
        bExtra.ensure(be, (size_t)capExtra * sizeof(u32));
        bFm.ensure(be, (size_t)capFm * sizeof(FMOccRec));
        bTx.ensure(be, (size_t)capTx * sizeof(TextOccRec));
        bTxCig.ensure(be, (size_t)capTx * B.cigW * 2);
        bOut.ensure(be, (size_t)capOut * sizeof(OutOcc));
        bOutCig.ensure(be, (size_t)capOut * B.cigW * 2);
        bCigJobs.ensure(be, (size_t)capOut * sizeof(u32));
        bMisc.ensure(be, sizeof(Misc));
        be.zero(bMisc.p, sizeof(Misc));
        Misc* M = (Misc*)bMisc.p;
        Misc hm;
        memset(&hm, 0, sizeof(hm));
        hm.nT = T;
        hm.nR = R;
        be.h2d(bMisc.p, &hm, sizeof(hm));
        B.tasks.data = (SearchTask*)bTasks.p; B.tasks.count = &M->nTasks; B.tasks.cap = capTasks;
        B.heavy.data = (u32*)(B.tasks.data + capTasks); B.heavy.count = &M->nHeavy; B.heavy.cap = capTasks;
        B.cands.data = (Cand*)bCands.p; B.cands.count = &M->nCands; B.cands.cap = capCands;
        B.candStart = (u32*)bCandAux.p;
        B.candMult = B.candStart + capCands;
        B.uniq.data = B.candMult + capCands; B.uniq.count = &M->nUniq; B.uniq.cap = capCands;
        B.surv.data = B.uniq.data + capCands; B.surv.count = &M->nSurv; B.surv.cap = capCands;
        B.extra.data = (u32*)bExtra.p; B.extra.count = &M->nExtra; B.extra.cap = capExtra;
        B.fmOccs.data = (FMOccRec*)bFm.p; B.fmOccs.count = &M->nFm; B.fmOccs.cap = capFm;
        B.textOccs.data = (TextOccRec*)bTx.p; B.textOccs.count = &M->nTx; B.textOccs.cap = capTx;
        B.textCig = (u16*)bTxCig.p;
        B.out.data = (OutOcc*)bOut.p; B.out.count = &M->nOut; B.out.cap = capOut;
        B.outCig = (u16*)bOutCig.p;
        B.cigJobs.data = (u32*)bCigJobs.p; B.cigJobs.count = &M->nCigJobs; B.cigJobs.cap = capOut;
        B.counters = M->counters;
        B.stats = M->stats;
        B.err = &M->err;
        be.setAbortWord(&M->err);
        // per-group arrays: 2 x (G+1)
        const size_t pr = (size_t)(G + 1);
        bPerRead.ensure(be, pr * 2 * 4 + R + 16);
        be.zero(bPerRead.p, pr * 2 * 4 + R + 16);
        B.outCount = (u32*)bPerRead.p;
        B.outStart = B.outCount + pr;
        B.readStatus = (u8*)(B.outStart + pr);
        // sort buffers of stages 4-6: two (u64 key, u32 value) arrays of capSort + 2 entries each
        const size_t capSort = (size_t)tmax(capFm, capTx) + 2;
        u32 groupBits = 1;
        while (groupBits < 32 && ((u64)1 << groupBits) < (u64)G)
            groupBits++;
        const u32 keyBits = 32 + groupBits;
        bSort.ensure(be, capSort * 2 * (sizeof(u64) + sizeof(u32)));
        B.sortKey[0] = (u64*)bSort.p;
        B.sortKey[1] = B.sortKey[0] + capSort;
        B.sortVal[0] = (u32*)(B.sortKey[1] + capSort);
        B.sortVal[1] = B.sortVal[0] + capSort;
        B.sortRes = BE::sortResult(keyBits);
        B.trimOf = (u32*)B.sortKey[1 - B.sortRes] + capSort; // second half of the scratch key array
        bFmCig.ensure(be, (size_t)capFm * (B.cigW * 2 + 1));
        B.fmCig = (u16*)bFmCig.p;
        B.fmNcig = (u8*)(B.fmCig + (size_t)capFm * B.cigW);
        B.fmSlots = M->fmSlots;
        const u32 capTrim = capOut;
        bTrim.ensure(be, (size_t)capTrim * (sizeof(u32) + sizeof(TrimRec) + B.cigW * 2));
        B.trimRecs = (TrimRec*)bTrim.p;
        B.trimJobs.data = (u32*)(B.trimRecs + capTrim); B.trimJobs.count = &M->nTrim; B.trimJobs.cap = capTrim;
        B.trimCig = (u16*)(B.trimJobs.data + capTrim);
        const DevStrategy* dst = (const DevStrategy*)bStrategy.p;

        // ---- stage 0/1
        {
            for (u32 sl = 0, r0 = 0; sl < nSlices; r0 = sliceEnd[sl], sl++) {
                if (!readsOnDevice)
                    be.copyWait(sl);
                PrepKernel k0{B, r0};
                be.launch(k0, (sliceEnd[sl] - r0) * (B.nw - 1), nullptr, WK_LIGHT);
            }
            be.mark(STG_PREP);
        }
        const u32 vRows = maxLen + 3 * kk + 8;
        B.verifyRowsCap = vRows;
        if (opt.k == 0) {
            // exact stratum: backward exact match + text compare, one task per read
            ExactKernel e0{ix, B};
            be.mark(STG_PARTITION);
            be.launch(e0, R, nullptr, WK_FILTER);
            be.mark(STG_SEARCH);
            be.mark(STG_VERIFY);
        } else {
        {
            // shared memory on the GPU; a slab for backends without it
            const u32 fw = be.fallbackWorkers();
            if (fw)
                bSlabPart.ensure(be, PartScratch::bytesPerWorker(P) * fw);
            PartBase pb{ix, dst, B, P, fw ? (u8*)bSlabPart.p : (u8*)0, 0};
            SeedKernel k1a{pb};
            be.launch(k1a, T, nullptr, WK_LIGHT);
            if (opt.partStrategy == PART_DYNAMIC) {
                // the greedy steps read one character of the sequence per step: stage it on chip
                PartBase pg = pb;
                pg.patWords = (!fw && maxLen <= 512) ? (maxLen + 15) / 16 : 0;
                GreedyKernel k1b{pg};
                be.launch(k1b, T, nullptr, WK_LIGHT);
            }
            EmitKernel k1c{pb};
            be.launch(k1c, T, nullptr, WK_LIGHT);
            be.mark(STG_PARTITION);
        }
        // ---- stage 2
        {
            // The search only appends to the candidate and index-occurrence lists, and an append beyond the capacity
            // still counts: when one of them overflows the search runs to its end, so that the re-run is sized from
            // the full count instead of from wherever the kernels stopped (one re-run instead of several, no
            // doubling past what the batch needs).  Every other overflow stops it at once, as it does all later stages.
            be.setAbortWord(nullptr);
            MaskWordKernel mw{&M->err, &M->errSearch, ~(u32)(ERR_CAND_OVERFLOW | ERR_FMOCC_OVERFLOW)};
            be.launch(mw, 1, nullptr, WK_LIGHT);
            be.setAbortWord(&M->errSearch);
            // one warp per task; a second launch with large scratch takes the tasks whose frontier
            // outgrew the first launch's
            const bool wide = opt.k > BPC<u64>::MAX_ED;
            // the first launch runs one task per group of be.searchGroupSize() lanes: a warp, or an aligned half or
            // quarter of one (the groups of a warp then work on different tasks independently of each other)
            const u32 sg = be.searchGroupSize();
            const u32 nGroups = be.workers(WK_SEARCH) / be.laneStride() * (32u / sg);
            const size_t minSlab = ((size_t)(maxLen + 128) * 4 * (P + 2) + 65536 + 255) & ~(size_t)255;
            const size_t sb = tmax<size_t>(searchSlabBytes, minSlab), hb = tmax<size_t>(heavySlabBytes, sb);
            // heavy tasks: one group of BE::kHeavyGroup lanes (a CTA on the GPU) per task
            const u32 nHeavyGroups = tmax<u32>(1u, be.workers(WK_SEARCH_HEAVY) / (be.heavyGroupSize() / 32u * be.laneStride()));
            bSlabSearch.ensure(be, tmax<size_t>(sb * nGroups, hb * nHeavyGroups));
            if (wide) {
                FrontierKernel<u128, 32> k2{ix, dst, B, (u8*)bSlabSearch.p, sb, 0, heavyWidth, rankPrefetch};
                be.launch(k2, capTasks, &M->nTasks, WK_SEARCH);
                FrontierKernel<u128, BE::kHeavyGroup> k2h{ix, dst, B, (u8*)bSlabSearch.p, hb, 1, heavyWidth, rankPrefetch};
                be.launch(k2h, capTasks, &M->nHeavy, WK_SEARCH_HEAVY);
            } else {
                if (sg == 8) {
                    FrontierKernel<u64, 8> k2{ix, dst, B, (u8*)bSlabSearch.p, sb, 0, heavyWidth, rankPrefetch};
                    be.launch(k2, capTasks, &M->nTasks, WK_SEARCH);
                } else if (sg == 16) {
                    FrontierKernel<u64, 16> k2{ix, dst, B, (u8*)bSlabSearch.p, sb, 0, heavyWidth, rankPrefetch};
                    be.launch(k2, capTasks, &M->nTasks, WK_SEARCH);
                } else {
                    FrontierKernel<u64, 32> k2{ix, dst, B, (u8*)bSlabSearch.p, sb, 0, heavyWidth, rankPrefetch};
                    be.launch(k2, capTasks, &M->nTasks, WK_SEARCH);
                }
                bool launched = false;
                if constexpr (BE::kHeavyGroup > 128) {
                    if (be.heavyGroupSize() == 128) {
                        FrontierKernel<u64, 128> k2s{ix, dst, B, (u8*)bSlabSearch.p, hb, 1, heavyWidth, rankPrefetch};
                        be.launch(k2s, capTasks, &M->nHeavy, WK_SEARCH_HEAVY);
                        launched = true;
                    }
                }
                if (!launched) {
                    FrontierKernel<u64, BE::kHeavyGroup> k2h{ix, dst, B, (u8*)bSlabSearch.p, hb, 1, heavyWidth, rankPrefetch};
                    be.launch(k2h, capTasks, &M->nHeavy, WK_SEARCH_HEAVY);
                }
            }
            be.setAbortWord(&M->err);
            be.mark(STG_SEARCH);
        }
        // ---- stage 3
        {
            const bool wide = !use64(2 * opt.k + 1, opt.k);
            const u32 nw = be.workers(WK_VERIFY), ls = be.laneStride();
            const size_t gb = VerifyKernel::groupSize(B.cigW, vRows, ls);
            bSlabVerify.ensure(be, gb * ((nw + ls - 1) / ls));
            const size_t wb = (size_t)vRows * sizeof(Row<u128>);
            if (wide)
                bSlabVerifyWide.ensure(be, wb * nw);
            VerifyKernel k3{ix, B, (u8*)bSlabVerify.p, gb, vRows, ls, wide ? (u8*)bSlabVerifyWide.p : (u8*)0, wb, 0, verify32 ? 1u : 0u};
            if (opt.metric == METRIC_EDIT) {
                LocateKernel l1{ix, B};
                be.launch(l1, capCands, &M->nCands, WK_LIGHT);
                // collapsing equal windows pays when seeds of a read share their diagonal (45 % of the candidates on
                // a small genome); on repeat-rich input it removes under 1 %, and the 32-neighbour compare is skipped
                // while the previous batch of this context says so
                dedupOn = dedupPays || dedupProbe == 0;
                DedupKernel l2{B, (wide || !dedupOn) ? 0u : 1u};
                be.launch(l2, capCands, &M->nCands, WK_LIGHT);
                ScreenKernel l3{ix, B, screen32 ? 1u : 0u};
                be.launch(l3, capCands, &M->nUniq, WK_LIGHT);
                be.launch(k3, capCands, &M->nSurv, WK_VERIFY);
            } else {
                be.launch(k3, capCands, &M->nCands, WK_VERIFY);
            }
            if (opt.metric == METRIC_EDIT) { // remaining centres of the candidates that have several
                VerifyKernel k3x = k3;
                k3x.pass = 1;
                be.launch(k3x, capExtra, &M->nExtra, WK_VERIFY);
            }
            be.mark(STG_VERIFY);
        }
        } // k > 0
        // ---- stages 4-6: global passes over the occurrence lists (clb_filter.h)
        {
            size_t sb = (((size_t)B.cigW * 2 * 2 + 63) & ~(size_t)63) + (size_t)vRows * sizeof(Row<u128>);
            sb = (sb + 63) & ~(size_t)63;
            const u32 nw = be.workers(WK_FILTER);
            bSlabFilter.ensure(be, sb * nw);
            u64 *ks = B.sortKey[B.sortRes], *kx = B.sortKey[1 - B.sortRes];
            u32 *vs = B.sortVal[B.sortRes], *vx = B.sortVal[1 - B.sortRes];
            (void)ks;
            (void)vs;
            // index occurrences -> text occurrences
            FmKeyKernel c1{ix, B};
            be.launch(c1, capFm, &M->nFm, WK_LIGHT);
            be.sortPairs(B.sortKey[0], B.sortVal[0], B.sortKey[1], B.sortVal[1], capFm, &M->nFm, keyBits);
            FmRunKernel c2{ix, B};
            be.launch(c2, capFm, &M->nFm, WK_LIGHT);
            be.exclusiveScan(vx, (u32*)kx, capFm, &M->nFm);
            FmReserveKernel c3{ix, B};
            be.launch(c3, 1, nullptr, WK_LIGHT);
            if (opt.metric == METRIC_EDIT && !opt.noCigar) {
                FmCigarKernel c4{ix, B, (u8*)bSlabFilter.p, sb, vRows};
                be.launch(c4, capFm, &M->nFm, WK_FILTER);
            }
            FmExpandKernel c5{ix, B};
            be.launch(c5, capTx, &M->fmSlots[1], WK_LIGHT);
            be.mark(STG_FM);
            // text occurrences -> output
            TxKeyKernel d1{ix, B};
            be.launch(d1, capTx, &M->nTx, WK_LIGHT);
            be.sortPairs(B.sortKey[0], B.sortVal[0], B.sortKey[1], B.sortVal[1], capTx, &M->nTx, keyBits);
            TxRunKernel d2{ix, B};
            be.launch(d2, capTx, &M->nTx, WK_LIGHT);
            TxChainKernel d3{ix, B};
            be.launch(d3, capTx, &M->nTx, WK_LIGHT);
            if (opt.metric == METRIC_EDIT) {
                TxTrimKernel d4{ix, B, (u8*)bSlabFilter.p, sb, vRows};
                be.launch(d4, capTrim, &M->nTrim, WK_FILTER);
            }
            be.exclusiveScan(vx, (u32*)kx, capTx, &M->nTx);
            TxReserveKernel d5{ix, B};
            be.launch(d5, 1, nullptr, WK_LIGHT);
            TxEmitKernel d6{ix, B};
            be.launch(d6, capTx, &M->nTx, WK_LIGHT);
            TxGroupKernel d7{ix, B};
            be.launch(d7, G, nullptr, WK_LIGHT);
            CigarJobKernel d8{ix, B, (u8*)bSlabFilter.p, sb, vRows};
            be.launch(d8, capOut, &M->nCigJobs, WK_FILTER);
            be.mark(STG_TEXT);
        }
        // ---- results
        be.setAbortWord(nullptr);
        be.d2h(&hm, bMisc.p, sizeof(hm));
        u32 err = hm.err;
        // the search stage counted every append unless something else stopped it (see stage 2): then the count is
        // what the batch needs, otherwise it is a lower bound
        const bool searchComplete = !(err & ERR_ABORT_MASK & ~(u32)(ERR_CAND_OVERFLOW | ERR_FMOCC_OVERFLOW));
        if (err & ERR_CAND_OVERFLOW)
            candPerRead = tmax(candPerRead * (searchComplete ? 1.0 : 2.0), (double)hm.nCands / R * 1.25);
        if (err & ERR_FMOCC_OVERFLOW)
            fmPerRead = tmax(fmPerRead * (searchComplete ? 1.0 : 2.0), (double)hm.nFm / R * 1.25);
        if (err & ERR_TEXTOCC_OVERFLOW)
            txPerRead = tmax(txPerRead * 2, (double)hm.nTx / R * 1.25 + 1);
        if (err & ERR_OUT_OVERFLOW)
            outPerRead = tmax(outPerRead * 2, (double)hm.nOut / R * 1.25);
        if ((err & ERR_STACK_OVERFLOW) && heavySlabBytes < ((size_t)2 << 30))
            heavySlabBytes *= 4; // a heavy search task outgrew even the large scratch
        if (minCap < (1u << 22) && (err & (ERR_CAND_OVERFLOW | ERR_FMOCC_OVERFLOW | ERR_TEXTOCC_OVERFLOW | ERR_OUT_OVERFLOW)))
            minCap *= 4;
        if (res) {
            res->err = err;
            memcpy(res->counters, hm.counters, sizeof(hm.counters));
            memcpy(res->stats, hm.stats, sizeof(hm.stats));
            res->stats[ST_UNIQUE_CANDS] = opt.metric == METRIC_EDIT ? tmin(hm.nUniq, capCands) : tmin(hm.nCands, capCands);
            res->stats[ST_HEAVY_TASKS] = hm.nHeavy;
        }
        if (opt.metric == METRIC_EDIT && !(err & ERR_ABORT_MASK) && hm.nCands > 100000) {
            if (dedupOn) {
                dedupPays = (double)hm.nUniq < 0.97 * (double)tmin(hm.nCands, capCands);
                dedupProbe = dedupPays ? 0u : 1u;
            } else {
                dedupProbe = (dedupProbe + 1) % 64; // try again every 64 batches
            }
        }
        if (res) {
            res->cigW = B.cigW;
            if (fetch && !(err & ~(u32)ERR_SHORT_READ)) {
                u32 n = hm.nOut;
                res->occ.resize(be, n);
                res->cig.resize(be, (size_t)n * B.cigW);
                res->outStart.resize(be, G);
                res->outCount.resize(be, G);
                if (n) {
                    be.d2hAsync(res->occ.data(), bOut.p, (size_t)n * sizeof(OutOcc));
                    be.d2hAsync(res->cig.data(), bOutCig.p, (size_t)n * B.cigW * 2);
                }
                res->readStatus.resize(be, R);
                res->nShort = 0;
                if (err & ERR_SHORT_READ)
                    be.d2hAsync(res->readStatus.data(), B.readStatus, R);
                else
                    memset(res->readStatus.data(), 0, R);
                be.d2hAsync(res->outStart.data(), B.outStart, (size_t)G * 4);
                be.d2hAsync(res->outCount.data(), B.outCount, (size_t)G * 4);
                be.sync();
                if (err & ERR_SHORT_READ)
                    for (u32 i = 0; i < R; i++)
                        res->nShort += res->readStatus.data()[i] != 0;
            }
        }
        return err;
    }

    // run with automatic re-run after list overflows
    u32 run(const DevStrategy& st, const u8* reads, const u32* off, u32 R, u32 maxLen, size_t totalBytes,
            bool readsOnDevice, const MapOptions& opt, MapResult<BE>* res, bool fetch) {
        const u32 retry = ERR_CAND_OVERFLOW | ERR_FMOCC_OVERFLOW | ERR_TEXTOCC_OVERFLOW | ERR_OUT_OVERFLOW | ERR_STACK_OVERFLOW;
        u32 err = 0;
        static const bool debug = getenv("COLUMBA_B200_DEBUG") != nullptr;
        for (int attempt = 0; attempt < 12; attempt++) {
            const auto t0 = std::chrono::steady_clock::now();
            err = runOnce(st, reads, off, R, maxLen, totalBytes, readsOnDevice, opt, res, fetch);
            if (debug)
                fprintf(stderr, "[clb] attempt %d: %.3f s, error bits %u, lists per read: cand %.1f fm %.1f tx %.1f out %.1f, heavy scratch %zu MB\n",
                        attempt, std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(), err, candPerRead,
                        fmPerRead, txPerRead, outPerRead, heavySlabBytes >> 20);
            if (!(err & retry))
                break;
        }
        return err;
    }
};

} // namespace clb
