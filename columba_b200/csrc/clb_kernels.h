// columba_b200 -- the kernels of the path as task functors.  A backend runs
// functor K over n tasks with W workers: worker w calls begin(), then run(i)
// for its share of the tasks, then end().  The CUDA backend (clb_cuda.cuh)
// launches `clb_kernel<K>` with one thread per worker; tests/hostemu runs the
// same functors in a loop on the CPU to unit-test the kernel logic.
#pragma once
#include "clb_filter.h"
#include "clb_frontier.h"

namespace clb {

struct WorkerCtr {
    LocalCtr lc;
};

// longest read of a batch whose offsets live on the device (the host sizes every per-read buffer from it)
struct MaxLenKernel {
    const u32* off;
    u32* out;
    struct State {
        u32 m;
    };
    CLB_HD void begin(State& s, u32) const { s.m = 0; }
    CLB_HD void run(State& s, u32 i) const { s.m = tmax(s.m, off[i + 1] - off[i]); }
    CLB_HD void end(State& s) const {
        if (s.m)
            atomMaxU32(out, s.m);
    }
};

// *dst = *src & mask (the error word the search stage aborts on, see Pipeline::runOnce)
struct MaskWordKernel {
    const u32* src;
    u32* dst;
    u32 mask;
    struct State {};
    CLB_HD void begin(State&, u32) const {}
    CLB_HD void run(State&, u32) const { *dst = *src & mask; }
    CLB_HD void end(State&) const {}
};

// stage 0
struct PrepKernel {
    Batch B;
    u32 firstRead; // the launch covers reads [firstRead, ...)
    struct State {};
    CLB_HD void begin(State&, u32) const {}
    CLB_HD void run(State&, u32 i) const { prepTask(B, i, firstRead); }
    CLB_HD void end(State&) const {}
};

// stage 1a / 1c: uniform kernels around the greedy steps (clb_partition.h)
struct PartBase {
    IndexView ix;
    const DevStrategy* st;
    Batch B;
    u32 P;       // parts (st->P, known to the host)
    u8* slabs;   // scratch of the workers when there is no shared memory (host emulation), else null
    u32 patWords; // words per worker for the sequence as 2-bit codes in shared memory (0: not staged)
};
struct PartWorkerState {
    LocalCtr lc;
    PartState ps;
};
#define CLB_PART_COMMON                                                                                        \
    PartBase a;                                                                                                \
    typedef PartWorkerState State;                                                                             \
    CLB_HD size_t smemBytes(u32 nThreads) const {                                                              \
        return (PartScratch::bytesPerWorker(a.P) + (size_t)a.patWords * sizeof(u32)) * nThreads;               \
    }                                                                                                          \
    CLB_HD void bindShared(State& s, u8* smem, u32 tid, u32 nThreads) const {                                  \
        s.ps.ps.bind(smem, nThreads, tid, a.P);                                                                \
        s.ps.patS = a.patWords ? (u32*)smem + (size_t)a.P * 7 * nThreads + tid : (u32*)0;                      \
        s.ps.patStride = nThreads;                                                                             \
    }                                                                                                          \
    CLB_HD void begin(State& s, u32 worker) const {                                                            \
        s.lc.clear();                                                                                          \
        s.ps.patS = (u32*)0;                                                                                   \
        s.ps.patStride = 1;                                                                                    \
        s.ps.ps.bind(a.slabs ? a.slabs + (size_t)worker * PartScratch::bytesPerWorker(a.P) : (u8*)0, 1, 0, a.P); \
    }                                                                                                          \
    CLB_HD void end(State& s) const { s.lc.flush(a.B.counters, a.B.stats); }

struct SeedKernel {
    CLB_PART_COMMON
    CLB_HD void run(State& s, u32 t) const { partStart(a.ix, *a.st, a.B, t, s.ps, s.lc); }
};
// stage 1b.  Stepped functor: besides run() (the whole task, used by the host
// emulation) it offers start / step / finish, which the CUDA launcher drives
// with per-lane task refill (clb_kernel_stepped).
struct GreedyKernel {
    CLB_PART_COMMON
    static const bool kStepped = true;
    CLB_HD void run(State& s, u32 t) const {
        start(s, t);
        while (!step(s)) {
        }
        finish(s);
    }
    CLB_HD void start(State& s, u32 t) const { greedyStart(*a.st, a.B, t, s.ps); }
    CLB_HD bool step(State& s) const { return partStep(a.ix, *a.st, s.ps, s.lc); }
    CLB_HD void finish(State& s) const { greedyFinish(*a.st, a.B, s.ps); }
};
struct EmitKernel {
    CLB_PART_COMMON
    CLB_HD void run(State& s, u32 t) const { partFinish(a.ix, *a.st, a.B, t, s.ps, s.lc, 1u); }
    CLB_HD void run(State& s, u32 t, unsigned act) const { partFinish(a.ix, *a.st, a.B, t, s.ps, s.lc, act); }
};

// stage 2: one group of GS lanes per (read, strand, search) task (clb_frontier.h).
// First launch: one warp per task (GS = 32).  A task whose frontier outgrows the
// warp's scratch is queued in B.heavy and run again by a second launch (heavy = 1)
// with large scratch and, on the GPU, a whole CTA of GS threads per task: such a
// task has levels of thousands of nodes, and the launch lasts as long as its
// longest task.
template <typename W, int GS> struct FrontierKernel {
    static const bool kWarpTask = true;
    static const int kGroup = GS;
    IndexView ix;
    const DevStrategy* st;
    Batch B;
    u8* slabs;
    size_t slabBytes; // per warp
    u32 heavy;
    u32 heavyWidth;   // first launch: tasks that start from a wider range are queued for the heavy launch at once
    u32 rankPrefetch; // ranking passes prefetch one level ahead (Frontier::rankPrefetch)
    struct State {
        LV<FCounters> ctr;
        LV<u64> started, immediate;
        u64 tasks;
        u8* scr;
    };
    CLB_HD void begin(State& s, u32 warp) const {
        CLB_LANES_G(l, GS) {
            s.ctr[l].nodes = s.ctr[l].lookups = s.ctr[l].rows = s.ctr[l].fmOccs = 0;
            s.started[l] = 0;
            s.immediate[l] = 0;
        }
        s.tasks = 0;
        s.scr = slabs + (size_t)warp * slabBytes; // warp = index of the group (warp or CTA)
    }
    CLB_HD void runWarp(State& s, u32 i) const {
        const u32 ti = heavy ? B.heavy.data[i] : i;
        const SearchTask task = B.tasks.data[ti];
        const PartInfo pi = B.parts.load(task.t, st->P);
        const DevSearch& se = st->sc[pi.scheme].s[task.search];
        Frontier<W, GS> F(ix, B, se, pi, task.t, task.search, s.scr, (u32)slabBytes);
        F.heavyWidth = heavy ? 0u : heavyWidth;
        F.rankPrefetch = rankPrefetch != 0;
        LV<u64> started, immediate;
        CLB_LANES_G(l, GS) {
            started[l] = 0;
            immediate[l] = 0;
        }
        const bool ok = F.run(started, immediate);
        if (!ok) {
            CLB_LANES_G(l, GS) {
                if (l == 0) {
                    if (F.err == ERR_STACK_OVERFLOW && !heavy)
                        B.heavy.push(ti, B.err, ERR_TASK_OVERFLOW);
                    else
                        atomOr(B.err, F.err);
                }
            }
            return;
        }
        s.tasks++;
        CLB_LANES_G(l, GS) {
            s.ctr[l].nodes += F.ctr[l].nodes;
            s.ctr[l].lookups += F.ctr[l].lookups;
            s.ctr[l].rows += F.ctr[l].rows;
            s.ctr[l].fmOccs += F.ctr[l].fmOccs;
            s.started[l] += started[l];
            s.immediate[l] += immediate[l];
        }
    }
    CLB_HD void end(State& s) const {
        CLB_LANES_G(l, GS) {
            if (s.ctr[l].nodes)
                atomAdd64(B.counters + NODE_COUNTER, s.ctr[l].nodes);
            if (s.ctr[l].lookups) {
                atomAdd64(B.stats + ST_RANK_LOOKUPS, s.ctr[l].lookups);
                atomAdd64(B.stats + ST_SEARCH_LOOKUPS, s.ctr[l].lookups);
            }
            if (s.ctr[l].rows)
                atomAdd64(B.stats + ST_SEARCH_ROWS, s.ctr[l].rows);
            if (s.ctr[l].fmOccs)
                atomAdd64(B.stats + ST_FM_OCCS, s.ctr[l].fmOccs);
            if (s.started[l])
                atomAdd64(B.counters + SEARCH_STARTED, s.started[l]);
            if (s.immediate[l])
                atomAdd64(B.counters + IMMEDIATE_SWITCH, s.immediate[l]);
            if (l == 0 && s.tasks)
                atomAdd64(B.stats + ST_SEARCH_TASKS, s.tasks);
        }
    }
};

// exact stratum (k = 0): one task per read, both strands
struct ExactKernel {
    IndexView ix;
    Batch B;
    typedef WorkerCtr State;
    CLB_HD void begin(State& s, u32) const { s.lc.clear(); }
    CLB_HD void run(State& s, u32 r) const { exactTask(ix, B, r, s.lc); }
    CLB_HD void end(State& s) const { s.lc.flush(B.counters, B.stats); }
};

// stage 3
struct VerifyKernel {
    IndexView ix;
    Batch B;
    u8* slabs;        // per lane group: ops[laneStride], cp[2 * cpCap][laneStride], ring[32][laneStride]
    size_t groupBytes;
    u32 rowsCap;
    u32 laneStride;   // 32 on the GPU (the lanes of a warp interleave), 1 in the host emulation
    u8* wideSlabs;    // Row<u128>[rowsCap] per worker, only when 128-bit matrices can occur
    size_t wideBytes;
    u32 pass;         // 0: all candidates, first centre; 1: the `extra` list, remaining centres
    u32 band32;       // 32-bit band where it fits (clb_verify.h)
    struct State {
        LocalCtr lc;
        VerifyScratch vs;
    };
    static CLB_HD size_t opsBytes(u32 cigW) { return ((size_t)cigW * sizeof(u16) + 15) & ~(size_t)15; }
    static CLB_HD u32 cpCapOf(u32 rowsCap) { return rowsCap / BPC<u64>::BLOCK + 2; }
    static CLB_HD size_t opsGroup(u32 cigW, u32 laneStride) { return (opsBytes(cigW) * laneStride + 255) & ~(size_t)255; }
    static CLB_HD size_t groupSize(u32 cigW, u32 rowsCap, u32 laneStride) {
        return opsGroup(cigW, laneStride) + (size_t)(2 * cpCapOf(rowsCap) + BPC<u64>::BLOCK) * laneStride * sizeof(u64);
    }
    // the CUDA launcher gives every CTA a ring in shared memory
    static CLB_HD size_t smemBytes(u32 nThreads) { return (size_t)BPC<u64>::BLOCK * nThreads * sizeof(u64); }
    CLB_HD void bindShared(State& s, u8* smem, u32 tid, u32 nThreads) const {
        s.vs.bs.ring = (u64*)smem + tid;
        s.vs.bs.ringStride = nThreads;
    }
    CLB_HD void begin(State& s, u32 worker) const {
        s.lc.clear();
        const u32 group = worker / laneStride, lane = worker % laneStride;
        u8* g = slabs + (size_t)group * groupBytes;
        s.vs.ops = (u16*)(g + lane * opsBytes(B.cigW));
        u64* cp = (u64*)(g + opsGroup(B.cigW, laneStride)) + lane;
        const u32 cpCap = cpCapOf(rowsCap);
        s.vs.bs.cp = cp;
        s.vs.bs.cpStride = laneStride;
        s.vs.bs.cpCap = cpCap;
        s.vs.bs.ring = cp + (size_t)2 * cpCap * laneStride;
        s.vs.bs.ringStride = laneStride;
        s.vs.rows = wideSlabs ? (void*)(wideSlabs + (size_t)worker * wideBytes) : (void*)0;
        s.vs.rowsCap = rowsCap;
        s.vs.warpMask = 1u;
        s.vs.band32 = band32;
    }
    // act: the lanes of the warp that run a task in this round
    CLB_HD void run(State& s, u32 i, unsigned act) const {
        s.vs.warpMask = act;
        run(s, i);
    }
    CLB_HD void run(State& s, u32 i) const {
        if (pass) {
            verifyEditTask(ix, B, B.extra.data[i], s.vs, s.lc, 1);
            return;
        }
        if (B.metric == METRIC_EDIT) {
            verifyEditTask(ix, B, B.surv.data[i], s.vs, s.lc, 0);
        } else {
            s.lc.s[ST_CANDIDATES]++;
            verifyHammingTask(ix, B, i, s.lc);
        }
    }
    CLB_HD void end(State& s) const { s.lc.flush(B.counters, B.stats); }
};

// stage 3c': screening pass (clb_verify.h), stepped with lane refill
struct ScreenKernel {
    IndexView ix;
    Batch B;
    u32 band32; // use the 32-bit band where it fits (clb_verify.h screenStep32)
    static const bool kStepped = true;
    struct State {
        LocalCtr lc;
        ScreenState s;
    };
    CLB_HD void begin(State& st, u32) const { st.lc.clear(); }
    CLB_HD void run(State& st, u32 i) const {
        start(st, i);
        while (!step(st)) {
        }
        finish(st);
    }
    CLB_HD void start(State& st, u32 i) const { screenStart(ix, B, B.uniq.data[i], st.s, band32 != 0); }
    CLB_HD bool step(State& st) const { return screenStep(ix, st.s); }
    CLB_HD void finish(State& st) const { screenFinish(B, st.s, st.lc); }
    CLB_HD void end(State& st) const { st.lc.flush(B.counters, B.stats); }
};

// stage 3a / 3b (edit distance)
struct LocateKernel {
    IndexView ix;
    Batch B;
    typedef WorkerCtr State;
    CLB_HD void begin(State& s, u32) const { s.lc.clear(); }
    CLB_HD void run(State& s, u32 i) const { locateTask(ix, B, i, s.lc); }
    CLB_HD void end(State& s) const { s.lc.flush(B.counters, B.stats); }
};
struct DedupKernel {
    Batch B;
    u32 enable;
    struct State {};
    CLB_HD void begin(State&, u32) const {}
    CLB_HD u32 total() const { return tmin(*B.cands.count, B.cands.cap); }
    CLB_HD void run(State&, u32 i) const { dedupTask(B, i, total(), enable != 0, 1u); }
    CLB_HD void run(State&, u32 i, unsigned act) const { dedupTask(B, i, total(), enable != 0, act); }
    CLB_HD void end(State&) const {}
};

// stages 4-6 as global passes (clb_filter.h)
struct FilterState {
    LocalCtr lc;
    FilterScratch fs;
};
CLB_HD void carveFilter(FilterState& s, u8* p, u32 cigW, u32 rowsCap) {
    s.lc.clear();
    s.fs.ops = (u16*)p;
    s.fs.rows = p + (((size_t)cigW * 2 * sizeof(u16) + 63) & ~(size_t)63);
    s.fs.rowsCap = rowsCap;
}
#define CLB_PLAIN_KERNEL(NAME, CALL)                       \
    struct NAME {                                          \
        IndexView ix;                                      \
        Batch B;                                           \
        struct State {};                                   \
        CLB_HD void begin(State&, u32) const {}            \
        CLB_HD void run(State&, u32 i) const { CALL; }     \
        CLB_HD void end(State&) const {}                   \
    };
#define CLB_COUNTING_KERNEL(NAME, CALL)                                        \
    struct NAME {                                                              \
        IndexView ix;                                                          \
        Batch B;                                                               \
        typedef WorkerCtr State;                                               \
        CLB_HD void begin(State& s, u32) const { s.lc.clear(); }               \
        CLB_HD void run(State& s, u32 i) const { CALL; }                       \
        CLB_HD void end(State& s) const { s.lc.flush(B.counters, B.stats); }   \
    };
#define CLB_SCRATCH_KERNEL(NAME, CALL)                                                                        \
    struct NAME {                                                                                             \
        IndexView ix;                                                                                         \
        Batch B;                                                                                              \
        u8* slabs;                                                                                            \
        size_t slabBytes;                                                                                     \
        u32 rowsCap;                                                                                          \
        typedef FilterState State;                                                                            \
        CLB_HD void begin(State& s, u32 worker) const {                                                       \
            carveFilter(s, slabs + (size_t)worker * slabBytes, B.cigW, rowsCap);                              \
        }                                                                                                     \
        CLB_HD void run(State& s, u32 i) const { CALL; }                                                      \
        CLB_HD void end(State& s) const { s.lc.flush(B.counters, B.stats); }                                  \
    };
CLB_PLAIN_KERNEL(FmKeyKernel, fmKeyTask(B, i))
CLB_COUNTING_KERNEL(FmRunKernel, fmRunTask(B, i, s.lc))
CLB_PLAIN_KERNEL(FmReserveKernel, fmReserveTask(B))
CLB_SCRATCH_KERNEL(FmCigarKernel, fmCigarTask(ix, B, i, s.fs, s.lc))
CLB_COUNTING_KERNEL(FmExpandKernel, fmExpandTask(ix, B, i, s.lc))
CLB_PLAIN_KERNEL(TxKeyKernel, txKeyTask(B, i))
CLB_PLAIN_KERNEL(TxRunKernel, txRunTask(B, i))
CLB_PLAIN_KERNEL(TxChainKernel, txChainTask(ix, B, i))
CLB_SCRATCH_KERNEL(TxTrimKernel, txTrimTask(ix, B, i, s.fs, s.lc))
CLB_PLAIN_KERNEL(TxReserveKernel, txReserveTask(B))
CLB_PLAIN_KERNEL(TxEmitKernel, txEmitTask(ix, B, i))
CLB_COUNTING_KERNEL(TxGroupKernel, txGroupTask(B, i, s.lc))

struct CigarJobKernel {
    IndexView ix;
    Batch B;
    u8* slabs;
    size_t slabBytes;
    u32 rowsCap;
    typedef FilterState State;
    CLB_HD void begin(State& s, u32 worker) const { carveFilter(s, slabs + (size_t)worker * slabBytes, B.cigW, rowsCap); }
    CLB_HD void run(State& s, u32 i) const { cigarJobTask(ix, B, i, s.fs, s.lc); }
    CLB_HD void end(State& s) const { s.lc.flush(B.counters, B.stats); }
};

// k-mer table construction (indexinterface.cpp:294-335), level by level:
// entry `code` of level d extends entry code>>2 of level d-1 by one character
// in the forward direction.
struct KmerLevelKernel {
    IndexView ix;
    const RangePair* prev;
    RangePair* cur;
    struct State {};
    CLB_HD void begin(State&, u32) const {}
    CLB_HD void run(State&, u32 code) const {
        RangePair r = prev[code >> 2];
        if (r.empty()) {
            cur[code] = emptyPair();
            return;
        }
        if (!extendOne(ix, DIR_FWD, false, code & 3, r))
            r = emptyPair();
        cur[code] = r;
    }
    CLB_HD void end(State&) const {}
};

// index parity micro-kernels (exposed through the C ABI for unit tests)
struct Occ4Kernel {
    IndexView ix;
    u32 reverse;
    const u32* pos;
    u32* out; // 4 per position
    struct State {};
    CLB_HD void begin(State&, u32) const {}
    CLB_HD void run(State&, u32 i) const { occ4(reverse ? ix.rev : ix.fwd, pos[i], out + 4 * (size_t)i); }
    CLB_HD void end(State&) const {}
};
struct ExtendKernel {
    IndexView ix;
    u32 mode; // 0 fwd, 1 bwd, 2 uni-directional bwd
    const RangePair* parents;
    RangePair* out; // 4 per parent
    struct State {};
    CLB_HD void begin(State&, u32) const {}
    CLB_HD void run(State&, u32 i) const {
        extendAll(ix, mode == 0 ? DIR_FWD : DIR_BWD, mode == 2, parents[i], out + 4 * (size_t)i);
    }
    CLB_HD void end(State&) const {}
};
struct FindSAKernel {
    IndexView ix;
    const u32* rows;
    u32* out;
    struct State {};
    CLB_HD void begin(State&, u32) const {}
    CLB_HD void run(State&, u32 i) const {
        u64 lf = 0;
        out[i] = findSA(ix, rows[i], lf);
    }
    CLB_HD void end(State&) const {}
};

} // namespace clb
