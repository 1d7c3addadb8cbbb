// columba_b200 -- the product library: CUDA backend (sm_100a) + C ABI.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -shared ...
// (see __graft_entry__.build()).  There is no CPU execution path in this
// translation unit: every task functor of clb_kernels.h runs inside
// clb_kernel<K> on the device.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "clb_pipeline.h"

namespace clb {

#define CLB_CUDA(x)                                                                                  \
    do {                                                                                             \
        cudaError_t e_ = (x);                                                                        \
        if (e_ != cudaSuccess)                                                                       \
            throw BackendError(std::string("CUDA error: ") + cudaGetErrorString(e_) + " at " #x);   \
    } while (0)

// One thread = one worker.  Static striding for uniform tasks; for tasks of
// very different cost (searches, verifications, per-read filters) each warp
// claims 32 tasks at a time from a global counter.
// optional per-CTA shared memory of a functor: K::smemBytes(nThreads) and
// K::bindShared(state, smem, tid, nThreads)
template <class K> auto smemBytesOf(const K& k, u32 nThreads, int) -> decltype(k.smemBytes(nThreads)) {
    return k.smemBytes(nThreads);
}
template <class K> size_t smemBytesOf(const K&, u32, long) { return 0; }
template <class K, class S>
__device__ __forceinline__ auto bindSharedOf(const K& k, S& st, u8* smem, int) -> decltype(k.bindShared(st, smem, 0u, 0u)) {
    k.bindShared(st, smem, threadIdx.x, blockDim.x);
}
template <class K, class S> __device__ __forceinline__ void bindSharedOf(const K&, S&, u8*, long) {}

// functors that want to know which lanes run a task next to them take the
// lane mask as a third argument of run()
template <class K, class S>
__device__ __forceinline__ auto runOf(const K& k, S& st, u32 i, unsigned act, int) -> decltype(k.run(st, i, act)) {
    k.run(st, i, act);
}
template <class K, class S> __device__ __forceinline__ void runOf(const K& k, S& st, u32 i, unsigned, long) { k.run(st, i); }

template <class K>
__global__ void __launch_bounds__(128) clb_kernel(const __grid_constant__ K k, u32 nMax, const u32* __restrict__ nDev, u32* workCtr, const u32* __restrict__ abortWord) {
    if (abortWord && (*abortWord & ERR_ABORT_MASK))
        return; // an earlier stage of this attempt overflowed a list: what follows would read unwritten records
    extern __shared__ __align__(16) u8 clb_smem[];
    u32 n = nMax;
    if (nDev) {
        u32 v = *nDev;
        n = v < nMax ? v : nMax;
    }
    const u32 worker = blockIdx.x * blockDim.x + threadIdx.x;
    const u32 nWorkers = gridDim.x * blockDim.x;
    typename K::State st;
    k.begin(st, worker);
    bindSharedOf(k, st, clb_smem, 0);
    if (workCtr == nullptr) {
        // warp-uniform trip count, so that the lanes of a warp can vote
        for (u32 i0 = worker - (threadIdx.x & 31); i0 < n; i0 += nWorkers) {
            const u32 i = i0 + (threadIdx.x & 31);
            const unsigned act = __ballot_sync(0xFFFFFFFFu, i < n);
            if (i < n)
                runOf(k, st, i, act, 0);
        }
    } else {
        const u32 lane = threadIdx.x & 31;
        for (;;) {
            u32 base = 0;
            if (lane == 0)
                base = atomicAdd(workCtr, 32u);
            base = __shfl_sync(0xFFFFFFFFu, base, 0);
            if (base >= n)
                break;
            u32 i = base + lane;
            const unsigned act = __ballot_sync(0xFFFFFFFFu, i < n);
            if (i < n)
                runOf(k, st, i, act, 0);
            __syncwarp();
        }
    }
    k.end(st);
}

// Stepped functors (K::kStepped): a task is start(), some number of step()s,
// finish().  Every lane owns one task at a time; lanes whose task is finished
// claim the next one (one atomic per warp and refill round) while the other
// lanes keep stepping, so the step loop runs with (almost) full warps even
// when the tasks differ a lot in their number of steps.
template <class K>
__global__ void __launch_bounds__(128) clb_kernel_stepped(const __grid_constant__ K k, u32 nMax, const u32* __restrict__ nDev, u32* workCtr, const u32* __restrict__ abortWord) {
    if (abortWord && (*abortWord & ERR_ABORT_MASK))
        return; // an earlier stage of this attempt overflowed a list: what follows would read unwritten records
    u32 n = nMax;
    if (nDev) {
        u32 v = *nDev;
        n = v < nMax ? v : nMax;
    }
    const u32 worker = blockIdx.x * blockDim.x + threadIdx.x;
    const u32 lane = threadIdx.x & 31;
    extern __shared__ __align__(16) u8 clb_smem[];
    typename K::State st;
    k.begin(st, worker);
    bindSharedOf(k, st, clb_smem, 0);
    bool have = false, done = false;
    for (;;) {
        const unsigned need = __ballot_sync(0xFFFFFFFFu, !have && !done);
        const unsigned busy = __ballot_sync(0xFFFFFFFFu, have);
        // start() and finish() run with only the refilling lanes: wait until a
        // quarter of the warp is idle (or nothing is left to step)
        if (need && (__popc(need) >= 8 || busy == 0)) {
            const int leader = __ffs(need) - 1;
            u32 base = 0;
            if ((int)lane == leader)
                base = atomicAdd(workCtr, (u32)__popc(need));
            base = __shfl_sync(0xFFFFFFFFu, base, leader);
            if (!have && !done) {
                const u32 i = base + __popc(need & ((1u << lane) - 1u));
                if (i < n) {
                    k.start(st, i);
                    have = true;
                } else {
                    done = true;
                }
            }
        }
        if (__all_sync(0xFFFFFFFFu, done))
            break;
        // a few steps per refill round: lanes that finish in between wait
        bool fin = false;
        if (have) {
#pragma unroll 1
            for (int q = 0; q < 4 && !fin; q++)
                fin = k.step(st);
        }
        if (fin) {
            k.finish(st);
            have = false;
        }
    }
    k.end(st);
}
template <class K> constexpr auto isStepped(int) -> decltype(K::kStepped) { return K::kStepped; }
template <class K> constexpr bool isStepped(long) { return false; }

// Warp-task functors (K::kWarpTask): a task is run by all 32 lanes of a warp
// together (K::runWarp); warps claim tasks one at a time from a global counter.
template <class K> constexpr auto groupOfK(int) -> decltype(K::kGroup) { return K::kGroup; }
template <class K> constexpr int groupOfK(long) { return 32; }
template <class K, int MINB, int G>
__global__ void __launch_bounds__(128, MINB) clb_kernel_warp(const __grid_constant__ K k, u32 nMax, const u32* __restrict__ nDev, u32* workCtr, const u32* __restrict__ abortWord) {
    if (abortWord && (*abortWord & ERR_ABORT_MASK))
        return; // an earlier stage of this attempt overflowed a list: what follows would read unwritten records
    u32 n = nMax;
    if (nDev) {
        u32 v = *nDev;
        n = v < nMax ? v : nMax;
    }
    extern __shared__ __align__(16) u8 clb_smem[];
    // a group is a warp or an aligned part of one (K::kGroup = 8, 16): the groups of a warp claim and run their
    // tasks independently, every collective below names the group's lanes only
    const u32 group = (blockIdx.x * blockDim.x + threadIdx.x) / (u32)G;
    const u32 lane = threadIdx.x & (u32)(G - 1);
    const unsigned gm = G == 32 ? 0xFFFFFFFFu : (((1u << (G & 31)) - 1u) << ((threadIdx.x & 31u) & ~(u32)(G - 1)));
    typename K::State st;
    k.begin(st, group);
    bindSharedOf(k, st, clb_smem, 0);
    for (;;) {
        u32 i = 0;
        if (lane == 0)
            i = atomicAdd(workCtr, 1u);
        i = __shfl_sync(gm, i, 0, G);
        if (i >= n)
            break;
        k.runWarp(st, i);
        __syncwarp(gm);
    }
    k.end(st);
}
template <class K> constexpr auto isWarpTask(int) -> decltype(K::kWarpTask) { return K::kWarpTask; }
template <class K> constexpr bool isWarpTask(long) { return false; }

// Group-task functors with K::kGroup > 32: a task is run by all threads of a CTA of K::kGroup threads
// together (the functor's collectives use __syncthreads); CTAs claim tasks one at a time.
static const int HEAVY_GROUP = 256, HEAVY_THREADS_PER_SM = 512;
template <class K, int THREADS, int PER_SM>
__global__ void __launch_bounds__(THREADS, PER_SM / THREADS) clb_kernel_cta(const __grid_constant__ K k, u32 nMax, const u32* __restrict__ nDev, u32* workCtr, const u32* __restrict__ abortWord) {
    if (abortWord && (*abortWord & ERR_ABORT_MASK))
        return; // an earlier stage of this attempt overflowed a list: what follows would read unwritten records
    __shared__ u32 claimed;
    u32 n = nMax;
    if (nDev) {
        u32 v = *nDev;
        n = v < nMax ? v : nMax;
    }
    typename K::State st;
    k.begin(st, blockIdx.x);
    for (;;) {
        if (threadIdx.x == 0)
            claimed = atomicAdd(workCtr, 1u);
        __syncthreads();
        const u32 i = claimed;
        __syncthreads();
        if (i >= n)
            break;
        k.runWarp(st, i);
        __syncthreads();
    }
    k.end(st);
}

// exclusive scan of n u32 values (out[n] = total) in three launches:
// per-tile sums, scan of the tile sums by one block, per-tile scan + offset.
// n = min(nMax, *nDev) when nDev is given (list sizes live on the device).
static const u32 SCAN_TILE = 2048; // 256 threads x 8 values
__device__ __forceinline__ u32 devCount(u32 nMax, const u32* nDev) {
    if (!nDev)
        return nMax;
    const u32 v = *nDev;
    return v < nMax ? v : nMax;
}
__global__ void __launch_bounds__(256) clb_scan_tile_sums(const u32* __restrict__ in, u32* __restrict__ sums, u32 nMax,
                                                          const u32* __restrict__ nDev) {
    __shared__ u32 red[8];
    const u32 n = devCount(nMax, nDev);
    const u32 nb = (n + SCAN_TILE - 1) / SCAN_TILE;
    for (u32 tile = blockIdx.x; tile < nb; tile += gridDim.x) {
        const u32 base = tile * SCAN_TILE;
        u32 s = 0;
        for (u32 q = 0; q < 8; q++) {
            u32 i = base + q * 256 + threadIdx.x;
            if (i < n)
                s += in[i];
        }
        for (u32 o = 16; o > 0; o >>= 1)
            s += __shfl_down_sync(0xFFFFFFFFu, s, o);
        if ((threadIdx.x & 31) == 0)
            red[threadIdx.x >> 5] = s;
        __syncthreads();
        if (threadIdx.x == 0) {
            u32 t = 0;
            for (u32 w = 0; w < 8; w++)
                t += red[w];
            sums[tile] = t;
        }
        __syncthreads();
    }
}
// in place exclusive scan of the tile sums by one block; total -> out[n]
__global__ void __launch_bounds__(1024) clb_scan_sums(u32* __restrict__ sums, u32 nMax, const u32* __restrict__ nDev,
                                                      u32* __restrict__ out) {
    __shared__ u32 part[1024];
    const u32 n = devCount(nMax, nDev);
    const u32 nb = (n + SCAN_TILE - 1) / SCAN_TILE;
    const u32 tid = threadIdx.x, nt = blockDim.x;
    const u32 per = (nb + nt - 1) / nt;
    const u32 b = tmin(tid * per, nb), e = tmin(b + per, nb);
    u32 sum = 0;
    for (u32 i = b; i < e; i++)
        sum += sums[i];
    part[tid] = sum;
    __syncthreads();
    for (u32 off = 1; off < nt; off <<= 1) {
        u32 v = (tid >= off) ? part[tid - off] : 0;
        __syncthreads();
        part[tid] += v;
        __syncthreads();
    }
    u32 run = part[tid] - sum;
    for (u32 i = b; i < e; i++) {
        u32 v = sums[i];
        sums[i] = run;
        run += v;
    }
    if (tid == nt - 1)
        out[n] = part[nt - 1];
}
__global__ void __launch_bounds__(256) clb_scan_tiles(const u32* __restrict__ in, u32* __restrict__ out,
                                                      const u32* __restrict__ sums, u32 nMax,
                                                      const u32* __restrict__ nDev) {
    __shared__ u32 wsum[8];
    const u32 n = devCount(nMax, nDev);
    const u32 nb = (n + SCAN_TILE - 1) / SCAN_TILE;
    for (u32 tile = blockIdx.x; tile < nb; tile += gridDim.x) {
        const u32 base = tile * SCAN_TILE + threadIdx.x * 8;
        u32 v[8], s = 0;
        for (u32 q = 0; q < 8; q++) {
            v[q] = (base + q < n) ? in[base + q] : 0;
            s += v[q];
        }
        // exclusive scan of s over the block
        u32 incl = s;
        const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        for (u32 o = 1; o < 32; o <<= 1) {
            u32 t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o)
                incl += t;
        }
        if (lane == 31)
            wsum[warp] = incl;
        __syncthreads();
        u32 woff = 0;
        for (u32 w = 0; w < warp; w++)
            woff += wsum[w];
        u32 run = sums[tile] + woff + incl - s;
        for (u32 q = 0; q < 8; q++) {
            if (base + q < n)
                out[base + q] = run;
            run += v[q];
        }
        __syncthreads();
    }
}

// Least-significant-digit radix sort of (u64 key, u32 value) pairs, 8 bits per
// pass, three launches per pass: per-tile digit histograms (digit-major, so
// that a row holds one digit's counts over the tiles), one CTA per digit turns
// its row into exclusive prefix sums and writes the digit's total, and the
// scatter pass ranks the keys of a tile stably -- warp by warp, eight rounds
// of 32 keys, the lanes that hold the same digit found with match.any -- and
// writes each pair to its final place of the pass.  n lives on the device.
static const u32 RS_TILE = 2048; // 256 threads, 8 warps x 8 rounds x 32 keys
__global__ void __launch_bounds__(256) clb_radix_hist(const u64* __restrict__ keys, u32 nMax, const u32* __restrict__ nDev,
                                                      u32 shift, u32* __restrict__ hist) {
    __shared__ u32 cnt[256];
    const u32 n = devCount(nMax, nDev);
    const u32 nTiles = (n + RS_TILE - 1) / RS_TILE;
    const u32 lane = threadIdx.x & 31;
    for (u32 tile = blockIdx.x; tile < nTiles; tile += gridDim.x) {
        cnt[threadIdx.x] = 0;
        __syncthreads();
        const u32 base = tile * RS_TILE;
#pragma unroll
        for (u32 q = 0; q < 8; q++) {
            const u32 i = base + q * 256 + threadIdx.x;
            const u32 d = i < n ? (u32)(keys[i] >> shift) & 255u : 0xFFFFFFFFu;
            const unsigned m = __match_any_sync(0xFFFFFFFFu, d);
            if (d != 0xFFFFFFFFu && (int)lane == __ffs(m) - 1)
                atomicAdd(&cnt[d], (u32)__popc(m));
        }
        __syncthreads();
        hist[(size_t)threadIdx.x * nTiles + tile] = cnt[threadIdx.x];
        __syncthreads();
    }
}
__global__ void __launch_bounds__(1024) clb_radix_scan(u32* __restrict__ hist, u32 nMax, const u32* __restrict__ nDev,
                                                       u32* __restrict__ digitTotal) {
    __shared__ u32 part[1024];
    const u32 n = devCount(nMax, nDev);
    const u32 nTiles = (n + RS_TILE - 1) / RS_TILE;
    u32* row = hist + (size_t)blockIdx.x * nTiles;
    const u32 tid = threadIdx.x, nt = blockDim.x;
    const u32 per = (nTiles + nt - 1) / nt;
    const u32 b = tmin(tid * per, nTiles), e = tmin(b + per, nTiles);
    u32 sum = 0;
    for (u32 i = b; i < e; i++)
        sum += row[i];
    part[tid] = sum;
    __syncthreads();
    for (u32 off = 1; off < nt; off <<= 1) {
        u32 v = (tid >= off) ? part[tid - off] : 0;
        __syncthreads();
        part[tid] += v;
        __syncthreads();
    }
    u32 run = part[tid] - sum;
    for (u32 i = b; i < e; i++) {
        u32 v = row[i];
        row[i] = run;
        run += v;
    }
    if (tid == nt - 1)
        digitTotal[blockIdx.x] = part[nt - 1];
}
__global__ void __launch_bounds__(256) clb_radix_scatter(const u64* __restrict__ kin, const u32* __restrict__ vin,
                                                         u64* __restrict__ kout, u32* __restrict__ vout, u32 nMax,
                                                         const u32* __restrict__ nDev, u32 shift,
                                                         const u32* __restrict__ hist,
                                                         const u32* __restrict__ digitTotal) {
    __shared__ u32 wcnt[8][256];
    __shared__ u32 dbase[256];
    __shared__ u32 wtot[8];
    const u32 n = devCount(nMax, nDev);
    const u32 nTiles = (n + RS_TILE - 1) / RS_TILE;
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    {
        // exclusive scan of the 256 digit totals
        const u32 v = digitTotal[threadIdx.x];
        u32 incl = v;
        for (u32 o = 1; o < 32; o <<= 1) {
            const u32 t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o)
                incl += t;
        }
        if (lane == 31)
            wtot[warp] = incl;
        __syncthreads();
        u32 woff = 0;
        for (u32 w = 0; w < warp; w++)
            woff += wtot[w];
        dbase[threadIdx.x] = woff + incl - v;
        __syncthreads();
    }
    for (u32 tile = blockIdx.x; tile < nTiles; tile += gridDim.x) {
#pragma unroll
        for (u32 w = 0; w < 8; w++)
            wcnt[w][threadIdx.x] = 0;
        __syncthreads();
        const u32 base = tile * RS_TILE + warp * 256;
        u64 key[8];
        u32 rank[8];
#pragma unroll
        for (u32 q = 0; q < 8; q++) {
            const u32 i = base + q * 32 + lane;
            key[q] = i < n ? kin[i] : 0;
            const u32 d = i < n ? (u32)(key[q] >> shift) & 255u : 0xFFFFFFFFu;
            const unsigned m = __match_any_sync(0xFFFFFFFFu, d);
            u32 r = 0;
            if (d != 0xFFFFFFFFu)
                r = wcnt[warp][d] + (u32)__popc(m & ((1u << lane) - 1u));
            __syncwarp();
            if (d != 0xFFFFFFFFu && (int)lane == __ffs(m) - 1)
                wcnt[warp][d] += (u32)__popc(m);
            __syncwarp();
            rank[q] = r;
        }
        __syncthreads();
        {
            // digit t: place of its first key of this tile, then per warp
            u32 run = dbase[threadIdx.x] + hist[(size_t)threadIdx.x * nTiles + tile];
#pragma unroll
            for (u32 w = 0; w < 8; w++) {
                const u32 c = wcnt[w][threadIdx.x];
                wcnt[w][threadIdx.x] = run;
                run += c;
            }
        }
        __syncthreads();
#pragma unroll
        for (u32 q = 0; q < 8; q++) {
            const u32 i = base + q * 32 + lane;
            if (i < n) {
                const u32 d = (u32)(key[q] >> shift) & 255u;
                const u32 dst = wcnt[warp][d] + rank[q];
                kout[dst] = key[q];
                vout[dst] = vin[i];
            }
        }
        __syncthreads();
    }
}

struct CudaBackend {
    int device;
    int numSMs = 0;
    cudaStream_t stream = nullptr;    // stream in use
    cudaStream_t ownStream = nullptr; // created by this backend
    cudaStream_t copyStream = nullptr; // sliced host-to-device copies
    cudaEvent_t evCopy[8] = {}, evCopyStart = nullptr;
    u32* workCtrs = nullptr; // pool of work counters, one per dynamic launch
    u32* scanSums = nullptr; // tile sums of exclusiveScan
    u32 scanCap = 0;
    u32* radixHist = nullptr; // per-tile digit histograms of sortPairs + 256 digit totals
    size_t radixCap = 0;
    u32 nextCtr = 0;
    static const u32 N_CTRS = 64;
    cudaEvent_t evBegin = nullptr, evEnd = nullptr, evStage[N_STAGES + 1];
    int lastStage = -1;
    u32 launches = 0;
    bool timed = false;

    static int deviceCount() {
        int n = 0;
        if (cudaGetDeviceCount(&n) != cudaSuccess)
            return 0;
        return n;
    }
    explicit CudaBackend(int dev) : device(dev) {
        int n = 0;
        cudaError_t e = cudaGetDeviceCount(&n);
        if (e != cudaSuccess || n == 0)
            throw BackendError(std::string("no CUDA device available (") +
                               (e != cudaSuccess ? cudaGetErrorString(e) : "device count 0") +
                               "); columba_b200 has no CPU fallback");
        if (dev < 0 || dev >= n)
            throw BackendError("invalid CUDA device ordinal " + std::to_string(dev));
        CLB_CUDA(cudaSetDevice(dev));
        CLB_CUDA(cudaDeviceGetAttribute(&numSMs, cudaDevAttrMultiProcessorCount, dev));
        CLB_CUDA(cudaStreamCreateWithFlags(&ownStream, cudaStreamNonBlocking));
        stream = ownStream;
        if (const char* e = getenv("COLUMBA_B200_SEARCH_OCC"))
            warpOcc = atoi(e) == 4 ? 4 : 8;
        if (const char* e = getenv("COLUMBA_B200_HEAVY_GROUP"))
            heavyGroup = atoi(e) == 128 ? 128 : 256;
        if (const char* e = getenv("COLUMBA_B200_HEAVY_PER_SM"))
            heavyPerSM = atoi(e) == 512 ? HEAVY_THREADS_PER_SM : 1024;
        if (const char* e = getenv("COLUMBA_B200_SEARCH_GROUP")) {
            const int v = atoi(e);
            searchGroup = v == 8 ? 8 : (v == 16 ? 16 : 32);
        }
        CLB_CUDA(cudaMalloc(&workCtrs, N_CTRS * sizeof(u32)));
        CLB_CUDA(cudaEventCreate(&evBegin));
        CLB_CUDA(cudaEventCreate(&evEnd));
        for (auto& ev : evStage)
            CLB_CUDA(cudaEventCreate(&ev));
    }
    ~CudaBackend() {
        cudaSetDevice(device);
        if (stream)
            cudaStreamSynchronize(stream);
        if (workCtrs)
            cudaFree(workCtrs);
        if (scanSums)
            cudaFree(scanSums);
        if (radixHist)
            cudaFree(radixHist);
        if (evBegin)
            cudaEventDestroy(evBegin);
        if (evEnd)
            cudaEventDestroy(evEnd);
        for (auto& ev : evStage)
            if (ev)
                cudaEventDestroy(ev);
        if (copyStream) {
            cudaStreamSynchronize(copyStream);
            for (auto& ev : evCopy)
                cudaEventDestroy(ev);
            cudaEventDestroy(evCopyStart);
            cudaStreamDestroy(copyStream);
        }
        if (ownStream)
            cudaStreamDestroy(ownStream);
    }
    void setStream(void* s) {
        CLB_CUDA(cudaSetDevice(device));
        CLB_CUDA(cudaStreamSynchronize(stream));
        stream = s ? (cudaStream_t)s : ownStream;
    }
    void* alloc(size_t n) {
        void* p = nullptr;
        CLB_CUDA(cudaSetDevice(device));
        const cudaError_t e = cudaMalloc(&p, n ? n : 16);
        if (e != cudaSuccess) {
            cudaGetLastError(); // not sticky: the context stays usable, the next launch check must not see it
            throw BackendError(std::string("CUDA error: ") + cudaGetErrorString(e) + " at cudaMalloc of " +
                               std::to_string(n >> 20) + " MB");
        }
        return p;
    }
    void free(void* p) {
        cudaSetDevice(device);
        cudaStreamSynchronize(stream);
        cudaFree(p);
    }
    void* hostAlloc(size_t n) {
        void* p = nullptr;
        CLB_CUDA(cudaSetDevice(device));
        CLB_CUDA(cudaHostAlloc(&p, n ? n : 16, cudaHostAllocDefault));
        return p;
    }
    void hostFree(void* p) { cudaFreeHost(p); }
    void zero(void* p, size_t n) { CLB_CUDA(cudaMemsetAsync(p, 0, n, stream)); }
    void h2d(void* d, const void* s, size_t n) { CLB_CUDA(cudaMemcpyAsync(d, s, n, cudaMemcpyHostToDevice, stream)); }
    void d2h(void* d, const void* s, size_t n) {
        CLB_CUDA(cudaMemcpyAsync(d, s, n, cudaMemcpyDeviceToHost, stream));
        CLB_CUDA(cudaStreamSynchronize(stream));
    }
    void d2hAsync(void* d, const void* s, size_t n) { CLB_CUDA(cudaMemcpyAsync(d, s, n, cudaMemcpyDeviceToHost, stream)); }
    void sync() { CLB_CUDA(cudaStreamSynchronize(stream)); }
    // sliced host-to-device copies on a second stream: the copy of slice s+1 runs
    // while the kernels of slice s do
    void copyBegin() {
        if (!copyStream) {
            CLB_CUDA(cudaStreamCreateWithFlags(&copyStream, cudaStreamNonBlocking));
            for (auto& ev : evCopy)
                CLB_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
            CLB_CUDA(cudaEventCreateWithFlags(&evCopyStart, cudaEventDisableTiming));
        }
        // the copies may overwrite buffers that earlier work on the main stream still reads
        CLB_CUDA(cudaEventRecord(evCopyStart, stream));
        CLB_CUDA(cudaStreamWaitEvent(copyStream, evCopyStart, 0));
    }
    void copySlice(void* d, const void* s, size_t n, u32 slice) {
        if (n)
            CLB_CUDA(cudaMemcpyAsync(d, s, n, cudaMemcpyHostToDevice, copyStream));
        CLB_CUDA(cudaEventRecord(evCopy[slice], copyStream));
    }
    void copyWait(u32 slice) { CLB_CUDA(cudaStreamWaitEvent(stream, evCopy[slice], 0)); }
    u32 blocksFor(int kind) const {
        switch (kind) {
        case WK_SEARCH: return (u32)numSMs * (warpOcc == 8 ? 8u : 4u); // x128 threads, one task per warp
        case WK_SEARCH_HEAVY: return (u32)numSMs * (u32)(heavyPerSM / heavyGroup); // CTAs of heavyGroup threads
        case WK_VERIFY: return (u32)numSMs * 8;
        case WK_FILTER: return (u32)numSMs * 4;
        default: return (u32)numSMs * 16;
        }
    }
    static const int kHeavyGroup = HEAVY_GROUP; // lanes per heavy search task (default)
    int heavyGroup = 256;                       // CTAs of 256 threads, four per SM (COLUMBA_B200_HEAVY_GROUP=128: 128 x 8)
    u32 heavyGroupSize() const { return (u32)heavyGroup; }
    // resident threads per SM of the heavy launch: 1024 (64 registers per thread) or, with
    // COLUMBA_B200_HEAVY_PER_SM=512, 512 (128 registers).  Measured on the 3.1 Gbp workload (search stage per
    // 1 M reads): 512 threads as 4 x 128: 163.4 ms, 1024 as 8 x 128: 156.7 ms, 1024 as 4 x 256: 154.8 ms
    int heavyPerSM = 1024;
    // lanes per task of the first search launch: a warp, or (COLUMBA_B200_SEARCH_GROUP = 16, 8) half or a quarter
    // of one -- most levels of a light task hold fewer nodes than a warp has lanes, and the kernel waits on
    // dependent HBM round trips: smaller groups put 2x / 4x as many tasks in flight with the same registers
    int searchGroup = 32;
    u32 searchGroupSize() const { return (u32)searchGroup; }
    // error word of the batch in flight: kernels return at once when an earlier one raised an overflow bit
    const u32* abortWord = nullptr;
    void setAbortWord(const u32* w) { abortWord = w; }
    // Resident CTAs per SM of the warp-per-task search launch: 8 (64 registers per thread) or, with
    // COLUMBA_B200_SEARCH_OCC=4, 4 (120 registers); 8 measured faster (profiles/r2k_*).  Staging the frontier of the
    // level under expansion in shared memory was built, measured 10 % slower (same profiles) and removed.
    int warpOcc = 8;
    u32 workers(int kind) { return blocksFor(kind) * (kind == WK_SEARCH_HEAVY ? (u32)heavyGroup : 128u); }
    u32 laneStride() const { return 32; }
    u32 fallbackWorkers() const { return 0; }
    template <class K> void launch(const K& k, u32 nMax, const u32* nDev, int kind) {
        if (nMax == 0)
            return;
        void (*fn)(const K, u32, const u32*, u32*, const u32*);
        if constexpr (isWarpTask<K>(0) && groupOfK<K>(0) > 32) {
            constexpr int G = groupOfK<K>(0);
            static_assert(G == 256 || G == 128, "CTA size of the group-task launcher");
            if (nextCtr >= N_CTRS)
                throw BackendError("work counter pool exhausted");
            if (heavyPerSM == 1024)
                clb_kernel_cta<K, G, 1024><<<(u32)numSMs * (u32)(1024 / G), G, 0, stream>>>(k, nMax, nDev, workCtrs + nextCtr++, abortWord);
            else
                clb_kernel_cta<K, G, HEAVY_THREADS_PER_SM><<<(u32)numSMs * (u32)(HEAVY_THREADS_PER_SM / G), G, 0, stream>>>(k, nMax, nDev, workCtrs + nextCtr++, abortWord);
            CLB_CUDA(cudaGetLastError());
            launches++;
            return;
        } else if constexpr (isWarpTask<K>(0))
            fn = warpOcc == 8 ? clb_kernel_warp<K, 8, (groupOfK<K>(0) < 32 ? groupOfK<K>(0) : 32)>
                              : clb_kernel_warp<K, 4, (groupOfK<K>(0) < 32 ? groupOfK<K>(0) : 32)>;
        else if constexpr (isStepped<K>(0))
            fn = clb_kernel_stepped<K>;
        else
            fn = clb_kernel<K>;
        u32 blocks = blocksFor(kind);
        const size_t smem = smemBytesOf(k, 128u, 0);
        if (smem) {
            // never more CTAs than can be resident
            int perSM = 0;
            if (smem > 48 * 1024)
                CLB_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            CLB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, fn, 128, smem));
            if (perSM < 1)
                throw BackendError("kernel does not fit on an SM");
            if ((u32)perSM * (u32)numSMs < blocks)
                blocks = (u32)perSM * (u32)numSMs;
        }
        u32* ctr = nullptr;
        if (kind != WK_LIGHT || isStepped<K>(0)) {
            if (nextCtr >= N_CTRS)
                throw BackendError("work counter pool exhausted");
            ctr = workCtrs + nextCtr++;
        }
        if (kind == WK_LIGHT && !nDev) {
            u32 need = (nMax + 127) / 128;
            if (need < blocks)
                blocks = need;
        }
        fn<<<blocks, 128, smem, stream>>>(k, nMax, nDev, ctr, abortWord);
        CLB_CUDA(cudaGetLastError());
        launches++;
    }
    // exclusive scan of the first n = min(nMax, *nDev) values; out[n] = total
    void exclusiveScan(const u32* in, u32* out, u32 nMax, const u32* nDev) {
        if (nMax == 0)
            return;
        const u32 nb = (nMax + SCAN_TILE - 1) / SCAN_TILE;
        if (nb > scanCap) {
            if (scanSums) {
                CLB_CUDA(cudaStreamSynchronize(stream));
                cudaFree(scanSums);
            }
            scanCap = nb + nb / 2 + 64;
            CLB_CUDA(cudaMalloc(&scanSums, scanCap * sizeof(u32)));
        }
        const u32 grid = tmin<u32>(nb, (u32)numSMs * 8);
        clb_scan_tile_sums<<<grid, 256, 0, stream>>>(in, scanSums, nMax, nDev);
        clb_scan_sums<<<1, 1024, 0, stream>>>(scanSums, nMax, nDev, out);
        clb_scan_tiles<<<grid, 256, 0, stream>>>(in, out, scanSums, nMax, nDev);
        CLB_CUDA(cudaGetLastError());
        launches += 3;
    }
    // radix sort of the first n = min(nMax, *nDev) (key, value) pairs on the low keyBits bits of the keys;
    // the pairs ping-pong between the two buffer pairs, the result is in pair sortResult(keyBits)
    static u32 sortPasses(u32 keyBits) { return (keyBits + 7) / 8; }
    static u32 sortResult(u32 keyBits) { return sortPasses(keyBits) & 1u; }
    void sortPairs(u64* k0, u32* v0, u64* k1, u32* v1, u32 nMax, const u32* nDev, u32 keyBits) {
        if (nMax == 0)
            return;
        const u32 nTiles = (nMax + RS_TILE - 1) / RS_TILE;
        const size_t need = (size_t)nTiles * 256 + 256;
        if (need > radixCap) {
            if (radixHist) {
                CLB_CUDA(cudaStreamSynchronize(stream));
                cudaFree(radixHist);
            }
            radixCap = need + need / 4;
            CLB_CUDA(cudaMalloc(&radixHist, radixCap * sizeof(u32)));
        }
        u32* digitTotal = radixHist + (size_t)nTiles * 256;
        const u32 grid = tmin<u32>(nTiles, (u32)numSMs * 8);
        u64* kk[2] = {k0, k1};
        u32* vv[2] = {v0, v1};
        const u32 passes = sortPasses(keyBits);
        for (u32 p = 0; p < passes; p++) {
            const u32 src = p & 1u, dst = src ^ 1u;
            clb_radix_hist<<<grid, 256, 0, stream>>>(kk[src], nMax, nDev, 8 * p, radixHist);
            clb_radix_scan<<<256, 1024, 0, stream>>>(radixHist, nMax, nDev, digitTotal);
            clb_radix_scatter<<<grid, 256, 0, stream>>>(kk[src], vv[src], kk[dst], vv[dst], nMax, nDev, 8 * p, radixHist,
                                                        digitTotal);
            launches += 3;
        }
        CLB_CUDA(cudaGetLastError());
    }
    void mark(int stage) {
        if (stage >= 0 && stage < N_STAGES) {
            CLB_CUDA(cudaEventRecord(evStage[stage + 1], stream));
            lastStage = stage;
        }
    }
    // a batch that is re-run after a list overflow times the stages of its last attempt
    void restartStages() {
        CLB_CUDA(cudaEventRecord(evStage[0], stream));
        lastStage = -1;
    }
    void beginBatch() {
        CLB_CUDA(cudaSetDevice(device));
        launches = 0;
        nextCtr = 0;
        lastStage = -1;
        CLB_CUDA(cudaMemsetAsync(workCtrs, 0, N_CTRS * sizeof(u32), stream));
        CLB_CUDA(cudaEventRecord(evBegin, stream));
        CLB_CUDA(cudaEventRecord(evStage[0], stream));
        timed = true;
    }
    void endBatch() {
        CLB_CUDA(cudaEventRecord(evEnd, stream));
        CLB_CUDA(cudaEventSynchronize(evEnd));
    }
    void fillTimings(float* stage, float* total, uint32_t* nl) {
        for (int i = 0; i < N_STAGES; i++)
            stage[i] = 0;
        *total = 0;
        *nl = launches;
        if (!timed)
            return;
        // a batch that was re-run after a list overflow records the events of
        // its last run; stage i spans evStage[i] .. evStage[i+1]
        for (int i = 0; i <= lastStage; i++) {
            float ms = 0;
            if (cudaEventElapsedTime(&ms, evStage[i], evStage[i + 1]) == cudaSuccess)
                stage[i] = ms;
        }
        float ms = 0;
        if (cudaEventElapsedTime(&ms, evBegin, evEnd) == cudaSuccess)
            *total = ms;
    }
};

} // namespace clb

#define CLB_BE clb::CudaBackend
#include "clb_api_impl.h"
#include "clb_index_build.h"

// ---- measured ceiling of the path's roofline unit: independent random 32-byte
// rank-record reads from the forward BWT records of the loaded index, 8 in flight per thread
namespace clb {
__global__ void __launch_bounds__(256) clb_gather_kernel(const BwtRec* __restrict__ rec, u32 nRec, u32 perThread,
                                                         u64 seed, u64* sink) {
    const u64 tid = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    u64 x = seed ^ (tid * 0x9E3779B97F4A7C15ull);
    u64 acc = 0;
    for (u32 i = 0; i < perThread; i += 8) {
        u32 idx[8];
#pragma unroll
        for (int q = 0; q < 8; q++) {
            x ^= x << 13;
            x ^= x >> 7;
            x ^= x << 17;
            idx[q] = (u32)(((x >> 32) * (u64)nRec) >> 32);
        }
#pragma unroll
        for (int q = 0; q < 8; q++) {
            const BwtRec r = loadRec(rec + idx[q]);
            acc += r.lo ^ r.hi ^ r.cnt[0] ^ r.cnt[3];
        }
    }
    if (acc == 0x1234567ull)
        *sink = acc;
}
} // namespace clb

extern "C" int clb_gather_ceiling(clb_ctx* c, uint64_t n_lookups, float* ms_out, uint64_t* done_out) {
    if (!c || !ms_out || !done_out || n_lookups == 0)
        return CLB_E_INVALID;
    try {
        using namespace clb;
        CudaBackend& be = *c->be;
        CLB_CUDA(cudaSetDevice(be.device));
        const IndexView ix = c->pipe->ix;
        const u32 nRec = ix.textLength / 64 + 1;
        const u32 blocks = (u32)be.numSMs * 8, threads = blocks * 256;
        u32 per = (u32)((n_lookups + threads - 1) / threads);
        per = (per + 7) & ~7u;
        u64* sink = (u64*)be.alloc(8);
        cudaEvent_t e0, e1;
        CLB_CUDA(cudaEventCreate(&e0));
        CLB_CUDA(cudaEventCreate(&e1));
        clb_gather_kernel<<<blocks, 256, 0, be.stream>>>(ix.fwd.rec, nRec, 64, 1, sink); // warm-up
        CLB_CUDA(cudaEventRecord(e0, be.stream));
        clb_gather_kernel<<<blocks, 256, 0, be.stream>>>(ix.fwd.rec, nRec, per, 0x5DEECE66Dull, sink);
        CLB_CUDA(cudaEventRecord(e1, be.stream));
        CLB_CUDA(cudaEventSynchronize(e1));
        CLB_CUDA(cudaGetLastError());
        CLB_CUDA(cudaEventElapsedTime(ms_out, e0, e1));
        *done_out = (uint64_t)per * threads;
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        be.free(sink);
    } catch (std::exception& e) {
        c->err = e.what();
        return CLB_E_CUDA;
    }
    return CLB_OK;
}

// ---- clb_find_sa_static: clb_find_sa for callers that must not allocate (include/columba_b200.h).
// clb_find_sa stages the rows and the results in two device buffers it allocates per call; this variant
// stages them in the file-scope array below instead: CLB_STATIC_MAX_ROWS rows in, as many results out,
// 8 MB that every GPU which loads this library holds until the process exits.  No counterpart in the
// reference.
//
// `clbStaticScratch` named in HOST code is the host-side stand-in of the symbol, not a device address:
// the address a kernel may use comes from cudaGetSymbolAddress, asked once per device with that device
// current (staticScratchOn), and is what the kernel below is given.
namespace clb {
static const u32 STATIC_MAX_ROWS = CLB_STATIC_MAX_ROWS;
static __device__ __align__(256) u32 clbStaticScratch[2 * (size_t)STATIC_MAX_ROWS];

__global__ void __launch_bounds__(256) clb_find_sa_static_kernel(IndexView ix, const u32* __restrict__ rows,
                                                                  u32* __restrict__ out, u32 n) {
    for (u32 i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        u64 lf = 0;
        out[i] = findSA(ix, rows[i], lf) + 1; // + 1: a slot still holding the 0 it was cleared to was never written
    }
}

static const int STATIC_MAX_DEVICES = 64;
static std::mutex staticScratchMutex;
static u32* staticScratchPtr[STATIC_MAX_DEVICES] = {};

// device address of clbStaticScratch on `device` (which the caller has made current)
static u32* staticScratchOn(int device) {
    if (device < 0 || device >= STATIC_MAX_DEVICES)
        throw BackendError("clb_find_sa_static: device ordinal beyond " + std::to_string(STATIC_MAX_DEVICES));
    std::lock_guard<std::mutex> lock(staticScratchMutex);
    if (!staticScratchPtr[device]) {
        u32* p = nullptr;
        CLB_CUDA(cudaGetSymbolAddress((void**)&p, clbStaticScratch));
        if (!p)
            throw BackendError("clb_find_sa_static: cudaGetSymbolAddress returned a null address");
        staticScratchPtr[device] = p;
    }
    return staticScratchPtr[device];
}
} // namespace clb

extern "C" int clb_find_sa_static(clb_ctx* c, const uint32_t* rows, uint32_t n, uint32_t* out) {
    if (!c)
        return CLB_E_INVALID;
    if (n > CLB_STATIC_MAX_ROWS) {
        c->err = "clb_find_sa_static: " + std::to_string(n) + " rows in one call, the static scratch holds " +
                 std::to_string((unsigned)CLB_STATIC_MAX_ROWS) + " (CLB_STATIC_MAX_ROWS); split the call or use clb_find_sa";
        return CLB_E_OVERFLOW;
    }
    if (n == 0)
        return CLB_OK;
    if (!rows || !out) {
        c->err = "clb_find_sa_static: null rows / out";
        return CLB_E_INVALID;
    }
    const clb::IndexView ix = c->pipe->ix;
    for (uint32_t i = 0; i < n; i++)
        if (rows[i] >= ix.textLength) {
            c->err = "row out of range";
            return CLB_E_INVALID;
        }
    try {
        using namespace clb;
        CudaBackend& be = *c->be;
        CLB_CUDA(cudaSetDevice(be.device));
        u32* dRows = staticScratchOn(be.device);
        u32* dOut = dRows + STATIC_MAX_ROWS;
        // the array keeps what the previous call left: the result slots of this call start at zero
        CLB_CUDA(cudaMemsetAsync(dOut, 0, (size_t)n * sizeof(u32), be.stream));
        CLB_CUDA(cudaMemcpyAsync(dRows, rows, (size_t)n * sizeof(u32), cudaMemcpyHostToDevice, be.stream));
        const u32 blocks = tmin((n + 255u) / 256u, (u32)be.numSMs * 8u);
        clb_find_sa_static_kernel<<<blocks, 256, 0, be.stream>>>(ix, dRows, dOut, n);
        CLB_CUDA(cudaGetLastError());
        CLB_CUDA(cudaMemcpyAsync(out, dOut, (size_t)n * sizeof(u32), cudaMemcpyDeviceToHost, be.stream));
        CLB_CUDA(cudaStreamSynchronize(be.stream));
        for (uint32_t i = 0; i < n; i++) {
            if (out[i] == 0) {
                c->err = "clb_find_sa_static: result slot " + std::to_string(i) + " was not written";
                return CLB_E_INTERNAL;
            }
            out[i]--;
        }
    } catch (std::exception& e) {
        c->err = e.what();
        return CLB_E_CUDA;
    }
    return CLB_OK;
}
