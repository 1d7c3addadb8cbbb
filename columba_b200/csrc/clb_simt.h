// columba_b200 -- warp-cooperative code that also compiles for the CPU
// unit-test harness (tests/hostemu).
//
// A warp-cooperative routine is written as a sequence of lane-parallel regions
//     CLB_LANES(l) { ... per-lane work, lane index l ... }
// separated by warp collectives (vote, scan, sync).  Values that live across
// regions and differ per lane are LV<T> variables, indexed by the lane: on the
// device an LV<T> is one register set per thread and a region runs once in
// every thread of the warp; in the host emulation an LV<T> is an array of 32
// and a region is a loop over the 32 lanes.  Everything outside a region is
// warp-uniform: the device executes it redundantly in all lanes (it must not
// have side effects other than writing the same value), the host once.
#pragma once
#include "clb_common.h"

namespace clb {

enum { GRP = 32 };

#if defined(__CUDA_ARCH__)
template <typename T> struct LV {
    T v;
    __device__ __forceinline__ T& operator[](int) { return v; }
    __device__ __forceinline__ const T& operator[](int) const { return v; }
};
#define CLB_LANES(l) for (int l = (int)(threadIdx.x & 31u), l##_go = 1; l##_go; l##_go = 0)
__device__ __forceinline__ void grpSync() { __syncwarp(); }
__device__ __forceinline__ u32 grpBallot(const LV<u32>& p) { return __ballot_sync(0xFFFFFFFFu, p.v != 0); }
// exclusive prefix sum over the lanes; returns the total
__device__ __forceinline__ u32 grpExScan(const LV<u32>& x, LV<u32>& out) {
    const u32 lane = threadIdx.x & 31u;
    u32 incl = x.v;
#pragma unroll
    for (u32 o = 1; o < 32; o <<= 1) {
        const u32 t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o)
            incl += t;
    }
    out.v = incl - x.v;
    return __shfl_sync(0xFFFFFFFFu, incl, 31);
}
__device__ __forceinline__ u32 grpSum(const LV<u32>& x) {
    u32 s = x.v;
#pragma unroll
    for (u32 o = 16; o > 0; o >>= 1)
        s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
    return s;
}
__device__ __forceinline__ u32 grpMax(const LV<u32>& x) {
    u32 s = x.v;
#pragma unroll
    for (u32 o = 16; o > 0; o >>= 1) {
        const u32 t = __shfl_xor_sync(0xFFFFFFFFu, s, o);
        s = t > s ? t : s;
    }
    return s;
}
__device__ __forceinline__ u32 grpBcast(const LV<u32>& x, int src) { return __shfl_sync(0xFFFFFFFFu, x.v, src); }
__device__ __forceinline__ u32 atomMinU32(u32* p, u32 v) { return atomicMin(p, v); }
// The same collectives for a group of GS lanes.  GS == 32 is a warp.  A larger GS is the whole CTA
// (blockDim.x == GS), whose collectives go through shared memory and __syncthreads.  A smaller GS (8, 16) is an
// aligned part of a warp: the warp then runs 32 / GS independent groups, each with its own task and its own
// control flow; their collectives name only the group's lanes (shuffles of width GS, __syncwarp on the group's
// mask), so the groups of a warp never wait for each other.  A routine written with these runs unchanged with
// any group.
template <int GS> __device__ __forceinline__ int grpLaneG() {
    return GS > 32 ? (int)threadIdx.x : (int)(threadIdx.x & (u32)((GS < 32 ? GS : 32) - 1));
}
template <int GS> __device__ __forceinline__ unsigned grpMaskG() {
    if (GS >= 32)
        return 0xFFFFFFFFu;
    return ((1u << (GS & 31)) - 1u) << ((threadIdx.x & 31u) & ~(u32)((GS & 31) - 1));
}
#define CLB_LANES_G(l, GS) for (int l = grpLaneG<GS>(), l##_go = 1; l##_go; l##_go = 0)
template <int GS> __device__ __forceinline__ void grpSyncG() {
    if (GS <= 32)
        __syncwarp(grpMaskG<GS>());
    else
        __syncthreads();
}
template <int GS> __device__ __forceinline__ u32 grpExScanG(const LV<u32>& x, LV<u32>& out) {
    if (GS == 32)
        return grpExScan(x, out);
    if (GS < 32) {
        constexpr int G = GS < 32 ? GS : 32;
        const unsigned m = grpMaskG<GS>();
        const u32 lane = threadIdx.x & (u32)(G - 1);
        u32 incl = x.v;
#pragma unroll
        for (u32 o = 1; o < (u32)G; o <<= 1) {
            const u32 t = __shfl_up_sync(m, incl, o, G);
            if (lane >= o)
                incl += t;
        }
        out.v = incl - x.v;
        return __shfl_sync(m, incl, G - 1, G);
    }
    __shared__ u32 wsum[GS / 32 + 1];
    const u32 lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    u32 incl = x.v;
#pragma unroll
    for (u32 o = 1; o < 32; o <<= 1) {
        const u32 t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o)
            incl += t;
    }
    if (lane == 31)
        wsum[warp] = incl;
    __syncthreads();
    u32 before = 0, total = 0;
#pragma unroll
    for (u32 w = 0; w < (u32)(GS / 32 + (GS < 32 ? 1 : 0)); w++) {
        const u32 v = wsum[w];
        before += w < warp ? v : 0u;
        total += v;
    }
    __syncthreads();
    out.v = before + incl - x.v;
    return total;
}
template <int GS> __device__ __forceinline__ u32 grpBcastG(const LV<u32>& x, int src) {
    if (GS == 32)
        return grpBcast(x, src);
    if (GS < 32)
        return __shfl_sync(grpMaskG<GS>(), x.v, src, GS < 32 ? GS : 32);
    __shared__ u32 slot;
    if ((int)threadIdx.x == src)
        slot = x.v;
    __syncthreads();
    const u32 v = slot;
    __syncthreads();
    return v;
}
#else
enum { GRP_MAX = 128 }; // largest group the host emulation runs (a CTA of 128 lanes)
template <typename T> struct LV {
    T v[GRP_MAX];
    T& operator[](int l) { return v[l]; }
    const T& operator[](int l) const { return v[l]; }
};
#define CLB_LANES(l) for (int l = 0; l < GRP; l++)
inline void grpSync() {}
inline u32 grpBallot(const LV<u32>& p) {
    u32 m = 0;
    for (int l = 0; l < GRP; l++)
        m |= (p.v[l] != 0 ? 1u : 0u) << l;
    return m;
}
inline u32 grpExScan(const LV<u32>& x, LV<u32>& out) {
    u32 run = 0;
    for (int l = 0; l < GRP; l++) {
        const u32 v = x.v[l];
        out.v[l] = run;
        run += v;
    }
    return run;
}
inline u32 grpSum(const LV<u32>& x) {
    u32 s = 0;
    for (int l = 0; l < GRP; l++)
        s += x.v[l];
    return s;
}
inline u32 grpMax(const LV<u32>& x) {
    u32 s = 0;
    for (int l = 0; l < GRP; l++)
        s = x.v[l] > s ? x.v[l] : s;
    return s;
}
inline u32 grpBcast(const LV<u32>& x, int src) { return x.v[src]; }
// the CPU harness emulates a group of GS <= GRP_MAX lanes
#define CLB_LANES_G(l, GS) for (int l = 0; l < (GS); l++)
template <int GS> inline void grpSyncG() { static_assert(GS <= GRP_MAX, "group size of the host emulation"); }
template <int GS> inline u32 grpExScanG(const LV<u32>& x, LV<u32>& out) {
    u32 run = 0;
    for (int l = 0; l < GS; l++) {
        const u32 v = x.v[l];
        out.v[l] = run;
        run += v;
    }
    return run;
}
template <int GS> inline u32 grpBcastG(const LV<u32>& x, int src) { return grpBcast(x, src); }
inline u32 atomMinU32(u32* p, u32 v) {
    const u32 o = *p;
    if (v < o)
        *p = v;
    return o;
}
#endif

} // namespace clb
