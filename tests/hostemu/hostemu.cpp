// TEST HARNESS (not product code): compiles the kernels' task functors of
// columba_b200/csrc for the CPU and runs them in a plain loop behind the same
// C ABI, so the kernel LOGIC can be unit-tested in the build container, which
// has no GPU.  The shipped library (columba_b200/libcolumba_b200.so) is built
// from clb_api.cu with the CUDA backend only and never contains this.
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <algorithm>
#include <string>
#include <utility>
#include <vector>

#include "../../columba_b200/csrc/clb_pipeline.h"

namespace clb {
struct HostEmuBackend {
    explicit HostEmuBackend(int) {}
    static int deviceCount() { return 1; }
    void* alloc(size_t n) {
        void* p = std::calloc(n ? n : 1, 1);
        if (!p)
            throw BackendError("host emulation: out of memory");
        return p;
    }
    void free(void* p) { std::free(p); }
    void zero(void* p, size_t n) { std::memset(p, 0, n); }
    void h2d(void* d, const void* s, size_t n) { std::memcpy(d, s, n); }
    void d2h(void* d, const void* s, size_t n) { std::memcpy(d, s, n); }
    void sync() {}
    void d2hAsync(void* d, const void* s, size_t n) { std::memcpy(d, s, n); }
    void copyBegin() {}
    void copySlice(void* d, const void* s, size_t n, u32) { std::memcpy(d, s, n); }
    void copyWait(u32) {}
    u32 workers(int) { return 1; }
    u32 laneStride() const { return 1; }
    void* hostAlloc(size_t n) { return std::malloc(n ? n : 16); }
    void hostFree(void* p) { std::free(p); }
    u32 fallbackWorkers() const { return 1; }
    template <class K> static constexpr auto isWarpTask(int) -> decltype(K::kWarpTask) { return K::kWarpTask; }
    template <class K> static constexpr bool isWarpTask(long) { return false; }
    template <class K> void launch(const K& k, u32 nMax, const u32* nDev, int) {
        u32 n = nDev ? (*nDev < nMax ? *nDev : nMax) : nMax;
        if (abortWord && (*abortWord & ERR_ABORT_MASK))
            return;
        typename K::State st;
        k.begin(st, 0);
        for (u32 i = 0; i < n; i++) {
            if constexpr (isWarpTask<K>(0))
                k.runWarp(st, i); // the 32 lanes of the warp are emulated inside (clb_simt.h)
            else
                k.run(st, i);
        }
        k.end(st);
        launches++;
    }
    static const int kHeavyGroup = 128; // lanes per heavy search task, as on the GPU
    u32 heavyGroupSize() const { return 128; }
    // lanes per task of the first search launch (8, 16 or 32), as in the CUDA backend
    u32 searchGroupSize() const {
        const char* e = getenv("COLUMBA_B200_SEARCH_GROUP");
        const int v = e ? atoi(e) : 0;
        return v == 8 ? 8u : (v == 16 ? 16u : 32u);
    }
    const u32* abortWord = nullptr;
    void setAbortWord(const u32* w) { abortWord = w; }
    static u32 sortResult(u32) { return 0; }
    void sortPairs(u64* k0, u32* v0, u64*, u32*, u32 nMax, const u32* nDev, u32) {
        const u32 n = nDev ? (*nDev < nMax ? *nDev : nMax) : nMax;
        std::vector<std::pair<u64, u32>> a(n);
        for (u32 i = 0; i < n; i++)
            a[i] = std::make_pair(k0[i], v0[i]);
        std::stable_sort(a.begin(), a.end(),
                         [](const std::pair<u64, u32>& x, const std::pair<u64, u32>& y) { return x.first < y.first; });
        for (u32 i = 0; i < n; i++) {
            k0[i] = a[i].first;
            v0[i] = a[i].second;
        }
    }
    void exclusiveScan(const u32* in, u32* out, u32 nMax, const u32* nDev) {
        const u32 n = nDev ? (*nDev < nMax ? *nDev : nMax) : nMax;
        u32 run = 0;
        for (u32 i = 0; i < n; i++) {
            u32 v = in[i];
            out[i] = run;
            run += v;
        }
        out[n] = run;
    }
    void mark(int) {}
    void restartStages() {}
    void setStream(void*) {}
    void beginBatch() { launches = 0; }
    void endBatch() {}
    void fillTimings(float* stage, float* total, uint32_t* nl) {
        for (int i = 0; i < N_STAGES; i++)
            stage[i] = 0;
        *total = 0;
        *nl = launches;
    }
    u32 launches = 0;
};
} // namespace clb

#define CLB_BE clb::HostEmuBackend
#include "../../columba_b200/csrc/clb_api_impl.h"
#include "../../columba_b200/csrc/clb_index_build.h"
