// TEST HARNESS (not product code): compiles the kernels' task functors of
// columba_b200/csrc for the CPU and runs them in a plain loop behind the same
// C ABI, so the kernel LOGIC can be unit-tested in the build container, which
// has no GPU.  The shipped library (columba_b200/libcolumba_b200.so) is built
// from clb_api.cu with the CUDA backend only and never contains this.
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>

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
    u32 sortBigSegments(const Batch&, u32) { return 0xFFFFFFFFu; }
    u32 sortBigMax() const { return 0xFFFFFFFFu; }
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
    void exclusiveScan(const u32* in, u32* out, u32 n) {
        u32 run = 0;
        for (u32 i = 0; i < n; i++) {
            u32 v = in[i];
            out[i] = run;
            run += v;
        }
        out[n] = run;
    }
    void mark(int) {}
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
