// columba_b200 -- implementation of the C ABI (include/columba_b200.h) on top
// of Pipeline<BE>.  Included exactly once by the translation unit that defines
// the backend type CLB_BE (clb_api.cu for the product; tests/hostemu for the
// CPU unit-test harness of the kernel logic).
#pragma once
#include "../../include/columba_b200.h"
#include "clb_index_load.h"
#include "clb_pipeline.h"
#include "clb_strategy.h"
#include <memory>
#include <string>

static_assert(sizeof(clb_occ) == sizeof(clb::OutOcc), "clb_occ layout");
static_assert((int)CLB_N_COUNTERS == (int)clb::N_COUNTERS && (int)CLB_N_STATS == (int)clb::N_STATS &&
                  (int)CLB_N_STAGES == (int)clb::N_STAGES,
              "enum sizes");

struct clb_index {
    std::unique_ptr<CLB_BE> be;
    std::unique_ptr<clb::DeviceIndex<CLB_BE>> dix;
    int device = 0;
};
struct clb_ctx {
    clb_index* index = nullptr;
    std::unique_ptr<CLB_BE> be;
    std::unique_ptr<clb::Pipeline<CLB_BE>> pipe;
    clb::HostStrategy strategy;
    bool haveStrategy = false;
    clb::DevStrategy flat[clb::CLB_MAX_K + 1];
    int uploadedK = -1;
    clb::MapOptions opt;
    clb::MapResult<CLB_BE> res;
    std::string err;
    ~clb_ctx() {
        if (be)
            res.release(*be);
    }
};
static thread_local std::string g_clbLoadError;

extern "C" {

int clb_abi_version(void) { return CLB_ABI_VERSION; }
size_t clb_sizeof_result(void) { return sizeof(clb_result); }
int clb_device_count(void) { return CLB_BE::deviceCount(); }

int clb_index_load(const char* base, int sa_sparse, int kmer, int in_text_switch, int no_cigar, int device,
                   clb_index** out) {
    if (!base || !out || sa_sparse <= 0 || kmer < 0 || in_text_switch < 0) {
        g_clbLoadError = "clb_index_load: invalid argument";
        return CLB_E_INVALID;
    }
    *out = nullptr;
    std::unique_ptr<clb_index> h(new clb_index());
    try {
        h->device = device;
        h->be.reset(new CLB_BE(device));
    } catch (std::exception& e) {
        g_clbLoadError = e.what();
        return CLB_E_CUDA;
    }
    try {
        h->dix.reset(new clb::DeviceIndex<CLB_BE>(*h->be));
        h->dix->load(base, (clb::u32)sa_sparse, (clb::u32)kmer, (clb::u32)in_text_switch, no_cigar != 0);
    } catch (clb::BackendError& e) {
        g_clbLoadError = e.what();
        return CLB_E_CUDA;
    } catch (std::exception& e) {
        g_clbLoadError = e.what();
        return CLB_E_IO;
    }
    *out = h.release();
    return CLB_OK;
}
void clb_index_free(clb_index* h) { delete h; }
uint32_t clb_index_text_length(const clb_index* h) { return h ? h->dix->meta.textLength : 0; }
uint32_t clb_index_num_seqs(const clb_index* h) { return h ? (uint32_t)h->dix->meta.seqStart.size() - 1 : 0; }
const char* clb_index_seq_name(const clb_index* h, uint32_t i) {
    if (!h || i >= h->dix->meta.seqNames.size())
        return "";
    return h->dix->meta.seqNames[i].c_str();
}
uint32_t clb_index_seq_start(const clb_index* h, uint32_t i) {
    if (!h || i >= h->dix->meta.seqStart.size())
        return 0;
    return h->dix->meta.seqStart[i];
}
uint64_t clb_index_device_bytes(const clb_index* h) { return h ? h->dix->meta.deviceBytes : 0; }

int clb_ctx_create(clb_index* ix, clb_ctx** out) {
    if (!ix || !out)
        return CLB_E_INVALID;
    *out = nullptr;
    std::unique_ptr<clb_ctx> c(new clb_ctx());
    try {
        c->index = ix;
        c->be.reset(new CLB_BE(ix->device));
        c->pipe.reset(new clb::Pipeline<CLB_BE>(*c->be));
        c->pipe->ix = ix->dix->view;
    } catch (std::exception& e) {
        g_clbLoadError = e.what();
        return CLB_E_CUDA;
    }
    *out = c.release();
    return CLB_OK;
}
void clb_ctx_free(clb_ctx* c) { delete c; }
const char* clb_last_error(const clb_ctx* c) { return c ? c->err.c_str() : g_clbLoadError.c_str(); }

int clb_ctx_set_strategy(clb_ctx* c, const char* kind, const char* arg, const char* data_dir) {
    if (!c || !kind || !arg)
        return CLB_E_INVALID;
    try {
        std::string k = kind, dd = data_dir ? data_dir : "";
        clb::HostStrategy st;
        if (k == "named")
            st = clb::namedStrategy(arg, dd);
        else if (k == "custom")
            st = clb::customStrategy(arg, 50);
        else if (k == "custom_dynamic")
            st = clb::dynamicFrom(clb::customStrategy(arg, 50));
        else if (k == "multi")
            st = clb::multiStrategy(arg);
        else {
            c->err = "unknown strategy kind " + k;
            return CLB_E_INVALID;
        }
        for (clb::u32 kk = 1; kk <= clb::CLB_MAX_K; kk++) {
            memset(&c->flat[kk], 0, sizeof(clb::DevStrategy));
            if (st.supports(kk))
                c->flat[kk] = st.flatten(kk);
        }
        c->strategy = st;
        c->haveStrategy = true;
        c->uploadedK = -1;
    } catch (std::exception& e) {
        c->err = e.what();
        return CLB_E_IO;
    }
    return CLB_OK;
}
int clb_ctx_set_options(clb_ctx* c, int metric, int partitioning) {
    if (!c || (metric != CLB_METRIC_EDIT && metric != CLB_METRIC_HAMMING) || partitioning < 0 || partitioning > 2)
        return CLB_E_INVALID;
    c->opt.metric = (clb::u32)metric;
    c->opt.partStrategy = (clb::u32)partitioning;
    return CLB_OK;
}
int clb_ctx_set_trim_max(clb_ctx* c, int largest_stratum) {
    if (!c || largest_stratum < 0 || largest_stratum > (int)clb::CLB_MAX_K)
        return CLB_E_INVALID;
    c->opt.trimMax = (clb::u32)largest_stratum;
    return CLB_OK;
}
int clb_ctx_set_filter_per_strand(clb_ctx* c, int per_strand) {
    if (!c)
        return CLB_E_INVALID;
    c->opt.perStrand = per_strand ? 1u : 0u;
    return CLB_OK;
}
int clb_ctx_set_stream(clb_ctx* c, void* cuda_stream) {
    if (!c)
        return CLB_E_INVALID;
    try {
        c->be->setStream(cuda_stream);
    } catch (std::exception& e) {
        c->err = e.what();
        return CLB_E_CUDA;
    }
    return CLB_OK;
}
int clb_ctx_supports_distance(const clb_ctx* c, int k) {
    return c && c->haveStrategy && k >= 1 && k <= clb::CLB_MAX_K && c->flat[k].supported;
}
int clb_ctx_num_parts(const clb_ctx* c, int k) { return clb_ctx_supports_distance(c, k) ? c->flat[k].P : 0; }

static int clbFinish(clb_ctx* c, clb::u32 err, uint32_t n_reads, clb_result* out, bool fetched) {
    using namespace clb;
    if (err & ~(u32)ERR_SHORT_READ) {
        c->err = "device error bits " + std::to_string(err) +
                 ((err & (ERR_CAND_OVERFLOW | ERR_FMOCC_OVERFLOW | ERR_TEXTOCC_OVERFLOW | ERR_OUT_OVERFLOW))
                      ? " (occurrence lists exceed the device capacity)"
                      : " (internal capacity exceeded in the search kernel)");
        return (err & ERR_INTERNAL) ? CLB_E_INTERNAL : CLB_E_OVERFLOW;
    }
    if (out) {
        memset(out, 0, sizeof(*out));
        out->n_reads = n_reads;
        out->n_groups = c->opt.perStrand ? 2 * n_reads : n_reads;
        memcpy(out->counters, c->res.counters, sizeof(out->counters));
        memcpy(out->stats, c->res.stats, sizeof(out->stats));
        out->cigar_stride = c->res.cigW;
        if (fetched) {
            out->n_occ = (uint32_t)c->res.occ.size();
            out->occ = (const clb_occ*)c->res.occ.data();
            out->cigar = c->res.cig.data();
            out->read_start = c->res.outStart.data();
            out->read_count = c->res.outCount.data();
            out->read_status = c->res.readStatus.data();
            out->n_short_reads = c->res.nShort;
        }
        c->be->fillTimings(out->stage_ms, &out->total_ms, &out->gpu_launches);
    }
    if (err & ERR_SHORT_READ) // not an error of the batch: read_status tells which reads were skipped
        c->err = "The pattern size is too short for the current number of parts (the naive fallback of "
                 "src/searchstrategy.cpp:442-459 is not implemented on the device); such reads got no hits";
    return CLB_OK;
}
static int clbPrepare(clb_ctx* c, int max_dist, int min_dist) {
    if (max_dist == 0) { // exact stratum: no search scheme involved (src/searchstrategy.cpp:499-510)
        memset(&c->flat[0], 0, sizeof(clb::DevStrategy));
        c->flat[0].supported = 1;
        c->flat[0].P = 1;
        c->opt.k = 0;
        c->opt.minD = 0;
        c->opt.noCigar = c->index->dix->view.noCigar;
        if (c->uploadedK != 0) {
            c->pipe->uploadStrategy(c->flat[0]);
            c->uploadedK = 0;
        }
        return CLB_OK;
    }
    if (!c->haveStrategy) {
        c->err = "no search strategy set (clb_ctx_set_strategy)";
        return CLB_E_INVALID;
    }
    if (max_dist < 1 || max_dist > clb::CLB_MAX_K || !c->flat[max_dist].supported) {
        c->err = "The search scheme " + c->strategy.name + " does not support distance " + std::to_string(max_dist);
        return CLB_E_UNSUPPORTED;
    }
    if (c->opt.metric == clb::METRIC_EDIT && max_dist > (int)clb::BPC<clb::u128>::MAX_ED) {
        c->err = "maximal edit distance too large";
        return CLB_E_UNSUPPORTED;
    }
    c->opt.k = (clb::u32)max_dist;
    c->opt.minD = (clb::u32)(min_dist < 0 ? 0 : min_dist);
    c->opt.noCigar = c->index->dix->view.noCigar;
    if (c->uploadedK != max_dist) {
        c->pipe->uploadStrategy(c->flat[max_dist]);
        c->uploadedK = max_dist;
    }
    return CLB_OK;
}

int clb_map_batch(clb_ctx* c, const uint8_t* reads, const uint32_t* offsets, uint32_t n_reads, int max_dist,
                  int min_dist, clb_result* out) {
    if (!c || !reads || !offsets || !out || n_reads == 0)
        return CLB_E_INVALID;
    try {
        int rc = clbPrepare(c, max_dist, min_dist);
        if (rc)
            return rc;
        clb::u32 maxLen = 0;
        for (uint32_t i = 0; i < n_reads; i++) {
            if (offsets[i + 1] < offsets[i]) {
                c->err = "offsets must be non-decreasing";
                return CLB_E_INVALID;
            }
            maxLen = std::max(maxLen, offsets[i + 1] - offsets[i]);
        }
        if (maxLen > 16000) {
            c->err = "reads longer than 16000 are not supported";
            return CLB_E_UNSUPPORTED;
        }
        c->be->beginBatch();
        clb::u32 err = c->pipe->run(c->flat[max_dist], reads, offsets, n_reads, maxLen, offsets[n_reads], false,
                                    c->opt, &c->res, true);
        c->be->endBatch();
        return clbFinish(c, err, n_reads, out, true);
    } catch (clb::BackendError& e) {
        c->err = e.what();
        return CLB_E_CUDA;
    } catch (std::exception& e) {
        c->err = e.what();
        return CLB_E_INTERNAL;
    }
}
int clb_map_batch_device(clb_ctx* c, const uint8_t* d_reads, const uint32_t* d_offsets, uint32_t n_reads,
                         uint32_t max_read_len, int max_dist, int min_dist, int fetch, clb_result* out) {
    if (!c || !d_reads || !d_offsets || !out || n_reads == 0 || max_read_len == 0)
        return CLB_E_INVALID;
    try {
        int rc = clbPrepare(c, max_dist, min_dist);
        if (rc)
            return rc;
        if (max_read_len > 16000) {
            c->err = "reads longer than 16000 are not supported";
            return CLB_E_UNSUPPORTED;
        }
        c->be->beginBatch();
        // every per-read buffer is sized from max_read_len: check it against the offsets before anything is written
        const clb::u32 longest = c->pipe->maxReadLen(d_offsets, n_reads);
        if (longest > max_read_len) {
            c->be->endBatch();
            c->err = "a read of " + std::to_string(longest) + " characters exceeds max_read_len = " + std::to_string(max_read_len);
            return CLB_E_INVALID;
        }
        clb::u32 err = c->pipe->run(c->flat[max_dist], d_reads, d_offsets, n_reads, max_read_len, 0, true, c->opt,
                                    &c->res, fetch != 0);
        c->be->endBatch();
        return clbFinish(c, err, n_reads, out, fetch != 0);
    } catch (clb::BackendError& e) {
        c->err = e.what();
        return CLB_E_CUDA;
    } catch (std::exception& e) {
        c->err = e.what();
        return CLB_E_INTERNAL;
    }
}

} // extern "C"

// ---- unit-level entry points ----------------------------------------------
template <typename K, typename Fill>
static int clbMicro(clb_ctx* c, const void* in, size_t inBytes, void* outp, size_t outBytes, uint32_t n, Fill fill) {
    if (!c || !in || !outp)
        return CLB_E_INVALID;
    if (n == 0)
        return CLB_OK;
    try {
        CLB_BE& be = *c->be;
        void* dIn = be.alloc(inBytes);
        void* dOut = be.alloc(outBytes);
        be.h2d(dIn, in, inBytes);
        K k = fill(dIn, dOut);
        be.launch(k, n, nullptr, clb::WK_LIGHT);
        be.d2h(outp, dOut, outBytes);
        be.free(dIn);
        be.free(dOut);
    } catch (std::exception& e) {
        c->err = e.what();
        return CLB_E_CUDA;
    }
    return CLB_OK;
}

extern "C" {

int clb_occ4(clb_ctx* c, int reverse, const uint32_t* pos, uint32_t n, uint32_t* out4) {
    if (!c)
        return CLB_E_INVALID;
    clb::IndexView ix = c->pipe->ix;
    for (uint32_t i = 0; pos && i < n; i++)
        if (pos[i] > ix.textLength) {
            c->err = "position out of range";
            return CLB_E_INVALID;
        }
    return clbMicro<clb::Occ4Kernel>(c, pos, (size_t)n * 4, out4, (size_t)n * 16, n, [&](void* a, void* b) {
        return clb::Occ4Kernel{ix, (clb::u32)(reverse != 0), (const clb::u32*)a, (clb::u32*)b};
    });
}
int clb_extend(clb_ctx* c, int mode, const uint32_t* parents, uint32_t n, uint32_t* out16) {
    if (!c || mode < 0 || mode > 2)
        return CLB_E_INVALID;
    clb::IndexView ix = c->pipe->ix;
    return clbMicro<clb::ExtendKernel>(c, parents, (size_t)n * 16, out16, (size_t)n * 64, n, [&](void* a, void* b) {
        return clb::ExtendKernel{ix, (clb::u32)mode, (const clb::RangePair*)a, (clb::RangePair*)b};
    });
}
int clb_find_sa(clb_ctx* c, const uint32_t* rows, uint32_t n, uint32_t* out) {
    if (!c)
        return CLB_E_INVALID;
    clb::IndexView ix = c->pipe->ix;
    for (uint32_t i = 0; rows && i < n; i++)
        if (rows[i] >= ix.textLength) {
            c->err = "row out of range";
            return CLB_E_INVALID;
        }
    return clbMicro<clb::FindSAKernel>(c, rows, (size_t)n * 4, out, (size_t)n * 4, n, [&](void* a, void* b) {
        return clb::FindSAKernel{ix, (const clb::u32*)a, (clb::u32*)b};
    });
}
int clb_kmer_lookup(clb_ctx* c, const uint32_t* codes, uint32_t n, uint32_t* out4) {
    if (!c || !codes || !out4)
        return CLB_E_INVALID;
    try {
        clb::IndexView ix = c->pipe->ix;
        for (uint32_t i = 0; i < n; i++) {
            if ((uint64_t)codes[i] >= ((uint64_t)1 << (2 * ix.K))) {
                c->err = "k-mer code out of range";
                return CLB_E_INVALID;
            }
            c->be->d2h(out4 + 4 * (size_t)i, ix.kmer + codes[i], 16);
        }
    } catch (std::exception& e) {
        c->err = e.what();
        return CLB_E_CUDA;
    }
    return CLB_OK;
}

} // extern "C"
