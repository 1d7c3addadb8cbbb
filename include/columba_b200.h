/* columba_b200 -- C ABI of the B200 approximate-matching path.
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++ or torch
 * types.  Everything the reference does between "a chunk of reads" and "a
 * vector of TextOcc per read" for ALL mode (src/parallel.cpp:67-78 ->
 * SearchStrategy::matchApprox, src/searchstrategy.h:2021) runs behind
 * clb_map_batch() on one GPU.  The library has NO CPU execution path: every
 * entry point that computes needs a CUDA device and fails with an error code
 * otherwise.
 *
 * Conventions: every function returns 0 on success or a negative CLB_E_* code;
 * clb_last_error() gives the message (the reference throws std::runtime_error
 * with the same texts; the C++ shim in columba_b200/host rethrows them).
 * Handles are not thread-safe; use one clb_ctx per host thread (the reference
 * keeps its scratch in thread_local members, src/indexinterface.h:62-109); a
 * clb_index is immutable and may be shared by any number of contexts.
 */
#ifndef COLUMBA_B200_H
#define COLUMBA_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CLB_ABI_VERSION 4

enum {
    CLB_OK = 0,
    CLB_E_INVALID = -1,   /* bad argument */
    CLB_E_IO = -2,        /* index / scheme file problem (message = reference's) */
    CLB_E_CUDA = -3,      /* no device, CUDA failure */
    CLB_E_UNSUPPORTED = -4,
    CLB_E_OVERFLOW = -5,  /* internal capacity exceeded */
    CLB_E_INTERNAL = -6
};

/* values follow the reference's enums (src/definitions.h) */
enum { CLB_METRIC_HAMMING = 0, CLB_METRIC_EDIT = 1 };
enum { CLB_PART_UNIFORM = 0, CLB_PART_STATIC = 1, CLB_PART_DYNAMIC = 2 };

/* indices into counters[] == Counters::CounterType, src/indexhelpers.h:1848-1883 */
enum {
    CLB_CTR_NODE_COUNTER, CLB_CTR_TOTAL_REPORTED_POSITIONS, CLB_CTR_CIGARS_IN_INDEX, CLB_CTR_IN_TEXT_STARTED,
    CLB_CTR_ABORTED_IN_TEXT_VERIF, CLB_CTR_CIGARS_IN_TEXT_VERIFICATION, CLB_CTR_IMMEDIATE_SWITCH,
    CLB_CTR_SEARCH_STARTED, CLB_CTR_DROPPED_UNIQUE_MATCHES, CLB_CTR_NUMBER_OF_READS, CLB_CTR_TOTAL_UNIQUE_MATCHES,
    CLB_CTR_MAPPED_READS, CLB_CTR_TOTAL_UNIQUE_PAIRS, CLB_CTR_MAPPED_PAIRS, CLB_CTR_DISCORDANTLY_MAPPED_PAIRS,
    CLB_CTR_MAPPED_HALF_PAIRS, CLB_CTR_UNPAIRED_BUT_MAPPED_PAIRS, CLB_N_COUNTERS
};
/* device statistics next to the reference's counters */
enum {
    CLB_STAT_RANK_LOOKUPS, CLB_STAT_LF_STEPS, CLB_STAT_VERIFY_ROWS, CLB_STAT_SEARCH_ROWS, CLB_STAT_SEARCH_TASKS,
    CLB_STAT_CANDIDATES, CLB_STAT_FM_OCCS, CLB_STAT_TEXT_OCCS, CLB_STAT_SEARCH_LOOKUPS, CLB_STAT_UNIQUE_CANDIDATES,
    CLB_STAT_HEAVY_TASKS, CLB_N_STATS
};
enum { CLB_N_STAGES = 6 }; /* prep, partition, search, verify, fm-expand, text-filter */

typedef struct clb_index clb_index;
typedef struct clb_ctx clb_ctx;

/* One occurrence == one TextOcc the reference would hand to the SAM writer
 * (src/indexhelpers.h:289-930) after findSeqName (src/indexinterface.cpp:799). */
typedef struct clb_occ {
    uint32_t read;        /* index of the read in the batch */
    uint32_t seq_id;      /* reference sequence (index into clb_index_seq_name) */
    uint32_t pos;         /* 0-based begin within the sequence (SAM POS - 1) */
    uint32_t end;         /* end (exclusive) within the sequence */
    uint32_t index_begin; /* begin in the concatenated text */
    uint8_t strand;       /* 0 forward, 1 reverse complement */
    uint8_t dist;         /* edit / Hamming distance (NM) */
    uint8_t n_cigar;      /* number of CIGAR runs in cigar[i * cigar_stride ...] */
    uint8_t trimmed;      /* 0, or 1 + the distance before clipping: FOUND_WITH_TRIMMING at a sequence
                             boundary (dist, pos, end, cigar describe the clipped alignment) */
} clb_occ;

/* Result of one batch; arrays are owned by the context and stay valid until
 * the next clb_map_batch* call on it.  The occurrences of read r are
 * occ[read_start[r] .. read_start[r] + read_count[r]) in text order (sorted by
 * index_begin; the reference sorts by distance just before writing SAM). */
typedef struct clb_result {
    uint32_t n_reads;
    uint32_t n_groups;            /* entries of read_start / read_count: n_reads, or 2 * n_reads with group
                                     2 * read + strand after clb_ctx_set_filter_per_strand(ctx, 1) */
    uint32_t n_occ;               /* entries in occ[]: read by read, no gaps */
    const clb_occ* occ;
    const uint16_t* cigar;        /* n_occ x cigar_stride, run = (len << 2) | op, op 0=M 1=I 2=D */
    uint32_t cigar_stride;
    const uint32_t* read_start;   /* n_groups */
    const uint32_t* read_count;   /* n_groups : hit counts */
    const uint8_t* read_status;   /* n_reads : 0 = mapped as asked; 1 = the read is not longer than the number of
                                     parts of the search scheme: the reference falls back to naive backtracking
                                     (src/searchstrategy.cpp:442-459), this library reports no hits for it */
    uint32_t n_short_reads;       /* number of reads with read_status 1 */
    uint64_t counters[CLB_N_COUNTERS];
    uint64_t stats[CLB_N_STATS];
    float stage_ms[CLB_N_STAGES]; /* device time per stage (CUDA events) */
    float total_ms;               /* device time of the whole batch incl. copies */
    uint32_t gpu_launches;        /* kernels launched for this batch */
} clb_result;

/* ---- index (replaces FMIndex::FMIndex, src/fmindex/fmindex.h:403-417) ----
 * base: basename of the reference's index files (.cct .brt .rev.brt .txt.bin
 * .sa.<s> .sa.bv.<s> .pos .sna .meta); kmer: word size of the seed table
 * (-K, default 10); in_text_switch: -i (default 4, src/parameters/alignparameters.h:90);
 * device: CUDA device ordinal. */
int clb_index_load(const char* base, int sa_sparse, int kmer, int in_text_switch, int no_cigar, int device,
                   clb_index** out);
void clb_index_free(clb_index*);
uint32_t clb_index_text_length(const clb_index*);
uint32_t clb_index_num_seqs(const clb_index*);
const char* clb_index_seq_name(const clb_index*, uint32_t i);
uint32_t clb_index_seq_start(const clb_index*, uint32_t i); /* i in [0, num_seqs] */
uint64_t clb_index_device_bytes(const clb_index*);

/* ---- context: stream + scratch + strategy ---- */
int clb_ctx_create(clb_index*, clb_ctx** out);
void clb_ctx_free(clb_ctx*);
const char* clb_last_error(const clb_ctx*); /* ctx may be NULL: last error of index loading */

/* Search strategy (replaces Parameters::createStrategy,
 * src/parameters/alignparameters.cpp:1312-1387):
 *   kind "named"          arg = columba|kuch1|kuch2|kianfar|pigeon|01*0|minU  (-S)
 *   kind "custom"         arg = search scheme directory, no dynamic selection (-c DIR -nD)
 *   kind "custom_dynamic" arg = search scheme directory + its mirror          (-c DIR)
 *   kind "multi"          arg = directory with <k>/scheme<x>.txt              (-d DIR)
 * data_dir: directory holding the built-in tables (columba_b200/schemes). */
int clb_ctx_set_strategy(clb_ctx*, const char* kind, const char* arg, const char* data_dir);
int clb_ctx_set_options(clb_ctx*, int metric, int partitioning);
/* Largest distance accepted for an alignment that is clipped at a sequence
 * boundary (largestStratum of IndexInterface::findSeqName,
 * src/indexinterface.cpp:799-897): 0 = the batch's max_dist (ALL mode); BEST
 * mode passes its cut-off (src/searchstrategy.cpp:537-568).  A clipped
 * alignment can therefore have a distance above max_dist. */
int clb_ctx_set_trim_max(clb_ctx*, int largest_stratum);
/* BEST mode maps and filters the two strands of a read separately
 * (SearchStrategy::processSeq -> mapRead, src/searchstrategy.cpp:777-811,
 * src/searchstrategy.h:490-517), ALL mode filters them together
 * (matchApproxAllMap, src/searchstrategy.cpp:495-535).  per_strand = 1 makes
 * the following batches group, de-duplicate and filter the occurrences per
 * (read, strand); clb_result then has 2 * n_reads groups. */
int clb_ctx_set_filter_per_strand(clb_ctx*, int per_strand);
/* Run this context's copies and kernels on the caller's CUDA stream
 * (cudaStream_t, e.g. torch.cuda.current_stream().cuda_stream) instead of the
 * context's own stream; NULL restores the own stream. */
int clb_ctx_set_stream(clb_ctx*, void* cuda_stream);
int clb_ctx_supports_distance(const clb_ctx*, int k);
int clb_ctx_num_parts(const clb_ctx*, int k); /* calculateNumParts */

/* ---- the hot path (replaces SearchStrategy::matchApproxAllMap for a whole
 * chunk, src/searchstrategy.cpp:495-535).  reads: concatenated ASCII reads,
 * offsets: n_reads+1 offsets into it.  HOST pointers; copies are part of the
 * call.  max_dist = -e; min_dist = minimal distance to report (0 in ALL mode): as in
 * SearchStrategy::mapRead (src/searchstrategy.h:490-517) the read is mapped and filtered over the whole
 * range 0..max_dist and occurrences below min_dist are removed afterwards.
 * max_dist = 0 is the exact stratum (IndexInterface::exactMatchesOutput,
 * src/indexinterface.cpp:947-1014): every exact occurrence of either strand,
 * no redundancy filter, no search scheme needed. */
int clb_map_batch(clb_ctx*, const uint8_t* reads, const uint32_t* offsets, uint32_t n_reads, int max_dist,
                  int min_dist, clb_result* out);
/* Same with DEVICE pointers for reads/offsets; with fetch == 0 the occurrence
 * arrays stay in HBM (counters, stats and timings are still returned). */
int clb_map_batch_device(clb_ctx*, const uint8_t* d_reads, const uint32_t* d_offsets, uint32_t n_reads,
                         uint32_t max_read_len, int max_dist, int min_dist, int fetch, clb_result* out);

/* ---- unit-level entry points (parity tests of the index primitives) ----
 * occ4: occ of A,C,G,T in BWT[0,p) (bwtrepr.h:80-89); reverse selects .rev.brt
 * extend: children of n parents for all 4 characters (fmindex.cpp:137-243);
 *         mode 0 forward, 1 backward, 2 backward uni-directional;
 *         parents: n x (sa.b, sa.e, rev.b, rev.e), out: n x 4 x same
 * find_sa: SA[row] through the sparse suffix array (fmindex.cpp:53-60) */
int clb_occ4(clb_ctx*, int reverse, const uint32_t* pos, uint32_t n, uint32_t* out4);
int clb_extend(clb_ctx*, int mode, const uint32_t* parents, uint32_t n, uint32_t* out16);
int clb_find_sa(clb_ctx*, const uint32_t* rows, uint32_t n, uint32_t* out);
int clb_kmer_lookup(clb_ctx*, const uint32_t* codes, uint32_t n, uint32_t* out4);

/* clb_find_sa for callers that cannot allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  No counterpart in the reference.
 * Same arguments, results and errors as clb_find_sa (HOST pointers), but the device staging of
 * rows and results lives in a static device array of the library -- 2 x CLB_STATIC_MAX_ROWS x 4 bytes = 8 MB
 * that every GPU which loads the library holds until the process exits -- and the call makes no cudaMalloc /
 * cudaFree or other allocating call.  n > CLB_STATIC_MAX_ROWS is refused with CLB_E_OVERFLOW (split the call,
 * or use clb_find_sa).  All calls on one GPU share the array, from every context and thread: calls that
 * overlap on the GPU race; serialise them.  The first call on a GPU resolves the array's device address
 * (cudaGetSymbolAddress), so make one call before starting a capture. */
#define CLB_STATIC_MAX_ROWS (1u << 20)
int clb_find_sa_static(clb_ctx*, const uint32_t* rows, uint32_t n, uint32_t* out);

/* Measured ceiling of the path's roofline unit: n_lookups independent random 32-byte rank-record reads from
 * the loaded index (8 in flight per thread); *ms = kernel time (CUDA events), *done = reads issued. */
int clb_gather_ceiling(clb_ctx*, uint64_t n_lookups, float* ms, uint64_t* done);

/* ---- index construction for synthetic benchmarks (host, not the hot path;
 * writes the reference's file formats, src/buildindex.cpp) ---- */
int clb_build_index(const char* fasta, const char* base, int sa_sparse, char* err, size_t err_len);

int clb_abi_version(void);
size_t clb_sizeof_result(void); /* sizeof(clb_result): lets a foreign-language binding check its struct layout */
int clb_device_count(void);

#ifdef __cplusplus
}
#endif
#endif
