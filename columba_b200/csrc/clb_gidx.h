// columba_b200 -- the .gidx sidecar: the HBM layout of clb_index.h as one file
// (SURVEY.md section 8 f3).  Written by clb_gpu_build (csrc/builder) or by
// clb_index_write_gidx from the reference's own index files; read by
// clb_index_load, which then skips the re-packing of the reference's files
// (src/fmindex/bwtrepr.h:113-135, src/bitvec.h:378-394,
// src/fmindex/suffixArray.h:229-243).  All sections start at multiples of 4096
// bytes; everything is little endian.
#pragma once
#include <stdint.h>

#define CLB_GIDX_MAGIC "CLBGIDX\0"
#define CLB_GIDX_VERSION 1u

namespace clb {

struct alignas(32) GidxBwtRec { // == BwtRec
    uint32_t cnt[4];
    uint64_t lo, hi;
};
struct alignas(16) GidxSsaRec { // == SsaRec
    uint64_t bits;
    uint32_t rank;
    uint32_t pad;
};

struct GidxHeader {
    char magic[8];
    uint32_t version;
    uint32_t textLength; // including '$'
    uint32_t saSparse;
    uint32_t C[5];
    uint32_t dollarFwd, dollarRev;
    uint32_t numSeq;
    uint32_t reserved0;
    uint64_t nRec;       // BwtRec per direction
    uint64_t nSsaRec;
    uint64_t nSsa;
    uint64_t nTextWords; // u32 words of 2-bit text (padded)
    uint64_t namesBytes; // zero-terminated sequence names
    uint64_t offSeqStart, offNames, offFwd, offRev, offSsaRec, offSsa, offText;
    uint64_t fileBytes;
};

} // namespace clb
