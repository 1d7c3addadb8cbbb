// Build shim (test infrastructure): 64-bit entry of the suffix sorter.
#pragma once
#include "libsais.h"
static inline int64_t libsais64(const uint8_t* T, int64_t* SA, int64_t n,
                                int64_t, int64_t*) {
    return clb_shim_suffix_sort<int64_t>(T, SA, n);
}
