// Build shim (test infrastructure, NOT product code).
// The reference's index builder calls libsais (a fetched dependency that is
// absent offline). The suffix array of a text is unique, so any correct
// builder yields byte-identical index files. This is a small self-written
// prefix-doubling suffix sorter with the libsais call signature.
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <vector>

template <typename I>
static inline int clb_shim_suffix_sort(const uint8_t* T, I* SA, I n) {
    if (n <= 0)
        return 0;
    const int K = 8; // initial prefix length (bytes)
    std::vector<std::pair<uint64_t, I>> keyed((size_t)n);
    for (I i = 0; i < n; i++) {
        uint64_t k = 0;
        for (int j = 0; j < K; j++)
            k = (k << 8) | ((i + j < n) ? T[i + j] : 0);
        keyed[i] = std::make_pair(k, i);
    }
    std::sort(keyed.begin(), keyed.end());
    std::vector<I> isa((size_t)n), nxt((size_t)n);
    for (I i = 0, g = 0; i < n; i++) {
        if (i > 0 && keyed[i].first != keyed[i - 1].first)
            g = i;
        SA[i] = keyed[i].second;
        isa[keyed[i].second] = g;
    }
    std::vector<std::pair<uint64_t, I>>().swap(keyed);
    for (int64_t h = K;; h *= 2) {
        bool any = false;
        std::copy(isa.begin(), isa.end(), nxt.begin());
        for (I i = 0; i < n;) {
            I j = i + 1;
            while (j < n && isa[SA[j]] == i)
                j++;
            if (j - i > 1) {
                any = true;
                auto key = [&](I x) -> int64_t {
                    return ((int64_t)x + h < (int64_t)n) ? (int64_t)isa[x + h]
                                                         : -1;
                };
                std::sort(SA + i, SA + j,
                          [&](I a, I b) { return key(a) < key(b); });
                I g = i;
                for (I t = i; t < j; t++) {
                    if (t > i && key(SA[t]) != key(SA[t - 1]))
                        g = t;
                    nxt[SA[t]] = g;
                }
            }
            i = j;
        }
        if (!any)
            break;
        isa.swap(nxt);
    }
    return 0;
}

static inline int32_t libsais(const uint8_t* T, int32_t* SA, int32_t n,
                              int32_t, int32_t*) {
    return clb_shim_suffix_sort<int32_t>(T, SA, n);
}
