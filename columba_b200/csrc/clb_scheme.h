// columba_b200 -- flattened search schemes as the kernels read them.
//
// Replaces the in-memory model of src/search.h:55-495 (Search) and :509-758
// (SearchScheme) plus the per-strategy parameters of src/searchstrategy.h
// (seeding positions :1825-1834, weights :283-289, static begins :245-253,
// dynamic scheme selection :2505-2537).  The host (clb_strategy.cpp) parses the
// search_schemes/ directory format and the built-in tables and fills one
// DevStrategy per maximal distance k; kernels only read it.
#pragma once
#include "clb_common.h"

namespace clb {

enum { MAXP = 16, MAX_SEARCHES = 16, MAX_SCHEMES = 6, CLB_MAX_K = 13 };

struct DevSearch {
    u8 P;
    u8 order[MAXP];
    u8 L[MAXP];
    u8 U[MAXP];
    u8 dir[MAXP];  // DIR_FWD / DIR_BWD per phase
    u8 dsw[MAXP];  // direction switch before phase i
    u8 low[MAXP];  // lowest part index covered after phase i
    u8 high[MAXP];
    u8 uniBackAll;
    u8 uniBackIdx;
    CLB_HD u32 maxED() const { return U[P - 1]; }
    CLB_HD u32 minED() const { return L[P - 1]; }
    CLB_HD bool isEdge(u32 idx) const { return order[idx] == 0 || order[idx] == P - 1; }
    CLB_HD bool isEnd(u32 idx) const { return idx == (u32)P - 1; }
    CLB_HD bool isUniBack(u32 idx) const { return uniBackAll || idx >= uniBackIdx; }
    CLB_HD u32 lowestBefore(u32 idx) const { return low[idx - 1]; }
};

struct DevScheme {
    u8 nSearches;
    u8 criticalPart;
    DevSearch s[MAX_SEARCHES];
};

struct DevStrategy {
    u8 supported;
    u8 nSchemes; // > 1: dynamic selection
    u8 P;
    u8 pad;
    u32 useKmerCutOff;
    double seeds[MAXP];   // P-2 seeding positions (dynamic partitioning)
    u64 weights[MAXP];    // P weights
    double begins[MAXP];  // P-1 static begin fractions
    DevScheme sc[MAX_SCHEMES];
};

} // namespace clb
