// columba_b200 -- host side of the search schemes: parses the search_schemes/
// directory format (unchanged from the reference, see
// further_info/advanced_options/README.md:36-134), validates it like the
// reference does, and flattens it into one DevStrategy per distance k.
//
// Replaces (reference): Search::makeSearch src/search.h:116-194, SearchScheme
// sanity checks :552-588, readScheme :684-711, mirrorPiStrings :745-753,
// CustomSearchStrategy src/searchstrategy.cpp:1990-2119 + sanity :2213-2264,
// MultipleSchemesStrategy src/searchstrategy.h:2624-2703, the built-in
// strategies :2829-3736 (their tables ship as data files under
// columba_b200/schemes/ in the same directory format), and the strategy factory
// src/parameters/alignparameters.cpp:1312-1387.
#pragma once
#include "clb_scheme.h"
#include <algorithm>
#include <fstream>
#include <sstream>
#include <string.h>
#include <stdexcept>
#include <string>
#include <sys/stat.h>
#include <vector>

namespace clb {

struct HostSearch {
    std::vector<u32> order, L, U;
    u32 sIdx = 0;
};
struct HostScheme {
    std::vector<HostSearch> searches;
    u32 k = 0;
};

inline DevSearch flattenSearch(const HostSearch& h) {
    const size_t P = h.order.size();
    if (P != h.L.size() || P != h.U.size())
        throw std::runtime_error("Could not create search, the sizes of all vectors are not equal");
    if (P < 2 || P > MAXP)
        throw std::runtime_error("A search must have between 2 and " + std::to_string((int)MAXP) + " parts");
    DevSearch d;
    memset(&d, 0, sizeof(d));
    d.P = (u8)P;
    for (size_t i = 0; i < P; i++) {
        d.order[i] = (u8)h.order[i];
        d.L[i] = (u8)h.L[i];
        d.U[i] = (u8)h.U[i];
    }
    d.dir[0] = (h.order[1] > h.order[0]) ? DIR_FWD : DIR_BWD;
    for (size_t i = 1; i < P; i++)
        d.dir[i] = (h.order[i] > h.order[i - 1]) ? DIR_FWD : DIR_BWD;
    d.dsw[0] = d.dsw[1] = 0;
    for (size_t i = 2; i < P; i++)
        d.dsw[i] = d.dir[i] != d.dir[i - 1];
    d.low[0] = d.high[0] = (u8)h.order[0];
    for (size_t i = 1; i < P; i++) {
        d.low[i] = std::min<u8>(d.low[i - 1], (u8)h.order[i]);
        d.high[i] = std::max<u8>(d.high[i - 1], (u8)h.order[i]);
    }
    d.uniBackAll = (h.order[0] == P - 1);
    d.uniBackIdx = (u8)P;
    if (h.order.back() == 0 && !d.uniBackAll) {
        for (size_t idx = 0; idx < P; idx++)
            if (h.order[idx] == P - 1) {
                d.uniBackIdx = (u8)(idx + 1);
                break;
            }
    } else if (h.order.back() == 0 && d.uniBackAll) {
        d.uniBackIdx = 0;
    }
    return d;
}

// search.h:422-441 (used to find the critical search)
inline bool searchLess(const HostSearch& a, const HostSearch& b) {
    for (size_t i = 0; i < a.U.size(); i++)
        if (a.U[i] != b.U[i])
            return a.U[i] > b.U[i];
    for (size_t i = 0; i < a.L.size(); i++)
        if (a.L[i] != b.L[i])
            return a.L[i] < b.L[i];
    return a.sIdx < b.sIdx;
}

inline void schemeSanity(const HostScheme& sc) { // search.h:552-588
    const size_t P = sc.searches.front().order.size();
    const std::string ks = std::to_string(sc.k);
    for (const auto& s : sc.searches) {
        if (s.order.size() != P)
            throw std::runtime_error("Not all searches for distance " + ks + " have the same number of parts");
        if (*std::min_element(s.order.begin(), s.order.end()) != 0)
            throw std::runtime_error("Not all searches are zero based for distance " + ks + "!");
        u32 hi = s.order[0], lo = s.order[0];
        for (size_t i = 1; i < P; i++) {
            if (s.order[i] == hi + 1)
                hi++;
            else if (s.order[i] + 1 == lo)
                lo--;
            else
                throw std::runtime_error("Connectivity property not satisfied for all searches with distance " + ks + "!");
        }
        bool ok = s.L[0] <= s.U[0];
        for (size_t i = 1; i < P && ok; i++)
            ok = !(s.L[i] > s.U[i] || s.L[i] < s.L[i - 1] || s.U[i] < s.U[i - 1]);
        if (!ok)
            throw std::runtime_error("Decreasing lower or upper bounds for a search for K  = " + ks);
    }
}

inline DevScheme flattenScheme(const HostScheme& sc) {
    if (sc.searches.empty())
        throw std::runtime_error("Empty scheme");
    if (sc.searches.size() > MAX_SEARCHES)
        throw std::runtime_error("Too many searches in one scheme (max " + std::to_string((int)MAX_SEARCHES) + ")");
    schemeSanity(sc);
    DevScheme d;
    memset(&d, 0, sizeof(d));
    d.nSearches = (u8)sc.searches.size();
    for (size_t i = 0; i < sc.searches.size(); i++)
        d.s[i] = flattenSearch(sc.searches[i]);
    auto it = std::min_element(sc.searches.begin(), sc.searches.end(), searchLess);
    d.criticalPart = (u8)it->order[0];
    return d;
}
inline HostScheme mirrorScheme(const HostScheme& sc) { // search.h:745-753
    HostScheme m = sc;
    for (auto& s : m.searches)
        for (auto& o : s.order)
            o = (u32)s.order.size() - 1 - o;
    return m;
}

inline void parseVector(const std::string& vs, std::vector<u32>& v) { // search.h:635-650
    if (vs.size() < 2)
        throw std::runtime_error(vs + " is not a valid vector for a search");
    std::stringstream ss(vs.substr(1, vs.size() - 2));
    std::string tok;
    while (std::getline(ss, tok, ','))
        v.push_back((u32)std::stoull(tok));
}
inline bool readSchemeFile(const std::string& fn, u32 k, HostScheme& out) { // search.h:599-623, :684-711
    std::ifstream f(fn);
    if (!f)
        return false;
    out = HostScheme();
    out.k = k;
    std::string line;
    u32 sIdx = 0;
    while (std::getline(f, line)) {
        if (line.empty())
            continue;
        std::stringstream ss(line);
        std::vector<std::string> t;
        std::string tok;
        while (ss >> tok)
            t.push_back(tok);
        if (t.empty())
            continue;
        if (t.size() != 3)
            throw std::runtime_error("A search should have 3 vectors: order, lower bound and upper bound!");
        HostSearch s;
        parseVector(t[0], s.order);
        parseVector(t[1], s.L);
        parseVector(t[2], s.U);
        if (s.order.size() != s.L.size() || s.order.size() != s.U.size())
            throw std::runtime_error("Could not create search, the sizes of all vectors are not equal");
        s.sIdx = sIdx++;
        out.searches.push_back(s);
    }
    if (out.searches.empty())
        throw std::runtime_error("Empty scheme in: " + fn);
    return true;
}

// everything a strategy contributes for one k, before flattening
struct HostKCfg {
    std::vector<HostScheme> schemes;
    bool customSeeds = false, customBegins = false;
    std::vector<double> seeds, begins;
    std::vector<u64> weights;
};
struct HostStrategy {
    std::string name;
    u32 useKmerCutOff = 20; // searchstrategy.h:222
    HostKCfg perK[CLB_MAX_K + 1];
    bool supports(u32 k) const { return k >= 1 && k <= CLB_MAX_K && !perK[k].schemes.empty(); }

    DevStrategy flatten(u32 k) const {
        if (!supports(k))
            throw std::runtime_error("The search scheme " + name + " does not support distance " + std::to_string(k));
        const HostKCfg& c = perK[k];
        if (c.schemes.size() > MAX_SCHEMES)
            throw std::runtime_error("Too many schemes for dynamic selection");
        DevStrategy d;
        memset(&d, 0, sizeof(d));
        d.supported = 1;
        d.nSchemes = (u8)c.schemes.size();
        d.useKmerCutOff = useKmerCutOff;
        for (size_t i = 0; i < c.schemes.size(); i++)
            d.sc[i] = flattenScheme(c.schemes[i]);
        const u32 P = d.sc[0].s[0].P;
        for (size_t i = 0; i < c.schemes.size(); i++)
            if (d.sc[i].s[0].P != P)
                throw std::runtime_error("Not all schemes have same number of parts");
        d.P = (u8)P;
        // seeding positions, searchstrategy.h:1825-1834
        if (c.customSeeds) {
            for (size_t i = 0; i < c.seeds.size() && i < MAXP; i++)
                d.seeds[i] = c.seeds[i];
            for (size_t i = 0; i < c.weights.size() && i < MAXP; i++)
                d.weights[i] = c.weights[i];
        } else {
            double u = 1.0 / (double)(int)(P - 1);
            for (u32 i = 1; i + 1 < P; i++)
                d.seeds[i - 1] = (double)(int)i * u;
            for (u32 i = 0; i < P; i++) // searchstrategy.h:283-289
                d.weights[i] = 1;
            d.weights[0] = 2;
            d.weights[P - 1] = 2;
        }
        if (c.customBegins) {
            for (size_t i = 0; i < c.begins.size() && i < MAXP; i++)
                d.begins[i] = c.begins[i];
        } else { // searchstrategy.h:245-253
            double u = 1.0 / (double)(int)P;
            for (u32 i = 1; i < P; i++)
                d.begins[i - 1] = (double)(int)i * u;
        }
        return d;
    }
};

inline std::string withSlash(std::string p) {
    if (!p.empty() && p.back() != '/')
        p += '/';
    return p;
}

// CustomSearchStrategy, searchstrategy.cpp:1990-2119
inline HostStrategy customStrategy(const std::string& folder_, u32 cutOff) {
    const std::string folder = withSlash(folder_);
    HostStrategy st;
    st.useKmerCutOff = cutOff;
    {
        std::ifstream f(folder + "name.txt");
        if (!f)
            throw std::runtime_error("Problem reading: " + folder +
                                     "name.txt\nDid you provide a directory to a search scheme without a name file?");
        std::getline(f, st.name);
    }
    for (u32 k = 1; k <= CLB_MAX_K; k++) {
        const std::string base = folder + std::to_string(k) + "/";
        HostScheme sc;
        if (!readSchemeFile(base + "searches.txt", k, sc))
            continue;
        schemeSanity(sc);
        HostKCfg& c = st.perK[k];
        c.schemes.push_back(sc);
        const size_t P = sc.searches.front().order.size();
        std::ifstream fs(base + "static_partitioning.txt");
        if (fs) {
            std::string line, tok;
            std::getline(fs, line);
            std::stringstream ss(line);
            while (ss >> tok)
                c.begins.push_back(std::stod(tok));
            if (c.begins.size() != P - 1)
                throw std::runtime_error("Not enough static positions provided in " + base + "static_partitioning.txt");
            for (size_t i = 0; i < c.begins.size(); i++) {
                if (c.begins[i] <= 0 || c.begins[i] >= 1)
                    throw std::runtime_error("One of the provided static positions is not between 0 and 1 (exclusive)");
                if (i + 1 < c.begins.size() && c.begins[i] - c.begins[i + 1] >= 0)
                    throw std::runtime_error("Provided static positions are not strictly increasing");
            }
            c.customBegins = true;
        }
        std::ifstream fd(base + "dynamic_partitioning.txt");
        if (fd) {
            std::string line, tok;
            std::getline(fd, line);
            std::stringstream ss(line);
            while (ss >> tok)
                c.seeds.push_back(std::stod(tok));
            if (c.seeds.size() != P - 2)
                throw std::runtime_error("Not enough seeding positions provided in " + base + "dynamic_partitioning.txt");
            for (size_t i = 0; i < c.seeds.size(); i++) {
                if (c.seeds[i] <= 0 || c.seeds[i] >= 1)
                    throw std::runtime_error("One of the provided seeding positions is not between 0 and 1 (exclusive)!");
                if (i + 1 < c.seeds.size() && c.seeds[i] - c.seeds[i + 1] >= 0)
                    throw std::runtime_error("Provided seeding positions are not strictly increasing");
            }
            std::getline(fd, line);
            std::stringstream sw(line);
            while (sw >> tok)
                c.weights.push_back((u64)std::stoi(tok));
            if (c.weights.size() != P)
                throw std::runtime_error("Not enough weights provided for max score " + std::to_string(k));
            c.customSeeds = true;
        }
    }
    return st;
}
// MultipleSchemesStrategy(SearchStrategy*), searchstrategy.h:2691-2703: the
// scheme and its mirror per k, default partitioning parameters
inline HostStrategy dynamicFrom(const HostStrategy& s) {
    HostStrategy d;
    d.name = s.name;
    d.useKmerCutOff = s.useKmerCutOff;
    for (u32 k = 1; k <= CLB_MAX_K; k++) {
        if (s.perK[k].schemes.empty())
            continue;
        d.perK[k].schemes.push_back(s.perK[k].schemes[0]);
        d.perK[k].schemes.push_back(mirrorScheme(s.perK[k].schemes[0]));
    }
    return d;
}
// MultipleSchemesStrategy(index, path), searchstrategy.h:2624-2682
inline HostStrategy multiStrategy(const std::string& folder_) {
    const std::string folder = withSlash(folder_);
    HostStrategy st;
    {
        std::ifstream f(folder + "name.txt");
        if (!f)
            throw std::runtime_error("Problem reading: " + folder + "name.txt");
        std::getline(f, st.name);
    }
    for (u32 k = 1; k <= CLB_MAX_K; k++) {
        const std::string dir = folder + std::to_string(k);
        for (int x = 1;; x++) {
            HostScheme sc;
            if (!readSchemeFile(dir + "/scheme" + std::to_string(x) + ".txt", k, sc))
                break;
            schemeSanity(sc);
            st.perK[k].schemes.push_back(sc);
        }
        auto& v = st.perK[k].schemes;
        for (auto& s : v)
            if (s.searches.front().order.size() != v.front().searches.front().order.size())
                throw std::runtime_error("Not all schemes have same number of parts in: " + dir);
    }
    return st;
}
// DynamicColumbaStrategy, searchstrategy.h:3666-3736
inline HostStrategy columbaStrategy(const std::string& dataDir_) {
    const std::string dataDir = withSlash(dataDir_);
    HostStrategy lo = customStrategy(dataDir + "minU/", 20), hi = customStrategy(dataDir + "columba_high/", 20),
                 mid = customStrategy(dataDir + "columba_mid/", 20);
    HostStrategy base;
    for (u32 k = 1; k <= CLB_MAX_K; k++) {
        const HostStrategy& src = k <= 7 ? lo : hi;
        if (!src.perK[k].schemes.empty())
            base.perK[k].schemes.push_back(src.perK[k].schemes[0]);
    }
    HostStrategy d = dynamicFrom(base);
    d.name = "Columba";
    d.useKmerCutOff = 20;
    for (u32 k : {2u, 4u, 6u})
        d.perK[k].schemes.push_back(mid.perK[k].schemes[0]);
    d.perK[6].schemes.push_back(mirrorScheme(mid.perK[6].schemes[0]));
    return d;
}
// alignparameters.cpp:1312-1387 (-S)
inline HostStrategy namedStrategy(const std::string& nm, const std::string& dataDir_) {
    const std::string dataDir = withSlash(dataDir_);
    if (nm == "columba")
        return columbaStrategy(dataDir);
    if (nm == "kuch1" || nm == "kuch2" || nm == "kianfar")
        return customStrategy(dataDir + nm + "/", 100);
    if (nm == "01*0")
        return customStrategy(dataDir + "01star0/", 100);
    if (nm == "pigeon" || nm == "minU")
        return customStrategy(dataDir + nm + "/", 20);
    throw std::runtime_error(nm + " is not an option as search scheme");
}

} // namespace clb
