// ORACLE (test infrastructure, NOT product code).
// CPU restatement of the reference's search / search-scheme model and of the
// strategy parameters (search.h, searchstrategy.h).  The on-disk format is the
// reference's search_schemes/ layout.
#pragma once
#include "orc_index.h"
#include <sstream>
#include <sys/stat.h>

namespace orc {

static const int MAX_K = 13; // definitions.h:54

// search.h:55-495
struct Search {
    std::vector<len_t> L, U, order;
    len_t sIdx = 0;
    std::vector<Dir> directions;
    std::vector<bool> dswitch;
    std::vector<std::pair<len_t, len_t>> lowHigh;
    bool uniBackAll = false;
    len_t uniBackIdx = 0;

    static Search make(std::vector<len_t> order, std::vector<len_t> L, std::vector<len_t> U,
                       len_t sIdx) { // search.h:116-194
        if (order.size() != L.size() || order.size() != U.size())
            throw std::runtime_error("Could not create search, the sizes of all vectors are not equal");
        Search s;
        s.order = order;
        s.L = L;
        s.U = U;
        s.sIdx = sIdx;
        s.directions.push_back((order[1] > order[0]) ? FWD : BWD);
        for (size_t i = 1; i < order.size(); i++)
            s.directions.push_back((order[i] > order[i - 1]) ? FWD : BWD);
        s.dswitch.push_back(false);
        s.dswitch.push_back(false);
        for (size_t i = 2; i < s.directions.size(); i++)
            s.dswitch.push_back(s.directions[i] != s.directions[i - 1]);
        s.lowHigh.emplace_back(order[0], order[0]);
        for (size_t i = 1; i < order.size(); i++) {
            auto pb = s.lowHigh.back();
            if (order[i] < pb.first)
                s.lowHigh.emplace_back(order[i], pb.second);
            else
                s.lowHigh.emplace_back(pb.first, order[i]);
        }
        s.uniBackAll = (order[0] == order.size() - 1);
        s.uniBackIdx = (len_t)order.size();
        if (order.back() != 0) {
            s.uniBackIdx = (len_t)order.size();
        } else if (s.uniBackAll) {
            s.uniBackIdx = 0;
        } else {
            for (size_t idx = 0; idx < order.size(); idx++)
                if (order[idx] == order.size() - 1) {
                    s.uniBackIdx = (len_t)idx + 1;
                    break;
                }
        }
        return s;
    }
    len_t numParts() const { return (len_t)order.size(); }
    len_t lowestBefore(len_t idx) const { return lowHigh[idx - 1].first; }
    len_t maxED() const { return U.back(); }
    len_t minED() const { return L.back(); }
    bool isEdge(len_t idx) const { return order[idx] == 0 || order[idx] == numParts() - 1; }
    bool isEnd(len_t idx) const { return idx == order.size() - 1; }
    bool isUniBack(len_t idx) const { return uniBackAll || idx >= uniBackIdx; }
    void setMinED(len_t d) { L.back() = std::max(L.back(), d); }
    Search mirror() const {
        std::vector<len_t> mo = order;
        for (size_t i = 0; i < order.size(); i++)
            mo[i] = (len_t)order.size() - 1 - order[i];
        return make(mo, L, U, sIdx);
    }
    bool operator<(const Search& r) const { // search.h:422-441
        for (len_t i = 0; i < numParts(); i++)
            if (U[i] != r.U[i])
                return U[i] > r.U[i];
        for (len_t i = 0; i < numParts(); i++)
            if (L[i] != r.L[i])
                return L[i] < r.L[i];
        return sIdx < r.sIdx;
    }
    bool connectivity() const {
        len_t hi = order[0], lo = order[0];
        for (size_t i = 1; i < order.size(); i++) {
            if (order[i] == hi + 1)
                hi++;
            else if (order[i] == lo - 1)
                lo--;
            else
                return false;
        }
        return true;
    }
    bool validBounds() const {
        if (L[0] > U[0])
            return false;
        for (size_t i = 1; i < order.size(); i++)
            if (L[i] > U[i] || L[i] < L[i - 1] || U[i] < U[i - 1])
                return false;
        return true;
    }
    bool zeroBased() const { return *std::min_element(order.begin(), order.end()) == 0; }
};

// search.h:509-758
struct Scheme {
    std::vector<Search> searches;
    unsigned k = 0;
    uint16_t criticalPart = 0;
    Scheme() {}
    Scheme(const std::vector<Search>& s, unsigned k_) : searches(s), k(k_) {
        sanity();
        auto it = std::min_element(searches.begin(), searches.end());
        criticalPart = (uint16_t)it->order[0];
    }
    void sanity() const { // search.h:552-588
        len_t P = searches.front().numParts();
        for (auto& s : searches) {
            if (s.numParts() != P)
                throw std::runtime_error("Not all searches for distance " + std::to_string(k) +
                                         " have the same number of parts");
            if (!s.zeroBased())
                throw std::runtime_error("Not all searches are zero based for distance " +
                                         std::to_string(k) + "!");
            if (!s.connectivity())
                throw std::runtime_error("Connectivity property not satisfied for all searches with distance " +
                                         std::to_string(k) + "!");
            if (!s.validBounds())
                throw std::runtime_error("Decreasing lower or upper bounds for a search for K  = " +
                                         std::to_string(k));
        }
    }
    len_t numParts() const { return searches.front().numParts(); }
    Scheme mirror() const {
        std::vector<Search> r;
        for (auto& s : searches)
            r.push_back(s.mirror());
        return Scheme(r, k);
    }
    static void getVector(const std::string& vs, std::vector<len_t>& v) { // search.h:635-650
        if (vs.size() < 2)
            throw std::runtime_error(vs + " is not a valid vector for a search");
        std::stringstream ss(vs.substr(1, vs.size() - 2));
        std::string tok;
        while (std::getline(ss, tok, ','))
            v.push_back((len_t)std::stoull(tok));
    }
    static Search parseLine(const std::string& line, len_t idx) { // search.h:599-623
        std::stringstream ss(line);
        std::vector<std::string> t;
        std::string tok;
        while (ss >> tok)
            t.push_back(tok);
        if (t.size() != 3)
            throw std::runtime_error("A search should have 3 vectors: order, lower bound and upper bound!");
        std::vector<len_t> o, l, u;
        getVector(t[0], o);
        getVector(t[1], l);
        getVector(t[2], u);
        return Search::make(o, l, u, idx);
    }
    static bool readFile(const std::string& fn, unsigned k, Scheme& out) { // search.h:684-711
        std::ifstream f(fn);
        if (!f)
            return false;
        std::vector<Search> rs;
        std::string line;
        len_t sIdx = 0;
        while (std::getline(f, line)) {
            if (line.empty())
                continue;
            rs.push_back(parseLine(line, sIdx++));
        }
        if (rs.empty())
            throw std::runtime_error("Empty scheme in: " + fn);
        out = Scheme(rs, k);
        return true;
    }
};

enum PartitionStrategy { UNIFORM, STATIC, DYNAMIC };
enum Metric { HAMMING, EDIT };

// Everything a SearchStrategy subclass contributes, as data
// (searchstrategy.h:2130-3775).
struct KCfg {
    std::vector<Scheme> schemes; // >1 => dynamic selection (searchstrategy.h:2505-2537)
    bool customSeeds = false, customWeights = false, customBegins = false;
    std::vector<double> seeds, begins;
    std::vector<uint64_t> weights;
    bool supported() const { return !schemes.empty(); }
};
struct Strategy {
    std::string name;
    len_t useKmerCutOff = 20; // searchstrategy.h:222
    KCfg perK[MAX_K + 1];     // index = k
    PartitionStrategy part = DYNAMIC;
    Metric metric = EDIT;

    bool supports(int k) const { return k >= 1 && k <= MAX_K && perK[k].supported(); }
    len_t numParts(int k) const { return perK[k].schemes.front().numParts(); }
    // searchstrategy.h:1825-1834 / custom
    std::vector<double> seedingPositions(int P, int k) const {
        if (perK[k].customSeeds)
            return perK[k].seeds;
        std::vector<double> s;
        double u = 1.0 / (P - 1);
        for (int i = 1; i < P - 1; i++)
            s.push_back(i * u);
        return s;
    }
    std::vector<uint64_t> weights(int P, int k) const { // searchstrategy.h:283-289
        if (perK[k].customWeights)
            return perK[k].weights;
        std::vector<uint64_t> w(P, 1);
        w.front() = 2;
        w.back() = 2;
        return w;
    }
    std::vector<double> begins(int P, int k) const { // searchstrategy.h:245-253
        if (perK[k].customBegins)
            return perK[k].begins;
        std::vector<double> b;
        double u = 1.0 / P;
        for (int i = 1; i < P; i++)
            b.push_back(i * u);
        return b;
    }
    // searchstrategy.h:2505-2537
    const std::vector<Search>& createSearches(int k, const std::vector<RangePair>& ranges) const {
        const auto& sc = perK[k].schemes;
        if (sc.size() == 1)
            return sc[0].searches;
        unsigned total = 0;
        for (auto& r : ranges)
            total += r.width();
        if (total <= numParts(k))
            return sc[0].searches;
        int minIndex = 0;
        unsigned minValue = ranges[sc[0].criticalPart].width();
        for (unsigned i = 1; i < sc.size(); ++i) {
            unsigned cp = sc[i].criticalPart;
            if (ranges[cp].width() < minValue) {
                minValue = ranges[cp].width();
                minIndex = i;
            }
        }
        return sc[minIndex].searches;
    }

    static bool isDir(const std::string& p) {
        struct stat st;
        return stat(p.c_str(), &st) == 0 && (st.st_mode & S_IFDIR);
    }
    static std::string slash(std::string p) {
        if (!p.empty() && p.back() != '/')
            p += '/';
        return p;
    }
    // CustomSearchStrategy (searchstrategy.cpp:1990-2119); cutOff 50 for -c,
    // 100 for the hard-coded kuch1/kuch2/kianfar/01*0 classes, 20 otherwise.
    static Strategy custom(const std::string& folder_, len_t cutOff) {
        std::string folder = slash(folder_);
        Strategy st;
        st.useKmerCutOff = cutOff;
        {
            std::ifstream f(folder + "name.txt");
            if (!f)
                throw std::runtime_error("Problem reading: " + folder +
                                         "name.txt\nDid you provide a directory to a search scheme without a name file?");
            std::getline(f, st.name);
        }
        for (int k = 1; k <= MAX_K; k++) {
            std::string base = folder + std::to_string(k) + "/";
            Scheme sc;
            if (!Scheme::readFile(base + "searches.txt", k, sc))
                continue;
            KCfg& c = st.perK[k];
            c.schemes.push_back(sc);
            len_t P = sc.numParts();
            std::ifstream fs(base + "static_partitioning.txt");
            if (fs) {
                std::string line, tok;
                std::getline(fs, line);
                std::stringstream ss(line);
                while (ss >> tok)
                    c.begins.push_back(std::stod(tok));
                if (c.begins.size() != P - 1)
                    throw std::runtime_error("Not enough static positions provided in " + base +
                                             "static_partitioning.txt");
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
                    throw std::runtime_error("Not enough seeding positions provided in " + base +
                                             "dynamic_partitioning.txt");
                for (size_t i = 0; i < c.seeds.size(); i++) {
                    if (c.seeds[i] <= 0 || c.seeds[i] >= 1)
                        throw std::runtime_error("One of the provided seeding positions is not between 0 and 1 (exclusive)!");
                    if (i + 1 < c.seeds.size() && c.seeds[i] - c.seeds[i + 1] >= 0)
                        throw std::runtime_error("Provided seeding positions are not strictly increasing");
                }
                std::getline(fd, line);
                std::stringstream sw(line);
                while (sw >> tok)
                    c.weights.push_back(std::stoi(tok));
                if (c.weights.size() != P)
                    throw std::runtime_error("Not enough weights provided for max score " + std::to_string(k));
                c.customSeeds = c.customWeights = true;
            }
        }
        return st;
    }
    // MultipleSchemesStrategy(SearchStrategy*) (searchstrategy.h:2691-2703):
    // scheme + mirrored scheme per k, base-class (default) partition params.
    static Strategy dynamicFrom(const Strategy& s) {
        Strategy d;
        d.name = s.name;
        d.useKmerCutOff = s.useKmerCutOff;
        for (int k = 1; k <= MAX_K; k++) {
            if (!s.perK[k].supported())
                continue;
            d.perK[k].schemes.push_back(s.perK[k].schemes[0]);
            d.perK[k].schemes.push_back(s.perK[k].schemes[0].mirror());
        }
        return d;
    }
    // MultipleSchemesStrategy(index, path, ...) (searchstrategy.h:2624-2682)
    static Strategy multi(const std::string& folder_) {
        std::string folder = slash(folder_);
        Strategy st;
        {
            std::ifstream f(folder + "name.txt");
            if (!f)
                throw std::runtime_error("Problem reading: " + folder + "name.txt");
            std::getline(f, st.name);
        }
        for (int k = 1; k <= MAX_K; k++) {
            std::string dir = folder + "/" + std::to_string(k);
            for (int x = 1;; x++) {
                Scheme sc;
                if (!Scheme::readFile(dir + "/scheme" + std::to_string(x) + ".txt", k, sc))
                    break;
                st.perK[k].schemes.push_back(sc);
            }
            auto& v = st.perK[k].schemes;
            for (auto& s : v)
                if (s.numParts() != v.front().numParts())
                    throw std::runtime_error("Not all schemes have same number of parts in: " + dir);
        }
        return st;
    }
    // DynamicColumbaStrategy (searchstrategy.h:3666-3736): minU (k<=7) and the
    // greedy tables (8..13), each with its mirror, plus the "mid" schemes.
    static Strategy columba(const std::string& dataDir_) {
        std::string dataDir = slash(dataDir_);
        Strategy lo = custom(dataDir + "minU/", 20), hi = custom(dataDir + "columba_high/", 20),
                 mid = custom(dataDir + "columba_mid/", 20);
        Strategy base;
        for (int k = 1; k <= MAX_K; k++) {
            const Strategy& src = k <= 7 ? lo : hi;
            if (src.perK[k].supported())
                base.perK[k].schemes.push_back(src.perK[k].schemes[0]);
        }
        Strategy d = dynamicFrom(base);
        d.name = "Columba";
        d.useKmerCutOff = 20;
        for (int k : {2, 4, 6})
            d.perK[k].schemes.push_back(mid.perK[k].schemes[0]);
        d.perK[6].schemes.push_back(mid.perK[6].schemes[0].mirror());
        return d;
    }
    // named strategies of alignparameters.cpp:1312-1387 (those with data tables)
    static Strategy named(const std::string& nm, const std::string& dataDir_) {
        std::string dataDir = slash(dataDir_);
        if (nm == "columba")
            return columba(dataDir);
        if (nm == "kuch1" || nm == "kuch2" || nm == "kianfar")
            return custom(dataDir + nm + "/", 100);
        if (nm == "01*0")
            return custom(dataDir + "01star0/", 100);
        if (nm == "pigeon" || nm == "minU")
            return custom(dataDir + nm + "/", 20);
        throw std::runtime_error(nm + " is not an option as search scheme");
    }
};

} // namespace orc
