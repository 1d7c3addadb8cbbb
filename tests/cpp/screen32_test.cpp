// Unit test of the screening pass on a 32-bit band (clb_verify.h: screenStep32)
// against the 64-bit window form (screenStep, the restatement of
// bitparallelmatrix.h:352-415): for every candidate window both forms must
// abort at the same row, keep the right-most active column on the same
// diagonal after every step, give the same final-column values wherever one of
// them is <= maxED + 1 (and values above that in the other), and take the same
// survivor decision with the same counters; the survivors then go through the
// full verification (verifyWindowFast) in both forms, which must report the same
// alignments with the same CIGAR strings.  Windows: true positions with
// shifted starts, junk positions, tandem repeats, homopolymers, the end of the
// text; reads with N; fixed and free start; maxED 0..7.
// Compiled with g++ (the kernel functors are host-callable).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../columba_b200/csrc/clb_partition.h"
#include "../../columba_b200/csrc/clb_verify.h"

using namespace clb;

static u64 rs = 0x9E3779B97F4A7C15ull;
static u32 rnd() {
    rs ^= rs << 13;
    rs ^= rs >> 7;
    rs ^= rs << 17;
    return (u32)(rs >> 16);
}
#define CHECK(cond)                                                                                           \
    do {                                                                                                      \
        if (!(cond)) {                                                                                        \
            std::printf("FAILED %s line %d (candidate %u read %u start %u maxED %u fixed %u)\n", #cond,       \
                        __LINE__, ci, (u32)c.t, start, (u32)c.maxED, (u32)c.fixedStart);                      \
            return 1;                                                                                         \
        }                                                                                                     \
    } while (0)

struct HitLog {
    std::vector<u32> v;
    void operator()(const VerifyHit& h, const u16* ops) {
        v.push_back(h.b);
        v.push_back(h.e);
        v.push_back(h.dist);
        v.push_back(h.ncig);
        for (u32 i = 0; i < h.ncig && h.ncig != 0xFFFFFFFFu; i++)
            v.push_back(ops[i]);
    }
};

// the walk of the right-most active column as the reference does it
template <typename W> static bool walkLoop(W HP, W HN, W& RAC, W stop) {
    u32 val = 1u;
    while (val > 0) {
        if (HP & RAC)
            val--;
        if (HN & RAC)
            val++;
        if (RAC == stop)
            return false;
        RAC >>= 1;
    }
    return true;
}

int main(int argc, char** argv) {
    // racWalk (closed form + fallback) against the loop: random delta words, every stop <= RAC
    for (u32 it = 0; it < 2000000; it++) {
        const u64 a = ((u64)rnd() << 40) ^ ((u64)rnd() << 20) ^ rnd(), b = ((u64)rnd() << 40) ^ ((u64)rnd() << 20) ^ rnd();
        const u64 dens = (it & 1) ? (((u64)rnd() << 40) ^ ((u64)rnd() << 20) ^ rnd()) : ~0ull;
        const u64 HP = a & ~b & dens, HN = b & ~a & dens & (((u64)rnd() << 32) | rnd());
        const u32 r = rnd() % 64, st = rnd() % (r + 1);
        u64 R1 = 1ull << r, R2 = R1;
        const bool o1 = racWalk(HP, HN, R1, 1ull << st), o2 = walkLoop<u64>(HP, HN, R2, 1ull << st);
        if (o1 != o2 || (o1 && R1 != R2)) {
            std::printf("FAILED racWalk 64 at %u\n", it);
            return 1;
        }
        const u32 r3 = r % 32, st3 = st % (r3 + 1);
        u32 Q1 = 1u << r3, Q2 = Q1;
        const bool p1 = racWalk((u32)HP, (u32)HN, Q1, 1u << st3), p2 = walkLoop<u32>((u32)HP, (u32)HN, Q2, 1u << st3);
        if (p1 != p2 || (p1 && Q1 != Q2)) {
            std::printf("FAILED racWalk 32 at %u\n", it);
            return 1;
        }
    }
    const u32 nReads = argc > 1 ? (u32)atoi(argv[1]) : 6000;
    // ---- text: random, with tandem repeats, homopolymers and copies
    const u32 N = 300000;
    std::vector<u8> txt(N);
    for (u32 i = 0; i < N;) {
        const u32 kind = rnd() % 16;
        if (kind == 0) { // tandem repeat of a short unit
            const u32 unit = 1 + rnd() % 6, len = 20 + rnd() % 300;
            u8 u[8];
            for (u32 j = 0; j < unit; j++)
                u[j] = (u8)(rnd() & 3);
            for (u32 j = 0; j < len && i < N; j++, i++)
                txt[i] = u[j % unit];
        } else if (kind == 1 && i > 2000) { // copy of an earlier stretch with a few changes
            const u32 from = rnd() % (i - 1000), len = 100 + rnd() % 600;
            for (u32 j = 0; j < len && i < N; j++, i++)
                txt[i] = (rnd() % 50 == 0) ? (u8)(rnd() & 3) : txt[from + j];
        } else {
            const u32 len = 50 + rnd() % 400;
            for (u32 j = 0; j < len && i < N; j++, i++)
                txt[i] = (u8)(rnd() & 3);
        }
    }
    std::vector<u32> packed(N / 16 + 16, 0);
    for (u32 i = 0; i < N; i++)
        packed[i >> 4] |= (u32)txt[i] << ((i & 15) * 2);
    IndexView ix;
    std::memset(&ix, 0, sizeof(ix));
    ix.text = packed.data();
    ix.textLength = N;
    // ---- reads
    const char acgt[] = "ACGT";
    const u32 maxLen = 260;
    std::vector<u8> ascii;
    std::vector<u32> off(1, 0), origin;
    for (u32 r = 0; r < nReads; r++) {
        const u32 len = 25 + rnd() % (maxLen - 25);
        const u32 pos = (r % 23 == 0) ? N - len - rnd() % 6 : rnd() % (N - len - 16);
        origin.push_back(pos);
        std::vector<u8> rd;
        const u32 kind = rnd() % 8;
        if (kind == 0) {
            for (u32 i = 0; i < len; i++)
                rd.push_back((u8)acgt[rnd() & 3]);
        } else {
            const u32 nEdits = rnd() % 10;
            for (u32 i = 0; i < len + 12 && pos + i < N; i++)
                rd.push_back((u8)acgt[txt[pos + i]]);
            for (u32 e = 0; e < nEdits && rd.size() > 8; e++) {
                const u32 at = rnd() % (u32)rd.size(), what = rnd() % 3;
                if (what == 0)
                    rd[at] = (u8)acgt[rnd() & 3];
                else if (what == 1)
                    rd.erase(rd.begin() + at);
                else
                    rd.insert(rd.begin() + at, (u8)acgt[rnd() & 3]);
            }
            if (rnd() % 20 == 0)
                rd[rnd() % (u32)rd.size()] = 'N';
            if (rd.size() > len)
                rd.resize(len);
        }
        ascii.insert(ascii.end(), rd.begin(), rd.end());
        off.push_back((u32)ascii.size());
    }
    ascii.resize(ascii.size() + 8);
    Batch B;
    std::memset(&B, 0, sizeof(B));
    B.seqStride = (maxLen + 63) & ~63u;
    B.nw = (maxLen + 63) / 64 + 1;
    std::vector<u8> codes((size_t)2 * nReads * B.seqStride, 0xEE);
    std::vector<u64> planes((size_t)2 * nReads * 2 * 4 * B.nw, 0);
    B.readsAscii = ascii.data();
    B.readOff = off.data();
    B.codes = codes.data();
    B.planes = planes.data();
    B.verifyRowsCap = maxLen + 64;
    for (u32 task = 0; task < nReads * (B.nw - 1); task++)
        prepTask(B, task, 0);
    // ---- candidates
    std::vector<Cand> cands;
    std::vector<u32> starts, mult;
    for (u32 r = 0; r < nReads; r++) {
        const u32 len = off[r + 1] - off[r];
        for (u32 q = 0; q < 40; q++) {
            Cand c;
            std::memset(&c, 0, sizeof(c));
            c.t = 2 * r; // forward strand
            c.maxED = (u8)(rnd() % 8);
            c.minED = (u8)((rnd() % 4 == 0) ? rnd() % (c.maxED + 1) : 0);
            c.fixedStart = (u8)(rnd() & 1);
            u32 s;
            const u32 kind = rnd() % 8;
            if (kind == 0)
                s = rnd() % (N - 8); // junk
            else if (kind == 1)
                s = N - len - rnd() % 24; // at the end of the text
            else {
                const int shift = c.fixedStart ? (int)(rnd() % 5) - 2 : (int)(rnd() % (4 * c.maxED + 7)) - (int)(2 * c.maxED + 3) - (int)c.maxED;
                const long v = (long)origin[r] + shift;
                s = v < 0 ? 0u : (u32)v;
            }
            if (s >= N - 1)
                s = N - 2;
            cands.push_back(c);
            starts.push_back(s);
            mult.push_back(1 + rnd() % 3);
        }
    }
    const u32 nC = (u32)cands.size();
    B.cands.data = cands.data();
    B.candStart = starts.data();
    B.candMult = mult.data();
    std::vector<u32> survA(nC), survB(nC);
    u32 nA = 0, nB = 0, errA = 0, errB = 0;
    u64 aborted = 0, survivors = 0, used32 = 0, rows = 0, hits = 0;
    for (u32 ci = 0; ci < nC; ci++) {
        const Cand c = cands[ci];
        const u32 start = starts[ci];
        ScreenState a, b;
        std::memset(&a, 0, sizeof(a));
        std::memset(&b, 0, sizeof(b));
        screenStart(ix, B, ci, a, true);
        screenStart(ix, B, ci, b, false);
        CHECK(b.b32 == 0);
        CHECK(a.done == b.done && a.run == b.run && a.considered == b.considered);
        used32 += a.b32;
        for (;;) {
            if (a.firstRow == 0 && a.i == 1 && !a.done) {
                const u32 lim = (u32)c.maxED + 2;
                CHECK(tmin<u32>(a.eds[0], lim) == tmin<u32>(b.eds[0], lim));
            }
            const bool fa = screenStep(ix, a), fb = screenStep(ix, b);
            CHECK(fa == fb);
            CHECK(a.i == b.i && a.ok == b.ok && a.done == b.done);
            if (a.done == 2 || !a.run)
                break;
            if (a.ok && a.b32 && a.i >= 2) { // same diagonal of the right-most active column
                const int da = ctz64(a.RAC) - (int)a.dB;
                const int db = ctz64(b.RAC) - (int)(((a.i - 1) % 32) + BPC<u64>::DIAG_R0);
                CHECK(da == db);
                CHECK(a.score == b.score || (a.score > (u32)c.maxED + 1 && b.score > (u32)c.maxED + 1));
            }
            if (fa)
                break;
        }
        if (a.run && a.done != 2) {
            const u32 lastRow = a.i - 1, lim = (u32)c.maxED + 2;
            for (u32 r = a.firstRow; r <= lastRow && a.ok; r++) {
                if (r == 0 && a.firstRow != 0)
                    continue;
                CHECK(tmin<u32>(a.eds[r - a.firstRow], lim) == tmin<u32>(b.eds[r - a.firstRow], lim));
            }
            rows += lastRow;
        }
        aborted += a.ok ? 0 : 1;
        LocalCtr la, lb;
        la.clear();
        lb.clear();
        B.surv.data = survA.data();
        B.surv.count = &nA;
        B.surv.cap = nC;
        B.err = &errA;
        const u32 beforeA = nA;
        screenFinish(B, a, la);
        B.surv.data = survB.data();
        B.surv.count = &nB;
        B.err = &errB;
        screenFinish(B, b, lb);
        CHECK(nA == nB);
        survivors += nA - beforeA;
        CHECK(la.c[IN_TEXT_STARTED] == lb.c[IN_TEXT_STARTED]);
        CHECK(la.c[ABORTED_IN_TEXT_VERIF] == lb.c[ABORTED_IN_TEXT_VERIF]);
        CHECK(la.s[ST_VERIFY_ROWS] == lb.s[ST_VERIFY_ROWS]);
        // the full verification (forward pass with band storage, traceback with block re-computation) in both forms
        if (nA != beforeA && a.done != 2) {
            const u32 nZeros = c.fixedStart ? 1u : 2 * (u32)c.maxED + 1;
            const u32 len = seqLen(B, c.t);
            const u32 nRows = nZeros - 1 + c.maxED + len + 1;
            const u32 end = tmin(ix.textLength - 1, start + nRows - 1);
            u64 ring[2][32], cp[2][2 * 16];
            u16 ops[2][64];
            HitLog hl[2];
            LocalCtr lv[2];
            bool cons[2], okv[2];
            u32 tot[2];
            for (u32 f = 0; f < 2; f++) {
                BandStore bs;
                bs.ring = ring[f];
                bs.ringStride = 1;
                bs.cp = cp[f];
                bs.cpStride = 1;
                bs.cpCap = 16;
                lv[f].clear();
                okv[f] = verifyWindowFast(ix, B, c.t, bs, start, end, nZeros, c.maxED, c.minED, ops[f], 2 * (u32)c.maxED + 2,
                                          lv[f], cons[f], hl[f], 1u, 0, 0xFFFFu, tot[f], 1, f == 0);
            }
            CHECK(okv[0] == okv[1] && cons[0] == cons[1] && tot[0] == tot[1]);
            CHECK(hl[0].v == hl[1].v);
            CHECK(lv[0].c[CIGARS_IN_TEXT_VERIFICATION] == lv[1].c[CIGARS_IN_TEXT_VERIFICATION]);
            CHECK(lv[0].s[ST_VERIFY_ROWS] == lv[1].s[ST_VERIFY_ROWS]);
            hits += hl[0].v.size() ? 1 : 0;
        }
    }
    for (u32 i = 0; i < nA; i++)
        if (survA[i] != survB[i]) {
            std::printf("FAILED survivor lists differ at %u\n", i);
            return 1;
        }
    if (used32 * 2 < nC || survivors == 0 || aborted == 0 || hits * 4 < survivors) {
        std::printf("FAILED the test does not exercise what it should: %llu of %u on the 32-bit band, %llu survivors, %llu aborted\n",
                    (unsigned long long)used32, nC, (unsigned long long)survivors, (unsigned long long)aborted);
        return 1;
    }
    std::fprintf(stderr, "%u candidates, %llu on the 32-bit band, %llu aborted, %llu survivors (%llu verified with hits), %llu rows\n", nC,
                 (unsigned long long)used32, (unsigned long long)aborted, (unsigned long long)survivors,
                 (unsigned long long)hits, (unsigned long long)rows);
    std::printf("OK\n");
    return 0;
}
