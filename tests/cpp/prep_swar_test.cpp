// Unit test of the preparation stage (clb_partition.h: asciiCodes4, codes64,
// prepTask) against a character-by-character restatement: every byte value,
// either case, characters other than ACGT, lengths around the 64-character
// chunk boundaries.  Compiled with g++ (the kernel functors are host-callable).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../columba_b200/csrc/clb_partition.h"

using namespace clb;

static u32 rnd() {
    static u64 s = 0x9E3779B97F4A7C15ull;
    s ^= s << 13;
    s ^= s >> 7;
    s ^= s << 17;
    return (u32)(s >> 16);
}
static u32 scalarCode(u8 c) { // nucleotide.h: ACGT in either case, anything else is "other"
    switch (c) {
    case 'A': case 'a': return 0;
    case 'C': case 'c': return 1;
    case 'G': case 'g': return 2;
    case 'T': case 't': return 3;
    default: return 4;
    }
}
#define CHECK(cond)                                                                                           \
    do {                                                                                                      \
        if (!(cond)) {                                                                                        \
            std::printf("FAILED %s line %d\n", #cond, __LINE__);                                              \
            return 1;                                                                                         \
        }                                                                                                     \
    } while (0)

int main() {
    // every byte value in every byte position
    for (u32 pos = 0; pos < 4; pos++)
        for (u32 v = 0; v < 256; v++)
            for (u32 rep = 0; rep < 8; rep++) {
                u32 w = rnd();
                w = (w & ~(0xFFu << (8 * pos))) | (v << (8 * pos));
                u32 c2, val;
                asciiCodes4(w, c2, val);
                for (u32 q = 0; q < 4; q++) {
                    const u8 ch = (u8)(w >> (8 * q));
                    const u32 want = scalarCode(ch);
                    CHECK(((val >> (8 * q + 7)) & 1u) == (want < 4 ? 1u : 0u));
                    CHECK(((val >> (8 * q)) & 0x7Fu) == 0);
                    if (want < 4)
                        CHECK(((c2 >> (8 * q)) & 0xFFu) == want);
                    CHECK(asciiCodeFast(ch) == want);
                }
            }
    // whole tasks: reads of all lengths 1..200 out of a mixed alphabet
    const char alphabet[] = "ACGTACGTACGTacgtNnRYKMxU-*.\x01\x7f\xff";
    const u32 nReads = 400, maxLen = 200;
    std::vector<u8> ascii;
    std::vector<u32> off(1, 0);
    for (u32 r = 0; r < nReads; r++) {
        const u32 len = 1 + r % maxLen;
        for (u32 i = 0; i < len; i++)
            ascii.push_back(r % 7 == 0 ? (u8)rnd() : (u8)alphabet[rnd() % (sizeof(alphabet) - 1)]);
        off.push_back((u32)ascii.size());
    }
    ascii.resize(ascii.size() + 8);
    Batch B;
    std::memset(&B, 0, sizeof(B));
    B.seqStride = (maxLen + 63) & ~63u;
    B.nw = (maxLen + 63) / 64 + 1;
    std::vector<u8> codes((size_t)2 * nReads * B.seqStride, 0xEE);
    std::vector<u64> planes((size_t)2 * nReads * 2 * 4 * B.nw, 0xEEEEEEEEEEEEEEEEull);
    B.readsAscii = ascii.data();
    B.readOff = off.data();
    B.codes = codes.data();
    B.planes = planes.data();
    const u32 nc = B.nw - 1;
    for (u32 task = 0; task < nReads * nc; task++)
        prepTask(B, task, 0);
    for (u32 r = 0; r < nReads; r++) {
        const u32 len = off[r + 1] - off[r];
        const u8* rd = ascii.data() + off[r];
        for (u32 strand = 0; strand < 2; strand++) {
            const u32 t = 2 * r + strand;
            std::vector<u32> want(B.seqStride, 4);
            for (u32 i = 0; i < len; i++) {
                const u32 c = scalarCode(strand ? rd[len - 1 - i] : rd[i]);
                want[i] = (strand && c < 4) ? 3 - c : c;
            }
            for (u32 i = 0; i < B.seqStride; i++)
                CHECK(codes[(size_t)t * B.seqStride + i] == want[i]);
            for (u32 d = 0; d < 2; d++)
                for (u32 c = 0; c < 4; c++)
                    for (u32 w = 0; w < B.nw; w++) {
                        u64 bits = 0;
                        for (u32 b = 0; b < 64; b++) {
                            const u32 i = 64 * w + b;
                            if (i >= len)
                                continue;
                            const u32 ch = (d == DIR_FWD) ? want[i] : want[len - 1 - i];
                            if (ch == c)
                                bits |= 1ull << b;
                        }
                        CHECK(planes[(((size_t)t * 2 + d) * 4 + c) * B.nw + w] == bits);
                    }
        }
    }
    std::printf("OK\n");
    return 0;
}
