// gw_staircase.cpp — host construction of the suffix ("staircase") filter tables used by the search kernel.
// Reference: align/src/main/java/org/utgenome/weaver/align/strategy/StaircaseFilter.java:47-73 (chunk layout with
// Java `byte` arithmetic) over BitVector (align/BitVector.java:180-192 _lshift, :217-224 _not), and
// AlignmentScoreConfig.getMaximumEditDistance (align/AlignmentScoreConfig.java:40-47).
// The masks are a pure function of (m, kk); the Java quirks that change them are kept on purpose:
//   * chunkStart / chunkLen are `byte`: values above 127 wrap negative;
//   * `~0L << (size % 64)` with size % 64 == 0 shifts by 0, so _not() zeroes the last block of a 64-multiple length;
//   * a negative shift length makes _lshift address blocks out of range -> the JVM throws (reported, not emulated); in the
//     `extended` domain such blocks read as 0 instead, as in the oracle's extended mode.
#include <math.h>

#include "gw_host.h"

namespace {
inline uint64_t jshl(uint64_t x, long long s) { return x << (unsigned)(s & 63); }
inline uint64_t jushr(uint64_t x, long long s) { return x >> (unsigned)(s & 63); }
inline int jbyte(int v) { return (int)(int8_t)(uint8_t)(v & 0xFF); }
}  // namespace

int gw_max_edit_distance(float k, int read_length) {
    if (k > 0.0f && k < 1.0f) {
        volatile float prod = (float)read_length * k;   // Java: int * float is float arithmetic, then Math.floor(double)
        return (int)floor((double)prod);
    }
    return (int)k;
}

bool gw_staircase_masks(int m, int kk, std::vector<uint64_t>& words, bool extended) {
    const int nb = (m + 63) / 64;
    words.assign((size_t)(kk + 1) * (size_t)nb, 0);
    if (nb == 0) return extended;   // BitVector._not on an empty block array throws
    const int lastChunk = (m - kk >= 6) ? m * 2 / (kk + 2) : m - kk;
    std::vector<int> chunkStart((size_t)kk + 2, 0);
    const int rest = jbyte(m - lastChunk);
    if (kk == 0) chunkStart[0] = rest;
    else for (int i = 0; i <= kk; ++i) chunkStart[(size_t)i] = jbyte(rest * i / kk);
    chunkStart[(size_t)kk + 1] = jbyte(m);
    for (int row = 0; row <= kk; ++row) {
        std::vector<uint64_t> blk((size_t)nb, 0);
        // new BitVector(m)._not()
        for (int b = 0; b + 1 < nb; ++b) blk[(size_t)b] = ~0ULL;
        blk[(size_t)nb - 1] = ~0ULL & ~jshl(~0ULL, m % 64);
        // ._lshift(chunkStart[row])
        const int len = chunkStart[(size_t)row];
        const int blockOffset = len / 64;          // truncates toward zero like Java
        const long long offset = len % 64;         // sign of the dividend like Java
        const uint64_t lowMask = ~jushr(~0ULL, offset);
        for (int i = nb - 1; i >= 0; --i) {
            const int b = i - blockOffset;
            uint64_t high = 0, low = 0;
            if (b >= 0) {
                if (b >= nb && !extended) return false;         // ArrayIndexOutOfBoundsException (extended: the block reads as 0)
                high = b >= nb ? 0 : jshl(blk[(size_t)b], offset);
            }
            if (b - 1 >= 0) {
                if (b - 1 >= nb && !extended) return false;
                low = b - 1 >= nb ? 0 : jushr(blk[(size_t)b - 1] & lowMask, 64 - offset);
            }
            blk[(size_t)i] = high | low;
        }
        for (int b = 0; b < nb; ++b) words[(size_t)row * nb + b] = blk[(size_t)b];
    }
    return true;
}
