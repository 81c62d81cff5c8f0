// gw_fm.cuh — device primitives of the FM-index layer: rank over the interleaved occurrence blocks, LF step,
// SA locate, contig translation. Replaces (reference, align/src/main/java/org/utgenome/weaver/align/):
//   OccurrenceCountTable.getOccACGTN :80-89, getOcc :99-107; ACGTSequence.fastCount(ACGTN) :456-553
//   FMIndexOnOccTable.backwardSearch :47-51, suffixLink :64-70; SparseSuffixArray.get :141-161
//   SequenceBoundary.translate :104-121
#pragma once
#include <cuda_runtime.h>

#include "gw_common.h"

typedef unsigned long long u64;

// Java shift semantics (count & 63)
__device__ __forceinline__ u64 jshl(u64 x, int s) { return x << (s & 63); }
__device__ __forceinline__ u64 jushr(u64 x, int s) { return x >> (s & 63); }

struct GwBlockRegs {
    uint4 c;   // cnt A,C,G,T
    u64 lo0, lo1, hi0, hi1, nf0, nf1;
};

// one 64-byte line as four 128-bit read-only loads
__device__ __forceinline__ GwBlockRegs gw_load_block(const GwOccBlock* blk) {
    const uint4* p = reinterpret_cast<const uint4*>(blk);
    uint4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2), d = __ldg(p + 3);
    GwBlockRegs r;
    r.c = a;
    r.lo0 = ((u64)b.y << 32) | b.x; r.lo1 = ((u64)b.w << 32) | b.z;
    r.hi0 = ((u64)c.y << 32) | c.x; r.hi1 = ((u64)c.w << 32) | c.z;
    r.nf0 = ((u64)d.y << 32) | d.x; r.nf1 = ((u64)d.w << 32) | d.z;
    return r;
}

// prefix masks for the first r (0..127) symbols of a block
__device__ __forceinline__ void gw_prefix_masks(uint32_t r, u64& m0, u64& m1) {
    m0 = r >= 64 ? ~0ULL : ((1ULL << r) - 1ULL);
    m1 = r > 64 ? ((1ULL << (r - 64)) - 1ULL) : 0ULL;
}

// occurrences of A,C,G,T (N-flagged symbols excluded) in BWT[0, i)  == getOccACGTN without the N column
__device__ __forceinline__ void gw_rank4(const GwBlockRegs& b, uint32_t r, uint32_t out[4]) {
    u64 m0, m1;
    gw_prefix_masks(r, m0, m1);
    u64 v0 = ~b.nf0 & m0, v1 = ~b.nf1 & m1;
    uint32_t t = __popcll(b.hi0 & b.lo0 & v0) + __popcll(b.hi1 & b.lo1 & v1);
    uint32_t g = __popcll(b.hi0 & ~b.lo0 & v0) + __popcll(b.hi1 & ~b.lo1 & v1);
    uint32_t c = __popcll(~b.hi0 & b.lo0 & v0) + __popcll(~b.hi1 & b.lo1 & v1);
    uint32_t a = __popcll(v0) + __popcll(v1) - t - g - c;
    out[0] = b.c.x + a; out[1] = b.c.y + c; out[2] = b.c.z + g; out[3] = b.c.w + t;
}

// occurrences of one symbol ch (0..3) in BWT[0, i)
__device__ __forceinline__ uint32_t gw_rank1(const GwBlockRegs& b, uint32_t r, int ch) {
    u64 m0, m1;
    gw_prefix_masks(r, m0, m1);
    u64 h0 = (ch & 2) ? b.hi0 : ~b.hi0, h1 = (ch & 2) ? b.hi1 : ~b.hi1;
    u64 l0 = (ch & 1) ? b.lo0 : ~b.lo0, l1 = (ch & 1) ? b.lo1 : ~b.lo1;
    uint32_t n = __popcll(h0 & l0 & ~b.nf0 & m0) + __popcll(h1 & l1 & ~b.nf1 & m1);
    uint32_t base = ch == 0 ? b.c.x : ch == 1 ? b.c.y : ch == 2 ? b.c.z : b.c.w;
    return base + n;
}

// number of N in BWT[0, i): block start minus the four counters, plus N flags in the prefix
__device__ __forceinline__ uint32_t gw_rankN(const GwBlockRegs& b, uint32_t blockStart, uint32_t r) {
    u64 m0, m1;
    gw_prefix_masks(r, m0, m1);
    return blockStart - (b.c.x + b.c.y + b.c.z + b.c.w) + __popcll(b.nf0 & m0) + __popcll(b.nf1 & m1);
}

__device__ __forceinline__ uint32_t gw_clampN(const GwIndexView& ix, uint32_t i) { return i > ix.N ? ix.N : i; }

// getOcc(ch, i) for ch in A,C,G,T
__device__ __forceinline__ uint32_t gw_occ1(const GwIndexView& ix, int fm, int ch, uint32_t i) {
    i = gw_clampN(ix, i);
    GwBlockRegs b = gw_load_block(ix.occ[fm] + (i >> 7));
    return gw_rank1(b, i & 127u, ch);
}

// FMIndexOnOccTable.suffixLink: one LF step
__device__ __forceinline__ uint32_t gw_suffix_link(const GwIndexView& ix, int fm, uint32_t index) {
    if (index >= ix.N) return 0;
    GwBlockRegs b = gw_load_block(ix.occ[fm] + (index >> 7));
    uint32_t r = index & 127u;
    u64 lo = r < 64 ? b.lo0 : b.lo1, hi = r < 64 ? b.hi0 : b.hi1, nf = r < 64 ? b.nf0 : b.nf1;
    uint32_t s = r & 63u;
    if ((nf >> s) & 1ULL) return ix.C[fm][4] + gw_rankN(b, index & ~127u, r);
    int ch = (int)(((hi >> s) & 1ULL) << 1 | ((lo >> s) & 1ULL));
    return ix.C[fm][ch] + gw_rank1(b, r, ch);
}

// SparseSuffixArray.get: text position of SA row `index`; *steps receives the number of LF steps (occ lookups)
__device__ __forceinline__ uint32_t gw_sa_get(const GwIndexView& ix, int fm, uint32_t index, uint32_t* steps) {
    uint32_t cursor = index, j = 0;
    while ((cursor % GW_SSA_L) != 0) {
        cursor = gw_suffix_link(ix, fm, cursor);
        ++j;
    }
    *steps = j;
    uint32_t s = __ldg(ix.ssa[fm] + cursor / GW_SSA_L);
    if (j == 0) return s;
    return (uint32_t)(((u64)s + j + ix.N) % ix.N);
}

// SequenceBoundary.translate(textIndex): greatest contig offset strictly below textIndex (TreeMap.headMap)
__device__ __forceinline__ bool gw_translate(const GwIndexView& ix, long long textIndex, int* chr, int* pos) {
    int lo = 0, hi = ix.n_ctg_off;  // first offset >= textIndex
    while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if ((long long)ix.ctg_off[mid] < textIndex) lo = mid + 1; else hi = mid;
    }
    if (lo == 0) return false;
    *chr = ix.ctg_id[lo - 1];
    *pos = (int)(textIndex - (long long)ix.ctg_off[lo - 1]);
    return true;
}

// symbol (0..4) of the forward reference text at position p
__device__ __forceinline__ int gw_ref_symbol(const GwIndexView& ix, uint32_t p) {
    uint32_t w = p >> 6, s = p & 63u;
    if ((__ldg(ix.ref_nf + w) >> s) & 1ULL) return 4;
    return (int)((((__ldg(ix.ref_hi + w) >> s) & 1ULL) << 1) | ((__ldg(ix.ref_lo + w) >> s) & 1ULL));
}
