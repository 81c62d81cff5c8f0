// gw_fm.cuh — device primitives of the FM-index layer: rank over the interleaved occurrence blocks, LF step,
// SA locate, contig translation. Replaces (reference, align/src/main/java/org/utgenome/weaver/align/):
//   OccurrenceCountTable.getOccACGTN :80-89, getOcc :99-107; ACGTSequence.fastCount(ACGTN) :456-553
//   FMIndexOnOccTable.backwardSearch :47-51, suffixLink :64-70; SparseSuffixArray.get :141-161
//   SequenceBoundary.translate :104-121
// These are the single-thread forms (one thread fetches the whole 64 B line with four 128-bit loads); the quick-scan
// and locate kernels in gw_kernels.cu also have 4-lane cooperative forms built on gw_quarter_*.
#pragma once
#include "gw_common.h"

struct GwBlk { GwU4 q[4]; };

GW_HD GwBlk gw_load_block(const GwOccBlock* blk) {
    GwBlk b;
    b.q[0] = gw_ld128(&blk->q[0]); b.q[1] = gw_ld128(&blk->q[1]);
    b.q[2] = gw_ld128(&blk->q[2]); b.q[3] = gw_ld128(&blk->q[3]);
    return b;
}

// mask of the symbols of quarter q that lie in the block prefix [0, r), r in 0..127
GW_HD uint32_t gw_quarter_mask(uint32_t r, int q) {
    int n = (int)r - 32 * q;
    return n <= 0 ? 0u : (n >= 32 ? 0xFFFFFFFFu : ((1u << n) - 1u));
}
// occurrences of symbol ch (0..3) among the prefix symbols of one quarter (N-flagged symbols excluded)
GW_HD uint32_t gw_quarter_count1(const GwU4& qd, uint32_t mask, int ch) {
    uint32_t h = (ch & 2) ? qd.z : ~qd.z, l = (ch & 1) ? qd.y : ~qd.y;
    return (uint32_t)gw_popc32(h & l & ~qd.w & mask);
}

// the single-thread forms below work on the block as two 64-symbol halves (quarters 0|1 and 2|3 as one 64-bit word per bit plane):
// two prefix masks and 64-bit popcounts instead of four masks and 32-bit popcounts
struct GwHalves { u64 lo[2], hi[2], nf[2]; };
GW_HD GwHalves gw_halves(const GwBlk& b) {
    GwHalves h;
    h.lo[0] = (u64)b.q[0].y | ((u64)b.q[1].y << 32); h.lo[1] = (u64)b.q[2].y | ((u64)b.q[3].y << 32);
    h.hi[0] = (u64)b.q[0].z | ((u64)b.q[1].z << 32); h.hi[1] = (u64)b.q[2].z | ((u64)b.q[3].z << 32);
    h.nf[0] = (u64)b.q[0].w | ((u64)b.q[1].w << 32); h.nf[1] = (u64)b.q[2].w | ((u64)b.q[3].w << 32);
    return h;
}
// masks of the symbols of the two halves that lie in the block prefix [0, r), r in 0..127
GW_HD u64 gw_half_mask0(uint32_t r) { return r >= 64u ? ~0ULL : ((1ULL << r) - 1ULL); }
GW_HD u64 gw_half_mask1(uint32_t r) { return r <= 64u ? 0ULL : ((1ULL << (r - 64u)) - 1ULL); }

// occurrences of A,C,G,T (N-flagged symbols excluded) in BWT[0, i)  == getOccACGTN without the N column
GW_HD void gw_rank4(const GwBlk& b, uint32_t r, uint32_t out[4]) {
    const GwHalves h = gw_halves(b);
    const u64 v0 = ~h.nf[0] & gw_half_mask0(r), v1 = ~h.nf[1] & gw_half_mask1(r);
    const uint32_t t = (uint32_t)(gw_popc64(h.hi[0] & h.lo[0] & v0) + gw_popc64(h.hi[1] & h.lo[1] & v1));
    const uint32_t g = (uint32_t)(gw_popc64(h.hi[0] & ~h.lo[0] & v0) + gw_popc64(h.hi[1] & ~h.lo[1] & v1));
    const uint32_t c = (uint32_t)(gw_popc64(~h.hi[0] & h.lo[0] & v0) + gw_popc64(~h.hi[1] & h.lo[1] & v1));
    const uint32_t a = (uint32_t)(gw_popc64(v0) + gw_popc64(v1)) - t - g - c;
    out[0] = b.q[0].x + a; out[1] = b.q[1].x + c; out[2] = b.q[2].x + g; out[3] = b.q[3].x + t;
}

// occurrences of one symbol ch (0..3) in BWT[0, i)
GW_HD uint32_t gw_rank1(const GwBlk& b, uint32_t r, int ch) {
    const GwHalves h = gw_halves(b);
    const u64 h0 = (ch & 2) ? h.hi[0] : ~h.hi[0], h1 = (ch & 2) ? h.hi[1] : ~h.hi[1];
    const u64 l0 = (ch & 1) ? h.lo[0] : ~h.lo[0], l1 = (ch & 1) ? h.lo[1] : ~h.lo[1];
    const uint32_t n = (uint32_t)(gw_popc64(h0 & l0 & ~h.nf[0] & gw_half_mask0(r)) + gw_popc64(h1 & l1 & ~h.nf[1] & gw_half_mask1(r)));
    const uint32_t base = ch == 0 ? b.q[0].x : ch == 1 ? b.q[1].x : ch == 2 ? b.q[2].x : b.q[3].x;
    return base + n;
}

// number of N in BWT[0, i): block start minus the four counters, plus N flags in the prefix
GW_HD uint32_t gw_rankN(const GwBlk& b, uint32_t blockStart, uint32_t r) {
    const GwHalves h = gw_halves(b);
    const uint32_t n = blockStart - (b.q[0].x + b.q[1].x + b.q[2].x + b.q[3].x);
    return n + (uint32_t)(gw_popc64(h.nf[0] & gw_half_mask0(r)) + gw_popc64(h.nf[1] & gw_half_mask1(r)));
}

GW_HD uint32_t gw_clampN(const GwIndexView& ix, uint32_t i) { return i > ix.N ? ix.N : i; }

// getOcc(ch, i) for ch in A,C,G,T (OccurrenceCountTable.java:99-107)
GW_HD uint32_t gw_occ1(const GwIndexView& ix, int fm, int ch, uint32_t i) {
    i = gw_clampN(ix, i);
    GwBlk b = gw_load_block(ix.occ[fm] + (i >> 7));
    return gw_rank1(b, i & 127u, ch);
}

// FMIndexOnOccTable.suffixLink: one LF step
GW_HD uint32_t gw_suffix_link(const GwIndexView& ix, int fm, uint32_t index) {
    if (index >= ix.N) return 0;
    GwBlk b = gw_load_block(ix.occ[fm] + (index >> 7));
    const uint32_t r = index & 127u;
    const GwU4 qd = b.q[r >> 5];
    const uint32_t s = r & 31u;
    if ((qd.w >> s) & 1u) return ix.C[fm][4] + gw_rankN(b, index & ~127u, r);
    const int ch = (int)((((qd.z >> s) & 1u) << 1) | ((qd.y >> s) & 1u));
    return ix.C[fm][ch] + gw_rank1(b, r, ch);
}

// SparseSuffixArray.get: text position of SA row `index`; *steps receives the number of LF steps (occ lookups) it took.
// With the locate table at its densest (loc_step == 1) no step is taken at all.
GW_HD uint32_t gw_sa_get(const GwIndexView& ix, int fm, uint32_t index, uint32_t* steps) {
    uint32_t cursor = index, j = 0;
    const uint32_t mask = ix.loc_step - 1u;
    while ((cursor & mask) != 0) {
        cursor = gw_suffix_link(ix, fm, cursor);
        ++j;
    }
    uint32_t s = gw_ld32(ix.loc[fm] + cursor / ix.loc_step);
    if (s == 0xFFFFFFFFu) {   // unknown to the table: the reference's own walk to a sampled row
        while ((cursor % GW_SSA_L) != 0) {
            cursor = gw_suffix_link(ix, fm, cursor);
            ++j;
        }
        s = gw_ld32(ix.ssa[fm] + cursor / GW_SSA_L);
    }
    *steps = j;
    if (j == 0) return s;
    return (uint32_t)(((u64)s + j + ix.N) % ix.N);
}

// SequenceBoundary.translate(textIndex): greatest contig offset strictly below textIndex (TreeMap.headMap)
GW_HD bool gw_translate(const GwIndexView& ix, long long textIndex, int* chr, int* pos) {
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
GW_HD int gw_ref_symbol(const GwIndexView& ix, uint32_t p) {
    const uint32_t w = p >> 6, s = p & 63u;
    if ((gw_ld64(ix.ref_nf + w) >> s) & 1ULL) return 4;
    return (int)((((gw_ld64(ix.ref_hi + w) >> s) & 1ULL) << 1) | ((gw_ld64(ix.ref_lo + w) >> s) & 1ULL));
}
