// gw_verify.cuh — bit-vector DP verifier with CIGAR traceback (one thread per candidate fragment).
// Replaces BitParallelSmithWaterman.AlignBlocksDetailed (reference:
// align/src/main/java/org/utgenome/weaver/align/BitParallelSmithWaterman.java): align :387-459 (block-wise Myers /
// Hyyro recurrence with adaptive number of active 64-row blocks), alignBlock :461-491, traceback :501-644
// (ops string, leading/trailing indel -> soft clip rewriting, run-length CIGAR).
//
// The reference keeps the VP/VN words of every (column, block) for the traceback: bMax x (N'+1) x 16 bytes per fragment. Here
// nothing of that is stored. The forward pass keeps one checkpoint of its state every GW_DP_SEG columns (a few hundred bytes in the
// lane's local memory). The traceback needs a VP/VN bit only where the reference and query symbols differ (a handful of cells per
// alignment: a matching cell is a diagonal step whatever the bits say); for such a cell the column is recomputed from the nearest
// checkpoint — at most GW_DP_SEG + 1 column steps — and kept until the traceback leaves it.
#pragma once
#include "gw_fm.cuh"

#define GW_RUN_CAP 64   // run-length encoded traceback ops kept per fragment (>= 2 * GW_MAX_K + 3)
#define GW_DP_SEG 32    // columns between two checkpoints of the forward pass
#define GW_DP_MAX_CK ((GW_MAX_WINDOW + GW_DP_SEG - 1) / GW_DP_SEG + 1)

struct GwVerifyOut {
    int valid;   // 0: reference returns null
    int pos, diff, cig_n;
    bool cig_overflow;
};

// state of the recurrence between two columns
struct GwDpCk {
    u64 vp[GW_MAX_NB], vn[GW_MAX_NB];
    int32_t D[GW_MAX_NB];
    int32_t bceil;
};

struct GwDp {
    int32_t m, ncols, bmax, has_best, best_tail, best_diff;
    GwDpCk ck[GW_DP_MAX_CK];   // ck[s]: the state before column s * GW_DP_SEG
};

GW_HD int gw_score_boundary(int m, int band_k, int r) { return gw_max(m - ((r + 1) * 64) + band_k, 0); }

// One column of the recurrence (:413-443) on the state `s` (NB: compile-time bound of the number of 64-row blocks, so that the state
// stays in registers and a 100 bp read does not pay for the blocks only 250 bp reads have). Returns what the reference's arrays hold
// for this column afterwards: blocks [0, *computed) carry the new VP/VN, blocks [*computed, bMax) keep Java's zero initial value
// (a block deactivated in this column was computed and keeps its values). *activated: the block that was switched on in this column
// (its entry of the PREVIOUS column was overwritten with VP = ~0, VN = 0 before its first step, :425-431), or -1.
template <int NB>
GW_HD void gw_dp_step(int (&D)[NB], u64 (&vp)[NB], u64 (&vn)[NB], int& bCeil, int m, int bMax, int band_k, int ch, const u64 (*peq)[GW_MAX_NB],
                      int* computed, int* activated) {
    const int w = 64;
    int carry = 0;
    int comp = bCeil;
#pragma unroll
    for (int r = 0; r < NB; ++r) {
        if (r < bCeil) {
            u64 x = ch < 4 ? peq[ch][r] : 0ULL;
            if (carry < 0) x |= 1ULL;
            const u64 d0 = (((x & vp[r]) + vp[r]) ^ vp[r]) | x | vn[r];
            const u64 hp = vn[r] | ~(d0 | vp[r]);
            const u64 hn = d0 & vp[r];
            const int hout = (int)((hp >> 63) & 1ULL) - (int)((hn >> 63) & 1ULL);
            u64 hp2 = hp << 1, hn2 = hn << 1;
            if (carry < 0) hn2 |= 1ULL;
            if (carry > 0) hp2 |= 1ULL;
            vp[r] = hn2 | ~(d0 | hp2);
            vn[r] = d0 & hp2;
            D[r] += hout;
            carry = hout;
        }
    }
    // activation of the next block (:425-431) — written without dynamic indexing so that D/vp/vn stay in registers
    int act = -1;
#pragma unroll
    for (int r = 1; r < NB; ++r) {
        if (act < 0 && r == bCeil && r < bMax) {
            u64 x = ch < 4 ? peq[ch][r] : 0ULL;
            if (D[r - 1] - carry <= gw_score_boundary(m, band_k, r - 1) && (((x & 1ULL) != 0ULL) | (carry < 0))) {
                // vp[r][j] = ~0, vn[r][j] = 0 ; then one block step
                const u64 vpr = ~0ULL, vnr = 0ULL;
                if (carry < 0) x |= 1ULL;
                const u64 d0 = (((x & vpr) + vpr) ^ vpr) | x | vnr;
                const u64 hp = vnr | ~(d0 | vpr);
                const u64 hn = d0 & vpr;
                const int hout = (int)((hp >> 63) & 1ULL) - (int)((hn >> 63) & 1ULL);
                u64 hp2 = hp << 1, hn2 = hn << 1;
                if (carry < 0) hn2 |= 1ULL;
                if (carry > 0) hp2 |= 1ULL;
                vp[r] = hn2 | ~(d0 | hp2);
                vn[r] = d0 & hp2;
                D[r] = D[r - 1] - carry + hout;
                bCeil = r + 1;
                comp = bCeil;
                act = r;
            }
        }
    }
    if (act < 0) {   // while (bCeil > 1 && D[bCeil - 1] > scoreBoundary[bCeil - 1] + w) --bCeil;  (:439-441)
#pragma unroll
        for (int r = NB - 1; r >= 1; --r)
            if (bCeil == r + 1 && D[r] > gw_score_boundary(m, band_k, r) + w) bCeil = r;
    }
    *computed = comp;
    *activated = act;
}

template <int NB>
GW_HD void gw_dp_load(const GwDpCk& c, int (&D)[NB], u64 (&vp)[NB], u64 (&vn)[NB], int& bCeil) {
#pragma unroll
    for (int r = 0; r < NB; ++r) { D[r] = c.D[r]; vp[r] = c.vp[r]; vn[r] = c.vn[r]; }
    bCeil = c.bceil;
}
template <int NB>
GW_HD void gw_dp_save(GwDpCk& c, const int (&D)[NB], const u64 (&vp)[NB], const u64 (&vn)[NB], int bCeil) {
#pragma unroll
    for (int r = 0; r < NB; ++r) { c.D[r] = D[r]; c.vp[r] = vp[r]; c.vn[r] = vn[r]; }
    c.bceil = bCeil;
}

// forward pass over all columns: best end column (:444-455) and one checkpoint per segment; nothing else is stored
// peq: [4][GW_MAX_NB] pattern masks of the query (bits >= m clear); refsym(j): reference symbol (0..4) of window column j
template <int NB, typename RefSym>
GW_HD void gw_dp_forward_nb(GwDp& d, int band_k, const u64 (*peq)[GW_MAX_NB], RefSym refsym) {
    const int m = d.m, bMax = d.bmax;
    int D[NB], bCeil;
    u64 vp[NB], vn[NB];
#pragma unroll
    for (int r = 0; r < NB; ++r) { D[r] = 0; vp[r] = ~0ULL; vn[r] = 0ULL; }
    D[0] = m;
    bCeil = gw_min(gw_max(1, (band_k + 63) / 64), bMax);   // band_k <= 64 in practice; keeps indices in range
    bool hasBest = false;
    int bestTail = -1, bestDiff = 0;
    for (int j = 0; j < d.ncols; ++j) {
        if ((j % GW_DP_SEG) == 0) gw_dp_save<NB>(d.ck[j / GW_DP_SEG], D, vp, vn, bCeil);
        int computed, activated;
        gw_dp_step<NB>(D, vp, vn, bCeil, m, bMax, band_k, refsym(j), peq, &computed, &activated);
        if (bCeil == bMax) {
            int dl = D[0];
#pragma unroll
            for (int r = 1; r < NB; ++r) if (r == bMax - 1) dl = D[r];
            if (!hasBest) { hasBest = true; bestTail = j; bestDiff = dl; }
            else if (bestDiff > dl) { bestTail = j; bestDiff = dl; }
        }
    }
    d.has_best = hasBest ? 1 : 0; d.best_tail = bestTail; d.best_diff = bestDiff;
}

// What the reference's VP / VN tables hold for window column `col` (entry col + 1), as blocks: the columns from the checkpoint before
// `col` up to col + 1 are stepped again (the column after `col` may switch a block on, which overwrites that block's entry of `col`)
template <int NB, typename RefSym>
GW_HD void gw_dp_column_nb(const GwDp& d, int col, int band_k, const u64 (*peq)[GW_MAX_NB], RefSym refsym, u64* sp_out, u64* sn_out) {
    const int m = d.m, bMax = d.bmax;
    int D[NB], bCeil;
    u64 vp[NB], vn[NB];
    const int s = col / GW_DP_SEG;
    gw_dp_load<NB>(d.ck[s], D, vp, vn, bCeil);
    u64 sp[NB], sn[NB];
#pragma unroll
    for (int r = 0; r < NB; ++r) { sp[r] = 0ULL; sn[r] = 0ULL; }
    const int j1 = gw_min(col + 2, d.ncols);
    for (int j = s * GW_DP_SEG; j < j1; ++j) {
        int computed, activated;
        gw_dp_step<NB>(D, vp, vn, bCeil, m, bMax, band_k, refsym(j), peq, &computed, &activated);
        if (j == col + 1) {
#pragma unroll
            for (int r = 0; r < NB; ++r) if (r == activated) { sp[r] = ~0ULL; sn[r] = 0ULL; }
        } else if (j == col) {
#pragma unroll
            for (int r = 0; r < NB; ++r) { const bool live = r < computed; sp[r] = live ? vp[r] : 0ULL; sn[r] = live ? vn[r] : 0ULL; }
        }
    }
#pragma unroll
    for (int r = 0; r < GW_MAX_NB; ++r) { sp_out[r] = 0ULL; sn_out[r] = 0ULL; }
#pragma unroll
    for (int r = 0; r < NB; ++r) { sp_out[r] = sp[r]; sn_out[r] = sn[r]; }
}

// nb: the instantiation to take (a caller whose lanes run together passes the largest block count among them, so that all of them take
// the same one: a lane with fewer blocks only skips the guarded parts)
template <typename RefSym>
GW_HD void gw_dp_forward(GwDp& d, int band_k, const u64 (*peq)[GW_MAX_NB], RefSym refsym, int nb) {
    if (nb <= 1) gw_dp_forward_nb<1>(d, band_k, peq, refsym);
    else if (nb <= 2) gw_dp_forward_nb<2>(d, band_k, peq, refsym);
    else gw_dp_forward_nb<GW_MAX_NB>(d, band_k, peq, refsym);
}
template <typename RefSym>
GW_HD void gw_dp_column(const GwDp& d, int col, int band_k, const u64 (*peq)[GW_MAX_NB], RefSym refsym, u64* sp, u64* sn, int nb) {
    if (nb <= 1) gw_dp_column_nb<1>(d, col, band_k, peq, refsym, sp, sn);
    else if (nb <= 2) gw_dp_column_nb<2>(d, col, band_k, peq, refsym, sp, sn);
    else gw_dp_column_nb<GW_MAX_NB>(d, col, band_k, peq, refsym, sp, sn);
}

// traceback (:529-595) + CIGAR (:599-644). refsym / qsym: symbols of the window columns / query rows; nb: instantiation of the
// recomputation (see gw_dp_forward)
template <typename RefSym, typename QSym>
GW_HD void gw_dp_traceback(const GwDp& d, int band_k, const u64 (*peq)[GW_MAX_NB], RefSym refsym, QSym qsym, int nb, uint32_t* cig_out, GwVerifyOut* out) {
    const int m = d.m;
    const bool hasBest = d.has_best != 0;
    const int bestTail = d.best_tail;
    out->cig_overflow = false;
    if (!hasBest) { out->valid = 0; out->pos = 0; out->diff = 0; out->cig_n = 0; return; }

    // ---- traceback: ops (0 M, 1 I, 2 D, 3 S) are produced right-to-left and run-length encoded on the fly
    uint32_t runs[GW_RUN_CAP];   // len << 2 | op, right-most run first
    int nruns = 0;
    uint32_t curOp = 4u, curLen = 0u;
    auto push = [&](uint32_t op) {
        if (op == curOp) { ++curLen; return; }
        if (curLen) { if (nruns < GW_RUN_CAP) runs[nruns] = (curLen << 2) | curOp; ++nruns; }
        curOp = op; curLen = 1u;
    };
    int leftMostPos = 0, diff = 0;
    int col = bestTail, row = m - 1;
    int have_col = -1;
    u64 sp[GW_MAX_NB], sn[GW_MAX_NB];   // the column the traceback is in, recomputed when a cell needs its bits
    for (;;) {
        int path = 0;
        if (col >= 0 && row >= 0) {
            if (refsym(col) == qsym(row)) path = 1;
            else {
                if (col != have_col) { gw_dp_column(d, col, band_k, peq, refsym, sp, sn, nb); have_col = col; }
                const int b = row >> 6, off = row & 63;
                u64 xp = 0ULL, xn = 0ULL;
#pragma unroll
                for (int r = 0; r < GW_MAX_NB; ++r) if (r == b) { xp = sp[r]; xn = sn[r]; }
                if ((xp >> off) & 1ULL) path = 3;
                else if (!((xn >> off) & 1ULL)) path = 2;
                else path = 4;
                diff++;
            }
        }
        if (path == 1 || path == 2) { push(0); leftMostPos = col; col--; row--; }
        else if (path == 3) { push(1); leftMostPos = col + 1; row--; }
        else if (path == 4) { push(2); col--; }
        else {
            // off the left / top edge: the remaining rows are soft-clipped
            if (row >= 0) {
                const uint32_t cnt = (uint32_t)(row + 1);
                if (curOp == 3u) curLen += cnt;
                else { push(3); curLen = cnt; }
            }
            break;
        }
    }
    if (curLen) { if (nruns < GW_RUN_CAP) runs[nruns] = (curLen << 2) | curOp; ++nruns; }
    bool overflow = nruns > GW_RUN_CAP;
    const int nr = overflow ? GW_RUN_CAP : nruns;
    // cigarStr = reverse(ops): its first run is runs[nr-1]. Leading / trailing S,I,D become S (diff-- per I/D base) (:599-636)
    int left = 0, right = 0, L = 0, R = 0;
    for (int i = nr - 1; i >= 0; --i) {
        const uint32_t t = runs[i] & 3u, len = runs[i] >> 2;
        if (t == 0u) break;
        left += (int)len; ++L;
        if (t != 3u) diff -= (int)len;
    }
    for (int i = 0; i < nr - L; ++i) {
        const uint32_t t = runs[i] & 3u, len = runs[i] >> 2;
        if (t == 0u) break;
        right += (int)len; ++R;
        if (t != 3u) diff -= (int)len;
    }
    int n = 0;
    if (left > 0) { cig_out[n] = ((uint32_t)left << 4) | GW_OP_S; ++n; }
    for (int i = nr - 1 - L; i >= R; --i) {
        const uint32_t t = runs[i] & 3u, len = runs[i] >> 2;
        if (n < GW_CIG_CAP) cig_out[n] = (len << 4) | (t == 0u ? GW_OP_M : t == 1u ? GW_OP_I : t == 2u ? GW_OP_D : GW_OP_S);
        else overflow = true;
        ++n;
    }
    if (right > 0) { if (n < GW_CIG_CAP) cig_out[n] = ((uint32_t)right << 4) | GW_OP_S; else overflow = true; ++n; }
    out->valid = 1; out->pos = leftMostPos; out->diff = diff; out->cig_n = n < GW_CIG_CAP ? n : GW_CIG_CAP; out->cig_overflow = overflow;
}

// alignBlockDetailed in one go (component entry point, host simulator)
template <typename RefSym, typename QSym>
GW_HD void gw_verify(int m, int Ncols, int band_k, const u64 (*peq)[GW_MAX_NB], RefSym refsym, QSym qsym, GwDp* d, uint32_t* cig_out, GwVerifyOut* out) {
    d->m = m; d->ncols = Ncols; d->bmax = gw_max(1, (m + 63) / 64);
    gw_dp_forward(*d, band_k, peq, refsym, d->bmax);
    gw_dp_traceback(*d, band_k, peq, refsym, qsym, d->bmax, cig_out, out);
}
