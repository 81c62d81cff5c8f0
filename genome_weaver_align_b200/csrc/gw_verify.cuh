// gw_verify.cuh — bit-vector DP verifier with CIGAR traceback (one thread per candidate fragment).
// Replaces BitParallelSmithWaterman.AlignBlocksDetailed (reference:
// align/src/main/java/org/utgenome/weaver/align/BitParallelSmithWaterman.java): align :387-459 (block-wise Myers /
// Hyyro recurrence with adaptive number of active 64-row blocks), alignBlock :461-491, traceback :501-644
// (ops string, leading/trailing indel -> soft clip rewriting, run-length CIGAR).
// The traceback needs the VP / VN bit of (row, column) only near the diagonal the alignment runs on: every row it drifts from the
// diagonal it started on costs one difference. So of the reference's history (bMax x (N'+1) words of VP and of VN) only a band of 32
// rows around the window's main diagonal is kept per column — 2 x 32 bits, one 8-byte word, laid out [column][request] so that the 32
// lanes of a warp write 256 contiguous bytes per column: a quarter of the bytes for a 100 bp read, an eighth for 250 bp. A cell
// outside the band (an alignment that wanders more than 16 rows off the diagonal: more differences than any accepted alignment
// has) is answered exactly by replaying the recurrence up to its column (gw_dp_replay, out of line).
#pragma once
#include "gw_fm.cuh"

#define GW_RUN_CAP 64   // run-length encoded traceback ops kept per fragment (>= 2 * GW_MAX_K + 3)

// The history is written once and read at most once, much later in cache terms: its stores and loads carry the streaming hint.
GW_HD void gw_hist_put(u64* p, uint32_t vpband, uint32_t vnband) {
    const u64 v = (u64)vpband | ((u64)vnband << 32);
#if defined(__CUDA_ARCH__)
    __stcs(p, v);
#else
    *p = v;
#endif
}
GW_HD u64 gw_hist_get(const u64* p) {
#if defined(__CUDA_ARCH__)
    return __ldcs(p);
#else
    return *p;
#endif
}
// 32 rows [lo, lo + 32) of a column given as NB blocks (rows outside [0, 64 * NB) read as 0)
template <int NB>
GW_HD uint32_t gw_dp_band(const u64 (&w)[NB], int lo) {
    if (lo <= -32 || lo >= 64 * NB) return 0u;
    if (lo < 0) return (uint32_t)(w[0] << (-lo));
    const int b = lo >> 6, off = lo & 63;
    u64 x0 = 0ULL, x1 = 0ULL;
#pragma unroll
    for (int r = 0; r < NB; ++r) { if (r == b) x0 = w[r]; if (r == b + 1) x1 = w[r]; }
    return (uint32_t)((x0 >> off) | ((x1 << 1) << (63 - off)));
}
// first row of the band kept for window column `col`: 16 rows above the window's main diagonal (a fragment of m rows in a window of
// ncols columns sits (ncols - m) / 2 columns in)
GW_HD int gw_dp_band_lo(int m, int ncols, int col) { return col - (ncols - m) / 2 - 16; }

struct GwVerifyOut {
    int valid;   // 0: reference returns null
    int pos, diff, cig_n;
    bool cig_overflow;
};


// Resumable form: the DP over the window columns can be advanced in slices (gw_dp_columns) so that the fused alignment kernel
// can interleave it with the search steps of other lanes; gw_verify runs the three stages back to back.
struct GwDp {
    int32_t m, ncols, bmax, j, bceil, has_best, best_tail, best_diff;
    int32_t D[GW_MAX_NB];
    u64 vp[GW_MAX_NB], vn[GW_MAX_NB];   // column j values of the active blocks
};

GW_HD int gw_score_boundary(int m, int band_k, int r) { return gw_max(m - ((r + 1) * 64) + band_k, 0); }

// hist: scratch base for this request; entry e (what the reference's tables hold at index e: the state after window column e - 1) at
// hist[e * stride]
GW_HD void gw_dp_begin(GwDp& d, int m, int Ncols, int band_k, u64* hist, size_t stride) {
    d.m = m; d.ncols = Ncols; d.bmax = gw_max(1, (m + 63) / 64); d.j = 0;
#pragma unroll
    for (int r = 0; r < GW_MAX_NB; ++r) { d.D[r] = 0; d.vp[r] = ~0ULL; d.vn[r] = 0ULL; }
    d.D[0] = m;
    (void)hist; (void)stride;   // entry 0 (VP = ~0, VN = 0) is never read: the traceback starts at column >= 0, i.e. entry >= 1
    d.bceil = gw_max(1, (band_k + 63) / 64);
    if (d.bceil > d.bmax) d.bceil = d.bmax;  // band_k <= 64 in practice; keeps indices in range
    d.has_best = 0; d.best_tail = -1; d.best_diff = 0;
}

// peq: [4][GW_MAX_NB] pattern masks of the query (bits >= m clear); refsym(j): reference symbol (0..4) of window column j
// NB: compile-time bound of the number of 64-row blocks (ceil(m / 64) <= NB): the loops over blocks are unrolled to NB, so a 100 bp
// read does not pay for the two blocks only 250 bp reads have
template <int NB, typename RefSym>
GW_HD void gw_dp_columns_nb(GwDp& d, int jEnd, int band_k, const u64 (*peq)[GW_MAX_NB], RefSym refsym, u64* hist, size_t stride) {
    const int w = 64;
    const int m = d.m, bMax = d.bmax;
    int bCeil = d.bceil;
    int D[NB];
    u64 vp[NB], vn[NB];
#pragma unroll
    for (int r = 0; r < NB; ++r) { D[r] = d.D[r]; vp[r] = d.vp[r]; vn[r] = d.vn[r]; }
    bool hasBest = d.has_best != 0;
    int bestTail = d.best_tail, bestDiff = d.best_diff;
    if (jEnd > d.ncols) jEnd = d.ncols;
    u64 pp[NB], pn[NB];   // what the tables hold for the previous column (needed again when this column switches a block on)
#pragma unroll
    for (int r = 0; r < NB; ++r) { pp[r] = 0ULL; pn[r] = 0ULL; }
    for (int j = d.j; j < jEnd; ++j) {
        const int ch = refsym(j);
        int carry = 0;
        int computed = bCeil;  // blocks whose column j+1 is written in this step
#pragma unroll
        for (int r = 0; r < NB; ++r) {
            if (r < bCeil) {
                u64 x = ch < 4 ? peq[ch][r] : 0ULL;
                if (carry < 0) x |= 1ULL;
                u64 d0 = (((x & vp[r]) + vp[r]) ^ vp[r]) | x | vn[r];
                u64 hp = vn[r] | ~(d0 | vp[r]);
                u64 hn = d0 & vp[r];
                int hout = (int)((hp >> 63) & 1ULL) - (int)((hn >> 63) & 1ULL);
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
        bool activated = false;
#pragma unroll
        for (int r = 1; r < NB; ++r) {
            if (!activated && r == bCeil && r < bMax) {
                u64 x = ch < 4 ? peq[ch][r] : 0ULL;
                if (D[r - 1] - carry <= gw_score_boundary(m, band_k, r - 1) && (((x & 1ULL) != 0ULL) | (carry < 0))) {
                    // vp[r][j] = ~0, vn[r][j] = 0 (the entry of the PREVIOUS column is overwritten: its band is stored again); then one block step
                    if (j > 0) {
#pragma unroll
                        for (int q = 0; q < NB; ++q) if (q == r) { pp[q] = ~0ULL; pn[q] = 0ULL; }
                        const int plo = gw_dp_band_lo(m, d.ncols, j - 1);
                        gw_hist_put(&hist[(size_t)j * stride], gw_dp_band<NB>(pp, plo), gw_dp_band<NB>(pn, plo));
                    }
                    u64 vpr = ~0ULL, vnr = 0ULL;
                    if (carry < 0) x |= 1ULL;
                    u64 d0 = (((x & vpr) + vpr) ^ vpr) | x | vnr;
                    u64 hp = vnr | ~(d0 | vpr);
                    u64 hn = d0 & vpr;
                    int hout = (int)((hp >> 63) & 1ULL) - (int)((hn >> 63) & 1ULL);
                    u64 hp2 = hp << 1, hn2 = hn << 1;
                    if (carry < 0) hn2 |= 1ULL;
                    if (carry > 0) hp2 |= 1ULL;
                    vp[r] = hn2 | ~(d0 | hp2);
                    vn[r] = d0 & hp2;
                    D[r] = D[r - 1] - carry + hout;
                    bCeil = r + 1;
                    computed = bCeil;
                    activated = true;
                }
            }
        }
        if (!activated) {   // while (bCeil > 1 && D[bCeil - 1] > scoreBoundary[bCeil - 1] + w) --bCeil;  (:439-441)
#pragma unroll
            for (int r = NB - 1; r >= 1; --r)
                if (bCeil == r + 1 && D[r] > gw_score_boundary(m, band_k, r) + w) bCeil = r;
        }
        // what the tables hold for this column: the computed blocks; blocks not computed in this step keep the zero initial value of the
        // Java arrays (a block deactivated just now was computed and keeps its values, as in the reference). Its band goes to the history.
#pragma unroll
        for (int r = 0; r < NB; ++r) { const bool live = r < computed; pp[r] = live ? vp[r] : 0ULL; pn[r] = live ? vn[r] : 0ULL; }
        {
            const int lo = gw_dp_band_lo(m, d.ncols, j);
            gw_hist_put(&hist[(size_t)(j + 1) * stride], gw_dp_band<NB>(pp, lo), gw_dp_band<NB>(pn, lo));
        }
        if (bCeil == bMax) {
            int dl = D[0];
#pragma unroll
            for (int r = 1; r < NB; ++r) if (r == bMax - 1) dl = D[r];
            if (!hasBest) { hasBest = true; bestTail = j; bestDiff = dl; }
            else if (bestDiff > dl) { bestTail = j; bestDiff = dl; }
        }
    }
#pragma unroll
    for (int r = 0; r < NB; ++r) { d.D[r] = D[r]; d.vp[r] = vp[r]; d.vn[r] = vn[r]; }
    d.j = jEnd; d.bceil = bCeil; d.has_best = hasBest ? 1 : 0; d.best_tail = bestTail; d.best_diff = bestDiff;
}

// nb_hint: a caller whose lanes run together passes the largest block count among them, so that all of them take the same
// instantiation (a lane with fewer blocks only skips the guarded parts) instead of running two instantiations one after the other
template <typename RefSym>
GW_HD void gw_dp_columns(GwDp& d, int jEnd, int band_k, const u64 (*peq)[GW_MAX_NB], RefSym refsym, u64* hist, size_t stride, int nb_hint = 0) {
    const int nb = d.bmax > nb_hint ? d.bmax : nb_hint;
    if (nb <= 1) gw_dp_columns_nb<1>(d, jEnd, band_k, peq, refsym, hist, stride);
    else if (nb <= 2) gw_dp_columns_nb<2>(d, jEnd, band_k, peq, refsym, hist, stride);
    else gw_dp_columns_nb<GW_MAX_NB>(d, jEnd, band_k, peq, refsym, hist, stride);
}

// What the reference's VP / VN tables hold for window column `col` at `row`, by replaying the recurrence from column 0 to col + 1 (the
// column after `col` may switch a block on, which overwrites that block's entry of `col`). Out of line and slow on purpose: only cells
// that left the band of the history need it.
template <typename RefSym>
GW_COLD void gw_dp_replay(int m, int ncols, int band_k, const u64 (*peq)[GW_MAX_NB], RefSym refsym, int col, int row, int* vpbit, int* vnbit) {
    const int w = 64, bMax = gw_max(1, (m + 63) / 64);
    int D[GW_MAX_NB], bCeil = gw_min(gw_max(1, (band_k + 63) / 64), bMax);
    u64 vp[GW_MAX_NB], vn[GW_MAX_NB], sp[GW_MAX_NB], sn[GW_MAX_NB];
    for (int r = 0; r < GW_MAX_NB; ++r) { D[r] = 0; vp[r] = ~0ULL; vn[r] = 0ULL; sp[r] = 0ULL; sn[r] = 0ULL; }
    D[0] = m;
    const int j1 = gw_min(col + 2, ncols);
    for (int j = 0; j < j1; ++j) {
        const int ch = refsym(j);
        int carry = 0, computed = bCeil, act = -1;
        for (int r = 0; r < bCeil; ++r) {
            u64 x = ch < 4 ? peq[ch][r] : 0ULL;
            if (carry < 0) x |= 1ULL;
            const u64 d0 = (((x & vp[r]) + vp[r]) ^ vp[r]) | x | vn[r];
            const u64 hp = vn[r] | ~(d0 | vp[r]), hn = d0 & vp[r];
            const int hout = (int)((hp >> 63) & 1ULL) - (int)((hn >> 63) & 1ULL);
            u64 hp2 = hp << 1, hn2 = hn << 1;
            if (carry < 0) hn2 |= 1ULL;
            if (carry > 0) hp2 |= 1ULL;
            vp[r] = hn2 | ~(d0 | hp2); vn[r] = d0 & hp2;
            D[r] += hout; carry = hout;
        }
        if (bCeil < bMax) {
            const int r = bCeil;
            u64 x = ch < 4 ? peq[ch][r] : 0ULL;
            if (D[r - 1] - carry <= gw_score_boundary(m, band_k, r - 1) && (((x & 1ULL) != 0ULL) | (carry < 0))) {
                const u64 vpr = ~0ULL, vnr = 0ULL;
                if (carry < 0) x |= 1ULL;
                const u64 d0 = (((x & vpr) + vpr) ^ vpr) | x | vnr;
                const u64 hp = vnr | ~(d0 | vpr), hn = d0 & vpr;
                const int hout = (int)((hp >> 63) & 1ULL) - (int)((hn >> 63) & 1ULL);
                u64 hp2 = hp << 1, hn2 = hn << 1;
                if (carry < 0) hn2 |= 1ULL;
                if (carry > 0) hp2 |= 1ULL;
                vp[r] = hn2 | ~(d0 | hp2); vn[r] = d0 & hp2;
                D[r] = D[r - 1] - carry + hout;
                bCeil = r + 1; computed = bCeil; act = r;
            }
        }
        if (act < 0) while (bCeil > 1 && D[bCeil - 1] > gw_score_boundary(m, band_k, bCeil - 1) + w) --bCeil;
        if (j == col) for (int r = 0; r < GW_MAX_NB; ++r) { const bool live = r < computed; sp[r] = live ? vp[r] : 0ULL; sn[r] = live ? vn[r] : 0ULL; }
        else if (j == col + 1 && act >= 0) { sp[act] = ~0ULL; sn[act] = 0ULL; }
    }
    const int b = row >> 6, off = row & 63;
    *vpbit = b < GW_MAX_NB ? (int)((sp[b] >> off) & 1ULL) : 0;
    *vnbit = b < GW_MAX_NB ? (int)((sn[b] >> off) & 1ULL) : 0;
}

// qsym(i): query symbol (0..4) of row i
template <typename RefSym, typename QSym>
GW_HD void gw_dp_traceback(const GwDp& d, int band_k, const u64 (*peq)[GW_MAX_NB], RefSym refsym, QSym qsym, const u64* hist, size_t stride, uint32_t* cig_out,
                           GwVerifyOut* out) {
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
    int hcol = -2;
    u64 v = 0ULL;
    for (;;) {
        int path = 0;
        if (col >= 0 && row >= 0) {
            if (refsym(col) == qsym(row)) path = 1;
            else {
                int vpb, vnb;
                const int rel = row - gw_dp_band_lo(m, d.ncols, col);
                if (rel >= 0 && rel < 32) {
                    if (col != hcol) { v = gw_hist_get(&hist[(size_t)(col + 1) * stride]); hcol = col; }
                    vpb = (int)((v >> rel) & 1ULL);
                    vnb = (int)((v >> (32 + rel)) & 1ULL);
                } else
                    gw_dp_replay(m, d.ncols, band_k, peq, refsym, col, row, &vpb, &vnb);
                if (vpb) path = 3;
                else if (!vnb) path = 2;
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

template <typename RefSym, typename QSym>
GW_HD void gw_verify(int m, int Ncols, int band_k, const u64 (*peq)[GW_MAX_NB], RefSym refsym, QSym qsym,
                                          u64* hist, size_t stride, uint32_t* cig_out, GwVerifyOut* out) {
    GwDp d;
    gw_dp_begin(d, m, Ncols, band_k, hist, stride);
    gw_dp_columns(d, Ncols, band_k, peq, refsym, hist, stride);
    gw_dp_traceback(d, band_k, peq, refsym, qsym, hist, stride, cig_out, out);
}
