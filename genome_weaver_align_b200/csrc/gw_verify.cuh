// gw_verify.cuh — bit-vector DP verifier with CIGAR traceback (one thread per candidate fragment).
// Replaces BitParallelSmithWaterman.AlignBlocksDetailed (reference:
// align/src/main/java/org/utgenome/weaver/align/BitParallelSmithWaterman.java): align :387-459 (block-wise Myers /
// Hyyro recurrence with adaptive number of active 64-row blocks), alignBlock :461-491, traceback :501-644
// (ops string, leading/trailing indel -> soft clip rewriting, run-length CIGAR).
// The VP/VN history needed by the traceback (bMax x (N'+1) words each) is staged in a scratch area laid out
// [column][block][request] so that the 32 lanes of a warp write 512 contiguous bytes per (column, block).
#pragma once
#include "gw_fm.cuh"

#define GW_RUN_CAP 64   // run-length encoded traceback ops kept per fragment (>= 2 * GW_MAX_K + 3)

// The history is written once and read at most once (by the traceback, minutes of GPU time later in cache terms): its stores and loads
// carry the streaming hint, so that they do not push the work areas of the searching reads out of the L2.
GW_HD void gw_hist_put(GwU64x2* p, u64 vp, u64 vn) {
#if defined(__CUDA_ARCH__)
    __stcs(reinterpret_cast<ulonglong2*>(p), make_ulonglong2(vp, vn));
#else
    p->x = vp; p->y = vn;
#endif
}
GW_HD GwU64x2 gw_hist_get(const GwU64x2* p) {
#if defined(__CUDA_ARCH__)
    const ulonglong2 v = __ldcs(reinterpret_cast<const ulonglong2*>(p));
    GwU64x2 r; r.x = v.x; r.y = v.y;
    return r;
#else
    return *p;
#endif
}

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

// hist: scratch base for this request; element (col, blk) at hist[((col * bMax + blk) * stride)] as GwU64x2 (vp, vn)
GW_HD void gw_dp_begin(GwDp& d, int m, int Ncols, int band_k, GwU64x2* hist, size_t stride) {
    d.m = m; d.ncols = Ncols; d.bmax = gw_max(1, (m + 63) / 64); d.j = 0;
#pragma unroll
    for (int r = 0; r < GW_MAX_NB; ++r) { d.D[r] = 0; d.vp[r] = ~0ULL; d.vn[r] = 0ULL; }
    d.D[0] = m;
    for (int r = 0; r < d.bmax; ++r) gw_hist_put(&hist[(size_t)r * stride], ~0ULL, 0ULL);  // column 0
    d.bceil = gw_max(1, (band_k + 63) / 64);
    if (d.bceil > d.bmax) d.bceil = d.bmax;  // band_k <= 64 in practice; keeps indices in range
    d.has_best = 0; d.best_tail = -1; d.best_diff = 0;
}

// peq: [4][GW_MAX_NB] pattern masks of the query (bits >= m clear); refsym(j): reference symbol (0..4) of window column j
// NB: compile-time bound of the number of 64-row blocks (ceil(m / 64) <= NB): the loops over blocks are unrolled to NB, so a 100 bp
// read does not pay for the two blocks only 250 bp reads have
template <int NB, typename RefSym>
GW_HD void gw_dp_columns_nb(GwDp& d, int jEnd, int band_k, const u64 (*peq)[GW_MAX_NB], RefSym refsym, GwU64x2* hist, size_t stride) {
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
    for (int j = d.j; j < jEnd; ++j) {
        const int ch = refsym(j);
        int carry = 0;
        GwU64x2* col = hist + (size_t)(j + 1) * bMax * stride;
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
                gw_hist_put(&col[(size_t)r * stride], vp[r], vn[r]);
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
                    // vp[r][j] = ~0, vn[r][j] = 0 ; then one block step
                    gw_hist_put(&hist[((size_t)j * bMax + r) * stride], ~0ULL, 0ULL);
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
                    gw_hist_put(&col[(size_t)r * stride], vp[r], vn[r]);
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
        // blocks not computed in this step: column j+1 keeps the zero initial value of the Java arrays
        // (a block deactivated just now was computed and keeps its values, as in the reference)
#pragma unroll
        for (int r = 1; r < NB; ++r)
            if (r >= computed && r < bMax) gw_hist_put(&col[(size_t)r * stride], 0ULL, 0ULL);
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
GW_HD void gw_dp_columns(GwDp& d, int jEnd, int band_k, const u64 (*peq)[GW_MAX_NB], RefSym refsym, GwU64x2* hist, size_t stride, int nb_hint = 0) {
    const int nb = d.bmax > nb_hint ? d.bmax : nb_hint;
    if (nb <= 1) gw_dp_columns_nb<1>(d, jEnd, band_k, peq, refsym, hist, stride);
    else if (nb <= 2) gw_dp_columns_nb<2>(d, jEnd, band_k, peq, refsym, hist, stride);
    else gw_dp_columns_nb<GW_MAX_NB>(d, jEnd, band_k, peq, refsym, hist, stride);
}

// qsym(i): query symbol (0..4) of row i
template <typename RefSym, typename QSym>
GW_HD void gw_dp_traceback(const GwDp& d, RefSym refsym, QSym qsym, const GwU64x2* hist, size_t stride, uint32_t* cig_out, GwVerifyOut* out) {
    const int m = d.m, bMax = d.bmax;
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
    int hcol = -2, hblk = -1;
    GwU64x2 v = GwU64x2{0ULL, 0ULL};
    for (;;) {
        int path = 0;
        if (col >= 0 && row >= 0) {
            const int block = row >> 6, offset = row & 63;
            if (col != hcol || block != hblk) { v = gw_hist_get(&hist[((size_t)(col + 1) * bMax + block) * stride]); hcol = col; hblk = block; }
            if (refsym(col) == qsym(row)) path = 1;
            else if ((v.x >> offset) & 1ULL) { path = 3; diff++; }
            else if (!((v.y >> offset) & 1ULL)) { path = 2; diff++; }
            else { path = 4; diff++; }
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
                                          GwU64x2* hist, size_t stride, uint32_t* cig_out, GwVerifyOut* out) {
    GwDp d;
    gw_dp_begin(d, m, Ncols, band_k, hist, stride);
    gw_dp_columns(d, Ncols, band_k, peq, refsym, hist, stride);
    gw_dp_traceback(d, refsym, qsym, hist, stride, cig_out, out);
}
