// gw_verify.cuh — bit-vector DP verifier with CIGAR traceback (one thread per candidate fragment).
// Replaces BitParallelSmithWaterman.AlignBlocksDetailed (reference:
// align/src/main/java/org/utgenome/weaver/align/BitParallelSmithWaterman.java): align :387-459 (block-wise Myers /
// Hyyro recurrence with adaptive number of active 64-row blocks), alignBlock :461-491, traceback :501-644
// (ops string, leading/trailing indel -> soft clip rewriting, run-length CIGAR).
// The VP/VN history needed by the traceback (bMax x (N'+1) words each) is staged in a scratch area laid out
// [column][block][request] so that the 32 lanes of a warp write 512 contiguous bytes per (column, block).
#pragma once
#include "gw_fm.cuh"

struct GwVerifyOut {
    int valid;   // 0: reference returns null
    int pos, diff, cig_n;
    bool cig_overflow;
};


// hist: scratch base for this request; element (col, blk) at hist[((col * bMax + blk) * stride)] as GwU64x2 (vp, vn)
// peq: [4][bMax] pattern masks of the query (bits >= m clear); refsym(j): reference symbol (0..4) of window column j
// qsym(i): query symbol (0..4) of row i
template <typename RefSym, typename QSym>
GW_HD void gw_verify(int m, int Ncols, int band_k, const u64 (*peq)[GW_MAX_NB], RefSym refsym, QSym qsym,
                                          GwU64x2* hist, size_t stride, uint32_t* cig_out, GwVerifyOut* out) {
    const int w = 64;
    const int bMax = gw_max(1, (m + w - 1) / w);
    int D[GW_MAX_NB];
    int scoreBoundary[GW_MAX_NB];
    u64 vp[GW_MAX_NB], vn[GW_MAX_NB];  // column j values of the active blocks
#pragma unroll
    for (int r = 0; r < GW_MAX_NB; ++r) {
        D[r] = 0; vp[r] = ~0ULL; vn[r] = 0ULL;
        scoreBoundary[r] = gw_max(m - ((r + 1) * w) + band_k, 0);
    }
    D[0] = m;
    for (int r = 0; r < bMax; ++r) hist[(size_t)r * stride] = GwU64x2{~0ULL, 0ULL};  // column 0
    int bCeil = gw_max(1, (band_k + w - 1) / w);
    if (bCeil > bMax) bCeil = bMax;  // band_k <= 64 in practice; keeps indices in range
    bool hasBest = false;
    int bestTail = -1, bestDiff = 0;
    for (int j = 0; j < Ncols; ++j) {
        const int ch = refsym(j);
        int carry = 0;
        GwU64x2* col = hist + (size_t)(j + 1) * bMax * stride;
        int computed = bCeil;  // blocks whose column j+1 is written in this step
#pragma unroll
        for (int r = 0; r < GW_MAX_NB; ++r) {
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
                col[(size_t)r * stride] = GwU64x2{vp[r], vn[r]};
                D[r] += hout;
                carry = hout;
            }
        }
        bool activated = false;
        if (bCeil < bMax) {
            const int r = bCeil;
            u64 x = ch < 4 ? peq[ch][r] : 0ULL;
            if (D[r - 1] - carry <= scoreBoundary[r - 1] && (((x & 1ULL) != 0ULL) | (carry < 0))) {
                // vp[r][j] = ~0, vn[r][j] = 0 ; then one block step
                hist[((size_t)j * bMax + r) * stride] = GwU64x2{~0ULL, 0ULL};
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
                col[(size_t)r * stride] = GwU64x2{vp[r], vn[r]};
                D[r] = D[r - 1] - carry + hout;
                bCeil++;
                computed = bCeil;
                activated = true;
            }
        }
        if (!activated) {
            while (bCeil > 1 && D[bCeil - 1] > scoreBoundary[bCeil - 1] + w) --bCeil;
        }
        // blocks not computed in this step: column j+1 keeps the zero initial value of the Java arrays
        // (a block deactivated just now was computed and keeps its values, as in the reference)
        for (int r = computed; r < bMax; ++r) col[(size_t)r * stride] = GwU64x2{0ULL, 0ULL};
        if (bCeil == bMax) {
            if (!hasBest) { hasBest = true; bestTail = j; bestDiff = D[bCeil - 1]; }
            else if (bestDiff > D[bCeil - 1]) { bestTail = j; bestDiff = D[bCeil - 1]; }
        }
    }
    out->cig_overflow = false;
    if (!hasBest) { out->valid = 0; out->pos = 0; out->diff = 0; out->cig_n = 0; return; }

    // ---- traceback: ops are produced right-to-left, 2 bits each (0 M, 1 I, 2 D, 3 S)
    uint32_t ops[(GW_MAX_READ_LEN + GW_MAX_WINDOW + 15) / 16];
    int nops = 0;
    auto push = [&](uint32_t op) { ops[nops >> 4] = (nops & 15) ? (ops[nops >> 4] | (op << ((nops & 15) * 2))) : op; ++nops; };
    int leftMostPos = 0, diff = 0;
    int col = bestTail, row = m - 1;
    for (;;) {
        int path = 0;
        if (col >= 0 && row >= 0) {
            int block = row >> 6, offset = row & 63;
            GwU64x2 v = hist[((size_t)(col + 1) * bMax + block) * stride];
            bool vpf = (v.x >> offset) & 1ULL, vnf = (v.y >> offset) & 1ULL;
            if (refsym(col) == qsym(row)) path = 1;
            else if (vpf) { path = 3; diff++; }
            else if (!vnf) { path = 2; diff++; }
            else { path = 4; diff++; }
        }
        if (path == 1 || path == 2) { push(0); leftMostPos = col; col--; row--; }
        else if (path == 3) { push(1); leftMostPos = col + 1; row--; }
        else if (path == 4) { push(2); col--; }
        else {
            while (col >= 0 || row >= 0) { if (row >= 0) push(3); col--; row--; }
            break;
        }
    }
    // cigarStr = reverse(ops); strip leading / trailing S,I,D into S (diff-- per I/D)
    auto opAt = [&](int i) -> uint32_t { int p = nops - 1 - i; return (ops[p >> 4] >> ((p & 15) * 2)) & 3u; };  // i in cigarStr order
    int left = 0, right = 0;
    for (int i = 0; i < nops; ++i) {
        uint32_t t = opAt(i);
        if (t == 3) left++;
        else if (t == 1 || t == 2) { left++; diff--; }
        else break;
    }
    for (int i = nops - 1; i >= left; --i) {
        uint32_t t = opAt(i);
        if (t == 3) right++;
        else if (t == 1 || t == 2) { right++; diff--; }
        else break;
    }
    // run-length encode: left x S, middle, right x S
    int n = 0;
    uint32_t curOp = 0xFFFFFFFFu, curLen = 0;
    bool overflow = false;
    auto emit = [&](uint32_t bamOp, uint32_t len) {
        if (len == 0) return;
        if (bamOp == curOp) { curLen += len; return; }
        if (curOp != 0xFFFFFFFFu) { if (n < GW_CIG_CAP) cig_out[n] = (curLen << 4) | curOp; else overflow = true; ++n; }
        curOp = bamOp; curLen = len;
    };
    emit(GW_OP_S, (uint32_t)left);
    for (int i = left; i < nops - right; ++i) {
        uint32_t t = opAt(i);
        emit(t == 0 ? GW_OP_M : t == 1 ? GW_OP_I : t == 2 ? GW_OP_D : GW_OP_S, 1);
    }
    emit(GW_OP_S, (uint32_t)right);
    if (curOp != 0xFFFFFFFFu) { if (n < GW_CIG_CAP) cig_out[n] = (curLen << 4) | curOp; else overflow = true; ++n; }
    out->valid = 1; out->pos = leftMostPos; out->diff = diff; out->cig_n = n < GW_CIG_CAP ? n : GW_CIG_CAP; out->cig_overflow = overflow;
}
