// gw_image.h — lays an index out in the memory image the kernels read (gw_common.h): interleaved occurrence blocks,
// character counts, bit-plane reference text and the contig table. Host code; gw_api.cu copies the arrays to HBM.
// Replaces the load-time work of FMIndexOnGenome.load / OccurrenceCountTable / CharacterCount of the reference
// (align/src/main/java/org/utgenome/weaver/align/FMIndexOnGenome.java:60-98, OccurrenceCountTable.java:51-72,
// CharacterCount.java:41-50, SequenceBoundary.java:104-121).
#pragma once
#include <algorithm>
#include <map>

#include "gw_common.h"
#include "gw_host.h"

struct GwIndexImage {
    std::vector<GwOccBlock> occ[2];
    std::vector<u64> ref_lo, ref_hi, ref_nf;
    std::vector<uint32_t> ctg_off;
    std::vector<int32_t> ctg_id, ctg_rank;
    GwIndexView view;   // scalar fields filled; pointer fields left null
};

inline void gw_layout_index(const GwHostIndex& h, GwIndexImage& img) {
    if (h.N >= 0xFFFFFFFFull) throw GwIoError("texts of 2^32-1 symbols or more are not supported");
    GwIndexView& v = img.view;
    memset(&v, 0, sizeof(v));
    v.N = (uint32_t)h.N;
    const size_t nblk = (size_t)(h.N / 128) + 1;   // like OccurrenceCountTable: (N + W) / W rows
    for (int s = 0; s < 2; ++s) {
        if (h.bwt[s].size() != h.N) throw GwIoError("BWT length differs from the text length");
        std::vector<GwOccBlock>& blk = img.occ[s];
        blk.resize(nblk);
        memset(blk.data(), 0, nblk * sizeof(GwOccBlock));
        uint32_t cnt[4] = {0, 0, 0, 0};
        const uint8_t* bw = h.bwt[s].data();
        // pieces of blocks are laid out side by side with counts relative to the start of the piece; the counts before each piece
        // are added in a second sweep
        const unsigned parts = gw_host_threads();
        std::vector<uint32_t> piece_cnt((size_t)parts * 4, 0);
        GwOccBlock* blocks = blk.data();
        const uint64_t N = h.N;
        gw_parallel_for(nblk, parts, [&](uint64_t b0, uint64_t b1, unsigned t) {
            uint32_t c4[4] = {0, 0, 0, 0};
            for (size_t b = (size_t)b0; b < (size_t)b1; ++b) {
                GwOccBlock& o = blocks[b];
                for (int q = 0; q < 4; ++q) o.q[q].x = c4[q];
                const uint64_t base = (uint64_t)b * 128;
                const unsigned lim = (unsigned)std::min<uint64_t>(128, N > base ? N - base : 0);
                for (unsigned j = 0; j < lim; ++j) {
                    const unsigned c = bw[base + j];
                    GwU4& qd = o.q[j >> 5];
                    const uint32_t bit = 1u << (j & 31);
                    if (c >= 4) qd.w |= bit;
                    else {
                        if (c & 1) qd.y |= bit;
                        if (c & 2) qd.z |= bit;
                        ++c4[c];
                    }
                }
            }
            for (int q = 0; q < 4; ++q) piece_cnt[(size_t)t * 4 + q] = c4[q];
        });
        std::vector<uint32_t> before((size_t)parts * 4, 0);
        for (unsigned t = 0; t < parts; ++t)
            for (int q = 0; q < 4; ++q) { before[(size_t)t * 4 + q] = cnt[q]; cnt[q] += piece_cnt[(size_t)t * 4 + q]; }
        gw_parallel_for(nblk, parts, [&](uint64_t b0, uint64_t b1, unsigned t) {
            for (size_t b = (size_t)b0; b < (size_t)b1; ++b)
                for (int q = 0; q < 4; ++q) blocks[b].q[q].x += before[(size_t)t * 4 + q];
        });
        uint32_t sum = 0;
        for (int c = 0; c < 4; ++c) { v.C[s][c] = sum; sum += cnt[c]; }
        v.C[s][4] = sum;
        if (s == 1) {   // FMIndexOnGenome.initRange = forwardSearch(FORWARD, [0,N)) on the reverse index (FMIndexOnGenome.java:97-98)
            v.init_valid = 0;
            for (int c = 0; c < 4; ++c) {
                v.init_lb[c] = v.C[1][c]; v.init_ub[c] = v.C[1][c] + cnt[c];
                if (cnt[c] > 0) v.init_valid |= 1u << c;
            }
        }
    }
    v.n_ssa = (uint32_t)h.ssa[0].size();
    v.loc_step = GW_SSA_L;   // the locate table is the row samples until the device densifies it (gw_api.cu)
    const size_t nw = (size_t)(h.N / 64) + 1;
    img.ref_lo.assign(nw, 0); img.ref_hi.assign(nw, 0); img.ref_nf.assign(nw, 0);
    {
        const uint8_t* text = h.text.data();
        u64 *rl = img.ref_lo.data(), *rh = img.ref_hi.data(), *rn = img.ref_nf.data();
        const uint64_t N = h.N;
        gw_parallel_for(nw, gw_host_threads(), [&](uint64_t w0, uint64_t w1, unsigned) {   // whole words per piece: no shared word
            for (uint64_t w = w0; w < w1; ++w) {
                const uint64_t lo = w * 64, hi = std::min<uint64_t>(N, lo + 64);
                u64 l = 0, hh = 0, nf = 0;
                for (uint64_t i = lo; i < hi; ++i) {
                    const unsigned c = text[(size_t)i];
                    const u64 bit = 1ULL << (i & 63);
                    if (c >= 4) nf |= bit;
                    else { if (c & 1) l |= bit; if (c & 2) hh |= bit; }
                }
                rl[w] = l; rh[w] = hh; rn[w] = nf;
            }
        });
    }
    // SequenceBoundary.translate: TreeMap offset -> name, later put wins
    std::map<int64_t, int> by_off;
    for (size_t i = 0; i < h.contigs.size(); ++i) by_off[h.contigs[i].offset] = (int)i;
    img.ctg_off.clear(); img.ctg_id.clear();
    for (auto& kv : by_off) { img.ctg_off.push_back((uint32_t)kv.first); img.ctg_id.push_back(kv.second); }
    std::vector<std::string> names;
    for (auto& c : h.contigs) names.push_back(c.name);
    std::sort(names.begin(), names.end());   // ASCII names: byte order == java.lang.String.compareTo order
    names.erase(std::unique(names.begin(), names.end()), names.end());
    img.ctg_rank.resize(h.contigs.size());
    for (size_t i = 0; i < h.contigs.size(); ++i)
        img.ctg_rank[i] = (int32_t)(std::lower_bound(names.begin(), names.end(), h.contigs[i].name) - names.begin());
    v.n_ctg_off = (int32_t)img.ctg_off.size();
}
