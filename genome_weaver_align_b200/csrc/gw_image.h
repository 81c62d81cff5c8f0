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
        for (size_t b = 0; b < nblk; ++b) {
            GwOccBlock& o = blk[b];
            for (int q = 0; q < 4; ++q) o.q[q].x = cnt[q];
            const uint64_t base = (uint64_t)b * 128;
            const unsigned lim = (unsigned)std::min<uint64_t>(128, h.N > base ? h.N - base : 0);
            for (unsigned j = 0; j < lim; ++j) {
                const unsigned c = bw[base + j];
                GwU4& qd = o.q[j >> 5];
                const uint32_t bit = 1u << (j & 31);
                if (c >= 4) qd.w |= bit;
                else {
                    if (c & 1) qd.y |= bit;
                    if (c & 2) qd.z |= bit;
                    ++cnt[c];
                }
            }
        }
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
    const size_t nw = (size_t)(h.N / 64) + 1;
    img.ref_lo.assign(nw, 0); img.ref_hi.assign(nw, 0); img.ref_nf.assign(nw, 0);
    for (uint64_t i = 0; i < h.N; ++i) {
        const unsigned c = h.text[(size_t)i];
        const u64 bit = 1ULL << (i & 63);
        if (c >= 4) img.ref_nf[(size_t)(i >> 6)] |= bit;
        else {
            if (c & 1) img.ref_lo[(size_t)(i >> 6)] |= bit;
            if (c & 2) img.ref_hi[(size_t)(i >> 6)] |= bit;
        }
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
