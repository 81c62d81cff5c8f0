// gw_bodies.cuh — per-thread bodies of the batch kernels (prep, quick scan, emit, settle).
// Each function is the work of ONE thread of the kernel of the same name in gw_kernels.cu; the search / locate / verify path is
// the flow kernel over gw_automaton.cuh and gw_job.cuh. Reference methods replaced (align/src/main/java/org/utgenome/weaver/align/):
//   prep     BidirectionalSuffixFilter.AlignmentProcess ctor + N handling  strategy/BidirectionalSuffixFilter.java:189-198, :281-291
//   scan     FMQuickScan.scanMismatchLocations                             strategy/FMQuickScan.java:66-94
//   emit     AlignmentProcess.align (choice of reported hits)              strategy/BidirectionalSuffixFilter.java:210-268
#pragma once
#include "gw_automaton.cuh"
#include "gw_verify.cuh"

// ACGT.java:37-44
GW_HD int gw_code_of(unsigned char ch) {
    switch (ch) {
    case 'A': case 'a': return 0;
    case 'C': case 'c': return 1;
    case 'G': case 'g': return 2;
    case 'T': case 't': case 'U': case 'u': return 3;
    default: return 4;
    }
}

struct GwBatchView {
    const char* bases;        // concatenated ASCII reads
    const u64* offsets;       // n+1
    const int32_t* k_by_m;    // [GW_MAX_READ_LEN+1] AlignmentScoreConfig.getMaximumEditDistance(m); -1: unsupported
    GwExtPool* ext_pool;      // arena extension blocks of this wave, or null
};

// ---- prep: encode one read into its slab -----------------------------------------------------------------
GW_HD void gw_prep_read(const GwBatchView& bv, u64* slab_base, const GwSlabParams& sp, uint32_t read_id) {
    GwSlab s(slab_base, sp);
    GwReadCtx* c = s.ctx();
    const u64 b0 = bv.offsets[read_id], b1 = bv.offsets[read_id + 1];
    const long long len = (long long)(b1 - b0);
    s.layout_init();
    c->ext_pool = bv.ext_pool;
    c->read_id = (int32_t)read_id;
    c->m = (int32_t)(len > 0x7fffffffLL ? 0x7fffffff : len);
    c->k = 0;
    c->max_match_len = 0; c->best_score = -1;
    c->n_fm = c->n_sw = c->n_cut = c->n_filt = 0;
    c->qn = 0; c->n_hits = 0; c->free1 = -1;
    c->top = (int32_t)s.arena_begin();
    c->pend_base = -1; c->pend_n = 0;
    c->sf_mm = -2; c->sf_off = -1;
    c->count_n = 0;
    c->occ_scan = c->occ_search = c->occ_locate = 0;
    c->status = GW_ST_OK;
    c->phase = GW_PH_INIT;
    for (int st = 0; st < 2; ++st) { c->scan_mm[st] = -1; c->scan_ls[st] = c->scan_le[st] = 0; c->scan_lb[st] = c->scan_ub[st] = 0; }
    if (len <= 0 || len > GW_MAX_READ_LEN || bv.k_by_m[len] < 0) {
        c->status = GW_ST_UNSUPPORTED; c->phase = GW_PH_DONE; c->min_mm = 1;
        return;
    }
    const int m = (int)len;
    const int k = bv.k_by_m[m];
    c->k = k;
    c->min_mm = k + 1;
    int countN = 0;
    u64 lo[GW_MAX_NB], hi[GW_MAX_NB];
    for (int b = 0; b < GW_MAX_NB; ++b) { lo[b] = 0; hi[b] = 0; }
    for (int i = 0; i < m; ++i) {
        int code = gw_code_of((unsigned char)bv.bases[b0 + (u64)i]);
        if (code == 4) { ++countN; code = 0; }  // replaceN_withA
        lo[i >> 6] |= (u64)(code & 1) << (i & 63);
        hi[i >> 6] |= (u64)(code >> 1) << (i & 63);
    }
    c->count_n = countN;
    const int nb = (m + 63) >> 6;
    for (int b = 0; b < GW_MAX_NB; ++b) {
        u64 valid = b < nb - 1 ? ~0ULL : (b == nb - 1 ? ((m & 63) ? ((1ULL << (m & 63)) - 1ULL) : ~0ULL) : 0ULL);
        c->qlo[0][b] = lo[b]; c->qhi[0][b] = hi[b];
        // q[1] = complement, N -> A in both: a replaced N is A on both strands, so complement every other base only
        c->qlo[1][b] = ~lo[b] & valid; c->qhi[1][b] = ~hi[b] & valid;
    }
    if (countN > 0) {
        for (int i = 0; i < m; ++i)
            if (gw_code_of((unsigned char)bv.bases[b0 + (u64)i]) == 4) {
                c->qlo[1][i >> 6] &= ~(1ULL << (i & 63));
                c->qhi[1][i >> 6] &= ~(1ULL << (i & 63));
            }
    }
    if (countN > k) { c->phase = GW_PH_DONE; }  // too many N: unmapped (:281-284)
}

// ---- quick scan (single-thread form): one (read, strand) task -----------------------------------------------
GW_HD void gw_scan_task(const GwIndexView& ix, GwReadCtx* c, int strand) {
    if (c->phase == GW_PH_DONE) return;
    const int m = c->m;
    const int fm = strand == 0 ? 1 : 0;  // FMIndexOnGenome.forwardSearch: FORWARD strand searches the reverse index
    uint32_t lb = 0, ub = ix.N;
    int numMismatches = 0, mark = 0, ls = 0, le = 0;
    bool hasLongest = false;
    const GwOccBlock* occ = ix.occ[fm];
    for (int i = 0; i < m; ++i) {
        const int ch = (int)((((c->qhi[strand][i >> 6] >> (i & 63)) & 1ULL) << 1) | ((c->qlo[strand][i >> 6] >> (i & 63)) & 1ULL));
        const uint32_t l = gw_clampN(ix, lb), u = gw_clampN(ix, ub);
        GwBlk bl = gw_load_block(occ + (l >> 7));
        uint32_t nl = ix.C[fm][ch] + gw_rank1(bl, l & 127u, ch), nu;
        if ((u >> 7) == (l >> 7)) nu = ix.C[fm][ch] + gw_rank1(bl, u & 127u, ch);
        else { GwBlk bu = gw_load_block(occ + (u >> 7)); nu = ix.C[fm][ch] + gw_rank1(bu, u & 127u, ch); }
        lb = nl; ub = nu;
        if (lb >= ub) {
            ++numMismatches;
            if (!hasLongest || (le - ls) < (i - mark)) { hasLongest = true; ls = mark; le = i; }
            lb = 0; ub = ix.N;
            mark = i + 1;
        }
    }
    if (!hasLongest || (le - ls) < (m - mark)) { ls = mark; le = m; }
    c->scan_mm[strand] = numMismatches; c->scan_ls[strand] = ls; c->scan_le[strand] = le;
    c->scan_lb[strand] = lb; c->scan_ub[strand] = ub;
    gw_atomic_add(&c->occ_scan, 2u * (uint32_t)m);  // the two strands of a read are scanned by different threads
}

// ---- emit ------------------------------------------------------------------------------------------------
// mirror of gw_hit_t in include/gw_align.h
struct GwHitRec {
    uint32_t read_index;
    int32_t chr_id, pos, match_length, diff, num_hits, q_start, q_end;
    uint32_t cigar_off, cigar_len;
    int8_t strand;
    uint8_t is_head, status, reserved;
};

// number of reported chains of a finished read (AlignmentProcess.align :215-261); 0 = the "*" record
GW_HD int gw_num_reported(const GwReadCtx* c, const GwDevConfig& cfg) {
    if (c->status != GW_ST_OK) return 0;
    const bool hasHit = c->min_mm <= c->k && c->n_hits > 0;
    if (!hasHit) return 0;
    if (cfg.report_type == 1) return c->n_hits;
    if (cfg.report_type == 0) return 1;
    return gw_min(cfg.top_l, c->n_hits);
}

GW_HD void gw_emit_count(u64* slab_base, const GwDevConfig& cfg, uint32_t* n_rec, uint32_t* n_cig) {
    GwSlab s(slab_base);
    const GwReadCtx* c = s.ctx();
    const int nrep = gw_num_reported(c, cfg);
    uint32_t rec = 0, cig = 0;
    if (nrep == 0) rec = 1;
    else {
        const int32_t* hl = s.hits();
        for (int i = 0; i < nrep; ++i)
            for (int32_t o = hl[i]; o >= 0; o = gw_hit(s, o)->next_split) { ++rec; cig += (uint32_t)gw_hit(s, o)->cig_n; }
    }
    *n_rec = rec; *n_cig = cig;
}

GW_HD void gw_emit_write(u64* slab_base, const GwDevConfig& cfg, uint32_t read_index, GwHitRec* recs, uint64_t rec_off, uint64_t rec_cap,
                         uint32_t* cigs, uint64_t cig_off, uint64_t cig_cap) {
    GwSlab s(slab_base);
    const GwReadCtx* c = s.ctx();
    const int nrep = gw_num_reported(c, cfg);
    if (nrep == 0) {
        if (rec_off >= rec_cap) return;
        GwHitRec r;
        r.read_index = read_index; r.chr_id = GW_CHR_STAR_; r.pos = 0; r.match_length = 0; r.diff = -1; r.num_hits = 0;
        r.q_start = 0; r.q_end = 0; r.cigar_off = (uint32_t)cig_off; r.cigar_len = 0; r.strand = 0; r.is_head = 1;
        r.status = (uint8_t)(c->status == GW_ST_OK ? GW_ST_UNMAPPED : (c->status == GW_ST_OVERFLOW ? GW_ST_UNSUPPORTED : c->status));
        r.reserved = 0;
        recs[rec_off] = r;
        return;
    }
    const int32_t* hl = s.hits();
    for (int i = 0; i < nrep; ++i) {
        bool head = true;
        for (int32_t o = hl[i]; o >= 0; o = gw_hit(s, o)->next_split) {
            GwHitNode* h = gw_hit(s, o);
            if (rec_off < rec_cap) {
                GwHitRec r;
                r.read_index = read_index; r.chr_id = h->chr; r.pos = h->pos; r.match_length = h->match_length; r.diff = h->diff;
                r.num_hits = h->num_hits; r.q_start = h->q_start; r.q_end = h->q_end; r.cigar_off = (uint32_t)cig_off;
                r.cigar_len = (uint32_t)h->cig_n; r.strand = (int8_t)h->strand; r.is_head = head ? 1 : 0; r.status = GW_ST_OK; r.reserved = 0;
                recs[rec_off] = r;
            }
            const uint32_t* cg = gw_hit_cigar(h);
            for (int j = 0; j < h->cig_n; ++j) if (cig_off + (uint64_t)j < cig_cap) cigs[cig_off + (uint64_t)j] = cg[j];
            cig_off += (uint64_t)h->cig_n;
            ++rec_off;
            head = false;
        }
    }
}

// ---- settle ----------------------------------------------------------------------------------------------
// When the reads that need a large work area do not fit the memory budget at once, they are run in chunks that share one buffer.
// After a chunk, the outcome of each read (context with status and counters + the chains that will be reported) is written back
// into the read's own small work area, which emit reads later; the large area is then free for the next chunk.
// A read whose reported chains do not fit the small area (ALLHITS with very many hits) ends as GW_ST_UNSUPPORTED.
GW_HD void gw_settle(u64* big_base, u64* small_base, const GwSlabParams& small_sp, const GwDevConfig& cfg) {
    GwSlab big(big_base);
    const GwReadCtx* bc = big.ctx();
    GwSlab sm(small_base, small_sp);
    GwReadCtx* sc = sm.ctx();
    for (uint32_t i = 0; i < GW_CTX_WORDS; ++i) small_base[i] = big_base[i];
    sm.layout_init();
    sc->ext_pool = nullptr;
    sc->qn = 0; sc->free1 = -1;
    sc->top = (int32_t)sm.arena_begin();
    const int nrep = gw_num_reported(bc, cfg);
    sc->n_hits = 0;
    if (nrep == 0) return;
    if ((uint32_t)nrep > small_sp.hit_cap) { sc->status = GW_ST_UNSUPPORTED; return; }
    const int32_t* bh = big.hits();
    int32_t* sh = sm.hits();
    for (int i = 0; i < nrep; ++i) {
        int32_t prev = -1;
        for (int32_t o = bh[i]; o >= 0; o = gw_hit(big, o)->next_split) {
            const GwHitNode* h = gw_hit(big, o);
            const uint32_t words = GW_HIT_WORDS + (uint32_t)((h->cig_n + 1) / 2);
            const int32_t n = gw_alloc(sm, words);
            if (n < 0) { sc->n_hits = 0; sc->status = GW_ST_UNSUPPORTED; return; }
            const u64* src = big.at(o);
            u64* dst = sm.at(n);
            for (uint32_t w = 0; w < words; ++w) dst[w] = src[w];
            gw_hit(sm, n)->next_split = -1;
            if (prev < 0) sh[i] = n; else gw_hit(sm, prev)->next_split = n;
            prev = n;
        }
    }
    sc->n_hits = nrep;
}
