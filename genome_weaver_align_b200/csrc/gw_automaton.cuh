// gw_automaton.cuh — resumable per-read driver around gw_search.cuh: AlignmentProcess.align_internal
// (reference: align/src/main/java/org/utgenome/weaver/align/strategy/BidirectionalSuffixFilter.java:278-477) cut at the
// points where the reference calls into the sampled suffix array and the verifier (verify :496-560). At those points the
// read parks with up to two fragment requests; the locate and verify kernels fill them and the read resumes.
#pragma once
#include "gw_search.cuh"

struct GwRoundQueues {
    uint32_t* next_active;   // slots that parked in this round
    uint32_t* locate_req;    // slot*2 + fragment
    uint32_t* verify_req;    // slot*2 + fragment
    uint32_t* counters;      // [0] n_next_active, [1] n_locate, [2] n_verify
};

GW_HD void gw_fail(GwReadCtx* c, int status) { c->status = status; c->phase = GW_PH_DONE; }

GW_HD int gw_q_symbol(const GwReadCtx* c, int strand, int i) {
    return (int)((((c->qhi[strand][i >> 6] >> (i & 63)) & 1ULL) << 1) | ((c->qlo[strand][i >> 6] >> (i & 63)) & 1ULL));
}

// reportAlignment request side: classify the fragments of the chain headed by `base`; returns true if the read must park
GW_HD bool gw_begin_report(const GwSearchEnv& env, const GwSlab& s, uint32_t slot, int32_t base,
                                                const GwRoundQueues& rq) {
    GwReadCtx* c = s.ctx();
    c->pend_base = base;
    int n = 0;
    bool need = false;
    for (int32_t o = base; o >= 0 && n < 2; o = gw_st(s, o)->next_split, ++n) {
        const GwState* st = gw_st(s, o);
        GwPendFrag* pf = &c->pend[n];
        pf->state_off = o;
        pf->res_abort = 0;
        if (gw_is_clipped(st)) pf->kind = 0;
        else if (!st->has_cur) pf->kind = 1;
        else {
            pf->kind = 2;
            pf->sa_index = st->cur_lb;
            // FMIndexOnGenome.toCoordinate: fm = ~(strand ^ (dir.isForward ? 0 : 1)) & 1
            pf->fm = ~(gw_strand(st) ^ (gw_is_fwd(st) ? 0 : 1)) & 1;
            if (rq.counters) {   // round mode; the fused kernel serves the request on the same lane (gw_fused.cuh)
                uint32_t id = slot * 2u + (uint32_t)n;
                rq.locate_req[gw_atomic_add(&rq.counters[1], 1u)] = id;
                rq.verify_req[gw_atomic_add(&rq.counters[2], 1u)] = id;
            }
            need = true;
        }
    }
    c->pend_n = n;
    return need;
}

// reportAlignment completion: ReadHits from the verified fragments, sortSplits, AlignmentResultHolder.add
GW_HD void gw_finish_report(const GwSearchEnv& env, const GwSlab& s) {
    GwReadCtx* c = s.ctx();
    const int k = c->k;
    int32_t hits[2] = {-1, -1};
    for (int f = 0; f < c->pend_n; ++f) {
        GwPendFrag* pf = &c->pend[f];
        const GwState* st = gw_st(s, pf->state_off);
        const int strand = gw_strand(st);
        int32_t h;
        if (pf->kind == 0) {
            uint32_t cg = ((uint32_t)gw_frag_len(st) << 4) | GW_OP_S;
            h = gw_new_hit(s, GW_CHR_NULL_, -1, 0, st->start, st->end, 0, strand, &cg, 1, 0);
        } else if (pf->kind == 1) {
            h = gw_new_hit(s, GW_CHR_EMPTY_, 0, 0, 0, 0, 0, strand, nullptr, 0, 0);
        } else {
            if (pf->res_abort) { gw_fail(c, pf->res_abort == 2 ? GW_ST_UNSUPPORTED : GW_ST_ABORT); return; }
            c->n_sw++;
            if (!pf->res_valid) h = gw_new_hit(s, GW_CHR_EMPTY_, 0, 0, 0, 0, 0, strand, nullptr, 0, 0);
            else {
                int chr, pos;
                if (!gw_translate(env.ix, pf->ref_start + pf->res_pos + 1, &chr, &pos)) { gw_fail(c, GW_ST_ABORT); return; }
                h = gw_new_hit(s, chr, pos, gw_frag_len(st), st->start, st->end, pf->res_diff, strand, pf->cig, pf->res_cig_n,
                               (int32_t)(st->cur_ub - st->cur_lb));
            }
        }
        if (h < 0) { gw_fail(c, GW_ST_OVERFLOW); return; }
        hits[f] = h;
    }
    if (c->pend_n == 2) gw_hit(s, hits[0])->next_split = hits[1];
    GwHitNode* head = gw_hit(s, hits[0]);
    if (gw_hit_total_match(s, head) == 0) return;
    const int newK = gw_hit_total_diff(s, head);
    if (newK > k) return;
    gw_set_lower_k(gw_st(s, c->pend_base), newK);
    int32_t first = hits[0];
    if (c->pend_n == 2) {  // ReadHit.sortSplits on two fragments
        GwHitNode* o1 = gw_hit(s, hits[0]);
        GwHitNode* o2 = gw_hit(s, hits[1]);
        int diff = 0;
        bool anyNull = o1->chr == GW_CHR_NULL_ || o2->chr == GW_CHR_NULL_;
        if (anyNull) {
            diff = o1->q_start - o2->q_start;
            if (head->strand != 0) diff = -diff;
        }
        if (diff == 0) {
            if (anyNull) { gw_fail(c, GW_ST_ABORT); return; }  // NullPointerException in the comparator
            int r1 = o1->chr >= 0 ? env.ix.ctg_rank[o1->chr] : -1, r2 = o2->chr >= 0 ? env.ix.ctg_rank[o2->chr] : -1;
            diff = r1 - r2;
            if (diff == 0) diff = o1->pos - o2->pos;
        }
        if (diff > 0) { o2->next_split = hits[0]; o1->next_split = -1; first = hits[1]; }
    }
    if (!gw_add_result(s, env.cfg, first)) gw_fail(c, GW_ST_OVERFLOW);
}

// staircase table of the current minMismatches; returns nullptr and fails the read where the reference constructor throws
GW_HD const u64* gw_staircase_for(const GwSearchEnv& env, GwReadCtx* c, int* kk) {
    const int m = c->m, mm = c->min_mm;
    *kk = mm;
    if (mm < 0 || ((env.sf.abort_by_m[m] >> mm) & 1u)) { gw_fail(c, GW_ST_ABORT); return nullptr; }
    const int nb = (m + 63) >> 6;
    return env.sf.words + env.sf.base_by_m[m] + nb * (mm * (mm + 1) / 2);
}

// ---- resumable driver ------------------------------------------------------------------------------------
// A read is advanced by gw_resume_read (once, when a thread picks the read up: consumes locate/verify results, seeds the
// queue) followed by gw_step_read calls, each of which is ONE iteration of the reference's `while (!queue.isEmpty())` loop
// (:352-475). Splitting the loop body out lets the search kernel keep all lanes of a warp busy: a lane whose read parks or
// finishes fetches the next read instead of idling until the slowest read of the warp is done.
#define GW_STEP_CONTINUE 0
#define GW_STEP_PARKED 1
#define GW_STEP_DONE 2

// returns GW_STEP_CONTINUE when the read is in the best-first loop and wants gw_step_read calls
GW_HD int gw_resume_read(const GwSearchEnv& env, const GwSlab& s, uint32_t slot, const GwRoundQueues& rq) {
    GwReadCtx* c = s.ctx();
    const int m = c->m, k = c->k;
    const GwDevConfig& cfg = env.cfg;
    if (c->phase == GW_PH_DONE) return GW_STEP_DONE;
    if (c->phase == GW_PH_WAIT_EXACT) {  // reportExactMatchAlignment after the locate
        const int strand = c->pend[0].state_off;
        long long pos = strand == 0 ? (long long)env.ix.N - (long long)c->pend[0].textpos : (long long)c->pend[0].textpos;
        if (strand == 0) pos -= m;
        pos += 1;
        int chr, p;
        if (pos < 0 || !gw_translate(env.ix, pos, &chr, &p)) { gw_fail(c, GW_ST_ABORT); return GW_STEP_DONE; }
        uint32_t cg = ((uint32_t)m << 4) | GW_OP_M;
        int32_t h = gw_new_hit(s, chr, p, m, 0, m, 0, strand, &cg, 1, (int32_t)(c->scan_ub[strand] - c->scan_lb[strand]));
        if (h < 0 || !gw_add_result(s, cfg, h)) { gw_fail(c, GW_ST_OVERFLOW); return GW_STEP_DONE; }
        c->phase = GW_PH_DONE;
        return GW_STEP_DONE;
    }
    if (c->phase == GW_PH_WAIT_VERIFY) {
        c->phase = GW_PH_RUN;
        gw_finish_report(env, s);
        if (c->phase == GW_PH_DONE) return GW_STEP_DONE;
    }
    if (c->phase == GW_PH_INIT) {
        for (int strand = 0; strand < 2; ++strand) {
            if (c->scan_mm[strand] == 0) {
                c->min_mm = 0;
                c->pend[0].state_off = strand;
                c->pend[0].kind = 2;
                c->pend[0].sa_index = c->scan_lb[strand];
                c->pend[0].fm = strand == 0 ? 1 : 0;  // FMIndexOnGenome.toCoordinate(saIndex, strand)
                c->pend_n = 1;
                if (rq.counters) {
                    rq.locate_req[gw_atomic_add(&rq.counters[1], 1u)] = slot * 2u;
                    rq.next_active[gw_atomic_add(&rq.counters[0], 1u)] = slot;
                }
                c->phase = GW_PH_WAIT_EXACT;
                return GW_STEP_PARKED;
            }
        }
        if (k == 0) { c->phase = GW_PH_DONE; return GW_STEP_DONE; }
        for (int strand = 0; strand < 2; ++strand) {
            if (c->scan_mm[strand] <= k) {
                int ls = c->scan_ls[strand];
                int32_t st = (ls != 0 && ls < m) ? gw_initial_state(env, s, strand, GW_DIR_BF, 0, m, ls, ls, k, 0, c->scan_mm[strand])
                                                 : gw_initial_state(env, s, strand, GW_DIR_F, 0, m, 0, 0, k, 0, c->scan_mm[strand]);
                if (st < 0 || !gw_queue_add(s, cfg, st)) { gw_fail(c, GW_ST_OVERFLOW); return GW_STEP_DONE; }
            }
        }
        c->phase = GW_PH_RUN;
    }
    return GW_STEP_CONTINUE;
}

// parks a running read without a request: it is simply continued in the next round (used to bound the tail of a launch)
GW_HD void gw_yield_read(const GwSlab& s, uint32_t slot, const GwRoundQueues& rq) {
    (void)s;
    rq.next_active[gw_atomic_add(&rq.counters[0], 1u)] = slot;
}

// one iteration of the best-first loop. The polled state is copied to registers once; all candidate extensions (the read's own
// base first, then A,C,G,T) go through ONE FM step + NFA step + insertion site, which keeps the kernel small and lets the lanes
// of a warp that are in the same loop trip run together.
GW_HD int gw_step_read(const GwSearchEnv& env, const GwSlab& s, uint32_t slot, const GwRoundQueues& rq) {
    GwReadCtx* c = s.ctx();
    const int m = c->m, k = c->k;
    const GwDevConfig& cfg = env.cfg;
    if (c->qn <= 0 || c->n_fm > m * 20) { c->phase = GW_PH_DONE; return GW_STEP_DONE; }   // queue empty / fmIndexSearchUpperBound
    const int32_t base = gw_queue_poll(s, cfg);
    int32_t c_off = base;
    GwState C = gw_load_state(s, base);
    for (;;) {   // first fragment of the chain that still has bases to process (:366-379)
        if (!(gw_has_hit(&C) || gw_is_clipped(&C) || gw_remaining(&C) == 0)) break;
        if (C.next_split < 0) {
            if (gw_begin_report(env, s, slot, base, rq)) {
                c->phase = GW_PH_WAIT_VERIFY;
                if (rq.counters) rq.next_active[gw_atomic_add(&rq.counters[0], 1u)] = slot;
                return GW_STEP_PARKED;
            }
            gw_finish_report(env, s);
            return c->phase == GW_PH_DONE ? GW_STEP_DONE : GW_STEP_CONTINUE;
        }
        c_off = C.next_split;
        C = gw_load_state(s, c_off);
    }
    if (gw_is_finished(&C)) return GW_STEP_CONTINUE;
    const int nm = gw_lower_k(&C);
    if (nm > c->min_mm) { c->n_cut++; return GW_STEP_CONTINUE; }
    if (c->min_mm - nm < 0) return GW_STEP_CONTINUE;
    {
        const int scoreUB = gw_score(s, cfg, gw_st(s, base), true);
        if (scoreUB < 0 || scoreUB < c->best_score) { c->n_cut++; return GW_STEP_CONTINUE; }
    }
    const int strand = gw_strand(&C);
    const int nextBase = gw_q_symbol(c, strand, gw_next_index(&C));
    // the staircase filter of the current minMismatches is needed on every path from here (:425, :439); where the reference's
    // constructor throws, the read ends here whichever path asked for it
    int kk;
    const u64* tbl = gw_staircase_for(env, c, &kk);
    if (!tbl) return GW_STEP_DONE;
    u64 prev[GW_MAX_ROWS];
    {
        const u64* rows = gw_rows(gw_st(s, c_off));
        for (int i = 0; i < C.height; ++i) prev[i] = rows[i];
    }
    int flags = C.state;
    bool extended = false;   // the read's own base went through: the parent is dropped (:426-429)
    int fail_status = 0;
#pragma unroll 1
    for (int t = 0; t < 5; ++t) {
        const int ch = t == 0 ? nextBase : t - 1;
        if ((flags >> ch) & 1) continue;
        flags |= 1 << ch;
        if (gw_si_is_empty(&C, ch)) continue;
        GwSiSet nsi;
        gw_next_si(env.ix, &C, ch, &nsi, &c->occ_search);
        c->n_fm++;
        const int32_t ns = gw_next_state(s, C, prev, ch, nsi, tbl, kk);
        if (ns == -2) { fail_status = GW_ST_OVERFLOW; break; }
        if (ns >= 0) {
            if (!gw_queue_add(s, cfg, gw_update(s, base, c_off, ns))) { fail_status = GW_ST_OVERFLOW; break; }
            if (t == 0) { extended = true; break; }
        } else
            c->n_filt++;
    }
    if (!extended) flags |= 1 << 4;
    gw_st(s, c_off)->state = flags;   // checked-base flags live on the (possibly shared) state object
    if (fail_status) { gw_fail(c, fail_status); return GW_STEP_DONE; }
    if (extended) return GW_STEP_CONTINUE;
    C.state = flags;
    if (gw_num_split(s, gw_st(s, base)) < cfg.num_split_allowed && nm + 1 <= c->min_mm) {
        const int index = gw_next_index(&C);
        if (index > cfg.indel_end_skip && m - index >= cfg.indel_end_skip) {
#pragma unroll 1
            for (int clip = 0; clip < 2; ++clip) {   // nextStateAfterSplit, then nextStateAfterClipping (:462-472)
                const int32_t ns = gw_split_state(env, s, C, k, clip != 0);
                if (ns == -2) { gw_fail(c, GW_ST_OVERFLOW); return GW_STEP_DONE; }
                if (ns == -3) { gw_fail(c, GW_ST_ABORT); return GW_STEP_DONE; }
                if (ns >= 0) { if (!gw_queue_add(s, cfg, gw_update(s, base, c_off, ns))) { gw_fail(c, GW_ST_OVERFLOW); return GW_STEP_DONE; } }
                else if (clip == 0) c->n_filt++;
            }
        }
    }
    return GW_STEP_CONTINUE;
}

// Runs one read until it finishes or parks (sequential form of the two functions above).
GW_HD void gw_run_read(const GwSearchEnv& env, const GwSlab& s, uint32_t slot, const GwRoundQueues& rq) {
    int r = gw_resume_read(env, s, slot, rq);
    while (r == GW_STEP_CONTINUE) r = gw_step_read(env, s, slot, rq);
}
