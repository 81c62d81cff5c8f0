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
GW_COLD bool gw_begin_report(const GwSearchEnv& env, const GwSlab& s, uint32_t slot, int32_t base,
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
GW_COLD void gw_finish_report(const GwSearchEnv& env, const GwSlab& s) {
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

// staircase table of the current minMismatches; returns nullptr and fails the read where the reference constructor throws.
// The table only changes when minMismatches does (a reported hit), so its offset is kept in the context: the usual call reads
// three words of the context header instead of walking m -> abort bitmap -> table base.
GW_HD const u64* gw_staircase_for(const GwSearchEnv& env, GwReadCtx* c, int* kk) {
    const int mm = c->min_mm;
    *kk = mm;
    if (c->sf_mm != mm) {
        const int m = c->m;
        int off = -1;
        if (mm >= 0 && !((env.sf.abort_by_m[m] >> mm) & 1u)) off = env.sf.base_by_m[m] + ((m + 63) >> 6) * (mm * (mm + 1) / 2);
        c->sf_mm = mm; c->sf_off = off;
    }
    const int off = c->sf_off;
    if (off < 0) { gw_fail(c, GW_ST_ABORT); return nullptr; }
    return env.sf.words + off;
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
GW_COLD int gw_resume_read(const GwSearchEnv& env, const GwSlab& s, uint32_t slot, const GwRoundQueues& rq) {
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

// One iteration of the best-first loop is cut in two kinds of slices so that the lanes of a warp run the expensive part together:
//   gw_search_poll   poll the queue, hit test / cut-offs (:352-409); leaves the lane "mid step" when there are candidates to try
//   gw_search_try    ONE candidate extension = one FM step + one NFA step + one queue insertion (:415-455); the candidates of a
//                    step (the read's own base first, then A,C,G,T) are tried by consecutive calls; after the last one the
//                    split / soft-clip children are generated (:457-474)
// Which candidates are done is recorded in the checked-base flags of the state itself (SearchState.state bits 0-4), exactly where
// the reference keeps them, so a lane only carries the two arena offsets below between slices.
struct GwSearchLane {
    int32_t base;    // polled queue entry (head of the split chain)
    int32_t c_off;   // fragment of the chain that is being extended
    int32_t mid;     // 1: candidates of the current step remain
};

// the chain headed by `base` is complete: classify its fragments (park for locate + verification) or report it at once (:366-379)
GW_COLD int gw_search_report(const GwSearchEnv& env, const GwSlab& s, uint32_t slot, const GwRoundQueues& rq, int32_t base) {
    GwReadCtx* c = s.ctx();
    if (gw_begin_report(env, s, slot, base, rq)) {
        c->phase = GW_PH_WAIT_VERIFY;
        if (rq.counters) rq.next_active[gw_atomic_add(&rq.counters[0], 1u)] = slot;
        return GW_STEP_PARKED;
    }
    gw_finish_report(env, s);
    return c->phase == GW_PH_DONE ? GW_STEP_DONE : GW_STEP_CONTINUE;
}

// L.mid after a slice: 0 step finished, GW_MID_TRY candidates remain; with DEFER (warp-pool kernel) the rarely taken, long
// continuations are not run inside the slice but left as GW_MID_SPLIT / GW_MID_REPORT for the caller, which runs them for
// several lanes of the warp at once (gw_search_split, gw_search_report)
#define GW_MID_TRY 1
#define GW_MID_SPLIT 2
#define GW_MID_REPORT 3

template <bool DEFER = false>
GW_HD int gw_search_poll(const GwSearchEnv& env, const GwSlab& s, uint32_t slot, const GwRoundQueues& rq, GwSearchLane& L) {
    GwReadCtx* c = s.ctx();
#if defined(__CUDA_ARCH__)
    // a step touches, one after the other, the context header, the head of the queue, the polled state and the read's bit planes:
    // the lines whose addresses are known now are requested together, so that their misses overlap
    asm volatile("prefetch.global.L1 [%0];" ::"l"(s.queue()));
    asm volatile("prefetch.global.L1 [%0];" ::"l"(&c->qlo[0][0]));
#endif
    const int m = c->m;
    const GwDevConfig& cfg = env.cfg;
    L.mid = 0;
    if (c->qn <= 0 || c->n_fm > m * 20) { c->phase = GW_PH_DONE; return GW_STEP_DONE; }   // queue empty / fmIndexSearchUpperBound
    const int32_t base = gw_queue_poll(s, cfg);
    int32_t c_off = base;
    const GwState* C = gw_st(s, base);
    for (;;) {   // first fragment of the chain that still has bases to process (:366-379)
        if (!(gw_has_hit(C) || gw_is_clipped(C) || gw_remaining(C) == 0)) break;
        if (C->next_split < 0) {
            if (DEFER) { L.base = base; L.mid = GW_MID_REPORT; return GW_STEP_CONTINUE; }
            return gw_search_report(env, s, slot, rq, base);
        }
        c_off = C->next_split;
        C = gw_st(s, c_off);
    }
    if (gw_is_finished(C)) return GW_STEP_CONTINUE;
    const int nm = gw_lower_k(C);
    if (nm > c->min_mm) { c->n_cut++; return GW_STEP_CONTINUE; }
    {
        const int scoreUB = gw_score(s, cfg, gw_st(s, base), true);
        if (scoreUB < 0 || scoreUB < c->best_score) { c->n_cut++; return GW_STEP_CONTINUE; }
    }
    // the staircase filter of the current minMismatches is needed on every path from here (:425, :439); where the reference's
    // constructor throws, the read ends here whichever path asked for it
    int kk;
    if (!gw_staircase_for(env, c, &kk)) return GW_STEP_DONE;
    L.base = base; L.c_off = c_off; L.mid = 1;
    return GW_STEP_CONTINUE;
}

// split / soft-clip children of the state whose candidates are exhausted (:457-474)
GW_COLD int gw_search_split(const GwSearchEnv& env, const GwSlab& s, const GwSearchLane& L) {
    GwReadCtx* c = s.ctx();
    const int m = c->m, k = c->k;
    const GwDevConfig& cfg = env.cfg;
    const GwState C = gw_load_state(s, L.c_off);
    const int nm = gw_lower_k(&C);
    if (gw_num_split(s, gw_st(s, L.base)) < cfg.num_split_allowed && nm + 1 <= c->min_mm) {
        const int index = gw_next_index(&C);
        if (index > cfg.indel_end_skip && m - index >= cfg.indel_end_skip) {
#pragma unroll 1
            for (int clip = 0; clip < 2; ++clip) {   // nextStateAfterSplit, then nextStateAfterClipping (:462-472)
                const int32_t ns = gw_split_state(env, s, C, k, clip != 0);
                if (ns == -2) { gw_fail(c, GW_ST_OVERFLOW); return GW_STEP_DONE; }
                if (ns == -3) { gw_fail(c, GW_ST_ABORT); return GW_STEP_DONE; }
                if (ns >= 0) { if (!gw_queue_add(s, cfg, gw_update(s, L.base, L.c_off, ns))) { gw_fail(c, GW_ST_OVERFLOW); return GW_STEP_DONE; } }
                else if (clip == 0) c->n_filt++;
            }
        }
    }
    return GW_STEP_CONTINUE;
}

template <bool DEFER = false, int MAXH = GW_MAX_ROWS>
GW_HD int gw_search_try(const GwSearchEnv& env, const GwSlab& s, GwSearchLane& L) {
    GwReadCtx* c = s.ctx();
    const GwDevConfig& cfg = env.cfg;
    GwState* Cp = gw_st(s, L.c_off);
    const GwState C = *Cp;
    const int strand = gw_strand(&C);
    const int nextBase = gw_q_symbol(c, strand, gw_next_index(&C));
    int flags = C.state;
    // next candidate in the reference's order: the read's own base, then A,C,G,T; checked or empty ones are passed over (and
    // marked checked) (:415-455). Written without branches so that the lanes of a warp stay together up to the FM step.
    int viable = 0;
#pragma unroll
    for (int cand = 0; cand < 4; ++cand)
        if (!((flags >> cand) & 1) && !gw_si_is_empty(&C, cand)) viable |= 1 << cand;
    const bool own = ((viable >> nextBase) & 1) != 0;
    int ch;
    if (own) { ch = nextBase; flags |= 1 << nextBase; }
    else if (viable) {
        ch = 0;
#pragma unroll
        for (int cand = 3; cand >= 0; --cand) if ((viable >> cand) & 1) ch = cand;
        flags |= ((2 << ch) - 1) | (1 << nextBase);
    } else { ch = -1; flags |= 0xF; }
    if (ch < 0) {   // candidates exhausted without extending with the read's own base
        flags |= 1 << 4;
        Cp->state = flags;
        if (DEFER) { L.mid = GW_MID_SPLIT; return GW_STEP_CONTINUE; }
        L.mid = 0;
        return gw_search_split(env, s, L);
    }
#if defined(__CUDA_ARCH__)
    {   // the new state will be sifted up from the end of the queue: request the lines of its parent and grandparent
        const int qn = c->qn;
        const GwQEntry* q = s.queue();
        asm volatile("prefetch.global.L1 [%0];" ::"l"(q + ((qn - 1) >> 1)));
        asm volatile("prefetch.global.L1 [%0];" ::"l"(q + ((((qn - 1) >> 1) - 1) >> 1)));
    }
#endif
    int kk;
    const u64* tbl = gw_staircase_for(env, c, &kk);
    u64 prev[MAXH];
    {
        const u64* rows = gw_rows(Cp);
        if constexpr (MAXH <= 8) {
#pragma unroll
            for (int i = 0; i < MAXH; ++i) prev[i] = i < C.height ? rows[i] : 0ULL;
        } else {
            for (int i = 0; i < C.height; ++i) prev[i] = rows[i];
        }
    }
    GwSiSet nsi;
    uint32_t lookups = 0;
    gw_next_si(env.ix, &C, ch, &nsi, &lookups);
    c->n_fm++;   // every bidirectionalSearch is two rank queries: occ_search is derived as 2 * n_fm when the read is emitted
    GwState built;
    const int32_t ns = gw_next_state<MAXH>(s, C, prev, ch, nsi, tbl, kk, &built);
    if (ns == -2) { Cp->state = flags; gw_fail(c, GW_ST_OVERFLOW); return GW_STEP_DONE; }
    if (ns >= 0) {
        // the usual case — a chain of one state — takes its queue key from the register copy of the state just built
        const bool ok = (L.c_off == L.base && built.next_split < 0) ? gw_queue_add_key(s, ns, gw_state_key_single(cfg, &built))
                                                                    : gw_queue_add(s, cfg, gw_update(s, L.base, L.c_off, ns));
        if (!ok) { Cp->state = flags; gw_fail(c, GW_ST_OVERFLOW); return GW_STEP_DONE; }
        if (own) {   // extended: the parent is dropped (:426-429)
            Cp->state = flags; L.mid = 0;
#if defined(__CUDA_ARCH__)
            // the step is over: the head of the queue is polled next; request its state now
            asm volatile("prefetch.global.L1 [%0];" ::"l"(s.at(s.queue()[0].off)));
#endif
            return GW_STEP_CONTINUE;
        }
    } else
        c->n_filt++;
    Cp->state = flags;
    return GW_STEP_CONTINUE;
}

// one whole iteration of the best-first loop (sequential form: the round-based search kernel and the host simulator)
GW_HD int gw_step_read(const GwSearchEnv& env, const GwSlab& s, uint32_t slot, const GwRoundQueues& rq) {
    GwSearchLane L;
    int r = gw_search_poll(env, s, slot, rq, L);
#pragma unroll 1
    while (r == GW_STEP_CONTINUE && L.mid) r = gw_search_try(env, s, L);
    return r;
}

// Runs one read until it finishes or parks (sequential form of the two functions above).
GW_HD void gw_run_read(const GwSearchEnv& env, const GwSlab& s, uint32_t slot, const GwRoundQueues& rq) {
    int r = gw_resume_read(env, s, slot, rq);
    while (r == GW_STEP_CONTINUE) r = gw_step_read(env, s, slot, rq);
}
