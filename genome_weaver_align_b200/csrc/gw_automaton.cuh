// gw_automaton.cuh — resumable per-read driver around gw_search.cuh: AlignmentProcess.align_internal
// (reference: align/src/main/java/org/utgenome/weaver/align/strategy/BidirectionalSuffixFilter.java:278-477) cut at the
// points where the reference calls into the sampled suffix array and the verifier (verify :496-560). At those points the
// read parks with up to two fragment requests; the verification batch of the flow kernel fills them and the read resumes.
//
// Two kinds of code live here. The HOT slices (gw_lane_poll, gw_lane_try) run on the lane's record of the read
// (GwLane, gw_search.cuh) and are written to stay small and branch-poor: they are the loop every search step goes through.
// The COLD continuations (taking a read up, reporting a chain, split / soft-clip children) are out-of-line functions that work on
// the read's work area in HBM; the caller writes the lane back before and re-reads it after them.
#pragma once
#include "gw_search.cuh"

GW_HD void gw_fail(GwReadCtx* c, int status) { c->status = status; c->phase = GW_PH_DONE; }

#define GW_STEP_CONTINUE 0
#define GW_STEP_PARKED 1
#define GW_STEP_DONE 2

// reportAlignment request side: classify the fragments of the chain headed by `base`; returns true if the read must park
GW_COLD bool gw_begin_report(const GwSearchEnv& env, const GwSlab& s, int32_t base) {
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

// ---- taking a read up ------------------------------------------------------------------------------------
// Consumes what the read was parked for (exact-hit locate, verification results) or seeds the queue from the quick-scan results
// (:296-349); returns GW_STEP_CONTINUE when the read is in the best-first loop and wants search slices.
GW_COLD int gw_resume_read(const GwSearchEnv& env, const GwSlab& s) {
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

// the chain headed by `base` is complete: classify its fragments (park for locate + verification) or report it at once (:366-379)
GW_COLD int gw_search_report(const GwSearchEnv& env, const GwSlab& s, int32_t base) {
    GwReadCtx* c = s.ctx();
    if (gw_begin_report(env, s, base)) {
        c->phase = GW_PH_WAIT_VERIFY;
        return GW_STEP_PARKED;
    }
    gw_finish_report(env, s);
    return c->phase == GW_PH_DONE ? GW_STEP_DONE : GW_STEP_CONTINUE;
}

// split / soft-clip children of the state whose candidates are exhausted (:457-474)
GW_COLD int gw_search_split(const GwSearchEnv& env, const GwSlab& s, int32_t base, int32_t c_off) {
    GwReadCtx* c = s.ctx();
    const int m = c->m, k = c->k;
    const GwDevConfig& cfg = env.cfg;
    const GwState C = gw_load_state(s, c_off);
    const int nm = gw_lower_k(&C);
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

// a new state of a split chain enters the queue through the chain's head (SearchState.update :861-875); rare: out of line
template <int MAXH>
GW_COLD bool gw_lane_chain_insert(const GwSearchEnv& env, const GwSlab& s, GwLane<MAXH>& ln, int32_t ns) {
    gw_lane_flush(s, ln);
    const bool ok = gw_queue_add(s, env.cfg, gw_update(s, ln.base, ln.c_off, ns));
    gw_lane_reload(s, ln);
    return ok;
}

// ---- hot slices ------------------------------------------------------------------------------------------
// One iteration of the reference's `while (!queue.isEmpty())` loop (:352-475) is cut in two kinds of slices so that the lanes of
// a warp run the expensive part together:
//   gw_lane_poll   poll the queue, hit test / cut-offs (:352-409); leaves the lane "mid step" when there are candidates to try
//   gw_lane_try    ONE candidate extension = one FM step + one NFA step + one queue insertion (:415-455); the candidates of a
//                  step (the read's own base first, then A,C,G,T) are tried by consecutive calls; after the last one the
//                  split / soft-clip children are generated (:457-474)
// After a slice ln.mid tells what remains of the step: 0 nothing, GW_MID_TRY candidates, GW_MID_SPLIT / GW_MID_REPORT one of the
// out-of-line continuations (gw_lane_split, gw_lane_report), which the caller runs for several lanes of the warp at once.

GW_HD bool gw_state_complete(const GwState* st) { return gw_has_hit(st) || gw_is_clipped(st) || gw_remaining(st) == 0; }

// staircase table of the current minMismatches (:425, :439); nullptr (and the read fails) where the reference's constructor throws.
// The table only changes when minMismatches does (a reported hit), so its offset is kept in the lane.
template <int MAXH>
GW_HD const u64* gw_lane_staircase(const GwSearchEnv& env, const GwSlab& s, GwLane<MAXH>& ln) {
    const int mm = ln.min_mm;
    if (ln.sf_mm != mm) {
        const int m = ln.m;
        int off = -1;
        if (mm >= 0 && !((env.sf.abort_by_m[m] >> mm) & 1u)) off = env.sf.base_by_m[m] + ((m + 63) >> 6) * (mm * (mm + 1) / 2);
        ln.sf_mm = mm; ln.sf_off = off;
    }
    if (ln.sf_off < 0) { gw_fail(s.ctx(), GW_ST_ABORT); return nullptr; }
    return env.sf.words + ln.sf_off;
}

GW_HD void gw_prefetch(const void* p) {
#if defined(__CUDA_ARCH__)
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
#else
    (void)p;
#endif
}
// what the next poll of this read will touch is known as soon as a slice ends: the root of the queue names the state, and a sift-down
// starts in the first entries of the queue. Both lines are requested now, a whole trip of the warp ahead of their use.
template <int MAXH>
GW_HD void gw_lane_prefetch_next(const GwSlab& s, const GwLane<MAXH>& ln) {
    if (ln.qn > 0 && ln.mid == 0) {
        if (ln.root_off != ln.cached) gw_prefetch(s.at(ln.root_off));
        if (ln.qn > 1) gw_prefetch(s.queue() + 1);
    }
}
template <int MAXH>
GW_HD bool gw_lane_is_exclusive(const GwLane<MAXH>& ln) { return ln.c_off == ln.base && ln.C.next_split < 0; }

template <int MAXH>
GW_HD int gw_lane_poll(const GwSearchEnv& env, const GwSlab& s, GwLane<MAXH>& ln) {
    GwReadCtx* c = s.ctx();
    const GwDevConfig& cfg = env.cfg;
    ln.mid = 0;
    if (ln.qn <= 0 || ln.n_fm > ln.m * 20) { c->phase = GW_PH_DONE; return GW_STEP_DONE; }   // queue empty / fmIndexSearchUpperBound
    const int32_t base = gw_lq_poll(s, ln);
    if (base != ln.cached) gw_lane_fetch(s, ln, base);
    ln.base = base;
    int32_t c_off = base;
    if (gw_state_complete(&ln.C)) {   // first fragment of the chain that still has bases to process (:366-379)
        int32_t o = ln.C.next_split;
        for (;;) {
            if (o < 0) { ln.mid = GW_MID_REPORT; return GW_STEP_CONTINUE; }
            const GwState* st = gw_st(s, o);
            if (!gw_state_complete(st)) break;
            o = st->next_split;
        }
        c_off = o;
        gw_lane_fetch(s, ln, o);
    }
    ln.c_off = c_off;
    const GwState& C = ln.C;
    const bool exclusive = c_off == base && C.next_split < 0;
    if (gw_is_finished(&C)) return GW_STEP_CONTINUE;
    const int nm = gw_lower_k(&C);
    bool cut = nm > ln.min_mm;
    if (!cut) {
        const int scoreUB = exclusive ? (gw_frag_len(&C) - nm) * cfg.match_score - nm * cfg.mismatch_penalty : gw_score(s, cfg, gw_st(s, base), true);
        cut = scoreUB < 0 || scoreUB < ln.best_score;
    }
    if (cut) {
        ln.n_cut++;
        if (exclusive) gw_lane_free_state(ln, base);
        return GW_STEP_CONTINUE;
    }
    // the staircase filter of the current minMismatches is needed on every path from here (:425, :439); where the reference's
    // constructor throws, the read ends here whichever path asked for it
    if (!gw_lane_staircase(env, s, ln)) return GW_STEP_DONE;
    ln.mid = GW_MID_TRY;
    return GW_STEP_CONTINUE;
}

// ONE candidate extension of the state in the lane's record: candidate choice (:415-455), Cursor.nextSi +
// FMIndexOnGenome.bidirectionalSearch (two rank queries on one index, FMIndexOnGenome.java:148-194), ReadAlignmentNFA.nextState +
// SearchState.nextState (ReadAlignmentNFA.java:146-215), queue insertion. MAXH: compile-time bound of the NFA height (k + 1): for
// MAXH <= 8 the rows stay in registers (every loop over rows is unrolled with constant indices).
// EXCL: the caller knows that the state heads no split chain and is the polled queue entry itself (nine steps out of ten): the code
// for chains (flags written back to the arena, insertion through the chain's head) is then not even compiled in, and a warp that runs
// such states only never passes through it.
template <int MAXH, bool EXCL = false>
GW_HD int gw_lane_try(const GwSearchEnv& env, const GwSlab& s, GwLane<MAXH>& ln) {
    GwReadCtx* c = s.ctx();
    const GwDevConfig& cfg = env.cfg;
    const GwIndexView& ix = env.ix;
    GwState& C = ln.C;
    const bool exclusive = EXCL || (ln.c_off == ln.base && C.next_split < 0);
    const int strand = gw_strand(&C);
    const int dir = gw_dir(&C);
    const bool fwdC = dir != GW_DIR_B;
    const int nextIdx = fwdC ? C.cursor : C.cursor - 1;
    const int nextBase = gw_q_symbol(c, strand, nextIdx);
    int flags = C.state;
    // next candidate in the reference's order: the read's own base, then A,C,G,T; checked or empty ones are passed over (and
    // marked checked). Written without branches so that the lanes of a warp stay together up to the FM step.
    const bool kindF = C.si_kind == GW_SI_F || C.si_kind == GW_SI_BD, kindB = C.si_kind == GW_SI_B || C.si_kind == GW_SI_BD;
    int viable = 0;
#pragma unroll
    for (int cand = 0; cand < 4; ++cand) {
        const bool nonempty = C.si_kind == GW_SI_B ? (((C.b_valid >> cand) & 1) && C.sib[cand][0] < C.sib[cand][1])
                                                   : (kindF && ((C.f_valid >> cand) & 1) && C.sif[cand][0] < C.sif[cand][1]);
        if (!((flags >> cand) & 1) && nonempty) viable |= 1 << cand;
    }
    const bool own = ((viable >> nextBase) & 1) != 0;
    int ch;
    if (own) { ch = nextBase; flags |= 1 << nextBase; }
    else if (viable) {
        ch = 0;
#pragma unroll
        for (int cand = 3; cand >= 0; --cand) if ((viable >> cand) & 1) ch = cand;
        flags |= ((2 << ch) - 1) | (1 << nextBase);
    } else { ch = -1; flags |= 0x1F; }
    C.state = flags;
    if (!exclusive) gw_st(s, ln.c_off)->state = flags;   // a fragment of a split chain can be reached through several heads
    if (ch < 0) { ln.mid = GW_MID_SPLIT; return GW_STEP_CONTINUE; }   // candidates exhausted without extending with the read's own base
    const u64* sf_tbl = gw_lane_staircase(env, s, ln);   // cannot fail here: gw_lane_poll checked it for this minMismatches
    const int sf_kk = ln.min_mm;

    // ---- Cursor.nextSi + bidirectionalSearch
    bool hasF = kindF && ((C.f_valid >> ch) & 1), hasB = kindB && ((C.b_valid >> ch) & 1);
    if (dir == GW_DIR_BF && C.cursor >= C.end - 1) hasF = false;
    const uint32_t f_lb = C.sif[ch][0], f_ub = C.sif[ch][1], b_lb = C.sib[ch][0], b_ub = C.sib[ch][1];   // intervals of `ch`
    GwState n;   // the child
#pragma unroll
    for (int i = 0; i < 4; ++i) { n.sif[i][0] = n.sif[i][1] = 0; n.sib[i][0] = n.sib[i][1] = 0; }
    n.f_valid = n.b_valid = 0;
    n.si_kind = GW_SI_EMPTY;
    if (hasF || hasB) {
        const int fm = hasF ? (strand == 0 ? 1 : 0) : (strand == 0 ? 0 : 1);
        const uint32_t lb = gw_clampN(ix, hasF ? f_lb : b_lb), ub = gw_clampN(ix, hasF ? f_ub : b_ub);
        // both 64 B lines are requested before the first use, so that their misses overlap (the same line twice is one L1 miss)
        const GwBlk bl = gw_load_block(ix.occ[fm] + (lb >> 7));
        const GwBlk bu = gw_load_block(ix.occ[fm] + (ub >> 7));
        uint32_t lo[4], hi[4];
        gw_rank4(bl, lb & 127u, lo);
        gw_rank4(bu, ub & 127u, hi);
        uint8_t valid = 0;
        uint32_t x = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t l = ix.C[fm][i] + lo[i], u = ix.C[fm][i] + hi[i];
            if (hasF) { n.sif[i][0] = l; n.sif[i][1] = u; } else { n.sib[i][0] = l; n.sib[i][1] = u; }
            if (l < u) valid |= (uint8_t)(1 << i);
            if (hasF && hasB) {
                const uint32_t y = hi[i] - lo[i];
                n.sib[i][0] = b_lb + x; n.sib[i][1] = b_lb + x + y;
                x += y;
            }
        }
        if (hasF) { n.f_valid = valid; n.b_valid = hasB ? valid : 0; n.si_kind = hasB ? GW_SI_BD : GW_SI_F; }
        else { n.b_valid = valid; n.si_kind = GW_SI_B; }
    }
    ln.n_fm++;   // every bidirectionalSearch is two rank queries: occ_search is derived as 2 * n_fm when the read is emitted

    // ---- ReadAlignmentNFA.nextState
    const int m_read = ln.m;
    const int height = C.height, kOffset = C.k_offset;
    const int k = kOffset + height - 1;
    const u64 qeq = gw_pattern_mask64(c, strand, m_read, fwdC, nextIdx, C.pivot, C.cursor, ch, k - kOffset);
    const int fragmentLength = C.end - C.start;
    const int progressIndex = fwdC ? C.cursor - C.pivot : C.end - C.cursor;
    const int sf_offset = progressIndex - k;
    u64 prev[MAXH], next[MAXH];
    if constexpr (MAXH <= 8) {
#pragma unroll
        for (int i = 0; i < MAXH; ++i) prev[i] = ln.row(i);
    } else {
        for (int i = 0; i < height; ++i) prev[i] = ln.row(i);
    }
    int minKwithMatch = k + 1, minKwithProgress = k + 1;
    next[0] = (prev[0] & qeq) << 1;
    if (next[0] != 0) { minKwithMatch = 0; minKwithProgress = 0; }
    next[0] &= gw_staircase64(sf_tbl, m_read, sf_kk, kOffset, sf_offset);
    const int mPos = k + fragmentLength - progressIndex;
    int remove = -1;
    bool hasMatch = false;
    if constexpr (MAXH <= 8) {
#pragma unroll
        for (int i = 1; i < MAXH; ++i) {
            if (i < height) {
                u64 v = (prev[i] & qeq) << 1;
                if (minKwithMatch > k && v != 0) minKwithMatch = i;
                v |= prev[i - 1] | (prev[i - 1] << 1) | (next[i - 1] << 1);
                v &= gw_staircase64(sf_tbl, m_read, sf_kk, kOffset + i, sf_offset);
                if (minKwithProgress > k && (v & jshl(1ULL, height)) != 0) minKwithProgress = i;
                next[i] = v;
            } else next[i] = 0ULL;
        }
        if (mPos < 64) {
#pragma unroll
            for (int nm = MAXH - 1; nm >= 0; --nm)
                if (nm < height && (next[nm] & jshl(1ULL, mPos)) != 0) { remove = nm; hasMatch = true; }
        }
    } else {
#pragma unroll 1
        for (int i = 1; i < height; ++i) {
            u64 v = (prev[i] & qeq) << 1;
            if (minKwithMatch > k && v != 0) minKwithMatch = i;
            v |= prev[i - 1] | (prev[i - 1] << 1) | (next[i - 1] << 1);
            v &= gw_staircase64(sf_tbl, m_read, sf_kk, kOffset + i, sf_offset);
            if (minKwithProgress > k && (v & jshl(1ULL, height)) != 0) minKwithProgress = i;
            next[i] = v;
        }
        if (mPos < 64) {
            for (int nm = 0; nm < height; ++nm)
                if ((next[nm] & jshl(1ULL, mPos)) != 0) { remove = nm; hasMatch = true; break; }
        }
    }
    if (remove < 0) {
        const int minK = minKwithMatch < minKwithProgress ? minKwithMatch : minKwithProgress;
        if (minK < k) remove = minK;
        else { ln.n_filt++; return GW_STEP_CONTINUE; }   // no match: the candidate is dropped, the step goes on (ln.mid stays GW_MID_TRY)
    }
    const int newHeight = height - remove;
    // the child of an extension with the read's own base replaces its parent (:426-429): when nothing else can refer to the parent,
    // the child is built in the parent's slot
    const int32_t off = (own && exclusive) ? ln.c_off : gw_lane_alloc_state(s, ln);
    if (off < 0) { gw_fail(c, GW_ST_OVERFLOW); return GW_STEP_DONE; }
    // Cursor.next
    int nextCursor = C.cursor, d = dir;
    if (d == GW_DIR_F) ++nextCursor;
    else if (d == GW_DIR_B) --nextCursor;
    else if (nextCursor + 1 < C.end) ++nextCursor;
    else { d = GW_DIR_B; nextCursor = C.pivot; }
    n.start = C.start; n.end = C.end; n.cursor = (int16_t)nextCursor; n.pivot = C.pivot;
    n.cflag = (uint8_t)(strand | (d << 1));
    // currentSi = nextCursor.isForwardSearch ? siTable.getForward(ch) : siTable.getBackward(ch)
    const bool fwd = d != GW_DIR_B;
    const bool has = fwd ? (kindF && ((C.f_valid >> ch) & 1)) : (kindB && ((C.b_valid >> ch) & 1));
    n.has_cur = has ? 1 : 0;
    n.cur_lb = has ? (fwd ? f_lb : b_lb) : 0;
    n.cur_ub = has ? (fwd ? f_ub : b_ub) : 0;
    n.height = (uint8_t)newHeight; n.k_offset = (uint8_t)(kOffset + remove); n.has_nfa = 1;
    n.state = gw_pack_state(ch, kOffset + remove, gw_priority(&C), hasMatch);
    n.next_split = C.next_split;
    u64 nrows[MAXH];
    if constexpr (MAXH <= 8) {
        // remove == 0 keeps the OLD automaton (sic, ReadAlignmentNFA.java:206-207)
#pragma unroll
        for (int j = 0; j < MAXH; ++j) {
            u64 v = 0ULL;
#pragma unroll
            for (int r = 0; r < MAXH - j; ++r)
                if (r == remove) v = remove == 0 ? prev[j] : (next[j + r] >> 1);
            nrows[j] = j < newHeight ? v : 0ULL;
        }
    } else {
        if (remove == 0) { for (int h = 0; h < newHeight; ++h) nrows[h] = prev[h]; }
        else { for (int h = 0; h < newHeight; ++h) nrows[h] = next[h + remove] >> 1; }
    }
    {   // written through to the arena
        GwState* np = gw_st(s, off);
        gw_state_put(np, n);
        u64* nr = gw_rows(np);
        if constexpr (MAXH <= 8) {
#pragma unroll
            for (int j = 0; j < MAXH; ++j) if (j < newHeight) nr[j] = nrows[j];
        } else {
            for (int h = 0; h < newHeight; ++h) nr[h] = nrows[h];
        }
    }
    // the usual case — a chain of one state — takes its queue key from the state just built, still in registers
    const bool ok = exclusive ? gw_lq_add(s, ln, off, gw_state_key_single(cfg, &n)) : gw_lane_chain_insert(env, s, ln, off);
    if (!ok) { gw_fail(c, GW_ST_OVERFLOW); return GW_STEP_DONE; }
    if (own) {   // extended: the parent is dropped (:426-429), the step is over; the child is most likely the next state polled
        ln.C = n;
        if constexpr (MAXH <= 8) {
#pragma unroll
            for (int j = 0; j < MAXH; ++j) ln.set_row(j, nrows[j]);
        } else {
            for (int j = 0; j < newHeight; ++j) ln.set_row(j, nrows[j]);
        }
        ln.cached = off;
        ln.mid = 0;
    }
    return GW_STEP_CONTINUE;
}

// the out-of-line continuations of a step, on the work area (the lane is written back before and re-read after)
template <int MAXH>
GW_HD int gw_lane_split(const GwSearchEnv& env, const GwSlab& s, GwLane<MAXH>& ln) {
    const bool exclusive = ln.c_off == ln.base && ln.C.next_split < 0;
    ln.mid = 0;
    gw_lane_flush(s, ln);
    const int r = gw_search_split(env, s, ln.base, ln.c_off);
    gw_lane_reload(s, ln);
    ln.cached = -1;
    if (exclusive) gw_lane_free_state(ln, ln.base);   // all its candidates are tried, its children are copies: unreachable from here on
    return r;
}
template <int MAXH>
GW_HD int gw_lane_report(const GwSearchEnv& env, const GwSlab& s, GwLane<MAXH>& ln) {
    ln.mid = 0;
    gw_lane_flush(s, ln);
    const int r = gw_search_report(env, s, ln.base);
    gw_lane_reload(s, ln);
    ln.cached = -1;
    return r;
}
