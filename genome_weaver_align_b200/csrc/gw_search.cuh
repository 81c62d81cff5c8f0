// gw_search.cuh — the per-read BSF automaton: best-first bidirectional FM search guided by the k-difference NFA
// and the staircase (suffix) filter, with split / soft-clip states. One thread drives one read; the automaton is
// strictly sequential per read because hit discovery order decides the reported alignment.
// Replaces BidirectionalSuffixFilter.AlignmentProcess (reference:
// align/src/main/java/org/utgenome/weaver/align/strategy/BidirectionalSuffixFilter.java):
//   align_internal :278-477, next :486-488, reportExactMatchAlignment :490-494, verify :496-560 (request side),
//   reportAlignment :562-586, AlignmentResultHolder.add :606-628, SearchState :658-877, StateQueue :133-165,
// plus FMIndexOnGenome.bidirectionalSearch (align/FMIndexOnGenome.java:148-194) and ReadAlignmentNFA.nextState
// (align/strategy/ReadAlignmentNFA.java:146-215).
// Verification itself (locate + bit-vector DP) runs in separate kernels: the automaton parks with a request and
// resumes in the next round with the results (gw_kernels.cu).
#pragma once
#include "gw_state.cuh"

struct GwSearchEnv {
    GwIndexView ix;
    GwDevConfig cfg;
    GwStaircaseView sf;
};


GW_HD GwState* gw_st(const GwSlab& s, int32_t off) { return reinterpret_cast<GwState*>(s.at(off)); }

// SearchState.getNumSplit / score / upperBoundOfScore
GW_HD int gw_num_split(const GwSlab& s, const GwState* st) {
    int n = 0;
    for (int32_t o = st->next_split; o >= 0; o = gw_st(s, o)->next_split) ++n;
    return n;
}
GW_HD int gw_score(const GwSlab& s, const GwDevConfig& cfg, const GwState* st, bool upper) {
    int total = 0;
    int nsplit = gw_num_split(s, st);
    for (const GwState* p = st;;) {
        int nm = gw_lower_k(p) - nsplit;
        int mm = gw_processed(p) + (upper ? gw_remaining(p) : 0) - nm;
        total += mm * cfg.match_score - nm * cfg.mismatch_penalty - nsplit * cfg.split_open_penalty;
        if (p->next_split < 0) break;
        p = gw_st(s, p->next_split);
        --nsplit;
    }
    return total;
}
// StateQueue comparator (:133-165): priority ascending, then score descending, then processed bases descending, packed into
// one integer so that "compare(a, b) < 0" is "key(a) < key(b)". A state's key cannot change while it is queued: cursor and
// lower bound of k are fixed at construction, the lower bound is only rewritten on a polled head (reportAlignment :582) and a
// head's split chain is only relinked (SearchState.update) between its poll and its re-insertion. So the key is computed once,
// at insertion, and the heap never has to touch the states again.
GW_HD u64 gw_state_key(const GwSlab& s, const GwDevConfig& cfg, int32_t off) {
    const GwState* o = gw_st(s, off);
    const int score = gw_score(s, cfg, o, false);
    return ((u64)(uint32_t)gw_priority(o) << 44) | ((u64)(uint32_t)(0x8000000 - score) << 16) | (u64)(uint32_t)(0xFFFF - gw_processed(o));
}
// the same key for a state without split chain, from a register copy (a state that was just built need not be read back)
GW_HD u64 gw_state_key_single(const GwDevConfig& cfg, const GwState* o) {
    const int nm = gw_lower_k(o);
    const int score = (gw_processed(o) - nm) * cfg.match_score - nm * cfg.mismatch_penalty;
    return ((u64)(uint32_t)gw_priority(o) << 44) | ((u64)(uint32_t)(0x8000000 - score) << 16) | (u64)(uint32_t)(0xFFFF - gw_processed(o));
}
// java.util.PriorityQueue.offer (siftUp) with the entry's key given; returns false on capacity overflow
GW_HD bool gw_queue_add_key(const GwSlab& s, int32_t e, const u64 ke) {
    GwReadCtx* c = s.ctx();
    GwQEntry* q = s.queue();
    int k = c->qn;
    if ((uint32_t)k >= s.sp.queue_cap) return false;
    c->qn = k + 1;
    while (k > 0) {
        const int parent = (k - 1) >> 1;
        const GwQEntry p = q[parent];
        if (ke >= p.key) break;
        q[k] = p;
        k = parent;
    }
    GwQEntry n;
    n.key = ke; n.off = e; n.pad_ = 0;
    q[k] = n;
    return true;
}
// java.util.PriorityQueue.offer (siftUp) ; returns false on capacity overflow
GW_HD bool gw_queue_add(const GwSlab& s, const GwDevConfig& cfg, int32_t e) {
    if (e < 0) return true;  // StateQueue.add(null) is a no-op
    return gw_queue_add_key(s, e, gw_state_key(s, cfg, e));
}
// java.util.PriorityQueue.poll (siftDown)
GW_HD int32_t gw_queue_poll(const GwSlab& s, const GwDevConfig& cfg) {
    (void)cfg;
    GwReadCtx* c = s.ctx();
    GwQEntry* q = s.queue();
    const int n = --c->qn;
    const int32_t result = q[0].off;
#if defined(__CUDA_ARCH__)
    // the caller reads this state next: request its line now, the sift-down below hides the latency
    asm volatile("prefetch.global.L1 [%0];" ::"l"(s.at(result)));
#endif
    if (n != 0) {
        const GwQEntry x = q[n];
        int k = 0;
        const int half = n >> 1;
        while (k < half) {
            int child = (k << 1) + 1;
            const int right = child + 1;
            GwQEntry cc = q[child];
            if (right < n) { const GwQEntry cr = q[right]; if (cc.key > cr.key) { cc = cr; child = right; } }
            if (x.key <= cc.key) break;
            q[k] = cc;
            k = child;
        }
        q[k] = x;
    }
    return result;
}

// allocate a state with `height` NFA rows; fields are filled by the caller
GW_HD int32_t gw_alloc_state(const GwSlab& s, int height) {
    return gw_alloc(s, GW_STATE_WORDS + (uint32_t)height);
}
// whole-state copy from / to the arena (six 128-bit accesses; states are 16-byte aligned)
GW_HD GwState gw_load_state(const GwSlab& s, int32_t off) { return *gw_st(s, off); }

GW_HD void gw_set_init_siset(const GwIndexView& ix, GwState* st, int dir) {
    st->f_valid = st->b_valid = 0;
    for (int i = 0; i < 4; ++i) { st->sif[i][0] = st->sif[i][1] = 0; st->sib[i][0] = st->sib[i][1] = 0; }
    if (dir == GW_DIR_F || dir == GW_DIR_BF) {
        st->f_valid = (uint8_t)ix.init_valid;
        for (int i = 0; i < 4; ++i) { st->sif[i][0] = ix.init_lb[i]; st->sif[i][1] = ix.init_ub[i]; }
    }
    if (dir == GW_DIR_B || dir == GW_DIR_BF) {
        st->b_valid = (uint8_t)ix.init_valid;
        for (int i = 0; i < 4; ++i) { st->sib[i][0] = ix.init_lb[i]; st->sib[i][1] = ix.init_ub[i]; }
    }
    st->si_kind = dir == GW_DIR_F ? GW_SI_F : dir == GW_DIR_B ? GW_SI_B : GW_SI_BD;
}

// SearchState(k, null, cursor, priority): initial state with diagonal NFA (:753-757)
GW_HD int32_t gw_initial_state(const GwSearchEnv& env, const GwSlab& s, int strand, int dir, int start, int end,
                               int cursor, int pivot, int k, int minK, int priority) {
    int32_t off = gw_alloc_state(s, k + 1);
    if (off < 0) return -1;
    GwState st;
    st.start = (int16_t)start; st.end = (int16_t)end; st.cursor = (int16_t)cursor; st.pivot = (int16_t)pivot;
    st.cflag = (uint8_t)(strand | (dir << 1));
    gw_set_init_siset(env.ix, &st, dir);
    st.height = (uint8_t)(k + 1); st.k_offset = 0; st.has_cur = 0; st.has_nfa = 1;
    st.state = gw_pack_state(4, minK, priority, false);
    st.next_split = -1;
    st.cur_lb = st.cur_ub = 0;
    *gw_st(s, off) = st;
    u64* rows = gw_rows(gw_st(s, off));
    for (int i = 0; i <= k; ++i) rows[i] = jshl(1ULL, k + i);
    return off;
}

GW_HD bool gw_si_is_empty(const GwState* c, int ch) {
    if (c->si_kind == GW_SI_EMPTY || c->si_kind == GW_SI_NULL) return true;
    if (c->si_kind == GW_SI_B) return !((c->b_valid >> ch) & 1) || c->sib[ch][0] >= c->sib[ch][1];
    return !((c->f_valid >> ch) & 1) || c->sif[ch][0] >= c->sif[ch][1];
}

struct GwSiSet {
    uint8_t kind, f_valid, b_valid;
    uint32_t sif[4][2], sib[4][2];
};

// Cursor.nextSi + FMIndexOnGenome.bidirectionalSearch: two rank queries on one index
GW_HD void gw_next_si(const GwIndexView& ix, const GwState* c, int ch, GwSiSet* out, uint32_t* occ_lookups) {
    bool hasF = (c->si_kind == GW_SI_F || c->si_kind == GW_SI_BD) && ((c->f_valid >> ch) & 1);
    bool hasB = (c->si_kind == GW_SI_B || c->si_kind == GW_SI_BD) && ((c->b_valid >> ch) & 1);
    if (gw_dir(c) == GW_DIR_BF && c->cursor >= c->end - 1) hasF = false;
    const int strand = gw_strand(c);
    out->f_valid = out->b_valid = 0;
    for (int i = 0; i < 4; ++i) { out->sif[i][0] = out->sif[i][1] = 0; out->sib[i][0] = out->sib[i][1] = 0; }
    if (!hasF && !hasB) { out->kind = GW_SI_EMPTY; return; }
    const int fm = hasF ? (strand == 0 ? 1 : 0) : (strand == 0 ? 0 : 1);
    uint32_t f_lb = 0, f_ub = 0, b_lb = 0, b_ub = 0;   // intervals of `ch`, selected without indexing by a run-time value
#pragma unroll
    for (int i = 0; i < 4; ++i)
        if (i == ch) { f_lb = c->sif[i][0]; f_ub = c->sif[i][1]; b_lb = c->sib[i][0]; b_ub = c->sib[i][1]; }
    const uint32_t lb = gw_clampN(ix, hasF ? f_lb : b_lb);
    const uint32_t ub = gw_clampN(ix, hasF ? f_ub : b_ub);
    uint32_t lo[4], hi[4];
    // both 64 B lines are requested before the first use, so that their misses overlap (the same line twice is one L1 miss)
    const GwBlk bl = gw_load_block(ix.occ[fm] + (lb >> 7));
    const GwBlk bu = gw_load_block(ix.occ[fm] + (ub >> 7));
    gw_rank4(bl, lb & 127u, lo);
    gw_rank4(bu, ub & 127u, hi);
    *occ_lookups += 2;
    uint8_t valid = 0;
    uint32_t x = 0;
    const uint32_t blb = b_lb;
    for (int i = 0; i < 4; ++i) {
        const uint32_t l = ix.C[fm][i] + lo[i], u = ix.C[fm][i] + hi[i];
        if (hasF) { out->sif[i][0] = l; out->sif[i][1] = u; } else { out->sib[i][0] = l; out->sib[i][1] = u; }
        if (l < u) valid |= (uint8_t)(1 << i);
        if (hasF && hasB) {
            const uint32_t y = hi[i] - lo[i];
            out->sib[i][0] = blb + x; out->sib[i][1] = blb + x + y;
            x += y;
        }
    }
    if (hasF) {
        out->f_valid = valid;
        if (hasB) { out->b_valid = valid; out->kind = GW_SI_BD; } else out->kind = GW_SI_F;
    } else {
        out->b_valid = valid;
        out->kind = GW_SI_B;
    }
}

// ReadAlignmentNFA.nextState + SearchState.nextState on a register copy `c` of the current state (rows in `prev`):
// returns the arena offset of the new state, -1 for "no match" (null), -2 on slab overflow
// MAXH: compile-time bound of the NFA height (k + 1). For MAXH <= 8 the rows are kept in registers (every loop over rows is
// unrolled with constant indices); the generic form (MAXH == GW_MAX_ROWS) keeps them in local memory.
template <int MAXH = GW_MAX_ROWS>
GW_HD int32_t gw_next_state(const GwSlab& s, const GwState& c, const u64* prev, int ch, const GwSiSet& nextSi,
                            const u64* sf_tbl, int sf_kk, GwState* built) {
    GwReadCtx* ctx = s.ctx();
    const int m_read = ctx->m;
    const int height = c.height;
    const int kOffset = c.k_offset;
    const int k = kOffset + height - 1;
    const int kr = k - kOffset;
    const u64 qeq = gw_pattern_mask64(ctx, gw_strand(&c), m_read, gw_is_fwd(&c), gw_next_index(&c), c.pivot, c.cursor, ch, kr);
    const int progressIndex = gw_processed(&c);
    const int fragmentLength = gw_frag_len(&c);
    u64 next[MAXH];
    int minKwithMatch = k + 1, minKwithProgress = k + 1;
    next[0] = (prev[0] & qeq) << 1;
    if (next[0] != 0) { minKwithMatch = 0; minKwithProgress = 0; }
    next[0] &= gw_staircase64(sf_tbl, m_read, sf_kk, kOffset, progressIndex - k);
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
                v &= gw_staircase64(sf_tbl, m_read, sf_kk, kOffset + i, progressIndex - k);
                if (minKwithProgress > k && (v & jshl(1ULL, height)) != 0) minKwithProgress = i;
                next[i] = v;
            } else next[i] = 0ULL;
        }
        if (mPos < 64) {
#pragma unroll
            for (int nm = 0; nm < MAXH; ++nm)
                if (nm < height && remove < 0 && (next[nm] & jshl(1ULL, mPos)) != 0) { remove = nm; hasMatch = true; }
        }
    } else {
#pragma unroll 1
        for (int i = 1; i < height; ++i) {
            u64 v = (prev[i] & qeq) << 1;
            if (minKwithMatch > k && v != 0) minKwithMatch = i;
            v |= prev[i - 1] | (prev[i - 1] << 1) | (next[i - 1] << 1);
            v &= gw_staircase64(sf_tbl, m_read, sf_kk, kOffset + i, progressIndex - k);
            if (minKwithProgress > k && (v & jshl(1ULL, height)) != 0) minKwithProgress = i;
            next[i] = v;
        }
        if (mPos < 64) {
            for (int nm = 0; nm < height; ++nm)
                if ((next[nm] & jshl(1ULL, mPos)) != 0) { remove = nm; hasMatch = true; break; }
        }
    }
    if (remove < 0) {
        int minK = minKwithMatch < minKwithProgress ? minKwithMatch : minKwithProgress;
        if (minK < k) remove = minK;
        else return -1;
    }
    const int newHeight = height - remove;
    int32_t off = gw_alloc_state(s, newHeight);
    if (off < 0) return -2;
    GwState n;
    // Cursor.next
    int nextCursor = c.cursor, d = gw_dir(&c);
    if (d == GW_DIR_F) ++nextCursor;
    else if (d == GW_DIR_B) --nextCursor;
    else if (nextCursor + 1 < c.end) ++nextCursor;
    else { d = GW_DIR_B; nextCursor = c.pivot; }
    n.start = c.start; n.end = c.end; n.cursor = (int16_t)nextCursor; n.pivot = c.pivot;
    n.cflag = (uint8_t)(gw_strand(&c) | (d << 1));
    // currentSi = nextCursor.isForwardSearch ? siTable.getForward(ch) : siTable.getBackward(ch)
    const bool fwd = d != GW_DIR_B;
    const bool has = fwd ? ((c.si_kind == GW_SI_F || c.si_kind == GW_SI_BD) && ((c.f_valid >> ch) & 1))
                         : ((c.si_kind == GW_SI_B || c.si_kind == GW_SI_BD) && ((c.b_valid >> ch) & 1));
    uint32_t clb = 0, cub = 0;   // selected without indexing the interval tables by a run-time value
#pragma unroll
    for (int i = 0; i < 4; ++i)
        if (i == ch) { clb = fwd ? c.sif[i][0] : c.sib[i][0]; cub = fwd ? c.sif[i][1] : c.sib[i][1]; }
    n.has_cur = has ? 1 : 0;
    n.cur_lb = has ? clb : 0;
    n.cur_ub = has ? cub : 0;
    n.si_kind = nextSi.kind; n.f_valid = nextSi.f_valid; n.b_valid = nextSi.b_valid;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        n.sif[i][0] = nextSi.sif[i][0]; n.sif[i][1] = nextSi.sif[i][1];
        n.sib[i][0] = nextSi.sib[i][0]; n.sib[i][1] = nextSi.sib[i][1];
    }
    n.height = (uint8_t)newHeight; n.k_offset = (uint8_t)(kOffset + remove); n.has_nfa = 1;
    n.state = gw_pack_state(ch, kOffset + remove, gw_priority(&c), hasMatch);
    n.next_split = c.next_split;
    *gw_st(s, off) = n;
    *built = n;
    u64* nr = gw_rows(gw_st(s, off));
    if constexpr (MAXH <= 8) {
        // remove == 0 keeps the OLD automaton (sic, ReadAlignmentNFA.java:206-207)
#pragma unroll
        for (int j = 0; j < MAXH; ++j)
            if (j >= remove && j < height) nr[j - remove] = remove == 0 ? prev[j] : (next[j] >> 1);
    } else {
        if (remove == 0) { for (int h = 0; h < newHeight; ++h) nr[h] = prev[h]; }  // sic: the OLD automaton (ReadAlignmentNFA.java:206-207)
        else { for (int h = 0; h < newHeight; ++h) nr[h] = next[h + remove] >> 1; }
    }
    return off;
}

// SearchState.nextStateAfterSplit (clip == false) / nextStateAfterClipping (clip == true) on a register copy of the current
// state; -1 null, -2 overflow, -3 abort
GW_COLD int32_t gw_split_state(const GwSearchEnv& env, const GwSlab& s, const GwState& c, int k, bool clip) {
    const int minK = gw_lower_k(&c);
    if (!(minK < k)) return -1;
    const int newHeight = (int)c.height - 1;
    if (!c.has_nfa || newHeight < 0) return -3;
    // Cursor.split
    const int d = gw_dir(&c), strand = gw_strand(&c);
    int ls, le, lc, lp, ld, rs, re, rc, rp, rd;
    if (d == GW_DIR_F) { ls = c.start; le = c.cursor; lc = c.cursor; lp = c.pivot; ld = d; rs = c.cursor; re = c.end; rc = c.cursor; rp = c.cursor; rd = GW_DIR_F; }
    else if (d == GW_DIR_B) { ls = c.cursor; le = c.end; lc = c.cursor; lp = c.pivot; ld = d; rs = c.start; re = c.cursor; rc = c.cursor; rp = c.start; rd = GW_DIR_B; }
    else {
        ls = c.start; le = c.cursor; lc = c.cursor; lp = c.pivot; ld = (c.cursor + 1 < c.end) ? d : GW_DIR_B;
        rs = c.cursor; re = c.end; rc = c.cursor; rp = c.cursor; rd = d;
    }
    int32_t loff = gw_alloc_state(s, newHeight);
    if (loff < 0) return -2;
    GwState l = c;   // currentSi, siTable are shared with the parent
    l.start = (int16_t)ls; l.end = (int16_t)le; l.cursor = (int16_t)lc; l.pivot = (int16_t)lp;
    l.cflag = (uint8_t)(strand | (ld << 1));
    l.height = (uint8_t)newHeight; l.k_offset = (uint8_t)(c.k_offset + 1); l.has_nfa = 1;
    l.state = gw_pack_state(gw_current_acgt(&c), minK, gw_priority(&c), gw_has_hit(&c));
    int32_t roff;
    if (!clip) {
        roff = gw_initial_state(env, s, strand, rd, rs, re, rc, rp, k, minK, gw_priority(&c));
        if (roff < 0) return -2;
    } else {
        roff = gw_alloc_state(s, 0);
        if (roff < 0) return -2;
        GwState r;
        r.start = (int16_t)rs; r.end = (int16_t)re; r.cursor = (int16_t)rc; r.pivot = (int16_t)rp;
        r.cflag = (uint8_t)(strand | (rd << 1));
        r.si_kind = GW_SI_NULL; r.f_valid = r.b_valid = 0;
        for (int i = 0; i < 4; ++i) { r.sif[i][0] = r.sif[i][1] = 0; r.sib[i][0] = r.sib[i][1] = 0; }
        r.has_cur = 0; r.cur_lb = r.cur_ub = 0;
        r.height = 0; r.k_offset = 0; r.has_nfa = 0;
        r.state = gw_pack_state(4, minK, gw_priority(&c), false) | (1 << 25);
        r.next_split = -1;
        *gw_st(s, roff) = r;
    }
    l.next_split = roff;
    *gw_st(s, loff) = l;
    u64* lr = gw_rows(gw_st(s, loff));
    for (int i = 0; i < newHeight; ++i) lr[i] = jshl(1ULL, newHeight + i - 1);  // ReadAlignmentNFA.nextStateAfterSplit
    return loff;
}

// SearchState.update(oldState, newState) on baseState: returns the queue entry (or -1 for null)
GW_HD int32_t gw_update(const GwSlab& s, int32_t base, int32_t oldState, int32_t newState) {
    if (oldState == base) return newState;
    int32_t prev = base;
    while (gw_st(s, prev)->next_split != oldState) {
        prev = gw_st(s, prev)->next_split;
        if (prev < 0) return -1;
    }
    gw_st(s, prev)->next_split = newState;
    return base;
}

// ---- hit list -----------------------------------------------------------------------------------------
GW_HD GwHitNode* gw_hit(const GwSlab& s, int32_t off) { return reinterpret_cast<GwHitNode*>(s.at(off)); }
GW_HD int gw_hit_total_match(const GwSlab& s, const GwHitNode* h) {
    int t = 0;
    for (;;) { t += h->match_length; if (h->next_split < 0) break; h = gw_hit(s, h->next_split); }
    return t;
}
GW_HD int gw_hit_total_diff(const GwSlab& s, const GwHitNode* h) {
    int t = 0;
    for (;;) { t += h->diff; if (h->next_split < 0) break; t += 1; h = gw_hit(s, h->next_split); }
    return t;
}
GW_HD int gw_hit_total_score(const GwSlab& s, const GwDevConfig& cfg, const GwHitNode* h) {
    int t = 0;
    for (;;) {
        t += h->match_length * cfg.match_score - h->diff * cfg.mismatch_penalty;
        if (h->next_split < 0) break;
        t -= cfg.split_open_penalty;
        h = gw_hit(s, h->next_split);
    }
    return t;
}
GW_HD int32_t gw_new_hit(const GwSlab& s, int chr, int pos, int matchLength, int qStart, int qEnd, int diff, int strand,
                                              const uint32_t* cig, int cigN, int numHits) {
    int32_t off = gw_alloc(s, GW_HIT_WORDS + (uint32_t)((cigN + 1) / 2));
    if (off < 0) return -1;
    GwHitNode* h = gw_hit(s, off);
    h->chr = chr; h->pos = pos; h->match_length = matchLength; h->diff = diff; h->strand = strand; h->num_hits = numHits;
    h->q_start = qStart; h->q_end = qEnd; h->next_split = -1; h->cig_n = cigN;
    uint32_t* c = gw_hit_cigar(h);
    for (int i = 0; i < cigN; ++i) c[i] = cig[i];
    return off;
}
// AlignmentResultHolder.add ; false on hit-list overflow
GW_HD bool gw_add_result(const GwSlab& s, const GwDevConfig& cfg, int32_t hit) {
    GwReadCtx* c = s.ctx();
    const GwHitNode* h = gw_hit(s, hit);
    int newK = gw_hit_total_diff(s, h), matchLen = gw_hit_total_match(s, h), newScore = gw_hit_total_score(s, cfg, h);
    if (newScore > c->best_score) {
        if (matchLen > 0 && newK <= c->min_mm) c->min_mm = newK;
        c->best_score = newScore;
    }
    if (c->max_match_len < matchLen) c->max_match_len = matchLen;
    int32_t* hl = s.hits();
    int w = 0;
    for (int i = 0; i < c->n_hits; ++i) {
        const GwHitNode* e = gw_hit(s, hl[i]);
        if (gw_hit_total_diff(s, e) <= c->min_mm && gw_hit_total_match(s, e) >= c->max_match_len) hl[w++] = hl[i];
    }
    if ((uint32_t)w >= s.sp.hit_cap) return false;
    hl[w++] = hit;
    c->n_hits = w;
    return true;
}
