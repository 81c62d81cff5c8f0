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

// allocate a state (room for the read's full NFA height whatever `height` is: see gw_state_words); fields are filled by the caller
GW_HD int32_t gw_alloc_state(const GwSlab& s, int height) {
    (void)height;
    GwReadCtx* c = s.ctx();
    if (c->free1 >= 0) { const int32_t o = c->free1; c->free1 = -1; return o; }
    return gw_alloc(s, gw_state_words(c->k));
}
// whole-state copy from / to the arena (six 128-bit accesses; states are 16-byte aligned)
GW_HD GwState gw_load_state(const GwSlab& s, int32_t off) { GwState st; gw_state_get(st, gw_st(s, off)); return st; }

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
    gw_state_put(gw_st(s, off), st);
    u64* rows = gw_rows(gw_st(s, off));
    for (int i = 0; i <= k; ++i) rows[i] = jshl(1ULL, k + i);
    return off;
}

GW_HD bool gw_si_is_empty(const GwState* c, int ch) {
    if (c->si_kind == GW_SI_EMPTY || c->si_kind == GW_SI_NULL) return true;
    if (c->si_kind == GW_SI_B) return !((c->b_valid >> ch) & 1) || c->sib[ch][0] >= c->sib[ch][1];
    return !((c->f_valid >> ch) & 1) || c->sif[ch][0] >= c->sif[ch][1];
}

// ---- search lane -------------------------------------------------------------------------------------------
// A lane (one thread) drives one read at a time. What every step of the best-first loop touches is kept in a small per-lane
// record while the lane owns the read: the counters and cursors of the context, the root of the priority queue, and the state
// that is being extended (`C`, with its NFA rows). In the flow kernel these records live in SHARED memory, one per thread at an odd
// word stride (all members are 32-bit, so the lanes of a warp never collide on a bank): the lane's hot data costs neither registers
// (which would cap the SM at a quarter of its warps) nor L1 lines, and the run-time indexed accesses of a step (the interval of the
// candidate base, the NFA row to cut at) are plain indexed loads. The work area in HBM stays the canonical copy for everything else;
// the lane's record is written back (gw_lane_flush) before any of the out-of-line continuations runs on the work area and re-read
// after it (gw_lane_reload). States are written through to the arena when they are built, so `cached` only ever saves the load: after
// an extension with the read's own base the child is the next state polled nearly every time, and it is already here.
template <int MAXH>
struct GwLane {
    int32_t m, k, min_mm, best_score;
    int32_t n_fm, n_cut, n_filt;
    int32_t qn, top, free1;
    int32_t sf_mm, sf_off;      // staircase table of minMismatches == sf_mm: word offset, -1: the reference throws
    uint32_t root_key_lo, root_key_hi;   // queue entry 0 (valid while qn > 0)
    int32_t root_off;
    int32_t base;               // polled queue entry (head of the split chain)
    int32_t c_off;              // fragment of the chain that is being extended
    int32_t mid;                // GW_MID_*: what remains of the current step
    int32_t cached;             // arena offset whose state C / rows hold, -1: none
    GwState C;
    uint32_t rows[2 * MAXH];    // NFA rows of C as (low, high) words
    GW_HD u64 root_key() const { return (u64)root_key_lo | ((u64)root_key_hi << 32); }
    GW_HD void set_root(u64 key, int32_t off) { root_key_lo = (uint32_t)key; root_key_hi = (uint32_t)(key >> 32); root_off = off; }
    GW_HD u64 row(int i) const { return (u64)rows[2 * i] | ((u64)rows[2 * i + 1] << 32); }
    GW_HD void set_row(int i, u64 v) { rows[2 * i] = (uint32_t)v; rows[2 * i + 1] = (uint32_t)(v >> 32); }
};
#define GW_LANE_WORDS(MAXH) ((uint32_t)(sizeof(GwLane<MAXH>) / 4))
static_assert(GW_LANE_WORDS(4) % 2 == 1 && GW_LANE_WORDS(8) % 2 == 1 && GW_LANE_WORDS(GW_MAX_ROWS) % 2 == 1, "odd word stride: conflict-free lanes");

#define GW_MID_TRY 1
#define GW_MID_SPLIT 2
#define GW_MID_REPORT 3

template <int MAXH>
GW_HD void gw_lane_reload(const GwSlab& s, GwLane<MAXH>& ln) {
    const GwReadCtx* c = s.ctx();
    ln.min_mm = c->min_mm; ln.best_score = c->best_score;
    ln.n_fm = c->n_fm; ln.n_cut = c->n_cut; ln.n_filt = c->n_filt;
    ln.qn = c->qn; ln.top = c->top; ln.free1 = c->free1;
    const GwQEntry r = s.queue()[0];
    ln.set_root(r.key, r.off);
}
template <int MAXH>
GW_HD void gw_lane_load(const GwSlab& s, GwLane<MAXH>& ln) {
    const GwReadCtx* c = s.ctx();
    ln.m = c->m; ln.k = c->k;
    ln.sf_mm = -2; ln.sf_off = -1;
    ln.base = ln.c_off = ln.cached = -1;
    ln.mid = 0;
    gw_lane_reload(s, ln);
}
template <int MAXH>
GW_HD void gw_lane_flush(const GwSlab& s, const GwLane<MAXH>& ln) {
    GwReadCtx* c = s.ctx();
    c->n_fm = ln.n_fm; c->n_cut = ln.n_cut; c->n_filt = ln.n_filt;
    c->qn = ln.qn; c->top = ln.top; c->free1 = ln.free1;
    if (ln.qn > 0) { GwQEntry r; r.key = ln.root_key(); r.off = ln.root_off; r.pad_ = 0; s.queue()[0] = r; }
}

// the state at `off` into the lane's record
template <int MAXH>
GW_HD void gw_lane_fetch(const GwSlab& s, GwLane<MAXH>& ln, int32_t off) {
    GwState* p = gw_st(s, off);
    gw_state_get(ln.C, p);
    const u64* rows = gw_rows(p);
    const int height = ln.C.height;
    if constexpr (MAXH <= 8) {
#pragma unroll
        for (int i = 0; i < MAXH; ++i) ln.set_row(i, i < height ? rows[i] : 0ULL);
    } else {
        for (int i = 0; i < height; ++i) ln.set_row(i, rows[i]);
    }
    ln.cached = off;
}

// java.util.PriorityQueue.offer / poll as above, with entry 0 in the lane's record
template <int MAXH>
GW_HD bool gw_lq_add(const GwSlab& s, GwLane<MAXH>& ln, int32_t e, const u64 ke) {
    GwQEntry* q = s.queue();
    int k = ln.qn;
    if ((uint32_t)k >= s.sp.queue_cap) return false;
    ln.qn = k + 1;
    while (k > 2) {
        const int parent = (k - 1) >> 1;
        const GwQEntry p = q[parent];
        if (ke >= p.key) break;
        q[k] = p;
        k = parent;
    }
    if (k > 0 && k <= 2 && ke < ln.root_key()) {   // the parent is the root
        GwQEntry p; p.key = ln.root_key(); p.off = ln.root_off; p.pad_ = 0;
        q[k] = p;
        k = 0;
    }
    if (k == 0) ln.set_root(ke, e);
    else { GwQEntry n; n.key = ke; n.off = e; n.pad_ = 0; q[k] = n; }
    return true;
}
template <int MAXH>
GW_HD int32_t gw_lq_poll(const GwSlab& s, GwLane<MAXH>& ln) {
    GwQEntry* q = s.queue();
    const int n = --ln.qn;
    const int32_t result = ln.root_off;
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
            if (k == 0) ln.set_root(cc.key, cc.off); else q[k] = cc;
            k = child;
        }
        if (k == 0) ln.set_root(x.key, x.off); else q[k] = x;
    }
    return result;
}

// a slot for a new state: the recycled one if there is one, else from the arena; -1 on overflow
template <int MAXH>
GW_HD int32_t gw_lane_alloc_state(const GwSlab& s, GwLane<MAXH>& ln) {
    if (ln.free1 >= 0) { const int32_t o = ln.free1; ln.free1 = -1; return o; }
    const uint32_t words = gw_state_words(ln.k), top = (uint32_t)ln.top;
    if (top + words > s.sp.slab_words) return -1;
    ln.top = (int32_t)(top + words);
    return (int32_t)top;
}
// the polled state of a finished step that heads no split chain is unreachable: its slot goes to the next state built
template <int MAXH>
GW_HD void gw_lane_free_state(GwLane<MAXH>& ln, int32_t off) {
    if (ln.free1 < 0) ln.free1 = off;
    if (ln.cached == off) ln.cached = -1;
}

GW_HD int gw_q_symbol(const GwReadCtx* c, int strand, int i) {
    return (int)((((c->qhi[strand][i >> 6] >> (i & 63)) & 1ULL) << 1) | ((c->qlo[strand][i >> 6] >> (i & 63)) & 1ULL));
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
        gw_state_put(gw_st(s, roff), r);
    }
    l.next_split = roff;
    gw_state_put(gw_st(s, loff), l);
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
    if ((uint32_t)w >= c->sp.hit_cap) {
        // the hit list is full (ALLHITS on repeats): it moves into the arena with twice the capacity — the arena is what grows when a
        // read moves to a larger work area — and the old list is abandoned. The context holds the place and the capacity (s.sp is a copy).
        const uint32_t ncap = c->sp.hit_cap * 2u;
        const int32_t at = gw_alloc(s, (ncap + 1u) / 2u);
        if (at < 0) return false;
        int32_t* nl = reinterpret_cast<int32_t*>(s.at(at));
        for (int i = 0; i < w; ++i) nl[i] = hl[i];
        c->hits_at = (uint32_t)at; c->sp.hit_cap = ncap;
        hl = nl;
    }
    hl[w++] = hit;
    c->n_hits = w;
    return true;
}
