// gw_state.cuh — per-read work area ("slab") in HBM and the value types of the BSF search:
// cursor, suffix-interval set, NFA rows, search state, priority queue, hit list.
// Replaces (reference, align/src/main/java/org/utgenome/weaver/align/): strategy/Cursor.java:40-198,
// SiSet.java:34-196, strategy/ReadAlignmentNFA.java:63-215, strategy/StaircaseFilter.java:91-102,
// QueryMask.java:73-97, BitVector.java:106-116, strategy/BidirectionalSuffixFilter.java:133-165 (StateQueue),
// :658-877 (SearchState), record/ReadHit.java:42-184.
//
// Object identity matters in the reference (states are mutated while aliased from several queue entries and
// split chains), so states live in a bump-allocated arena and are referred to by word offsets, never copied.
#pragma once
#include "gw_fm.cuh"
#include "gw_verify.cuh"

// ---- slab layout (8-byte words) -------------------------------------------------------------------------
struct GwPendFrag {
    int32_t state_off;   // arena offset of the fragment's SearchState
    int32_t kind;        // 0 clipped, 1 no hit (currentSi == null), 2 needs locate+verify
    uint32_t sa_index;   // currentSi.lowerBound
    int32_t fm;          // index used by toCoordinate: 0 forward, 1 reverse
    uint32_t textpos;    // SparseSuffixArray.get result (written by the locate kernel)
    int32_t res_valid;   // 0: alignBlockDetailed returned null
    int32_t res_pos, res_diff, res_cig_n;
    int32_t res_abort;   // the reference would throw while verifying this fragment
    long long ref_start; // window start (verify kernel)
    uint32_t cig[GW_CIG_CAP];
};

// verification of one pending fragment (gw_job.cuh): lives in the local memory of the lane that runs it
#define GW_RW_WORDS ((GW_MAX_WINDOW + 63) / 64)
struct GwJob {
    int32_t frag;            // index into pend[]
    int32_t pad_;
    GwDp dp;
    u64 peq[4][GW_MAX_NB];   // pattern masks of the query fragment
    u64 rw_lo[GW_RW_WORDS], rw_hi[GW_RW_WORDS], rw_nf[GW_RW_WORDS];   // reference window planes, column 0 at bit 0
};

struct GwReadCtx {
    uint32_t area_words;   // size of the work area the read lives in (sp.slab_words is where its arena ends)
    uint32_t hits_at;      // word offset of the hit list (it stays where gw_layout_init put it)
    GwExtPool* ext_pool;   // pool of such blocks, or null
    int32_t m, k;
    int32_t min_mm, max_match_len, best_score;
    int32_t n_fm, n_sw, n_cut, n_filt;
    int32_t qn, top, phase, status;
    int32_t n_hits;
    int32_t free1;            // one recycled state slot (arena offset) or -1: see gw_state_words
    int32_t scan_mm[2], scan_ls[2], scan_le[2];
    uint32_t scan_lb[2], scan_ub[2];
    int32_t pend_base, pend_n;
    int32_t sf_mm, sf_off;    // staircase table of minMismatches == sf_mm: word offset in the table set, -1: the reference throws (cache)
    int32_t count_n, read_id;
    GwSlabParams sp;          // geometry of the slab this context lives in
    uint32_t occ_scan, occ_search, occ_locate;   // occurrence-table lookups done for this read (roofline work unit)
    GwPendFrag pend[2];
    u64 qlo[2][GW_MAX_NB], qhi[2][GW_MAX_NB];  // q[0] = read, q[1] = its complement, N replaced by A in both
};

#define GW_CTX_WORDS ((uint32_t)((sizeof(GwReadCtx) + 7) / 8))
#define GW_QUEUE_BASE ((GW_CTX_WORDS + 1u) & ~1u)   // 16-byte aligned

struct alignas(16) GwQEntry {
    u64 key;
    int32_t off;
    int32_t pad_;
};

struct GwSlab {
    u64* base;
    GwSlabParams sp;
    GW_HD GwSlab() : base(nullptr) {}
    GW_HD explicit GwSlab(u64* b) : base(b) { sp = reinterpret_cast<GwReadCtx*>(b)->sp; }
    GW_HD GwSlab(u64* b, const GwSlabParams& p) : base(b), sp(p) { if (sp.queue_at == 0u) sp.queue_at = GW_QUEUE_BASE; }
    GW_HD GwReadCtx* ctx() const { return reinterpret_cast<GwReadCtx*>(base); }
    // priority queue: one 16-byte entry per queued state — the sort key (comparator of StateQueue packed into one integer, cached at
    // insertion) next to the state offset, so that a sift step is ONE 128-bit load per level
    GW_HD GwQEntry* queue() const { return reinterpret_cast<GwQEntry*>(base + sp.queue_at); }
    GW_HD int32_t* hits() const { return reinterpret_cast<int32_t*>(base + ctx()->hits_at); }   // cold paths only
    // a fresh work area with the standard layout: context, queue, hit list, arena
    GW_HD void layout_init() {
        sp.queue_at = GW_QUEUE_BASE;
        GwReadCtx* c = ctx();
        c->sp = sp; c->area_words = sp.slab_words; c->hits_at = GW_QUEUE_BASE + 2u * sp.queue_cap;
    }
    // 128-byte aligned: a state of a read with k <= 3 (16 words) is exactly one line
    GW_HD uint32_t arena_begin() const { return (GW_QUEUE_BASE + 2u * sp.queue_cap + (sp.hit_cap + 1) / 2 + 15u) & ~15u; }
    GW_HD u64* at(int32_t off) const { return base + off; }
};


// ---- search state node --------------------------------------------------------------------------------
// si_kind: SiSet subclass
#define GW_SI_NULL 0
#define GW_SI_EMPTY 1
#define GW_SI_F 2
#define GW_SI_B 3
#define GW_SI_BD 4

#define GW_DIR_F 0
#define GW_DIR_B 1
#define GW_DIR_BF 2

// 4-byte alignment only: the search lanes keep a copy of the state they extend in shared memory at an odd word stride (no bank
// conflicts between the lanes of a warp); in the arena states sit on 16-byte boundaries and are moved with 128-bit accesses
// (gw_state_put / gw_state_get).
struct GwState {
    int16_t start, end, cursor, pivot;   // Cursor
    uint8_t cflag;                       // strand | dir << 1
    uint8_t si_kind;
    uint8_t f_valid, b_valid;            // non-null entries of siF / siB (bit per base)
    uint8_t height;                      // NFA rows; 0 with has_nfa == 0 means null automaton
    uint8_t k_offset;
    uint8_t has_cur;                     // currentSi != null
    uint8_t has_nfa;
    int32_t state;                       // packed flags as in SearchState.state
    int32_t next_split;                  // arena offset, -1 == null
    uint32_t cur_lb, cur_ub;
    uint32_t sif[4][2], sib[4][2];
    // followed by `height` u64 NFA rows
};
#define GW_STATE_WORDS ((uint32_t)((sizeof(GwState) + 7) / 8))

static_assert(sizeof(GwState) == 96, "GwState is 24 words");
// whole-state copies between the arena (16-byte aligned) and a lane's copy (4-byte aligned)
GW_HD void gw_state_put(GwState* arena, const GwState& src) { memcpy(__builtin_assume_aligned(arena, 16), &src, sizeof(GwState)); }
GW_HD void gw_state_get(GwState& dst, const GwState* arena) { memcpy(&dst, __builtin_assume_aligned(arena, 16), sizeof(GwState)); }
GW_HD u64* gw_rows(GwState* s) { return reinterpret_cast<u64*>(s) + GW_STATE_WORDS; }
GW_HD int gw_strand(const GwState* s) { return s->cflag & 1; }
GW_HD int gw_dir(const GwState* s) { int d = (s->cflag >> 1) & 3; return d == 3 ? GW_DIR_F : d; }
GW_HD bool gw_is_fwd(const GwState* s) { return gw_dir(s) != GW_DIR_B; }
GW_HD int gw_frag_len(const GwState* s) { return s->end - s->start; }
GW_HD int gw_processed(const GwState* s) { return gw_is_fwd(s) ? s->cursor - s->pivot : s->end - s->cursor; }
GW_HD int gw_remaining(const GwState* s) { return gw_frag_len(s) - gw_processed(s); }
GW_HD int gw_next_index(const GwState* s) { return gw_is_fwd(s) ? s->cursor : s->cursor - 1; }
GW_HD int gw_head_offset(const GwState* s) {
    int off = s->cursor - s->start;
    if (gw_strand(s) == 1) off = gw_frag_len(s) - off;
    return off;
}
GW_HD int gw_lower_k(const GwState* s) { return ((uint32_t)s->state >> 8) & 0xFF; }
GW_HD void gw_set_lower_k(GwState* s, int d) { s->state = (s->state & ~(0xFF << 8)) | ((d & 0xFF) << 8); }
GW_HD int gw_priority(const GwState* s) { return ((uint32_t)s->state >> 16) & 0xFF; }
GW_HD bool gw_has_hit(const GwState* s) { return (((uint32_t)s->state >> 24) & 1) != 0; }
GW_HD bool gw_is_clipped(const GwState* s) { return (((uint32_t)s->state >> 25) & 1) != 0; }
GW_HD bool gw_is_finished(const GwState* s) { return (s->state & 0x1F) == 0x1F; }
GW_HD int gw_current_acgt(const GwState* s) { int c = (s->state >> 5) & 7; return c > 4 ? 4 : c; }
GW_HD int gw_pack_state(int ch, int minK, int prio, bool hasMatch) {
    return ((ch & 7) << 5) | ((minK & 0xFF) << 8) | ((prio & 0xFF) << 16) | ((hasMatch ? 1 : 0) << 24);
}

// ---- ReadHit node -------------------------------------------------------------------------------------
struct GwHitNode {
    int32_t chr;          // contig id or GW_CHR_*
    int32_t pos;
    int32_t match_length, diff, strand, num_hits, q_start, q_end;
    int32_t next_split;   // arena offset or -1
    int32_t cig_n;
    // followed by cig_n u32 cigar elements (BAM encoding)
};
#define GW_HIT_WORDS ((uint32_t)((sizeof(GwHitNode) + 7) / 8))
GW_HD uint32_t* gw_hit_cigar(GwHitNode* h) { return reinterpret_cast<uint32_t*>(reinterpret_cast<u64*>(h) + GW_HIT_WORDS); }

// Every search state of a read is allocated with room for the full NFA height of the read (k + 1 rows), whatever its own height:
// states then have ONE size per read, so the slot of a state that can no longer be reached (the polled state of a finished step
// that heads no split chain: nothing but its queue entry ever referred to it) can be handed to the next state that is built.
// The usual step — extension with the read's own base — even rebuilds the child in the parent's slot, so a read that follows its
// greedy path keeps rewriting the same 128 bytes instead of walking through the arena.
GW_HD uint32_t gw_state_words(int k) { return (GW_STATE_WORDS + (uint32_t)k + 2u) & ~1u; }

// bump allocation; returns -1 (and the caller flags the read for the large-slab pass) on overflow
GW_HD int32_t gw_alloc(const GwSlab& s, uint32_t words) {
    GwReadCtx* c = s.ctx();
    words = (words + 1u) & ~1u;   // nodes stay 16-byte aligned (128-bit loads/stores of whole states)
    uint32_t top = (uint32_t)c->top;
    if (top + words > s.sp.slab_words) return -1;
    c->top = (int32_t)(top + words);
    return (int32_t)top;
}

// Moving a read to a larger work area while it runs. All references inside a work area are word offsets from its base and the
// layout in front of the arena (context, queue, hit list) depends on queue_cap / hit_cap only, so a work area can be copied to the
// start of a larger block and simply go on there with a larger slab_words. This is done between two steps of the best-first loop
// (no pointer into the old area is live then), when the free space falls below what one step can allocate at most:
// up to 5 candidate states + 4 states of the split / clip children + 2 hit nodes.
GW_HD uint32_t gw_step_alloc_bound(int k) {
    return 9u * gw_state_words(k) + 2u * ((GW_HIT_WORDS + GW_CIG_CAP / 2u + 1u) & ~1u);
}
// `top`: the read's allocation cursor (kept in a register by the search lanes)
#define GW_STEP_QUEUE_BOUND 8u   // queue entries one search slice can add (candidate, split and soft-clip children), generously
GW_HD bool gw_wants_promotion(const GwSlab& s, uint32_t top, int k, int qn) {
    if (top + gw_step_alloc_bound(k) <= s.sp.slab_words && (uint32_t)qn + GW_STEP_QUEUE_BOUND <= s.sp.queue_cap) return false;
    return s.ctx()->ext_pool != nullptr;   // cleared when the pool has nothing larger left for this read (gw_promote_take)
}
// the move in three parts, so that the copy can be shared by the lanes of a warp: take a block from the pool (null: none left) ...
GW_COLD u64* gw_promote_take(const GwSlab& s) {
    GwReadCtx* c = s.ctx();
    GwExtPool* pool = c->ext_pool;
    const uint32_t cur = c->area_words;
    if (pool->block_words > cur && pool->next < pool->n_blocks) {
        const uint32_t idx = gw_atomic_add(&pool->next, 1u);
        if (idx < pool->n_blocks) return pool->base + (size_t)idx * pool->block_words;
    }
    // the second size: for reads that filled a block of the first, and for those that found none left
    if (pool->n_blocks2 != 0u && pool->block_words2 > cur && pool->next2 < pool->n_blocks2) {
        const uint32_t idx = gw_atomic_add(&pool->next2, 1u);
        if (idx < pool->n_blocks2) return pool->base2 + (size_t)idx * pool->block_words2;
    }
    c->ext_pool = nullptr;   // nothing larger for this read: it runs until its work area is full and is run again in a later pass
    return nullptr;
}
// ... copy the used part of the work area [0, top) (top is even: the arena starts at an even word, nodes are multiples of two
// words), `part` of `parts` ...
GW_HD void gw_promote_copy(const u64* from, u64* to, uint32_t top, uint32_t part, uint32_t parts) {
    const GwU64x2* src = reinterpret_cast<const GwU64x2*>(from);
    GwU64x2* dst = reinterpret_cast<GwU64x2*>(to);
    for (uint32_t i = part; i < top / 2u; i += parts) dst[i] = src[i];
}
// ... and make the block the read's work area: the queue goes to the end of the block with the capacity a work area of that size has
// (its entries are state offsets, which do not change: the arena stays where it is and now ends where the queue begins)
GW_HD void gw_promote_commit(GwSlab& s, u64* nb, u64** slab_ptr) {
    GwReadCtx* nc = reinterpret_cast<GwReadCtx*>(nb);
    const GwExtPool* pool = nc->ext_pool;
    const bool second = pool->n_blocks2 != 0u && nb >= pool->base2 && nb < pool->base2 + (size_t)pool->n_blocks2 * pool->block_words2;
    const uint32_t words = second ? pool->block_words2 : pool->block_words;
    uint32_t ncap = words / 16u;
    if (ncap > 8192u) ncap = 8192u;
    if (ncap < s.sp.queue_cap) ncap = s.sp.queue_cap;
    const uint32_t nat = words - 2u * ncap;
    const GwQEntry* from = s.queue();   // in the work area the read leaves (at its end if that was a block already: not part of the copy)
    GwQEntry* to = reinterpret_cast<GwQEntry*>(nb + nat);
    for (int32_t i = 0; i < nc->qn; ++i) to[i] = from[i];
    nc->sp.queue_at = nat; nc->sp.queue_cap = ncap; nc->sp.slab_words = nat;
    nc->area_words = words;
    *slab_ptr = nb;
    s.base = nb;
    s.sp = nc->sp;
}
GW_COLD bool gw_promote(GwSlab& s, u64** slab_ptr) {
    u64* nb = gw_promote_take(s);
    if (!nb) return false;
    gw_promote_copy(s.base, nb, (uint32_t)s.ctx()->top, 0u, 1u);
    gw_promote_commit(s, nb, slab_ptr);
    return true;
}

// ---- bit-vector windows ------------------------------------------------------------------------------------
// All of these are written without data-dependent branches (clamped indices + selects): they sit in the hottest loop of the
// search lanes, where a branch costs both divergence and instruction-cache footprint.

// (hi:lo) >> off for off in 0..63, as BitVector.substring64 combines two blocks (BitVector.java:106-116)
GW_HD u64 gw_funnel64(u64 lo, u64 hi, int off) { return (lo >> off) | ((hi << 1) << (63 - off)); }

// BitVector.substring64(start, start + 64) over nb blocks, start >= 0
GW_HD u64 gw_substring64(const u64* block, int nb, int start) {
    const int pos = start >> 6, off = start & 63;
    const u64 lo = block[gw_min(pos, nb - 1)], hi = block[gw_min(pos + 1, nb - 1)];
    return gw_funnel64(pos < nb ? lo : 0ULL, pos + 1 < nb ? hi : 0ULL, off);
}

// QueryMask pattern vectors are not materialised: bit i of patternMaskF[ch] is (q[i] == ch), which is read straight from the
// 2-bit planes of the read, and patternMaskR is patternMaskF mirrored (bit m-1-i), i.e. a bit reversal of a forward window.
// One 64-symbol word of patternMaskF[ch] (bits at positions >= m are zero; w outside [0, nb) gives zero)
GW_HD u64 gw_match_word(const u64* qlo, const u64* qhi, int m, int nb, int ch, int w) {
    const int wi = gw_min(gw_max(w, 0), nb - 1);
    const u64 lo = qlo[wi], hi = qhi[wi];
    const int rem = m - (wi << 6);
    const u64 valid = rem >= 64 ? ~0ULL : ((1ULL << rem) - 1ULL);
    const u64 f = ((ch & 2) ? hi : ~hi) & ((ch & 1) ? lo : ~lo) & valid;
    return w == wi ? f : 0ULL;
}
// 64 bits of patternMaskF[ch] starting at bit `start` (any integer; bits outside [0, m) are zero)
GW_HD u64 gw_match_window(const u64* qlo, const u64* qhi, int m, int ch, int start) {
    const int nb = (m + 63) >> 6;
    if (start <= -64) return 0ULL;
    const int w = start >> 6, off = start & 63;   // arithmetic shift: w == -1 for start in [-63, -1]
    return gw_funnel64(gw_match_word(qlo, qhi, m, nb, ch, w), gw_match_word(qlo, qhi, m, nb, ch, w + 1), off);
}

// QueryMask.getBidirectionalPatternMask64 (QueryMask.java:73-97) on the planes of q[strand]
GW_HD u64 gw_pattern_mask64(const GwReadCtx* c, int strand, int m, bool dirIsForward, int nextIdx, int pivot, int cursor, int ch, int margin) {
    const u64* qlo = c->qlo[strand];
    const u64* qhi = c->qhi[strand];
    // forward: substring64(pos, m) with pos = nextIdx - margin, or substring64(0, 64) << -pos for negative pos (the bits shifted in
    // from below position 0 are zero either way). backward: patternMaskR.substring64(m - pivot, m - pivot + 64), whose bit j is
    // F[pivot - 1 - j], shifted by pivot - cursor - margin (right if positive)
    const u64 win = gw_match_window(qlo, qhi, m, ch, dirIsForward ? gw_max(nextIdx - margin, 0) : pivot - 64);
    const int sh = dirIsForward ? gw_min(nextIdx - margin, 0) : pivot - cursor - margin;
    const u64 p = dirIsForward ? win : gw_brev64(win);
    return sh >= 0 ? jushr(p, sh) : jshl(p, -sh);
}

// StaircaseFilter.getStairCaseMask64bit for the filter (m, kk); tbl points at the (kk+1) x nb words of that filter
GW_HD u64 gw_staircase64(const u64* tbl, int m, int kk, int row, int offset) {
    const int nb = (m + 63) >> 6;
    const u64* bv = tbl + gw_min(row, kk) * nb;
    const u64 sub = gw_substring64(bv, nb, gw_max(offset, 0));
    const u64 v = offset >= 0 ? (jshl(~0ULL, m - offset) | sub) : jshl(sub, -offset);
    return row >= kk + 1 ? 0ULL : v;
}
