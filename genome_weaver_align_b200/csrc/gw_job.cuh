// gw_job.cuh — verification of one pending fragment of a parked read: reference window and pattern masks (AlignmentProcess.verify,
// reference: align/src/main/java/org/utgenome/weaver/align/strategy/BidirectionalSuffixFilter.java:496-560), the bit-vector DP
// columns and the traceback (align/BitParallelSmithWaterman.java:387-644, gw_verify.cuh), as three calls so that the lanes of a
// verification batch run each part together.
#pragma once
#include "gw_bodies.cuh"

// 64 window columns starting at text position p (bit j = position p + j; positions >= N read as 0)
GW_HD u64 gw_plane_window(const u64* plane, uint32_t n_words, uint32_t p) {
    const uint32_t w = p >> 6, off = p & 63u;
    if (w >= n_words) return 0ULL;
    u64 v = gw_ld64(plane + w) >> off;
    if (off && w + 1 < n_words) v |= gw_ld64(plane + w + 1) << (64 - off);
    return v;
}

// AlignmentProcess.verify, window + query set-up (:517-538); returns false when the fragment is finished without a DP
GW_COLD bool gw_job_begin_dp(const GwSearchEnv& env, const GwSlab& s, GwJob* job, int frag, GwU64x2* hist, size_t stride) {
    GwReadCtx* c = s.ctx();
    job->frag = frag;
    GwPendFrag* pf = &c->pend[frag];
    const GwIndexView& ix = env.ix;
    const GwState* st = gw_st(s, pf->state_off);
    const int strand = gw_strand(st);
    const int start = st->start, end = st->end;
    const int fragLen = end - start;
    const int k = c->k;
    const long long seqIndex = pf->fm == 0 ? (long long)pf->textpos : (long long)ix.N - (long long)pf->textpos;
    const long long x = seqIndex - gw_head_offset(st);
    long long refStart = x - k; if (refStart < 0) refStart = 0;
    long long refEnd = x + fragLen + k; if (refEnd > (long long)ix.N) refEnd = (long long)ix.N;
    pf->ref_start = refStart;
    pf->res_abort = 0;
    if (refStart > refEnd) { pf->res_abort = 1; pf->res_valid = 0; return false; }  // ACGTSequence.subString throws
    const int Ncols = (int)(refEnd - refStart);
#ifdef GW_VERIFY_STATS
    fprintf(stderr, "VS %d %d %d %d %lld %d\n", c->read_id, strand, start, end, refStart, Ncols);
#endif
    // pattern masks of the fragment q[strand][start, end), reversed on the reverse strand (:529-534), from the read's bit planes
    const int m = c->m;
    const int nbf = gw_max(1, (fragLen + 63) >> 6);
    for (int b = 0; b < GW_MAX_NB; ++b) {
        const int rem = fragLen - 64 * b;
        const u64 keep = rem >= 64 ? ~0ULL : (rem <= 0 ? 0ULL : ((1ULL << rem) - 1ULL));
        for (int a = 0; a < 4; ++a) {
            u64 v = 0ULL;
            if (b < nbf && rem > 0) {
                if (strand == 0) v = gw_match_window(c->qlo[0], c->qhi[0], m, a, start + 64 * b);
                else v = gw_brev64(gw_match_window(c->qlo[1], c->qhi[1], m, a, end - 64 * (b + 1)));
            }
            job->peq[a][b] = v & keep;
        }
    }
    const uint32_t n_words = (ix.N >> 6) + 1u;
    for (int w = 0; w < GW_RW_WORDS; ++w) {
        const int rem = Ncols - 64 * w;
        if (rem <= 0) { job->rw_lo[w] = job->rw_hi[w] = job->rw_nf[w] = 0ULL; continue; }
        const uint32_t p = (uint32_t)refStart + 64u * (uint32_t)w;
        job->rw_lo[w] = gw_plane_window(ix.ref_lo, n_words, p);
        job->rw_hi[w] = gw_plane_window(ix.ref_hi, n_words, p);
        job->rw_nf[w] = gw_plane_window(ix.ref_nf, n_words, p);
    }
    GwDp d;
    gw_dp_begin(d, fragLen, Ncols, env.cfg.band_width, hist, stride);
    job->dp = d;
    return true;
}

GW_HD int gw_job_ref_symbol(const GwJob* job, int j) {
    const int w = j >> 6, sft = j & 63;
    if ((job->rw_nf[w] >> sft) & 1ULL) return 4;
    return (int)((((job->rw_hi[w] >> sft) & 1ULL) << 1) | ((job->rw_lo[w] >> sft) & 1ULL));
}

// traceback + CIGAR of the fragment whose DP is complete; results go to its pending-fragment record
GW_HD void gw_job_trace(const GwSlab& s, GwJob* job, const GwU64x2* hist, size_t stride) {
    GwReadCtx* c = s.ctx();
    GwPendFrag* pf = &c->pend[job->frag];
    const GwState* st = gw_st(s, pf->state_off);
    const int strand = gw_strand(st), start = st->start, end = st->end;
    const u64* qlo = c->qlo[strand];
    const u64* qhi = c->qhi[strand];
    // the traceback moves through rows and columns one at a time: keep the 64-symbol words it is in
    int qw = -1; u64 ql = 0, qh = 0;
    auto qsym = [&](int i) -> int {
        const int p = strand == 0 ? start + i : end - 1 - i;
        if ((p >> 6) != qw) { qw = p >> 6; ql = qlo[qw]; qh = qhi[qw]; }
        return (int)((((qh >> (p & 63)) & 1ULL) << 1) | ((ql >> (p & 63)) & 1ULL));
    };
    int rw = -1; u64 rl = 0, rh = 0, rn = 0;
    auto refsym = [&](int j) -> int {
        if ((j >> 6) != rw) { rw = j >> 6; rl = job->rw_lo[rw]; rh = job->rw_hi[rw]; rn = job->rw_nf[rw]; }
        const int sft = j & 63;
        if ((rn >> sft) & 1ULL) return 4;
        return (int)((((rh >> sft) & 1ULL) << 1) | ((rl >> sft) & 1ULL));
    };
    GwVerifyOut out;
    gw_dp_traceback(job->dp, refsym, qsym, hist, stride, pf->cig, &out);
    pf->res_valid = out.valid; pf->res_pos = out.pos; pf->res_diff = out.diff; pf->res_cig_n = out.cig_n;
    if (out.cig_overflow) pf->res_abort = 2;  // cannot happen within GW_MAX_K; reported as unsupported
}

// DP columns [d.j, jEnd) of a job
GW_HD void gw_job_columns(const GwSearchEnv& env, GwJob* job, int n_cols, GwU64x2* hist, size_t stride, int nb_hint = 0) {
    GwDp d = job->dp;
    int rw = -1; u64 rl = 0, rh = 0, rn = 0;
    auto refsym = [&](int j) -> int {   // columns advance one by one: reload the window words every 64 columns only
        if ((j >> 6) != rw) { rw = j >> 6; rl = job->rw_lo[rw]; rh = job->rw_hi[rw]; rn = job->rw_nf[rw]; }
        const int sft = j & 63;
        if ((rn >> sft) & 1ULL) return 4;
        return (int)((((rh >> sft) & 1ULL) << 1) | ((rl >> sft) & 1ULL));
    };
    gw_dp_columns(d, d.j + n_cols, env.cfg.band_width, job->peq, refsym, hist, stride, nb_hint);
    job->dp = d;
}

// whole verification of one pending fragment in one go (pool kernel: the job lives in the lane's local memory)
GW_HD void gw_verify_frag(const GwSearchEnv& env, const GwSlab& s, GwJob* job, int frag, GwU64x2* hist, size_t stride) {
    if (!gw_job_begin_dp(env, s, job, frag, hist, stride)) return;
    gw_job_columns(env, job, job->dp.ncols, hist, stride);
    gw_job_trace(s, job, hist, stride);
}

