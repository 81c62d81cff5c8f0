// TEST INFRASTRUCTURE ONLY — host simulator of the device automaton.
// Compiles the per-thread kernel bodies of genome_weaver_align_b200/csrc (the GW_HD functions in gw_*.cuh) with g++ and
// runs them sequentially, one "thread" after another, through the same prep / scan / search-locate-verify rounds / emit
// phases as the CUDA driver in gw_api.cu. It exists so that the automaton logic can be checked against the oracle on a
// machine without a GPU (`pytest -m "not gpu"`). It is never linked into libgwalign.so and never used by the product:
// the product path is CUDA only and fails loudly without a device.
#include <stdlib.h>
#include <string.h>

#include <string>
#include <type_traits>
#include <vector>

#include "../../include/gw_align.h"
#include "../../genome_weaver_align_b200/csrc/gw_job.cuh"
#include "../../genome_weaver_align_b200/csrc/gw_host.h"
#include "../../genome_weaver_align_b200/csrc/gw_image.h"

bool gw_sam_one(const std::vector<GwContig>& ctg, const std::string& name, const char* bases, size_t m, const char* qual, size_t qlen,
                const gw_hit_t* chain, int n, const uint32_t* ops, std::string& out);

namespace {
struct SimIndex {
    GwHostIndex host;
    GwIndexImage img;
    GwIndexView view;
};
thread_local std::string g_err;

GwSlabParams slab_params(uint32_t words) {   // same rule as gw_api_align.inc
    GwSlabParams sp;
    sp.slab_words = words;
    sp.queue_cap = std::max<uint32_t>(64, std::min<uint32_t>(words / 16, 8192));
    sp.hit_cap = std::max<uint32_t>(32, std::min<uint32_t>(words / 64, 2048));
    if (const char* hc = getenv("GWSIM_HITCAP")) sp.hit_cap = (uint32_t)std::max(1, atoi(hc));   // tiny hit lists: they must grow (gw_add_result)
    sp.queue_at = 0;
    return sp;
}
}  // namespace

extern "C" {

const char* gwsim_last_error() { return g_err.c_str(); }

void* gwsim_open(const char* fasta_path) {
    try {
        SimIndex* s = new SimIndex();
        gw_load_index_files(fasta_path, s->host);
        gw_layout_index(s->host, s->img);
        s->view = s->img.view;
        for (int i = 0; i < 2; ++i) { s->view.occ[i] = s->img.occ[i].data(); s->view.ssa[i] = s->host.ssa[i].data(); s->view.loc[i] = s->view.ssa[i]; }
        s->view.ref_lo = s->img.ref_lo.data(); s->view.ref_hi = s->img.ref_hi.data(); s->view.ref_nf = s->img.ref_nf.data();
        s->view.ctg_off = s->img.ctg_off.data(); s->view.ctg_id = s->img.ctg_id.data(); s->view.ctg_rank = s->img.ctg_rank.data();
        return s;
    } catch (std::exception& e) { g_err = e.what(); return nullptr; }
}
void gwsim_close(void* h) { delete (SimIndex*)h; }

void gwsim_rank_acgtn(void* h, int which, uint32_t pos, uint32_t* out5) {
    SimIndex* s = (SimIndex*)h;
    uint32_t p = gw_clampN(s->view, pos);
    GwBlk b = gw_load_block(s->view.occ[which] + (p >> 7));
    gw_rank4(b, p & 127u, out5);
    out5[4] = gw_rankN(b, p & ~127u, p & 127u);
}
uint32_t gwsim_locate(void* h, int which, uint32_t row) {
    uint32_t steps;
    return gw_sa_get(((SimIndex*)h)->view, which, row, &steps);
}

// cfg: k, numSplitAllowed, matchScore, mismatchPenalty, splitOpenPenalty, indelEndSkip, bandWidth, reportType, topL, extended
// returns malloc'ed SAM text; status[n], counters[n*4], lines[n]; *tiers_used = reads that needed a larger work area
char* gwsim_align(void* h, const double* cfgv, uint32_t n_reads, const char* bases, const uint64_t* offsets, const char* names,
                  const uint64_t* name_off, const char* quals, const uint64_t* qual_off, uint32_t slab_words, int* status, int* counters,
                  int* lines, uint32_t* tiers_used) {
    try {
        SimIndex* S = (SimIndex*)h;
        GwSearchEnv env;
        env.ix = S->view;
        env.cfg.k = (float)cfgv[0]; env.cfg.num_split_allowed = (int)cfgv[1]; env.cfg.match_score = (int)cfgv[2];
        env.cfg.mismatch_penalty = (int)cfgv[3]; env.cfg.split_open_penalty = (int)cfgv[4]; env.cfg.indel_end_skip = (int)cfgv[5];
        env.cfg.band_width = (int)cfgv[6]; env.cfg.report_type = (int)cfgv[7]; env.cfg.top_l = (int)cfgv[8];
        // tables (same construction as build_tables in gw_api_align.inc)
        int32_t k_by_m[GW_MAX_READ_LEN + 1];
        int kmax = 0;
        std::vector<u64> words;
        std::vector<int32_t> base(GW_MAX_READ_LEN + 1, -1);
        std::vector<uint32_t> abortm(GW_MAX_READ_LEN + 1, 0);
        for (int m = 0; m <= GW_MAX_READ_LEN; ++m) {
            int k = m == 0 ? -1 : gw_max_edit_distance(env.cfg.k, m);
            if (k < 0 || k > GW_MAX_K) k = -1;
            k_by_m[m] = k;
            kmax = std::max(kmax, k);
            if (k < 0) continue;
            base[(size_t)m] = (int32_t)words.size();
            for (int kk = 0; kk <= k + 1; ++kk) {
                std::vector<uint64_t> w;
                if (!gw_staircase_masks(m, kk, w, cfgv[9] != 0)) abortm[(size_t)m] |= 1u << kk;
                for (uint64_t x : w) words.push_back((u64)x);
            }
        }
        env.sf.words = words.data(); env.sf.base_by_m = base.data(); env.sf.abort_by_m = abortm.data();
        GwBatchView bv;
        bv.bases = bases; bv.offsets = (const u64*)offsets; bv.k_by_m = k_by_m; bv.ext_pool = nullptr;

        std::vector<std::vector<u64>> slab_store(n_reads);
        std::vector<u64*> slabs(n_reads, nullptr);
        uint32_t max_m = 1;
        for (uint32_t i = 0; i < n_reads; ++i) max_m = std::max<uint32_t>(max_m, (uint32_t)std::min<uint64_t>(offsets[i + 1] - offsets[i], GW_MAX_READ_LEN));
        std::vector<GwU64x2> hist((size_t)(max_m + 2 * kmax + 2) * ((max_m + 63) / 64), GwU64x2{0xDEADBEEFDEADBEEFULL, 0xFEEDFACEFEEDFACEULL});   // poison
        // GWSIM_NOCACHE=1: the lane re-reads its state from the work area before every slice; GWSIM_SPILL=n: the lane is written back
        // and re-read every n slices (as the flow kernel does around its out-of-line continuations). Neither may change any result.
        // GWSIM_PROMOTE="n1,n2": running reads whose work area is nearly full move into one of n1 blocks of 8 x slab_words words and from
        // there into one of n2 blocks of 64 x slab_words words, as in the flow kernel (gw_promote_*); when the blocks run out the reads
        // fall back to the passes with larger work areas. No result may depend on any of this.
        GwExtPool pool;
        memset(&pool, 0, sizeof pool);
        std::vector<u64> pool_mem1, pool_mem2;
        if (const char* pe = getenv("GWSIM_PROMOTE")) {
            unsigned n1 = 0, n2 = 0;
            sscanf(pe, "%u,%u", &n1, &n2);
            pool.block_words = slab_words * 8; pool.n_blocks = n1;
            pool.block_words2 = slab_words * 64; pool.n_blocks2 = n2;
            pool_mem1.assign((size_t)n1 * pool.block_words + 2, 0xDEADBEEFDEADBEEFULL);
            pool_mem2.assign((size_t)n2 * pool.block_words2 + 2, 0xDEADBEEFDEADBEEFULL);
            pool.base = pool_mem1.data(); pool.base2 = pool_mem2.data();
            bv.ext_pool = &pool;
        }
        const bool nocache = getenv("GWSIM_NOCACHE") && atoi(getenv("GWSIM_NOCACHE"));
        const int spill = getenv("GWSIM_SPILL") ? atoi(getenv("GWSIM_SPILL")) : 0;
        auto verify_pending = [&](const GwSlab& s) {   // what a verification batch of the flow kernel does for one parked read
            GwReadCtx* c = s.ctx();
            for (int f = 0; f < c->pend_n; ++f) {
                if (c->pend[f].kind != 2) continue;
                uint32_t steps = 0;
                c->pend[f].textpos = gw_sa_get(env.ix, c->pend[f].fm, c->pend[f].sa_index, &steps);
                c->occ_locate += steps;
            }
            if (c->phase == GW_PH_WAIT_EXACT) return;
            for (int f = 0; f < c->pend_n; ++f) {
                if (c->pend[f].kind != 2) continue;
                GwJob job;
                gw_verify_frag(env, s, &job, f, hist.data(), 1);
            }
        };
        auto drive = [&](uint32_t slot, auto maxh_tag) {   // like one lane of the flow kernel: the read is driven to completion slice by slice
            constexpr int MAXH = decltype(maxh_tag)::value;
            GwSlab s(slabs[slot]);
            for (;;) {
                int r = gw_resume_read(env, s);
                if (r == GW_STEP_DONE) return;
                if (r == GW_STEP_PARKED) { verify_pending(s); continue; }
                GwLane<MAXH> ln;
                gw_lane_load(s, ln);
                int slices = 0;
                while (r == GW_STEP_CONTINUE) {
                    if (nocache && ln.mid == 0) ln.cached = -1;
                    if (spill && ++slices % spill == 0) { gw_lane_flush(s, ln); gw_lane_reload(s, ln); }
                    if (ln.mid == 0 && gw_wants_promotion(s, (uint32_t)ln.top, ln.k, ln.qn)) {   // as the flow kernel does before a slice
                        gw_lane_flush(s, ln);
                        if (u64* nb = gw_promote_take(s)) {
                            gw_promote_copy(s.base, nb, (uint32_t)ln.top, 0u, 1u);
                            gw_promote_commit(s, nb, &slabs[slot]);
                        }
                    }
                    if (ln.mid == 0) r = gw_lane_poll(env, s, ln);
                    if (r == GW_STEP_CONTINUE && ln.mid == GW_MID_TRY)   // as the flow kernel: the specialised form for states that head no split chain
                        r = gw_lane_is_exclusive(ln) ? gw_lane_try<MAXH, true>(env, s, ln) : gw_lane_try<MAXH, false>(env, s, ln);
                    if (r == GW_STEP_CONTINUE && ln.mid == GW_MID_SPLIT) r = gw_lane_split(env, s, ln);
                    if (r == GW_STEP_CONTINUE && ln.mid == GW_MID_REPORT) r = gw_lane_report(env, s, ln);
                }
                gw_lane_flush(s, ln);
                if (r == GW_STEP_DONE) return;
                verify_pending(s);
            }
        };
        auto rounds = [&](const std::vector<uint32_t>& cur) {
            for (uint32_t slot : cur) {
                if (kmax + 1 <= 4) drive(slot, std::integral_constant<int, 4>());
                else if (kmax + 1 <= 8) drive(slot, std::integral_constant<int, 8>());
                else drive(slot, std::integral_constant<int, GW_MAX_ROWS>());
            }
        };
        auto pass = [&](const std::vector<uint32_t>& slots, uint32_t words_per_slab) {
            const GwSlabParams sp = slab_params(words_per_slab);
            for (uint32_t slot : slots) {
                slab_store[slot].assign(words_per_slab, 0xDEADBEEFDEADBEEFULL);   // poison: the slab is not zero-initialised on the device
                slabs[slot] = slab_store[slot].data();
                gw_prep_read(bv, slabs[slot], sp, slot);
                gw_scan_task(env.ix, reinterpret_cast<GwReadCtx*>(slabs[slot]), 0);
                gw_scan_task(env.ix, reinterpret_cast<GwReadCtx*>(slabs[slot]), 1);
            }
            rounds(slots);
        };
        std::vector<uint32_t> all(n_reads);
        for (uint32_t i = 0; i < n_reads; ++i) all[i] = i;
        pass(all, slab_words);
        *tiers_used = 0;
        for (int tier = 1; tier < 3; ++tier) {
            std::vector<uint32_t> over;
            for (uint32_t i = 0; i < n_reads; ++i) if (reinterpret_cast<GwReadCtx*>(slabs[i])->status == GW_ST_OVERFLOW) over.push_back(i);
            if (over.empty()) break;
            *tiers_used += (uint32_t)over.size();
            pass(over, tier == 1 ? slab_words * 8 : 262144);
        }
        // emit + SAM
        std::string text;
        for (uint32_t i = 0; i < n_reads; ++i) {
            uint32_t nrec = 0, ncig = 0;
            gw_emit_count(slabs[i], env.cfg, &nrec, &ncig);
            std::vector<GwHitRec> recs(nrec);
            std::vector<uint32_t> cigs(ncig + 1);
            gw_emit_write(slabs[i], env.cfg, i, recs.data(), 0, nrec, cigs.data(), 0, ncig);
            const GwReadCtx* c = reinterpret_cast<GwReadCtx*>(slabs[i]);
            status[i] = recs.empty() ? -1 : recs[0].status;
            const bool ok = status[i] == GW_ST_OK || status[i] == GW_ST_UNMAPPED;
            counters[i * 4 + 0] = ok ? c->n_fm : -1; counters[i * 4 + 1] = ok ? c->n_sw : -1;
            counters[i * 4 + 2] = ok ? c->n_cut : -1; counters[i * 4 + 3] = ok ? c->n_filt : -1;
            lines[i] = 0;
            if (!ok) continue;
            const std::string name(names + name_off[i], (size_t)(name_off[i + 1] - name_off[i]));
            size_t hidx = 0;
            while (hidx < recs.size()) {
                int n = 1;
                while (hidx + (size_t)n < recs.size() && !recs[hidx + (size_t)n].is_head) ++n;
                const size_t before = text.size();
                gw_sam_one(S->host.contigs, name, bases + offsets[i], (size_t)(offsets[i + 1] - offsets[i]), quals ? quals + qual_off[i] : nullptr,
                           quals ? (size_t)(qual_off[i + 1] - qual_off[i]) : 0, reinterpret_cast<const gw_hit_t*>(&recs[hidx]), std::min(n, 2), cigs.data(), text);
                for (size_t p = before; p < text.size(); ++p) if (text[p] == '\n') lines[i]++;
                hidx += (size_t)n;
            }
        }
        char* o = (char*)malloc(text.size() + 1);
        memcpy(o, text.c_str(), text.size() + 1);
        return o;
    } catch (std::exception& e) { g_err = e.what(); return nullptr; }
}
void gwsim_free(void* p) { free(p); }

// alignBlockDetailed on an explicit pair through the device verifier body
int gwsim_verify_pair(const char* ref, const char* query, int band_k, int* pos, int* nm, uint32_t* cig, int* cig_n) {
    const int Ncols = (int)strlen(ref), m = (int)strlen(query);
    auto qsym = [&](int j) -> int { return gw_code_of((unsigned char)query[j]); };
    auto refsym = [&](int j) -> int { return gw_code_of((unsigned char)ref[j]); };
    u64 peq[4][GW_MAX_NB];
    memset(peq, 0, sizeof(peq));
    for (int j = 0; j < m; ++j) {
        int c = qsym(j);
        if (c == 4) { for (int a = 0; a < 4; ++a) peq[a][j >> 6] |= 1ULL << (j & 63); }
        else peq[c][j >> 6] |= 1ULL << (j & 63);
    }
    std::vector<GwU64x2> hist((size_t)(Ncols + 2) * std::max(1, (m + 63) / 64), GwU64x2{0xDEADBEEFDEADBEEFULL, 0xFEEDFACEFEEDFACEULL});   // poison: device scratch is not zeroed
    GwVerifyOut out;
    gw_verify(m, Ncols, band_k, peq, refsym, qsym, hist.data(), 1, cig, &out);
    *pos = out.pos; *nm = out.diff; *cig_n = out.cig_n;
    return out.valid;
}

}  // extern "C"
