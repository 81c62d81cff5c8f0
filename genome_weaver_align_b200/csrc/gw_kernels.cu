// gw_kernels.cu — the sm_100a kernels of the read-alignment path and their launchers.
//
//   kernel 1   gw_scan_coop_kernel   rank/occ + exact quick scan; 4 lanes share one 64 B occurrence block (one 128-bit
//                                    load per lane), popcounts combined with warp shuffles   (FMQuickScan.java:66-94)
//              gw_scan_kernel        single-thread form of the same scan (kept for A/B measurements)
//   kernels 2-4 gw_flow_kernel       per-read best-first bidirectional search (BidirectionalSuffixFilter.java:278-477), SA locate
//                                    (SparseSuffixArray.java:141-161) and bit-vector DP verification + CIGAR traceback
//                                    (BitParallelSmithWaterman.java:387-644) as warp-wide work of one persistent kernel
//   plus prep (read encoding), emit (hit records) and small utility kernels.
// Every kernel is HBM/L2 random-access or integer-pipe bound (no tensor cores: nothing here is a dense contraction).
#include <cuda_runtime.h>

#include "gw_job.cuh"
#include "gw_kernels.h"
#include <stdlib.h>

#include <algorithm>


// ---------------------------------------------------------------------------------------------------------
__global__ void gw_prep_kernel(GwBatchView bv, u64* const* slabs, const uint32_t* slots, uint32_t n, uint32_t read_base, GwSlabParams sp) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t slot = slots ? slots[i] : i;
    gw_prep_read(bv, slabs[slot], sp, read_base + slot);
}

__global__ void gw_scan_kernel(GwIndexView ix, u64* const* slabs, const uint32_t* slots, uint32_t n_tasks) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_tasks) return;
    uint32_t slot = slots ? slots[t >> 1] : (t >> 1);
    gw_scan_task(ix, reinterpret_cast<GwReadCtx*>(slabs[slot]), (int)(t & 1u));
}

// 4 lanes per (read, strand) task: lane q of the group loads quarter q of the lb block and of the ub block
__global__ void __launch_bounds__(256) gw_scan_coop_kernel(GwIndexView ix, u64* const* slabs, const uint32_t* slots, uint32_t n_tasks) {
    const uint32_t gtid = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t task = gtid >> 2;
    const int q = (int)(gtid & 3u);
    if (task >= n_tasks) return;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t gbase = lane & ~3u;
    const uint32_t gmask = 0xFu << gbase;
    const uint32_t slot = slots ? slots[task >> 1] : (task >> 1);
    const int strand = (int)(task & 1u);
    GwReadCtx* c = reinterpret_cast<GwReadCtx*>(slabs[slot]);
    if (c->phase == GW_PH_DONE) return;
    const int m = c->m;
    const int fm = strand == 0 ? 1 : 0;
    const GwOccBlock* occ = ix.occ[fm];
    const uint32_t N = ix.N;
    const uint32_t C0 = ix.C[fm][0], C1 = ix.C[fm][1], C2 = ix.C[fm][2], C3 = ix.C[fm][3];
    uint32_t lb = 0, ub = N;
    int numMismatches = 0, mark = 0, ls = 0, le = 0;
    bool hasLongest = false;
    u64 wlo = 0, whi = 0;
    for (int i = 0; i < m; ++i) {
        if ((i & 63) == 0) { wlo = c->qlo[strand][i >> 6]; whi = c->qhi[strand][i >> 6]; }
        const int ch = (int)(((whi & 1ULL) << 1) | (wlo & 1ULL));
        wlo >>= 1; whi >>= 1;
        const uint32_t l = lb > N ? N : lb, u = ub > N ? N : ub;
        const GwU4 ql = gw_ld128(&occ[l >> 7].q[q]);
        const GwU4 qu = gw_ld128(&occ[u >> 7].q[q]);
        uint32_t packed = gw_quarter_count1(ql, gw_quarter_mask(l & 127u, q), ch) | (gw_quarter_count1(qu, gw_quarter_mask(u & 127u, q), ch) << 16);
        packed += __shfl_xor_sync(gmask, packed, 1);
        packed += __shfl_xor_sync(gmask, packed, 2);
        const uint32_t cl = __shfl_sync(gmask, ql.x, gbase + ch);
        const uint32_t cu = __shfl_sync(gmask, qu.x, gbase + ch);
        const uint32_t Cc = ch == 0 ? C0 : ch == 1 ? C1 : ch == 2 ? C2 : C3;
        lb = Cc + cl + (packed & 0xFFFFu);
        ub = Cc + cu + (packed >> 16);
        if (lb >= ub) {
            ++numMismatches;
            if (!hasLongest || (le - ls) < (i - mark)) { hasLongest = true; ls = mark; le = i; }
            lb = 0; ub = N;
            mark = i + 1;
        }
    }
    if (!hasLongest || (le - ls) < (m - mark)) { ls = mark; le = m; }
    if (q == 0) {
        c->scan_mm[strand] = numMismatches; c->scan_ls[strand] = ls; c->scan_le[strand] = le;
        c->scan_lb[strand] = lb; c->scan_ub[strand] = ub;
        atomicAdd(&c->occ_scan, 2u * (uint32_t)m);
    }
}

// ---------------------------------------------------------------------------------------------------------
// Flow kernel: warp-level pools without rounds. Every warp owns two small FIFOs in shared memory — reads ready to (re)start
// searching and fragments waiting for locate + verification — and loops over three kinds of warp-wide work:
//   * search slices: one poll + one candidate extension per lane and trip, on the lane's register copy of its read (GwLane); the
//     long, rarely taken continuations (taking a read up / consuming verification results, split + soft-clip children, chain
//     reporting) are deferred until `batch` lanes wait for one (or `age` trips passed, or nothing else can run), so that they run
//     for several lanes at once instead of one lane while 31 wait,
//   * pick-up: idle lanes take reads from the ready FIFO (refilled from the global work list, `cap` reads at a time),
//   * a verification batch as soon as 32 fragments wait: every lane walks one fragment to its SA sample and runs its DP + traceback;
//     the reads whose fragments are all done go back to the ready FIFO.
// There is no per-round tail: a lane that parks or finishes a read takes the next one at once, and locate / DP always run with
// (nearly) full warps.
#ifndef GW_FLOW_THREADS
#define GW_FLOW_THREADS 256
#endif
#define GW_RING 128     // entries per FIFO; a warp owns at most GW_RING reads (refill stops above GW_RING - 32), at most 95 fragments wait
#define GW_WARPS (GW_FLOW_THREADS / 32)

struct GwFlowParams {
    uint32_t cap;        // reads a warp takes from the work list at a time
    uint32_t batch;      // lanes that must wait for a deferred continuation before the warp runs it
    uint32_t age;        // ... or this many trips since the first lane started waiting
    uint32_t flush;      // once the work list is drained, a partial verification batch runs when this many lanes idle
};

// dynamic shared memory of a CTA: the two FIFOs of every warp, then one GwLane record per thread (odd word stride)
extern __shared__ uint32_t gw_flow_smem[];
template <int MAXH>
static size_t gw_flow_smem_bytes() { return ((size_t)2 * GW_WARPS * GW_RING + (size_t)GW_FLOW_THREADS * GW_LANE_WORDS(MAXH)) * 4; }

template <int MAXH>
__global__ void __launch_bounds__(GW_FLOW_THREADS, 2) gw_flow_kernel(GwSearchEnv env, u64* const* slabs, const uint32_t* active, uint32_t n_active,
                                                                  uint32_t* work_counter, GwU64x2* hist, GwFlowParams pp) {
    const uint32_t FULL = 0xFFFFFFFFu, MASK = GW_RING - 1u;
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t* ready = gw_flow_smem + wid * GW_RING;
    uint32_t* frags = gw_flow_smem + (GW_WARPS + wid) * GW_RING;
    GwLane<MAXH>& ln = *reinterpret_cast<GwLane<MAXH>*>(gw_flow_smem + 2 * GW_WARPS * GW_RING + threadIdx.x * GW_LANE_WORDS(MAXH));
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    GwU64x2* my_hist = hist + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t r_head = 0, r_tail = 0, f_head = 0, f_tail = 0, age_pick = 0, age_cold = 0, owned = 0;
    bool drained = false;
    bool have = false;
    GwSlab s;
    ln.mid = 0; ln.top = 0; ln.k = 0;
    uint32_t slot = 0;
    for (;;) {
        const uint32_t busy = __ballot_sync(FULL, have);
        const uint32_t idle = ~busy;
        const uint32_t n_ready = r_tail - r_head, n_fr = f_tail - f_head;
        // ---- verification batch
        if (n_fr >= 32u || (n_fr != 0u && n_ready == 0u && drained && (busy == 0u || (uint32_t)__popc(idle) >= pp.flush))) {
            uint32_t nb = n_fr < 32u ? n_fr : 32u;
            if (n_fr > 32u && (frags[(f_head + 31u) & MASK] >> 1) == (frags[(f_head + 32u) & MASK] >> 1)) nb = 31u;   // both fragments of a read in one batch
            const uint32_t id = lane < nb ? frags[(f_head + lane) & MASK] : 0xFFFFFFFFu;
            f_head += nb;
            GwReadCtx* c = nullptr;
            if (lane < nb) {   // locate
                c = reinterpret_cast<GwReadCtx*>(slabs[id >> 1]);
                GwPendFrag* pf = &c->pend[id & 1u];
                uint32_t j = 0;
                pf->textpos = gw_sa_get(env.ix, pf->fm, pf->sa_index, &j);
                if (j) atomicAdd(&c->occ_locate, j);   // the two fragments of a read are walked by different lanes
            }
            __syncwarp();
            {   // DP + traceback; the lanes take the DP instantiation of the largest block count among them
                GwJob job;
                GwSlab sv;
                bool dp = false;
                if (lane < nb && c->phase != GW_PH_WAIT_EXACT) {   // exact hit: no verification (:490-494)
                    sv = GwSlab(slabs[id >> 1]);
                    dp = gw_job_begin_dp(env, sv, &job, (int)(id & 1u), my_hist, stride);
                }
                const int hint = (int)__reduce_max_sync(FULL, dp ? (unsigned)job.dp.bmax : 0u);
                if (dp) gw_job_columns(env, &job, job.dp.ncols, my_hist, stride, hint);
                __syncwarp();
                if (dp) gw_job_trace(sv, &job, my_hist, stride);
            }
            __syncwarp();
            const uint32_t nid = __shfl_down_sync(FULL, id, 1);
            const bool last = lane < nb && (lane == nb - 1u || (nid >> 1) != (id >> 1));
            const uint32_t bl = __ballot_sync(FULL, last);
            if (last) ready[(r_tail + (uint32_t)__popc(bl & lt)) & MASK] = id >> 1;
            r_tail += (uint32_t)__popc(bl);
            __syncwarp();
            continue;
        }
        // ---- pick-up (idle lanes take reads from the ready FIFO; refill from the global work list)
        if (idle != 0u && (n_ready != 0u || !drained)) {
            if ((uint32_t)__popc(idle) >= pp.batch || busy == 0u || ++age_pick >= pp.age) {
                age_pick = 0;
                uint32_t nrd = n_ready;
                if (nrd < (uint32_t)__popc(idle) && !drained && owned + pp.cap <= GW_RING) {
                    uint32_t first = 0;
                    if (lane == 0) first = atomicAdd(work_counter, pp.cap);
                    first = __shfl_sync(FULL, first, 0);
                    uint32_t got = 0;
                    if (first < n_active) got = min(pp.cap, n_active - first);
                    if (first + pp.cap >= n_active) drained = true;
                    if (lane < got) ready[(r_tail + lane) & MASK] = active ? active[first + lane] : first + lane;
                    r_tail += got; nrd += got; owned += got;
                    __syncwarp();
                }
                int r0 = -1;
                if (!have) {
                    const uint32_t i = (uint32_t)__popc(idle & lt);
                    if (i < nrd) {
                        slot = ready[(r_head + i) & MASK];
                        s = GwSlab(slabs[slot]);
                        r0 = gw_resume_read(env, s);
                        have = r0 == GW_STEP_CONTINUE;
                        if (have) gw_lane_load(s, ln);
                    }
                }
                r_head += min((uint32_t)__popc(idle), nrd);
                const uint32_t pk = __ballot_sync(FULL, r0 == GW_STEP_PARKED);   // exact hit: needs its locate only
                if (r0 == GW_STEP_PARKED) frags[(f_tail + (uint32_t)__popc(pk & lt)) & MASK] = slot * 2u;
                f_tail += (uint32_t)__popc(pk);
                owned -= (uint32_t)__popc(__ballot_sync(FULL, r0 == GW_STEP_DONE));
                __syncwarp();
                continue;
            }
        } else if (busy == 0u) {
            if (n_fr == 0u && n_ready == 0u && drained) break;
        }
        // ---- one search slice per lane
        int r = GW_STEP_CONTINUE;
        {   // work areas that are nearly full: the reads go on in a larger one; the lanes of the warp share each copy
            const bool wantp = have && ln.mid == 0 && gw_wants_promotion(s, (uint32_t)ln.top, ln.k, ln.qn);
            if (__ballot_sync(FULL, wantp) != 0u) {
                if (wantp) gw_lane_flush(s, ln);   // the copy takes the work area as it is in HBM
                u64* nb = wantp ? gw_promote_take(s) : nullptr;
                uint32_t pm = __ballot_sync(FULL, nb != nullptr);
                const uint32_t my_top = nb ? (uint32_t)ln.top : 0u;
                while (pm) {
                    const int src = __ffs(pm) - 1;
                    pm &= pm - 1u;
                    const u64* from = reinterpret_cast<const u64*>(__shfl_sync(FULL, (unsigned long long)s.base, src));
                    u64* to = reinterpret_cast<u64*>(__shfl_sync(FULL, (unsigned long long)nb, src));
                    const uint32_t top = __shfl_sync(FULL, my_top, src);
                    gw_promote_copy(from, to, top, lane, 32u);
                }
                __syncwarp();
                if (nb) gw_promote_commit(s, nb, const_cast<u64**>(slabs) + slot);
            }
        }
        // the lanes are brought together before each part: the compiler lets a lane that skips the poll run ahead into the extension
        // on its own, and the warp would then pass through the same code twice with part of its lanes each time
        __syncwarp();
        if (have && ln.mid == 0) r = gw_lane_poll(env, s, ln);
        __syncwarp();
        if (have && r == GW_STEP_CONTINUE && ln.mid == GW_MID_TRY) r = gw_lane_try<MAXH, false>(env, s, ln);
        __syncwarp();
        // ---- deferred continuations
        {
            const bool wants = have && r == GW_STEP_CONTINUE && ln.mid >= GW_MID_SPLIT;
            const uint32_t cold = __ballot_sync(FULL, wants);
            if (cold != 0u) {
                const uint32_t hot = __ballot_sync(FULL, have && r == GW_STEP_CONTINUE && ln.mid < GW_MID_SPLIT);
                if ((uint32_t)__popc(cold) >= pp.batch || hot == 0u || ++age_cold >= pp.age) {
                    age_cold = 0;
                    if (wants) r = ln.mid == GW_MID_SPLIT ? gw_lane_split(env, s, ln) : gw_lane_report(env, s, ln);
                }
            }
        }
        // ---- parked reads queue their fragments (both fragments of a read next to each other); finished reads just leave
        bool f0 = false, f1 = false;
        if (have && r != GW_STEP_CONTINUE) {
            gw_lane_flush(s, ln);
            if (r == GW_STEP_PARKED) {
                const GwReadCtx* c = s.ctx();
                f0 = c->pend_n > 0 && c->pend[0].kind == 2;
                f1 = c->pend_n > 1 && c->pend[1].kind == 2;
            }
        }
        const uint32_t b_f0 = __ballot_sync(FULL, f0), b_f1 = __ballot_sync(FULL, f1);
        if (b_f0 | b_f1) {
            const uint32_t o = f_tail + (uint32_t)__popc(b_f0 & lt) + (uint32_t)__popc(b_f1 & lt);
            if (f0) frags[o & MASK] = slot * 2u;
            if (f1) frags[(o + (f0 ? 1u : 0u)) & MASK] = slot * 2u + 1u;
            f_tail += (uint32_t)__popc(b_f0) + (uint32_t)__popc(b_f1);
            __syncwarp();
        }
        owned -= (uint32_t)__popc(__ballot_sync(FULL, have && r == GW_STEP_DONE));
        if (have && r != GW_STEP_CONTINUE) have = false;
    }
}

// Densifies the locate table of one index. The LF walk that starts at a sampled row (a multiple of GW_SSA_L) passes through every row
// whose own walk ends at the NEXT sampled row it reaches — the rows SparseSuffixArray.get resolves with that sample
// (SparseSuffixArray.java:141-161): a row i steps into the walk, len - i steps before its end, gets (sample_end + len - i) % N. Since
// one LF step moves one position back in the text that is sample_start - i, and the walk checks it when it arrives: sample_start ==
// (sample_end + len) % N. (It always holds for the suffix array of a primitive text; should it not, the walk is done a second time
// with the reference's own formula.) The walks have geometric lengths, so the lanes do not own one walk each: every lane takes the
// next sampled rows from a shared counter as soon as its walk ends, and the warp's loop body is ONE LF step for all of its lanes.
__global__ void __launch_bounds__(256) gw_densify_kernel(GwIndexView ix, int fm, uint32_t step, uint32_t* table, uint32_t n_samples, uint32_t* counter) {
    const uint32_t CHUNK = 8;
    const u64 N = ix.N;
    uint32_t next = 0, chunk_end = 0;
    uint32_t cursor = 0, start = 0, len = 0, redo_len = 0, redo_smp = 0;
    u64 pos = 0;
    int mode = 0;   // 0: no walk, 1: walking, 2: walking again with the end sample known
    for (;;) {
        if (mode == 0) {
            if (next == chunk_end) { next = atomicAdd(counter, CHUNK); chunk_end = next + CHUNK; }
            if (next >= n_samples) break;
            start = next++ * GW_SSA_L;
            pos = gw_ld32(ix.ssa[fm] + start / GW_SSA_L);
            table[start / step] = (uint32_t)pos;
            cursor = start; len = 0; mode = 1;
        }
        cursor = gw_suffix_link(ix, fm, cursor);
        ++len;
        if ((cursor % GW_SSA_L) == 0) {   // the end of the walk
            const uint32_t smp = gw_ld32(ix.ssa[fm] + cursor / GW_SSA_L);
            if (mode == 1 && (u64)gw_ld32(ix.ssa[fm] + start / GW_SSA_L) != ((u64)smp + len) % N) {   // not a plain text walk: once more, the reference's way
                redo_len = len; redo_smp = smp; cursor = start; len = 0; mode = 2;
            } else
                mode = 0;
            continue;
        }
        if ((cursor & (step - 1u)) == 0) {
            if (mode == 1) { pos = pos == 0 ? N - 1 : pos - 1; table[cursor / step] = (uint32_t)pos; }
            else table[cursor / step] = (uint32_t)(((u64)redo_smp + (u64)(redo_len - len) + N) % N);
        } else if (mode == 1)
            pos = pos == 0 ? N - 1 : pos - 1;
    }
}

// collects the slots whose slab overflowed (rerun with the large slab)
__global__ void gw_collect_overflow_kernel(u64* const* slabs, uint32_t n, uint32_t* list, uint32_t* count) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const GwReadCtx* c = reinterpret_cast<const GwReadCtx*>(slabs[i]);
    if (c->status == GW_ST_OVERFLOW) list[atomicAdd(count, 1u)] = i;
}

// ---- work order of a wave ---------------------------------------------------------------------------------
// The flow kernel hands reads to its warps 32 at a time in the order of the work list. Ordered by the quick scan's verdict — the reads
// with the most mismatching stretches first, reads with an exact hit last — the 32 reads a warp takes up need about the same amount of
// search (its lanes stay in step), and the kernel ends on the cheapest reads instead of waiting for a few expensive ones. The order
// changes no result: reads are independent and every output is addressed by read. Stable counting sort over GW_ORDER_CLASSES classes:
// per-CTA class counts, one scan over (class, CTA), stable scatter.
#define GW_ORDER_CLASSES 32
#define GW_ORDER_T 256
// Measured on the bench configs (tools/order_study.py): a searched read costs about one FM search per base whatever its mismatch count
// (98 / 101 / 111 searches for 1 / 2 / 3 mismatching stretches at 100 bp), so the length decides, then the mismatches; a read with an
// exact hit costs one locate. Classes: 8 length groups of 32 bases (longest first) x 3 mismatch groups, then the reads that need no search.
GW_HD int gw_order_class(const GwReadCtx* c) {
    if (c->phase == GW_PH_DONE) return 31;                      // nothing to do (empty / unsupported read)
    const int mm = gw_min(c->scan_mm[0], c->scan_mm[1]);
    if (mm == 0) return 30;                                     // exact hit: one locate
    if (mm > c->k) return 29;                                   // neither strand can be within k: finishes when taken up
    const int len = 7 - gw_min(7, (c->m - 1) >> 5);             // 0: 225..256 bases ... 7: 1..32
    return len * 3 + (mm >= 3 ? 0 : 3 - mm);
}
// ranks of the CTA's reads within their class (stable): warp part by match_any, CTA part by a table of per-warp counts
__device__ __forceinline__ uint32_t gw_order_rank(int cls, bool valid, uint32_t (*wcnt)[GW_ORDER_CLASSES], uint32_t* cta_total_of_class) {
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    const uint32_t same = __match_any_sync(0xFFFFFFFFu, valid ? cls : GW_ORDER_CLASSES);
    const uint32_t in_warp = (uint32_t)__popc(same & ((1u << lane) - 1u));
    if (threadIdx.x < (GW_ORDER_T / 32) * GW_ORDER_CLASSES) wcnt[threadIdx.x / GW_ORDER_CLASSES][threadIdx.x % GW_ORDER_CLASSES] = 0;
    __syncthreads();
    if (valid && in_warp == 0) wcnt[wid][cls] = (uint32_t)__popc(same);
    __syncthreads();
    uint32_t before = 0, total = 0;
    if (valid) {
        for (uint32_t w = 0; w < GW_ORDER_T / 32; ++w) { const uint32_t v = wcnt[w][cls]; if (w < wid) before += v; total += v; }
    }
    *cta_total_of_class = total;
    return before + in_warp;
}
__global__ void __launch_bounds__(GW_ORDER_T) gw_order_count_kernel(u64* const* slabs, uint32_t n, uint8_t* cls_out, uint32_t* counts /* [class][CTA] */) {
    __shared__ uint32_t wcnt[GW_ORDER_T / 32][GW_ORDER_CLASSES];
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = i < n;
    const int cls = valid ? gw_order_class(reinterpret_cast<const GwReadCtx*>(slabs[i])) : 0;
    if (valid) cls_out[i] = (uint8_t)cls;
    uint32_t total = 0;
    const uint32_t rank = gw_order_rank(cls, valid, wcnt, &total);
    if (valid && rank == 0) counts[(size_t)cls * gridDim.x + blockIdx.x] = total;
}
// exclusive scan of `len` counts in place (one CTA; len = classes x CTAs, tens of thousands at most)
__global__ void __launch_bounds__(1024) gw_order_scan_kernel(uint32_t* counts, uint32_t len) {
    __shared__ uint32_t sh[1024];
    const uint32_t t = threadIdx.x, chunk = (len + 1023u) / 1024u;
    const uint32_t lo = min(len, t * chunk), hi = min(len, lo + chunk);
    uint32_t sum = 0;
    for (uint32_t i = lo; i < hi; ++i) sum += counts[i];
    sh[t] = sum;
    __syncthreads();
    for (uint32_t d = 1; d < 1024; d <<= 1) {
        const uint32_t y = t >= d ? sh[t - d] : 0;
        __syncthreads();
        sh[t] += y;
        __syncthreads();
    }
    uint32_t run = sh[t] - sum;
    for (uint32_t i = lo; i < hi; ++i) { const uint32_t v = counts[i]; counts[i] = run; run += v; }
}
__global__ void __launch_bounds__(GW_ORDER_T) gw_order_scatter_kernel(const uint8_t* cls_in, uint32_t n, const uint32_t* offsets /* [class][CTA] */, uint32_t* order) {
    __shared__ uint32_t wcnt[GW_ORDER_T / 32][GW_ORDER_CLASSES];
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = i < n;
    const int cls = valid ? (int)cls_in[i] : 0;
    uint32_t total = 0;
    const uint32_t rank = gw_order_rank(cls, valid, wcnt, &total);
    if (valid) order[offsets[(size_t)cls * gridDim.x + blockIdx.x] + rank] = i;
}

__global__ void gw_settle_kernel(u64** slabs, u64* small_base, size_t small_words, GwSlabParams small_sp, const uint32_t* list, uint32_t n, GwDevConfig cfg) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t slot = list[i];
    u64* sm = small_base + (size_t)slot * small_words;
    gw_settle(slabs[slot], sm, small_sp, cfg);
    slabs[slot] = sm;
}
__global__ void gw_emit_count_kernel(u64* const* slabs, uint32_t n, GwDevConfig cfg, uint32_t* n_rec, uint32_t* n_cig, unsigned long long* totals,
                                     uint32_t* read_ctr) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long v[7] = {0, 0, 0, 0, 0, 0, 0};
    if (i < n) {
        gw_emit_count(slabs[i], cfg, &n_rec[i], &n_cig[i]);
        const GwReadCtx* c = reinterpret_cast<const GwReadCtx*>(slabs[i]);
        v[0] = (unsigned long long)c->n_fm; v[1] = (unsigned long long)c->n_sw; v[2] = (unsigned long long)c->n_cut; v[3] = (unsigned long long)c->n_filt;
        v[4] = c->occ_scan; v[5] = 2u * (uint32_t)c->n_fm; v[6] = c->occ_locate;
        // the reference's per-read debug counters (BidirectionalSuffixFilter.java:588-592), kept for gw_aligner_read_counters
        if (read_ctr) reinterpret_cast<uint4*>(read_ctr)[i] = make_uint4((uint32_t)c->n_fm, (uint32_t)c->n_sw, (uint32_t)c->n_cut, (uint32_t)c->n_filt);
    }
#pragma unroll
    for (int j = 0; j < 7; ++j) {
        unsigned long long x = v[j];
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xFFFFFFFFu, x, o);
        if ((threadIdx.x & 31u) == 0 && x) atomicAdd(&totals[j], x);
    }
}

// exclusive scan of two u32 arrays into u64 offsets (+ base), totals to out_tot[0..1], in three launches: per-block sums of chunks
// of `chunk` elements (a multiple of the block size, at most 1024 chunks), a single-block scan of those sums, and the scan inside
// every chunk. All loads and stores are coalesced.
#define GW_SCAN_T 256
__device__ __forceinline__ void gw_block_sum2(unsigned long long& xa, unsigned long long& xb, unsigned long long* sh) {
    for (int d = 16; d > 0; d >>= 1) { xa += __shfl_xor_sync(0xFFFFFFFFu, xa, d); xb += __shfl_xor_sync(0xFFFFFFFFu, xb, d); }
    const uint32_t w = threadIdx.x >> 5, l = threadIdx.x & 31u;
    if (l == 0) { sh[w] = xa; sh[8 + w] = xb; }
    __syncthreads();
    xa = 0; xb = 0;
    for (uint32_t i = 0; i < GW_SCAN_T / 32; ++i) { xa += sh[i]; xb += sh[8 + i]; }
}
__global__ void __launch_bounds__(GW_SCAN_T) gw_scan2_partial_kernel(const uint32_t* a, const uint32_t* b, uint32_t n, uint32_t chunk,
                                                                      unsigned long long* pa, unsigned long long* pb) {
    __shared__ unsigned long long sh[16];
    const uint32_t lo = blockIdx.x * chunk, hi = min(n, lo + chunk);
    unsigned long long xa = 0, xb = 0;
    for (uint32_t i = lo + threadIdx.x; i < hi; i += GW_SCAN_T) { xa += a[i]; xb += b[i]; }
    gw_block_sum2(xa, xb, sh);
    if (threadIdx.x == 0) { pa[blockIdx.x] = xa; pb[blockIdx.x] = xb; }
}
__global__ void __launch_bounds__(1024) gw_scan2_offsets_kernel(unsigned long long* pa, unsigned long long* pb, uint32_t nb, unsigned long long base_a,
                                                                 unsigned long long base_b, unsigned long long* out_tot) {
    __shared__ unsigned long long sa[1024], sb[1024];
    const uint32_t t = threadIdx.x;
    const unsigned long long xa = t < nb ? pa[t] : 0, xb = t < nb ? pb[t] : 0;
    sa[t] = xa; sb[t] = xb;
    __syncthreads();
    for (uint32_t d = 1; d < 1024; d <<= 1) {
        const unsigned long long ya = t >= d ? sa[t - d] : 0, yb = t >= d ? sb[t - d] : 0;
        __syncthreads();
        sa[t] += ya; sb[t] += yb;
        __syncthreads();
    }
    if (t < nb) { pa[t] = base_a + sa[t] - xa; pb[t] = base_b + sb[t] - xb; }
    if (t == 1023) { out_tot[0] = sa[t]; out_tot[1] = sb[t]; }
}
__global__ void __launch_bounds__(GW_SCAN_T) gw_scan2_final_kernel(const uint32_t* a, const uint32_t* b, uint32_t n, uint32_t chunk,
                                                                    const unsigned long long* pa, const unsigned long long* pb,
                                                                    unsigned long long* oa, unsigned long long* ob) {
    __shared__ unsigned long long sa[GW_SCAN_T], sb[GW_SCAN_T];
    const uint32_t t = threadIdx.x, per = chunk / GW_SCAN_T;
    const uint32_t lo = min(n, blockIdx.x * chunk + t * per), hi = min(n, lo + per);
    unsigned long long xa = 0, xb = 0;
    for (uint32_t i = lo; i < hi; ++i) { xa += a[i]; xb += b[i]; }
    sa[t] = xa; sb[t] = xb;
    __syncthreads();
    for (uint32_t d = 1; d < GW_SCAN_T; d <<= 1) {
        const unsigned long long ya = t >= d ? sa[t - d] : 0, yb = t >= d ? sb[t - d] : 0;
        __syncthreads();
        sa[t] += ya; sb[t] += yb;
        __syncthreads();
    }
    unsigned long long qa = pa[blockIdx.x] + sa[t] - xa, qb = pb[blockIdx.x] + sb[t] - xb;
    for (uint32_t i = lo; i < hi; ++i) { oa[i] = qa; ob[i] = qb; qa += a[i]; qb += b[i]; }
}

__global__ void gw_emit_write_kernel(u64* const* slabs, uint32_t n, GwDevConfig cfg, uint32_t read_base, const unsigned long long* rec_off,
                                     const unsigned long long* cig_off, GwHitRec* recs, unsigned long long rec_cap, uint32_t* cigs,
                                     unsigned long long cig_cap) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    gw_emit_write(slabs[i], cfg, read_base + i, recs, rec_off[i], rec_cap, cigs, cig_off[i], cig_cap);
}

__global__ void gw_fill_slab_ptrs_kernel(u64** slabs, u64* base, size_t slab_words, const uint32_t* slots, uint32_t n) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t slot = slots ? slots[i] : i;
    slabs[slot] = base + (size_t)i * slab_words;
}

// ---- component kernels (parity tests of single reference methods) ------------------------------------------
__global__ void gw_rank_acgtn_kernel(GwIndexView ix, int which, const uint32_t* pos, uint32_t n, uint32_t* out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t p = gw_clampN(ix, pos[i]);
    GwBlk b = gw_load_block(ix.occ[which] + (p >> 7));
    uint32_t r[4];
    gw_rank4(b, p & 127u, r);
    out[i * 5 + 0] = r[0]; out[i * 5 + 1] = r[1]; out[i * 5 + 2] = r[2]; out[i * 5 + 3] = r[3];
    out[i * 5 + 4] = gw_rankN(b, p & ~127u, p & 127u);
}

__global__ void gw_locate_rows_kernel(GwIndexView ix, int which, const uint32_t* rows, uint32_t n, uint32_t* out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t steps;
    out[i] = gw_sa_get(ix, which, rows[i], &steps);
}

__global__ void gw_scan_export_kernel(u64* const* slabs, uint32_t n, int strand, uint32_t* out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const GwReadCtx* c = reinterpret_cast<const GwReadCtx*>(slabs[i]);
    out[i * 5 + 0] = (uint32_t)c->scan_mm[strand]; out[i * 5 + 1] = (uint32_t)c->scan_ls[strand]; out[i * 5 + 2] = (uint32_t)c->scan_le[strand];
    out[i * 5 + 3] = c->scan_lb[strand]; out[i * 5 + 4] = c->scan_ub[strand];
}

// alignBlockDetailed over explicit (ref, query) ASCII pairs
__global__ void gw_verify_pairs_kernel(const char* refs, const u64* ref_off, const char* queries, const u64* query_off, uint32_t n, int band_k,
                                       int32_t* result, uint32_t* cigar_ops, uint32_t cigar_stride, GwU64x2* hist, uint32_t max_m) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const char* r = refs + ref_off[i];
    const char* qs = queries + query_off[i];
    const int Ncols = (int)(ref_off[i + 1] - ref_off[i]);
    const int m = (int)(query_off[i + 1] - query_off[i]);
    auto qsym = [&](int j) -> int { return gw_code_of((unsigned char)qs[j]); };
    auto refsym = [&](int j) -> int { return gw_code_of((unsigned char)r[j]); };
    u64 peq[4][GW_MAX_NB];
    for (int a = 0; a < 4; ++a) for (int b = 0; b < GW_MAX_NB; ++b) peq[a][b] = 0;
    for (int j = 0; j < m; ++j) {
        int cde = qsym(j);
        if (cde == 4) { for (int a = 0; a < 4; ++a) peq[a][j >> 6] |= 1ULL << (j & 63); }  // QueryMask.java:50-55: N matches every base
        else peq[cde][j >> 6] |= 1ULL << (j & 63);
    }
    uint32_t cig[GW_CIG_CAP];
    GwVerifyOut out;
    gw_verify(m, Ncols, band_k, peq, refsym, qsym, hist + i, (size_t)n, cig, &out);
    result[i * 4 + 0] = out.valid; result[i * 4 + 1] = out.pos; result[i * 4 + 2] = out.diff;
    result[i * 4 + 3] = out.cig_overflow ? -1 : out.cig_n;
    for (int j = 0; j < out.cig_n && j < (int)cigar_stride; ++j) cigar_ops[(size_t)i * cigar_stride + j] = cig[j];
}

// ---------------------------------------------------------------------------------------------------------
static inline unsigned gw_grid(uint64_t n, unsigned block) { return (unsigned)((n + block - 1) / block); }


void gwk_fill_slab_ptrs(cudaStream_t st, u64** slabs, u64* base, size_t slab_words, const uint32_t* slots, uint32_t n) {
    if (n) gw_fill_slab_ptrs_kernel<<<gw_grid(n, 256), 256, 0, st>>>(slabs, base, slab_words, slots, n);
}
void gwk_prep(cudaStream_t st, const GwBatchView& bv, u64* const* slabs, const uint32_t* slots, uint32_t n, uint32_t read_base, const GwSlabParams& sp) {
    if (n) gw_prep_kernel<<<gw_grid(n, 128), 128, 0, st>>>(bv, slabs, slots, n, read_base, sp);
}
void gwk_scan(cudaStream_t st, const GwIndexView& ix, u64* const* slabs, const uint32_t* slots, uint32_t n_slots, int coop) {
    if (!n_slots) return;
    const uint64_t n_tasks = 2ull * n_slots;
    if (coop) gw_scan_coop_kernel<<<gw_grid(n_tasks * 4, 256), 256, 0, st>>>(ix, slabs, slots, (uint32_t)n_tasks);
    else gw_scan_kernel<<<gw_grid(n_tasks, 128), 128, 0, st>>>(ix, slabs, slots, (uint32_t)n_tasks);
}
template <int MAXH>
static void gw_flow_attr() {
    // the lane records take most of the shared memory the two CTAs of an SM ask for; the rest of the SM's 256 KB is L1 for the work areas
    const size_t bytes = gw_flow_smem_bytes<MAXH>();
    cudaFuncSetAttribute(gw_flow_kernel<MAXH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    const int pct = (int)((bytes + 1024) * 2 * 100 / (228 * 1024)) + 1;
    cudaFuncSetAttribute(gw_flow_kernel<MAXH>, cudaFuncAttributePreferredSharedMemoryCarveout, pct > 100 ? 100 : pct);
}
void gwk_flow_setup(void) {   // function attributes are per device: called once per aligner (after cudaSetDevice)
    gw_flow_attr<4>(); gw_flow_attr<8>(); gw_flow_attr<GW_MAX_ROWS>();
}
void gwk_flow(cudaStream_t st, const GwSearchEnv& env, u64* const* slabs, const uint32_t* active, uint32_t n_active, uint32_t* work_counter, void* hist,
              uint32_t fetch, uint32_t batch, uint32_t age, uint32_t flush, uint32_t blocks, int maxh) {
    if (!n_active) return;
    GwFlowParams pp;
    pp.cap = fetch < 1 ? 1 : (fetch > 32 ? 32 : fetch);
    pp.batch = batch < 1 ? 1 : batch; pp.age = age < 1 ? 1 : age; pp.flush = flush < 1 ? 1 : flush;
    GwU64x2* h = (GwU64x2*)hist;
    if (maxh <= 4) gw_flow_kernel<4><<<blocks, GW_FLOW_THREADS, gw_flow_smem_bytes<4>(), st>>>(env, slabs, active, n_active, work_counter, h, pp);
    else if (maxh <= 8) gw_flow_kernel<8><<<blocks, GW_FLOW_THREADS, gw_flow_smem_bytes<8>(), st>>>(env, slabs, active, n_active, work_counter, h, pp);
    else gw_flow_kernel<GW_MAX_ROWS><<<blocks, GW_FLOW_THREADS, gw_flow_smem_bytes<GW_MAX_ROWS>(), st>>>(env, slabs, active, n_active, work_counter, h, pp);
}
void gwk_densify(cudaStream_t st, const GwIndexView& ix, int fm, uint32_t step, uint32_t* table, uint32_t n_samples, uint32_t* counter) {
    if (!n_samples) return;
    const uint64_t want = ((uint64_t)n_samples + 255) / 256;   // persistent lanes: 8 CTAs of 256 per SM are plenty
    gw_densify_kernel<<<(unsigned)std::min<uint64_t>(want, 148 * 8), 256, 0, st>>>(ix, fm, step, table, n_samples, counter);
}
void gwk_collect_overflow(cudaStream_t st, u64* const* slabs, uint32_t n, uint32_t* list, uint32_t* count) {
    if (n) gw_collect_overflow_kernel<<<gw_grid(n, 256), 256, 0, st>>>(slabs, n, list, count);
}
// order[0..n) = the slots 0..n-1 sorted by gw_order_class (stable); cls: n bytes, counts: GW_ORDER_CLASSES x ceil(n / 256) words of scratch
size_t gwk_order_scratch_words(uint32_t n) { return (size_t)GW_ORDER_CLASSES * gw_grid(n, GW_ORDER_T); }
void gwk_order(cudaStream_t st, u64* const* slabs, uint32_t n, uint8_t* cls, uint32_t* counts, uint32_t* order) {
    if (!n) return;
    const unsigned blocks = gw_grid(n, GW_ORDER_T);
    cudaMemsetAsync(counts, 0, (size_t)GW_ORDER_CLASSES * blocks * 4, st);
    gw_order_count_kernel<<<blocks, GW_ORDER_T, 0, st>>>(slabs, n, cls, counts);
    gw_order_scan_kernel<<<1, 1024, 0, st>>>(counts, GW_ORDER_CLASSES * blocks);
    gw_order_scatter_kernel<<<blocks, GW_ORDER_T, 0, st>>>(cls, n, counts, order);
}
void gwk_settle(cudaStream_t st, u64** slabs, u64* small_base, size_t small_words, const GwSlabParams& small_sp, const uint32_t* list, uint32_t n, const GwDevConfig& cfg) {
    if (n) gw_settle_kernel<<<gw_grid(n, 128), 128, 0, st>>>(slabs, small_base, small_words, small_sp, list, n, cfg);
}
void gwk_emit_count(cudaStream_t st, u64* const* slabs, uint32_t n, const GwDevConfig& cfg, uint32_t* n_rec, uint32_t* n_cig, unsigned long long* totals,
                    uint32_t* read_ctr) {
    if (n) gw_emit_count_kernel<<<gw_grid(n, 256), 256, 0, st>>>(slabs, n, cfg, n_rec, n_cig, totals, read_ctr);
}
void gwk_scan2(cudaStream_t st, const uint32_t* a, const uint32_t* b, uint32_t n, unsigned long long base_a, unsigned long long base_b,
               unsigned long long* oa, unsigned long long* ob, unsigned long long* out_tot, unsigned long long* part) {
    if (!n) { gw_scan2_offsets_kernel<<<1, 1024, 0, st>>>(part, part + 1024, 0, base_a, base_b, out_tot); return; }
    uint32_t chunk = (n + 1023) / 1024;
    chunk = (chunk + GW_SCAN_T - 1) / GW_SCAN_T * GW_SCAN_T;
    const uint32_t nb = (n + chunk - 1) / chunk;
    gw_scan2_partial_kernel<<<nb, GW_SCAN_T, 0, st>>>(a, b, n, chunk, part, part + 1024);
    gw_scan2_offsets_kernel<<<1, 1024, 0, st>>>(part, part + 1024, nb, base_a, base_b, out_tot);
    gw_scan2_final_kernel<<<nb, GW_SCAN_T, 0, st>>>(a, b, n, chunk, part, part + 1024, oa, ob);
}
void gwk_emit_write(cudaStream_t st, u64* const* slabs, uint32_t n, const GwDevConfig& cfg, uint32_t read_base, const unsigned long long* rec_off,
                    const unsigned long long* cig_off, void* recs, unsigned long long rec_cap, uint32_t* cigs, unsigned long long cig_cap) {
    if (n) gw_emit_write_kernel<<<gw_grid(n, 128), 128, 0, st>>>(slabs, n, cfg, read_base, rec_off, cig_off, (GwHitRec*)recs, rec_cap, cigs, cig_cap);
}
// Measurement aid (SURVEY.md 8d "random-access roofline"): every thread does `per_thread` INDEPENDENT rank queries at pseudo-random
// BWT positions (one 64 B occurrence block each, four 128-bit loads, all issued before the first use). This is the rate the memory
// system delivers for the access pattern of an FM step when nothing serialises the lookups; the search kernels, whose lookups
// depend on each other through the automaton, are reported against it as well as against the streaming peak.
__global__ void __launch_bounds__(256) gw_gather_bench_kernel(GwIndexView ix, int which, uint32_t per_thread, uint32_t seed, uint32_t* sink) {
    const uint32_t gtid = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t nblk = (uint32_t)((ix.N + 127) >> 7);
    uint32_t x = (gtid + 1u) * 2654435761u ^ seed, acc = 0;
    for (uint32_t j = 0; j < per_thread; j += 4) {
        GwBlk b[4];
        uint32_t r[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            x ^= x << 13; x ^= x >> 17; x ^= x << 5;   // xorshift32
            const uint32_t blk = (uint32_t)(((u64)x * nblk) >> 32);
            r[u] = x & 127u;
            b[u] = gw_load_block(ix.occ[which] + blk);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) { uint32_t o[4]; gw_rank4(b[u], r[u], o); acc += o[0] ^ o[1] ^ o[2] ^ o[3]; }
    }
    if (acc == 0x9E3779B9u) sink[0] = acc;   // keeps the loads alive
}
void gwk_gather_bench(cudaStream_t st, const GwIndexView& ix, int which, uint32_t blocks, uint32_t per_thread, uint32_t seed, uint32_t* sink) {
    gw_gather_bench_kernel<<<blocks, 256, 0, st>>>(ix, which, per_thread, seed, sink);
}
void gwk_rank_acgtn(cudaStream_t st, const GwIndexView& ix, int which, const uint32_t* pos, uint32_t n, uint32_t* out) {
    if (n) gw_rank_acgtn_kernel<<<gw_grid(n, 128), 128, 0, st>>>(ix, which, pos, n, out);
}
void gwk_locate_rows(cudaStream_t st, const GwIndexView& ix, int which, const uint32_t* rows, uint32_t n, uint32_t* out) {
    if (n) gw_locate_rows_kernel<<<gw_grid(n, 128), 128, 0, st>>>(ix, which, rows, n, out);
}
void gwk_scan_export(cudaStream_t st, u64* const* slabs, uint32_t n, int strand, uint32_t* out) {
    if (n) gw_scan_export_kernel<<<gw_grid(n, 128), 128, 0, st>>>(slabs, n, strand, out);
}
void gwk_verify_pairs(cudaStream_t st, const char* refs, const u64* ref_off, const char* queries, const u64* query_off, uint32_t n, int band_k,
                      int32_t* result, uint32_t* cigar_ops, uint32_t cigar_stride, void* hist, uint32_t max_m) {
    if (n) gw_verify_pairs_kernel<<<gw_grid(n, 128), 128, 0, st>>>(refs, ref_off, queries, query_off, n, band_k, result, cigar_ops, cigar_stride, (GwU64x2*)hist, max_m);
}
