// gw_kernels.h — host-callable launchers of the kernels in gw_kernels.cu (all asynchronous on the given stream).
#pragma once
#include <cuda_runtime.h>

#include "gw_bodies.cuh"

void gwk_fill_slab_ptrs(cudaStream_t st, u64** slabs, u64* base, size_t slab_words, const uint32_t* slots, uint32_t n);
void gwk_prep(cudaStream_t st, const GwBatchView& bv, u64* const* slabs, const uint32_t* slots, uint32_t n, uint32_t read_base, const GwSlabParams& sp);
void gwk_scan(cudaStream_t st, const GwIndexView& ix, u64* const* slabs, const uint32_t* slots, uint32_t n_slots, int coop);
void gwk_flow_setup(void);   // per-device function attributes of the flow kernel
void gwk_flow(cudaStream_t st, const GwSearchEnv& env, u64* const* slabs, const uint32_t* active, uint32_t n_active, uint32_t* work_counter, void* hist,
              uint32_t fetch, uint32_t batch, uint32_t age, uint32_t flush, uint32_t blocks, int maxh);
void gwk_densify(cudaStream_t st, const GwIndexView& ix, int fm, uint32_t step, uint32_t* table /* ceil(N / step), preset to 0xFF */, uint32_t n_samples,
                 uint32_t* counter /* device word, zero */);
void gwk_collect_overflow(cudaStream_t st, u64* const* slabs, uint32_t n, uint32_t* list, uint32_t* count);
void gwk_settle(cudaStream_t st, u64** slabs, u64* small_base, size_t small_words, const GwSlabParams& small_sp, const uint32_t* list, uint32_t n, const GwDevConfig& cfg);
void gwk_emit_count(cudaStream_t st, u64* const* slabs, uint32_t n, const GwDevConfig& cfg, uint32_t* n_rec, uint32_t* n_cig, unsigned long long* totals,
                    uint32_t* read_ctr /* 4 x n: fm searches, sw calls, cut-offs, filtered; may be null */);
void gwk_scan2(cudaStream_t st, const uint32_t* a, const uint32_t* b, uint32_t n, unsigned long long base_a, unsigned long long base_b,
               unsigned long long* oa, unsigned long long* ob, unsigned long long* out_tot, unsigned long long* part /* 2048 words of scratch */);
void gwk_emit_write(cudaStream_t st, u64* const* slabs, uint32_t n, const GwDevConfig& cfg, uint32_t read_base, const unsigned long long* rec_off,
                    const unsigned long long* cig_off, void* recs, unsigned long long rec_cap, uint32_t* cigs, unsigned long long cig_cap);
void gwk_gather_bench(cudaStream_t st, const GwIndexView& ix, int which, uint32_t blocks, uint32_t per_thread, uint32_t seed, uint32_t* sink);
void gwk_rank_acgtn(cudaStream_t st, const GwIndexView& ix, int which, const uint32_t* pos, uint32_t n, uint32_t* out);
void gwk_locate_rows(cudaStream_t st, const GwIndexView& ix, int which, const uint32_t* rows, uint32_t n, uint32_t* out);
void gwk_scan_export(cudaStream_t st, u64* const* slabs, uint32_t n, int strand, uint32_t* out);
size_t gwk_order_scratch_words(uint32_t n);
void gwk_order(cudaStream_t st, u64* const* slabs, uint32_t n, uint8_t* cls, uint32_t* counts, uint32_t* order);
void gwk_verify_pairs(cudaStream_t st, const char* refs, const u64* ref_off, const char* queries, const u64* query_off, uint32_t n, int band_k,
                      int32_t* result, uint32_t* cigar_ops, uint32_t cigar_stride, void* hist, uint32_t max_m);
// gw_noalloc.cu: the pair verifier with its DP history in a static __device__ array (no allocation anywhere on the call path)
#define GW_NOALLOC_SCRATCH_BYTES (64u << 20)
#ifndef GW_NOALLOC_PAIR_TOO_LONG
#define GW_NOALLOC_PAIR_TOO_LONG (-2)
#endif
size_t gwk_verify_pairs_noalloc_need(uint32_t n, uint32_t max_cols, uint32_t max_m);
cudaError_t gwk_verify_pairs_noalloc(cudaStream_t st, const char* refs, const u64* ref_off, const char* queries, const u64* query_off, uint32_t n,
                                     uint32_t max_cols, uint32_t max_m, int band_k, int32_t* result, uint32_t* cigar_ops, uint32_t cigar_stride);
