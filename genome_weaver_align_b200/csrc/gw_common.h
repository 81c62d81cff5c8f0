// gw_common.h — structures shared by host and device code of libgwalign.so.
//
// Data layout in HBM (DESIGN.md "Data layout"):
//   occ block (64 B, one per 128 BWT symbols, 64 B aligned) = 4 quarters of 16 B; quarter q holds
//     u32 cnt   occurrences of symbol q (A,C,G,T for q = 0..3) in BWT[0, 128*b)
//     u32 lo    low  bit of the 2-bit code of symbols 128*b+32*q .. +31 (bit j = symbol 32*q+j)
//     u32 hi    high bit of the 2-bit code
//     u32 nf    N flag (N symbols are excluded from the A/C/G/T counts)
//   Every quarter is self-contained, so a rank query can be served either by one thread with four 128-bit loads
//   or by four lanes with one 128-bit load each (one 64 B line per query either way).
//   This replaces OccurrenceCountTable.occTable + the BWT ACGTSequence of the reference
//   (align/src/main/java/org/utgenome/weaver/align/OccurrenceCountTable.java:51-72, ACGTSequence.java:47-58).
//
// The per-thread bodies of all kernels are written as GW_HD functions: nvcc compiles them for sm_100a; the test-only
// host simulator (tests/hostsim) compiles the same headers with g++ to check the automaton logic without a GPU.
#pragma once
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define GW_HD __host__ __device__ __forceinline__
// rarely executed, large bodies (seeding, reporting, split states, DP set-up, traceback): ONE out-of-line copy keeps the hot loops
// of the fused kernel small enough for the instruction cache
#define GW_COLD __attribute__((unused)) static __host__ __device__ __noinline__
#else
#define GW_HD inline
#define GW_COLD __attribute__((unused)) static inline
#endif

typedef unsigned long long u64;

#define GW_MAX_READ_LEN 256
#define GW_MAX_NB 4          // ceil(GW_MAX_READ_LEN / 64)
#define GW_MAX_K 25          // NFA rows <= 26: covers the default -k 0.1 for every supported read length
#define GW_MAX_ROWS (GW_MAX_K + 1)
#define GW_SSA_L 32          // SparseSuffixArray sampling (BWTransform.java:141)
#define GW_CIG_CAP 56        // cigar elements kept per verified fragment (>= 2*GW_MAX_K + 3)
#define GW_MAX_WINDOW (GW_MAX_READ_LEN + 2 * GW_MAX_K + 2)

// special chr ids / status codes (mirrored in include/gw_align.h)
#define GW_CHR_NULL_ (-1)
#define GW_CHR_EMPTY_ (-2)
#define GW_CHR_STAR_ (-3)
#define GW_ST_OK 0
#define GW_ST_UNMAPPED 1
#define GW_ST_ABORT 2
#define GW_ST_UNSUPPORTED 3
#define GW_ST_OVERFLOW 100   // internal: slab too small, rerun in the large-slab pass

#define GW_OP_M 0u
#define GW_OP_I 1u
#define GW_OP_D 2u
#define GW_OP_S 4u

struct GwU4 { uint32_t x, y, z, w; };
struct GwOccBlock { GwU4 q[4]; };   // q[i] = {cnt_i, lo, hi, nf}
struct GwU64x2 { u64 x, y; };

// device view of an opened index
struct GwIndexView {
    const GwOccBlock* occ[2];   // [0] BWT of the text ("forward index"), [1] BWT of the reversed text ("reverse index")
    const uint32_t* ssa[2];     // SA samples every GW_SSA_L rows (the reference's .ssa files)
    // locate table: the suffix-array value of every loc_step-th row (loc_step a power of two <= GW_SSA_L). It is the .ssa samples
    // themselves (loc_step == GW_SSA_L) until gw_index_open densifies it on the device: HBM has room for the whole suffix array of a
    // human-scale genome, and a locate then costs one 4-byte gather instead of a walk of ~31 dependent occurrence-block gathers.
    // 0xFFFFFFFF marks a row the table does not know (only possible for rows no LF walk from a sampled row passes through).
    const uint32_t* loc[2];
    uint32_t loc_step;
    uint32_t C[2][5];           // CharacterCount.C per index
    uint32_t N;                 // text length (all contigs concatenated)
    uint32_t n_ssa;
    // forward reference text as bit planes (verification windows), one u64 per 64 bases, bit j = base 64*w+j
    const u64* ref_lo;
    const u64* ref_hi;
    const u64* ref_nf;
    // contig table for SequenceBoundary.translate: distinct offsets ascending, the contig id owning each
    const uint32_t* ctg_off;
    const int32_t* ctg_id;
    const int32_t* ctg_rank;    // by contig id: rank of its name under java.lang.String.compareTo (equal names share a rank)
    int32_t n_ctg_off;
    // FMIndexOnGenome.initRange (forwardSearch(FORWARD, wholeRange)): valid mask + bounds for A,C,G,T
    uint32_t init_lb[4], init_ub[4];
    uint32_t init_valid;
};

// AlignmentConfig on the device
struct GwDevConfig {
    float k;
    int32_t num_split_allowed, match_score, mismatch_penalty, split_open_penalty, indel_end_skip, band_width, report_type, top_l;
};

// per-wave slab geometry (one slab per read; see gw_state.cuh)
struct GwSlabParams {
    uint32_t slab_words;   // size of one read slab in 8-byte words
    uint32_t queue_cap;    // max live states in the priority queue
    uint32_t hit_cap;      // max entries of the hit list
    uint32_t queue_at;     // word offset of the priority queue in the work area; 0 in parameters made on the host = the standard place
                           // (GW_QUEUE_BASE, gw_state.cuh). A read that moves into a larger block takes its queue to the end of the
                           // block, with a larger capacity (gw_promote_commit); slab_words is then where the arena ends.
};

// pool of larger work areas (device memory): a read whose work area is nearly full is moved into ONE block of the pool between two
// search steps and goes on there, instead of being run again from the start. `next` is the allocation cursor (atomic, reset by
// the host for every wave).
// blocks a running read can move its work area into when the arena is nearly full (gw_promote_*): two sizes; a read moves from its
// first work area into a block of the first size and, if that fills up too, from there into one of the second size
struct GwExtPool {
    u64* base;
    uint32_t n_blocks;
    uint32_t block_words;
    uint32_t next;
    uint32_t pad_;
    u64* base2;            // second size (block_words2 > block_words), or n_blocks2 == 0
    uint32_t n_blocks2;
    uint32_t block_words2;
    uint32_t next2;
    uint32_t pad2_;
};

// staircase filter tables for every read length present in a batch
struct GwStaircaseView {
    const u64* words;           // all tables
    const int32_t* base_by_m;   // [GW_MAX_READ_LEN+1] word offset of the table set of length m, -1 if absent
    const uint32_t* abort_by_m; // [GW_MAX_READ_LEN+1] bit kk set: the reference throws when constructing StaircaseFilter(m, kk)
};

enum GwPhase : int32_t {
    GW_PH_INIT = 0,        // quick scan done, search not started
    GW_PH_RUN = 1,         // in the best-first loop
    GW_PH_WAIT_EXACT = 2,  // parked: exact hit waiting for its locate
    GW_PH_WAIT_VERIFY = 3, // parked: a chain waits for locate + verification of its fragments
    GW_PH_DONE = 4
};

// ---- portability shims (device intrinsics / host builtins) ------------------------------------------------
GW_HD int gw_popc32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __popc(x);
#else
    return __builtin_popcount(x);
#endif
}
GW_HD int gw_popc64(u64 x) {
#if defined(__CUDA_ARCH__)
    return __popcll(x);
#else
    return __builtin_popcountll(x);
#endif
}
GW_HD u64 gw_brev64(u64 x) {
#if defined(__CUDA_ARCH__)
    return __brevll(x);
#else
    u64 r = 0;
    for (int i = 0; i < 64; ++i) r |= ((x >> i) & 1ULL) << (63 - i);
    return r;
#endif
}
GW_HD GwU4 gw_ld128(const GwU4* p) {
#if defined(__CUDA_ARCH__)
    uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
    GwU4 r; r.x = v.x; r.y = v.y; r.z = v.z; r.w = v.w;
    return r;
#else
    return *p;
#endif
}
GW_HD uint32_t gw_ld32(const uint32_t* p) {
#if defined(__CUDA_ARCH__)
    return __ldg(p);
#else
    return *p;
#endif
}
GW_HD u64 gw_ld64(const u64* p) {
#if defined(__CUDA_ARCH__)
    return __ldg(p);
#else
    return *p;
#endif
}
GW_HD uint32_t gw_atomic_add(uint32_t* p, uint32_t v) {
#if defined(__CUDA_ARCH__)
    return atomicAdd(p, v);
#else
    return __atomic_fetch_add(p, v, __ATOMIC_RELAXED);
#endif
}
GW_HD int gw_max(int a, int b) { return a > b ? a : b; }
GW_HD int gw_min(int a, int b) { return a < b ? a : b; }

// Java shift semantics (count & 63)
GW_HD u64 jshl(u64 x, int s) { return x << (s & 63); }
GW_HD u64 jushr(u64 x, int s) { return x >> (s & 63); }
