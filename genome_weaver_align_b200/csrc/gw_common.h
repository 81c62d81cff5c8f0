// gw_common.h — structures shared by host and device code of libgwalign.so.
// Data layout in HBM (DESIGN.md "Data layout"):
//   occ block (64 B, one per 128 BWT symbols, 64 B aligned):
//     u32 cnt[4]   occurrences of A,C,G,T in BWT[0, 128*b)           (16 B)
//     u64 lo[2]    low  bit of the 2-bit code, bit j = symbol 128*b+j (16 B)
//     u64 hi[2]    high bit of the 2-bit code                          (16 B)
//     u64 nf[2]    N flag (N symbols are excluded from A/C/G/T counts) (16 B)
//   This replaces OccurrenceCountTable.occTable + the BWT ACGTSequence (reference: OccurrenceCountTable.java:51-72,
//   ACGTSequence.java:47-58) by one cache-line-aligned record, so a rank query touches exactly one 64 B line.
#pragma once
#include <stdint.h>

#define GW_MAX_READ_LEN 256
#define GW_MAX_NB 4          // ceil(GW_MAX_READ_LEN / 64)
#define GW_MAX_K 15          // NFA rows <= 16
#define GW_MAX_ROWS (GW_MAX_K + 1)
#define GW_SSA_L 32          // SparseSuffixArray sampling (BWTransform.java:141)
#define GW_CIG_CAP 36        // cigar elements kept per verified fragment (>= 2*GW_MAX_K + 3)
#define GW_MAX_WINDOW (GW_MAX_READ_LEN + 2 * GW_MAX_K + 2)

struct GwOccBlock {
    uint32_t cnt[4];
    uint64_t lo[2], hi[2], nf[2];
};

// device view of an opened index
struct GwIndexView {
    const GwOccBlock* occ[2];   // [0] BWT of the text ("forward index"), [1] BWT of the reversed text ("reverse index")
    const uint32_t* ssa[2];     // SA samples every GW_SSA_L rows
    uint32_t C[2][5];           // CharacterCount.C per index
    uint32_t N;                 // text length (all contigs concatenated)
    uint32_t n_ssa;
    // forward reference text as bit planes (verification windows), one u64 per 64 bases
    const uint64_t* ref_lo;
    const uint64_t* ref_hi;
    const uint64_t* ref_nf;
    // contig table for SequenceBoundary.translate: distinct offsets ascending, the contig id owning each, name order rank
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

// per-wave slab geometry (one slab per read; see gw_device.cuh)
struct GwSlabParams {
    uint32_t slab_words;   // size of one read slab in 8-byte words
    uint32_t queue_cap;    // max live states in the priority queue
    uint32_t hit_cap;      // max entries of the hit list
};

// staircase filter tables for every read length present in a batch
struct GwStaircaseView {
    const uint64_t* words;      // all tables
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
