/* gw_align.h — C-ABI of the B200-native single-end read-alignment path (libgwalign.so).
 *
 * Drop-in boundary for genome-weaver-align's `Aligner` interface. The reference has no FFI (it is 100% JVM);
 * each entry point below names the JVM interface or method it replaces, with paths relative to
 * /root/reference/align/src/main/java/org/utgenome/weaver/ . The JVM-side stub (`CudaAligner implements Aligner`,
 * JNI or Panama FFM) a maintainer would add is in INTEGRATION.md.
 *
 * Rules: plain pointers and sizes only; all buffers are caller-owned and nothing is retained after return;
 * every function returns 0 on success or a negative GW_E_* code (never throws, never exits);
 * gw_last_error() gives the thread-local message of the last failure. There is no CPU fallback: if no CUDA
 * device is usable the calls fail with GW_E_CUDA.
 */
#ifndef GW_ALIGN_H
#define GW_ALIGN_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GW_OK 0
#define GW_E_ARG (-1)     /* bad argument / unsupported configuration */
#define GW_E_IO (-2)      /* index or read file missing / malformed */
#define GW_E_CUDA (-3)    /* CUDA runtime failure or no device */
#define GW_E_CAPACITY (-4) /* caller's output buffers too small (required sizes are written back) */

/* read status codes in gw_hit_t.status */
#define GW_READ_MAPPED 0
#define GW_READ_UNMAPPED 1
#define GW_READ_REF_ABORT 2   /* the reference implementation throws on this read */
#define GW_READ_UNSUPPORTED 3 /* outside the supported domain (read length 0 or > 256, k > 25) */

/* special chr_id values (ReadHit.chr): */
#define GW_CHR_NULL (-1)  /* Java null  (soft-clipped fragment, BidirectionalSuffixFilter.java:505-508) */
#define GW_CHR_EMPTY (-2) /* ""         (ReadHit.noHit, record/ReadHit.java:55-57) */
#define GW_CHR_STAR (-3)  /* "*"        (unmapped read, BidirectionalSuffixFilter.java:260) */

typedef struct gw_index gw_index_t;     /* FMIndexOnGenome + forward reference (align/FMIndexOnGenome.java:42-86, align/Align.java:159-170);
                                           immutable after open, lives in HBM of one device */
typedef struct gw_aligner gw_aligner_t; /* BidirectionalSuffixFilter instance (align/strategy/BidirectionalSuffixFilter.java:97-102):
                                           owns a stream, device work areas and staircase-filter tables; one per host thread */

/* AlignmentConfig / AlignmentScoreConfig (align/AlignmentScoreConfig.java:37-77, align/AlignmentConfig.java:40-72) */
typedef struct {
    float k;                  /* -k : 0<k<1 => floor(m*k), else (int)k ; default 0.1 */
    int32_t num_split_allowed;  /* -s : 0 or 1 (default 1) */
    int32_t match_score;        /* -M : 1 */
    int32_t mismatch_penalty;   /* -N : 3 */
    int32_t split_open_penalty; /* -S : 11 */
    int32_t indel_end_skip;     /* -P : 5 */
    int32_t band_width;         /* -W : 31 */
    int32_t report_type;        /* -R : 0 BESTHIT, 1 ALLHITS, 2 TOPL */
    int32_t top_l;              /* -L : 5 */
    int32_t strategy;           /* -m : 0 BSF (only BSF is implemented) */
    int32_t extended;           /* 0 (default): reads on which the reference's StaircaseFilter constructor throws (m >= 156) come back
                                   GW_READ_REF_ABORT; 1: such filters are built with out-of-range blocks read as 0 and the read is aligned */
} gw_config_t;

/* One ReadHit fragment (record/ReadHit.java:42-73). A reported alignment is a chain of 1 or 2 fragments:
 * the head has is_head=1; a following record with is_head=0 is its nextSplit. */
typedef struct {
    uint32_t read_index;  /* index of the read within the batch */
    int32_t chr_id;       /* contig index into the index's contig table, or GW_CHR_* */
    int32_t pos;          /* 1-based position on the contig (ReadHit.pos) */
    int32_t match_length; /* ReadHit.matchLength */
    int32_t diff;         /* ReadHit.diff (NM of this fragment) */
    int32_t num_hits;     /* ReadHit.numHits (X0) */
    int32_t q_start, q_end;
    uint32_t cigar_off;   /* offset into the cigar_ops output (BAM encoding: len<<4 | op, op: 0 M,1 I,2 D,4 S) */
    uint32_t cigar_len;   /* number of cigar elements */
    int8_t strand;        /* 0 '+', 1 '-' */
    uint8_t is_head;      /* 1: first fragment of a reported alignment */
    uint8_t status;       /* GW_READ_* (meaningful on head records) */
    uint8_t reserved;
} gw_hit_t;

/* per-batch work counters (the reference's per-read debug counters summed, BidirectionalSuffixFilter.java:588-592,
 * plus the roofline work unit: FM occurrence-table lookups, 64 algorithmic bytes each — SURVEY.md 8d) */
typedef struct {
    uint64_t fm_searches, sw_calls, cutoffs, filtered;
    uint64_t occ_lookups_scan, occ_lookups_search, occ_lookups_locate;
    uint64_t rounds;        /* search/locate/verify rounds executed */
    uint64_t gpu_launches;  /* kernels launched for this batch */
} gw_counters_t;

const char* gw_last_error(void);

/* `genome-weaver bwt <fasta>` (align/BWTransform.java:72-151, align/PackFasta.java:61-138): writes <prefix>.i.silk,
 * .f.3bit, .r.3bit, .f.ssa, .r.ssa, .f.bwt, .r.bwt next to the FASTA; suffix sorting runs on `device`. */
int gw_index_build(const char* fasta_path, int device);

/* FMIndexOnGenome.load + ACGTSequence.loadFrom(.f.3bit) (align/FMIndexOnGenome.java:60-86, align/Align.java:159-170):
 * reads the reference's index files and lays them out in HBM of `device`. */
int gw_index_open(const char* fasta_path, int device, gw_index_t** out);
/* Locate table in HBM: the suffix-array value of every step-th row (step 1, 2, 4, 8, 16; 32 = the .ssa samples alone, as
 * SparseSuffixArray.get uses them, align/SparseSuffixArray.java:141-161; 0 = the densest table that fits a fifth of the free device
 * memory, which gw_index_open applies). Built on the device from the opened index; results of every call are unchanged by it, a
 * locate just takes fewer (with step 1: no) LF steps. */
int gw_index_locate_table(gw_index_t*, int step);
int gw_index_locate_step(const gw_index_t*);
/* FMIndexOnGenome.buildFromSequence (align/FMIndexOnGenome.java:104-119): single contig from an ACGTN string, no files. */
int gw_index_from_sequence(const char* name, const char* bases, int device, gw_index_t** out);
void gw_index_close(gw_index_t*);
int64_t gw_index_text_size(const gw_index_t*);
int32_t gw_index_num_contigs(const gw_index_t*);
const char* gw_index_contig_name(const gw_index_t*, int32_t i);
int64_t gw_index_contig_length(const gw_index_t*, int32_t i);
int64_t gw_index_contig_offset(const gw_index_t*, int32_t i);
int gw_index_device(const gw_index_t*);

void gw_config_default(gw_config_t*);
/* new BidirectionalSuffixFilter(fmIndex, reference, config) (align/strategy/BidirectionalSuffixFilter.java:97-102).
 * Device memory: at its first batch an aligner measures the free memory of its device and plans its work areas within it (up to about
 * 90 %; buffers are allocated as batches need them and kept until gw_aligner_destroy). Option "mem_budget_mb" sets a smaller budget, e.g.
 * for several aligners on one device. */
int gw_aligner_create(const gw_index_t*, const gw_config_t*, gw_aligner_t** out);
void gw_aligner_destroy(gw_aligner_t*);

/* Aligner.align(Read, Reporter) for a whole batch (align/Aligner.java:30-33, align/Align.java:188-195).
 *   bases    ASCII read sequences, concatenated; read i is bases[offsets[i] .. offsets[i+1])
 *   hits     out: ReadHit records sorted by read_index, then report order; *n_hits in: capacity, out: used
 *   cigar_ops out: BAM-encoded cigar elements; *n_cigar in: capacity, out: used
 *   counters out (may be NULL)
 * Every read yields at least one head record (an unmapped read yields the reference's "*" record with status 1).
 * With GW_E_CAPACITY the required sizes are written to *n_hits / *n_cigar and nothing else is valid. */
int gw_align_batch(gw_aligner_t*, const char* bases, const uint64_t* offsets, uint32_t n_reads, gw_hit_t* hits,
                   uint64_t* n_hits, uint32_t* cigar_ops, uint64_t* n_cigar, gw_counters_t* counters);
/* No counterpart in the reference (its Reporter is called once per alignment): the hit records of this aligner's LAST gw_align_batch call,
 * copied out again. For the caller that got GW_E_CAPACITY: enlarge the buffers to the sizes written back and fetch, instead of aligning the
 * batch a second time. Same meaning of n_hits / n_cigar (in: capacity, out: count). GW_E_ARG when nothing is retained (no batch yet, the
 * last batch failed, or gw_align_batch_device was called since). */
int gw_align_batch_fetch(gw_aligner_t*, gw_hit_t* hits, uint64_t* n_hits, uint32_t* cigar_ops, uint64_t* n_cigar);

/* Same path with inputs already resident in HBM of the aligner's device (device pointers), results left on the device:
 * used to time the kernels without host<->device copies. d_hits/d_cigar capacities in elements. Returns counts via host ptrs. */
int gw_align_batch_device(gw_aligner_t*, const char* d_bases, const uint64_t* d_offsets, uint32_t n_reads, uint64_t total_bases,
                          uint32_t max_read_len, gw_hit_t* d_hits, uint64_t hits_cap, uint32_t* d_cigar, uint64_t cigar_cap,
                          uint64_t* n_hits, uint64_t* n_cigar, gw_counters_t* counters);

/* milliseconds spent in device work of the last gw_align_batch* call per kernel family, measured with CUDA events on the
 * aligner's stream: [0] prep + quick scan + work order, [1] flow kernel (search + locate + verify, all passes), [2] and [3] unused (0),
 * [4] emit, [5] whole batch */
int gw_aligner_last_timing(const gw_aligner_t*, float ms[6]);

/* the reference's per-read debug counters (align/strategy/BidirectionalSuffixFilter.java:224-233, :588-592) of reads
 * [first, first+n) of the last gw_align_batch* call: out[4*i..] = numFMIndexSearches, numSW, numCutOff, numFiltered */
int gw_aligner_read_counters(const gw_aligner_t*, uint32_t first, uint32_t n, uint32_t* out);

/* tunables of an aligner (all optional; none changes a result — INTEGRATION.md has the table): "scan_coop" 0/1 (4-lane cooperative quick
 * scan, default 1), "slab_words" (8-byte words of a read's first work area; 0 = chosen from the batch), "big_slab_words" (largest work
 * area), "ext_pct" / "ext_mult" / "ext2_pm" (blocks running reads can move into), "mem_budget_mb" / "tier_budget_mb" (device memory for
 * work areas; 0 = from the free memory), "wave_reads" (upper bound of the reads resident per wave), "flow_order", "flow_fetch",
 * "flow_batch", "flow_age", "flow_flush", "flow_blocks" (flow kernel), "l2_persist_mb". Unknown names return GW_E_ARG. */
int gw_aligner_set_option(gw_aligner_t*, const char* name, int64_t value);

/* AlignmentRecord.convert + toSAMLine + SAMOutput.emit (align/record/AlignmentRecord.java:92-276, align/SAMOutput.java:72-82) for the hit
 * records of one batch, on the host. names/quals are concatenated like bases (offset arrays of n_reads+1); quals may be NULL ("*").
 * *out_len in: capacity of out, out: bytes needed/used (GW_E_CAPACITY when too small). lines_per_read (may be NULL) receives the
 * number of SAM lines each read produced (0 where convert returns null or the reference throws on the read, GW_READ_REF_ABORT; a read
 * outside the supported domain, GW_READ_UNSUPPORTED, gets the unmapped "*" record so that the file keeps one entry per read). */
int gw_sam_format(const gw_index_t*, const char* names, const uint64_t* name_off, const char* bases, const uint64_t* base_off,
                  const char* quals, const uint64_t* qual_off, uint32_t n_reads, const gw_hit_t* hits, uint64_t n_hits,
                  const uint32_t* cigar_ops, char* out, uint64_t* out_len, uint32_t* lines_per_read);
/* SAMOutput.init: "@SQ\tSN:%s\tLN:%d\n" per contig (align/SAMOutput.java:51-59, align/SequenceBoundary.java:81-87) */
int gw_sam_header(const gw_index_t*, char* out, uint64_t* out_len);

/* ---- component entry points (used by the parity tests; each replaces one reference method) ---- */
/* FMIndexOnOccTable.rankACGTN (align/FMIndexOnOccTable.java:53-55) for n positions; which: 0 forward BWT, 1 reverse BWT; out n x 5 */
int gw_rank_acgtn(const gw_index_t*, int which, const uint32_t* pos, uint32_t n, uint32_t* out);
/* measurement aid, no reference counterpart: rate of INDEPENDENT random 64 B occurrence-block gathers (rank queries at
 * random BWT positions of index `which`), the bound a dependent FM step can approach; about n_lookups gathers per iteration */
int gw_gather_benchmark(const gw_index_t*, int which, uint64_t n_lookups, int iters, double* gb_per_s, double* ms_per_iter);
/* SparseSuffixArray.get (align/SparseSuffixArray.java:141-161) for n SA rows */
int gw_locate(const gw_index_t*, int which, const uint32_t* sa_rows, uint32_t n, uint32_t* out);
/* FMQuickScan.scanMismatchLocations (align/strategy/FMQuickScan.java:66-94) for a read batch and strand (0 '+': scans the read,
 * 1 '-': scans its complement): out n x 5 = numMismatches, longestMatch.start, longestMatch.end, si.lb, si.ub.
 * cooperative != 0: the 4-lanes-per-task kernel the aligner uses; 0: the one-thread-per-task form */
int gw_quick_scan(const gw_index_t*, const char* bases, const uint64_t* offsets, uint32_t n_reads, int strand, int cooperative, uint32_t* out);
/* BitParallelSmithWaterman.alignBlockDetailed(ref, query, k) (align/BitParallelSmithWaterman.java:141-147) for n pairs of ACGTN
 * strings: result[i*4..] = valid, pos, numMismatches, cigar_len ; cigars at cigar_ops[i*cigar_stride..] */
int gw_verify_pairs(int device, const char* refs, const uint64_t* ref_off, const char* queries, const uint64_t* query_off, uint32_t n,
                    int band_k, int32_t* result, uint32_t* cigar_ops, uint32_t cigar_stride);
/* No counterpart in the reference: added for callers that cannot allocate device memory (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc). Same computation and same outputs as gw_verify_pairs, but
 *  - every pointer is a DEVICE pointer on `device` (refs/queries: ACGTN characters; *_off: n + 1 ascending byte offsets; d_result: n x 4
 *    int32; d_cigar_ops: n x cigar_stride uint32, rows are zero-filled after the cigar as gw_verify_pairs leaves them);
 *  - the DP history (the verifier's only scratch memory, 16 bytes per pair, reference column and 64-base block of the query) lives in a
 *    static device array of
 *    GW_VERIFY_PAIRS_NOALLOC_MAX_BYTES that every GPU which loads the library holds until the process exits. The call allocates and frees
 *    nothing (no cudaMalloc / cudaFree, no allocating library routine) and is ASYNCHRONOUS on `stream` (a cudaStream_t, NULL = default
 *    stream): synchronise the stream before reading the outputs;
 *  - max_ref_len / max_query_len: upper bounds of the reference (window) and query lengths of the batch, stated by the caller because the
 *    offsets are on the device. n * (max_ref_len + 2) * ceil(max_query_len / 64) * 16 bytes must fit the static array (about 19 400 pairs of
 *    100-base queries with 106-base windows), else GW_E_ARG with a message that says so; a pair longer than stated gets
 *    result[i*4] = GW_NOALLOC_PAIR_TOO_LONG and is not aligned.
 * All calls share the one array per GPU, which keeps the last call's contents (nothing in it has to start at zero: every word read was
 * written by the same launch). Two calls whose kernels overlap on one GPU (different streams or host threads) RACE on it: serialise them. */
#define GW_VERIFY_PAIRS_NOALLOC_MAX_BYTES (64u << 20)
#define GW_NOALLOC_PAIR_TOO_LONG (-2)
int gw_verify_pairs_noalloc(int device, void* stream, const char* d_refs, const uint64_t* d_ref_off, const char* d_queries,
                            const uint64_t* d_query_off, uint32_t n, uint32_t max_ref_len, uint32_t max_query_len, int band_k, int32_t* d_result,
                            uint32_t* d_cigar_ops, uint32_t cigar_stride);
/* StaircaseFilter(m, kk) masks (align/strategy/StaircaseFilter.java:47-73): (kk+1) x ceil(m/64) words; returns 2 if the reference throws
 * (never with extended != 0: gw_config_t.extended) */
int gw_staircase_table(int m, int kk, int extended, uint64_t* out, int cap_words);
/* CyclicSAIS.SAIS (align/sais/CyclicSAIS.java:223-382): suffix array of the cyclic rotations of an ACGTN string, computed on `device` */
int gw_cyclic_suffix_array(const char* bases, uint64_t n, int device, uint32_t* sa_out);
/* AlignmentScoreConfig.getMaximumEditDistance (align/AlignmentScoreConfig.java:40-47) */
int gw_max_edit_distance_c(float k, int read_length);

#ifdef __cplusplus
}
#endif
#endif
