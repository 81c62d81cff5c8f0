// gw_host.h — host-side data structures of libgwalign.so: the `genome-weaver bwt` index files, FASTA packing,
// staircase-filter tables and SAM record formatting. No CUDA types here.
#pragma once
#include <stdint.h>

#include <algorithm>
#include <stdexcept>
#include <string>
#include <vector>

struct GwContig {
    std::string name, desc;
    int64_t length = 0, offset = 0;
};

// an index in host memory, one symbol code (A0 C1 G2 T3 N4) per byte
struct GwHostIndex {
    uint64_t N = 0;
    std::vector<uint8_t> text;     // forward reference text (.f.3bit)
    std::vector<uint8_t> bwt[2];   // [0] BWT of the text (.f.bwt), [1] BWT of the reversed text (.r.bwt)
    std::vector<uint32_t> ssa[2];  // SA samples every 32 rows (.f.ssa / .r.ssa)
    std::vector<GwContig> contigs; // .i.silk
};

struct GwIoError : public std::runtime_error {
    explicit GwIoError(const std::string& w) : std::runtime_error(w) {}
};

// host threads for opening an index (file decoding and laying out the HBM image): fn(lo, hi, part) over [0, n) in `parts` pieces;
// the first exception thrown by a piece is rethrown in the caller
#include <exception>
#include <thread>
template <typename F>
inline void gw_parallel_for(uint64_t n, unsigned parts, F fn) {
    if (parts < 1) parts = 1;
    if (parts == 1 || n < 2) { fn((uint64_t)0, n, 0u); return; }
    std::vector<std::thread> th;
    std::vector<std::exception_ptr> err(parts);
    const uint64_t per = (n + parts - 1) / parts;
    for (unsigned t = 0; t < parts; ++t) {
        const uint64_t lo = std::min<uint64_t>(n, (uint64_t)t * per), hi = std::min<uint64_t>(n, lo + per);
        th.emplace_back([&, lo, hi, t] { try { fn(lo, hi, t); } catch (...) { err[t] = std::current_exception(); } });
    }
    for (auto& x : th) x.join();
    for (auto& e : err) if (e) std::rethrow_exception(e);
}
inline unsigned gw_host_threads() { unsigned h = std::thread::hardware_concurrency(); return h < 1 ? 1 : (h > 32 ? 32 : h); }

// ---- files (gw_files.cpp) ----------------------------------------------------------------------------------
std::string gw_index_prefix(const std::string& fasta_path);                       // BWTFiles.java:40-47
void gw_pack_fasta(const std::string& fasta_path, std::vector<uint8_t>& codes, std::vector<GwContig>& contigs);  // PackFasta.java:81-130
void gw_save_acgt(const std::string& path, const uint8_t* codes, uint64_t n);      // ACGTSequence.saveTo :391-397
void gw_load_acgt(const std::string& path, std::vector<uint8_t>& codes);           // ACGTSequence.loadFrom :399-408
void gw_save_ssa(const std::string& path, uint64_t N, const std::vector<uint32_t>& samples);  // SparseSuffixArray.saveTo :91-99
void gw_load_ssa(const std::string& path, uint64_t* N, int* L, std::vector<uint32_t>& samples);
void gw_save_silk(const std::string& path, const std::vector<GwContig>& contigs, uint64_t total);  // PackFasta.java:120-127
void gw_load_silk(const std::string& path, std::vector<GwContig>& contigs, uint64_t* total);       // SequenceBoundary.loadSilk :123-139
void gw_save_index_files(const std::string& fasta_path, const GwHostIndex& ix);   // all seven files (needs text reversed for .r.3bit)
void gw_load_index_files(const std::string& fasta_path, GwHostIndex& ix);         // FMIndexOnGenome.load :60-86 + Align.java:166

// ---- staircase filter tables (gw_staircase.cpp) ------------------------------------------------------------
// StaircaseFilter(m, kk) masks (strategy/StaircaseFilter.java:47-73) with Java byte/shift semantics.
// words: (kk+1) x ceil(m/64). Returns false where the reference constructor throws (ArrayIndexOutOfBoundsException).
// extended: out-of-range block reads give 0 instead of the exception (the oracle's `extended` domain)
bool gw_staircase_masks(int m, int kk, std::vector<uint64_t>& words, bool extended = false);
// AlignmentScoreConfig.getMaximumEditDistance (AlignmentScoreConfig.java:40-47)
int gw_max_edit_distance(float k, int read_length);

// ---- index construction on the device (gw_build.cu) --------------------------------------------------------
// BWT + SA samples (every 32 rows) of the cyclic rotations of codes[0..n); sa_out (optional) receives the full suffix array
void gw_build_bwt_device(int device, const uint8_t* codes, uint64_t n, std::vector<uint8_t>& bwt, std::vector<uint32_t>& ssa,
                         std::vector<uint32_t>* sa_out);
void gw_build_host_index(int device, GwHostIndex& ix);   // fills bwt[2], ssa[2] from text

// ---- SAM formatting (gw_sam.cpp) ---------------------------------------------------------------------------
struct gw_hit_t_tag;
std::string gw_sam_header_text(const std::vector<GwContig>& ctg);
