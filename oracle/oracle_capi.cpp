// TEST INFRASTRUCTURE ONLY (see java_sem.hpp): C entry points over the CPU oracle, loaded with ctypes by
// tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs. Never linked into the product.
// Mirrors the reference's command flow: Align.execute / Align.query / PassReadToAligner
// (align/src/main/java/org/utgenome/weaver/align/Align.java:76-140, :188-195) and SAMOutput (:51-82).
#include <atomic>
#include <chrono>
#include <cstdlib>
#include <thread>

#include "bsf.hpp"
#include "index_io.hpp"

using namespace gwo;

namespace {
struct OracleIndex {
    std::unique_ptr<FMIndexOnGenome> fm;
    std::unique_ptr<ACGTSequence> reference;
};
std::vector<std::string> splitLines(const char* s) {
    std::vector<std::string> out;
    if (!s) return out;
    std::string cur;
    for (const char* p = s; *p; ++p) { if (*p == '\n') { out.push_back(cur); cur.clear(); } else cur += *p; }
    out.push_back(cur);
    return out;
}
thread_local std::string g_err;
}  // namespace

extern "C" {

const char* gwo_last_error() { return g_err.c_str(); }

void* gwo_index_from_sequence(const char* name, const char* seq) {
    try {
        auto* h = new OracleIndex();
        ACGTSequence ref{std::string(seq)};
        h->fm = buildFromSequence(name, ref);
        h->reference.reset(new ACGTSequence(ref));
        h->fm->index.translate(1);
        return h;
    } catch (std::exception& e) { g_err = e.what(); return nullptr; }
}
int gwo_index_build(const char* fasta_path) {
    try { buildIndexFiles(fasta_path); return 0; } catch (std::exception& e) { g_err = e.what(); return -1; }
}
void* gwo_index_load(const char* fasta_path) {
    try {
        auto* h = new OracleIndex();
        h->fm = loadIndex(fasta_path);
        h->reference.reset(new ACGTSequence(loadACGTSequence(indexPrefix(fasta_path) + ".f.3bit")));
        if (!h->fm->index.index.empty()) h->fm->index.translate(1);
        return h;
    } catch (std::exception& e) { g_err = e.what(); return nullptr; }
}
void gwo_index_free(void* h) { delete (OracleIndex*)h; }
void gwo_free(void* p) { free(p); }
long long gwo_index_text_size(void* h) { return ((OracleIndex*)h)->fm->textSize(); }
char* gwo_sam_header(void* h) {
    std::string s = ((OracleIndex*)h)->fm->index.toSAMHeader();
    char* o = (char*)malloc(s.size() + 1);
    memcpy(o, s.c_str(), s.size() + 1);
    return o;
}

// cfg: [k(float as double), numSplitAllowed, matchScore, mismatchPenalty, splitOpenPenalty, indelEndSkip, bandWidth, reportType, topL, extended]
// status per read: 0 mapped, 1 unmapped, 2 reference aborts here. counters per read: FM searches, SW, cutoff, filtered.
// lines per read: number of SAM lines that read produced (0 when AlignmentRecord.convert returns null).
int gwo_align(void* handle, const double* cfg, int n_reads, const char* names_joined, const char* seqs_joined,
              const char* quals_joined, int nthreads, char** out_sam, int* status, int* counters, int* lines_per_read,
              unsigned long long* occ_lookups, double* elapsed_sec) {
    try {
        OracleIndex* h = (OracleIndex*)handle;
        AlignmentConfig c;
        c.k = (float)cfg[0]; c.numSplitAlowed = (jint)cfg[1]; c.matchScore = (jint)cfg[2]; c.mismatchPenalty = (jint)cfg[3];
        c.splitOpenPenalty = (jint)cfg[4]; c.indelEndSkip = (jint)cfg[5]; c.bandWidth = (jint)cfg[6]; c.reportType = (int)cfg[7];
        c.topL = (jint)cfg[8]; c.extended = cfg[9] != 0;
        std::vector<std::string> names = splitLines(names_joined), seqs = splitLines(seqs_joined), quals = splitLines(quals_joined);
        if ((int)names.size() < n_reads || (int)seqs.size() < n_reads) { g_err = "too few names/seqs"; return -1; }
        bool haveQual = quals_joined != nullptr && (int)quals.size() >= n_reads;
        std::vector<BidirectionalSuffixFilter::Read> reads((size_t)n_reads);
        for (int i = 0; i < n_reads; ++i) {
            reads[(size_t)i].name = names[(size_t)i];
            reads[(size_t)i].seq = ACGTSequence(seqs[(size_t)i]);
            reads[(size_t)i].qualIsNull = !haveQual;
            if (haveQual) reads[(size_t)i].qual = quals[(size_t)i];
        }
        if (nthreads < 1) nthreads = 1;
        std::vector<std::string> sam((size_t)n_reads);
        std::vector<uint64_t> occ((size_t)nthreads, 0);
        std::atomic<int> next(0);
        auto t0 = std::chrono::steady_clock::now();
        auto work = [&](int tid) {
            BidirectionalSuffixFilter bsf(*h->fm, h->reference.get(), c);
            OccCounter::lookups = 0;
            for (;;) {
              int lo = next.fetch_add(64), hi = std::min(n_reads, lo + 64);  // dynamic chunks: per-read work has a heavy tail
              if (lo >= n_reads) break;
              for (int i = lo; i < hi; ++i) {
                std::string& out = sam[(size_t)i];
                int nl = 0;
                try {
                    BidirectionalSuffixFilter::ReadResult r = bsf.align(reads[(size_t)i]);
                    status[i] = r.unmapped ? 1 : 0;
                    counters[i * 4 + 0] = r.numFMIndexSearches; counters[i * 4 + 1] = r.numSW;
                    counters[i * 4 + 2] = r.numCutOff; counters[i * 4 + 3] = r.numFiltered;
                    for (auto& hit : r.reported) {
                        AlignmentRecordPtr rec = BidirectionalSuffixFilter::convert(*hit, reads[(size_t)i], 0);
                        if (rec) {  // SAMOutput.emit drops null
                            std::string line = rec->toSAMLine();
                            for (char ch : line) if (ch == '\n') nl++;
                            nl++;
                            out += line; out += '\n';
                        }
                    }
                } catch (JavaAbort& e) {
                    status[i] = 2;
                    counters[i * 4 + 0] = counters[i * 4 + 1] = counters[i * 4 + 2] = counters[i * 4 + 3] = -1;
                    out.clear(); nl = 0;
                }
                lines_per_read[i] = nl;
              }
            }
            occ[(size_t)tid] = OccCounter::lookups;
        };
        if (nthreads == 1) work(0);
        else {
            std::vector<std::thread> th;
            for (int t = 0; t < nthreads; ++t) th.emplace_back(work, t);
            for (auto& t : th) t.join();
        }
        auto t1 = std::chrono::steady_clock::now();
        if (elapsed_sec) *elapsed_sec = std::chrono::duration<double>(t1 - t0).count();
        if (occ_lookups) { *occ_lookups = 0; for (auto v : occ) *occ_lookups += v; }
        if (out_sam) {
            size_t total = 0;
            for (auto& s : sam) total += s.size();
            char* o = (char*)malloc(total + 1);
            size_t p = 0;
            for (auto& s : sam) { memcpy(o + p, s.data(), s.size()); p += s.size(); }
            o[p] = 0;
            *out_sam = o;
        }
        return 0;
    } catch (std::exception& e) { g_err = e.what(); return -1; }
}

// BitParallelSmithWaterman.alignBlockDetailed(String ref, String query, int k); returns 0 when the reference returns null
int gwo_align_block_detailed(const char* ref, const char* query, int k, char* cigar_out, int cap, int* pos, int* nm, int* score) {
    Alignment a = alignBlockDetailed(ACGTSequence(std::string(ref)), ACGTSequence(std::string(query)), k);
    if (!a.valid) return 0;
    std::string c = a.cigar.toString();
    snprintf(cigar_out, (size_t)cap, "%s", c.c_str());
    *pos = a.pos; *nm = a.numMismatches; *score = a.score;
    return 1;
}
// FMIndexOnOccTable.backwardSearch on the forward (which=0) or reverse (1) BWT
void gwo_backward_search(void* handle, int which, int ch, long long lb, long long ub, long long* out2) {
    OracleIndex* h = (OracleIndex*)handle;
    const FMIndexOnOccTable& fm = which == 0 ? *h->fm->forwardIndex : *h->fm->reverseIndex;
    SuffixInterval si = fm.backwardSearch(ch, SuffixInterval(lb, ub));
    out2[0] = si.lowerBound; out2[1] = si.upperBound;
}
void gwo_rank_acgtn(void* handle, int which, long long i, long long* out5) {
    OracleIndex* h = (OracleIndex*)handle;
    (which == 0 ? *h->fm->forwardIndex : *h->fm->reverseIndex).rankACGTN(i, (jlong*)out5);
}
long long gwo_locate(void* handle, int which, long long saIndex) {
    OracleIndex* h = (OracleIndex*)handle;
    return which == 0 ? h->fm->forwardSA.get(saIndex, *h->fm->forwardIndex) : h->fm->reverseSA.get(saIndex, *h->fm->reverseIndex);
}
// BWT string of the forward/reverse index as ACGTN text (caller frees)
char* gwo_bwt_string(void* handle, int which) {
    OracleIndex* h = (OracleIndex*)handle;
    std::string s = (which == 0 ? h->fm->forwardIndex : h->fm->reverseIndex)->seq.toString();
    char* o = (char*)malloc(s.size() + 1);
    memcpy(o, s.c_str(), s.size() + 1);
    return o;
}
void gwo_cyclic_sais_int(const int* t, int n, int* sa) {
    std::vector<uint32_t> v((size_t)n);
    for (int i = 0; i < n; ++i) v[(size_t)i] = (uint32_t)t[i];
    std::vector<uint32_t> r = cyclicSuffixSort(v);
    for (int i = 0; i < n; ++i) sa[i] = (int)r[(size_t)i];
}
void gwo_fast_count_acgtn(const char* seq, long long start, long long end, long long* out5) {
    ACGTSequence s{std::string(seq)};
    s.fastCountACGTN(start, end, (jlong*)out5);
}
long long gwo_fast_count(const char* seq, int ch, long long start, long long end) {
    ACGTSequence s{std::string(seq)};
    return s.fastCount(ch, start, end);
}
// QueryMask.getBidirectionalPatternMask64 (QueryMaskTest vectors)
unsigned long long gwo_query_mask64(const char* query, int dir, int nextIdx, int pivot, int cursor, int ch, int margin) {
    QueryMask qm{ACGTSequence(std::string(query))};
    return qm.getBidirectionalPatternMask64(dirIsForward(dir), nextIdx, pivot, cursor, ch, margin);
}
// StaircaseFilter table: for (m, kk) writes (kk+1) x ceil(m/64) words; returns 0 ok, 2 if the reference would throw
int gwo_staircase_table(int m, int kk, int extended, unsigned long long* out, int cap_words) {
    try {
        StaircaseFilter f(m, kk, extended != 0);
        int nb = (m + 63) / 64, p = 0;
        for (auto& bv : f.staircaseMask) for (int b = 0; b < nb; ++b) { if (p < cap_words) out[p] = bv.block[(size_t)b]; p++; }
        return 0;
    } catch (JavaAbort&) { return 2; }
}
int gwo_compress_roundtrip(const unsigned char* in, int n) {
    std::vector<uint8_t> c, d;
    snappyCompressLiteral(in, (size_t)n, c);
    if (!snappyUncompress(c.data(), c.size(), d)) return -1;
    return d.size() == (size_t)n && memcmp(d.data(), in, (size_t)n) == 0 ? 0 : -2;
}
// ---- re-hosted unit-test surface (BitVectorTest, CIGARTest, SparseSuffixArrayTest, BWTransformTest) ----
// BitVector.parseString(bits) [optionally .not() first] .lshift(len) .toString(): most significant bit first, as BitVector.toString
// prints it (BitVector.java:73-104, :180-192, :217-224). Only the members on the alignment path (_not, _lshift) exist in the oracle.
int gwo_bitvector_lshift(const char* bits, int len, int not_first, char* out, int cap) {
    try {
        const int n = (int)strlen(bits);
        BitVector v(n);
        for (int i = 0; i < n; ++i) if (bits[i] == '1') v.set(n - 1 - i);
        if (not_first) v._not();
        v._lshift(len);
        if (cap < n + 1) return -1;
        for (int i = 0; i < n; ++i) out[i] = ((v.block[(size_t)((n - 1 - i) / 64)] >> ((n - 1 - i) % 64)) & 1ULL) ? '1' : '0';
        out[n] = 0;
        return 0;
    } catch (JavaAbort&) { return 2; }
}
// new CIGAR(a).add(new CIGAR(b)).toString() (CIGAR.java:121-140, :221-241); b may be empty
int gwo_cigar_merge(const char* a, const char* b, char* out, int cap) {
    CIGAR c1{std::string(a)};
    CIGAR c2{std::string(b)};
    c1.add(c2);
    const std::string s = c1.toString();
    if ((int)s.size() + 1 > cap) return -1;
    memcpy(out, s.c_str(), s.size() + 1);
    return (int)c1.cigar.size();
}
// new CIGAR() with add(length, type) calls
int gwo_cigar_build(const int* lengths, const char* types, int n, char* out, int cap) {
    CIGAR c;
    for (int i = 0; i < n; ++i) c.add(lengths[i], types[i]);
    const std::string s = c.toString();
    if ((int)s.size() + 1 > cap) return -1;
    memcpy(out, s.c_str(), s.size() + 1);
    return (int)c.cigar.size();
}
// SparseSuffixArray.buildFromSuffixArray(CyclicSAIS(seq), L).saveTo -> loadFrom (SparseSuffixArray.java:91-131): 0 when the reloaded
// samples equal SA[i * L]; sa_out receives the suffix array
int gwo_ssa_roundtrip(const char* seq, int L, const char* tmp_path, int* sa_out) {
    try {
        ACGTSequence s{std::string(seq)};
        std::vector<uint32_t> codes((size_t)s.textSize());
        for (jlong i = 0; i < s.textSize(); ++i) codes[(size_t)i] = (uint32_t)s.lookup(i);
        std::vector<uint32_t> SA = cyclicSuffixSort(codes);
        for (size_t i = 0; i < SA.size(); ++i) sa_out[i] = (int)SA[i];
        SparseSuffixArray a = SparseSuffixArray::buildFromSuffixArray(SA, L);
        saveSSA(a, tmp_path);
        SparseSuffixArray b = loadSSA(tmp_path);
        if (b.N != a.N || b.L != a.L || b.sparseSA.size() != a.sparseSA.size()) return 1;
        for (size_t i = 0; i < a.sparseSA.size(); ++i) if (a.sparseSA[i] != b.sparseSA[i] || a.sparseSA[i] != SA[i * (size_t)L]) return 2;
        return 0;
    } catch (std::exception& e) { g_err = e.what(); return -1; }
}
// ACGTSequence.loadFrom(path).toString() (BWTransformTest.transform reads .f.3bit / .r.3bit back)
char* gwo_load_acgt_string(const char* path) {
    try {
        std::string s = loadACGTSequence(path).toString();
        char* o = (char*)malloc(s.size() + 1);
        memcpy(o, s.c_str(), s.size() + 1);
        return o;
    } catch (std::exception& e) { g_err = e.what(); return nullptr; }
}
}  // extern "C"
