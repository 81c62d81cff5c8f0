// gw_cli.cpp — `genome-weaver bwt <fasta>` and `genome-weaver align -r ref.fa reads.fastq > out.sam` over the C-ABI of
// include/gw_align.h (libgwalign.so). Same sub-commands, flags and output as the reference's command line
// (reference, align/src/main/java/org/utgenome/weaver/: GenomeWeaver.java:64-91, align/BWTransform.java:68-90,
// align/Align.java:76-104,188-195, align/AlignmentConfig.java:40-72, align/AlignmentScoreConfig.java:37-77,
// align/record/ReadReaderFactory.java:126-257, align/SAMOutput.java:51-82).
//
// The reference aligns one read per call on one thread; here the host is a pipeline
//   reader (FASTQ/FASTA -> batches) -> one worker per GPU (gw_align_batch + gw_sam_format) -> writer (batch order)
// with every GPU holding its own replica of the index (reads shard, nothing is exchanged between devices).
// SAM lines come out in input order, exactly as the reference's sequential loop prints them.
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include <zlib.h>

#include "../../include/gw_align.h"

namespace {

struct Batch {
    uint64_t seq = 0;           // position of the batch in the input
    uint32_t n = 0;
    std::string names, bases, quals;
    std::vector<uint64_t> name_off{0}, base_off{0}, qual_off{0};
    bool has_qual = true;       // FASTA reads and -q carry no quality (qual == null -> "*")
    std::vector<gw_hit_t> hits; // hit records and cigar elements of the batch (filled by the align stage of a worker)
    std::vector<uint32_t> cig;
    uint64_t nh = 0, nc = 0;
    bool aligned = false;
    std::string sam;            // SAM text of the batch (filled by the format stage of a worker)
    int64_t abort_read = -1;    // first read on which the reference throws (its run stops there), -1: none
    uint64_t n_unsupported = 0; // reads outside the supported domain (length 0 or > 256, k > 25, work area exhausted)
    int64_t first_unsupported = -1;
    gw_counters_t counters{};
    void add(const std::string& name, const std::string& seq_, const std::string* qual) {
        names += name; name_off.push_back(names.size());
        bases += seq_; base_off.push_back(bases.size());
        if (qual) quals += *qual; else has_qual = false;
        qual_off.push_back(quals.size());
        ++n;
    }
};

// ---- read files (ReadReaderFactory.createReader :126-153) ---------------------------------------------------------
bool ends_with(const std::string& s, const char* suf) {
    const size_t n = strlen(suf);
    return s.size() >= n && s.compare(s.size() - n, n, suf) == 0;
}

// lines of a block of text held in memory; views are valid as long as the block is
class LineReader {
  public:
    LineReader(const char* p, size_t n) : p_(p), end_(p + n) {}
    // next line without its terminator
    bool next(const char*& p, size_t& n) {
        if (p_ >= end_) return false;
        const char* e = (const char*)memchr(p_, '\n', (size_t)(end_ - p_));
        p = p_;
        n = (size_t)((e ? e : end_) - p_);
        p_ = e ? e + 1 : end_;
        if (n && p[n - 1] == '\r') --n;
        return true;
    }
  private:
    const char* p_;
    const char* end_;
};

std::string trim(const std::string& s) {
    size_t a = 0, b = s.size();
    while (a < b && (unsigned char)s[a] <= ' ') ++a;
    while (b > a && (unsigned char)s[b - 1] <= ' ') --b;
    return s.substr(a, b - a);
}

// FASTQ: 4 lines per record; read name = header line minus '@' (lens FASTQ.scala:35-52). Lines go straight from the
// block into the batch arrays.
void parse_fastq(const char* data, size_t len, Batch& b) {
    LineReader lr(data, len);
    const char* p; size_t n;
    for (;;) {
        for (;;) {
            if (!lr.next(p, n)) return;
            if (n) break;
        }
        const size_t n0 = b.names.size(), b0 = b.bases.size();
        if (p[0] == '@') { ++p; --n; }
        b.names.append(p, n);
        bool ok = lr.next(p, n);
        if (ok) {
            while (n && (unsigned char)p[0] <= ' ') { ++p; --n; }
            while (n && (unsigned char)p[n - 1] <= ' ') --n;
            b.bases.append(p, n);
            ok = lr.next(p, n) && lr.next(p, n);
        }
        if (!ok) { b.names.resize(n0); b.bases.resize(b0); return; }   // truncated record: stop like FastqReader does at EOF
        b.quals.append(p, n);
        b.name_off.push_back(b.names.size()); b.base_off.push_back(b.bases.size()); b.qual_off.push_back(b.quals.size());
        ++b.n;
    }
}

// FASTA reads: name = first whitespace-delimited token of the description line, sequence lines concatenated, no quality
void parse_fasta(const char* data, size_t len, Batch& b) {
    LineReader lr(data, len);
    const char* p; size_t n;
    bool have = false;
    std::string name, seq;
    auto pick_name = [](const char* q, size_t m) {
        size_t a = 0;
        while (a < m && (unsigned char)q[a] <= ' ') ++a;
        size_t e = a;
        while (e < m && (unsigned char)q[e] > ' ') ++e;
        return std::string(q + a, e - a);
    };
    while (lr.next(p, n)) {
        if (n && p[0] == '>') {
            if (have) b.add(name, seq, nullptr);
            have = true; name = pick_name(p + 1, n - 1); seq.clear();
        } else if (have)
            seq += trim(std::string(p, n));
    }
    if (have) b.add(name, seq, nullptr);
}

// ---- byte sources (ReadReaderFactory.createReader :126-153: plain, .gz through GZIPInputStream, .snap through SnappyInputStream) --
struct ByteSource {
    virtual ~ByteSource() {}
    virtual size_t read(char* dst, size_t cap) = 0;   // 0: end of the stream
    virtual const char* error() const { return nullptr; }
};
struct PlainSource : ByteSource {
    FILE* f;
    explicit PlainSource(FILE* f_) : f(f_) {}
    ~PlainSource() override { if (f) fclose(f); }
    size_t read(char* dst, size_t cap) override { return fread(dst, 1, cap, f); }
};
struct GzSource : ByteSource {
    gzFile g;
    std::string err;
    explicit GzSource(gzFile g_) : g(g_) { gzbuffer(g, 1u << 20); }
    ~GzSource() override { if (g) gzclose(g); }
    size_t read(char* dst, size_t cap) override {
        const int n = gzread(g, dst, (unsigned)std::min<size_t>(cap, 1u << 30));
        if (n < 0) { int e = 0; err = gzerror(g, &e); return 0; }
        return (size_t)n;
    }
    const char* error() const override { return err.empty() ? nullptr : err.c_str(); }
};
// snappy-java stream (SnappyOutputStream): 8-byte magic 82 'S' 'N' 'A' 'P' 'P' 'Y' 00, BE int version, BE int compatible version, then
// chunks [BE int32 compressed length][raw Snappy block]
struct SnapSource : ByteSource {
    FILE* f;
    std::string err;
    std::vector<char> out;
    size_t pos = 0;
    bool started = false;
    explicit SnapSource(FILE* f_) : f(f_) {}
    ~SnapSource() override { if (f) fclose(f); }
    static bool uncompress(const unsigned char* in, size_t n, std::vector<char>& o) {   // raw Snappy block format
        size_t p = 0, ulen = 0;
        int shift = 0;
        for (;;) {
            if (p >= n || shift > 35) return false;
            const unsigned char c = in[p++];
            ulen |= (size_t)(c & 0x7F) << shift;
            if (!(c & 0x80)) break;
            shift += 7;
        }
        o.clear();
        o.reserve(ulen);
        while (p < n) {
            const unsigned char tag = in[p++];
            size_t len, off;
            if ((tag & 3) == 0) {   // literal
                len = (size_t)(tag >> 2) + 1;
                if (len > 60) {
                    const size_t nb = len - 60;
                    if (p + nb > n) return false;
                    len = 0;
                    for (size_t i = 0; i < nb; ++i) len |= (size_t)in[p + i] << (8 * i);
                    len += 1; p += nb;
                }
                if (p + len > n) return false;
                o.insert(o.end(), (const char*)in + p, (const char*)in + p + len);
                p += len;
                continue;
            }
            if ((tag & 3) == 1) { if (p + 1 > n) return false; len = (size_t)((tag >> 2) & 7) + 4; off = ((size_t)(tag >> 5) << 8) | in[p]; p += 1; }
            else if ((tag & 3) == 2) { if (p + 2 > n) return false; len = (size_t)(tag >> 2) + 1; off = (size_t)in[p] | ((size_t)in[p + 1] << 8); p += 2; }
            else { if (p + 4 > n) return false; len = (size_t)(tag >> 2) + 1; off = (size_t)in[p] | ((size_t)in[p + 1] << 8) | ((size_t)in[p + 2] << 16) | ((size_t)in[p + 3] << 24); p += 4; }
            if (off == 0 || off > o.size()) return false;
            const size_t from = o.size() - off;
            for (size_t i = 0; i < len; ++i) o.push_back(o[from + i]);   // may overlap its own output
        }
        return o.size() == ulen;
    }
    bool next_chunk() {
        unsigned char h[16];
        if (!started) {
            started = true;
            static const unsigned char magic[8] = {0x82, 'S', 'N', 'A', 'P', 'P', 'Y', 0};
            if (fread(h, 1, 16, f) != 16 || memcmp(h, magic, 8) != 0) { err = "not a snappy-java stream"; return false; }
        }
        if (fread(h, 1, 4, f) != 4) return false;   // end of the stream
        const size_t clen = ((size_t)h[0] << 24) | ((size_t)h[1] << 16) | ((size_t)h[2] << 8) | (size_t)h[3];
        std::vector<unsigned char> c(clen);
        if (fread(c.data(), 1, clen, f) != clen || !uncompress(c.data(), clen, out)) { err = "corrupt snappy chunk"; return false; }
        pos = 0;
        return true;
    }
    size_t read(char* dst, size_t cap) override {
        while (pos >= out.size()) if (!next_chunk()) return 0;
        const size_t n = std::min(cap, out.size() - pos);
        memcpy(dst, out.data() + pos, n);
        pos += n;
        return n;
    }
    const char* error() const override { return err.empty() ? nullptr : err.c_str(); }
};

// a block of the read file that holds whole records: parsed into one batch by one of the parser threads
struct Chunk {
    uint64_t seq = 0;
    std::vector<char> data;
};

// Cuts the byte stream into chunks of whole records — at most `max_reads` records and about `target` bytes each — so that several
// threads can parse them side by side; the batches keep the order of the file through their sequence numbers.
class RecordSplitter {
  public:
    RecordSplitter(ByteSource* src, bool fasta, uint32_t max_reads, size_t target) : src_(src), fasta_(fasta), max_reads_(max_reads), target_(target) {}
    bool next(std::vector<char>& out) {
        for (;;) {
            // scan what is buffered: `cut_` is the end of the last whole record seen, `recs_` how many records end there
            while (scan_ < buf_.size()) {
                const char* base = buf_.data();
                const char* e = (const char*)memchr(base + scan_, '\n', buf_.size() - scan_);
                if (!e) break;
                const size_t ls = scan_, le = (size_t)(e - base);   // the line is [ls, le)
                scan_ = le + 1;
                if (fasta_) {
                    if (le > ls && base[ls] == '>') { if (seen_header_) { ++recs_; cut_ = ls; } seen_header_ = true; }
                } else {
                    const bool blank = le == ls || (le == ls + 1 && base[ls] == '\r');
                    if (line_in_rec_ == 0 && blank) { cut_ = scan_; continue; }   // blank lines between records belong to nobody
                    if (++line_in_rec_ == 4) { line_in_rec_ = 0; ++recs_; cut_ = scan_; }
                }
                if (recs_ >= max_reads_ || (cut_ >= target_ && recs_ > 0)) return emit(out, cut_);
            }
            if (eof_) {
                if (buf_.empty()) return false;
                return emit(out, buf_.size());   // the tail (a FASTA record ends at the end of the file; a cut FASTQ record is dropped by the parser)
            }
            const size_t old = buf_.size();
            buf_.resize(old + (8u << 20));
            const size_t got = src_->read(buf_.data() + old, 8u << 20);
            buf_.resize(old + got);
            if (got == 0) eof_ = true;
        }
    }
  private:
    bool emit(std::vector<char>& out, size_t upto) {
        out.assign(buf_.begin(), buf_.begin() + (long)upto);
        buf_.erase(buf_.begin(), buf_.begin() + (long)upto);
        scan_ -= std::min(scan_, upto);
        cut_ = 0; recs_ = 0;
        if (fasta_) seen_header_ = scan_ > 0 ? seen_header_ : false;
        return true;
    }
    ByteSource* src_;
    bool fasta_;
    uint32_t max_reads_;
    size_t target_;
    std::vector<char> buf_;
    size_t scan_ = 0, cut_ = 0;
    uint32_t recs_ = 0;
    int line_in_rec_ = 0;
    bool seen_header_ = false, eof_ = false;
};

// ---- bounded queues between the pipeline stages ---------------------------------------------------------------------
template <class T>
class Channel {
  public:
    explicit Channel(size_t cap) : cap_(cap) {}
    void push(T v) {
        std::unique_lock<std::mutex> g(m_);
        not_full_.wait(g, [&] { return q_.size() < cap_ || closed_; });
        q_.push_back(std::move(v));
        not_empty_.notify_one();
    }
    bool pop(T& v) {
        std::unique_lock<std::mutex> g(m_);
        not_empty_.wait(g, [&] { return !q_.empty() || closed_; });
        if (q_.empty()) return false;
        v = std::move(q_.front());
        q_.pop_front();
        not_full_.notify_one();
        return true;
    }
    void close() {
        std::lock_guard<std::mutex> g(m_);
        closed_ = true;
        not_empty_.notify_all();
        not_full_.notify_all();
    }
  private:
    std::mutex m_;
    std::condition_variable not_full_, not_empty_;
    std::deque<T> q_;
    size_t cap_;
    bool closed_ = false;
};

typedef std::unique_ptr<Batch> BatchPtr;

struct Options {
    std::string cmd, ref, query;
    std::vector<std::string> files;
    gw_config_t cfg;
    bool silent = false, stats = false, keep_going = false;
    int gpus = 1, device = 0;
    uint32_t batch = 1u << 20;
    std::vector<std::pair<std::string, long long>> tunables;
};

int usage() {
    fprintf(stderr,
            "usage: genome-weaver bwt <fasta> [--device D]\n"
            "       genome-weaver align -r <ref.fa> [-q <query> | <reads.fastq|.fq|.fa|.fasta|.fan[.gz|.snap]>] [-k K] [-s S] [-M n] [-N n] [-S n] [-P n] [-W n]\n"
            "                           [-R besthit|allhits|topL] [-L n] [-m bsf] [--silent] [--gpus N] [--device D] [--batch reads] [--stats]\n"
            "                           [--keep-going] [--extended] [--opt name=value]\n"
            "       reads of 1..256 bases with at most 25 differences are aligned; others are written as unmapped records and make the exit\n"
            "       status 2 unless --keep-going is given; --extended aligns the long reads (156 bases and more) on which the reference's\n"
            "       suffix filter throws, instead of ending the run there\n");
    return 1;
}

// one GPU: its own index replica and aligner; takes batches as they come, hands SAM text to the writer
void worker(int device, const Options* o, Channel<BatchPtr>* in, Channel<BatchPtr>* out, std::string* err, double* busy_s, double* open_s) {
    const auto t_open = std::chrono::steady_clock::now();
    gw_index_t* ix = nullptr;
    gw_aligner_t* al = nullptr;
    std::mutex err_m;   // the align and the format stage may both report
    auto fail = [&](const char* what) { std::lock_guard<std::mutex> g(err_m); if (err->empty()) *err = std::string(what) + ": " + gw_last_error(); };
    auto failed = [&] { std::lock_guard<std::mutex> g(err_m); return !err->empty(); };
    if (gw_index_open(o->ref.c_str(), device, &ix) != GW_OK) fail("index");
    else if (gw_aligner_create(ix, &o->cfg, &al) != GW_OK) fail("aligner");
    if (err->empty()) for (auto& t : o->tunables) if (gw_aligner_set_option(al, t.first.c_str(), t.second) != GW_OK) { fail("option"); break; }
    *open_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_open).count();
    // two stages per GPU, so that the device aligns the next batch while the host threads format the previous one:
    // align (this thread) -> format (SAM text, gw_sam_format uses the host threads) -> writer
    Channel<BatchPtr> mid(2);
    std::thread formatter([&] {
        std::vector<uint32_t> lines;
        BatchPtr b;
        while (mid.pop(b)) {
            if (b->aligned) {
                lines.assign(b->n, 0);
                uint64_t len = b->bases.size() * 3 + b->names.size() * 2 + (size_t)b->n * 96 + 64;
                int rc;
                for (;;) {
                    b->sam.resize(len);
                    rc = gw_sam_format(ix, b->names.data(), b->name_off.data(), b->bases.data(), b->base_off.data(),
                                       b->has_qual ? b->quals.data() : nullptr, b->has_qual ? b->qual_off.data() : nullptr, b->n, b->hits.data(), b->nh,
                                       b->cig.data(), &b->sam[0], &len, lines.data());
                    if (rc != GW_E_CAPACITY) break;
                    len += 64;
                }
                if (rc != GW_OK) { b->sam.clear(); fail("sam"); }
                else {
                    b->sam.resize(len);
                    if (b->abort_read >= 0 && !o->keep_going) {   // keep the lines of the reads before the aborting one
                        uint64_t keep = 0;
                        for (int64_t r = 0; r < b->abort_read; ++r) keep += lines[(size_t)r];
                        size_t p = 0;
                        while (keep && p < b->sam.size()) { if (b->sam[p] == '\n') --keep; ++p; }
                        b->sam.resize(p);
                    }
                }
                std::vector<gw_hit_t>().swap(b->hits);
                std::vector<uint32_t>().swap(b->cig);
            }
            out->push(std::move(b));
        }
    });
    BatchPtr b;
    // a failed worker still drains its share so the pipeline ends
    while (in->pop(b)) {
        if (!failed() && b->seq == 0 && b->n == 0) {   // start token: SAMOutput.init, "@SQ" lines from this worker's index
            uint64_t n = 0;
            gw_sam_header(ix, nullptr, &n);
            b->sam.resize(n);
            if (n && gw_sam_header(ix, &b->sam[0], &n) != GW_OK) fail("header");
        }
        if (!failed() && b->n) {
            const auto t0 = std::chrono::steady_clock::now();
            b->hits.resize((size_t)b->n * 2 + 16);
            b->cig.resize((size_t)b->n * 8 + 64);
            int rc;
            for (;;) {
                b->nh = b->hits.size(); b->nc = b->cig.size();
                rc = gw_align_batch(al, b->bases.data(), b->base_off.data(), b->n, b->hits.data(), &b->nh, b->cig.data(), &b->nc, &b->counters);
                if (rc != GW_E_CAPACITY) break;
                // the batch is aligned, only the buffers were too small: fetch the retained records instead of aligning again
                b->hits.resize(b->nh + 16); b->cig.resize(b->nc + 64);
                b->nh = b->hits.size(); b->nc = b->cig.size();
                rc = gw_align_batch_fetch(al, b->hits.data(), &b->nh, b->cig.data(), &b->nc);
                break;
            }
            if (rc != GW_OK) fail("align");
            else {
                for (uint64_t i = 0; i < b->nh; ++i) {
                    if (!b->hits[i].is_head) continue;
                    if (b->hits[i].status == GW_READ_UNSUPPORTED) { if (!b->n_unsupported++) b->first_unsupported = b->hits[i].read_index; }
                    // Align.java has no recovery: the first read the reference throws on ends its run
                    if (b->hits[i].status == GW_READ_REF_ABORT && b->abort_read < 0) b->abort_read = b->hits[i].read_index;
                }
                b->aligned = true;
            }
            *busy_s += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        }
        mid.push(std::move(b));
    }
    mid.close();
    formatter.join();
    if (al) gw_aligner_destroy(al);
    if (ix) gw_index_close(ix);
}

bool parse_ll(const char* s, long long* v) {
    char* e = nullptr;
    *v = strtoll(s, &e, 10);
    return e && *e == 0 && e != s;
}

int cmd_bwt(int argc, char** argv) {
    std::string fasta;
    long long device = 0;
    for (int i = 0; i < argc; ++i) {
        if (!strcmp(argv[i], "--device") && i + 1 < argc) { if (!parse_ll(argv[++i], &device)) return usage(); }
        else if (argv[i][0] == '-') return usage();
        else fasta = argv[i];
    }
    if (fasta.empty()) return usage();
    if (gw_index_build(fasta.c_str(), (int)device) != GW_OK) {
        fprintf(stderr, "genome-weaver bwt: %s\n", gw_last_error());
        return 1;
    }
    return 0;
}

int cmd_align(int argc, char** argv) {
    Options o;
    gw_config_default(&o.cfg);
    for (int i = 0; i < argc; ++i) {
        const std::string a = argv[i];
        auto need = [&](long long* v) { return i + 1 < argc && parse_ll(argv[++i], v); };
        long long v = 0;
        if (a == "-r" && i + 1 < argc) o.ref = argv[++i];
        else if (a == "-q" && i + 1 < argc) o.query = argv[++i];
        else if (a == "-k" && i + 1 < argc) o.cfg.k = strtof(argv[++i], nullptr);
        else if (a == "-s") { if (!need(&v)) return usage(); o.cfg.num_split_allowed = (int)v; }
        else if (a == "-M") { if (!need(&v)) return usage(); o.cfg.match_score = (int)v; }
        else if (a == "-N") { if (!need(&v)) return usage(); o.cfg.mismatch_penalty = (int)v; }
        else if (a == "-S") { if (!need(&v)) return usage(); o.cfg.split_open_penalty = (int)v; }
        else if (a == "-P") { if (!need(&v)) return usage(); o.cfg.indel_end_skip = (int)v; }
        else if (a == "-W") { if (!need(&v)) return usage(); o.cfg.band_width = (int)v; }
        else if (a == "-L") { if (!need(&v)) return usage(); o.cfg.top_l = (int)v; }
        else if ((a == "-g" || a == "-e" || a == "-G" || a == "-E" || a == "--bssN") && i + 1 < argc) ++i;   // parsed by the reference, unused on the BSF path
        else if (a == "--bss") {}
        else if (a == "-R" && i + 1 < argc) {
            std::string r = argv[++i];
            for (char& c : r) c = (char)tolower((unsigned char)c);
            if (r == "besthit") o.cfg.report_type = 0;
            else if (r == "allhits") o.cfg.report_type = 1;
            else if (r == "topl") o.cfg.report_type = 2;
            else { fprintf(stderr, "unknown report type: %s\n", argv[i]); return 1; }
        } else if (a == "-m" && i + 1 < argc) {
            std::string m = argv[++i];
            for (char& c : m) c = (char)tolower((unsigned char)c);
            if (m != "bsf") { fprintf(stderr, "%s mode is not supported (this build accelerates -m bsf only)\n", argv[i]); return 1; }
        } else if (a == "--silent") o.silent = true;
        else if (a == "--stats") o.stats = true;
        else if (a == "--keep-going") o.keep_going = true;
        else if (a == "--extended") o.cfg.extended = 1;
        else if (a == "--gpus") { if (!need(&v) || v < 1) return usage(); o.gpus = (int)v; }
        else if (a == "--device") { if (!need(&v)) return usage(); o.device = (int)v; }
        else if (a == "--batch") { if (!need(&v) || v < 1) return usage(); o.batch = (uint32_t)v; }
        else if (a == "--opt" && i + 1 < argc) {
            const std::string kv = argv[++i];
            const size_t eq = kv.find('=');
            if (eq == std::string::npos || !parse_ll(kv.c_str() + eq + 1, &v)) return usage();
            o.tunables.push_back({kv.substr(0, eq), v});
        } else if (!a.empty() && a[0] == '-') return usage();
        else o.files.push_back(a);
    }
    if (o.ref.empty()) { fprintf(stderr, "no reference (-r) is given\n"); return 1; }
    if (o.query.empty() && o.files.empty()) { fprintf(stderr, "no query is given\n"); return 1; }   // Align.java:79-80

    // open the read source before any device work (ReadReaderFactory.createReader :126-153: unsupported types fail up front;
    // ".gz" / ".snap" wrap the stream, the extension underneath picks the record format)
    std::unique_ptr<ByteSource> src;
    bool fasta_reads = false;
    if (o.query.empty()) {
        const std::string& p = o.files[0];
        std::string prefix = p;
        const bool gz = ends_with(p, ".gz"), snap = ends_with(p, ".snap");
        if (gz) prefix = p.substr(0, p.size() - 3);
        if (snap) prefix = p.substr(0, p.size() - 5);
        if (ends_with(prefix, ".fa") || ends_with(prefix, ".fasta") || ends_with(prefix, ".fan")) fasta_reads = true;
        else if (!(ends_with(prefix, ".fastq") || ends_with(prefix, ".fq"))) { fprintf(stderr, "Unsupported file type: %s\n", p.c_str()); return 1; }
        if (gz) {
            gzFile g = gzopen(p.c_str(), "rb");
            if (!g) { fprintf(stderr, "cannot open %s\n", p.c_str()); return 1; }
            src.reset(new GzSource(g));
        } else {
            FILE* f = fopen(p.c_str(), "rb");
            if (!f) { fprintf(stderr, "cannot open %s\n", p.c_str()); return 1; }
            if (snap) src.reset(new SnapSource(f)); else src.reset(new PlainSource(f));
        }
    }

    const auto t_start = std::chrono::steady_clock::now();
    const int G = o.gpus;
    Channel<BatchPtr> to_gpu((size_t)G * 2), done((size_t)G * 4);
    std::vector<std::thread> workers;
    std::vector<std::string> errs((size_t)G);
    std::vector<double> busy((size_t)G, 0.0), opened((size_t)G, 0.0);
    for (int g = 0; g < G; ++g)
        workers.emplace_back(worker, o.device + g, &o, &to_gpu, &done, &errs[(size_t)g], &busy[(size_t)g], &opened[(size_t)g]);

    // writer: batches in input order; stops printing at the read the reference would have died on
    uint64_t n_batches = 0, total_reads = 0;
    gw_counters_t tot{};
    bool aborted = false;
    uint64_t unsupported = 0;
    std::thread writer([&] {
        std::map<uint64_t, BatchPtr> held;
        uint64_t next = 0;
        std::vector<char> obuf(8u << 20);
        setvbuf(stdout, obuf.data(), _IOFBF, obuf.size());
        BatchPtr b;
        while (done.pop(b)) {
            held[b->seq] = std::move(b);
            for (auto it = held.find(next); it != held.end(); it = held.find(next)) {
                Batch& cur = *it->second;
                if (!aborted) {
                    if (!o.silent) fwrite(cur.sam.data(), 1, cur.sam.size(), stdout);
                    total_reads += cur.n;
                    tot.fm_searches += cur.counters.fm_searches; tot.sw_calls += cur.counters.sw_calls;
                    tot.cutoffs += cur.counters.cutoffs; tot.filtered += cur.counters.filtered;
                    tot.occ_lookups_scan += cur.counters.occ_lookups_scan; tot.occ_lookups_search += cur.counters.occ_lookups_search;
                    tot.occ_lookups_locate += cur.counters.occ_lookups_locate; tot.gpu_launches += cur.counters.gpu_launches;
                    if (cur.n_unsupported) {
                        if (!unsupported)
                            fprintf(stderr, "genome-weaver align: read #%llu (%s) is outside the supported domain (1..256 bases, at most 25 differences); it is written as an unmapped record\n",
                                    (unsigned long long)(total_reads - cur.n + (uint64_t)cur.first_unsupported),
                                    cur.names.substr(cur.name_off[(size_t)cur.first_unsupported], cur.name_off[(size_t)cur.first_unsupported + 1] - cur.name_off[(size_t)cur.first_unsupported]).c_str());
                        unsupported += cur.n_unsupported;
                    }
                    if (cur.abort_read >= 0 && !o.keep_going) {
                        aborted = true;
                        fprintf(stderr, "genome-weaver align: the reference aborts on read #%llu (%s); output ends there like the reference's (use --keep-going to skip such reads)\n",
                                (unsigned long long)(total_reads - cur.n + (uint64_t)cur.abort_read),
                                cur.names.substr(cur.name_off[(size_t)cur.abort_read], cur.name_off[(size_t)cur.abort_read + 1] - cur.name_off[(size_t)cur.abort_read]).c_str());
                    }
                }
                held.erase(it);
                ++next;
            }
        }
        fflush(stdout);
    });

    // reader (this thread): batch 0 is the start token that carries the SAM header back
    {
        BatchPtr tok(new Batch());
        tok->seq = n_batches++;
        to_gpu.push(std::move(tok));
    }
    std::string read_err;
    if (!o.query.empty()) {   // ReadReaderFactory.singleQueryReader :165-190: one read named "read", no quality
        BatchPtr b(new Batch());
        b->seq = n_batches++;
        b->add("read", o.query, nullptr);
        to_gpu.push(std::move(b));
    } else {
        // this thread cuts the stream into chunks of whole records; the parser threads turn chunks into batches side by side (a batch
        // carries the sequence number of its chunk, the writer puts the SAM text back in that order)
        const unsigned P = std::max(2u, std::min(8u, std::thread::hardware_concurrency() / 4u));
        Channel<std::unique_ptr<Chunk>> chunks(P * 2);
        std::vector<std::thread> parsers;
        for (unsigned t = 0; t < P; ++t)
            parsers.emplace_back([&] {
                std::unique_ptr<Chunk> c;
                while (chunks.pop(c)) {
                    BatchPtr b(new Batch());
                    b->seq = c->seq;
                    b->bases.reserve(c->data.size() / 2);
                    if (fasta_reads) parse_fasta(c->data.data(), c->data.size(), *b); else parse_fastq(c->data.data(), c->data.size(), *b);
                    c.reset();
                    to_gpu.push(std::move(b));   // (an empty batch still travels: the writer waits for every sequence number)
                }
            });
        RecordSplitter split(src.get(), fasta_reads, o.batch, (size_t)64 << 20);
        for (;;) {
            std::unique_ptr<Chunk> c(new Chunk());
            if (!split.next(c->data)) break;
            c->seq = n_batches++;
            chunks.push(std::move(c));
        }
        if (src->error()) read_err = src->error();
        chunks.close();
        for (auto& t : parsers) t.join();
    }
    to_gpu.close();
    for (auto& w : workers) w.join();
    done.close();
    writer.join();
    int rc = 0;
    if (!read_err.empty()) { fprintf(stderr, "genome-weaver align: %s: %s\n", o.files[0].c_str(), read_err.c_str()); rc = 1; }
    for (int g = 0; g < G; ++g)
        if (!errs[(size_t)g].empty()) { fprintf(stderr, "genome-weaver align (device %d): %s\n", o.device + g, errs[(size_t)g].c_str()); rc = 1; }
    if (unsupported) {
        fprintf(stderr, "genome-weaver align: %llu read(s) outside the supported domain were written as unmapped records%s\n", (unsigned long long)unsupported,
                o.keep_going ? "" : " (exit status 2; --keep-going accepts them)");
        if (!o.keep_going && rc == 0) rc = 2;
    }
    if (o.stats) {
        const double wall = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_start).count();
        double b = 0, op = 0;
        for (double x : busy) b += x;
        for (double x : opened) op = std::max(op, x);
        fprintf(stderr,
                "{\"reads\": %llu, \"wall_s\": %.3f, \"reads_per_s\": %.0f, \"index_open_s\": %.3f, \"reads_per_s_after_open\": %.0f, \"gpus\": %d, \"worker_busy_s\": %.3f, \"fm_searches\": %llu, \"sw_calls\": %llu, "
                "\"cutoffs\": %llu, \"filtered\": %llu, \"occ_lookups\": %llu, \"gpu_launches\": %llu, \"aborted\": %s}\n",
                (unsigned long long)total_reads, wall, wall > 0 ? total_reads / wall : 0.0, op, wall > op ? total_reads / (wall - op) : 0.0, G, b, (unsigned long long)tot.fm_searches,
                (unsigned long long)tot.sw_calls, (unsigned long long)tot.cutoffs, (unsigned long long)tot.filtered,
                (unsigned long long)(tot.occ_lookups_scan + tot.occ_lookups_search + tot.occ_lookups_locate), (unsigned long long)tot.gpu_launches,
                aborted ? "true" : "false");
    }
    return rc;
}

}  // namespace

int main(int argc, char** argv) {
    if (argc < 2) return usage();
    const std::string cmd = argv[1];
    if (cmd == "bwt") return cmd_bwt(argc - 2, argv + 2);
    if (cmd == "align") return cmd_align(argc - 2, argv + 2);
    return usage();
}
