// gw_files.cpp — the on-disk layout of `genome-weaver bwt` (reference: align/src/main/java/org/utgenome/weaver/align/):
//   <p>.f.3bit .r.3bit .f.bwt .r.bwt   ACGTSequence.saveTo/loadFrom  ACGTSequence.java:391-408
//        BE int64 numBases, then a snappy-java stream of long[3*ceil(n/64)] in little-endian byte order;
//        per 64 bases: word0 = N flags (base i at bit 63-i), word1 = bases 0-31, word2 = bases 32-63,
//        2 bits per base MSB first (ACGTSequence.java:47-58, :165-192)
//   <p>.f.ssa .r.ssa                   SparseSuffixArray.saveTo/loadFrom  SparseSuffixArray.java:91-122
//   <p>.i.silk                         PackFasta.encodeFASTA :110-130 / SequenceBoundary.loadSilk :123-139
//   <p> = FASTA path minus .tar.gz / .gz   BWTFiles.java:40-56
// snappy-java (org.xerial.snappy) is an un-vendored dependency of the reference with no pinned version; its stream
// container is restated from its published format: 8-byte magic 82 'S' 'N' 'A' 'P' 'P' 'Y' 00, BE int32 version (1),
// BE int32 compatible version (1), then chunks [BE int32 compressed length][raw Snappy block] of at most 32 KiB of input.
// The writer emits literal-only Snappy blocks (valid for any Snappy decoder); the reader decodes full Snappy.
#include <stdio.h>
#include <string.h>

#include <algorithm>

#include "gw_host.h"

namespace {

struct ByteSink {
    std::vector<uint8_t> d;
    void be32(uint32_t v) { for (int s = 24; s >= 0; s -= 8) d.push_back((uint8_t)(v >> s)); }
    void be64(uint64_t v) { for (int s = 56; s >= 0; s -= 8) d.push_back((uint8_t)(v >> s)); }
    void raw(const void* p, size_t n) { const uint8_t* b = (const uint8_t*)p; d.insert(d.end(), b, b + n); }
};

uint32_t rd_be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }
uint64_t rd_be64(const uint8_t* p) { return ((uint64_t)rd_be32(p) << 32) | rd_be32(p + 4); }

std::vector<uint8_t> slurp(const std::string& path) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) throw GwIoError("cannot open " + path);
    std::vector<uint8_t> d;
    fseek(f, 0, SEEK_END);
    long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    if (sz > 0) {
        d.resize((size_t)sz);
        if (fread(d.data(), 1, (size_t)sz, f) != (size_t)sz) { fclose(f); throw GwIoError("short read on " + path); }
    }
    fclose(f);
    return d;
}
void spill(const std::string& path, const std::vector<uint8_t>& d) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) throw GwIoError("cannot write " + path);
    if (!d.empty() && fwrite(d.data(), 1, d.size(), f) != d.size()) { fclose(f); throw GwIoError("short write on " + path); }
    fclose(f);
}

const uint8_t kSnappyMagic[8] = {0x82, 'S', 'N', 'A', 'P', 'P', 'Y', 0x00};
const size_t kSnappyChunk = 32 * 1024;

// one raw Snappy block holding `n` bytes as literals
void snappy_literal_block(const uint8_t* src, size_t n, ByteSink& out) {
    for (size_t v = n;;) {  // varint length preamble
        if (v < 0x80) { out.d.push_back((uint8_t)v); break; }
        out.d.push_back((uint8_t)(v & 0x7F) | 0x80);
        v >>= 7;
    }
    for (size_t done = 0; done < n;) {
        size_t len = std::min<size_t>(n - done, 1u << 16);
        size_t enc = len - 1;
        if (enc < 60) out.d.push_back((uint8_t)(enc << 2));
        else if (enc < 0x100) { out.d.push_back(60u << 2); out.d.push_back((uint8_t)enc); }
        else { out.d.push_back(61u << 2); out.d.push_back((uint8_t)(enc & 0xFF)); out.d.push_back((uint8_t)(enc >> 8)); }
        out.raw(src + done, len);
        done += len;
    }
}

// decodes one raw Snappy block, appending to out
void snappy_decode_block(const uint8_t* p, size_t n, std::vector<uint8_t>& out, const std::string& what) {
    size_t i = 0, want = 0;
    for (int shift = 0;; shift += 7) {
        if (i >= n || shift > 35) throw GwIoError("bad snappy preamble in " + what);
        uint8_t b = p[i++];
        want |= (size_t)(b & 0x7F) << shift;
        if (!(b & 0x80)) break;
    }
    const size_t base = out.size();
    while (i < n) {
        const uint8_t tag = p[i++];
        size_t len, off;
        switch (tag & 3) {
        case 0: {
            len = tag >> 2;
            if (len >= 60) {
                size_t nb = len - 59;
                if (i + nb > n) throw GwIoError("truncated snappy literal in " + what);
                len = 0;
                for (size_t j = 0; j < nb; ++j) len |= (size_t)p[i + j] << (8 * j);
                i += nb;
            }
            ++len;
            if (i + len > n) throw GwIoError("truncated snappy literal in " + what);
            out.insert(out.end(), p + i, p + i + len);
            i += len;
            continue;
        }
        case 1:
            if (i + 1 > n) throw GwIoError("truncated snappy copy in " + what);
            len = ((tag >> 2) & 7) + 4; off = ((size_t)(tag >> 5) << 8) | p[i]; i += 1;
            break;
        case 2:
            if (i + 2 > n) throw GwIoError("truncated snappy copy in " + what);
            len = (tag >> 2) + 1; off = (size_t)p[i] | ((size_t)p[i + 1] << 8); i += 2;
            break;
        default:
            if (i + 4 > n) throw GwIoError("truncated snappy copy in " + what);
            len = (tag >> 2) + 1; off = (size_t)p[i] | ((size_t)p[i + 1] << 8) | ((size_t)p[i + 2] << 16) | ((size_t)p[i + 3] << 24); i += 4;
        }
        if (off == 0 || off > out.size() - base) throw GwIoError("bad snappy offset in " + what);
        for (size_t j = 0, from = out.size() - off; j < len; ++j) out.push_back(out[from + j]);  // may overlap: byte by byte
    }
    if (out.size() - base != want) throw GwIoError("snappy length mismatch in " + what);
}

std::string strip(const std::string& s) {
    size_t a = 0, b = s.size();
    while (a < b && (s[a] == ' ' || s[a] == '\t' || s[a] == '\r' || s[a] == '\n')) ++a;
    while (b > a && (s[b - 1] == ' ' || s[b - 1] == '\t' || s[b - 1] == '\r' || s[b - 1] == '\n')) --b;
    return s.substr(a, b - a);
}

inline int code_of(unsigned char ch) {  // ACGT.java:37-44
    switch (ch) {
    case 'A': case 'a': return 0;
    case 'C': case 'c': return 1;
    case 'G': case 'g': return 2;
    case 'T': case 't': case 'U': case 'u': return 3;
    default: return 4;
    }
}

}  // namespace

std::string gw_index_prefix(const std::string& fasta) {
    auto ends = [&](const char* suf) { size_t n = strlen(suf); return fasta.size() >= n && fasta.compare(fasta.size() - n, n, suf) == 0; };
    if (ends(".tar.gz")) return fasta.substr(0, fasta.size() - 7);
    if (ends(".gz")) return fasta.substr(0, fasta.size() - 3);
    return fasta;
}

void gw_pack_fasta(const std::string& fasta_path, std::vector<uint8_t>& codes, std::vector<GwContig>& contigs) {
    std::vector<uint8_t> d = slurp(fasta_path);
    codes.clear();
    contigs.clear();
    codes.reserve(d.size());
    bool open = false;
    GwContig cur;
    auto close = [&]() {
        if (!open) return;
        cur.length = (int64_t)codes.size() - cur.offset;
        contigs.push_back(cur);
    };
    size_t i = 0;
    while (i < d.size()) {
        size_t e = i;
        while (e < d.size() && d[e] != '\n') ++e;
        if (e > i && d[i] == '>') {
            close();
            open = true;
            cur = GwContig();
            cur.desc = strip(std::string((const char*)d.data() + i + 1, e - i - 1));
            size_t sp = cur.desc.find_first_of(" \t");   // CompactFASTA.pickSequenceName: first token of the description line
            cur.name = sp == std::string::npos ? cur.desc : cur.desc.substr(0, sp);
            cur.offset = (int64_t)codes.size();
        } else if (open) {
            size_t a = i, b = e;   // sequence lines are trimmed (PackFasta.java:100-104)
            while (a < b && (d[a] == ' ' || d[a] == '\t' || d[a] == '\r')) ++a;
            while (b > a && (d[b - 1] == ' ' || d[b - 1] == '\t' || d[b - 1] == '\r')) --b;
            for (size_t j = a; j < b; ++j) codes.push_back((uint8_t)code_of(d[j]));
        }
        i = e + 1;
    }
    close();
}

void gw_save_acgt(const std::string& path, const uint8_t* codes, uint64_t n) {
    const uint64_t groups = (n + 63) / 64;
    std::vector<uint64_t> w((size_t)groups * 3, 0);
    for (uint64_t i = 0; i < n; ++i) {
        const uint64_t g = i >> 6;
        const unsigned off = (unsigned)(i & 63);
        const uint64_t c = codes[i];
        w[g * 3] |= ((c >> 2) & 1ULL) << (63 - off);
        w[g * 3 + 1 + (off >> 5)] |= (c & 3ULL) << (62 - 2 * (off & 31));
    }
    ByteSink o;
    o.be64(n);
    o.raw(kSnappyMagic, 8);
    o.be32(1);
    o.be32(1);
    const uint8_t* raw = (const uint8_t*)w.data();   // x86: little-endian, the JVM's native order there
    const size_t total = w.size() * 8;
    for (size_t pos = 0; pos < total; pos += kSnappyChunk) {
        ByteSink blk;
        snappy_literal_block(raw + pos, std::min(kSnappyChunk, total - pos), blk);
        o.be32((uint32_t)blk.d.size());
        o.raw(blk.d.data(), blk.d.size());
    }
    spill(path, o.d);
}

void gw_load_acgt(const std::string& path, std::vector<uint8_t>& codes) {
    std::vector<uint8_t> d = slurp(path);
    if (d.size() < 8) throw GwIoError("truncated " + path);
    const uint64_t n = rd_be64(d.data());
    if (n >= 0xFFFFFFFFull) throw GwIoError("sequences of 2^32 - 1 symbols or more are not supported: " + path);
    size_t p = 8;
    if (d.size() >= p + 16 && memcmp(d.data() + p, kSnappyMagic, 8) == 0) p += 16;
    std::vector<uint8_t> raw;
    raw.reserve((size_t)std::min<uint64_t>(((n + 63) / 64) * 24, (uint64_t)d.size() * 64));
    while (p + 4 <= d.size()) {
        const uint32_t clen = rd_be32(d.data() + p);
        p += 4;
        if (p + clen > d.size()) throw GwIoError("truncated snappy chunk in " + path);
        snappy_decode_block(d.data() + p, clen, raw, path);
        p += clen;
    }
    const uint64_t groups = (n + 63) / 64;
    if (raw.size() < groups * 24) throw GwIoError("short payload in " + path);
    codes.resize((size_t)n);
    uint8_t* out = codes.data();
    const uint8_t* in = raw.data();
    gw_parallel_for(groups, gw_host_threads(), [&](uint64_t g0, uint64_t g1, unsigned) {
        for (uint64_t g = g0; g < g1; ++g) {
            uint64_t w[3];
            memcpy(w, in + g * 24, 24);
            const unsigned cnt = (unsigned)std::min<uint64_t>(64, n - g * 64);
            for (unsigned off = 0; off < cnt; ++off) {
                const unsigned nf = (unsigned)((w[0] >> (63 - off)) & 1ULL);
                const unsigned c2 = (unsigned)((w[1 + (off >> 5)] >> (62 - 2 * (off & 31))) & 3ULL);
                out[(size_t)(g * 64 + off)] = (uint8_t)(nf ? 4 : c2);
            }
        }
    });
}

void gw_save_ssa(const std::string& path, uint64_t N, const std::vector<uint32_t>& samples) {
    ByteSink o;
    o.d.reserve(20 + samples.size() * 4);
    o.be64(N);
    o.be32(32);
    o.be64(samples.size());
    for (uint32_t v : samples) o.be32(v);
    spill(path, o.d);
}

void gw_load_ssa(const std::string& path, uint64_t* N, int* L, std::vector<uint32_t>& samples) {
    std::vector<uint8_t> d = slurp(path);
    if (d.size() < 20) throw GwIoError("truncated " + path);
    *N = rd_be64(d.data());
    *L = (int)rd_be32(d.data() + 8);
    const uint64_t cnt = rd_be64(d.data() + 12);
    if (cnt > (d.size() - 20) / 4) throw GwIoError("truncated samples in " + path);   // (also rejects counts whose byte size would wrap)
    if (*L <= 0 || cnt != (*N + (uint64_t)*L - 1) / (uint64_t)*L) throw GwIoError("sample count of " + path + " does not match its text size");
    if (*N >= 0xFFFFFFFFull) throw GwIoError("texts of 2^32 - 1 symbols or more are not supported: " + path);
    samples.resize((size_t)cnt);
    for (uint64_t i = 0; i < cnt; ++i) samples[(size_t)i] = rd_be32(d.data() + 20 + 4 * i);
}

void gw_save_silk(const std::string& path, const std::vector<GwContig>& contigs, uint64_t total) {
    std::string s;
    for (const GwContig& c : contigs) {
        s += "-index\n -name:" + c.name + "\n -desc:" + c.desc + "\n -length:" + std::to_string(c.length) + "\n -offset:" +
             std::to_string(c.offset) + "\n";
    }
    s += "-total size:" + std::to_string(total) + "\n";
    spill(path, std::vector<uint8_t>(s.begin(), s.end()));
}

void gw_load_silk(const std::string& path, std::vector<GwContig>& contigs, uint64_t* total) {
    std::vector<uint8_t> d = slurp(path);
    contigs.clear();
    *total = 0;
    auto assign = [&](const std::string& key, const std::string& val) {
        if (contigs.empty()) return;
        GwContig& c = contigs.back();
        if (key == "name") c.name = val;
        else if (key == "desc") c.desc = val;
        else if (key == "length") c.length = atoll(val.c_str());
        else if (key == "offset") c.offset = atoll(val.c_str());
    };
    size_t i = 0;
    while (i < d.size()) {
        size_t e = i;
        while (e < d.size() && d[e] != '\n') ++e;
        std::string t = strip(std::string((const char*)d.data() + i, e - i));
        i = e + 1;
        if (t.empty() || t[0] != '-') continue;
        t.erase(0, 1);
        if (t.compare(0, 5, "index") == 0) {
            contigs.push_back(GwContig());
            size_t lp = t.find('('), rp = t.rfind(')');
            if (lp != std::string::npos && rp != std::string::npos && rp > lp) {   // inline form -index(name:.., desc:.., ...)
                std::string body = t.substr(lp + 1, rp - lp - 1);
                size_t a = 0;
                while (a <= body.size()) {
                    size_t b = body.find(',', a);
                    if (b == std::string::npos) b = body.size();
                    std::string kv = body.substr(a, b - a);
                    size_t c = kv.find(':');
                    if (c != std::string::npos) assign(strip(kv.substr(0, c)), strip(kv.substr(c + 1)));
                    a = b + 1;
                }
            }
        } else if (t.compare(0, 10, "total size") == 0) {
            size_t c = t.find(':');
            if (c != std::string::npos) *total = (uint64_t)atoll(strip(t.substr(c + 1)).c_str());
        } else {
            size_t c = t.find(':');
            if (c != std::string::npos) assign(strip(t.substr(0, c)), strip(t.substr(c + 1)));
        }
    }
}

void gw_save_index_files(const std::string& fasta_path, const GwHostIndex& ix) {
    const std::string p = gw_index_prefix(fasta_path);
    gw_save_silk(p + ".i.silk", ix.contigs, ix.N);
    gw_save_acgt(p + ".f.3bit", ix.text.data(), ix.N);
    {
        std::vector<uint8_t> rev(ix.text.rbegin(), ix.text.rend());   // ACGTSequence.reverse (not complemented), PackFasta.java:132-138
        gw_save_acgt(p + ".r.3bit", rev.data(), ix.N);
    }
    gw_save_ssa(p + ".f.ssa", ix.N, ix.ssa[0]);
    gw_save_ssa(p + ".r.ssa", ix.N, ix.ssa[1]);
    gw_save_acgt(p + ".f.bwt", ix.bwt[0].data(), ix.N);
    gw_save_acgt(p + ".r.bwt", ix.bwt[1].data(), ix.N);
}

void gw_load_index_files(const std::string& fasta_path, GwHostIndex& ix) {
    const std::string p = gw_index_prefix(fasta_path);
    uint64_t total = 0;
    gw_load_silk(p + ".i.silk", ix.contigs, &total);
    ix.N = total;
    {   // the contig table must tile [0, N): offsets ascending, every contig inside the text, lengths summing to N
        uint64_t sum = 0, prev_off = 0;
        for (size_t i = 0; i < ix.contigs.size(); ++i) {
            const GwContig& c = ix.contigs[i];
            if (c.length < 0 || c.offset < 0 || (uint64_t)c.offset + (uint64_t)c.length > total || (i && (uint64_t)c.offset < prev_off))
                throw GwIoError("contig table of " + p + ".i.silk is inconsistent (contig " + c.name + ")");
            prev_off = (uint64_t)c.offset;
            sum += (uint64_t)c.length;
        }
        if (ix.contigs.empty() || sum != total) throw GwIoError("contig lengths of " + p + ".i.silk do not sum to the total size");
    }
    // the five files are independent: read and decode them side by side
    gw_parallel_for(5, 5, [&](uint64_t lo, uint64_t hi, unsigned) {
        for (uint64_t f = lo; f < hi; ++f) {
            if (f < 2) {
                const int s = (int)f;
                uint64_t n = 0;
                int L = 0;
                gw_load_ssa(p + (s == 0 ? ".f.ssa" : ".r.ssa"), &n, &L, ix.ssa[s]);
                if (L != 32) throw GwIoError("unsupported suffix array sampling (L != 32) in " + p);
                if (n != ix.N) throw GwIoError("index files of " + p + " disagree on the text size");
            } else if (f < 4) {
                const int s = (int)f - 2;
                gw_load_acgt(p + (s == 0 ? ".f.bwt" : ".r.bwt"), ix.bwt[s]);
                if (ix.bwt[s].size() != ix.N) throw GwIoError("index files of " + p + " disagree on the text size");
            } else {
                gw_load_acgt(p + ".f.3bit", ix.text);
                if (ix.text.size() != ix.N) throw GwIoError("reference text of " + p + " disagrees on the text size");
            }
        }
    });
}
