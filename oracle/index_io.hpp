// TEST INFRASTRUCTURE ONLY (see java_sem.hpp). On-disk formats of `genome-weaver bwt` and a small index builder:
//   .3bit/.bwt   ACGTSequence.saveTo/loadFrom   align/src/main/java/org/utgenome/weaver/align/ACGTSequence.java:391-408
//                (BE int64 numBases, then a snappy-java stream of the long[] in little-endian memory order)
//   .ssa         SparseSuffixArray.saveTo/loadFrom   .../align/SparseSuffixArray.java:91-122
//   .i.silk      PackFasta.encodeFASTA   .../align/PackFasta.java:110-130 ; SequenceBoundary.loadSilk :123-139
//   file names   BWTFiles   .../align/BWTFiles.java:40-80
//   build        BWTransform.pac2bwt / bwt   .../align/BWTransform.java:110-151, :163-179 (SA of CYCLIC rotations, L=32)
// snappy-java (org.xerial.snappy, un-vendored, version unpinned — "parity unpinned"): the stream framing is
// restated from its published format: magic 82 53 4E 41 50 50 59 00, BE int version=1, BE int compatibleVersion=1, then
// chunks [BE int32 compressedLength][raw Snappy block], each chunk <= 32 KiB uncompressed.
#pragma once
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>

#include "fm_index.hpp"

namespace gwo {

// ---------- raw Snappy block format (published format description) ----------
static inline void snappyCompressLiteral(const uint8_t* in, size_t n, std::vector<uint8_t>& out) {
    size_t v = n;  // preamble: uncompressed length varint
    while (v >= 0x80) { out.push_back((uint8_t)(v | 0x80)); v >>= 7; }
    out.push_back((uint8_t)v);
    size_t pos = 0;
    while (pos < n) {  // literal-only encoding: valid Snappy, no matching
        size_t len = std::min<size_t>(n - pos, 65536);
        size_t l1 = len - 1;
        if (l1 < 60) out.push_back((uint8_t)(l1 << 2));
        else if (l1 < 256) { out.push_back(60 << 2); out.push_back((uint8_t)l1); }
        else { out.push_back(61 << 2); out.push_back((uint8_t)(l1 & 0xFF)); out.push_back((uint8_t)(l1 >> 8)); }
        out.insert(out.end(), in + pos, in + pos + len);
        pos += len;
    }
}
static inline bool snappyUncompress(const uint8_t* in, size_t n, std::vector<uint8_t>& out) {
    size_t p = 0, ulen = 0;
    int shift = 0;
    for (;;) {
        if (p >= n) return false;
        uint8_t b = in[p++];
        ulen |= (size_t)(b & 0x7F) << shift;
        if (!(b & 0x80)) break;
        shift += 7;
    }
    size_t base = out.size();
    if (out.capacity() < base + ulen) out.reserve(std::max(base + ulen, out.capacity() * 2));   // geometric: the loaders append chunk after chunk
    while (p < n) {
        uint8_t tag = in[p++];
        int type = tag & 3;
        if (type == 0) {
            size_t len = (tag >> 2);
            if (len >= 60) {
                int nb = (int)len - 59;
                if (p + nb > n) return false;
                len = 0;
                for (int i = 0; i < nb; ++i) len |= (size_t)in[p + i] << (8 * i);
                p += nb;
            }
            len += 1;
            if (p + len > n) return false;
            out.insert(out.end(), in + p, in + p + len);
            p += len;
        } else {
            size_t len, off;
            if (type == 1) { if (p + 1 > n) return false; len = ((tag >> 2) & 7) + 4; off = ((size_t)(tag >> 5) << 8) | in[p]; p += 1; }
            else if (type == 2) { if (p + 2 > n) return false; len = (tag >> 2) + 1; off = in[p] | ((size_t)in[p + 1] << 8); p += 2; }
            else { if (p + 4 > n) return false; len = (tag >> 2) + 1; off = in[p] | ((size_t)in[p + 1] << 8) | ((size_t)in[p + 2] << 16) | ((size_t)in[p + 3] << 24); p += 4; }
            if (off == 0 || off > out.size() - base) return false;
            size_t from = out.size() - off;
            for (size_t i = 0; i < len; ++i) out.push_back(out[from + i]);
        }
    }
    return out.size() - base == ulen;
}

static inline void putBE64(std::vector<uint8_t>& o, uint64_t v) { for (int i = 7; i >= 0; --i) o.push_back((uint8_t)(v >> (8 * i))); }
static inline void putBE32(std::vector<uint8_t>& o, uint32_t v) { for (int i = 3; i >= 0; --i) o.push_back((uint8_t)(v >> (8 * i))); }
static inline uint64_t getBE64(const uint8_t* p) { uint64_t v = 0; for (int i = 0; i < 8; ++i) v = (v << 8) | p[i]; return v; }
static inline uint32_t getBE32(const uint8_t* p) { uint32_t v = 0; for (int i = 0; i < 4; ++i) v = (v << 8) | p[i]; return v; }

static inline std::vector<uint8_t> readFile(const std::string& path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("cannot open " + path);
    return std::vector<uint8_t>((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
}
static inline void writeFile(const std::string& path, const std::vector<uint8_t>& d) {
    std::ofstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("cannot write " + path);
    f.write((const char*)d.data(), (std::streamsize)d.size());
}

static const uint8_t SNAPPY_MAGIC[8] = {0x82, 'S', 'N', 'A', 'P', 'P', 'Y', 0};

static inline void saveACGTSequence(const ACGTSequence& s, const std::string& path) {
    std::vector<uint8_t> o;
    putBE64(o, (uint64_t)s.numBases);
    o.insert(o.end(), SNAPPY_MAGIC, SNAPPY_MAGIC + 8);
    putBE32(o, 1); putBE32(o, 1);
    size_t nLongs = (size_t)ACGTSequence::minArraySize(s.numBases);
    const uint8_t* raw = (const uint8_t*)s.seq.data();  // host is little-endian like the JVM's native order on x86
    size_t total = nLongs * 8, pos = 0;
    while (pos < total) {
        size_t len = std::min<size_t>(32 * 1024, total - pos);
        std::vector<uint8_t> c;
        snappyCompressLiteral(raw + pos, len, c);
        putBE32(o, (uint32_t)c.size());
        o.insert(o.end(), c.begin(), c.end());
        pos += len;
    }
    writeFile(path, o);
}
static inline ACGTSequence loadACGTSequence(const std::string& path) {
    std::vector<uint8_t> d = readFile(path);
    if (d.size() < 8) throw std::runtime_error("truncated " + path);
    jlong numBases = (jlong)getBE64(d.data());
    size_t p = 8;
    if (d.size() >= p + 16 && memcmp(d.data() + p, SNAPPY_MAGIC, 8) == 0) p += 16;
    std::vector<uint8_t> raw;
    raw.reserve((size_t)ACGTSequence::minArraySize(numBases) * 8 + 64);
    while (p + 4 <= d.size()) {
        uint32_t clen = getBE32(d.data() + p);
        p += 4;
        if (p + clen > d.size() || !snappyUncompress(d.data() + p, clen, raw)) throw std::runtime_error("bad snappy chunk in " + path);
        p += clen;
    }
    ACGTSequence s(numBases);
    size_t nLongs = (size_t)ACGTSequence::minArraySize(numBases);
    if (raw.size() < nLongs * 8) throw std::runtime_error("short payload in " + path);
    memcpy(s.seq.data(), raw.data(), nLongs * 8);
    return s;
}
static inline void saveSSA(const SparseSuffixArray& s, const std::string& path) {
    std::vector<uint8_t> o;
    putBE64(o, (uint64_t)s.N); putBE32(o, (uint32_t)s.L); putBE64(o, (uint64_t)s.sparseSA.size());
    for (uint64_t v : s.sparseSA) putBE32(o, (uint32_t)v);
    writeFile(path, o);
}
static inline SparseSuffixArray loadSSA(const std::string& path) {
    std::vector<uint8_t> d = readFile(path);
    SparseSuffixArray s;
    s.N = (jlong)getBE64(d.data()); s.L = (jint)getBE32(d.data() + 8);
    uint64_t len = getBE64(d.data() + 12);
    s.sparseSA.resize((size_t)len);
    for (uint64_t i = 0; i < len; ++i) s.sparseSA[(size_t)i] = getBE32(d.data() + 20 + 4 * i);
    return s;
}
static inline void saveSilk(const SequenceBoundary& b, const std::string& path) {
    std::ostringstream o;
    for (auto& e : b.index) o << "-index\n -name:" << e.name << "\n -desc:" << e.desc << "\n -length:" << e.length << "\n -offset:" << e.offset << "\n";
    o << "-total size:" << b.totalSize << "\n";
    std::string s = o.str();
    writeFile(path, std::vector<uint8_t>(s.begin(), s.end()));
}
static inline std::string trim(const std::string& s) {
    size_t a = s.find_first_not_of(" \t\r\n"), b = s.find_last_not_of(" \t\r\n");
    return a == std::string::npos ? "" : s.substr(a, b - a + 1);
}
static inline SequenceBoundary loadSilk(const std::string& path) {
    std::vector<uint8_t> d = readFile(path);
    std::istringstream in(std::string(d.begin(), d.end()));
    SequenceBoundary b;
    std::string line;
    auto setField = [&](const std::string& key, const std::string& val) {
        if (b.index.empty()) return;
        SequenceIndex& e = b.index.back();
        if (key == "name") e.name = val; else if (key == "desc") e.desc = val;
        else if (key == "length") e.length = std::stoll(val); else if (key == "offset") e.offset = std::stoll(val);
    };
    while (std::getline(in, line)) {
        std::string t = trim(line);
        if (t.empty() || t[0] != '-') continue;
        t = t.substr(1);
        if (t.compare(0, 5, "index") == 0) {
            b.index.push_back(SequenceIndex{"", "", 0, 0});
            size_t lp = t.find('(');
            if (lp != std::string::npos) {  // inline form -index(name:.., desc:.., length:.., offset:..)
                std::string body = t.substr(lp + 1, t.rfind(')') - lp - 1);
                std::istringstream bs(body);
                std::string kv;
                while (std::getline(bs, kv, ',')) {
                    size_t c = kv.find(':');
                    if (c != std::string::npos) setField(trim(kv.substr(0, c)), trim(kv.substr(c + 1)));
                }
            }
        } else if (t.compare(0, 10, "total size") == 0) {
            b.totalSize = std::stoll(trim(t.substr(t.find(':') + 1)));
        } else {
            size_t c = t.find(':');
            if (c != std::string::npos) setField(trim(t.substr(0, c)), trim(t.substr(c + 1)));
        }
    }
    return b;
}

static inline std::string indexPrefix(const std::string& fasta) {  // BWTFiles.java:40-47
    auto ends = [&](const char* suf) { size_t n = strlen(suf); return fasta.size() >= n && fasta.compare(fasta.size() - n, n, suf) == 0; };
    if (ends(".tar.gz")) return fasta.substr(0, fasta.size() - 7);
    if (ends(".gz")) return fasta.substr(0, fasta.size() - 3);
    return fasta;
}

// ---------- suffix array of CYCLIC rotations (what CyclicSAIS computes), by prefix doubling ----------
// ties that never resolve (non-primitive text) are broken by position.
static inline std::vector<uint32_t> cyclicSuffixSort(const std::vector<uint32_t>& sym) {
    const size_t n = sym.size();
    std::vector<uint32_t> sa(n), rank(sym), nrank(n);
    if (n == 0) return sa;
    std::vector<std::pair<uint64_t, uint32_t>> keyed(n);
    for (size_t h = 1;; h *= 2) {
        for (size_t i = 0; i < n; ++i) keyed[i] = {((uint64_t)rank[i] << 32) | rank[(i + h) % n], (uint32_t)i};
        std::sort(keyed.begin(), keyed.end());
        bool distinct = true;
        uint32_t r = 0;
        for (size_t j = 0; j < n; ++j) {
            if (j > 0 && keyed[j].first != keyed[j - 1].first) r = (uint32_t)j; else if (j > 0) distinct = false;
            nrank[keyed[j].second] = r;
        }
        rank.swap(nrank);
        if (distinct || h >= n) break;
    }
    for (size_t j = 0; j < n; ++j) sa[j] = keyed[j].second;
    return sa;
}
static inline std::vector<uint32_t> cyclicSuffixSort(const ACGTSequence& seq) {
    const size_t n = (size_t)seq.textSize();
    // pack 8 symbols (base 5) as the initial rank == prefix doubling started at h=8
    std::vector<uint32_t> rank(n), nrank(n), sa(n);
    if (n == 0) return sa;
    std::vector<uint8_t> s(n);
    for (size_t i = 0; i < n; ++i) s[i] = (uint8_t)seq.lookup((jlong)i);
    const size_t P = 8;
    for (size_t i = 0; i < n; ++i) { uint32_t v = 0; for (size_t j = 0; j < P; ++j) v = v * 5 + s[(i + j) % n]; rank[i] = v; }
    std::vector<std::pair<uint64_t, uint32_t>> keyed(n);
    for (size_t h = P;; h *= 2) {
        for (size_t i = 0; i < n; ++i) keyed[i] = {((uint64_t)rank[i] << 32) | rank[(i + h) % n], (uint32_t)i};
        std::sort(keyed.begin(), keyed.end());
        bool distinct = true;
        uint32_t r = 0;
        for (size_t j = 0; j < n; ++j) {
            if (j > 0 && keyed[j].first != keyed[j - 1].first) r = (uint32_t)j; else if (j > 0) distinct = false;
            nrank[keyed[j].second] = r;
        }
        rank.swap(nrank);
        if (distinct || h >= n) break;
    }
    for (size_t j = 0; j < n; ++j) sa[j] = keyed[j].second;
    return sa;
}
static inline ACGTSequence bwtOf(const ACGTSequence& seq, const std::vector<uint32_t>& SA) {  // BWTransform.java:172-179
    ACGTSequence bwt((jlong)SA.size());
    const jlong N = seq.textSize();
    for (size_t i = 0; i < SA.size(); ++i) bwt.set((jlong)i, seq.lookup(((jlong)SA[i] - 1 + N) % N));
    return bwt;
}

// FASTA -> (packed forward text, boundaries); PackFasta.encode :81-108
static inline void packFasta(const std::string& fastaText, ACGTSequence& packed, SequenceBoundary& b) {
    std::istringstream in(fastaText);
    std::string line;
    std::string bases;
    bool have = false;
    std::string name, desc;
    jlong offset = 0;
    std::vector<uint8_t> codes;
    auto flush = [&]() {
        if (!have) return;
        jlong pos = (jlong)codes.size();
        b.index.push_back(SequenceIndex{name, desc, pos - offset, offset});
        offset = pos;
    };
    while (std::getline(in, line)) {
        if (!line.empty() && line[0] == '>') {
            flush();
            have = true;
            desc = trim(line.substr(1));
            size_t e = desc.find_first_of(" \t");
            name = e == std::string::npos ? desc : desc.substr(0, e);
        } else if (have) {
            std::string t = trim(line);
            for (char c : t) codes.push_back((uint8_t)to3bitCode((unsigned char)c));
        }
    }
    flush();
    packed = ACGTSequence((jlong)codes.size());
    for (size_t i = 0; i < codes.size(); ++i) packed.set((jlong)i, codes[i]);
    b.totalSize = (jlong)codes.size();
}

// `genome-weaver bwt <fasta>`: writes <p>.i.silk, <p>.f.3bit, <p>.r.3bit, <p>.{f,r}.ssa, <p>.{f,r}.bwt
static inline void buildIndexFiles(const std::string& fastaPath) {
    std::vector<uint8_t> d = readFile(fastaPath);
    ACGTSequence fwd;
    SequenceBoundary b;
    packFasta(std::string(d.begin(), d.end()), fwd, b);
    std::string p = indexPrefix(fastaPath);
    saveSilk(b, p + ".i.silk");
    saveACGTSequence(fwd, p + ".f.3bit");
    ACGTSequence rev = fwd.reverse();
    saveACGTSequence(rev, p + ".r.3bit");
    for (int s = 0; s < 2; ++s) {
        const ACGTSequence& seq = s == 0 ? fwd : rev;
        std::vector<uint32_t> SA = cyclicSuffixSort(seq);
        saveSSA(SparseSuffixArray::buildFromSuffixArray(SA, 32), p + (s == 0 ? ".f.ssa" : ".r.ssa"));
        saveACGTSequence(bwtOf(seq, SA), p + (s == 0 ? ".f.bwt" : ".r.bwt"));
    }
}
static inline std::unique_ptr<FMIndexOnGenome> loadIndex(const std::string& fastaPath) {  // FMIndexOnGenome.load :60-86
    std::string p = indexPrefix(fastaPath);
    SequenceBoundary b = loadSilk(p + ".i.silk");
    SparseSuffixArray f = loadSSA(p + ".f.ssa"), r = loadSSA(p + ".r.ssa");
    ACGTSequence bf = loadACGTSequence(p + ".f.bwt"), br = loadACGTSequence(p + ".r.bwt");
    jlong n = b.totalSize;
    return std::unique_ptr<FMIndexOnGenome>(new FMIndexOnGenome(std::move(bf), std::move(br), std::move(f), std::move(r), std::move(b), n));
}
static inline std::unique_ptr<FMIndexOnGenome> buildFromSequence(const std::string& name, const ACGTSequence& refF) {  // :104-119
    ACGTSequence refR = refF.reverse();
    std::vector<uint32_t> saF = cyclicSuffixSort(refF), saR = cyclicSuffixSort(refR);
    SequenceBoundary b;
    b.totalSize = refF.textSize();
    b.index.push_back(SequenceIndex{name, name, refF.textSize(), 0});
    return std::unique_ptr<FMIndexOnGenome>(new FMIndexOnGenome(bwtOf(refF, saF), bwtOf(refR, saR), SparseSuffixArray::buildFromSuffixArray(saF, 32),
                                                                SparseSuffixArray::buildFromSuffixArray(saR, 32), std::move(b), refF.textSize()));
}

}  // namespace gwo
