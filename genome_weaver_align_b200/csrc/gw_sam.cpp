// gw_sam.cpp — host formatting of hit records into SAM text: the reference-side half of the boundary, provided natively so
// that the native CLI and the benchmarks can produce `out.sam` without a JVM.
// Reference: align/src/main/java/org/utgenome/weaver/align/record/AlignmentRecord.java:92-170 (toSAMLine), :181-276 (convert),
// record/ReadHit.java:82-132, align/SAMOutput.java:51-82, align/SequenceBoundary.java:81-87 (header).
#include <string.h>

#include "../../include/gw_align.h"
#include "gw_host.h"

namespace {

struct Rec {   // AlignmentRecord
    std::string chr;
    int strand = 0, start = 0, end = 0, nm = 0, numBestHits = 0;
    std::string cigar, seq, qual, state;
    bool qualNull = true;
    bool hasSplit = false;
};

inline char norm_base(char c) {   // ACGT.java:37-44 then ACGTSequence.toString
    switch (c) {
    case 'A': case 'a': return 'A';
    case 'C': case 'c': return 'C';
    case 'G': case 'g': return 'G';
    case 'T': case 't': case 'U': case 'u': return 'T';
    default: return 'N';
    }
}
inline char comp_base(char c) { return c == 'A' ? 'T' : c == 'C' ? 'G' : c == 'G' ? 'C' : c == 'T' ? 'A' : 'N'; }

const char kOps[] = "MIDNSHPX";   // BAM op codes

struct Cig {   // CIGAR.java:154-182
    std::vector<std::pair<char, int>> e;
    void add(int len, char t) {
        if (!e.empty() && e.back().first == t) e.back().second += len;
        else e.push_back({t, len});
    }
    void add(const Cig& o) { for (auto& x : o.e) add(x.second, x.first); }
    int unclipped() const {   // CIGAR.getUnclippedSize :133-147
        int l = 0;
        for (auto& x : e) if (x.first == 'M' || x.first == 'X' || x.first == 'D' || x.first == 'P' || x.first == 'N') l += x.second;
        return l;
    }
    std::string str() const {
        std::string s;
        for (auto& x : e) { s += std::to_string(x.second); s += x.first; }
        return s;
    }
};

Cig cig_of(const gw_hit_t& h, const uint32_t* ops) {
    Cig c;
    for (uint32_t i = 0; i < h.cigar_len; ++i) {
        uint32_t v = ops[h.cigar_off + i];
        c.e.push_back({kOps[v & 7u], (int)(v >> 4)});   // elements are kept as produced (a zero-length clip prints as "0S")
    }
    return c;
}

std::string chr_name(const std::vector<GwContig>& ctg, int id) {
    if (id >= 0 && id < (int)ctg.size()) return ctg[(size_t)id].name;
    if (id == GW_CHR_STAR) return "*";
    return "";
}

inline const char* state_letter(int numHits) { return numHits > 0 ? (numHits == 1 ? "U" : "R") : "N"; }

// ReadHit.getAlignmentStateOf(head) evaluated on fragment `self`
std::string state_of(const gw_hit_t* chain, int n, int self) {
    std::string s;
    for (int i = 0; i < n; ++i) {
        char c = state_letter(chain[i].num_hits)[0];
        if (i != self) c = (char)(c - 'A' + 'a');
        s += c;
    }
    return s;
}
std::string state_all(const gw_hit_t* chain, int n) {
    std::string s;
    for (int i = 0; i < n; ++i) s += state_letter(chain[i].num_hits);
    return s;
}

void sam_line(std::string& out, const std::string& name, const Rec& r, const Rec* split, bool hasSegments, bool isFirst, bool eachMapped) {
    int flag = 0;
    if (hasSegments) flag |= 0x1;
    if (r.strand == 1) flag |= 0x10;
    if (isFirst) {
        flag |= 0x40;
        if (r.numBestHits <= 0 || (split && split->numBestHits <= 0)) eachMapped = false;
    } else if (!split) flag |= 0x80;
    if (eachMapped) flag |= 0x2;
    if (r.numBestHits <= 0) flag |= 0x4;
    if (split && split->numBestHits <= 0) flag |= 0x8;
    out += name; out += '\t';
    out += std::to_string(flag); out += '\t';
    out += r.chr; out += '\t';
    out += std::to_string(r.start); out += "\t1\t";   // the score column is the constant 1 (AlignmentRecord.java:196-198)
    out += r.cigar; out += '\t';
    if (!split) out += "*\t0\t0";
    else {
        if (r.chr != "*" && r.chr == split->chr) out += "=";
        else out += split->chr;
        out += '\t'; out += std::to_string(split->start);
        out += '\t'; out += std::to_string(split->end - r.start);
    }
    out += '\t'; out += r.seq;
    out += '\t'; out += r.qualNull ? std::string("*") : r.qual;
    if (r.numBestHits > 0) {
        if (r.nm >= 0) { out += "\tNM:i:"; out += std::to_string(r.nm); }
        out += "\tXP:Z:"; out += r.state;
        out += "\tX0:i:"; out += std::to_string(r.numBestHits);
    }
    out += '\n';
    if (split) sam_line(out, name, *split, nullptr, hasSegments, false, eachMapped);
}

}  // namespace

// Formats one reported chain (1 or 2 fragments) of one read. Returns false where AlignmentRecord.convert returns null or throws.
bool gw_sam_one(const std::vector<GwContig>& ctg, const std::string& name, const char* bases, size_t m, const char* qual, size_t qlen,
                const gw_hit_t* chain, int n, const uint32_t* ops, std::string& out) {
    const gw_hit_t& hit = chain[0];
    std::string query(m, 'N'), q;
    for (size_t i = 0; i < m; ++i) query[i] = norm_base(bases[i]);
    const bool qualNull = qual == nullptr;
    if (!qualNull) q.assign(qual, qlen);
    if (hit.strand != 0) {
        std::string rc(m, 'N');
        for (size_t i = 0; i < m; ++i) rc[i] = comp_base(query[m - 1 - i]);
        query.swap(rc);
        if (!qualNull) q.assign(q.rbegin(), q.rend());
    }
    Rec r;
    r.qualNull = qualNull;
    if (n == 1) {
        r.chr = chr_name(ctg, hit.chr_id); r.strand = hit.strand; r.start = hit.pos; r.end = hit.pos + hit.match_length; r.nm = hit.diff;
        r.cigar = cig_of(hit, ops).str(); r.seq = query; r.qual = q; r.numBestHits = hit.num_hits; r.state = state_all(chain, 1);
        sam_line(out, name, r, nullptr, false, true, true);
        return true;
    }
    const gw_hit_t& split = chain[1];
    const int mi = (int)m;
    const int qualLen = !qualNull ? (int)q.size() : hit.match_length;
    auto sub = [&](int a, int b) -> std::string {
        if (a < 0 || b > mi || a > b) throw GwIoError("substring out of range");
        return query.substr((size_t)a, (size_t)(b - a));
    };
    std::string s1, s2, q1, q2;
    try {
        s1 = sub(hit.q_start, hit.q_end);
        s2 = sub(split.q_start, split.q_end);
        const int b1 = std::min(qualLen, hit.match_length), b2 = std::min(qualLen, mi);
        if (!qualNull) {
            if (b1 < 0 || b1 > b2 || b2 > (int)q.size()) return false;
            q1 = q.substr(0, (size_t)b1);
            q2 = q.substr((size_t)b1, (size_t)(b2 - b1));
        }
    } catch (GwIoError&) { return false; }
    const std::string hchr = chr_name(ctg, hit.chr_id), schr = chr_name(ctg, split.chr_id);
    const Cig hc = cig_of(hit, ops), sc = cig_of(split, ops);
    const bool hu = hit.num_hits == 1, su = split.num_hits == 1;
    if (hu) {
        if (su) {
            if (hchr == schr) {   // two records on one line pair (AlignmentRecord.java:209-223)
                Rec a, b;
                a.qualNull = b.qualNull = qualNull;
                a.chr = hchr; a.strand = hit.strand; a.start = hit.pos; a.end = hit.pos + hit.match_length; a.nm = hit.diff; a.cigar = hc.str();
                a.seq = s1; a.qual = q1; a.numBestHits = 1; a.state = state_of(chain, 2, 0);
                b.chr = schr; b.strand = split.strand; b.start = split.pos; b.end = split.pos + split.match_length; b.nm = split.diff;
                b.cigar = sc.str(); b.seq = s2; b.qual = q2; b.numBestHits = 1; b.state = state_of(chain, 2, 1);
                sam_line(out, name, a, &b, true, true, true);
                return true;
            }
            if (hit.match_length >= split.match_length) {
                Cig c = hc; c.add(split.match_length, 'S');
                r.chr = hchr; r.strand = hit.strand; r.start = hit.pos; r.end = hit.pos + mi; r.nm = hit.diff; r.cigar = c.str();
                r.state = state_of(chain, 2, 0);
            } else {
                Cig c = hc; c.add(sc);
                r.chr = schr; r.strand = split.strand; r.start = split.pos - hit.match_length; r.end = r.start + mi; r.nm = split.diff;
                r.cigar = c.str(); r.state = state_of(chain, 2, 1);
            }
        } else {
            Cig c = hc; c.add(split.q_end - split.q_start, 'S');
            r.chr = hchr; r.strand = hit.strand; r.start = hit.pos; r.end = hit.pos + c.unclipped(); r.nm = hit.diff; r.cigar = c.str();
            r.state = state_of(chain, 2, 0);
        }
    } else if (su) {
        Cig c; c.add(hit.match_length, 'S'); c.add(sc);
        r.chr = schr; r.strand = split.strand; r.start = split.pos - hit.match_length; r.end = r.start + mi; r.nm = split.diff;
        r.cigar = c.str(); r.state = state_of(chain, 2, 1);
    } else
        return false;   // both fragments repeat: convert returns null, SAMOutput prints nothing
    r.seq = query; r.qual = q; r.numBestHits = hit.num_hits;
    sam_line(out, name, r, nullptr, false, true, true);
    return true;
}

std::string gw_sam_header_text(const std::vector<GwContig>& ctg) {
    std::string s;
    for (const GwContig& c : ctg) s += "@SQ\tSN:" + c.name + "\tLN:" + std::to_string(c.length) + "\n";
    return s;
}
