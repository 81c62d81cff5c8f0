// TEST INFRASTRUCTURE ONLY (see java_sem.hpp). CPU restatement of the reference's verifier:
//   CIGAR                     align/src/main/java/org/utgenome/weaver/align/CIGAR.java:133-147, :154-182, :209-215
//   BitParallelSmithWaterman  .../align/BitParallelSmithWaterman.java:141-147 alignBlockDetailed,
//                             :354-361 ctor, :368-385 peq, :387-459 align, :461-491 alignBlock, :501-644 traceback
//   Alignment                 .../align/SmithWatermanAligner.java:42-58
#pragma once
#include <string>
#include <vector>

#include "sequence.hpp"

namespace gwo {

struct CigarElement { char type; jint length; };

class CIGAR {
public:
    std::vector<CigarElement> cigar;
    CIGAR() {}
    explicit CIGAR(const std::string& s) {  // parse, CIGAR.java:221-241
        size_t startNum = 0;
        for (size_t cur = 0; cur < s.size(); ++cur) {
            char c = s[cur];
            if (c >= '0' && c <= '9') continue;
            if (startNum == cur) break;
            cigar.push_back({normalise(c), (jint)std::stoi(s.substr(startNum, cur - startNum))});
            startNum = cur + 1;
        }
    }
    static char normalise(char c) {  // Type.convert: unknown letters become SkippedRegion "N"
        switch (c) { case 'M': case 'I': case 'D': case 'S': case 'H': case 'P': case 'X': return c; default: return 'N'; }
    }
    CIGAR& add(char ch) {
        char t = normalise(ch);
        if (!cigar.empty() && cigar.back().type == t) cigar.back().length++;
        else cigar.push_back({t, 1});
        return *this;
    }
    CIGAR& add(jint length, char type) {
        if (!cigar.empty() && cigar.back().type == type) cigar.back().length += length;
        else cigar.push_back({type, length});
        return *this;
    }
    CIGAR& add(const CIGAR& other) { for (auto& e : other.cigar) add(e.length, e.type); return *this; }
    jint getUnclippedSize() const {
        jint l = 0;
        for (auto& e : cigar) if (e.type == 'M' || e.type == 'X' || e.type == 'D' || e.type == 'P' || e.type == 'N') l += e.length;
        return l;
    }
    std::string toString() const {
        std::string s;
        for (auto& e : cigar) { s += std::to_string(e.length); s += e.type; }
        return s;
    }
};

struct Alignment {
    bool valid = false;  // false == Java null
    CIGAR cigar;
    jint score = 0, numMismatches = 0, pos = 0;
};

class AlignBlocksDetailed {
public:
    static const int w = 64;
    jint k, m, bMax;
    std::vector<u64> peq;  // [5][bMax]
    std::vector<jint> D;
    std::vector<u64> vp, vn;  // [bMax][N+1]
    jint Ncols = 0;
    AlignBlocksDetailed(jint m_, jint k_) : k(k_), m(m_) {
        bMax = std::max(1, (m + w - 1) / w);
        peq.assign((size_t)5 * bMax, 0);
        D.assign((size_t)bMax, 0);
    }
    u64& VP(jint r, jint j) { return vp[(size_t)r * (Ncols + 1) + j]; }
    u64& VN(jint r, jint j) { return vn[(size_t)r * (Ncols + 1) + j]; }

    // returns tailPos or -1 (Java null)
    jint align(const ACGTSequence& ref, const ACGTSequence& query) {
        QueryMask qm(query);
        for (int ch = 0; ch < 4; ++ch)
            for (jint i = 0; i < bMax; ++i) peq[(size_t)ch * bMax + i] = qm.getPatternMaskIn64bit(ch, i, w);
        return alignRef(ref);
    }
    jint alignRef(const ACGTSequence& ref) {
        bool hasBest = false;
        jint bestTail = -1, bestDiff = 0;
        const jint N = (jint)ref.textSize();
        Ncols = N;
        vp.assign((size_t)bMax * (N + 1), 0);
        vn.assign((size_t)bMax * (N + 1), 0);
        for (jint r = 0; r < bMax; ++r) { VP(r, 0) = ~0ULL; VN(r, 0) = 0; }
        D[0] = m;
        std::vector<jint> scoreBoundary((size_t)bMax);
        for (jint i = 0; i < bMax; ++i) scoreBoundary[(size_t)i] = std::max(m - ((i + 1) * w) + k, 0);
        jint bCeil = std::max(1, (k + w - 1) / w);
        for (jint j = 0; j < N; ++j) {
            int ch = ref.getACGT(j);
            jint carry = 0;
            for (jint b = 0; b < bCeil; ++b) {
                jint nextScore = alignBlock(j, ch, b, carry);
                D[(size_t)b] += nextScore;
                carry = nextScore;
            }
            if (bCeil < bMax && D[(size_t)bCeil - 1] - carry <= scoreBoundary[(size_t)bCeil - 1]
                && (((peq[(size_t)ch * bMax + bCeil] & 1ULL) != 0) | (carry < 0))) {
                VP(bCeil, j) = ~0ULL;
                VN(bCeil, j) = 0;
                jint nextScore = alignBlock(j, ch, bCeil, carry);
                D[(size_t)bCeil] = D[(size_t)bCeil - 1] - carry + nextScore;
                bCeil++;
            } else {
                while (bCeil > 1 && D[(size_t)bCeil - 1] > scoreBoundary[(size_t)bCeil - 1] + w) --bCeil;
            }
            if (bCeil == bMax) {
                if (!hasBest) { hasBest = true; bestTail = j; bestDiff = D[(size_t)bCeil - 1]; continue; }
                if (bestDiff > D[(size_t)bCeil - 1]) { bestTail = j; bestDiff = D[(size_t)bCeil - 1]; }
            }
        }
        return hasBest ? bestTail : -1;
    }
    jint alignBlock(jint j, int ch, jint r, jint hin) {
        u64 vp_ = VP(r, j), vn_ = VN(r, j);
        u64 x = peq[(size_t)ch * bMax + r];
        if (hin < 0) x |= 1ULL;
        u64 d0 = (((x & vp_) + vp_) ^ vp_) | x | vn_;
        u64 hp = vn_ | ~(d0 | vp_);
        u64 hn = d0 & vp_;
        jint hout = 0;
        hout += (jint)((hp >> (w - 1)) & 1ULL);
        hout -= (jint)((hn >> (w - 1)) & 1ULL);
        u64 hp2 = hp << 1, hn2 = hn << 1;
        if (hin < 0) hn2 |= 1ULL;
        if (hin > 0) hp2 |= 1ULL;
        VP(r, j + 1) = hn2 | ~(d0 | hp2);
        VN(r, j + 1) = d0 & hp2;
        return hout;
    }
    Alignment traceback(const ACGTSequence& ref, const ACGTSequence& query, jint tailPos) {
        jint maxRow = m - 1, maxCol = tailPos;
        std::string cigar;  // built right-to-left, reversed below
        jint leftMostPos = 0;
        jint diff = 0;
        jint col = maxCol, row = maxRow;
        for (;;) {
            int path = 0;  // 0 NONE, 1 DIAGONAL, 2 DIAGONAL_MISMATCH, 3 UP, 4 LEFT
            if (col >= 0 && row >= 0) {
                jint block = row / w, offset = row % w;
                u64 vpf = VP(block, col + 1) & (1ULL << offset);
                u64 vnf = VN(block, col + 1) & (1ULL << offset);
                if (ref.getACGT(col) == query.getACGT(row)) path = 1;
                else if (vpf != 0) { path = 3; diff++; }
                else if (vnf == 0) { path = 2; diff++; }
                else { path = 4; diff++; }
            }
            bool done = false;
            switch (path) {
            case 1: case 2: cigar += 'M'; leftMostPos = col; col--; row--; break;
            case 3: cigar += 'I'; leftMostPos = col + 1; row--; break;
            case 4: cigar += 'D'; col--; break;
            default:
                while (col >= 0 || row >= 0) { if (row >= 0) cigar += 'S'; col--; row--; }
                done = true;
            }
            if (done) break;
        }
        std::string cigarStr(cigar.rbegin(), cigar.rend());
        {
            jint left = 0, right = 0;
            for (jint i = 0; i < (jint)cigarStr.size(); ++i) {
                char t = cigarStr[(size_t)i];
                if (t == 'S') left++;
                else if (t == 'I' || t == 'D') { left++; diff--; }
                else break;
            }
            for (jint i = (jint)cigarStr.size() - 1; i >= left; --i) {
                char t = cigarStr[(size_t)i];
                if (t == 'S') right++;
                else if (t == 'I' || t == 'D') { right++; diff--; }
                else break;
            }
            std::string nc((size_t)left, 'S');
            nc += cigarStr.substr((size_t)left, cigarStr.size() - (size_t)right - (size_t)left);
            nc += std::string((size_t)right, 'S');
            cigarStr = nc;
        }
        Alignment a;
        a.valid = true;
        for (char c : cigarStr) a.cigar.add(c);
        a.score = m - diff;
        a.numMismatches = diff;
        a.pos = leftMostPos;
        return a;
    }
};

// BitParallelSmithWaterman.alignBlockDetailed, :141-147
static inline Alignment alignBlockDetailed(const ACGTSequence& ref, const ACGTSequence& query, jint k) {
    AlignBlocksDetailed a((jint)query.textSize(), k);
    jint tail = a.align(ref, query);
    if (tail < 0) return Alignment();
    return a.traceback(ref, query, tail);
}

}  // namespace gwo
