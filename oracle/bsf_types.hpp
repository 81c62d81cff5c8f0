// TEST INFRASTRUCTURE ONLY (see java_sem.hpp). CPU restatement of the BSF strategy's helper types:
//   StaircaseFilter    align/src/main/java/org/utgenome/weaver/align/strategy/StaircaseFilter.java:47-73, :91-102, :113-160
//   ReadAlignmentNFA   .../strategy/ReadAlignmentNFA.java:63-84, :107-114, :127-203, :205-215
//   Cursor             .../strategy/Cursor.java:40-198
//   FMQuickScan        .../strategy/FMQuickScan.java:66-94
//   ReadHit            .../record/ReadHit.java:42-184
//   AlignmentRecord    .../record/AlignmentRecord.java:92-170 toSAMLine, :181-276 convert
//   AlignmentConfig    .../AlignmentScoreConfig.java:37-77, .../AlignmentConfig.java:40-72
#pragma once
#include <cmath>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "fm_index.hpp"
#include "verifier.hpp"

namespace gwo {

struct AlignmentConfig {
    float k = 0.1f;
    jint numSplitAlowed = 1, matchScore = 1, mismatchPenalty = 3, splitOpenPenalty = 11, indelEndSkip = 5, bandWidth = 31;
    int reportType = 0;  // 0 BESTHIT, 1 ALLHITS, 2 TOPL
    jint topL = 5;
    bool extended = false;  // false: strict (abort where the reference aborts); true: SURVEY.md A.6 extended semantics
    jint getMaximumEditDistance(jint readLength) const {
        if (k > 0 && k < 1) return (jint)std::floor(readLength * k);  // float*int in Java is float arithmetic
        return (jint)k;
    }
};

class StaircaseFilter {
public:
    jint m, k;
    std::vector<int8_t> chunkStart, chunkLen;
    std::vector<BitVector> staircaseMask;
    StaircaseFilter(jint m_, jint k_, bool extended) : m(m_), k(k_) {
        jint lastChunkSize = (m - k >= 6) ? m * 2 / (k + 2) : m - k;
        chunkStart.assign((size_t)k + 2, 0);
        int8_t rest = jbyte(m - lastChunkSize);
        if (k == 0) chunkStart[0] = rest;
        else for (jint i = 0; i <= k; ++i) chunkStart[(size_t)i] = jbyte((jint)rest * i / k);
        chunkStart[(size_t)k + 1] = jbyte(m);
        chunkLen.assign((size_t)k + 1, 0);
        for (jint i = 0; i <= k; ++i) chunkLen[(size_t)i] = jbyte((jint)chunkStart[(size_t)i + 1] - (jint)chunkStart[(size_t)i]);
        for (jint i = 0; i <= k; ++i) {
            BitVector bv(m, extended);
            bv._not()._lshift((jint)chunkStart[(size_t)i]);
            staircaseMask.push_back(bv);
        }
    }
    u64 getStairCaseMask64bit(jint kk, jint offset) const {
        if (kk >= (jint)staircaseMask.size()) return 0;
        if (offset >= 0) {
            u64 base = shl(~0ULL, m - offset);
            return base | staircaseMask[(size_t)kk].substring64(offset, offset + 64);
        }
        return shl(staircaseMask[(size_t)kk].substring64(0, 64), -offset);
    }
};

class StaircaseFilterHolder {
public:
    std::map<std::pair<jint, jint>, std::unique_ptr<StaircaseFilter>> holder;
    const StaircaseFilter& get(jint queryLength, jint kk, bool extended) {
        auto key = std::make_pair(kk, queryLength);
        auto it = holder.find(key);
        if (it == holder.end()) it = holder.emplace(key, std::unique_ptr<StaircaseFilter>(new StaircaseFilter(queryLength, kk, extended))).first;
        return *it->second;
    }
};

struct Cursor {
    jint flag = 0, start = 0, end = 0, cursor = 0, pivot = 0;
    Cursor() {}
    Cursor(int strand, int dir, jint s, jint e, jint c, jint p) : flag(strand | (dir << 1)), start(s), end(e), cursor(c), pivot(p) {}
    jint getFragmentLength() const { return end - start; }
    int getStrand() const { return flag & 1; }
    int getSearchDirection() const { int d = (flag >> 1) & 3; return d == 3 ? DIR_FORWARD : d; }
    bool isForwardSearch() const { return dirIsForward(getSearchDirection()); }
    jint getProcessedBases() const { return isForwardSearch() ? cursor - pivot : end - cursor; }
    jint getRemainingBases() const { return getFragmentLength() - getProcessedBases(); }
    jint getOffsetOfSearchHead() const {
        jint offset = cursor - start;
        if (getStrand() == REVERSE) offset = getFragmentLength() - offset;
        return offset;
    }
    jint getNextACGTIndex() const { return isForwardSearch() ? cursor : cursor - 1; }
    void split(Cursor& left, Cursor& right) const {
        int d = getSearchDirection();
        switch (d) {
        case DIR_FORWARD:
            left = Cursor(getStrand(), d, start, cursor, cursor, pivot);
            right = Cursor(getStrand(), DIR_FORWARD, cursor, end, cursor, cursor);
            break;
        case DIR_BACKWARD:
            left = Cursor(getStrand(), d, cursor, end, cursor, pivot);
            right = Cursor(getStrand(), DIR_BACKWARD, start, cursor, cursor, start);
            break;
        default:
            if (cursor + 1 < end) {
                left = Cursor(getStrand(), d, start, cursor, cursor, pivot);
                right = Cursor(getStrand(), d, cursor, end, cursor, cursor);
            } else {
                left = Cursor(getStrand(), DIR_BACKWARD, start, cursor, cursor, pivot);
                right = Cursor(getStrand(), d, cursor, end, cursor, cursor);
            }
        }
    }
    Cursor next() const {
        jint nextCursor = cursor;
        int d = getSearchDirection();
        switch (d) {
        case DIR_FORWARD: ++nextCursor; break;
        case DIR_BACKWARD: --nextCursor; break;
        default:
            if (nextCursor + 1 < end) ++nextCursor;
            else { d = DIR_BACKWARD; nextCursor = pivot; }
        }
        return Cursor(getStrand(), d, start, end, nextCursor, pivot);
    }
    SiSetPtr nextSi(const FMIndexOnGenome& fm, const SiSet& si, int currentBase) const {
        SuffixInterval siF = si.getForward(currentBase), siB = si.getBackward(currentBase);
        if (getSearchDirection() == DIR_BIDIRECTIONAL_FORWARD && cursor >= end - 1) siF = SuffixInterval();
        return fm.bidirectionalSearch(getStrand(), siF, siB);
    }
};

class ReadAlignmentNFA {
public:
    std::vector<u64> automaton;
    jint kOffset = 0;
    explicit ReadAlignmentNFA(jint numAllowedMismatches) : automaton((size_t)numAllowedMismatches + 1, 0) {}
    ReadAlignmentNFA(std::vector<u64> a, jint ko) : automaton(std::move(a)), kOffset(ko) {}
    ReadAlignmentNFA& activateDiagonalStates() {
        const jint k = (jint)automaton.size() - 1;
        for (jint i = 0; i < (jint)automaton.size(); ++i) automaton[(size_t)i] = shl(1ULL, k + i);
        return *this;
    }
    ReadAlignmentNFA nextStateAfterSplit(jint /*k*/) const {
        jint height = (jint)automaton.size() - 1;
        if (height < 0) throw JavaAbort("NegativeArraySizeException in nextStateAfterSplit");
        std::vector<u64> next((size_t)height);
        for (jint i = 0; i < height; ++i) next[(size_t)i] = shl(1ULL, height + i - 1);
        return ReadAlignmentNFA(next, kOffset + 1);
    }
    struct NextState { bool valid = false; std::vector<u64> automaton; jint kOffset = 0; bool hasMatch = false; };
    NextState nextState(const Cursor& cursor, int ch, const QueryMask& queryMask, const StaircaseFilter& sf) const {
        const jint height = (jint)automaton.size();
        const jint k = kOffset + height - 1;
        const jint kr = k - kOffset;
        const u64 qeq = queryMask.getBidirectionalPatternMask64(cursor.isForwardSearch(), cursor.getNextACGTIndex(),
                                                                 cursor.pivot, cursor.cursor, ch, kr);
        return nextState(qeq, cursor.getProcessedBases(), cursor.getFragmentLength(), sf);
    }
    NextState nextState(u64 qeq, jint progressIndex, jint fragmentLength, const StaircaseFilter& sf) const {
        const jint height = (jint)automaton.size();
        const jint k = kOffset + height - 1;
        NextState out;
        if (height == 0) throw JavaAbort("ArrayIndexOutOfBoundsException: empty automaton");
        const std::vector<u64>& prev = automaton;
        std::vector<u64> next((size_t)height, 0);
        jint minKwithMatch = k + 1, minKwithProgress = k + 1;
        next[0] = (prev[0] & qeq) << 1;
        if (next[0] != 0) { minKwithMatch = 0; minKwithProgress = 0; }
        next[0] &= sf.getStairCaseMask64bit(kOffset, progressIndex - k);
        for (jint i = 1; i < height; ++i) {
            next[(size_t)i] = (prev[(size_t)i] & qeq) << 1;
            if (minKwithMatch > k && next[(size_t)i] != 0) minKwithMatch = i;
            next[(size_t)i] |= prev[(size_t)i - 1] | (prev[(size_t)i - 1] << 1) | (next[(size_t)i - 1] << 1);
            next[(size_t)i] &= sf.getStairCaseMask64bit(kOffset + i, progressIndex - k);
            if (minKwithProgress > k && (next[(size_t)i] & shl(1ULL, height)) != 0) minKwithProgress = i;
        }
        const jint m = fragmentLength;
        const jint mPos = k + m - progressIndex;
        if (mPos < 64) {
            for (jint nm = 0; nm < height; ++nm) {
                if ((next[(size_t)nm] & shl(1ULL, mPos)) != 0) {
                    out.valid = true; out.hasMatch = true; out.kOffset = kOffset + nm;
                    out.automaton = removeLayers(next, nm);
                    return out;
                }
            }
        }
        jint minK = std::min(minKwithMatch, minKwithProgress);
        if (minK < k) {
            out.valid = true; out.hasMatch = false; out.kOffset = kOffset + minK;
            out.automaton = removeLayers(next, minK);
            return out;
        }
        return out;  // null
    }
    std::vector<u64> removeLayers(const std::vector<u64>& next, jint n) const {
        if (n == 0) return automaton;  // sic: the OLD automaton (ReadAlignmentNFA.java:206-207)
        jint nextHeight = (jint)automaton.size() - n;
        std::vector<u64> t((size_t)nextHeight);
        for (jint h = 0; h < nextHeight; ++h) t[(size_t)h] = next[(size_t)(h + n)] >> 1;
        return t;
    }
};

struct FMQuickScan {
    int strand;
    SuffixInterval si;
    jint numMismatches;
    jint longestStart, longestEnd;
    static FMQuickScan scan(const FMIndexOnGenome& fm, const ACGTSequence& query, int strand) {
        jint qLen = (jint)query.textSize();
        jint numMismatches = 0;
        SuffixInterval si = fm.wholeRange;
        jint mark = 0;
        bool hasLongest = false;
        jint ls = 0, le = 0;
        jint i = 0;
        for (; i < qLen; ++i) {
            int ch = query.getACGT(i);
            si = fm.forwardSearch(strand, ch, si);
            if (si.isEmpty()) {
                numMismatches++;
                if (!hasLongest || (le - ls) < (i - mark)) { hasLongest = true; ls = mark; le = i; }
                si = fm.wholeRange;
                mark = i + 1;
            }
        }
        if (!hasLongest || (le - ls) < (i - mark)) { ls = mark; le = i; }
        return FMQuickScan{strand, si, numMismatches, ls, le};
    }
};

struct ReadHit;
typedef std::shared_ptr<ReadHit> ReadHitPtr;
struct ReadHit {
    bool chrIsNull = false;
    std::string chr;
    jlong pos = 0;
    jint matchLength = 0, diff = 0;
    int strand = FORWARD;
    CIGAR cigar;
    jint numHits = 0, qStart = 0, qEnd = 0;
    ReadHitPtr nextSplit;
    static ReadHitPtr make(bool chrNull, const std::string& chr, jlong pos, jint matchLength, jint qStart, jint qEnd, jint diff,
                           int strand, const CIGAR& cigar, jint numHits) {
        auto h = std::make_shared<ReadHit>();
        h->chrIsNull = chrNull; h->chr = chr; h->pos = pos; h->matchLength = matchLength; h->qStart = qStart; h->qEnd = qEnd;
        h->diff = diff; h->strand = strand; h->cigar = cigar; h->numHits = numHits;
        return h;
    }
    static ReadHitPtr noHit(int strand) { return make(false, "", 0, 0, 0, 0, 0, strand, CIGAR(), 0); }
    jint getFragmentLength() const { return qEnd - qStart; }
    jint getTotalMatchLength() const { return nextSplit ? matchLength + nextSplit->getTotalMatchLength() : matchLength; }
    jint getTotalDifferences() const { return nextSplit ? diff + 1 + nextSplit->getTotalDifferences() : diff; }
    bool isUnique() const { return numHits == 1; }
    jint getTotalScore(const AlignmentConfig& c) const {
        jint score = matchLength * c.matchScore - diff * c.mismatchPenalty;
        return nextSplit ? score - c.splitOpenPenalty + nextSplit->getTotalScore(c) : score;
    }
    std::string stateSingle() const { return numHits > 0 ? (numHits == 1 ? "U" : "R") : "N"; }
    std::string getAlignmentState() const { return nextSplit ? stateSingle() + nextSplit->getAlignmentState() : stateSingle(); }
    std::string getAlignmentStateOf(const ReadHit* head) const {
        std::string s;
        for (const ReadHit* c = head; c != nullptr; c = c->nextSplit.get()) {
            std::string st = c->stateSingle();
            if (c != this) st[0] = (char)tolower(st[0]);
            s += st;
        }
        return s;
    }
};

// Java String.compareTo
static inline jint javaCompareTo(const std::string& a, const std::string& b) {
    size_t n = std::min(a.size(), b.size());
    for (size_t i = 0; i < n; ++i) if (a[i] != b[i]) return (jint)(unsigned char)a[i] - (jint)(unsigned char)b[i];
    return (jint)a.size() - (jint)b.size();
}

// ReadHit.sortSplits, ReadHit.java:146-182 (stable sort, comparator restated)
static inline ReadHitPtr sortSplits(ReadHitPtr self) {
    if (!self->nextSplit) return self;
    std::vector<ReadHitPtr> r;
    for (ReadHitPtr h = self; h; h = h->nextSplit) r.push_back(h);
    const bool fwd = self->strand == FORWARD;
    auto cmp = [&](const ReadHitPtr& o1, const ReadHitPtr& o2) -> jint {
        jint diff = 0;
        if (o1->chrIsNull || o2->chrIsNull) {
            diff = o1->qStart - o2->qStart;
            if (!fwd) diff = -diff;
        }
        if (diff != 0) return diff;
        if (o1->chrIsNull || o2->chrIsNull) throw JavaAbort("NullPointerException in sortSplits");
        diff = javaCompareTo(o1->chr, o2->chr);
        if (diff != 0) return diff;
        return (jint)(o1->pos - o2->pos);
    };
    // stable insertion sort == java.util.Collections.sort for these tiny lists (legacy merge sort uses insertion sort below 7)
    for (size_t i = 1; i < r.size(); ++i)
        for (size_t j = i; j > 0 && cmp(r[j - 1], r[j]) > 0; --j) std::swap(r[j - 1], r[j]);
    for (size_t i = 0; i + 1 < r.size(); ++i) r[i]->nextSplit = r[i + 1];
    r.back()->nextSplit = nullptr;
    return r[0];
}

struct AlignmentRecord;
typedef std::shared_ptr<AlignmentRecord> AlignmentRecordPtr;
struct AlignmentRecord {
    std::string readName, chr;
    int strand = FORWARD;
    jint start = 0, end = 0, numMismatches = 0;
    CIGAR cigar;
    std::string querySeq, qual;
    bool qualIsNull = true;
    jint score = 0, numBestHits = 0;
    std::string alignmentState;
    AlignmentRecordPtr split;

    std::string toSAMLine() const { return toSAMLine(split != nullptr, true, true); }
    std::string toSAMLine(bool hasSegments, bool isFirst, bool eachFragmentIsMapped) const {
        jint flag = 0;
        if (hasSegments) flag |= 0x1;
        if (strand == REVERSE) flag |= 0x10;
        if (isFirst) {
            flag |= 0x40;
            for (const AlignmentRecord* r = this; r != nullptr; r = r->split.get())
                if (r->numBestHits <= 0) { eachFragmentIsMapped = false; break; }
        } else if (!split) flag |= 0x80;
        if (eachFragmentIsMapped) flag |= 0x2;
        if (numBestHits <= 0) flag |= 0x4;
        if (split && split->numBestHits <= 0) flag |= 0x8;
        std::string line = readName + "\t" + std::to_string(flag) + "\t" + chr + "\t" + std::to_string(start) + "\t" +
                           std::to_string(score) + "\t" + cigar.toString() + "\t";
        if (!split) line += "*\t0\t0";
        else {
            if (chr != "*" && chr == split->chr) line += "=";
            else line += split->chr;
            line += "\t" + std::to_string(split->start) + "\t" + std::to_string(split->end - start);
        }
        line += "\t" + querySeq + "\t" + (qualIsNull ? std::string("*") : qual);
        if (numBestHits > 0) {
            if (numMismatches >= 0) line += "\tNM:i:" + std::to_string(numMismatches);
            line += "\tXP:Z:" + alignmentState;
            line += "\tX0:i:" + std::to_string(numBestHits);
        }
        if (split) line += "\n" + split->toSAMLine(hasSegments, false, eachFragmentIsMapped);
        return line;
    }
};

}  // namespace gwo
