// TEST INFRASTRUCTURE ONLY (see java_sem.hpp). CPU restatement of the reference's FM-index layer:
//   OccurrenceCountTable   align/src/main/java/org/utgenome/weaver/align/OccurrenceCountTable.java:51-72, :80-89, :99-107
//   CharacterCount         .../align/CharacterCount.java:41-50, :74-76
//   FMIndexOnOccTable      .../align/FMIndexOnOccTable.java:47-51 backwardSearch, :53-55 rankACGTN, :64-70 suffixLink
//   SuffixInterval, SiSet  .../align/SuffixInterval.java:33-64, .../align/SiSet.java:34-196
//   SparseSuffixArray      .../align/SparseSuffixArray.java:124-131 buildFromSuffixArray, :141-161 get
//   SequenceBoundary       .../align/SequenceBoundary.java:81-87 toSAMHeader, :104-121 translate
//   FMIndexOnGenome        .../align/FMIndexOnGenome.java:121-132 initSet, :138-141 forwardSearch, :148-167 backwardSearch,
//                          :169-194 bidirectionalSearch, :211-225 forwardSearch(fm,lo,hi), :227-269 toCoordinate/toGenomeCoordinate
#pragma once
#include <map>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "sequence.hpp"

namespace gwo {

struct SuffixInterval {
    jlong lowerBound = 0, upperBound = 0;
    bool valid = false;  // false == Java null
    SuffixInterval() {}
    SuffixInterval(jlong lb, jlong ub) : lowerBound(lb), upperBound(ub), valid(true) {}
    jlong range() const { return upperBound - lowerBound; }
    bool isEmpty() const { return lowerBound >= upperBound; }
};

// global counter of occ lookups (getOcc / getOccACGTN calls): the roofline work unit
struct OccCounter { static thread_local uint64_t lookups; };
thread_local uint64_t OccCounter::lookups = 0;

class OccurrenceCountTable {
public:
    const ACGTSequence* seq;
    jlong W;
    std::vector<jlong> occTable;  // [numRows][5]
    OccurrenceCountTable(const ACGTSequence* seq_, int windowSize) : seq(seq_), W(windowSize) {
        const jint numRows = (jint)((seq->textSize() + W) / W);
        occTable.assign((size_t)numRows * 5, 0);
        // The reference fills the table with one running sum (OccurrenceCountTable.java:51-72). The per-window counts do not depend on
        // each other, so they are taken by several threads first and summed afterwards: same table, a human-scale index opens in minutes.
        unsigned T = std::max(1u, std::min(32u, std::thread::hardware_concurrency()));
        if (numRows < 4096) T = 1;
        std::vector<std::thread> th;
        for (unsigned t = 0; t < T; ++t) {
            th.emplace_back([this, t, T, numRows]() {
                const jint lo = 1 + (jint)((jlong)(numRows - 1) * t / T), hi = 1 + (jint)((jlong)(numRows - 1) * (t + 1) / T);
                for (jint i = lo; i < hi; ++i) {
                    jlong start = (jlong)(i - 1) * W;
                    jlong end = std::min(start + W, seq->textSize());
                    jlong count[5];
                    seq->fastCountACGTN(start, end, count);
                    for (int k = 0; k < 5; ++k) occTable[(size_t)i * 5 + k] = count[k];
                }
            });
        }
        for (auto& x : th) x.join();
        for (jint i = 1; i < numRows; ++i)
            for (int k = 0; k < 5; ++k) occTable[(size_t)i * 5 + k] += occTable[(size_t)(i - 1) * 5 + k];
    }
    void getOccACGTN(jlong index, jlong occ[5]) const {
        OccCounter::lookups++;
        if (index > seq->textSize()) index = seq->textSize();
        jint blockPos = (jint)(index / W);
        seq->fastCountACGTN(blockPos * W, index, occ);
        for (int i = 0; i < 5; ++i) occ[i] += occTable[(size_t)blockPos * 5 + i];
    }
    jlong getOcc(int ch, jlong index) const {
        OccCounter::lookups++;
        if (index > seq->textSize()) index = seq->textSize();
        jint blockPos = (jint)(index / W);
        return occTable[(size_t)blockPos * 5 + ch] + seq->fastCount(ch, blockPos * W, index);
    }
};

class FMIndexOnOccTable {
public:
    ACGTSequence seq;  // the BWT string
    OccurrenceCountTable occ;
    jlong C[5], count[5];
    FMIndexOnOccTable(ACGTSequence bwt, int windowSize) : seq(std::move(bwt)), occ(&seq, windowSize) {
        for (int ch = 0; ch < 5; ++ch) count[ch] = occ.getOcc(ch, seq.textSize());  // CharacterCount.java:41-50
        jlong sum = 0;
        for (int i = 0; i < 5; ++i) { C[i] = sum; sum += count[i]; }
    }
    FMIndexOnOccTable(const FMIndexOnOccTable&) = delete;
    jlong textSize() const { return seq.textSize(); }
    SuffixInterval backwardSearch(int ch, const SuffixInterval& cur) const {
        jlong lb = C[ch] + occ.getOcc(ch, cur.lowerBound);
        jlong ub = C[ch] + occ.getOcc(ch, cur.upperBound);
        return SuffixInterval(lb, ub);
    }
    void rankACGTN(jlong i, jlong out[5]) const { occ.getOccACGTN(i, out); }
    jlong suffixLink(jlong index) const {
        if (index >= seq.textSize()) return 0;
        int c = seq.lookup(index);
        return C[c] + occ.getOcc(c, index);
    }
};

class SparseSuffixArray {
public:
    std::vector<uint64_t> sparseSA;
    jlong N = 0;
    jint L = 32;
    static SparseSuffixArray buildFromSuffixArray(const std::vector<uint32_t>& SA, jint L) {
        SparseSuffixArray s;
        s.N = (jlong)SA.size();
        s.L = L;
        jlong len = ((jlong)SA.size() + L - 1) / L;
        s.sparseSA.resize((size_t)len);
        for (jlong i = 0; i < len; ++i) s.sparseSA[(size_t)i] = SA[(size_t)(i * L)];
        return s;
    }
    jlong get(jlong index, const FMIndexOnOccTable& fm) const {
        jlong pos = index / L, offset = index % L;
        if (offset == 0) return (jlong)sparseSA.at((size_t)pos);
        jlong cursor = index;
        const jlong n = fm.textSize();
        for (jlong j = 1; j <= n; j++) {
            cursor = fm.suffixLink(cursor);
            if (cursor % L == 0) return ((jlong)sparseSA.at((size_t)(cursor / L)) + j + n) % n;
        }
        throw JavaAbort("IllegalStateException: cannot reach here: SparseSuffixArray.get");
    }
};

struct SequenceIndex { std::string name, desc; jlong length, offset; };
struct PosOnGenome { std::string chr; jint pos; bool valid; };

class SequenceBoundary {
public:
    std::vector<SequenceIndex> index;
    jlong totalSize = 0;
    mutable std::map<jlong, std::string> indexToChrTable;
    std::string toSAMHeader() const {
        std::string s;
        for (auto& e : index) s += "@SQ\tSN:" + e.name + "\tLN:" + std::to_string(e.length) + "\n";
        return s;
    }
    PosOnGenome translate(jlong textIndex) const {
        if (indexToChrTable.empty()) for (auto& e : index) indexToChrTable[e.offset] = e.name;  // later put wins
        auto it = indexToChrTable.lower_bound(textIndex);  // headMap(textIndex): keys strictly less
        if (it == indexToChrTable.begin()) throw JavaAbort("UTGBException: invalid index in translate");
        --it;
        return PosOnGenome{it->second, (jint)(textIndex - it->first), true};
    }
};

enum SearchDirection : int { DIR_FORWARD = 0, DIR_BACKWARD = 1, DIR_BIDIRECTIONAL_FORWARD = 2 };
static inline bool dirIsForward(int d) { return d != DIR_BACKWARD; }
enum Strand : int { FORWARD = 0, REVERSE = 1 };

struct SiSet {
    enum Kind { EMPTY, F, B, BD } kind = EMPTY;
    SuffixInterval siF[5], siB[5];
    SuffixInterval getForward(int ch) const { return (kind == F || kind == BD) ? siF[ch] : SuffixInterval(); }
    SuffixInterval getBackward(int ch) const { return (kind == B || kind == BD) ? siB[ch] : SuffixInterval(); }
    bool isEmpty(int ch) const {
        if (kind == EMPTY) return true;
        const SuffixInterval& s = (kind == B) ? siB[ch] : siF[ch];
        return !s.valid || s.isEmpty();
    }
};
typedef std::shared_ptr<const SiSet> SiSetPtr;

class FMIndexOnGenome {
public:
    std::unique_ptr<FMIndexOnOccTable> forwardIndex, reverseIndex;
    SparseSuffixArray forwardSA, reverseSA;
    SequenceBoundary index;
    jlong N;
    SuffixInterval wholeRange;
    SuffixInterval initRange[5];

    FMIndexOnGenome(ACGTSequence bwtF, ACGTSequence bwtR, SparseSuffixArray ssaF, SparseSuffixArray ssaR,
                    SequenceBoundary b, jlong n)
        : forwardIndex(new FMIndexOnOccTable(std::move(bwtF), 128)),
          reverseIndex(new FMIndexOnOccTable(std::move(bwtR), 128)),
          forwardSA(std::move(ssaF)), reverseSA(std::move(ssaR)), index(std::move(b)), N(n) {
        wholeRange = SuffixInterval(0, N);
        forwardSearchAll(FORWARD, wholeRange, initRange);  // FMIndexOnGenome.java:97-98
    }
    jlong textSize() const { return N; }
    SiSetPtr initSet(int d) const {
        auto s = std::make_shared<SiSet>();
        if (d == DIR_FORWARD) { s->kind = SiSet::F; for (int i = 0; i < 5; ++i) s->siF[i] = initRange[i]; }
        else if (d == DIR_BACKWARD) { s->kind = SiSet::B; for (int i = 0; i < 5; ++i) s->siB[i] = initRange[i]; }
        else { s->kind = SiSet::BD; for (int i = 0; i < 5; ++i) { s->siF[i] = initRange[i]; s->siB[i] = initRange[i]; } }
        return s;
    }
    SuffixInterval forwardSearch(int strand, int nextBase, const SuffixInterval& si) const {
        const FMIndexOnOccTable& fm = strand == FORWARD ? *reverseIndex : *forwardIndex;
        return fm.backwardSearch(nextBase, si);
    }
    void forwardSearch(const FMIndexOnOccTable& fm, const jlong lo[5], const jlong hi[5], SuffixInterval out[5]) const {
        for (int i = 0; i < 5; ++i) {
            jlong lb = fm.C[i] + lo[i], ub = fm.C[i] + hi[i];
            out[i] = (lb < ub) ? SuffixInterval(lb, ub) : SuffixInterval();
        }
    }
    void forwardSearchAll(int strand, const SuffixInterval& si, SuffixInterval out[5]) const {
        const FMIndexOnOccTable& fm = strand == FORWARD ? *reverseIndex : *forwardIndex;
        jlong lo[5], hi[5];
        fm.rankACGTN(si.lowerBound, lo);
        fm.rankACGTN(si.upperBound, hi);
        forwardSearch(fm, lo, hi, out);
    }
    SiSetPtr bidirectionalSearch(int strand, const SuffixInterval& siF, const SuffixInterval& siB) const {
        auto s = std::make_shared<SiSet>();
        if (siF.valid) {
            const FMIndexOnOccTable& fm = strand == FORWARD ? *reverseIndex : *forwardIndex;
            jlong lo[5], hi[5];
            fm.rankACGTN(siF.lowerBound, lo);
            fm.rankACGTN(siF.upperBound, hi);
            forwardSearch(fm, lo, hi, s->siF);
            if (!siB.valid) { s->kind = SiSet::F; return s; }
            for (int i = 0; i < 5; ++i) {  // :148-167
                if (!s->siF[i].valid) continue;
                jlong x = 0;
                for (int j = 0; j < i; ++j) x += hi[j] - lo[j];
                jlong y = hi[i] - lo[i];
                s->siB[i] = SuffixInterval(siB.lowerBound + x, siB.lowerBound + x + y);
            }
            s->kind = SiSet::BD;
            return s;
        } else if (siB.valid) {
            const FMIndexOnOccTable& fm = strand == FORWARD ? *forwardIndex : *reverseIndex;
            jlong lo[5], hi[5];
            fm.rankACGTN(siB.lowerBound, lo);
            fm.rankACGTN(siB.upperBound, hi);
            forwardSearch(fm, lo, hi, s->siB);
            s->kind = SiSet::B;
            return s;
        }
        return s;  // SiSet.empty
    }
    jlong toCoordinate(jlong saIndex, int strand, int searchDirection) const {  // :227-238
        int fm = ~(strand ^ (dirIsForward(searchDirection) ? 0 : 1)) & 1;
        if (fm == 0) return forwardSA.get(saIndex, *forwardIndex);
        return N - reverseSA.get(saIndex, *reverseIndex);
    }
    jlong toCoordinate(jlong saIndex, int strand) const {  // :240-252
        if (strand == FORWARD) return reverseIndex->textSize() - reverseSA.get(saIndex, *reverseIndex);
        return forwardSA.get(saIndex, *forwardIndex);
    }
    PosOnGenome toGenomeCoordinate(jlong saIndex, jlong querySize, int strand) const {  // :258-269
        jlong pos = toCoordinate(saIndex, strand);
        if (strand == FORWARD) pos -= querySize;
        pos += 1;
        if (pos >= 0) return index.translate(pos);
        return PosOnGenome{"", 0, false};  // Java null
    }
};

}  // namespace gwo
