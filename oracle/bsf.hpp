// TEST INFRASTRUCTURE ONLY (see java_sem.hpp). CPU restatement of the reference's default aligner:
//   BidirectionalSuffixFilter  align/src/main/java/org/utgenome/weaver/align/strategy/BidirectionalSuffixFilter.java
//     StateQueue (java.util.PriorityQueue + comparator)  :133-165
//     AlignmentProcess  ctor :189-198, align :210-268, align_internal :278-477, next :486-488,
//                       reportExactMatchAlignment :490-494, verify :496-560, reportAlignment :562-586,
//                       AlignmentResultHolder.add :606-628
//     SearchState       :658-877
//   AlignmentRecord.convert    .../record/AlignmentRecord.java:181-276
#pragma once
#include "bsf_types.hpp"

namespace gwo {

struct SearchState;
typedef std::shared_ptr<SearchState> StatePtr;

struct SearchState {
    Cursor cursor;
    SuffixInterval currentSi;  // nullable
    SiSetPtr siTable;          // nullable (clipped fragment)
    std::shared_ptr<ReadAlignmentNFA> automaton;  // nullable (clipped fragment)
    jint state = 0;
    StatePtr nextSplit;

    SearchState(const SuffixInterval& cur, int ch, const Cursor& c, SiSetPtr si, std::shared_ptr<ReadAlignmentNFA> a,
                bool hasMatch, jint minK, jint priority, StatePtr split)
        : cursor(c), currentSi(cur), siTable(std::move(si)), automaton(std::move(a)), nextSplit(std::move(split)) {
        state = ((ch & 0x7) << 5) | ((minK & 0xFF) << 8) | ((priority & 0xFF) << 16) | ((hasMatch ? 1 : 0) << 24);
    }
    jint getLowerBoundOfK() const { return ((uint32_t)state >> 8) & 0xFF; }
    void setLowerBoundOfK(jint diff) { state &= ~(0xFF << 8); state |= (diff & 0xFF) << 8; }
    jint getNumSplit() const { return nextSplit ? 1 + nextSplit->getNumSplit() : 0; }
    jint getPriority() const { return ((uint32_t)state >> 16) & 0xFF; }
    bool hasHit() const { return (((uint32_t)state >> 24) & 1) != 0; }
    bool isClipped() const { return (((uint32_t)state >> 25) & 1) != 0; }
    void updateClippedFlag() { state |= 1 << 25; }
    bool isFinished() const { return (state & 0x1F) == 0x1F; }
    int currentACGT() const { int c = (state >> 5) & 7; return c > 4 ? 4 : c; }
    void updateFlag(int ch) { state |= 1 << ch; }
    void updateSplitFlag() { state |= 1 << 4; }
    bool isChecked(int ch) const { return (state & (1 << ch)) != 0; }
};

class BidirectionalSuffixFilter {
public:
    const FMIndexOnGenome& fmIndex;
    const ACGTSequence* reference;  // may be null (exact-only use, BWAlignTest.align3)
    AlignmentConfig config;
    StaircaseFilterHolder staircaseFilterHolder;
    // counters accumulated over all reads (the reference's per-read debug counters, :588-592)
    uint64_t totalFMIndexSearches = 0, totalSW = 0, totalCutOff = 0, totalFiltered = 0;

    BidirectionalSuffixFilter(const FMIndexOnGenome& fm, const ACGTSequence* ref, const AlignmentConfig& cfg)
        : fmIndex(fm), reference(ref), config(cfg) {}

    struct Read { std::string name; ACGTSequence seq; std::string qual; bool qualIsNull; };

    // result of one read: the ReadHit chains handed to report(), in order
    struct ReadResult {
        std::vector<ReadHitPtr> reported;
        bool unmapped = false;
        jint numFMIndexSearches = 0, numSW = 0, numCutOff = 0, numFiltered = 0;
    };

    class AlignmentProcess {
    public:
        BidirectionalSuffixFilter& P;
        const Read& read;
        jint m;
        ACGTSequence q[2];
        std::unique_ptr<QueryMask> queryMask[2];
        std::vector<StatePtr> queue;  // java.util.PriorityQueue storage
        std::vector<ReadHitPtr> hitList;
        jint k, minMismatches, maxMatchLength = 0, bestScore = -1;
        jint numFMIndexSearches = 0, numSW = 0, numCutOff = 0, numFiltered = 0;

        AlignmentProcess(BidirectionalSuffixFilter& p, const Read& r) : P(p), read(r) {
            m = (jint)r.seq.textSize();
            q[0] = r.seq;
            q[1] = q[0].complement();
            k = P.config.getMaximumEditDistance(m);
            minMismatches = k + 1;
        }
        const AlignmentConfig& config() const { return P.config; }
        const StaircaseFilter& getStairCaseFilter(jint queryLength) {
            if (minMismatches < 0) throw JavaAbort("NegativeArraySizeException in StaircaseFilter");
            return P.staircaseFilterHolder.get(queryLength, minMismatches, P.config.extended);
        }

        // ---- state score helpers (SearchState.score / upperBoundOfScore, :781-801)
        jint score(const SearchState& s) const {
            jint numSplits = s.getNumSplit();
            jint nm = s.getLowerBoundOfK() - numSplits;
            jint mm = s.cursor.getProcessedBases() - nm;
            jint sc = mm * config().matchScore - nm * config().mismatchPenalty - numSplits * config().splitOpenPenalty;
            return s.nextSplit ? sc + score(*s.nextSplit) : sc;
        }
        jint upperBoundOfScore(const SearchState& s) const {
            jint numSplits = s.getNumSplit();
            jint nm = s.getLowerBoundOfK() - numSplits;
            jint mm = s.cursor.getProcessedBases() + s.cursor.getRemainingBases() - nm;
            jint sc = mm * config().matchScore - nm * config().mismatchPenalty - numSplits * config().splitOpenPenalty;
            return s.nextSplit ? sc + upperBoundOfScore(*s.nextSplit) : sc;
        }
        jint compare(const SearchState& o1, const SearchState& o2) const {
            jint diff = o1.getPriority() - o2.getPriority();
            if (diff == 0) diff = -(score(o1) - score(o2));
            if (diff == 0) diff = -(o1.cursor.getProcessedBases() - o2.cursor.getProcessedBases());
            return diff;
        }
        // ---- java.util.PriorityQueue.offer / poll with the comparator above
        void queueAdd(const StatePtr& e) {
            if (!e) return;  // StateQueue.add(null) is a no-op, :154-159
            size_t kk = queue.size();
            queue.push_back(e);
            while (kk > 0) {
                size_t parent = (kk - 1) >> 1;
                if (compare(*e, *queue[parent]) >= 0) break;
                queue[kk] = queue[parent];
                kk = parent;
            }
            queue[kk] = e;
        }
        StatePtr queuePoll() {
            size_t s = queue.size() - 1;
            StatePtr result = queue[0];
            StatePtr x = queue[s];
            queue.pop_back();
            if (s != 0) {
                size_t kk = 0, half = s >> 1;
                while (kk < half) {
                    size_t child = (kk << 1) + 1;
                    StatePtr c = queue[child];
                    size_t right = child + 1;
                    if (right < s && compare(*c, *queue[right]) > 0) c = queue[child = right];
                    if (compare(*x, *c) <= 0) break;
                    queue[kk] = c;
                    kk = child;
                }
                queue[kk] = x;
            }
            return result;
        }
        // ---- SearchState factories
        StatePtr initialState(const Cursor& cursor, jint priority) {  // :753-757
            auto nfa = std::make_shared<ReadAlignmentNFA>(k);
            nfa->activateDiagonalStates();
            return std::make_shared<SearchState>(SuffixInterval(), N_, cursor, P.fmIndex.initSet(cursor.getSearchDirection()),
                                                 nfa, false, 0, priority, nullptr);
        }
        StatePtr nextStateAfterSplit(SearchState& c) {  // :803-819
            c.updateSplitFlag();
            jint minK = c.getLowerBoundOfK();
            if (minK < k) {
                Cursor l, r;
                c.cursor.split(l, r);
                auto ns = std::make_shared<SearchState>(c.currentSi, c.currentACGT(), l, c.siTable,
                        std::make_shared<ReadAlignmentNFA>(c.automaton->nextStateAfterSplit(k)), c.hasHit(), minK, c.getPriority(), nullptr);
                auto nfa = std::make_shared<ReadAlignmentNFA>(k);
                nfa->activateDiagonalStates();
                ns->nextSplit = std::make_shared<SearchState>(SuffixInterval(), N_, r, P.fmIndex.initSet(r.getSearchDirection()), nfa,
                                                              false, minK, c.getPriority(), nullptr);
                return ns;
            }
            return nullptr;
        }
        StatePtr nextStateAfterClipping(SearchState& c) {  // :821-838
            c.updateSplitFlag();
            jint minK = c.getLowerBoundOfK();
            if (minK < k) {
                Cursor l, r;
                c.cursor.split(l, r);
                auto ns = std::make_shared<SearchState>(c.currentSi, c.currentACGT(), l, c.siTable,
                        std::make_shared<ReadAlignmentNFA>(c.automaton->nextStateAfterSplit(k)), c.hasHit(), minK, c.getPriority(), nullptr);
                ns->nextSplit = std::make_shared<SearchState>(SuffixInterval(), N_, r, nullptr, nullptr, false, minK, c.getPriority(), nullptr);
                ns->nextSplit->updateClippedFlag();
                return ns;
            }
            return nullptr;
        }
        StatePtr nextState(SearchState& c, int ch, const SiSetPtr& nextSi, const QueryMask& qm, const StaircaseFilter& sf) {  // :840-852
            ReadAlignmentNFA::NextState next = c.automaton->nextState(c.cursor, ch, qm, sf);
            if (!next.valid) return nullptr;
            jint nextMinK = next.kOffset;
            Cursor nextCursor = c.cursor.next();
            SuffixInterval si = nextCursor.isForwardSearch() ? c.siTable->getForward(ch) : c.siTable->getBackward(ch);
            return std::make_shared<SearchState>(si, ch, nextCursor, nextSi,
                    std::make_shared<ReadAlignmentNFA>(std::move(next.automaton), next.kOffset), next.hasMatch, nextMinK,
                    c.getPriority(), c.nextSplit);
        }
        // SearchState.update, :861-875
        static StatePtr update(const StatePtr& self, const StatePtr& oldState, const StatePtr& newState) {
            if (oldState == self) return newState;
            StatePtr prev = self;
            while (prev->nextSplit != oldState) {
                prev = prev->nextSplit;
                if (!prev) return nullptr;
            }
            prev->nextSplit = newState;
            return self;
        }

        void addResult(const ReadHitPtr& hit) {  // AlignmentResultHolder.add, :606-628
            jint newK = hit->getTotalDifferences();
            jint matchLen = hit->getTotalMatchLength();
            jint newScore = hit->getTotalScore(config());
            if (newScore > bestScore) {
                if (matchLen > 0 && newK <= minMismatches) minMismatches = newK;
                bestScore = newScore;
            }
            if (maxMatchLength < matchLen) maxMatchLength = matchLen;
            std::vector<ReadHitPtr> newList;
            for (auto& each : hitList)
                if (each->getTotalDifferences() <= minMismatches && each->getTotalMatchLength() >= maxMatchLength) newList.push_back(each);
            newList.push_back(hit);
            hitList.swap(newList);
        }
        void reportExactMatchAlignment(const FMQuickScan& f) {  // :490-494
            PosOnGenome pos = P.fmIndex.toGenomeCoordinate(f.si.lowerBound, m, f.strand);
            if (!pos.valid) throw JavaAbort("NullPointerException: exact hit wraps the cyclic boundary");
            addResult(ReadHit::make(false, pos.chr, pos.pos, m, 0, m, 0, f.strand, CIGAR(std::to_string(m) + "M"), (jint)f.si.range()));
        }
        ReadHitPtr verify(const SearchState& s) {  // :496-560
            const Cursor& cursor = s.cursor;
            if (s.isClipped())
                return ReadHit::make(true, "", -1, 0, cursor.start, cursor.end, 0, cursor.getStrand(),
                                     CIGAR(std::to_string(cursor.getFragmentLength()) + "S"), 0);
            if (!s.currentSi.valid) return ReadHit::noHit(cursor.getStrand());
            const SuffixInterval& si = s.currentSi;
            jlong seqIndex = P.fmIndex.toCoordinate(si.lowerBound, cursor.getStrand(), cursor.getSearchDirection());
            jlong x = seqIndex - cursor.getOffsetOfSearchHead();
            jint fragmentLength = cursor.getFragmentLength();
            jlong refStart = std::max<jlong>(0, x - k);
            jlong refEnd = std::min<jlong>(x + fragmentLength + k, P.fmIndex.textSize());
            if (!P.reference) throw JavaAbort("NullPointerException: reference is null");
            ACGTSequence ref = P.reference->subString(refStart, refEnd);
            ACGTSequence query = q[cursor.getStrand()].subString(cursor.start, cursor.end);
            if (cursor.getStrand() == REVERSE) query = query.reverse();
            Alignment alignment = alignBlockDetailed(ref, query, config().bandWidth);
            ++numSW;
            if (!alignment.valid) return ReadHit::noHit(cursor.getStrand());
            PosOnGenome p = P.fmIndex.index.translate(refStart + alignment.pos + 1);
            return ReadHit::make(false, p.chr, p.pos, fragmentLength, cursor.start, cursor.end, alignment.numMismatches,
                                 cursor.getStrand(), alignment.cigar, (jint)si.range());
        }
        void reportAlignment(SearchState& c) {  // :562-586
            ReadHitPtr alignment = verify(c);
            {
                ReadHitPtr nextHit = alignment;
                for (StatePtr next = c.nextSplit; next; next = next->nextSplit) {
                    ReadHitPtr result = verify(*next);
                    nextHit->nextSplit = result;
                }
            }
            if (alignment->getTotalMatchLength() == 0) return;
            jint newK = alignment->getTotalDifferences();
            if (newK > k) return;
            c.setLowerBoundOfK(newK);
            addResult(sortSplits(alignment));
        }

        void align_internal() {  // :278-477
            {
                jlong countN = q[0].fastCount(N_, 0, m);
                if (countN > k) return;
                if (countN > 0) { q[0] = q[0].replaceN_withA(); q[1] = q[1].replaceN_withA(); }
            }
            {
                FMQuickScan scanF = FMQuickScan::scan(P.fmIndex, q[0], FORWARD);
                if (scanF.numMismatches == 0) { minMismatches = 0; reportExactMatchAlignment(scanF); return; }
                FMQuickScan scanR = FMQuickScan::scan(P.fmIndex, q[1], REVERSE);
                if (scanR.numMismatches == 0) { minMismatches = 0; reportExactMatchAlignment(scanR); return; }
                if (k == 0) return;
                queryMask[0].reset(new QueryMask(q[0]));
                queryMask[1].reset(new QueryMask(q[1]));
                StatePtr sF, sR;
                if (scanF.numMismatches <= k) {
                    if (scanF.longestStart != 0 && scanF.longestStart < m)
                        sF = initialState(Cursor(FORWARD, DIR_BIDIRECTIONAL_FORWARD, 0, m, scanF.longestStart, scanF.longestStart), scanF.numMismatches);
                    else
                        sF = initialState(Cursor(FORWARD, DIR_FORWARD, 0, m, 0, 0), scanF.numMismatches);
                }
                if (scanR.numMismatches <= k) {
                    if (scanR.longestStart != 0 && scanR.longestStart < m)
                        sR = initialState(Cursor(REVERSE, DIR_BIDIRECTIONAL_FORWARD, 0, m, scanR.longestStart, scanR.longestStart), scanR.numMismatches);
                    else
                        sR = initialState(Cursor(REVERSE, DIR_FORWARD, 0, m, 0, 0), scanR.numMismatches);
                }
                queueAdd(sF);
                queueAdd(sR);
            }
            const jint fmIndexSearchUpperBound = m * 20;
            while (!queue.empty()) {
                if (numFMIndexSearches > fmIndexSearchUpperBound) break;
                StatePtr baseState = queuePoll();
                StatePtr c = baseState;
                {
                    StatePtr next = c;
                    bool reported = false;
                    while (next->hasHit() || next->isClipped() || next->cursor.getRemainingBases() == 0) {
                        if (!next->nextSplit) { reportAlignment(*c); reported = true; break; }
                        next = next->nextSplit;
                    }
                    if (reported) continue;
                    c = next;
                }
                if (c->isFinished()) continue;
                jint nm = c->getLowerBoundOfK();
                if (nm > minMismatches) { ++numCutOff; continue; }
                jint allowedMismatches = minMismatches - nm;
                if (allowedMismatches < 0) continue;
                {
                    jint scoreUB = upperBoundOfScore(*baseState);
                    if (scoreUB < 0 || scoreUB < bestScore) { ++numCutOff; continue; }
                }
                const int strandIndex = c->cursor.getStrand();
                int nextBase = q[strandIndex].getACGT(c->cursor.getNextACGTIndex());
                {
                    int ch = nextBase;
                    if (!c->isChecked(ch)) {
                        c->updateFlag(ch);
                        if (!c->siTable->isEmpty(ch)) {
                            SiSetPtr nextSi = c->cursor.nextSi(P.fmIndex, *c->siTable, ch);
                            ++numFMIndexSearches;
                            StatePtr ns = nextState(*c, ch, nextSi, *queryMask[strandIndex], getStairCaseFilter(m));
                            if (ns) { queueAdd(update(baseState, c, ns)); continue; }
                            else ++numFiltered;
                        }
                    }
                }
                {
                    const StaircaseFilter& sf = getStairCaseFilter(m);
                    for (int ch = 0; ch < 4; ++ch) {
                        if (!c->isChecked(ch)) {
                            c->updateFlag(ch);
                            if (!c->siTable->isEmpty(ch)) {
                                SiSetPtr nextSi = c->cursor.nextSi(P.fmIndex, *c->siTable, ch);
                                ++numFMIndexSearches;
                                StatePtr ns = nextState(*c, ch, nextSi, *queryMask[strandIndex], sf);
                                if (ns) queueAdd(update(baseState, c, ns));
                                else ++numFiltered;
                            }
                        }
                    }
                }
                c->updateSplitFlag();
                if (baseState->getNumSplit() < config().numSplitAlowed && nm + 1 <= minMismatches) {
                    const jint index = c->cursor.getNextACGTIndex();
                    if (index > config().indelEndSkip && m - index >= config().indelEndSkip) {
                        StatePtr ns = nextStateAfterSplit(*c);
                        if (ns) queueAdd(update(baseState, c, ns));
                        else ++numFiltered;
                        StatePtr clipped = nextStateAfterClipping(*c);
                        if (clipped) queueAdd(update(baseState, c, clipped));
                    }
                }
            }
        }

        ReadResult align() {  // :210-268
            ReadResult res;
            align_internal();
            bool hasHit = minMismatches <= k && !hitList.empty();
            if (hasHit) {
                switch (config().reportType) {
                case 1: for (auto& h : hitList) res.reported.push_back(h); break;
                case 0: res.reported.push_back(hitList[0]); break;
                default: {
                    jint mx = std::min<jint>(config().topL, (jint)hitList.size());
                    for (jint i = 0; i < mx; ++i) res.reported.push_back(hitList[(size_t)i]);
                }
                }
            } else {
                res.unmapped = true;
                res.reported.push_back(ReadHit::make(false, "*", 0, 0, 0, 0, -1, FORWARD, CIGAR(), 0));
            }
            res.numFMIndexSearches = numFMIndexSearches; res.numSW = numSW; res.numCutOff = numCutOff; res.numFiltered = numFiltered;
            return res;
        }
    };

    ReadResult align(const Read& read) {
        AlignmentProcess p(*this, read);
        ReadResult r = p.align();
        totalFMIndexSearches += (uint64_t)r.numFMIndexSearches; totalSW += (uint64_t)r.numSW;
        totalCutOff += (uint64_t)r.numCutOff; totalFiltered += (uint64_t)r.numFiltered;
        return r;
    }

    // AlignmentRecord.convert, AlignmentRecord.java:181-276 ; returns nullptr where the reference returns null
    static AlignmentRecordPtr convert(const ReadHit& hit, const Read& read, jint numOtherBestHits) {
        const jint numHits = hit.numHits + numOtherBestHits;
        ACGTSequence query = read.seq;
        const jint m = (jint)query.textSize();
        std::string qual = read.qual;
        bool qualNull = read.qualIsNull;
        if (hit.strand != FORWARD) {
            query = query.reverseComplement();
            if (!qualNull) qual = std::string(qual.rbegin(), qual.rend());
        }
        auto mk = [&](const std::string& chr, int strand, jint start, jint end, jint nm, const CIGAR& cigar, const std::string& qs,
                      const std::string& ql, bool qn, jint nbest, const std::string& st) {
            auto r = std::make_shared<AlignmentRecord>();
            r->readName = read.name; r->chr = chr; r->strand = strand; r->start = start; r->end = end; r->numMismatches = nm;
            r->cigar = cigar; r->querySeq = qs; r->qual = ql; r->qualIsNull = qn; r->score = 1; r->numBestHits = nbest; r->alignmentState = st;
            return r;
        };
        if (!hit.nextSplit)
            return mk(hit.chr, hit.strand, (jint)hit.pos, (jint)hit.pos + hit.matchLength, hit.getTotalDifferences(), hit.cigar,
                      query.toString(), qual, qualNull, numHits, hit.getAlignmentState());
        jint qualLen = !qualNull ? (jint)qual.size() : hit.matchLength;
        const ReadHit& split = *hit.nextSplit;
        ACGTSequence s1 = query.subString(hit.qStart, hit.qEnd);
        ACGTSequence s2 = query.subString(split.qStart, split.qEnd);
        jint b1 = std::min(qualLen, hit.matchLength);
        jint b2 = std::min(qualLen, m);
        std::string q1, q2;
        if (!qualNull) {
            if (b1 > b2) throw JavaAbort("StringIndexOutOfBoundsException in convert");
            q1 = qual.substr(0, (size_t)b1);
            q2 = qual.substr((size_t)b1, (size_t)(b2 - b1));
        }
        if (hit.isUnique()) {
            if (split.isUnique()) {
                if (hit.chr == split.chr) {
                    auto rec = mk(hit.chr, hit.strand, (jint)hit.pos, (jint)hit.pos + hit.matchLength, hit.diff, hit.cigar, s1.toString(),
                                  q1, qualNull, 1, hit.getAlignmentStateOf(&hit));
                    rec->split = mk(split.chr, split.strand, (jint)split.pos, (jint)split.pos + split.matchLength, split.diff, split.cigar,
                                    s2.toString(), q2, qualNull, 1, split.getAlignmentStateOf(&hit));
                    return rec;
                }
                if (hit.matchLength >= split.matchLength) {
                    CIGAR cigar(hit.cigar);
                    cigar.add(split.matchLength, 'S');
                    return mk(hit.chr, hit.strand, (jint)hit.pos, (jint)hit.pos + m, hit.diff, cigar, query.toString(), qual, qualNull,
                              numHits, hit.getAlignmentStateOf(&hit));
                }
                CIGAR cigar(hit.cigar);
                cigar.add(split.cigar);
                return mk(split.chr, split.strand, (jint)split.pos - hit.matchLength, (jint)split.pos - hit.matchLength + m, split.diff,
                          cigar, query.toString(), qual, qualNull, numHits, split.getAlignmentStateOf(&hit));
            }
            CIGAR cigar(hit.cigar);
            cigar.add(split.getFragmentLength(), 'S');
            return mk(hit.chr, hit.strand, (jint)hit.pos, (jint)hit.pos + cigar.getUnclippedSize(), hit.diff, cigar, query.toString(),
                      qual, qualNull, numHits, hit.getAlignmentStateOf(&hit));
        }
        if (split.isUnique()) {
            CIGAR cigar;
            cigar.add(hit.matchLength, 'S');
            cigar.add(split.cigar);
            return mk(split.chr, split.strand, (jint)split.pos - hit.matchLength, (jint)split.pos - hit.matchLength + m, split.diff, cigar,
                      query.toString(), qual, qualNull, numHits, split.getAlignmentStateOf(&hit));
        }
        return nullptr;
    }
};

}  // namespace gwo
