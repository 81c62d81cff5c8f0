// TEST INFRASTRUCTURE ONLY (see java_sem.hpp). CPU restatement of the reference's packed
// sequence and bit-vector types:
//   ACGT            align/src/main/java/org/utgenome/weaver/align/ACGT.java:34-87
//   ACGTSequence    .../align/ACGTSequence.java:47-58 (layout), :165-192 lookup/set, :270-332 subString,
//                   :346-379 reverse/complement, :456-553 fastCount/fastCountACGTN, :570-577 interleave
//   BitVector       .../align/BitVector.java:46-55, :106-116 substring64, :132-136 set, :180-192 _lshift, :217-224 _not
//   QueryMask       .../align/QueryMask.java:41-63 ctor, :73-97 bidirectional mask, :107-109 block mask
#pragma once
#include <algorithm>
#include <string>
#include <vector>

#include "java_sem.hpp"

namespace gwo {

enum : int { A_ = 0, C_ = 1, G_ = 2, T_ = 3, N_ = 4 };

// ACGT.java:37-44 charToACGTCodeTable: A/a 0, C/c 1, G/g 2, T/t/U/u 3, everything else 4
static inline int to3bitCode(unsigned char ch) {
    switch (ch) {
    case 'A': case 'a': return 0;
    case 'C': case 'c': return 1;
    case 'G': case 'g': return 2;
    case 'T': case 't': case 'U': case 'u': return 3;
    default: return 4;
    }
}
static inline char acgtChar(int code) { return "ACGTN"[code > 4 ? 4 : code]; }

class ACGTSequence {
public:
    std::vector<u64> seq;  // 3 longs per 64 bases: N flags | bases 0-31 | bases 32-63
    jlong numBases = 0;

    ACGTSequence() { seq.assign(3, 0); }
    explicit ACGTSequence(jlong n) : numBases(n) { seq.assign((size_t)minArraySize(n), 0); }
    explicit ACGTSequence(const std::string& s) {  // ACGTSequence.java:86-98 (skips ' ')
        jlong cnt = 0;
        for (char c : s) if (c != ' ') cnt++;
        numBases = cnt;
        seq.assign((size_t)minArraySize(cnt), 0);
        jlong idx = 0;
        for (char c : s) { if (c == ' ') continue; set(idx++, to3bitCode((unsigned char)c)); }
    }
    // ACGTSequence.java:129-138
    static jlong minArraySize(jlong numBases) {
        jlong bitSize = numBases * 3;
        jlong blockBitSize = 8 * 3 * 8;
        return ((bitSize + blockBitSize - 1) / blockBitSize) * 3;
    }
    jlong textSize() const { return numBases; }
    // :165-173
    int lookup(jlong index) const {
        size_t pos = (size_t)(index >> 6);
        int offset = (int)(index & 0x3F);
        int shift = 62 - ((int)(index & 0x1F) << 1);
        u64 nFlag = seq[pos * 3] & (1ULL << (63 - offset));
        int code = (int)(seq[pos * 3 + (offset >> 5) + 1] >> shift) & 0x03;
        return nFlag == 0 ? code : 4;
    }
    int getACGT(jlong index) const { return lookup(index); }
    // :180-192
    void set(jlong index, jlong val) {
        size_t pos = (size_t)(index >> 6);
        int offset = (int)(index & 0x3F);
        int shift = (offset & 0x1F) << 1;
        seq[pos * 3] &= ~(1ULL << (63 - offset));
        seq[pos * 3] |= (u64)((val >> 2) & 1) << (63 - offset);
        size_t bPos = pos * 3 + (offset >> 5) + 1;
        seq[bPos] &= ~(0xC000000000000000ULL >> shift);
        seq[bPos] |= (u64)(val & 3) << (62 - shift);
    }
    void append(int code) {
        jlong index = numBases++;
        size_t need = (size_t)minArraySize(numBases);
        if (seq.size() < need) {
            size_t ns = std::max(need, seq.size() * 2);
            seq.resize((ns + 2) / 3 * 3, 0);
        }
        set(index, code);
    }
    // :99-107
    ACGTSequence replaceN_withA() const {
        ACGTSequence n(*this);
        for (jlong i = 0; i < numBases; ++i) if (lookup(i) == N_) n.set(i, A_);
        return n;
    }
    // :270-332 (semantics: copy of [start,end); restated base by base — the word-shuffling there is an optimisation)
    ACGTSequence subString(jlong start, jlong end) const {
        if (start > end) throw JavaAbort("IllegalArgumentException: invalid range in subString");
        ACGTSequence ss(end - start);
        for (jlong i = 0; i < end - start; ++i) {
            // the reference copies raw words, so an N-flagged base keeps its underlying 2-bit code
            size_t pos = (size_t)((start + i) >> 6);
            int offset = (int)((start + i) & 0x3F);
            int shift = 62 - ((offset & 0x1F) << 1);
            int code2 = (int)(seq[pos * 3 + (offset >> 5) + 1] >> shift) & 3;
            int nf = (int)((seq[pos * 3] >> (63 - offset)) & 1);
            ss.setRaw(i, nf, code2);
        }
        return ss;
    }
    void setRaw(jlong index, int nflag, int code2) {
        size_t pos = (size_t)(index >> 6);
        int offset = (int)(index & 0x3F);
        int shift = (offset & 0x1F) << 1;
        seq[pos * 3] |= (u64)nflag << (63 - offset);
        seq[pos * 3 + (offset >> 5) + 1] |= (u64)code2 << (62 - shift);
    }
    // :346-352
    ACGTSequence reverse() const {
        ACGTSequence rev(numBases);
        for (jlong i = 0; i < numBases; ++i) rev.set(i, lookup(numBases - i - 1));
        return rev;
    }
    // :359-369 (N flags kept, base words inverted)
    ACGTSequence complement() const {
        ACGTSequence c(numBases);
        size_t nb = seq.size() / 3;
        c.seq.assign(seq.size(), 0);
        for (size_t i = 0; i < nb; ++i) {
            c.seq[i * 3] = seq[i * 3];
            c.seq[i * 3 + 1] = ~seq[i * 3 + 1];
            c.seq[i * 3 + 2] = ~seq[i * 3 + 2];
        }
        return c;
    }
    ACGTSequence reverseComplement() const {  // :371-379
        ACGTSequence rc = reverse();
        size_t nb = rc.seq.size() / 3;
        for (size_t i = 0; i < nb; ++i) { rc.seq[i * 3 + 1] = ~rc.seq[i * 3 + 1]; rc.seq[i * 3 + 2] = ~rc.seq[i * 3 + 2]; }
        return rc;
    }
    std::string toString() const {
        std::string s((size_t)numBases, 'N');
        for (jlong i = 0; i < numBases; ++i) s[(size_t)i] = acgtChar(lookup(i));
        return s;
    }
    // :570-577
    static u64 interleave32With0(u64 v) {
        v = ((v & 0xFFFF0000ULL) << 16) | (v & 0x0000FFFFULL);
        v = ((v << 8) | v) & 0x00FF00FF00FF00FFULL;
        v = ((v << 4) | v) & 0x0F0F0F0F0F0F0F0FULL;
        v = ((v << 2) | v) & 0x3333333333333333ULL;
        v = ((v << 1) | v) & 0x5555555555555555ULL;
        return v;
    }
    // :456-511
    jlong fastCount(int base, jlong start, jlong end) const {
        jlong count = 0;
        if (base == N_) {
            jlong sPos = (jlong)((u64)start >> 6);
            int sOffset = (int)(start & 0x3F);
            jlong ePos = (jlong)((u64)(end + 64 - 1) >> 6);
            for (; sPos < ePos; ++sPos) {
                u64 mask = ~0ULL;
                if (sOffset != 0) { mask = ushr(mask, sOffset); sOffset = 0; }
                if (sPos == ePos - 1) {
                    int eOffset = (int)(end & 0x3F);
                    u64 rMask = (eOffset == 0) ? ~0ULL : ~(shl(1ULL, 64 - eOffset) - 1);
                    mask &= rMask;
                }
                count += popcount64(seq[(size_t)sPos * 3] & mask);
            }
        } else {
            jlong sPos = (jlong)((u64)start >> 5);
            int sOffset = (int)(start & 0x1F);
            jlong ePos = (jlong)((u64)(end + 32 - 1) >> 5);
            for (; sPos < ePos; ++sPos) {
                u64 mask = ~0ULL;
                if (sOffset != 0) { mask = ushr(mask, sOffset * 2); sOffset = 0; }
                size_t bIndex = (size_t)(sPos / 2 * 3);
                int block = (int)(sPos % 2);
                u64 v = seq[bIndex + 1 + block];
                u64 nFlag = interleave32With0(ushr(seq[bIndex], 32 * (1 - block)));
                if (sPos == ePos - 1) {
                    int eOffset = (int)(end & 0x1F);
                    u64 rMask = (eOffset == 0) ? ~0ULL : ~(shl(1ULL, (32 - eOffset) * 2) - 1);
                    mask &= rMask;
                }
                u64 r = ~0ULL;
                r &= ((base & 2) == 0 ? ~v : v) >> 1;
                r &= ((base & 1) == 0 ? ~v : v);
                r &= 0x5555555555555555ULL;
                r &= ~nFlag;
                r &= mask;
                count += popcount64(r);
            }
        }
        return count;
    }
    // :513-553
    void fastCountACGTN(jlong start, jlong end, jlong count[5]) const {
        for (int i = 0; i < 5; ++i) count[i] = 0;
        jlong sPos = (jlong)((u64)start >> 5);
        int sOffset = (int)(start & 0x1F);
        jlong ePos = (jlong)((u64)(end + 32 - 1) >> 5);
        for (; sPos < ePos; ++sPos) {
            u64 mask = ~0ULL;
            if (sOffset != 0) { mask = ushr(mask, sOffset * 2); sOffset = 0; }
            size_t bIndex = (size_t)(sPos / 2 * 3);
            int block = (int)(sPos % 2);
            u64 v = seq[bIndex + 1 + block];
            u64 nFlag = interleave32With0(ushr(seq[bIndex], 32 * (1 - block)));
            if (sPos == ePos - 1) {
                int eOffset = (int)(end & 0x1F);
                u64 rMask = (eOffset == 0) ? ~0ULL : ~(shl(1ULL, (32 - eOffset) * 2) - 1);
                mask &= rMask;
            }
            for (int base = 0; base < 4; ++base) {
                u64 r = ~0ULL;
                r &= ((base & 2) == 0 ? ~v : v) >> 1;
                r &= ((base & 1) == 0 ? ~v : v);
                r &= 0x5555555555555555ULL;
                r &= ~nFlag;
                r &= mask;
                count[base] += popcount64(r);
            }
            count[N_] += popcount64(nFlag & mask);
        }
    }
};

// BitVector.java — only the members on the hot path; block index checks mirror Java's
// ArrayIndexOutOfBoundsException (extended mode reads out-of-range blocks as 0)
class BitVector {
public:
    std::vector<u64> block;
    jlong size;
    bool extended;  // false: throw JavaAbort on out-of-range block access like the JVM would
    explicit BitVector(jlong size_, bool extended_ = false) : size(size_), extended(extended_) {
        block.assign((size_t)((size_ + 63) / 64), 0);
    }
    u64 at(jlong i) const {
        if (i < 0 || i >= (jlong)block.size()) {
            if (extended) return 0;
            throw JavaAbort("ArrayIndexOutOfBoundsException in BitVector");
        }
        return block[(size_t)i];
    }
    void set(jlong index) { block[(size_t)(index / 64)] |= shl(1ULL, index % 64); }  // :132-136
    // :106-116
    u64 substring64(jlong start, jlong end) const {
        jint pos = (jint)(start / 64);
        if (pos >= (jint)block.size()) return 0;
        jlong range = end - start;
        u64 mask = (range >= 64) ? ~0ULL : ~shl(~0ULL, range);
        jlong offset = start % 64;
        u64 low = ushr(at(pos), offset);
        u64 high = (pos + 1 < (jint)block.size()) ? shl(at(pos + 1) & ~shl(~0ULL, offset), 64 - offset) : 0ULL;
        return (high | low) & mask;
    }
    // :180-192 (len may be negative through (byte) overflow in StaircaseFilter)
    BitVector& _lshift(jint len) {
        jint blockOffset = len / 64;  // C++ '/' and '%' truncate toward zero like Java
        jlong offset = len % 64;
        u64 lowMask = ~ushr(~0ULL, offset);
        for (jint i = (jint)block.size() - 1; i >= 0; --i) {
            jint b = i - blockOffset;
            u64 high = b >= 0 ? shl(at(b), offset) : 0ULL;
            u64 low = (b - 1 >= 0) ? ushr(at(b - 1) & lowMask, 64 - offset) : 0ULL;
            block[(size_t)i] = high | low;
        }
        return *this;
    }
    // :217-224 ( (int) size % B ; shift count 0 when size%64==0 => last block becomes 0 )
    BitVector& _not() {
        if (block.empty()) {
            if (extended) return *this;
            throw JavaAbort("ArrayIndexOutOfBoundsException in BitVector._not");
        }
        for (size_t i = 0; i + 1 < block.size(); ++i) block[i] = ~block[i];
        jint offset = (jint)size % 64;
        block[block.size() - 1] = (~block[block.size() - 1]) & ~shl(~0ULL, offset);
        return *this;
    }
};

class QueryMask {
public:
    jint m;
    std::vector<BitVector> patternMaskF, patternMaskR;
    explicit QueryMask(const ACGTSequence& query) {  // QueryMask.java:41-63
        m = (jint)query.textSize();
        for (int i = 0; i < 4; ++i) { patternMaskF.emplace_back(m, true); patternMaskR.emplace_back(m, true); }
        for (jint i = 0; i < m; ++i) {
            int ch = query.getACGT(i);
            if (ch == N_) {
                for (int e = 0; e < 4; ++e) { patternMaskF[e].set(i); patternMaskR[e].set(m - i - 1); }
            } else {
                patternMaskF[ch].set(i);
                patternMaskR[ch].set(m - i - 1);
            }
        }
    }
    // :73-97
    u64 getBidirectionalPatternMask64(bool dIsForward, jint nextACGTIndex, jint pivot, jint cursor, int ch, jint margin) const {
        u64 p;
        if (dIsForward) {
            jint pos = nextACGTIndex - margin;
            if (pos < 0) {
                p = patternMaskF[ch].substring64(0, 64);
                p = shl(p, -pos);
            } else
                p = patternMaskF[ch].substring64(pos, m);
        } else {
            jint b = m - pivot;
            jint rshift = pivot - cursor - margin;
            p = patternMaskR[ch].substring64(b, b + 64);
            if (rshift >= 0) p = ushr(p, rshift);
            else p = shl(p, -rshift);
        }
        return p;
    }
    // :107-109
    u64 getPatternMaskIn64bit(int ch, jint blockIndex, jint w) const {
        return patternMaskF[ch].substring64((jlong)blockIndex * w, (jlong)(blockIndex + 1) * w);
    }
};

}  // namespace gwo
