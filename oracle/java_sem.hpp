// TEST INFRASTRUCTURE ONLY — CPU oracle for genome-weaver-align's BSF read-alignment path.
// Nothing under oracle/ is linked into, imported by, or executed from the product library
// (genome_weaver_align_b200/csrc). Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may use it, and only as the checker / CPU baseline.
//
// java_sem.hpp: the handful of Java language semantics that change results when restated in C++
// (SURVEY.md appendix A.5): shift-count masking, truncating division, (byte)/(int) narrowing.
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>

namespace gwo {

typedef int64_t jlong;
typedef int32_t jint;
typedef uint64_t u64;

// Java: long << s, >> s, >>> s use (s & 63)
static inline u64 shl(u64 x, jlong s) { return x << (unsigned)(s & 63); }
static inline u64 ushr(u64 x, jlong s) { return x >> (unsigned)(s & 63); }
static inline jlong sshr(jlong x, jlong s) { return x >> (unsigned)(s & 63); }
// Java: int << s uses (s & 31)
static inline uint32_t shl32(uint32_t x, jint s) { return x << (unsigned)(s & 31); }
static inline int8_t jbyte(jlong v) { return (int8_t)(uint8_t)(v & 0xFF); }
static inline int popcount64(u64 x) { return __builtin_popcountll(x); }

// thrown where the reference would throw (run aborts)
struct JavaAbort : public std::runtime_error {
    explicit JavaAbort(const std::string& what) : std::runtime_error(what) {}
};

}  // namespace gwo
