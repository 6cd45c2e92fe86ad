// tk_swar.cuh -- SWAR (SIMD-within-a-register) helpers shared by the JSON kernels:
// carry-free byte classification on 32-bit words, flag gathering, 8-digit multiply-shift
// reductions.  Device-only; plain 32-bit integer arithmetic (the __v*4 video intrinsics are
// emulated with ~10 instructions each on sm_100, these forms take 4-6).
#pragma once
#include "tk_internal.h"
#include "tk_json.cuh"

// ---------------------------------------------------------------- SWAR helpers
// Token-character class, four bytes at a time with plain 32-bit arithmetic (the __v*4 video
// intrinsics are emulated with ~10 instructions each on sm_100).  A byte is a token character
// iff  0x2B <= c <= 0x7F,  c != ','  and  (c | 6) != 0x5F  (which removes '[' and ']' -- and
// 'Y' and '_', neither of which occurs in a JSON number).  tk_is_numclass() (tk_number.cuh) is the scalar
// twin used by the separator checks; the two must agree for the parallel validation to tile
// the record.  Result: 0x80 in every byte lane that is a token character.
__device__ __forceinline__ uint32_t tokclass4(uint32_t x) {
    const uint32_t lo7 = x & 0x7F7F7F7Fu;
    const uint32_t ge = (lo7 + 0x55555555u) & ~x;                          // bit7: 0x2B <= c <= 0x7F
    const uint32_t yc = x ^ 0x2C2C2C2Cu;                                   // zero byte where c == ','
    const uint32_t nc = ((yc & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | yc;           // bit7: c != ','
    const uint32_t yb = (x | 0x06060606u) ^ 0x5F5F5F5Fu;                   // zero byte where (c|6) == 0x5F
    const uint32_t nb = ((yb & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | yb;           // bit7: not a bracket
    return ge & nc & nb & 0x80808080u;
}
// bits 7,15,23,31 of m (nothing else set) -> bits 0..3.  One multiply puts copies of the four flags
// at bits 28..31 (7+21, 15+14, 23+7, 31+0; all partial products land on distinct bits, no carries).
__device__ __forceinline__ uint32_t hibits_to_nibble(uint32_t m) {
    return (m * 0x00204081u) >> 28;
}
// a * b + c issued as an integer multiply-add (FMA pipe); the kernels in this file are bound by
// the ALU pipe (logic / shift / add), so merges of disjoint bit fields and plain additions in the
// hottest helpers are expressed this way
__device__ __forceinline__ uint32_t imad(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ uint32_t class16(const uint4 v) {
    return hibits_to_nibble(tokclass4(v.x)) | (hibits_to_nibble(tokclass4(v.y)) << 4) |
           (hibits_to_nibble(tokclass4(v.z)) << 8) | (hibits_to_nibble(tokclass4(v.w)) << 12);
}
// bits [lo, hi) of a 16-bit mask, lo/hi clamped to [0,16]
__device__ __forceinline__ uint32_t range16(int lo, int hi) {
    lo = max(lo, 0); hi = min(hi, 16);
    if (hi <= lo) return 0u;
    return ((1u << hi) - 1u) & ~((1u << lo) - 1u);
}

__device__ __forceinline__ void report_error(tk_result_t* res, uint32_t rec, int status) {
    atomicAdd(&res->n_errors, 1u);
    const uint32_t prev = atomicMin(&res->first_error_rec, rec);
    // status of the lowest-indexed bad record: last writer among decreasing minima wins;
    // the host re-reads rec_status[first_error_rec] when it needs the exact code.
    if (rec < prev) res->first_error_status = (uint32_t)status;
}

// 10^m for the digit-group accumulation  w = w * 10^m + group
static __constant__ uint32_t c_pow10[9] = {1u, 10u, 100u, 1000u, 10000u, 100000u, 1000000u, 10000000u, 100000000u};

// 8 bytes of text, first character in the lowest byte of .lo
struct Window8 { uint32_t lo, hi; };

// value of 4 ASCII-digit bytes already reduced to 0..9, first character in the lowest byte
__device__ __forceinline__ uint32_t parse4(uint32_t v) {
    const uint32_t d = v * 2561u;                              // byte 1 = 10*d0 + d1, byte 3 = 10*d2 + d3
    return imad(__byte_perm(d, 0, 0x4441), 100u, __byte_perm(d, 0, 0x4443));
}
// value of the first m (1..8) digits of the window
__device__ __forceinline__ uint32_t parse_upto8(const Window8 w, uint32_t m) {
    unsigned long long x = (((unsigned long long)(w.hi ^ 0x30303030u)) << 32) | (w.lo ^ 0x30303030u);
    x <<= 8u * (8u - m);                                        // digits to the top, zero digits below
    return parse4((uint32_t)x) * 10000u + parse4((uint32_t)(x >> 32));
}
__device__ __forceinline__ uint32_t byte_of(const Window8 w, uint32_t k) {       // k in 0..7
    const unsigned long long x = ((unsigned long long)w.hi << 32) | w.lo;
    return (uint32_t)(x >> (8u * k)) & 0xFFu;
}

