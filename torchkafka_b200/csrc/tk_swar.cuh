// tk_swar.cuh -- SWAR (SIMD-within-a-register) helpers shared by the JSON kernels: flag gathering,
// bit ranges, the error report.  Device-only; plain 32-bit integer arithmetic (the __v*4 video
// intrinsics are emulated with ~10 instructions each on sm_100).  The number fast lane has its own
// host+device helpers (tk_fastlane.cuh).
#pragma once
#include "tk_internal.h"
#include "tk_json.cuh"

// ---------------------------------------------------------------- SWAR helpers
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
