// tk_json_parse.cuh -- device helpers shared by the JSON array kernels (tk_json_kernels.cu: scan /
// parse of the three-kernel chain; tk_json_fused.cu: the single-pass kernel): comma classification,
// token flags, the word sources of the fast lane (tk_fastlane.cuh) and the row writer.
#pragma once
#include "tk_swar.cuh"
#include "tk_fastlane.cuh"

struct GlobalReader {
    const uint8_t* bytes;
    uint32_t end;
    __device__ __forceinline__ int operator()(uint32_t pos) const { return pos < end ? (int)bytes[pos] : -1; }
};

// word sources of the fast lane (tk_fastlane.cuh): the slot buffer through the read-only path, a
// shared-memory window with plain loads
struct LdgWords {
    const uint8_t* p;
    __device__ __forceinline__ uint32_t word(uint32_t a, int k) const { return __ldg(reinterpret_cast<const uint32_t*>(p + a) + k); }
};
struct SmemWords {
    const uint8_t* p;
    __device__ __forceinline__ uint32_t word(uint32_t a, int k) const { return reinterpret_cast<const uint32_t*>(p + a)[k]; }
};

// 0x80 in every byte lane of x that holds ','
__device__ __forceinline__ uint32_t comma4(uint32_t x) {
    const uint32_t y = x ^ 0x2C2C2C2Cu;
    return ~(((y & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | y) & 0x80808080u;
}

// flags in the record field of a token entry
#define TOK_LAST 0x80000000u     // last element of its record
#define TOK_TAIL 0x40000000u     // ... and the record's final byte is ']'
#define TOK_REC_MASK 0x3FFFFFFFu


__device__ __forceinline__ void flag_record(int32_t* rec_status, tk_result_t* res, uint32_t rec, int code) {
    if (atomicCAS(&rec_status[rec], 0, code) == 0) report_error(res, rec, code);   // count a record once
}

// 4 mask bytes (1 = real element) for elements e .. e+3 of a record with `len` elements
__device__ __forceinline__ uint32_t mask_word(uint32_t len, uint32_t e) {
    const int left = (int)len - (int)e;
    return left >= 4 ? 0x01010101u : (left <= 0 ? 0u : (0x01010101u >> (8 * (4 - left))));
}


// Pad, length and mask of one kept record's row (row = its index in the output planes, len = its
// element count), whole warp: store-only work, 16 bytes per lane where the planes allow it.
template <int ELEM_BYTES>
__device__ __forceinline__ void json_write_row(const tk_plane_t& values, const tk_plane_t& lengths, const tk_plane_t& mask,
                                               uint32_t max_len, uint64_t row, uint32_t len, uint32_t lane,
                                               bool vec_values, bool vec_mask) {
    constexpr uint32_t PER16 = 16 / ELEM_BYTES;                 // elements per 16-byte store
    const uint32_t fill_from = min(len, max_len);
    uint8_t* vrow = (uint8_t*)values.base + row * values.row_bytes;
    // pad [fill_from, max_len): scalar stores up to the next 16-byte boundary, then 16 bytes per lane
    uint32_t e0 = fill_from;
    if (vec_values) {
        const uint32_t up = min((fill_from + PER16 - 1) & ~(PER16 - 1), max_len);
        if (fill_from + lane < up) {
            if (ELEM_BYTES == 4) reinterpret_cast<uint32_t*>(vrow)[fill_from + lane] = 0u;
            else reinterpret_cast<unsigned long long*>(vrow)[fill_from + lane] = 0ull;
        }
        const uint32_t vend = max_len & ~(PER16 - 1);
        for (uint32_t e = up + lane * PER16; e < vend; e += 32 * PER16)
            *reinterpret_cast<uint4*>(vrow + (size_t)e * ELEM_BYTES) = make_uint4(0, 0, 0, 0);
        e0 = max(up, vend);
    }
    if (ELEM_BYTES == 4) {
        uint32_t* p = reinterpret_cast<uint32_t*>(vrow);
        for (uint32_t e = e0 + lane; e < max_len; e += 32) p[e] = 0u;
    } else {
        unsigned long long* p = reinterpret_cast<unsigned long long*>(vrow);
        for (uint32_t e = e0 + lane; e < max_len; e += 32) p[e] = 0ull;
    }
    if (lengths.base && lane == 0)
        *reinterpret_cast<int32_t*>((uint8_t*)lengths.base + row * lengths.row_bytes) = (int32_t)len;
    if (mask.base) {
        uint8_t* m = (uint8_t*)mask.base + row * mask.row_bytes;
        if (vec_mask) {
            for (uint32_t e = lane * 16; e < max_len; e += 32 * 16)
                *reinterpret_cast<uint4*>(m + e) = make_uint4(mask_word(len, e), mask_word(len, e + 4),
                                                              mask_word(len, e + 8), mask_word(len, e + 12));
        } else if ((max_len & 3u) == 0 && ((uintptr_t)m & 3u) == 0) {
            uint32_t* m32 = reinterpret_cast<uint32_t*>(m);
            for (uint32_t w = lane; w < max_len / 4; w += 32) m32[w] = mask_word(len, w * 4);
        } else {
            for (uint32_t e = lane; e < max_len; e += 32) m[e] = e < len ? 1 : 0;
        }
    }
}
