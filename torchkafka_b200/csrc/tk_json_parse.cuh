// tk_json_parse.cuh -- device helpers shared by the JSON array kernels (tk_json_kernels.cu: scan /
// parse of the three-kernel chain; tk_json_fused.cu: the single-pass kernel): comma classification,
// token flags, the parse kernel's shared-memory tables, the Eisel-Lemire specialisation for
// w * 10^-k and the fast lane that covers what json.dumps writes.
#pragma once
#include "tk_swar.cuh"

struct GlobalReader {
    const uint8_t* bytes;
    uint32_t end;
    __device__ __forceinline__ int operator()(uint32_t pos) const { return pos < end ? (int)bytes[pos] : -1; }
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


// ---- per-block lookup tables of the parse kernel.  Shared memory, not __constant__: the index
// differs from lane to lane (number of fraction digits), which a constant-bank read serialises.
struct parse_tables {
    ulonglong2 pow5[20];     // 128-bit normalised 5^-k (k = 0..19): rows 342-k of the Eisel-Lemire table
    uint64_t   pow10w[17];   // 10^k, k = 0..16
    double     inv10[17];    // 10^-k correctly rounded, k = 0..16
    uint32_t   pow10[9];     // 10^m for the digit-group accumulation  F = g1 * 10^m + g2
};
static __device__ const double g_inv10[17] = {1e0, 1e-1, 1e-2, 1e-3, 1e-4, 1e-5, 1e-6, 1e-7, 1e-8,
                                              1e-9, 1e-10, 1e-11, 1e-12, 1e-13, 1e-14, 1e-15, 1e-16};
static __device__ const uint64_t g_pow10w[17] = {
    1ull, 10ull, 100ull, 1000ull, 10000ull, 100000ull, 1000000ull, 10000000ull, 100000000ull,
    1000000000ull, 10000000000ull, 100000000000ull, 1000000000000ull, 10000000000000ull,
    100000000000000ull, 1000000000000000ull, 10000000000000000ull};

__device__ __forceinline__ void parse_tables_init(parse_tables& tb) {
    const uint32_t i = threadIdx.x;
    if (i < 20) tb.pow5[i] = make_ulonglong2(TK_POW5_TABLE[342 - i][0], TK_POW5_TABLE[342 - i][1]);
    else if (i >= 32 && i < 49) { tb.pow10w[i - 32] = g_pow10w[i - 32]; tb.inv10[i - 32] = g_inv10[i - 32]; }
    else if (i >= 64 && i < 73) tb.pow10[i - 64] = c_pow10[i - 64];
    __syncthreads();
}

// Eisel-Lemire for  w * 10^-k,  w != 0,  0 <= k <= 19.  Same arithmetic as tk_eisel_lemire
// (tk_number.cuh) with the arms that cannot be reached here removed: 1 <= w < 2^64 and
// 10^-19 <= 10^-k <= 1 keep the result a normal double, far from overflow and underflow.
__device__ __forceinline__ uint64_t el_small_neg_exp(uint64_t w, uint32_t k, const ulonglong2* pow5) {
    const int lz = __clzll((long long)w);
    w <<= lz;
    const ulonglong2 p = pow5[k];
    uint64_t hi = __umul64hi(w, p.x), lo = w * p.x;
    if ((hi & 0x1FFu) == 0x1FFu) {                              // 55 bits not settled: second product
        const uint64_t second = __umul64hi(w, p.y);
        lo += second;
        if (second > lo) ++hi;
    }
    const uint32_t upperbit = (uint32_t)(hi >> 63);
    const uint32_t shift = upperbit + 9u;                      // 64 - 52 - 3
    uint64_t mant = hi >> shift;
    // exactly halfway between two doubles: only when 5^k fits 64 bits (k <= 4 on this side)
    if (k <= 4u && lo <= 1u && (mant & 3u) == 1u && (mant << shift) == hi) mant &= ~1ull;
    mant = (mant + (mant & 1u)) >> 1;                           // 2^52 <= mant <= 2^53
    // floor(log2(10^-k)) + 63 + 1023, then the implicit bit of mant carries into the exponent
    // (and a mantissa that rounded up to 2^53 carries once more)
    const int32_t power2 = ((-(int32_t)(217706u * k)) >> 16) + 63 + 1023 + (int32_t)upperbit - lz;
    return ((uint64_t)(uint32_t)(power2 - 1) << 52) + mant;
}

// Fast lane of the parse kernel.  Accepts exactly
//     [' '] ['-'] d{1,8} ['.' d{1,16}]        at most 19 digits, no leading zero,
// ending at byte `stop` (the ',' the scan found, or the record's final ']'), which is what
// json.dumps writes for every int below 10^8 and every float without exponent.  Everything
// else (exponents, other white space, long literals, specials, garbage) returns false and is
// handled -- or rejected -- by the scalar grammar.  Out: I = the integer digits, F = the L2
// digits after the point (as integers), neg.
// SMEM: `bytes` is a shared-memory window (plain loads) instead of the slot buffer (read-only path).
template <bool SMEM = false>
__device__ __forceinline__ bool fast_number(const uint8_t* bytes, uint32_t pos, uint32_t stop,
                                            const uint32_t* pow10, uint32_t* I_out, uint64_t* F_out,
                                            uint32_t* L2_out, uint32_t* neg_out) {
    // 28 bytes from the aligned word the token starts in (the slot buffer is padded: tk_internal.h)
    const uint32_t* w32 = reinterpret_cast<const uint32_t*>(bytes + (pos & ~3u));
    uint32_t r[7];
#pragma unroll
    for (int k = 0; k < 7; ++k) r[k] = SMEM ? w32[k] : __ldg(w32 + k);
    uint32_t sh = (pos & 3u) * 8u;
    const uint32_t sp = ((r[0] >> sh) & 0xFFu) == ' ' ? 1u : 0u;      // json.dumps: ", "
    sh += 8u * sp;                                                    // 0..32: clamped funnel shift
    pos += sp;
    uint32_t W[6];                                                    // 24 bytes of token text
#pragma unroll
    for (int k = 0; k < 6; ++k) W[k] = __funnelshift_rc(r[k], r[k + 1], sh);
    uint32_t nd = 0xFF000000u;                                        // bit i: byte i is not a digit
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        const uint32_t u = W[k] ^ 0x30303030u;
        const uint32_t t = (imad(u & 0x7F7F7F7Fu, 1u, 0x76767676u) | u) & 0x80808080u;
        nd = imad(hibits_to_nibble(t), 1u << (4 * k), nd);
    }
    const uint32_t neg = (W[0] & 0xFFu) == '-' ? 1u : 0u;
    const uint32_t L1 = __ffs(nd >> neg) - 1;                         // integer digits
    if (L1 - 1u > 7u) return false;
    const Window8 wi = {__funnelshift_r(W[0], W[1], 8 * neg), __funnelshift_r(W[1], W[2], 8 * neg)};
    if (L1 > 1 && (wi.lo & 0xFFu) == '0') return false;
    const uint32_t tpos = neg + L1;                                   // 1..9: byte after the integer part
    const uint32_t tc = (tpos < 8 ? __byte_perm(W[0], W[1], tpos) : __byte_perm(W[2], 0u, tpos - 8u)) & 0xFFu;
    const uint32_t I = parse_upto8(wi, L1);
    uint64_t F = 0;
    uint32_t L2 = 0, endp = tpos;
    if (tc == '.') {
        const uint32_t q0 = tpos + 1;
        L2 = __ffs(nd >> q0) - 1;                                     // fraction digits
        if (L2 - 1u > 15u || L1 + L2 > 19u) return false;
        endp = q0 + L2;                                               // <= 21: inside the 24 bytes
        // the fraction as two 8-byte windows, re-read at its own alignment (L1 hits)
        const uint32_t fpos = pos + q0;
        const uint32_t* f32 = reinterpret_cast<const uint32_t*>(bytes + (fpos & ~3u));
        uint32_t f[5];
#pragma unroll
        for (int k = 0; k < 5; ++k) f[k] = SMEM ? f32[k] : __ldg(f32 + k);
        const uint32_t fsh = (fpos & 3u) * 8u;
        const Window8 x0 = {__funnelshift_r(f[0], f[1], fsh), __funnelshift_r(f[1], f[2], fsh)};
        const uint32_t m0 = min(L2, 8u), m1 = L2 - m0;
        const uint32_t g1 = parse_upto8(x0, m0);
        F = g1;
        if (m1) {
            const Window8 x1 = {__funnelshift_r(f[2], f[3], fsh), __funnelshift_r(f[3], f[4], fsh)};
            F = (uint64_t)g1 * pow10[m1] + parse_upto8(x1, m1);
        }
    }
    if (pos + endp != stop) return false;
    *I_out = I; *F_out = F; *L2_out = L2; *neg_out = neg;
    return true;
}

// Value of a literal the fast lane accepted (I integer digits, F the L2 fraction digits): the same
// bits as json.loads + torch.tensor give.  False: float literal with an integer target (scalar path).
template <int KIND>
__device__ __forceinline__ bool fast_convert(uint32_t I, uint64_t F, uint32_t L2, uint32_t neg, const parse_tables& tb,
                                             typename tk_elem<KIND>::type* out) {
    typedef typename tk_elem<KIND>::type T;
    T v;
    // torch.tensor(json.loads(..)): an int literal is a Python int ("-0" is +0.0), a float
    // literal keeps its sign ("-0.0" is -0.0)
    const bool minus = neg && (I | L2 | (uint32_t)F | (uint32_t)(F >> 32));
    if (KIND == TK_F32) {
        // The reference rounds twice: text -> double (json.loads), double -> float
        // (torch.tensor).  d = I + F * 10^-L2 in one fused multiply-add is within 2.5 ulp
        // (double) of the exact value x: F -> double and the rounded 10^-L2 contribute a
        // relative 2^-52 of the fraction (< 1), the final rounding half an ulp.  The
        // correctly rounded double y is within 3.5 ulp of d, so unless a float rounding
        // boundary (low 29 mantissa bits = 0x10000000) lies within 8 ulp of d, float(d)
        // == float(y) bit for bit.  Otherwise (2^-25 of the inputs): exact conversion.
        double d = (double)I;
        if (L2) d = fma((double)F, tb.inv10[L2], d);
        const uint32_t low = (uint32_t)__double_as_longlong(d) & 0x1FFFFFFFu;
        if (low - (0x10000000u - 8u) <= 16u) {
            const uint64_t w = (uint64_t)I * tb.pow10w[L2] + F;
            d = __longlong_as_double((long long)(w ? el_small_neg_exp(w, L2, tb.pow5) : 0ull));
        }
        const float f = (float)d;
        v = (T)(minus ? -f : f);
    } else if (KIND == TK_F64) {
        const uint64_t w = (uint64_t)I * tb.pow10w[L2] + F;
        uint64_t bits = w ? el_small_neg_exp(w, L2, tb.pow5) : 0ull;
        if (minus) bits |= 0x8000000000000000ull;
        v = (T)__longlong_as_double((long long)bits);
    } else if (L2 == 0) {
        v = (T)(neg ? -(int32_t)I : (int32_t)I);         // < 10^8: fits every integer target
    } else {
        return false;                                    // float literal, integer target
    }
    *out = v;
    return true;
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
