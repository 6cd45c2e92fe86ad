// tk_fastlane.cuh -- the fast lane of the JSON array kernels: one number literal, whose last byte
// is known, to its value with straight-line word arithmetic.
//
// Replaces, per element, what the reference does with
//     json.loads(record.value)            README.md:74       (text -> Python float / int)
//     torch.tensor(elements, dtype=...)   README.md:41-43    (-> the tensor's element type)
// for the literals json.dumps writes; everything else goes to the scalar grammar (tk_json.cuh).
//
// The kernels know where an element ends (`stop`: the next ',' or the record's final ']'), so the
// lane anchors the literal at its END: the 20 bytes in front of `stop` are loaded as five words,
// the bytes in front of the literal are masked off, and the last digit always sits in the top byte
// of the last word.  That makes every later step position-independent:
//     point     the one byte equal to '.', found with a zero-byte test on five words;
//     digits    the point is taken out by shifting the bytes in front of it up by one (a select
//               between each word and its funnel-shifted neighbour), which leaves at most 19
//               digits right-aligned in 20 bytes; one range test on the five words proves that
//               every byte is a digit (which also rejects a second point, exponents, signs,
//               white space: anything but  ['-'] digits ['.' digits] );
//     value     w = the digits as one integer (five 4-digit reductions), K = digits behind the
//               point; the result is w * 10^-K: Eisel-Lemire for float64, one multiplication plus
//               a distance test to the nearest float32 rounding boundary for float32.
// Host and device: the same source is compiled with g++ for the differential tests against
// CPython (tests/test_fastlane_host.py).
#pragma once
#include "tk_json.cuh"

#define TK_FL_WIN 20u
#if defined(__CUDA_ARCH__)
#define TK_FL_UNROLL _Pragma("unroll")
#else
#define TK_FL_UNROLL
#endif

struct alignas(16) tk_fl_row { uint32_t w[8]; };          // five words used; 32-byte rows
struct alignas(16) tk_fl_u128 { uint64_t x, y; };

struct tk_fl_tables {
    tk_fl_row  lead[TK_FL_WIN];        // [L]: 0xFF in the bytes L..19 of the window (the literal's)
    tk_fl_row  del[TK_FL_WIN + 1];     // [d]: byte selectors (tk_fl_prmt) that move the bytes 0..d up by one; [20]: identity
    tk_fl_u128 pow5[21];               // 128-bit normalised 5^-k, k = 0..20 (rows 342-k of the Eisel-Lemire table)
    double     inv10[21];              // 10^-k correctly rounded
};

// Filled by the threads of a block (device) or one caller (host: tid 0 of 1).
TK_HD void tk_fl_tables_fill(tk_fl_tables& tb, uint32_t tid, uint32_t nthreads) {
    for (uint32_t i = tid; i < (TK_FL_WIN + TK_FL_WIN + 1u) * 5u; i += nthreads) {
        const uint32_t row = i / 5u, k = i % 5u;
        uint32_t m = 0;
        for (uint32_t j = 0; j < 4u; ++j) {
            const uint32_t byte = 4u * k + j;
            if (row < TK_FL_WIN) {
                if (byte >= row) m |= 0xFFu << (8u * j);
            } else {
                // selector nibble of byte j: 4 + j = the word's own byte, 3 + j = the byte in front of it
                const uint32_t d = row - TK_FL_WIN;
                m |= ((d < TK_FL_WIN && byte <= d) ? 3u + j : 4u + j) << (4u * j);
            }
        }
        if (row < TK_FL_WIN) tb.lead[row].w[k] = m;
        else tb.del[row - TK_FL_WIN].w[k] = m;
    }
    const double inv[21] = {1e0,   1e-1,  1e-2,  1e-3,  1e-4,  1e-5,  1e-6,  1e-7,  1e-8,  1e-9,  1e-10,
                            1e-11, 1e-12, 1e-13, 1e-14, 1e-15, 1e-16, 1e-17, 1e-18, 1e-19, 1e-20};
    for (uint32_t i = tid; i < 21u; i += nthreads) {
        tb.pow5[i].x = TK_POW5_TABLE[342 - i][0];
        tb.pow5[i].y = TK_POW5_TABLE[342 - i][1];
        tb.inv10[i] = inv[i];
    }
}

// the five mask words of a table row
TK_HD void tk_fl_row_load(const tk_fl_row& row, uint32_t* m) {
#if defined(__CUDA_ARCH__)
    const uint4 v = *reinterpret_cast<const uint4*>(row.w);     // one 16-byte and one 4-byte read
    m[0] = v.x; m[1] = v.y; m[2] = v.z; m[3] = v.w;
    m[4] = row.w[4];
#else
    for (int k = 0; k < 5; ++k) m[k] = row.w[k];
#endif
}

// low word of (hi:lo) >> s, s in 0..31
TK_HD uint32_t tk_fl_funnel_r(uint32_t lo, uint32_t hi, uint32_t s) {
#if defined(__CUDA_ARCH__)
    return __funnelshift_r(lo, hi, s);
#else
    return s ? (lo >> s) | (hi << (32u - s)) : lo;
#endif
}

// a + b on the multiply-add pipe (the lane is bound by the logic / shift pipe)
TK_HD uint32_t tk_fl_add(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
    uint32_t r;
    asm("mad.lo.u32 %0, %1, 1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
#else
    return a + b;
#endif
}

// sum of the four byte products of a and b, plus c
TK_HD uint32_t tk_fl_dp4a(uint32_t a, uint32_t b, uint32_t c) {
#if defined(__CUDA_ARCH__)
    return __dp4a(a, b, c);
#else
    for (int j = 0; j < 4; ++j) c += ((a >> (8 * j)) & 0xFFu) * ((b >> (8 * j)) & 0xFFu);
    return c;
#endif
}

// byte j of the result = byte (nibble j of sel) of the 8 bytes lo (0..3), hi (4..7); nibbles 0..7 only
TK_HD uint32_t tk_fl_prmt(uint32_t lo, uint32_t hi, uint32_t sel) {
#if defined(__CUDA_ARCH__)
    return __byte_perm(lo, hi, sel);
#else
    const uint64_t v = ((uint64_t)hi << 32) | lo;
    uint32_t r = 0;
    for (int j = 0; j < 4; ++j) r |= (uint32_t)((v >> (8 * ((sel >> (4 * j)) & 7u))) & 0xFFu) << (8 * j);
    return r;
#endif
}

// Eisel-Lemire for  w * 10^-k,  w != 0,  0 <= k <= 20 (Lemire, "Number Parsing at a Gigabyte per
// Second"; the fast_float library): the arithmetic of tk_eisel_lemire (tk_number.cuh) with the arms
// that cannot be reached here removed -- 1 <= w < 2^64 and 10^-20 <= 10^-k <= 1 keep the result a
// normal double, far from overflow and underflow.
TK_HD uint64_t tk_fl_eisel_lemire(uint64_t w, uint32_t k, const tk_fl_u128* pow5) {
    const int lz = tk_clz64(w);
    w <<= lz;
    const tk_fl_u128 p = pow5[k];
    tk_u128 first = tk_mul64(w, p.x);
    uint64_t hi = first.hi, lo = first.lo;
    if ((hi & 0x1FFu) == 0x1FFu) {                              // 55 bits not settled: second product
        const uint64_t second = tk_mul64(w, p.y).hi;
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

// Accepts exactly
//     [' '] ['-'] digits ['.' digits]      at most 19 digits (8 without a point; 20 behind "0.0"),
//                                          no leading zero,
// in the bytes [pos, stop): what json.dumps writes for every int below 10^8 and every float
// without an exponent.  Everything else returns false: exponents go to tk_fl_scan_long (below),
// other white space, long literals, specials and garbage to the scalar grammar.
// `src.word(a, k)` returns the 32-bit word at the byte address a + 4k (a 4-byte aligned); words up to
// 7 bytes past `pos` and 3 bytes past `stop` are read (the buffers are padded: tk_internal.h).
// Out: w = the digits as one integer, K = how many of them follow the point, dot, neg.
// WIDE = false leaves the bodies of 21 and 22 bytes ("0.00...") to tk_fl_scan_long: the leanest
// form, for data that has none of them (the kernels choose: tk_json_kernels.cu).
template <bool WIDE = true, class Src>
TK_HD bool tk_fl_scan(const Src& src, uint32_t pos, uint32_t stop, const tk_fl_tables& tb, uint64_t* w_out,
                      uint32_t* K_out, uint32_t* dot_out, uint32_t* neg_out) {
    const uint32_t n = stop - pos;                                   // the element's bytes
    if (n - 1u > (WIDE ? 23u : 21u) || stop < TK_FL_WIN) return false;
    // ---- head: optional space (json.dumps writes ", "), optional sign, the first two characters
    const uint32_t ha = pos & ~3u;
    const uint32_t hw = tk_fl_funnel_r(src.word(ha, 0), src.word(ha, 1), (pos & 3u) * 8u);
    const uint32_t sp = (hw & 0xFFu) == 0x20u ? 1u : 0u;
    const uint32_t h1 = hw >> (8u * sp);
    const uint32_t neg = (h1 & 0xFFu) == 0x2Du ? 1u : 0u;
    const uint32_t h2 = h1 >> (8u * neg);
    const uint32_t m = n - sp - neg;                                 // digits and the point
    if (m - 1u > (WIDE ? TK_FL_WIN + 1u : TK_FL_WIN - 1u)) return false;
    // a leading zero is the whole integer part
    if ((h2 & 0xFFu) == 0x30u && m > 1u && ((h2 >> 8) & 0xFFu) != 0x2Eu) return false;
    // A body of 21 or 22 bytes must start with "0." (json.dumps writes up to three zeros behind it
    // before it switches to exponents: magnitudes below 0.01); `extra` of these bytes lie in front
    // of the 20-byte window.
    const uint32_t extra = WIDE && m > TK_FL_WIN ? m - TK_FL_WIN : 0u;
    if (extra && (h2 & 0xFFFFu) != 0x2E30u) return false;
    const uint32_t lead = extra ? 0u : TK_FL_WIN - m;                // bytes of the window in front of the body
    // ---- the 20 bytes in front of stop; u = digit values, the bytes in front of the literal cleared
    const uint32_t base = stop - TK_FL_WIN, ba = base & ~3u, bs = (base & 3u) * 8u;
    uint32_t r[6];
TK_FL_UNROLL
    for (int k = 0; k < 6; ++k) r[k] = src.word(ba, k);
    uint32_t lm[5];
    tk_fl_row_load(tb.lead[lead], lm);
    uint32_t u[5];
TK_FL_UNROLL
    for (int k = 0; k < 5; ++k) u[k] = (tk_fl_funnel_r(r[k], r[k + 1], bs) ^ 0x30303030u) & lm[k];
    // ---- the point and the digit test, word by word:
    //   z  0x80 in every byte equal to '.' (zero-byte test of u ^ 0x1E; a byte 0x1F above a point is
    //      flagged too -- then two bytes are counted and the literal leaves the lane)
    //   t  0x80 in every byte that is not 0..9
    // every non-digit must be the point (t == z), and the points are counted and located with two
    // byte dot products per word (weights 1 and the byte's window index): 128 per point.
    uint32_t cnt = 0, psum = 0, bad = 0;
TK_FL_UNROLL
    for (int k = 0; k < 5; ++k) {
        const uint32_t y = u[k] ^ 0x1E1E1E1Eu;
        const uint32_t z = tk_fl_add(y, 0xFEFEFEFFu) & ~y & 0x80808080u;
        const uint32_t t = (tk_fl_add(u[k], 0x76767676u) | u[k]) & 0x80808080u;
        bad |= t ^ z;
        cnt = tk_fl_dp4a(z, 0x01010101u, cnt);
        psum = tk_fl_dp4a(z, 0x03020100u + 0x04040404u * (uint32_t)k, psum);
    }
    if (bad | (cnt & ~128u)) return false;                          // a non-digit that is no point; two points
    const uint32_t g = cnt;                                          // != 0: there is a point in the window
    uint32_t d = TK_FL_WIN;                                          // its window index
    if (g) {
        d = psum >> 7;
        // no digit in front of the point (extra == 1: the '0' in front of the window is) / behind it;
        // extra == 2: the point lies in front of the window, this would be a second one
        if ((d == lead && !extra) || d == TK_FL_WIN - 1u || extra == 2u) return false;
    } else if (m > 8u && extra != 2u) {
        return false;                                                // long integers: scalar grammar
    }
    // ---- take the point out: bytes 0..d move up by one (a byte permutation of each word and the
    // one in front of it, selector from the table)
    uint32_t sel[5];
    tk_fl_row_load(tb.del[d], sel);
    uint32_t D[5];
TK_FL_UNROLL
    for (int k = 0; k < 5; ++k) D[k] = tk_fl_prmt(k ? u[k - 1] : 0u, u[k], sel[k]);
    if (WIDE && (D[0] & 0xFFu)) return false;                        // a 20th digit (only behind "0."): scalar grammar
    // ---- 19 digits in the bytes 1..19: pairs of digits by byte dot products (weights 10, 1)
    const uint32_t WH = 0x0000010Au, WL = 0x010A0000u;
    const uint32_t C = tk_fl_dp4a(D[0], WH, 0) * 100u + tk_fl_dp4a(D[0], WL, 0);
    uint32_t B = tk_fl_dp4a(D[2], WL, 0);
    B = tk_fl_dp4a(D[2], WH, 0) * 100u + B;
    B = tk_fl_dp4a(D[1], WL, 0) * 10000u + B;
    B = tk_fl_dp4a(D[1], WH, 0) * 1000000u + B;
    uint32_t A = tk_fl_dp4a(D[4], WL, 0);
    A = tk_fl_dp4a(D[4], WH, 0) * 100u + A;
    A = tk_fl_dp4a(D[3], WL, 0) * 10000u + A;
    A = tk_fl_dp4a(D[3], WH, 0) * 1000000u + A;
    const uint64_t w = ((uint64_t)C * 100000000u + B) * 100000000ull + A;
    *w_out = w;
    *K_out = g ? TK_FL_WIN - 1u - d : extra == 2u ? TK_FL_WIN : 0u;
    *dot_out = (g | extra) ? 1u : 0u;
    *neg_out = neg;
    return true;
}

// Value of a literal the lane accepted: the same bits as json.loads + torch.tensor give.
// False: float literal with an integer target (scalar grammar).
template <int KIND>
TK_HD bool tk_fl_value(uint64_t w, uint32_t K, uint32_t dot, uint32_t neg, const tk_fl_tables& tb,
                       typename tk_elem<KIND>::type* out) {
    typedef typename tk_elem<KIND>::type T;
    // torch.tensor(json.loads(..)): an int literal is a Python int ("-0" is +0.0), a float literal
    // keeps its sign ("-0.0" is -0.0)
    const bool minus = neg && (dot || w);
    if (KIND == TK_F32) {
        // The reference rounds twice: text -> double (json.loads), double -> float (torch.tensor).
        // d = double(w) * 10^-K is within 3 ulp (double) of the exact value x: the conversion of w,
        // the rounded table entry and the product contribute a relative 2^-53 each.  The correctly
        // rounded double y is within half an ulp of x, so unless a float rounding boundary (low 29
        // mantissa bits = 0x10000000) lies within 8 ulp of d, float(d) == float(y) bit for bit.
        // Otherwise (2^-25 of the inputs): exact conversion.
        double d = (double)w * tb.inv10[K];
#if defined(__CUDA_ARCH__)
        const uint32_t low = (uint32_t)__double_as_longlong(d) & 0x1FFFFFFFu;
#else
        uint64_t dbits;
        memcpy(&dbits, &d, 8);
        const uint32_t low = (uint32_t)dbits & 0x1FFFFFFFu;
#endif
        if (low - (0x10000000u - 8u) <= 16u) {
            const uint64_t bits = w ? tk_fl_eisel_lemire(w, K, tb.pow5) : 0ull;
#if defined(__CUDA_ARCH__)
            d = __longlong_as_double((long long)bits);
#else
            memcpy(&d, &bits, 8);
#endif
        }
        const float f = (float)d;
        *out = (T)(minus ? -f : f);
    } else if (KIND == TK_F64) {
        uint64_t bits = w ? tk_fl_eisel_lemire(w, K, tb.pow5) : 0ull;
        if (minus) bits |= 0x8000000000000000ull;
        double d;
#if defined(__CUDA_ARCH__)
        d = __longlong_as_double((long long)bits);
#else
        memcpy(&d, &bits, 8);
#endif
        *out = (T)d;
    } else {
        if (dot) return false;                                   // float literal, integer target
        *out = (T)(neg ? -(int64_t)w : (int64_t)w);              // < 10^8: fits every integer target
    }
    return true;
}

// ---------------------------------------------------------------- the long forms
// What json.dumps writes and the inline lane above does not take:
//     a body of 21 or 22 bytes that starts with "0."   (magnitudes below 0.01: 0.00123...)
//     an exponent at the end                            [('e' | 'E') ['+' | '-'] d{1,3}]: e-05, e+16
// Same window arithmetic, but without the shared-memory tables (masks and selectors are computed)
// and with the general Eisel-Lemire of tk_number.cuh, so that the kernels can call it from their
// slow branches with nothing but the bytes: the element loop keeps its registers and its call
// signature.  On data of small or large magnitude this is what keeps the elements away from the
// byte-by-byte grammar (gauss(0, 1e-3): 2.8 x on the chain).
struct tk_fl_num {
    uint64_t w;          // the digits as one integer
    uint32_t K;          // how many of them follow the point
    int32_t  e10;        // the exponent
    uint32_t ok, dot, neg;
};

template <class Src>
TK_HD uint32_t tk_fl_byte(const Src& src, uint32_t a) {
    return (src.word(a & ~3u, 0) >> ((a & 3u) * 8u)) & 0xFFu;
}

// digits and point of the body [stop - m, stop), 1 <= m <= 22; first2 = its first two characters
template <class Src>
TK_HD bool tk_fl_core_long(const Src& src, uint32_t stop, uint32_t m, uint32_t first2, uint64_t* w_out,
                           uint32_t* K_out, uint32_t* dot_out) {
    if (m - 1u > TK_FL_WIN + 1u || stop < TK_FL_WIN) return false;
    if ((first2 & 0xFFu) == 0x30u && m > 1u && ((first2 >> 8) & 0xFFu) != 0x2Eu) return false;   // leading zero
    uint32_t extra = 0;                                              // bytes of the body in front of the window
    if (m > TK_FL_WIN) {
        extra = m - TK_FL_WIN;
        if ((first2 & 0xFFFFu) != 0x2E30u) return false;             // "0."
    }
    const uint32_t lead = extra ? 0u : TK_FL_WIN - m;
    const uint32_t base = stop - TK_FL_WIN, ba = base & ~3u, bs = (base & 3u) * 8u;
    uint32_t r[6], u[5];
    for (int k = 0; k < 6; ++k) r[k] = src.word(ba, k);
    uint32_t cnt = 0, psum = 0, bad = 0;
    for (uint32_t k = 0; k < 5u; ++k) {
        // bytes lead.. of the window belong to the body
        const uint32_t keep = lead <= 4u * k ? 0xFFFFFFFFu : lead >= 4u * k + 4u ? 0u : 0xFFFFFFFFu << (8u * (lead - 4u * k));
        u[k] = (tk_fl_funnel_r(r[k], r[k + 1], bs) ^ 0x30303030u) & keep;
        const uint32_t y = u[k] ^ 0x1E1E1E1Eu;
        const uint32_t z = (y - 0x01010101u) & ~y & 0x80808080u;
        const uint32_t t = ((u[k] + 0x76767676u) | u[k]) & 0x80808080u;
        bad |= t ^ z;
        cnt = tk_fl_dp4a(z, 0x01010101u, cnt);
        psum = tk_fl_dp4a(z, 0x03020100u + 0x04040404u * k, psum);
    }
    if (bad | (cnt & ~128u)) return false;
    uint32_t d = TK_FL_WIN, K, dot = cnt ? 1u : 0u;
    if (extra == 2u) {                                               // "0." in front of the window: 20 fraction digits
        if (cnt) return false;
        dot = 1u;
        K = TK_FL_WIN;
    } else if (cnt) {
        d = psum >> 7;
        // (extra == 1: the '0' in front of the window is the digit in front of the point)
        if (extra ? d != 0u : d == lead) return false;
        if (d == TK_FL_WIN - 1u) return false;
        K = TK_FL_WIN - 1u - d;
    } else {
        if (extra | (m > 8u)) return false;
        K = 0u;
    }
    uint32_t D[5];
    for (uint32_t k = 0; k < 5u; ++k) {
        // bytes 0..d move up by one: nibble j of the selector is 3 + j for them, 4 + j for the others
        const uint32_t moved = d >= TK_FL_WIN || d < 4u * k ? 0u : (d - 4u * k + 1u < 4u ? d - 4u * k + 1u : 4u);
        const uint32_t sel = 0x7654u - (0x1111u & ((1u << (4u * moved)) - 1u));
        D[k] = tk_fl_prmt(k ? u[k - 1] : 0u, u[k], sel);
    }
    if (D[0] & 0xFFu) return false;                                  // a 20th digit
    const uint32_t WH = 0x0000010Au, WL = 0x010A0000u;
    const uint32_t C = tk_fl_dp4a(D[0], WH, 0) * 100u + tk_fl_dp4a(D[0], WL, 0);
    const uint32_t B = (tk_fl_dp4a(D[1], WH, 0) * 100u + tk_fl_dp4a(D[1], WL, 0)) * 10000u +
                       tk_fl_dp4a(D[2], WH, 0) * 100u + tk_fl_dp4a(D[2], WL, 0);
    const uint32_t A = (tk_fl_dp4a(D[3], WH, 0) * 100u + tk_fl_dp4a(D[3], WL, 0)) * 10000u +
                       tk_fl_dp4a(D[4], WH, 0) * 100u + tk_fl_dp4a(D[4], WL, 0);
    *w_out = ((uint64_t)C * 100000000u + B) * 100000000ull + A;
    *K_out = K;
    *dot_out = dot;
    return true;
}

// The element [pos, stop):  [' '] ['-'] body [exponent]
template <class Src>
TK_HD tk_fl_num tk_fl_scan_long(const Src& src, uint32_t pos, uint32_t stop) {
    tk_fl_num r;
    r.w = 0; r.K = 0; r.e10 = 0; r.ok = 0; r.dot = 0; r.neg = 0;
    if (stop - pos - 1u > 29u) return r;
    const uint32_t ha = pos & ~3u;
    const uint32_t hw = tk_fl_funnel_r(src.word(ha, 0), src.word(ha, 1), (pos & 3u) * 8u);
    const uint32_t sp = (hw & 0xFFu) == 0x20u ? 1u : 0u;
    const uint32_t h1 = hw >> (8u * sp);
    r.neg = (h1 & 0xFFu) == 0x2Du ? 1u : 0u;
    const uint32_t first2 = h1 >> (8u * r.neg);
    const uint32_t first = pos + sp + r.neg, m = stop - first;
    if (tk_fl_core_long(src, stop, m, first2, &r.w, &r.K, &r.dot)) { r.ok = 1u; return r; }
    // an exponent at the end?
    uint32_t q = stop, e = 0, scale = 1, nd = 0;
    while (nd < 3u && q > first + 1u) {
        const uint32_t c = tk_fl_byte(src, q - 1u) - 0x30u;
        if (c > 9u) break;
        e += c * scale;
        scale *= 10u;
        --q;
        ++nd;
    }
    if (nd == 0u || q <= first + 1u) return r;
    uint32_t c = tk_fl_byte(src, q - 1u);
    const bool eneg = c == 0x2Du;
    if (eneg || c == 0x2Bu) {
        --q;
        if (q <= first + 1u) return r;
        c = tk_fl_byte(src, q - 1u);
    }
    if ((c | 0x20u) != 0x65u) return r;
    --q;                                                             // q: the 'e', i.e. the end of the mantissa
    if (!tk_fl_core_long(src, q, q - first, first2, &r.w, &r.K, &r.dot)) return r;
    r.dot = 1u;                                                      // a float literal
    r.e10 = eneg ? -(int32_t)e : (int32_t)e;
    r.ok = 1u;
    return r;
}

// Value of a literal tk_fl_scan_long accepted.  False: float literal with an integer target.
template <int KIND>
TK_HD bool tk_fl_value_long(const tk_fl_num& r, typename tk_elem<KIND>::type* out) {
    typedef typename tk_elem<KIND>::type T;
    const bool minus = r.neg && (r.dot || r.w);
    if (KIND == TK_F32 || KIND == TK_F64) {
        const int32_t q = r.e10 - (int32_t)r.K;
        uint64_t bits = 0;
        if (r.w == 0 || q < -342) bits = 0;                          // (below 10^-342: rounds to zero, tk_number.cuh)
        else if (q > 308) bits = 0x7FF0000000000000ull;              // float('1e400') is inf
        else bits = tk_eisel_lemire(r.w, q);
        double d;
#if defined(__CUDA_ARCH__)
        d = __longlong_as_double((long long)bits);
#else
        memcpy(&d, &bits, 8);
#endif
        if (KIND == TK_F32) {
            const float f = (float)d;                                // double-rounded, as torch.tensor does
            *out = (T)(minus ? -f : f);
        } else {
            *out = (T)(minus ? -d : d);
        }
    } else {
        if (r.dot) return false;
        *out = (T)(r.neg ? -(int64_t)r.w : (int64_t)r.w);
    }
    return true;
}
