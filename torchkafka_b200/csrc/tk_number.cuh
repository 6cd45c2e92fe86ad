// tk_number.cuh -- JSON number token -> IEEE-754 double / int64, bit-exact with
// CPython's json C scanner + PyOS_string_to_double (what the reference's
// `_process` bodies call, README.md:54,74; behaviour table in SURVEY.md §4.1-11).
//
// Everything here is plain integer code usable from device and host
// (`TK_HD`), so the same functions are differential-tested against CPython on a
// CPU-only machine (tests/test_number_host.py) and run inside the sm_100a
// decode kernels.  Decimal -> binary uses the Eisel-Lemire algorithm with the
// 128-bit power-of-five table in tk_pow5_table.inc; inputs with more than 19
// significant digits take the "w and w+1 agree" shortcut and otherwise fall to
// an exact big-integer comparison (tk_bigcmp below).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define TK_HD __host__ __device__ __forceinline__
#define TK_HD_NOINLINE __host__ __device__ __noinline__
#else
#define TK_HD inline
#define TK_HD_NOINLINE inline
#endif

// ---- per-record status codes (also in include/torchkafka_b200.h) -----------
enum {
    TK_OK = 0,
    TK_ERR_SYNTAX = 1,       // not a JSON array of numbers (json.JSONDecodeError in the reference)
    TK_ERR_OVERFLOW = 2,     // value does not fit the target dtype (RuntimeError/OverflowError/ValueError)
    TK_ERR_TOO_LONG = 3,     // more than max_len elements
    TK_ERR_UNSUPPORTED = 4,  // valid JSON the schema does not cover (true/false/null, strings, nesting)
    TK_ERR_SIZE = 5,         // raw record has the wrong byte count
    TK_ERR_RAGGED = 6,       // fixed_len schema and a kept record of another length (torch.stack raises)
};

#if defined(__CUDACC__)
static __device__ const uint64_t tk_pow5_dev[651][2] = {
#include "tk_pow5_table.inc"
};
#endif
#if defined(__CUDA_ARCH__)
#define TK_POW5_TABLE tk_pow5_dev
#else
#define TK_POW5_TABLE tk_pow5_host
static const uint64_t tk_pow5_host[651][2] = {
#include "tk_pow5_table.inc"
};
#endif

struct tk_u128 { uint64_t hi, lo; };

TK_HD tk_u128 tk_mul64(uint64_t a, uint64_t b) {
    tk_u128 r;
#if defined(__CUDA_ARCH__)
    r.lo = a * b;
    r.hi = __umul64hi(a, b);
#else
    unsigned __int128 p = (unsigned __int128)a * b;
    r.lo = (uint64_t)p;
    r.hi = (uint64_t)(p >> 64);
#endif
    return r;
}

TK_HD int tk_clz64(uint64_t x) {
#if defined(__CUDA_ARCH__)
    return __clzll((long long)x);
#else
    return __builtin_clzll(x);
#endif
}

// Binary result of the decimal conversion: `bits` is the IEEE-754 double bit
// pattern of |value| (sign applied by the caller).  ok == false means the
// 128-bit approximation could not decide (only possible for truncated inputs,
// handled by the caller).
TK_HD uint64_t tk_eisel_lemire(uint64_t w, int32_t q) {
    // value = w * 10^q, w != 0, q in [-342, 308]
    int lz = tk_clz64(w);
    w <<= lz;
    const uint64_t p_hi = TK_POW5_TABLE[q + 342][0];
    const uint64_t p_lo = TK_POW5_TABLE[q + 342][1];
    tk_u128 first = tk_mul64(w, p_hi);
    // 52 explicit bits + 3 => need 55 bits of precision; refine when the low 9
    // bits of the high word are all ones.
    if ((first.hi & 0x1FF) == 0x1FF) {
        tk_u128 second = tk_mul64(w, p_lo);
        first.lo += second.hi;
        if (second.hi > first.lo) first.hi++;
    }
    const int upperbit = (int)(first.hi >> 63);
    const int shift = upperbit + 64 - 52 - 3;
    uint64_t mant = first.hi >> shift;
    // floor(log2(10^q)) + 63, via the 152170/65536 approximation of log2(10)-... (valid on [-342,308])
    int32_t power2 = (int32_t)((((int64_t)(152170 + 65536) * q) >> 16) + 63) + upperbit - lz + 1023;
    if (power2 <= 0) {                       // subnormal (or zero)
        if (-power2 + 1 >= 64) return 0;
        mant >>= (-power2 + 1);
        mant += (mant & 1);
        mant >>= 1;
        // a rounded-up subnormal can become the smallest normal
        const uint64_t e = (mant < (1ull << 52)) ? 0ull : 1ull;
        return (e << 52) | (mant & ((1ull << 52) - 1));
    }
    // exactly-halfway inputs need round-to-even; only possible when 5^q fits 64 bits
    if (first.lo <= 1 && q >= -4 && q <= 23 && (mant & 3) == 1) {
        if ((mant << shift) == first.hi) mant &= ~1ull;
    }
    mant += (mant & 1);
    mant >>= 1;
    if (mant >= (2ull << 52)) { mant = (1ull << 52); power2++; }
    mant &= ~(1ull << 52);
    if (power2 >= 0x7FF) return 0x7FF0000000000000ull;
    return ((uint64_t)power2 << 52) | mant;
}

// ---------------------------------------------------------------------------
// Exact fallback for > 19 significant digits when w and w+1 round differently.
// Decides whether   D * 10^e10   (D = the full digit string)   is below, equal
// to or above the midpoint between double `lo_bits` and its successor, using
// multi-word integers.  Rare by construction (needs a 20+ digit literal within
// 1e-19 relative distance of a rounding boundary); never on the json.dumps
// output of a double, which has at most 17 digits.
// ---------------------------------------------------------------------------
#define TK_BIG_LIMBS 192   // 6144 bits: 800 digits (2658 bits) + 5^1142 (2652 bits) + slack

struct tk_big {
    uint32_t v[TK_BIG_LIMBS];
    int n;
};

TK_HD void tk_big_set(tk_big& b, uint64_t x) {
    b.n = 0;
    if (x) { b.v[b.n++] = (uint32_t)x; if (x >> 32) b.v[b.n++] = (uint32_t)(x >> 32); }
}
// b = b * m + add ; returns false on capacity overflow
TK_HD bool tk_big_muladd(tk_big& b, uint32_t m, uint32_t add) {
    uint64_t carry = add;
    for (int i = 0; i < b.n; ++i) {
        uint64_t t = (uint64_t)b.v[i] * m + carry;
        b.v[i] = (uint32_t)t;
        carry = t >> 32;
    }
    if (carry) { if (b.n >= TK_BIG_LIMBS) return false; b.v[b.n++] = (uint32_t)carry; }
    return true;
}
TK_HD bool tk_big_mul_pow5(tk_big& b, int k) {
    while (k >= 13) { if (!tk_big_muladd(b, 1220703125u, 0)) return false; k -= 13; }   // 5^13
    uint32_t m = 1;
    while (k-- > 0) m *= 5;
    return m == 1 ? true : tk_big_muladd(b, m, 0);
}
TK_HD bool tk_big_shl(tk_big& b, int bits) {
    if (b.n == 0) return true;
    const int words = bits >> 5, r = bits & 31;
    if (b.n + words + 1 > TK_BIG_LIMBS) return false;
    if (r) {
        uint32_t carry = 0;
        for (int i = 0; i < b.n; ++i) {
            uint32_t nv = (b.v[i] << r) | carry;
            carry = b.v[i] >> (32 - r);
            b.v[i] = nv;
        }
        if (carry) b.v[b.n++] = carry;
    }
    if (words) {
        for (int i = b.n - 1; i >= 0; --i) b.v[i + words] = b.v[i];
        for (int i = 0; i < words; ++i) b.v[i] = 0;
        b.n += words;
    }
    return true;
}
TK_HD int tk_big_cmp(const tk_big& a, const tk_big& b) {
    if (a.n != b.n) return a.n < b.n ? -1 : 1;
    for (int i = a.n - 1; i >= 0; --i)
        if (a.v[i] != b.v[i]) return a.v[i] < b.v[i] ? -1 : 1;
    return 0;
}

// ---------------------------------------------------------------------------
// Token scanner.  `R` is any reader with `int operator()(uint32_t pos) const`
// returning the byte at pos, or -1 at/after the end of the record.
// ---------------------------------------------------------------------------
struct tk_token {
    uint64_t w;          // first <= 19 significant digits
    int32_t  q;          // value ~= w * 10^q
    uint32_t end;        // position one past the number
    uint32_t dig_begin;  // position of the first significant digit (for the exact fallback)
    int32_t  exp10;      // explicit exponent (saturated)
    uint8_t  neg;        // leading '-'
    uint8_t  is_int;     // no fraction, no exponent  (a Python int in the reference)
    uint8_t  truncated;  // non-zero digits were dropped from w
    uint8_t  special;    // 0 finite, 1 NaN, 2 Infinity
    uint8_t  err;        // TK_OK / TK_ERR_SYNTAX / TK_ERR_UNSUPPORTED
    int32_t  nsig;       // count of significant digits seen (may exceed 19)
    int32_t  frac_digits;// digits after '.', total
};

TK_HD bool tk_is_digit(int c) { return (unsigned)(c - '0') <= 9u; }
TK_HD bool tk_is_ws(int c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r'; }
// Characters that can belong to a token (number-like run).  Deliberately a cheap superset of
// the characters of JSON numbers / NaN / Infinity / true / false: 0x2B..0x7F without ',' and
// without the four codes with (c | 6) == 0x5F ('[', ']', 'Y', '_').
TK_HD bool tk_is_numclass(int c) {
    return c >= 0x2B && c <= 0x7F && c != ',' && (c | 6) != 0x5F;
}

template <class R>
TK_HD bool tk_match_word(const R& rd, uint32_t pos, const char* word, int n) {
    for (int i = 0; i < n; ++i)
        if (rd(pos + i) != (int)word[i]) return false;
    return true;
}

// Follows CPython Modules/_json.c _match_number_unicode + the constant branch of
// scan_once_unicode ('N' -> NaN, 'I' -> Infinity, '-' 'I' -> -Infinity).
template <class R>
TK_HD tk_token tk_scan_number(const R& rd, uint32_t pos) {
    tk_token t;
    t.w = 0; t.q = 0; t.end = pos; t.dig_begin = pos; t.exp10 = 0;
    t.neg = 0; t.is_int = 1; t.truncated = 0; t.special = 0; t.err = TK_OK;
    t.nsig = 0; t.frac_digits = 0;

    int c = rd(pos);
    if (c == '-') {
        t.neg = 1;
        c = rd(++pos);
        if (c == 'I') {
            if (tk_match_word(rd, pos, "Infinity", 8)) { t.special = 2; t.is_int = 0; t.end = pos + 8; return t; }
            t.err = TK_ERR_SYNTAX; return t;
        }
    } else if (c == 'N') {
        if (tk_match_word(rd, pos, "NaN", 3)) { t.special = 1; t.is_int = 0; t.end = pos + 3; return t; }
        t.err = TK_ERR_SYNTAX; return t;
    } else if (c == 'I') {
        if (tk_match_word(rd, pos, "Infinity", 8)) { t.special = 2; t.is_int = 0; t.end = pos + 8; return t; }
        t.err = TK_ERR_SYNTAX; return t;
    } else if (c == 't' || c == 'f' || c == 'n') {
        // json.loads yields Python bools for true/false, which torch.tensor stores as 1/0 for
        // every numeric dtype; null becomes None, which torch.tensor rejects (TypeError)
        if (tk_match_word(rd, pos, "true", 4)) { t.w = 1; t.nsig = 1; t.end = pos + 4; return t; }
        if (tk_match_word(rd, pos, "false", 5)) { t.w = 0; t.nsig = 1; t.end = pos + 5; return t; }
        if (tk_match_word(rd, pos, "null", 4)) { t.err = TK_ERR_UNSUPPORTED; return t; }
        t.err = TK_ERR_SYNTAX; return t;
    }

    uint64_t w = 0;
    int nsig = 0;          // significant digits accumulated or dropped
    int dropped_int = 0;   // integer-part digits not folded into w
    bool seen_nonzero = false;
    if (c == '0') {
        c = rd(++pos);
    } else if (c >= '1' && c <= '9') {
        t.dig_begin = pos;
        seen_nonzero = true;
        do {
            if (nsig < 19) w = w * 10 + (uint64_t)(c - '0');
            else { dropped_int++; if (c != '0') t.truncated = 1; }
            nsig++;
            c = rd(++pos);
        } while (tk_is_digit(c));
    } else {
        t.err = TK_ERR_SYNTAX; return t;
    }

    int frac_used = 0;     // fraction digits folded into w (affect q)
    if (c == '.' && tk_is_digit(rd(pos + 1))) {
        t.is_int = 0;
        c = rd(++pos);
        do {
            t.frac_digits++;
            if (!seen_nonzero) {
                if (c != '0') { seen_nonzero = true; t.dig_begin = pos; }
            }
            if (seen_nonzero) {
                if (nsig < 19) { w = w * 10 + (uint64_t)(c - '0'); frac_used = t.frac_digits; }
                else if (c != '0') t.truncated = 1;
                nsig++;
            } else {
                frac_used = t.frac_digits;      // leading fractional zeros shift the exponent
            }
            c = rd(++pos);
        } while (tk_is_digit(c));
    }

    int32_t e10 = 0;
    if (c == 'e' || c == 'E') {
        uint32_t p2 = pos + 1;
        int c2 = rd(p2);
        bool eneg = false;
        if (c2 == '-' || c2 == '+') { eneg = (c2 == '-'); c2 = rd(++p2); }
        if (tk_is_digit(c2)) {
            t.is_int = 0;
            do {
                if (e10 < 100000) e10 = e10 * 10 + (c2 - '0');
                c2 = rd(++p2);
            } while (tk_is_digit(c2));
            if (eneg) e10 = -e10;
            pos = p2;
        }
        // otherwise the number ends before the 'e' (and the caller's separator
        // check will reject the record, as json.loads does)
    }
    t.w = w;
    t.nsig = nsig;
    t.exp10 = e10;
    t.q = e10 - frac_used + dropped_int;
    t.end = pos;
    return t;
}

// Exact decision for truncated inputs; returns the correctly rounded bits.
// lo_bits = conversion of (w, q) (round-down candidate side), hi_bits = of (w+1, q).
template <class R>
TK_HD_NOINLINE uint64_t tk_exact_round(const R& rd, const tk_token& t, uint64_t lo_bits, uint64_t hi_bits) {
    // Candidates differ: the true value lies in (lo-ish, hi-ish).  Take b = the
    // smaller candidate's double and test against the midpoint m = b + ulp/2.
    uint64_t b = lo_bits;
    if (b >= 0x7FF0000000000000ull) return b;
    // midpoint mantissa/exponent: m = (2*M + 1) * 2^(E-1) with b = M * 2^E
    uint64_t frac = b & ((1ull << 52) - 1);
    int bexp = (int)(b >> 52);
    uint64_t M; int E;
    if (bexp == 0) { M = frac; E = -1074; } else { M = frac | (1ull << 52); E = bexp - 1075; }
    uint64_t M2 = 2 * M + 1; int E2 = E - 1;

    // left = D (all significant digits, up to 800 then a sticky 1) ; scale exponent accordingly
    tk_big left; left.n = 0;
    int taken = 0, skipped = 0;
    bool sticky = false;
    uint32_t p = t.dig_begin;
    int total_sig = t.nsig;
    for (int seen = 0; seen < total_sig; ) {
        int c = rd(p++);
        if (c == '.') continue;
        if (taken < 800) { if (!tk_big_muladd(left, 10, (uint32_t)(c - '0'))) return hi_bits; taken++; }
        else { skipped++; if (c != '0') sticky = true; }
        seen++;
    }
    // The significant digits d1..dn read as one integer D end at the last mantissa digit, so
    // value = D * 10^(exp10 - frac_digits).  Dropping the last `skipped` digits raises that
    // exponent by `skipped`; a non-zero remainder is represented by one appended '1' digit.
    int e10 = t.exp10 - t.frac_digits + skipped;
    if (sticky) { if (!tk_big_muladd(left, 10, 1)) return hi_bits; e10 -= 1; }

    tk_big right; tk_big_set(right, M2);
    // compare left * 10^e10  vs  right * 2^E2
    int p5 = e10, p2l = e10, p2r = E2;
    bool ok = true;
    if (p5 >= 0) ok = ok && tk_big_mul_pow5(left, p5); else ok = ok && tk_big_mul_pow5(right, -p5);
    int sh = p2l - p2r;      // left * 2^p2l vs right * 2^p2r  ->  shift the side with the larger power
    if (sh >= 0) ok = ok && tk_big_shl(left, sh); else ok = ok && tk_big_shl(right, -sh);
    if (!ok) return hi_bits;  // capacity exceeded: cannot happen for |exp| <= ~1100 with 800 digits
    int cmp = tk_big_cmp(left, right);
    if (cmp < 0) return b;
    if (cmp > 0) return b + 1;          // next double (works across exponent and into infinity)
    return (b & 1) ? b + 1 : b;         // tie -> even
}

// Decimal token -> double bits (without sign).  Returns bits of |value|.
template <class R>
TK_HD uint64_t tk_token_to_double_bits(const R& rd, const tk_token& t) {
    if (t.special == 1) return 0x7FF8000000000000ull;
    if (t.special == 2) return 0x7FF0000000000000ull;
    if (t.w == 0) return 0;
    int32_t q = t.q;
    // w < 10^19 (and, if truncated, the true value < (w+1) <= 10^19), so for q < -342 the value is
    // below 10^-324 < 2^-1075 (half the smallest subnormal): it rounds to zero.
    if (q < -342) return 0;
    if (q > 308) return 0x7FF0000000000000ull;
    uint64_t lo = tk_eisel_lemire(t.w, q);
    if (!t.truncated) return lo;
    uint64_t hi = tk_eisel_lemire(t.w + 1, q);   // w <= 10^19 - 1 so w + 1 cannot wrap
    if (lo == hi) return lo;
    return tk_exact_round(rd, t, lo, hi);
}

TK_HD double tk_bits_to_double(uint64_t bits) {
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long)bits);
#else
    union { uint64_t u; double d; } x; x.u = bits; return x.d;
#endif
}

// Element conversion for float targets: what torch.tensor(json.loads(v), dtype=float{32,64})
// stores.  Integer literals are Python ints in the reference: "-0" is +0.0, and an int that
// rounds to infinity raises OverflowError.
template <class R>
TK_HD int tk_token_to_f64(const R& rd, const tk_token& t, double* out) {
    uint64_t bits = tk_token_to_double_bits(rd, t);
    if (t.is_int) {
        if (bits == 0x7FF0000000000000ull) return TK_ERR_OVERFLOW;
        if (bits == 0) { *out = 0.0; return TK_OK; }       // int 0 / -0 -> +0.0
    }
    if (t.neg) bits |= 0x8000000000000000ull;
    *out = tk_bits_to_double(bits);
    return TK_OK;
}

// Element conversion for integer targets (bits = 32 or 64).
template <class R>
TK_HD int tk_token_to_i64(const R& rd, const tk_token& t, int bits, int64_t* out) {
    const int64_t tmax = bits == 32 ? 2147483647ll : 9223372036854775807ll;
    const int64_t tmin = -tmax - 1;
    if (t.is_int) {
        // Python int -> int64 (ValueError if it does not fit) -> range check of the target
        if (t.nsig > 19 || t.w > 9223372036854775808ull) return TK_ERR_OVERFLOW;
        if (t.w == 9223372036854775808ull && !t.neg) return TK_ERR_OVERFLOW;
        int64_t v = t.neg ? (int64_t)(0 - t.w) : (int64_t)t.w;
        if (v < tmin || v > tmax) return TK_ERR_OVERFLOW;
        *out = v;
        return TK_OK;
    }
    // Python float -> checked truncation (c10::overflows<intN, double>)
    uint64_t ub = tk_token_to_double_bits(rd, t);
    if (ub >= 0x7FF0000000000000ull) return TK_ERR_OVERFLOW;   // inf / nan
    double d = tk_bits_to_double(ub);
    if (t.neg) d = -d;
    if (d < (double)tmin || d > (double)tmax) return TK_ERR_OVERFLOW;
    if (bits == 64 && d >= 9223372036854775808.0) { *out = tmin; return TK_OK; }  // x86 cvttsd2si result
    *out = (int64_t)d;
    return TK_OK;
}
