// tk_object_fast.cuh -- fast lane of the JSON object decoder: the records a
// producer writes with json.dumps -- one object, plain member names, numbers for the members the
// schema names -- decoded member by member with straight-line word arithmetic, so that the lanes of
// a warp, one record each, stay in step.
//
// Replaces, for those records, the same reference lines as tk_json_object.cuh
//     d = json.loads(record.value); torch.tensor([d[k] for k in fields], dtype=...)
// and accepts exactly
//     '{' member (',' member)* '}'                       nothing in front, nothing behind
//     member = [' '] '"' name '"' ':' [' '] value        name: 1..22 bytes 0x20..0x7E without '"' and '\'
//     value  = number the array fast lane takes (tk_fastlane.cuh), for every member, or
//              any other JSON value, for members the schema does not name (short plain strings
//              are checked here, everything else by the scanner's tk_json_skip_value)
// with every named member present.  Anything else -- other white space, escaped or long member
// names, a named member that is not such a number (true / false, exponents, long literals), a
// missing member -- returns false and the record goes to the sequential scanner (tk_json_object_record), which decodes or reports it; duplicates are
// handled here as there (the last one wins).  Host and device; the host build is compared with
// CPython through the same tests as the scanner (tests/test_object_host.py).
#pragma once
#include "tk_json_object.cuh"
#include "tk_fastlane.cuh"

#define TK_OF_WIN 24u             // bytes one search looks at

// W[0..5] = the 24 bytes at p (`src.word(a, k)`: the word at byte address a + 4k, a 4-byte aligned)
template <class Src>
TK_HD void tk_of_window(const Src& src, uint32_t p, uint32_t* W) {
    const uint32_t a = p & ~3u, sh = (p & 3u) * 8u;
    uint32_t r[7];
TK_FL_UNROLL
    for (int k = 0; k < 7; ++k) r[k] = src.word(a, k);
TK_FL_UNROLL
    for (int k = 0; k < 6; ++k) W[k] = tk_fl_funnel_r(r[k], r[k + 1], sh);
}

// bit i set: byte i of the window equals the byte replicated in c4 (or, inside one word, follows
// such a byte and exceeds it by one: the lowest set bit is always a true match)
TK_HD uint32_t tk_of_eq_mask(const uint32_t* W, uint32_t c4) {
    uint32_t m = 0;
TK_FL_UNROLL
    for (int k = 0; k < 6; ++k) {
        const uint32_t y = W[k] ^ c4;
        const uint32_t z = tk_fl_add(y, 0xFEFEFEFFu) & ~y & 0x80808080u;
        m |= ((z * 0x00204081u) >> 28) << (4 * k);          // bits 7, 15, 23, 31 -> one nibble
    }
    return m;
}

TK_HD uint32_t tk_of_first(uint32_t mask) {                // index of the lowest set bit, TK_OF_WIN if none
    if (!mask) return TK_OF_WIN;
#if defined(__CUDA_ARCH__)
    return (uint32_t)__ffs((int)mask) - 1u;
#else
    return (uint32_t)__builtin_ctz(mask);
#endif
}

// bit i set: byte i of the window is not a plain string character (below 0x20, above 0x7E, '\')
TK_HD uint32_t tk_of_unplain_mask(const uint32_t* W) {
    uint32_t m = 0;
TK_FL_UNROLL
    for (int k = 0; k < 6; ++k) {
        const uint32_t x = W[k];
        const uint32_t lo7 = x & 0x7F7F7F7Fu;
        // bit 7 of a byte: below 0x20 (lo7 + 0x60 does not reach 0x80) / 0x7F / high bit / backslash
        const uint32_t ctrl = ~tk_fl_add(lo7, 0x60606060u);
        const uint32_t del = tk_fl_add(lo7, 0x01010101u);
        const uint32_t yb = x ^ 0x5C5C5C5Cu;
        const uint32_t bs = tk_fl_add(yb, 0xFEFEFEFFu) & ~yb;
        const uint32_t z = (ctrl | del | x | bs) & 0x80808080u;
        m |= ((z * 0x00204081u) >> 28) << (4 * k);
    }
    return m;
}

// The record [pos, end): true = decoded, row[f] holds every field; false = not a record of the
// fast lane (row is unspecified; the caller runs tk_json_object_record).  `rd` and `src` address
// the same bytes (byte reader / word source); `lo` = lowest byte address src may be asked for.
template <int KIND, class Src, class R>
TK_HD bool tk_json_object_fast(const Src& src, const R& rd, uint32_t lo, uint32_t pos, uint32_t end,
                               const tk_obj_fields& ft, const tk_fl_tables& tb, typename tk_elem<KIND>::type* row) {
    typedef typename tk_elem<KIND>::type T;
    if (end - pos < 7u || rd(pos) != '{') return false;
    uint32_t p = pos + 1;
    uint64_t found = 0;
    uint32_t expect = 0;
    uint32_t W[6];
    for (uint32_t guard = 0; guard < 256u; ++guard) {
        // ---- '"' name '"' ':'
        tk_of_window(src, p, W);
        uint32_t sp = (W[0] & 0xFFu) == 0x20u ? 1u : 0u;                       // json.dumps: ", "
        if (((W[0] >> (8u * sp)) & 0xFFu) != 0x22u) return false;
        const uint32_t kb = p + sp + 1u;
        tk_of_window(src, kb, W);
        const uint32_t klen = tk_of_first(tk_of_eq_mask(W, 0x22222222u));     // closing quote
        if (klen == 0u || klen > TK_OF_WIN - 2u) return false;
        if (tk_of_unplain_mask(W) & ((1u << klen) - 1u)) return false;
        // (bytes klen + 1, klen + 2 of the window: ':' and what follows it)
        const uint32_t c_colon = klen + 1u < TK_OF_WIN ? (W[(klen + 1u) >> 2] >> (8u * ((klen + 1u) & 3u))) & 0xFFu : 0u;
        if (c_colon != 0x3Au) return false;
        const int f = tk_obj_lookup(rd, kb, kb + klen, ft, expect);
        p = kb + klen + 2u;                                                    // behind the ':'
        // ---- value
        tk_of_window(src, p, W);
        sp = (W[0] & 0xFFu) == 0x20u ? 1u : 0u;                                // json.dumps: ": "
        const uint32_t c = (W[0] >> (8u * sp)) & 0xFFu;
        const uint32_t vb = p + sp;
        uint32_t ve;                                                           // first byte behind the value
        const bool digit = c - 0x30u <= 9u || c == 0x2Du;
        if (!digit) {
            // not a number: fine for a member the schema does not name -- a short plain string is
            // checked here, anything else (long or escaped strings, arrays, objects, true / false /
            // null) by the scanner's own skip, for this one value
            if (f >= 0) return false;
            bool done = false;
            ve = vb;
            if (c == 0x22u) {
                tk_of_window(src, vb + 1u, W);
                const uint32_t slen = tk_of_first(tk_of_eq_mask(W, 0x22222222u));
                if (slen <= TK_OF_WIN - 2u && !(tk_of_unplain_mask(W) & ((1u << slen) - 1u))) {
                    ve = vb + slen + 2u;
                    done = true;
                }
            }
            if (!done && tk_json_skip_value(rd, vb, &ve) != TK_OK) return false;
        } else {
            // a number: it ends at the first ',' or '}' (the fast lane rejects what else may lie in between)
            if (sp) tk_of_window(src, vb, W);
            const uint32_t stop_at = tk_of_first(tk_of_eq_mask(W, 0x2C2C2C2Cu) | tk_of_eq_mask(W, 0x7D7D7D7Du));
            if (stop_at == 0u || stop_at >= TK_OF_WIN - 1u) return false;
            ve = vb + stop_at;
            // (the lane reads 20 bytes in front of the number's end, or of its mantissa's)
            if (ve < lo + TK_FL_WIN + 5u) return false;
            uint64_t w;
            uint32_t K, dot, neg;
            T v;
            if (tk_fl_scan(src, vb, ve, tb, &w, &K, &dot, &neg)) {
                if (f >= 0 && !tk_fl_value<KIND>(w, K, dot, neg, tb, &v)) return false;
            } else {
                const tk_fl_num num = tk_fl_scan_long(src, vb, ve);              // "0.00…", exponents
                if (!num.ok) return false;
                if (f >= 0 && !tk_fl_value_long<KIND>(num, &v)) return false;
            }
            if (f >= 0) {
                row[f] = v;
                found |= 1ull << f;
                expect = (uint32_t)f + 1u == ft.n ? 0u : (uint32_t)f + 1u;
            }
        }
        if (ve >= end) return false;
        const int d = rd(ve);
        p = ve + 1u;
        if (d == '}') break;
        if (d != ',') return false;
    }
    if (p != end) return false;                                                // (white space behind the object: scanner)
    const uint64_t all = ft.n >= 64u ? ~0ull : (1ull << ft.n) - 1ull;
    return found == all;
}
