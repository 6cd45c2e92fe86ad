// tk_json_object.cuh -- one JSON object with named numeric fields, as the
// reference user's hook
//     d = json.loads(record.value)
//     return torch.tensor([d[k] for k in fields], dtype=...)
// reads it (CPython Modules/_json.c: `_parse_object_unicode`, `scanstring_unicode`,
// `scan_once_unicode`; json/__init__.py `loads` / `detect_encoding`).  Sequential scanner, one
// record per caller; host+device code (TK_HD), differential-tested against CPython on the host
// (tests/test_object_host.py).
//
// What the whole document must satisfy is what json.loads checks: strict JSON plus NaN / Infinity,
// strings without raw control characters and with well-formed escapes, valid UTF-8 (the payload is
// decoded with errors='surrogatepass' first), one value and nothing but white space behind it.
// What the named fields must satisfy is what torch.tensor checks: a number or true / false
// (stored as 1 / 0); duplicate keys: the last one wins, as in a Python dict.
// Not covered (reported, never silently different): UTF-16/32 payloads, keys written with
// backslash escapes, values nested deeper than TK_OBJ_MAX_DEPTH.
#pragma once
#include "tk_json.cuh"

#define TK_ERR_MISSING_KEY 8      // a named field is absent (KeyError in the reference)
#define TK_OBJ_MAX_FIELDS 64
#define TK_OBJ_MAX_DEPTH 32
#define TK_OBJ_MAX_INT_DIGITS 4300   // sys.int_info.default_max_str_digits: longer int literals raise ValueError

struct tk_obj_fields {
    const uint8_t*  names;        // field names (UTF-8) back to back
    const uint32_t* name_off;     // [n + 1]
    uint32_t        n;
};

template <class R>
TK_HD uint32_t tk_skip_ws(const R& rd, uint32_t pos) {
    while (tk_is_ws(rd(pos))) ++pos;
    return pos;
}

TK_HD bool tk_is_hex(int c) {
    return (unsigned)(c - '0') <= 9u || (unsigned)((c | 0x20) - 'a') <= 5u;
}

// String body starting right behind the opening quote.  *end = position behind the closing quote,
// *plain = no backslash inside (the bytes between the quotes are the string's UTF-8).
template <class R>
TK_HD int tk_json_string(const R& rd, uint32_t pos, uint32_t* end, bool* plain) {
    *plain = true;
    for (;;) {
        int c = rd(pos);
        if (c == '"') { *end = pos + 1; return TK_OK; }
        if (c < 0x20) return TK_ERR_SYNTAX;              // end of record ("Unterminated string") / control character
        if (c == '\\') {
            *plain = false;
            c = rd(pos + 1);
            if (c == 'u') {
                for (uint32_t k = 2; k < 6; ++k)
                    if (!tk_is_hex(rd(pos + k))) return TK_ERR_SYNTAX;
                pos += 6;
            } else if (c == '"' || c == '\\' || c == '/' || c == 'b' || c == 'f' || c == 'n' || c == 'r' || c == 't') {
                pos += 2;
            } else {
                return TK_ERR_SYNTAX;                    // "Invalid \escape"
            }
            continue;
        }
        if (c < 0x80) { ++pos; continue; }
        // UTF-8 sequence; encoded surrogates (ED A0..BF xx) pass, as with 'surrogatepass'
        int more, lo = 0x80, hi = 0xBF;
        if (c >= 0xC2 && c <= 0xDF) more = 1;
        else if (c == 0xE0) { more = 2; lo = 0xA0; }
        else if (c >= 0xE1 && c <= 0xEF) more = 2;
        else if (c == 0xF0) { more = 3; lo = 0x90; }
        else if (c >= 0xF1 && c <= 0xF3) more = 3;
        else if (c == 0xF4) { more = 3; hi = 0x8F; }
        else return TK_ERR_SYNTAX;
        c = rd(pos + 1);
        if (c < lo || c > hi) return TK_ERR_SYNTAX;
        for (int k = 2; k <= more; ++k) {
            c = rd(pos + k);
            if (c < 0x80 || c > 0xBF) return TK_ERR_SYNTAX;
        }
        pos += 1 + more;
    }
}

// Member name at pos (which must hold the opening quote), ws* ':' ws*.  [*kb, *ke) = bytes of the
// name, *value = first byte of the member's value.
template <class R>
TK_HD int tk_json_key(const R& rd, uint32_t pos, uint32_t* kb, uint32_t* ke, bool* plain, uint32_t* value) {
    if (rd(pos) != '"') return TK_ERR_SYNTAX;            // "Expecting property name enclosed in double quotes"
    uint32_t end;
    const int st = tk_json_string(rd, pos + 1, &end, plain);
    if (st) return st;
    *kb = pos + 1;
    *ke = end - 1;
    pos = tk_skip_ws(rd, end);
    if (rd(pos) != ':') return TK_ERR_SYNTAX;
    *value = tk_skip_ws(rd, pos + 1);
    return TK_OK;
}

// Scalar that is not a string: number, NaN, [-]Infinity, true, false, null.
template <class R>
TK_HD int tk_json_skip_scalar(const R& rd, uint32_t pos, uint32_t* end) {
    if (rd(pos) == 'n') {
        if (!tk_match_word(rd, pos, "null", 4)) return TK_ERR_SYNTAX;
        *end = pos + 4;
        return TK_OK;
    }
    const tk_token t = tk_scan_number(rd, pos);
    if (t.err) return t.err;
    if (t.is_int && t.nsig > TK_OBJ_MAX_INT_DIGITS) return TK_ERR_SYNTAX;
    *end = t.end;
    return TK_OK;
}

// Any JSON value starting at pos (white space already skipped): checks it, *end = position behind it.
// Containers are followed with a bit stack (bit 0: the innermost open container is an object).
template <class R>
TK_HD int tk_json_skip_value(const R& rd, uint32_t pos, uint32_t* end) {
    uint32_t is_obj = 0;
    int depth = 0, st;
    uint32_t kb, ke;
    bool plain;
    for (;;) {
        // ---- a value starts at pos
        int c = rd(pos);
        if (c == '{' || c == '[') {
            if (depth == TK_OBJ_MAX_DEPTH) return TK_ERR_UNSUPPORTED;
            is_obj = (is_obj << 1) | (c == '{' ? 1u : 0u);
            ++depth;
            pos = tk_skip_ws(rd, pos + 1);
            if (rd(pos) != ((is_obj & 1u) ? '}' : ']')) {
                if (is_obj & 1u) {
                    st = tk_json_key(rd, pos, &kb, &ke, &plain, &pos);
                    if (st) return st;
                }
                continue;                                   // first member / element
            }
            ++pos;                                          // empty container
            is_obj >>= 1;
            --depth;
        } else if (c == '"') {
            st = tk_json_string(rd, pos + 1, &pos, &plain);
            if (st) return st;
        } else {
            st = tk_json_skip_scalar(rd, pos, &pos);
            if (st) return st;
        }
        // ---- a value ended at pos: next member / element, or close containers
        for (;;) {
            if (depth == 0) { *end = pos; return TK_OK; }
            pos = tk_skip_ws(rd, pos);
            c = rd(pos);
            if (c == ',') {
                pos = tk_skip_ws(rd, pos + 1);
                if (is_obj & 1u) {
                    st = tk_json_key(rd, pos, &kb, &ke, &plain, &pos);
                    if (st) return st;
                }
                break;
            }
            if (c != ((is_obj & 1u) ? '}' : ']')) return TK_ERR_SYNTAX;
            ++pos;
            is_obj >>= 1;
            --depth;
        }
    }
}

// Index of the field whose name is the bytes [kb, ke), -1 if none.  Members usually come in the
// order of the schema: the search starts at `expect`.
template <class R>
TK_HD int tk_obj_lookup(const R& rd, uint32_t kb, uint32_t ke, const tk_obj_fields& ft, uint32_t expect) {
    const uint32_t klen = ke - kb;
    for (uint32_t k = 0; k < ft.n; ++k) {
        uint32_t f = expect + k;
        if (f >= ft.n) f -= ft.n;
        const uint32_t o = ft.name_off[f];
        if (ft.name_off[f + 1] - o != klen) continue;
        uint32_t j = 0;
        while (j < klen && rd(kb + j) == (int)ft.names[o + j]) ++j;
        if (j == klen) return (int)f;
    }
    return -1;
}

// The record at pos: row[f] = value of field f for every field of `ft`.  Returns a status; `row`
// holds unspecified values when it is not TK_OK.
template <int KIND, class R>
TK_HD int tk_json_object_record(const R& rd, uint32_t pos, const tk_obj_fields& ft,
                                typename tk_elem<KIND>::type* row) {
    typedef typename tk_elem<KIND>::type T;
    if (rd(pos) == 0xEF && rd(pos + 1) == 0xBB && rd(pos + 2) == 0xBF) pos += 3;   // utf-8-sig
    pos = tk_skip_ws(rd, pos);
    int c = rd(pos);
    if (c != '{') {
        // another JSON value: d[k] raises TypeError -- or not JSON at all
        return (c == '[' || c == '"' || tk_is_numclass(c)) ? TK_ERR_UNSUPPORTED : TK_ERR_SYNTAX;
    }
    pos = tk_skip_ws(rd, pos + 1);
    // per field: seen at all / last occurrence is not a number / last occurrence does not fit T
    uint64_t found = 0, bad_type = 0, bad_range = 0;
    uint32_t expect = 0;
    int st;
    if (rd(pos) == '}') {
        ++pos;
    } else {
        for (;;) {
            uint32_t kb, ke;
            bool plain;
            st = tk_json_key(rd, pos, &kb, &ke, &plain, &pos);
            if (st) return st;
            if (!plain) return TK_ERR_UNSUPPORTED;          // escaped member name: not compared
            const int f = tk_obj_lookup(rd, kb, ke, ft, expect);
            c = rd(pos);
            if (f < 0) {
                st = tk_json_skip_value(rd, pos, &pos);
                if (st) return st;
            } else {
                const uint64_t bit = 1ull << f;
                expect = (uint32_t)f + 1u == ft.n ? 0u : (uint32_t)f + 1u;
                found |= bit;
                if (c == '"' || c == '{' || c == '[' || c == 'n') {
                    st = tk_json_skip_value(rd, pos, &pos);
                    if (st) return st;
                    bad_type |= bit;
                    bad_range &= ~bit;
                } else {
                    const tk_token t = tk_scan_number(rd, pos);
                    if (t.err) return t.err;
                    if (t.is_int && t.nsig > TK_OBJ_MAX_INT_DIGITS) return TK_ERR_SYNTAX;
                    if (KIND == TK_F32 || KIND == TK_F64) {
                        double d = 0.0;
                        st = tk_token_to_f64(rd, t, &d);
                        row[f] = (T)d;                      // float32: double-rounded, as torch.tensor does
                    } else {
                        int64_t v = 0;
                        st = tk_token_to_i64(rd, t, KIND == TK_I32 ? 32 : 64, &v);
                        row[f] = (T)v;
                    }
                    bad_type &= ~bit;
                    bad_range = st ? (bad_range | bit) : (bad_range & ~bit);
                    pos = t.end;
                }
            }
            pos = tk_skip_ws(rd, pos);
            c = rd(pos);
            if (c == '}') { ++pos; break; }
            if (c != ',') return TK_ERR_SYNTAX;              // "Expecting ',' delimiter"
            pos = tk_skip_ws(rd, pos + 1);
        }
    }
    if (rd(tk_skip_ws(rd, pos)) != -1) return TK_ERR_SYNTAX;    // "Extra data"
    const uint64_t all = ft.n >= 64u ? ~0ull : (1ull << ft.n) - 1ull;
    if (found != all) return TK_ERR_MISSING_KEY;
    if (bad_type) return TK_ERR_UNSUPPORTED;
    if (bad_range) return TK_ERR_OVERFLOW;
    return TK_OK;
}

// ---------------------------------------------------------------- nested arrays of fixed shape
// The record at pos must be arrays nested `ndim` deep with extents dims[0..ndim), numbers (or
// true / false) at the leaves: what `torch.tensor(json.loads(v), dtype=...)` turns into a tensor
// of that shape.  row[] receives the leaves in C order.  Another nesting depth
// or another extent anywhere is TK_ERR_RAGGED (torch.tensor raises on ragged lists, torch.stack
// on samples of different shapes).
#define TK_NESTED_MAX_DIMS 4
template <int KIND, class R>
TK_HD int tk_json_nested_record(const R& rd, uint32_t pos, const uint32_t* dims, uint32_t ndim,
                                typename tk_elem<KIND>::type* row) {
    typedef typename tk_elem<KIND>::type T;
    if (rd(pos) == 0xEF && rd(pos + 1) == 0xBB && rd(pos + 2) == 0xBF) pos += 3;   // utf-8-sig
    uint32_t idx[TK_NESTED_MAX_DIMS];
    uint32_t d = 0, n_out = 0;
    int bad_range = 0;
    for (;;) {
        pos = tk_skip_ws(rd, pos);
        int c = rd(pos);
        if (d < ndim) {                                     // an array opens here
            if (c != '[') return (tk_is_numclass(c) || c == '{' || c == '"') ? TK_ERR_RAGGED : TK_ERR_SYNTAX;
            ++pos;
            idx[d++] = 0;
            continue;
        }
        if (c == '[' || c == '{' || c == '"') return TK_ERR_RAGGED;      // deeper than declared / not a number
        const tk_token t = tk_scan_number(rd, pos);
        if (t.err) return t.err;
        int st;
        if (KIND == TK_F32 || KIND == TK_F64) {
            double v = 0.0;
            st = tk_token_to_f64(rd, t, &v);
            row[n_out] = (T)v;
        } else {
            int64_t v = 0;
            st = tk_token_to_i64(rd, t, KIND == TK_I32 ? 32 : 64, &v);
            row[n_out] = (T)v;
        }
        ++n_out;
        bad_range |= st;
        pos = t.end;
        // a value ended: next element of the enclosing array, or close the arrays that are full
        for (;;) {
            if (d == 0) {
                if (rd(tk_skip_ws(rd, pos)) != -1) return TK_ERR_SYNTAX;      // "Extra data"
                return bad_range ? TK_ERR_OVERFLOW : TK_OK;
            }
            pos = tk_skip_ws(rd, pos);
            c = rd(pos);
            if (++idx[d - 1] < dims[d - 1]) {
                if (c != ',') return c == ']' ? TK_ERR_RAGGED : TK_ERR_SYNTAX;     // too few elements
                ++pos;
                break;
            }
            if (c != ']') return c == ',' ? TK_ERR_RAGGED : TK_ERR_SYNTAX;         // too many
            ++pos;
            --d;
        }
    }
}
