// tk_json.cuh -- record-level grammar of a flat JSON array of numbers, as the
// reference's `json.loads(record.value)` accepts it (CPython Modules/_json.c:
// `_parse_array_unicode`, json/decoder.py `JSONDecoder.decode`).
//
// The decode kernels validate a record *in parallel*: number-like runs are
// located by character class, then the lane that owns run j checks
//     number  ws*  ( ','  ws*  <start of a number-like run>  |  ']'  ws*  END )
// and the lane that owns the record checks the prefix
//     [BOM] ws* '[' ws* ( <start of a run> | ']' ws* END ).
// By induction over j the checked ranges tile the whole record, so a record is
// accepted iff a sequential scanner would accept it.  Host+device code (TK_HD).
#pragma once
#include "tk_number.cuh"

// Prefix check.  On success *first is the position of the first number-like
// run, or, for an empty array, *empty is set.
template <class R>
TK_HD int tk_json_prefix(const R& rd, uint32_t pos, uint32_t* first, bool* empty) {
    *empty = false;
    // json.loads(bytes) strips one UTF-8 BOM (json/__init__.py detect_encoding -> 'utf-8-sig')
    if (rd(pos) == 0xEF && rd(pos + 1) == 0xBB && rd(pos + 2) == 0xBF) pos += 3;
    while (tk_is_ws(rd(pos))) ++pos;
    int c = rd(pos);
    if (c != '[') {
        // a bare scalar / object / string is valid JSON but not an array of numbers
        if (c == '{' || c == '"' || tk_is_numclass(c)) return TK_ERR_UNSUPPORTED;
        return TK_ERR_SYNTAX;
    }
    ++pos;
    while (tk_is_ws(rd(pos))) ++pos;
    c = rd(pos);
    if (c == ']') {
        ++pos;
        while (tk_is_ws(rd(pos))) ++pos;
        if (rd(pos) != -1) return TK_ERR_SYNTAX;      // "Extra data"
        *empty = true;
        return TK_OK;
    }
    if (c == '[' || c == '{' || c == '"') return TK_ERR_UNSUPPORTED;   // nesting / strings
    if (!tk_is_numclass(c)) return TK_ERR_SYNTAX;
    *first = pos;
    return TK_OK;
}

// Separator check after a number that ended at `pos`.
template <class R>
TK_HD int tk_json_gap(const R& rd, uint32_t pos) {
    while (tk_is_ws(rd(pos))) ++pos;
    int c = rd(pos);
    if (c == ',') {
        ++pos;
        while (tk_is_ws(rd(pos))) ++pos;
        c = rd(pos);
        if (tk_is_numclass(c)) return TK_OK;           // next run starts here; its lane validates it
        if (c == '[' || c == '{' || c == '"') return TK_ERR_UNSUPPORTED;
        return TK_ERR_SYNTAX;                          // "[1,]" , "[1,,2]" , "[1, "
    }
    if (c == ']') {
        ++pos;
        while (tk_is_ws(rd(pos))) ++pos;
        return rd(pos) == -1 ? TK_OK : TK_ERR_SYNTAX;  // "Extra data"
    }
    return TK_ERR_SYNTAX;                              // "Expecting ',' delimiter"
}

// Separator check of the comma-split decomposition (flat decode path): the record is cut at
// every ',' ; segment j must read  ws* number ws*  and end at the next ',' -- or, for the last
// segment, at ']' ws* END.  Together with the prefix check this tiles the record as well: the
// only bytes not looked at by a number scan or one of the two checks are the commas themselves.
template <class R>
TK_HD int tk_json_gap_flat(const R& rd, uint32_t pos) {
    while (tk_is_ws(rd(pos))) ++pos;
    int c = rd(pos);
    if (c == ',') return TK_OK;
    if (c == ']') {
        ++pos;
        while (tk_is_ws(rd(pos))) ++pos;
        return rd(pos) == -1 ? TK_OK : TK_ERR_SYNTAX;
    }
    if (c == '[' || c == '{' || c == '"') return TK_ERR_UNSUPPORTED;
    return TK_ERR_SYNTAX;
}

// element kinds (tk_dtype_t in include/torchkafka_b200.h)
enum { TK_F32 = 0, TK_F64 = 1, TK_I32 = 2, TK_I64 = 3 };

template <int KIND> struct tk_elem;
template <> struct tk_elem<TK_F32> { typedef float type; };
template <> struct tk_elem<TK_F64> { typedef double type; };
template <> struct tk_elem<TK_I32> { typedef int32_t type; };
template <> struct tk_elem<TK_I64> { typedef int64_t type; };

// Parse + validate the number-like run at `pos` and its trailing separator,
// convert to the element type.  Returns a status code.
template <int KIND, class R>
TK_HD int tk_json_element(const R& rd, uint32_t pos, typename tk_elem<KIND>::type* out) {
    tk_token t = tk_scan_number(rd, pos);
    if (t.err) return t.err;
    int st = tk_json_gap(rd, t.end);
    if (st) return st;
    if (KIND == TK_F32 || KIND == TK_F64) {
        double d;
        st = tk_token_to_f64(rd, t, &d);
        if (st) return st;
        // float32: decimal -> nearest double -> RNE to float (double rounding, as torch.tensor does)
        *out = (typename tk_elem<KIND>::type)d;
        return TK_OK;
    } else {
        int64_t v;
        st = tk_token_to_i64(rd, t, KIND == TK_I32 ? 32 : 64, &v);
        if (st) return st;
        *out = (typename tk_elem<KIND>::type)v;
        return TK_OK;
    }
}

// Segment of the comma-split decomposition starting right after a ',' (or after '['):
// ws* number, then tk_json_gap_flat.
template <int KIND, class R>
TK_HD int tk_json_element_flat(const R& rd, uint32_t pos, typename tk_elem<KIND>::type* out) {
    while (tk_is_ws(rd(pos))) ++pos;
    tk_token t = tk_scan_number(rd, pos);
    if (t.err) return t.err;
    int st = tk_json_gap_flat(rd, t.end);
    if (st) return st;
    if (KIND == TK_F32 || KIND == TK_F64) {
        double d;
        st = tk_token_to_f64(rd, t, &d);
        if (st) return st;
        *out = (typename tk_elem<KIND>::type)d;
    } else {
        int64_t v;
        st = tk_token_to_i64(rd, t, KIND == TK_I32 ? 32 : 64, &v);
        if (st) return st;
        *out = (typename tk_elem<KIND>::type)v;
    }
    return TK_OK;
}

// ---------------------------------------------------------------------------------------------
// Nested arrays of a fixed shape in the comma-split decomposition: what
// `torch.tensor(json.loads(record.value))` accepts for a declared shape dims[0..ndim).  Every comma of
// such a record separates two leaves (whatever the depth), so cutting at the commas yields one
// segment per leaf, and what surrounds leaf e is a function of e and the shape alone:
//     ws* ('[' ws*){opens(e)}  number  ws* (']' ws*){opens(e + 1)}   then ','
// with opens(e) = the number of dimensions j >= 1 whose sub-array starts at leaf e (e is a
// multiple of dims[j] * ... * dims[ndim-1]); the first leaf follows the head `[BOM] ws* ('[' ws*){ndim}`
// and the last one is followed by (']' ws*){ndim} END.  Together with the count check (leaves ==
// product of the extents) the segments tile the record, so a record is accepted iff the
// sequential scanner of tk_json_object.cuh (tk_json_nested_record) accepts it.
#define TK_JSON_NESTED_MAX_DIMS 4

// suffix products of the shape: span[j] = dims[j] * ... * dims[ndim - 1]  (span[0] = leaves per record)
struct tk_nested_shape {
    uint32_t ndim;
    uint32_t span[TK_JSON_NESTED_MAX_DIMS];
};

TK_HD uint32_t tk_nested_opens(const tk_nested_shape& sh, uint32_t e) {
    uint32_t n = 0;
    for (uint32_t j = sh.ndim; j-- > 1;) {           // innermost dimension first
        if (e % sh.span[j]) break;
        ++n;
    }
    return n;
}

template <class R>
TK_HD uint32_t tk_json_skip_ws(const R& rd, uint32_t pos) {
    while (tk_is_ws(rd(pos))) ++pos;
    return pos;
}

// Head of a nested record.  On success *first is the position of the first leaf.
template <class R>
TK_HD int tk_json_prefix_nested(const R& rd, uint32_t pos, uint32_t ndim, uint32_t* first) {
    if (rd(pos) == 0xEF && rd(pos + 1) == 0xBB && rd(pos + 2) == 0xBF) pos += 3;   // utf-8-sig
    for (uint32_t d = 0; d < ndim; ++d) {
        pos = tk_json_skip_ws(rd, pos);
        const int c = rd(pos);
        if (c != '[') return (tk_is_numclass(c) || c == '{' || c == '"') ? TK_ERR_RAGGED : TK_ERR_SYNTAX;
        ++pos;
    }
    pos = tk_json_skip_ws(rd, pos);
    const int c = rd(pos);
    if (c == '[' || c == '{' || c == '"' || c == ']') return TK_ERR_RAGGED;     // deeper / not a number / empty
    if (!tk_is_numclass(c)) return TK_ERR_SYNTAX;
    *first = pos;
    return TK_OK;
}

// One leaf segment starting at `pos` (right behind a ',' -- or, for the first leaf, at the leaf).
template <int KIND, class R>
TK_HD int tk_json_element_nested(const R& rd, uint32_t pos, uint32_t n_open, uint32_t n_close, bool last,
                                 typename tk_elem<KIND>::type* out) {
    for (uint32_t k = 0; k < n_open; ++k) {
        pos = tk_json_skip_ws(rd, pos);
        const int c = rd(pos);
        if (c != '[') return (tk_is_numclass(c) || c == '{' || c == '"') ? TK_ERR_RAGGED : TK_ERR_SYNTAX;
        ++pos;
    }
    pos = tk_json_skip_ws(rd, pos);
    {
        const int c = rd(pos);
        if (c == '[' || c == '{' || c == '"') return TK_ERR_RAGGED;           // deeper than declared / not a number
    }
    const tk_token t = tk_scan_number(rd, pos);
    if (t.err) return t.err;
    pos = t.end;
    for (uint32_t k = 0; k < n_close; ++k) {
        pos = tk_json_skip_ws(rd, pos);
        const int c = rd(pos);
        if (c != ']') return c == ',' ? TK_ERR_RAGGED : TK_ERR_SYNTAX;           // the sub-array goes on
        ++pos;
    }
    pos = tk_json_skip_ws(rd, pos);
    {
        const int c = rd(pos);
        if (last) { if (c != -1) return TK_ERR_SYNTAX; }                        // "Extra data"
        else if (c != ',') return c == ']' ? TK_ERR_RAGGED : TK_ERR_SYNTAX;      // the sub-array ends early
    }
    int st;
    if (KIND == TK_F32 || KIND == TK_F64) {
        double d;
        st = tk_token_to_f64(rd, t, &d);
        if (st) return st;
        *out = (typename tk_elem<KIND>::type)d;
    } else {
        int64_t v;
        st = tk_token_to_i64(rd, t, KIND == TK_I32 ? 32 : 64, &v);
        if (st) return st;
        *out = (typename tk_elem<KIND>::type)v;
    }
    return TK_OK;
}

