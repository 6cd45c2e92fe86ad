// Host build of the device-side record decoder (tk_number.cuh / tk_json.cuh).
//
// TEST INFRASTRUCTURE ONLY.  The production path runs these same functions
// inside CUDA kernels; this translation unit compiles them with g++ so that the
// grammar and the decimal->binary conversion can be differential-tested against
// CPython (json.loads + float()) on a machine without a GPU
// (tests/test_number_host.py).  Nothing in torchkafka_b200/ loads this library.
#include <stdint.h>
#include <string.h>
#include "../../torchkafka_b200/csrc/tk_json_object.cuh"
#include "../../torchkafka_b200/csrc/tk_inflate.cuh"
#include "../../torchkafka_b200/csrc/tk_zstd.cuh"

// word source of the fast lanes over a host buffer
struct HostWords {
    const uint8_t* p;
    uint32_t word(uint32_t a, int k) const { uint32_t v; memcpy(&v, p + a + 4 * k, 4); return v; }
};

struct HostReader {
    const uint8_t* p;
    uint32_t end;
    int operator()(uint32_t pos) const { return pos < end ? (int)p[pos] : -1; }
};

template <int KIND>
static int decode(const uint8_t* buf, uint32_t len, int max_len, void* out, int* n_out) {
    typedef typename tk_elem<KIND>::type T;
    HostReader rd{buf, len};
    uint32_t first = 0;
    bool empty = false;
    *n_out = 0;
    int st = tk_json_prefix(rd, 0, &first, &empty);
    if (st) return st;
    int n = 0;
    int prev_class = 0;
    int worst = TK_OK;
    for (uint32_t pos = 0; pos < len; ++pos) {
        int cls = tk_is_numclass(buf[pos]) ? 1 : 0;
        if (cls && !prev_class) {
            T v;
            st = tk_json_element<KIND>(rd, pos, &v);
            if (st && !worst) worst = st;
            if (n < max_len) ((T*)out)[n] = st ? (T)0 : v;
            ++n;
        }
        prev_class = cls;
    }
    *n_out = n;
    if (worst) return worst;
    if (n > max_len) return TK_ERR_TOO_LONG;
    return TK_OK;
}

// Mirror of the flat (comma-split) decode path: prefix check, then one segment per comma.
template <int KIND>
static int decode_flat(const uint8_t* buf, uint32_t len, int max_len, void* out, int* n_out) {
    typedef typename tk_elem<KIND>::type T;
    HostReader rd{buf, len};
    uint32_t first = 0;
    bool empty = false;
    *n_out = 0;
    int st = tk_json_prefix(rd, 0, &first, &empty);
    if (st) return st;
    if (empty) return TK_OK;
    int n = 0, worst = TK_OK;
    uint32_t seg = first;
    for (;;) {
        T v;
        st = tk_json_element_flat<KIND>(rd, seg, &v);
        if (st && !worst) worst = st;
        if (n < max_len) ((T*)out)[n] = st ? (T)0 : v;
        ++n;
        uint32_t p = seg;
        while (p < len && buf[p] != ',') ++p;      // next split point
        if (p >= len) break;
        seg = p + 1;
    }
    *n_out = n;
    if (worst) return worst;
    if (n > max_len) return TK_ERR_TOO_LONG;
    return TK_OK;
}

extern "C" int tk_host_decode_flat(const uint8_t* buf, uint32_t len, int kind, int max_len,
                                   void* out, int* n_out) {
    switch (kind) {
        case TK_F32: return decode_flat<TK_F32>(buf, len, max_len, out, n_out);
        case TK_F64: return decode_flat<TK_F64>(buf, len, max_len, out, n_out);
        case TK_I32: return decode_flat<TK_I32>(buf, len, max_len, out, n_out);
        case TK_I64: return decode_flat<TK_I64>(buf, len, max_len, out, n_out);
    }
    return -1;
}

extern "C" int tk_host_decode(const uint8_t* buf, uint32_t len, int kind, int max_len,
                              void* out, int* n_out) {
    switch (kind) {
        case TK_F32: return decode<TK_F32>(buf, len, max_len, out, n_out);
        case TK_F64: return decode<TK_F64>(buf, len, max_len, out, n_out);
        case TK_I32: return decode<TK_I32>(buf, len, max_len, out, n_out);
        case TK_I64: return decode<TK_I64>(buf, len, max_len, out, n_out);
    }
    return -1;
}

// Single decimal string -> double bits (status in return value), for bulk float tests.
extern "C" int tk_host_strtod(const uint8_t* buf, uint32_t len, uint64_t* bits) {
    HostReader rd{buf, len};
    tk_token t = tk_scan_number(rd, 0);
    if (t.err) return t.err;
    if (t.end != len) return TK_ERR_SYNTAX;
    double d;
    int st = tk_token_to_f64(rd, t, &d);
    if (st) return st;
    memcpy(bits, &d, 8);
    return TK_OK;
}

// Many newline-separated decimal strings at once (fast path for million-case tests).
extern "C" int tk_host_strtod_many(const uint8_t* buf, const uint32_t* starts, uint32_t n,
                                   uint64_t* bits, uint8_t* status) {
    for (uint32_t i = 0; i < n; ++i) {
        bits[i] = 0;
        status[i] = (uint8_t)tk_host_strtod(buf + starts[i], starts[i + 1] - starts[i], &bits[i]);
    }
    return 0;
}

// One JSON object with named numeric fields: row[f] per field, status -- as the device decodes it:
// the fast lane (tk_object_fast.cuh) first, the sequential scanner (tk_json_object.cuh) for what it
// does not take.  *fast (optional) = 1 if the fast lane decoded the record.
#include "../../torchkafka_b200/csrc/tk_object_fast.cuh"
#include <vector>
template <int KIND>
static int decode_object(const uint8_t* buf, uint32_t len, const tk_obj_fields& ft, void* row, int* fast) {
    typedef typename tk_elem<KIND>::type T;
    static tk_fl_tables tb;
    static bool ready = false;
    if (!ready) { tk_fl_tables_fill(tb, 0, 1); ready = true; }
    // the record with bytes in front of it and behind it, as in the device's window
    std::vector<uint8_t> pad(32 + (size_t)len + 64, 0);
    memcpy(pad.data() + 32, buf, len);
    HostReader rd{pad.data(), 32 + len};
    HostWords src{pad.data()};
    if (tk_json_object_fast<KIND>(src, rd, 0u, 32u, 32u + len, ft, tb, (T*)row)) {
        if (fast) *fast = 1;
        return TK_OK;
    }
    if (fast) *fast = 0;
    return tk_json_object_record<KIND>(rd, 32, ft, (T*)row);
}
extern "C" int tk_host_decode_object_ex(const uint8_t* buf, uint32_t len, int kind, const uint8_t* names,
                                        const uint32_t* name_off, uint32_t n_fields, void* row, int* fast) {
    tk_obj_fields ft{names, name_off, n_fields};
    switch (kind) {
        case TK_F32: return decode_object<TK_F32>(buf, len, ft, row, fast);
        case TK_F64: return decode_object<TK_F64>(buf, len, ft, row, fast);
        case TK_I32: return decode_object<TK_I32>(buf, len, ft, row, fast);
        case TK_I64: return decode_object<TK_I64>(buf, len, ft, row, fast);
    }
    return -1;
}
extern "C" int tk_host_decode_object(const uint8_t* buf, uint32_t len, int kind, const uint8_t* names,
                                     const uint32_t* name_off, uint32_t n_fields, void* row) {
    return tk_host_decode_object_ex(buf, len, kind, names, name_off, n_fields, row, nullptr);
}
// The sequential scanner alone (the fast lane's results are compared with it).
extern "C" int tk_host_decode_object_scanner(const uint8_t* buf, uint32_t len, int kind, const uint8_t* names,
                                             const uint32_t* name_off, uint32_t n_fields, void* row) {
    HostReader rd{buf, len};
    tk_obj_fields ft{names, name_off, n_fields};
    switch (kind) {
        case TK_F32: return tk_json_object_record<TK_F32>(rd, 0, ft, (float*)row);
        case TK_F64: return tk_json_object_record<TK_F64>(rd, 0, ft, (double*)row);
        case TK_I32: return tk_json_object_record<TK_I32>(rd, 0, ft, (int32_t*)row);
        case TK_I64: return tk_json_object_record<TK_I64>(rd, 0, ft, (int64_t*)row);
    }
    return -1;
}

// One raw DEFLATE stream (tk_inflate.cuh) with the match copy done serially: 0 ok, 1 error.
extern "C" int tk_host_inflate(const uint8_t* in, uint32_t len, uint8_t* out, uint32_t cap, uint32_t* out_len,
                               uint32_t* consumed) {
    tk_inflate_state s;
    tk_inflate_tables tb;
    tk_inflate_init(s, in, 0, len);
    uint32_t op = 0, mlen = 0, mdist = 0;
    for (;;) {
        const int ev = tk_inflate_next(s, tb, out, cap, op, &mlen, &mdist, 12);
        if (ev == TK_INF_ERROR) return 1;
        if (ev == TK_INF_DONE) break;
        for (uint32_t k = 0; k < mlen; ++k) out[op + k] = out[op - mdist + k];
        op += mlen;
    }
    *out_len = op;
    *consumed = tk_inf_consumed(s);
    return 0;
}

// The same with the matches collected in batches (tk_inflate_next with a batch) and copied in the
// least favourable order the device may use: the matches that read nothing an earlier match of the
// batch writes first, last one first, then the others in order.
extern "C" int tk_host_inflate_batched(const uint8_t* in, uint32_t len, uint8_t* out, uint32_t cap, uint32_t* out_len,
                                       uint32_t* consumed) {
    tk_inflate_state s;
    static tk_inflate_tables tb;
    tk_inflate_init(s, in, 0, len);
    uint32_t op = 0, mlen = 0, mdist = 0;
    for (;;) {
        const int ev = tk_inflate_next(s, tb, out, cap, op, &mlen, &mdist, 0, &tb.batch);
        if (ev == TK_INF_ERROR) return 1;
        const tk_inflate_match* b = tb.batch.m;
        bool dep[TK_INF_BATCH_MAX];
        for (uint32_t k = 0; k < s.nb; ++k) dep[k] = b[k].dst - b[k].dist + b[k].len > b[0].dst || b[k].len > 32u;
        for (uint32_t k = s.nb; k-- > 0;)
            if (!dep[k]) {
                uint8_t tmp[32];
                memcpy(tmp, out + b[k].dst - b[k].dist, b[k].len);          // every load before the first store
                memcpy(out + b[k].dst, tmp, b[k].len);
            }
        for (uint32_t k = 0; k < s.nb; ++k)
            if (dep[k])
                for (uint32_t j = 0; j < b[k].len; ++j) out[b[k].dst + j] = out[b[k].dst - b[k].dist + j];
        s.nb = 0;
        if (ev == TK_INF_DONE) break;
    }
    *out_len = op;
    *consumed = tk_inf_consumed(s);
    return 0;
}

// Zstandard payload (tk_zstd.cuh) with the copies done serially.  emit = 0: size pass only.
// Returns 0 ok, else the decoder's status; *out_len = bytes produced.
extern "C" int tk_host_zstd(const uint8_t* in, uint32_t len, uint8_t* out, uint32_t cap, int emit, uint32_t* out_len) {
    static tk_zstd_tables tb;
    tk_zstd_state s;
    tk_zstd_init(s, tb, in, 0, len);
    uint32_t op = 0;
    tk_zstd_event ev;
    for (;;) {
        const int st = tk_zstd_next(s, tb, out, emit ? cap : 0x7FFFFFFFu, op, emit != 0, &ev);
        if (st == TK_ZS_DONE) break;
        if (st == TK_ZS_HUF) {
            for (uint32_t k = 0; k < ev.h_streams; ++k)
                if (!tk_huf_stream(in, ev.h_lo[k], ev.h_hi[k], tb.huf, (int)ev.h_log, out + ev.h_dst[k], ev.h_n[k])) return TK_ZS_ERROR;
            continue;
        }
        if (st != TK_ZS_COPY) return st;
        if (emit) {
            if (ev.lit_kind == TK_ZL_INPUT) memmove(out + op, in + ev.lit_src, ev.lit_len);
            else if (ev.lit_kind == TK_ZL_RLE) memset(out + op, (int)ev.lit_src, ev.lit_len);
            else for (uint32_t k = 0; k < ev.lit_len; ++k) out[op + k] = out[ev.lit_src + k];
            for (uint32_t k = 0; k < ev.mlen; ++k) out[op + ev.lit_len + k] = out[op + ev.lit_len - ev.offset + k];
        }
        op += ev.lit_len + ev.mlen;
    }
    *out_len = op;
    return 0;
}

// The same through the batched sequence decoder (tk_zstd_next_batch), the copies of a batch executed
// in the order the device executes them: every literal run, then the matches that do not depend on
// an earlier match of the batch, then the dependent ones in sequence order.
extern "C" int tk_host_zstd_batched(const uint8_t* in, uint32_t len, uint8_t* out, uint32_t cap, uint32_t* out_len) {
    static tk_zstd_tables tb;
    static tk_zstd_batch b;
    tk_zstd_state s;
    tk_zstd_init(s, tb, in, 0, len);
    uint32_t op = 0;
    tk_zstd_event ev;
    for (;;) {
        const int n = tk_zstd_next_batch(s, tb, cap, op, true, &b);
        if (n < 0) return -n;
        if (n > 0) {
            uint32_t dst[TK_ZS_BATCH], pos = op, first_dst = 0;
            bool have_first = false;
            for (int k = 0; k < n; ++k) { dst[k] = pos; pos += b.q[k].lit_len + b.q[k].mlen; }
            if (b.lit_kind == TK_ZL_OUTPUT && pos > b.q[0].lit_src) {
                // the batch writes into parked literals it still reads: sequence order
                for (int k = 0; k < n; ++k) {
                    for (uint32_t j = 0; j < b.q[k].lit_len; ++j) out[dst[k] + j] = out[b.q[k].lit_src + j];
                    const uint32_t mdst = dst[k] + b.q[k].lit_len;
                    for (uint32_t j = 0; j < b.q[k].mlen; ++j) out[mdst + j] = out[mdst - b.q[k].offset + j];
                }
                op = pos;
                continue;
            }
            // literals, last run first: the device copies them side by side, so no run may write
            // what another one reads -- the reverse order makes such a dependency visible here
            for (int k = n - 1; k >= 0; --k) {
                if (b.lit_kind == TK_ZL_INPUT) memmove(out + dst[k], in + b.q[k].lit_src, b.q[k].lit_len);
                else if (b.lit_kind == TK_ZL_RLE) memset(out + dst[k], (int)b.q[k].lit_src, b.q[k].lit_len);
                else for (uint32_t j = 0; j < b.q[k].lit_len; ++j) out[dst[k] + j] = out[b.q[k].lit_src + j];
            }
            bool dep[TK_ZS_BATCH];
            for (int k = 0; k < n; ++k) {
                const uint32_t mdst = dst[k] + b.q[k].lit_len;
                if (b.q[k].mlen && !have_first) { have_first = true; first_dst = mdst; }
                dep[k] = b.q[k].mlen && (mdst - b.q[k].offset + b.q[k].mlen > first_dst || b.q[k].offset < b.q[k].mlen || b.q[k].mlen > 32u);
            }
            for (int pass = 0; pass < 2; ++pass)                 // independent matches first, then the dependent ones
                for (int k = 0; k < n; ++k) {
                    if ((int)dep[k] != pass) continue;
                    const uint32_t mdst = dst[k] + b.q[k].lit_len;
                    for (uint32_t j = 0; j < b.q[k].mlen; ++j) out[mdst + j] = out[mdst - b.q[k].offset + j];
                }
            op = pos;
            continue;
        }
        const int st = tk_zstd_next(s, tb, out, cap, op, true, &ev);
        if (st == TK_ZS_DONE) break;
        if (st == TK_ZS_HUF) {
            for (uint32_t k = 0; k < ev.h_streams; ++k)
                if (!tk_huf_stream(in, ev.h_lo[k], ev.h_hi[k], tb.huf, (int)ev.h_log, out + ev.h_dst[k], ev.h_n[k])) return TK_ZS_ERROR;
            continue;
        }
        if (st != TK_ZS_COPY) return st;
        if (ev.lit_kind == TK_ZL_INPUT) memmove(out + op, in + ev.lit_src, ev.lit_len);
        else if (ev.lit_kind == TK_ZL_RLE) memset(out + op, (int)ev.lit_src, ev.lit_len);
        else for (uint32_t k = 0; k < ev.lit_len; ++k) out[op + k] = out[ev.lit_src + k];
        for (uint32_t k = 0; k < ev.mlen; ++k) out[op + ev.lit_len + k] = out[op + ev.lit_len - ev.offset + k];
        op += ev.lit_len + ev.mlen;
    }
    *out_len = op;
    return 0;
}

// Nested arrays of a fixed shape (tk_json_object.cuh): leaves in C order, status.
extern "C" int tk_host_decode_nested(const uint8_t* buf, uint32_t len, int kind, const uint32_t* dims, uint32_t ndim,
                                     void* row) {
    HostReader rd{buf, len};
    switch (kind) {
        case TK_F32: return tk_json_nested_record<TK_F32>(rd, 0, dims, ndim, (float*)row);
        case TK_F64: return tk_json_nested_record<TK_F64>(rd, 0, dims, ndim, (double*)row);
        case TK_I32: return tk_json_nested_record<TK_I32>(rd, 0, dims, ndim, (int32_t*)row);
        case TK_I64: return tk_json_nested_record<TK_I64>(rd, 0, dims, ndim, (int64_t*)row);
    }
    return -1;
}

// Mirror of the flat (comma-split) path for nested arrays of a fixed shape (tk_json.cuh): head,
// one segment per comma, count check.  Same contract as tk_host_decode_nested.
template <int KIND>
static int decode_nested_flat(const uint8_t* buf, uint32_t len, const uint32_t* dims, uint32_t ndim, void* row) {
    typedef typename tk_elem<KIND>::type T;
    HostReader rd{buf, len};
    tk_nested_shape sh;
    sh.ndim = ndim;
    uint32_t prod = 1;
    for (uint32_t j = ndim; j-- > 0;) { prod *= dims[j]; sh.span[j] = prod; }
    uint32_t first = 0;
    int st = tk_json_prefix_nested(rd, 0, ndim, &first);
    if (st) return st;
    uint32_t commas = 0;
    for (uint32_t p = 0; p < len; ++p) commas += buf[p] == ',';
    const uint32_t n = commas + 1;
    if (n != prod) return TK_ERR_RAGGED;
    int worst = TK_OK;
    uint32_t seg = first;
    for (uint32_t e = 0; e < n; ++e) {
        const bool last = e + 1 == n;
        T v = (T)0;
        st = tk_json_element_nested<KIND>(rd, seg, e ? tk_nested_opens(sh, e) : 0u,
                                          last ? ndim : tk_nested_opens(sh, e + 1), last, &v);
        if (st && !worst) worst = st;
        ((T*)row)[e] = st ? (T)0 : v;
        uint32_t p = seg;
        while (p < len && buf[p] != ',') ++p;
        seg = p + 1;
    }
    return worst;
}

extern "C" int tk_host_decode_nested_flat(const uint8_t* buf, uint32_t len, int kind, const uint32_t* dims,
                                          uint32_t ndim, void* row) {
    switch (kind) {
        case TK_F32: return decode_nested_flat<TK_F32>(buf, len, dims, ndim, row);
        case TK_F64: return decode_nested_flat<TK_F64>(buf, len, dims, ndim, row);
        case TK_I32: return decode_nested_flat<TK_I32>(buf, len, dims, ndim, row);
        case TK_I64: return decode_nested_flat<TK_I64>(buf, len, dims, ndim, row);
    }
    return -1;
}


// The fast lane of the JSON array kernels (tk_fastlane.cuh) on one element [pos, stop) of a padded
// buffer.  Returns 1 and the value when the lane accepts the literal, 0 when it leaves it to the
// scalar grammar.
#include "../../torchkafka_b200/csrc/tk_fastlane.cuh"
extern "C" int tk_host_fastlane(const uint8_t* buf, uint32_t pos, uint32_t stop, int kind, void* out) {
    static tk_fl_tables tb;
    static bool ready = false;
    if (!ready) { tk_fl_tables_fill(tb, 0, 1); ready = true; }
    HostWords src{buf};
    uint64_t w;
    uint32_t K, dot, neg;
    if (!tk_fl_scan(src, pos, stop, tb, &w, &K, &dot, &neg)) {
        // the kernels' slow branch: the long forms, before the grammar
        const tk_fl_num num = tk_fl_scan_long(src, pos, stop);
        if (!num.ok) return 0;
        switch (kind) {
            case TK_F32: return tk_fl_value_long<TK_F32>(num, (float*)out) ? 1 : 0;
            case TK_F64: return tk_fl_value_long<TK_F64>(num, (double*)out) ? 1 : 0;
            case TK_I32: return tk_fl_value_long<TK_I32>(num, (int32_t*)out) ? 1 : 0;
            case TK_I64: return tk_fl_value_long<TK_I64>(num, (int64_t*)out) ? 1 : 0;
        }
        return -1;
    }
    switch (kind) {
        case TK_F32: return tk_fl_value<TK_F32>(w, K, dot, neg, tb, (float*)out) ? 1 : 0;
        case TK_F64: return tk_fl_value<TK_F64>(w, K, dot, neg, tb, (double*)out) ? 1 : 0;
        case TK_I32: return tk_fl_value<TK_I32>(w, K, dot, neg, tb, (int32_t*)out) ? 1 : 0;
        case TK_I64: return tk_fl_value<TK_I64>(w, K, dot, neg, tb, (int64_t*)out) ? 1 : 0;
    }
    return -1;
}

// Many elements at once (the Python loop would dominate): n elements, element i = [pos[i], stop[i]).
extern "C" void tk_host_fastlane_many(const uint8_t* buf, const uint32_t* pos, const uint32_t* stop, uint32_t n, int kind,
                                      void* out, uint8_t* accepted) {
    const size_t sz = (kind == TK_F32 || kind == TK_I32) ? 4 : 8;
    for (uint32_t i = 0; i < n; ++i)
        accepted[i] = (uint8_t)tk_host_fastlane(buf, pos[i], stop[i], kind, (uint8_t*)out + sz * i);
}
