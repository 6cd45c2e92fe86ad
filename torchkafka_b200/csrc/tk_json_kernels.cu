// tk_json_kernels.cu -- K1a / K1b (JSON numeric array decode) and K3 (collate, pad, length,
// mask) of the flat decode path.
//
// Replaces, per record, the reference's user hook
//     elements = json.loads(record.value)             README.md:74
//     if len(elements) < self.min_size: return None   README.md:76-77
//     torch.tensor(elements, dtype=...)               README.md:41-43 / BASELINE.md §3
// and, per batch, default_collate -> torch.stack (collate.py:275) or the pad collate of
// SURVEY.md §8c.  Grammar and conversions: tk_json.cuh / tk_number.cuh.
//
//   k_json_scan      warp per record: element count, head check, keep flag, token array
//   (compaction and commit offsets: tk_plan.cu)
//   k_json_rows      warp per kept record: pad, length, mask
//   k_json_parse<T>  thread per element: text -> value, stored in the record's row
#include "tk_swar.cuh"

// ---------------------------------------------------------------- K1 phase 1
// One warp per record, two passes over the record's bytes (the second one hits L1/L2):
//   pass 1  count the commas -> element count (commas + 1, or 0 for "[]"), validate the head of
//           the record ([BOM] ws* '[' ws*), decide the keep flag (rec_kidx = 0 kept / -1 skipped)
//           and reserve the record's range in the slot-wide token array with one atomicAdd;
//   pass 2  cut the record at its commas: one (byte position, record) pair per element.
// The token array feeds the thread-per-element parse kernel (k_json_parse); the order of records
// inside it is the order of the atomicAdds, which does not matter.
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

struct scan_args {
    const uint8_t*  bytes;
    const tk_rec_t* recs;
    uint32_t        n, max_len, min_size;
    int32_t*        rec_len;
    int32_t*        rec_status;
    int32_t*        rec_kidx;
    uint32_t*       rec_tok;
    uint32_t*       tok;          // uint2 {byte position, record} per element
    uint32_t*       tok_count;
    uint32_t        tok_cap;
};

// comma bits of one 16-byte chunk restricted to the record
__device__ __forceinline__ uint32_t comma_mask16(const uint4 v, uint32_t pos, uint32_t start, uint32_t end) {
    const uint32_t c0 = comma4(v.x), c1 = comma4(v.y), c2 = comma4(v.z), c3 = comma4(v.w);
    if ((c0 | c1 | c2 | c3) == 0) return 0;
    uint32_t cm = hibits_to_nibble(c0);
    cm = imad(hibits_to_nibble(c1), 16u, cm);
    cm = imad(hibits_to_nibble(c2), 256u, cm);
    cm = imad(hibits_to_nibble(c3), 4096u, cm);
    if (pos < start || pos + 16 > end) cm &= range16((int)start - (int)pos, (int)end - (int)pos);
    return cm;
}

#define SCAN_KEEP 12     // 512-byte steps whose comma masks stay in registers between the passes

__global__ void __launch_bounds__(256) k_json_scan(const scan_args a) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    uint2* tok = reinterpret_cast<uint2*>(a.tok);
    for (uint32_t i = warp; i < a.n; i += nwarps) {
        const uint32_t start = a.recs[i].start, end = a.recs[i + 1].start;
        const uint32_t base0 = start & ~15u;
        // ---- pass 1: comma masks (kept in registers for short records) and their count
        uint32_t commas = 0;
        uint32_t kept_mask[SCAN_KEEP / 2];
#pragma unroll
        for (int k = 0; k < SCAN_KEEP / 2; ++k) kept_mask[k] = 0;
        uint4 head_chunk = make_uint4(0, 0, 0, 0);             // lane 0: the 16 bytes the record starts in
        // loads are issued four at a time (predicated, no branches in between) so that each warp
        // keeps 2 KiB in flight: the kernel is latency-bound otherwise
#pragma unroll
        for (int g = 0; g < SCAN_KEEP; g += 4) {
            if (base0 + g * 512 < end) {                       // warp-uniform
                uint4 v[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const uint32_t pos = base0 + (g + j) * 512 + lane * 16;
                    v[j] = make_uint4(0, 0, 0, 0);
                    if (pos < end) v[j] = __ldg(reinterpret_cast<const uint4*>(a.bytes + pos));
                }
                if (g == 0) head_chunk = v[0];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int k = g + j;
                    const uint32_t pos = base0 + k * 512 + lane * 16;
                    const uint32_t cm = pos < end ? comma_mask16(v[j], pos, start, end) : 0u;
                    commas += __popc(cm);
                    kept_mask[k >> 1] |= cm << (16 * (k & 1));
                }
            }
        }
        for (uint32_t pos = base0 + SCAN_KEEP * 512 + lane * 16; pos < end; pos += 512)
            commas += __popc(comma_mask16(__ldg(reinterpret_cast<const uint4*>(a.bytes + pos)), pos, start, end));
        commas = __reduce_add_sync(0xffffffffu, commas);
        uint32_t len = 0, t0 = 0;
        if (lane == 0) {
            uint32_t first = start;
            bool empty = false;
            int head;
            // common head "[<token character>": decided from the chunk already in registers,
            // without the dependent byte loads of the general check
            const uint32_t off = start & 15u;
            const uint32_t w0 = off < 4 ? head_chunk.x : off < 8 ? head_chunk.y : off < 12 ? head_chunk.z : head_chunk.w;
            const uint32_t w1 = off < 4 ? head_chunk.y : off < 8 ? head_chunk.z : off < 12 ? head_chunk.w : 0u;
            const uint32_t two = __funnelshift_r(w0, w1, 8 * (off & 3u));     // bytes start, start + 1
            if (off < 15 && start + 1 < end && (two & 0xFFu) == '[' && tk_is_numclass((int)((two >> 8) & 0xFFu))) {
                head = TK_OK;
                first = start + 1;
            } else {
                GlobalReader rd{a.bytes, end};
                head = end > start ? tk_json_prefix(rd, start, &first, &empty) : TK_ERR_SYNTAX;
            }
            len = (!head && empty) ? 0u : commas + 1u;
            const int status = head ? head : (len > a.max_len ? TK_ERR_TOO_LONG : TK_OK);
            a.rec_len[i] = (int32_t)len;
            a.rec_status[i] = status;
            a.rec_kidx[i] = (status == TK_OK && len >= a.min_size) ? 0 : -1;
            if (len) {
                t0 = atomicAdd(a.tok_count, len);
                if (t0 < a.tok_cap) tok[t0] = make_uint2(head ? start : first, i);
            }
            a.rec_tok[i] = t0;
        }
        len = __shfl_sync(0xffffffffu, len, 0);
        t0 = __shfl_sync(0xffffffffu, t0, 0);
        if (len <= 1) continue;
        // ---- pass 2: element e (>= 1) starts right after the e-th comma
        uint32_t seen = 0;
        auto emit = [&](uint32_t cm, uint32_t pos) {
            // exclusive prefix of the per-lane comma counts: one ballot per count level
            const uint32_t cnt = __popc(cm);
            uint32_t before = 0, total = 0;
            for (uint32_t level = 1;; ++level) {
                const uint32_t has = __ballot_sync(0xffffffffu, cnt >= level);
                if (!has) break;
                before += __popc(has & lt_mask);
                total += __popc(has);
            }
            uint32_t t = t0 + 1 + seen + before;
            while (cm) {
                const int b = __ffs(cm) - 1;
                cm &= cm - 1;
                if (t < a.tok_cap) tok[t] = make_uint2(pos + b + 1, i);
                ++t;
            }
            seen += total;
        };
#pragma unroll
        for (int k = 0; k < SCAN_KEEP; ++k) {                  // masks still in registers
            if (base0 + k * 512 < end)
                emit((kept_mask[k >> 1] >> (16 * (k & 1))) & 0xFFFFu, base0 + k * 512 + lane * 16);
        }
        for (uint32_t base = base0 + SCAN_KEEP * 512; base < end; base += 512) {
            const uint32_t pos = base + lane * 16;
            emit(pos < end ? comma_mask16(__ldg(reinterpret_cast<const uint4*>(a.bytes + pos)), pos, start, end) : 0u,
                 pos);
        }
    }
}

extern "C" int tk_json_scan(tk_ctx* ctx, int slot, uint32_t max_len, uint32_t min_size, void* stream) {
    TK_ENTER(ctx, slot);
    tk_slot& s = ctx->slots[slot];
    const uint32_t n = s.n_records;
    int rc = tk_ensure_token_arrays(ctx);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    TK_CUDA(cudaMemsetAsync(s.d.tok_count, 0, sizeof(uint32_t), st));
    if (!n) return 0;
    scan_args a;
    a.bytes = s.d.bytes; a.recs = s.d.recs; a.n = n; a.max_len = max_len; a.min_size = min_size;
    a.rec_len = s.d.rec_len; a.rec_status = s.d.rec_status; a.rec_kidx = s.d.rec_kidx; a.rec_tok = s.d.rec_tok;
    a.tok = s.d.tok_pos; a.tok_count = s.d.tok_count;
    a.tok_cap = (uint32_t)(ctx->tok_cap > 0xFFFFFFFFull ? 0xFFFFFFFFull : ctx->tok_cap);
    const uint32_t blocks = (n + 7) / 8;                   // one warp per record
    k_json_scan<<<blocks, 256, 0, st>>>(a);
    TK_LAUNCH_CHECK();
    return 0;
}

// ================================================================ flat decode path
// After k_json_scan (token array) and the compaction (row of every kept record):
//   k_json_rows      (warp per record) reports the records the scan rejected, checks fixed_len,
//                    and writes pad, length and mask of every kept record's row;
//   k_json_parse<T>  (thread per element) parses one token each: perfectly balanced, every
//                    lane busy whatever the record lengths are, no shared memory.
// The grammar is checked with the comma-split decomposition (tk_json_gap_flat in tk_json.cuh).
struct flat_args {
    const uint8_t*  bytes;
    const tk_rec_t* recs;
    uint32_t        n;
    uint32_t        max_len;
    uint32_t        fixed_len;
    const int32_t*  rec_kidx;
    const int32_t*  rec_len;
    int32_t*        rec_status;
    const uint32_t* rec_tok;
    tk_result_t*    res;
    const uint2*    tok;          // {byte position, record} per element
    const uint32_t* tok_count;
    uint32_t        tok_cap;
    tk_plane_t      values, lengths, mask;
};

__device__ __forceinline__ void flag_record(int32_t* rec_status, tk_result_t* res, uint32_t rec, int code) {
    if (atomicCAS(&rec_status[rec], 0, code) == 0) report_error(res, rec, code);   // count a record once
}

template <int ELEM_BYTES>
__global__ void __launch_bounds__(256) k_json_rows(const flat_args a) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t B = a.res->batch_size;
    for (uint32_t i = warp; i < a.n; i += nwarps) {
        const int32_t kidx = a.rec_kidx[i];
        const uint32_t len = (uint32_t)a.rec_len[i];
        if (lane == 0) {
            const int pre = a.rec_status[i];                  // head / TOO_LONG verdict of the scan
            if (pre) report_error(a.res, i, pre);
            else if (kidx >= 0 && a.fixed_len && len != a.fixed_len)
                flag_record(a.rec_status, a.res, i, TK_ERR_RAGGED);
        }
        if (kidx < 0) continue;
        const uint64_t row = (uint64_t)B + (uint32_t)kidx;
        const uint32_t fill_from = min(len, a.max_len);
        uint8_t* vrow = (uint8_t*)a.values.base + row * a.values.row_bytes;
        if (ELEM_BYTES == 4) {
            uint32_t* p = reinterpret_cast<uint32_t*>(vrow);
            for (uint32_t e = fill_from + lane; e < a.max_len; e += 32) p[e] = 0u;
        } else {
            unsigned long long* p = reinterpret_cast<unsigned long long*>(vrow);
            for (uint32_t e = fill_from + lane; e < a.max_len; e += 32) p[e] = 0ull;
        }
        if (a.lengths.base && lane == 0)
            *reinterpret_cast<int32_t*>((uint8_t*)a.lengths.base + row * a.lengths.row_bytes) = (int32_t)len;
        if (a.mask.base) {
            uint8_t* m = (uint8_t*)a.mask.base + row * a.mask.row_bytes;
            if ((a.max_len & 3u) == 0 && ((uintptr_t)m & 3u) == 0) {
                uint32_t* m32 = reinterpret_cast<uint32_t*>(m);
                for (uint32_t w = lane; w < a.max_len / 4; w += 32) {
                    const int left = (int)len - (int)(w * 4);
                    m32[w] = left >= 4 ? 0x01010101u : (left <= 0 ? 0u : (0x01010101u >> (8 * (4 - left))));
                }
            } else {
                for (uint32_t e = lane; e < a.max_len; e += 32) m[e] = e < len ? 1 : 0;
            }
        }
    }
}

// unaligned 8-byte window out of global memory (L1-resident: neighbouring threads read
// neighbouring tokens)
__device__ __forceinline__ Window8 global_load8(const uint8_t* bytes, uint32_t pos) {
    const uint32_t* w32 = reinterpret_cast<const uint32_t*>(bytes) + (pos >> 2);
    const uint32_t w0 = __ldg(w32), w1 = __ldg(w32 + 1), w2 = __ldg(w32 + 2);
    const uint32_t sh = (pos & 3u) * 8u;
    Window8 w;
    w.lo = __funnelshift_r(w0, w1, sh);
    w.hi = __funnelshift_r(w1, w2, sh);
    return w;
}

// fast_token24 for the flat path: bytes come from global memory, one leading space is skipped,
// and the separator that ends the segment must simply be ','.
__device__ __forceinline__ bool flat_token24(const uint8_t* bytes, uint32_t pos, uint32_t end, tk_token* out) {
    if (pos >= end) return false;
    if (bytes[pos] == ' ') ++pos;
    if (pos >= end) return false;
    const uint32_t avail = end - pos;
    const uint32_t* w32 = reinterpret_cast<const uint32_t*>(bytes) + (pos >> 2);
    const uint32_t sh = (pos & 3u) * 8u;
    uint32_t a[7];
#pragma unroll
    for (int k = 0; k < 7; ++k) a[k] = __ldg(w32 + k);
    uint32_t W[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) W[k] = __funnelshift_r(a[k], a[k + 1], sh);
    uint32_t nd = 0xFF000000u;
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        const uint32_t u = W[k] ^ 0x30303030u;
        const uint32_t t = (imad(u & 0x7F7F7F7Fu, 1u, 0x76767676u) | u) & 0x80808080u;
        nd = imad(hibits_to_nibble(t), 1u << (4 * k), nd);
    }
    if (avail < 24) nd |= 0xFFFFFFFFu << avail;
    const uint32_t neg = (W[0] & 0xFFu) == '-' ? 1u : 0u;
    const uint32_t p = neg;
    const uint32_t L1 = __ffs(nd >> p) - 1;
    if (L1 == 0 || L1 > 8) return false;
    const Window8 wi = {__funnelshift_r(W[0], W[1], 8 * p), __funnelshift_r(W[1], W[2], 8 * p)};
    if (L1 > 1 && (wi.lo & 0xFFu) == '0') return false;
    const uint32_t tpos = p + L1;
    const uint32_t tword = tpos < 4 ? W[0] : (tpos < 8 ? W[1] : W[2]);
    const uint32_t tc = (tword >> (8 * (tpos & 3u))) & 0xFFu;
    uint32_t L2 = 0, endp = tpos;
    uint64_t w = parse_upto8(wi, L1);
    if (tc == '.') {
        const uint32_t q0 = tpos + 1;
        L2 = __ffs(nd >> q0) - 1;
        if (L2 == 0 || L2 > 16 || L1 + L2 > 19) return false;
        endp = q0 + L2;
        if (endp > 23) return false;
        const uint32_t m0 = L2 < 8 ? L2 : 8, m1 = L2 - m0;
        w = w * c_pow10[m0] + parse_upto8(global_load8(bytes, pos + q0), m0);
        if (m1) w = w * c_pow10[m1] + parse_upto8(global_load8(bytes, pos + q0 + 8), m1);
    }
    if (endp >= avail) return false;
    {
        // the segment ends at the next ',' -- or, for the last element, at the closing ']' that is
        // also the last byte of the record; anything else (ws, exponent, garbage): general path
        const uint32_t g = bytes[pos + endp];
        if (g != ',' && !(g == ']' && endp + 1 == avail)) return false;
    }
    tk_token t;
    t.w = w; t.q = -(int32_t)L2; t.end = pos + endp; t.dig_begin = pos + p; t.exp10 = 0;
    t.neg = (uint8_t)neg; t.is_int = L2 == 0; t.truncated = 0; t.special = 0; t.err = TK_OK;
    t.nsig = (int32_t)(L1 + L2); t.frac_digits = (int32_t)L2;
    *out = t;
    return true;
}

template <int KIND>
__global__ void __launch_bounds__(256) k_json_parse(const flat_args a) {
    typedef typename tk_elem<KIND>::type T;
    const uint32_t n_tok = min(*a.tok_count, a.tok_cap);
    const uint64_t B = a.res->batch_size;
    T* values = reinterpret_cast<T*>(a.values.base);
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < n_tok; t += gridDim.x * blockDim.x) {
        const uint2 tk = __ldg(a.tok + t);
        const uint32_t pos = tk.x, rec = tk.y;
        const uint32_t end = a.recs[rec + 1].start;
        GlobalReader rd{a.bytes, end};
        T v;
        int st;
        tk_token tok;
        if (flat_token24(a.bytes, pos, end, &tok)) {
            if (KIND == TK_F32 || KIND == TK_F64) {
                double d;
                st = tk_token_to_f64(rd, tok, &d);
                v = (T)d;
            } else {
                int64_t x = 0;
                st = tk_token_to_i64(rd, tok, KIND == TK_I32 ? 32 : 64, &x);
                v = (T)x;
            }
        } else {
            st = tk_json_element_flat<KIND>(rd, pos, &v);
        }
        if (st) { flag_record(a.rec_status, a.res, rec, st); continue; }
        const int32_t kidx = a.rec_kidx[rec];
        const uint32_t e = t - a.rec_tok[rec];               // element index inside the record
        if (kidx >= 0 && e < a.max_len) values[(B + (uint32_t)kidx) * a.max_len + e] = v;
    }
}

extern "C" int tk_json_decode(tk_ctx* ctx, int slot, int dtype, uint32_t max_len, uint32_t fixed_len,
                              const tk_plane_t* values, const tk_plane_t* lengths, const tk_plane_t* mask,
                              void* stream) {
    TK_ENTER(ctx, slot);
    tk_slot& s = ctx->slots[slot];
    cudaStream_t st = (cudaStream_t)stream;
    TK_REQUIRE(values && values->base, "values plane is required");
    TK_REQUIRE(max_len > 0, "max_len must be positive");
    TK_REQUIRE(dtype >= TK_DTYPE_F32 && dtype <= TK_DTYPE_I64, "unknown dtype %d", dtype);
    TK_REQUIRE(ctx->tok_cap, "tk_json_scan must run before tk_json_decode");
    const uint32_t n = s.n_records;
    if (!n) return 0;
    flat_args a;
    a.bytes = s.d.bytes; a.recs = s.d.recs; a.n = n; a.max_len = max_len; a.fixed_len = fixed_len;
    a.rec_kidx = s.d.rec_kidx; a.rec_len = s.d.rec_len; a.rec_status = s.d.rec_status;
    a.rec_tok = s.d.rec_tok; a.res = s.d.result;
    a.tok = reinterpret_cast<const uint2*>(s.d.tok_pos); a.tok_count = s.d.tok_count;
    a.tok_cap = (uint32_t)(ctx->tok_cap > 0xFFFFFFFFull ? 0xFFFFFFFFull : ctx->tok_cap);
    tk_plane_t z; z.base = nullptr; z.row_bytes = 0;
    a.values = *values;
    a.lengths = lengths ? *lengths : z;
    a.mask = mask ? *mask : z;
    const uint32_t rblocks = min((n + 7) / 8, 148u * 16u);
    if (dtype == TK_DTYPE_F32 || dtype == TK_DTYPE_I32) k_json_rows<4><<<rblocks, 256, 0, st>>>(a);
    else k_json_rows<8><<<rblocks, 256, 0, st>>>(a);
    TK_LAUNCH_CHECK();
    // the element count lives on the device; size the grid for the worst case this slot allows
    // (an element needs two bytes) and let surplus blocks exit at once
    const uint64_t max_tok = (uint64_t)s.n_bytes / 2 + n;
    const uint32_t pblocks = (uint32_t)min((max_tok + 255) / 256, (uint64_t)148 * 32);
    switch (dtype) {
        case TK_DTYPE_F32: k_json_parse<TK_F32><<<pblocks, 256, 0, st>>>(a); break;
        case TK_DTYPE_F64: k_json_parse<TK_F64><<<pblocks, 256, 0, st>>>(a); break;
        case TK_DTYPE_I32: k_json_parse<TK_I32><<<pblocks, 256, 0, st>>>(a); break;
        default:           k_json_parse<TK_I64><<<pblocks, 256, 0, st>>>(a); break;
    }
    TK_LAUNCH_CHECK();
    return 0;
}
