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
//   k_json_parse<T>  rows part (pad, length, mask of the kept records) + thread per element:
//                    text -> value, stored in the record's row
#include "tk_json_parse.cuh"

// ---------------------------------------------------------------- K1 phase 1
// One warp per record, two passes over the record's bytes (the second one hits L1/L2):
//   pass 1  count the commas -> element count (commas + 1, or 0 for "[]"), validate the head of
//           the record ([BOM] ws* '[' ws*), decide the keep flag (rec_kidx = 0 kept / -1 skipped)
//           and reserve the record's range in the slot-wide token array with one atomicAdd;
//   pass 2  cut the record at its commas: one (byte position, record) pair per element.
// The token array feeds the thread-per-element parse kernel (k_json_parse); the order of records
// inside it is the order of the atomicAdds, which does not matter.
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
    uint32_t*       tok_count_next;   // counter of the slot's next scan: zeroed here
    uint32_t        tok_cap;
    uint32_t        ndim;         // > 1: nested arrays of a fixed shape (head = ndim brackets, every comma a leaf)
    tk_plan_reset   reset;
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
    tk_pdl_launch_dependents();
    tk_pdl_wait();
    if (blockIdx.x == 0 && threadIdx.x == 0) *a.tok_count_next = 0;
    tk_plan_reset_run(a.reset, blockIdx.x * blockDim.x + threadIdx.x, gridDim.x * blockDim.x);
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    uint2* tok = reinterpret_cast<uint2*>(a.tok);
    for (uint32_t i = warp; i < a.n; i += nwarps) {
        const uint32_t start = a.recs[i].start, end = a.recs[i + 1].start;
        // records the record-batch split flagged (tk_recordbatch.cu): skipped ones were never
        // consumed, broken ones carry their batch's error; neither has a payload to look at
        const int32_t pflags = a.recs[i].partition & TK_REC_FLAGS;
        if (pflags) {
            if (lane == 0) {
                a.rec_len[i] = 0;
                a.rec_status[i] = (pflags & TK_REC_OVERFLOW) == TK_REC_OVERFLOW ? TK_STATUS_SLOT_OVERFLOW : (pflags & TK_REC_CORRUPT) ? TK_STATUS_CORRUPT
                                : (pflags & TK_REC_UNSUPPORTED) ? TK_STATUS_UNSUPPORTED : TK_STATUS_OK;
                a.rec_kidx[i] = -1;
                a.rec_tok[i] = 0;
            }
            continue;
        }
        const uint32_t base0 = start & ~15u;
        // what the last element's entry carries: the parse kernel takes its fast lane only when the
        // record ends in ']' right behind the last number (trailing white space: scalar grammar)
        const uint32_t last_flags =
            TOK_LAST | ((end > start && __ldg(a.bytes + end - 1) == ']') ? TOK_TAIL : 0u);
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
            if (a.ndim > 1) {
                GlobalReader rd{a.bytes, end};
                head = end > start ? tk_json_prefix_nested(rd, start, a.ndim, &first) : TK_ERR_SYNTAX;
            } else if (off < 15 && start + 1 < end && (two & 0xFFu) == '[' && tk_is_numclass((int)((two >> 8) & 0xFFu))) {
                head = TK_OK;
                first = start + 1;
            } else {
                GlobalReader rd{a.bytes, end};
                head = end > start ? tk_json_prefix(rd, start, &first, &empty) : TK_ERR_SYNTAX;
            }
            len = (!head && empty) ? 0u : commas + 1u;
            const int status = head ? head : (len > a.max_len ? (a.ndim > 1 ? TK_ERR_RAGGED : TK_ERR_TOO_LONG) : TK_OK);
            a.rec_len[i] = (int32_t)len;
            a.rec_status[i] = status;
            a.rec_kidx[i] = (status == TK_OK && len >= a.min_size) ? 0 : -1;
            if (len) {
                t0 = atomicAdd(a.tok_count, len);
                if (t0 < a.tok_cap) tok[t0] = make_uint2(head ? start : first, i | (len == 1 ? last_flags : 0u));
            }
            a.rec_tok[i] = t0;
        }
        len = __shfl_sync(0xffffffffu, len, 0);
        t0 = __shfl_sync(0xffffffffu, t0, 0);
        if (len <= 1) continue;
        // ---- pass 2: element e (>= 1) starts right after the e-th comma
        uint32_t seen = 0;
        const uint32_t t_last = t0 + len - 1;
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
                if (t < a.tok_cap) tok[t] = make_uint2(pos + b + 1, i | (t == t_last ? last_flags : 0u));
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

static int json_scan_impl(tk_ctx* ctx, int slot, uint32_t max_len, uint32_t min_size, uint32_t ndim, void* stream);

extern "C" int tk_json_scan(tk_ctx* ctx, int slot, uint32_t max_len, uint32_t min_size, void* stream) {
    return json_scan_impl(ctx, slot, max_len, min_size, 0, stream);
}

// leaves per record of a [dims] array; 0 if the shape is not acceptable
static uint64_t nested_leaves(uint32_t ndim, const uint32_t* dims) {
    if (ndim < 1 || ndim > TK_JSON_NESTED_MAX_DIMS || !dims) return 0;
    uint64_t n = 1;
    for (uint32_t k = 0; k < ndim; ++k) {
        if (!dims[k]) return 0;
        n *= dims[k];
        if (n > (1u << 24)) return 0;
    }
    return n;
}

extern "C" int tk_json_nested_scan(tk_ctx* ctx, int slot, uint32_t ndim, const uint32_t* dims, void* stream) {
    const uint64_t leaves = nested_leaves(ndim, dims);
    TK_REQUIRE(leaves, "1..%d dimensions of at least 1, at most 2^24 elements", TK_JSON_NESTED_MAX_DIMS);
    return json_scan_impl(ctx, slot, (uint32_t)leaves, 0, ndim, stream);
}

static int json_scan_impl(tk_ctx* ctx, int slot, uint32_t max_len, uint32_t min_size, uint32_t ndim, void* stream) {
    TK_ENTER(ctx, slot);
    tk_slot& s = ctx->slots[slot];
    const uint32_t n = s.n_records;
    int rc = tk_ensure_token_arrays(ctx);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    scan_args a;                                           // (an empty slot still resets the plan outputs)
    a.bytes = s.d.bytes; a.recs = s.d.recs; a.n = n; a.max_len = max_len; a.min_size = min_size;
    a.rec_len = s.d.rec_len; a.rec_status = s.d.rec_status; a.rec_kidx = s.d.rec_kidx; a.rec_tok = s.d.rec_tok;
    // the counter of this use was zeroed by the previous scan of the slot (or at allocation)
    s.tok_use++;
    a.tok = s.d.tok_pos; a.tok_count = s.d.tok_count + (s.tok_use & 1u); a.tok_count_next = s.d.tok_count + ((s.tok_use + 1) & 1u);
    a.tok_cap = (uint32_t)(ctx->tok_cap > 0xFFFFFFFFull ? 0xFFFFFFFFull : ctx->tok_cap);
    a.ndim = ndim;
    s.json_row_elems = max_len;                            // k_plan writes rec_base for rows of max_len elements
    a.reset = tk_plan_reset_args(ctx, slot);
    // one warp per record; 4-warp blocks: records differ in length, and a block holds its
    // slots until its longest record is done
    const uint32_t threads = 128, wpb = threads / 32;
    const uint32_t blocks = max((n + wpb - 1) / wpb, 1u);
    TK_LAUNCH_PDL(k_json_scan, blocks, threads, st, a);
    return 0;
}

// ================================================================ flat decode path
// After k_json_scan (token array) and the compaction (row of every kept record):
// k_json_parse<T>, one persistent grid, two parts per block:
//   rows   (warp per record) reports the records the scan rejected, checks fixed_len, and writes
//          pad, length and mask of every kept record's row;
//   parse  (thread per element) parses one token each: perfectly balanced, every lane busy
//          whatever the record lengths are.
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
    const long long* rec_base;    // values index of token t of record r = rec_base[r] + t (tk_internal.h)
    tk_result_t*    res;
    const uint2*    tok;          // {byte position, record} per element
    const uint32_t* tok_count;
    uint32_t        tok_cap;
    tk_nested_shape shape;        // ndim > 1: nested arrays of a fixed shape
    tk_plane_t      values, lengths, mask;
};

// Rows part of the decode kernel: reports what the scan rejected, checks fixed_len, and writes
// pad, length and mask of every kept record's row -- store-only work, 16 bytes per lane where
// the planes allow it.  A warp takes ROWS_GROUP records at a time: their metadata comes in with
// one load per array (lanes 0..ROWS_GROUP-1), then the rows are written one after the other
// without a dependent load in between, so the warp spends its time issuing stores that drain
// while the block goes on to parse numbers.
#define ROWS_GROUP 8
template <int ELEM_BYTES>
__device__ __forceinline__ void json_rows_part(const flat_args& a, uint32_t warp, uint32_t nwarps, uint32_t lane) {
    const uint32_t B = a.res->batch_size;
    const bool vec_values = ((a.values.row_bytes | (uintptr_t)a.values.base) & 15u) == 0;
    const bool vec_mask = a.mask.base && ((a.mask.row_bytes | (uintptr_t)a.mask.base | a.max_len) & 15u) == 0;
    for (uint32_t g = warp * ROWS_GROUP; g < a.n; g += nwarps * ROWS_GROUP) {
        int32_t my_kidx = -1;
        uint32_t my_len = 0;
        if (lane < ROWS_GROUP && g + lane < a.n) {
            const uint32_t i = g + lane;
            my_kidx = a.rec_kidx[i];
            my_len = (uint32_t)a.rec_len[i];
            const int pre = a.rec_status[i];                  // head / TOO_LONG verdict of the scan
            if (pre) report_error(a.res, i, pre);
            else if (my_kidx >= 0 && a.fixed_len && my_len != a.fixed_len)
                flag_record(a.rec_status, a.res, i, TK_ERR_RAGGED);
        }
#pragma unroll 1
        for (uint32_t r = 0; r < ROWS_GROUP; ++r) {
            const int32_t kidx = __shfl_sync(0xffffffffu, my_kidx, r);
            const uint32_t len = __shfl_sync(0xffffffffu, my_len, r);
            if (kidx < 0) continue;
            json_write_row<ELEM_BYTES>(a.values, a.lengths, a.mask, a.max_len, (uint64_t)B + (uint32_t)kidx, len, lane,
                                       vec_values, vec_mask);
        }
    }
}

// Elements the fast lane does not take: the scalar grammar.  Out of line: it is the rare case and
// would otherwise sit inside the (unrolled) element loop.
// (Arguments by value: a reference to the kernel's parameter block would force a copy of it into
// local memory, and every parameter read of the element loop with it.)
// (LONG: only the wide variant of the kernel carries the long lane -- what the callee needs in
// registers the element loop of the caller spills.)
template <int KIND, bool LONG>
__device__ __noinline__ int json_element_slow(const uint8_t* bytes, uint32_t end, uint32_t pos, uint32_t e, uint32_t flags,
                                              const tk_nested_shape shape, typename tk_elem<KIND>::type* v) {
    GlobalReader rd{bytes, end};
    if (LONG && shape.ndim <= 1) {
        // the fast lane's long forms first ("0.00..." with 20 digits, exponents: tk_fl_scan_long): the
        // element ends at the next ',' or at the ']' that closes the record
        uint32_t stop = pos;
        while (stop < end && stop - pos < 31u && bytes[stop] != ',' && bytes[stop] != ']') ++stop;
        // (']' must be the record's last byte: what follows it is the grammar's business)
        if (stop < end && (bytes[stop] == ',' || stop + 1u == end)) {
            const tk_fl_num num = tk_fl_scan_long(LdgWords{bytes}, pos, stop);
            if (num.ok && tk_fl_value_long<KIND>(num, v)) return TK_OK;
        }
        return tk_json_element_flat<KIND>(rd, pos, v);
    }
    if (shape.ndim <= 1) return tk_json_element_flat<KIND>(rd, pos, v);
    // nested arrays: the brackets around leaf e follow from e and the shape
    const uint32_t n_open = e ? tk_nested_opens(shape, e) : 0u;                  // (the head took the first leaf's)
    const uint32_t n_close = flags ? shape.ndim : tk_nested_opens(shape, e + 1);
    return tk_json_element_nested<KIND>(rd, pos, n_open, n_close, flags != 0, v);
}

// One element: token entry tk (.x = byte position, .y = record index with TOK_LAST on the record's
// last element and TOK_TAIL when the record's final byte is ']'), next_pos = position of the
// element behind it in the token array.
template <int KIND, bool WIDE>
__device__ __forceinline__ void json_parse_one(const flat_args& a, const tk_fl_tables& tb, uint32_t t, uint2 tk,
                                               uint32_t next_pos, typename tk_elem<KIND>::type* values) {
    typedef typename tk_elem<KIND>::type T;
    const uint32_t pos = tk.x, rec = tk.y & TOK_REC_MASK, flags = tk.y & ~TOK_REC_MASK;
    const long long base = a.rec_base[rec];
    // the element ends at the next ',' -- or, for the last one, at the ']' that closes the record
    uint32_t stop = next_pos - 1u;
    if (flags) stop = a.recs[rec + 1].start - 1u;
    T v;
    uint64_t w;
    uint32_t K, dot, neg;
    // nested arrays: leaves without brackets around them take the same fast lane as the elements of
    // a flat array
    bool fast = flags != TOK_LAST;
    if (a.shape.ndim > 1) {
        const uint32_t e = t - a.rec_tok[rec];
        fast = fast && !((e ? tk_nested_opens(a.shape, e) : 0u) | (flags ? a.shape.ndim : tk_nested_opens(a.shape, e + 1)));
    }
    fast = fast && tk_fl_scan<WIDE>(LdgWords{a.bytes}, pos, stop, tb, &w, &K, &dot, &neg);
    if (fast) fast = tk_fl_value<KIND>(w, K, dot, neg, tb, &v);
    if (!fast) {
        const uint32_t e = a.shape.ndim > 1 ? t - a.rec_tok[rec] : 0u;
        const int st = json_element_slow<KIND, WIDE>(a.bytes, a.recs[rec + 1].start, pos, e, flags, a.shape, &v);
        if (st) { flag_record(a.rec_status, a.res, rec, st); return; }
    }
    // (a record with a row has at most max_len elements: longer ones were rejected by the scan)
    if (base != TK_REC_BASE_SKIP) values[base + (long long)t] = v;
}

// 6 blocks per SM (40 registers): measured best of 4 / 5 / 6 (profiles/r02_parse_variants.md).
// Two variants: WIDE = false, the leanest element loop, for data whose numbers the 20-byte lane
// takes (magnitudes of 0.01 and more: BASELINE's uniform(-1e3, 1e3)); WIDE = true with the "0.00..."
// bodies inline (3 us slower per 94 MB slot on such data, 15 % faster on uniform(-1, 1) or
// gauss(0, 1), where one element in a hundred is below 0.01 and would take the out-of-line call,
// 1.5 x faster on gauss(0, 1e-3)).  tk_ctx_set_json_wide chooses.
template <int KIND, bool WIDE>
__global__ void __launch_bounds__(256, 6) k_json_parse(const flat_args a) {
    typedef typename tk_elem<KIND>::type T;
    __shared__ tk_fl_tables tb;
    tk_fl_tables_fill(tb, threadIdx.x, blockDim.x);   // constants only: overlaps the tail of the previous kernel
    __syncthreads();
    tk_pdl_wait();
    tk_pdl_launch_dependents();             // after the wait: see k_commit_final (tk_plan.cu)
    json_rows_part<sizeof(T)>(a, (blockIdx.x * blockDim.x + threadIdx.x) >> 5, (gridDim.x * blockDim.x) >> 5,
                              threadIdx.x & 31);
    const uint32_t n_tok = min(*a.tok_count, a.tok_cap);
    T* values = reinterpret_cast<T*>(a.values.base);
    // Persistent grid, software-pipelined two elements deep: the token entry of iteration i + 2
    // is requested and the text of element i + 1 is prefetched into L1 while element i is parsed,
    // so neither of the two dependent DRAM round trips (entry -> text) is waited for.
    const uint32_t stride = gridDim.x * blockDim.x;
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    uint2 tk = make_uint2(0, 0), tk1 = make_uint2(0, 0), tk2 = make_uint2(0, 0);
    uint32_t np = 0, np1 = 0, np2 = 0;
    if (t < n_tok) { tk = __ldg(a.tok + t); np = __ldg(&a.tok[t + 1].x); }   // one spare entry behind the array
    if (t + stride < n_tok) { tk1 = __ldg(a.tok + t + stride); np1 = __ldg(&a.tok[t + stride + 1].x); }
#pragma unroll 1
    for (; t < n_tok; t += stride, tk = tk1, np = np1, tk1 = tk2, np1 = np2) {
        const uint32_t t2 = t + 2 * stride;
        const uint8_t* ahead = a.bytes + tk1.x;
        asm volatile("prefetch.global.L1 [%0];" ::"l"(ahead));
        asm volatile("prefetch.global.L1 [%0];" ::"l"(ahead + 27));
        if (t2 < n_tok) { tk2 = __ldg(a.tok + t2); np2 = __ldg(&a.tok[t2 + 1].x); }
        json_parse_one<KIND, WIDE>(a, tb, t, tk, np, values);
    }
}

// one wave of resident blocks (no tail wave): SMs x blocks per SM of the kernel, asked once
static uint32_t parse_grid(int dtype) {
    static uint32_t cached[4] = {0, 0, 0, 0};
    const int k = dtype - TK_DTYPE_F32;
    if (!cached[k]) {
        int per_sm = 0, sms = 0, dev = 0;
        const void* fn = k == 0 ? (const void*)k_json_parse<TK_F32, false> : k == 1 ? (const void*)k_json_parse<TK_F64, false>
                       : k == 2 ? (const void*)k_json_parse<TK_I32, false> : (const void*)k_json_parse<TK_I64, false>;
        if (cudaGetDevice(&dev) != cudaSuccess ||
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, 256, 0) != cudaSuccess || per_sm < 1 || sms < 1) {
            cudaGetLastError();
            per_sm = 6; sms = 148;
        }
        cached[k] = (uint32_t)(per_sm * sms);
    }
    return cached[k];
}

static int json_decode_impl(tk_ctx* ctx, int slot, int dtype, uint32_t max_len, uint32_t fixed_len, uint32_t ndim,
                            const uint32_t* dims, const tk_plane_t* values, const tk_plane_t* lengths,
                            const tk_plane_t* mask, void* stream);

extern "C" int tk_json_decode(tk_ctx* ctx, int slot, int dtype, uint32_t max_len, uint32_t fixed_len,
                              const tk_plane_t* values, const tk_plane_t* lengths, const tk_plane_t* mask,
                              void* stream) {
    return json_decode_impl(ctx, slot, dtype, max_len, fixed_len, 0, nullptr, values, lengths, mask, stream);
}

extern "C" int tk_json_nested_decode_flat(tk_ctx* ctx, int slot, int dtype, uint32_t ndim, const uint32_t* dims,
                                          const tk_plane_t* values, void* stream) {
    const uint64_t leaves = nested_leaves(ndim, dims);
    TK_REQUIRE(leaves, "1..%d dimensions of at least 1, at most 2^24 elements", TK_JSON_NESTED_MAX_DIMS);
    return json_decode_impl(ctx, slot, dtype, (uint32_t)leaves, (uint32_t)leaves, ndim, dims, values, nullptr, nullptr, stream);
}

static int json_decode_impl(tk_ctx* ctx, int slot, int dtype, uint32_t max_len, uint32_t fixed_len, uint32_t ndim,
                            const uint32_t* dims, const tk_plane_t* values, const tk_plane_t* lengths,
                            const tk_plane_t* mask, void* stream) {
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
    a.rec_tok = s.d.rec_tok; a.rec_base = s.d.rec_base; a.res = s.d.result;
    TK_REQUIRE(s.json_row_elems == max_len,
               "tk_json_scan + tk_compact must run before tk_json_decode, with the same max_len (scan %u, decode %u)",
               s.json_row_elems, max_len);
    a.tok = reinterpret_cast<const uint2*>(s.d.tok_pos); a.tok_count = s.d.tok_count + (s.tok_use & 1u);
    a.tok_cap = (uint32_t)(ctx->tok_cap > 0xFFFFFFFFull ? 0xFFFFFFFFull : ctx->tok_cap);
    a.shape.ndim = ndim;
    {
        uint32_t prod = 1;
        for (uint32_t j = TK_JSON_NESTED_MAX_DIMS; j-- > 0;) {
            if (j < ndim) prod *= dims[j];
            a.shape.span[j] = prod;
        }
    }
    tk_plane_t z; z.base = nullptr; z.row_bytes = 0;
    a.values = *values;
    a.lengths = lengths ? *lengths : z;
    a.mask = mask ? *mask : z;
    // the element count lives on the device; size the grid for the worst case this slot allows
    // (an element needs two bytes) and let surplus blocks exit at once
    const uint64_t max_tok = (uint64_t)s.n_bytes / 2 + n;
    // (the rows part wants a warp per ~record too: never fewer blocks than records / 8 would use)
    const uint32_t pblocks = (uint32_t)min(max((max_tok + 255) / 256, ((uint64_t)n + 7) / 8), (uint64_t)parse_grid(dtype));
    // the variant is the caller's choice (tk_ctx_set_json_wide: the Python engine samples the payload)
    if (ctx->json_wide) {
        switch (dtype) {
            case TK_DTYPE_F32: TK_LAUNCH_PDL((k_json_parse<TK_F32, true>), pblocks, 256, st, a); break;
            case TK_DTYPE_F64: TK_LAUNCH_PDL((k_json_parse<TK_F64, true>), pblocks, 256, st, a); break;
            case TK_DTYPE_I32: TK_LAUNCH_PDL((k_json_parse<TK_I32, true>), pblocks, 256, st, a); break;
            default:           TK_LAUNCH_PDL((k_json_parse<TK_I64, true>), pblocks, 256, st, a); break;
        }
    } else {
        switch (dtype) {
            case TK_DTYPE_F32: TK_LAUNCH_PDL((k_json_parse<TK_F32, false>), pblocks, 256, st, a); break;
            case TK_DTYPE_F64: TK_LAUNCH_PDL((k_json_parse<TK_F64, false>), pblocks, 256, st, a); break;
            case TK_DTYPE_I32: TK_LAUNCH_PDL((k_json_parse<TK_I32, false>), pblocks, 256, st, a); break;
            default:           TK_LAUNCH_PDL((k_json_parse<TK_I64, false>), pblocks, 256, st, a); break;
        }
    }
    return 0;
}
