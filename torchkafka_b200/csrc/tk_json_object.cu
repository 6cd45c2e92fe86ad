// tk_json_object.cu -- K1 + K3 for JSON objects with named numeric fields and for nested arrays of
// a fixed shape.
//
// Replaces, per record, a user hook of the shape of README.md:72-79 with a dict payload
//     d = json.loads(record.value)
//     return torch.tensor([d[k] for k in fields], dtype=...)
// and, per batch, default_collate -> torch.stack (collate.py:275): row k of the batch holds the
// named members of kept record k in schema order.  Grammar and conversions: tk_json_object.cuh.
//
// Nested arrays: `torch.tensor(json.loads(record.value), dtype=...)` of a declared shape
// (dims[0..ndim)), batches [B, *shape]; same kernel, other record grammar (tk_json_nested_record).
//
//   (scan: tk_raw_scan with TK_ANY_RECORD_BYTES -- every record is kept; compaction and commit
//    offsets: tk_plan.cu)
//   k_json_object<T>  thread per record.  An object is a chain of dependent decisions (string
//                     ends, member names, nesting), so a record is scanned sequentially; the
//                     parallelism is across records.  A warp takes the next run of consecutive
//                     records that fits its shared-memory window (they lie back to back in the
//                     slot's byte buffer), brings the run in with coalesced 16-byte loads, and
//                     every lane decodes its own record out of shared memory and writes its row:
//                     first with the fast lane of tk_object_fast.cuh (flat objects as json.dumps
//                     writes them: member by member in straight-line code, so the lanes stay in
//                     step), else with the sequential scanner.
//                     Records larger than the window are left to a second loop, one record per
//                     warp, scanned from global memory by lane 0 (through L1).
#include "tk_swar.cuh"
#include "tk_json_object.cuh"
#include "tk_object_fast.cuh"

#define OBJ_THREADS 128
#define OBJ_WARPS (OBJ_THREADS / 32)
#define OBJ_WINDOW 8192u                                   // bytes of shared memory per warp
#define OBJ_TABLE_OFFSETS ((TK_OBJ_MAX_FIELDS + 1) * 4)    // device table: offsets, then names

struct ObjSharedReader {
    const uint8_t* win;
    uint32_t base, end;
    __device__ __forceinline__ int operator()(uint32_t pos) const { return pos < end ? (int)win[pos - base] : -1; }
};
// word source of the fast lane (tk_object_fast.cuh) over the same window
struct ObjSharedWords {
    const uint8_t* win;
    uint32_t base;
    __device__ __forceinline__ uint32_t word(uint32_t a, int k) const {
        return reinterpret_cast<const uint32_t*>(win + (a - base))[k];
    }
};
struct ObjGlobalReader {
    const uint8_t* bytes;
    uint32_t end;
    __device__ __forceinline__ int operator()(uint32_t pos) const { return pos < end ? (int)bytes[pos] : -1; }
};

struct obj_args {
    const uint8_t*  bytes;
    const tk_rec_t* recs;
    uint32_t        n;
    const int32_t*  rec_kidx;
    int32_t*        rec_status;
    tk_result_t*    res;
    const uint8_t*  table;
    uint32_t        n_fields, name_bytes;
    uint32_t        ndim, dims[TK_NESTED_MAX_DIMS];   // ndim > 0: nested arrays instead of an object
    uint32_t        big_records;  // the slot may hold records larger than a window (mean size says so)
    tk_plane_t      values;
};

template <int KIND, class R>
__device__ __forceinline__ int obj_parse_record(const obj_args& a, const R& rd, uint32_t pos, const tk_obj_fields& ft,
                                                typename tk_elem<KIND>::type* row) {
    if (a.ndim) return tk_json_nested_record<KIND>(rd, pos, a.dims, a.ndim, row);
    return tk_json_object_record<KIND>(rd, pos, ft, row);
}

template <int KIND>
__global__ void __launch_bounds__(OBJ_THREADS) k_json_object(const obj_args a) {
    typedef typename tk_elem<KIND>::type T;
    // (slack behind the window: the fast lane's word reads may run 28 bytes past a record's end)
    __shared__ __align__(16) uint8_t win[OBJ_WARPS][OBJ_WINDOW + 64];
    __shared__ tk_fl_tables fl;
    tk_fl_tables_fill(fl, threadIdx.x, OBJ_THREADS);
    __shared__ uint32_t s_off[TK_OBJ_MAX_FIELDS + 1];
    __shared__ uint8_t s_names[TK_MAX_OBJECT_NAME_BYTES];
    // the field table was written by tk_ctx_set_object_fields (synchronous): not data of the chain
    for (uint32_t k = threadIdx.x; a.n_fields && k <= a.n_fields; k += OBJ_THREADS) s_off[k] = reinterpret_cast<const uint32_t*>(a.table)[k];
    for (uint32_t k = threadIdx.x; k < a.name_bytes; k += OBJ_THREADS) s_names[k] = a.table[OBJ_TABLE_OFFSETS + k];
    __syncthreads();
    tk_pdl_wait();
    tk_pdl_launch_dependents();             // after the wait: see k_commit_final (tk_plan.cu)
    const tk_obj_fields ft = {s_names, s_off, a.n_fields};
    const uint32_t B = a.res->batch_size;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * OBJ_THREADS + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * OBJ_THREADS) >> 5;
    uint8_t* w = win[threadIdx.x >> 5];
    for (uint32_t g0 = warp * 32; g0 < a.n; g0 += nwarps * 32) {
        const uint32_t gend = min(g0 + 32u, a.n);
        uint32_t g = g0;
        while (g < gend) {                                  // warp-uniform
            const uint32_t i = g + lane;
            const bool have = i < gend;
            uint32_t my_start = 0, my_end = 0;
            int32_t kidx = -1;
            if (have) {
                my_start = a.recs[i].start;
                my_end = a.recs[i + 1].start;
                kidx = a.rec_kidx[i];
            }
            // (32 bytes in front of the first record come along: the fast lane reads 20 bytes in
            // front of a number's end)
            const uint32_t first16 = __shfl_sync(0xffffffffu, my_start, 0) & ~15u;
            const uint32_t base = first16 >= 32u ? first16 - 32u : 0u;
            const uint32_t fits = __ballot_sync(0xffffffffu, have && my_end - base <= OBJ_WINDOW);
            const uint32_t k = fits == 0xffffffffu ? 32u : (uint32_t)__ffs((int)~fits) - 1u;   // leading records that fit
            int st = TK_OK;
            bool mine = false;
            if (k == 0) {
                // the first record alone exceeds the window: second loop below
                if (lane == 0 && kidx < 0) mine = true;     // (only its status is left to report)
                g += 1;
            } else {
                const uint32_t span_end = __shfl_sync(0xffffffffu, my_end, (int)k - 1);
                for (uint32_t off = lane * 16; base + off < span_end; off += 512)      // (the slot buffer is padded)
                    *reinterpret_cast<uint4*>(w + off) = __ldg(reinterpret_cast<const uint4*>(a.bytes + base + off));
                __syncwarp();
                mine = lane < k;
                if (mine && kidx >= 0) {
                    const ObjSharedReader rd = {w, base, my_end};
                    T* row = reinterpret_cast<T*>((uint8_t*)a.values.base + ((uint64_t)B + (uint32_t)kidx) * a.values.row_bytes);
                    if (a.ndim || !tk_json_object_fast<KIND>(ObjSharedWords{w, base}, rd, base, my_start, my_end, ft, fl, row))
                        st = obj_parse_record<KIND>(a, rd, my_start, ft, row);
                }
                __syncwarp();                               // the window is reused by the next run
                g += k;
            }
            if (mine) {
                if (st) { a.rec_status[i] = st; report_error(a.res, i, st); }
                else if (kidx < 0) {                        // flagged by the record-batch split
                    const int pre = a.rec_status[i];
                    if (pre) report_error(a.res, i, pre);
                }
            }
        }
    }
    // records larger than a window: one per warp
    if (a.big_records) {
        for (uint32_t i = warp; i < a.n; i += nwarps) {
            if (lane != 0) continue;
            const uint32_t start = a.recs[i].start, end = a.recs[i + 1].start;
            const int32_t kidx = a.rec_kidx[i];
            if (end - (start & ~15u) <= OBJ_WINDOW || kidx < 0) continue;
            const ObjGlobalReader rd = {a.bytes, end};
            T* row = reinterpret_cast<T*>((uint8_t*)a.values.base + ((uint64_t)B + (uint32_t)kidx) * a.values.row_bytes);
            const int st = obj_parse_record<KIND>(a, rd, start, ft, row);
            if (st) { a.rec_status[i] = st; report_error(a.res, i, st); }
        }
    }
}

// Field names of the object schema: `names` holds the UTF-8 names back to back, name f is
// [name_offsets[f], name_offsets[f + 1]).  Synchronous (set once per context, before the first slot).
extern "C" int tk_ctx_set_object_fields(tk_ctx* ctx, const uint8_t* names, const uint32_t* name_offsets,
                                        uint32_t n_fields, void* stream) {
    TK_REQUIRE(ctx, "null context");
    TK_CUDA(cudaSetDevice(ctx->device));
    TK_REQUIRE(names && name_offsets, "names / name_offsets are null");
    TK_REQUIRE(n_fields >= 1 && n_fields <= TK_OBJ_MAX_FIELDS, "1..%d fields", TK_OBJ_MAX_FIELDS);
    TK_REQUIRE(name_offsets[0] == 0, "name_offsets must start at 0");
    for (uint32_t f = 0; f < n_fields; ++f)
        TK_REQUIRE(name_offsets[f + 1] >= name_offsets[f], "name_offsets must not decrease");
    const uint32_t nb = name_offsets[n_fields];
    TK_REQUIRE(nb <= TK_MAX_OBJECT_NAME_BYTES, "field names take %u bytes, at most %d", nb, TK_MAX_OBJECT_NAME_BYTES);
    for (uint32_t f = 0; f < n_fields; ++f)
        for (uint32_t g = f + 1; g < n_fields; ++g) {
            const uint32_t lf = name_offsets[f + 1] - name_offsets[f], lg = name_offsets[g + 1] - name_offsets[g];
            TK_REQUIRE(lf != lg || memcmp(names + name_offsets[f], names + name_offsets[g], lf) != 0,
                       "fields %u and %u have the same name", f, g);
        }
    uint8_t host[OBJ_TABLE_OFFSETS + TK_MAX_OBJECT_NAME_BYTES];
    memset(host, 0, sizeof(host));
    memcpy(host, name_offsets, sizeof(uint32_t) * (n_fields + 1));
    memcpy(host + OBJ_TABLE_OFFSETS, names, nb);
    cudaStream_t st = (cudaStream_t)stream;
    TK_CUDA(cudaStreamSynchronize(st));                      // kernels still reading the previous table
    if (!ctx->d_obj_table) TK_CUDA(cudaMalloc(&ctx->d_obj_table, sizeof(host)));
    TK_CUDA(cudaMemcpyAsync(ctx->d_obj_table, host, sizeof(host), cudaMemcpyHostToDevice, st));
    TK_CUDA(cudaStreamSynchronize(st));
    ctx->obj_fields = n_fields;
    ctx->obj_name_bytes = nb;
    return 0;
}

static int obj_launch(tk_ctx* ctx, int slot, int dtype, const tk_plane_t* values, uint32_t row_elems, uint32_t ndim,
                      const uint32_t* dims, void* stream) {
    tk_slot& s = ctx->slots[slot];
    TK_REQUIRE(values && values->base, "values plane is required");
    TK_REQUIRE(dtype >= TK_DTYPE_F32 && dtype <= TK_DTYPE_I64, "unknown dtype %d", dtype);
    const uint32_t esize = (dtype == TK_DTYPE_F32 || dtype == TK_DTYPE_I32) ? 4u : 8u;
    TK_REQUIRE(values->row_bytes == (uint64_t)row_elems * esize && ((uintptr_t)values->base & (esize - 1)) == 0,
               "values rows must hold %u elements of %u bytes", row_elems, esize);
    const uint32_t n = s.n_records;
    if (!n) return 0;
    obj_args a;
    a.bytes = s.d.bytes; a.recs = s.d.recs; a.n = n;
    a.rec_kidx = s.d.rec_kidx; a.rec_status = s.d.rec_status; a.res = s.d.result;
    a.table = ctx->d_obj_table; a.n_fields = ndim ? 0u : ctx->obj_fields;
    a.name_bytes = ndim ? 0u : ctx->obj_name_bytes;
    a.ndim = ndim;
    for (uint32_t k = 0; k < TK_NESTED_MAX_DIMS; ++k) a.dims[k] = k < ndim ? dims[k] : 1u;
    a.values = *values;
    // a warp per 32 records -- or, when records can exceed the window, per window of bytes
    const uint64_t by_bytes = s.n_bytes / OBJ_WINDOW + 1;
    a.big_records = s.n_bytes > OBJ_WINDOW ? 1u : 0u;
    const uint64_t warps = max((uint64_t)(n + 31) / 32, min(by_bytes, (uint64_t)n));
    const uint32_t blocks = (uint32_t)min((warps + OBJ_WARPS - 1) / OBJ_WARPS, (uint64_t)(148u * 8u));
    cudaStream_t st = (cudaStream_t)stream;
    switch (dtype) {
        case TK_DTYPE_F32: TK_LAUNCH_PDL(k_json_object<TK_F32>, blocks, OBJ_THREADS, st, a); break;
        case TK_DTYPE_F64: TK_LAUNCH_PDL(k_json_object<TK_F64>, blocks, OBJ_THREADS, st, a); break;
        case TK_DTYPE_I32: TK_LAUNCH_PDL(k_json_object<TK_I32>, blocks, OBJ_THREADS, st, a); break;
        default:           TK_LAUNCH_PDL(k_json_object<TK_I64>, blocks, OBJ_THREADS, st, a); break;
    }
    return 0;
}

extern "C" int tk_json_object_decode(tk_ctx* ctx, int slot, int dtype, const tk_plane_t* values, void* stream) {
    TK_ENTER(ctx, slot);
    TK_REQUIRE(ctx->obj_fields, "tk_ctx_set_object_fields must be called first");
    return obj_launch(ctx, slot, dtype, values, ctx->obj_fields, 0, nullptr, stream);
}

extern "C" int tk_json_nested_decode(tk_ctx* ctx, int slot, int dtype, uint32_t ndim, const uint32_t* dims,
                                     const tk_plane_t* values, void* stream) {
    TK_ENTER(ctx, slot);
    TK_REQUIRE(ndim >= 1 && ndim <= TK_NESTED_MAX_DIMS && dims, "1..%d dimensions", TK_NESTED_MAX_DIMS);
    uint64_t elems = 1;
    for (uint32_t k = 0; k < ndim; ++k) {
        TK_REQUIRE(dims[k] >= 1, "dimension %u is empty", k);
        elems *= dims[k];
        TK_REQUIRE(elems <= (1u << 24), "more than 2^24 elements per record");
    }
    return obj_launch(ctx, slot, dtype, values, (uint32_t)elems, ndim, dims, stream);
}
