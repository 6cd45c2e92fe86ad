// tk_internal.h -- structures shared by the translation units of libtorchkafka_b200.so
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/torchkafka_b200.h"

// ---------------------------------------------------------------- error channel
void tk_set_error(const char* fmt, ...);

#define TK_CUDA(call)                                                                     \
    do {                                                                                  \
        cudaError_t e__ = (call);                                                         \
        if (e__ != cudaSuccess) {                                                         \
            tk_set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, \
                         __LINE__);                                                       \
            return -1;                                                                    \
        }                                                                                 \
    } while (0)

#define TK_REQUIRE(cond, ...)                 \
    do {                                      \
        if (!(cond)) {                        \
            tk_set_error(__VA_ARGS__);        \
            return -2;                        \
        }                                     \
    } while (0)

// ---------------------------------------------------------------- device state
// Persistent across slots: the open batch and the per-partition positions.
struct tk_state {
    uint32_t carry_rows;   // rows of the batch left open by the previous slot
    uint32_t pad;
};

// One RecordBatch of a raw slot (tk_recordbatch.cu)
struct tk_rb_batch {
    uint32_t start, end;       // byte range of the batch in the raw buffer
    int32_t  partition;
    uint32_t status;           // tk_status_t: the batch could not be split
    int64_t  base_offset;      // + blob offset_delta
    int64_t  fetch_offset;
    uint32_t n_rec;            // records (1 pseudo-record if status != 0)
    uint32_t value_bytes;      // bytes the batch contributes to the slot
    uint32_t rec_base;         // first record / first value byte of the batch in the slot
    uint32_t byte_base;
    uint32_t codec;            // attributes & 7: 0 plain, 2 snappy / 3 LZ4 (decompressed on the device), others unsupported
    uint32_t rec_start, rec_end;   // the records: [start + 61, end) of the raw bytes, or the batch's
                                   // region of the decompressed buffer
    uint32_t unit_first, unit_count;   // its decompression units (tk_rb.h)
    uint32_t reserved;
};
struct tk_rb_unit;

// Per-slot device scratch ("plan"): everything the kernels hand to one another.
struct tk_slot_dev {
    uint8_t*     bytes;        // [slot_bytes + pad]
    tk_rec_t*    recs;         // [slot_records + 1]
    int32_t*     rec_len;      // element count (JSON) / byte count (raw)
    int32_t*     rec_status;   // tk_status_t
    int32_t*     rec_kidx;     // index among kept records of this slot, -1 if skipped
    // single-pass compaction scan (tk_plan.cu): per tile of CMP_TILE records the packed sum
    // (kept << 40 | elements) of the tile alone and of everything up to and including it, a
    // status word (epoch << 2 | 1: tile sum valid, | 2: inclusive sum valid), and two counters
    // (tile ticket, finished tiles) that the last block resets
    unsigned long long* tile_agg;
    unsigned long long* tile_incl;
    uint32_t*    tile_flag;
    uint32_t*    plan_ctr;
    uint32_t*    rec_tok;      // [slot_records] index of each record's first element in the token array
    long long*   rec_base;     // [slot_records] flat JSON decode: (output row of the record) * max_len - rec_tok,
                               // so that element t of the token array goes to values[rec_base + t];
                               // TK_REC_BASE_SKIP for a record without a row (written by k_plan)
    uint32_t*    tok_pos;      // [tok_cap] token array of the flat decode path: uint2 {byte position, record}
    uint32_t*    tok_count;    // [2] elements emitted so far (token array allocation cursor): two
                               // counters used alternately, each scan zeroes the one of the next
    // single-pass JSON kernel (tk_json_fused.cu; allocated on first use): first record of every
    // tile of the byte buffer, look-back arrays over the tiles, counters (ticket, finished, exited)
    uint32_t*    fz_tile_first;
    unsigned long long* fz_agg;
    unsigned long long* fz_incl;
    uint32_t*    fz_flag;
    uint32_t*    fz_ctr;
    // raw record-batch input (allocated on first tk_slot_submit_recordbatches)
    uint8_t*     raw;          // [slot_bytes + pad]
    uint8_t*     raw2;         // [raw2_bytes + pad] decompressed records of compressed batches (first use)
    uint64_t     raw2_bytes;   // slot_bytes, more once zstd units were placed by their upper bounds
    tk_rb_batch* rb_batches;   // [slot_records]
    tk_rb_unit*  rb_units;     // decompression units (first compressed batch)
    tk_rb_summary_t* rb_summary;
    uint4*       rb_temp;      // [slot_records] per-record copy descriptors
    uint32_t*    rb_bigq;      // records copied piecewise
    uint8_t*     block;        // one allocation: header | batch_end | batch_lmax | commit
    tk_result_t* result;       // header
    uint32_t*    batch_end;    // [max_batches + 1]
    uint32_t*    batch_lmax;   // [max_batches + 1]
    int64_t*     commit;       // [n_partitions][commit_pitch]
};

struct tk_slot {
    // host pinned staging
    uint8_t*  h_bytes;
    tk_rec_t* h_recs;
    // host pinned result block (one allocation mirroring d.block; the pointers below point into it)
    uint8_t*     h_block;
    tk_result_t* h_result;
    uint32_t*    h_batch_end;
    uint32_t*    h_batch_lmax;
    int64_t*     h_commit;
    tk_slot_dev  d;
    uint32_t     n_records;
    uint64_t     n_bytes;
    cudaEvent_t  ev_h2d;      // H2D landed
    cudaEvent_t  ev_free;     // kernels finished reading the device copy
    cudaEvent_t  ev_result;   // result block is in pinned memory
    cudaEvent_t  ev_raw;      // raw record-batch bytes landed (created with the raw buffers)
    cudaStream_t split_stream;   // this slot's decompression + split kernels: they overlap the next
                                 // slot's H2D and the kernels of the other slots (latency-bound decoders)
    tk_rb_batch* h_rb_batches;   // pinned, [slot_records]: written by the host's header walk
    tk_rb_unit*  h_rb_units;     // pinned: decompression units of compressed batches
    tk_rb_summary_t* h_rb_summary;   // pinned
    cudaEvent_t  ev_dec0, ev_dec1;   // timing events around the decode / collate kernels (tk_ctx_set_timing)
    bool         dec_timed;
    uint32_t     tok_use;     // tk_json_scan calls on this slot: selects the token counter
    uint32_t     json_row_elems;   // max_len of the slot's last scan if that was tk_json_scan (k_plan then
                                   // writes rec_base), 0 after any other scan
    uint32_t     plan_epoch;  // tk_compact calls on this slot: tags the tile status words
    uint32_t     fz_epoch;    // tk_json_ingest calls on this slot: tags the fused kernel's tile status words
    bool         free_recorded;
};

struct tk_ctx {
    int       device;
    uint64_t  slot_bytes;
    uint32_t  slot_records;
    int       n_slots;
    uint32_t  max_batches;     // batches one slot can close (+1 open)
    uint32_t  commit_pitch;    // max_batches + 1 rounded up to 4
    uint32_t  n_partitions;
    size_t    block_bytes;     // size of a slot's result block
    size_t    tok_cap;         // entries of the token array (a token needs >= 2 bytes)
    size_t    fz_tiles;        // entries of the fused kernel's tile arrays (0: not allocated)
    int       json_fused;      // 1: tk_process_slot runs JSON arrays through the single-pass kernel (tk_json_fused.cu)
    int       json_wide;       // 1: k_json_parse with the "0.00..." bodies inline (tk_ctx_set_json_wide)
    uint32_t  batch_size;      // of the last tk_compact
    cudaStream_t copy_stream;
    tk_state* d_state;
    int64_t*  d_positions;     // [n_partitions] running max(offset)+1, -1 = never seen
    uint16_t* d_image_lut;     // [4][256] bf16 bit patterns
    int       image_channels;
    int       image_fma;       // 1: fused multiply-add reproduces the exact table for all 256 x C inputs
    float     image_a[4], image_b[4];   // y = fma(float(x), a[c], b[c])
    int       timing;          // tk_ctx_set_timing
    uint8_t*  d_obj_table;     // JSON-object schema: name offsets [65] then the names (tk_json_object.cu)
    uint32_t  obj_fields;      // 0: no field table set
    uint32_t  obj_name_bytes;
    tk_slot*  slots;
};

static inline int tk_check_slot(tk_ctx* ctx, int slot) {
    if (!ctx) { tk_set_error("null context"); return -2; }
    if (slot < 0 || slot >= ctx->n_slots) { tk_set_error("slot %d out of range [0,%d)", slot, ctx->n_slots); return -2; }
    return 0;
}

#define TK_ENTER(ctx, slot)                                   \
    do {                                                      \
        int rc__ = tk_check_slot((ctx), (slot));              \
        if (rc__) return rc__;                                \
        TK_CUDA(cudaSetDevice((ctx)->device));                \
    } while (0)

#define TK_LAUNCH_CHECK() TK_CUDA(cudaGetLastError())

// Programmatic dependent launch (sm_90+).  The kernels of a slot form a chain on one stream, each
// a few microseconds to a few tens of microseconds long, so the launch latency between them is
// a visible share of the slot.  Every kernel of the chain
//   - calls tk_pdl_launch_dependents() first: the next kernel of the stream may be scheduled (its
//     blocks become resident, read their parameters, run any prologue that touches no data of
//     earlier kernels) as soon as every block of this one has started;
//   - calls tk_pdl_wait() before its first access to global memory: returns once every earlier
//     kernel of the chain has completed and flushed its writes (transitively: the predecessor
//     itself only completes after its own wait).
// Both are no-ops in a kernel launched without the attribute.
#if defined(__CUDACC__)
__device__ __forceinline__ void tk_pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;"); }
__device__ __forceinline__ void tk_pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// Outputs of the plan kernel (tk_plan.cu) that accumulate with atomics start from a clean state:
// commit matrix -1 ("partition not seen"), per-batch longest record 0, closing record n - 1.
// The scan kernels do this (grid-stride, every thread a few entries) because they run before
// the plan and every thread of a large grid is at hand; tk_compact relies on it.
struct tk_plan_reset {
    int64_t*  commit;
    uint32_t* batch_end;
    uint32_t* batch_lmax;
    uint32_t  commit_elems;    // n_partitions * commit_pitch
    uint32_t  n_batches;       // max_batches + 1
    uint32_t  last_record;     // n ? n - 1 : 0
};
__device__ __forceinline__ void tk_plan_reset_run(const tk_plan_reset& r, uint32_t gtid, uint32_t nthreads) {
    for (uint32_t x = gtid; x < r.commit_elems; x += nthreads) r.commit[x] = -1;
    for (uint32_t j = gtid; j < r.n_batches; j += nthreads) { r.batch_end[j] = r.last_record; r.batch_lmax[j] = 0; }
}

#define TK_LAUNCH_PDL(kernel, grid, block, strm__, ...)                                         \
    do {                                                                                        \
        cudaLaunchConfig_t cfg__ = {};                                                          \
        cfg__.gridDim = dim3(grid);                                                             \
        cfg__.blockDim = dim3(block);                                                           \
        cfg__.dynamicSmemBytes = 0;                                                             \
        cfg__.stream = (strm__);                                                                \
        cudaLaunchAttribute attr__[1];                                                          \
        attr__[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;                      \
        attr__[0].val.programmaticStreamSerializationAllowed = 1;                               \
        cfg__.attrs = attr__;                                                                   \
        cfg__.numAttrs = 1;                                                                     \
        TK_CUDA(cudaLaunchKernelEx(&cfg__, kernel, __VA_ARGS__));                               \
    } while (0)
#endif

int tk_ensure_token_arrays(tk_ctx* ctx);
int tk_ensure_fused(tk_ctx* ctx);
tk_plan_reset tk_plan_reset_args(tk_ctx* ctx, int slot);
// K4 second half; wait_for_prior = 0 only behind a kernel that released its dependents after its own wait
int tk_commit_final(tk_ctx* ctx, int slot, int wait_for_prior, void* stream);

// bytes the kernels may over-read past n_bytes (aligned vector loads of the last tile)
#define TK_BYTES_PAD 4096
#define TK_REC_BASE_SKIP (-0x7FFFFFFFFFFFFFFFLL - 1)
