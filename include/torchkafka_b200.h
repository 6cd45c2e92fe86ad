/* torchkafka_b200.h -- C ABI of the B200 ingest library (libtorchkafka_b200.so).
 *
 * The reference (Bendabir/torch-kafka) is pure Python and has no native boundary
 * at all; its per-record hot path is
 *
 *     KafkaDataset.__iter__        src/kafka_dataset.py:147-171
 *       _process(record)           src/kafka_dataset.py:173-186, bodies README.md:41-43,72-79
 *       None-skip filter           src/kafka_dataset.py:161-162
 *     DataLoader fetch + collate   torch/utils/data/_utils/fetch.py:27-45,
 *                                  torch/utils/data/_utils/collate.py:246-275
 *     auto_commit -> commit()      src/auto_commit.py:55-58, src/kafka_dataset.py:120-145
 *
 * Each entry point below names the reference lines it replaces.  All functions
 * are plain C: POD arguments, raw host/device pointers, an int return code
 * (0 = ok, negative = error, text via tk_last_error()).  No C++ exception and no
 * torch type crosses this boundary; CUDA streams are passed as void*
 * (cudaStream_t).  Output tensors are allocated by the caller (PyTorch) and
 * passed as device pointers.
 *
 * Data layout of one staging slot (host pinned copy and device copy alike):
 *     bytes[n_bytes]        record values back to back (CSR payload)
 *     recs[n_records + 1]   tk_rec_t: byte offset of the value, partition, offset;
 *                           recs[n_records].start == n_bytes closes the CSR
 * Threading: one producer thread may fill/submit slots while one consumer thread
 * waits/launches/fetches; a slot is owned by one side at a time.
 */
#ifndef TORCHKAFKA_B200_H
#define TORCHKAFKA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TK_ABI_VERSION 1

typedef struct tk_ctx tk_ctx;

/* One record header, 16 bytes.  (.partition, .offset) are the ConsumerRecord
 * fields the reference reads implicitly through consumer.commit()
 * (src/kafka_dataset.py:130); .start indexes into the slot's byte buffer. */
typedef struct tk_rec {
    uint32_t start;
    int32_t  partition;   /* dense partition index, 0 <= partition < n_partitions */
    int64_t  offset;
} tk_rec_t;

/* element kinds of JSON schemas */
typedef enum { TK_DTYPE_F32 = 0, TK_DTYPE_F64 = 1, TK_DTYPE_I32 = 2, TK_DTYPE_I64 = 3 } tk_dtype_t;

/* per-record status (what _process would have raised in the reference) */
typedef enum {
    TK_STATUS_OK = 0,
    TK_STATUS_SYNTAX = 1,       /* json.JSONDecodeError */
    TK_STATUS_OVERFLOW = 2,     /* torch.tensor overflow / OverflowError */
    TK_STATUS_TOO_LONG = 3,     /* more than max_len elements */
    TK_STATUS_UNSUPPORTED = 4,  /* valid JSON outside the numeric-array schema */
    TK_STATUS_SIZE = 5,         /* raw record with the wrong byte count */
    TK_STATUS_RAGGED = 6,       /* fixed_len schema, other length (torch.stack raises) */
    TK_STATUS_CORRUPT = 7,      /* record batch failed its CRC or framing (kafka CorruptRecordException) */
    TK_STATUS_MISSING_KEY = 8,  /* JSON object without one of the named fields (KeyError) */
    TK_STATUS_SLOT_OVERFLOW = 9 /* record batches whose (decompressed) values do not fit the slot's byte buffer */
} tk_status_t;

/* Per-slot result header, written by the kernels, fetched with tk_slot_fetch_result. */
typedef struct tk_result {
    uint32_t n_records;          /* records in the slot */
    uint32_t n_kept;             /* records that passed the filter (K2) */
    uint32_t carry_in;           /* rows the open batch held before this slot */
    uint32_t carry_out;          /* rows the open batch holds after this slot */
    uint32_t n_closed;           /* batches completed by this slot */
    uint32_t batch_size;
    uint32_t n_errors;           /* records with status != OK */
    uint32_t first_error_rec;    /* lowest such record index, 0xFFFFFFFF if none */
    uint32_t first_error_status;
    uint32_t reserved[7];
} tk_result_t;

/* One output plane (values / lengths / mask) of the tensor a slot writes into.
 * The tensor holds batch_size "head" rows followed by one "body" row per kept
 * record of the slot: kept record k lands in row batch_size + k; the rows of the
 * batch left open by the previous slot are copied into the end of the head region
 * (tk_carry_rows), so every batch is one contiguous row range.  See DESIGN.md
 * "Output layout". */
typedef struct tk_plane {
    void*    base;
    uint64_t row_bytes;
} tk_plane_t;

const char* tk_last_error(void);
int tk_abi_version(void);

/* ---- context, staging slots, H2D (K0; no reference counterpart: the reference
 *      never leaves host memory) ------------------------------ */
int tk_ctx_create(int device, uint64_t slot_bytes, uint32_t slot_records, int n_slots,
                  uint32_t max_batches, uint32_t n_partitions, tk_ctx** out);
int tk_ctx_destroy(tk_ctx* ctx);
/* forget the open batch and all partition positions (new iteration of the dataset) */
int tk_ctx_reset(tk_ctx* ctx, void* stream);
int tk_slot_host_ptrs(tk_ctx* ctx, int slot, uint8_t** bytes, tk_rec_t** recs);
int tk_slot_device_ptrs(tk_ctx* ctx, int slot, uint8_t** d_bytes, tk_rec_t** d_recs);
/* async copy of the filled slot on the library's copy stream; returns at once */
int tk_slot_submit(tk_ctx* ctx, int slot, uint32_t n_records, uint64_t n_bytes);
/* declare a slot whose device copy was filled by other means (benchmarks: data already in HBM) */
int tk_slot_assume_resident(tk_ctx* ctx, int slot, uint32_t n_records, uint64_t n_bytes);
/* make `stream` wait until the slot's H2D copy has landed */
int tk_slot_wait(tk_ctx* ctx, int slot, void* stream);
/* pack helper: copy `n` values given as (pointer, length) pairs into the slot,
 * multi-threaded, and fill recs[] (replaces the per-record Python loop of
 * src/kafka_dataset.py:156 on the host side) */
int tk_slot_pack(tk_ctx* ctx, int slot, uint32_t first_record, uint64_t first_byte,
                 const uint8_t* const* values, const uint32_t* lengths,
                 const int32_t* partitions, const int64_t* offsets, uint32_t n,
                 int n_threads, uint64_t* bytes_after);
/* pack helper for a run that is already contiguous: values[starts[0] .. starts[n]) is copied
 * (multi-threaded) behind first_byte and recs[first_record ..] are filled from the arrays */
int tk_slot_fill(tk_ctx* ctx, int slot, uint32_t first_record, uint64_t first_byte,
                 const uint8_t* values, const uint32_t* starts, const int32_t* partitions,
                 const int64_t* offsets, uint32_t n, int n_threads, uint64_t* bytes_after);
int tk_h2d_peak_gbs(int device, uint64_t bytes, int iters, double* gbs_out);
/* pin the pack helper's threads to the given cores (n = 0: no pinning): the staging slots should be
 * filled by cores of the GPU's NUMA node (hard part 1) */
int tk_pack_set_affinity(const int* cpus, int n);

/* ---- K1a/K1b phase 1: element count per record + keep flag -----------------
 * replaces `len(json.loads(record.value)) < min_size -> None`
 * (README.md:74-77; src/kafka_dataset.py:161-162) */
int tk_json_scan(tk_ctx* ctx, int slot, uint32_t max_len, uint32_t min_size, void* stream);
/* raw payloads: every record is kept; wrong byte count -> TK_STATUS_SIZE
 * (record_bytes == TK_ANY_RECORD_BYTES: any size; the scan of the JSON-object schema) */
#define TK_ANY_RECORD_BYTES 0xFFFFFFFFFFFFFFFFull
int tk_raw_scan(tk_ctx* ctx, int slot, uint64_t record_bytes, void* stream);

/* ---- K2: stream compaction of the keep flags -> row of every kept record,
 *      batch boundaries (a batch closes on its batch_size-th kept record)
 * replaces `if data is not None: yield data` (src/kafka_dataset.py:161-162) and
 * the batch_size loop of _IterableDatasetFetcher.fetch (fetch.py:33-38) */
/* (the same pass also accumulates, per (batch, partition), the largest offset + 1 it sees:
 * the segmented-max half of K4, done here because the batch index of each record is at hand)
 * Call order per slot pass: one scan (tk_json_scan or tk_raw_scan -- it also resets the
 * per-batch outputs this call accumulates into), then tk_compact, then tk_commit_offsets and
 * the decode / collate call in either order. */
int tk_compact(tk_ctx* ctx, int slot, uint32_t batch_size, void* stream);

/* ---- K4: per batch, per partition: max(offset)+1 over every record consumed up
 *      to the record that closed the batch (segmented max, see tk_compact, + running max
 *      along the batches and across slots, done here; call after tk_compact)
 * replaces KafkaConsumer.commit() position bookkeeping reached from
 * src/auto_commit.py:58 -> src/kafka_dataset.py:130 */
int tk_commit_offsets(tk_ctx* ctx, int slot, void* stream);

/* ---- K3 helper: move the rows of the still-open batch from the tensor of the
 *      previous slot to the head region of the tensor of the current slot, so the
 *      batch that straddles the two slots is contiguous.  Reads the row counts
 *      from the previous slot's device-side result (no host round trip).
 * replaces the list `data` that fetch.py:33-38 keeps filling across consumer polls */
int tk_carry_rows(tk_ctx* ctx, int prev_slot, int n_planes, const tk_plane_t* prev,
                  const tk_plane_t* cur, void* stream);

/* ---- K1a/K1b phase 2 + K3: parse, convert, collate, pad, mask --------------
 * replaces `torch.tensor(json.loads(record.value), dtype=...)` (README.md:74) and
 * default_collate/torch.stack (collate.py:275) or the pad collate of SURVEY.md §8c.
 * planes: [0] values (row = max_len elements), [1] lengths (int32, may be NULL
 * pointers), [2] mask (bool[max_len], may be NULL pointers) */
int tk_json_decode(tk_ctx* ctx, int slot, int dtype, uint32_t max_len, uint32_t fixed_len,
                   const tk_plane_t* values, const tk_plane_t* lengths, const tk_plane_t* mask,
                   void* stream);

/* ---- the JSON array path of a slot in one pass over the payload ---------------
 * Same reference lines as tk_json_scan + tk_compact + tk_json_decode together (README.md:74-77,
 * src/kafka_dataset.py:161-162, fetch.py:27-45, collate.py:275): element counts, keep flags,
 * compaction, batch boundaries, the per-(batch, partition) commit maxima, pad / length / mask rows
 * and the parsed values, by one persistent kernel that brings every tile of the byte buffer into
 * shared memory once (bulk asynchronous copy) and chains the tiles' kept counts with a decoupled
 * look-back.  Follow with tk_commit_offsets and tk_slot_fetch_result.  HBM traffic is the algorithmic
 * minimum (payload read once), but the kernel is bound by its per-tile synchronisation and is the
 * slower of the two paths today; tk_process_slot runs the three-kernel chain for TK_SCHEMA_JSON unless
 * tk_ctx_set_json_fused(ctx, 1) (or TK_JSON_FUSED=1 in the environment when the context is created). */
int tk_json_ingest(tk_ctx* ctx, int slot, int dtype, uint32_t max_len, uint32_t min_size,
                   uint32_t fixed_len, uint32_t batch_size, const tk_plane_t* values,
                   const tk_plane_t* lengths, const tk_plane_t* mask, void* stream);
int tk_ctx_set_json_fused(tk_ctx* ctx, int enabled);

/* Variant of the parse kernel of tk_json_decode (the per-element part of README.md:74,
 * json.loads, and README.md:41-43, torch.tensor).  0 (default; TK_JSON_WIDE=1 in the environment
 * changes it): the leanest element loop -- number literals of at most 20 bytes are decoded inline,
 * "0.00..." literals of 21 / 22 bytes and literals with an exponent by an out-of-line call.  1: the
 * 21 / 22-byte literals inline as well: a few percent slower on data without them, faster as soon as
 * one number in a few hundred is below 0.01.  Results are identical.  The Python engine chooses from a
 * sample of the payload. */
int tk_ctx_set_json_wide(tk_ctx* ctx, int enabled);

/* ---- K1c + K3: fixed-layout bytes -> rows (torch.frombuffer + torch.stack) -- */
int tk_collate_raw(tk_ctx* ctx, int slot, uint64_t record_bytes, const tk_plane_t* out,
                   void* stream);
/* fixed-layout struct records: bytes [field_offset, field_offset + field_bytes) of
 * every kept record -> row of `out` (row_bytes == field_bytes); one call per field.  What a
 * `_process` returning a tuple of torch.frombuffer views + default_collate's per-field
 * torch.stack (collate.py:275 via the tuple branch) compute. */
int tk_collate_field(tk_ctx* ctx, int slot, uint64_t record_bytes, uint32_t field_offset, uint32_t field_bytes,
                     const tk_plane_t* out, void* stream);
/* ---- K1 + K3: JSON objects with named numeric fields ---------
 * replaces a `_process` of the shape of README.md:72-79 over a dict payload,
 *     d = json.loads(record.value); torch.tensor([d[k] for k in fields], dtype=...)
 * and default_collate/torch.stack (collate.py:275): row = the named members in schema order.
 * The whole record must be what json.loads accepts (other members may hold any JSON value);
 * a named member must be a number or true/false, the last of duplicate members counts,
 * a missing one is TK_STATUS_MISSING_KEY.  Reported as TK_STATUS_UNSUPPORTED rather than
 * decoded: member names written with backslash escapes, nesting deeper than 32, UTF-16/32.
 * tk_ctx_set_object_fields (synchronous, once per context): `names` = the UTF-8 field names
 * back to back, name f = [name_offsets[f], name_offsets[f + 1]).  Scan: tk_raw_scan with
 * TK_ANY_RECORD_BYTES.  values: row_bytes = n_fields * element size. */
#define TK_MAX_OBJECT_FIELDS 64
#define TK_MAX_OBJECT_NAME_BYTES 2048
int tk_ctx_set_object_fields(tk_ctx* ctx, const uint8_t* names, const uint32_t* name_offsets,
                             uint32_t n_fields, void* stream);
int tk_json_object_decode(tk_ctx* ctx, int slot, int dtype, const tk_plane_t* values, void* stream);
/* nested arrays of a fixed shape: the record is arrays nested ndim deep with
 * extents dims[0..ndim), numbers at the leaves -- `torch.tensor(json.loads(record.value),
 * dtype=...)` of that shape; row = the leaves in C order (row_bytes = prod(dims) * element
 * size), batches [B, *dims] by torch.stack.  Another depth or extent: TK_STATUS_RAGGED.
 * Scan: tk_raw_scan with TK_ANY_RECORD_BYTES. */
#define TK_MAX_NESTED_DIMS 4
int tk_json_nested_decode(tk_ctx* ctx, int slot, int dtype, uint32_t ndim, const uint32_t* dims,
                          const tk_plane_t* values, void* stream);
/* The same contract through the flat chain (what tk_process_slot runs for two or more dimensions):
 * every comma of such a record separates two leaves, so the comma split of tk_json_scan cuts it into
 * one segment per leaf, and the brackets around leaf e follow from e and the shape -- each leaf's
 * thread checks its own.  Large records (a 64 x 24 matrix is 30 KB of text) are decoded by as many
 * threads as they have leaves instead of one.  Scan: tk_json_nested_scan (in place of tk_raw_scan),
 * then tk_compact, then tk_json_nested_decode_flat. */
int tk_json_nested_scan(tk_ctx* ctx, int slot, uint32_t ndim, const uint32_t* dims, void* stream);
int tk_json_nested_decode_flat(tk_ctx* ctx, int slot, int dtype, uint32_t ndim, const uint32_t* dims,
                               const tk_plane_t* values, void* stream);

/* HWC uint8 -> normalised bfloat16 CHW:  bf16(((x / 255) - mean[c]) / std[c]) in
 * float32 with IEEE division, the op order of SURVEY.md §8c */
int tk_image_set_norm(tk_ctx* ctx, int channels, const float* mean, const float* std, void* stream);
int tk_collate_image_u8_bf16_nchw(tk_ctx* ctx, int slot, uint32_t height, uint32_t width,
                                  uint32_t channels, const tk_plane_t* out, void* stream);

/* ---- the whole hot path of one slot in one call ------------------------------
 * tk_slot_wait, tk_carry_rows, scan, tk_compact, tk_commit_offsets, decode/collate,
 * tk_slot_fetch_result, enqueued on `stream`.  What a host plugin binds in place of
 * the per-record loop KafkaDataset.__iter__ (src/kafka_dataset.py:147-171) + the
 * DataLoader fetch/collate (fetch.py:27-45) + auto_commit's commit (auto_commit.py:55-58). */
typedef enum { TK_SCHEMA_JSON = 0, TK_SCHEMA_RAW = 1, TK_SCHEMA_IMAGE = 2, TK_SCHEMA_STRUCT = 3,
               TK_SCHEMA_OBJECT = 4, TK_SCHEMA_NESTED = 5 } tk_schema_t;
#define TK_MAX_PLANES 8
typedef struct tk_job {
    int32_t    schema;        /* tk_schema_t */
    int32_t    dtype;         /* JSON array / object: tk_dtype_t */
    uint32_t   max_len, min_size, fixed_len;            /* JSON arrays; nested arrays: max_len = number of
                                                           dimensions, their extents in field_bytes[] */
    uint32_t   height, width, channels;                 /* image */
    uint64_t   record_bytes;                            /* raw / image input bytes per record */
    uint32_t   batch_size;
    int32_t    prev_slot;     /* slot processed just before this one, -1 if none */
    int32_t    n_planes;      /* 1 (values), 3 (values, lengths, mask), or one per struct field */
    int32_t    resident;      /* non-zero: device copy already valid, skip the H2D wait */
    tk_plane_t prev[TK_MAX_PLANES];       /* planes of the previous slot's tensors (carry source) */
    tk_plane_t cur[TK_MAX_PLANES];        /* planes of this slot's tensors */
    uint32_t   field_offset[TK_MAX_PLANES];   /* struct: byte offset of field k inside the record ... */
    uint32_t   field_bytes[TK_MAX_PLANES];    /* ... and its size (== row_bytes of plane k) */
} tk_job_t;
int tk_process_slot(tk_ctx* ctx, int slot, const tk_job_t* job, void* stream);

/* ---- K2 without any allocation ------------------------------------------------
 * No counterpart in the reference (it has no device code and no allocator to avoid).
 * Same computation as tk_compact -- from per-record keep flags (>= 0 kept, < 0 skipped)
 * to the index of every record among the kept ones (k for kept records, -1 - k for
 * skipped ones, k = kept records before it) plus the batch bookkeeping in *d_result
 * (n_records, n_kept, carry_in, carry_out, n_closed, batch_size) -- for callers that
 * cannot call cudaMalloc where they need it: hosts whose whole HBM budget is already
 * owned by their own pool, plugin threads on which driver allocations are forbidden,
 * and latency-critical sections that cannot take the device-wide synchronisation an
 * allocation implies.  The scratch between the three kernels lives in a fixed
 * __device__ array inside the library (about 4 KiB, reserved on every GPU that loads
 * the library until the process exits); the caller passes every input and output as
 * device pointers and needs no tk_ctx.
 *   - at most TK_NOALLOC_MAX_RECORDS records per call; larger n is rejected with an error
 *   - the first call on a device looks up the address of the scratch array
 *     (cudaGetSymbolAddress); make that first call before any stream capture, and during
 *     initialisation: with lazy module loading it is also when the driver loads this
 *     library's device code, which may allocate inside the driver.  Host threads may make
 *     it concurrently
 *   - all calls on a device share the scratch: calls that may overlap on the GPU
 *     (different streams without a dependency) race; issue them on one stream
 *   - no cudaMalloc / cudaFree / allocating library call is made */
#define TK_NOALLOC_MAX_RECORDS (1u << 20)
int tk_compact_noalloc(const int32_t* d_keep, uint32_t n, uint32_t batch_size, uint32_t carry_in,
                       int32_t* d_kidx, tk_result_t* d_result, void* stream);

/* ---- Kafka RecordBatch v2 blobs -> slot -----------------------
 * The step before the path on real topics.  The host copies each fetched partition's
 * `records` field (a sequence of RecordBatch v2, magic 2, as kafka-python's fetcher receives
 * it; kafka/record/default_records.py and kafka/consumer/fetcher.py of the reference's
 * dependency kafka-python >= 2.0.2, setup.py:9) unparsed into the slot's pinned byte buffer
 * (tk_slot_host_ptrs) and describes the blobs; the device validates every batch (CRC-32C),
 * splits it at its varint record framing and builds the slot's CSR.  Replaces, per record,
 * the ConsumerRecord that `for record in self._consumer` (src/kafka_dataset.py:156) yields.
 *   - a batch that is cut off by the end of its blob is ignored (fetch responses end so)
 *   - records with offset < fetch_offset are skipped (batches are returned whole)
 *   - compressed batches (attributes & 7: 1 gzip, 2 snappy, 3 LZ4, 4 zstd; kafka/codec.py
 *     gzip_decode, snappy_decode, lz4_decode, zstd_decode) are decompressed on the device: gzip
 *     member + DEFLATE; snappy blocks, raw or in snappy-java's xerial framing; LZ4 frame + block
 *     format; Zstandard frames without dictionary (one with a dictionary id surfaces as
 *     TK_STATUS_UNSUPPORTED).  The inner checksums (CRC-32, xxHash) are not verified: the
 *     batch's CRC-32C covers the same bytes
 *   - control batches are skipped; null values surface as TK_STATUS_UNSUPPORTED, CRC /
 *     framing / compressed-stream errors as TK_STATUS_CORRUPT on the first record of the batch, when
 *     the batch that contains it is delivered
 *   - offset_delta is added to every offset (0 for real fetch responses; lets a test log be
 *     replayed at higher offsets without re-encoding) */
typedef struct tk_blob {
    uint32_t start;          /* first byte of the blob in the slot's byte buffer */
    uint32_t length;
    int32_t  partition;      /* dense partition index, as in tk_rec_t */
    uint32_t reserved;
    int64_t  fetch_offset;
    int64_t  offset_delta;
} tk_blob_t;
typedef struct tk_rb_summary {
    uint32_t n_records;      /* records in the slot, including skipped ones */
    uint32_t n_batches;
    uint64_t n_bytes;        /* value bytes in the slot */
    uint32_t status;         /* TK_RB_OK */
    uint32_t reserved;
} tk_rb_summary_t;
enum { TK_RB_OK = 0, TK_RB_OVERFLOW = 1 /* some batches' values did not fit slot_bytes: their first
                                           record carries TK_STATUS_SLOT_OVERFLOW, nothing was written
                                           outside the slot */ };
#define TK_MAX_BLOBS 1024u
/* flags the split leaves in tk_rec_t.partition (the scans act on them) */
#define TK_REC_DROPPED     0x40000000   /* below the fetch offset / control batch: not consumed */
#define TK_REC_CORRUPT     0x20000000
#define TK_REC_UNSUPPORTED 0x10000000
#define TK_REC_OVERFLOW    0x30000000   /* both of the above: values did not fit the slot */
#define TK_REC_FLAGS       0x70000000
/* Raw bytes are in the slot's pinned buffer [0, n_bytes).  The call follows the chain of batch
 * headers of every blob on the host (one read per batch) and returns the number of record
 * entries the slot will hold in *n_records (skipped records included: their entries stay, flagged);
 * the H2D copy and the split kernels (CRC-32C, record framing, value gather) are asynchronous on
 * the library's copy stream; tk_process_slot / tk_slot_wait order behind them as behind
 * tk_slot_submit.  Fails at once if the slot geometry (slot_records) is too small. */
int tk_slot_submit_recordbatches(tk_ctx* ctx, int slot, uint64_t n_bytes, const tk_blob_t* blobs,
                                 uint32_t n_blobs, uint32_t* n_records);
/* Same, for a consumer that keeps the fetched bytes in buffers of its own (a receive arena): blob k
 * lies at host_ptrs[k] (blobs[k].start is still its place in the slot's device buffer, blobs back
 * to back) and is copied to the device straight from there -- no gather into the slot's pinned
 * buffer.  Page-lock the arena once with tk_host_register so the copies are asynchronous DMA
 * transfers; the memory must stay valid until the slot's split has run (tk_slot_recordbatch_result,
 * or the slot's result). */
int tk_slot_submit_recordbatches_ext(tk_ctx* ctx, int slot, uint64_t n_bytes, const tk_blob_t* blobs,
                                     const uint8_t* const* host_ptrs, uint32_t n_blobs, uint32_t* n_records);
int tk_host_register(void* ptr, uint64_t bytes);
int tk_host_unregister(void* ptr);
/* optional: wait for the split and read its totals (value bytes, batches) */
int tk_slot_recordbatch_result(tk_ctx* ctx, int slot, tk_rb_summary_t* summary);
/* one record header of the device twin (error reports) */
int tk_slot_record(tk_ctx* ctx, int slot, uint32_t index, tk_rec_t* out);
/* host-side CRC-32C (Castagnoli) for integrators that build or check batches themselves */
uint32_t tk_crc32c(const void* data, uint64_t n, uint32_t crc);

/* Measurement aid (no counterpart in the reference): with timing enabled, tk_process_slot
 * brackets the decode / collate kernels of every slot with two CUDA events on `stream`;
 * tk_slot_decode_ms waits for the second one and returns the elapsed milliseconds.  This is
 * how bench.py measures the dominant kernel inside the timed region without issuing the
 * sequence kernel by kernel from the host. */
int tk_ctx_set_timing(tk_ctx* ctx, int enabled);
int tk_slot_decode_ms(tk_ctx* ctx, int slot, float* ms);

/* ---- results ---------------------------------------------------------------- */
/* async D2H of the slot's result block; also marks the slot's device buffers free
 * for the next tk_slot_submit once `stream` reaches this point */
int tk_slot_fetch_result(tk_ctx* ctx, int slot, void* stream);
/* block until the fetch finished, then expose the pinned copies:
 *   batch_end[j]   index of the record that closed batch j (j < n_closed); entry
 *                  n_closed describes the open batch (n_records - 1)
 *   batch_lmax[j]  longest kept record of batch j (JSON schemas)
 *   commit         int64 matrix, commit[p * pitch + j] = offset to commit for
 *                  partition p after batch j, -1 if the partition was never seen */
int tk_slot_result(tk_ctx* ctx, int slot, const tk_result_t** hdr, const uint32_t** batch_end,
                   const uint32_t** batch_lmax, const int64_t** commit, uint32_t* commit_pitch);
/* per-record scratch of the last run on this slot, copied to host (tests, error reports) */
int tk_slot_debug_records(tk_ctx* ctx, int slot, int32_t* rec_len, int32_t* rec_status,
                          int32_t* rec_kidx, uint32_t n);

#ifdef __cplusplus
}
#endif
#endif /* TORCHKAFKA_B200_H */
