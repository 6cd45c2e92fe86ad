// tk_recordbatch.cu -- the step before the hot path on real topics: Kafka
// RecordBatch v2 blobs, as they sit in a fetch response, are shipped over the link unparsed and
// split into the slot's CSR (bytes back to back + tk_rec_t per record) on the device.
//
// Replaces, on the host, kafka-python's per-record work for every fetched partition:
//   DefaultRecordBatch header parse + CRC-32C validation + varint record framing
//   (kafka/record/default_records.py), the "offset below the fetch position -> skip" rule of
//   Fetcher._message_generator (kafka/consumer/fetcher.py), and the ConsumerRecord object that
//   KafkaDataset.__iter__ (src/kafka_dataset.py:156) would then read `.value` from.
// kafka-python is a dependency of the reference (setup.py:9, >= 2.0.2) and not part of its tree;
// the format implemented here is the published one (Kafka protocol guide, "Record Batch", magic 2).
//
// Kernels (all on the slot's copy stream, behind the H2D of the raw bytes):
//   k_rb_headers  one block: thread per partition blob walks the chain of batch headers (twice:
//                 count, block scan, emit) -> batch descriptors in consumption order
//   k_rb_check    warp per batch: CRC-32C of the batch (strided slicing-by-4, lanes own one 16-byte
//                 chunk of every 512-byte row; lane states combined by carry-less multiplication),
//                 then lane 0 walks the records (bytes now in L1/L2): count, value bytes, framing
//   k_rb_prefix   one block: exclusive scan over the batches -> first record / first value byte of
//                 each batch, totals, capacity checks
//   k_rb_emit     warp per batch: lane 0 walks the records again, the warp copies each value to
//                 its place in the slot's byte buffer (aligned destination words, funnel-shifted
//                 source) and writes the record's tk_rec_t
// Big-endian header fields, zig-zag varints.  Uncompressed batches only (codecs: SURVEY §8f-2).
#include "tk_internal.h"
#include <mutex>

// ---------------------------------------------------------------- CRC-32C tables
#define TK_CRC_QUAL const
#define TK_CRC_NAME(x) tk_crc32c_host_##x
#include "tk_crc32c_tables.inc"
#undef TK_CRC_QUAL
#undef TK_CRC_NAME
#define TK_CRC_QUAL __device__ const
#define TK_CRC_NAME(x) tk_crc32c_dev_##x
#include "tk_crc32c_tables.inc"
#undef TK_CRC_QUAL
#undef TK_CRC_NAME

// Host-side CRC-32C (slicing-by-8): for integrators and the test encoder; the device validates
// fetched batches itself.  `crc` continues a previous value (0 to start).
extern "C" uint32_t tk_crc32c(const void* data, uint64_t n, uint32_t crc) {
    const uint8_t* p = static_cast<const uint8_t*>(data);
    uint32_t c = ~crc;
    while (n && (reinterpret_cast<uintptr_t>(p) & 7u)) { c = tk_crc32c_host_t[0][(c ^ *p++) & 0xFF] ^ (c >> 8); --n; }
    while (n >= 8) {
        uint32_t lo, hi;
        memcpy(&lo, p, 4); memcpy(&hi, p + 4, 4);
        lo ^= c;
        c = tk_crc32c_host_t[7][lo & 0xFF] ^ tk_crc32c_host_t[6][(lo >> 8) & 0xFF] ^
            tk_crc32c_host_t[5][(lo >> 16) & 0xFF] ^ tk_crc32c_host_t[4][lo >> 24] ^
            tk_crc32c_host_t[3][hi & 0xFF] ^ tk_crc32c_host_t[2][(hi >> 8) & 0xFF] ^
            tk_crc32c_host_t[1][(hi >> 16) & 0xFF] ^ tk_crc32c_host_t[0][hi >> 24];
        p += 8; n -= 8;
    }
    while (n) { c = tk_crc32c_host_t[0][(c ^ *p++) & 0xFF] ^ (c >> 8); --n; }
    return ~c;
}

// ---------------------------------------------------------------- device helpers
#define RB_HEADER 61u
#define RB_CRC_FROM 21u

__device__ __forceinline__ uint32_t rb_be32(const uint8_t* p) {
    return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3];
}
__device__ __forceinline__ uint64_t rb_be64(const uint8_t* p) {
    return ((uint64_t)rb_be32(p) << 32) | rb_be32(p + 4);
}

// zig-zag varint at p (p < end); returns false on overrun / more than 10 bytes
__device__ __forceinline__ bool rb_varint(const uint8_t* raw, uint32_t& p, uint32_t end, long long& out) {
    unsigned long long u = 0;
    for (int shift = 0; shift < 70; shift += 7) {
        if (p >= end) return false;
        const uint32_t b = raw[p++];
        u |= (unsigned long long)(b & 0x7Fu) << shift;
        if (!(b & 0x80u)) {
            out = (long long)(u >> 1) ^ -(long long)(u & 1);
            return true;
        }
    }
    return false;
}

struct rb_args {
    const uint8_t*   raw;
    const tk_blob_t* blobs;
    uint32_t         n_blobs;
    tk_rb_batch*     batches;
    uint32_t         max_batches;
    tk_rb_summary_t* summary;
    uint8_t*         bytes;           // slot byte buffer (destination)
    tk_rec_t*        recs;            // slot record headers (destination)
    uint32_t         slot_records;
    uint64_t         slot_bytes;
};

// ---------------------------------------------------------------- k_rb_headers
// One step of the header chain of a blob.  A blob ends at its first incomplete batch (fetch
// responses are cut at max_bytes); a header that cannot be trusted ends the blob with one
// descriptor of status CORRUPT, because the next boundary cannot be found.
struct rb_hdr {
    uint32_t start, end, status;
};
__device__ __forceinline__ bool rb_next_header(const uint8_t* raw, uint32_t pos, uint32_t blob_end, rb_hdr& h) {
    if (blob_end - pos < 12u) return false;
    const uint32_t blen = rb_be32(raw + pos + 8);
    h.start = pos;
    if ((int32_t)blen < (int32_t)(RB_HEADER - 12u)) { h.end = blob_end; h.status = TK_STATUS_CORRUPT; return true; }
    if (blen > blob_end - pos - 12u) return false;                    // partial trailing batch
    h.end = pos + 12u + blen;
    h.status = raw[pos + 16] == 2 ? TK_STATUS_OK : TK_STATUS_CORRUPT;
    if (h.status) h.end = blob_end;
    return true;
}

__global__ void __launch_bounds__(1024) k_rb_headers(const rb_args a) {
    __shared__ uint32_t warp_tot[32];
    __shared__ uint32_t s_total;
    const uint32_t t = threadIdx.x;
    uint32_t count = 0, blob_start = 0, blob_end = 0;
    if (t < a.n_blobs) {
        blob_start = a.blobs[t].start;
        blob_end = blob_start + a.blobs[t].length;
        rb_hdr h;
        for (uint32_t pos = blob_start; rb_next_header(a.raw, pos, blob_end, h); pos = h.end) ++count;
    }
    // exclusive scan of the per-blob batch counts (blobs <= 1024: one value per thread)
    uint32_t incl = count;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
        if ((t & 31) >= d) incl += o;
    }
    if ((t & 31) == 31) warp_tot[t >> 5] = incl;
    __syncthreads();
    if (t < 32) {
        uint32_t w = warp_tot[t], wi = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t o = __shfl_up_sync(0xffffffffu, wi, d);
            if (t >= d) wi += o;
        }
        warp_tot[t] = wi - w;
        if (t == 31) s_total = wi;
    }
    __syncthreads();
    uint32_t idx = warp_tot[t >> 5] + incl - count;
    if (t == 0) {
        a.summary->n_batches = min(s_total, a.max_batches);
        a.summary->status = s_total > a.max_batches ? TK_RB_TOO_MANY_BATCHES : TK_RB_OK;
    }
    if (t < a.n_blobs) {
        const tk_blob_t blob = a.blobs[t];
        rb_hdr h;
        for (uint32_t pos = blob_start; rb_next_header(a.raw, pos, blob_end, h); pos = h.end, ++idx) {
            if (idx >= a.max_batches) break;
            tk_rb_batch b;
            b.start = h.start; b.end = h.end; b.partition = blob.partition; b.status = h.status;
            b.base_offset = h.status ? 0 : (int64_t)rb_be64(a.raw + h.start) + blob.offset_delta;
            b.fetch_offset = blob.fetch_offset;
            b.n_rec = 0; b.value_bytes = 0; b.rec_base = 0; b.byte_base = 0;
            a.batches[idx] = b;
        }
    }
}

// ---------------------------------------------------------------- CRC-32C on the device
struct rb_tables {
    uint32_t t[4][256];          // slicing-by-4: fold one little-endian word into the state
    uint32_t a496[4][256];       // advance the state by 496 zero bytes
};
__device__ __forceinline__ void rb_tables_init(rb_tables& tb) {
    for (uint32_t x = threadIdx.x; x < 1024; x += blockDim.x) {
        (&tb.t[0][0])[x] = (&tk_crc32c_dev_t[0][0])[x];
        (&tb.a496[0][0])[x] = (&tk_crc32c_dev_a496[0][0])[x];
    }
    __syncthreads();
}
__device__ __forceinline__ uint32_t crc_word(const rb_tables& tb, uint32_t c, uint32_t w) {
    c ^= w;
    return tb.t[3][c & 0xFF] ^ tb.t[2][(c >> 8) & 0xFF] ^ tb.t[1][(c >> 16) & 0xFF] ^ tb.t[0][c >> 24];
}
__device__ __forceinline__ uint32_t crc_byte(const rb_tables& tb, uint32_t c, uint32_t b) {
    return tb.t[0][(c ^ b) & 0xFF] ^ (c >> 8);
}
__device__ __forceinline__ uint32_t crc_skip496(const rb_tables& tb, uint32_t c) {
    return tb.a496[0][c & 0xFF] ^ tb.a496[1][(c >> 8) & 0xFF] ^ tb.a496[2][(c >> 16) & 0xFF] ^ tb.a496[3][c >> 24];
}
// a(x) * b(x) mod P in the reflected representation (bit 31 = x^0): advances state a by the
// number of bytes b = x^(8n) stands for
__device__ __forceinline__ uint32_t crc_mul(uint32_t a, uint32_t b) {
    uint32_t r = 0;
#pragma unroll 8
    for (int i = 0; i < 32; ++i) {
        r ^= (uint32_t)((int32_t)a >> 31) & b;
        a <<= 1;
        b = (b >> 1) ^ ((0u - (b & 1u)) & 0x82F63B78u);
    }
    return r;
}

// CRC-32C of raw[from, to), whole warp.  Lane 0 folds the bytes up to the first 16-byte boundary;
// then every lane owns one 16-byte chunk of each 512-byte row (coalesced loads), skipping the 31
// chunks in between with the a496 operator; the 32 lane states are combined pairwise (advance by
// 16, 32, .. 256 bytes); lane 0 folds what is left behind the last full row.
__device__ __forceinline__ uint32_t rb_crc32c_warp(const rb_tables& tb, const uint8_t* raw, uint32_t from,
                                                   uint32_t to, uint32_t lane) {
    uint32_t c = 0xFFFFFFFFu;
    uint32_t pos = from;
    const uint32_t body = min((from + 15u) & ~15u, to);
    if (lane == 0) for (; pos < body; ++pos) c = crc_byte(tb, c, raw[pos]);
    pos = body;
    const uint32_t rows = (to - pos) / 512u;
    if (rows) {
        uint32_t s = lane == 0 ? c : 0u;
        const uint4* src = reinterpret_cast<const uint4*>(raw + pos) + lane;
        uint4 v = __ldg(src);
        for (uint32_t r = 0; r < rows; ++r) {
            const uint4 cur = v;
            if (r + 1 < rows) v = __ldg(src + (size_t)(r + 1) * 32);
            if (r) s = crc_skip496(tb, s);
            s = crc_word(tb, s, cur.x);
            s = crc_word(tb, s, cur.y);
            s = crc_word(tb, s, cur.z);
            s = crc_word(tb, s, cur.w);
        }
        // pairwise: the lower lane of each pair is advanced over the bytes of the upper one
#pragma unroll
        for (int j = 0; j < 5; ++j) {
            const uint32_t d = 1u << j;
            const uint32_t adv = (lane & (2 * d - 1)) == d - 1 ? crc_mul(s, tk_crc32c_dev_xpow[j]) : 0u;
            const uint32_t from_low = __shfl_up_sync(0xffffffffu, adv, d);
            if ((lane & (2 * d - 1)) == 2 * d - 1) s ^= from_low;
        }
        c = __shfl_sync(0xffffffffu, s, 31);
        pos += rows * 512u;
    } else {
        c = __shfl_sync(0xffffffffu, c, 0);
    }
    if (lane == 0) {
        for (; pos + 4 <= to && !(pos & 3u); pos += 4) c = crc_word(tb, c, *reinterpret_cast<const uint32_t*>(raw + pos));
        for (; pos < to; ++pos) c = crc_byte(tb, c, raw[pos]);
    }
    return ~__shfl_sync(0xffffffffu, c, 0);
}

// ---------------------------------------------------------------- record walk
// One record of a batch, decoded by one lane.  Layout (all varints zig-zag):
//   length | attributes(1) | timestampDelta | offsetDelta | keyLength | key | valueLength | value | headers
struct rb_record {
    uint32_t value_pos;      // byte position of the value in the raw buffer
    int32_t  value_len;      // -1: null value
    int64_t  offset_delta;
    uint32_t next;           // position of the next record
};
__device__ __forceinline__ bool rb_decode_record(const uint8_t* raw, uint32_t p, uint32_t end, rb_record& r) {
    long long length, ts, od, klen, vlen;
    if (!rb_varint(raw, p, end, length) || length < 1 || (unsigned long long)length > end - p) return false;
    const uint32_t rec_end = p + (uint32_t)length;
    uint32_t q = p + 1;                                       // attributes
    if (!rb_varint(raw, q, rec_end, ts) || !rb_varint(raw, q, rec_end, od) || !rb_varint(raw, q, rec_end, klen))
        return false;
    if (klen > 0) {
        if ((unsigned long long)klen > rec_end - q) return false;
        q += (uint32_t)klen;
    }
    if (!rb_varint(raw, q, rec_end, vlen)) return false;
    if (vlen > 0 && (unsigned long long)vlen > rec_end - q) return false;
    if (od < 0 || od > 0x7FFFFFFF) return false;
    r.value_pos = q;
    r.value_len = vlen < 0 ? -1 : (int32_t)vlen;
    r.offset_delta = od;
    r.next = rec_end;
    return true;
}

// ---------------------------------------------------------------- k_rb_check
__global__ void __launch_bounds__(256) k_rb_check(const rb_args a) {
    __shared__ rb_tables tb;
    rb_tables_init(tb);
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t n_batches = a.summary->n_batches;
    for (uint32_t i = warp; i < n_batches; i += nwarps) {
        tk_rb_batch b = a.batches[i];
        uint32_t status = b.status, n_rec = 1, value_bytes = 0;
        if (!status) {
            const uint32_t want = rb_be32(a.raw + b.start + 17);
            const uint32_t got = rb_crc32c_warp(tb, a.raw, b.start + RB_CRC_FROM, b.end, lane);
            if (got != want) status = TK_STATUS_CORRUPT;
        }
        if (!status && lane == 0) {
            const uint32_t attributes = ((uint32_t)a.raw[b.start + 21] << 8) | a.raw[b.start + 22];
            const int32_t count = (int32_t)rb_be32(a.raw + b.start + 57);
            const bool control = attributes & 0x20u;
            if (attributes & 0x07u) status = TK_STATUS_UNSUPPORTED;              // compressed
            else if (count < 0 || (uint32_t)count > (b.end - b.start)) status = TK_STATUS_CORRUPT;
            else {
                uint32_t p = b.start + RB_HEADER;
                rb_record r;
                for (int32_t k = 0; k < count; ++k) {
                    if (!rb_decode_record(a.raw, p, b.end, r)) { status = TK_STATUS_CORRUPT; break; }
                    const bool dropped = control || b.base_offset + r.offset_delta < b.fetch_offset;
                    if (!dropped && r.value_len > 0) value_bytes += (uint32_t)r.value_len;
                    p = r.next;
                }
                if (!status && p != b.end) status = TK_STATUS_CORRUPT;
                if (!status) n_rec = (uint32_t)count;
            }
        }
        status = __shfl_sync(0xffffffffu, status, 0);
        if (lane == 0) {
            // a batch that cannot be split is carried as one pseudo-record that holds the error
            a.batches[i].status = status;
            a.batches[i].n_rec = status ? 1u : n_rec;
            a.batches[i].value_bytes = status ? 0u : value_bytes;
        }
    }
}

// ---------------------------------------------------------------- k_rb_prefix
__global__ void __launch_bounds__(1024) k_rb_prefix(const rb_args a) {
    __shared__ unsigned long long warp_tot[32];
    __shared__ unsigned long long s_run;
    const uint32_t t = threadIdx.x, lane = t & 31;
    const uint32_t n_batches = a.summary->n_batches;
    if (t == 0) s_run = 0;
    __syncthreads();
    for (uint32_t base = 0; base < n_batches; base += blockDim.x) {
        const uint32_t i = base + t;
        unsigned long long v = 0;
        if (i < n_batches) v = ((unsigned long long)a.batches[i].n_rec << 40) | a.batches[i].value_bytes;
        unsigned long long incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned long long o = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= (uint32_t)d) incl += o;
        }
        if (lane == 31) warp_tot[t >> 5] = incl;
        __syncthreads();
        if (t < 32) {
            unsigned long long w = warp_tot[t], wi = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const unsigned long long o = __shfl_up_sync(0xffffffffu, wi, d);
                if (t >= (uint32_t)d) wi += o;
            }
            warp_tot[t] = wi - w;
        }
        __syncthreads();
        const unsigned long long excl = s_run + warp_tot[t >> 5] + incl - v;
        if (i < n_batches) {
            a.batches[i].rec_base = (uint32_t)(excl >> 40);
            a.batches[i].byte_base = (uint32_t)(excl & 0xFFFFFFFFFFull);
        }
        __syncthreads();
        if (t == blockDim.x - 1) s_run = excl + v;
        __syncthreads();
    }
    if (t == 0) {
        const unsigned long long n_rec = s_run >> 40, n_bytes = s_run & 0xFFFFFFFFFFull;
        a.summary->n_records = (uint32_t)min(n_rec, 0xFFFFFFFFull);
        a.summary->n_bytes = n_bytes;
        if (a.summary->status == TK_RB_OK) {
            if (n_rec > a.slot_records) a.summary->status = TK_RB_TOO_MANY_RECORDS;
            else if (n_bytes > a.slot_bytes) a.summary->status = TK_RB_TOO_MANY_BYTES;
            else a.recs[n_rec].start = (uint32_t)n_bytes;          // closes the CSR
        }
    }
}

// ---------------------------------------------------------------- k_rb_emit
// dst[0, len) = src[0, len), whole warp; destination words aligned, source funnel-shifted
__device__ __forceinline__ void rb_warp_copy(uint8_t* dst, const uint8_t* src, uint32_t len, uint32_t lane) {
    const uint32_t head = min((4u - ((uint32_t)(uintptr_t)dst & 3u)) & 3u, len);
    if (lane < head) dst[lane] = src[lane];
    const uint32_t words = (len - head) >> 2;
    const uint8_t* s = src + head;
    uint32_t* d32 = reinterpret_cast<uint32_t*>(dst + head);
    const uint32_t sh = ((uint32_t)(uintptr_t)s & 3u) * 8u;
    const uint32_t* s32 = reinterpret_cast<const uint32_t*>(s - ((uintptr_t)s & 3u));
    for (uint32_t w = lane; w < words; w += 32) d32[w] = __funnelshift_r(__ldg(s32 + w), __ldg(s32 + w + 1), sh);
    const uint32_t done = head + words * 4u;
    if (done + lane < len) dst[done + lane] = src[done + lane];
}

__global__ void __launch_bounds__(256) k_rb_emit(const rb_args a) {
    if (a.summary->status != TK_RB_OK) return;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t n_batches = a.summary->n_batches;
    for (uint32_t i = warp; i < n_batches; i += nwarps) {
        const tk_rb_batch b = a.batches[i];
        if (b.status) {
            if (lane == 0) {
                tk_rec_t rec;
                rec.start = b.byte_base;
                rec.partition = b.partition | (b.status == TK_STATUS_UNSUPPORTED ? TK_REC_UNSUPPORTED : TK_REC_CORRUPT);
                rec.offset = b.base_offset;
                a.recs[b.rec_base] = rec;
            }
            continue;
        }
        const bool control = a.raw[b.start + 22] & 0x20u;
        uint32_t p = b.start + RB_HEADER, dst = b.byte_base;
        for (uint32_t k = 0; k < b.n_rec; ++k) {
            rb_record r;
            r.value_pos = 0; r.value_len = 0; r.offset_delta = 0; r.next = 0;
            if (lane == 0) rb_decode_record(a.raw, p, b.end, r);       // validated by k_rb_check
            const uint32_t vpos = __shfl_sync(0xffffffffu, r.value_pos, 0);
            const int32_t vlen = __shfl_sync(0xffffffffu, r.value_len, 0);
            p = __shfl_sync(0xffffffffu, r.next, 0);
            const long long offset = b.base_offset + __shfl_sync(0xffffffffu, (long long)r.offset_delta, 0);
            const bool dropped = control || offset < b.fetch_offset;
            const uint32_t n = (!dropped && vlen > 0) ? (uint32_t)vlen : 0u;
            if (n) rb_warp_copy(a.bytes + dst, a.raw + vpos, n, lane);
            if (lane == 0) {
                tk_rec_t rec;
                rec.start = dst;
                rec.partition = b.partition | (dropped ? TK_REC_DROPPED : (vlen < 0 ? TK_REC_UNSUPPORTED : 0));
                rec.offset = offset;
                a.recs[b.rec_base + k] = rec;
            }
            dst += n;
        }
    }
}

// ---------------------------------------------------------------- entry points
// (producer threads submit concurrently: the first-use allocation is serialised)
static std::mutex g_rb_mutex;
static int rb_ensure(tk_ctx* ctx) {
    std::lock_guard<std::mutex> lock(g_rb_mutex);
    for (int i = 0; i < ctx->n_slots; ++i) {
        tk_slot& s = ctx->slots[i];
        if (s.d.raw) continue;
        TK_CUDA(cudaMalloc(&s.d.raw, ctx->slot_bytes + TK_BYTES_PAD));
        TK_CUDA(cudaMemset(s.d.raw, 0, ctx->slot_bytes + TK_BYTES_PAD));
        TK_CUDA(cudaMalloc(&s.d.blobs, sizeof(tk_blob_t) * TK_MAX_BLOBS));
        TK_CUDA(cudaMalloc(&s.d.rb_batches, sizeof(tk_rb_batch) * (size_t)ctx->slot_records));
        TK_CUDA(cudaMalloc(&s.d.rb_summary, sizeof(tk_rb_summary_t)));
        TK_CUDA(cudaHostAlloc(&s.h_blobs, sizeof(tk_blob_t) * TK_MAX_BLOBS, cudaHostAllocDefault));
        TK_CUDA(cudaHostAlloc(&s.h_rb_summary, sizeof(tk_rb_summary_t), cudaHostAllocDefault));
    }
    return 0;
}

extern "C" int tk_slot_submit_recordbatches(tk_ctx* ctx, int slot, uint64_t n_bytes, const tk_blob_t* blobs,
                                            uint32_t n_blobs) {
    TK_ENTER(ctx, slot);
    TK_REQUIRE(n_bytes <= ctx->slot_bytes, "n_bytes %llu exceeds slot_bytes %llu", (unsigned long long)n_bytes,
               (unsigned long long)ctx->slot_bytes);
    TK_REQUIRE(n_blobs <= TK_MAX_BLOBS && (blobs || !n_blobs), "at most %u blobs per slot", TK_MAX_BLOBS);
    for (uint32_t k = 0; k < n_blobs; ++k)
        TK_REQUIRE((uint64_t)blobs[k].start + blobs[k].length <= n_bytes && blobs[k].partition >= 0 &&
                       !(blobs[k].partition & TK_REC_FLAGS),
                   "blob %u outside the buffer or bad partition", k);
    int rc = rb_ensure(ctx); if (rc) return rc;
    tk_slot& s = ctx->slots[slot];
    cudaStream_t st = ctx->copy_stream;
    // until the result is read the slot's geometry is unknown: an upper bound keeps checks sane
    s.n_records = 0;
    s.n_bytes = 0;
    if (n_blobs) memcpy(s.h_blobs, blobs, sizeof(tk_blob_t) * n_blobs);
    if (s.free_recorded) TK_CUDA(cudaStreamWaitEvent(st, s.ev_free, 0));
    if (n_bytes) TK_CUDA(cudaMemcpyAsync(s.d.raw, s.h_bytes, n_bytes, cudaMemcpyHostToDevice, st));
    if (n_blobs) TK_CUDA(cudaMemcpyAsync(s.d.blobs, s.h_blobs, sizeof(tk_blob_t) * n_blobs, cudaMemcpyHostToDevice, st));
    rb_args a;
    a.raw = s.d.raw; a.blobs = s.d.blobs; a.n_blobs = n_blobs; a.batches = s.d.rb_batches;
    a.max_batches = ctx->slot_records; a.summary = s.d.rb_summary; a.bytes = s.d.bytes; a.recs = s.d.recs;
    a.slot_records = ctx->slot_records; a.slot_bytes = ctx->slot_bytes;
    k_rb_headers<<<1, 1024, 0, st>>>(a);
    k_rb_check<<<148 * 4, 256, 0, st>>>(a);
    k_rb_prefix<<<1, 1024, 0, st>>>(a);
    k_rb_emit<<<148 * 4, 256, 0, st>>>(a);
    TK_LAUNCH_CHECK();
    TK_CUDA(cudaMemcpyAsync(s.h_rb_summary, s.d.rb_summary, sizeof(tk_rb_summary_t), cudaMemcpyDeviceToHost, st));
    TK_CUDA(cudaEventRecord(s.ev_h2d, st));
    return 0;
}

extern "C" int tk_slot_recordbatch_result(tk_ctx* ctx, int slot, tk_rb_summary_t* summary) {
    TK_ENTER(ctx, slot);
    tk_slot& s = ctx->slots[slot];
    TK_REQUIRE(s.h_rb_summary, "no record batches were submitted on this context");
    TK_CUDA(cudaEventSynchronize(s.ev_h2d));
    const tk_rb_summary_t r = *s.h_rb_summary;
    if (summary) *summary = r;
    if (r.status == TK_RB_OK) {
        s.n_records = r.n_records;
        s.n_bytes = r.n_bytes;
    }
    return 0;
}

extern "C" int tk_slot_record(tk_ctx* ctx, int slot, uint32_t index, tk_rec_t* out) {
    TK_ENTER(ctx, slot);
    TK_REQUIRE(out && index <= ctx->slot_records, "bad record index %u", index);
    TK_CUDA(cudaDeviceSynchronize());
    TK_CUDA(cudaMemcpy(out, ctx->slots[slot].d.recs + index, sizeof(tk_rec_t), cudaMemcpyDeviceToHost));
    return 0;
}
