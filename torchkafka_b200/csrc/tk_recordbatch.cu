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
// Host (tk_slot_submit_recordbatches): follows the chain of batch headers of every blob -> batch
// descriptors in consumption order, each with the index of its first record; returns the record
// count of the slot.  Kernels (on the slot's copy stream, behind the H2D of the raw bytes):
//   k_rb_check    warp per batch: CRC-32C of the batch (strided slicing-by-4, lanes own one 16-byte
//                 chunk of every 512-byte row; lane states combined by carry-less multiplication),
//                 then the record walk (bytes now in L1/L2): lane 0 follows the chain of record
//                 lengths, 32 at a time, and every lane decodes one record: framing checks,
//                 record header (partition, offset), copy descriptor
//   k_rb_prefix   one block: exclusive scan of the batches' value bytes, capacity check
//   k_rb_copy     warp per record: the value moves to its place in the slot's byte buffer (aligned
//                 destination words, funnel-shifted source); values above 16 KiB are queued and
//                 copied in 4 KiB pieces by all warps (k_rb_copy_big)
// Big-endian header fields, zig-zag varints.  LZ4 and snappy batches are decompressed first
// (tk_codecs.cu), gzip and zstd batches too (tk_inflate.cuh, tk_zstd.cuh).
#include "tk_rb.h"
#include <mutex>
#include <stdlib.h>

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

// ---------------------------------------------------------------- batch headers (host)
// The chain of batch headers of a blob is followed on the host, in tk_slot_submit_recordbatches:
// it is one dependent read per batch (tens of kilobytes apart) of memory the host has just
// written, a few dozen per blob -- a latency chain that would occupy one GPU thread per blob for
// tens of microseconds, and whose result (the record count of the slot) the host needs before it
// can size the output tensors.  Everything that is proportional to the bytes or to the records
// (checksum, record framing, value gather) runs on the device.
//
// A blob ends at its first incomplete batch (fetch responses are cut at max_bytes); a header that
// cannot be trusted ends the blob with one descriptor of status CORRUPT, because the next boundary
// cannot be found.  The header's record count gives every batch its first record index in the
// slot; a batch that fails later keeps its record slots (first one carries the error).
static inline uint32_t rb_host_be32(const uint8_t* p) {
    return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3];
}
// `raw` addresses the slot's pinned buffer; with `ext` (one host pointer per blob) blob t's bytes are
// read from ext[t] instead, wherever the caller keeps them (positions stay slot offsets).
static int rb_walk_host(const uint8_t* raw_slot, const uint8_t* const* ext, const tk_blob_t* blobs, uint32_t n_blobs, tk_rb_batch* out,
                        uint32_t max_batches, uint32_t max_records, uint32_t* n_batches_out, uint32_t* n_records_out,
                        tk_rb_unit* units, uint32_t max_units, uint32_t* n_units_out, uint32_t* n_coded_out) {
    uint32_t nb = 0;
    *n_units_out = 0;
    *n_coded_out = 0;
    uint64_t nrec = 0;
    for (uint32_t t = 0; t < n_blobs; ++t) {
        const tk_blob_t& blob = blobs[t];
        const uint32_t blob_end = blob.start + blob.length;
        uint32_t pos = blob.start;
        const uint8_t* raw = ext ? ext[t] - blob.start : raw_slot;       // raw + slot offset = the byte
        while (blob_end - pos >= 12u) {
            const uint32_t blen = rb_host_be32(raw + pos + 8);
            uint32_t end, status = TK_STATUS_OK, count = 1, codec = 0;
            if ((int32_t)blen < (int32_t)(RB_HEADER - 12u)) {
                end = blob_end; status = TK_STATUS_CORRUPT;
            } else {
                if (blen > blob_end - pos - 12u) break;                   // partial trailing batch
                end = pos + 12u + blen;
                const uint32_t magic = raw[pos + 16];
                const uint32_t c = rb_host_be32(raw + pos + 57);
                codec = raw[pos + 22] & 7u;
                // a record is at least 7 bytes (LZ4 and snappy expand at most ~255-fold, DEFLATE 1032-fold, zstd's RLE blocks far more): a larger count cannot
                // be true (the checksum would say so too, but the count sizes the slot before the
                // checksum is looked at)
                const uint64_t room = (uint64_t)(end - pos - RB_HEADER) * (codec == 4u ? 65536u : codec == 1u ? 1032u : codec ? 255u : 1u);
                if (magic != 2) { status = TK_STATUS_CORRUPT; end = blob_end; }
                else if (c > room / 7u + 1u) status = TK_STATUS_CORRUPT;
                else count = c;
            }
            TK_REQUIRE(nb < max_batches, "more record batches in the slot than slot_records (%u)", max_batches);
            tk_rb_batch& b = out[nb++];
            b.start = pos; b.end = end; b.partition = blob.partition; b.status = status;
            int64_t base = 0;
            if (!status) {
                uint64_t u = ((uint64_t)rb_host_be32(raw + pos) << 32) | rb_host_be32(raw + pos + 4);
                base = (int64_t)u + blob.offset_delta;
            }
            b.base_offset = base;
            b.fetch_offset = blob.fetch_offset;
            b.n_rec = count; b.value_bytes = 0; b.rec_base = (uint32_t)nrec; b.byte_base = 0;
            b.codec = status ? 0u : codec; b.rec_start = pos + RB_HEADER; b.rec_end = end;
            b.unit_first = 0; b.unit_count = 0; b.reserved = 0;
            if (b.codec >= 1u && b.codec <= 4u) {            // gzip, snappy, LZ4, zstd: cut into decodable units
                ++*n_coded_out;
                const int rc = tk_codec_units_host(raw, b, nb - 1, units, max_units, n_units_out);
                if (rc) return rc;
            }
            nrec += count;
            TK_REQUIRE(nrec <= max_records, "more records in the slot than slot_records (%u)", max_records);
            pos = end;
        }
    }
    *n_batches_out = nb;
    *n_records_out = (uint32_t)nrec;
    return 0;
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
        uint32_t r = 0;
        if (rows >= 8) {
            // four row streams per lane (quarters of the rows): the table walk of one stream is a
            // chain of dependent shared-memory lookups, four of them in flight hide its latency
            const uint32_t q = rows / 4;
            uint32_t s1 = 0, s2 = 0, s3 = 0;
            const uint4 *p0 = src, *p1 = src + (size_t)q * 32, *p2 = src + (size_t)2 * q * 32, *p3 = src + (size_t)3 * q * 32;
            for (uint32_t i = 0; i < q; ++i) {
                const uint4 v0 = __ldg(p0 + (size_t)i * 32), v1 = __ldg(p1 + (size_t)i * 32);
                const uint4 v2 = __ldg(p2 + (size_t)i * 32), v3 = __ldg(p3 + (size_t)i * 32);
                if (i) { s = crc_skip496(tb, s); s1 = crc_skip496(tb, s1); s2 = crc_skip496(tb, s2); s3 = crc_skip496(tb, s3); }
                s = crc_word(tb, s, v0.x); s1 = crc_word(tb, s1, v1.x); s2 = crc_word(tb, s2, v2.x); s3 = crc_word(tb, s3, v3.x);
                s = crc_word(tb, s, v0.y); s1 = crc_word(tb, s1, v1.y); s2 = crc_word(tb, s2, v2.y); s3 = crc_word(tb, s3, v3.y);
                s = crc_word(tb, s, v0.z); s1 = crc_word(tb, s1, v1.z); s2 = crc_word(tb, s2, v2.z); s3 = crc_word(tb, s3, v3.z);
                s = crc_word(tb, s, v0.w); s1 = crc_word(tb, s1, v1.w); s2 = crc_word(tb, s2, v2.w); s3 = crc_word(tb, s3, v3.w);
            }
            // join: every stream is advanced over the q rows of each stream behind it
            uint32_t mq = 0x80000000u;                                  // x^0
            for (uint32_t j = 0, k = q; k; ++j, k >>= 1)
                if (k & 1u) mq = crc_mul(mq, tk_crc32c_dev_xrow[j]);
            s = crc_mul(s, mq) ^ s1;
            s = crc_mul(s, mq) ^ s2;
            s = crc_mul(s, mq) ^ s3;
            r = 4 * q;
        }
        for (; r < rows; ++r) {
            const uint4 cur = __ldg(src + (size_t)r * 32);
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
// Per record, between the kernels: where its value sits in the raw buffer, how long it is, where
// it goes inside its batch's share of the slot, and which batch that is.
struct rb_temp {
    uint32_t src, len, rel, batch;
};

// Warp per batch: checksum, then the record walk (the batch's bytes are in L1/L2 now).  Only the
// chain of record lengths is sequential: lane 0 follows it for 32 records at a time (one varint and
// one add per record), then every lane decodes one record completely and writes its header
// (partition, offset) and copy descriptor.  A batch that fails keeps its record slots: the first
// carries the error, the others are marked skipped.
__global__ void __launch_bounds__(256) k_rb_check(const rb_args a) {
    __shared__ rb_tables tb;
    __shared__ uint32_t s_pos[8][33];
    rb_tables_init(tb);
    const uint32_t lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t n_batches = a.n_batches;
    rb_temp* temp = reinterpret_cast<rb_temp*>(a.temp);
    for (uint32_t i = warp; i < n_batches; i += nwarps) {
        const tk_rb_batch b = a.batches[i];
        uint32_t status = b.status, value_bytes = 0;
        if (!status) {
            const uint32_t want = rb_be32(a.raw + b.start + 17);
            const uint32_t got = rb_crc32c_warp(tb, a.raw, b.start + RB_CRC_FROM, b.end, lane);
            if (got != want) status = TK_STATUS_CORRUPT;
        }
        bool control = false;
        if (!status) {
            const uint32_t attributes = ((uint32_t)a.raw[b.start + 21] << 8) | a.raw[b.start + 22];
            control = attributes & 0x20u;
            if (b.codec > 4u) status = TK_STATUS_UNSUPPORTED;  // no such codec
        }
        // the records: behind the header, or (compressed) in the batch's region of the decompressed buffer
        const uint8_t* rbase = b.codec ? a.raw2 : a.raw;
        const uint32_t rend = b.rec_end, src_flag = b.codec ? RB_SRC2 : 0u;
        uint32_t p = b.rec_start;                                                // lane 0's cursor
        for (uint32_t k0 = 0; !status && k0 < b.n_rec; k0 += 32) {
            const uint32_t m = min(32u, b.n_rec - k0);
            if (lane == 0) {
                uint32_t bad = 0;
                for (uint32_t j = 0; j < m; ++j) {
                    s_pos[wib][j] = p;
                    long long length;
                    uint32_t q = p;
                    if (!rb_varint(rbase, q, rend, length) || length < 1 || (unsigned long long)length > rend - q) {
                        bad = 1;
                        break;
                    }
                    p = q + (uint32_t)length;
                }
                s_pos[wib][32] = bad;
            }
            __syncwarp();
            if (s_pos[wib][32]) status = TK_STATUS_CORRUPT;
            uint32_t n = 0;
            if (!status && lane < m) {
                rb_record r;
                if (!rb_decode_record(rbase, s_pos[wib][lane], rend, r)) {
                    status = TK_STATUS_CORRUPT;
                } else {
                    const long long offset = b.base_offset + r.offset_delta;
                    const bool dropped = control || offset < b.fetch_offset;
                    n = (!dropped && r.value_len > 0) ? (uint32_t)r.value_len : 0u;
                    tk_rec_t rec;
                    rec.start = 0;
                    rec.partition = b.partition | (dropped ? TK_REC_DROPPED : (r.value_len < 0 ? TK_REC_UNSUPPORTED : 0));
                    rec.offset = offset;
                    a.recs[b.rec_base + k0 + lane] = rec;
                    // .src now, .rel after the warp scan below
                    temp[b.rec_base + k0 + lane].src = r.value_pos | src_flag;
                }
            }
            if (__any_sync(0xffffffffu, status != 0)) { status = TK_STATUS_CORRUPT; break; }
            uint32_t incl = n;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= (uint32_t)d) incl += o;
            }
            if (lane < m) {
                rb_temp* t = &temp[b.rec_base + k0 + lane];
                t->len = n; t->rel = value_bytes + incl - n; t->batch = i;
            }
            value_bytes += __shfl_sync(0xffffffffu, incl, 31);
            __syncwarp();
        }
        // the cursor must land on the end of the batch
        const uint32_t p_end = __shfl_sync(0xffffffffu, p, 0);
        if (!status && p_end != rend) status = TK_STATUS_CORRUPT;
        if (status) {
            value_bytes = 0;
            for (uint32_t k = lane; k < b.n_rec; k += 32) {
                rb_temp tmp;
                tmp.src = 0; tmp.len = 0; tmp.rel = 0; tmp.batch = i;
                temp[b.rec_base + k] = tmp;
                tk_rec_t rec;
                rec.start = 0;
                rec.partition = b.partition | (k ? TK_REC_DROPPED
                                                 : status == TK_STATUS_UNSUPPORTED ? TK_REC_UNSUPPORTED
                                                 : status == TK_STATUS_SLOT_OVERFLOW ? TK_REC_OVERFLOW : TK_REC_CORRUPT);
                rec.offset = b.base_offset;
                a.recs[b.rec_base + k] = rec;
            }
        }
        if (lane == 0) {
            a.batches[i].status = status;
            a.batches[i].value_bytes = value_bytes;
        }
    }
}

// ---------------------------------------------------------------- k_rb_prefix
// One block: exclusive scan of the batches' value bytes -> first byte of each batch in the slot.
__global__ void __launch_bounds__(1024) k_rb_prefix(const rb_args a) {
    __shared__ unsigned long long warp_tot[32];
    __shared__ unsigned long long s_total, s_run, s_fit;
    const uint32_t t = threadIdx.x;
    const uint32_t n_batches = a.n_batches;
    if (t == 0) { s_run = 0; s_fit = ~0ull; }
    __syncthreads();
    for (uint32_t base = 0; base < n_batches; base += blockDim.x) {
        const uint32_t i = base + t;
        const unsigned long long v = i < n_batches ? a.batches[i].value_bytes : 0ull;
        const unsigned long long excl = s_run + rb_block_scan(v, warp_tot, &s_total);
        if (i < n_batches) {
            // Values of compressed batches come out of the decompressed buffer, which can hold more
            // than the slot: a batch whose values end behind the slot's byte buffer (and, the prefix
            // being monotone, every batch after it) is not copied; its records surface as
            // TK_STATUS_SLOT_OVERFLOW.  s_fit = value bytes of the batches in front of the first such one.
            if (excl + v > a.slot_bytes) {
                a.batches[i].status = TK_STATUS_SLOT_OVERFLOW;
                a.batches[i].value_bytes = 0;
                a.batches[i].byte_base = 0;
                atomicMin(&s_fit, excl);
            } else {
                a.batches[i].byte_base = (uint32_t)excl;
            }
        }
        __syncthreads();
        if (t == 0) s_run += s_total;
        __syncthreads();
    }
    if (t == 0) {
        const unsigned long long used = s_fit < s_run ? s_fit : s_run;
        a.summary->n_records = a.n_records;
        a.summary->n_batches = n_batches;
        a.summary->n_bytes = used;
        a.summary->status = s_fit < s_run ? TK_RB_OVERFLOW : TK_RB_OK;
        a.summary->reserved = 0;                              // length of the big-record queue
        a.recs[a.n_records].start = (uint32_t)used;           // closes the CSR
    }
}

// ---------------------------------------------------------------- k_rb_copy
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

// Warp per record: the value moves to its place in the slot's byte buffer, the record header gets
// its start.  Values longer than RB_BIG are only queued here: one warp per multi-megabyte value
// would leave the rest of the GPU idle, so k_rb_copy_big spreads their RB_PIECE-byte pieces over
// all warps.
#define RB_BIG 16384u
#define RB_PIECE 4096u
__global__ void __launch_bounds__(256) k_rb_copy(const rb_args a) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t n = a.n_records;
    const uint4* temp = reinterpret_cast<const uint4*>(a.temp);
    for (uint32_t i = warp; i < n; i += nwarps) {
        const uint4 t = __ldg(temp + i);                      // {src, len, rel, batch}
        if (a.batches[t.w].status == TK_STATUS_SLOT_OVERFLOW) {
            // did not fit the slot (k_rb_prefix): no bytes; the batch's first record carries the error
            if (lane == 0) {
                tk_rec_t rec;
                rec.start = (uint32_t)a.summary->n_bytes;
                rec.partition = a.batches[t.w].partition | (i == a.batches[t.w].rec_base ? TK_REC_OVERFLOW : TK_REC_DROPPED);
                rec.offset = a.batches[t.w].base_offset;
                a.recs[i] = rec;
            }
            continue;
        }
        const uint32_t dst = a.batches[t.w].byte_base + t.z;
        if (lane == 0) {
            a.recs[i].start = dst;
            if (t.y > RB_BIG) a.bigq[atomicAdd(&a.summary->reserved, 1u)] = i;
        }
        if (t.y && t.y <= RB_BIG)
            rb_warp_copy(a.bytes + dst, ((t.x & RB_SRC2) ? a.raw2 : a.raw) + (t.x & ~RB_SRC2), t.y, lane);
    }
}

__global__ void __launch_bounds__(256) k_rb_copy_big(const rb_args a) {
    const uint32_t count = a.summary->reserved;
    if (!count) return;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    const uint4* temp = reinterpret_cast<const uint4*>(a.temp);
    uint32_t piece0 = 0;                                      // global index of the entry's first piece
    for (uint32_t e = 0; e < count; ++e) {
        const uint4 t = __ldg(temp + a.bigq[e]);
        const uint32_t pieces = (t.y + RB_PIECE - 1) / RB_PIECE;
        // first piece of this entry that belongs to this warp: global index == warp (mod nwarps)
        uint32_t k = (warp + nwarps - piece0 % nwarps) % nwarps;
        if (k < pieces) {
            const uint32_t dst = a.batches[t.w].byte_base + t.z;
            // pieces are cut at 4-byte boundaries of the destination so that every piece but the
            // first starts on an aligned word
            const uint32_t lead = (4u - (dst & 3u)) & 3u;
            for (; k < pieces; k += nwarps) {
                const uint32_t from = k ? k * RB_PIECE + lead : 0u;
                const uint32_t to = min((k + 1) * RB_PIECE + lead, t.y);
                if (from < to)
                    rb_warp_copy(a.bytes + dst + from, ((t.x & RB_SRC2) ? a.raw2 : a.raw) + (t.x & ~RB_SRC2) + from,
                                 to - from, lane);
            }
        }
        piece0 += pieces;
    }
}

// ---------------------------------------------------------------- entry points
// compressed blocks a slot can hold: one per record plus one per 4 KiB (snappy-java chunks are 32 KiB,
// LZ4 blocks 64 KiB and more)
static inline size_t rb_max_units(const tk_ctx* ctx) { return (size_t)ctx->slot_records + ctx->slot_bytes / 4096 + 16; }

// (producer threads submit concurrently: the first-use allocation is serialised)
static std::mutex g_rb_mutex;
static int rb_ensure(tk_ctx* ctx) {
    std::lock_guard<std::mutex> lock(g_rb_mutex);
    for (int i = 0; i < ctx->n_slots; ++i) {
        tk_slot& s = ctx->slots[i];
        if (s.d.raw) continue;
        TK_CUDA(cudaMalloc(&s.d.raw, ctx->slot_bytes + TK_BYTES_PAD));
        TK_CUDA(cudaMemset(s.d.raw, 0, ctx->slot_bytes + TK_BYTES_PAD));
        TK_CUDA(cudaMalloc(&s.d.rb_batches, sizeof(tk_rb_batch) * (size_t)ctx->slot_records));
        TK_CUDA(cudaMalloc(&s.d.rb_summary, sizeof(tk_rb_summary_t)));
        TK_CUDA(cudaMalloc(&s.d.rb_temp, sizeof(uint4) * (size_t)ctx->slot_records));
        TK_CUDA(cudaMalloc(&s.d.rb_bigq, sizeof(uint32_t) * ((size_t)(ctx->slot_bytes / RB_BIG) + 2)));
        TK_CUDA(cudaHostAlloc(&s.h_rb_batches, sizeof(tk_rb_batch) * (size_t)ctx->slot_records, cudaHostAllocDefault));
        TK_CUDA(cudaHostAlloc(&s.h_rb_summary, sizeof(tk_rb_summary_t), cudaHostAllocDefault));
        TK_CUDA(cudaHostAlloc(&s.h_rb_units, sizeof(tk_rb_unit) * rb_max_units(ctx), cudaHostAllocDefault));
        TK_CUDA(cudaEventCreateWithFlags(&s.ev_raw, cudaEventDisableTiming));
        TK_CUDA(cudaStreamCreateWithFlags(&s.split_stream, cudaStreamNonBlocking));
    }
    return 0;
}

// the decompressed buffer and the device unit table: first compressed batch
static int rb_ensure_codecs(tk_ctx* ctx) {
    std::lock_guard<std::mutex> lock(g_rb_mutex);
    TK_REQUIRE(ctx->slot_bytes < (1ull << 31), "compressed record batches need slot_bytes < 2 GiB");
    if (ctx->n_slots && ctx->slots[ctx->n_slots - 1].d.raw2) return 0;
    for (int i = 0; i < ctx->n_slots; ++i) {
        tk_slot& s = ctx->slots[i];
        if (s.d.raw2) continue;
        TK_CUDA(cudaMalloc(&s.d.raw2, ctx->slot_bytes + TK_BYTES_PAD));
        TK_CUDA(cudaMemset(s.d.raw2, 0, ctx->slot_bytes + TK_BYTES_PAD));
        s.d.raw2_bytes = ctx->slot_bytes;
        TK_CUDA(cudaMalloc(&s.d.rb_units, sizeof(tk_rb_unit) * rb_max_units(ctx)));
    }
    return 0;
}

// zstd frames without a content size and LZ4 frames that are one unit are placed in the decompressed
// buffer by their upper bound (128 KiB per compressed zstd block; the LZ4 frame's maximal block size),
// which saves the size pass -- as long as the bounds of the slot
// fit a buffer of at most RB_BOUND_FACTOR x slot_bytes (below the 2 GiB the copy descriptors
// address); the buffer grows to what the largest slot so far needed.  Otherwise (or with
// TK_ZSTD_SIZE_PASS=1 in the environment, for tests and comparisons) the units go back to the
// size pass.
#define RB_BOUND_FACTOR 16ull
static int rb_place_by_bounds(tk_ctx* ctx, tk_slot& s, tk_rb_unit* units, uint32_t n_units) {
    uint64_t need = 0;
    bool any = false;
    for (uint32_t i = 0; i < n_units; ++i) {
        if (units[i].kind & (RB_UNIT_SIZED | RB_UNIT_BOUND)) need += units[i].unc_len;
        else need += 255ull * units[i].src_len;              // LZ4 unit before its size pass: worst case
        any |= (units[i].kind & RB_UNIT_BOUND) != 0;
    }
    if (!any) return 0;
    const char* force = getenv("TK_ZSTD_SIZE_PASS");
    const bool size_pass = force && force[0] == '1';
    if (!size_pass && need <= s.d.raw2_bytes) return 0;
    const uint64_t limit = min(RB_BOUND_FACTOR * ctx->slot_bytes, 0x7F000000ull);
    if (!size_pass && need <= limit) {
        std::lock_guard<std::mutex> lock(g_rb_mutex);
        const uint64_t bytes = min(limit, need + need / 4);
        TK_CUDA(cudaFree(s.d.raw2));                          // (waits for the device: nobody reads the old buffer)
        s.d.raw2 = nullptr;
        TK_CUDA(cudaMalloc(&s.d.raw2, bytes + TK_BYTES_PAD));
        TK_CUDA(cudaMemset(s.d.raw2, 0, bytes + TK_BYTES_PAD));
        s.d.raw2_bytes = bytes;
        return 0;
    }
    for (uint32_t i = 0; i < n_units; ++i)
        if (units[i].kind & RB_UNIT_BOUND) { units[i].kind &= ~RB_UNIT_BOUND; units[i].unc_len = 0; }
    return 0;
}

static int submit_recordbatches(tk_ctx* ctx, int slot, uint64_t n_bytes, const tk_blob_t* blobs, uint32_t n_blobs,
                                const uint8_t* const* ext, uint32_t* n_records);

extern "C" int tk_slot_submit_recordbatches(tk_ctx* ctx, int slot, uint64_t n_bytes, const tk_blob_t* blobs,
                                            uint32_t n_blobs, uint32_t* n_records) {
    return submit_recordbatches(ctx, slot, n_bytes, blobs, n_blobs, nullptr, n_records);
}

extern "C" int tk_slot_submit_recordbatches_ext(tk_ctx* ctx, int slot, uint64_t n_bytes, const tk_blob_t* blobs,
                                                const uint8_t* const* host_ptrs, uint32_t n_blobs,
                                                uint32_t* n_records) {
    TK_REQUIRE(host_ptrs || !n_blobs, "host_ptrs is null");
    for (uint32_t k = 0; k < n_blobs; ++k) TK_REQUIRE(host_ptrs[k] || !blobs[k].length, "blob %u has no host pointer", k);
    return submit_recordbatches(ctx, slot, n_bytes, blobs, n_blobs, host_ptrs, n_records);
}

static int submit_recordbatches(tk_ctx* ctx, int slot, uint64_t n_bytes, const tk_blob_t* blobs, uint32_t n_blobs,
                                const uint8_t* const* ext, uint32_t* n_records) {
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
    uint32_t n_batches = 0, n_rec = 0, n_units = 0, n_coded = 0;
    rc = rb_walk_host(s.h_bytes, ext, blobs, n_blobs, s.h_rb_batches, ctx->slot_records, ctx->slot_records, &n_batches,
                      &n_rec, s.h_rb_units, (uint32_t)rb_max_units(ctx), &n_units, &n_coded);
    if (rc) return rc;
    if (n_coded) {
        rc = rb_ensure_codecs(ctx); if (rc) return rc;
        rc = rb_place_by_bounds(ctx, s, s.h_rb_units, n_units); if (rc) return rc;
    }
    s.n_records = n_rec;
    s.n_bytes = n_bytes;                     // upper bound of the value bytes (sizes grids only)
    if (n_records) *n_records = n_rec;
    // copies on the copy stream, kernels on the slot's own stream: the H2D of the next slot and
    // the kernels of the other slots run while this one is decompressed and split
    cudaStream_t ks = s.split_stream;
    if (s.free_recorded) {
        TK_CUDA(cudaStreamWaitEvent(st, s.ev_free, 0));
        TK_CUDA(cudaStreamWaitEvent(ks, s.ev_free, 0));
    }
    if (ext) {
        // every blob straight out of the caller's (page-locked) buffer to its place in the device buffer;
        // blobs that follow one another in the caller's buffer as they do in the slot (a consumer
        // that receives a whole fetch response into one arena) go as one copy
        for (uint32_t k = 0; k < n_blobs;) {
            uint32_t j = k + 1;
            uint64_t len = blobs[k].length;
            while (j < n_blobs && ext[k] + len == ext[j] && (uint64_t)blobs[k].start + len == blobs[j].start) {
                len += blobs[j].length;
                ++j;
            }
            if (len) TK_CUDA(cudaMemcpyAsync(s.d.raw + blobs[k].start, ext[k], len, cudaMemcpyHostToDevice, st));
            k = j;
        }
    } else if (n_bytes) {
        TK_CUDA(cudaMemcpyAsync(s.d.raw, s.h_bytes, n_bytes, cudaMemcpyHostToDevice, st));
    }
    if (n_batches)
        TK_CUDA(cudaMemcpyAsync(s.d.rb_batches, s.h_rb_batches, sizeof(tk_rb_batch) * n_batches, cudaMemcpyHostToDevice, st));
    if (n_units)
        TK_CUDA(cudaMemcpyAsync(s.d.rb_units, s.h_rb_units, sizeof(tk_rb_unit) * n_units, cudaMemcpyHostToDevice, st));
    TK_CUDA(cudaEventRecord(s.ev_raw, st));
    TK_CUDA(cudaStreamWaitEvent(ks, s.ev_raw, 0));
    rb_args a;
    a.raw = s.d.raw; a.raw2 = s.d.raw2; a.raw2_bytes = s.d.raw2_bytes; a.batches = s.d.rb_batches; a.n_batches = n_batches; a.n_records = n_rec;
    a.units = s.d.rb_units; a.n_units = n_units;
    a.summary = s.d.rb_summary; a.temp = s.d.rb_temp; a.bigq = s.d.rb_bigq; a.bytes = s.d.bytes; a.recs = s.d.recs;
    a.slot_records = ctx->slot_records; a.slot_bytes = ctx->slot_bytes;
    if (n_coded) { rc = tk_codec_launch(a, s.h_rb_units, ks); if (rc) return rc; }
    k_rb_check<<<148 * 4, 256, 0, ks>>>(a);
    k_rb_prefix<<<1, 1024, 0, ks>>>(a);
    k_rb_copy<<<148 * 8, 256, 0, ks>>>(a);
    k_rb_copy_big<<<148 * 8, 256, 0, ks>>>(a);
    TK_LAUNCH_CHECK();
    TK_CUDA(cudaMemcpyAsync(s.h_rb_summary, s.d.rb_summary, sizeof(tk_rb_summary_t), cudaMemcpyDeviceToHost, ks));
    st = ks;
    TK_CUDA(cudaEventRecord(s.ev_h2d, st));
    return 0;
}

extern "C" int tk_slot_recordbatch_result(tk_ctx* ctx, int slot, tk_rb_summary_t* summary) {
    TK_ENTER(ctx, slot);
    tk_slot& s = ctx->slots[slot];
    TK_REQUIRE(s.h_rb_summary, "no record batches were submitted on this context");
    TK_CUDA(cudaEventSynchronize(s.ev_h2d));
    if (summary) *summary = *s.h_rb_summary;
    return 0;
}

extern "C" int tk_slot_record(tk_ctx* ctx, int slot, uint32_t index, tk_rec_t* out) {
    TK_ENTER(ctx, slot);
    TK_REQUIRE(out && index <= ctx->slot_records, "bad record index %u", index);
    TK_CUDA(cudaDeviceSynchronize());
    TK_CUDA(cudaMemcpy(out, ctx->slots[slot].d.recs + index, sizeof(tk_rec_t), cudaMemcpyDeviceToHost));
    return 0;
}
