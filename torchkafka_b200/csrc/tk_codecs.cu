// tk_codecs.cu -- compressed record batches: LZ4 (attributes & 7 == 3), snappy
// (== 2), gzip (== 1) and zstd (== 4) are decompressed on the device, between the H2D copy of the raw fetch
// bytes and the record split of tk_recordbatch.cu.
//
// Replaces kafka-python's `lz4_decode` / `snappy_decode` / `gzip_decode` / `zstd_decode` (kafka/codec.py, called from
// DefaultRecordBatch._maybe_uncompress in kafka/record/default_records.py; dependency of the
// reference, setup.py:9, not in its tree).  Formats: lz4_Frame_format.md and lz4_Block_format.md of
// the LZ4 project; format_description.txt of snappy and the xerial framing of snappy-java
// (8-byte magic "\x82SNAPPY\0", two int32, then {int32 length, snappy block} chunks).  The xxHash
// checksums of LZ4 frames are not verified: the batch's CRC-32C covers the same bytes.  gzip: RFC
// 1952 member header, DEFLATE (RFC 1951, tk_inflate.cuh), CRC-32 (not verified, same reason) and
// ISIZE trailer; one member per batch (what java.util.zip.GZIPOutputStream writes).  zstd: RFC 8878
// (tk_zstd.cuh); frames usually carry no content size, so they take the size pass.
//
// Work split:
//   host    tk_codec_units_host: follows the codec's framing (a few header reads per batch) and
//           cuts every batch into independently decodable *units*: LZ4 blocks of a
//           block-independent frame (what Kafka's Java client writes), the whole block chain of a
//           block-dependent frame (liblz4's default), stored blocks, snappy blocks (xerial chunks),
//           the DEFLATE stream of a gzip member (its size is the ISIZE trailer), the zstd frames
//           of a batch (block headers walked; sized when every frame states its content size)
//   device  k_unit_size    warp per LZ4 unit: walks the sequences without copying -> decompressed
//                          size (LZ4 writes it nowhere; snappy blocks start with it)
//           k_unit_place   one block: exclusive scan of the sizes -> place of every unit and of
//                          every batch's records in the decompressed buffer
//           k_unit_decode  warp per unit
// DEFLATE and zstd are chains of dependent bit fields: lane 0 walks them (tables in shared memory),
// the warp copies literal runs and matches (k_unit_size_zstd / k_unit_decode_serial).  LZ4 / snappy decoders: every lane runs the same element parse on bytes fetched by shuffle from a 128-byte
// window of the compressed stream held by the warp (one coalesced load per ~100 bytes consumed:
// the fields of an element depend on one another, a cache round trip each would dominate);
// literals and matches are copied by the whole warp, overlapping matches by repeating the source
// with period `offset`; matches are read back from the unit's own output through L1.
#include <algorithm>
#include <type_traits>
#include "tk_rb.h"
#include "tk_inflate.cuh"
#include "tk_zstd.cuh"

// ---------------------------------------------------------------- host: units
static inline uint32_t cu_le32(const uint8_t* p) {
    return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
}
static inline uint32_t cu_be32(const uint8_t* p) {
    return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3];
}

struct unit_sink {
    tk_rb_unit* units;
    uint32_t max_units, n, batch;
    bool full;
    void add(uint32_t src, uint32_t len, uint32_t kind, uint32_t unc_len) {
        if (n >= max_units) { full = true; return; }
        tk_rb_unit& u = units[n++];
        u.src = src; u.src_len = len; u.batch = batch; u.kind = kind; u.unc_len = unc_len; u.out = 0;
    }
};

// LZ4 frame at [p, end) -> units; returns a tk_status_t
static uint32_t lz4_units(const uint8_t* raw, uint32_t p, uint32_t end, unit_sink& sink) {
    if (end - p < 7u || cu_le32(raw + p) != 0x184D2204u) return TK_STATUS_CORRUPT;
    const uint32_t flg = raw[p + 4];
    if ((flg >> 6) != 1u) return TK_STATUS_CORRUPT;
    if (flg & 1u) return TK_STATUS_UNSUPPORTED;                              // dictionary
    const bool independent = flg & 0x20u, block_checksum = flg & 0x10u, content_checksum = flg & 0x04u;
    // BD byte: the largest block the frame may hold decompressed (64 KB .. 4 MB).  A batch whose
    // frame is one unit (one block, or a chain) is placed by that bound and needs no size pass.
    const uint32_t bd = (raw[p + 5] >> 4) & 7u;
    if (bd < 4u) return TK_STATUS_CORRUPT;
    const uint64_t block_max = 1ull << (8u + 2u * bd);
    p += 6u + ((flg & 0x08u) ? 8u : 0u) + 1u;                                // content size, header checksum
    if (p > end) return TK_STATUS_CORRUPT;
    if (!independent) {
        // blocks may refer to the output of earlier blocks: one unit decodes the chain in order;
        // its bound follows from a walk over the block headers
        uint64_t bound = 0;
        bool walked = false;
        for (uint32_t q = p; end - q >= 4u;) {
            uint32_t size = cu_le32(raw + q);
            q += 4;
            if (size == 0) { walked = true; break; }
            const bool stored = size >> 31;
            size &= 0x7FFFFFFFu;
            if (size > end - q) break;
            bound += stored ? size : std::min<uint64_t>(block_max, 255ull * size);
            q += size + (block_checksum ? 4u : 0u);
            if (q > end) break;
        }
        const bool bounded = walked && bound <= 0x7FFFFFFFull;     // (a broken chain is found by the decoder)
        sink.add(p, end - p, RB_UNIT_LZ4_CHAIN | (block_checksum ? RB_UNIT_BLOCK_CHECKSUM : 0u) |
                                 (content_checksum ? RB_UNIT_CONTENT_CHECKSUM : 0u) | (bounded ? RB_UNIT_BOUND : 0u),
                 bounded ? (uint32_t)bound : 0u);
        return TK_STATUS_OK;
    }
    const uint32_t first_unit = sink.n;
    for (;;) {
        if (end - p < 4u) return TK_STATUS_CORRUPT;
        uint32_t size = cu_le32(raw + p);
        p += 4;
        if (size == 0) break;
        const bool stored = size >> 31;
        size &= 0x7FFFFFFFu;
        if (size > end - p) return TK_STATUS_CORRUPT;
        if (stored) sink.add(p, size, RB_UNIT_STORED | RB_UNIT_SIZED, size);
        else sink.add(p, size, RB_UNIT_LZ4_BLOCK, 0);
        p += size + (block_checksum ? 4u : 0u);
        if (p > end) return TK_STATUS_CORRUPT;
    }
    // Every compressed block is placed by its bound, the frame's maximal block size: writers fill
    // every block but the last (Kafka's KafkaLZ4BlockOutputStream, LZ4F_compressUpdate), so the blocks
    // land back to back; k_unit_close moves them together where one was shorter.
    if (!sink.full)
        for (uint32_t i = first_unit; i < sink.n; ++i) {
            tk_rb_unit& u = sink.units[i];
            if ((u.kind & 0xFFu) != RB_UNIT_LZ4_BLOCK) continue;
            u.kind |= RB_UNIT_BOUND;
            u.unc_len = (uint32_t)std::min<uint64_t>(block_max, 255ull * u.src_len);
        }
    return p + (content_checksum ? 4u : 0u) == end ? TK_STATUS_OK : TK_STATUS_CORRUPT;
}

// decompressed length a snappy block announces (varint32 preamble); false if malformed
static bool snappy_preamble(const uint8_t* raw, uint32_t p, uint32_t end, uint32_t* len) {
    uint64_t v = 0;
    for (int shift = 0; shift < 35; shift += 7) {
        if (p >= end) return false;
        const uint32_t b = raw[p++];
        v |= (uint64_t)(b & 0x7Fu) << shift;
        if (!(b & 0x80u)) { *len = (uint32_t)v; return v <= 0x7FFFFFFFull; }
    }
    return false;
}

static uint32_t snappy_units(const uint8_t* raw, uint32_t p, uint32_t end, unit_sink& sink) {
    static const uint8_t magic[8] = {0x82, 'S', 'N', 'A', 'P', 'P', 'Y', 0};
    uint32_t len;
    if (end - p >= 16u && memcmp(raw + p, magic, 8) == 0) {                   // xerial framing
        p += 16;
        while (p < end) {
            if (end - p < 4u) return TK_STATUS_CORRUPT;
            const uint32_t size = cu_be32(raw + p);
            p += 4;
            if (size > end - p || !snappy_preamble(raw, p, p + size, &len)) return TK_STATUS_CORRUPT;
            sink.add(p, size, RB_UNIT_SNAPPY | RB_UNIT_SIZED, len);
            p += size;
        }
        return TK_STATUS_OK;
    }
    if (!snappy_preamble(raw, p, end, &len)) return TK_STATUS_CORRUPT;
    sink.add(p, end - p, RB_UNIT_SNAPPY | RB_UNIT_SIZED, len);
    return TK_STATUS_OK;
}

// gzip member at [p, end): header (RFC 1952), DEFLATE stream, CRC-32, ISIZE
static uint32_t gzip_units(const uint8_t* raw, uint32_t p, uint32_t end, unit_sink& sink) {
    if (end - p < 18u || raw[p] != 0x1Fu || raw[p + 1] != 0x8Bu) return TK_STATUS_CORRUPT;
    if (raw[p + 2] != 8u) return TK_STATUS_UNSUPPORTED;                      // compression method
    const uint32_t flg = raw[p + 3];
    if (flg & 0xE0u) return TK_STATUS_CORRUPT;                               // reserved bits
    p += 10;
    if (flg & 4u) {                                                          // FEXTRA
        if (end - p < 2u) return TK_STATUS_CORRUPT;
        const uint32_t xlen = raw[p] | ((uint32_t)raw[p + 1] << 8);
        if (xlen > end - p - 2u) return TK_STATUS_CORRUPT;
        p += 2u + xlen;
    }
    for (uint32_t bit = 8u; bit <= 16u; bit <<= 1)                           // FNAME, FCOMMENT: zero-terminated
        if (flg & bit) {
            while (p < end && raw[p]) ++p;
            if (p >= end) return TK_STATUS_CORRUPT;
            ++p;
        }
    if (flg & 2u) p += 2;                                                    // FHCRC
    if (p > end || end - p < 8u) return TK_STATUS_CORRUPT;
    sink.add(p, end - 8u - p, RB_UNIT_DEFLATE | RB_UNIT_SIZED, cu_le32(raw + end - 4));
    return TK_STATUS_OK;
}

// zstd frames at [p, end): frame headers and block headers are followed (a few reads per 128 KiB);
// the unit is the whole payload, sized when every frame announces its content size
static uint32_t zstd_units(const uint8_t* raw, uint32_t p, uint32_t end, unit_sink& sink) {
    const uint32_t start = p;
    uint64_t total = 0, bound = 0;
    bool sized = true;
    uint32_t frames = 0;
    while (p < end) {
        if (end - p < 4u) return TK_STATUS_CORRUPT;
        const uint32_t magic = cu_le32(raw + p);
        if ((magic & 0xFFFFFFF0u) == 0x184D2A50u) {                          // skippable frame
            if (end - p < 8u || cu_le32(raw + p + 4) > end - p - 8u) return TK_STATUS_CORRUPT;
            p += 8u + cu_le32(raw + p + 4);
            continue;
        }
        if (magic != 0xFD2FB528u || end - p < 5u) return TK_STATUS_CORRUPT;
        const uint32_t fhd = raw[p + 4];
        if (fhd & 0x08u) return TK_STATUS_CORRUPT;
        const uint32_t single = (fhd >> 5) & 1u, fcs = fhd >> 6, did = fhd & 3u;
        const uint32_t did_bytes = did == 3u ? 4u : did;
        const uint32_t fcs_bytes = fcs == 0 ? single : fcs == 1u ? 2u : fcs == 2u ? 4u : 8u;
        p += 5u;
        const uint32_t head = (single ? 0u : 1u) + did_bytes + fcs_bytes;
        if (head > end - p) return TK_STATUS_CORRUPT;
        const uint8_t* f = raw + p + (single ? 0u : 1u);
        for (uint32_t k = 0; k < did_bytes; ++k)
            if (f[k]) return TK_STATUS_UNSUPPORTED;                          // dictionary
        f += did_bytes;
        if (!fcs_bytes) sized = false;
        else {
            uint64_t v = 0;
            for (uint32_t k = 0; k < fcs_bytes; ++k) v |= (uint64_t)f[k] << (8u * k);
            total += v + (fcs_bytes == 2u ? 256u : 0u);
        }
        p += head;
        for (;;) {                                                           // blocks
            if (end - p < 3u) return TK_STATUS_CORRUPT;
            const uint32_t bh = raw[p] | ((uint32_t)raw[p + 1] << 8) | ((uint32_t)raw[p + 2] << 16);
            p += 3u;
            const uint32_t type = (bh >> 1) & 3u, size = bh >> 3;
            if (type == 3u) return TK_STATUS_CORRUPT;
            const uint32_t body = type == 1u ? 1u : size;
            if (body > end - p) return TK_STATUS_CORRUPT;
            p += body;
            bound += type == 2u ? TK_ZS_BLOCK_MAX : size;                    // a compressed block holds at most 128 KiB
            if (bh & 1u) break;
        }
        if (fhd & 0x04u) {                                                   // content checksum
            if (end - p < 4u) return TK_STATUS_CORRUPT;
            p += 4u;
        }
        ++frames;
    }
    if (total > 0x7FFFFFFFull) return TK_STATUS_UNSUPPORTED;
    sized = sized && frames > 0;
    // without a content size: the upper bound (the caller decides between it and the size pass)
    const bool bounded = !sized && bound <= 0x7FFFFFFFull;
    sink.add(start, end - start, RB_UNIT_ZSTD | (sized ? RB_UNIT_SIZED : bounded ? RB_UNIT_BOUND : 0u),
             sized ? (uint32_t)total : bounded ? (uint32_t)bound : 0u);
    return TK_STATUS_OK;
}

int tk_codec_units_host(const uint8_t* raw, tk_rb_batch& b, uint32_t batch_index, tk_rb_unit* units,
                        uint32_t max_units, uint32_t* n_units) {
    unit_sink sink = {units, max_units, *n_units, batch_index, false};
    const uint32_t first = sink.n;
    const uint32_t status = b.codec == 3u ? lz4_units(raw, b.start + RB_HEADER, b.end, sink)
                          : b.codec == 2u ? snappy_units(raw, b.start + RB_HEADER, b.end, sink)
                          : b.codec == 1u ? gzip_units(raw, b.start + RB_HEADER, b.end, sink)
                                          : zstd_units(raw, b.start + RB_HEADER, b.end, sink);
    TK_REQUIRE(!sink.full, "more compressed blocks in the slot than the unit table holds (%u)", max_units);
    if (status) {
        b.status = status;
        sink.n = first;
    }
    b.unit_first = first;
    b.unit_count = sink.n - first;
    *n_units = sink.n;
    return 0;
}

// ---------------------------------------------------------------- device: the shuffled window
struct cz_window {
    const uint8_t* raw;
    uint32_t at;          // position of the window's first byte (multiple of 4)
    uint32_t word;        // this lane's 4 bytes
};
__device__ __forceinline__ void cz_load(cz_window& w, uint32_t pos, uint32_t lane) {
    w.at = pos & ~3u;
    w.word = *reinterpret_cast<const uint32_t*>(w.raw + w.at + 4u * lane);    // inside the padded buffer
}
// byte at `pos` (same in every lane)
__device__ __forceinline__ uint32_t cz_byte(cz_window& w, uint32_t pos, uint32_t lane) {
    if (pos - w.at >= 128u) cz_load(w, pos, lane);
    const uint32_t j = pos - w.at;
    return (__shfl_sync(0xffffffffu, w.word, j >> 2) >> ((j & 3u) * 8u)) & 0xFFu;
}
// out[op, op + n) = raw[from, from + n): from the window when it holds them, else from memory
__device__ __forceinline__ void cz_literals(cz_window& w, uint32_t lane, uint32_t from, uint32_t n, uint8_t* out,
                                            uint32_t op) {
    if (n <= 32u && from - w.at + 32u <= 128u) {
        const uint32_t j = from - w.at + lane;
        const uint8_t v = (uint8_t)(__shfl_sync(0xffffffffu, w.word, (j >> 2) & 31u) >> ((j & 3u) * 8u));
        if (lane < n) out[op + lane] = v;
    } else {
        for (uint32_t k = lane; k < n; k += 32) out[op + k] = w.raw[from + k];
    }
}
// out[op + k] = out[op - offset + k], k < n.  Where the match overlaps its own output the source
// repeats with period `offset`, so every byte comes from before `op`: the lanes never read what
// this match writes.  The bytes were written by other lanes of this warp: __syncwarp orders them,
// and they are read back through the L1 of the same SM (the region belongs to this warp alone).
__device__ __forceinline__ void cz_match(uint8_t* out, uint32_t op, uint32_t offset, uint32_t n, uint32_t lane) {
    __syncwarp();
    const uint8_t* src = out + op - offset;
    if (offset >= n) {
        for (uint32_t k = lane; k < n; k += 32) out[op + k] = src[k];
    } else {
        for (uint32_t k = lane; k < n; k += 32) out[op + k] = src[k % offset];
    }
    __syncwarp();
}

// ---------------------------------------------------------------- LZ4
// length with its 255-extension bytes; false if it runs past `end`
__device__ __forceinline__ bool lz4_length(cz_window& w, uint32_t lane, uint32_t& p, uint32_t end, uint32_t nibble,
                                           uint32_t& len) {
    len = nibble;
    if (nibble == 15u) {
        uint32_t b;
        do {
            if (p >= end) return false;
            b = cz_byte(w, p++, lane);
            len += b;
            if (len > 0x7FFFFFFFu) return false;
        } while (b == 255u);
    }
    return true;
}

// A run of at most 32 bytes whose source is not written by anything that runs beside it: all loads
// of a chunk are issued before its stores (the output is read back through L2: one round trip per
// chunk, not per byte).
__device__ __forceinline__ void cz_copy_small(uint8_t* dst, const uint8_t* src, uint32_t len) {
    for (uint32_t k0 = 0; k0 < len; k0 += 8) {
        uint8_t v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = k0 + j < len ? src[k0 + j] : (uint8_t)0;
#pragma unroll
        for (int j = 0; j < 8; ++j) if (k0 + j < len) dst[k0 + j] = v[j];
    }
}

// The copies of up to 32 sequences, lane per sequence (ll literal bytes from raw[lsrc..] to out[dst..],
// then a match of ml bytes at distance off): every literal run side by side, then the matches that
// read nothing an earlier match of the batch writes (source end not behind the first match
// destination of the batch; most of them), then the others in sequence order, each by the whole warp.
__device__ __forceinline__ void cz_exec_seqs(const uint8_t* raw, uint8_t* out, uint32_t lane, uint32_t ll, uint32_t lsrc,
                                             uint32_t dst, uint32_t ml, uint32_t off) {
    uint32_t longs = __ballot_sync(0xffffffffu, ll > 32u);
    if (ll && ll <= 32u) cz_copy_small(out + dst, raw + lsrc, ll);
    while (longs) {
        const int L = __ffs((int)longs) - 1;
        longs &= longs - 1u;
        const uint32_t sL = __shfl_sync(0xffffffffu, lsrc, L), dL = __shfl_sync(0xffffffffu, dst, L);
        const uint32_t nL = __shfl_sync(0xffffffffu, ll, L);
        for (uint32_t k = lane; k < nL; k += 32) out[dL + k] = raw[sL + k];
    }
    __syncwarp();
    const uint32_t mdst = dst + ll;
    const uint32_t has = __ballot_sync(0xffffffffu, ml > 0u);
    if (!has) return;
    const uint32_t first_dst = __shfl_sync(0xffffffffu, mdst, __ffs((int)has) - 1);
    // (a match that overlaps its own destination has its source end behind its destination too)
    const bool dep = ml && (mdst - off + ml > first_dst || ml > 32u);
    if (ml && !dep) cz_copy_small(out + mdst, out + mdst - off, ml);
    uint32_t deps = __ballot_sync(0xffffffffu, dep);
    while (deps) {
        const int L = __ffs((int)deps) - 1;
        deps &= deps - 1u;
        cz_match(out, __shfl_sync(0xffffffffu, mdst, L), __shfl_sync(0xffffffffu, off, L), __shfl_sync(0xffffffffu, ml, L), lane);
    }
    __syncwarp();
}

// One sequence at p, general form (every field through the checked window): literal run [*lit_at,
// + *lit), then -- unless the block ends with the literals (*last) -- offset and match length.
__device__ __forceinline__ uint32_t lz4_parse_general(cz_window& w, uint32_t lane, uint32_t& p, uint32_t bend, uint32_t* lit,
                                                      uint32_t* lit_at, uint32_t* offset, uint32_t* mlen, bool* last) {
    const uint32_t token = cz_byte(w, p, lane);
    uint32_t at = p + 1, n;
    if (!lz4_length(w, lane, at, bend, token >> 4, n) || n > bend - at) return TK_STATUS_CORRUPT;
    *lit = n; *lit_at = at;
    p = at + n;
    *offset = 1; *mlen = 0;
    if (p >= bend) { *last = true; return TK_STATUS_OK; }       // the last sequence has no match
    if (bend - p < 2u) return TK_STATUS_CORRUPT;
    *offset = cz_byte(w, p, lane) | (cz_byte(w, p + 1, lane) << 8);
    p += 2;
    uint32_t m;
    if (!lz4_length(w, lane, p, bend, token & 15u, m)) return TK_STATUS_CORRUPT;
    *mlen = m + 4u;
    return TK_STATUS_OK;
}

// The sequences of one block [p, bend).  EMIT = false: only counts (op advances); EMIT = true: writes
// out[op...] (capacity cap); matches may reach back to out[window].
// The fields of a sequence depend on one another but on nothing the decoder writes, so the warp
// first parses up to 32 sequences (every lane the same parse on bytes taken from the shuffled
// window; lane e keeps sequence e), then copies them side by side (cz_exec_seqs).  Fast lane of the
// parse (literal run < 15, the usual case in text): lane i takes byte p + i out of the window with
// one shuffle; the token is lane 0's byte, the offset the bytes of lanes lit + 1, lit + 2.
template <bool EMIT>
__device__ __forceinline__ uint32_t lz4_block_run(cz_window& w, uint32_t lane, uint32_t p, uint32_t bend, uint8_t* out,
                                                  uint32_t cap, uint32_t window, uint32_t& op) {
    bool block_done = false;
    while (p < bend && !block_done) {
        uint32_t my_ll = 0, my_lsrc = 0, my_dst = 0, my_ml = 0, my_off = 1;
        uint32_t n = 0;
        for (; n < 32u && p < bend; ++n) {
            uint32_t lit, lit_at, offset, mlen;
            bool last = false;
            if (p - w.at > 96u) cz_load(w, p, lane);              // [p, p + 32) is inside the window
            const uint32_t jj = p - w.at + lane;
            const uint32_t mine = (__shfl_sync(0xffffffffu, w.word, jj >> 2) >> ((jj & 3u) * 8u)) & 0xFFu;
            const uint32_t token = __shfl_sync(0xffffffffu, mine, 0);
            if ((token >> 4) == 15u || bend - p < 32u) {          // long literal run / end of the block
                const uint32_t st = lz4_parse_general(w, lane, p, bend, &lit, &lit_at, &offset, &mlen, &last);
                if (st) return st;
            } else {
                lit = token >> 4;
                lit_at = p + 1u;
                offset = __shfl_sync(0xffffffffu, mine, lit + 1u) | (__shfl_sync(0xffffffffu, mine, lit + 2u) << 8);
                p += lit + 3u;
                if (!lz4_length(w, lane, p, bend, token & 15u, mlen)) return TK_STATUS_CORRUPT;
                mlen += 4u;
            }
            if (lit > cap - op) return EMIT ? TK_STATUS_CORRUPT : TK_STATUS_UNSUPPORTED;
            if (!last) {
                if (offset == 0 || offset > op + lit - window) return TK_STATUS_CORRUPT;
                if (mlen > cap - op - lit) return EMIT ? TK_STATUS_CORRUPT : TK_STATUS_UNSUPPORTED;
            }
            if (lane == n) { my_ll = lit; my_lsrc = lit_at; my_dst = op; my_ml = mlen; my_off = offset; }
            op += lit + mlen;
            if (last) { block_done = true; ++n; break; }
        }
        if (EMIT) cz_exec_seqs(w.raw, out, lane, my_ll, my_lsrc, my_dst, my_ml, my_off);
    }
    return TK_STATUS_OK;
}

// The block chain of a block-dependent frame: {u32 size, data, [checksum]} ... end mark, [checksum]
template <bool EMIT>
__device__ __forceinline__ uint32_t lz4_chain_run(cz_window& w, uint32_t lane, uint32_t p, uint32_t end, uint32_t kind,
                                                  uint8_t* out, uint32_t cap, uint32_t& op) {
    for (;;) {
        if (end - p < 4u) return TK_STATUS_CORRUPT;
        uint32_t size = cz_byte(w, p, lane) | (cz_byte(w, p + 1, lane) << 8) | (cz_byte(w, p + 2, lane) << 16) |
                        (cz_byte(w, p + 3, lane) << 24);
        p += 4;
        if (size == 0) break;
        const bool stored = size >> 31;
        size &= 0x7FFFFFFFu;
        if (size > end - p) return TK_STATUS_CORRUPT;
        if (stored) {
            if (size > cap - op) return EMIT ? TK_STATUS_CORRUPT : TK_STATUS_UNSUPPORTED;
            if (EMIT) {
                for (uint32_t k = lane; k < size; k += 32) out[op + k] = w.raw[p + k];
                __syncwarp();
            }
            op += size;
        } else {
            const uint32_t st = lz4_block_run<EMIT>(w, lane, p, p + size, out, cap, 0u, op);
            if (st) return st;
        }
        p += size + ((kind & RB_UNIT_BLOCK_CHECKSUM) ? 4u : 0u);
        if (p > end) return TK_STATUS_CORRUPT;
    }
    return p + ((kind & RB_UNIT_CONTENT_CHECKSUM) ? 4u : 0u) == end ? TK_STATUS_OK : TK_STATUS_CORRUPT;
}

// ---------------------------------------------------------------- snappy
// One block: varint32 length, then elements; tag & 3: 0 literal (length - 1 in the upper 6 bits, 60..63:
// in the next 1..4 bytes), 1 copy (length 4..11, 11-bit offset), 2 copy (length 1..64, 16-bit offset),
// 3 copy (32-bit offset).
__device__ __forceinline__ uint32_t snappy_run(cz_window& w, uint32_t lane, uint32_t p, uint32_t end, uint8_t* out,
                                               uint32_t cap, uint32_t& op) {
    uint32_t b;
    do {                                                    // preamble (validated on the host)
        if (p >= end) return TK_STATUS_CORRUPT;
        b = cz_byte(w, p++, lane);
    } while (b & 0x80u);
    // As for LZ4: the warp parses up to 32 elements (lane e keeps element e: a literal run or a
    // copy), then copies them side by side (cz_exec_seqs).
    while (p < end) {
        uint32_t my_ll = 0, my_lsrc = 0, my_dst = 0, my_ml = 0, my_off = 1;
        uint32_t n = 0;
        for (; n < 32u && p < end; ++n) {
            uint32_t lit = 0, lit_at = 0, len = 0, offset = 1;
            bool parsed = false;
            if (end - p >= 32u) {
                // fast lane: lane i takes byte p + i with one shuffle; tag = lane 0's byte, a copy's
                // offset the bytes of lanes 1 (, 2)
                if (p - w.at > 96u) cz_load(w, p, lane);
                const uint32_t jj = p - w.at + lane;
                const uint32_t mine = (__shfl_sync(0xffffffffu, w.word, jj >> 2) >> ((jj & 3u) * 8u)) & 0xFFu;
                const uint32_t tag = __shfl_sync(0xffffffffu, mine, 0);
                const uint32_t kind = tag & 3u;
                if (kind == 0u && (tag >> 2) < 60u) {
                    lit = (tag >> 2) + 1u;
                    lit_at = p + 1u;
                    if (lit > end - lit_at) return TK_STATUS_CORRUPT;
                    p = lit_at + lit;
                    parsed = true;
                } else if (kind == 1u || kind == 2u) {
                    const uint32_t b1 = __shfl_sync(0xffffffffu, mine, 1), b2 = __shfl_sync(0xffffffffu, mine, 2);
                    len = kind == 1u ? 4u + ((tag >> 2) & 7u) : 1u + (tag >> 2);
                    offset = kind == 1u ? (((tag >> 5) << 8) | b1) : (b1 | (b2 << 8));
                    p += kind == 1u ? 2u : 3u;
                    parsed = true;
                }
            }
            if (!parsed) {
                const uint32_t tag = cz_byte(w, p++, lane);
                if ((tag & 3u) == 0u) {
                    lit = tag >> 2;
                    if (lit >= 60u) {
                        const uint32_t nb = lit - 59u;
                        if (end - p < nb) return TK_STATUS_CORRUPT;
                        lit = 0;
                        for (uint32_t k = 0; k < nb; ++k) lit |= cz_byte(w, p + k, lane) << (8u * k);
                        p += nb;
                    }
                    if (lit >= 0x7FFFFFFFu) return TK_STATUS_CORRUPT;
                    lit += 1u;
                    if (lit > end - p) return TK_STATUS_CORRUPT;
                    lit_at = p;
                    p += lit;
                } else if ((tag & 3u) == 1u) {
                    if (p >= end) return TK_STATUS_CORRUPT;
                    len = 4u + ((tag >> 2) & 7u);
                    offset = ((tag >> 5) << 8) | cz_byte(w, p, lane);
                    p += 1;
                } else if ((tag & 3u) == 2u) {
                    if (end - p < 2u) return TK_STATUS_CORRUPT;
                    len = 1u + (tag >> 2);
                    offset = cz_byte(w, p, lane) | (cz_byte(w, p + 1, lane) << 8);
                    p += 2;
                } else {
                    if (end - p < 4u) return TK_STATUS_CORRUPT;
                    len = 1u + (tag >> 2);
                    offset = cz_byte(w, p, lane) | (cz_byte(w, p + 1, lane) << 8) | (cz_byte(w, p + 2, lane) << 16) |
                             (cz_byte(w, p + 3, lane) << 24);
                    p += 4;
                }
            }
            if (lit > cap - op) return TK_STATUS_CORRUPT;
            if (len && (offset == 0 || offset > op || len > cap - op)) return TK_STATUS_CORRUPT;
            if (lane == n) { my_ll = lit; my_lsrc = lit_at; my_dst = op; my_ml = len; my_off = offset; }
            op += lit + len;
        }
        cz_exec_seqs(w.raw, out, lane, my_ll, my_lsrc, my_dst, my_ml, my_off);
    }
    return TK_STATUS_OK;
}

#define CZ_SMALL_COPY 12u       // DEFLATE matches of at most this many bytes stay with lane 0

// ---------------------------------------------------------------- DEFLATE (gzip)
// Lane 0 owns the decoder state and runs tk_inflate_next up to the next match; its literals are
// plain stores of lane 0 (ordered before the warp's reads by the __syncwarp in cz_match).
// The matches of a batch (tk_inflate_next), lane per match; the dependent ones by the whole warp in order.
__device__ __forceinline__ void inflate_exec_batch(const tk_inflate_batch& b, uint32_t n, uint8_t* out, uint32_t lane) {
    if (!n) return;
    uint32_t dst = 0, len = 0, dist = 1;
    if (lane < n) { const tk_inflate_match e = b.m[lane]; dst = e.dst; len = e.len; dist = e.dist; }
    const uint32_t first_dst = __shfl_sync(0xffffffffu, dst, 0), src = dst - dist;
    // (a match that overlaps its own destination has its source end behind its destination too)
    const bool dep = len && (src + len > first_dst || len > 32u);
    if (len && !dep) cz_copy_small(out + dst, out + src, len);
    uint32_t deps = __ballot_sync(0xffffffffu, dep);
    while (deps) {
        const int L = __ffs((int)deps) - 1;
        deps &= deps - 1u;
        cz_match(out, __shfl_sync(0xffffffffu, dst, L), __shfl_sync(0xffffffffu, dist, L), __shfl_sync(0xffffffffu, len, L), lane);
    }
    __syncwarp();
}

// Lane 0 walks the stream: literals go straight to their place, matches are collected in batches
// of 32 and copied by the warp.
__device__ __forceinline__ uint32_t inflate_run(const uint8_t* raw, uint32_t lane, uint32_t p, uint32_t end, uint8_t* out,
                                                uint32_t cap, uint32_t& op, tk_inflate_tables& tb) {
    tk_inflate_state s;
    tk_inflate_init(s, raw, p, end);
    for (;;) {
        uint32_t ev = TK_INF_DONE, mlen = 0, mdist = 0;
        if (lane == 0) ev = (uint32_t)tk_inflate_next(s, tb, out, cap, op, &mlen, &mdist, 0, &tb.batch);
        __syncwarp();                                       // the batch and lane 0's literals are written
        ev = __shfl_sync(0xffffffffu, ev, 0);
        if (ev == TK_INF_ERROR) return TK_STATUS_CORRUPT;
        inflate_exec_batch(tb.batch, __shfl_sync(0xffffffffu, s.nb, 0), out, lane);
        s.nb = 0;
        if (ev == TK_INF_DONE) break;
    }
    op = __shfl_sync(0xffffffffu, op, 0);
    // the stream must end where the trailer begins (a second member behind it is not followed)
    const uint32_t used = __shfl_sync(0xffffffffu, tk_inf_consumed(s), 0);
    return used == end ? TK_STATUS_OK : TK_STATUS_CORRUPT;
}

// ---------------------------------------------------------------- kernels
__global__ void __launch_bounds__(256) k_unit_size(const rb_args a) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t i = warp; i < a.n_units; i += nwarps) {
        const tk_rb_unit u = a.units[i];
        if ((u.kind & (RB_UNIT_SIZED | RB_UNIT_BOUND)) || (u.kind & 0xFFu) > RB_UNIT_LZ4_CHAIN) continue;
        cz_window w;
        w.raw = a.raw;
        cz_load(w, u.src, lane);
        uint32_t op = 0, st;
        if ((u.kind & 0xFFu) == RB_UNIT_LZ4_BLOCK)
            st = lz4_block_run<false>(w, lane, u.src, u.src + u.src_len, nullptr, 0x7FFFFFFFu, 0u, op);
        else
            st = lz4_chain_run<false>(w, lane, u.src, u.src + u.src_len, u.kind, nullptr, 0x7FFFFFFFu, op);
        if (lane == 0) {
            a.units[i].unc_len = st ? 0u : op;
            if (st) a.units[i].kind = u.kind | RB_UNIT_FAILED;
        }
    }
}

// One block: place of every unit (exclusive scan of the sizes), then of every batch's records.
__global__ void __launch_bounds__(1024) k_unit_place(const rb_args a) {
    __shared__ unsigned long long warp_tot[32];
    __shared__ unsigned long long s_total, s_run;
    const uint32_t t = threadIdx.x;
    if (t == 0) s_run = 0;
    __syncthreads();
    for (uint32_t base = 0; base < a.n_units; base += blockDim.x) {
        const uint32_t i = base + t;
        const unsigned long long v = i < a.n_units ? a.units[i].unc_len : 0ull;
        const unsigned long long excl = s_run + rb_block_scan(v, warp_tot, &s_total);
        // (past the buffer: the batch is rejected below, the unit is never decoded)
        if (i < a.n_units) a.units[i].out = (uint32_t)min(excl, 0xFFFFFFFFull);
        __syncthreads();
        if (t == 0) s_run += s_total;
        __syncthreads();
    }
    for (uint32_t i = t; i < a.n_batches; i += blockDim.x) {
        tk_rb_batch& b = a.batches[i];
        if (b.status || !b.codec || !b.unit_count) {
            if (!b.status && b.codec) { b.rec_start = 0; b.rec_end = 0; }     // a frame without blocks
            continue;
        }
        unsigned long long total = 0;
        bool failed = false;
        for (uint32_t k = 0; k < b.unit_count; ++k) {
            const tk_rb_unit u = a.units[b.unit_first + k];
            failed |= (u.kind & RB_UNIT_FAILED) != 0;
            total += u.unc_len;
        }
        const unsigned long long first = a.units[b.unit_first].out;
        if (failed) b.status = TK_STATUS_CORRUPT;
        else if (first + total > a.raw2_bytes) b.status = TK_STATUS_SLOT_OVERFLOW;   // does not fit the buffer
        else { b.rec_start = (uint32_t)first; b.rec_end = (uint32_t)(first + total); }
    }
}

__global__ void __launch_bounds__(128) k_unit_decode(const rb_args a) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t i = warp; i < a.n_units; i += nwarps) {
        const tk_rb_unit u = a.units[i];
        const uint32_t kind = u.kind & 0xFFu;
        if (kind >= RB_UNIT_DEFLATE) continue;             // k_unit_decode_serial
        if (a.batches[u.batch].status) continue;           // rejected by the size pass or the placement
        uint8_t* out = a.raw2 + u.out;
        uint32_t op = 0, st = TK_STATUS_OK;
        if (kind == RB_UNIT_STORED) {
            for (uint32_t k = lane; k < u.src_len; k += 32) out[k] = a.raw[u.src + k];
            op = u.src_len;
        } else {
            cz_window w;
            w.raw = a.raw;
            cz_load(w, u.src, lane);
            if (kind == RB_UNIT_LZ4_BLOCK) st = lz4_block_run<true>(w, lane, u.src, u.src + u.src_len, out, u.unc_len, 0u, op);
            else if (kind == RB_UNIT_LZ4_CHAIN) st = lz4_chain_run<true>(w, lane, u.src, u.src + u.src_len, u.kind, out, u.unc_len, op);
            else st = snappy_run(w, lane, u.src, u.src + u.src_len, out, u.unc_len, op);
        }
        if (u.kind & RB_UNIT_BOUND) {                       // placed by its upper bound
            if (!st && lane == 0) {
                // the batch's only unit: the records end here; one of several: k_unit_close adds them up
                if (a.batches[u.batch].unit_count == 1u) a.batches[u.batch].rec_end = a.batches[u.batch].rec_start + op;
                else a.units[i].unc_len = op;
            }
        } else if (!st && op != u.unc_len) {
            st = TK_STATUS_CORRUPT;
        }
        // (k_rb_check runs behind this kernel and turns the status into the batch's error entry)
        if (st && lane == 0) a.batches[u.batch].status = TK_STATUS_CORRUPT;
    }
}

// Batches of several units placed by their bounds (LZ4 frames of several blocks): the units decoded
// to their real sizes (k_unit_decode left them in unc_len); where one was shorter than its bound
// the units behind it move down (destination below source: chunks of 32 x 16 bytes, loaded before
// they are stored), and the batch's record range ends where the last one ends.  Warp per batch.
__global__ void __launch_bounds__(128) k_unit_close(const rb_args a) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t i = warp; i < a.n_batches; i += nwarps) {
        const tk_rb_batch b = a.batches[i];
        if (b.status || b.unit_count < 2u) continue;
        bool bounded = false;                               // (a stored block among them has its exact size)
        for (uint32_t k = 0; k < b.unit_count; ++k) bounded |= (a.units[b.unit_first + k].kind & RB_UNIT_BOUND) != 0;
        if (!bounded) continue;
        uint32_t pos = b.rec_start;
        for (uint32_t k = 0; k < b.unit_count; ++k) {
            const tk_rb_unit u = a.units[b.unit_first + k];
            if (u.out != pos) {
                for (uint32_t base = 0; base < u.unc_len; base += 32u * 16u) {
                    uint8_t v[16];
                    const uint32_t at = base + lane * 16u;
#pragma unroll
                    for (int j = 0; j < 16; ++j) v[j] = at + j < u.unc_len ? a.raw2[u.out + at + j] : (uint8_t)0;
                    __syncwarp();
#pragma unroll
                    for (int j = 0; j < 16; ++j) if (at + j < u.unc_len) a.raw2[pos + at + j] = v[j];
                    __syncwarp();
                }
            }
            pos += u.unc_len;
        }
        if (lane == 0) a.batches[i].rec_end = pos;
    }
}

// ---------------------------------------------------------------- serial formats: gzip, zstd
// Lane 0 owns the decoder state and produces copy events; the warp executes them.  Literals that
// lane 0 decoded itself (Huffman) are parked at the end of the unit's output and move down from
// there: the destination is below the source, so chunks of 32 bytes go load, barrier, store.
// The copies of a batch of sequences (tk_zstd_next_batch), lane per sequence: places by a warp prefix
// sum, every literal run, then the matches that read nothing an earlier match of the batch writes
// (most of them), then the dependent ones in sequence order, each by the whole warp.
__device__ __forceinline__ void zstd_exec_batch(const tk_zstd_batch& b, uint32_t n, const uint8_t* raw, uint8_t* out,
                                                uint32_t& op, uint32_t lane) {
    uint32_t ll = 0, ml = 0, off = 0, src = 0;
    if (lane < n) { const tk_zstd_seq q = b.q[lane]; ll = q.lit_len; ml = q.mlen; off = q.offset; src = q.lit_src; }
    const uint32_t kind = b.lit_kind;
    uint32_t incl = ll + ml;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= (uint32_t)d) incl += o;
    }
    const uint32_t dst = op + incl - (ll + ml);
    const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
    // Parked literals (TK_ZL_OUTPUT) lie ahead of the write position of the sequence that reads
    // them, not of the later sequences' of the batch: near the end of a unit whose output region
    // is exactly its size, sequence e may write where sequence j < e still has to read.  Such a
    // batch (its writes reach the first parked byte it reads) goes in sequence order.
    if (kind == TK_ZL_OUTPUT && op + total > __shfl_sync(0xffffffffu, src, 0)) {
        for (uint32_t e = 0; e < n; ++e) {
            const uint32_t sE = __shfl_sync(0xffffffffu, src, e), dE = __shfl_sync(0xffffffffu, dst, e);
            const uint32_t nE = __shfl_sync(0xffffffffu, ll, e), mE = __shfl_sync(0xffffffffu, ml, e);
            const uint32_t oE = __shfl_sync(0xffffffffu, off, e);
            if (sE != dE) {
                for (uint32_t base = 0; base < nE; base += 32) {      // destination below the source
                    const uint32_t k = base + lane;
                    const uint8_t v = k < nE ? out[sE + k] : 0;
                    __syncwarp();
                    if (k < nE) out[dE + k] = v;
                }
            }
            __syncwarp();
            if (mE) cz_match(out, dE + nE, oE, mE, lane);
            __syncwarp();
        }
        op += total;
        return;
    }
    // literal runs (their sources are the input, a byte, or parked literals ahead of every write
    // position of the batch); long ones by the whole warp
    uint32_t longs = __ballot_sync(0xffffffffu, ll > 32u);
    if (ll <= 32u) {
        if (kind == TK_ZL_INPUT) { for (uint32_t k = 0; k < ll; ++k) out[dst + k] = raw[src + k]; }
        else if (kind == TK_ZL_RLE) { for (uint32_t k = 0; k < ll; ++k) out[dst + k] = (uint8_t)src; }
        else { for (uint32_t k = 0; k < ll; ++k) out[dst + k] = out[src + k]; }
    }
    while (longs) {
        const int L = __ffs((int)longs) - 1;
        longs &= longs - 1u;
        const uint32_t sL = __shfl_sync(0xffffffffu, src, L), dL = __shfl_sync(0xffffffffu, dst, L);
        const uint32_t nL = __shfl_sync(0xffffffffu, ll, L);
        for (uint32_t k = lane; k < nL; k += 32)
            out[dL + k] = kind == TK_ZL_INPUT ? raw[sL + k] : kind == TK_ZL_RLE ? (uint8_t)sL : out[sL + k];
    }
    __syncwarp();
    // matches: one depends on earlier matches of the batch only if its source reaches behind the
    // first match destination of the batch (or into its own destination)
    const uint32_t mdst = dst + ll;
    const uint32_t has = __ballot_sync(0xffffffffu, ml > 0u);
    const uint32_t first_dst = __shfl_sync(0xffffffffu, mdst, has ? __ffs((int)has) - 1 : 0);
    const bool dep = ml && (mdst - off + ml > first_dst || off < ml || ml > 32u);
    if (ml && !dep) cz_copy_small(out + mdst, out + mdst - off, ml);
    uint32_t deps = __ballot_sync(0xffffffffu, dep);
    while (deps) {
        const int L = __ffs((int)deps) - 1;
        deps &= deps - 1u;
        cz_match(out, __shfl_sync(0xffffffffu, mdst, L), __shfl_sync(0xffffffffu, off, L), __shfl_sync(0xffffffffu, ml, L), lane);
    }
    op += __shfl_sync(0xffffffffu, incl, 31);
    __syncwarp();
}

__device__ __forceinline__ uint32_t zstd_run(const uint8_t* raw, uint32_t lane, uint32_t p, uint32_t end, uint8_t* out,
                                             uint32_t cap, uint32_t& op, tk_zstd_tables& tb, tk_zstd_event& ev) {
    tk_zstd_state s;
    if (lane == 0) tk_zstd_init(s, tb, raw, p, end);
    tk_zstd_batch* batch = tk_zstd_batch_of(tb);
    for (;;) {
        // inside a sequence section: up to 32 sequences decoded by lane 0 in one tight loop, their
        // copies executed by the warp side by side
        int nb = 0;
        if (lane == 0) nb = tk_zstd_next_batch(s, tb, cap, op, true, batch);
        __syncwarp();
        nb = __shfl_sync(0xffffffffu, nb, 0);
        if (nb < 0) return nb == -(int)TK_ZS_UNSUPPORTED ? TK_STATUS_UNSUPPORTED : TK_STATUS_CORRUPT;
        if (nb > 0) { zstd_exec_batch(*batch, (uint32_t)nb, raw, out, op, lane); continue; }
        uint32_t st = TK_ZS_DONE;
        if (lane == 0) st = (uint32_t)tk_zstd_next(s, tb, out, cap, op, true, &ev);
        __syncwarp();                                       // the event (shared memory) and the tables are written
        st = __shfl_sync(0xffffffffu, st, 0);
        if (st == TK_ZS_DONE) return TK_STATUS_OK;
        if (st == TK_ZS_HUF) {                              // the block's literals: one Huffman stream per lane
            bool ok = true;
            if (lane < ev.h_streams)
                ok = tk_huf_stream(raw, ev.h_lo[lane], ev.h_hi[lane], tb.huf, (int)ev.h_log, out + ev.h_dst[lane], ev.h_n[lane]);
            if (!__all_sync(0xffffffffu, ok)) return TK_STATUS_CORRUPT;
            __syncwarp();                                   // literals stored, event read by every lane
            continue;
        }
        if (st != TK_ZS_COPY) return st == TK_ZS_UNSUPPORTED ? TK_STATUS_UNSUPPORTED : TK_STATUS_CORRUPT;
        const uint32_t n = ev.lit_len, kind = ev.lit_kind, src = ev.lit_src, mlen = ev.mlen, offset = ev.offset;
        if (n + mlen <= 32u && offset >= n + mlen && (kind != TK_ZL_OUTPUT || src >= op + n + mlen)) {
            // small sequence whose sources (literals, match) all lie outside what it writes: one
            // load and one store per lane, no barrier in between (most sequences of text)
            uint8_t v = 0;
            if (lane < n) v = kind == TK_ZL_INPUT ? raw[src + lane] : kind == TK_ZL_RLE ? (uint8_t)src : out[src + lane];
            else if (lane < n + mlen) v = out[op + lane - offset];
            if (lane < n + mlen) out[op + lane] = v;
            op += n + mlen;
            __syncwarp();
            continue;
        }
        if (kind == TK_ZL_INPUT) {
            for (uint32_t k = lane; k < n; k += 32) out[op + k] = raw[src + k];
        } else if (kind == TK_ZL_RLE) {
            for (uint32_t k = lane; k < n; k += 32) out[op + k] = (uint8_t)src;
        } else if (src != op) {
            for (uint32_t base = 0; base < n; base += 32) {
                const uint32_t k = base + lane;
                const uint8_t v = k < n ? out[src + k] : 0;
                __syncwarp();
                if (k < n) out[op + k] = v;
            }
        }
        op += n;
        if (mlen) cz_match(out, op, offset, mlen, lane);
        op += mlen;
        __syncwarp();                                       // every lane has read the event
    }
}

// Size pass of zstd units without a content size: lane 0 walks the sequences, nothing is written.
#define SERIAL_WARPS 2       // units per block of the serial decoders: small blocks, so that the shared
                             // memory of an SM divides into as many tables as possible
__global__ void __launch_bounds__(32 * SERIAL_WARPS) k_unit_size_zstd(const rb_args a) {
    __shared__ tk_zstd_tables tables[SERIAL_WARPS];
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t i = warp; i < a.n_units; i += nwarps) {
        const tk_rb_unit u = a.units[i];
        if ((u.kind & 0xFFu) != RB_UNIT_ZSTD || (u.kind & (RB_UNIT_SIZED | RB_UNIT_BOUND))) continue;
        if (lane == 0) {
            tk_zstd_tables& tb = tables[threadIdx.x >> 5];
            tk_zstd_state s;
            tk_zstd_init(s, tb, a.raw, u.src, u.src + u.src_len);
            tk_zstd_event ev;
            uint32_t op = 0, st;
            while ((st = (uint32_t)tk_zstd_next(s, tb, nullptr, 0x7FFFFFFFu, op, false, &ev)) == TK_ZS_COPY)
                op += ev.lit_len + ev.mlen;
            a.units[i].unc_len = st == TK_ZS_DONE ? op : 0u;
            if (st != TK_ZS_DONE) a.units[i].kind = u.kind | RB_UNIT_FAILED;
        }
        __syncwarp();
    }
}

// One kernel per format: a DEFLATE unit needs 3.9 KB of tables, a zstd unit 10.8 KB, and the number of
// units an SM works on at a time is what its shared memory divides into.
// (zstd: eleven 2-warp blocks on an SM -- the tables are sized for that, the registers capped to it)
template <bool ZSTD>
__global__ void __launch_bounds__(32 * SERIAL_WARPS, ZSTD ? 11 : 1) k_unit_decode_serial(const rb_args a) {
    __shared__ typename std::conditional<ZSTD, tk_zstd_tables, tk_inflate_tables>::type tables[SERIAL_WARPS];
    __shared__ tk_zstd_event events[ZSTD ? SERIAL_WARPS : 1];
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t i = warp; i < a.n_units; i += nwarps) {
        const tk_rb_unit u = a.units[i];
        const uint32_t kind = u.kind & 0xFFu;
        if (kind != (ZSTD ? RB_UNIT_ZSTD : RB_UNIT_DEFLATE) || a.batches[u.batch].status) continue;
        uint8_t* out = a.raw2 + u.out;
        uint32_t op = 0, st;
        if constexpr (ZSTD)
            st = zstd_run(a.raw, lane, u.src, u.src + u.src_len, out, u.unc_len, op, tables[threadIdx.x >> 5],
                          events[threadIdx.x >> 5]);
        else
            st = inflate_run(a.raw, lane, u.src, u.src + u.src_len, out, u.unc_len, op, tables[threadIdx.x >> 5]);
        if (u.kind & RB_UNIT_BOUND) {                       // placed by its upper bound: the records end here
            if (!st && lane == 0) a.batches[u.batch].rec_end = a.batches[u.batch].rec_start + op;
        } else if (!st && op != u.unc_len) {
            st = TK_STATUS_CORRUPT;
        }
        if (st && lane == 0) a.batches[u.batch].status = st;
        __syncwarp();                                       // the tables are reused by the warp's next unit
    }
}

// resident blocks of a kernel on this device (asked once per kernel)
template <class K>
static uint32_t serial_grid(K kernel, uint32_t* cache) {
    if (!*cache) {
        int per_sm = 0, sms = 0, dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess ||
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, 32 * SERIAL_WARPS, 0) != cudaSuccess || per_sm < 1) {
            (void)cudaGetLastError();
            per_sm = 8; sms = 148;
        }
        *cache = (uint32_t)(per_sm * sms);
    }
    return *cache;
}

int tk_codec_launch(const rb_args& a, const tk_rb_unit* h_units, cudaStream_t stream) {
    // which kernels this slot needs
    uint32_t lz_unsized = 0, zs_unsized = 0, warp_units = 0, deflate_units = 0, zstd_units = 0, lz_bound = 0;
    for (uint32_t i = 0; i < a.n_units; ++i) {
        const uint32_t kind = h_units[i].kind & 0xFFu;
        const bool sized = h_units[i].kind & (RB_UNIT_SIZED | RB_UNIT_BOUND);
        if (kind == RB_UNIT_ZSTD) { ++zstd_units; if (!sized) ++zs_unsized; }
        else if (kind >= RB_UNIT_DEFLATE) ++deflate_units;
        else { ++warp_units; if (!sized) ++lz_unsized; if (h_units[i].kind & RB_UNIT_BOUND) ++lz_bound; }
    }
    if (lz_unsized) k_unit_size<<<min((a.n_units + 7) / 8, 148u * 8u), 256, 0, stream>>>(a);
    static uint32_t g_size = 0, g_inflate = 0, g_zstd = 0;
    const uint32_t blocks_wanted = (a.n_units + SERIAL_WARPS - 1) / SERIAL_WARPS;
    if (zs_unsized)
        k_unit_size_zstd<<<min(blocks_wanted, serial_grid(k_unit_size_zstd, &g_size)), 32 * SERIAL_WARPS, 0, stream>>>(a);
    // (also sets the empty record ranges of compressed batches without a single block)
    k_unit_place<<<1, 1024, 0, stream>>>(a);
    if (warp_units) k_unit_decode<<<min((a.n_units + 3) / 4, 148u * 16u), 128, 0, stream>>>(a);
    // (batches of one bounded unit were closed by the decode kernel itself)
    if (lz_bound) k_unit_close<<<min((a.n_batches + 3) / 4, 148u * 8u), 128, 0, stream>>>(a);
    if (deflate_units)
        k_unit_decode_serial<false><<<min(blocks_wanted, serial_grid(k_unit_decode_serial<false>, &g_inflate)), 32 * SERIAL_WARPS, 0, stream>>>(a);
    if (zstd_units)
        k_unit_decode_serial<true><<<min(blocks_wanted, serial_grid(k_unit_decode_serial<true>, &g_zstd)), 32 * SERIAL_WARPS, 0, stream>>>(a);
    TK_LAUNCH_CHECK();
    return 0;
}
