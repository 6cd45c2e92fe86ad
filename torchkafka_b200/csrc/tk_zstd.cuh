// tk_zstd.cuh -- Zstandard (RFC 8878) frame decoder for zstd-compressed record batches
// (attributes & 7 == 4).  Replaces kafka-python's `zstd_decode` (kafka/codec.py,
// the `zstandard` package), called from DefaultRecordBatch._maybe_uncompress; dependency of the
// reference (setup.py:9), not in its tree.  Format restated from RFC 8878: frames, raw / RLE /
// compressed blocks, Huffman-coded literals (one or four streams, direct or FSE-coded weights,
// tree reuse), FSE-coded sequences (predefined, RLE, described and repeated tables), repeat
// offsets.  Dictionaries are not handled (Kafka producers use none); the optional content
// checksum (xxHash64) is skipped: the batch's CRC-32C covers the same bytes.
//
// Like DEFLATE, a zstd frame is chains of dependent bit fields, so one thread walks it and hands
// the copies -- a run of literals, then a match -- to its caller as events: on the device the
// whole warp executes them (tk_codecs.cu), on the host a plain loop (tests/native, differential
// test against libzstd through pyarrow).  With emit = false nothing is written and the literals
// are not decoded: the events then only carry the sizes (the size pass; Kafka's Java client
// writes frames without a content size).  Host+device code (TK_HD).
#pragma once
#include <stddef.h>
#include <string.h>
#include "tk_number.cuh"      // TK_HD

enum { TK_ZS_COPY = 0, TK_ZS_DONE = 1, TK_ZS_ERROR = 2, TK_ZS_UNSUPPORTED = 3, TK_ZS_HUF = 4 };
enum { TK_ZL_INPUT = 0, TK_ZL_OUTPUT = 1, TK_ZL_RLE = 2 };

// One step of the output: lit_len literal bytes (from the input at lit_src, from the output
// buffer at lit_src -- Huffman-decoded literals are parked at its end --, or lit_len copies of the
// byte lit_src), then mlen bytes from `offset` bytes back.
// TK_ZS_HUF instead: the block's literals are h_streams (1 or 4) independent Huffman streams,
// stream k = in[h_lo[k], h_hi[k]) -> h_n[k] bytes at out + h_dst[k] (tk_huf_stream with the
// decoder's current tree, longest code h_log); the caller decodes them -- the device one per lane.
struct tk_zstd_event {
    uint32_t lit_len, lit_kind, lit_src, mlen, offset;
    uint32_t h_streams, h_log, h_lo[4], h_hi[4], h_dst[4], h_n[4];
};

struct tk_fse_entry {
    uint8_t  symbol, nbits;
    uint16_t base;
};

// code -> baseline << 5 | extra bits of literal lengths (36 codes) and match lengths (53 codes): RFC 8878
// 3.1.1.3.2.1.1.  Constants: read by the one lane that walks a sequence section (constant bank on
// the device).
#define TK_ZS_LL_CODES { 0x0u, 0x20u, 0x40u, 0x60u, 0x80u, 0xA0u, 0xC0u, 0xE0u, 0x100u, 0x120u, 0x140u, 0x160u, 0x180u, 0x1A0u, 0x1C0u, 0x1E0u, 0x201u, 0x241u, 0x281u, 0x2C1u, 0x302u, 0x382u, 0x403u, 0x503u, 0x604u, 0x806u, 0x1007u, 0x2008u, 0x4009u, 0x800Au, 0x1000Bu, 0x2000Cu, 0x4000Du, 0x8000Eu, 0x10000Fu, 0x200010u }
#define TK_ZS_ML_CODES { 0x60u, 0x80u, 0xA0u, 0xC0u, 0xE0u, 0x100u, 0x120u, 0x140u, 0x160u, 0x180u, 0x1A0u, 0x1C0u, 0x1E0u, 0x200u, 0x220u, 0x240u, 0x260u, 0x280u, 0x2A0u, 0x2C0u, 0x2E0u, 0x300u, 0x320u, 0x340u, 0x360u, 0x380u, 0x3A0u, 0x3C0u, 0x3E0u, 0x400u, 0x420u, 0x440u, 0x461u, 0x4A1u, 0x4E1u, 0x521u, 0x562u, 0x5E2u, 0x663u, 0x763u, 0x864u, 0xA64u, 0xC65u, 0x1067u, 0x2068u, 0x4069u, 0x806Au, 0x1006Bu, 0x2006Cu, 0x4006Du, 0x8006Eu, 0x10006Fu, 0x200070u }
#if defined(__CUDACC__)
static __constant__ uint32_t tk_zs_ll_dev[36] = TK_ZS_LL_CODES;
static __constant__ uint32_t tk_zs_ml_dev[53] = TK_ZS_ML_CODES;
#endif
#if !defined(__CUDA_ARCH__)
static const uint32_t tk_zs_ll_host[36] = TK_ZS_LL_CODES;
static const uint32_t tk_zs_ml_host[53] = TK_ZS_ML_CODES;
#endif
TK_HD uint32_t tk_zs_ll_code(uint32_t c) {
#if defined(__CUDA_ARCH__)
    return tk_zs_ll_dev[c];
#else
    return tk_zs_ll_host[c];
#endif
}
TK_HD uint32_t tk_zs_ml_code(uint32_t c) {
#if defined(__CUDA_ARCH__)
    return tk_zs_ml_dev[c];
#else
    return tk_zs_ml_host[c];
#endif
}

// One decoder's tables.  Their size decides how many payloads an SM decodes side by side (shared
// memory), so what is dead when something else is alive shares its place:
//  * the FSE table of the Huffman weights (tk_zstd_wt: 64 entries) lives at the head of `huf`, which
//    is written only after the weights are decoded;
//  * the batch of decoded sequences (tk_zstd_batch_of) lives in the scratch of the table builds.
#define TK_ZS_MAX_SYMBOLS 64       // symbols of a table description: 36 / 53 / 32 for sequences, weights 0..11
struct tk_zstd_tables {
    tk_fse_entry ll[512], ml[512], of[256];   // sequences (accuracy <= 9, 9, 8)
    uint16_t huf[2048];            // literals: symbol | code length << 8, indexed by the next huf_log bits
    // scratch of the table builds (528 bytes, 16-byte aligned)
    int16_t  norm[TK_ZS_MAX_SYMBOLS];   // normalised counts of a table description
    uint16_t desc[TK_ZS_MAX_SYMBOLS];   // per-symbol state counter of the table build
    uint8_t  weights[256];
    uint8_t  scratch_tail[16];
};
TK_HD tk_fse_entry* tk_zstd_wt(tk_zstd_tables& tb) { return reinterpret_cast<tk_fse_entry*>(tb.huf); }

TK_HD int tk_zs_highbit(uint32_t x) {                  // index of the highest set bit, x != 0
#if defined(__CUDA_ARCH__)
    return 31 - __clz((int)x);
#else
    return 31 - __builtin_clz(x);
#endif
}

// n <= 32 bits starting at bit `bitpos` of the little-endian bit array in[lo, hi); bits outside
// the array (bitpos may be negative) read as zero
TK_HD uint32_t tk_zs_bits(const uint8_t* in, uint32_t lo, uint32_t hi, int32_t bitpos, int n) {
    if (n == 0) return 0;
    const int32_t first = bitpos >> 3;
    const int sh = bitpos & 7;
    uint64_t v = 0;
    for (int k = 0; k < 5; ++k) {
        const int32_t b = first + k;
        if (b >= 0 && (uint32_t)b < hi - lo) v |= (uint64_t)in[lo + (uint32_t)b] << (8 * k);
    }
    v >>= sh;
    return n >= 32 ? (uint32_t)v : (uint32_t)v & ((1u << n) - 1u);
}

// Backward bit stream over in[lo, hi): *bitpos = unread bits (the last byte's highest set bit is
// the end mark).  false: empty stream or no mark.
TK_HD bool tk_zs_bwd_init(const uint8_t* in, uint32_t lo, uint32_t hi, int32_t* bitpos) {
    if (hi <= lo || in[hi - 1] == 0) return false;
    *bitpos = (int32_t)(8u * (hi - lo - 1u)) + tk_zs_highbit(in[hi - 1]);
    return true;
}
TK_HD uint32_t tk_zs_bwd_read(const uint8_t* in, uint32_t lo, uint32_t hi, int32_t* bitpos, int n) {
    *bitpos -= n;
    return tk_zs_bits(in, lo, hi, *bitpos, n);
}

// The backward streams are read through a cached 64-bit window of the bit array (one reload per
// ~50 bits instead of five byte loads per field).
struct tk_zs_win {
    uint64_t bits;                 // bits [lo, lo + 64) of the array
    int32_t  lo;
};
TK_HD void tk_zs_win_init(tk_zs_win& w) { w.bits = 0; w.lo = -(1 << 30); }
// bytes [b, b + 8) of the array in[lo, hi) as a little-endian word, bytes outside it read as zero.
// Inside the array (all reloads but the ones at its two ends): three aligned word loads and a
// funnel shift on the device (the compressed bytes sit in a padded, word-aligned buffer), one
// unaligned load on the host; byte by byte only at the ends.
TK_HD uint64_t tk_zs_load64(const uint8_t* in, uint32_t lo, uint32_t hi, int32_t b) {
    if (b >= 0 && (uint32_t)b + 8u <= hi - lo) {
        const uint8_t* a = in + lo + (uint32_t)b;
#if defined(__CUDA_ARCH__)
        const uint32_t* w = reinterpret_cast<const uint32_t*>(reinterpret_cast<uintptr_t>(a) & ~(uintptr_t)3);
        const uint32_t sh = ((uint32_t)reinterpret_cast<uintptr_t>(a) & 3u) * 8u;
        const uint32_t w0 = w[0], w1 = w[1], w2 = sh ? w[2] : 0u;
        return ((uint64_t)__funnelshift_r(w1, w2, sh) << 32) | __funnelshift_r(w0, w1, sh);
#else
        uint64_t v;
        memcpy(&v, a, 8);
        return v;
#endif
    }
    uint64_t v = 0;
    for (int k = 0; k < 8; ++k) {
        const int32_t bk = b + k;
        if (bk >= 0 && (uint32_t)bk < hi - lo) v |= (uint64_t)in[lo + (uint32_t)bk] << (8 * k);
    }
    return v;
}
TK_HD uint32_t tk_zs_win_get(tk_zs_win& w, const uint8_t* in, uint32_t lo, uint32_t hi, int32_t bitpos, int n) {
    if (n == 0) return 0;
    if ((uint32_t)(bitpos - w.lo) > (uint32_t)(64 - n)) {
        // reload so that the window ends just above these bits: the readers move downwards
        const int32_t b = ((bitpos + n + 7) >> 3) - 8;
        w.bits = tk_zs_load64(in, lo, hi, b);
        w.lo = 8 * b;
    }
    const uint32_t v = (uint32_t)(w.bits >> (bitpos - w.lo));
    return n >= 32 ? v : v & ((1u << n) - 1u);
}
// n <= 56 bits at bitpos, with garbage above them (the caller peels its fields off the bottom)
TK_HD uint64_t tk_zs_win_get64(tk_zs_win& w, const uint8_t* in, uint32_t lo, uint32_t hi, int32_t bitpos, int n) {
    if ((uint32_t)(bitpos - w.lo) > (uint32_t)(64 - n)) {
        const int32_t b = ((bitpos + n + 7) >> 3) - 8;
        w.bits = tk_zs_load64(in, lo, hi, b);
        w.lo = 8 * b;
    }
    return w.bits >> (bitpos - w.lo);
}
TK_HD uint32_t tk_zs_win_read(tk_zs_win& w, const uint8_t* in, uint32_t lo, uint32_t hi, int32_t* bitpos, int n) {
    *bitpos -= n;
    return tk_zs_win_get(w, in, lo, hi, *bitpos, n);
}

// ---------------------------------------------------------------- FSE
// Table description at in[pos, end) (forward bit stream): normalised counts -> norm[0, *nsym), the
// accuracy log, the position behind the description.  max_sym = number of symbols allowed.
TK_HD bool tk_fse_read_norm(const uint8_t* in, uint32_t pos, uint32_t end, int max_log, int max_sym, int16_t* norm,
                            int* log_out, int* nsym_out, uint32_t* next) {
    int32_t bit = 0;
    const int al = 5 + (int)tk_zs_bits(in, pos, end, bit, 4);
    bit += 4;
    if (al > max_log) return false;
    int remaining = 1 << al, symb = 0;
    while (remaining > 0 && symb < max_sym) {
        const int nb = tk_zs_highbit((uint32_t)remaining + 1u) + 1;
        int val = (int)tk_zs_bits(in, pos, end, bit, nb);
        bit += nb;
        const int lower_mask = (1 << (nb - 1)) - 1;
        const int threshold = (1 << nb) - 1 - (remaining + 1);
        if ((val & lower_mask) < threshold) {
            bit -= 1;
            val &= lower_mask;
        } else if (val > lower_mask) {
            val -= threshold;
        }
        const int proba = val - 1;                      // -1: "less than one", takes one cell
        remaining -= proba < 0 ? -proba : proba;
        norm[symb++] = (int16_t)proba;
        if (proba == 0) {
            int rep = (int)tk_zs_bits(in, pos, end, bit, 2);
            bit += 2;
            for (;;) {
                for (int k = 0; k < rep && symb < max_sym; ++k) norm[symb++] = 0;
                if (rep != 3) break;
                rep = (int)tk_zs_bits(in, pos, end, bit, 2);
                bit += 2;
            }
        }
    }
    if (remaining != 0) return false;
    const uint32_t bytes = (uint32_t)(bit + 7) >> 3;
    if (bytes > end - pos) return false;
    *log_out = al;
    *nsym_out = symb;
    *next = pos + bytes;
    return true;
}

// Decoding table of 1 << log entries from normalised counts.
TK_HD bool tk_fse_build(tk_fse_entry* table, int log, const int16_t* norm, int nsym, uint16_t* desc) {
    const int size = 1 << log;
    int high = size;
    for (int s = 0; s < nsym; ++s)
        if (norm[s] == -1) {
            if (high == 0) return false;
            table[--high].symbol = (uint8_t)s;
            desc[s] = 1;
        }
    const int step = (size >> 1) + (size >> 3) + 3, mask = size - 1;
    int pos = 0;
    for (int s = 0; s < nsym; ++s) {
        if (norm[s] <= 0) continue;
        desc[s] = (uint16_t)norm[s];
        for (int k = 0; k < norm[s]; ++k) {
            table[pos].symbol = (uint8_t)s;
            do { pos = (pos + step) & mask; } while (pos >= high);
        }
    }
    if (pos != 0) return false;
    for (int k = 0; k < size; ++k) {
        const uint32_t next = desc[table[k].symbol]++;
        const int nbits = log - tk_zs_highbit(next);
        table[k].nbits = (uint8_t)nbits;
        table[k].base = (uint16_t)((next << nbits) - (uint32_t)size);
    }
    return true;
}

// ---------------------------------------------------------------- Huffman (literals)
// Weights coded with FSE: two interleaved states over the backward stream in[lo, hi), until the
// stream runs dry; returns the number of weights, -1 on error.
TK_HD int tk_huf_fse_weights(const uint8_t* in, uint32_t lo, uint32_t hi, tk_zstd_tables& tb) {
    int log, nsym;
    uint32_t body;
    // (weights are 0..11: a description that gives a value of TK_ZS_MAX_SYMBOLS or more a probability is
    // refused here, whether or not the stream ever produces it)
    if (!tk_fse_read_norm(in, lo, hi, 6, TK_ZS_MAX_SYMBOLS, tb.norm, &log, &nsym, &body)) return -1;
    tk_fse_entry* wt = tk_zstd_wt(tb);                   // at the head of tb.huf, which is built from the weights afterwards
    if (!tk_fse_build(wt, log, tb.norm, nsym, tb.desc)) return -1;
    int32_t bitpos;
    if (!tk_zs_bwd_init(in, body, hi, &bitpos)) return -1;
    uint32_t s1 = tk_zs_bwd_read(in, body, hi, &bitpos, log);
    uint32_t s2 = tk_zs_bwd_read(in, body, hi, &bitpos, log);
    if (bitpos < 0) return -1;
    int n = 0;
    for (;;) {                                           // at most 255 explicit weights (one more is implied)
        if (n > 252) return -1;
        tb.weights[n++] = wt[s1].symbol;
        s1 = wt[s1].base + tk_zs_bwd_read(in, body, hi, &bitpos, wt[s1].nbits);
        if (bitpos < 0) { tb.weights[n++] = wt[s2].symbol; break; }
        tb.weights[n++] = wt[s2].symbol;
        s2 = wt[s2].base + tk_zs_bwd_read(in, body, hi, &bitpos, wt[s2].nbits);
        if (bitpos < 0) { tb.weights[n++] = wt[s1].symbol; break; }
    }
    return n;
}

// Tree description at in[pos, end): weights -> decoding table.  *log = longest code, *used =
// bytes of the description.
TK_HD bool tk_huf_read_tree(const uint8_t* in, uint32_t pos, uint32_t end, tk_zstd_tables& tb, int* log, uint32_t* used) {
    if (pos >= end) return false;
    const uint32_t hb = in[pos];
    int n;
    if (hb >= 128u) {                                    // direct: 4 bits per weight
        n = (int)hb - 127;
        const uint32_t bytes = ((uint32_t)n + 1u) / 2u;
        if (bytes > end - pos - 1u) return false;
        for (int k = 0; k < n; ++k) {
            const uint32_t b = in[pos + 1u + (uint32_t)(k >> 1)];
            tb.weights[k] = (uint8_t)((k & 1) ? (b & 15u) : (b >> 4));
        }
        *used = 1u + bytes;
    } else {
        if (hb == 0 || hb > end - pos - 1u) return false;
        n = tk_huf_fse_weights(in, pos + 1u, pos + 1u + hb, tb);
        if (n < 0) return false;
        *used = 1u + hb;
    }
    uint32_t sum = 0;
    for (int k = 0; k < n; ++k) {
        if (tb.weights[k] > 11) return false;
        if (tb.weights[k]) sum += 1u << (tb.weights[k] - 1);
    }
    if (sum == 0) return false;
    const int maxbits = tk_zs_highbit(sum) + 1;
    if (maxbits > 11) return false;
    const uint32_t left = (1u << maxbits) - sum;          // the last symbol's share: a power of two
    if (left & (left - 1u)) return false;
    tb.weights[n++] = (uint8_t)(tk_zs_highbit(left) + 1);
    // codes sorted by weight (longest codes first), symbols of one weight in increasing order
    uint32_t start[13];
    uint32_t idx = 0;
    for (int w = 1; w <= maxbits; ++w) {
        start[w] = idx;
        for (int k = 0; k < n; ++k)
            if (tb.weights[k] == w) idx += 1u << (w - 1);
    }
    if (idx != (1u << maxbits)) return false;
    for (int k = 0; k < n; ++k) {
        const int w = tb.weights[k];
        if (!w) continue;
        const uint32_t len = 1u << (w - 1);
        const uint16_t e = (uint16_t)((uint32_t)k | ((uint32_t)(maxbits + 1 - w) << 8));
        for (uint32_t j = 0; j < len; ++j) tb.huf[start[w] + j] = e;
        start[w] += len;
    }
    *log = maxbits;
    return true;
}

// One Huffman stream in[lo, hi) (backward) -> n symbols at out; the stream must end exactly.
TK_HD bool tk_huf_stream(const uint8_t* in, uint32_t lo, uint32_t hi, const uint16_t* huf, int log, uint8_t* out,
                         uint32_t n) {
    int32_t bitpos;
    if (!tk_zs_bwd_init(in, lo, hi, &bitpos)) return false;
    tk_zs_win w;
    tk_zs_win_init(w);
    for (uint32_t k = 0; k < n; ++k) {
        const uint32_t e = huf[tk_zs_win_get(w, in, lo, hi, bitpos - log, log)];
        out[k] = (uint8_t)e;
        bitpos -= (int32_t)(e >> 8);
        if (bitpos < 0) return false;
    }
    return bitpos == 0;
}

// ---------------------------------------------------------------- frames, blocks, sequences
enum { TK_ZM_FRAME = 0, TK_ZM_BLOCK = 1, TK_ZM_SEQ = 2, TK_ZM_SEQHDR = 3 };
#define TK_ZS_BLOCK_MAX (128u << 10)

struct tk_zstd_state {
    const uint8_t* in;
    uint32_t pos, end;             // cursor in the batch payload
    uint32_t rep0, rep1, rep2;     // repeat offsets of the frame (scalars: the state stays in registers)
    int      ll_log, ml_log, of_log;   // accuracy logs of the current sequence tables, -1: none yet
    int      huf_log;              // longest code of the current literals tree, 0: none yet
    int      mode, last_block, checksum;
    uint32_t block_end;
    uint32_t seq_left, seq_lo;     // sequences to go; their bit stream is in[seq_lo, block_end)
    int32_t  bitpos;
    tk_zs_win win;
    uint32_t st_ll, st_ml, st_of;
    uint32_t lit_kind, lit_src, lit_left;
};

TK_HD uint32_t tk_zs_le(const uint8_t* p, int n) {
    uint32_t v = 0;
    for (int k = 0; k < n; ++k) v |= (uint32_t)p[k] << (8 * k);
    return v;
}

TK_HD void tk_zstd_init(tk_zstd_state& s, tk_zstd_tables& tb, const uint8_t* in, uint32_t pos, uint32_t end) {
    s.in = in; s.pos = pos; s.end = end; s.mode = TK_ZM_FRAME;
    s.last_block = 0; s.checksum = 0; s.seq_left = 0; s.lit_left = 0; s.huf_log = 0;
    s.ll_log = s.ml_log = s.of_log = -1;
}

// One of the three sequence tables, as the block's compression mode says.
TK_HD int tk_zstd_seq_table(tk_zstd_state& s, tk_zstd_tables& tb, uint32_t mode, tk_fse_entry* table, int* log,
                            int max_log, int max_sym, const char* def_norm, int def_log, int def_sym) {
    if (mode == 0) {                                     // predefined distribution
        for (int k = 0; k < def_sym; ++k) tb.norm[k] = (int16_t)(signed char)def_norm[k];
        if (!tk_fse_build(table, def_log, tb.norm, def_sym, tb.desc)) return TK_ZS_ERROR;
        *log = def_log;
    } else if (mode == 1) {                              // one symbol
        if (s.pos >= s.block_end || s.in[s.pos] >= max_sym) return TK_ZS_ERROR;
        table[0].symbol = s.in[s.pos++];
        table[0].nbits = 0;
        table[0].base = 0;
        *log = 0;
    } else if (mode == 2) {                              // described in the block
        int nsym;
        if (!tk_fse_read_norm(s.in, s.pos, s.block_end, max_log, max_sym, tb.norm, log, &nsym, &s.pos)) return TK_ZS_ERROR;
        if (!tk_fse_build(table, *log, tb.norm, nsym, tb.desc)) return TK_ZS_ERROR;
    } else if (*log < 0) {                               // repeat: there must be a previous one
        return TK_ZS_ERROR;
    }
    return TK_ZS_COPY;
}

// Literals section of the compressed block at s.pos .. s.block_end.  TK_ZS_HUF: *ev describes the
// Huffman streams the caller has to decode before it goes on.
TK_HD int tk_zstd_literals(tk_zstd_state& s, tk_zstd_tables& tb, uint32_t cap, uint32_t op, bool emit, tk_zstd_event* ev) {
    const uint8_t* in = s.in;
    const uint32_t bend = s.block_end;
    if (bend - s.pos < 2u) return TK_ZS_ERROR;
    const uint32_t b0 = in[s.pos], type = b0 & 3u, sf = (b0 >> 2) & 3u;
    uint32_t regen, hl;
    if (type < 2u) {
        if ((sf & 1u) == 0) { regen = b0 >> 3; hl = 1; }
        else if (sf == 1u) { if (bend - s.pos < 2u) return TK_ZS_ERROR; regen = tk_zs_le(in + s.pos, 2) >> 4; hl = 2; }
        else { if (bend - s.pos < 3u) return TK_ZS_ERROR; regen = tk_zs_le(in + s.pos, 3) >> 4; hl = 3; }
        if (regen > TK_ZS_BLOCK_MAX) return TK_ZS_ERROR;
        const uint32_t body = type == 0 ? regen : 1u;
        if (body > bend - s.pos - hl) return TK_ZS_ERROR;
        s.lit_kind = type == 0 ? TK_ZL_INPUT : TK_ZL_RLE;
        s.lit_src = type == 0 ? s.pos + hl : in[s.pos + hl];
        s.pos += hl + body;
    } else {
        uint32_t comp, streams = sf == 0 ? 1u : 4u;
        if (sf < 2u) {
            if (bend - s.pos < 3u) return TK_ZS_ERROR;
            const uint32_t h = tk_zs_le(in + s.pos, 3);
            regen = (h >> 4) & 0x3FFu; comp = (h >> 14) & 0x3FFu; hl = 3;
        } else if (sf == 2u) {
            if (bend - s.pos < 4u) return TK_ZS_ERROR;
            const uint32_t h = tk_zs_le(in + s.pos, 4);
            regen = (h >> 4) & 0x3FFFu; comp = h >> 18; hl = 4;
        } else {
            if (bend - s.pos < 5u) return TK_ZS_ERROR;
            const uint32_t h = tk_zs_le(in + s.pos, 4), h4 = in[s.pos + 4];
            regen = (h >> 4) & 0x3FFFFu; comp = (h >> 22) | (h4 << 10); hl = 5;
        }
        if (regen > TK_ZS_BLOCK_MAX || comp > bend - s.pos - hl) return TK_ZS_ERROR;
        uint32_t lo = s.pos + hl;
        const uint32_t hi = lo + comp;
        s.pos = hi;
        s.lit_kind = TK_ZL_OUTPUT;
        s.lit_src = 0;
        if (emit) {
            if (type == 2u) {
                uint32_t used;
                if (!tk_huf_read_tree(in, lo, hi, tb, &s.huf_log, &used)) return TK_ZS_ERROR;
                lo += used;
            } else if (s.huf_log == 0) {
                return TK_ZS_ERROR;                      // tree reuse without a tree
            }
            // the decoded literals are parked at the end of the output buffer: the block's output
            // ends at or before it, so the write position never passes the unread literals
            if (regen > cap - op) return TK_ZS_ERROR;
            const uint32_t dst = cap - regen;
            s.lit_src = dst;
            ev->h_streams = streams;
            ev->h_log = (uint32_t)s.huf_log;
            if (streams == 1u) {
                ev->h_lo[0] = lo; ev->h_hi[0] = hi; ev->h_dst[0] = dst; ev->h_n[0] = regen;
            } else {
                if (hi - lo < 6u) return TK_ZS_ERROR;
                const uint32_t z1 = tk_zs_le(in + lo, 2), z2 = tk_zs_le(in + lo + 2, 2), z3 = tk_zs_le(in + lo + 4, 2);
                lo += 6;
                if ((uint64_t)z1 + z2 + z3 > hi - lo) return TK_ZS_ERROR;
                const uint32_t per = (regen + 3u) / 4u;
                if (3u * per > regen) return TK_ZS_ERROR;
                ev->h_lo[0] = lo;           ev->h_hi[0] = lo + z1;
                ev->h_lo[1] = lo + z1;      ev->h_hi[1] = lo + z1 + z2;
                ev->h_lo[2] = lo + z1 + z2; ev->h_hi[2] = lo + z1 + z2 + z3;
                ev->h_lo[3] = lo + z1 + z2 + z3; ev->h_hi[3] = hi;
                for (uint32_t k = 0; k < 4u; ++k) { ev->h_dst[k] = dst + k * per; ev->h_n[k] = k < 3u ? per : regen - 3u * per; }
            }
            s.lit_left = regen;
            return regen ? TK_ZS_HUF : TK_ZS_COPY;
        }
    }
    s.lit_left = regen;
    return TK_ZS_COPY;
}

// Sequences header of the block (behind its literals section): count, tables, initial states.
TK_HD int tk_zstd_seq_header(tk_zstd_state& s, tk_zstd_tables& tb) {
    const uint8_t* in = s.in;
    const uint32_t bend = s.block_end;
    if (s.pos >= bend) return TK_ZS_ERROR;
    uint32_t nseq = in[s.pos++];
    if (nseq >= 128u) {
        if (nseq < 255u) {
            if (s.pos >= bend) return TK_ZS_ERROR;
            nseq = ((nseq - 128u) << 8) + in[s.pos++];
        } else {
            if (bend - s.pos < 2u) return TK_ZS_ERROR;
            nseq = tk_zs_le(in + s.pos, 2) + 0x7F00u;
            s.pos += 2;
        }
    }
    s.seq_left = nseq;
    if (nseq == 0) return s.pos == bend ? TK_ZS_COPY : TK_ZS_ERROR;
    if (s.pos >= bend) return TK_ZS_ERROR;
    const uint32_t modes = in[s.pos++];
    if (modes & 3u) return TK_ZS_ERROR;
    const char* ll_def = "\4\3\2\2\2\2\2\2\2\2\2\2\2\1\1\1\2\2\2\2\2\2\2\2\2\3\2\1\1\1\1\1\377\377\377\377";
    const char* of_def = "\1\1\1\1\1\1\2\2\2\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1\377\377\377\377\377";
    const char* ml_def = "\1\4\3\2\2\2\2\2\2\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1\1"
                         "\377\377\377\377\377\377\377";
    int st = tk_zstd_seq_table(s, tb, modes >> 6, tb.ll, &s.ll_log, 9, 36, ll_def, 6, 36);
    if (st) return st;
    st = tk_zstd_seq_table(s, tb, (modes >> 4) & 3u, tb.of, &s.of_log, 8, 32, of_def, 5, 29);
    if (st) return st;
    st = tk_zstd_seq_table(s, tb, (modes >> 2) & 3u, tb.ml, &s.ml_log, 9, 53, ml_def, 6, 53);
    if (st) return st;
    s.seq_lo = s.pos;
    if (!tk_zs_bwd_init(in, s.seq_lo, bend, &s.bitpos)) return TK_ZS_ERROR;
    tk_zs_win_init(s.win);
    s.st_ll = tk_zs_win_read(s.win, in, s.seq_lo, bend, &s.bitpos, s.ll_log);
    s.st_of = tk_zs_win_read(s.win, in, s.seq_lo, bend, &s.bitpos, s.of_log);
    s.st_ml = tk_zs_win_read(s.win, in, s.seq_lo, bend, &s.bitpos, s.ml_log);
    return s.bitpos < 0 ? TK_ZS_ERROR : TK_ZS_COPY;
}

// Advances to the next piece of output.  TK_ZS_COPY: *ev describes it (the caller writes
// lit_len + mlen bytes at `op` and adds them to op before the next call); TK_ZS_DONE: the payload
// is exhausted; otherwise an error.  `op` = bytes produced so far, `cap` = size of the output
// buffer (emit) or a bound on the total (size pass).
TK_HD int tk_zstd_next(tk_zstd_state& s, tk_zstd_tables& tb, uint8_t* out, uint32_t cap, uint32_t op, bool emit,
                       tk_zstd_event* ev) {
    const uint8_t* in = s.in;
    for (;;) {
        if (s.mode == TK_ZM_FRAME) {
            if (s.pos == s.end) return TK_ZS_DONE;
            if (s.end - s.pos < 4u) return TK_ZS_ERROR;
            const uint32_t magic = tk_zs_le(in + s.pos, 4);
            if ((magic & 0xFFFFFFF0u) == 0x184D2A50u) {              // skippable frame
                if (s.end - s.pos < 8u) return TK_ZS_ERROR;
                const uint32_t size = tk_zs_le(in + s.pos + 4, 4);
                if (size > s.end - s.pos - 8u) return TK_ZS_ERROR;
                s.pos += 8u + size;
                continue;
            }
            if (magic != 0xFD2FB528u || s.end - s.pos < 5u) return TK_ZS_ERROR;
            const uint32_t fhd = in[s.pos + 4];
            s.pos += 5;
            if (fhd & 0x08u) return TK_ZS_ERROR;                     // reserved bit
            const uint32_t single = (fhd >> 5) & 1u, fcs = fhd >> 6, did = fhd & 3u;
            s.checksum = (int)((fhd >> 2) & 1u);
            const uint32_t did_bytes = did == 3u ? 4u : did;
            const uint32_t fcs_bytes = fcs == 0 ? single : fcs == 1u ? 2u : fcs == 2u ? 4u : 8u;
            const uint32_t head = (single ? 0u : 1u) + did_bytes + fcs_bytes;
            if (head > s.end - s.pos) return TK_ZS_ERROR;
            if (did_bytes && tk_zs_le(in + s.pos + (single ? 0u : 1u), (int)did_bytes) != 0) return TK_ZS_UNSUPPORTED;
            s.pos += head;
            s.rep0 = 1; s.rep1 = 4; s.rep2 = 8;
            s.ll_log = s.ml_log = s.of_log = -1;
            s.huf_log = 0;
            s.last_block = 0;
            s.mode = TK_ZM_BLOCK;
        }
        if (s.mode == TK_ZM_BLOCK) {
            if (s.last_block) {
                if (s.checksum) {
                    if (s.end - s.pos < 4u) return TK_ZS_ERROR;
                    s.pos += 4;
                }
                s.mode = TK_ZM_FRAME;
                continue;
            }
            if (s.end - s.pos < 3u) return TK_ZS_ERROR;
            const uint32_t bh = tk_zs_le(in + s.pos, 3);
            s.pos += 3;
            s.last_block = (int)(bh & 1u);
            const uint32_t type = (bh >> 1) & 3u, size = bh >> 3;
            if (type == 3u || size > TK_ZS_BLOCK_MAX) return TK_ZS_ERROR;
            ev->mlen = 0; ev->offset = 0;
            if (type < 2u) {                                         // raw / RLE block
                const uint32_t body = type == 0 ? size : 1u;
                if (body > s.end - s.pos || size > cap - op) return TK_ZS_ERROR;
                ev->lit_len = size;
                ev->lit_kind = type == 0 ? TK_ZL_INPUT : TK_ZL_RLE;
                ev->lit_src = type == 0 ? s.pos : in[s.pos];
                s.pos += body;
                if (size == 0) continue;
                return TK_ZS_COPY;
            }
            if (size > s.end - s.pos) return TK_ZS_ERROR;
            s.block_end = s.pos + size;
            const int st = tk_zstd_literals(s, tb, cap, op, emit, ev);
            s.mode = TK_ZM_SEQHDR;
            if (st) return st;                                       // TK_ZS_HUF: back here once the caller decoded them
        }
        if (s.mode == TK_ZM_SEQHDR) {
            const int st = tk_zstd_seq_header(s, tb);
            if (st) return st;
            s.mode = TK_ZM_SEQ;
        }
        // ---- TK_ZM_SEQ: the next sequence, or the literals behind the last one
        if (s.seq_left == 0) {
            s.mode = TK_ZM_BLOCK;
            s.pos = s.block_end;
            if (s.lit_left == 0) continue;
            if (s.lit_left > cap - op) return TK_ZS_ERROR;
            ev->lit_len = s.lit_left; ev->lit_kind = s.lit_kind; ev->lit_src = s.lit_src;
            ev->mlen = 0; ev->offset = 0;
            s.lit_left = 0;
            return TK_ZS_COPY;
        }
        const uint32_t lo = s.seq_lo, hi = s.block_end;
        // the three table entries (symbol, state bits, state base: one 32-bit word each)
        const uint32_t e_of = reinterpret_cast<const uint32_t*>(tb.of)[s.st_of];
        const uint32_t e_ml = reinterpret_cast<const uint32_t*>(tb.ml)[s.st_ml];
        const uint32_t e_ll = reinterpret_cast<const uint32_t*>(tb.ll)[s.st_ll];
        const uint32_t ofc = e_of & 0xFFu, mlc = e_ml & 0xFFu, llc = e_ll & 0xFFu;
        if (ofc > 31u || mlc > 52u || llc > 35u) return TK_ZS_ERROR;
        const uint32_t mlp = tk_zs_ml_code(mlc), llp = tk_zs_ll_code(llc), mlb = mlp & 31u, llb = llp & 31u;
        const bool more = s.seq_left > 1u;                           // the states move on unless this is the last one
        const uint32_t nb_ll = (e_ll >> 8) & 0xFFu, nb_ml = (e_ml >> 8) & 0xFFu, nb_of = (e_of >> 8) & 0xFFu;
        const uint32_t total = ofc + mlb + llb + (more ? nb_ll + nb_ml + nb_of : 0u);
        uint32_t ov, mlen, llen;
        if (total <= 56u) {
            // all fields of the sequence in one read: offset bits on top, then match length, literal
            // length, and the bits of the three state updates at the bottom
            s.bitpos -= (int32_t)total;
            uint64_t v = total ? tk_zs_win_get64(s.win, in, lo, hi, s.bitpos, (int)total) : 0ull;
            if (more) {
                s.st_of = (e_of >> 16) + ((uint32_t)v & ((1u << nb_of) - 1u)); v >>= nb_of;
                s.st_ml = (e_ml >> 16) + ((uint32_t)v & ((1u << nb_ml) - 1u)); v >>= nb_ml;
                s.st_ll = (e_ll >> 16) + ((uint32_t)v & ((1u << nb_ll) - 1u)); v >>= nb_ll;
            }
            llen = (llp >> 5) + ((uint32_t)v & ((1u << llb) - 1u)); v >>= llb;
            mlen = (mlp >> 5) + ((uint32_t)v & ((1u << mlb) - 1u)); v >>= mlb;
            ov = (1u << ofc) + ((uint32_t)v & ((1u << ofc) - 1u));
            if (s.bitpos < 0) return TK_ZS_ERROR;
        } else {
            ov = (1u << ofc) + tk_zs_win_read(s.win, in, lo, hi, &s.bitpos, (int)ofc);
            mlen = (mlp >> 5) + tk_zs_win_read(s.win, in, lo, hi, &s.bitpos, (int)mlb);
            llen = (llp >> 5) + tk_zs_win_read(s.win, in, lo, hi, &s.bitpos, (int)llb);
            if (s.bitpos < 0) return TK_ZS_ERROR;
            if (more) {
                s.st_ll = (e_ll >> 16) + tk_zs_win_read(s.win, in, lo, hi, &s.bitpos, (int)nb_ll);
                s.st_ml = (e_ml >> 16) + tk_zs_win_read(s.win, in, lo, hi, &s.bitpos, (int)nb_ml);
                s.st_of = (e_of >> 16) + tk_zs_win_read(s.win, in, lo, hi, &s.bitpos, (int)nb_of);
                if (s.bitpos < 0) return TK_ZS_ERROR;
            }
        }
        uint32_t offset;
        if (ov > 3u) {
            offset = ov - 3u;
            s.rep2 = s.rep1; s.rep1 = s.rep0; s.rep0 = offset;
        } else {
            const uint32_t idx = ov - 1u + (llen == 0 ? 1u : 0u);
            if (idx == 0) {
                offset = s.rep0;
            } else {
                offset = idx == 1u ? s.rep1 : idx == 2u ? s.rep2 : s.rep0 - 1u;
                if (idx > 1u) s.rep2 = s.rep1;
                s.rep1 = s.rep0;
                s.rep0 = offset;
            }
        }
        --s.seq_left;
        if (!more && s.bitpos != 0) return TK_ZS_ERROR;              // the bit stream must end exactly
        if (llen > s.lit_left || llen > cap - op || mlen > cap - op - llen) return TK_ZS_ERROR;
        if (offset == 0 || offset > op + llen) return TK_ZS_ERROR;   // no dictionary: only this payload's output
        ev->lit_len = llen; ev->lit_kind = s.lit_kind; ev->lit_src = s.lit_src;
        ev->mlen = mlen; ev->offset = offset;
        s.lit_left -= llen;
        if (s.lit_kind != TK_ZL_RLE) s.lit_src += llen;
        // parked literals not yet read must stay ahead of the write position
        if (emit && s.lit_kind == TK_ZL_OUTPUT && s.lit_left && op + llen + mlen > s.lit_src) return TK_ZS_ERROR;
        return TK_ZS_COPY;
    }
}

// ---------------------------------------------------------------------------------------------
// Sequences in batches.  The sequence chain of a block is serial (three FSE states), but nothing
// else about a sequence is: tk_zstd_next_batch decodes up to TK_ZS_BATCH sequences in one tight
// loop (state in registers, no dispatch, no event per call) and leaves their copies to the caller,
// who can place them with a prefix sum and execute them side by side:
//   * the literal runs of a batch never depend on one another (their sources are the input, a
//     byte, or the block's parked literals, which stay ahead of every write position);
//   * a match depends on earlier matches of the batch only when its source reaches into what
//     they write, i.e. ends behind the first match destination of the batch (or overlaps its own
//     destination); on text that is one sequence in twenty.  The others can be copied in any
//     order once the literals are in place; the dependent ones follow in sequence order.
// Returns the number of sequences decoded (>= 1), 0 when the decoder is not inside a sequence
// section (the caller goes on with tk_zstd_next), or -status on a broken stream.  `op` is the
// output position in front of the first sequence of the batch.
#define TK_ZS_BATCH 32
struct alignas(16) tk_zstd_seq { uint32_t lit_len, lit_src, mlen, offset; };
struct tk_zstd_batch {
    tk_zstd_seq q[TK_ZS_BATCH];
    uint32_t lit_kind;             // of every literal run of the batch (the block's)
};

// The batch's place inside the tables' scratch (dead while a block's sequences are decoded).
TK_HD tk_zstd_batch* tk_zstd_batch_of(tk_zstd_tables& tb) {
    static_assert(sizeof(tk_zstd_batch) <= sizeof(tb.norm) + sizeof(tb.desc) + sizeof(tb.weights) + sizeof(tb.scratch_tail), "the batch must fit the scratch");
    static_assert(offsetof(tk_zstd_tables, norm) % 16 == 0, "the batch is read with 16-byte loads");
    return reinterpret_cast<tk_zstd_batch*>(tb.norm);
}

TK_HD int tk_zstd_next_batch(tk_zstd_state& s, const tk_zstd_tables& tb, uint32_t cap, uint32_t op, bool emit,
                             tk_zstd_batch* b) {
    if (s.mode != TK_ZM_SEQ || s.seq_left == 0) return 0;
    const uint8_t* in = s.in;
    const uint32_t lo = s.seq_lo, hi = s.block_end;
    // the state of the chain, in registers for the duration of the batch
    uint32_t st_ll = s.st_ll, st_ml = s.st_ml, st_of = s.st_of;
    uint32_t rep0 = s.rep0, rep1 = s.rep1, rep2 = s.rep2;
    uint32_t seq_left = s.seq_left, lit_left = s.lit_left, lit_src = s.lit_src;
    const uint32_t lit_kind = s.lit_kind;
    int32_t bitpos = s.bitpos;
    tk_zs_win win = s.win;
    b->lit_kind = lit_kind;
    // Errors are collected, not branched on: nothing inside the loop reads memory through a value
    // a broken stream controls (the states stay inside their tables by construction, the window
    // reads zeros outside the stream, the batch index is the loop counter), so a broken sequence
    // only produces garbage that is thrown away when the loop is left.
    uint32_t bad = 0;
    int n = 0;
    for (; n < TK_ZS_BATCH && seq_left; ++n) {
        const uint32_t e_of = reinterpret_cast<const uint32_t*>(tb.of)[st_of];
        const uint32_t e_ml = reinterpret_cast<const uint32_t*>(tb.ml)[st_ml];
        const uint32_t e_ll = reinterpret_cast<const uint32_t*>(tb.ll)[st_ll];
        // (the symbols are below 32 / 53 / 36: the table builds admit no others)
        const uint32_t ofc = e_of & 0xFFu, mlc = e_ml & 0xFFu, llc = e_ll & 0xFFu;
        const uint32_t mlp = tk_zs_ml_code(mlc), llp = tk_zs_ll_code(llc), mlb = mlp & 31u, llb = llp & 31u;
        const bool more = seq_left > 1u;
        const uint32_t nb_ll = (e_ll >> 8) & 0xFFu, nb_ml = (e_ml >> 8) & 0xFFu, nb_of = (e_of >> 8) & 0xFFu;
        const uint32_t total = ofc + mlb + llb + (more ? nb_ll + nb_ml + nb_of : 0u);
        uint32_t ov, mlen, llen;
        if (total <= 56u) {
            bitpos -= (int32_t)total;
            uint64_t v = total ? tk_zs_win_get64(win, in, lo, hi, bitpos, (int)total) : 0ull;
            if (more) {
                st_of = (e_of >> 16) + ((uint32_t)v & ((1u << nb_of) - 1u)); v >>= nb_of;
                st_ml = (e_ml >> 16) + ((uint32_t)v & ((1u << nb_ml) - 1u)); v >>= nb_ml;
                st_ll = (e_ll >> 16) + ((uint32_t)v & ((1u << nb_ll) - 1u)); v >>= nb_ll;
            }
            llen = (llp >> 5) + ((uint32_t)v & ((1u << llb) - 1u)); v >>= llb;
            mlen = (mlp >> 5) + ((uint32_t)v & ((1u << mlb) - 1u)); v >>= mlb;
            ov = (1u << ofc) + ((uint32_t)v & ((1u << ofc) - 1u));
        } else {
            ov = (1u << ofc) + tk_zs_win_read(win, in, lo, hi, &bitpos, (int)ofc);
            mlen = (mlp >> 5) + tk_zs_win_read(win, in, lo, hi, &bitpos, (int)mlb);
            llen = (llp >> 5) + tk_zs_win_read(win, in, lo, hi, &bitpos, (int)llb);
            bad |= (uint32_t)bitpos >> 31;
            if (more) {
                st_ll = (e_ll >> 16) + tk_zs_win_read(win, in, lo, hi, &bitpos, (int)nb_ll);
                st_ml = (e_ml >> 16) + tk_zs_win_read(win, in, lo, hi, &bitpos, (int)nb_ml);
                st_of = (e_of >> 16) + tk_zs_win_read(win, in, lo, hi, &bitpos, (int)nb_of);
            }
        }
        bad |= (uint32_t)bitpos >> 31;                                 // ran off the front of the stream
        uint32_t offset;
        if (ov > 3u) {
            offset = ov - 3u;
            rep2 = rep1; rep1 = rep0; rep0 = offset;
        } else {
            const uint32_t idx = ov - 1u + (llen == 0 ? 1u : 0u);
            if (idx == 0) {
                offset = rep0;
            } else {
                offset = idx == 1u ? rep1 : idx == 2u ? rep2 : rep0 - 1u;
                if (idx > 1u) rep2 = rep1;
                rep1 = rep0;
                rep0 = offset;
            }
        }
        --seq_left;
        const uint32_t room = cap - op;
        bad |= (uint32_t)(!more & (bitpos != 0))                        // the bit stream must end exactly
             | (uint32_t)(llen > lit_left) | (uint32_t)(llen > room) | (uint32_t)(mlen > room - llen)
             | (uint32_t)(offset == 0) | (uint32_t)(offset > op + llen);   // no dictionary: only this payload's output
        tk_zstd_seq q;
        q.lit_len = llen; q.lit_src = lit_src; q.mlen = mlen; q.offset = offset;
        b->q[n] = q;
        lit_left -= llen;
        if (lit_kind != TK_ZL_RLE) lit_src += llen;
        op += llen + mlen;
        // parked literals not yet read must stay ahead of the write position
        bad |= (uint32_t)(emit & (lit_kind == TK_ZL_OUTPUT) & (lit_left != 0) & (op > lit_src));
    }
    if (bad) return -TK_ZS_ERROR;
    s.st_ll = st_ll; s.st_ml = st_ml; s.st_of = st_of;
    s.rep0 = rep0; s.rep1 = rep1; s.rep2 = rep2;
    s.seq_left = seq_left; s.lit_left = lit_left; s.lit_src = lit_src;
    s.bitpos = bitpos;
    s.win = win;
    return n;
}
