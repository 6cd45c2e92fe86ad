// tk_inflate.cuh -- DEFLATE (RFC 1951) decoder for gzip-compressed record batches
// (attributes & 7 == 1).  Replaces kafka-python's `gzip_decode`
// (kafka/codec.py: gzip.GzipFile(...).read(), i.e. zlib's inflate), called from
// DefaultRecordBatch._maybe_uncompress; dependency of the reference (setup.py:9), not in its tree.
//
// A DEFLATE stream is one chain of dependent bit fields, so one thread walks it: it decodes
// symbols, writes literals itself, and hands every match (length, distance) to its caller, which
// copies it -- on the device with the whole warp (tk_codecs.cu), on the host in a plain loop
// (tests/native, differential test against zlib).  Host+device code (TK_HD).
// Huffman decoding is canonical: per code length the number of codes and the symbols in code
// order.  Codes of up to TK_INF_LIT_FAST / TK_INF_DIST_FAST bits are found with one lookup of the
// next bits in a table; longer ones are walked one bit at a time until they fall into the range of
// their length.  The tables of the literal / length and the distance code hold what the symbol
// means (tk_inf_lit_entry, tk_inf_dist_entry: the literal, or the baseline and the number of extra
// bits), so that a match costs two table lookups and no further ones.
#pragma once
#include <string.h>
#include "tk_number.cuh"      // TK_HD

enum { TK_INF_MATCH = 0, TK_INF_DONE = 1, TK_INF_ERROR = 2, TK_INF_BATCH = 3 };
enum { TK_INF_HEADER = 0, TK_INF_CODES = 1, TK_INF_END = 2 };

#ifndef TK_INF_LIT_FAST
#define TK_INF_LIT_FAST 10
#endif
#define TK_INF_DIST_FAST 9
// Matches whose copy is left to the caller (tk_inflate_next with a batch): destination, length and
// distance of each; the literals around them are already in place.  The copies may run side by side
// except where a match reads what an earlier match of the batch writes (source end behind the first
// destination of the batch): those follow in order.
#define TK_INF_BATCH_MAX 32
struct alignas(16) tk_inflate_match { uint32_t dst, len, dist, reserved; };
struct tk_inflate_batch {
    tk_inflate_match m[TK_INF_BATCH_MAX];      // the first tk_inflate_state::nb are valid
};

// length code - 257 / distance code -> baseline << 4 | extra bits (RFC 1951, 3.2.5).  Constants:
// read by the one lane that walks the stream (constant bank on the device).
#define TK_INF_LEN_CODES { 0x30u, 0x40u, 0x50u, 0x60u, 0x70u, 0x80u, 0x90u, 0xA0u, 0xB1u, 0xD1u, 0xF1u, 0x111u, 0x132u, 0x172u, 0x1B2u, 0x1F2u, 0x233u, 0x2B3u, 0x333u, 0x3B3u, 0x434u, 0x534u, 0x634u, 0x734u, 0x835u, 0xA35u, 0xC35u, 0xE35u, 0x1020u }
#define TK_INF_DIST_CODES { 0x10u, 0x20u, 0x30u, 0x40u, 0x51u, 0x71u, 0x92u, 0xD2u, 0x113u, 0x193u, 0x214u, 0x314u, 0x415u, 0x615u, 0x816u, 0xC16u, 0x1017u, 0x1817u, 0x2018u, 0x3018u, 0x4019u, 0x6019u, 0x801Au, 0xC01Au, 0x1001Bu, 0x1801Bu, 0x2001Cu, 0x3001Cu, 0x4001Du, 0x6001Du }
#if defined(__CUDACC__)
static __constant__ uint32_t tk_inf_len_dev[29] = TK_INF_LEN_CODES;
static __constant__ uint32_t tk_inf_dist_dev[30] = TK_INF_DIST_CODES;
#endif
#if !defined(__CUDA_ARCH__)
static const uint32_t tk_inf_len_host[29] = TK_INF_LEN_CODES;
static const uint32_t tk_inf_dist_host[30] = TK_INF_DIST_CODES;
#endif
TK_HD uint32_t tk_inf_len_code(int c) {
#if defined(__CUDA_ARCH__)
    return tk_inf_len_dev[c];
#else
    return tk_inf_len_host[c];
#endif
}
TK_HD uint32_t tk_inf_dist_code(int c) {
#if defined(__CUDA_ARCH__)
    return tk_inf_dist_dev[c];
#else
    return tk_inf_dist_host[c];
#endif
}

// Where the bit-by-bit walk through a canonical code stands after `len` bits that matched no code:
// the walk for a code the table of `len` index bits does not hold starts there.
struct tk_huff_walk { uint32_t len, first, index; };

struct tk_inflate_tables {
    uint16_t lit_count[16], lit_symbol[288];
    uint16_t dist_count[16], dist_symbol[32];
    uint16_t lengths[320];         // scratch: code lengths of a dynamic block
    uint32_t lit_fast[1 << TK_INF_LIT_FAST], dist_fast[1 << TK_INF_DIST_FAST];   // entries (below); 0: longer code
    tk_huff_walk lit_walk, dist_walk;               // where the walk for a longer code starts
    uint16_t code_fast[128];       // the code of the code lengths (dynamic block header): symbol << 4 | length
    tk_inflate_batch batch;
};

struct tk_inflate_state {
    const uint8_t* in;
    uint32_t pos, end;             // next input byte; pos may run past end (zeros are fed, checked later)
    uint64_t buf;                  // bit buffer, least significant bit first
    int      cnt;                  // bits in buf
    int      mode, last;
    uint32_t nb;                   // matches waiting in the caller's batch
};

// The 4 input bytes at pos (4-byte aligned); zeros behind the end of the input.
TK_HD uint32_t tk_inf_word(const tk_inflate_state& s, uint32_t pos) {
    uint32_t w = 0;
    if (pos < s.end && s.end - pos >= 4u) {
#if defined(__CUDA_ARCH__)
        w = *reinterpret_cast<const uint32_t*>(s.in + pos);        // aligned: tk_inflate_init
#else
        memcpy(&w, s.in + pos, 4);
#endif
    } else {
        for (uint32_t k = 0; k < 4; ++k)
            if (pos + k < s.end) w |= (uint32_t)s.in[pos + k] << (8 * k);
    }
    return w;
}
// Brings the buffer to at least 33 bits with one aligned 4-byte word (every field group of the
// format needs at most 32: see the callers).  Past the end of the input zeros are fed; the position
// check (tk_inf_consumed) catches a stream that used them.  (Loading the word one refill ahead,
// with or without a prefetch of the line behind it, made the device decoder slower.)
TK_HD void tk_inf_refill(tk_inflate_state& s) {
    if (s.cnt > 32) return;
    const uint32_t w = tk_inf_word(s, s.pos);
    s.pos += 4;
    s.buf |= (uint64_t)w << s.cnt;
    s.cnt += 32;
}
TK_HD void tk_inflate_init(tk_inflate_state& s, const uint8_t* in, uint32_t pos, uint32_t end) {
    s.in = in; s.pos = pos; s.end = end; s.buf = 0; s.cnt = 0; s.mode = TK_INF_HEADER; s.last = 0; s.nb = 0;
    while (((uintptr_t)(in + s.pos) & 3u) && s.pos < end) {         // up to the first aligned word
        s.buf |= (uint64_t)in[s.pos++] << s.cnt;
        s.cnt += 8;
    }
}
TK_HD uint32_t tk_inf_bits(tk_inflate_state& s, int n) {          // n <= 16, buffer holds them
    const uint32_t v = (uint32_t)s.buf & ((1u << n) - 1u);
    s.buf >>= n;
    s.cnt -= n;
    return v;
}
// bytes of input consumed so far (whole bytes still in the buffer do not count)
TK_HD uint32_t tk_inf_consumed(const tk_inflate_state& s) { return s.pos - (uint32_t)(s.cnt >> 3); }

// Canonical code from code lengths.  Returns 0 for a complete code, > 0 for an incomplete one
// (that many codes of the longest length are unused), < 0 for an over-subscribed one.
TK_HD int tk_huff_build(uint16_t* count, uint16_t* symbol, const uint16_t* length, int n) {
    for (int l = 0; l < 16; ++l) count[l] = 0;
    for (int i = 0; i < n; ++i) count[length[i]]++;
    if (count[0] == n) return 0;                       // no codes at all: nothing can be decoded
    int left = 1;
    for (int l = 1; l < 16; ++l) {
        left <<= 1;
        left -= count[l];
        if (left < 0) return left;
    }
    uint16_t offs[16];
    offs[1] = 0;
    for (int l = 1; l < 15; ++l) offs[l + 1] = (uint16_t)(offs[l] + count[l]);
    for (int i = 0; i < n; ++i)
        if (length[i]) symbol[offs[length[i]]++] = (uint16_t)i;
    return left;
}

// Lookup table of the codes of at most `fast_bits` bits (call right after tk_huff_build, same
// lengths): entry at every index whose low bits are the code, first bit of the code lowest.
TK_HD void tk_huff_fast(uint16_t* fast, int fast_bits, const uint16_t* count, const uint16_t* length, int n) {
    uint32_t next[16];
    next[1] = 0;
    for (int l = 1; l < 15; ++l) next[l + 1] = (next[l] + count[l]) << 1;
    for (int k = 0; k < (1 << fast_bits); ++k) fast[k] = 0;
    for (int i = 0; i < n; ++i) {
        const int l = length[i];
        if (!l) continue;
        const uint32_t code = next[l]++;
        if (l > fast_bits) continue;
        uint32_t rev = 0;
        for (int b = 0; b < l; ++b) rev |= ((code >> b) & 1u) << (l - 1 - b);
        for (uint32_t k = rev; k < (1u << fast_bits); k += 1u << l) fast[k] = (uint16_t)((i << 4) | l);
    }
}

// Table entries of the literal / length code and of the distance code:
//   bits 0-3    length of the code (0 in an entry made for a symbol the bit-by-bit walk found)
//   bits 4-7    number of extra bits that follow the code
//   bits 8-9    TK_INF_E_LITERAL / _MATCH (length or distance) / _END (end of block) / _BAD (a
//               symbol the format does not define, or no symbol at all)
//   bits 16-31  the literal, or the baseline of the length / distance
enum { TK_INF_E_LITERAL = 0u << 8, TK_INF_E_MATCH = 1u << 8, TK_INF_E_END = 2u << 8, TK_INF_E_BAD = 3u << 8, TK_INF_E_KIND = 3u << 8 };
TK_HD uint32_t tk_inf_lit_entry(int sym, uint32_t l) {
    if (sym < 0 || sym > 285) return TK_INF_E_BAD | l;
    if (sym < 256) return TK_INF_E_LITERAL | l | ((uint32_t)sym << 16);
    if (sym == 256) return TK_INF_E_END | l;
    const uint32_t lc = tk_inf_len_code(sym - 257);
    return TK_INF_E_MATCH | l | ((lc & 15u) << 4) | ((lc >> 4) << 16);
}
TK_HD uint32_t tk_inf_dist_entry(int sym, uint32_t l) {
    if (sym < 0 || sym > 29) return TK_INF_E_BAD | l;
    const uint32_t dc = tk_inf_dist_code(sym);
    return TK_INF_E_MATCH | l | ((dc & 15u) << 4) | ((dc >> 4) << 16);
}
// Like tk_huff_fast, with such entries (DIST: of the distance code).
template <bool DIST>
TK_HD void tk_huff_fast_entries(uint32_t* fast, int fast_bits, const uint16_t* count, const uint16_t* length, int n) {
    uint32_t next[16];
    next[1] = 0;
    for (int l = 1; l < 15; ++l) next[l + 1] = (next[l] + count[l]) << 1;
    for (int k = 0; k < (1 << fast_bits); ++k) fast[k] = 0;
    for (int i = 0; i < n; ++i) {
        const int l = length[i];
        if (!l) continue;
        const uint32_t code = next[l]++;
        if (l > fast_bits) continue;
        uint32_t rev = 0;
        for (int b = 0; b < l; ++b) rev |= ((code >> b) & 1u) << (l - 1 - b);
        const uint32_t e = DIST ? tk_inf_dist_entry(i, (uint32_t)l) : tk_inf_lit_entry(i, (uint32_t)l);
        for (uint32_t k = rev; k < (1u << fast_bits); k += 1u << l) fast[k] = e;
    }
}

TK_HD tk_huff_walk tk_huff_walk_from(const uint16_t* count, int len) {
    tk_huff_walk w;
    w.len = (uint32_t)len; w.first = 0; w.index = 0;
    for (int l = 1; l <= len; ++l) {
        w.index += count[l];
        w.first = (w.first + count[l]) << 1;
    }
    return w;
}
TK_HD uint32_t tk_inf_reverse(uint32_t bits, uint32_t n) {         // the low n bits, first one highest (n <= 15)
#if defined(__CUDA_ARCH__)
    return n ? __brev(bits) >> (32u - n) : 0u;
#else
    uint32_t r = 0;
    for (uint32_t b = 0; b < n; ++b) r |= ((bits >> b) & 1u) << (n - 1u - b);
    return r;
#endif
}

// Next symbol of a code the table does not hold: bit by bit through the canonical code, from
// where the table's index bits end (`from`; {0, 0, 0}: from the first bit); -1 if the bits are no
// code.  The buffer holds at least 15 bits.
TK_HD int tk_huff_decode_slow(tk_inflate_state& s, const uint16_t* count, const uint16_t* symbol, const tk_huff_walk& from) {
    uint32_t bits = (uint32_t)s.buf;
    int code = (int)(tk_inf_reverse(bits, from.len) << 1), first = (int)from.first, index = (int)from.index;
    bits >>= from.len;
    for (int len = (int)from.len + 1; len < 16; ++len) {
        code |= (int)(bits & 1u);
        bits >>= 1;
        const int c = count[len];
        if (code - c < first) {
            s.buf >>= len;
            s.cnt -= len;
            return symbol[index + (code - first)];
        }
        index += c;
        first = (first + c) << 1;
        code <<= 1;
    }
    return -1;
}

// Next symbol (the buffer holds at least 15 bits), -1 if the bits are no code.
TK_HD int tk_huff_decode(tk_inflate_state& s, const uint16_t* count, const uint16_t* symbol, const uint16_t* fast,
                         int fast_bits) {
    const uint32_t e = fast[(uint32_t)s.buf & ((1u << fast_bits) - 1u)];
    if (e) {
        s.buf >>= (e & 15u);
        s.cnt -= (int)(e & 15u);
        return (int)(e >> 4);
    }
    const tk_huff_walk from = {0u, 0u, 0u};
    return tk_huff_decode_slow(s, count, symbol, from);
}

TK_HD bool tk_inf_fixed_tables(tk_inflate_tables& tb) {
    for (int i = 0; i < 288; ++i) tb.lengths[i] = (uint16_t)(i < 144 ? 8 : i < 256 ? 9 : i < 280 ? 7 : 8);
    tk_huff_build(tb.lit_count, tb.lit_symbol, tb.lengths, 288);
    tk_huff_fast_entries<false>(tb.lit_fast, TK_INF_LIT_FAST, tb.lit_count, tb.lengths, 288);
    tb.lit_walk = tk_huff_walk_from(tb.lit_count, TK_INF_LIT_FAST);
    for (int i = 0; i < 30; ++i) tb.lengths[i] = 5;
    tk_huff_build(tb.dist_count, tb.dist_symbol, tb.lengths, 30);
    tk_huff_fast_entries<true>(tb.dist_fast, TK_INF_DIST_FAST, tb.dist_count, tb.lengths, 30);
    tb.dist_walk = tk_huff_walk_from(tb.dist_count, TK_INF_DIST_FAST);
    return true;
}

TK_HD bool tk_inf_dynamic_tables(tk_inflate_state& s, tk_inflate_tables& tb) {
    tk_inf_refill(s);
    const int nlen = (int)tk_inf_bits(s, 5) + 257, ndist = (int)tk_inf_bits(s, 5) + 1, ncode = (int)tk_inf_bits(s, 4) + 4;
    if (nlen > 286 || ndist > 30) return false;
    const char* order = "\x10\x11\x12\x00\x08\x07\x09\x06\x0a\x05\x0b\x04\x0c\x03\x0d\x02\x0e\x01\x0f";
    for (int i = 0; i < 19; ++i) tb.lengths[i] = 0;
    for (int i = 0; i < ncode; ++i) {
        tk_inf_refill(s);
        tb.lengths[(int)order[i]] = (uint16_t)tk_inf_bits(s, 3);
    }
    // the code of the code lengths lives in the literal table for the moment (codes of <= 7 bits)
    if (tk_huff_build(tb.lit_count, tb.lit_symbol, tb.lengths, 19) != 0) return false;      // must be complete
    tk_huff_fast(tb.code_fast, 7, tb.lit_count, tb.lengths, 19);
    int i = 0;
    while (i < nlen + ndist) {
        tk_inf_refill(s);
        const int sym = tk_huff_decode(s, tb.lit_count, tb.lit_symbol, tb.code_fast, 7);
        if (sym < 0) return false;
        if (sym < 16) { tb.lengths[i++] = (uint16_t)sym; continue; }
        int rep, val = 0;
        if (sym == 16) {
            if (i == 0) return false;
            val = tb.lengths[i - 1];
            rep = 3 + (int)tk_inf_bits(s, 2);
        } else if (sym == 17) {
            rep = 3 + (int)tk_inf_bits(s, 3);
        } else {
            rep = 11 + (int)tk_inf_bits(s, 7);
        }
        if (i + rep > nlen + ndist) return false;
        while (rep--) tb.lengths[i++] = (uint16_t)val;
    }
    if (tk_inf_consumed(s) > s.end) return false;
    if (tb.lengths[256] == 0) return false;                    // no end-of-block code
    // distance lengths first (building the literal table overwrites nothing they need, but the
    // literal build reads lengths[0, nlen) which must stay intact: build dist from its own range)
    int err = tk_huff_build(tb.dist_count, tb.dist_symbol, tb.lengths + nlen, ndist);
    if (err && (err < 0 || ndist != tb.dist_count[0] + tb.dist_count[1])) return false;   // incomplete: only a single 1-bit code
    tk_huff_fast_entries<true>(tb.dist_fast, TK_INF_DIST_FAST, tb.dist_count, tb.lengths + nlen, ndist);
    tb.dist_walk = tk_huff_walk_from(tb.dist_count, TK_INF_DIST_FAST);
    err = tk_huff_build(tb.lit_count, tb.lit_symbol, tb.lengths, nlen);
    if (err && (err < 0 || nlen != tb.lit_count[0] + tb.lit_count[1])) return false;
    tk_huff_fast_entries<false>(tb.lit_fast, TK_INF_LIT_FAST, tb.lit_count, tb.lengths, nlen);
    tb.lit_walk = tk_huff_walk_from(tb.lit_count, TK_INF_LIT_FAST);
    return true;
}

// Decodes on from the current state: literals go to out[op++] (capacity cap), matches of at most
// `self_copy` bytes are copied here, byte by byte; returns at the next longer match with *mlen /
// *mdist set (the caller copies out[op - dist ..] and advances op by mlen; both are checked
// against op and cap here), at the end of the stream, or on an error.
// With `batch`: no match is copied or returned; each is appended to the batch and op moves past it.
// Returns TK_INF_BATCH when the batch is full; at the end of the stream (or on an error) the batch
// may hold matches too.  The caller copies the first s.nb entries and sets s.nb = 0.
TK_HD int tk_inflate_next(tk_inflate_state& s, tk_inflate_tables& tb, uint8_t* out, uint32_t cap, uint32_t& op,
                          uint32_t* mlen, uint32_t* mdist, uint32_t self_copy = 0, tk_inflate_batch* batch = nullptr) {
    for (;;) {
        if (s.mode == TK_INF_HEADER) {
            if (s.last) {
                s.mode = TK_INF_END;
                return tk_inf_consumed(s) > s.end ? TK_INF_ERROR : TK_INF_DONE;   // ended on bits that are not there
            }
            tk_inf_refill(s);
            s.last = (int)tk_inf_bits(s, 1);
            const uint32_t type = tk_inf_bits(s, 2);
            if (type == 0) {                                    // stored
                tk_inf_bits(s, s.cnt & 7);                      // to the byte boundary
                tk_inf_refill(s);
                const uint32_t len = tk_inf_bits(s, 16), nlen = tk_inf_bits(s, 16);
                if ((len ^ nlen) != 0xFFFFu || len > cap - op) return TK_INF_ERROR;
                for (uint32_t k = 0; k < len; ++k) {
                    tk_inf_refill(s);
                    out[op++] = (uint8_t)tk_inf_bits(s, 8);
                }
                if (tk_inf_consumed(s) > s.end) return TK_INF_ERROR;
                continue;
            }
            if (type == 1) tk_inf_fixed_tables(tb);
            else if (type == 2) { if (!tk_inf_dynamic_tables(s, tb)) return TK_INF_ERROR; }
            else return TK_INF_ERROR;
            s.mode = TK_INF_CODES;
        }
        if (s.mode == TK_INF_END) return TK_INF_DONE;
        // (the error conditions of a symbol are collected and branched on once: with a single lane
        // active every branch costs its convergence bookkeeping)
        for (;;) {
            tk_inf_refill(s);                                   // >= 33 bits: literal/length code 15 + 5 extra
            // (the input position is only checked once the input is all loaded)
            bool bad = s.pos > s.end && tk_inf_consumed(s) > s.end;
            uint32_t e = tb.lit_fast[(uint32_t)s.buf & ((1u << TK_INF_LIT_FAST) - 1u)];
            if (e) {
                s.buf >>= (e & 15u);
                s.cnt -= (int)(e & 15u);
            } else {
                e = tk_inf_lit_entry(tk_huff_decode_slow(s, tb.lit_count, tb.lit_symbol, tb.lit_walk), 0u);
            }
            const uint32_t kind = e & TK_INF_E_KIND;
            if (kind == TK_INF_E_LITERAL) {
                if (bad | (op >= cap)) return TK_INF_ERROR;
                out[op++] = (uint8_t)(e >> 16);
                continue;
            }
            if (kind == TK_INF_E_END) {
                if (bad) return TK_INF_ERROR;
                s.mode = TK_INF_HEADER;
                break;
            }
            bad |= kind == TK_INF_E_BAD;
            const uint32_t len = (e >> 16) + tk_inf_bits(s, (int)((e >> 4) & 15u));    // 0..5 extra bits
            tk_inf_refill(s);                                   // distance code 15 + 13 extra
            uint32_t d = tb.dist_fast[(uint32_t)s.buf & ((1u << TK_INF_DIST_FAST) - 1u)];
            if (d) {
                s.buf >>= (d & 15u);
                s.cnt -= (int)(d & 15u);
            } else {
                d = tk_inf_dist_entry(tk_huff_decode_slow(s, tb.dist_count, tb.dist_symbol, tb.dist_walk), 0u);
            }
            bad |= (d & TK_INF_E_KIND) == TK_INF_E_BAD;
            const uint32_t dist = (d >> 16) + tk_inf_bits(s, (int)((d >> 4) & 15u));   // 0..13 extra bits
            bad |= s.pos > s.end && tk_inf_consumed(s) > s.end;
            if (bad | (dist > op) | (len > cap - op)) return TK_INF_ERROR;
            if (batch) {
                tk_inflate_match m;
                m.dst = op; m.len = len; m.dist = dist; m.reserved = 0;
                batch->m[s.nb++] = m;
                op += len;
                if (s.nb == TK_INF_BATCH_MAX) return TK_INF_BATCH;
                continue;
            }
            if (len <= self_copy) {
                const uint8_t* src = out + op - dist;
                for (uint32_t k = 0; k < len; ++k) out[op + k] = src[k];
                op += len;
                continue;
            }
            *mlen = len;
            *mdist = dist;
            return TK_INF_MATCH;
        }
    }
}
