// tk_json_warp.cu -- the first decode design, kept as an independent cross-check of the flat
// path (tests/test_gpu_parity.py::test_two_decode_decompositions_agree) and for comparison runs:
// K1 phase 2 + K3 in ONE kernel, one warp per record.
//
//   tiles of 2 KiB of the record are staged in a 4 KiB shared-memory ring with coalesced 16-byte
//   loads; while a tile is in registers its bytes are classified (SWAR) into a "number-like" bit
//   mask; run starts are mask & ~(mask << 1); a warp prefix sum turns them into element indices
//   and a per-tile list of token positions in shared memory; then the 32 lanes parse 32 tokens at
//   a time out of the ring (a token may run into the next tile, which is already resident) and
//   store element e of the record at row[e]; the rest of the row is padded.
//
// ncu (profiles/r01_kernels_ncu.md, history table): 24.8 of 32 lanes active, 37 % occupancy,
// 2.4x slower than the flat path for the same work -- which is why tk_json_decode no longer
// uses it.  The grammar is validated with the number-run decomposition (tk_json_gap).
#include "tk_swar.cuh"

#define JD_WARPS 4
#define JD_TILE 2048u
#define JD_RING 4096u
#define JD_MAXTOK (JD_TILE / 2 + 8)

// ---------------------------------------------------------------- K1 phase 2 + K3
struct RingReader {
    const uint8_t* ring;    // this warp's 4 KiB ring (index = absolute position & 4095)
    const uint8_t* gbytes;  // slot byte buffer, for the rare token that outruns the ring
    uint32_t lo, hi;        // absolute positions resident in the ring: [lo, hi)
    uint32_t end;           // end of the record
    __device__ __forceinline__ int operator()(uint32_t pos) const {
        if (pos >= end) return -1;
        if (pos - lo < hi - lo) return ring[pos & (JD_RING - 1)];
        return gbytes[pos];
    }
};

// ---------------------------------------------------------------- SWAR fast path
// The common token shape  [-]digits[.digits]  with at most 19 digits and no exponent is
// scanned 8 bytes at a time straight out of the shared-memory ring: one unaligned 8-byte
// window (3 aligned 32-bit loads + 2 funnel shifts), a carry-free "is this byte a digit"
// mask, count-trailing-zeros for the run length, and the classic multiply-shift reduction
// of 4 ASCII digits to a number.  It produces the same tk_token the scalar scanner
// (tk_scan_number) would, so the conversion code after it is shared; anything else
// (exponents, > 19 digits, NaN/Infinity, true/false, malformed text, tokens that are not
// fully resident in the ring) returns false and takes the scalar, fully general path.
__device__ __forceinline__ Window8 ring_load8(const uint8_t* ring, uint32_t pos) {
    const uint32_t* r32 = reinterpret_cast<const uint32_t*>(ring);
    const uint32_t i = (pos >> 2) & (JD_RING / 4 - 1);
    const uint32_t w0 = r32[i], w1 = r32[(i + 1) & (JD_RING / 4 - 1)], w2 = r32[(i + 2) & (JD_RING / 4 - 1)];
    const uint32_t sh = (pos & 3u) * 8u;
    Window8 w;
    w.lo = __funnelshift_r(w0, w1, sh);
    w.hi = __funnelshift_r(w1, w2, sh);
    return w;
}
// number of leading (lowest-address) bytes of the window that are ASCII digits: 0..8
__device__ __forceinline__ uint32_t leading_digits(const Window8 w) {
    const uint32_t ul = w.lo ^ 0x30303030u, uh = w.hi ^ 0x30303030u;
    const uint32_t nl = (((ul & 0x7F7F7F7Fu) + 0x76767676u) | ul) & 0x80808080u;   // 0x80 where not a digit
    const uint32_t nh = (((uh & 0x7F7F7F7Fu) + 0x76767676u) | uh) & 0x80808080u;
    const unsigned long long nd = ((unsigned long long)nh << 32) | nl;
    return (uint32_t)__clzll((long long)__brevll(nd)) >> 3;                          // 64 >> 3 = 8 if none
}
// Scan one run of digits starting at *cur (absolute position); accumulates into *w, counts into
// *ndig; on return `last` is the window that contains the first non-digit at index `m_last`.
__device__ __forceinline__ bool digit_run(const uint8_t* ring, uint32_t end, uint32_t* cur, uint64_t* w,
                                          uint32_t* ndig, uint32_t* term) {
    uint32_t n_here = 0;
    for (;;) {
        const Window8 win = ring_load8(ring, *cur);
        uint32_t m = leading_digits(win);
        const uint32_t left = end - *cur;                       // bytes of the record from *cur on
        if (m > left) m = left;
        if (m) {
            if (*ndig + m > 19) return false;
            *w = *w * c_pow10[m] + parse_upto8(win, m);
            *ndig += m;
            n_here += m;
            *cur += m;
        }
        if (m < 8) {
            *term = (m < left) ? byte_of(win, m) : 0xFFFFFFFFu;  // 0xFFFFFFFF: the record ends here
            return n_here > 0;
        }
    }
}

template <class R>
__device__ __forceinline__ bool fast_scan_number(const R& rd, uint32_t pos, tk_token* out) {
    // the token, its separator and the start of the next token must be ring resident
    if (pos < rd.lo || pos + 64 > rd.hi) return false;
    const uint8_t* ring = rd.ring;
    const uint32_t end = rd.end;
    uint32_t cur = pos;
    const Window8 first = ring_load8(ring, cur);
    const uint32_t c0 = first.lo & 0xFFu;
    uint32_t neg = 0;
    if (c0 == '-') { neg = 1; cur += 1; }
    uint64_t w = 0;
    uint32_t ndig = 0, term = 0;
    const uint32_t int_begin = cur;
    if (!digit_run(ring, end, &cur, &w, &ndig, &term)) return false;
    const uint32_t int_digits = ndig;
    // "01" is not a JSON number (the scalar path reports it); a lone "0" is
    if (int_digits > 1 && rd.ring[int_begin & (JD_RING - 1)] == '0') return false;
    uint32_t frac_digits = 0;
    if (term == '.') {
        cur += 1;
        if (!digit_run(ring, end, &cur, &w, &ndig, &term)) return false;
        frac_digits = ndig - int_digits;
    }
    if (term == 'e' || term == 'E') return false;
    tk_token t;
    t.w = w; t.q = -(int32_t)frac_digits; t.end = cur; t.dig_begin = int_begin; t.exp10 = 0;
    t.neg = (uint8_t)neg; t.is_int = frac_digits == 0; t.truncated = 0; t.special = 0; t.err = TK_OK;
    t.nsig = (int32_t)ndig; t.frac_digits = (int32_t)frac_digits;
    *out = t;
    return true;
}

// separator after a number that ended at `pos`: fast accept of ",<token>" / ", <token>"
template <class R>
__device__ __forceinline__ int fast_gap(const R& rd, uint32_t pos) {
    if (pos >= rd.lo && pos + 16 <= rd.hi && pos + 3 < rd.end) {
        const Window8 g = ring_load8(rd.ring, pos);
        if ((g.lo & 0xFFu) == ',') {
            const uint32_t c1 = (g.lo >> 8) & 0xFFu;
            const uint32_t c = c1 == ' ' ? ((g.lo >> 16) & 0xFFu) : c1;
            if (tk_is_numclass((int)c)) return TK_OK;
        }
    }
    return tk_json_gap(rd, pos);
}

// One-shot variant of the fast path for the dominant token shape: sign, at most 8 integer
// digits, optional '.', at most 16 fraction digits, everything including the terminator inside
// the first 24 bytes.  The digit/non-digit mask of those 24 bytes is computed once (six words),
// run lengths come from find-first-set on that mask, and the separator that follows is checked
// from the same neighbourhood.  On success *gap is TK_OK if the separator was accepted too, or
// -1 if the caller still has to run the general separator check at t->end.
template <class R>
__device__ __forceinline__ bool fast_token24(const R& rd, uint32_t pos, tk_token* out, int* gap) {
    if (pos < rd.lo || pos + 64 > rd.hi) return false;
    const uint32_t avail = rd.end - pos;                       // bytes of the record from pos on (>= 1)
    const uint32_t* r32 = reinterpret_cast<const uint32_t*>(rd.ring);
    const uint32_t i = pos >> 2, sh = (pos & 3u) * 8u;
    constexpr uint32_t M = JD_RING / 4 - 1;
    uint32_t a[7];
#pragma unroll
    for (int k = 0; k < 7; ++k) a[k] = r32[(i + k) & M];
    uint32_t W[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) W[k] = __funnelshift_r(a[k], a[k + 1], sh);
    uint32_t nd = 0xFF000000u;                                 // bytes 24.. count as "not a digit"
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        const uint32_t u = W[k] ^ 0x30303030u;
        const uint32_t t = (((u & 0x7F7F7F7Fu) + 0x76767676u) | u) & 0x80808080u;
        nd |= hibits_to_nibble(t) << (4 * k);
    }
    if (avail < 24) nd |= 0xFFFFFFFFu << avail;
    const uint32_t neg = (W[0] & 0xFFu) == '-' ? 1u : 0u;
    const uint32_t p = neg;
    const uint32_t L1 = __ffs(nd >> p) - 1;                    // integer digits
    if (L1 == 0 || L1 > 8) return false;
    const Window8 wi = {__funnelshift_r(W[0], W[1], 8 * p), __funnelshift_r(W[1], W[2], 8 * p)};
    if (L1 > 1 && (wi.lo & 0xFFu) == '0') return false;        // "01": the general path reports it
    const uint32_t tpos = p + L1;                              // <= 9
    const uint32_t tword = tpos < 4 ? W[0] : (tpos < 8 ? W[1] : W[2]);
    const uint32_t tc = (tword >> (8 * (tpos & 3u))) & 0xFFu;
    uint32_t L2 = 0, endp = tpos;
    uint64_t w = parse_upto8(wi, L1);
    if (tc == '.') {
        const uint32_t q0 = tpos + 1;
        L2 = __ffs(nd >> q0) - 1;                              // fraction digits
        if (L2 == 0 || L2 > 16 || L1 + L2 > 19) return false;
        endp = q0 + L2;
        if (endp > 23) return false;                           // terminator not inside the mask
        const uint32_t m0 = L2 < 8 ? L2 : 8, m1 = L2 - m0;
        w = w * c_pow10[m0] + parse_upto8(ring_load8(rd.ring, pos + q0), m0);
        if (m1) w = w * c_pow10[m1] + parse_upto8(ring_load8(rd.ring, pos + q0 + 8), m1);
    }
    if (endp >= avail) return false;                           // record ends inside the token
    // separator: the common ", x" / ",x" shapes are accepted here
    const uint32_t gi = (pos + endp) >> 2, gs = ((pos + endp) & 3u) * 8u;
    const uint32_t g = __funnelshift_r(r32[gi & M], r32[(gi + 1) & M], gs);
    const uint32_t g0 = g & 0xFFu;
    if (g0 == 'e' || g0 == 'E') return false;                  // exponent: general path
    *gap = -1;
    if (g0 == ',') {
        const uint32_t g1 = (g >> 8) & 0xFFu;
        if (g1 == ' ') { if (endp + 2 < avail && tk_is_numclass((int)((g >> 16) & 0xFFu))) *gap = TK_OK; }
        else if (endp + 1 < avail && tk_is_numclass((int)g1)) *gap = TK_OK;
    }
    tk_token t;
    t.w = w; t.q = -(int32_t)L2; t.end = pos + endp; t.dig_begin = pos + p; t.exp10 = 0;
    t.neg = (uint8_t)neg; t.is_int = L2 == 0; t.truncated = 0; t.special = 0; t.err = TK_OK;
    t.nsig = (int32_t)(L1 + L2); t.frac_digits = (int32_t)L2;
    *out = t;
    return true;
}

template <int KIND, class R>
__device__ __forceinline__ int json_element_fast(const R& rd, uint32_t pos, typename tk_elem<KIND>::type* out) {
    tk_token t;
    int st = -1;
    if (!fast_token24(rd, pos, &t, &st)) {
        if (!fast_scan_number(rd, pos, &t)) return tk_json_element<KIND>(rd, pos, out);
    }
    if (st) st = fast_gap(rd, t.end);
    if (st) return st;
    if (KIND == TK_F32 || KIND == TK_F64) {
        double d;
        st = tk_token_to_f64(rd, t, &d);
        if (st) return st;
        *out = (typename tk_elem<KIND>::type)d;
    } else {
        int64_t v;
        st = tk_token_to_i64(rd, t, KIND == TK_I32 ? 32 : 64, &v);
        if (st) return st;
        *out = (typename tk_elem<KIND>::type)v;
    }
    return TK_OK;
}

struct decode_args {
    const uint8_t*  bytes;
    const tk_rec_t* recs;
    uint32_t        n;
    uint32_t        max_len;
    uint32_t        fixed_len;     // 0 = variable length
    const int32_t*  rec_kidx;
    int32_t*        rec_len;
    int32_t*        rec_status;
    tk_result_t*    res;
    tk_plane_t      values, lengths, mask;
};

template <int KIND>
__global__ void __launch_bounds__(JD_WARPS * 32, 8) k_json_decode(const decode_args a) {
    typedef typename tk_elem<KIND>::type T;
    __shared__ __align__(16) uint8_t s_ring[JD_WARPS][JD_RING];
    __shared__ uint16_t s_tok[JD_WARPS][JD_MAXTOK];

    const uint32_t lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const uint32_t i = blockIdx.x * JD_WARPS + wib;
    if (i >= a.n) return;                         // whole warp exits together; no block barriers below
    uint8_t* ring = s_ring[wib];
    uint16_t* tok = s_tok[wib];

    const uint32_t start = a.recs[i].start, end = a.recs[i + 1].start;
    const int32_t kidx = a.rec_kidx[i];

    // destination row (nullptr when the record was filtered out: still validated, never written)
    T* dst = nullptr;
    int32_t* dst_len = nullptr;
    uint8_t* dst_mask = nullptr;
    if (kidx >= 0) {
        const uint64_t row = (uint64_t)a.res->batch_size + (uint32_t)kidx;
        dst = reinterpret_cast<T*>((uint8_t*)a.values.base + row * a.values.row_bytes);
        if (a.lengths.base) dst_len = reinterpret_cast<int32_t*>((uint8_t*)a.lengths.base + row * a.lengths.row_bytes);
        if (a.mask.base) dst_mask = (uint8_t*)a.mask.base + row * a.mask.row_bytes;
    }

    int err = TK_OK;
    uint32_t elem_base = 0;

    if (end > start) {
        const uint32_t tile0 = start / JD_TILE, tileN = (end - 1) / JD_TILE;
        uint32_t cls_cur[4], cls_next[4] = {0, 0, 0, 0};
        // stage one tile: coalesced 16-byte loads -> ring, and classify while in registers
        auto stage = [&](uint32_t t, uint32_t (&cls)[4]) {
            const uint32_t tbase = t * JD_TILE;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t pos = tbase + k * 512 + lane * 16;
                const uint4 v = __ldg(reinterpret_cast<const uint4*>(a.bytes + pos));
                *reinterpret_cast<uint4*>(ring + (pos & (JD_RING - 1))) = v;
                cls[k] = class16(v) & range16((int)start - (int)pos, (int)end - (int)pos);
            }
        };
        stage(tile0, cls_cur);
        uint32_t prev_tile_last = 0;      // class of the byte just before the current tile
        for (uint32_t t = tile0; t <= tileN; ++t) {
            if (t < tileN) stage(t + 1, cls_next);
            __syncwarp();
            RingReader rd;
            rd.ring = ring; rd.gbytes = a.bytes; rd.end = end;
            rd.lo = t * JD_TILE;
            rd.hi = (t < tileN) ? (t + 2) * JD_TILE : (t + 1) * JD_TILE;

            if (t == tile0 && lane == 0) {
                uint32_t first; bool empty;
                const int st = tk_json_prefix(rd, start, &first, &empty);
                if (st) err = st;
            }

            // run starts of this tile -> token list
            uint32_t ntok = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t m = cls_cur[k];
                uint32_t before = __shfl_up_sync(0xffffffffu, m >> 15, 1);       // last byte of the chunk to the left
                const uint32_t wrap = __shfl_sync(0xffffffffu, k ? cls_cur[k ? k - 1 : 0] >> 15 : prev_tile_last, 31);
                if (lane == 0) before = wrap;
                uint32_t starts = m & ~((m << 1) | (before & 1u)) & 0xFFFFu;
                const uint32_t cnt = __popc(starts);
                uint32_t incl = cnt;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
                    if (lane >= (uint32_t)d) incl += o;
                }
                uint32_t at = ntok + incl - cnt;
                const uint32_t rel = k * 512 + lane * 16;
                while (starts) {
                    const int b = __ffs(starts) - 1;
                    starts &= starts - 1;
                    if (at < JD_MAXTOK) tok[at] = (uint16_t)(rel + b);
                    ++at;
                }
                ntok += __shfl_sync(0xffffffffu, incl, 31);
            }
            prev_tile_last = __shfl_sync(0xffffffffu, cls_cur[3] >> 15, 31);
            __syncwarp();

            // parse 32 tokens at a time
            const uint32_t tbase = t * JD_TILE;
            for (uint32_t j = lane; j < ntok; j += 32) {
                const uint32_t pos = tbase + tok[j];
                const uint32_t e = elem_base + j;
                T v;
                const int st = json_element_fast<KIND>(rd, pos, &v);
                if (st) { if (!err) err = st; }
                else if (dst && e < a.max_len) dst[e] = v;
            }
            elem_base += ntok;
            __syncwarp();
#pragma unroll
            for (int k = 0; k < 4; ++k) cls_cur[k] = cls_next[k];
        }
    } else {
        err = TK_ERR_SYNTAX;      // empty payload: json.loads(b"") raises
    }

    // record-level status
    const uint32_t ntotal = elem_base;
    const int pre = a.rec_status[i];
    if (!err && pre) err = pre;
    if (!err && ntotal > a.max_len) err = TK_ERR_TOO_LONG;
    if (!err && kidx >= 0 && a.fixed_len && ntotal != a.fixed_len) err = TK_ERR_RAGGED;
    const uint32_t bad = __ballot_sync(0xffffffffu, err != TK_OK);
    if (bad) {
        const int code = __shfl_sync(0xffffffffu, err, __ffs(bad) - 1);
        if (lane == 0) {
            a.rec_status[i] = code;
            report_error(a.res, i, code);
        }
        return;
    }
    if (lane == 0) a.rec_len[i] = (int32_t)ntotal;

    if (dst) {
        for (uint32_t e = ntotal + lane; e < a.max_len; e += 32) dst[e] = (T)0;
        if (dst_len && lane == 0) *dst_len = (int32_t)ntotal;
        if (dst_mask) {
            if ((a.max_len & 3u) == 0 && ((uintptr_t)dst_mask & 3u) == 0) {
                uint32_t* m32 = reinterpret_cast<uint32_t*>(dst_mask);
                for (uint32_t w = lane; w < a.max_len / 4; w += 32) {
                    const int left = (int)ntotal - (int)(w * 4);
                    m32[w] = left >= 4 ? 0x01010101u : (left <= 0 ? 0u : (0x01010101u >> (8 * (4 - left))));
                }
            } else {
                for (uint32_t e = lane; e < a.max_len; e += 32) dst_mask[e] = e < ntotal ? 1 : 0;
            }
        }
    }
}

extern "C" int tk_json_decode_warp(tk_ctx* ctx, int slot, int dtype, uint32_t max_len, uint32_t fixed_len,
                                   const tk_plane_t* values, const tk_plane_t* lengths, const tk_plane_t* mask,
                                   void* stream) {
    TK_ENTER(ctx, slot);
    tk_slot& s = ctx->slots[slot];
    cudaStream_t st = (cudaStream_t)stream;
    TK_REQUIRE(values && values->base, "values plane is required");
    TK_REQUIRE(max_len > 0, "max_len must be positive");
    const uint32_t n = s.n_records;
    decode_args a;
    a.bytes = s.d.bytes; a.recs = s.d.recs; a.n = n; a.max_len = max_len; a.fixed_len = fixed_len;
    a.rec_kidx = s.d.rec_kidx; a.rec_len = s.d.rec_len; a.rec_status = s.d.rec_status; a.res = s.d.result;
    tk_plane_t z; z.base = nullptr; z.row_bytes = 0;
    a.values = *values;
    a.lengths = lengths ? *lengths : z;
    a.mask = mask ? *mask : z;
    if (n) {
        const uint32_t blocks = (n + JD_WARPS - 1) / JD_WARPS;
        switch (dtype) {
            case TK_DTYPE_F32: k_json_decode<TK_F32><<<blocks, JD_WARPS * 32, 0, st>>>(a); break;
            case TK_DTYPE_F64: k_json_decode<TK_F64><<<blocks, JD_WARPS * 32, 0, st>>>(a); break;
            case TK_DTYPE_I32: k_json_decode<TK_I32><<<blocks, JD_WARPS * 32, 0, st>>>(a); break;
            case TK_DTYPE_I64: k_json_decode<TK_I64><<<blocks, JD_WARPS * 32, 0, st>>>(a); break;
            default: TK_REQUIRE(false, "unknown dtype %d", dtype);
        }
        TK_LAUNCH_CHECK();
    }
    return 0;
}

