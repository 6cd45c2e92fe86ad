// tk_json_kernels.cu -- K1a / K1b (JSON numeric array decode) fused with K3
// (collate, pad, length, mask).
//
// Replaces, per record, the reference's user hook
//     elements = json.loads(record.value)             README.md:74
//     if len(elements) < self.min_size: return None   README.md:76-77
//     torch.tensor(elements, dtype=...)               README.md:41-43 / BASELINE.md §3
// and, per batch, default_collate -> torch.stack (collate.py:275) or the pad
// collate of SURVEY.md §8c.  Grammar and conversions: tk_json.cuh / tk_number.cuh.
//
// Work decomposition (one warp per record):
//   tiles of 2 KiB of the record are staged in a 4 KiB shared-memory ring with
//   coalesced 16-byte loads; while a tile is in registers its bytes are
//   classified (SWAR byte compares) into a "number-like" bit mask; run starts are
//   mask & ~(mask << 1); a warp prefix sum turns them into element indices and a
//   per-tile list of token positions in shared memory; then the 32 lanes parse 32
//   tokens at a time out of the ring (a token may run into the next tile, which
//   is already resident) and store element e of the record at row[e] -- lanes
//   hold consecutive e, so stores coalesce.  The rest of the row is padded.
#include "tk_internal.h"
#include "tk_json.cuh"

#define JD_WARPS 4
#define JD_TILE 2048u
#define JD_RING 4096u
#define JD_MAXTOK (JD_TILE / 2 + 8)

// ---------------------------------------------------------------- SWAR helpers
// 0xFF in every byte lane of x that is number-like: [0-9], [A-Za-z], '+', '-', '.'
__device__ __forceinline__ uint32_t numclass4(uint32_t x) {
    const uint32_t dig = __vcmpleu4(__vsub4(x, 0x30303030u), 0x09090909u);
    const uint32_t let = __vcmpleu4(__vsub4(x | 0x20202020u, 0x61616161u), 0x19191919u);
    const uint32_t pun = __vcmpleu4(__vsub4(x, 0x2b2b2b2bu), 0x03030303u) & ~__vcmpeq4(x, 0x2c2c2c2cu);
    return dig | let | pun;
}
// one bit per byte lane (bit b set iff byte b of m is 0xFF)
__device__ __forceinline__ uint32_t bytemask_to_bits(uint32_t m) {
    const uint32_t t = m & 0x01010101u;   // bits 0, 8, 16, 24
    return (t | (t >> 7) | (t >> 14) | (t >> 21)) & 0xFu;
}
__device__ __forceinline__ uint32_t class16(const uint4 v) {
    return bytemask_to_bits(numclass4(v.x)) | (bytemask_to_bits(numclass4(v.y)) << 4) |
           (bytemask_to_bits(numclass4(v.z)) << 8) | (bytemask_to_bits(numclass4(v.w)) << 12);
}
// bits [lo, hi) of a 16-bit mask, lo/hi clamped to [0,16]
__device__ __forceinline__ uint32_t range16(int lo, int hi) {
    lo = max(lo, 0); hi = min(hi, 16);
    if (hi <= lo) return 0u;
    return ((1u << hi) - 1u) & ~((1u << lo) - 1u);
}

__device__ __forceinline__ void report_error(tk_result_t* res, uint32_t rec, int status) {
    atomicAdd(&res->n_errors, 1u);
    const uint32_t prev = atomicMin(&res->first_error_rec, rec);
    // status of the lowest-indexed bad record: last writer among decreasing minima wins;
    // the host re-reads rec_status[first_error_rec] when it needs the exact code.
    if (rec < prev) res->first_error_status = (uint32_t)status;
}

// ---------------------------------------------------------------- K1 phase 1
// Element count per record = commas + 1 for a non-empty array.  (Only valid
// records need a correct count: malformed ones are flagged by the decode kernel
// and poison their batch.)  keep flag: rec_kidx = 0 (kept) / -1 (skipped).
__global__ void __launch_bounds__(256) k_json_scan(const uint8_t* __restrict__ bytes,
                                                    const tk_rec_t* __restrict__ recs, uint32_t n,
                                                    uint32_t max_len, uint32_t min_size,
                                                    int32_t* __restrict__ rec_len,
                                                    int32_t* __restrict__ rec_status,
                                                    int32_t* __restrict__ rec_kidx) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t i = warp; i < n; i += nwarps) {
        const uint32_t start = recs[i].start, end = recs[i + 1].start;
        uint32_t commas = 0, content = 0;
        for (uint32_t pos = (start & ~15u) + lane * 16; pos < end; pos += 512) {
            const uint4 v = __ldg(reinterpret_cast<const uint4*>(bytes + pos));
            const uint32_t valid = range16((int)start - (int)pos, (int)end - (int)pos);
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
            uint32_t cm = 0, ct = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                cm |= bytemask_to_bits(__vcmpeq4(w[k], 0x2c2c2c2cu)) << (4 * k);
                const uint32_t dig = __vcmpleu4(__vsub4(w[k], 0x30303030u), 0x09090909u);
                // N(aN), I(nfinity), t(rue), f(alse): the non-digit ways an element can start
                const uint32_t word = __vcmpeq4(w[k], 0x4e4e4e4eu) | __vcmpeq4(w[k], 0x49494949u) |
                                      __vcmpeq4(w[k], 0x74747474u) | __vcmpeq4(w[k], 0x66666666u);
                ct |= bytemask_to_bits(dig | word) << (4 * k);
            }
            commas += __popc(cm & valid);
            content |= ct & valid;
        }
        commas = __reduce_add_sync(0xffffffffu, commas);
        content = __reduce_or_sync(0xffffffffu, content);
        if (lane == 0) {
            const uint32_t len = content ? commas + 1 : 0;
            rec_len[i] = (int32_t)len;
            const bool too_long = len > max_len;
            rec_status[i] = too_long ? TK_ERR_TOO_LONG : TK_OK;
            rec_kidx[i] = (!too_long && len >= min_size) ? 0 : -1;
        }
    }
}

extern "C" int tk_json_scan(tk_ctx* ctx, int slot, uint32_t max_len, uint32_t min_size, void* stream) {
    TK_ENTER(ctx, slot);
    tk_slot& s = ctx->slots[slot];
    const uint32_t n = s.n_records;
    if (!n) return 0;
    const uint32_t blocks = min((n + 7) / 8, 148u * 16u);
    k_json_scan<<<blocks, 256, 0, (cudaStream_t)stream>>>(s.d.bytes, s.d.recs, n, max_len, min_size,
                                                         s.d.rec_len, s.d.rec_status, s.d.rec_kidx);
    TK_LAUNCH_CHECK();
    return 0;
}

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
__global__ void __launch_bounds__(JD_WARPS * 32) k_json_decode(const decode_args a) {
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
                const int st = tk_json_element<KIND>(rd, pos, &v);
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

extern "C" int tk_json_decode(tk_ctx* ctx, int slot, int dtype, uint32_t max_len, uint32_t fixed_len,
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
