// tk_json_fused.cu -- the JSON array path of a slot in ONE pass over the payload.
//
// Replaces, per record, the same reference lines as tk_json_kernels.cu / tk_plan.cu
//     elements = json.loads(record.value)             README.md:74
//     if len(elements) < self.min_size: return None   README.md:76-77      (src/kafka_dataset.py:161-162)
//     torch.tensor(elements, dtype=...)               README.md:41-43 / BASELINE.md §3
//     batch of batch_size kept samples                torch/utils/data/_utils/fetch.py:27-45
//     torch.stack / pad collate                       collate.py:275 / SURVEY.md §8c
//     commit offsets after each batch                 src/auto_commit.py:55-58 -> src/kafka_dataset.py:130
// but where the three-kernel chain (k_json_scan -> k_plan -> k_json_parse) reads the payload from
// HBM twice and hands an 8-byte token per element from one kernel to the next, this kernel reads
// every payload byte once and keeps everything in between in shared memory:
//
//   tile      the records whose first byte lies in one TILE_BYTES stretch of the slot's byte buffer
//             (k_fused_prep finds the first record and first byte of every tile).  Persistent blocks
//             take tiles in order (ticket counter).
//   load      the tile's bytes come in with one bulk asynchronous copy (cp.async.bulk, completion
//             on an mbarrier) into a shared-memory window; record starts come in meanwhile.
//   cut       every thread classifies 32-byte chunks of the window: comma masks, their prefix
//             sums (block scan) -> the element count of every record (commas + 1), its keep flag
//             (min_size), the head check ([BOM] ws* '['), and a list of segments (one per element:
//             window position, record) scattered into shared memory.
//   parse     thread per element out of the window: the fast lane of tk_json_parse.cuh (what
//             json.dumps writes), else the scalar grammar.  The values of the first round of
//             segments are staged in shared memory: their rows are not known yet.
//   place     single-pass scan with decoupled look-back over the tiles (as tk_plan.cu does over
//             blocks of records): the tile publishes its kept count right after the cut and looks
//             back for its exclusive prefix only after the parse, when the tiles in front of it
//             have long published theirs -> output row of every kept record, batch boundaries,
//             longest record per batch, and the segmented max of (offset + 1) per (batch,
//             partition) for the commit.
//   rows      pad / length / mask of every kept record's row (warp per record); the staged values
//             go to their rows.
//
// A tile that does not fit the window (a record longer than the window's slack) or holds more
// records than the per-tile tables takes a scalar path, thread per record, out of global memory:
// correct, slow, and rare by construction of the window size (tk_json_ingest).
//
// HBM traffic per slot: payload + record headers read once, rows written once (SURVEY.md §8d's
// algorithmic bytes), plus 12 bytes of per-record scratch.
#include "tk_json_parse.cuh"

#define FZ_THREADS 256
#define FZ_WARPS (FZ_THREADS / 32)
#define FZ_TILE_BYTES 16384u
#define FZ_TOKCAP 1536u              // segments per round
#define FZ_CHUNK 32u                 // bytes per comma-mask word

// record info bits kept in shared memory
#define FR_KEEP 1u                   // passes the filter: gets a row
#define FR_TAIL 2u                   // the record's final byte is ']'
#define FR_PARSE 4u                  // status OK and at least one element: its segments are parsed

struct fused_args {
    const uint8_t*  bytes;
    const tk_rec_t* recs;
    uint32_t        n, n_bytes;
    uint32_t        max_len, min_size, fixed_len, batch_size;
    int32_t*        rec_len;
    int32_t*        rec_status;
    int32_t*        rec_kidx;
    const uint2*    tile_info;       // [n_tiles + 1] {first record, its first byte}
    uint32_t        n_tiles;
    uint32_t        win_bytes;       // shared-memory window (multiple of FZ_CHUNK)
    uint32_t        max_recs;        // per-tile record tables
    unsigned long long* tile_agg;
    unsigned long long* tile_incl;
    uint32_t*       tile_flag;
    uint32_t*       ctr;             // [0] tile ticket, [2] blocks that have left
    uint32_t        epoch;
    tk_state*       state;
    tk_result_t*    res;
    uint32_t*       batch_end;
    uint32_t*       batch_lmax;
    int64_t*        commit;
    uint32_t        pitch, n_partitions;
    tk_plane_t      values, lengths, mask;
};

// ---------------------------------------------------------------- prep
// Grid-stride over the records: first record of every tile, reset of what the main kernel
// accumulates into (commit matrix, per-batch arrays, the header's error fields).
struct prep_args {
    const tk_rec_t* recs;
    uint32_t        n, n_tiles;
    uint2*          tile_info;       // [n_tiles + 1] {first record of the tile, its first byte}
    tk_result_t*    res;
    tk_plan_reset   reset;
};

__device__ __forceinline__ uint32_t fz_tile_of(uint32_t start, uint32_t n_tiles) {
    return min(start / FZ_TILE_BYTES, n_tiles - 1u);
}

__global__ void __launch_bounds__(256) k_fused_prep(const prep_args a) {
    tk_pdl_launch_dependents();
    tk_pdl_wait();
    const uint32_t gtid = blockIdx.x * blockDim.x + threadIdx.x, nthr = gridDim.x * blockDim.x;
    tk_plan_reset_run(a.reset, gtid, nthr);
    if (gtid == 0) {
        a.res->n_errors = 0;
        a.res->first_error_rec = 0xFFFFFFFFu;
        a.res->first_error_status = 0;
    }
    // record i is the first of every tile in (tile of record i-1, tile of record i]; the tiles up to
    // the first record's get 0, the ones behind the last record's get n
    for (uint32_t i = gtid; i <= a.n; i += nthr) {
        const uint32_t lo = i == 0 ? 0u : fz_tile_of(a.recs[i - 1].start, a.n_tiles) + 1u;
        const uint32_t hi = i == a.n ? a.n_tiles : fz_tile_of(a.recs[i].start, a.n_tiles);
        const uint32_t byte = a.recs[i].start;            // (i == n: the closing entry, start = n_bytes)
        for (uint32_t t = lo; t <= hi; ++t) a.tile_info[t] = make_uint2(i, byte);
    }
}

// ---------------------------------------------------------------- small helpers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t phase) {
    uint32_t done;
    do {
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(done)
                     : "r"(smem_u32(bar)), "r"(phase)
                     : "memory");
    } while (!done);
}

__device__ __forceinline__ uint32_t fz_ld_acquire(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void fz_st_release(uint32_t* p, uint32_t v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// reads of the window / of the slot buffer for the scalar grammar (tk_json.cuh): byte at `pos`, -1 at `end`
struct WindowReader {
    const uint8_t* win;
    uint32_t end;
    __device__ __forceinline__ int operator()(uint32_t pos) const { return pos < end ? (int)win[pos] : -1; }
};

// comma bits of one aligned 32-byte chunk of the window
__device__ __forceinline__ uint32_t fz_comma_mask32(const uint8_t* p) {
    const uint4 a = *reinterpret_cast<const uint4*>(p), b = *reinterpret_cast<const uint4*>(p + 16);
    uint32_t m = hibits_to_nibble(comma4(a.x));
    m = imad(hibits_to_nibble(comma4(a.y)), 1u << 4, m);
    m = imad(hibits_to_nibble(comma4(a.z)), 1u << 8, m);
    m = imad(hibits_to_nibble(comma4(a.w)), 1u << 12, m);
    m = imad(hibits_to_nibble(comma4(b.x)), 1u << 16, m);
    m = imad(hibits_to_nibble(comma4(b.y)), 1u << 20, m);
    m = imad(hibits_to_nibble(comma4(b.z)), 1u << 24, m);
    m = imad(hibits_to_nibble(comma4(b.w)), 1u << 28, m);
    return m;
}
// bits [lo, hi) of a 32-bit mask, lo / hi relative to the chunk and clamped to [0, 32]
__device__ __forceinline__ uint32_t fz_range32(int lo, int hi) {
    lo = max(lo, 0); hi = min(hi, 32);
    if (hi <= lo) return 0u;
    const uint32_t upto_hi = hi >= 32 ? 0xFFFFFFFFu : ((1u << hi) - 1u);
    return upto_hi & ~((1u << lo) - 1u);
}

// block-wide exclusive scan of one 32-bit value per thread (FZ_THREADS threads); *total = block sum
__device__ __forceinline__ uint32_t fz_block_scan(uint32_t v, uint32_t* warp_tot, uint32_t* total) {
    const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= (uint32_t)d) incl += o;
    }
    if (lane == 31) warp_tot[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        const uint32_t w = lane < FZ_WARPS ? warp_tot[lane] : 0u;
        uint32_t wi = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t o = __shfl_up_sync(0xffffffffu, wi, d);
            if (lane >= (uint32_t)d) wi += o;
        }
        if (lane < FZ_WARPS) warp_tot[lane] = wi - w;
        if (lane == 31) *total = wi;
    }
    __syncthreads();
    const uint32_t r = warp_tot[wid] + incl - v;
    __syncthreads();
    return r;
}

// Exclusive prefix (packed kept << 40 | elements) of tile `tile`, whole block, result in every
// thread.  Decoupled look-back, FZ_THREADS predecessors per round: with hundreds of tiles in flight
// the nearest predecessor that already knows its inclusive sum is typically a few hundred tiles
// back, and a round costs one acquire-load latency whatever its width.
__device__ __forceinline__ unsigned long long fz_lookback(const fused_args& a, uint32_t tile, uint32_t* s_first,
                                                          unsigned long long* s_sum) {
    const uint32_t tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    unsigned long long excl = 0;
    int p = (int)tile - 1;
    while (p >= 0) {
        const int q = p - (int)tid;                       // thread 0 looks at the nearest predecessor
        uint32_t st = 2;                                  // tiles before 0: "inclusive sum 0"
        if (q >= 0) {
            uint32_t f;
            do { f = fz_ld_acquire(a.tile_flag + q); } while ((f >> 2) != a.epoch);
            st = f & 3u;
        }
        const uint32_t full = __ballot_sync(0xffffffffu, st == 2);
        if (lane == 0) s_first[wid] = full ? wid * 32u + (uint32_t)__ffs(full) - 1u : 0xFFFFFFFFu;
        __syncthreads();
        uint32_t first = 0xFFFFFFFFu;                     // nearest inclusive sum of this round
#pragma unroll
        for (int w = 0; w < FZ_WARPS; ++w) first = min(first, s_first[w]);
        unsigned long long v = 0;
        if (q >= 0 && tid <= first) v = tid == first ? a.tile_incl[q] : a.tile_agg[q];
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
        if (lane == 0) s_sum[wid] = v;
        __syncthreads();
#pragma unroll
        for (int w = 0; w < FZ_WARPS; ++w) excl += s_sum[w];
        __syncthreads();                                  // s_first / s_sum are reused by the next round
        if (first != 0xFFFFFFFFu) break;
        p -= FZ_THREADS;
    }
    return excl;
}

#define FZ_PACK_SHIFT 40

// ---------------------------------------------------------------- shared-memory layout
// Dynamic shared memory, carved by the kernel:
//   win     [win_bytes + 64]   the tile's bytes (16-byte aligned, padded for the fast lane's window reads)
//   val     [TOKCAP]   T       values of the first round of segments, until their rows are known
//   cmask   [nch]      u32     comma bits per 32-byte chunk
//   pre     [nch + 1]  u32     exclusive prefix per chunk: commas (low 16 bits) | record starts (high 16)
//   rcnt    [nch]      u16     records starting in the chunk
//   tok     [TOKCAP+1] u32     segment: window position (low 16) | local record (high 16)
//   rst     [max_recs + 1] u16 record start in the window
//   fseg    [max_recs + 1] u16 first segment of the record
//   rfirst  [max_recs] u16     where the record's first element starts (behind '[')
//   rlen    [max_recs] u16     element count
//   rrow    [max_recs] u32     output row, 0xFFFFFFFF if skipped
//   rpart   [max_recs] u32     partition | flags of the record header
//   rinfo   [max_recs] u8      FR_*
static size_t fz_smem_bytes(uint32_t win_bytes, uint32_t max_recs, size_t elem_bytes) {
    const size_t nch = win_bytes / FZ_CHUNK;
    size_t b = win_bytes + 64 + elem_bytes * FZ_TOKCAP;
    b += 4 * nch + 4 * (nch + 1) + 4 * (FZ_TOKCAP + 1) + 8 * (size_t)max_recs;       // u32 arrays
    b += 2 * nch + 2 * (size_t)(max_recs + 1) * 2 + 2 * (size_t)max_recs * 2;         // u16 arrays
    b += max_recs;                                                                    // u8
    return (b + 64 + 15) & ~(size_t)15;
}

// Segment list of one round [sbase, sbase + FZ_TOKCAP]: the first segment of every record and one
// behind every comma, window position | local record << 16, by `nthr` threads (`t` = this one).
__device__ __forceinline__ void fz_scatter(uint32_t sbase, uint32_t n_seg, uint32_t m, uint32_t nch, uint32_t t, uint32_t nthr,
                                           const uint32_t* cmask, const uint32_t* pre, const uint16_t* rcnt,
                                           const uint16_t* rst, const uint16_t* fseg, const uint16_t* rfirst, uint32_t* tok) {
    for (uint32_t k = t; k < m; k += nthr) {
        const uint32_t s = (uint32_t)fseg[k] - sbase;
        if (s <= FZ_TOKCAP) tok[s] = (uint32_t)rfirst[k] | (k << 16);
    }
    for (uint32_t c = t; c < nch; c += nthr) {
        uint32_t mk = cmask[c];
        if (!mk) continue;
        const uint32_t pc = pre[c];
        uint32_t kc = pc & 0xFFFFu;                         // commas before the chunk
        const uint32_t rbefore = pc >> 16;                   // records started before the chunk
        uint32_t cand = rbefore;
        const uint32_t cend = rbefore + rcnt[c];
        uint32_t k = rbefore - 1u;                           // the record the chunk starts in
        while (mk) {
            const uint32_t b = __ffs(mk) - 1;
            mk &= mk - 1;
            const uint32_t pos = c * FZ_CHUNK + b;
            while (cand < cend && rst[cand] <= pos) { k = cand; ++cand; }
            const uint32_t s = kc + k + 1u - sbase;
            if (s <= FZ_TOKCAP) tok[s] = (pos + 1u) | (k << 16);
            ++kc;
        }
    }
}

// One round of segments, thread per element out of the window.  STAGE: the value goes to the
// staging array (its row is not known yet); else straight to the record's row.
template <int KIND, bool STAGE>
__device__ __forceinline__ void fz_parse_round(const fused_args& a, const tk_fl_tables& tb, const uint8_t* win,
                                               const uint32_t* tok, const uint8_t* rinfo, const uint16_t* fseg,
                                               const uint16_t* rlen, const uint16_t* rst, const uint32_t* rrow,
                                               typename tk_elem<KIND>::type* val, typename tk_elem<KIND>::type* values,
                                               uint32_t sbase, uint32_t cnt, uint32_t r0, uint32_t tid) {
    typedef typename tk_elem<KIND>::type T;
    for (uint32_t s = tid; s < cnt; s += FZ_THREADS) {
        const uint32_t t = tok[s];
        const uint32_t pos = t & 0xFFFFu, k = t >> 16;
        const uint32_t info = rinfo[k];
        if (!(info & FR_PARSE)) continue;
        const uint32_t e = sbase + s - fseg[k];             // element index inside the record
        const uint32_t end = rst[k + 1];
        const bool last = e + 1 == rlen[k];
        // the element ends at the next ',' -- or, for the last one, at the ']' that closes the record
        const uint32_t stop = (last ? end : (tok[s + 1] & 0xFFFFu)) - 1u;
        T v;
        int st = TK_OK;
        uint64_t w;
        uint32_t K, dot, neg;
        // (the last element takes the fast lane only when the record's final byte is ']'; an element
        // that ends in the window's first 20 bytes cannot be anchored at its end: scalar grammar)
        bool fast = !(last && !(info & FR_TAIL)) && tk_fl_scan(SmemWords{win}, pos, stop, tb, &w, &K, &dot, &neg);
        if (fast) fast = tk_fl_value<KIND>(w, K, dot, neg, tb, &v);
        if (!fast) {
            // the lane's long forms ("0.00…" with 20 digits, exponents), then the grammar
            bool done = false;
            if (!(last && !(info & FR_TAIL))) {
                const tk_fl_num num = tk_fl_scan_long(SmemWords{win}, pos, stop);
                done = num.ok && tk_fl_value_long<KIND>(num, &v);
            }
            if (!done) {
                WindowReader rd{win, end};
                st = tk_json_element_flat<KIND>(rd, pos, &v);
            }
        }
        if (st) { flag_record(a.rec_status, a.res, r0 + k, st); continue; }
        if (STAGE) {
            val[s] = v;
        } else {
            const uint32_t row = rrow[k];
            if (row != 0xFFFFFFFFu && e < a.max_len) values[(uint64_t)row * a.max_len + e] = v;
        }
    }
}

template <int KIND>
__global__ void __launch_bounds__(FZ_THREADS, 4) k_json_fused(const fused_args a) {
    typedef typename tk_elem<KIND>::type T;
    extern __shared__ __align__(128) uint8_t fz_smem[];
    __shared__ tk_fl_tables tb;
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ uint32_t s_warp_tot[32];
    __shared__ uint32_t s_total, s_kept, s_elems;
    __shared__ uint32_t s_next[5];          // the block's next tile: index, first record, first byte, end record, end byte
    __shared__ uint32_t s_first[FZ_WARPS];
    __shared__ unsigned long long s_sum[FZ_WARPS];

    const uint32_t tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t nch = a.win_bytes / FZ_CHUNK;
    uint8_t* win = fz_smem;
    T* val = reinterpret_cast<T*>(win + a.win_bytes + 64);               // staged values of the first round
    uint32_t* cmask = reinterpret_cast<uint32_t*>(val + FZ_TOKCAP);
    uint32_t* pre = cmask + nch;
    uint32_t* tok = pre + nch + 1;
    uint32_t* rrow = tok + FZ_TOKCAP + 1;
    uint32_t* rpart = rrow + a.max_recs;
    uint16_t* rcnt = reinterpret_cast<uint16_t*>(rpart + a.max_recs);
    uint16_t* rst = rcnt + nch;
    uint16_t* fseg = rst + a.max_recs + 1;
    uint16_t* rfirst = fseg + a.max_recs + 1;
    uint16_t* rlen = rfirst + a.max_recs;
    uint8_t* rinfo = reinterpret_cast<uint8_t*>(rlen + a.max_recs);

    tk_pdl_launch_dependents();             // (the commit kernel behind this one waits for all of it)
    tk_fl_tables_fill(tb, threadIdx.x, blockDim.x);   // constants only: overlaps the tail of the previous kernel
    __syncthreads();
    if (tid == 0) {
        mbar_init(&s_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    tk_pdl_wait();

    const uint32_t B = a.batch_size;
    const uint32_t carry = a.state->carry_rows;
    const bool vec_values = ((a.values.row_bytes | (uintptr_t)a.values.base) & 15u) == 0;
    const bool vec_mask = a.mask.base && ((a.mask.row_bytes | (uintptr_t)a.mask.base | a.max_len) & 15u) == 0;
    T* values = reinterpret_cast<T*>(a.values.base);
    uint32_t phase = 0;

    // Tiles are taken from a ticket counter when the block is ready for one (not earlier: a ticket
    // held by a busy block would stall the look-back of every tile behind it).
    for (;;) {
        if (tid == 0) {
            const uint32_t t = atomicAdd(&a.ctr[0], 1u);
            s_next[0] = t;
            if (t < a.n_tiles) {
                const uint2 x = a.tile_info[t], y = a.tile_info[t + 1];
                s_next[1] = x.x; s_next[2] = x.y; s_next[3] = y.x; s_next[4] = y.y;
            }
            s_kept = 0; s_elems = 0;
        }
        __syncthreads();
        const uint32_t tile = s_next[0];
        if (tile >= a.n_tiles) break;
        const uint32_t r0 = s_next[1], lo = s_next[2], r1 = s_next[3], hi = s_next[4];
        const uint32_t m = r1 - r0;
        const uint32_t base = lo & ~15u;                             // window origin in the slot buffer
        const uint32_t span = m ? ((hi + 15u) & ~15u) - base : 0u;   // bytes to bring in
        // (strictly inside the window: positions up to `span` index the per-chunk tables)
        const bool fast_tile = m && span + FZ_CHUNK <= a.win_bytes && m <= a.max_recs;
        if (tid == 0 && fast_tile && span) {
            mbar_expect_tx(&s_bar, span);
            bulk_load(win, a.bytes + base, span, &s_bar);
        }

        // ---------------- cut
        if (fast_tile) {
            // meanwhile: record headers (start, partition | flags), chunk tables cleared
            for (uint32_t k = tid; k <= m; k += FZ_THREADS) {
                const uint2 h = __ldg(reinterpret_cast<const uint2*>(a.recs + r0 + k));
                rst[k] = (uint16_t)(h.x - base);
                if (k < m) rpart[k] = h.y;
            }
            for (uint32_t c = tid; c < nch; c += FZ_THREADS) rcnt[c] = 0;
            __syncthreads();
            // records starting in each chunk (several may: short or empty records)
            for (uint32_t k = tid; k < m; k += FZ_THREADS) {
                const uint32_t c = rst[k] / FZ_CHUNK;
                uint32_t* word = reinterpret_cast<uint32_t*>(rcnt) + (c >> 1);
                atomicAdd(word, 1u << (16u * (c & 1u)));
            }
            if (span) mbar_wait(&s_bar, phase);
            phase ^= span ? 1u : 0u;
            // comma masks of the chunks that hold the tile's bytes
            const uint32_t lo_rel = lo - base, hi_rel = hi - base;
            const uint32_t used = (hi_rel + FZ_CHUNK - 1) / FZ_CHUNK;
            for (uint32_t c = tid; c < nch; c += FZ_THREADS) {
                uint32_t mk = 0;
                if (c < used) {
                    mk = fz_comma_mask32(win + c * FZ_CHUNK);
                    const uint32_t cb = c * FZ_CHUNK;
                    if (cb < lo_rel || cb + FZ_CHUNK > hi_rel) mk &= fz_range32((int)lo_rel - (int)cb, (int)hi_rel - (int)cb);
                }
                cmask[c] = mk;
            }
            __syncthreads();
            // exclusive prefix per chunk, thread t owns chunks [t * per, (t + 1) * per)
            {
                const uint32_t per = (nch + FZ_THREADS - 1) / FZ_THREADS;
                const uint32_t c0 = min(tid * per, nch), c1 = min(c0 + per, nch);
                uint32_t sum = 0;
                for (uint32_t c = c0; c < c1; ++c) sum += __popc(cmask[c]) + ((uint32_t)rcnt[c] << 16);
                uint32_t run = fz_block_scan(sum, s_warp_tot, &s_total);
                for (uint32_t c = c0; c < c1; ++c) {
                    pre[c] = run;
                    run += __popc(cmask[c]) + ((uint32_t)rcnt[c] << 16);
                }
                if (tid == FZ_THREADS - 1) pre[nch] = run;
            }
            __syncthreads();
            // per record: element count, head check, keep flag, first segment
            uint32_t my_kept = 0, my_elems = 0;
            for (uint32_t k = tid; k < m; k += FZ_THREADS) {
                const uint32_t i = r0 + k;
                const uint32_t ra = rst[k], rb = rst[k + 1];
                const int32_t pflags = (int32_t)rpart[k] & TK_REC_FLAGS;
                const uint32_t ca = (pre[ra / FZ_CHUNK] & 0xFFFFu) + __popc(cmask[ra / FZ_CHUNK] & ((1u << (ra & 31u)) - 1u));
                uint32_t len = 0, first = ra, info = 0;
                int status;
                if (pflags) {
                    // flagged by the record-batch split: skipped records were never consumed, broken
                    // ones carry their batch's error; neither has a payload to look at
                    status = (pflags & TK_REC_OVERFLOW) == TK_REC_OVERFLOW ? TK_STATUS_SLOT_OVERFLOW
                           : (pflags & TK_REC_CORRUPT) ? TK_STATUS_CORRUPT
                           : (pflags & TK_REC_UNSUPPORTED) ? TK_STATUS_UNSUPPORTED : TK_STATUS_OK;
                } else {
                    const uint32_t cb = (pre[rb / FZ_CHUNK] & 0xFFFFu) + __popc(cmask[rb / FZ_CHUNK] & ((1u << (rb & 31u)) - 1u));
                    bool empty = false;
                    int head;
                    if (rb > ra + 1 && win[ra] == '[' && tk_is_numclass((int)win[ra + 1])) {
                        head = TK_OK;
                        first = ra + 1;
                    } else {
                        WindowReader rd{win, rb};
                        head = rb > ra ? tk_json_prefix(rd, ra, &first, &empty) : TK_ERR_SYNTAX;
                        if (head) first = ra;
                    }
                    len = (!head && empty) ? 0u : cb - ca + 1u;
                    status = head ? head : (len > a.max_len ? TK_ERR_TOO_LONG : TK_OK);
                    if (status == TK_OK && len >= a.min_size) info |= FR_KEEP;
                    if (status == TK_OK && len) info |= FR_PARSE;
                    if (rb > ra && win[rb - 1] == ']') info |= FR_TAIL;
                }
                a.rec_len[i] = (int32_t)len;
                a.rec_status[i] = status;
                if (status) report_error(a.res, i, status);
                else if ((info & FR_KEEP) && a.fixed_len && len != a.fixed_len) {
                    a.rec_status[i] = TK_ERR_RAGGED;
                    report_error(a.res, i, TK_ERR_RAGGED);
                }
                rfirst[k] = (uint16_t)first;
                rlen[k] = (uint16_t)len;
                rinfo[k] = (uint8_t)info;
                fseg[k] = (uint16_t)(ca + k);
                my_kept += info & FR_KEEP;
                my_elems += len;
            }
            if (tid == 0) fseg[m] = (uint16_t)((pre[nch] & 0xFFFFu) + m);
            my_kept = __reduce_add_sync(0xffffffffu, my_kept);
            my_elems = __reduce_add_sync(0xffffffffu, my_elems);
            if (lane == 0 && (my_kept | my_elems)) { atomicAdd(&s_kept, my_kept); atomicAdd(&s_elems, my_elems); }
        } else if (m) {
            // scalar path: thread per record out of global memory
            uint32_t my_kept = 0, my_elems = 0;
            for (uint32_t k = tid; k < m; k += FZ_THREADS) {
                const uint32_t i = r0 + k;
                const uint32_t start = a.recs[i].start, end = a.recs[i + 1].start;
                const int32_t pflags = a.recs[i].partition & TK_REC_FLAGS;
                uint32_t len = 0;
                int status;
                bool keep = false;
                if (pflags) {
                    status = (pflags & TK_REC_OVERFLOW) == TK_REC_OVERFLOW ? TK_STATUS_SLOT_OVERFLOW
                           : (pflags & TK_REC_CORRUPT) ? TK_STATUS_CORRUPT
                           : (pflags & TK_REC_UNSUPPORTED) ? TK_STATUS_UNSUPPORTED : TK_STATUS_OK;
                } else {
                    uint32_t commas = 0;
                    for (uint32_t p = start; p < end; ++p) commas += a.bytes[p] == ',';
                    GlobalReader rd{a.bytes, end};
                    uint32_t first = start;
                    bool empty = false;
                    const int head = end > start ? tk_json_prefix(rd, start, &first, &empty) : TK_ERR_SYNTAX;
                    len = (!head && empty) ? 0u : commas + 1u;
                    status = head ? head : (len > a.max_len ? TK_ERR_TOO_LONG : TK_OK);
                    keep = status == TK_OK && len >= a.min_size;
                }
                a.rec_len[i] = (int32_t)len;
                a.rec_status[i] = status;
                if (status) report_error(a.res, i, status);
                else if (keep && a.fixed_len && len != a.fixed_len) {
                    a.rec_status[i] = TK_ERR_RAGGED;
                    report_error(a.res, i, TK_ERR_RAGGED);
                }
                a.rec_kidx[i] = keep ? 0 : -1;                    // provisional: the keep flag
                my_kept += keep ? 1u : 0u;
                my_elems += len;
            }
            my_kept = __reduce_add_sync(0xffffffffu, my_kept);
            my_elems = __reduce_add_sync(0xffffffffu, my_elems);
            if (lane == 0 && (my_kept | my_elems)) { atomicAdd(&s_kept, my_kept); atomicAdd(&s_elems, my_elems); }
        }
        __syncthreads();

        // ---------------- publish the tile's sum (the tiles behind this one wait for it), then do
        // everything that does not need the tile's place among the kept records: scatter the first
        // round of segments and parse them into the staging array.  By the time the block looks
        // back, the tiles in front of it have long published theirs.
        const uint32_t n_seg = fast_tile ? (uint32_t)fseg[m] : 0u;
        const unsigned long long tot = ((unsigned long long)s_kept << FZ_PACK_SHIFT) + s_elems;
        if (tid == 0) {
            a.tile_agg[tile] = tot;
            if (tile == 0) a.tile_incl[0] = tot;
            fz_st_release(a.tile_flag + tile, (a.epoch << 2) | (tile == 0 ? 2u : 1u));
        }
        const uint32_t cnt0 = min(FZ_TOKCAP, n_seg);
        if (fast_tile) {
            fz_scatter(0, n_seg, m, nch, tid, FZ_THREADS, cmask, pre, rcnt, rst, fseg, rfirst, tok);
            __syncthreads();
            fz_parse_round<KIND, true>(a, tb, win, tok, rinfo, fseg, rlen, rst, rrow, val, values, 0, cnt0, r0, tid);
        }
        const unsigned long long excl = tile ? fz_lookback(a, tile, s_first, s_sum) : 0ull;
        if (tid == 0 && tile) {
            a.tile_incl[tile] = excl + tot;
            fz_st_release(a.tile_flag + tile, (a.epoch << 2) | 2u);
        }
        __syncthreads();
        const uint32_t excl_kept = (uint32_t)(excl >> FZ_PACK_SHIFT);
        // rows of the kept records, batch boundaries, longest record per batch, commit maxima:
        // thread per record in record order (uniform trip count: the loop body uses warp collectives)
        uint32_t running = 0;
        for (uint32_t k0 = 0; k0 < m; k0 += FZ_THREADS) {
            const uint32_t k = k0 + tid;
            const bool live = k < m;
            const uint32_t i = r0 + k;
            uint32_t keep = 0, len = 0;
            if (live) {
                if (fast_tile) { keep = rinfo[k] & FR_KEEP; len = rlen[k]; }
                else { keep = a.rec_kidx[i] >= 0 ? 1u : 0u; len = (uint32_t)a.rec_len[i]; }
            }
            const uint32_t ex = excl_kept + running + fz_block_scan(keep, s_warp_tot, &s_total);
            running += s_total;
            const uint32_t j = (carry + ex) / B;               // batch of this record (kept or skipped)
            uint32_t klen = 0;
            if (live) {
                if (keep) {
                    a.rec_kidx[i] = (int32_t)ex;
                    klen = len;
                    if ((carry + ex + 1) % B == 0) a.batch_end[j] = i;
                } else {
                    a.rec_kidx[i] = -1 - (int32_t)ex;
                }
                if (fast_tile) rrow[k] = keep ? B + ex : 0xFFFFFFFFu;
            }
            // longest kept record of the batch: one atomic per warp when the warp lies inside one batch
            const uint32_t j0 = __shfl_sync(0xffffffffu, j, 0);
            if (__all_sync(0xffffffffu, j == j0)) {
                const uint32_t mx = __reduce_max_sync(0xffffffffu, klen);
                if (lane == 0 && mx) atomicMax(&a.batch_lmax[j0], mx);
            } else if (klen) {
                atomicMax(&a.batch_lmax[j], klen);
            }
            // commit: lanes that share a (batch, partition) key elect one lane for the group's atomic
            uint32_t key = 0xFFFFFFFFu, p = 0;
            long long val = 0;
            if (live) {
                const uint4 raw = __ldg(reinterpret_cast<const uint4*>(a.recs) + i);
                p = raw.y;
                val = (long long)(((uint64_t)raw.w << 32) | raw.z) + 1;
                if (p < a.n_partitions) key = j * a.n_partitions + p;
            }
            const unsigned peers = __match_any_sync(0xffffffffu, key);
            const uint32_t vhi = (uint32_t)((unsigned long long)val >> 32), vlo = (uint32_t)val;
            const uint32_t mh = __reduce_max_sync(peers, vhi);
            const uint32_t ml = __reduce_max_sync(peers, vhi == mh ? vlo : 0u);
            if (key != 0xFFFFFFFFu && (int)lane == __ffs(peers) - 1)
                atomicMax(reinterpret_cast<long long*>(&a.commit[(size_t)p * a.pitch + j]),
                          (long long)(((unsigned long long)mh << 32) | ml));
        }
        __syncthreads();

        if (fast_tile) {
            // ---------------- rows: pad, length, mask of every kept record (warp per record)
            for (uint32_t k = wid; k < m; k += FZ_WARPS) {
                const uint32_t row = rrow[k];
                if (row != 0xFFFFFFFFu)
                    json_write_row<sizeof(T)>(a.values, a.lengths, a.mask, a.max_len, row, rlen[k], lane, vec_values, vec_mask);
            }
            // ---------------- the staged values of the first round go to their rows
            for (uint32_t s = tid; s < cnt0; s += FZ_THREADS) {
                const uint32_t k = tok[s] >> 16;
                if (!(rinfo[k] & FR_PARSE)) continue;
                const uint32_t row = rrow[k], e = s - fseg[k];
                if (row != 0xFFFFFFFFu && e < a.max_len) values[(uint64_t)row * a.max_len + e] = val[s];
            }
            // ---------------- further rounds (records with very many short elements): rows are known
            for (uint32_t sbase = FZ_TOKCAP; sbase < n_seg; sbase += FZ_TOKCAP) {
                __syncthreads();                                        // tok is reused
                fz_scatter(sbase, n_seg, m, nch, tid, FZ_THREADS, cmask, pre, rcnt, rst, fseg, rfirst, tok);
                __syncthreads();
                fz_parse_round<KIND, false>(a, tb, win, tok, rinfo, fseg, rlen, rst, rrow, val, values, sbase,
                                            min(FZ_TOKCAP, n_seg - sbase), r0, tid);
            }
        } else if (m) {
            // scalar path: warp per record for the row, thread per record for the elements
            for (uint32_t k = wid; k < m; k += FZ_WARPS) {
                const int32_t kidx = a.rec_kidx[r0 + k];
                if (kidx >= 0)
                    json_write_row<sizeof(T)>(a.values, a.lengths, a.mask, a.max_len, (uint64_t)B + (uint32_t)kidx,
                                              (uint32_t)a.rec_len[r0 + k], lane, vec_values, vec_mask);
            }
            __syncthreads();
            for (uint32_t k = tid; k < m; k += FZ_THREADS) {
                const uint32_t i = r0 + k;
                const uint32_t len = (uint32_t)a.rec_len[i];
                if (a.rec_status[i] != TK_STATUS_OK || !len) continue;
                const uint32_t start = a.recs[i].start, end = a.recs[i + 1].start;
                const int32_t kidx = a.rec_kidx[i];
                GlobalReader rd{a.bytes, end};
                uint32_t pos = start;
                bool empty = false;
                tk_json_prefix(rd, start, &pos, &empty);          // accepted above: pos = first element
                for (uint32_t e = 0; e < len; ++e) {
                    T v;
                    const int st = tk_json_element_flat<KIND>(rd, pos, &v);
                    if (st) { flag_record(a.rec_status, a.res, i, st); break; }
                    if (kidx >= 0 && e < a.max_len) values[(uint64_t)(B + (uint32_t)kidx) * a.max_len + e] = v;
                    while (pos < end && a.bytes[pos] != ',') ++pos;   // the segment behind the next comma
                    ++pos;
                }
            }
        }
        __syncthreads();                    // the window, the tables and s_next are free for the next tile
    }

    // Every block leaves through this counter; all tiles are done when the last one does, and it
    // writes the slot's result header and resets the counters for the slot's next pass.
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        const uint32_t left = atomicAdd(&a.ctr[2], 1u);
        if (left == gridDim.x - 1) {
            __threadfence();
            const unsigned long long total = *(volatile unsigned long long*)&a.tile_incl[a.n_tiles - 1];
            const uint32_t n_kept = (uint32_t)(total >> FZ_PACK_SHIFT);
            const unsigned long long n_tokens = total & ((1ull << FZ_PACK_SHIFT) - 1);
            tk_result_t* res = a.res;
            res->n_records = a.n;
            res->n_kept = n_kept;
            res->carry_in = carry;
            res->carry_out = (carry + n_kept) % B;
            res->n_closed = (carry + n_kept) / B;
            res->batch_size = B;
            res->reserved[0] = (uint32_t)(n_tokens > 0xFFFFFFFFull ? 0xFFFFFFFFu : n_tokens);
            a.state->carry_rows = (carry + n_kept) % B;
            a.ctr[0] = 0;
            a.ctr[2] = 0;
        }
    }
}

// ---------------------------------------------------------------- host side
// Per-slot scratch of the fused path (first use): tile table and look-back arrays sized by the
// slot's bytes, not by its record capacity.
int tk_ensure_fused(tk_ctx* ctx) {
    if (ctx->fz_tiles) return 0;
    const size_t tiles = (size_t)(ctx->slot_bytes / FZ_TILE_BYTES) + 2;
    for (int i = 0; i < ctx->n_slots; ++i) {
        tk_slot& s = ctx->slots[i];
        TK_CUDA(cudaMalloc(&s.d.fz_tile_first, sizeof(uint2) * (tiles + 1)));
        TK_CUDA(cudaMalloc(&s.d.fz_agg, sizeof(unsigned long long) * tiles));
        TK_CUDA(cudaMalloc(&s.d.fz_incl, sizeof(unsigned long long) * tiles));
        TK_CUDA(cudaMalloc(&s.d.fz_flag, sizeof(uint32_t) * tiles));
        TK_CUDA(cudaMemset(s.d.fz_flag, 0, sizeof(uint32_t) * tiles));
        TK_CUDA(cudaMalloc(&s.d.fz_ctr, 4 * sizeof(uint32_t)));
        TK_CUDA(cudaMemset(s.d.fz_ctr, 0, 4 * sizeof(uint32_t)));
        s.fz_epoch = 0;
    }
    ctx->fz_tiles = tiles;
    return 0;
}

template <int KIND>
static int fz_launch(const fused_args& a, size_t smem, cudaStream_t st) {
    static int configured_smem = 0, per_sm = 0, sms = 0;
    if ((int)smem > configured_smem) {
        TK_CUDA(cudaFuncSetAttribute(k_json_fused<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured_smem = (int)smem;
        per_sm = 0;
    }
    if (!per_sm) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess ||
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_json_fused<KIND>, FZ_THREADS, smem) != cudaSuccess ||
            per_sm < 1 || sms < 1) {
            cudaGetLastError();
            per_sm = 1; sms = 148;
        }
    }
    // persistent blocks: one wave, never more blocks than tiles
    const uint32_t blocks = (uint32_t)min((uint64_t)per_sm * sms, (uint64_t)max(a.n_tiles, 1u));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(blocks);
    cfg.blockDim = dim3(FZ_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    TK_CUDA(cudaLaunchKernelEx(&cfg, k_json_fused<KIND>, a));
    return 0;
}

extern "C" int tk_json_ingest(tk_ctx* ctx, int slot, int dtype, uint32_t max_len, uint32_t min_size,
                              uint32_t fixed_len, uint32_t batch_size, const tk_plane_t* values,
                              const tk_plane_t* lengths, const tk_plane_t* mask, void* stream) {
    TK_ENTER(ctx, slot);
    tk_slot& s = ctx->slots[slot];
    cudaStream_t st = (cudaStream_t)stream;
    TK_REQUIRE(values && values->base, "values plane is required");
    TK_REQUIRE(max_len > 0 && batch_size > 0, "max_len and batch_size must be positive");
    TK_REQUIRE(dtype >= TK_DTYPE_F32 && dtype <= TK_DTYPE_I64, "unknown dtype %d", dtype);
    const uint32_t n = s.n_records;
    TK_REQUIRE(((uint64_t)n + batch_size - 1) / batch_size <= ctx->max_batches,
               "slot with %u records can close more than max_batches=%u batches of %u", n, ctx->max_batches, batch_size);
    int rc = tk_ensure_fused(ctx);
    if (rc) return rc;
    ctx->batch_size = batch_size;
    const uint32_t n_bytes = (uint32_t)s.n_bytes;
    const uint32_t n_tiles = max((n_bytes + FZ_TILE_BYTES - 1) / FZ_TILE_BYTES, 1u);
    TK_REQUIRE((size_t)n_tiles + 1 <= ctx->fz_tiles, "slot bytes exceed the context geometry");

    prep_args p;
    p.recs = s.d.recs; p.n = n; p.n_tiles = n_tiles; p.tile_info = reinterpret_cast<uint2*>(s.d.fz_tile_first); p.res = s.d.result;
    p.reset = tk_plan_reset_args(ctx, slot);
    const uint32_t pblocks = min(max((n + 256u) / 256u, 1u), 148u * 8u);
    TK_LAUNCH_PDL(k_fused_prep, pblocks, 256, st, p);

    // The window holds a tile (records starting within TILE_BYTES) plus the tail of its last record:
    // room for the longest record the schema allows as json.dumps writes it (~26 bytes per float,
    // ~8 per small integer), within what two blocks per SM can have.
    const uint32_t per_elem = (dtype == TK_DTYPE_F32 || dtype == TK_DTYPE_F64) ? 26u : 12u;
    uint64_t over = (uint64_t)max_len * per_elem + 64;
    over = (over + 4095) & ~4095ull;
    if (over < 8192) over = 8192;
    if (over > 40960) over = 40960;
    fused_args a;
    a.bytes = s.d.bytes; a.recs = s.d.recs; a.n = n; a.n_bytes = n_bytes;
    a.max_len = max_len; a.min_size = min_size; a.fixed_len = fixed_len; a.batch_size = batch_size;
    a.rec_len = s.d.rec_len; a.rec_status = s.d.rec_status; a.rec_kidx = s.d.rec_kidx;
    a.tile_info = reinterpret_cast<const uint2*>(s.d.fz_tile_first); a.n_tiles = n_tiles;
    a.win_bytes = FZ_TILE_BYTES + (uint32_t)over;                 // <= 57344: positions fit 16 bits
    a.max_recs = 512u;
    a.tile_agg = s.d.fz_agg; a.tile_incl = s.d.fz_incl; a.tile_flag = s.d.fz_flag; a.ctr = s.d.fz_ctr;
    s.fz_epoch = (s.fz_epoch + 1) & 0x3FFFFFFFu;
    if (!s.fz_epoch) s.fz_epoch = 1;                              // 0 is the state of a fresh allocation
    a.epoch = s.fz_epoch;
    a.state = ctx->d_state; a.res = s.d.result; a.batch_end = s.d.batch_end; a.batch_lmax = s.d.batch_lmax;
    a.commit = s.d.commit; a.pitch = ctx->commit_pitch; a.n_partitions = ctx->n_partitions;
    tk_plane_t z; z.base = nullptr; z.row_bytes = 0;
    a.values = *values;
    a.lengths = lengths ? *lengths : z;
    a.mask = mask ? *mask : z;
    const size_t smem = fz_smem_bytes(a.win_bytes, a.max_recs, (dtype == TK_DTYPE_F64 || dtype == TK_DTYPE_I64) ? 8 : 4);
    switch (dtype) {
        case TK_DTYPE_F32: return fz_launch<TK_F32>(a, smem, st);
        case TK_DTYPE_F64: return fz_launch<TK_F64>(a, smem, st);
        case TK_DTYPE_I32: return fz_launch<TK_I32>(a, smem, st);
        default:           return fz_launch<TK_I64>(a, smem, st);
    }
}
