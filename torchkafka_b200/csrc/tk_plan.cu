// tk_plan.cu -- K2 (filter compaction, batch boundaries) and K4 (commit offsets).
//
// K2 replaces the reference's sequential "drop None, count to batch_size" loop
// (src/kafka_dataset.py:161-162 + torch/utils/data/_utils/fetch.py:33-38): a
// batch closes on its batch_size-th *kept* record; skipped records before the
// closing record belong to that batch, skipped records after it to the next
// (SURVEY.md §4.1-3).  On the GPU that is an exclusive prefix sum of the keep
// flags: record i with k kept records before it lies in batch (carry + k) / B.
//
// K4 replaces the consumer-position bookkeeping behind consumer.commit()
// (src/kafka_dataset.py:130): after batch j, partition p commits
// max(offset) + 1 over every record consumed up to the closing record of j
// (SURVEY.md §4.1-5).  Segmented max by (batch, partition), then a running max
// along the batch axis seeded with the positions of earlier slots.
#include "tk_internal.h"

#define CMP_THREADS 128
#define CMP_ITEMS 2
#define CMP_TILE (CMP_THREADS * CMP_ITEMS)

// The compaction scans two quantities at once, packed as
// (kept records << 40) | elements: the second sum is the number of elements in the slot, which
// sizes the thread-per-element parse kernel of the flat decode path.
#define PACK_SHIFT 40
__device__ __forceinline__ unsigned long long warp_incl_scan64(unsigned long long v) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        unsigned long long t = __shfl_up_sync(0xffffffffu, v, d);
        if ((threadIdx.x & 31) >= d) v += t;
    }
    return v;
}
__device__ __forceinline__ unsigned long long block_excl_scan64(unsigned long long v, unsigned long long* total) {
    __shared__ unsigned long long warp_tot[32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned long long incl = warp_incl_scan64(v);
    if (lane == 31) warp_tot[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        unsigned long long w = lane < (int)((blockDim.x + 31) >> 5) ? warp_tot[lane] : 0;
        unsigned long long wi = warp_incl_scan64(w);
        warp_tot[lane] = wi - w;
        if (lane == 31) *total = wi;
    }
    __syncthreads();
    unsigned long long r = warp_tot[wid] + incl - v;
    __syncthreads();
    return r;
}

// ------------------------------------------------------------------------ K2 + K4, one kernel
// Single-pass scan with decoupled look-back: every block takes the next tile of CMP_TILE
// records (ticket counter, so a tile's predecessors are always running or done), publishes the
// tile's packed sum, looks back over the predecessors' published sums for its exclusive prefix,
// publishes its inclusive sum, and applies:
//   kept record    -> rec_kidx = k            (k = kept records before it in this slot)
//   skipped record -> rec_kidx = -1 - k
//   per batch: longest kept record, index of the closing record
//   K4 first half: segmented max of (offset + 1) per (batch, partition) into the commit matrix.
// The block that finishes last writes the result header and advances the carry.  K4's second
// half (running max along the batch axis) is k_commit_final below.
// The per-batch arrays and the commit matrix were reset by the scan kernel of the slot
// (tk_plan_reset, tk_internal.h).
struct plan_args {
    int32_t*        rec_kidx;
    const int32_t*  rec_len;
    const tk_rec_t* recs;
    uint32_t        n, n_tiles;
    uint32_t        batch_size, max_batches;
    uint32_t        pitch, n_partitions;
    uint32_t        epoch;
    tk_state*       state;
    tk_result_t*    res;
    uint32_t*       batch_end;
    uint32_t*       batch_lmax;
    int64_t*        commit;
    int64_t*        positions;
    unsigned long long* tile_agg;
    unsigned long long* tile_incl;
    uint32_t*       tile_flag;
    uint32_t*       ctr;          // [0] tile ticket, [1] finished tiles
    // flat JSON slots (row_elems != 0): where the record's elements go (tk_internal.h: rec_base)
    const uint32_t* rec_tok;
    long long*      rec_base;
    uint32_t        row_elems;
};

__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long warp_sum64(unsigned long long v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}

// exclusive prefix of tile `tile` (warp 0 of the block; result valid in every lane)
__device__ __forceinline__ unsigned long long plan_lookback(const plan_args& a, uint32_t tile, uint32_t lane) {
    unsigned long long excl = 0;
    int p = (int)tile - 1;
    while (p >= 0) {
        const int q = p - (int)lane;                      // lane 0 looks at the nearest predecessor
        uint32_t st = 2;                                  // tiles before 0: "inclusive sum 0"
        if (q >= 0) {
            uint32_t f;
            do { f = ld_acquire_u32(a.tile_flag + q); } while ((f >> 2) != a.epoch);
            st = f & 3u;
        }
        const uint32_t full = __ballot_sync(0xffffffffu, st == 2);
        const uint32_t first = full ? (uint32_t)__ffs(full) - 1 : 32u;      // nearest inclusive sum
        unsigned long long v = 0;
        if (q >= 0 && lane <= first) v = lane == first ? a.tile_incl[q] : a.tile_agg[q];
        excl += warp_sum64(v);
        if (full) break;
        p -= 32;
    }
    return excl;
}

__global__ void __launch_bounds__(CMP_THREADS) k_plan(const plan_args a) {
    __shared__ unsigned long long tot, s_excl;
    __shared__ uint32_t s_tile, s_last;
    tk_pdl_launch_dependents();
    tk_pdl_wait();
    if (threadIdx.x == 0) s_tile = atomicAdd(&a.ctr[0], 1u);
    __syncthreads();
    const uint32_t tile = s_tile, lane = threadIdx.x & 31;
    const uint32_t base = tile * CMP_TILE + threadIdx.x * CMP_ITEMS;
    const uint32_t n = a.n, B = a.batch_size;
    int keep[CMP_ITEMS];
    uint32_t lens[CMP_ITEMS];
    unsigned long long c = 0;
#pragma unroll
    for (int k = 0; k < CMP_ITEMS; ++k) {
        const uint32_t i = base + k;
        keep[k] = (i < n && a.rec_kidx[i] >= 0) ? 1 : 0;
        lens[k] = i < n ? (uint32_t)a.rec_len[i] : 0u;
        c += ((unsigned long long)keep[k] << PACK_SHIFT) + lens[k];
    }
    const unsigned long long in_tile = block_excl_scan64(c, &tot);
    if (threadIdx.x < 32) {
        if (lane == 0) {
            a.tile_agg[tile] = tot;
            if (tile == 0) a.tile_incl[0] = tot;
            st_release_u32(a.tile_flag + tile, (a.epoch << 2) | (tile == 0 ? 2u : 1u));
        }
        const unsigned long long excl = tile ? plan_lookback(a, tile, lane) : 0ull;
        if (lane == 0) {
            if (tile) {
                a.tile_incl[tile] = excl + tot;
                st_release_u32(a.tile_flag + tile, (a.epoch << 2) | 2u);
            }
            s_excl = excl;
        }
    }
    __syncthreads();
    uint32_t ex = (uint32_t)((in_tile + s_excl) >> PACK_SHIFT);
    const uint32_t carry = a.state->carry_rows;
#pragma unroll
    for (int k = 0; k < CMP_ITEMS; ++k) {
        const uint32_t i = base + k;
        const bool live = i < n;
        const uint32_t j = (carry + ex) / B;               // batch of this record (kept or skipped)
        uint32_t len = 0;
        if (live) {
            if (keep[k]) {
                a.rec_kidx[i] = (int32_t)ex;
                len = lens[k];
                if ((carry + ex + 1) % B == 0) a.batch_end[j] = i;
                // (the first B rows of the output planes hold the batch left open by the previous slot)
                if (a.row_elems) a.rec_base[i] = (long long)((unsigned long long)(B + ex) * a.row_elems) - (long long)a.rec_tok[i];
            } else {
                a.rec_kidx[i] = -1 - (int32_t)ex;
                if (a.row_elems) a.rec_base[i] = TK_REC_BASE_SKIP;
            }
        }
        // longest kept record of the batch: one atomic per warp when the warp lies inside one batch
        const uint32_t j0 = __shfl_sync(0xffffffffu, j, 0);
        if (__all_sync(0xffffffffu, j == j0)) {
            const uint32_t m = __reduce_max_sync(0xffffffffu, len);
            if (lane == 0 && m) atomicMax(&a.batch_lmax[j0], m);
        } else if (len) {
            atomicMax(&a.batch_lmax[j], len);
        }
        // K4, first half.  Lanes of the warp that share a (batch, partition) key elect one lane
        // that issues the group's only atomic.
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
        if (live && keep[k]) ++ex;
    }
    // ---- the last block to get here finishes the plan
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(&a.ctr[1], 1u) == a.n_tiles - 1 ? 1u : 0u;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    const unsigned long long total = *(volatile unsigned long long*)&a.tile_incl[a.n_tiles - 1];
    const uint32_t n_kept = (uint32_t)(total >> PACK_SHIFT);
    const unsigned long long n_tokens = total & ((1ull << PACK_SHIFT) - 1);
    const uint32_t n_closed = (carry + n_kept) / B;
    if (threadIdx.x == 0) {
        tk_result_t* res = a.res;
        res->n_records = n;
        res->n_kept = n_kept;
        res->carry_in = carry;
        res->carry_out = (carry + n_kept) % B;
        res->n_closed = n_closed;
        res->batch_size = B;
        res->n_errors = 0;
        res->first_error_rec = 0xFFFFFFFFu;
        res->first_error_status = 0;
        res->reserved[0] = (uint32_t)(n_tokens > 0xFFFFFFFFull ? 0xFFFFFFFFu : n_tokens);   // elements in the slot
        a.state->carry_rows = (carry + n_kept) % B;
        a.ctr[0] = 0;
        a.ctr[1] = 0;
    }
}

extern "C" int tk_compact(tk_ctx* ctx, int slot, uint32_t batch_size, void* stream) {
    TK_ENTER(ctx, slot);
    tk_slot& s = ctx->slots[slot];
    cudaStream_t st = (cudaStream_t)stream;
    TK_REQUIRE(batch_size > 0, "batch_size must be positive");
    const uint32_t n = s.n_records;
    // worst case: every record kept and the open batch almost full
    TK_REQUIRE(((uint64_t)n + batch_size - 1) / batch_size <= ctx->max_batches,
               "slot with %u records can close more than max_batches=%u batches of %u", n, ctx->max_batches, batch_size);
    ctx->batch_size = batch_size;
    plan_args a;
    a.rec_kidx = s.d.rec_kidx; a.rec_len = s.d.rec_len; a.recs = s.d.recs;
    a.n = n; a.n_tiles = n ? (n + CMP_TILE - 1) / CMP_TILE : 1;
    a.batch_size = batch_size; a.max_batches = ctx->max_batches;
    a.pitch = ctx->commit_pitch; a.n_partitions = ctx->n_partitions;
    s.plan_epoch = (s.plan_epoch + 1) & 0x3FFFFFFFu;
    if (!s.plan_epoch) s.plan_epoch = 1;                   // 0 is the state of a fresh allocation
    a.epoch = s.plan_epoch;
    a.state = ctx->d_state; a.res = s.d.result; a.batch_end = s.d.batch_end; a.batch_lmax = s.d.batch_lmax;
    a.commit = s.d.commit; a.positions = ctx->d_positions;
    a.tile_agg = s.d.tile_agg; a.tile_incl = s.d.tile_incl; a.tile_flag = s.d.tile_flag; a.ctr = s.d.plan_ctr;
    a.rec_tok = s.d.rec_tok; a.rec_base = s.d.rec_base; a.row_elems = s.json_row_elems;
    TK_LAUNCH_PDL(k_plan, a.n_tiles, CMP_THREADS, st, a);
    return 0;
}

// ------------------------------------------------------------------------ K4, second half
// Running max along the batch axis of every partition, seeded with the positions earlier slots
// left behind.  One block; the (partition, batch) matrix is staged through shared memory in
// chunks: one round of independent loads, then thread p runs the max along the batch axis of
// partition p out of shared memory, then one round of stores -- three memory latencies per
// chunk instead of one per partition.
//
// Nothing on the device consumes the result (the host reads it with the slot's result block), so
// tk_process_slot launches this kernel *behind* the decode kernel with wait_for_prior = 0: the
// decode kernel releases its dependents only after its own griddepcontrol.wait, i.e. once the
// plan is complete and visible, and this block then runs next to the decode kernel's blocks
// instead of in front of them.
#define PLAN_STAGE 4096u      // int64 entries of the staging buffer (32 KiB)
__global__ void __launch_bounds__(256) k_commit_final(int64_t* __restrict__ commit, uint32_t pitch,
                                                       uint32_t n_partitions, const tk_result_t* __restrict__ res,
                                                       int64_t* __restrict__ positions, int wait_for_prior) {
    __shared__ long long s_stage[PLAN_STAGE];
    tk_pdl_launch_dependents();
    if (wait_for_prior) tk_pdl_wait();
    const uint32_t nb = __ldcg(&res->n_closed) + 1;                  // closed batches + the open one
    for (uint32_t p0 = 0; p0 < n_partitions; p0 += blockDim.x) {
        const uint32_t pg = min(blockDim.x, n_partitions - p0);
        const uint32_t J = max(1u, PLAN_STAGE / pg - 1u);   // batches per chunk (row stride J | 1 <= J + 1)
        long long run = threadIdx.x < pg ? positions[p0 + threadIdx.x] : -1;
        for (uint32_t j0 = 0; j0 < nb; j0 += J) {
            const uint32_t jn = min(J, nb - j0), stride = jn | 1u;
            // (ld.cg: the maxima were produced by atomics of other SMs and live in L2.)  Eight loads
            // per thread are issued before the first is used.
            const uint32_t total = pg * jn;
            for (uint32_t x0 = threadIdx.x; x0 < total; x0 += blockDim.x * 8) {
                long long v[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const uint32_t x = x0 + k * blockDim.x;
                    const uint32_t pp = x / jn, jj = x - pp * jn;
                    v[k] = x < total ? __ldcg(reinterpret_cast<const long long*>(&commit[(size_t)(p0 + pp) * pitch + j0 + jj])) : 0;
                }
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const uint32_t x = x0 + k * blockDim.x;
                    const uint32_t pp = x / jn, jj = x - pp * jn;
                    if (x < total) s_stage[pp * stride + jj] = v[k];
                }
            }
            __syncthreads();
            if (threadIdx.x < pg) {
                long long* row = s_stage + threadIdx.x * stride;
                for (uint32_t jj = 0; jj < jn; ++jj) {
                    const long long v = row[jj];
                    run = v > run ? v : run;
                    row[jj] = run;
                }
            }
            __syncthreads();
            for (uint32_t x = threadIdx.x; x < pg * jn; x += blockDim.x) {
                const uint32_t pp = x / jn, jj = x - pp * jn;
                commit[(size_t)(p0 + pp) * pitch + j0 + jj] = s_stage[pp * stride + jj];
            }
            __syncthreads();
        }
        if (threadIdx.x < pg) positions[p0 + threadIdx.x] = run;
    }
}

// Many batches per slot (small batch sizes): one block per partition, every thread a contiguous
// chunk of the batch axis, two passes (chunk maxima, then the running max).
__global__ void __launch_bounds__(128) k_commit_rows(int64_t* __restrict__ commit, uint32_t pitch,
                                                      const tk_result_t* __restrict__ res,
                                                      int64_t* __restrict__ positions, int wait_for_prior) {
    __shared__ long long chunk_max[128];
    tk_pdl_launch_dependents();
    if (wait_for_prior) tk_pdl_wait();
    const uint32_t p = blockIdx.x;
    const uint32_t nb = __ldcg(&res->n_closed) + 1;
    long long* row = reinterpret_cast<long long*>(commit) + (size_t)p * pitch;
    const uint32_t per = (nb + blockDim.x - 1) / blockDim.x;
    const uint32_t a = threadIdx.x * per;
    const uint32_t b = min(a + per, nb);
    long long m = -1;
    for (uint32_t j = a; j < b; ++j) { const long long v = __ldcg(row + j); m = v > m ? v : m; }
    chunk_max[threadIdx.x] = m;
    __syncthreads();
    long long run = positions[p];
    for (uint32_t t = 0; t < threadIdx.x; ++t) { const long long v = chunk_max[t]; run = v > run ? v : run; }
    for (uint32_t j = a; j < b; ++j) {
        const long long v = __ldcg(row + j);
        run = v > run ? v : run;
        row[j] = run;
    }
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) {
        long long g = positions[p];
        for (uint32_t t = 0; t < blockDim.x; ++t) { const long long v = chunk_max[t]; g = v > g ? v : g; }
        positions[p] = g;
    }
}

int tk_commit_final(tk_ctx* ctx, int slot, int wait_for_prior, void* stream) {
    TK_ENTER(ctx, slot);
    tk_slot& s = ctx->slots[slot];
    // upper bound of the batches this slot can touch: carry < B, kept <= n
    const uint32_t B = ctx->batch_size ? ctx->batch_size : 1;
    const uint64_t nb_bound = ((uint64_t)s.n_records + B - 1) / B + 1;
    if (nb_bound <= 64)
        TK_LAUNCH_PDL(k_commit_final, 1, 256, (cudaStream_t)stream, s.d.commit, ctx->commit_pitch, ctx->n_partitions,
                      (const tk_result_t*)s.d.result, ctx->d_positions, wait_for_prior);
    else
        TK_LAUNCH_PDL(k_commit_rows, ctx->n_partitions, 128, (cudaStream_t)stream, s.d.commit, ctx->commit_pitch,
                      (const tk_result_t*)s.d.result, ctx->d_positions, wait_for_prior);
    return 0;
}

tk_plan_reset tk_plan_reset_args(tk_ctx* ctx, int slot) {
    tk_slot& s = ctx->slots[slot];
    tk_plan_reset r;
    r.commit = s.d.commit; r.batch_end = s.d.batch_end; r.batch_lmax = s.d.batch_lmax;
    r.commit_elems = (uint32_t)(ctx->n_partitions * ctx->commit_pitch);
    r.n_batches = ctx->max_batches + 1;
    r.last_record = s.n_records ? s.n_records - 1 : 0;
    return r;
}

extern "C" int tk_commit_offsets(tk_ctx* ctx, int slot, void* stream) {
    return tk_commit_final(ctx, slot, 1, stream);
}

// ------------------------------------------------------------------------ carry
// The batch left open by the previous slot occupies the last carry_out rows of
// that slot's tensor: rows [B + n_kept - carry_out, B + n_kept).  Copy them to
// rows [B - carry_out, B) of the current tensor (the end of its head region).
struct carry_args {
    tk_plane_t prev[TK_MAX_PLANES];
    tk_plane_t cur[TK_MAX_PLANES];
    int n_planes;
};

__global__ void __launch_bounds__(256) k_carry_rows(const tk_result_t* __restrict__ prev_res, const carry_args a) {
    tk_pdl_launch_dependents();
    tk_pdl_wait();
    const uint64_t co = prev_res->carry_out;
    if (co == 0) return;
    const uint64_t B = prev_res->batch_size, nk = prev_res->n_kept;
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t nthr = (uint64_t)gridDim.x * blockDim.x;
    for (int k = 0; k < a.n_planes; ++k) {
        const uint64_t rb = a.prev[k].row_bytes;
        if (!a.prev[k].base || !a.cur[k].base) continue;
        const uint8_t* src = (const uint8_t*)a.prev[k].base + (B + nk - co) * rb;
        uint8_t* dst = (uint8_t*)a.cur[k].base + (B - co) * rb;
        const uint64_t nbytes = co * rb;
        if ((((uintptr_t)src | (uintptr_t)dst | nbytes) & 15u) == 0) {
            for (uint64_t x = tid; x < nbytes / 16; x += nthr)
                reinterpret_cast<uint4*>(dst)[x] = reinterpret_cast<const uint4*>(src)[x];
        } else if ((((uintptr_t)src | (uintptr_t)dst | nbytes) & 3u) == 0) {
            for (uint64_t x = tid; x < nbytes / 4; x += nthr)
                reinterpret_cast<uint32_t*>(dst)[x] = reinterpret_cast<const uint32_t*>(src)[x];
        } else {
            for (uint64_t x = tid; x < nbytes; x += nthr) dst[x] = src[x];
        }
    }
}

extern "C" int tk_carry_rows(tk_ctx* ctx, int prev_slot, int n_planes, const tk_plane_t* prev,
                             const tk_plane_t* cur, void* stream) {
    TK_ENTER(ctx, prev_slot);
    TK_REQUIRE(n_planes >= 1 && n_planes <= TK_MAX_PLANES && prev && cur, "1..%d planes", TK_MAX_PLANES);
    carry_args a;
    a.n_planes = n_planes;
    for (int k = 0; k < n_planes; ++k) {
        TK_REQUIRE(prev[k].row_bytes == cur[k].row_bytes, "plane %d: row_bytes differ", k);
        a.prev[k] = prev[k];
        a.cur[k] = cur[k];
    }
    TK_LAUNCH_PDL(k_carry_rows, 148 * 2, 256, (cudaStream_t)stream, ctx->slots[prev_slot].d.result, a);
    return 0;
}
