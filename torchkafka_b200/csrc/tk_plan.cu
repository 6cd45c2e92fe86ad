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

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, v, d);
        if ((threadIdx.x & 31) >= d) v += t;
    }
    return v;
}

// block-wide exclusive scan of one value per thread (blockDim.x <= 1024); returns total via *total
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* total) {
    __shared__ uint32_t warp_tot[32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t incl = warp_incl_scan(v);
    if (lane == 31) warp_tot[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        uint32_t w = lane < (int)((blockDim.x + 31) >> 5) ? warp_tot[lane] : 0;
        uint32_t wi = warp_incl_scan(w);
        warp_tot[lane] = wi - w;                       // exclusive warp offsets
        if (lane == 31) warp_tot[31] = wi - w, *total = wi;
    }
    __syncthreads();
    uint32_t r = warp_tot[wid] + incl - v;
    __syncthreads();
    return r;
}

// 64-bit variants: the compaction scans two quantities at once, packed as
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

// pass 1: kept count per tile of CMP_TILE records
__global__ void __launch_bounds__(CMP_THREADS) k_compact_count(const int32_t* __restrict__ rec_kidx,
                                                                 const int32_t* __restrict__ rec_len, uint32_t n,
                                                                 unsigned long long* __restrict__ blk_sum) {
    __shared__ unsigned long long tot;
    tk_pdl_launch_dependents();
    tk_pdl_wait();
    const uint32_t base = blockIdx.x * CMP_TILE + threadIdx.x * CMP_ITEMS;
    unsigned long long c = 0;
#pragma unroll
    for (int k = 0; k < CMP_ITEMS; ++k) {
        uint32_t i = base + k;
        if (i < n) c += ((unsigned long long)(rec_kidx[i] >= 0 ? 1 : 0) << PACK_SHIFT) + (uint32_t)rec_len[i];
    }
    block_excl_scan64(c, &tot);
    if (threadIdx.x == 0) blk_sum[blockIdx.x] = tot;
}

// pass 2: one block scans the tile sums, fills the result header, advances the carry
__global__ void __launch_bounds__(1024) k_compact_scan(unsigned long long* __restrict__ blk_sum, uint32_t nblk,
                                                       uint32_t n, uint32_t batch_size,
                                                       uint32_t max_batches, tk_state* state,
                                                       tk_result_t* __restrict__ res,
                                                       uint32_t* __restrict__ batch_end,
                                                       uint32_t* __restrict__ batch_lmax,
                                                       int64_t* __restrict__ commit, uint32_t commit_elems) {
    __shared__ unsigned long long tot;
    tk_pdl_launch_dependents();
    tk_pdl_wait();
    // commit matrix starts at -1 ("partition not seen"); filled by the apply pass
    for (uint32_t x = threadIdx.x; x < commit_elems; x += blockDim.x) commit[x] = -1;
    unsigned long long running = 0;
    for (uint32_t base = 0; base < nblk; base += blockDim.x) {
        uint32_t i = base + threadIdx.x;
        unsigned long long v = i < nblk ? blk_sum[i] : 0;
        unsigned long long ex = block_excl_scan64(v, &tot);
        if (i < nblk) blk_sum[i] = running + ex;
        running += tot;
        __syncthreads();
    }
    const uint32_t n_kept = (uint32_t)(running >> PACK_SHIFT);
    const unsigned long long n_tokens = running & ((1ull << PACK_SHIFT) - 1);
    const uint32_t c = state->carry_rows;
    const uint32_t n_closed = (c + n_kept) / batch_size;
    for (uint32_t j = threadIdx.x; j <= max_batches; j += blockDim.x) {
        batch_end[j] = n ? n - 1 : 0;
        batch_lmax[j] = 0;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        res->n_records = n;
        res->n_kept = n_kept;
        res->carry_in = c;
        res->carry_out = (c + n_kept) % batch_size;
        res->n_closed = n_closed;
        res->batch_size = batch_size;
        res->n_errors = 0;
        res->first_error_rec = 0xFFFFFFFFu;
        res->first_error_status = 0;
        res->reserved[0] = (uint32_t)(n_tokens > 0xFFFFFFFFull ? 0xFFFFFFFFu : n_tokens);   // elements in the slot
        state->carry_rows = (c + n_kept) % batch_size;
    }
}

// pass 3: final index of every record.
//   kept record    -> rec_kidx = k            (k = kept records before it in this slot)
//   skipped record -> rec_kidx = -1 - k
// and per batch: the longest kept record (one atomic per warp when the warp lies inside one
// batch, which is the common case) and the index of the record that closed it.
__global__ void __launch_bounds__(CMP_THREADS) k_compact_apply(int32_t* __restrict__ rec_kidx,
                                                                 const int32_t* __restrict__ rec_len,
                                                                 uint32_t n, const unsigned long long* __restrict__ blk_sum,
                                                                 const tk_result_t* __restrict__ res,
                                                                 uint32_t* __restrict__ batch_end,
                                                                 uint32_t* __restrict__ batch_lmax,
                                                                 const tk_rec_t* __restrict__ recs,
                                                                 int64_t* __restrict__ commit, uint32_t pitch,
                                                                 uint32_t n_partitions) {
    __shared__ unsigned long long tot;
    tk_pdl_launch_dependents();
    tk_pdl_wait();
    const uint32_t base = blockIdx.x * CMP_TILE + threadIdx.x * CMP_ITEMS;
    int keep[CMP_ITEMS];
    uint32_t lens[CMP_ITEMS];
    unsigned long long c = 0;
#pragma unroll
    for (int k = 0; k < CMP_ITEMS; ++k) {
        uint32_t i = base + k;
        keep[k] = (i < n && rec_kidx[i] >= 0) ? 1 : 0;
        lens[k] = i < n ? (uint32_t)rec_len[i] : 0u;
        c += ((unsigned long long)keep[k] << PACK_SHIFT) + lens[k];
    }
    const unsigned long long ex64 = block_excl_scan64(c, &tot) + blk_sum[blockIdx.x];
    uint32_t ex = (uint32_t)(ex64 >> PACK_SHIFT);
    const uint32_t carry = res->carry_in, B = res->batch_size;
#pragma unroll
    for (int k = 0; k < CMP_ITEMS; ++k) {
        const uint32_t i = base + k;
        const bool live = i < n;
        const uint32_t j = (carry + ex) / B;               // batch of this record (kept or skipped)
        uint32_t len = 0;
        if (live) {
            if (keep[k]) {
                rec_kidx[i] = (int32_t)ex;
                len = lens[k];
                if ((carry + ex + 1) % B == 0) batch_end[j] = i;
            } else {
                rec_kidx[i] = -1 - (int32_t)ex;
            }
        }
        // longest kept record of the batch
        const uint32_t j0 = __shfl_sync(0xffffffffu, j, 0);
        if (__all_sync(0xffffffffu, j == j0)) {
            const uint32_t m = __reduce_max_sync(0xffffffffu, len);
            if ((threadIdx.x & 31) == 0 && m > *(volatile uint32_t*)&batch_lmax[j0]) atomicMax(&batch_lmax[j0], m);
        } else if (len > *(volatile uint32_t*)&batch_lmax[j]) {
            atomicMax(&batch_lmax[j], len);
        }
        // K4, first half: segmented max of (offset + 1) per (batch, partition).  Lanes of the warp
        // that share a key elect one lane that issues the group's only atomic.
        uint32_t key = 0xFFFFFFFFu, p = 0;
        long long val = 0;
        if (live) {
            const uint4 raw = __ldg(reinterpret_cast<const uint4*>(recs) + i);
            p = raw.y;
            val = (long long)(((uint64_t)raw.w << 32) | raw.z) + 1;
            if (p < n_partitions) key = j * n_partitions + p;
        }
        const unsigned peers = __match_any_sync(0xffffffffu, key);
        const uint32_t vhi = (uint32_t)((unsigned long long)val >> 32), vlo = (uint32_t)val;
        const uint32_t mh = __reduce_max_sync(peers, vhi);
        const uint32_t ml = __reduce_max_sync(peers, vhi == mh ? vlo : 0u);
        if (key != 0xFFFFFFFFu && (int)(threadIdx.x & 31) == __ffs(peers) - 1)
            atomicMax(reinterpret_cast<long long*>(&commit[(size_t)p * pitch + j]),
                      (long long)(((unsigned long long)mh << 32) | ml));
        if (live && keep[k]) ++ex;
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
    const uint32_t nblk = (n + CMP_TILE - 1) / CMP_TILE;
    if (nblk) TK_LAUNCH_PDL(k_compact_count, nblk, CMP_THREADS, st, s.d.rec_kidx, s.d.rec_len, n, s.d.blk_sum);
    TK_LAUNCH_PDL(k_compact_scan, 1, 1024, st, s.d.blk_sum, nblk, n, batch_size, ctx->max_batches, ctx->d_state,
                  s.d.result, s.d.batch_end, s.d.batch_lmax, s.d.commit,
                  (uint32_t)(ctx->n_partitions * ctx->commit_pitch));
    if (nblk)
        TK_LAUNCH_PDL(k_compact_apply, nblk, CMP_THREADS, st, s.d.rec_kidx, s.d.rec_len, n, s.d.blk_sum, s.d.result,
                      s.d.batch_end, s.d.batch_lmax, s.d.recs, s.d.commit, ctx->commit_pitch, ctx->n_partitions);
    return 0;
}

// ------------------------------------------------------------------------ K4
// First half (segmented max per (batch, partition)) runs inside k_compact_apply, where the
// batch index of every record is at hand.  Second half:
// Running max along the batch axis, one block per partition.
__global__ void __launch_bounds__(128) k_commit_scan(int64_t* __restrict__ commit, uint32_t pitch,
                                                      const tk_result_t* __restrict__ res,
                                                      int64_t* __restrict__ positions) {
    __shared__ long long chunk_max[128];
    tk_pdl_launch_dependents();
    tk_pdl_wait();
    const uint32_t p = blockIdx.x;
    const uint32_t nb = res->n_closed + 1;                 // closed batches + the open one
    int64_t* row = commit + (size_t)p * pitch;
    const uint32_t per = (nb + blockDim.x - 1) / blockDim.x;
    const uint32_t a = threadIdx.x * per;
    const uint32_t b = min(a + per, nb);
    long long m = -1;
    for (uint32_t j = a; j < b; ++j) { long long v = row[j]; m = v > m ? v : m; }
    chunk_max[threadIdx.x] = m;
    __syncthreads();
    long long run = positions[p];
    for (uint32_t t = 0; t < threadIdx.x; ++t) { long long v = chunk_max[t]; run = v > run ? v : run; }
    for (uint32_t j = a; j < b; ++j) {
        long long v = row[j];
        run = v > run ? v : run;
        row[j] = run;
    }
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) {
        // last thread's run covers every chunk only if it owned the tail; recompute the global max
        long long g = positions[p];
        for (uint32_t t = 0; t < blockDim.x; ++t) { long long v = chunk_max[t]; g = v > g ? v : g; }
        positions[p] = g;
    }
}

extern "C" int tk_commit_offsets(tk_ctx* ctx, int slot, void* stream) {
    TK_ENTER(ctx, slot);
    tk_slot& s = ctx->slots[slot];
    cudaStream_t st = (cudaStream_t)stream;
    TK_LAUNCH_PDL(k_commit_scan, ctx->n_partitions, 128, st, s.d.commit, ctx->commit_pitch, s.d.result, ctx->d_positions);
    return 0;
}

// ------------------------------------------------------------------------ carry
// The batch left open by the previous slot occupies the last carry_out rows of
// that slot's tensor: rows [B + n_kept - carry_out, B + n_kept).  Copy them to
// rows [B - carry_out, B) of the current tensor (the end of its head region).
struct carry_args {
    tk_plane_t prev[4];
    tk_plane_t cur[4];
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
    TK_REQUIRE(n_planes >= 1 && n_planes <= 4 && prev && cur, "1..4 planes");
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
