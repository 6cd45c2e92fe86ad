// tk_ctx.cu -- context, staging slots, H2D submission (K0), result fetch.
//
// No reference counterpart: the reference never leaves host memory
// (SURVEY.md §2.2, K0).  A slot is a CSR-packed run of polled records in pinned
// host memory with a device twin; tk_slot_submit() copies it on the library's own
// copy stream so the transfer overlaps the kernels working on the previous slot.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include <new>
#include <thread>
#include <vector>
#include "tk_internal.h"

#include <atomic>
#include <condition_variable>
#include <functional>
#include <memory>
#include <mutex>
#if defined(__linux__)
#include <pthread.h>
#include <sched.h>
#endif

static thread_local char g_err[1024] = "";

// ---------------------------------------------------------------- host copy pool
// Persistent worker threads for the pinned-memory pack: one core's memcpy (~10 GB/s)
// is far below the H2D link (~55 GB/s), so the payload copy is split across threads
// (SURVEY.md §7 hard part 1).  Callers run with the GIL released (ctypes).
class CopyPool {
public:
    explicit CopyPool(int n) : stop_(false), pending_(0), gen_(0) {
        for (int i = 0; i < n; ++i) workers_.emplace_back([this, i]() { loop(i); });
    }
    ~CopyPool() {
        { std::lock_guard<std::mutex> g(m_); stop_ = true; ++gen_; }
        cv_.notify_all();
        for (auto& t : workers_) t.join();
    }
    int size() const { return (int)workers_.size(); }
    // run fn(part) for part in [0, parts) on the pool (+ the calling thread), wait for all
    void run(int parts, const std::function<void(int)>& fn) {
        if (parts <= 1 || workers_.empty()) { for (int p = 0; p < parts; ++p) fn(p); return; }
        std::lock_guard<std::mutex> serial(run_m_);
        {
            std::lock_guard<std::mutex> g(m_);
            fn_ = &fn; parts_ = parts; next_ = 0; pending_ = parts; ++gen_;
        }
        cv_.notify_all();
        work();
        std::unique_lock<std::mutex> g(m_);
        done_cv_.wait(g, [this]() { return pending_ == 0; });
        fn_ = nullptr;
    }
private:
    void work() {
        for (;;) {
            int p;
            {
                std::lock_guard<std::mutex> g(m_);
                if (!fn_ || next_ >= parts_) return;
                p = next_++;
            }
            (*fn_)(p);
            {
                std::lock_guard<std::mutex> g(m_);
                if (--pending_ == 0) done_cv_.notify_all();
            }
        }
    }
    void loop(int);
    std::vector<std::thread> workers_;
    std::mutex m_, run_m_;
    std::condition_variable cv_, done_cv_;
    const std::function<void(int)>* fn_ = nullptr;
    int parts_ = 0, next_ = 0;
    bool stop_;
    int pending_;
    unsigned long gen_;
};

// The pool is process-wide (one process drives one GPU).  A caller asking for more threads than
// the pool has gets a new, larger pool; the old one is destroyed by whoever holds its last
// reference, so a run() in flight on it finishes undisturbed.
static std::shared_ptr<CopyPool> g_pool;
static std::mutex g_pool_m;

// Cores the pack threads run on (tk_pack_set_affinity): applied by every worker when it wakes up
// and finds a new generation of the list.
static std::vector<int> g_pack_cpus;
static std::atomic<unsigned> g_pack_cpus_gen{0};

static void apply_pack_affinity(unsigned* seen) {
#if defined(__linux__)
    const unsigned gen = g_pack_cpus_gen.load(std::memory_order_acquire);
    if (gen == *seen) return;
    *seen = gen;
    std::vector<int> cpus;
    { std::lock_guard<std::mutex> g(g_pool_m); cpus = g_pack_cpus; }
    if (cpus.empty()) return;
    cpu_set_t set;
    CPU_ZERO(&set);
    for (int c : cpus) if (c >= 0 && c < CPU_SETSIZE) CPU_SET(c, &set);
    pthread_setaffinity_np(pthread_self(), sizeof(set), &set);
#else
    (void)seen;
#endif
}

void CopyPool::loop(int) {
    unsigned long seen = 0;
    unsigned cpus_seen = 0;
    for (;;) {
        {
            std::unique_lock<std::mutex> g(m_);
            cv_.wait(g, [&]() { return gen_ != seen; });
            seen = gen_;
            if (stop_) return;
        }
        apply_pack_affinity(&cpus_seen);
        work();
    }
}

static std::shared_ptr<CopyPool> copy_pool(int threads) {
    std::lock_guard<std::mutex> g(g_pool_m);
    if (!g_pool || g_pool->size() < threads - 1)
        g_pool = std::make_shared<CopyPool>(threads > 1 ? threads - 1 : 0);
    return g_pool;
}

// Pin the pack threads to `cpus` (n = 0: leave them where they are).  The pinned staging slots
// should be filled by cores of the NUMA node the GPU hangs off (torchkafka_b200/affinity.py).
extern "C" int tk_pack_set_affinity(const int* cpus, int n) {
    TK_REQUIRE(n >= 0 && (cpus || n == 0), "bad cpu list");
    {
        std::lock_guard<std::mutex> g(g_pool_m);
        g_pack_cpus.assign(cpus, cpus + n);
    }
    g_pack_cpus_gen.fetch_add(1, std::memory_order_release);
    return 0;
}

// Page-lock a host buffer the caller owns (a consumer's receive arena) so that the H2D copies of
// tk_slot_submit_recordbatches_ext are true asynchronous DMA transfers out of it.
extern "C" int tk_host_register(void* ptr, uint64_t bytes) {
    TK_REQUIRE(ptr && bytes, "bad buffer");
    cudaError_t e = cudaHostRegister(ptr, bytes, cudaHostRegisterPortable);
    if (e == cudaErrorHostMemoryAlreadyRegistered) { cudaGetLastError(); return 0; }
    if (e != cudaSuccess) { cudaGetLastError(); tk_set_error("cudaHostRegister failed: %s", cudaGetErrorString(e)); return -1; }
    return 0;
}
extern "C" int tk_host_unregister(void* ptr) {
    TK_REQUIRE(ptr, "bad buffer");
    cudaError_t e = cudaHostUnregister(ptr);
    if (e != cudaSuccess) { cudaGetLastError(); tk_set_error("cudaHostUnregister failed: %s", cudaGetErrorString(e)); return -1; }
    return 0;
}

// Gather copy with non-temporal stores: the destination (pinned staging) is written once and
// next read by the DMA engine, so it should not be pulled into the cache first (saves the
// read-for-ownership traffic of ordinary stores; host DRAM bandwidth is the pack's limit).
#if defined(__x86_64__)
#include <immintrin.h>
static inline void stream_copy(uint8_t* dst, const uint8_t* src, size_t n) {
    if (n < 256) { memcpy(dst, src, n); return; }
    const size_t head = (16 - ((uintptr_t)dst & 15)) & 15;
    if (head) { memcpy(dst, src, head); dst += head; src += head; n -= head; }
    size_t i = 0;
    for (; i + 64 <= n; i += 64) {
        const __m128i a = _mm_loadu_si128((const __m128i*)(src + i));
        const __m128i b = _mm_loadu_si128((const __m128i*)(src + i + 16));
        const __m128i c = _mm_loadu_si128((const __m128i*)(src + i + 32));
        const __m128i d = _mm_loadu_si128((const __m128i*)(src + i + 48));
        _mm_stream_si128((__m128i*)(dst + i), a);
        _mm_stream_si128((__m128i*)(dst + i + 16), b);
        _mm_stream_si128((__m128i*)(dst + i + 32), c);
        _mm_stream_si128((__m128i*)(dst + i + 48), d);
    }
    if (i < n) memcpy(dst + i, src + i, n - i);
}
static inline void stream_fence() { _mm_sfence(); }
#else
static inline void stream_copy(uint8_t* dst, const uint8_t* src, size_t n) { memcpy(dst, src, n); }
static inline void stream_fence() {}
#endif

static void parallel_memcpy(uint8_t* dst, const uint8_t* src, uint64_t n, int threads) {
    if (threads < 1) threads = 1;
    if (threads > 64) threads = 64;
    if (n < (4u << 20) || threads == 1) { memcpy(dst, src, n); return; }
    const uint64_t chunk = ((n + threads - 1) / threads + 4095) & ~4095ull;
    const int parts = (int)((n + chunk - 1) / chunk);
    copy_pool(threads)->run(parts, [=](int p) {
        const uint64_t a = (uint64_t)p * chunk;
        const uint64_t b = a + chunk < n ? a + chunk : n;
        memcpy(dst + a, src + a, b - a);
    });
}

void tk_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* tk_last_error(void) { return g_err; }
extern "C" int tk_abi_version(void) { return TK_ABI_VERSION; }

__global__ void tk_fill_i64(int64_t* p, int64_t v, uint32_t n) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

static void free_slot(tk_slot& s) {
    if (s.h_bytes) cudaFreeHost(s.h_bytes);
    if (s.h_recs) cudaFreeHost(s.h_recs);
    if (s.h_block) cudaFreeHost(s.h_block);
    if (s.h_rb_batches) cudaFreeHost(s.h_rb_batches);
    if (s.h_rb_units) cudaFreeHost(s.h_rb_units);
    if (s.h_rb_summary) cudaFreeHost(s.h_rb_summary);
    cudaFree(s.d.raw); cudaFree(s.d.raw2); cudaFree(s.d.rb_batches); cudaFree(s.d.rb_units); cudaFree(s.d.rb_summary); cudaFree(s.d.rb_temp); cudaFree(s.d.rb_bigq);
    cudaFree(s.d.bytes); cudaFree(s.d.recs); cudaFree(s.d.rec_len); cudaFree(s.d.rec_status);
    cudaFree(s.d.rec_kidx); cudaFree(s.d.tile_agg); cudaFree(s.d.tile_incl); cudaFree(s.d.tile_flag);
    cudaFree(s.d.plan_ctr); cudaFree(s.d.block);
    cudaFree(s.d.rec_tok); cudaFree(s.d.rec_base); cudaFree(s.d.tok_pos); cudaFree(s.d.tok_count);
    cudaFree(s.d.fz_tile_first); cudaFree(s.d.fz_agg); cudaFree(s.d.fz_incl); cudaFree(s.d.fz_flag); cudaFree(s.d.fz_ctr);
    if (s.ev_h2d) cudaEventDestroy(s.ev_h2d);
    if (s.ev_free) cudaEventDestroy(s.ev_free);
    if (s.ev_result) cudaEventDestroy(s.ev_result);
    if (s.ev_raw) cudaEventDestroy(s.ev_raw);
    if (s.split_stream) cudaStreamDestroy(s.split_stream);
    if (s.ev_dec0) cudaEventDestroy(s.ev_dec0);
    if (s.ev_dec1) cudaEventDestroy(s.ev_dec1);
}

extern "C" int tk_ctx_destroy(tk_ctx* ctx) {
    if (!ctx) return 0;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    if (ctx->slots) {
        for (int i = 0; i < ctx->n_slots; ++i) free_slot(ctx->slots[i]);
        delete[] ctx->slots;
    }
    cudaFree(ctx->d_state);
    cudaFree(ctx->d_positions);
    cudaFree(ctx->d_image_lut);
    cudaFree(ctx->d_obj_table);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    delete ctx;
    return 0;
}

static int create_impl(tk_ctx* ctx) {
    TK_CUDA(cudaSetDevice(ctx->device));
    TK_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    TK_CUDA(cudaMalloc(&ctx->d_state, sizeof(tk_state)));
    TK_CUDA(cudaMalloc(&ctx->d_positions, sizeof(int64_t) * ctx->n_partitions));
    TK_CUDA(cudaMalloc(&ctx->d_image_lut, sizeof(uint16_t) * 4 * 256 + 16));      // + a spare word
    TK_CUDA(cudaMemset(ctx->d_image_lut, 0, sizeof(uint16_t) * 4 * 256 + 16));
    ctx->slots = new (std::nothrow) tk_slot[ctx->n_slots];
    TK_REQUIRE(ctx->slots, "out of host memory");
    memset(ctx->slots, 0, sizeof(tk_slot) * ctx->n_slots);
    const size_t nrec = ctx->slot_records;
    const size_t nb = ctx->max_batches + 1;
    const size_t commit_elems = (size_t)ctx->n_partitions * ctx->commit_pitch;
    const size_t nblk = (nrec + 255) / 256 + 1;      // one entry per compaction tile (tk_plan.cu CMP_TILE)
    for (int i = 0; i < ctx->n_slots; ++i) {
        tk_slot& s = ctx->slots[i];
        TK_CUDA(cudaHostAlloc(&s.h_bytes, ctx->slot_bytes + TK_BYTES_PAD, cudaHostAllocDefault));
        TK_CUDA(cudaHostAlloc(&s.h_recs, sizeof(tk_rec_t) * (nrec + 1), cudaHostAllocDefault));
        // result block: header | batch_end[nb] | batch_lmax[nb] | commit[P][pitch], one D2H per slot
        const size_t off_end = sizeof(tk_result_t);
        const size_t off_lmax = off_end + ((sizeof(uint32_t) * nb + 7) & ~(size_t)7);
        const size_t off_commit = off_lmax + ((sizeof(uint32_t) * nb + 7) & ~(size_t)7);
        ctx->block_bytes = off_commit + sizeof(int64_t) * commit_elems;
        TK_CUDA(cudaHostAlloc(&s.h_block, ctx->block_bytes, cudaHostAllocDefault));
        s.h_result = reinterpret_cast<tk_result_t*>(s.h_block);
        s.h_batch_end = reinterpret_cast<uint32_t*>(s.h_block + off_end);
        s.h_batch_lmax = reinterpret_cast<uint32_t*>(s.h_block + off_lmax);
        s.h_commit = reinterpret_cast<int64_t*>(s.h_block + off_commit);
        TK_CUDA(cudaMalloc(&s.d.bytes, ctx->slot_bytes + TK_BYTES_PAD));
        TK_CUDA(cudaMemset(s.d.bytes, 0, ctx->slot_bytes + TK_BYTES_PAD));
        TK_CUDA(cudaMalloc(&s.d.recs, sizeof(tk_rec_t) * (nrec + 1)));
        TK_CUDA(cudaMalloc(&s.d.rec_len, sizeof(int32_t) * nrec));
        TK_CUDA(cudaMalloc(&s.d.rec_status, sizeof(int32_t) * nrec));
        TK_CUDA(cudaMalloc(&s.d.rec_kidx, sizeof(int32_t) * nrec));
        TK_CUDA(cudaMalloc(&s.d.tile_agg, sizeof(unsigned long long) * nblk));
        TK_CUDA(cudaMalloc(&s.d.tile_incl, sizeof(unsigned long long) * nblk));
        TK_CUDA(cudaMalloc(&s.d.tile_flag, sizeof(uint32_t) * nblk));
        TK_CUDA(cudaMemset(s.d.tile_flag, 0, sizeof(uint32_t) * nblk));
        TK_CUDA(cudaMalloc(&s.d.plan_ctr, 2 * sizeof(uint32_t)));
        TK_CUDA(cudaMemset(s.d.plan_ctr, 0, 2 * sizeof(uint32_t)));
        TK_CUDA(cudaMalloc(&s.d.rec_tok, sizeof(uint32_t) * (nrec + 1)));
        TK_CUDA(cudaMalloc(&s.d.rec_base, sizeof(long long) * (nrec + 1)));
        // token arrays of the JSON path are allocated on first use (tk_ensure_token_arrays)
        TK_CUDA(cudaMalloc(&s.d.block, ctx->block_bytes));
        s.d.result = reinterpret_cast<tk_result_t*>(s.d.block);
        s.d.batch_end = reinterpret_cast<uint32_t*>(s.d.block + off_end);
        s.d.batch_lmax = reinterpret_cast<uint32_t*>(s.d.block + off_lmax);
        s.d.commit = reinterpret_cast<int64_t*>(s.d.block + off_commit);
        TK_CUDA(cudaEventCreateWithFlags(&s.ev_h2d, cudaEventDisableTiming));
        TK_CUDA(cudaEventCreateWithFlags(&s.ev_free, cudaEventDisableTiming));
        TK_CUDA(cudaEventCreateWithFlags(&s.ev_result, cudaEventDisableTiming));
        TK_CUDA(cudaEventCreate(&s.ev_dec0));
        TK_CUDA(cudaEventCreate(&s.ev_dec1));
    }
    return tk_ctx_reset(ctx, nullptr);
}

extern "C" int tk_ctx_create(int device, uint64_t slot_bytes, uint32_t slot_records, int n_slots,
                             uint32_t max_batches, uint32_t n_partitions, tk_ctx** out) {
    TK_REQUIRE(out, "out is null");
    *out = nullptr;
    TK_REQUIRE(slot_bytes > 0 && slot_bytes < (1ull << 32) - TK_BYTES_PAD,
               "slot_bytes must be in (0, 4 GiB): CSR offsets are 32-bit");
    TK_REQUIRE(slot_records > 0 && slot_records < (1u << 30) && n_slots > 0 && n_slots <= 64 && max_batches > 0 && n_partitions > 0,
               "bad context geometry");
    tk_ctx* ctx = new (std::nothrow) tk_ctx;
    TK_REQUIRE(ctx, "out of host memory");
    memset(ctx, 0, sizeof(*ctx));
    ctx->device = device;
    ctx->slot_bytes = slot_bytes;
    ctx->slot_records = slot_records;
    ctx->n_slots = n_slots;
    ctx->max_batches = max_batches;
    ctx->commit_pitch = (max_batches + 1 + 3) & ~3u;
    ctx->n_partitions = n_partitions;
    { const char* e = getenv("TK_JSON_FUSED"); ctx->json_fused = (e && e[0] == '1') ? 1 : 0; }
    { const char* e = getenv("TK_JSON_WIDE"); ctx->json_wide = (e && e[0] == '1') ? 1 : 0; }   // (the default; tk_ctx_set_json_wide)
    int rc = create_impl(ctx);
    if (rc) { char keep[1024]; strncpy(keep, g_err, sizeof(keep)); keep[1023] = 0; tk_ctx_destroy(ctx); tk_set_error("%s", keep); return rc; }
    *out = ctx;
    return 0;
}

// The flat JSON decode keeps one (position, record) pair per element of the slot.
// An element needs at least two bytes ("1,"), so slot_bytes / 2 + slot_records entries always
// suffice.  Raw / image contexts never pay for it.
int tk_ensure_token_arrays(tk_ctx* ctx) {
    if (ctx->tok_cap) return 0;
    const size_t cap = (size_t)(ctx->slot_bytes / 2) + ctx->slot_records + 64;
    for (int i = 0; i < ctx->n_slots; ++i) {
        tk_slot& s = ctx->slots[i];
        TK_CUDA(cudaMalloc(&s.d.tok_pos, sizeof(uint32_t) * 2 * cap));       // uint2 {position, record}
        TK_CUDA(cudaMalloc(&s.d.tok_count, 2 * sizeof(uint32_t)));
        TK_CUDA(cudaMemset(s.d.tok_count, 0, 2 * sizeof(uint32_t)));
        s.tok_use = 0;
    }
    ctx->tok_cap = cap;
    return 0;
}

extern "C" int tk_ctx_reset(tk_ctx* ctx, void* stream) {
    TK_REQUIRE(ctx, "null context");
    TK_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    TK_CUDA(cudaMemsetAsync(ctx->d_state, 0, sizeof(tk_state), st));
    tk_fill_i64<<<(ctx->n_partitions + 255) / 256, 256, 0, st>>>(ctx->d_positions, -1, ctx->n_partitions);
    TK_LAUNCH_CHECK();
    TK_CUDA(cudaStreamSynchronize(st));
    return 0;
}

extern "C" int tk_slot_host_ptrs(tk_ctx* ctx, int slot, uint8_t** bytes, tk_rec_t** recs) {
    int rc = tk_check_slot(ctx, slot); if (rc) return rc;
    if (bytes) *bytes = ctx->slots[slot].h_bytes;
    if (recs) *recs = ctx->slots[slot].h_recs;
    return 0;
}

extern "C" int tk_slot_device_ptrs(tk_ctx* ctx, int slot, uint8_t** d_bytes, tk_rec_t** d_recs) {
    int rc = tk_check_slot(ctx, slot); if (rc) return rc;
    if (d_bytes) *d_bytes = ctx->slots[slot].d.bytes;
    if (d_recs) *d_recs = ctx->slots[slot].d.recs;
    return 0;
}

static int check_fill(tk_ctx* ctx, uint32_t n_records, uint64_t n_bytes) {
    TK_REQUIRE(n_records <= ctx->slot_records, "slot holds %u records, %u submitted", ctx->slot_records, n_records);
    TK_REQUIRE(n_bytes <= ctx->slot_bytes, "slot holds %llu bytes, %llu submitted",
               (unsigned long long)ctx->slot_bytes, (unsigned long long)n_bytes);
    return 0;
}

extern "C" int tk_slot_submit(tk_ctx* ctx, int slot, uint32_t n_records, uint64_t n_bytes) {
    TK_ENTER(ctx, slot);
    int rc = check_fill(ctx, n_records, n_bytes); if (rc) return rc;
    tk_slot& s = ctx->slots[slot];
    TK_REQUIRE(s.h_recs[n_records].start == n_bytes, "recs[n_records].start (%u) must equal n_bytes (%llu)",
               s.h_recs[n_records].start, (unsigned long long)n_bytes);
    s.n_records = n_records;
    s.n_bytes = n_bytes;
    // the device twin may still be read by kernels of the slot's previous round
    if (s.free_recorded) TK_CUDA(cudaStreamWaitEvent(ctx->copy_stream, s.ev_free, 0));
    if (n_bytes) TK_CUDA(cudaMemcpyAsync(s.d.bytes, s.h_bytes, n_bytes, cudaMemcpyHostToDevice, ctx->copy_stream));
    TK_CUDA(cudaMemcpyAsync(s.d.recs, s.h_recs, sizeof(tk_rec_t) * ((size_t)n_records + 1),
                            cudaMemcpyHostToDevice, ctx->copy_stream));
    TK_CUDA(cudaEventRecord(s.ev_h2d, ctx->copy_stream));
    return 0;
}

extern "C" int tk_slot_assume_resident(tk_ctx* ctx, int slot, uint32_t n_records, uint64_t n_bytes) {
    TK_ENTER(ctx, slot);
    int rc = check_fill(ctx, n_records, n_bytes); if (rc) return rc;
    tk_slot& s = ctx->slots[slot];
    s.n_records = n_records;
    s.n_bytes = n_bytes;
    TK_CUDA(cudaEventRecord(s.ev_h2d, ctx->copy_stream));
    return 0;
}

extern "C" int tk_slot_wait(tk_ctx* ctx, int slot, void* stream) {
    TK_ENTER(ctx, slot);
    TK_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, ctx->slots[slot].ev_h2d, 0));
    return 0;
}

extern "C" int tk_slot_pack(tk_ctx* ctx, int slot, uint32_t first_record, uint64_t first_byte,
                            const uint8_t* const* values, const uint32_t* lengths,
                            const int32_t* partitions, const int64_t* offsets, uint32_t n,
                            int n_threads, uint64_t* bytes_after) {
    int rc = tk_check_slot(ctx, slot); if (rc) return rc;
    tk_slot& s = ctx->slots[slot];
    TK_REQUIRE((uint64_t)first_record + n <= ctx->slot_records, "pack: record capacity exceeded");
    // pass 1: headers + total size (serial, 16 bytes per record)
    uint64_t pos = first_byte;
    for (uint32_t i = 0; i < n; ++i) {
        tk_rec_t& r = s.h_recs[first_record + i];
        r.start = (uint32_t)pos;
        r.partition = partitions[i];
        r.offset = offsets[i];
        pos += lengths[i];
    }
    TK_REQUIRE(pos <= ctx->slot_bytes, "pack: byte capacity exceeded");
    s.h_recs[first_record + n].start = (uint32_t)pos;
    s.h_recs[first_record + n].partition = 0;
    s.h_recs[first_record + n].offset = 0;
    // pass 2: payload copies, split by record ranges of similar byte volume
    const uint64_t total = pos - first_byte;
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 64) n_threads = 64;
    if (total < (1u << 20) || n_threads == 1) {
        for (uint32_t i = 0; i < n; ++i) memcpy(s.h_bytes + s.h_recs[first_record + i].start, values[i], lengths[i]);
    } else {
        // equal byte ranges, one per part; a range may start and end inside a value (a fetch
        // response can be a handful of multi-megabyte blobs)
        tk_slot* sp = &s;
        copy_pool(n_threads)->run(n_threads, [=](int t) {
            const uint64_t lo = first_byte + total * (uint64_t)t / n_threads;
            const uint64_t hi = first_byte + total * (uint64_t)(t + 1) / n_threads;
            if (hi <= lo) return;
            // first record whose end lies behind lo
            uint32_t a = 0, b = n;
            while (a < b) {
                const uint32_t m = (a + b) / 2;
                if ((uint64_t)sp->h_recs[first_record + m + 1].start <= lo) a = m + 1; else b = m;
            }
            for (uint32_t i = a; i < n; ++i) {
                const uint64_t rs = sp->h_recs[first_record + i].start;
                if (rs >= hi) break;
                const uint64_t re = rs + lengths[i];
                const uint64_t from = rs > lo ? rs : lo, to = re < hi ? re : hi;
                if (to > from) stream_copy(sp->h_bytes + from, values[i] + (from - rs), to - from);
            }
            stream_fence();
        });
    }
    if (bytes_after) *bytes_after = pos;
    return 0;
}

extern "C" int tk_slot_fill(tk_ctx* ctx, int slot, uint32_t first_record, uint64_t first_byte,
                            const uint8_t* values, const uint32_t* starts, const int32_t* partitions,
                            const int64_t* offsets, uint32_t n, int n_threads, uint64_t* bytes_after) {
    int rc = tk_check_slot(ctx, slot); if (rc) return rc;
    tk_slot& s = ctx->slots[slot];
    TK_REQUIRE((uint64_t)first_record + n <= ctx->slot_records, "fill: record capacity exceeded");
    const uint64_t base = starts[0];
    const uint64_t total = (uint64_t)starts[n] - base;
    TK_REQUIRE(first_byte + total <= ctx->slot_bytes, "fill: byte capacity exceeded");
    for (uint32_t i = 0; i < n; ++i) {
        tk_rec_t& r = s.h_recs[first_record + i];
        r.start = (uint32_t)(first_byte + (starts[i] - base));
        r.partition = partitions[i];
        r.offset = offsets[i];
    }
    tk_rec_t& last = s.h_recs[first_record + n];
    last.start = (uint32_t)(first_byte + total);
    last.partition = 0;
    last.offset = 0;
    parallel_memcpy(s.h_bytes + first_byte, values + base, total, n_threads);
    if (bytes_after) *bytes_after = first_byte + total;
    return 0;
}

extern "C" int tk_h2d_peak_gbs(int device, uint64_t bytes, int iters, double* gbs_out) {
    TK_REQUIRE(gbs_out && bytes > 0 && iters > 0, "bad arguments");
    TK_CUDA(cudaSetDevice(device));
    void *h = nullptr, *d = nullptr;
    cudaStream_t st;
    cudaEvent_t e0, e1;
    TK_CUDA(cudaHostAlloc(&h, bytes, cudaHostAllocDefault));
    memset(h, 1, bytes);
    TK_CUDA(cudaMalloc(&d, bytes));
    TK_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    TK_CUDA(cudaEventCreate(&e0));
    TK_CUDA(cudaEventCreate(&e1));
    double best = 0.0;
    for (int i = 0; i < iters + 1; ++i) {
        TK_CUDA(cudaEventRecord(e0, st));
        TK_CUDA(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, st));
        TK_CUDA(cudaEventRecord(e1, st));
        TK_CUDA(cudaEventSynchronize(e1));
        float ms = 0.f;
        TK_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        if (i > 0 && ms > 0.f) { double g = (double)bytes / (ms * 1e6); if (g > best) best = g; }
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaStreamDestroy(st); cudaFree(d); cudaFreeHost(h);
    *gbs_out = best;
    return 0;
}

extern "C" int tk_slot_fetch_result(tk_ctx* ctx, int slot, void* stream) {
    TK_ENTER(ctx, slot);
    tk_slot& s = ctx->slots[slot];
    cudaStream_t st = (cudaStream_t)stream;
    TK_CUDA(cudaMemcpyAsync(s.h_block, s.d.block, ctx->block_bytes, cudaMemcpyDeviceToHost, st));
    TK_CUDA(cudaEventRecord(s.ev_result, st));
    TK_CUDA(cudaEventRecord(s.ev_free, st));
    s.free_recorded = true;
    return 0;
}

extern "C" int tk_slot_result(tk_ctx* ctx, int slot, const tk_result_t** hdr, const uint32_t** batch_end,
                              const uint32_t** batch_lmax, const int64_t** commit, uint32_t* commit_pitch) {
    TK_ENTER(ctx, slot);
    tk_slot& s = ctx->slots[slot];
    TK_CUDA(cudaEventSynchronize(s.ev_result));
    if (hdr) *hdr = s.h_result;
    if (batch_end) *batch_end = s.h_batch_end;
    if (batch_lmax) *batch_lmax = s.h_batch_lmax;
    if (commit) *commit = s.h_commit;
    if (commit_pitch) *commit_pitch = ctx->commit_pitch;
    return 0;
}

extern "C" int tk_slot_debug_records(tk_ctx* ctx, int slot, int32_t* rec_len, int32_t* rec_status,
                                     int32_t* rec_kidx, uint32_t n) {
    TK_ENTER(ctx, slot);
    tk_slot& s = ctx->slots[slot];
    TK_REQUIRE(n <= ctx->slot_records, "n too large");
    TK_CUDA(cudaDeviceSynchronize());
    if (rec_len) TK_CUDA(cudaMemcpy(rec_len, s.d.rec_len, sizeof(int32_t) * n, cudaMemcpyDeviceToHost));
    if (rec_status) TK_CUDA(cudaMemcpy(rec_status, s.d.rec_status, sizeof(int32_t) * n, cudaMemcpyDeviceToHost));
    if (rec_kidx) TK_CUDA(cudaMemcpy(rec_kidx, s.d.rec_kidx, sizeof(int32_t) * n, cudaMemcpyDeviceToHost));
    return 0;
}
