// tk_raw_kernels.cu -- K1c + K3 for fixed-layout binary payloads.
//
// Replaces `torch.frombuffer(record.value, dtype).view(shape)` per record and
// torch.stack per batch (collate.py:275); for images additionally the
// HWC->CHW permute, /255, (x-mean)/std in float32 and the bfloat16 cast of
// BASELINE.md §3 / SURVEY.md §8c.  These payloads need no parsing, so the
// kernels are pure HBM streams: coalesced 16-byte loads, 16-byte stores.
#include <cuda_bf16.h>
#include "tk_internal.h"

__device__ __forceinline__ void raw_report_error(tk_result_t* res, uint32_t rec, int status) {
    if (status == TK_STATUS_OK) return;      // not kept and no error: skipped by the record-batch split
    atomicAdd(&res->n_errors, 1u);
    const uint32_t prev = atomicMin(&res->first_error_rec, rec);
    if (rec < prev) res->first_error_status = (uint32_t)status;
}

// every record is kept; a record of the wrong size is an error
__global__ void __launch_bounds__(256) k_raw_scan(const tk_rec_t* __restrict__ recs, uint32_t n,
                                                   uint64_t record_bytes, int32_t* __restrict__ rec_len,
                                                   int32_t* __restrict__ rec_status, int32_t* __restrict__ rec_kidx,
                                                   const tk_plan_reset reset) {
    tk_pdl_launch_dependents();
    tk_pdl_wait();
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    tk_plan_reset_run(reset, i, gridDim.x * blockDim.x);
    if (i >= n) return;
    const uint32_t len = recs[i + 1].start - recs[i].start;
    rec_len[i] = (int32_t)len;
    const int32_t pflags = recs[i].partition & TK_REC_FLAGS;    // set by the record-batch split
    if (pflags) {
        rec_status[i] = (pflags & TK_REC_OVERFLOW) == TK_REC_OVERFLOW ? TK_STATUS_SLOT_OVERFLOW : (pflags & TK_REC_CORRUPT) ? TK_STATUS_CORRUPT
                      : (pflags & TK_REC_UNSUPPORTED) ? TK_STATUS_UNSUPPORTED : TK_STATUS_OK;
        rec_kidx[i] = -1;
        return;
    }
    const bool ok = len == record_bytes || record_bytes == TK_ANY_RECORD_BYTES;
    rec_status[i] = ok ? TK_STATUS_OK : TK_STATUS_SIZE;
    rec_kidx[i] = ok ? 0 : -1;
}

extern "C" int tk_raw_scan(tk_ctx* ctx, int slot, uint64_t record_bytes, void* stream) {
    TK_ENTER(ctx, slot);
    tk_slot& s = ctx->slots[slot];
    const uint32_t n = s.n_records;                        // (an empty slot still resets the plan outputs)
    s.json_row_elems = 0;
    TK_LAUNCH_PDL(k_raw_scan, max((n + 255) / 256, 1u), 256, (cudaStream_t)stream, s.d.recs, n, record_bytes, s.d.rec_len,
                  s.d.rec_status, s.d.rec_kidx, tk_plan_reset_args(ctx, slot));
    return 0;
}

// destination row of kept record `kidx`: head rows first, then one body row per kept record
__device__ __forceinline__ uint8_t* row_ptr(const tk_result_t* res, const tk_plane_t& p, uint32_t kidx) {
    return (uint8_t*)p.base + ((uint64_t)res->batch_size + kidx) * p.row_bytes;
}

// One thread per 16-byte chunk of a record; records of `record_bytes` each.  32-bit index math
// (a slot holds < 2^28 chunks): a 64-bit division per thread would dominate this copy kernel.
__global__ void __launch_bounds__(256) k_collate_raw(const uint8_t* __restrict__ bytes,
                                                      const tk_rec_t* __restrict__ recs, uint32_t n,
                                                      uint32_t record_bytes, uint32_t chunks_per_rec,
                                                      const int32_t* __restrict__ rec_kidx,
                                                      const int32_t* __restrict__ rec_status,
                                                      tk_result_t* res, tk_plane_t out) {
    tk_pdl_wait();
    tk_pdl_launch_dependents();             // after the wait: see k_commit_final (tk_plan.cu)
    const uint32_t total = n * chunks_per_rec;
    for (uint32_t x = blockIdx.x * blockDim.x + threadIdx.x; x < total; x += gridDim.x * blockDim.x) {
        const uint32_t i = x / chunks_per_rec;
        const uint32_t ck = x - i * chunks_per_rec;
        const int32_t kidx = rec_kidx[i];
        if (kidx < 0) {
            if (ck == 0) raw_report_error(res, i, rec_status[i]);
            continue;
        }
        const uint8_t* src = bytes + recs[i].start + ck * 16;
        uint8_t* dst = row_ptr(res, out, (uint32_t)kidx) + ck * 16;
        const uint32_t left = record_bytes - ck * 16;
        if (left >= 16 && ((((uintptr_t)src) | ((uintptr_t)dst)) & 15u) == 0) {
            *reinterpret_cast<uint4*>(dst) = __ldg(reinterpret_cast<const uint4*>(src));
        } else {
            const uint32_t m = left < 16 ? left : 16u;
            for (uint32_t b = 0; b < m; ++b) dst[b] = src[b];
        }
    }
}

extern "C" int tk_collate_raw(tk_ctx* ctx, int slot, uint64_t record_bytes, const tk_plane_t* out, void* stream) {
    TK_ENTER(ctx, slot);
    tk_slot& s = ctx->slots[slot];
    cudaStream_t st = (cudaStream_t)stream;
    TK_REQUIRE(out && out->base, "output plane is required");
    TK_REQUIRE(record_bytes > 0 && out->row_bytes == record_bytes, "row_bytes must equal record_bytes");
    const uint32_t n = s.n_records;
    if (n) {
        const uint64_t cpr = (record_bytes + 15) / 16;
        const uint64_t total = (uint64_t)n * cpr;
        TK_REQUIRE(total < (1ull << 32) && record_bytes < (1ull << 32), "slot too large for 32-bit chunk indexing");
        const uint64_t want = (total + 255) / 256;
        const uint32_t blocks = (uint32_t)(want < 148ull * 32 ? want : 148ull * 32);
        TK_LAUNCH_PDL(k_collate_raw, blocks, 256, st, s.d.bytes, s.d.recs, n, (uint32_t)record_bytes, (uint32_t)cpr,
                      s.d.rec_kidx, s.d.rec_status, s.d.result, *out);
    }
    return 0;
}

// One field of fixed-layout struct records: thread per 4-byte word of the destination row (aligned
// store; the source is read as two aligned words and funnel-shifted: records and fields sit at
// arbitrary byte offsets), thread per byte when the field is not a multiple of 4 bytes.
__global__ void __launch_bounds__(256) k_collate_field(const uint8_t* __restrict__ bytes,
                                                        const tk_rec_t* __restrict__ recs, uint32_t n,
                                                        uint32_t field_offset, uint32_t field_bytes, uint32_t units_per_rec,
                                                        uint32_t unit_bytes, const int32_t* __restrict__ rec_kidx,
                                                        const int32_t* __restrict__ rec_status, int report,
                                                        tk_result_t* res, tk_plane_t out) {
    tk_pdl_wait();
    tk_pdl_launch_dependents();             // after the wait: see k_commit_final (tk_plan.cu)
    const uint32_t total = n * units_per_rec;
    for (uint32_t x = blockIdx.x * blockDim.x + threadIdx.x; x < total; x += gridDim.x * blockDim.x) {
        const uint32_t i = x / units_per_rec;
        const uint32_t u = x - i * units_per_rec;
        const int32_t kidx = rec_kidx[i];
        if (kidx < 0) {
            if (u == 0 && report) raw_report_error(res, i, rec_status[i]);
            continue;
        }
        const uint8_t* src = bytes + recs[i].start + field_offset + u * unit_bytes;
        uint8_t* dst = row_ptr(res, out, (uint32_t)kidx) + u * unit_bytes;
        if (unit_bytes == 4u) {
            const uint32_t sh = ((uint32_t)(uintptr_t)src & 3u) * 8u;
            const uint32_t* s32 = reinterpret_cast<const uint32_t*>(src - ((uintptr_t)src & 3u));
            *reinterpret_cast<uint32_t*>(dst) = __funnelshift_r(__ldg(s32), __ldg(s32 + 1), sh);
        } else {
            *dst = *src;
        }
    }
}

extern "C" int tk_collate_field(tk_ctx* ctx, int slot, uint64_t record_bytes, uint32_t field_offset, uint32_t field_bytes,
                                const tk_plane_t* out, void* stream) {
    TK_ENTER(ctx, slot);
    tk_slot& s = ctx->slots[slot];
    TK_REQUIRE(out && out->base, "output plane is required");
    TK_REQUIRE(field_bytes > 0 && (uint64_t)field_offset + field_bytes <= record_bytes, "field outside the record");
    TK_REQUIRE(out->row_bytes == field_bytes, "row_bytes must equal field_bytes");
    const uint32_t n = s.n_records;
    if (n) {
        const bool words = (field_bytes & 3u) == 0 && ((uintptr_t)out->base & 3u) == 0;
        const uint32_t unit = words ? 4u : 1u, upr = field_bytes / unit;
        const uint64_t total = (uint64_t)n * upr;
        TK_REQUIRE(total < (1ull << 32), "slot too large for 32-bit indexing");
        const uint64_t want = (total + 255) / 256;
        const uint32_t blocks = (uint32_t)(want < 148ull * 32 ? want : 148ull * 32);
        // (records of the wrong size are reported once, by the field at offset 0)
        TK_LAUNCH_PDL(k_collate_field, blocks, 256, (cudaStream_t)stream, s.d.bytes, s.d.recs, n, field_offset,
                      field_bytes, upr, unit, s.d.rec_kidx, s.d.rec_status, field_offset == 0 ? 1 : 0, s.d.result, *out);
    }
    return 0;
}

// ---------------------------------------------------------------------- images
// The normalisation has only 256 x C distinct inputs, so it is evaluated once,
// exactly (IEEE float32 division and subtraction, round-to-nearest-even bf16),
// into a table; the collate kernel is then a byte gather + 2-byte scatter.
__global__ void k_image_lut(uint16_t* __restrict__ lut, int channels, float m0, float m1, float m2, float m3,
                            float s0, float s1, float s2, float s3) {
    const int x = threadIdx.x;       // 0..255
    const float mean[4] = {m0, m1, m2, m3}, stdv[4] = {s0, s1, s2, s3};
    for (int c = 0; c < channels; ++c) {
        const float v = __fdiv_rn(__fsub_rn(__fdiv_rn((float)x, 255.0f), mean[c]), stdv[c]);
        lut[c * 256 + x] = __bfloat16_as_ushort(__float2bfloat16_rn(v));
    }
}

// Fast form of the same normalisation: one fused multiply-add per element,
//     y = fma(float(x), a, b),  a = 1 / (255 * std),  b = -mean / std,
// then round to bfloat16.  It is only used when it reproduces the exact table (two IEEE
// divisions, one subtraction) for all 256 inputs of every channel -- checked on the device,
// with the very same instructions, whenever the constants are set (true for the ImageNet
// mean/std); otherwise the table kernel runs.
__device__ __forceinline__ float norm_fma(uint32_t x, float a, float b) { return __fmaf_rn((float)x, a, b); }

__global__ void k_image_check_fma(const uint16_t* __restrict__ lut, int channels, float a0, float a1, float a2,
                                  float a3, float b0, float b1, float b2, float b3, int* __restrict__ mismatches) {
    const int x = threadIdx.x;
    const float a[4] = {a0, a1, a2, a3}, b[4] = {b0, b1, b2, b3};
    int bad = 0;
    for (int c = 0; c < channels; ++c)
        bad += __bfloat16_as_ushort(__float2bfloat16_rn(norm_fma((uint32_t)x, a[c], b[c]))) != lut[c * 256 + x];
    if (bad) atomicAdd(mismatches, bad);
}

extern "C" int tk_image_set_norm(tk_ctx* ctx, int channels, const float* mean, const float* stdv, void* stream) {
    TK_REQUIRE(ctx, "null context");
    TK_CUDA(cudaSetDevice(ctx->device));
    TK_REQUIRE(channels >= 1 && channels <= 4 && mean && stdv, "1..4 channels with mean and std");
    cudaStream_t st = (cudaStream_t)stream;
    float m[4] = {0, 0, 0, 0}, s[4] = {1, 1, 1, 1};
    for (int c = 0; c < channels; ++c) { m[c] = mean[c]; s[c] = stdv[c]; }
    k_image_lut<<<1, 256, 0, st>>>(ctx->d_image_lut, channels, m[0], m[1], m[2], m[3], s[0], s[1], s[2], s[3]);
    TK_LAUNCH_CHECK();
    // candidate constants of the one-FMA form, then the exhaustive check against the table
    for (int c = 0; c < 4; ++c) {
        ctx->image_a[c] = (float)(1.0 / (255.0 * (double)s[c]));
        ctx->image_b[c] = (float)(-(double)m[c] / (double)s[c]);
    }
    int* d_bad = reinterpret_cast<int*>(ctx->d_image_lut + 4 * 256);     // spare word behind the table
    TK_CUDA(cudaMemsetAsync(d_bad, 0, sizeof(int), st));
    k_image_check_fma<<<1, 256, 0, st>>>(ctx->d_image_lut, channels, ctx->image_a[0], ctx->image_a[1],
                                         ctx->image_a[2], ctx->image_a[3], ctx->image_b[0], ctx->image_b[1],
                                         ctx->image_b[2], ctx->image_b[3], d_bad);
    TK_LAUNCH_CHECK();
    int bad = 1;
    TK_CUDA(cudaMemcpyAsync(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, st));
    TK_CUDA(cudaStreamSynchronize(st));
    ctx->image_fma = bad == 0;
    ctx->image_channels = channels;
    return 0;
}

#define IMG_PIX 4096           // pixels per tile
#define IMG_THREADS 256
#define IMG_LUT_GROUPS 8       // lane groups with private shared-memory banks (4 banks each)

// Persistent blocks loop over (image, tile) pairs.  Per tile: coalesced 16-byte loads into shared
// memory; then every thread takes 8 consecutive pixels, reads their 8*C bytes as C 8-byte words,
// looks every byte up in the bf16 table and issues one 16-byte store per channel plane.
// The table lives in shared memory in IMG_LUT_GROUPS copies laid out so that lanes of different
// groups (lane & 7) can never hit the same bank: entry x of channel c sits in row c*32 + x/8,
// bank 4*group + (x/2 & 3); random pixel values then conflict only among the 4 lanes of a group.
template <int C>
__global__ void __launch_bounds__(IMG_THREADS) k_collate_image(const uint8_t* __restrict__ bytes,
                                                               const tk_rec_t* __restrict__ recs, uint32_t n,
                                                               uint32_t npix, uint32_t tiles_per_img,
                                                               const int32_t* __restrict__ rec_kidx,
                                                               const int32_t* __restrict__ rec_status,
                                                               const uint16_t* __restrict__ lut_g,
                                                               tk_result_t* res, tk_plane_t out) {
    __shared__ __align__(16) uint8_t s_in[IMG_PIX * C];
    __shared__ uint32_t s_lut[C * 32 * 32];
    tk_pdl_wait();
    tk_pdl_launch_dependents();             // after the wait: see k_commit_final (tk_plan.cu)
    // table: word (c*32 + row)*32 + 4*g + q holds entries x = 8*row + 2*q (low half) and x + 1 (high half)
    for (uint32_t w = threadIdx.x; w < C * 32 * 32; w += IMG_THREADS) {
        const uint32_t col = w & 31u, row = (w >> 5) & 31u, c = w >> 10;
        const uint32_t x = 8 * row + 2 * (col & 3u);
        s_lut[w] = (uint32_t)lut_g[c * 256 + x] | ((uint32_t)lut_g[c * 256 + x + 1] << 16);
    }
    const uint32_t group_col = 4u * (threadIdx.x & (IMG_LUT_GROUPS - 1));
    const uint32_t total_tiles = n * tiles_per_img;
    for (uint32_t job = blockIdx.x; job < total_tiles; job += gridDim.x) {
        const uint32_t i = job / tiles_per_img;
        const uint32_t tile = job - i * tiles_per_img;
        const int32_t kidx = rec_kidx[i];
        if (kidx < 0) {                                    // block-uniform
            if (tile == 0 && threadIdx.x == 0) raw_report_error(res, i, rec_status[i]);
            continue;
        }
        const uint32_t p0 = tile * IMG_PIX;
        const uint32_t np = min((uint32_t)IMG_PIX, npix - p0);
        const uint8_t* src = bytes + recs[i].start + (uint64_t)p0 * C;
        const uint32_t nbytes = np * C;
        __syncthreads();                                   // previous tile fully consumed (and table ready)
        if ((((uintptr_t)src) & 15u) == 0) {
            const uint32_t nv = nbytes / 16;
            for (uint32_t x = threadIdx.x; x < nv; x += IMG_THREADS)
                reinterpret_cast<uint4*>(s_in)[x] = __ldg(reinterpret_cast<const uint4*>(src) + x);
            for (uint32_t x = nv * 16 + threadIdx.x; x < nbytes; x += IMG_THREADS) s_in[x] = src[x];
        } else {
            for (uint32_t x = threadIdx.x; x < nbytes; x += IMG_THREADS) s_in[x] = src[x];
        }
        __syncthreads();
        uint16_t* dst_img = reinterpret_cast<uint16_t*>(row_ptr(res, out, (uint32_t)kidx));
        const uint32_t groups = (np + 7) / 8;
        for (uint32_t g = threadIdx.x; g < groups; g += IMG_THREADS) {
            const uint32_t px = g * 8;
            if (px + 8 <= np && ((((uintptr_t)(dst_img + p0 + px)) | ((uint64_t)npix * 2)) & 15u) == 0) {
                // 8 pixels x C channels = 8*C bytes, word aligned in shared memory
                uint32_t in[2 * C];
                const uint2* sp = reinterpret_cast<const uint2*>(s_in + px * C);
#pragma unroll
                for (int k = 0; k < C; ++k) { const uint2 t = sp[k]; in[2 * k] = t.x; in[2 * k + 1] = t.y; }
#pragma unroll
                for (int c = 0; c < C; ++c) {
                    uint32_t w[4];
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const int bidx = k * C + c;                          // byte index inside in[]
                        const uint32_t x = (in[bidx >> 2] >> (8 * (bidx & 3))) & 0xFFu;
                        const uint32_t e = s_lut[(c * 32 + (x >> 3)) * 32 + group_col + ((x >> 1) & 3u)];
                        const uint32_t v = (x & 1u) ? (e >> 16) : (e & 0xFFFFu);
                        if (k & 1) w[k >> 1] |= v << 16; else w[k >> 1] = v;
                    }
                    *reinterpret_cast<uint4*>(dst_img + (uint64_t)c * npix + p0 + px) = make_uint4(w[0], w[1], w[2], w[3]);
                }
            } else {
                for (int c = 0; c < C; ++c)
                    for (uint32_t k = 0; k < 8 && px + k < np; ++k) {
                        const uint32_t x = s_in[(px + k) * C + c];
                        const uint32_t e = s_lut[(c * 32 + (x >> 3)) * 32 + group_col + ((x >> 1) & 3u)];
                        dst_img[(uint64_t)c * npix + p0 + px + k] = (uint16_t)((x & 1u) ? (e >> 16) : (e & 0xFFFFu));
                    }
            }
        }
    }
}

// One-FMA variant (see norm_fma): no table, ~3 instructions per element, so the kernel is a
// pure HBM stream: 16-byte loads -> shared memory (HWC), 8 pixels x C channels per thread,
// one 16-byte store per channel plane (CHW).
struct img_consts { float a[4], b[4]; };

template <int C>
__global__ void __launch_bounds__(IMG_THREADS) k_collate_image_fma(const uint8_t* __restrict__ bytes,
                                                                   const tk_rec_t* __restrict__ recs, uint32_t n,
                                                                   uint32_t npix, uint32_t tiles_per_img,
                                                                   const int32_t* __restrict__ rec_kidx,
                                                                   const int32_t* __restrict__ rec_status,
                                                                   const img_consts k,
                                                                   tk_result_t* res, tk_plane_t out) {
    __shared__ __align__(16) uint8_t s_in[IMG_PIX * C];
    tk_pdl_wait();
    tk_pdl_launch_dependents();             // after the wait: see k_commit_final (tk_plan.cu)
    const uint32_t total_tiles = n * tiles_per_img;
    for (uint32_t job = blockIdx.x; job < total_tiles; job += gridDim.x) {
        const uint32_t i = job / tiles_per_img;
        const uint32_t tile = job - i * tiles_per_img;
        const int32_t kidx = rec_kidx[i];
        if (kidx < 0) {                                    // block-uniform
            if (tile == 0 && threadIdx.x == 0) raw_report_error(res, i, rec_status[i]);
            continue;
        }
        const uint32_t p0 = tile * IMG_PIX;
        const uint32_t np = min((uint32_t)IMG_PIX, npix - p0);
        const uint8_t* src = bytes + recs[i].start + (uint64_t)p0 * C;
        const uint32_t nbytes = np * C;
        __syncthreads();                                   // previous tile fully consumed
        if ((((uintptr_t)src) & 15u) == 0) {
            const uint32_t nv = nbytes / 16;
            for (uint32_t x = threadIdx.x; x < nv; x += IMG_THREADS)
                reinterpret_cast<uint4*>(s_in)[x] = __ldg(reinterpret_cast<const uint4*>(src) + x);
            for (uint32_t x = nv * 16 + threadIdx.x; x < nbytes; x += IMG_THREADS) s_in[x] = src[x];
        } else {
            for (uint32_t x = threadIdx.x; x < nbytes; x += IMG_THREADS) s_in[x] = src[x];
        }
        __syncthreads();
        __nv_bfloat16* dst_img = reinterpret_cast<__nv_bfloat16*>(row_ptr(res, out, (uint32_t)kidx));
        const uint32_t groups = (np + 7) / 8;
        for (uint32_t g = threadIdx.x; g < groups; g += IMG_THREADS) {
            const uint32_t px = g * 8;
            if (px + 8 <= np && ((((uintptr_t)(dst_img + p0 + px)) | ((uint64_t)npix * 2)) & 15u) == 0) {
                uint32_t in[2 * C];
                const uint2* sp = reinterpret_cast<const uint2*>(s_in + px * C);
#pragma unroll
                for (int q = 0; q < C; ++q) { const uint2 t = sp[q]; in[2 * q] = t.x; in[2 * q + 1] = t.y; }
#pragma unroll
                for (int c = 0; c < C; ++c) {
                    uint32_t w[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int b0 = (2 * q) * C + c, b1 = (2 * q + 1) * C + c;      // byte indices inside in[]
                        const float y0 = norm_fma((in[b0 >> 2] >> (8 * (b0 & 3))) & 0xFFu, k.a[c], k.b[c]);
                        const float y1 = norm_fma((in[b1 >> 2] >> (8 * (b1 & 3))) & 0xFFu, k.a[c], k.b[c]);
                        const __nv_bfloat162 pr = __floats2bfloat162_rn(y0, y1);
                        w[q] = *reinterpret_cast<const uint32_t*>(&pr);
                    }
                    *reinterpret_cast<uint4*>(dst_img + (uint64_t)c * npix + p0 + px) = make_uint4(w[0], w[1], w[2], w[3]);
                }
            } else {
                for (int c = 0; c < C; ++c)
                    for (uint32_t q = 0; q < 8 && px + q < np; ++q)
                        dst_img[(uint64_t)c * npix + p0 + px + q] =
                            __float2bfloat16_rn(norm_fma(s_in[(px + q) * C + c], k.a[c], k.b[c]));
            }
        }
    }
}

extern "C" int tk_collate_image_u8_bf16_nchw(tk_ctx* ctx, int slot, uint32_t height, uint32_t width,
                                             uint32_t channels, const tk_plane_t* out, void* stream) {
    TK_ENTER(ctx, slot);
    tk_slot& s = ctx->slots[slot];
    cudaStream_t st = (cudaStream_t)stream;
    TK_REQUIRE(out && out->base, "output plane is required");
    TK_REQUIRE((int)channels == ctx->image_channels, "call tk_image_set_norm for %u channels first", channels);
    const uint32_t npix = height * width;
    TK_REQUIRE(npix > 0 && out->row_bytes == (uint64_t)npix * channels * 2, "row_bytes must be H*W*C*2");
    const uint32_t n = s.n_records;
    if (n) {
        const uint32_t tiles = (npix + IMG_PIX - 1) / IMG_PIX;
        const uint32_t blocks = min(n * tiles, 148u * 6u);          // persistent: 6 blocks per SM
        img_consts kc;
        for (int c = 0; c < 4; ++c) { kc.a[c] = ctx->image_a[c]; kc.b[c] = ctx->image_b[c]; }
#define TK_IMG_LAUNCH(CC)                                                                                   \
    if (ctx->image_fma)                                                                                     \
        TK_LAUNCH_PDL(k_collate_image_fma<CC>, blocks, IMG_THREADS, st, s.d.bytes, s.d.recs, n, npix, tiles, \
                      s.d.rec_kidx, s.d.rec_status, kc, s.d.result, *out);                                  \
    else                                                                                                    \
        TK_LAUNCH_PDL(k_collate_image<CC>, blocks, IMG_THREADS, st, s.d.bytes, s.d.recs, n, npix, tiles,    \
                      s.d.rec_kidx, s.d.rec_status, (const uint16_t*)ctx->d_image_lut, s.d.result, *out)
        switch (channels) {
            case 1: TK_IMG_LAUNCH(1); break;
            case 2: TK_IMG_LAUNCH(2); break;
            case 3: TK_IMG_LAUNCH(3); break;
            case 4: TK_IMG_LAUNCH(4); break;
            default: TK_REQUIRE(false, "1..4 channels supported");
        }
#undef TK_IMG_LAUNCH
    }
    return 0;
}
