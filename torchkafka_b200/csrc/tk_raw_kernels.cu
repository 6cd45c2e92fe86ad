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
    atomicAdd(&res->n_errors, 1u);
    const uint32_t prev = atomicMin(&res->first_error_rec, rec);
    if (rec < prev) res->first_error_status = (uint32_t)status;
}

// every record is kept; a record of the wrong size is an error
__global__ void __launch_bounds__(256) k_raw_scan(const tk_rec_t* __restrict__ recs, uint32_t n,
                                                   uint64_t record_bytes, int32_t* __restrict__ rec_len,
                                                   int32_t* __restrict__ rec_status, int32_t* __restrict__ rec_kidx) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t len = recs[i + 1].start - recs[i].start;
    rec_len[i] = (int32_t)len;
    const bool ok = len == record_bytes;
    rec_status[i] = ok ? TK_STATUS_OK : TK_STATUS_SIZE;
    rec_kidx[i] = ok ? 0 : -1;
}

extern "C" int tk_raw_scan(tk_ctx* ctx, int slot, uint64_t record_bytes, void* stream) {
    TK_ENTER(ctx, slot);
    tk_slot& s = ctx->slots[slot];
    const uint32_t n = s.n_records;
    if (!n) return 0;
    k_raw_scan<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(s.d.recs, n, record_bytes, s.d.rec_len,
                                                                  s.d.rec_status, s.d.rec_kidx);
    TK_LAUNCH_CHECK();
    return 0;
}

// destination row of kept record `kidx`: head rows first, then one body row per kept record
__device__ __forceinline__ uint8_t* row_ptr(const tk_result_t* res, const tk_plane_t& p, uint32_t kidx) {
    return (uint8_t*)p.base + ((uint64_t)res->batch_size + kidx) * p.row_bytes;
}

// One thread per 16-byte chunk of a record; records of `record_bytes` each.
__global__ void __launch_bounds__(256) k_collate_raw(const uint8_t* __restrict__ bytes,
                                                      const tk_rec_t* __restrict__ recs, uint32_t n,
                                                      uint64_t record_bytes, const int32_t* __restrict__ rec_kidx,
                                                      const int32_t* __restrict__ rec_status,
                                                      tk_result_t* res, tk_plane_t out) {
    const uint64_t chunks_per_rec = (record_bytes + 15) / 16;
    const uint64_t total = (uint64_t)n * chunks_per_rec;
    for (uint64_t x = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; x < total;
         x += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t i = (uint32_t)(x / chunks_per_rec);
        const uint64_t ck = x - (uint64_t)i * chunks_per_rec;
        const int32_t kidx = rec_kidx[i];
        if (kidx < 0) {
            if (ck == 0) raw_report_error(res, i, rec_status[i]);
            continue;
        }
        const uint8_t* src = bytes + recs[i].start + ck * 16;
        uint8_t* dst = row_ptr(res, out, (uint32_t)kidx) + ck * 16;
        const uint64_t left = record_bytes - ck * 16;
        if (left >= 16 && ((((uintptr_t)src) | ((uintptr_t)dst)) & 15u) == 0) {
            *reinterpret_cast<uint4*>(dst) = __ldg(reinterpret_cast<const uint4*>(src));
        } else {
            const uint32_t m = left < 16 ? (uint32_t)left : 16u;
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
        const uint64_t total = (uint64_t)n * ((record_bytes + 15) / 16);
        const uint64_t want = (total + 255) / 256;
        const uint32_t blocks = (uint32_t)(want < 148ull * 32 ? want : 148ull * 32);
        k_collate_raw<<<blocks, 256, 0, st>>>(s.d.bytes, s.d.recs, n, record_bytes, s.d.rec_kidx, s.d.rec_status,
                                              s.d.result, *out);
        TK_LAUNCH_CHECK();
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

extern "C" int tk_image_set_norm(tk_ctx* ctx, int channels, const float* mean, const float* stdv, void* stream) {
    TK_REQUIRE(ctx, "null context");
    TK_CUDA(cudaSetDevice(ctx->device));
    TK_REQUIRE(channels >= 1 && channels <= 4 && mean && stdv, "1..4 channels with mean and std");
    float m[4] = {0, 0, 0, 0}, s[4] = {1, 1, 1, 1};
    for (int c = 0; c < channels; ++c) { m[c] = mean[c]; s[c] = stdv[c]; }
    k_image_lut<<<1, 256, 0, (cudaStream_t)stream>>>(ctx->d_image_lut, channels, m[0], m[1], m[2], m[3],
                                                     s[0], s[1], s[2], s[3]);
    TK_LAUNCH_CHECK();
    ctx->image_channels = channels;
    return 0;
}

#define IMG_PIX 4096           // pixels per block tile
#define IMG_THREADS 256

template <int C>
__global__ void __launch_bounds__(IMG_THREADS) k_collate_image(const uint8_t* __restrict__ bytes,
                                                               const tk_rec_t* __restrict__ recs, uint32_t n,
                                                               uint32_t npix, uint32_t tiles_per_img,
                                                               const int32_t* __restrict__ rec_kidx,
                                                               const int32_t* __restrict__ rec_status,
                                                               const uint16_t* __restrict__ lut_g,
                                                               tk_result_t* res, tk_plane_t out) {
    __shared__ __align__(16) uint8_t s_in[IMG_PIX * C];
    __shared__ uint16_t s_lut[C * 256];
    const uint32_t i = blockIdx.x / tiles_per_img;
    const uint32_t tile = blockIdx.x - i * tiles_per_img;
    const int32_t kidx = rec_kidx[i];
    if (kidx < 0) {
        if (tile == 0 && threadIdx.x == 0) raw_report_error(res, i, rec_status[i]);
        return;
    }
    for (int x = threadIdx.x; x < C * 256; x += IMG_THREADS) s_lut[x] = lut_g[x];
    const uint32_t p0 = tile * IMG_PIX;
    const uint32_t np = min((uint32_t)IMG_PIX, npix - p0);
    const uint8_t* src = bytes + recs[i].start + (uint64_t)p0 * C;
    const uint32_t nbytes = np * C;
    if ((((uintptr_t)src) & 15u) == 0) {
        const uint32_t nv = nbytes / 16;
        for (uint32_t x = threadIdx.x; x < nv; x += IMG_THREADS)
            reinterpret_cast<uint4*>(s_in)[x] = __ldg(reinterpret_cast<const uint4*>(src) + x);
        for (uint32_t x = nv * 16 + threadIdx.x; x < nbytes; x += IMG_THREADS) s_in[x] = src[x];
    } else {
        for (uint32_t x = threadIdx.x; x < nbytes; x += IMG_THREADS) s_in[x] = src[x];
    }
    __syncthreads();
    // each thread: 8 consecutive pixels of one channel -> one 16-byte store
    uint16_t* dst_img = reinterpret_cast<uint16_t*>(row_ptr(res, out, (uint32_t)kidx));
    const uint32_t groups = (np + 7) / 8;
    for (uint32_t x = threadIdx.x; x < groups * C; x += IMG_THREADS) {
        const uint32_t c = x / groups, g = x - c * groups;
        const uint32_t px = g * 8;
        uint16_t* d = dst_img + (uint64_t)c * npix + p0 + px;
        const uint16_t* lut = s_lut + c * 256;
        if (px + 8 <= np && (((uintptr_t)d) & 15u) == 0) {
            uint32_t w[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t a = lut[s_in[(px + 2 * k) * C + c]];
                const uint32_t b = lut[s_in[(px + 2 * k + 1) * C + c]];
                w[k] = a | (b << 16);
            }
            *reinterpret_cast<uint4*>(d) = make_uint4(w[0], w[1], w[2], w[3]);
        } else {
            for (uint32_t k = 0; k < 8 && px + k < np; ++k) d[k] = lut[s_in[(px + k) * C + c]];
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
        const uint32_t blocks = n * tiles;
#define TK_IMG_LAUNCH(CC)                                                                              \
    k_collate_image<CC><<<blocks, IMG_THREADS, 0, st>>>(s.d.bytes, s.d.recs, n, npix, tiles, s.d.rec_kidx, \
                                                        s.d.rec_status, ctx->d_image_lut, s.d.result, *out)
        switch (channels) {
            case 1: TK_IMG_LAUNCH(1); break;
            case 2: TK_IMG_LAUNCH(2); break;
            case 3: TK_IMG_LAUNCH(3); break;
            case 4: TK_IMG_LAUNCH(4); break;
            default: TK_REQUIRE(false, "1..4 channels supported");
        }
#undef TK_IMG_LAUNCH
        TK_LAUNCH_CHECK();
    }
    return 0;
}
