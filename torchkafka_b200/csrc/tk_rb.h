// tk_rb.h -- shared by tk_recordbatch.cu (RecordBatch v2 split) and tk_codecs.cu (decompression
// of compressed batches): kernel arguments, the decompression work unit, a block scan.
#pragma once
#include "tk_internal.h"

#define RB_HEADER 61u             // baseOffset .. recordCount
#define RB_CRC_FROM 21u           // the checksum covers attributes .. end of batch
#define RB_SRC2 0x80000000u       // copy descriptor: the value lies in the decompressed buffer

// One independently decodable piece of a compressed batch, found by the host's walk over the
// codec's framing (tk_codec_units_host): an LZ4 block of a block-independent frame, the whole
// block chain of a block-dependent frame, a stored (uncompressed) block, a snappy block (the
// whole batch, or one chunk of xerial framing), the DEFLATE stream of a gzip member, the
// Zstandard frames of a batch.  Units of a batch are consecutive and decode
// into consecutive ranges of the decompressed buffer.
enum { RB_UNIT_LZ4_BLOCK = 1, RB_UNIT_LZ4_CHAIN = 2, RB_UNIT_STORED = 3, RB_UNIT_SNAPPY = 4, RB_UNIT_DEFLATE = 5, RB_UNIT_ZSTD = 6 };
#define RB_UNIT_BLOCK_CHECKSUM 0x100u    // LZ4 chain: 4 checksum bytes behind every block
#define RB_UNIT_CONTENT_CHECKSUM 0x200u  // LZ4 chain: 4 checksum bytes behind the end mark
#define RB_UNIT_SIZED 0x400u             // unc_len is known from the framing (no size pass)
#define RB_UNIT_FAILED 0x800u            // the size pass found the stream broken
#define RB_UNIT_BOUND 0x1000u            // unc_len is an upper bound (zstd frames without content size, LZ4 frames of one unit):
                                         // no size pass, the decode sets the batch's record range
struct tk_rb_unit {
    uint32_t src, src_len;     // compressed bytes in the raw buffer
    uint32_t batch;            // index of the batch it belongs to
    uint32_t kind;             // RB_UNIT_* | flags
    uint32_t unc_len;          // decompressed size
    uint32_t out;              // where it goes in the decompressed buffer
};

struct rb_args {
    const uint8_t*   raw;
    uint8_t*         raw2;            // decompressed records of compressed batches
    uint64_t         raw2_bytes;
    tk_rb_batch*     batches;
    uint32_t         n_batches, n_records;    // from the host's header walk
    tk_rb_unit*      units;
    uint32_t         n_units;
    tk_rb_summary_t* summary;
    uint4*           temp;            // [slot_records] per-record copy descriptors (rb_temp)
    uint32_t*        bigq;            // records longer than RB_BIG, copied piecewise by k_rb_copy_big
    uint8_t*         bytes;           // slot byte buffer (destination)
    tk_rec_t*        recs;            // slot record headers (destination)
    uint32_t         slot_records;
    uint64_t         slot_bytes;
};

// host: units of batch `b` appended to units[*n_units...]; sets b.unit_first /
// b.unit_count, or b.status when the framing is broken or not handled (codec 1 .. 4)
int tk_codec_units_host(const uint8_t* raw, tk_rb_batch& b, uint32_t batch_index, tk_rb_unit* units,
                        uint32_t max_units, uint32_t* n_units);
// device: size pass, placement, decode of a.units on `stream`; h_units = the host's copy (which
// kernels are needed)
int tk_codec_launch(const rb_args& a, const tk_rb_unit* h_units, cudaStream_t stream);

#if defined(__CUDACC__)
// exclusive scan of one 64-bit value per thread of a 1024-thread block; *total = block sum
__device__ __forceinline__ unsigned long long rb_block_scan(unsigned long long v, unsigned long long* warp_tot,
                                                            unsigned long long* total) {
    const uint32_t t = threadIdx.x, lane = t & 31;
    unsigned long long incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned long long o = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= (uint32_t)d) incl += o;
    }
    if (lane == 31) warp_tot[t >> 5] = incl;
    __syncthreads();
    if (t < 32) {
        const unsigned long long w = warp_tot[t];
        unsigned long long wi = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned long long o = __shfl_up_sync(0xffffffffu, wi, d);
            if (t >= (uint32_t)d) wi += o;
        }
        warp_tot[t] = wi - w;
        if (t == 31) *total = wi;
    }
    __syncthreads();
    return warp_tot[t >> 5] + incl - v;
}
#endif
