// tk_process.cu -- one call that enqueues the whole kernel sequence of a slot.
//
// This is the entry point a host-side plugin binds for the hot path: it stands for one
// pass of `KafkaDataset.__iter__ -> _process -> None filter -> collate -> commit offsets`
// (src/kafka_dataset.py:147-171, fetch.py:27-45, collate.py:246-275, auto_commit.py:55-58)
// over every record of the slot.  Everything is asynchronous on `stream`.
#include "tk_internal.h"

extern "C" int tk_process_slot(tk_ctx* ctx, int slot, const tk_job_t* job, void* stream) {
    TK_ENTER(ctx, slot);
    TK_REQUIRE(job, "job is null");
    int rc;
    if (!job->resident) { rc = tk_slot_wait(ctx, slot, stream); if (rc) return rc; }
    if (job->prev_slot >= 0) {
        rc = tk_carry_rows(ctx, job->prev_slot, job->n_planes, job->prev, job->cur, stream);
        if (rc) return rc;
    }
    // JSON arrays: the three-kernel chain (scan, plan, parse), or -- tk_ctx_set_json_fused /
    // TK_JSON_FUSED=1 -- the single-pass kernel of tk_json_fused.cu, which reads the payload from HBM
    // once but is the slower of the two today
    const bool fused = job->schema == TK_SCHEMA_JSON && ctx->json_fused;
    tk_slot& s = ctx->slots[slot];
    if (fused) {
        s.dec_timed = false;
        if (ctx->timing) TK_CUDA(cudaEventRecord(s.ev_dec0, (cudaStream_t)stream));
        rc = tk_json_ingest(ctx, slot, job->dtype, job->max_len, job->min_size, job->fixed_len, job->batch_size,
                            &job->cur[0], job->n_planes > 1 ? &job->cur[1] : nullptr,
                            job->n_planes > 2 ? &job->cur[2] : nullptr, stream);
        if (rc) return rc;
        if (ctx->timing) { TK_CUDA(cudaEventRecord(s.ev_dec1, (cudaStream_t)stream)); s.dec_timed = true; }
        // the per-(batch, partition) maxima are produced inside the single-pass kernel: the running
        // max has to wait for all of it
        rc = tk_commit_final(ctx, slot, 1, stream);
        if (rc) return rc;
        return tk_slot_fetch_result(ctx, slot, stream);
    }
    // nested arrays of two or more dimensions go through the flat chain too: every comma separates two
    // leaves, the brackets around a leaf follow from its index (tk_json.cuh); the object kernel's
    // sequential scanner (tk_json_nested_decode) stays the staged entry point and the cross-check
    const bool nested_flat = job->schema == TK_SCHEMA_NESTED && job->max_len >= 2;
    if (job->schema == TK_SCHEMA_JSON) rc = tk_json_scan(ctx, slot, job->max_len, job->min_size, stream);
    else if (nested_flat) rc = tk_json_nested_scan(ctx, slot, job->max_len, job->field_bytes, stream);
    else rc = tk_raw_scan(ctx, slot, (job->schema == TK_SCHEMA_OBJECT || job->schema == TK_SCHEMA_NESTED) ? TK_ANY_RECORD_BYTES : job->record_bytes, stream);
    if (rc) return rc;
    rc = tk_compact(ctx, slot, job->batch_size, stream); if (rc) return rc;
    s.dec_timed = false;
    if (ctx->timing) TK_CUDA(cudaEventRecord(s.ev_dec0, (cudaStream_t)stream));
    switch (job->schema) {
        case TK_SCHEMA_JSON:
            rc = tk_json_decode(ctx, slot, job->dtype, job->max_len, job->fixed_len, &job->cur[0],
                                job->n_planes > 1 ? &job->cur[1] : nullptr,
                                job->n_planes > 2 ? &job->cur[2] : nullptr, stream);
            break;
        case TK_SCHEMA_RAW:
            rc = tk_collate_raw(ctx, slot, job->record_bytes, &job->cur[0], stream);
            break;
        case TK_SCHEMA_STRUCT:
            TK_REQUIRE(job->n_planes >= 1 && job->n_planes <= TK_MAX_PLANES, "1..%d struct fields", TK_MAX_PLANES);
            for (int k = 0; k < job->n_planes && !rc; ++k)
                rc = tk_collate_field(ctx, slot, job->record_bytes, job->field_offset[k], job->field_bytes[k],
                                      &job->cur[k], stream);
            break;
        case TK_SCHEMA_OBJECT:
            rc = tk_json_object_decode(ctx, slot, job->dtype, &job->cur[0], stream);
            break;
        case TK_SCHEMA_NESTED:
            rc = nested_flat ? tk_json_nested_decode_flat(ctx, slot, job->dtype, job->max_len, job->field_bytes, &job->cur[0], stream)
                             : tk_json_nested_decode(ctx, slot, job->dtype, job->max_len, job->field_bytes, &job->cur[0], stream);
            break;
        case TK_SCHEMA_IMAGE:
            rc = tk_collate_image_u8_bf16_nchw(ctx, slot, job->height, job->width, job->channels,
                                               &job->cur[0], stream);
            break;
        default:
            TK_REQUIRE(false, "unknown schema %d", job->schema);
    }
    if (rc) return rc;
    if (ctx->timing) { TK_CUDA(cudaEventRecord(s.ev_dec1, (cudaStream_t)stream)); s.dec_timed = true; }
    // commit offsets, second half: next to the decode kernel, not in front of it (tk_plan.cu).  With
    // no records there was no decode kernel to order it behind the plan, and with timing events in
    // between the chain is cut anyway: wait then.
    rc = tk_commit_final(ctx, slot, (s.n_records == 0 || ctx->timing) ? 1 : 0, stream);
    if (rc) return rc;
    return tk_slot_fetch_result(ctx, slot, stream);
}

extern "C" int tk_ctx_set_json_fused(tk_ctx* ctx, int enabled) {
    TK_REQUIRE(ctx, "null context");
    ctx->json_fused = enabled ? 1 : 0;
    return 0;
}

extern "C" int tk_ctx_set_json_wide(tk_ctx* ctx, int enabled) {
    TK_REQUIRE(ctx, "null context");
    ctx->json_wide = enabled ? 1 : 0;
    return 0;
}

extern "C" int tk_ctx_set_timing(tk_ctx* ctx, int enabled) {
    TK_REQUIRE(ctx, "null context");
    ctx->timing = enabled ? 1 : 0;
    return 0;
}

extern "C" int tk_slot_decode_ms(tk_ctx* ctx, int slot, float* ms) {
    TK_ENTER(ctx, slot);
    TK_REQUIRE(ms, "ms is null");
    tk_slot& s = ctx->slots[slot];
    TK_REQUIRE(s.dec_timed, "slot %d was not processed with timing enabled", slot);
    TK_CUDA(cudaEventSynchronize(s.ev_dec1));
    TK_CUDA(cudaEventElapsedTime(ms, s.ev_dec0, s.ev_dec1));
    return 0;
}
