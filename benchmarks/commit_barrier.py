#!/usr/bin/env python3
"""Config 5: 64 partitions x 64-byte records, per-batch commit offsets (K4) and the cross-rank
commit barrier.  Latency-bound by design: reports batches/s of the ingest
loop and p50/p99 of the barrier, not GB/s.

    python benchmarks/commit_barrier.py                       # 1 GPU: K4 only
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 \
        benchmarks/commit_barrier.py                          # 8 ranks, 8 partitions each

Every rank owns the partitions p % W == r (torchkafka_b200.distributed), ingests its own 64-byte
records through the public API (RawBytes schema) and, after every batch, takes part in the barrier:
all_gather of its [64] int64 commit vector, element-wise max.  Both transports are timed: host
tensors over gloo and a device tensor over NCCL.
"""
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch                      # noqa: E402
import torch.distributed as dist  # noqa: E402
from torch.utils.data import DataLoader  # noqa: E402

from torchkafka_b200 import KafkaDataset, RawBytes, auto_commit      # noqa: E402
from torchkafka_b200 import distributed as tkd                       # noqa: E402
from torchkafka_b200.testing import stub_kafka as sk                 # noqa: E402
from torchkafka_b200.testing import synthetic as syn                 # noqa: E402

P, BATCH, N_BATCHES = 64, 256, 400


def main():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("cpu:gloo,cuda:nccl")
    full = syn.make_log(5, (N_BATCHES + 8) * BATCH * world, seed=5, partitions=P)
    mine = full.select_partitions(lambda p: tkd.owns(p, rank, world)) if world > 1 else full

    class DS(KafkaDataset):
        payload_schema = RawBytes((64,), torch.uint8)
        device_slot_bytes = 64 * BATCH * 8          # small slots: latency, not throughput
        device_index = local
        max_partitions = P

        @classmethod
        def new_consumer(cls, *a, **k):
            k.pop("enable_auto_commit", None)
            return sk.KafkaConsumer(*a, log=mine, **k)

    ds = DS("topic", group_id="cfg5")
    lat_gloo, lat_nccl = [], []
    seen = 0
    t0 = None
    for batch in auto_commit(DataLoader(ds, batch_size=BATCH)):
        if seen == 8:
            torch.cuda.synchronize()
            t0 = time.perf_counter()
        # the offsets this rank is about to commit (what K4 produced for the batch)
        local_vec = tkd.offsets_vector({tp.partition: om.offset for tp, om in ds._explicit_offsets.items()}, P)
        if world > 1:
            a = time.perf_counter()
            g = tkd.commit_barrier(local_vec)
            b = time.perf_counter()
            dv = local_vec.cuda()
            torch.cuda.synchronize()
            c = time.perf_counter()
            g2 = tkd.commit_barrier(dv)
            torch.cuda.synchronize()
            d = time.perf_counter()
            assert torch.equal(g, g2.cpu())
            if seen >= 8:
                lat_gloo.append((b - a) * 1e6)
                lat_nccl.append((d - c) * 1e6)
        seen += 1
        if seen == 8 + N_BATCHES:
            break
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    ds.close()

    def pct(v, q):
        v = sorted(v)
        return v[min(len(v) - 1, int(q * len(v)))] if v else None

    if rank == 0:
        print(json.dumps({
            "config": 5, "world_size": world, "partitions": P, "batch": BATCH, "batches": N_BATCHES,
            "batches_per_s_per_rank": N_BATCHES / dt, "records_per_s_per_rank": N_BATCHES * BATCH / dt,
            "barrier_us": {"gloo_p50": pct(lat_gloo, 0.5), "gloo_p99": pct(lat_gloo, 0.99),
                           "nccl_p50": pct(lat_nccl, 0.5), "nccl_p99": pct(lat_nccl, 0.99)},
            "note": "loop includes both barrier transports per batch when world_size > 1",
        }), flush=True)
    if world > 1:
        dist.all_reduce(torch.zeros(1))
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
