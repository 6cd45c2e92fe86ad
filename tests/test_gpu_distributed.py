"""Two ranks (gloo) sharing one GPU drive the product pipeline with the commit barrier
(SURVEY.md §8e): rank r ingests the partitions p % 2 == r through
auto_commit(DataLoader(dataset), commit_barrier=True); batches and commits are compared with the
oracle over the rank's own partitions, the group's offset vector with the union of both."""
import json
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from torch.utils.data import DataLoader
    from helpers import assert_same, make_dataset, run_oracle, to_cpu
    from torchkafka_b200 import auto_commit
    from torchkafka_b200 import distributed as tkd
    from torchkafka_b200.schema import JsonArray
    from torchkafka_b200.testing import synthetic as syn
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.cuda.set_device(0)
    schema = JsonArray(torch.float32, max_len=256, min_size=32)
    # rank 1 gets a longer stream: rank 0 finishes first and has to keep answering the barrier
    full = syn.make_log(2, 1500 if rank == 0 else 2600, seed=90 + rank, partitions=8, layout="chunked", chunk=25)
    mine = full.select_partitions(lambda p: tkd.owns(p, rank, world))
    ds = make_dataset(mine, schema, slot_bytes=300_000, device_index=0, max_partitions=8)
    got, vectors = [], []
    gen = auto_commit(DataLoader(ds, batch_size=64), commit_barrier=True)
    pending = None
    for batch in gen:
        if pending is not None:
            got.append((pending, dict(ds._consumer.committed_log[-1])))
            vectors.append(ds.group_offsets.tolist())
        pending = to_cpu(batch)
    got.append((pending, dict(ds._consumer.committed_log[-1])))
    vectors.append(ds.group_offsets.tolist())
    assert_same(got, run_oracle(mine, schema, 64), schema)
    with open(os.path.join(out_dir, f"r{rank}.json"), "w") as f:
        json.dump({"batches": len(got), "final_local": {str(k): v for k, v in got[-1][1].items()},
                   "final_group": vectors[-1], "barriers": len(ds.commit_barrier_us)}, f)
    ds.close()
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_share_a_gpu_and_commit_through_the_barrier(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    a, b = [json.load(open(tmp_path / f"r{r}.json")) for r in range(world)]
    assert a["batches"] != b["batches"]                      # uneven streams: the join protocol ran
    assert a["barriers"] == b["barriers"] >= max(a["batches"], b["batches"])
    # both ranks end with the same group vector = every partition's last committed offset
    assert a["final_group"] == b["final_group"]
    union = {**a["final_local"], **b["final_local"]}
    assert set(int(p) % 2 for p in a["final_local"]) == {0} and set(int(p) % 2 for p in b["final_local"]) == {1}
    assert a["final_group"] == [union.get(str(p), -1) for p in range(8)]
