"""World-size-2 gloo test of the partition sharding and the commit barrier (host logic of the
multi-GPU path; runs on CPU)."""
import json
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from golden_cases import CASES, build_log
from oracle import reference_path as ref
from torchkafka_b200 import distributed as tkd
from torchkafka_b200.testing import stub_kafka as sk


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    case = CASES["cfg2_varlen_filter"]
    log = build_log(case).select_partitions(lambda p: tkd.owns(p, rank, world))
    consumer = sk.KafkaConsumer("topic", log=log)
    process = ref.process_json(torch.float32, case["min_size"])
    results = []
    for _batch, commit in ref.run(consumer, process, case["batch_size"], False, ref.pad_collate):
        # every rank reaches the barrier once per local batch; ranks have different batch counts,
        # so the barrier is only exercised for the common prefix and once more at the end
        results.append(commit)
    final_local = results[-1] if results else {}
    merged = tkd.global_commit(final_local, case["partitions"])
    with open(os.path.join(out_dir, f"rank{rank}.json"), "w") as f:
        json.dump({"local": {str(k): v for k, v in final_local.items()},
                   "global": {str(k): v for k, v in merged.items()},
                   "owned": tkd.owned_partitions(case["partitions"], rank, world)}, f)
    dist.barrier()
    dist.destroy_process_group()


def test_partition_sharding_and_commit_barrier_gloo(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    outs = [json.load(open(tmp_path / f"rank{r}.json")) for r in range(world)]
    # disjoint ownership covering every partition
    owned = [set(o["owned"]) for o in outs]
    assert owned[0].isdisjoint(owned[1]) and owned[0] | owned[1] == set(range(64))
    for r, o in enumerate(outs):
        assert all(int(p) % world == r for p in o["local"])
    # both ranks hold the same global vector = union of the local ones
    assert outs[0]["global"] == outs[1]["global"]
    union = {}
    for o in outs:
        union.update(o["local"])
    assert outs[0]["global"] == union
    # and it equals what one process consuming every partition would have committed at the end
    case = CASES["cfg2_varlen_filter"]
    full = build_log(case)
    expect = ref.positions_after(full.partitions, full.offsets, len(full) - 1)
    assert {int(k): v for k, v in outs[0]["global"].items()} == expect


def test_commit_barrier_without_process_group_is_identity():
    v = torch.tensor([3, -1, 7], dtype=torch.int64)
    assert torch.equal(tkd.commit_barrier(v), v)
    assert tkd.owned_partitions(8, 1, 4) == [1, 5]


def _barrier_worker(rank, world, port, out_dir):
    """Two ranks with different batch counts drive auto_commit's per-batch barrier object."""
    import collections
    from torchkafka_b200.auto_commit import _CommitBarrier
    from torchkafka_b200.testing.stub_kafka import OffsetAndMetadata, TopicPartition
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)

    class FakeDataset:
        max_partitions = 8
        device_index = None

    ds = FakeDataset()
    barrier = _CommitBarrier(ds, True)
    seen = []
    n_batches = 3 if rank == 0 else 5                     # rank 0 finishes first and keeps answering
    for k in range(n_batches):
        barrier.note({TopicPartition("t", rank + 2 * (k % 2)): OffsetAndMetadata(10 * (k + 1) + rank, "")})
        assert barrier.exchange(False) is False
        seen.append(ds.group_offsets.tolist())
    rounds = 0
    while True:
        done = barrier.exchange(True)
        seen.append(ds.group_offsets.tolist())
        if done:
            break
        rounds += 1
    with open(os.path.join(out_dir, f"b{rank}.json"), "w") as f:
        json.dump({"seen": seen, "rounds": rounds, "lat": len(ds.commit_barrier_us)}, f)
    dist.barrier()
    dist.destroy_process_group()


def test_auto_commit_barrier_object_with_uneven_ranks(tmp_path):
    world = 2
    mp.spawn(_barrier_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    a, b = [json.load(open(tmp_path / f"b{r}.json")) for r in range(world)]
    # both ranks took part in the same number of exchanges and saw the same vectors in each
    assert a["lat"] == b["lat"] == 6
    assert a["seen"] == b["seen"]
    assert a["rounds"] == 2 and b["rounds"] == 0           # rank 0 answered rank 1's last two batches
    # step 0: rank 0 reported partition 0 -> 10, rank 1 partition 1 -> 11; others unseen
    assert a["seen"][0] == [10, 11, -1, -1, -1, -1, -1, -1]
    # final vector: the last offset every rank reported per partition
    assert a["seen"][-1] == [30, 51, 20, 41, -1, -1, -1, -1]
