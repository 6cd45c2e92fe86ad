"""Drop-in behaviour of the Python API on the CPU (non-schema) path, checked against facts
recorded from the unmodified reference (tests/golden/api_facts.json)."""
import json
import logging
import os

import numpy as np
import pytest
import torch
from torch.utils.data import DataLoader, TensorDataset

import torchkafka
from golden_cases import CASES, build_log
from torchkafka_b200 import KafkaDataset, auto_commit
from torchkafka_b200.testing import stub_kafka as sk

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
FACTS = json.load(open(os.path.join(GOLD, "api_facts.json")))


def make(log, process=None, consumer_kw=None):
    class DS(KafkaDataset):
        @classmethod
        def new_consumer(cls, *a, **k):
            k.pop("enable_auto_commit", None)
            return sk.KafkaConsumer(*a, log=log, **(consumer_kw or {}), **k)

        def _process(self, record):
            return process(record) if process else torch.tensor(json.loads(record.value))

    return DS


def raised(fn):
    try:
        fn()
    except Exception as exc:                              # noqa: BLE001
        return [type(exc).__name__, str(exc)]
    return None


def test_import_name_and_exports():
    assert torchkafka.KafkaDataset is KafkaDataset and torchkafka.auto_commit is auto_commit


def test_exceptions_match_reference():
    empty = sk.RecordLog([], [], [])
    D = make(empty)
    assert raised(lambda: KafkaDataset()) == FACTS["no_topic"]
    assert raised(lambda: KafkaDataset.new_consumer()) == FACTS["new_consumer_no_topic"]
    assert raised(lambda: next(iter(D.placeholder()))) == FACTS["placeholder_iter"]
    assert raised(lambda: D.placeholder().commit()) == FACTS["placeholder_commit"]
    assert raised(lambda: KafkaDataset._process(None, None)) == FACTS["process_not_implemented"]
    assert raised(lambda: next(auto_commit([1, 2, 3]))) == FACTS["auto_commit_type"]

    def direct():
        d = D("t")
        d._worker_id = 0
        d.commit()

    def bad_signal():
        d = D("t")
        d._worker_id = 0
        d.commit(999, None)

    assert raised(direct) == FACTS["worker_direct_commit"]
    assert raised(bad_signal) == FACTS["worker_bad_signal"]


def test_consumer_receives_user_kwargs_plus_auto_commit_off(monkeypatch):
    sk.install()          # the default new_consumer looks the consumer class up in `kafka`
    import torchkafka_b200.kafka_dataset as kd
    monkeypatch.setattr(kd, "_KafkaConsumer", None)
    d = KafkaDataset("t", group_id="g", foo=1)
    assert {k: d._consumer.config[k] for k in sorted(d._consumer.config)} == FACTS["consumer_kwargs"]
    assert list(d._consumer.topics) == FACTS["consumer_topics"]
    assert "_is_placeholder" not in d._consumer.config


def test_cpu_path_batches_and_commits_match_reference_golden():
    case = CASES["cfg1_readme_8floats"]
    g = np.load(os.path.join(GOLD, "cfg1_readme_8floats.npz"))
    D = make(build_log(case), lambda r: torch.tensor(json.loads(r.value), dtype=torch.float32))
    ds = D("topic", group_id="g")
    batches = list(auto_commit(DataLoader(ds, batch_size=case["batch_size"])))
    assert len(batches) == int(g["n_batches"])
    for k, b in enumerate(batches):
        assert np.array_equal(b.numpy(), g[f"b{k}_values"])
    assert len(ds._consumer.committed_log) == len(batches)
    for k, snap in enumerate(ds._consumer.committed_log):
        assert snap == {p: int(o) for p, o in enumerate(g["commits"][k]) if o >= 0}


def test_commit_happens_when_next_batch_is_requested():
    D = make(build_log(CASES["cfg1_readme_8floats"]))
    ds = D("topic")
    gen = auto_commit(DataLoader(ds, batch_size=4))
    next(gen)
    assert len(ds._consumer.committed_log) == 0      # at-least-once
    next(gen)
    assert len(ds._consumer.committed_log) == 1
    gen.close()
    assert len(ds._consumer.committed_log) == 1      # breaking out leaves the last batch uncommitted


def test_commit_failure_is_logged_not_raised(caplog):
    D = make(build_log(CASES["cfg1_readme_8floats"]), consumer_kw={"fail_commits": True})
    ds = D("topic")
    with caplog.at_level(logging.ERROR, logger="torchkafka_b200.kafka_dataset"):
        ds.commit()
    assert "Commit failed." in caplog.text
    assert ds._commit_required is False


def test_close_does_not_commit():
    D = make(build_log(CASES["cfg1_readme_8floats"]))
    ds = D("topic")
    consumer = ds._consumer
    ds.close()
    assert consumer.closed and consumer.close_autocommit is False and consumer.committed_log == []


def test_non_kafka_dataset_passes_through():
    data = TensorDataset(torch.arange(10))
    out = list(auto_commit(DataLoader(data, batch_size=5)))
    assert len(out) == 2


def test_ragged_samples_raise_like_default_collate():
    log = sk.RecordLog([b"[1, 2]", b"[1, 2, 3]"], [0, 0], [0, 1])
    D = make(log, lambda r: json.loads(r.value))          # list-returning _process (README.md:72-79)
    with pytest.raises(RuntimeError, match="equal size"):
        list(DataLoader(D("t"), batch_size=2))


def test_multi_worker_signal_commit_path():
    """placeholder + init_worker + auto_commit with num_workers=2 (README.md:104-132).  The stub
    blocks after its last record, as a real consumer does, and we stop after a fixed number of
    batches (a finite stream kills the reference's own workers)."""
    base = build_log(CASES["cfg5_raw64"])

    class DS(KafkaDataset):
        @classmethod
        def new_consumer(cls, *a, **k):
            from torch.utils.data import get_worker_info
            info = get_worker_info()
            k.pop("enable_auto_commit", None)
            sub = base.select_partitions(lambda p: p % info.num_workers == info.id)
            return sk.KafkaConsumer(*a, log=sub, block_at_end=True, **k)

        def _process(self, record):
            return torch.frombuffer(bytearray(record.value), dtype=torch.uint8)

    loader = DataLoader(DS.placeholder(), batch_size=16, num_workers=2,
                        worker_init_fn=DS.init_worker("topic", group_id="g"))
    seen = 0
    gen = auto_commit(loader)
    for batch in gen:
        assert batch.shape == (16, 64)
        seen += 1
        if seen == 6:
            break
    assert seen == 6


def test_schema_dataset_without_gpu_fails_loudly():
    if torch.cuda.is_available():
        pytest.skip("only meaningful on a machine without CUDA")
    from torchkafka_b200.schema import JsonArray
    log = build_log(CASES["cfg1_readme_8floats"])

    class DS(KafkaDataset):
        payload_schema = JsonArray(torch.float32, fixed_len=8)

        @classmethod
        def new_consumer(cls, *a, **k):
            return sk.KafkaConsumer(*a, log=log)

    with pytest.raises(RuntimeError, match="no CPU fallback"):
        next(auto_commit(DataLoader(DS("t"), batch_size=4)))


def test_finite_stream_with_workers_does_not_kill_the_workers():
    """with the reference a finite stream ends in 'DataLoader worker is killed by
    signal: User defined signal 1'.  Here the run completes and every batch is delivered."""
    base = build_log(CASES["cfg5_raw64"])

    class DS(KafkaDataset):
        @classmethod
        def new_consumer(cls, *a, **k):
            from torch.utils.data import get_worker_info
            info = get_worker_info()
            k.pop("enable_auto_commit", None)
            sub = base.select_partitions(lambda p: p % info.num_workers == info.id)
            return sk.KafkaConsumer(*a, log=sub, **k)           # finite: StopIteration at the end

        def _process(self, record):
            return torch.frombuffer(bytearray(record.value), dtype=torch.uint8)

    loader = DataLoader(DS.placeholder(), batch_size=50, num_workers=2,
                        worker_init_fn=DS.init_worker("topic", group_id="g"))
    rows = sum(b.shape[0] for b in auto_commit(loader))
    assert rows == len(base)


def test_struct_schema_layout():
    import torch
    from torchkafka_b200.schema import StructBytes
    sc = StructBytes((("x", torch.float32, (3,)), ("flag", torch.uint8, ()), ("y", torch.int64, 1), ("h", torch.float16, (2, 2))))
    assert sc.field_bytes == (12, 1, 8, 8) and sc.field_offsets == (0, 12, 13, 21) and sc.record_bytes == 29
    assert sc.fields[2][2] == (1,)
    b = sc.wrap(1, 2, 3, 4)
    assert b._fields == ("x", "flag", "y", "h")
    import pytest
    with pytest.raises(ValueError):
        StructBytes(tuple(("f%d" % i, torch.uint8, ()) for i in range(9)))


def test_workers_commit_what_was_delivered_not_their_read_ahead(tmp_path):
    """with the reference a worker commits its consumer's position, which is
    ahead of the delivered batches by the DataLoader's prefetch, and signals go round-robin even
    after a worker's stream ended.  Here every commit a worker makes is the position at the end of
    one of *its* delivered batches -- the table single-process iteration over its partitions
    commits -- and never beyond what the main process received from it."""
    import json as _json
    from oracle import reference_path as ref
    base = build_log(CASES["cfg5_raw64"])
    B = 8

    def process(record):
        return torch.frombuffer(bytearray(record.value), dtype=torch.uint8)

    class Logged(sk.KafkaConsumer):
        def commit(self, offsets=None):
            super().commit(offsets)
            with open(self.log_path, "a") as f:
                f.write(_json.dumps({str(k): v for k, v in self.committed_log[-1].items()}) + "\n")

    def sub_log(worker, n_workers):
        # uneven split: worker 0 owns one partition in four, so its stream ends long before worker 1's
        return base.select_partitions(lambda p: (p % 4 == 0) == (worker == 0))

    class DS(KafkaDataset):
        @classmethod
        def new_consumer(cls, *a, **k):
            from torch.utils.data import get_worker_info
            info = get_worker_info()
            k.pop("enable_auto_commit", None)
            c = Logged(*a, log=sub_log(info.id, info.num_workers), **k)
            c.log_path = str(tmp_path / f"worker{info.id}.jsonl")
            return c

        def _process(self, record):
            return process(record)

    loader = DataLoader(DS.placeholder(), batch_size=B, num_workers=2, prefetch_factor=2,
                        worker_init_fn=DS.init_worker("topic", group_id="g"))
    rows = sum(b.shape[0] for b in auto_commit(loader))
    assert rows == len(base)
    total = 0
    for w in range(2):
        want = [{str(k): v for k, v in committed.items()}
                for _, committed in ref.run(sk.KafkaConsumer("topic", log=sub_log(w, 2)), process, B)]
        path = tmp_path / f"worker{w}.jsonl"
        got = [_json.loads(line) for line in path.read_text().splitlines()] if path.exists() else []
        assert got, f"worker {w} never committed"
        k = -1
        for table in got:                                  # each one a batch boundary, in order
            k = want.index(table, k + 1)
        # the first commit of a worker: the reference is already two prefetched batches ahead
        assert got[0] in want[:2]
        total += len(got)
    assert total >= 4
