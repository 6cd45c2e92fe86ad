#!/usr/bin/env python3
"""Generate tests/golden/*.npz|json from the UNMODIFIED reference.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

The reference ships no tests or golden vectors, so the parity pin is its own
output: this script installs the in-memory `kafka` stub, imports `/root/reference/src` as the
package `src` (the only way its absolute import in auto_commit.py:19 resolves),
defines the README-style subclasses for the five BASELINE.json configs, runs
`src.auto_commit(DataLoader(dataset, batch_size))` on small seeded logs and stores what came out:
every batch and the offsets committed after it.  It also records the reference's exception
types/messages and CPython/torch's answers on the numeric known-answer set
(SURVEY §4.1-11).  Inputs are regenerated from seeds by the tests; their SHA-256 is stored so a
drifting generator is noticed.
"""
import hashlib
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from torchkafka_b200.testing import stub_kafka as sk          # noqa: E402
from torchkafka_b200.testing import synthetic as syn          # noqa: E402
from golden_cases import CASES, KNOWN_ANSWER_PAYLOADS, build_log   # noqa: E402

sk.install()
sys.path.insert(0, "/root/reference")
import src as ref                                               # noqa: E402  (the reference itself)
from torch.utils.data import DataLoader                         # noqa: E402

IMAGENET_MEAN = torch.tensor(syn.IMAGENET_MEAN).view(3, 1, 1)
IMAGENET_STD = torch.tensor(syn.IMAGENET_STD).view(3, 1, 1)


def dataset_class(case, log):
    kind = case["kind"]

    class DS(ref.KafkaDataset):
        @classmethod
        def new_consumer(cls, *args, **kwargs):          # README.md:46-57 override point
            kwargs.pop("enable_auto_commit", None)
            return sk.KafkaConsumer(*args, log=log, **kwargs)

        def _process(self, record):
            if kind == "json":
                elements = json.loads(record.value)                       # README.md:74
                if len(elements) < case.get("min_size", 0):               # README.md:76-77
                    return None
                return torch.tensor(elements, dtype=getattr(torch, case["dtype"]))
            if kind == "raw":
                return torch.frombuffer(bytearray(record.value), dtype=torch.uint8)
            x = torch.frombuffer(bytearray(record.value), dtype=torch.uint8).view(*case["hwc"])
            x = x.permute(2, 0, 1).float().div(255)
            return ((x - IMAGENET_MEAN) / IMAGENET_STD).to(torch.bfloat16)

    return DS


def pad_collate(samples):
    lengths = torch.tensor([s.shape[0] for s in samples], dtype=torch.int32)
    padded = torch.nn.utils.rnn.pad_sequence(samples, batch_first=True, padding_value=0)
    return padded, lengths


def run_case(name, case):
    log = build_log(case)
    ds = dataset_class(case, log)("topic", group_id="golden")
    kw = {"collate_fn": pad_collate} if case.get("pad") else {}
    loader = DataLoader(ds, batch_size=case["batch_size"], drop_last=case.get("drop_last", False), **kw)
    out = {}
    k = 0
    for batch in ref.auto_commit(loader):
        if case.get("pad"):
            vals, lengths = batch
            out[f"b{k}_values"] = vals.numpy()
            out[f"b{k}_lengths"] = lengths.numpy()
        elif batch.dtype == torch.bfloat16:
            out[f"b{k}_values"] = batch.view(torch.int16).numpy()          # raw bf16 bit patterns
        else:
            out[f"b{k}_values"] = batch.numpy()
        k += 1
    commits = ds._consumer.committed_log
    assert len(commits) == k
    nparts = int(log.partitions.max()) + 1
    table = np.full((k, nparts), -1, dtype=np.int64)
    for j, snap in enumerate(commits):
        for p, o in snap.items():
            table[j, p] = o
    out["commits"] = table
    out["n_batches"] = np.array(k)
    out["input_sha256"] = np.frombuffer(hashlib.sha256(log.values.tobytes()).digest(), dtype=np.uint8)
    np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **out)
    print(f"{name}: {k} batches, {len(log)} records")


def known_answers():
    res = []
    for payload in KNOWN_ANSWER_PAYLOADS:
        entry = {"payload": payload}
        for dt in ("float32", "float64", "int32", "int64"):
            try:
                t = torch.tensor(json.loads(payload.encode("latin-1")), dtype=getattr(torch, dt))
                if t.dim() != 1:
                    entry[dt] = {"shape": list(t.shape)}
                elif dt.startswith("float"):
                    it = torch.int32 if dt == "float32" else torch.int64
                    entry[dt] = {"bits": [int(x) for x in t.view(it).tolist()]}
                else:
                    entry[dt] = {"values": [int(x) for x in t.tolist()]}
            except Exception as exc:                      # noqa: BLE001
                entry[dt] = {"error": type(exc).__name__}
        res.append(entry)
    return res


def api_facts():
    """Exception types / messages of the reference API."""
    facts = {}

    def grab(key, fn):
        try:
            fn()
            facts[key] = None
        except Exception as exc:                          # noqa: BLE001
            facts[key] = [type(exc).__name__, str(exc)]

    class D(ref.KafkaDataset):
        @classmethod
        def new_consumer(cls, *a, **k):
            return sk.KafkaConsumer(*a, log=sk.RecordLog([], [], []))

    grab("no_topic", lambda: ref.KafkaDataset())
    grab("new_consumer_no_topic", lambda: ref.KafkaDataset.new_consumer())
    grab("placeholder_iter", lambda: next(iter(D.placeholder())))
    grab("placeholder_commit", lambda: D.placeholder().commit())
    grab("process_not_implemented", lambda: ref.KafkaDataset._process(None, None))
    grab("auto_commit_type", lambda: next(ref.auto_commit([1, 2, 3])))

    def worker_direct_commit():
        d = D("t")
        d._worker_id = 0
        d.commit()

    def worker_bad_signal():
        d = D("t")
        d._worker_id = 0
        d.commit(999, None)

    grab("worker_direct_commit", worker_direct_commit)
    grab("worker_bad_signal", worker_bad_signal)
    # default new_consumer: the consumer sees the user's kwargs + enable_auto_commit=False (§4.1-2)
    d = ref.KafkaDataset("t", group_id="g", foo=1)
    facts["consumer_kwargs"] = {k: d._consumer.config[k] for k in sorted(d._consumer.config)}
    facts["consumer_topics"] = list(d._consumer.topics)
    return facts


def main():
    for name, case in CASES.items():
        run_case(name, case)
    with open(os.path.join(HERE, "known_answers.json"), "w") as f:
        json.dump({"python": sys.version.split()[0], "torch": torch.__version__,
                   "answers": known_answers()}, f, indent=1)
    with open(os.path.join(HERE, "api_facts.json"), "w") as f:
        json.dump(api_facts(), f, indent=1)
    print("known answers + api facts written")


if __name__ == "__main__":
    main()
