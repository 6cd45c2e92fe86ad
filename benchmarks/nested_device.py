import sys, json, numpy as np, torch
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/benchmarks'); sys.path.insert(0, '/root/repo/tests')
import all_configs as ac
from torchkafka_b200.schema import JsonNested
from torchkafka_b200.testing.stub_kafka import RecordLog
for shape, n in [((4, 3), 65536), ((64, 24), 2048)]:
    rng = np.random.default_rng(1)
    vals = [json.dumps(rng.uniform(-1e3, 1e3, shape).tolist()).encode() for _ in range(min(n, 4096))]
    vals = (vals * (n // len(vals) + 1))[:n]
    i = np.arange(n)
    log = RecordLog(vals, i % 64, i // 64)
    cfg = dict(schema=JsonNested(shape, torch.float32), batch=256, slot_bytes=log.nbytes + (1 << 20))
    dev, _ = ac.device_resident(cfg, log, passes=8)
    print(shape, n, log.nbytes // n, dev)
