import sys, time, os
sys.path.insert(0, '/root/repo')
import numpy as np, torch
from torchkafka_b200.engine import SlotEngine
from torchkafka_b200.schema import JsonArray
from torchkafka_b200.testing import synthetic as syn
log = syn.make_log(2, 17600, seed=1, layout="chunked", chunk=500)
eng = SlotEngine(JsonArray(torch.float32, max_len=256, min_size=32), 1024, slot_bytes=48<<20, slot_records=20000, n_slots=2)
vals, starts, parts, offs = log.values, log.starts.astype(np.uint32), log.partitions, log.offsets
print("cores", len(os.sched_getaffinity(0)))
for th in (1, 2, 4, 7, 8, 12, 14, 16):
    eng.pack_threads = th
    eng.fill_packed(0, vals, starts, parts, offs)
    t = time.perf_counter()
    for k in range(10): eng.fill_packed(k & 1, vals, starts, parts, offs)
    dt = (time.perf_counter() - t) / 10
    print(th, "threads:", round(log.nbytes / dt / 1e9, 1), "GB/s copy", round(dt * 1e3, 2), "ms/slot")
# plain numpy copy for reference
dst = np.empty_like(vals)
t = time.perf_counter()
for k in range(5): np.copyto(dst, vals)
print("numpy single thread:", round(log.nbytes * 5 / (time.perf_counter() - t) / 1e9, 1), "GB/s")
