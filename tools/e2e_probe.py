import sys, time
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pytorch_distributed_example_b200 as ga
from pytorch_distributed_example_b200 import _lib
torch.cuda.set_device(0)
W, S = 4, 64 << 20
avg = ga.GradAverager(local_world=W, device=0, capacity_bytes=S)
host = [[torch.empty(S // 4).pin_memory().normal_()] for _ in range(W)]
for chunk in (1 << 20, 2 << 20, 4 << 20, 8 << 20, 16 << 20, 64 << 20):
    avg.set_option(_lib.OPT_HOST_CHUNK_BYTES, chunk)
    for _ in range(3): avg.average_host(host)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): avg.average_host(host)
    e1.record(); torch.cuda.synchronize()
    print("max chunk %3d MiB: %.3f ms/step" % (chunk >> 20, e0.elapsed_time(e1) / 10))
# plain copies for reference
d = torch.empty(W * S // 4, device="cuda"); h = torch.empty(W * S // 4).pin_memory()
for name, fn in (("H2D 256MiB", lambda: d.copy_(h, non_blocking=True)), ("D2H 256MiB", lambda: h.copy_(d, non_blocking=True))):
    fn(); torch.cuda.synchronize(); e0.record()
    for _ in range(5): fn()
    e1.record(); torch.cuda.synchronize(); print(name, "%.3f ms" % (e0.elapsed_time(e1) / 5))
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream(); h2 = torch.empty(W * S // 4).pin_memory(); d2 = torch.empty(W * S // 4, device="cuda")
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(5):
    with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
torch.cuda.synchronize(); print("H2D + D2H concurrently, 256 MiB each: %.3f ms" % ((time.perf_counter() - t0) / 5 * 1e3))
