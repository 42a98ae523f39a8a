"""ga_average_host pipeline: time per call against the copy-in / reduce / copy-out chunk size (development tool).
    python tools/e2e_chunk_sweep.py            # 4 hosted ranks x 64 MiB on one GPU"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import pytorch_distributed_example_b200 as ga
from pytorch_distributed_example_b200 import _lib

S = 64 << 20
multi = "LOCAL_RANK" in os.environ                       # torchrun: one rank per GPU; else 4 hosted ranks on one GPU
if multi:
    import torch.distributed as dist
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    world, hosted = dist.get_world_size(), 1
    avg = ga.GradAverager(capacity_bytes=S)
else:
    world = hosted = 4
    torch.cuda.set_device(0)
    avg = ga.GradAverager(local_world=world, device=0, capacity_bytes=S)
host = [[torch.empty(S // 4, dtype=torch.float32).pin_memory().normal_() for _ in range(hosted)] for _ in range(2)]
for chunk in (0, 1 << 20, 2 << 20, 4 << 20, 8 << 20, 16 << 20, 32 << 20, 64 << 20):
    if chunk:
        avg.set_option(_lib.OPT_HOST_CHUNK_BYTES, chunk)
    for variant in ("auto",):
        for i in range(3):
            avg.average_host([[t] for t in host[i % 2]], variant=variant)
        torch.cuda.synchronize()
        if multi:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(10):
            avg.average_host([[t] for t in host[i % 2]], variant=variant)
        e1.record()
        torch.cuda.synchronize()
        if not multi or dist.get_rank() == 0:
            print(json.dumps({"world": world, "chunk_bytes": chunk or "default", "variant": variant, "ms_per_call": round(e0.elapsed_time(e1) / 10, 3)}), flush=True)
avg.check_status()
if multi:
    dist.barrier()
avg.close()
if multi:
    dist.destroy_process_group()
