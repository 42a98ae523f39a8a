"""Does the averaging kernel run faster straight on pinned host gradients (UVA zero-copy: PCIe reads and writes
issued by the SMs) than ga_average_host's copy-in / reduce / copy-out pipeline?  One GPU, 4 hosted ranks, or torchrun."""
import ctypes
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import pytorch_distributed_example_b200 as ga

S = 64 << 20
multi = "LOCAL_RANK" in os.environ
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
lib = ga.load_library()
host = [[torch.empty(S // 4, dtype=torch.float32).pin_memory().normal_() for _ in range(hosted)] for _ in range(2)]
ref = [t.clone() for t in host[0]]
avg.average_host([[t] for t in host[0]])           # creates the context; result for the parity check below
torch.cuda.synchronize()
want = host[0][0].clone()
for t, r in zip(host[0], ref):
    t.copy_(r)


def zero_copy(ts, variant):
    n = 1
    ptrs = (ctypes.c_void_p * (n * len(ts)))(*[t.data_ptr() for t in ts])
    numel = (ctypes.c_int64 * n)(ts[0].numel())
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.ga_average(avg._ctx, ptrs, numel, n, st, variant)
    assert rc == 0, lib.ga_last_error()


for name, variant in (("auto", 0), ("twoshot_push", 4), ("twoshot", 2), ("oneshot", 1)):
    if not multi and variant == 4:
        continue
    for t, r in zip(host[0], ref):
        t.copy_(r)
    zero_copy(host[0], variant)
    torch.cuda.synchronize()
    avg.check_status()
    ok = bool(torch.equal(host[0][0], want))
    for i in range(2):
        zero_copy(host[i % 2], variant)
    torch.cuda.synchronize()
    if multi:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(10):
        zero_copy(host[i % 2], variant)
    e1.record()
    torch.cuda.synchronize()
    if not multi or dist.get_rank() == 0:
        print(json.dumps({"world": world, "zero_copy_variant": name, "ms_per_call": round(e0.elapsed_time(e1) / 10, 3), "same_bits_as_pipeline": ok}), flush=True)
for i in range(3):
    avg.average_host([[t] for t in host[i % 2]])
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(10):
    avg.average_host([[t] for t in host[i % 2]])
e1.record()
torch.cuda.synchronize()
if not multi or dist.get_rank() == 0:
    print(json.dumps({"world": world, "pipeline_ms_per_call": round(e0.elapsed_time(e1) / 10, 3)}), flush=True)
avg.check_status()
if multi:
    dist.barrier()
avg.close()
if multi:
    dist.destroy_process_group()
