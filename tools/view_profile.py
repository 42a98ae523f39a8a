"""Bare loop for ncu: ga_view_kernel on a hosted world (4 ranks x 64 MiB on one GPU, cooperative launch --
ncu must not wrap a multi-rank run).  python tools/view_profile.py [calls]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import pytorch_distributed_example_b200 as ga

world, S = 4, 64 << 20
torch.cuda.set_device(0)
avg = ga.GradAverager(local_world=world, device=0, capacity_bytes=S)
avg.create_view(S, multicast=False)
src = [torch.randn(S // 4, device="cuda") for _ in range(world)]
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 6):
    for l in range(world):
        avg.view_tensor(S, l).copy_(src[l])
    avg.allreduce_view(S, variant="view")
avg.check_status()
avg.close()
print("ok")
