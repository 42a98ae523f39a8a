#!/usr/bin/env python
"""The toy all-reduce loop on GPUs: every step each rank draws an int, all ranks sum it through the
fused kernel (int32, one-shot) and print the sum.  Same flags and output line as the reference
(`/root/reference/toy/main.py`: flags `:36-56`, loop `:9-25`)."""
import argparse
import os
import random
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pytorch_distributed_example_b200 import distributed as dist   # noqa: E402  (instead of torch.distributed)


def run(rank, world_size, steps, pause):
    everyone = dist.new_group(ranks=list(range(world_size)))
    for step in range(1, steps + 1):
        value = random.randint(0, 10)
        t = torch.IntTensor([value])
        dist.all_reduce(t, op=dist.reduce_op.SUM, group=everyone)
        print("rank: {}, step: {}, value: {}, reduced sum: {}.".format(rank, step, value, float(t)))
        time.sleep(pause)


def main(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("--backend", type=str, default="gloo", help="rendezvous backend (carries the IPC handles)")
    p.add_argument("--init-method", "-i", type=str, default="tcp://127.0.0.1:23456")
    p.add_argument("--rank", "-r", type=int, required=True)
    p.add_argument("--world-size", "-s", type=int, required=True)
    p.add_argument("--steps", type=int, default=20)
    p.add_argument("--pause", type=float, default=1.0, help="seconds between steps (the reference sleeps 1 s)")
    p.add_argument("--seed", type=int, default=None, help="random.seed(seed + rank), for reproducible runs")
    args = p.parse_args(argv)
    print(args)
    if args.seed is not None:
        random.seed(args.seed + args.rank)
    dist.init_process_group(backend=args.backend, init_method=args.init_method, rank=args.rank,
                            world_size=args.world_size)
    run(args.rank, args.world_size, args.steps, args.pause)
    import pytorch_distributed_example_b200 as ga
    ga.reset_default_averager()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
