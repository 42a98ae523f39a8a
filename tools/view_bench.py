"""Quick per-variant timing across GPUs (development tool): torchrun --nproc-per-node N tools/view_bench.py
Fused tensor call (drop-in: pack + exchange + unpack), in-place view kernel, in-switch NVLS kernel,
NCCL all_reduce alone and + div_, per size; device-timed, median of per-iteration max over ranks."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pytorch_distributed_example_b200 as ga  # noqa: E402
from pytorch_distributed_example_b200 import _lib  # noqa: E402


def main():
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    lr = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    world, rank = dist.get_world_size(), dist.get_rank()
    sizes = [int(s) for s in os.environ.get("VB_SIZES", "1048576,1724320,4194304,16777216,46758048,67108864,268435456").split(",")]
    chunks = [int(c) for c in os.environ.get("VB_CHUNKS", "0").split(",")]
    cap = max(sizes) + 4096
    if os.environ.get("VB_ONESHOT"):
        cap = max(cap, min(max(sizes), 64 << 20) * (world - 1))
    avg = ga.GradAverager(capacity_bytes=cap)
    avg.create_view(cap, multicast=True)
    dev = torch.device("cuda", lr)

    def timed(fn, iters, warm=5):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
        evs[0].record()
        for i in range(iters):
            fn()
            evs[i + 1].record()
        torch.cuda.synchronize()
        t = torch.tensor([evs[i].elapsed_time(evs[i + 1]) for i in range(iters)], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t = t.sort().values
        return round(float(t[iters // 2]) * 1e3, 1)

    for S in sizes:
        n = S // 4
        iters = 40 if S < (256 << 20) else 10
        bufs = [torch.randn(n, device=dev) for _ in range(2 if S >= (32 << 20) else 6)]
        k = [0]
        row = {"bytes": S, "world": world, "multicast": avg.has_multicast}

        def fused():
            k[0] += 1
            avg.all_reduce_([[bufs[k[0] % len(bufs)]]], average=True)
        row["dropin_auto_us"] = timed(fused, iters)
        v = avg.view_tensor(S)
        v.copy_(bufs[0])
        for c in chunks:
            avg.set_option(_lib.OPT_VIEW_CHUNK_TILES, c)
            row["view_c%d_us" % c] = timed(lambda: avg.allreduce_view(S, variant="view"), iters)
            row["view_c%d_grid" % c] = avg.last_grid[0]
        avg.set_option(_lib.OPT_VIEW_CHUNK_TILES, 0)
        if os.environ.get("VB_ONESHOT") and S * (world - 1) <= cap:
            avg.set_option(_lib.OPT_VIEW_ONESHOT_MAX_BYTES, 1 << 40)
            row["view_oneshot_us"] = timed(lambda: avg.allreduce_view(S, variant="view"), iters)
            avg.set_option(_lib.OPT_VIEW_ONESHOT_MAX_BYTES, 0)
            row["view_twoshot_us"] = timed(lambda: avg.allreduce_view(S, variant="view"), iters)
            avg.set_option(_lib.OPT_VIEW_ONESHOT_MAX_BYTES, -1)
        if avg.has_multicast:
            row["nvls_us"] = timed(lambda: avg.allreduce_view(S, variant="nvls"), iters)
        b = bufs[0]
        row["nccl_allreduce_us"] = timed(lambda: dist.all_reduce(b), iters)
        row["nccl_allreduce_div_us"] = timed(lambda: (dist.all_reduce(b), b.div_(float(world))), iters)
        f = 2.0 * (world - 1) / world
        best = min(x for key, x in row.items() if key.startswith("view_c") and key.endswith("_us"))
        row["view_bus_gbps"] = round(S / best / 1e3 * f, 1)
        if "nvls_us" in row:
            row["nvls_bus_gbps"] = round(S / row["nvls_us"] / 1e3 * f, 1)
        avg.check_status()
        if rank == 0:
            print(json.dumps(row), flush=True)
        del bufs
    dist.barrier()
    avg.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
