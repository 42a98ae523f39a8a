"""Per-phase timing of the push kernels across real GPUs (ncu cannot attach to a multi-rank run).
    torchrun --nproc-per-node N tools/phase_trace.py [bytes ...]
Prints, per size and variant, the median over launches of: the mean and the max over blocks of each
phase's duration on rank 0, plus the whole-kernel span (first block start -> last block end)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

import pytorch_distributed_example_b200 as ga
from pytorch_distributed_example_b200 import _lib

PHASES = ["wait_prev_done", "scatter_issue", "flags1_raise(acks)", "flags1_wait", "fold(+result stores)",
          "flags2_raise(acks)", "flags2_wait", "unpack", "done_signal"]


def main():
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    world = dist.get_world_size()
    sizes = [int(a) for a in sys.argv[1:]] or [4 << 10, 1 << 20, 1724320 // 4 * 4, 16 << 20, 64 << 20]
    avg = ga.GradAverager(capacity_bytes=max(max(sizes), 8 << 20))
    avg.set_option(_lib.OPT_TRACE, 1)
    out = []
    for S in sizes:
        bufs = [torch.randn(S // 4, device="cuda") for _ in range(4)]
        for vname, v in (("oneshot_push", ga.ONESHOT_PUSH), ("twoshot_push", ga.TWOSHOT_PUSH)):
            if vname == "oneshot_push" and S * world > avg.capacity:
                continue
            rows = []
            for it in range(12):
                dist.barrier()
                torch.cuda.synchronize()
                avg.all_reduce_([[bufs[it % 4]]], average=True, variant=v)
                t = avg.read_trace()[0]                     # [blocks, slots]
                if it < 2:
                    continue
                d = np.diff(t[:, :10], axis=1).astype(np.float64) / 1e3       # us per phase per block
                if vname == "oneshot_push":
                    d[:, 4:8] = 0.0                                           # slots 5-7 unused; phase "fold" = 4 -> 8
                    d[:, 4] = (t[:, 8] - t[:, 4]) / 1e3
                rows.append({"mean": d.mean(axis=0), "max": d.max(axis=0),
                             "span": (t[:, 9].max() - t[:, 0].min()) / 1e3, "blocks": t.shape[0]})
            rec = {"bytes": S, "variant": vname, "world": world, "blocks": rows[0]["blocks"],
                   "kernel_span_us": float(np.median([r["span"] for r in rows])),
                   "phase_us_mean_over_blocks": {p: round(float(np.median([r["mean"][i] for r in rows])), 2) for i, p in enumerate(PHASES)},
                   "phase_us_max_over_blocks": {p: round(float(np.median([r["max"][i] for r in rows])), 2) for i, p in enumerate(PHASES)}}
            out.append(rec)
            if rank == 0:
                print(json.dumps(rec), flush=True)
        del bufs
    avg.check_status()
    dist.barrier()
    avg.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
