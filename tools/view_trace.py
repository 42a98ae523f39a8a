"""Per-phase timing of ga_view_kernel / ga_nvls_kernel across real GPUs (ncu cannot wrap a multi-rank run).
    torchrun --nproc-per-node N tools/view_trace.py [bytes ...]      (VT_CHUNKS=0,28 to sweep chunk sizes)
Per size: median over launches of each timestamp relative to the first block's start on rank 0
(mean and max over the blocks of a role), in microseconds."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

import pytorch_distributed_example_b200 as ga
from pytorch_distributed_example_b200 import _lib

SCATTER = {1: "staging_released", 5: "data_warp_issued_all", 3: "first_flags_raised", 2: "last_flags_raised", 9: "end"}
FOLD = {3: "first_chunk_arrived", 6: "last_chunk_arrived", 5: "data_warp_issued_all", 2: "last_flags_raised", 4: "all_results_seen", 9: "end"}


def main():
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    world = dist.get_world_size()
    sizes = [int(a) for a in sys.argv[1:]] or [1724320, 16 << 20, 64 << 20]
    chunks = [int(c) for c in os.environ.get("VT_CHUNKS", "0").split(",")]
    cap = max(max(sizes), 8 << 20)
    avg = ga.GradAverager(capacity_bytes=cap)
    avg.create_view(cap, multicast=True)
    avg.set_option(_lib.OPT_TRACE, 1)
    for S in sizes:
        src = torch.randn(S // 4, device="cuda")
        v = avg.view_tensor(S)
        for variant in ["view"] + (["nvls"] if avg.has_multicast else []):
            for chunk in (chunks if variant == "view" else [0]):
                avg.set_option(_lib.OPT_VIEW_CHUNK_TILES, chunk)
                recs = []
                for it in range(12):
                    v.copy_(src)
                    dist.barrier()
                    torch.cuda.synchronize()
                    # back-to-back launches pace one another on the GPUs: the LAST one is traced free of the
                    # tens of microseconds of host launch skew that follow a barrier
                    for _ in range(int(os.environ.get("VT_TRAIN", "6"))):
                        avg.allreduce_view(S, variant=variant)
                    t = avg.read_trace()[0].astype(np.float64)        # [blocks, slots]
                    if it >= 2:
                        recs.append((t - t[:, 0].min()) / 1e3)
                t = np.median(np.stack(recs), axis=0)
                rec = {"bytes": S, "variant": variant, "world": world, "chunk_tiles": chunk, "blocks": int(t.shape[0]),
                       "kernel_span_us": round(float(t[:, 9].max()), 1), "start_skew_us": round(float(t[:, 0].max()), 1)}
                if variant == "view":
                    for name, rows, slots in (("scatter", t[0::2], SCATTER), ("fold", t[1::2], FOLD)):
                        rec[name] = {lbl: [round(float(rows[:, s].mean()), 1), round(float(rows[:, s].max()), 1)] for s, lbl in slots.items()}
                else:
                    rec["nvls"] = {lbl: [round(float(t[:, s].mean()), 1), round(float(t[:, s].max()), 1)]
                                   for s, lbl in {1: "all_ranks_started", 2: "reduce_and_stores_issued", 4: "all_stores_seen", 9: "end"}.items()}
                if rank == 0:
                    print(json.dumps(rec), flush=True)
    avg.set_option(_lib.OPT_VIEW_CHUNK_TILES, 0)
    avg.check_status()
    dist.barrier()
    avg.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
