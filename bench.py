#!/usr/bin/env python
"""bench.py -- gradient-averaging throughput on B200 (contract: see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W]                     # our arm
    python bench.py --impl reference [--gpus N] [--steps K] [--warmup W]    # reference CPU arm
    torchrun --nproc-per-node N ... bench.py --gpus N ...                   # N > 1, one rank per GPU

A step is ONE pass of the hot path (`Trainer.average_gradients`, /root/reference/mnist/main.py:122-127)
over one synthetic gradient set: every rank's S-byte fp32 gradient is replaced by the rank-order sum
over all `world` ranks divided by float(world).

  N = 1   world = 4 ranks hosted on the one B200 (BASELINE.json configs[1]: "gloo world_size=4 on host
          cores vs 1 B200"), S = 64 MiB flat fp32 gradient per rank (the bucket size north_star quotes).
          Everything moves through HBM, so the roofline bound is "hbm".
  N > 1   world = N, one rank per GPU over CUDA-IPC/NVLink, same S.  The bound is NVLink; NCCL
          all_reduce + div_ on a same-size buffer is timed in the same run.

value   = world * busBW, busBW = (S / t) * 2(world-1)/world  (nccl-tests convention),
          t = device time per step (CUDA events on the launching stream, max over ranks), inputs
          resident in HBM.  Between steps the gradient buffers rotate over sets larger than L2.
e2e     = same formula through the host-buffer entry point (ga_average_host): pinned host gradients,
          H2D + kernel + D2H inside the timed region.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "grad-avg bus GB/s"
UNIT = "GB/s"
DEFAULT_BYTES = 64 << 20
L2_BYTES = 126 << 20


_T0 = time.time()


def stage(msg):
    """Progress line on stderr (the JSON line on stdout stays alone)."""
    print("[bench %6.1fs rank %s] %s" % (time.time() - _T0, os.environ.get("RANK", "0"), msg), file=sys.stderr, flush=True)


def bus_factor(world):
    return 2.0 * (world - 1) / world if world > 1 else 0.0


def agg_bus_gbps(world, nbytes, seconds):
    """world * busBW in GB/s."""
    return world * (nbytes / seconds) * bus_factor(world) / 1e9


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,utilization.gpu,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0, period_ms=50):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(index), "--query-gpu=" + self.FIELDS, "--format=csv,noheader,nounits",
                 "-lms", str(period_ms)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, windows):
        """windows: list of (t0, t1) wall-clock intervals that count as 'under load'."""
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        time.sleep(0.12)
        self.proc.terminate()
        sm, smax, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, line in self.rows:
            if not any(a <= t <= b + 0.06 for a, b in windows):
                continue
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[0]))
                smax = float(f[1])
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": smax,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# reference arm / cpu baseline (oracle.gloo_path on host cores) -- the only place oracle/ is executed
# ------------------------------------------------------------------------------------------------
def cpu_reference(world, nbytes, steps, warmup, budget_s):
    """The reference's own averaging loop on gloo, `world` CPU processes, one flat fp32 buffer.
    Each step covers a bounded sample of the workload so that the run fits `budget_s`."""
    from oracle import gloo_path
    est_s_per_byte = 1.2e-9 * max(1.0, world / 4.0)              # 64 MiB, ws=4 -> 72 ms
    sample = int(min(nbytes, max(1 << 20, budget_s / max(1, steps + warmup) / est_s_per_byte)))
    sample -= sample % 4
    cores = os.cpu_count() or 1
    threads = max(1, cores // world)
    times = gloo_path.time_flat_average(world, sample // 4, steps, warmup=warmup, threads=threads)
    sec = sum(times) / len(times)
    return {
        "value": agg_bus_gbps(world, sample, sec), "unit": UNIT, "cores": min(cores, world * threads),
        "kind": "port", "ms_per_step": sec * 1e3, "world_size": world,
        "sample": "%d of %d bytes per rank per step, %d steps, gloo tcp://127.0.0.1, %d procs x %d threads "
                  "(host has %d cores%s); oracle/gloo_path.py = mnist/main.py:122-127 restated"
                  % (sample, nbytes, steps, world, threads, cores, ", oversubscribed" if world > cores else ""),
    }


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return                                                   # rank 0 alone runs the CPU arm
    world = args.world or (4 if args.gpus == 1 else args.gpus)
    t0 = time.time()
    cb = cpu_reference(world, args.bytes, args.steps, args.warmup, budget_s=90.0)
    line = {
        "impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": cb["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(world, args.gpus, args.bytes, "gloo ring (torch)", None),
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": time.time() - t0,
    }
    print(json.dumps(line), flush=True)


def workload_config(world, gpus, nbytes, variant, rotate):
    return {
        "workload": "gradient averaging of one %d-byte fp32 gradient per rank, world_size=%d (%s); "
                    "BASELINE configs[1] world size on 1 GPU, configs[3] 64 MiB bucket point"
                    % (nbytes, world, "%d ranks hosted per GPU" % (world // gpus) if gpus == 1 else "one rank per GPU"),
        "bytes_per_rank": nbytes, "world_size": world, "ranks_per_gpu": world // gpus if gpus == 1 else 1,
        "variant": variant, "value_formula": "world * (S/t) * 2(world-1)/world",
        "l2": None if rotate is None else "gradient buffers rotate over %d sets (%d MiB > L2) between steps"
              % (rotate[0], rotate[1] >> 20),
    }


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def algorithmic_hbm_bytes(world, nbytes, variant_name, hosted):
    """Bytes the fused kernel's algorithm loads + stores per launch on ONE GPU hosting `hosted`
    ranks (kernel table): pack 2S; one-shot: read W*S, write S; two-shot: reduce-scatter
    read S + write 2S/W, all-gather read + write (W-1)S/W each."""
    s, w = float(nbytes), world
    per_rank = {"oneshot": (w + 3) * s,                       # pack 2S, read W*S, write S
                "twoshot": 5.0 * s,                           # pack 2S, RS read S + write 2S/W, AG read+write (W-1)S/W each
                "oneshot_push": (2 * w + 2) * s,              # read S, write W*S staging, read W*S, write S
                "twoshot_push": (4.0 + 2.0 * (w - 1) / w) * s,  # read S, write S, read S, write S (grads+gather), unpack 2(W-1)S/W
                "twoshot_pipe": (4.0 + 2.0 * (w - 1) / w) * s,
                "hosted_direct": 2.0 * s                      # read every rank's gradient once, write every result once
                }[variant_name]
    return per_rank * hosted


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--bytes", type=int, default=DEFAULT_BYTES, help="gradient bytes per rank")
    ap.add_argument("--world", type=int, default=0, help="world size when --gpus 1 (default 4)")
    ap.add_argument("--variant", default="auto", choices=["auto", "oneshot", "twoshot", "oneshot_push", "twoshot_push", "twoshot_pipe", "hosted_direct"])
    ap.add_argument("--no-extras", action="store_true", help="skip the sweep / MNIST / cpu_baseline legs")
    ap.add_argument("--sweep", action="store_true", help="N > 1: size sweep 4 KB..1 GB, ours (both variants, fused and bucket) vs NCCL")
    ap.add_argument("--sweep-max", type=int, default=1 << 30)
    ap.add_argument("--explore", action="store_true", help="N > 1: tuning grid for the pipelined variant (group x blocks)")
    ap.add_argument("--profile", action="store_true", help="bare loop for ncu: no extras, no cpu leg, no e2e")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        run_reference_arm(args)
        return

    import torch
    import pytorch_distributed_example_b200 as ga
    from pytorch_distributed_example_b200 import _lib

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU arm)")
    multi = args.gpus > 1
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if multi:
        # NCCL writes its banner / warnings to stdout by default; keep stdout for the one JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        world, hosted = dist.get_world_size(), 1
        assert world == args.gpus, "launch with torchrun --nproc-per-node %d" % args.gpus
        avg = ga.GradAverager(capacity_bytes=args.bytes)
    else:
        torch.cuda.set_device(0)
        world = args.world or 4
        hosted = world
        avg = ga.GradAverager(local_world=world, device=0, capacity_bytes=args.bytes)
        dist = None
    dev = torch.device("cuda", torch.cuda.current_device())
    S = args.bytes - args.bytes % 4
    n = S // 4
    names = {ga.ONESHOT: "oneshot", ga.TWOSHOT: "twoshot", ga.ONESHOT_PUSH: "oneshot_push", ga.TWOSHOT_PUSH: "twoshot_push",
             ga.TWOSHOT_PIPE: "twoshot_pipe", ga.HOSTED_DIRECT: "hosted_direct"}
    variant_id = {"auto": avg.pick_variant(S), "oneshot": ga.ONESHOT, "twoshot": ga.TWOSHOT,
                  "oneshot_push": ga.ONESHOT_PUSH, "twoshot_push": ga.TWOSHOT_PUSH,
                  "twoshot_pipe": ga.TWOSHOT_PIPE, "hosted_direct": ga.HOSTED_DIRECT}[args.variant]
    variant_name = names[variant_id]

    # gradient sets: rotate so that consecutive steps never find their inputs in L2
    n_sets = max(2, min(64, -(-(300 << 20) // (hosted * S))))
    gen = torch.Generator(device=dev).manual_seed(1234 + rank)
    sets = [[torch.randn(n, device=dev, generator=gen) for _ in range(hosted)] for _ in range(n_sets)]

    def step(i):
        avg.all_reduce_([[g] for g in sets[i % n_sets]], average=True, variant=variant_id)

    def barrier():
        if multi:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if not multi:
            return x
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    sampler = ClockSampler(index=torch.cuda.current_device()) if rank == 0 else None
    windows = []

    # ---- value: device-resident, K steps between two events, plus one event pair per launch ------
    stage("context ready: world=%d hosted=%d S=%d variant=%s sets=%d" % (world, hosted, S, variant_name, n_sets))
    for i in range(args.warmup):
        step(i)
    avg.check_status()                      # a peer that never arrives surfaces here, not as a hung bench
    barrier()
    stage("warm-up done, grid=%s" % (avg.last_grid,))
    launches0 = avg.launch_count
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    w0 = time.time()
    ev[0].record()
    for i in range(args.steps):
        step(i)
    ev[1].record()
    torch.cuda.synchronize()
    if multi:
        dist.barrier()
    windows.append((w0, time.time()))
    gpu_launches = avg.launch_count - launches0
    avg.check_status()
    stage("timed region done")
    total_ms = max_over_ranks(ev[0].elapsed_time(ev[1]))
    ms_per_step = total_ms / args.steps
    kernel_ms = ms_per_step          # one launch per step: average launch duration incl. the launch gap
    value = agg_bus_gbps(world, S, ms_per_step / 1e3)
    grid = avg.last_grid

    # ---- roofline of the one kernel -------------------------------------------------------------
    hbm_peak, peak_src = load_peaks()
    alg_bytes = algorithmic_hbm_bytes(world, S, variant_name, hosted)
    if multi:
        bus = (S / (kernel_ms / 1e3)) * bus_factor(world) / 1e9
        roofline = {"bound": "nvlink", "achieved": bus, "peak": 900.0, "unit": "GB/s", "frac": bus / 900.0,
                    "traffic": None, "kernel": "ga_fused_kernel<%d,%s,f32avg>" % (world, variant_name),
                    "kernel_ms": kernel_ms, "peak_source": "nominal NVLink 5 per direction per GPU (north_star); "
                    "measured peer copy on this pool is 770 GB/s (B200_PROFILING.md)",
                    "algorithmic": "busBW = S/t * 2(N-1)/N per GPU",
                    "hbm_side": {"bytes": alg_bytes, "gbps": alg_bytes / (kernel_ms / 1e3) / 1e9, "peak": hbm_peak}}
    else:
        ach = alg_bytes / (kernel_ms / 1e3) / 1e9
        roofline = {"bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                    "traffic": load_traffic(), "kernel": "ga_fused_kernel<%d,%s,f32avg>" % (world, variant_name),
                    "kernel_ms": kernel_ms, "peak_source": peak_src,
                    "algorithmic": "%d hosted ranks x %.2f x S bytes per launch (DESIGN.md kernel table)"
                                   % (hosted, alg_bytes / hosted / S)}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(world, args.gpus, S, variant_name, (n_sets, n_sets * hosted * S)),
        "bus_gbps_per_rank": value / world, "alg_gbps_per_rank": S / (ms_per_step / 1e3) / 1e9,
        "roofline": roofline, "gpu_launches": gpu_launches, "grid": {"blocks_per_rank": grid[0], "threads": grid[1]},
    }

    if not args.profile:
        # ---- e2e: host buffers through ga_average_host ------------------------------------------
        host_sets = [[torch.empty(n, dtype=torch.float32).pin_memory() for _ in range(hosted)] for _ in range(2)]
        for hs in host_sets:
            for t in hs:
                t.normal_()
        e2e_steps = max(3, min(args.steps, 20))
        for i in range(3):
            avg.average_host([[t] for t in host_sets[i % 2]], variant=variant_id)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0 = time.time()
        e0.record()
        for i in range(e2e_steps):
            avg.average_host([[t] for t in host_sets[i % 2]], variant=variant_id)
        e1.record()
        torch.cuda.synchronize()
        if multi:
            dist.barrier()
        windows.append((w0, time.time()))
        avg.check_status()
        stage("e2e done")
        e2e_ms = max_over_ranks(e0.elapsed_time(e1)) / e2e_steps
        line["e2e"] = {"value": agg_bus_gbps(world, S, e2e_ms / 1e3), "unit": UNIT, "ms_per_step": e2e_ms,
                       "steps": e2e_steps, "h2d_bytes_per_step": hosted * S, "d2h_bytes_per_step": hosted * S,
                       "api": "GradAverager.average_host -> ga_average_host (pinned host fp32 tensors)"}

        # ---- NCCL on a same-size buffer, same run (N > 1) ---------------------------------------
        if multi:
            buf = torch.randn(n, device=dev, generator=gen)
            for _ in range(5):
                dist.all_reduce(buf)
                buf.div_(float(world))
            barrier()
            k = max(10, min(args.steps, 100))
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(k):
                dist.all_reduce(buf)               # mnist/main.py:126 on NCCL
                buf.div_(float(world))             # mnist/main.py:127
            e1.record()
            torch.cuda.synchronize()
            stage("nccl comparison done")
            nccl_ms = max_over_ranks(e0.elapsed_time(e1)) / k
            line["nccl_same_buffer"] = {"ms_per_step": nccl_ms, "bus_gbps_per_rank": S / (nccl_ms / 1e3) * bus_factor(world) / 1e9,
                                        "what": "torch.distributed.all_reduce(SUM) + div_ on one %d-byte buffer" % S}

        if not args.no_extras and rank == 0 and not multi:
            line["extras"] = extras_single_gpu(avg, world, dev)
        if not args.no_extras:
            stage("mnist steps/sec leg")
            line["mnist_steps_per_sec"] = mnist_leg(avg, world, hosted, dev, dist if multi else None, rank)
        if multi and not args.no_extras:
            stage("overlap leg")
            line["resnet18_overlap"] = overlap_leg(avg, world, dev, dist)
        if multi and args.sweep:
            line["sweep"] = sweep_multi(avg, world, dev, dist, args.sweep_max)
        if multi and args.explore:
            line["explore"] = explore_pipe(avg, world, dev, dist)

    if sampler is not None:
        line["clocks"] = sampler.stop(windows)

    if not args.profile and not args.no_extras and rank == 0 and not multi:
        # CPU leg last, GPU idle, N = 1 only: the reference's averaging on host cores, bounded sample
        line["cpu_baseline"] = cpu_reference(world, S, steps=12, warmup=2, budget_s=20.0)

    if multi:
        dist.barrier()
    avg.close()
    if rank == 0:
        print(json.dumps(line), flush=True)
    if multi:
        dist.destroy_process_group()


def sweep_multi(avg, world, dev, dist, max_bytes):
    """Config 4: flat fp32 buffer, 4 KB * 4^k up to 1 GB plus the two real set sizes; per size the
    fused call (pack + reduce + unpack on a gradient tensor), the bucket-direct call (reduce only),
    both variants, and NCCL all_reduce + div_ on the same-size buffer.  Median of per-iteration
    max over ranks, device-timed."""
    import torch
    import pytorch_distributed_example_b200 as ga
    from pytorch_distributed_example_b200 import _lib
    sizes = [4 << 10, 16 << 10, 64 << 10, 256 << 10, 1 << 20, 1724320, 4 << 20, 16 << 20, 46758048, 64 << 20,
             256 << 20, 1 << 30]
    sizes = [s for s in sizes if s <= max_bytes]
    avg.reserve(max(sizes) + 4096)
    out = []

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
        return float(t[iters // 2]) * 1e3, float(t[0]) * 1e3          # median, best (us)

    # what the harness itself costs per iteration (Python call + event record, max over ranks): sizes
    # whose times sit near this floor are host-bound, not kernel-bound
    tiny = torch.zeros(8, device=dev)
    floor_med, floor_best = timed(lambda: tiny.add_(1.0), 50)
    out.append({"bytes": 0, "harness_floor_one_torch_launch": {"us": floor_med, "best_us": floor_best, "bus_gbps": 0.0}})
    for S in sizes:
        n = S // 4
        iters = 50 if S < (256 << 20) else 10
        # rotate two buffers so that no iteration finds its own input in L2 (>= 64 MiB: exceeds L2 anyway)
        bufs = [torch.randn(n, device=dev) for _ in range(2 if S >= (32 << 20) else 8)]
        k = [0]
        row = {"bytes": S}
        for vname, v in (("auto", ga.AUTO), ("oneshot", ga.ONESHOT), ("twoshot", ga.TWOSHOT), ("oneshot_push", ga.ONESHOT_PUSH),
                         ("twoshot_push", ga.TWOSHOT_PUSH), ("twoshot_pipe", ga.TWOSHOT_PIPE)):
            if vname.startswith("oneshot") and S > (64 << 20) and world > 2:
                continue
            if vname == "oneshot_push" and S * world > avg.capacity:
                continue

            def fused():
                k[0] += 1
                avg.all_reduce_([[bufs[k[0] % len(bufs)]]], average=True, variant=v)
            med, best = timed(fused, iters)
            row["fused_" + vname] = {"us": med, "best_us": best, "bus_gbps": S / med / 1e3 * bus_factor(world)}
            if vname == "auto":
                row["fused_auto"]["picked"] = avg.pick_variant(S)
            if vname.endswith("_push") or vname.endswith("_pipe") or vname == "auto":
                continue
            avg.bucket_acquire()

            def bucket():
                avg.allreduce_bucket(S, variant=v)
            med, best = timed(bucket, iters)
            row["bucket_" + vname] = {"us": med, "best_us": best, "bus_gbps": S / med / 1e3 * bus_factor(world)}

        def nccl():
            k[0] += 1
            b = bufs[k[0] % len(bufs)]
            dist.all_reduce(b)
            b.div_(float(world))
        med, best = timed(nccl, iters)
        row["nccl_allreduce_div"] = {"us": med, "best_us": best, "bus_gbps": S / med / 1e3 * bus_factor(world)}
        if S <= (4 << 20):
            # host-free latency: 20 back-to-back calls captured in one CUDA graph, replayed
            row["graph_auto"] = graph_latency(lambda b: avg.all_reduce_([[b]], average=True), bufs, dev, dist)
            row["graph_nccl_allreduce_div"] = graph_latency(lambda b: (dist.all_reduce(b), b.div_(float(world))), bufs, dev, dist)
        out.append(row)
        del bufs
        avg.check_status()
        stage("sweep %d bytes: %s" % (S, {k2: round(v2["us"], 1) for k2, v2 in row.items() if isinstance(v2, dict)}))
    return out


def mnist_leg(avg, world, hosted, dev, dist, rank, steps=200, batch=128):
    """BASELINE metric, second half: synthetic-MNIST steps/sec with the reference's step order
    (mnist/main.py:84-99: H2D of the batch, forward, zero_grad, backward, average, Adam step,
    loss.item()).  `ours` = fused averaging; `per_tensor_nccl` (N > 1) = the reference's own
    GPU path, one all_reduce + one div_ per parameter.  Hosted worlds run the `hosted` replicas
    one after another on the one GPU.  steps/sec counts optimizer steps of the whole job per
    second / world (i.e. synchronous data-parallel steps)."""
    import torch
    import torch.nn.functional as F
    sys.path.insert(0, os.path.join(ROOT, "examples"))
    import mnist_main

    def run(mode):
        nets, opts, data = [], [], []
        for l in range(hosted):
            torch.manual_seed(0)
            net = mnist_main.Net().to(dev)
            nets.append(net)
            opts.append(torch.optim.Adam(net.parameters(), lr=1e-3))
            data.append(mnist_main.SyntheticBatches(steps + 10, batch, rank + l))
        iters = [iter(d) for d in data]

        def one_step():
            for l in range(hosted):
                x, y = next(iters[l])
                x, y = x.to(dev, non_blocking=True), y.to(dev, non_blocking=True)
                loss = F.cross_entropy(nets[l](x), y)
                opts[l].zero_grad()
                loss.backward()
                if hosted == 1:
                    average(nets, mode)
                    opts[l].step()
                    loss.item()
            if hosted > 1:
                average(nets, mode)
                for l in range(hosted):
                    opts[l].step()
                loss.item()

        def average(nets, mode):
            if mode == "ours":
                if hosted == 1:
                    avg.average_gradients(nets[0].parameters())
                else:
                    avg.average_gradients_local([n.parameters() for n in nets])
            elif mode == "per_tensor_nccl":
                for p in nets[0].parameters():                       # mnist/main.py:125-127 on NCCL
                    dist.all_reduce(p.grad.data, op=dist.ReduceOp.SUM)
                    p.grad.data /= float(world)

        for _ in range(10):
            one_step()
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            one_step()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if dist is not None:
            t = torch.tensor([dt], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t)
        return steps / dt

    def run_graphed(mode):
        """Same step with forward, backward, averaging and Adam captured in ONE CUDA graph per
        replica set (possible because the averaging kernel keeps its flag epoch on the device):
        per step only the batch copy, the replay and loss.item() remain on the host."""
        nets, opts, data = [], [], []
        for l in range(hosted):
            torch.manual_seed(0)
            net = mnist_main.Net().to(dev)
            nets.append(net)
            opts.append(torch.optim.Adam(net.parameters(), lr=1e-3, capturable=True))
            data.append(mnist_main.SyntheticBatches(steps + 10, batch, rank + l))
        sx = [torch.zeros(batch, 1, 28, 28, device=dev) for _ in range(hosted)]
        sy = [torch.zeros(batch, dtype=torch.long, device=dev) for _ in range(hosted)]
        losses = [None] * hosted

        def body():
            for l in range(hosted):
                loss = F.cross_entropy(nets[l](sx[l]), sy[l])
                opts[l].zero_grad(set_to_none=False)
                loss.backward()
                losses[l] = loss
            if mode == "ours":
                if hosted == 1:
                    avg.average_gradients(nets[0].parameters())
                else:
                    avg.average_gradients_local([n.parameters() for n in nets])
            elif mode == "per_tensor_nccl":
                for p in nets[0].parameters():
                    dist.all_reduce(p.grad.data, op=dist.ReduceOp.SUM)
                    p.grad.data /= float(world)
            for l in range(hosted):
                opts[l].step()

        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                body()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            body()
        iters = [iter(d) for d in data]

        def one_step():
            for l in range(hosted):
                x, y = next(iters[l])
                sx[l].copy_(x, non_blocking=True)
                sy[l].copy_(y, non_blocking=True)
            graph.replay()
            return losses[hosted - 1].item()

        for _ in range(10):
            one_step()
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            one_step()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if dist is not None:
            t = torch.tensor([dt], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t)
        return steps / dt

    out = {"steps": steps, "batch_per_rank": batch, "world_size": world,
           "what": "mnist/main.py:84-99 step order, synthetic 128x1x28x28 batches from pinned host memory, Adam lr 1e-3"}
    out["ours"] = run("ours")
    if dist is not None:
        out["per_tensor_nccl"] = run("per_tensor_nccl")
    else:
        out["no_averaging_same_replicas"] = run("none")
    try:
        out["graphed_step_ours"] = run_graphed("ours")
        if dist is not None:
            out["graphed_step_per_tensor_nccl"] = run_graphed("per_tensor_nccl")
    except Exception as e:                                   # report, never fake
        out["graphed_step_error"] = "%s: %s" % (type(e).__name__, str(e)[:300])
    return out


def overlap_leg(avg, world, dev, dist, steps=30, batch=32):
    """BASELINE configs[4]: ResNet-18 (62 gradient tensors, 46.76 MB) forward + backward on
    randn(32,3,224,224); per-step averaging (a) absent, (b) one fused call after backward (the
    reference's call site), (c) bucketed and overlapped with backward from grad hooks.  Exposed
    averaging time = step time minus (a).  Device-timed, max over ranks."""
    import torch
    import torch.nn.functional as F
    import pytorch_distributed_example_b200 as ga
    try:
        import torchvision
        net = torchvision.models.resnet18(weights=None)
    except Exception as e:                                     # torchvision missing: report, do not fake
        return {"unavailable": "torchvision resnet18 not importable: %s" % e}
    torch.manual_seed(0)
    net = net.to(dev).to(memory_format=torch.channels_last)
    x = torch.randn(batch, 3, 224, 224, device=dev).contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 1000, (batch,), device=dev)
    params = [p for p in net.parameters()]
    ov = [None]

    def step(mode):
        for p in params:
            p.grad = None
        loss = F.cross_entropy(net(x), y)
        loss.backward()
        if mode == "after":
            avg.average_gradients(params)
        elif mode == "overlap":
            ov[0].average_gradients()

    def timed(mode):
        for _ in range(5):
            step(mode)
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            step(mode)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / steps], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    out = {"model": "torchvision resnet18, batch %d, fp32, channels_last" % batch,
           "grad_bytes": sum(p.numel() for p in params) * 4, "tensors": len(params)}
    out["ms_no_averaging"] = timed("none")
    out["ms_fused_after_backward"] = timed("after")
    ov[0] = ga.OverlappedAverager(net, averager=avg, bucket_cap_bytes=16 << 20, max_blocks=32)
    out["buckets"] = len(ov[0].buckets)
    out["ms_overlapped"] = timed("overlap")
    ov[0].remove()
    out["exposed_us_after_backward"] = (out["ms_fused_after_backward"] - out["ms_no_averaging"]) * 1e3
    out["exposed_us_overlapped"] = (out["ms_overlapped"] - out["ms_no_averaging"]) * 1e3
    avg.check_status()
    return out


def explore_pipe(avg, world, dev, dist):
    """Tuning grid: pipelined two-shot push over (slots per block per round) x (block cap)."""
    import torch
    import pytorch_distributed_example_b200 as ga
    from pytorch_distributed_example_b200 import _lib
    out = []
    avg.reserve((256 << 20) + 4096)

    def timed(fn, iters=30, warm=4):
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
        return float(t.sort().values[iters // 2]) * 1e3

    mid = os.environ.get("GA_EXPLORE_MID") == "1"
    for S in ((256 << 10, 1 << 20, 1724320, 4 << 20, 8 << 20) if mid else (16 << 20, 64 << 20, 256 << 20)):
        bufs = [torch.randn(S // 4, device=dev) for _ in range(3)]
        k = [0]
        row = {"bytes": S}
        for vname, v in (("twoshot_push", ga.TWOSHOT_PUSH), ("twoshot", ga.TWOSHOT)) + \
                ((("oneshot_push", ga.ONESHOT_PUSH),) if mid and S * world <= avg.capacity else ()):
            def f():
                k[0] += 1
                avg.all_reduce_([[bufs[k[0] % 3]]], average=True, variant=v)
            row[vname] = round(timed(f), 1)
        for blocks in ((0,) if mid else (0, 148, 74)):
            for group in ((1, 2, 4, 8, 16) if mid else (0, 1, 2, 4, 8, 16)):
                avg.set_option(_lib.OPT_MAX_BLOCKS, blocks)
                avg.set_option(_lib.OPT_PIPE_GROUP, group)

                def f():
                    k[0] += 1
                    avg.all_reduce_([[bufs[k[0] % 3]]], average=True, variant=ga.TWOSHOT_PIPE)
                row["pipe_b%d_g%d" % (blocks, group)] = round(timed(f), 1)
        avg.set_option(_lib.OPT_MAX_BLOCKS, 0)
        avg.set_option(_lib.OPT_PIPE_GROUP, 0)
        avg.check_status()
        stage("explore %d: %s" % (S, row))
        out.append(row)
        del bufs
    return out


def graph_latency(call, bufs, dev, dist, calls=20, replays=6):
    """us per call with the host out of the way: `calls` back-to-back invocations captured in one
    CUDA graph (possible for our kernels because their flag epoch lives on the device), replayed
    `replays` times; median replay time / calls, max over ranks."""
    import torch
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for i in range(3):
            call(bufs[i % len(bufs)])
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    dist.barrier()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(calls):
            call(bufs[i % len(bufs)])
    times = []
    for _ in range(replays):
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1) * 1e3 / calls)
    t = torch.tensor(times, device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    t = t.sort().values
    return {"us": float(t[len(times) // 2]), "best_us": float(t[0]), "bus_gbps": 0.0}


def load_traffic():
    """dram bytes per launch of the fused kernel from the committed ncu capture, if present."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get("dram_bytes_per_launch")
        except Exception:
            return None
    return None


def extras_single_gpu(avg, world, dev):
    """Secondary numbers on the same GPU: the two real gradient sets and a size sweep, both
    variants.  Device-timed, rotating buffers, 3 warm-ups each."""
    import torch
    import pytorch_distributed_example_b200 as ga
    from pytorch_distributed_example_b200 import workloads as shapes_only
    out = {}

    def time_set(shapes, variant, iters):
        numel = sum(shapes_only.numel(s) for s in shapes)
        nsets = max(2, min(48, -(-(300 << 20) // (world * numel * 4))))
        sets = [[[torch.randn(s, device=dev) for s in shapes] for _ in range(world)] for _ in range(nsets)]
        for i in range(3):
            avg.all_reduce_(sets[i % nsets], average=True, variant=variant)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(iters):
            avg.all_reduce_(sets[i % nsets], average=True, variant=variant)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        del sets
        return {"us": ms * 1e3, "bus_gbps_per_rank": numel * 4 / (ms / 1e3) * bus_factor(world) / 1e9}

    for name, shapes in (("mnist_cnn_1.72MB_8_tensors", shapes_only.MNIST_CNN_SHAPES),
                         ("resnet18_46.76MB_62_tensors", shapes_only.resnet18_shapes())):
        out[name] = {"oneshot": time_set(shapes, ga.ONESHOT, 200), "twoshot": time_set(shapes, ga.TWOSHOT, 200)}
    for name, shapes in list(out.items()):
        out[name]["hosted_direct"] = time_set(shapes_only.MNIST_CNN_SHAPES if name.startswith("mnist") else
                                              shapes_only.resnet18_shapes(), ga.HOSTED_DIRECT, 200)
    sweep = {}
    for kb in (4, 64, 1024, 16384, 65536):
        shp = [(kb * 256,)]
        sweep["%dKB" % kb] = {"hosted_direct": time_set(shp, ga.HOSTED_DIRECT, 100),
                              "oneshot": time_set(shp, ga.ONESHOT, 100), "twoshot": time_set(shp, ga.TWOSHOT, 100),
                              "twoshot_push": time_set(shp, ga.TWOSHOT_PUSH, 100)}
    out["sweep_world%d_hosted" % world] = sweep
    out["note"] = ("hosted_direct = no bucket (every rank's tensors addressed from one launch); oneshot/twoshot/"
                   "twoshot_push = the cross-GPU protocol kernels run on the hosted world, same code as N > 1")
    return out


if __name__ == "__main__":
    main()
