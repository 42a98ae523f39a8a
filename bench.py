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

_JSON_OUT = sys.stdout
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


def cpu_train_reference(world, steps=40):
    """BASELINE configs[1]: mnist/main.py:84-99 on gloo, `world` CPU processes on the box's host cores,
    synthetic 128 x 1 x 28 x 28 batches; also the same step at world_size 1 (no averaging, :193)."""
    from oracle import gloo_path
    cores = os.cpu_count() or 1
    threads = max(1, cores // world)
    t0 = time.time()
    out = {"world_size": world, "cores": min(cores, world * threads), "threads_per_rank": threads, "steps": steps, "batch_per_rank": 128,
           "steps_per_sec": gloo_path.time_train_loop(world, steps=steps, warmup=4, batch=128, threads=threads),
           "steps_per_sec_world1_no_averaging": gloo_path.time_train_loop(1, steps=steps, warmup=4, batch=128, threads=threads),
           "what": "oracle/gloo_path.py:_train_job = mnist/main.py:84-99 restated (RefNet = :130-147), gloo tcp://127.0.0.1, "
                   "host has %d cores%s" % (cores, ", oversubscribed" if world > cores else "")}
    out["wall_s"] = time.time() - t0
    return out


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
        "config": workload_config(world, args.gpus, args.bytes), "variant": "gloo ring (torch)", "l2": None,
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": time.time() - t0,
    }
    print(json.dumps(line), file=_JSON_OUT, flush=True)


def workload_config(world, gpus, nbytes):
    """The same dict in both arms (the driver compares it): arm-specific facts -- which kernel ran,
    how L2 was defeated -- are top-level keys of the line, not part of the workload."""
    return {
        "workload": "gradient averaging of one %d-byte fp32 gradient per rank, world_size=%d (%s); "
                    "BASELINE configs[1] world size on 1 GPU, configs[3] 64 MiB bucket point"
                    % (nbytes, world, "%d ranks hosted per GPU" % (world // gpus) if gpus == 1 else "one rank per GPU"),
        "bytes_per_rank": nbytes, "world_size": world, "ranks_per_gpu": world // gpus if gpus == 1 else 1,
        "value_formula": "world * (S/t) * 2(world-1)/world",
        "timed_call": "Trainer.average_gradients (mnist/main.py:122-127) on one gradient tensor per rank",
    }


def l2_note(n_sets, total_bytes):
    return "gradient buffers rotate over %d sets (%d MiB > L2) between steps" % (n_sets, total_bytes >> 20)


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


NVLINK_NOMINAL = 900.0
# profiles/r02_nvlink_probe.md: what this pool's fabric carries, per GPU and direction, both directions busy
NVLINK_MEASURED = {"sm_stores_all_to_all_8gpu_gbps": 649.0, "sm_stores_2gpu_gbps": 665.0, "copy_engine_2gpu_gbps": 709.0,
                   "copy_engine_all_to_all_8gpu_gbps": 404.0, "multimem_st_ingress_8gpu_gbps": 697.0,
                   "guide_peer_copy_gbps": 770.0, "source": "tools/nvlink_probe.cu (profiles/r02_nvlink_probe.md); B200_PROFILING.md for the guide figure"}


class _P:
    """What `Trainer.average_gradients` iterates over: an object with `.grad` (mnist/main.py:125-127)."""
    __slots__ = ("grad",)

    def __init__(self, g):
        self.grad = g


def nvlink_counters(index):
    """(tx_bytes, rx_bytes) summed over the GPU's links from `nvidia-smi nvlink -gt d`, or None."""
    import re
    try:
        out = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d", "-i", str(index)], capture_output=True, text=True, timeout=20).stdout
        tx = [int(x) for x in re.findall(r"Data Tx:\s*(\d+)\s*KiB", out)]
        rx = [int(x) for x in re.findall(r"Data Rx:\s*(\d+)\s*KiB", out)]
        if not tx or not rx:
            return None
        return sum(tx) * 1024, sum(rx) * 1024
    except Exception:
        return None


def view_averager(nbytes, local_world):
    """A context with a view region of `nbytes`: with an NVSwitch multicast mapping where the box allows it
    (GPU support and descriptor passing between the rank processes), else shared through CUDA-IPC alone.
    Every rank sees the same failure, so the retry stays collective."""
    import pytorch_distributed_example_b200 as ga
    for multicast in (True, False):
        a = ga.GradAverager(local_world=local_world, device=0, capacity_bytes=nbytes) if local_world else ga.GradAverager(capacity_bytes=nbytes)
        try:
            a.create_view(nbytes, multicast=multicast)
            return a
        except ga.GradAvgError as e:
            stage("view region with multicast=%s failed (%s)%s" % (multicast, e, ": retrying without" if multicast else ""))
            try:
                a.close()
            except Exception:
                pass
            if not multicast:
                raise


def modes_leg(avg, vavg, world, hosted, S, sets, dev, multi, timed_loop, nvlink_roofline, hbm_peak):
    """The same averaging through the library's other modes, 64 MiB per rank, each with its own roofline:
      dropin_*   Trainer.average_gradients on ordinary gradient tensors (pack + exchange + unpack), other variants
      view       gradients adopted into the peer-mapped region: in place, no pack / unpack; bit-exact
      nvls       opt-in NVSwitch multimem: in-switch sum (correctly rounded, NOT rank order)"""
    import torch
    import pytorch_distributed_example_b200 as ga
    from pytorch_distributed_example_b200 import _lib
    out, more = {}, []
    n_sets = len(sets)
    k, w = 60, 5
    params = [[[_P(g)] for g in st] for st in sets]

    def call(variant):
        def f(i):
            ps = params[i % n_sets]
            if hosted == 1:
                avg.average_gradients(ps[0], variant=variant)
            else:
                avg.average_gradients_local(ps, variant=variant)
        return f

    if multi:
        for v in ("twoshot_pipe", "twoshot_push", "twoshot"):
            ms, _ = timed_loop(call(v), k, w)
            out["dropin_" + v] = {"ms_per_step": ms, "bus_gbps_per_rank": S / (ms / 1e3) * bus_factor(world) / 1e9, "bit_exact": True}
    else:
        for v in ("twoshot", "oneshot", "twoshot_pipe"):
            ms, _ = timed_loop(call(v), k, w)
            alg = algorithmic_hbm_bytes(world, S, v, hosted)
            out["protocol_" + v] = {"ms_per_step": ms, "hbm_gbps": alg / (ms / 1e3) / 1e9, "bit_exact": True}
            more.append({"bound": "hbm", "achieved": alg / (ms / 1e3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                         "frac": alg / (ms / 1e3) / 1e9 / hbm_peak, "traffic": None,
                         "kernel": {"twoshot": "ga_fused_kernel<%d,2,f32avg>", "oneshot": "ga_fused_kernel<%d,1,f32avg>",
                                    "twoshot_pipe": "ga_push_pipe_kernel<%d,f32avg>"}[v] % world,
                         "kernel_ms": ms, "what": "the cross-GPU protocol kernel (pack -> exchange -> unpack, flags) run by the "
                                                  "hosted world on this one GPU: the code path N > 1 ships; a fraction above 1 means "
                                                  "the 126 MB L2 served part of the algorithmic bytes (hosted ranks re-read each other's buckets)",
                         "algorithmic": "%d hosted ranks x %.2f x S bytes per launch" % (hosted, alg / hosted / S)})
    # view mode: every hosted rank's gradient lives in its view region
    views = [vavg.view_tensor(S, l) for l in range(hosted)]
    for v_, g in zip(views, sets[0]):
        v_.copy_(g)
    ms, _ = timed_loop(lambda i: vavg.allreduce_view(S, variant="view"), k, w)
    out["view"] = {"ms_per_step": ms, "bit_exact": True, "grid": vavg.last_grid[0],
                   "what": "GradAverager.adopt(): p.grad are views of the peer-mapped region, ga_average_view in place"}
    if multi:
        out["view"]["bus_gbps_per_rank"] = S / (ms / 1e3) * bus_factor(world) / 1e9
        more.append(nvlink_roofline(ms, "ga_view_kernel<%d,f32avg>" % world, None))
    else:
        alg = hosted * (2.0 * (world - 1) / world + 2.0) * S         # scatter: read + write (W-1)/W * S; fold: read S, write S
        out["view"]["hbm_gbps"] = alg / (ms / 1e3) / 1e9
        vt = None
        try:
            vt = json.load(open(os.path.join(ROOT, "profiles", "traffic_view.json"))).get("dram_bytes_per_launch")
        except Exception:
            pass
        more.append({"bound": "hbm", "achieved": alg / (ms / 1e3) / 1e9, "peak": hbm_peak, "unit": "GB/s", "frac": alg / (ms / 1e3) / 1e9 / hbm_peak,
                     "traffic": vt, "kernel": "ga_view_kernel<%d,f32avg>" % world, "kernel_ms": ms,
                     "algorithmic": "%d hosted ranks x (2(W-1)/W scatter + 2 fold) x S bytes" % hosted})
    if multi and vavg.has_multicast:
        ms, _ = timed_loop(lambda i: vavg.allreduce_view(S, variant="nvls"), k, w)
        r = nvlink_roofline(ms, "ga_nvls_kernel<f32avg>", None)
        r["algorithmic"] += "; the in-switch form moves (N+1)/N * S per direction, the bus formula is kept for comparison with nccl-tests"
        more.append(r)
        out["nvls"] = {"ms_per_step": ms, "bus_gbps_per_rank": S / (ms / 1e3) * bus_factor(world) / 1e9, "bit_exact": False,
                       "what": "opt-in GA_VARIANT_NVLS: multimem.ld_reduce + multimem.st; correctly rounded exact sum, not the rank-order fold"}
    avg.check_status()
    vavg.check_status()
    return out, more


def parity_leg(avg, vavg, world, rank, S, headline_grad, dev, dist):
    """N > 1, outside the timed region: every variant AUTO can pick (and the view / NVLS modes) on the
    seeded MNIST-CNN gradient set against the committed sha256 of the rank-order fold
    (tests/golden/mnist_cnn.json: written by tests/golden/make_golden.py next to the REAL reference's
    output), and one 64 MiB step against a sequential fp32 fold of the all-gathered inputs computed with
    torch on the GPU.  Digests are all-gathered: every rank is checked, and must hold the same bytes."""
    import hashlib
    import numpy as np
    import torch
    from pytorch_distributed_example_b200 import workloads
    out = {"world": world}
    shapes = workloads.MNIST_CNN_SHAPES
    g = torch.Generator().manual_seed(1234 + rank)                 # the seeding of oracle/gradient_sets.py and the golden files
    host = [torch.randn(s, generator=g, dtype=torch.float32) for s in shapes]
    golden = None
    try:
        golden = json.load(open(os.path.join(ROOT, "tests", "golden", "mnist_cnn.json")))["worlds"][str(world)]["signed"]["fold_sha256"]
    except Exception:
        pass

    def digest(tensors):
        flat = np.concatenate([t.detach().cpu().numpy().ravel() for t in tensors])
        return hashlib.sha256(flat.tobytes()).digest()

    def gather_digests(d):
        t = torch.tensor(list(d), dtype=torch.uint8, device=dev)
        outl = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(outl, t)
        return [bytes(o.cpu().tolist()) for o in outl]

    def torch_fold(x, wide=False):
        """((g0 + g1) + g2) + ... then one IEEE division, on the GPU; wide=True: binary64 sum rounded once (NVLS oracle)."""
        gathered = [torch.empty_like(x) for _ in range(world)]
        dist.all_gather(gathered, x)
        acc = gathered[0].double() if wide else gathered[0].clone()
        for t in gathered[1:]:
            acc = acc + (t.double() if wide else t)
        acc = acc.float() if wide else acc
        scale = torch.stack([t.abs() for t in gathered]).mean(0)
        return acc / torch.full((), float(world), device=dev), scale

    ok_all, identical = True, True
    per_variant = {}
    flat_in = torch.cat([t.ravel() for t in host]).to(dev)
    want_fold, _ = torch_fold(flat_in)
    for variant in ("auto", "oneshot_push", "twoshot_push", "twoshot_pipe", "oneshot", "twoshot"):
        params = [_P(t.to(dev)) for t in host]
        avg.average_gradients(params, variant=variant)
        avg.check_status()
        ds = gather_digests(digest([p.grad for p in params]))
        same = all(d == ds[0] for d in ds)
        good = (ds[0].hex() == golden) if golden else bool(torch.equal(torch.cat([p.grad.ravel() for p in params]).view(torch.int32), want_fold.view(torch.int32)))
        per_variant[variant] = bool(good and same)
        ok_all &= good
        identical &= same
    # view mode on the same set (flat layout), then NVLS against its own oracle
    nb = flat_in.numel() * 4
    v = vavg.view_tensor(nb)
    v.copy_(flat_in)
    vavg.allreduce_view(nb, variant="view")
    vavg.check_status()
    ds = gather_digests(digest([v]))
    per_variant["view"] = bool(all(d == ds[0] for d in ds) and ((ds[0].hex() == golden) if golden else torch.equal(v.view(torch.int32), want_fold.view(torch.int32))))
    ok_all &= per_variant["view"]
    out["mnist_sha_ok"] = bool(ok_all)
    out["mnist_per_variant"] = per_variant
    out["golden"] = "tests/golden/mnist_cnn.json fold_sha256" if golden else "torch rank-order fold on the GPU (world size not in the golden file)"
    # headline buffer: one step, AUTO and view, bit for bit against the sequential fold
    x = headline_grad.clone()
    want, scale = torch_fold(x)
    p = [_P(x.clone())]
    avg.average_gradients(p)
    exact_auto = bool(torch.equal(p[0].grad.view(torch.int32), want.view(torch.int32)))
    hv = vavg.view_tensor(S)
    hv.copy_(x)
    vavg.allreduce_view(S, variant="view")
    exact_view = bool(torch.equal(hv.view(torch.int32), want.view(torch.int32)))
    ds = gather_digests(hashlib.sha256(hv.cpu().numpy().tobytes()).digest())
    identical &= all(d == ds[0] for d in ds)
    out["headline_bit_exact"] = bool(exact_auto and exact_view)
    out["headline_bit_exact_by_mode"] = {"auto": exact_auto, "view": exact_view}
    out["ranks_identical"] = bool(identical)
    if vavg.has_multicast:
        wide, _ = torch_fold(x, wide=True)
        hv.copy_(x)
        vavg.allreduce_view(S, variant="nvls")
        vavg.check_status()
        eq = float((hv.view(torch.int32) == wide.view(torch.int32)).float().mean())
        err = float(((hv - want).abs() / scale.clamp_min(1e-30)).max())
        ds = gather_digests(hashlib.sha256(hv.cpu().numpy().tobytes()).digest())
        out["nvls"] = {"equal_to_wide_sum_oracle": eq, "max_abs_err_vs_rank_order_over_mean_abs_input": err,
                       "ranks_identical": all(d == ds[0] for d in ds), "gate": "err <= 1e-6 (north_star's fp32 tolerance, SURVEY F3 metric); opt-in, not bit-exact"}
        out["nvls_ok"] = bool(err <= 1e-6 and out["nvls"]["ranks_identical"])
    out["ok"] = bool(out["mnist_sha_ok"] and out["headline_bit_exact"] and out["ranks_identical"] and out.get("nvls_ok", True))
    return out


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
    # stdout carries the ONE JSON line: whatever libraries print (NCCL's version banner, warnings) is sent to stderr
    global _JSON_OUT
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
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
    n_sets = max(2, min(16, -(-(300 << 20) // (hosted * S))))
    gen = torch.Generator(device=dev).manual_seed(1234 + rank)
    sets = [[torch.randn(n, device=dev, generator=gen) for _ in range(hosted)] for _ in range(n_sets)]
    # what Trainer.average_gradients receives: parameters whose .grad is the gradient (mnist/main.py:125-127)
    param_sets = [[[_P(g)] for g in st] for st in sets]

    def step(i):
        ps = param_sets[i % n_sets]
        if hosted == 1:
            avg.average_gradients(ps[0], variant=variant_id)          # GradAverager._run_params -> ga_average
        else:
            avg.average_gradients_local(ps, variant=variant_id)

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

    def timed_loop(fn, steps, warmup):
        """K calls between two CUDA events on the launching stream, barrier + synchronize on both sides,
        max over ranks; returns (ms per call, wall-clock window)."""
        for i in range(warmup):
            fn(i)
        avg.check_status()                  # a peer that never arrives surfaces here, not as a hung bench
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0 = time.time()
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        if multi:
            dist.barrier()
        w1 = time.time()
        avg.check_status()
        return max_over_ranks(e0.elapsed_time(e1)) / steps, (w0, w1)

    sampler = ClockSampler(index=torch.cuda.current_device()) if rank == 0 else None
    windows = []

    # ---- value: device-resident, K steps between two events ---------------------------------------
    stage("context ready: world=%d hosted=%d S=%d variant=%s sets=%d" % (world, hosted, S, variant_name, n_sets))
    launches0 = avg.launch_count
    nv0 = nvlink_counters(torch.cuda.current_device()) if multi and rank == 0 else None
    ms_per_step, win = timed_loop(step, args.steps, args.warmup)
    nv1 = nvlink_counters(torch.cuda.current_device()) if multi and rank == 0 else None
    windows.append(win)
    gpu_launches = avg.launch_count - launches0 - args.warmup
    stage("timed region done: %.1f us per step, grid=%s" % (ms_per_step * 1e3, avg.last_grid))
    kernel_ms = ms_per_step          # one launch per step: average launch duration incl. the launch gap
    value = agg_bus_gbps(world, S, ms_per_step / 1e3)
    grid = avg.last_grid

    # ---- roofline of the one kernel -------------------------------------------------------------
    hbm_peak, peak_src = load_peaks()
    alg_bytes = algorithmic_hbm_bytes(world, S, variant_name, hosted)
    kernel_names = {"oneshot": "ga_fused_kernel<%d,1,f32avg>", "twoshot": "ga_fused_kernel<%d,2,f32avg>",
                    "oneshot_push": "ga_push1_kernel<%d,f32avg>", "twoshot_push": "ga_push_kernel<%d,f32avg>",
                    "twoshot_pipe": "ga_push_pipe_kernel<%d,f32avg>", "hosted_direct": "ga_hosted_direct_kernel<%d,f32avg>",
                    "view": "ga_view_kernel<%d,f32avg>"}

    def nvlink_roofline(ms, kernel, traffic):
        bus = (S / (ms / 1e3)) * bus_factor(world) / 1e9
        return {"bound": "nvlink", "achieved": bus, "peak": NVLINK_NOMINAL, "unit": "GB/s", "frac": bus / NVLINK_NOMINAL,
                "traffic": traffic, "kernel": kernel, "kernel_ms": ms,
                "peak_source": "nominal NVLink 5 per direction per GPU (north_star)",
                "peak_measured": NVLINK_MEASURED, "frac_of_measured": bus / NVLINK_MEASURED["sm_stores_all_to_all_8gpu_gbps"],
                "algorithmic": "busBW = S/t * 2(N-1)/N per GPU; link bytes per GPU and direction 2(N-1)/N * S = %d"
                               % int(S * bus_factor(world))}

    if multi:
        traffic = None
        if nv0 and nv1:
            k = args.steps + args.warmup
            traffic = {"tx_bytes_per_launch": (nv1[0] - nv0[0]) / k, "rx_bytes_per_launch": (nv1[1] - nv0[1]) / k,
                       "source": "nvidia-smi nvlink -gt d on rank 0's GPU before / after the warm-up + timed launches",
                       "algorithmic_per_direction": S * bus_factor(world)}
        roofline = nvlink_roofline(kernel_ms, kernel_names[variant_name] % world, traffic)
        roofline["hbm_side"] = {"bytes": alg_bytes, "gbps": alg_bytes / (kernel_ms / 1e3) / 1e9, "peak": hbm_peak}
    else:
        ach = alg_bytes / (kernel_ms / 1e3) / 1e9
        roofline = {"bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                    "traffic": load_traffic(), "kernel": kernel_names[variant_name] % world,
                    "kernel_ms": kernel_ms, "peak_source": peak_src,
                    "algorithmic": "%d hosted ranks x %.2f x S bytes per launch (DESIGN.md kernel table)"
                                   % (hosted, alg_bytes / hosted / S)}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(world, args.gpus, S), "variant": variant_name, "l2": l2_note(n_sets, n_sets * hosted * S),
        "bus_gbps_per_rank": value / world, "alg_gbps_per_rank": S / (ms_per_step / 1e3) / 1e9,
        "roofline": roofline, "gpu_launches": gpu_launches, "grid": {"blocks_per_rank": grid[0], "threads": grid[1]},
    }

    if not args.profile:
        # ---- the other ways to make the same call (same gradients, same result bits unless said otherwise) ----
        stage("modes leg")
        # the view region pins its context's size, so the in-place modes get a context of their own
        vavg = view_averager(S, world if not multi else 0)
        line["modes"], more = modes_leg(avg, vavg, world, hosted, S, sets, dev, multi, timed_loop, nvlink_roofline, hbm_peak)
        line["roofline_more"] = more
        if multi:
            stage("parity leg")
            line["parity"] = parity_leg(avg, vavg, world, rank, S, sets[0][0], dev, dist)
        gpu_launches_other = vavg.launch_count
        if multi:
            dist.barrier()
        vavg.close()
        line["gpu_launches_outside_timed_region"] = gpu_launches_other

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
        # the floor under it: the same bytes copied in and out at once (two streams), no kernel, every rank at the same time
        scratch = [torch.empty(n, device=dev) for _ in range(hosted)]
        s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()

        def bare(i):
            cur = torch.cuda.current_stream()
            s_in.wait_stream(cur)
            s_out.wait_stream(cur)
            with torch.cuda.stream(s_in):
                for t, d_ in zip(host_sets[0], scratch):
                    d_.copy_(t, non_blocking=True)
            with torch.cuda.stream(s_out):
                for t, g_ in zip(host_sets[1], sets[0]):
                    t.copy_(g_, non_blocking=True)
            cur.wait_stream(s_in)
            cur.wait_stream(s_out)
        bare_ms, _ = timed_loop(bare, 6, 2)
        del scratch
        line["e2e"] = {"value": agg_bus_gbps(world, S, e2e_ms / 1e3), "unit": UNIT, "ms_per_step": e2e_ms,
                       "steps": e2e_steps, "h2d_bytes_per_step": hosted * S, "d2h_bytes_per_step": hosted * S,
                       "api": "GradAverager.average_host -> ga_average_host (pinned host fp32 tensors)",
                       "bare_copy_floor_ms": bare_ms,
                       "floor_what": "H2D and D2H of the same bytes at once on two streams, no kernel, all %d rank(s) at the same "
                                     "time (PCIe and host memory are shared by the GPUs of the box)" % world}

        # ---- NCCL on a same-size buffer, same run (N > 1): all_reduce alone, and with the reference's div_ ----
        if multi:
            buf = torch.randn(n, device=dev, generator=gen)
            k = max(10, min(args.steps, 100))
            nccl_ms, _ = timed_loop(lambda i: dist.all_reduce(buf), k, 5)                          # mnist/main.py:126 on NCCL
            nccl_div_ms, _ = timed_loop(lambda i: (dist.all_reduce(buf), buf.div_(float(world))), k, 5)   # + mnist/main.py:127
            stage("nccl comparison done")
            line["nccl_same_buffer"] = {
                "allreduce_only": {"ms_per_step": nccl_ms, "bus_gbps_per_rank": S / (nccl_ms / 1e3) * bus_factor(world) / 1e9},
                "allreduce_plus_div": {"ms_per_step": nccl_div_ms, "bus_gbps_per_rank": S / (nccl_div_ms / 1e3) * bus_factor(world) / 1e9},
                "ms_per_step": nccl_div_ms,
                "what": "torch.distributed.all_reduce(SUM) on one %d-byte buffer (in place), alone and followed by div_ "
                        "(what the reference's loop does per tensor); NCCL's sum order is its own (NVLS / ring), not rank order" % S}

        if not args.no_extras and rank == 0 and not multi:
            line["extras"] = extras_single_gpu(avg, world, dev)
        if not args.no_extras:
            stage("mnist steps/sec leg")
            line["mnist_steps_per_sec"] = mnist_leg(avg, world, hosted, dev, dist if multi else None, rank)
        if multi and not args.no_extras:
            stage("overlap leg")
            line["resnet18_overlap"] = overlap_leg(avg, world, dev, dist)
        if multi and args.sweep:
            line["sweep"] = sweep_multi(avg, world, dev, dist, args.sweep_max, rank)
        if multi and args.explore:
            line["explore"] = explore_pipe(avg, world, dev, dist)

    if sampler is not None:
        line["clocks"] = sampler.stop(windows)

    if not args.profile and not args.no_extras and rank == 0 and not multi:
        # CPU leg last, GPU idle, N = 1 only: the reference's averaging on host cores, bounded sample
        line["cpu_baseline"] = cpu_reference(world, S, steps=12, warmup=2, budget_s=20.0)
        # second half of the metric on the same host cores: the reference's training step on gloo (configs[1])
        if isinstance(line.get("mnist_steps_per_sec"), dict):
            line["mnist_steps_per_sec"]["cpu_reference_gloo"] = cpu_train_reference(world)

    if multi:
        dist.barrier()
    avg.close()
    if rank == 0:
        print(json.dumps(line), file=_JSON_OUT, flush=True)
    if multi:
        dist.destroy_process_group()
    if line.get("parity") is not None and not line["parity"]["ok"]:
        raise SystemExit("bench.py: PARITY FAILED at N=%d: %s" % (args.gpus, json.dumps(line["parity"])))


def sweep_multi(avg, world, dev, dist, max_bytes, rank=0):
    """Config 4: flat fp32 buffer, 4 KB * 4^k up to 1 GB plus the two real set sizes.  Per size:
      fused_*    Trainer.average_gradients on a gradient tensor (pack + exchange + unpack), every variant and AUTO
      bucket_*   the bucket-direct call (reduce only, input already in the bucket)
      view*      in place on the peer-mapped view region (GA_VARIANT_VIEW; view_auto = what adopt() runs)
      nvls       opt-in in-switch variant (not rank order)
      nccl_*     torch.distributed.all_reduce alone, and + div_ (the reference's per-tensor pair), same size, in place
      gloo       the reference's CPU path on the box's host cores (one gloo session over all sizes, rank 0's host)
    Median of per-iteration max over ranks, device-timed."""
    import torch
    import pytorch_distributed_example_b200 as ga
    from pytorch_distributed_example_b200 import _lib
    sizes = [4 << 10, 16 << 10, 64 << 10, 256 << 10, 1 << 20, 1724320, 4 << 20, 16 << 20, 46758048, 64 << 20,
             256 << 20, 1 << 30]
    sizes = [s for s in sizes if s <= max_bytes]
    avg.reserve(max(sizes) + 4096)
    vavg = view_averager(max(sizes) + 4096, 0)
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

    def cell(med, best, S):
        return {"us": med, "best_us": best, "bus_gbps": S / med / 1e3 * bus_factor(world)}

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
            row["fused_" + vname] = cell(*timed(fused, iters), S)
            if vname == "auto":
                row["fused_auto"]["picked"] = avg.pick_variant(S)
            if vname.endswith("_push") or vname.endswith("_pipe") or vname == "auto":
                continue
            avg.bucket_acquire()

            def bucket():
                avg.allreduce_bucket(S, variant=v)
            row["bucket_" + vname] = cell(*timed(bucket, iters), S)
        vt = vavg.view_tensor(S)
        vt.copy_(bufs[0])
        row["view"] = cell(*timed(lambda: vavg.allreduce_view(S, variant="view"), iters), S)
        row["view_auto"] = cell(*timed(lambda: vavg.allreduce_view(S), iters), S)
        if vavg.has_multicast:
            row["nvls"] = cell(*timed(lambda: vavg.allreduce_view(S, variant="nvls"), iters), S)

        def nccl():
            k[0] += 1
            dist.all_reduce(bufs[k[0] % len(bufs)])

        def nccl_div():
            k[0] += 1
            b = bufs[k[0] % len(bufs)]
            dist.all_reduce(b)
            b.div_(float(world))
        row["nccl_allreduce"] = cell(*timed(nccl, iters), S)
        row["nccl_allreduce_div"] = cell(*timed(nccl_div, iters), S)
        if S <= (4 << 20):
            # host-free latency: 20 back-to-back calls captured in one CUDA graph, replayed
            row["graph_auto"] = graph_latency(lambda b: avg.all_reduce_([[b]], average=True), bufs, dev, dist)
            row["graph_nccl_allreduce"] = graph_latency(lambda b: dist.all_reduce(b), bufs, dev, dist)
            row["graph_nccl_allreduce_div"] = graph_latency(lambda b: (dist.all_reduce(b), b.div_(float(world))), bufs, dev, dist)
        out.append(row)
        del bufs
        avg.check_status()
        vavg.check_status()
        stage("sweep %d bytes: %s" % (S, {k2: round(v2["us"], 1) for k2, v2 in row.items() if isinstance(v2, dict)}))
    dist.barrier()
    vavg.close()
    # the reference's CPU path on the host cores, every size, while the GPUs idle (rank 0's process only)
    if rank == 0:
        try:
            from oracle import gloo_path
            cores = os.cpu_count() or 1
            threads = max(1, cores // world)
            gl = [s for s in sizes if s <= (256 << 20) or world <= 4]
            secs = gloo_path.time_flat_average_sizes(world, [s // 4 for s in gl], iters=5, warmup=1, threads=threads)
            for row in out:
                if row["bytes"] in gl:
                    t = secs[gl.index(row["bytes"])]
                    row["gloo_cpu"] = {"us": t * 1e6, "bus_gbps": row["bytes"] / t / 1e9 * bus_factor(world), "cores": min(cores, world * threads),
                                       "what": "oracle/gloo_path.py (mnist/main.py:122-127 restated), %d procs x %d threads" % (world, threads)}
        except Exception as e:                                  # report, never fake
            out.append({"gloo_cpu_error": "%s: %s" % (type(e).__name__, str(e)[:300])})
    dist.barrier()
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


def overlap_leg(avg, world, dev, dist, steps=20, batch=128):
    """BASELINE configs[4]: ResNet-18 (62 gradient tensors, 46.76 MB) forward + backward on
    randn(128,3,224,224) per rank (batch 128 so that backward, not the host, sets the step time);
    per-step averaging (a) absent, (b) one fused call after backward (the reference's call site),
    (c) bucketed and overlapped with backward from grad hooks, forced on, (d) the same object left to
    decide for itself (it refuses when the exchange it could hide is shorter than what the hooks cost).
    Exposed averaging time = step time minus (a).  Device-timed, max over ranks."""
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
        for _ in range(4):
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

    out = {"model": "torchvision resnet18, batch %d per rank, fp32, channels_last" % batch,
           "grad_bytes": sum(p.numel() for p in params) * 4, "tensors": len(params)}
    out["ms_no_averaging"] = timed("none")
    out["ms_fused_after_backward"] = timed("after")
    ov[0] = ga.OverlappedAverager(net, averager=avg, bucket_cap_bytes=16 << 20, max_blocks=32, force=True)
    out["buckets"] = len(ov[0].buckets)
    out["ms_overlapped_forced"] = timed("overlap")
    ov[0].remove()
    ov[0] = ga.OverlappedAverager(net, averager=avg, bucket_cap_bytes=16 << 20, max_blocks=32)
    out["self_gating"] = {"enabled": ov[0].enabled, "reason": ov[0].reason}
    out["ms_overlapped"] = timed("overlap")
    ov[0].remove()
    out["ms_no_averaging_again"] = timed("none")               # the step itself drifts: bracket it
    base = 0.5 * (out["ms_no_averaging"] + out["ms_no_averaging_again"])
    out["exposed_us_after_backward"] = (out["ms_fused_after_backward"] - base) * 1e3
    out["exposed_us_overlapped_forced"] = (out["ms_overlapped_forced"] - base) * 1e3
    out["exposed_us_overlapped"] = (out["ms_overlapped"] - base) * 1e3
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
