"""The reference hot path on its own CPU backend (TEST INFRASTRUCTURE).

`average_gradients_ref` restates `/root/reference/mnist/main.py:122-127` statement for
statement on `torch.distributed`; `toy_allreduce_ref` restates `/root/reference/toy/main.py:15-20`.
`run_world` starts `world` CPU processes on 127.0.0.1 with the gloo backend, the way the
reference READMEs do by hand (`mnist/README.md:18-33`), and returns every rank's result.

The arithmetic below `all_reduce` is gloo's, statically linked into the installed torch wheel
(`third_party/gloo/gloo/allreduce.cc`; effective pin: torch 2.11.0+cu128 of this image -- the
reference itself pins only `pytorch>=0.4.0`, README.md:5).  This file is also what
`bench.py` times as `cpu_baseline` (kind "port") and under `--impl reference`.
"""
import os
import socket
import time

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def average_gradients_ref(params, world_size=None):
    """mnist/main.py:122-127."""
    if world_size is None:
        world_size = dist.get_world_size()            # :123
    for p in params:                                  # :125
        dist.all_reduce(p.grad.data, op=dist.ReduceOp.SUM)   # :126 (reduce_op is the old alias)
        p.grad.data /= float(world_size)              # :127


def toy_allreduce_ref(value, world_size):
    """toy/main.py:15-20: new group of all ranks, IntTensor([value]), all_reduce SUM."""
    group = dist.new_group(ranks=list(range(world_size)))   # :15-16
    tensor = torch.IntTensor([value])                        # :19
    dist.all_reduce(tensor, op=dist.ReduceOp.SUM, group=group)  # :20
    return float(tensor)                                     # :23


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


_LAUNCHER_ENV = ("TORCHELASTIC_USE_AGENT_STORE", "TORCHELASTIC_RUN_ID", "TORCHELASTIC_RESTART_COUNT",
                 "TORCHELASTIC_MAX_RESTARTS", "RANK", "LOCAL_RANK", "WORLD_SIZE", "LOCAL_WORLD_SIZE", "GROUP_RANK",
                 "ROLE_RANK", "MASTER_ADDR", "MASTER_PORT", "OMP_NUM_THREADS")


def _worker(rank, world, port, fn, args, out_dir, threads):
    # When the parent runs under torchrun its variables leak into this CPU world; with
    # TORCHELASTIC_USE_AGENT_STORE set, rank 0 would wait for an agent-owned store instead of
    # serving tcp://127.0.0.1:port itself and every rank would hang.
    for k in _LAUNCHER_ENV:
        os.environ.pop(k, None)
    torch.set_num_threads(threads)
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    try:
        res = fn(rank, world, *args)
        if res is not None:
            torch.save(res, os.path.join(out_dir, "rank%d.pt" % rank))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def run_world(world, fn, args=(), threads=1):
    """Run fn(rank, world, *args) on `world` gloo CPU ranks; returns [result of rank 0, 1, ...]."""
    import tempfile
    port = free_port()
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_worker, args=(world, port, fn, args, d, threads), nprocs=world, join=True)
        out = []
        for r in range(world):
            f = os.path.join(d, "rank%d.pt" % r)
            out.append(torch.load(f, weights_only=False) if os.path.exists(f) else None)
    return out


class _P:  # minimal stand-in for a parameter: only `.grad.data` is touched by the path
    def __init__(self, g):
        self.grad = g


def _avg_job(rank, world, shapes, positive, base_seed):
    from .gradient_sets import seeded_grads, seeded_positive_grads
    grads = (seeded_positive_grads if positive else seeded_grads)(shapes, rank, base_seed)
    params = [_P(g) for g in grads]
    average_gradients_ref(params, world)
    return [p.grad.numpy() for p in params]


def gloo_average(world, shapes, positive=False, base_seed=1234):
    """Average the seeded gradient set over `world` gloo ranks; returns rank 0's tensors and
    asserts that every rank ended with identical bytes."""
    res = run_world(world, _avg_job, (shapes, positive, base_seed))
    for r in range(1, world):
        for a, b in zip(res[0], res[r]):
            assert a.tobytes() == b.tobytes(), "gloo ranks disagree"
    return res[0]


def _flat_job(rank, world, numel, iters, warmup, base_seed):
    g = torch.Generator().manual_seed(base_seed + rank)
    buf = torch.randn(numel, generator=g, dtype=torch.float32)
    p = [_P(buf)]
    times = []
    for i in range(warmup + iters):
        dist.barrier()
        t0 = time.perf_counter()
        average_gradients_ref(p, world)
        t1 = time.perf_counter()
        t = torch.tensor([t1 - t0], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if i >= warmup:
            times.append(float(t))
    return times


def time_flat_average(world, numel, iters, warmup=1, threads=1, base_seed=1234):
    """Wall-clock seconds per reference-style averaging of one flat fp32 buffer (max over
    ranks per iteration), BASELINE.md section 4."""
    return run_world(world, _flat_job, (numel, iters, warmup, base_seed), threads=threads)[0]


def _set_job(rank, world, shapes, iters, warmup, base_seed):
    from .gradient_sets import seeded_grads
    params = [_P(g) for g in seeded_grads(shapes, rank, base_seed)]
    times = []
    for i in range(warmup + iters):
        dist.barrier()
        t0 = time.perf_counter()
        average_gradients_ref(params, world)
        t1 = time.perf_counter()
        t = torch.tensor([t1 - t0], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if i >= warmup:
            times.append(float(t))
    return times


def time_set_average(world, shapes, iters, warmup=1, threads=1, base_seed=1234):
    return run_world(world, _set_job, (shapes, iters, warmup, base_seed), threads=threads)[0]


def _flat_multi_job(rank, world, numels, iters, warmup, base_seed):
    out = []
    for numel in numels:
        g = torch.Generator().manual_seed(base_seed + rank)
        buf = torch.randn(numel, generator=g, dtype=torch.float32)
        p = [_P(buf)]
        times = []
        n_it = iters if numel * 4 < (256 << 20) else max(2, iters // 3)
        for i in range(warmup + n_it):
            dist.barrier()
            t0 = time.perf_counter()
            average_gradients_ref(p, world)
            t1 = time.perf_counter()
            t = torch.tensor([t1 - t0], dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            if i >= warmup:
                times.append(float(t))
        out.append(sorted(times)[len(times) // 2])
        del buf, p
    return out


def time_flat_average_sizes(world, numels, iters=5, warmup=1, threads=1, base_seed=1234):
    """Median wall-clock seconds of the reference-style averaging for each flat buffer size, in ONE
    gloo session (the size sweep of BASELINE.json configs[3])."""
    return run_world(world, _flat_multi_job, (list(numels), iters, warmup, base_seed), threads=threads)[0]


class RefNet(torch.nn.Module):
    """The reference model restated (`/root/reference/mnist/main.py:130-147`): conv(1->20,5) -> pool ->
    conv(20->50,5) -> pool -> fc(800->500) -> fc(500->10) -> log_softmax."""

    def __init__(self):
        super().__init__()
        self.conv1 = torch.nn.Conv2d(1, 20, 5, 1)
        self.conv2 = torch.nn.Conv2d(20, 50, 5, 1)
        self.fc1 = torch.nn.Linear(4 * 4 * 50, 500)
        self.fc2 = torch.nn.Linear(500, 10)

    def forward(self, x):
        F = torch.nn.functional
        x = F.max_pool2d(F.relu(self.conv1(x)), 2, 2)
        x = F.max_pool2d(F.relu(self.conv2(x)), 2, 2)
        x = F.relu(self.fc1(x.view(-1, 4 * 4 * 50)))
        return F.log_softmax(self.fc2(x), dim=1)


def _train_job(rank, world, steps, warmup, batch):
    """The reference step on the CPU (`mnist/main.py:84-99`): forward, cross_entropy, zero_grad, backward,
    average_gradients (when world > 1, :94-95), Adam step, loss.item(); synthetic 28x28 batches."""
    F = torch.nn.functional
    torch.manual_seed(0)
    net = RefNet()
    opt = torch.optim.Adam(net.parameters(), lr=1e-3)         # :187, :229
    g = torch.Generator().manual_seed(777 + rank)
    data = torch.randn(batch, 1, 28, 28, generator=g)
    label = torch.randint(0, 10, (batch,), generator=g)
    t0 = None
    for i in range(warmup + steps):
        if i == warmup:
            if world > 1:
                dist.barrier()
            t0 = time.perf_counter()
        output = net(data)                                    # :88
        loss = F.cross_entropy(output, label)                 # :89
        opt.zero_grad()                                       # :91
        loss.backward()                                       # :92
        if world > 1:
            average_gradients_ref(net.parameters(), world)    # :94-95
        opt.step()                                            # :96
        loss.item()                                           # :98
    dt = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([dt], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t)
    return steps / dt


def time_train_loop(world, steps=60, warmup=5, batch=128, threads=1):
    """Synchronous data-parallel steps per second of the reference training step on `world` gloo CPU ranks."""
    return run_world(world, _train_job, (steps, warmup, batch), threads=threads)[0]
