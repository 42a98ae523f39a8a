"""CPU: host-side logic of the shim -- layout arithmetic, the `distributed` stand-in, the Trainer
patch -- including a world_size-2 gloo run of the multi-process code path up to the point where
a GPU would be needed."""
import os
import sys

import pytest
import torch

import pytorch_distributed_example_b200 as ga
from pytorch_distributed_example_b200 import averager, distributed

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_bucket_layout_arithmetic():
    assert averager.bucket_bytes([1]) == 2048
    assert averager.bucket_bytes([512, 513, 0]) == 2048 + 4096
    import oracle
    mn = averager.bucket_bytes(oracle.numel_of(oracle.MNIST_CNN_SHAPES))
    assert mn % 2048 == 0 and 1724320 <= mn <= 1724320 + 8 * 2048


def test_reduce_op_alias_and_passthrough():
    assert distributed.reduce_op.SUM == torch.distributed.ReduceOp.SUM
    assert distributed.ReduceOp is torch.distributed.ReduceOp
    assert distributed.is_available is torch.distributed.is_available       # everything else: torch's


def test_install_patches_class_and_instance():
    class Trainer:                       # same surface as mnist/main.py:54-63,122
        def __init__(self, net):
            self.net = net

        def average_gradients(self):
            raise AssertionError("reference path must not run")

    calls = []
    orig = ga.trainer.average_gradients
    ga.trainer.average_gradients = lambda params, variant=None: calls.append((list(params), variant))
    try:
        net = torch.nn.Linear(3, 2)
        ga.install(Trainer, variant="twoshot")
        Trainer(net).average_gradients()
        assert len(calls) == 1 and len(calls[0][0]) == 2 and calls[0][1] == "twoshot"
        t = Trainer(net)
        ga.install(t)
        t.average_gradients()
        assert len(calls) == 2
    finally:
        ga.trainer.average_gradients = orig


def test_missing_grad_raises_like_reference():
    net = torch.nn.Linear(2, 2)          # no backward yet: p.grad is None -> `.data` AttributeError in the reference
    with pytest.raises(AttributeError):
        ga.GradAverager._grads_of(net.parameters())


def _w2(rank, world, port, q):
    sys.path.insert(0, ROOT)
    from pytorch_distributed_example_b200 import distributed as dist
    import pytorch_distributed_example_b200 as ga2
    dist.init_process_group(backend="gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    out = {"ws": dist.get_world_size(), "rank": dist.get_rank()}
    tok = dist.new_group(ranks=list(range(world)))           # toy/main.py:15-16 -> token, no new torch group
    out["token"] = type(tok).__name__
    avg = ga2.GradAverager()                                  # binds to the default group, no GPU needed yet
    out["avg_ws"], out["avg_rank"] = avg.world_size, avg.rank
    try:
        dist.all_reduce(torch.IntTensor([rank + 1]), op=dist.reduce_op.SUM, group=tok)
        out["err"] = None
    except ga2.GradAvgError as e:                             # no GPU here: must fail, not fall back to gloo
        out["err"] = str(e)
    try:
        dist.all_reduce(torch.zeros(2, dtype=torch.float64))
    except ga2.GradAvgError as e:
        out["dtype_err"] = str(e)
    dist.barrier()
    dist.destroy_process_group()
    q.put(out)


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour of the multi-process path")
def test_world2_gloo_host_path():
    import torch.multiprocessing as mp
    from oracle.gloo_path import free_port
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=_w2, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    res = sorted([q.get(timeout=120) for _ in procs], key=lambda d: d["rank"])
    [p.join(60) for p in procs]
    for r, d in enumerate(res):
        assert d["ws"] == 2 and d["rank"] == r and d["avg_ws"] == 2 and d["avg_rank"] == r
        assert d["token"] == "WorldToken"
        assert d["err"] is not None and "no CPU fallback" in d["err"]
        assert "float64" in d["dtype_err"]


def test_workload_shapes_agree_with_oracle_tables():
    import oracle
    from pytorch_distributed_example_b200 import workloads
    assert workloads.MNIST_CNN_SHAPES == oracle.MNIST_CNN_SHAPES
    assert workloads.resnet18_shapes() == oracle.resnet18_shapes()


def test_bucket_plan_reverse_order_and_cap():
    from pytorch_distributed_example_b200 import plan_buckets, workloads
    numels = [workloads.numel(s) for s in workloads.resnet18_shapes()]
    buckets = plan_buckets(numels, 25 << 20)
    flat = [i for b in buckets for i in b]
    assert flat == list(reversed(range(len(numels))))                  # last parameter first, nothing lost
    for b in buckets:
        assert averager.bucket_bytes([numels[i] for i in b]) <= (25 << 20) or len(b) == 1
    assert len(buckets) >= 2
    assert plan_buckets([10, 10], 1) == [[1], [0]]                      # a bucket always takes one tensor
    assert plan_buckets([], 1 << 20) == []


def test_plan_cache_survives_id_reuse():
    """The per-parameter-list layout cache is keyed by object identity; it must keep the objects
    alive (ids of dead objects are recycled) and re-check sizes on every call."""
    import ctypes
    import threading
    calls = []

    class FakeLib:
        def ga_average(self, ctx, ptrs, numel, n, stream, variant):
            calls.append([numel[i] for i in range(n)])
            return 0

    class G:
        def __init__(self, n):
            self.n, self.dtype = n, torch.float32

        def numel(self):
            return self.n

        def get_device(self):
            return 0

        def is_contiguous(self):
            return True

        def data_ptr(self):
            return 4096

    class P:
        def __init__(self, n):
            self.grad = G(n)

    avg = ga.GradAverager.__new__(ga.GradAverager)
    avg._lib, avg._ctx, avg._lock = FakeLib(), ctypes.c_void_p(1), threading.Lock()
    avg._variant, avg._capacity, avg._local, avg.world_size, avg.rank = 0, 1 << 30, 0, 2, 0
    avg._device, avg._dev_obj, avg._plans, avg._ptr_cache, avg._group = 0, torch.device("cuda", 0), {}, {}, None
    avg._launch_base, avg._adopted = 0, None
    orig = averager.torch.cuda.current_stream
    averager.torch.cuda.current_stream = lambda d=None: type("S", (), {"cuda_stream": 0})()
    try:
        for step in range(200):                      # fresh objects every step: ids would be recycled
            sizes = [3 + step % 7, 11]
            avg.average_gradients([P(s) for s in sizes])
            assert calls[-1] == sizes
        ps = [P(5)]
        avg.average_gradients(ps)
        ps[0].grad = G(6)                            # same parameter object, gradient changed size
        with pytest.raises(ga.GradAvgError):
            avg.average_gradients(ps)
    finally:
        averager.torch.cuda.current_stream = orig


def test_init_process_group_selects_the_ranks_gpu(monkeypatch):
    """the reference never selects a device, so every rank would sit on GPU 0; the
    stand-in picks LOCAL_RANK (torchrun) or the rank, modulo the visible devices, before the
    rendezvous, and forwards the reference's arguments unchanged."""
    chosen, forwarded = [], {}
    monkeypatch.setattr(torch.cuda, "is_available", lambda: True)
    monkeypatch.setattr(torch.cuda, "device_count", lambda: 4)
    monkeypatch.setattr(torch.cuda, "set_device", lambda d: chosen.append(d))
    monkeypatch.setattr(distributed._td, "init_process_group", lambda **kw: forwarded.update(kw))
    monkeypatch.delenv("LOCAL_RANK", raising=False)
    distributed.init_process_group(backend="nccl", init_method="tcp://127.0.0.1:23456", rank=6, world_size=8)
    assert chosen == [2]                                     # 6 % 4
    assert forwarded == {"backend": "nccl", "init_method": "tcp://127.0.0.1:23456", "rank": 6, "world_size": 8}
    monkeypatch.setenv("LOCAL_RANK", "3")
    distributed.init_process_group(backend="gloo", rank=11, world_size=16)
    assert chosen[-1] == 3


def test_overlap_decision_is_a_pure_function_of_sizes():
    """configs[4]: hook-driven overlap is only worth its host cost when the exchange it could hide is longer;
    the model (fitted to the round-2 bench) declines ResNet-18-sized sets on NVLink and accepts GB-sized ones."""
    from pytorch_distributed_example_b200 import workloads
    from pytorch_distributed_example_b200.overlap import overlap_pays, plan_buckets
    numels = [workloads.numel(s) for s in workloads.resnet18_shapes()]
    buckets = plan_buckets(numels, 16 << 20)
    assert sorted(i for b in buckets for i in b) == list(range(62)) and buckets[0][0] == 61     # reverse order, every tensor once
    for world in (2, 4, 8):
        pays, exchange_us, overhead_us = overlap_pays(numels, world, len(buckets))
        assert not pays and 50 < exchange_us < 200 and overhead_us > 2 * exchange_us
    pays, exchange_us, overhead_us = overlap_pays([50_000_000] * 30, 8, 30)                     # 6 GB of gradients
    assert pays and exchange_us > 5 * overhead_us
    assert overlap_pays(numels, 1, 4)[0] is False                                               # one rank: nothing to exchange


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_new_entry_points_fail_loudly_without_gpu():
    """The round-2 entry points (views, broadcast, the no-allocation variant and its status call) have no
    CPU fallback either: without a device they return an error code and a message, and compute nothing."""
    import ctypes
    import pytorch_distributed_example_b200 as ga
    lib = ga.load_library()
    assert lib.ga_multicast_supported(0) == 0
    assert lib.ga_view_create(None, 1 << 20, 0) == -1 and lib.ga_average_view(None, 1024, 0, None, 0) == -1
    assert lib.ga_view_bytes(None) == 0 and lib.ga_view_has_multicast(None) == 0 and lib.ga_view_destroy(None) == -1
    assert lib.ga_broadcast_tensors(None, None, None, 0, 0, None) == -1
    numel = (ctypes.c_int64 * 1)(16)
    ptrs = (ctypes.c_void_p * 2)(4096, 8192)
    assert lib.ga_average_static(2, 0, ptrs, numel, 1, None, 0) == -2 and b"cuda" in lib.ga_last_error().lower()
    assert lib.ga_static_status(0, None) == -2
    assert lib.ga_average_static(2, 0, ptrs, (ctypes.c_int64 * 1)(3_000_000), 1, None, 0) == -3     # size check comes before any CUDA call
    avg = ga.GradAverager(local_world=2)
    with pytest.raises(ga.GradAvgError):
        avg.adopt([torch.nn.Linear(3, 3), torch.nn.Linear(3, 3)])
    with pytest.raises(ga.GradAvgError):
        avg.broadcast_([[torch.zeros(4)], [torch.zeros(4)]])
