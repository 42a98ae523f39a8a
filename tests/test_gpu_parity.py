"""Parity of the CUDA path against the oracle, through the C ABI (ctypes) -- GPU box only.

All ranks of a `world`-rank job are hosted on the one GPU (`ga_create_local`): same kernels,
same flag protocol, one cooperative launch.  Gate: BIT-EXACT against the rank-order fold
(oracle/fold.c, numpy) -- the tolerance is zero and is written as `tobytes() ==`.
"""
import ctypes
import hashlib
import json
import os

import numpy as np
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu

ONESHOT, TWOSHOT, ONESHOT_PUSH, TWOSHOT_PUSH, TWOSHOT_PIPE, HOSTED_DIRECT = 1, 2, 3, 4, 5, 6
ALL_VARIANTS = (ONESHOT, TWOSHOT, ONESHOT_PUSH, TWOSHOT_PUSH, TWOSHOT_PIPE, HOSTED_DIRECT)
F32_AVG, I32_SUM, F32_SUM = 0, 1, 2


def _check(lib, rc):
    assert rc == 0, lib.ga_last_error().decode()


class LocalWorld:
    """Thin ctypes driver of the C ABI for a world hosted on cuda:0."""

    def __init__(self, lib, world, capacity):
        self.lib, self.world = lib, world
        self.ctx = ctypes.c_void_p()
        _check(lib, lib.ga_create_local(world, 0, capacity, ctypes.byref(self.ctx)))
        _check(lib, lib.ga_set_option(self.ctx, 2, 5000))      # GA_OPT_TIMEOUT_MS: fail fast, never hang the box

    def tensors(self, per_rank, op=F32_AVG, variant=ONESHOT):
        n = len(per_rank[0])
        ptrs = (ctypes.c_void_p * (n * self.world))(*[t.data_ptr() for ts in per_rank for t in ts])
        numel = (ctypes.c_int64 * n)(*[t.numel() for t in per_rank[0]])
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        if op == F32_AVG:
            rc = self.lib.ga_average(self.ctx, ptrs, numel, n, st, variant)
        else:
            rc = self.lib.ga_allreduce_tensors(self.ctx, ptrs, numel, n, op, st, variant)
        _check(self.lib, rc)
        torch.cuda.synchronize()
        _check(self.lib, self.lib.ga_status(self.ctx))

    def close(self):
        self.lib.ga_destroy(self.ctx)


@pytest.fixture(scope="module")
def lib(built):
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    return built


def _dev_view(ptr, n, typestr="<f4"):
    class _H:
        pass
    h = _H()
    h.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (int(ptr), False), "version": 2}
    return torch.as_tensor(h, device="cuda:0")


def _host_sets(shapes, world, gen=oracle.seeded_grads):
    return [[g.numpy() for g in gen(shapes, r)] for r in range(world)]


def _fold(host, world):
    return [oracle.fold_average_c([host[r][i] for r in range(world)]) for i in range(len(host[0]))]


def _to_dev(host):
    return [[torch.from_numpy(a).cuda() for a in rank] for rank in host]


def _assert_all_ranks_equal(dev, want, what):
    for r, rank in enumerate(dev):
        for i, (got, w) in enumerate(zip(rank, want)):
            assert got.cpu().numpy().tobytes() == np.ascontiguousarray(w).tobytes(), \
                "%s: rank %d tensor %d differs from the rank-order fold" % (what, r, i)


@pytest.mark.parametrize("world", [1, 2, 3, 4, 5, 8])
@pytest.mark.parametrize("variant", ALL_VARIANTS)
@pytest.mark.parametrize("set_name", ["mini", "mnist"])
def test_bit_exact_vs_fold(lib, world, variant, set_name):
    shapes = oracle.MINI_SHAPES if set_name == "mini" else oracle.MNIST_CNN_SHAPES
    host = _host_sets(shapes, world)
    want = _fold(host, world)
    w = LocalWorld(lib, world, 4 << 20)
    try:
        dev = _to_dev(host)
        w.tensors(dev, variant=variant)
        _assert_all_ranks_equal(dev, want, "%s ws=%d v=%d" % (set_name, world, variant))
    finally:
        w.close()


@pytest.mark.parametrize("world", [2, 3, 4, 8])
def test_golden_mnist_sha(lib, golden_dir, world):
    """Full 431 080-element set: sha256 of our output == sha256 of the fold recorded next to the
    REAL reference's gloo output (tests/golden/mnist_cnn.json); at ws=2 that is also gloo's sha."""
    meta = json.load(open(os.path.join(golden_dir, "mnist_cnn.json")))["worlds"][str(world)]
    for name, gen in (("signed", oracle.seeded_grads), ("positive", oracle.seeded_positive_grads)):
        host = _host_sets(oracle.MNIST_CNN_SHAPES, world, gen)
        w = LocalWorld(lib, world, 4 << 20)
        try:
            for variant in ALL_VARIANTS:
                dev = _to_dev(host)
                w.tensors(dev, variant=variant)
                flat = np.concatenate([t.cpu().numpy().ravel() for t in dev[0]])
                sha = hashlib.sha256(flat.tobytes()).hexdigest()
                assert sha == meta[name]["fold_sha256"]
                if world == 2:
                    assert sha == meta[name]["sha256"]          # bit-identical to the reference on gloo
        finally:
            w.close()


@pytest.mark.parametrize("world", [2, 3, 4, 8])
def test_golden_mini_vs_reference_gloo(lib, golden_dir, world):
    """Against the reference's own output (real `Trainer.average_gradients` on gloo, stored
    arrays): bit-exact at ws=2; at ws>=3 gloo's ring order differs in the last bits,
    so the gate is max|d| / mean_r|g_r| <= 1e-6 (tolerance from north_star, metric from F3)."""
    z = np.load(os.path.join(golden_dir, "mini_ws%d.npz" % world))
    n = len(oracle.MINI_SHAPES)
    host = [[z["in_r%d_t%d" % (r, i)] for i in range(n)] for r in range(world)]
    ref = [z["out_t%d" % i] for i in range(n)]
    w = LocalWorld(lib, world, 1 << 20)
    try:
        for variant in ALL_VARIANTS:
            dev = _to_dev(host)
            w.tensors(dev, variant=variant)
            for i in range(n):
                got = dev[world - 1][i].cpu().numpy()
                if world == 2:
                    assert got.tobytes() == ref[i].tobytes()
                else:
                    err = oracle.backward_error(got, ref[i], [host[r][i] for r in range(world)])
                    assert err <= 1e-6, err
    finally:
        w.close()


def test_edge_cases(lib):
    """Empty list, zero-element tensors, 1 element, sizes around the 16-byte vector and the
    2 KiB tile, a misaligned (4-byte aligned only) view, and a tensor count above the inline
    pointer-table size (device-side table path)."""
    world = 3
    w = LocalWorld(lib, world, 8 << 20)
    try:
        # empty list: no launch, no error
        st = ctypes.c_void_p(0)
        _check(lib, lib.ga_average(w.ctx, None, None, 0, st, ONESHOT))
        sizes = [0, 1, 2, 3, 4, 5, 511, 512, 513, 1023, 1025, 0, 7, 4096 + 3]
        for variant in ALL_VARIANTS:
            host = [[np.random.default_rng(100 * r + i).standard_normal(s).astype(np.float32)
                     for i, s in enumerate(sizes)] for r in range(world)]
            want = _fold(host, world)
            dev = _to_dev(host)
            w.tensors(dev, variant=variant)
            _assert_all_ranks_equal(dev, want, "ragged v=%d" % variant)
            # misaligned views: element offset 1 inside a larger buffer; neighbours must survive
            big = [torch.full((3000,), 7.0, device="cuda") for _ in range(world)]
            hv = [np.random.default_rng(7 + r).standard_normal(2001).astype(np.float32) for r in range(world)]
            views = []
            for r in range(world):
                v = big[r][1:2002]
                v.copy_(torch.from_numpy(hv[r]))
                views.append([v])
            w.tensors(views, variant=variant)
            wv = oracle.fold_average(hv)
            for r in range(world):
                assert views[r][0].cpu().numpy().tobytes() == wv.tobytes()
                assert float(big[r][0]) == 7.0 and bool((big[r][2002:] == 7.0).all())
            # 200 tensors per rank -> 600 table entries > 128 inline
            many = [[np.random.default_rng(1000 * r + i).standard_normal(17 + i).astype(np.float32)
                     for i in range(200)] for r in range(world)]
            dev = _to_dev(many)
            w.tensors(dev, variant=variant)
            _assert_all_ranks_equal(dev, _fold(many, world), "many v=%d" % variant)
    finally:
        w.close()


@pytest.mark.parametrize("world", [3, 4, 8])
def test_exact_capacity(lib, world):
    """Bucket capacity == bytes of the set (what the Python layer allocates on first use): the
    two-shot staging layout needs world * ceil(T / world) >= T tiles and must not spill."""
    shapes = oracle.MNIST_CNN_SHAPES
    numel = (ctypes.c_int64 * len(shapes))(*oracle.numel_of(shapes))
    need = lib.ga_bucket_offset(numel, len(shapes), len(shapes))
    host = _host_sets(shapes, world)
    want = _fold(host, world)
    w = LocalWorld(lib, world, need)
    try:
        for rep in range(3):
            for variant in ALL_VARIANTS:
                dev = _to_dev(host)
                w.tensors(dev, variant=variant)
                _assert_all_ranks_equal(dev, want, "exact capacity ws=%d v=%d" % (world, variant))
    finally:
        w.close()


def test_special_values(lib):
    """inf / nan / denormals / signed zero go through the same IEEE adds and divide as the fold."""
    world = 3
    vals = np.array([np.inf, -np.inf, np.nan, 0.0, -0.0, 1e-45, -1e-45, 1.1754942e-38, 3.4e38, -3.4e38, 1.0, 1 / 3],
                    dtype=np.float32)
    rng = np.random.default_rng(5)
    host = [[rng.choice(vals, size=4099).astype(np.float32)] for _ in range(world)]
    want = _fold(host, world)[0]
    w = LocalWorld(lib, world, 1 << 20)
    try:
        for variant in ALL_VARIANTS:
            dev = _to_dev(host)
            w.tensors(dev, variant=variant)
            got = dev[1][0].cpu().numpy()
            nan = np.isnan(want)
            assert (np.isnan(got) == nan).all()
            assert got[~nan].tobytes() == want[~nan].tobytes()
    finally:
        w.close()


def test_int32_sum_toy(lib, golden_dir):
    """toy/main.py:19-20: one int32 per rank, SUM.  Values and expected sums are the ones the real
    reference printed (tests/golden/toy.json)."""
    import re
    toy = json.load(open(os.path.join(golden_dir, "toy.json")))["worlds"]
    for ws_s, outs in toy.items():
        world = int(ws_s)
        rows = [re.findall(r"value: (\d+), reduced sum: (\d+\.\d+)", o) for o in outs]
        w = LocalWorld(lib, world, 1 << 20)
        try:
            for step in range(len(rows[0])):
                dev = [[torch.tensor([int(rows[r][step][0])], dtype=torch.int32, device="cuda")] for r in range(world)]
                w.tensors(dev, op=I32_SUM, variant=ONESHOT)
                for r in range(world):
                    assert float(dev[r][0]) == float(rows[r][step][1])
        finally:
            w.close()
    # wrap-around and a larger vector, both variants
    world = 4
    host = [[np.random.default_rng(r).integers(-2**31, 2**31 - 1, size=70001, dtype=np.int64).astype(np.int32)]
            for r in range(world)]
    want = oracle.fold_sum_i32_c([host[r][0] for r in range(world)])
    w = LocalWorld(lib, world, 1 << 20)
    try:
        for variant in ALL_VARIANTS:
            dev = _to_dev(host)
            w.tensors(dev, op=I32_SUM, variant=variant)
            for r in range(world):
                assert dev[r][0].cpu().numpy().tobytes() == want.tobytes()
    finally:
        w.close()


def test_f32_sum_then_divide_matches(lib):
    """all_reduce alone (GA_OP_F32_SUM, mnist/main.py:126) followed by a host-side true division
    equals the fused average."""
    world = 3
    host = _host_sets(oracle.MINI_SHAPES, world)
    want = _fold(host, world)
    w = LocalWorld(lib, world, 1 << 20)
    try:
        dev = _to_dev(host)
        w.tensors(dev, op=F32_SUM, variant=TWOSHOT)
        for i, t in enumerate(dev[0]):
            assert (t.cpu().numpy() / np.float32(3.0)).tobytes() == want[i].tobytes()
    finally:
        w.close()


def test_capacity_and_argument_errors(lib):
    w = LocalWorld(lib, 2, 64 << 10)
    try:
        big = [[torch.zeros(100000, device="cuda")] for _ in range(2)]
        ptrs = (ctypes.c_void_p * 2)(*[t[0].data_ptr() for t in big])
        numel = (ctypes.c_int64 * 1)(100000)
        rc = lib.ga_average(w.ctx, ptrs, numel, 1, None, ONESHOT)
        assert rc == -3 and b"capacity" in lib.ga_last_error()
        numel[0] = 100
        assert lib.ga_average(w.ctx, ptrs, numel, 1, None, 7) == -1
        numel[0] = -1
        assert lib.ga_average(w.ctx, ptrs, numel, 1, None, ONESHOT) == -1
        assert lib.ga_allreduce_bucket(w.ctx, 1 << 30, 0, None, ONESHOT) == -3
    finally:
        w.close()
    ctx = ctypes.c_void_p()
    assert lib.ga_create_local(9, 0, 1 << 20, ctypes.byref(ctx)) == -1
    assert lib.ga_create(3, 2, 0, 1 << 20, ctypes.byref(ctx)) == -1
    assert lib.ga_create_local(2, 99, 1 << 20, ctypes.byref(ctx)) == -1


def test_repeated_steps_changing_pointers(lib):
    """SURVEY F6 + flag-reuse hazards: 300 back-to-back calls without host synchronisation,
    fresh gradient allocations every step, sizes and variants alternating."""
    world = 4
    w = LocalWorld(lib, world, 8 << 20)
    try:
        sets = [oracle.MINI_SHAPES, oracle.MNIST_CNN_SHAPES, [(70001,)], [(3,), (5,)]]
        expected, results = [], []
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        for step in range(300):
            shapes = sets[step % len(sets)]
            variant = ALL_VARIANTS[(step // 2) % 6]
            dev = [[torch.randn(s, device="cuda", generator=None) for s in shapes] for _ in range(world)]
            if step % 50 == 0:
                expected.append((step, [[t.cpu().numpy().copy() for t in rank] for rank in dev]))
            n = len(shapes)
            ptrs = (ctypes.c_void_p * (n * world))(*[t.data_ptr() for ts in dev for t in ts])
            numel = (ctypes.c_int64 * n)(*[t.numel() for t in dev[0]])
            _check(lib, lib.ga_average(w.ctx, ptrs, numel, n, st, variant))
            if step % 50 == 0:
                results.append(dev)
        torch.cuda.synchronize()
        _check(lib, lib.ga_status(w.ctx))
        for (step, host), dev in zip(expected, results):
            _assert_all_ranks_equal(dev, _fold(host, world), "step %d" % step)
    finally:
        w.close()


def test_bucket_direct_and_host_forms(lib):
    """ga_allreduce_bucket (send -> recv) and ga_average_host (pinned host tensors) against the fold,
    including the tensor-call -> bucket-call transition (ga_bucket_acquire)."""
    world = 4
    n = 300007
    host = [np.random.default_rng(r).standard_normal(n).astype(np.float32) for r in range(world)]
    want = oracle.fold_average_c(host)
    w = LocalWorld(lib, world, 4 << 20)
    try:
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        for variant in (ONESHOT, TWOSHOT):
            # a tensor-list call first, so that the transition path runs
            dev = _to_dev([[h] for h in host])
            w.tensors(dev, variant=variant)
            _check(lib, lib.ga_bucket_acquire(w.ctx, st))
            for r in range(world):
                _dev_view(lib.ga_send_ptr(w.ctx, r), n).copy_(torch.from_numpy(host[r]).cuda())
            _check(lib, lib.ga_allreduce_bucket(w.ctx, n * 4, F32_AVG, st, variant))
            torch.cuda.synchronize()
            _check(lib, lib.ga_status(w.ctx))
            for r in range(world):
                out = _dev_view(lib.ga_recv_ptr(w.ctx, r), n)
                assert out.cpu().numpy().tobytes() == want.tobytes()
            # host form
            pinned = [[torch.from_numpy(h.copy()).pin_memory()] for h in host]
            ptrs = (ctypes.c_void_p * world)(*[p[0].data_ptr() for p in pinned])
            numel = (ctypes.c_int64 * 1)(n)
            _check(lib, lib.ga_average_host(w.ctx, ptrs, numel, 1, st, variant))
            torch.cuda.synchronize()
            for r in range(world):
                assert pinned[r][0].numpy().tobytes() == want.tobytes()
    finally:
        w.close()


def test_full_size_properties(lib):
    """64 MiB per rank x 4 ranks (BASELINE's bucket size; the oracle still finishes in seconds in
    C): bit-exact vs fold.c, every rank identical, and averaging W copies of one gradient returns
    it unchanged (W a power of two => exact)."""
    world = 4
    n = (64 << 20) // 4
    w = LocalWorld(lib, world, 64 << 20)
    try:
        g = torch.Generator(device="cuda").manual_seed(99)
        base = [torch.randn(n, device="cuda", generator=g) for _ in range(world)]
        host = [b.cpu().numpy() for b in base]
        want = oracle.fold_average_c(host)
        for variant in ALL_VARIANTS:
            dev = [[b.clone()] for b in base]
            w.tensors(dev, variant=variant)
            got0 = dev[0][0].cpu().numpy()
            assert got0.tobytes() == want.tobytes()
            for r in range(1, world):
                assert torch.equal(dev[r][0], dev[0][0])
            same = [[base[0].clone()] for _ in range(world)]
            w.tensors(same, variant=variant)
            assert torch.equal(same[2][0], base[0])
    finally:
        w.close()


def test_resnet18_set(lib):
    """62 tensors, 46.76 MB: both variants, 8 ranks hosted."""
    world = 8
    shapes = oracle.resnet18_shapes()
    g = torch.Generator(device="cuda").manual_seed(3)
    base = [[torch.randn(s, device="cuda", generator=g) for s in shapes] for _ in range(world)]
    host = [[t.cpu().numpy() for t in rank] for rank in base]
    want = _fold(host, world)
    w = LocalWorld(lib, world, 48 << 20)
    try:
        for variant in ALL_VARIANTS:
            dev = [[t.clone() for t in rank] for rank in base]
            w.tensors(dev, variant=variant)
            _assert_all_ranks_equal(dev, want, "resnet18 v=%d" % variant)
    finally:
        w.close()


def test_python_layer_layouts_and_errors(built):
    """GradAverager: channels_last gradients are accepted (dense, elementwise op), strided views
    and wrong dtypes / devices are rejected loudly, a parameter without .grad raises AttributeError."""
    import pytorch_distributed_example_b200 as ga
    torch.cuda.set_device(0)
    world = 2
    avg = ga.GradAverager(local_world=world, device=0)
    try:
        ps = []
        for r in range(world):
            p = torch.nn.Parameter(torch.zeros(8, 4, 3, 3, device="cuda").to(memory_format=torch.channels_last))
            p.grad = torch.randn(8, 4, 3, 3, device="cuda", generator=torch.Generator(device="cuda").manual_seed(r)) \
                .to(memory_format=torch.channels_last)
            ps.append([p])
        want = (ps[0][0].grad + ps[1][0].grad) / 2.0
        avg.average_gradients_local(ps)
        assert torch.equal(ps[0][0].grad, want) and torch.equal(ps[1][0].grad, want)
        assert ps[0][0].grad.is_contiguous(memory_format=torch.channels_last)
        bad = torch.nn.Parameter(torch.zeros(16, device="cuda"))
        bad.grad = torch.zeros(32, device="cuda")[::2]
        with pytest.raises(ga.GradAvgError):
            avg.average_gradients_local([[bad], [bad]])
        half = torch.nn.Parameter(torch.zeros(4, device="cuda"))
        half.grad = torch.zeros(4, device="cuda")
        half.grad.data = half.grad.data  # no-op; dtype check below
        dbl = torch.nn.Parameter(torch.zeros(4, device="cuda", dtype=torch.float64))
        dbl.grad = torch.zeros(4, device="cuda", dtype=torch.float64)
        with pytest.raises(ga.GradAvgError):
            avg.average_gradients_local([[dbl], [dbl]])
        nog = torch.nn.Parameter(torch.zeros(4, device="cuda"))
        with pytest.raises(AttributeError):
            avg.average_gradients_local([[nog], [nog]])
    finally:
        avg.close()


@pytest.mark.parametrize("variant", ["auto", "twoshot", "twoshot_push", "twoshot_pipe", "oneshot_push"])
def test_cuda_graph_replay(built, variant):
    """The flag epoch lives in device memory, so a captured launch can be replayed: capture one
    averaging of static gradient tensors in a CUDA graph, replay it on fresh contents five times,
    interleaved with an eager call, and compare every result with the rank-order fold."""
    import pytorch_distributed_example_b200 as ga
    torch.cuda.set_device(0)
    world = 3
    shapes = oracle.MINI_SHAPES
    avg = ga.GradAverager(local_world=world, device=0, capacity_bytes=1 << 20)

    class P:
        def __init__(self, g):
            self.grad = g
    static = [[P(torch.zeros(s, device="cuda")) for s in shapes] for _ in range(world)]
    try:
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(2):                                   # warm-up outside capture (allocations, occupancy queries)
                avg.average_gradients_local(static, variant=variant)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            avg.average_gradients_local(static, variant=variant)
        for rep in range(5):
            host = [[t.numpy() for t in oracle.seeded_grads(shapes, r, 500 + 10 * rep)] for r in range(world)]
            for r in range(world):
                for p, h in zip(static[r], host[r]):
                    p.grad.copy_(torch.from_numpy(h))
            g.replay()
            torch.cuda.synchronize()
            want = _fold(host, world)
            for r in range(world):
                for p, w in zip(static[r], want):
                    assert p.grad.cpu().numpy().tobytes() == w.tobytes(), "replay %d" % rep
            if rep == 2:                                          # an eager call between replays keeps the epochs in step
                dev = _to_dev(host)
                avg.all_reduce_(dev, average=True, variant=variant)
                _assert_all_ranks_equal(dev, want, "eager between replays")
        avg.check_status()
    finally:
        avg.close()
