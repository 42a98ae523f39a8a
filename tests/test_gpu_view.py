"""GPU: gradients as views of the peer-mapped region, parameter broadcast (8f-3)
and the stream-order guard, on a hosted world (every rank on the one GPU, the same kernels and
flag protocol as across GPUs).  Every comparison is bit-exact against oracle/fold.c."""
import ctypes

import numpy as np
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu


def _fold(host):
    return oracle.fold_average_c(host)


@pytest.mark.parametrize("world", [1, 2, 3, 4, 5, 8])
def test_view_region_in_place_is_bit_exact(built, world):
    """ga_average_view through the C ABI on a flat region, odd sizes, every op, several chunkings."""
    lib = built
    torch.cuda.set_device(0)
    cap = 6 << 20
    ctx = ctypes.c_void_p()
    assert lib.ga_create_local(world, 0, cap, ctypes.byref(ctx)) == 0, lib.ga_last_error()
    try:
        assert lib.ga_view_create(ctx, cap, 0) == 0, lib.ga_last_error()
        assert lib.ga_view_bytes(ctx) == cap and lib.ga_view_has_multicast(ctx) == 0
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

        def view(l, n, dtype=torch.float32):
            class H:
                pass
            h = H()
            h.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4" if dtype == torch.float32 else "<i4",
                                          "data": (int(lib.ga_view_ptr(ctx, l)), False), "version": 2}
            return torch.as_tensor(h, device="cuda:0")

        rng = np.random.default_rng(7)
        # (elements, chunk tiles [0 = automatic], one-shot limit [-1 = automatic, 0 = two-shot only, big = one-shot only])
        cases = ((1, 0, -1), (511, 0, 0), (512 * 37 + 5, 1, -1), (512 * 37 + 5, 2, 1 << 40), (431080, 0, -1), (431080, 3, 0),
                 (431080, 5, 1 << 40), (1_400_003, 0, -1), (1_400_003, 64, 0), (cap // 4 // max(1, world - 1), 0, 1 << 40), (cap // 4, 0, -1))
        for n, chunk, one in cases:
            assert lib.ga_set_option(ctx, 7, chunk) == 0 and lib.ga_set_option(ctx, 8, one) == 0
            host = [rng.standard_normal(n).astype(np.float32) for _ in range(world)]
            for l in range(world):
                view(l, n).copy_(torch.from_numpy(host[l]))
            assert lib.ga_average_view(ctx, n * 4, 0, st, 0) == 0, lib.ga_last_error()
            torch.cuda.synchronize()
            assert lib.ga_status(ctx) == 0
            want = _fold(host)
            for l in range(world):
                assert view(l, n).cpu().numpy().tobytes() == want.tobytes(), "n=%d chunk=%d one=%d rank %d" % (n, chunk, one, l)
        assert lib.ga_set_option(ctx, 7, 0) == 0 and lib.ga_set_option(ctx, 8, -1) == 0
        # fp32 sum and int32 sum (toy/main.py:19-20) in place
        n = 70001
        host = [rng.standard_normal(n).astype(np.float32) for _ in range(world)]
        for l in range(world):
            view(l, n).copy_(torch.from_numpy(host[l]))
        assert lib.ga_average_view(ctx, n * 4, 2, st, 7) == 0, lib.ga_last_error()
        torch.cuda.synchronize()
        acc = host[0].copy()
        for h in host[1:]:
            acc = acc + h
        assert view(0, n).cpu().numpy().tobytes() == acc.tobytes()
        ih = [rng.integers(-2**31, 2**31 - 1, n, dtype=np.int64).astype(np.int32) for _ in range(world)]
        for l in range(world):
            view(l, n, torch.int32).copy_(torch.from_numpy(ih[l]))
        assert lib.ga_average_view(ctx, n * 4, 1, st, 0) == 0, lib.ga_last_error()
        torch.cuda.synchronize()
        assert view(world - 1, n, torch.int32).cpu().numpy().tobytes() == oracle.fold_sum_i32_c(ih).tobytes()
        # refusals: too large, NVLS without a multicast mapping, bucket variants
        assert lib.ga_average_view(ctx, cap + 4, 0, st, 0) == -3
        assert lib.ga_average_view(ctx, 1024, 0, st, 8) == -4 and b"multicast" in lib.ga_last_error()
        assert lib.ga_average_view(ctx, 1024, 0, st, 2) == -1
        assert lib.ga_view_create(ctx, cap, 0) == -4
    finally:
        lib.ga_destroy(ctx)


def test_view_region_needs_a_context_of_that_capacity(built):
    lib = built
    ctx = ctypes.c_void_p()
    assert lib.ga_create_local(2, 0, 1 << 20, ctypes.byref(ctx)) == 0
    try:
        assert lib.ga_view_create(ctx, 2 << 20, 0) == -3
        assert lib.ga_average_view(ctx, 1024, 0, None, 0) == -4          # no view region yet
    finally:
        lib.ga_destroy(ctx)


def _replicas(world, seed0=0):
    nets = []
    for r in range(world):
        torch.manual_seed(seed0 + r)
        nets.append(torch.nn.Sequential(torch.nn.Conv2d(1, 6, 3), torch.nn.ReLU(), torch.nn.Flatten(),
                                        torch.nn.Linear(6 * 26 * 26, 33), torch.nn.Linear(33, 10)).cuda())
    return nets


@pytest.mark.parametrize("world", [2, 3, 4, 8])
def test_adopted_gradients_train_like_the_drop_in_mode(built, world):
    """adopt(): p.grad are views, averaging is in place.  Same bits as the oracle on the gathered
    gradients, and a 200-step loss curve identical to the drop-in (pack/unpack) mode."""
    import pytorch_distributed_example_b200 as ga
    import torch.nn.functional as F
    torch.cuda.set_device(0)
    torch.backends.cudnn.deterministic = True
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    curves = {}
    for mode in ("view", "dropin"):
        nets = _replicas(world, seed0=0)
        avg = ga.GradAverager(local_world=world, device=0, capacity_bytes=8 << 20)
        avg.broadcast_parameters(nets, root=0)                       # replicas start as rank 0's
        opts = [torch.optim.Adam(n.parameters(), lr=1e-3) for n in nets]
        if mode == "view":
            avg.adopt(nets)
        losses = []
        steps = 200 if world <= 4 else 60
        for step in range(steps):
            g = torch.Generator(device="cuda").manual_seed(777 + step)
            for l, net in enumerate(nets):
                x = torch.randn(16, 1, 28, 28, device="cuda", generator=g)
                y = torch.randint(0, 10, (16,), device="cuda", generator=g)
                loss = F.cross_entropy(net(x), y)
                opts[l].zero_grad(set_to_none=(mode != "view"))
                loss.backward()
                if l == 0:
                    losses.append(loss)
            if step in (3, 4):
                host = [[p.grad.cpu().numpy().copy() for p in n.parameters()] for n in nets]
            # view mode: the in-place kernel on even steps, AUTO (bucket kernels on the region below 32 MiB) on odd ones
            avg.average_gradients_local(nets, variant="view" if (mode == "view" and step % 2 == 0) else None)
            if step in (3, 4):
                for i in range(len(host[0])):
                    want = _fold([host[r][i].ravel() for r in range(world)])
                    for r in range(world):
                        got = list(nets[r].parameters())[i].grad.cpu().numpy().ravel()
                        assert got.tobytes() == want.tobytes(), "%s tensor %d rank %d" % (mode, i, r)
            for o in opts:
                o.step()
        avg.check_status()
        curves[mode] = [float(v) for v in losses]
        if mode == "view":
            # every call went through the in-place kernel: no pack/unpack launch after adopt()
            assert all(p.grad.data_ptr() == a for p, a in zip([p for n in nets for p in n.parameters()], avg._adopted[1]))
            # zero_grad() with the torch default drops the views: the next call silently takes the drop-in path
            host = []
            for l, net in enumerate(nets):
                opts[l].zero_grad()
                F.cross_entropy(net(torch.randn(4, 1, 28, 28, device="cuda")), torch.zeros(4, dtype=torch.long, device="cuda")).backward()
                host.append([p.grad.cpu().numpy().copy() for p in net.parameters()])
            avg.average_gradients_local(nets)
            want = _fold([host[r][0].ravel() for r in range(world)])
            assert list(nets[1].parameters())[0].grad.cpu().numpy().tobytes() == want.tobytes()
        before = [p.grad.clone() for p in nets[0].parameters()]
        avg.close()
        if mode == "view":
            # close() frees the view region: the gradients must have moved out of it, values intact
            torch.zeros(32 << 20, device="cuda")                      # reuse whatever was freed
            for b, p in zip(before, nets[0].parameters()):
                assert torch.equal(b, p.grad)
    assert curves["view"] == curves["dropin"]


def test_view_calls_mixed_with_tensor_calls_unsynchronised(built):
    """The view kernel shares the context's staging area and flag epoch with the bucket kernels:
    300 back-to-back launches alternating kinds and sizes, no host sync, every result checked."""
    import pytorch_distributed_example_b200 as ga
    torch.cuda.set_device(0)
    world = 4
    avg = ga.GradAverager(local_world=world, device=0, capacity_bytes=4 << 20)
    avg.create_view(4 << 20, multicast=False)
    kept = []
    gen = torch.Generator(device="cuda").manual_seed(5)
    kinds = ("view", "oneshot_push", "view", "twoshot_pipe", "twoshot", "view", "twoshot_push", "oneshot")
    for step in range(300):
        n = (70001, 5, 431080, 1 << 20)[step % 4]
        kind = kinds[step % len(kinds)]
        ins = [torch.randn(n, device="cuda", generator=gen) for _ in range(world)]
        if kind == "view":
            outs = [avg.view_tensor(n * 4, l) for l in range(world)]
            for o, i in zip(outs, ins):
                o.copy_(i)
            avg.allreduce_view(n * 4)
            res = [o.clone() for o in outs] if step % 10 == 0 else None
        else:
            work = [i.clone() for i in ins]
            avg.all_reduce_([[w] for w in work], average=True, variant=kind)
            res = work if step % 10 == 0 else None
        if res is not None:
            kept.append((step, kind, ins, res))
    avg.check_status()
    for step, kind, ins, res in kept:
        want = _fold([i.cpu().numpy() for i in ins])
        for l in range(world):
            assert res[l].cpu().numpy().tobytes() == want.tobytes(), "step %d (%s) rank %d" % (step, kind, l)
    avg.close()


@pytest.mark.parametrize("world", [2, 5, 8])
def test_broadcast_makes_replicas_identical_to_the_root(built, world):
    """Replicas built from different seeds (what the reference does, mnist/main.py:185, with no seed at
    all) become bit-identical to rank root's, checked against a host copy of the root."""
    import pytorch_distributed_example_b200 as ga
    torch.cuda.set_device(0)
    for root in (0, world - 1):
        nets = _replicas(world, seed0=100)
        want = [t.detach().cpu().numpy().copy() for t in nets[root].parameters()]
        assert not np.array_equal(want[0], list(nets[(root + 1) % world].parameters())[0].detach().cpu().numpy())
        avg = ga.GradAverager(local_world=world, device=0)
        avg.broadcast_parameters(nets, root=root)
        avg.check_status()
        for r in range(world):
            for w, p in zip(want, nets[r].parameters()):
                assert p.detach().cpu().numpy().tobytes() == w.tobytes()
        # odd sizes and an int64 buffer (moved as 32-bit words), mixed with an averaging call on the same context
        ts = [[torch.full((3,), r, dtype=torch.int64, device="cuda"), torch.randn(70001, device="cuda"),
               torch.randn(1, device="cuda")] for r in range(world)]
        ref = [t.cpu().numpy().copy() for t in ts[root]]
        avg.all_reduce_([[torch.ones(9, device="cuda")] for _ in range(world)], average=True, variant="twoshot_pipe")
        avg.broadcast_(ts, root=root)
        avg.check_status()
        for r in range(world):
            for w, t in zip(ref, ts[r]):
                assert t.cpu().numpy().tobytes() == w.tobytes()
        avg.close()


def test_calls_on_different_streams_are_ordered_by_the_library(built):
    """two launches of one context on different streams with no event between them
    shared the flag epoch and the bucket.  The library now makes the new stream wait."""
    import pytorch_distributed_example_b200 as ga
    torch.cuda.set_device(0)
    world = 4
    avg = ga.GradAverager(local_world=world, device=0, capacity_bytes=64 << 20)
    n = 4 << 20
    streams = [torch.cuda.Stream() for _ in range(3)]
    gen = torch.Generator(device="cuda").manual_seed(11)
    ins = [[torch.randn(n, device="cuda", generator=gen) for _ in range(world)] for _ in range(6)]
    work = [[t.clone() for t in s] for s in ins]
    torch.cuda.synchronize()
    for k, w in enumerate(work):
        with torch.cuda.stream(streams[k % 3]):
            avg.all_reduce_([[t] for t in w], average=True, variant=("twoshot", "oneshot", "twoshot_pipe")[k % 3])
    torch.cuda.synchronize()
    avg.check_status()
    for s, w in zip(ins, work):
        want = _fold([t.cpu().numpy() for t in s])
        for l in range(world):
            assert w[l].cpu().numpy().tobytes() == want.tobytes()
    avg.close()


def test_overlap_refuses_itself_when_it_cannot_pay(built):
    """configs[4]: hook-driven overlap only switches itself on when the exchange it could hide is
    longer than what the hooks cost; otherwise average_gradients() is the single fused call.  Both
    modes give the oracle's bits."""
    import pytorch_distributed_example_b200 as ga
    from pytorch_distributed_example_b200.overlap import overlap_pays
    torch.cuda.set_device(0)
    torch.backends.cudnn.deterministic = True
    world = 2
    assert overlap_pays([11_689_512 // 62] * 62, 8, 3)[0] is False          # ResNet-18-sized set on 8 GPUs: ~170 us exposed
    assert overlap_pays([25_000_000] * 40, 8, 40)[0] is True                # 4 GB of gradients
    for force in (None, True):
        nets = _replicas(world, seed0=3)
        avg = ga.GradAverager(local_world=world, device=0, capacity_bytes=8 << 20)
        ov = ga.OverlappedAverager(nets, averager=avg, bucket_cap_bytes=64 << 10, max_blocks=8, force=force)
        assert ov.enabled is (force is True) and "us" in ov.reason
        for l, net in enumerate(nets):
            net(torch.randn(8, 1, 28, 28, device="cuda", generator=torch.Generator(device="cuda").manual_seed(l))).sum().backward()
        # the hooks have already averaged some buckets in place when overlap is on: rebuild the inputs
        host = []
        for l in range(world):
            ref = _replicas(world, seed0=3)[l]
            ref(torch.randn(8, 1, 28, 28, device="cuda", generator=torch.Generator(device="cuda").manual_seed(l))).sum().backward()
            host.append([p.grad.cpu().numpy().ravel().copy() for p in ref.parameters()])
        ov.average_gradients()
        avg.check_status()
        for i, p in enumerate(nets[1].parameters()):
            assert p.grad.cpu().numpy().tobytes() == _fold([host[r][i] for r in range(world)]).tobytes(), "force=%s tensor %d" % (force, i)
        ov.remove()
        avg.close()
