"""Worker bodies for the multi-process tests (spawn needs importable top-level functions)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def ipc_world(rank, world, port, q, backend):
    """One rank per GPU over real CUDA-IPC peer memory: parity vs the rank-order fold."""
    import numpy as np
    import torch
    import oracle
    import pytorch_distributed_example_b200 as ga
    from pytorch_distributed_example_b200 import distributed as dist, _lib

    out = {"rank": rank, "fail": []}
    import time as _time
    _t0 = _time.time()

    def note(msg):
        print("[ipc_world rank %d %5.1fs] %s" % (rank, _time.time() - _t0, msg), file=sys.stderr, flush=True)
    try:
        os.environ.setdefault("GA_TIMEOUT_MS", "8000")
        dist.init_process_group(backend=backend, init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
        assert torch.cuda.current_device() == rank          # F5 fix applied by the shim
        avg = ga.default_averager()

        def fold(shapes, gen=oracle.seeded_grads, seed=1234):
            host = [[g.numpy() for g in gen(shapes, r, seed)] for r in range(world)]
            return host, [oracle.fold_average_c([host[r][i] for r in range(world)]) for i in range(len(shapes))]

        class P:
            def __init__(self, g):
                self.grad = g

        note("group up, averager ready")
        # 1. the reference call on the MNIST set and the mini set, both variants
        for name, shapes in (("mini", oracle.MINI_SHAPES), ("mnist", oracle.MNIST_CNN_SHAPES)):
            host, want = fold(shapes)
            for variant in ("oneshot", "twoshot", "oneshot_push", "twoshot_push", "twoshot_pipe"):
                params = [P(torch.from_numpy(a).cuda()) for a in host[rank]]
                ga.average_gradients(params, variant=variant)
                avg.check_status()
                for i, (p, w) in enumerate(zip(params, want)):
                    if p.grad.cpu().numpy().tobytes() != w.tobytes():
                        out["fail"].append("%s %s tensor %d" % (name, variant, i))

        note("part 1 done")
        # 2. skew stress: random extra GPU work before the call on a random rank, fresh pointers,
        #    alternating sizes and variants, no host sync between steps
        rng = np.random.default_rng(42)                       # same stream on every rank
        sets = [oracle.MINI_SHAPES, [(70001,)], oracle.MNIST_CNN_SHAPES, [(5,)]]
        keep = []
        for step in range(200):
            shapes = sets[step % len(sets)]
            variant = ("oneshot", "twoshot_push", "twoshot", "oneshot_push", "twoshot_pipe")[(step // 3) % 5]
            slow = int(rng.integers(0, world))
            cycles = int(rng.integers(0, 3_000_000))
            if rank == slow:
                torch.cuda._sleep(cycles)
            g = torch.Generator(device="cuda").manual_seed(10_000 * step + rank)
            params = [P(torch.randn(s, device="cuda", generator=g)) for s in shapes]
            ins = [p.grad.clone() for p in params] if step % 20 == 0 else None
            ga.average_gradients(params, variant=variant)
            if ins is not None:
                keep.append((step, shapes, ins, params))
        avg.check_status()
        import torch.distributed as tdist
        for step, shapes, ins, params in keep:
            for i, (x, p) in enumerate(zip(ins, params)):
                gathered = [torch.empty_like(x) for _ in range(world)]
                if backend == "gloo":
                    xs = x.cpu()
                    gathered = [torch.empty_like(xs) for _ in range(world)]
                    tdist.all_gather(gathered, xs)
                else:
                    tdist.all_gather(gathered, x)
                want = oracle.fold_average_c([t.cpu().numpy() for t in gathered])
                if p.grad.cpu().numpy().tobytes() != want.tobytes():
                    out["fail"].append("skew step %d tensor %d" % (step, i))

        note("part 2 done")
        # 3. toy path: int32 SUM of a host IntTensor through distributed.all_reduce (toy/main.py:15-20)
        group = dist.new_group(ranks=list(range(world)))
        t = torch.IntTensor([rank + 3])
        dist.all_reduce(t, op=dist.reduce_op.SUM, group=group)
        if int(t) != sum(r + 3 for r in range(world)):
            out["fail"].append("toy int32 sum %d" % int(t))

        note("part 3 done")
        # 4. bucket-direct and host forms
        n = 262144 + 3
        host = [np.random.default_rng(r).standard_normal(n).astype(np.float32) for r in range(world)]
        want = oracle.fold_average_c(host)
        avg.bucket_acquire()
        for variant in ("oneshot", "twoshot"):
            avg.send_view(n * 4).copy_(torch.from_numpy(host[rank]).cuda())
            avg.allreduce_bucket(n * 4, variant=variant)
            avg.check_status()
            if avg.recv_view(n * 4).cpu().numpy().tobytes() != want.tobytes():
                out["fail"].append("bucket-direct " + variant)
            pinned = torch.from_numpy(host[rank].copy()).pin_memory()
            avg.average_host([pinned], variant=variant)
            torch.cuda.synchronize()
            if pinned.numpy().tobytes() != want.tobytes():
                out["fail"].append("host form " + variant)
        note("part 4 done")
        # 5. overlapped, bucketed averaging from grad hooks == the single call after backward
        from pytorch_distributed_example_b200 import workloads
        shapes5 = workloads.resnet18_shapes()[-20:]

        class Bag(torch.nn.Module):
            def __init__(self):
                super().__init__()
                self.ps = torch.nn.ParameterList([torch.nn.Parameter(torch.zeros(s)) for s in shapes5])

            def forward(self, x):
                acc = x
                for p in self.ps:
                    acc = torch.tanh(acc + p.reshape(-1)[:4].sum() * x)
                return acc.sum()

        nets = []
        for mode in ("overlap", "after"):
            torch.manual_seed(1)
            net = Bag().cuda()
            with torch.no_grad():
                for p in net.parameters():
                    p.normal_()
            x = torch.randn(64, device="cuda", generator=torch.Generator(device="cuda").manual_seed(9 + rank))
            ov = ga.OverlappedAverager(net, averager=avg, bucket_cap_bytes=4 << 20, max_blocks=16, force=True) if mode == "overlap" else None
            note("part 5 %s: averager built" % mode)
            for step in range(3):
                net.zero_grad()
                net(x * (step + 1)).backward()
                if ov is not None:
                    ov.average_gradients()
                else:
                    avg.average_gradients(net.parameters())
            if ov is not None:
                ov.remove()
            note("part 5 %s: steps enqueued" % mode)
            avg.check_status()
            note("part 5 %s: done" % mode)
            nets.append(net)
        for pa, pb in zip(nets[0].parameters(), nets[1].parameters()):
            if pa.grad.cpu().numpy().tobytes() != pb.grad.cpu().numpy().tobytes():
                out["fail"].append("overlap != after-backward")
                break
        note("part 5 done")
        # 6. gradients as views of the peer-mapped region: in place, bit-exact; with NVSwitch multicast
        #    where the GPUs have it (cuMem allocation + descriptors fetched with pidfd_getfd), else CUDA-IPC
        nview = 3 * 431080 + 7
        avg.create_view(8 << 20, multicast=True)
        out["multicast"] = avg.has_multicast
        rng6 = np.random.default_rng(6)
        allh = [rng6.standard_normal(nview).astype(np.float32) for _ in range(world)]
        want6 = oracle.fold_average_c(allh)
        for rep, chunk in enumerate((0, 1, 5, 0)):
            avg.set_option(_lib.OPT_VIEW_CHUNK_TILES, chunk)
            if rank == rep % world:
                torch.cuda._sleep(2_000_000)                    # launch skew
            v = avg.view_tensor(nview * 4)
            v.copy_(torch.from_numpy(allh[rank]))
            avg.allreduce_view(nview * 4, variant="view")
            avg.check_status()
            if v.cpu().numpy().tobytes() != want6.tobytes():
                out["fail"].append("view in place, chunk %d" % chunk)
        # view calls alternating with tensor-list calls (shared staging and epoch), no host sync in between
        keep6 = []
        for step in range(40):
            n6 = (70001, 5, 431080, 1 << 20)[step % 4]
            g = torch.Generator(device="cuda").manual_seed(20_000 * step + rank)
            x = torch.randn(n6, device="cuda", generator=g)
            if step % 2 == 0:
                v = avg.view_tensor(n6 * 4)
                v.copy_(x)
                avg.allreduce_view(n6 * 4)
                res = v.clone()
            else:
                res = x.clone()
                avg.all_reduce_([[res]], average=True, variant=("twoshot_pipe", "oneshot_push", "twoshot_push")[step % 3])
            keep6.append((step, x, res))
        avg.check_status()
        for step, x, res in keep6:
            gathered = [torch.empty_like(x) for _ in range(world)]
            if backend == "gloo":
                gathered = [torch.empty_like(x.cpu()) for _ in range(world)]
                tdist.all_gather(gathered, x.cpu())
            else:
                tdist.all_gather(gathered, x)
            if res.cpu().numpy().tobytes() != oracle.fold_average_c([t.cpu().numpy() for t in gathered]).tobytes():
                out["fail"].append("view/tensor mix step %d" % step)
        note("part 6 done (multicast=%s)" % out["multicast"])
        # 7. opt-in in-switch variant: the switch's wide sum rounded once, then the division -- checked
        #    against its own oracle (fold.c: oracle_fold_wide_average_f32), never against the rank-order fold
        if avg.has_multicast:
            v = avg.view_tensor(nview * 4)
            for rep in range(3):
                v.copy_(torch.from_numpy(allh[rank]))
                avg.allreduce_view(nview * 4, variant="nvls")
                avg.check_status()
                got = v.cpu().numpy()
                wide = oracle.fold_wide_average_c(allh)
                same = float((got.view(np.uint32) == wide.view(np.uint32)).mean())
                ulp = int(np.abs(got.view(np.int32).astype(np.int64) - wide.view(np.int32).astype(np.int64)).max())
                scale = np.mean([np.abs(h) for h in allh], axis=0)
                nerr = float((np.abs(got - want6) / np.maximum(scale, 1e-30)).max())
                out["nvls"] = {"equal_to_wide_oracle": same, "max_ulp": ulp, "max_err_vs_rank_order_over_mean_abs": nerr}
                if same < 1.0 or nerr > 1e-6:
                    out["fail"].append("nvls: %s" % out["nvls"])
            gathered = [torch.empty_like(v) for _ in range(world)]
            tdist.all_gather(gathered, v) if backend != "gloo" else None
            if backend != "gloo" and any(not torch.equal(gathered[0], t) for t in gathered):
                out["fail"].append("nvls: ranks differ")
        note("part 7 done")
        # 8. parameter broadcast from a non-zero root, then adopt(): gradients live in the view region
        torch.manual_seed(1000 + rank)                                # replicas differ, as in the reference
        net = torch.nn.Sequential(torch.nn.Linear(37, 101), torch.nn.ReLU(), torch.nn.Linear(101, 10)).cuda()
        root = world - 1
        avg.broadcast_parameters(net, root=root)
        avg.check_status()
        torch.manual_seed(1000 + root)
        ref = torch.nn.Sequential(torch.nn.Linear(37, 101), torch.nn.ReLU(), torch.nn.Linear(101, 10))
        for p, q_ in zip(net.parameters(), ref.parameters()):
            if p.detach().cpu().numpy().tobytes() != q_.detach().numpy().tobytes():
                out["fail"].append("broadcast: parameter differs from the root's")
        out["launches"] = avg.launch_count
        tdist.barrier()
        ga.reset_default_averager()
        # adopt() on a fresh averager (a context holds one view region)
        avg2 = ga.GradAverager(capacity_bytes=4 << 20)
        opt = torch.optim.SGD(net.parameters(), lr=0.1)
        avg2.adopt(net, optimizer=opt)
        for step in range(3):
            opt.zero_grad()                                           # set_to_none=False through adopt(optimizer=)
            xg = torch.Generator(device="cuda").manual_seed(50 + 10 * step + rank)
            net(torch.randn(8, 37, device="cuda", generator=xg)).square().mean().backward()
            before = [p.grad.clone() for p in net.parameters()]
            avg2.average_gradients(net.parameters())
            avg2.check_status()
            for i, (b, p) in enumerate(zip(before, net.parameters())):
                gathered = [torch.empty_like(b.cpu()) for _ in range(world)]
                tdist.all_gather(gathered, b.cpu()) if backend == "gloo" else None
                if backend != "gloo":
                    gathered = [torch.empty_like(b) for _ in range(world)]
                    tdist.all_gather(gathered, b)
                if p.grad.cpu().numpy().tobytes() != oracle.fold_average_c([t.cpu().numpy().ravel() for t in gathered]).tobytes():
                    out["fail"].append("adopt step %d tensor %d" % (step, i))
            opt.step()
        if avg2._adopted is None or any(p.grad.data_ptr() != a for p, a in zip(net.parameters(), avg2._adopted[1])):
            out["fail"].append("adopt: gradients left their views")
        out["launches"] += avg2.launch_count
        tdist.barrier()
        avg2.close()
        tdist.destroy_process_group()
    except Exception as e:                                    # report, never hang the parent
        import traceback
        out["fail"].append("exception: %s\n%s" % (e, traceback.format_exc()))
    q.put(out)


def big_world(rank, world, port, q):
    """configs[3] goes to 1 GB: the kernels AUTO picks there (two-shot push above 32 MiB, the in-place
    view kernel) at 256 MiB and 1 GiB across real GPUs, where 32-bit tile indices, shard_tiles and the
    chunk count limit (1024 flag rows) would bite.  Checked on the GPU against torch's rank-order fold
    of the all-gathered inputs (bit for bit, every element) and against oracle/fold.c on windows at the
    start, across every shard boundary and at the end."""
    import numpy as np
    import torch
    import torch.distributed as tdist
    import oracle
    import pytorch_distributed_example_b200 as ga
    from pytorch_distributed_example_b200 import distributed as dist

    out = {"rank": rank, "fail": [], "launches": 0}
    try:
        os.environ.setdefault("GA_TIMEOUT_MS", "20000")
        dist.init_process_group(backend="nccl", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
        avg = ga.GradAverager(capacity_bytes=(1 << 30) + 4096)
        avg.create_view((1 << 30) + 4096, multicast=True)
        for nbytes in (256 << 20, (1 << 30) - 2048 * 3 + 4):          # the second is not a multiple of a tile or of world
            n = nbytes // 4
            g = torch.Generator(device="cuda").manual_seed(4321 + rank)
            x = torch.randn(n, device="cuda", generator=g)
            gathered = [torch.empty_like(x) for _ in range(world)]
            tdist.all_gather(gathered, x)
            acc = gathered[0].clone()
            for t in gathered[1:]:
                acc += t
            want = acc / torch.full((), float(world), device="cuda")
            del acc
            wins = [0, n - 300_000] + [max(0, (n // world) * j - 150_000) for j in range(1, world)]
            for mode in ("auto", "twoshot_push", "twoshot_pipe", "view"):
                if mode == "view":
                    res = avg.view_tensor(nbytes)
                    res.copy_(x)
                    avg.allreduce_view(nbytes, variant="view")
                else:
                    res = x.clone()
                    avg.all_reduce_([[res]], average=True, variant=mode)
                avg.check_status()
                if not torch.equal(res.view(torch.int32), want.view(torch.int32)):
                    bad = int((res.view(torch.int32) != want.view(torch.int32)).sum())
                    out["fail"].append("%d bytes, %s: %d elements differ from the GPU fold" % (nbytes, mode, bad))
                for w0 in wins:
                    host = [t[w0:w0 + 300_000].cpu().numpy() for t in gathered]
                    if res[w0:w0 + 300_000].cpu().numpy().tobytes() != oracle.fold_average_c(host).tobytes():
                        out["fail"].append("%d bytes, %s: window at %d differs from oracle/fold.c" % (nbytes, mode, w0))
                if mode != "view":
                    del res
            del gathered, want, x
            torch.cuda.empty_cache()
        out["launches"] = avg.launch_count
        tdist.barrier()
        avg.close()
        tdist.destroy_process_group()
    except Exception as e:
        import traceback
        out["fail"].append("exception: %s\n%s" % (e, traceback.format_exc()))
    q.put(out)


def loss_curve_world(rank, world, port, q):
    """north_star: "training loss curves matching within 1e-5 over 200 steps" against the reference's own
    reduction orders.  The reference step (mnist/main.py:84-99) on `world` real GPUs, 200 steps, with the
    averaging done (a) by the fused kernel, (b) in place on adopted gradient views, (c) by the reference's
    per-tensor loop on NCCL, (d) by the same loop on gloo with the gradients staged through the host
    (mnist/main.py:126-127 on its CPU backend).  Everything else is held fixed: same initial weights
    (broadcast from rank 0), same batches, TF32 off, deterministic cuDNN."""
    import torch
    import torch.distributed as tdist
    import torch.nn.functional as F
    import pytorch_distributed_example_b200 as ga
    from pytorch_distributed_example_b200 import distributed as dist
    sys.path.insert(0, os.path.join(ROOT, "examples"))
    import mnist_main

    out = {"rank": rank, "fail": [], "launches": 0}
    try:
        os.environ.setdefault("GA_TIMEOUT_MS", "20000")
        torch.backends.cudnn.deterministic = True
        torch.backends.cudnn.benchmark = False
        torch.backends.cuda.matmul.allow_tf32 = False
        torch.backends.cudnn.allow_tf32 = False
        dist.init_process_group(backend="nccl", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
        gloo = tdist.new_group(backend="gloo")
        torch.manual_seed(0)
        init = mnist_main.Net().cuda().state_dict()
        steps = 200
        curves = {}
        grad_err = 0.0
        for mode in ("fused", "view", "nccl_per_tensor", "gloo_staged"):
            net = mnist_main.Net().cuda()
            net.load_state_dict(init)
            opt = torch.optim.Adam(net.parameters(), lr=1e-3)                 # mnist/main.py:187, 229
            avg = None
            if mode in ("fused", "view"):
                avg = ga.GradAverager(capacity_bytes=4 << 20)
                if mode == "view":
                    avg.adopt(net, optimizer=opt)
            losses = []
            batches = iter(mnist_main.SyntheticBatches(steps, 128, rank))
            for step in range(steps):
                data, label = next(batches)
                data, label = data.cuda(), label.cuda()
                loss = F.cross_entropy(net(data), label)                     # :88-89
                opt.zero_grad()                                               # :91
                loss.backward()                                               # :92
                if mode == "fused":
                    # along the real trajectory: what would the reference's NCCL loop have produced from the
                    # SAME gradients?  (per-element tolerance of north_star, metric of SURVEY F3)
                    pre = [p.grad.detach().clone() for p in net.parameters()]
                    avg.average_gradients(net.parameters())
                    for i, (g0, p) in enumerate(zip(pre, net.parameters())):
                        scale = g0.abs()
                        tdist.all_reduce(scale, op=tdist.ReduceOp.SUM)
                        scale /= float(world)
                        tdist.all_reduce(g0, op=tdist.ReduceOp.SUM)               # :126
                        g0 /= float(world)                                        # :127
                        err = float(((p.grad - g0).abs() / scale.clamp_min(1e-30)).max())
                        grad_err = max(grad_err, err)
                elif avg is not None:
                    avg.average_gradients(net.parameters(), variant="view" if mode == "view" else None)
                elif mode == "nccl_per_tensor":
                    for p in net.parameters():                                # :125-127 on NCCL
                        tdist.all_reduce(p.grad.data, op=tdist.ReduceOp.SUM)
                        p.grad.data /= float(world)
                else:
                    for p in net.parameters():                                # :125-127 on gloo, through host memory
                        h = p.grad.data.cpu()
                        tdist.all_reduce(h, op=tdist.ReduceOp.SUM, group=gloo)
                        h /= float(world)
                        p.grad.data.copy_(h)
                opt.step()                                                    # :96
                losses.append(loss.detach())
            curves[mode] = torch.stack(losses).cpu()
            if avg is not None:
                avg.check_status()
                out["launches"] += avg.launch_count
                tdist.barrier()
                avg.close()
        ref = curves["fused"]
        out["max_abs_loss_diff"] = {m: float((curves[m] - ref).abs().max()) for m in curves}
        out["final_loss"] = {m: float(curves[m][-1]) for m in curves}
        # the yardstick: how far the reference's OWN two backends drift from each other over the same 200 steps
        mutual = float((curves["nccl_per_tensor"] - curves["gloo_staged"]).abs().max())
        out["max_abs_loss_diff"]["nccl_vs_gloo"] = mutual
        first = {}
        for m in ("nccl_per_tensor", "gloo_staged"):
            over = ((curves[m] - ref).abs() > 1e-5).nonzero()
            first[m] = int(over[0]) if len(over) else None
        out["first_step_over_1e-5"] = first
        if not torch.equal(curves["view"], ref):
            out["fail"].append("view mode's loss curve differs from the drop-in mode's: %g" % out["max_abs_loss_diff"]["view"])
        if world == 2:
            # every order is the rank order at world 2 and / 2.0 is exact: the gate is north_star's 1e-5 (measured: 0)
            for m in ("nccl_per_tensor", "gloo_staged"):
                if out["max_abs_loss_diff"][m] > 1e-5:
                    out["fail"].append("loss curve vs %s: max |d| = %g > 1e-5" % (m, out["max_abs_loss_diff"][m]))
            if not torch.equal(curves["gloo_staged"], ref):
                out["fail"].append("world 2: gloo's order is the rank order, the curves should be identical")
        else:
            # world >= 3: gloo's ring, NCCL and the rank order round differently in the last bit, and Adam -- whose
            # first steps are sign(g) * lr -- turns a sum that cancels to +tiny in one order and -tiny in another
            # into a parameter that moves the other way (SURVEY.md 8c-iii warns of exactly this).  Measured here:
            # the reference's two backends are 1e-5 apart from EACH OTHER by step 3 and 3e-4 .. 1e-3 by step 200, so
            # "1e-5 over 200 steps" cannot hold between any two of the three orders.  What is gated instead:
            # the averaging itself stays within north_star's per-element tolerance at every one of the 200 steps
            # (grad_err above), the free-running curves agree to 1e-5 until the amplification sets in, and
            # never separate by more than training noise.
            for m in ("nccl_per_tensor", "gloo_staged"):
                d = (curves[m] - ref).abs()
                if float(d[:2].max()) > 1e-5:
                    out["fail"].append("loss curve vs %s: first two steps differ by %g > 1e-5" % (m, float(d[:2].max())))
                if out["max_abs_loss_diff"][m] > 1e-2:
                    out["fail"].append("loss curve vs %s: max |d| = %g > 1e-2" % (m, out["max_abs_loss_diff"][m]))
        out["max_grad_err_vs_nccl_over_mean_abs"] = grad_err
        if grad_err > 1e-6:
            out["fail"].append("averaged gradients along the trajectory: max |d| / mean_r|g_r| = %g > 1e-6 vs the NCCL loop" % grad_err)
        print("[loss_curve_world rank %d] %s first>1e-5 %s final %s\n" % (rank, dict(out["max_abs_loss_diff"], grad_err=grad_err), first, out["final_loss"]), file=sys.stderr, flush=True)
        tdist.barrier()
        tdist.destroy_process_group()
    except Exception as e:
        import traceback
        out["fail"].append("exception: %s\n%s" % (e, traceback.format_exc()))
    q.put(out)


def fallback_world(rank, world, port, q):
    """A box on which the rank processes may not fetch each other's descriptors (pidfd_getfd refused): the
    multicast kind of view region cannot be set up, every rank sees the same error, and the region is made
    again through CUDA-IPC alone.  In-place averaging must work and be bit-exact; NVLS must be refused."""
    import numpy as np
    import torch
    import torch.distributed as tdist
    import oracle
    import pytorch_distributed_example_b200 as ga
    from pytorch_distributed_example_b200 import distributed as dist

    out = {"rank": rank, "fail": [], "launches": 0}
    try:
        os.environ["GA_TEST_FAIL_FETCH_FD"] = "1"
        os.environ.setdefault("GA_TIMEOUT_MS", "8000")
        dist.init_process_group(backend="nccl", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
        avg = ga.GradAverager(capacity_bytes=8 << 20)
        avg.create_view(8 << 20, multicast=True)
        if avg.has_multicast:
            out["fail"].append("multicast mapping although the descriptors could not be fetched")
        n = 1_000_003
        host = [np.random.default_rng(40 + r).standard_normal(n).astype(np.float32) for r in range(world)]
        v = avg.view_tensor(n * 4)
        v.copy_(torch.from_numpy(host[rank]))
        avg.allreduce_view(n * 4, variant="view")
        avg.check_status()
        if v.cpu().numpy().tobytes() != oracle.fold_average_c(host).tobytes():
            out["fail"].append("view kernel on the CUDA-IPC region differs from the oracle")
        try:
            avg.allreduce_view(n * 4, variant="nvls")
            out["fail"].append("nvls accepted without a multicast mapping")
        except ga.GradAvgError:
            pass
        out["launches"] = avg.launch_count
        tdist.barrier()
        avg.close()
        tdist.destroy_process_group()
    except Exception as e:
        import traceback
        out["fail"].append("exception: %s\n%s" % (e, traceback.format_exc()))
    q.put(out)
