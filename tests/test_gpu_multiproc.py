"""Real multi-GPU path (one process per GPU, CUDA-IPC peer buffers over NVLink).  Needs >= 2
GPUs: runs under `gpurun --gpus N`, skipped on a one-GPU box where the same kernels are covered
by the hosted-world tests in test_gpu_parity.py."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(world, backend, worker="ipc_world", timeout=150, min_launches=200):
    import torch.multiprocessing as mp
    from oracle.gloo_path import free_port
    from tests import _mp_workers
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    target = getattr(_mp_workers, worker)
    args = (lambda r: (r, world, port, q, backend)) if worker == "ipc_world" else (lambda r: (r, world, port, q))
    procs = [ctx.Process(target=target, args=args(r)) for r in range(world)]
    [p.start() for p in procs]
    res = []
    try:
        for _ in procs:
            res.append(q.get(timeout=timeout))
            if res[-1]["fail"]:
                break                                   # one rank failed: the others may be stuck in a collective
    finally:
        for p in procs:
            p.join(5 if (res and res[-1]["fail"]) or len(res) < len(procs) else 60)
            if p.is_alive():
                p.terminate()
    for d in res:
        assert not d["fail"], d["fail"]
        assert d["launches"] > min_launches
    return res


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("backend", ["gloo", "nccl"])
def test_two_ranks_ipc(backend):
    _run(2, backend)


@pytest.mark.skipif(torch.cuda.device_count() < 4, reason="needs 4 GPUs")
def test_four_ranks_ipc():
    _run(4, "nccl")


@pytest.mark.skipif(torch.cuda.device_count() < 8, reason="needs 8 GPUs")
def test_eight_ranks_ipc():
    _run(8, "nccl")


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_view_region_falls_back_to_cuda_ipc_two_ranks():
    """Descriptor passing refused (test hook): the view region is re-made without multicast, still bit-exact."""
    _run(2, "nccl", worker="fallback_world", timeout=200, min_launches=0)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_large_sets_two_ranks():
    """256 MiB and ~1 GiB per rank (configs[3]'s upper end) through two-shot push, AUTO and the view kernel."""
    _run(2, "nccl", worker="big_world", timeout=400, min_launches=7)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_loss_curves_two_ranks():
    """200 steps of the reference training step on 2 GPUs: ours == view mode == gloo-staged (rank order at
    world 2), and within 1e-5 of the per-tensor NCCL loop."""
    _run(2, "nccl", worker="loss_curve_world", timeout=400, min_launches=399)


@pytest.mark.skipif(torch.cuda.device_count() < 4, reason="needs 4 GPUs")
def test_loss_curves_four_ranks():
    """at world 4 gloo's ring and NCCL add in another order than the rank order (and than
    each other).  Identical to the view mode; the averaged gradients within 1e-6 (normalised) of the NCCL
    loop's at every step of the trajectory; free-running curves within 1e-5 until Adam's amplification of
    last-bit differences sets in (the two reference backends are 1e-5 apart from each other by step 3)."""
    _run(4, "nccl", worker="loss_curve_world", timeout=400, min_launches=399)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_reference_style_script_two_ranks(tmp_path):
    """A script in the reference's style, run by the drop-in runner on 2 GPUs: once with the fused
    `average_gradients` installed, once with its own per-tensor loop over our `distributed.all_reduce`.
    At world_size 2 `/ 2.0` is exact either way, so both runs print the same final losses."""
    import os
    import subprocess
    import sys
    from oracle.gloo_path import free_port
    from tests.test_run_script import STYLE_SCRIPT, ROOT
    script = tmp_path / "main.py"
    script.write_text(STYLE_SCRIPT)
    losses = {}
    for mode in ("fused", "per-tensor"):
        port = free_port()
        procs = []
        for r in range(2):
            cmd = [sys.executable, "-m", "pytorch_distributed_example_b200.run_script", str(script), "--ga-synthetic", "6"]
            if mode == "per-tensor":
                cmd.append("--ga-per-tensor")
            cmd += ["--", "--rank", str(r), "--world-size", "2", "--backend", "gloo", "-i", "tcp://127.0.0.1:%d" % port]
            env = dict(os.environ, GA_TIMEOUT_MS="8000")
            procs.append(subprocess.Popen(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, env=env))
        outs = []
        for p in procs:
            try:
                o, e = p.communicate(timeout=240)
            except subprocess.TimeoutExpired:
                for q in procs:
                    q.kill()
                raise
            assert p.returncode == 0, e[-2000:]
            outs.append(o)
        losses[mode] = [[l for l in o.splitlines() if l.startswith("final loss")][0] for o in outs]
        assert all("done" in o for o in outs)
    assert losses["fused"] == losses["per-tensor"]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_entry_points_restore_the_callers_device(built):
    """A process driving several GPUs: a context on cuda:0 used while cuda:1 is current works and
    leaves cuda:1 current (every entry point switches to its device and back)."""
    import ctypes
    import oracle
    lib = built
    torch.cuda.set_device(1)
    world = 2
    ctx = ctypes.c_void_p()
    assert lib.ga_create_local(world, 0, 1 << 20, ctypes.byref(ctx)) == 0, lib.ga_last_error()
    try:
        assert torch.cuda.current_device() == 1
        host = [[g.numpy() for g in oracle.seeded_grads(oracle.MINI_SHAPES, r)] for r in range(world)]
        want = [oracle.fold_average_c([host[r][i] for r in range(world)]) for i in range(len(oracle.MINI_SHAPES))]
        dev = [[torch.from_numpy(a).to("cuda:0") for a in rank] for rank in host]
        n = len(dev[0])
        ptrs = (ctypes.c_void_p * (n * world))(*[t.data_ptr() for ts in dev for t in ts])
        numel = (ctypes.c_int64 * n)(*[t.numel() for t in dev[0]])
        st = ctypes.c_void_p(torch.cuda.current_stream(torch.device("cuda", 0)).cuda_stream)
        for variant in (1, 2, 6):
            assert lib.ga_average(ctx, ptrs, numel, n, st, variant) == 0, lib.ga_last_error()
            x = torch.ones(4, device="cuda")                   # allocated on the CURRENT device
            assert x.device.index == 1 and torch.cuda.current_device() == 1
        torch.cuda.synchronize(0)
        del want                                               # three averages in a row: the values are covered elsewhere
        assert lib.ga_status(ctx) == 0
    finally:
        lib.ga_destroy(ctx)
        assert torch.cuda.current_device() == 1
        torch.cuda.set_device(0)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_toy_example_prints_the_reference_lines(golden_dir):
    """examples/toy_main.py on 2 GPUs (int32 SUM through the fused kernel) prints, line for line,
    what the REAL reference toy/main.py printed on gloo with the same seeds (tests/golden/toy.json)."""
    import json
    import os
    import subprocess
    import sys
    from oracle.gloo_path import free_port
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    want = json.load(open(os.path.join(golden_dir, "toy.json")))["worlds"]["2"]
    port = free_port()
    procs = [subprocess.Popen([sys.executable, os.path.join(root, "examples", "toy_main.py"), "-s", "2", "-r", str(r),
                               "--steps", "3", "--pause", "0", "--seed", "100", "-i", "tcp://127.0.0.1:%d" % port],
                              stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True,
                              env=dict(os.environ, GA_TIMEOUT_MS="8000")) for r in range(2)]
    for r, p in enumerate(procs):
        try:
            out, err = p.communicate(timeout=200)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        assert p.returncode == 0, err[-1500:]
        got = "".join(l + "\n" for l in out.splitlines() if l.startswith("rank:"))
        assert got == want[r]
