"""Real multi-GPU path (one process per GPU, CUDA-IPC peer buffers over NVLink).  Needs >= 2
GPUs: runs under `gpurun --gpus N`, skipped on a one-GPU box where the same kernels are covered
by the hosted-world tests in test_gpu_parity.py."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(world, backend):
    import torch.multiprocessing as mp
    from oracle.gloo_path import free_port
    from tests._mp_workers import ipc_world
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=ipc_world, args=(r, world, port, q, backend)) for r in range(world)]
    [p.start() for p in procs]
    res = []
    try:
        for _ in procs:
            res.append(q.get(timeout=240))
            if res[-1]["fail"]:
                break                                   # one rank failed: the others may be stuck in a collective
    finally:
        for p in procs:
            p.join(5 if (res and res[-1]["fail"]) or len(res) < len(procs) else 60)
            if p.is_alive():
                p.terminate()
    for d in res:
        assert not d["fail"], d["fail"]
        assert d["launches"] > 200


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
