"""GPU: the allocation-free entry point `ga_average_static` against `ga_average` and the oracle.

Some users of this library may not call cudaMalloc when they need an average (their host
application owns the allocator, or they sit in a section where cudaMalloc's implicit device
synchronisation is not acceptable), so `ga_average_static` keeps its bucket and flags in an
array built into the library.  This file checks, through the C ABI, that it returns the same
bytes as the context-based call on the same inputs and as oracle/fold.c -- the tolerance is zero,
as everywhere in these tests -- twice in a row (the second call finds the first call's leftovers in
the array), that its outcome can be read back without allocating (`ga_static_status`), that no
device memory is taken after the first call, and that it refuses gradient sets larger than the
built-in array.  It is named so that pytest
collects it last; it must pass alone in a fresh process as well as after the whole suite.
"""
import ctypes

import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu

ONESHOT, TWOSHOT = 1, 2


def _table(per_rank):
    n = len(per_rank[0])
    ptrs = (ctypes.c_void_p * (n * len(per_rank)))(*[t.data_ptr() for ts in per_rank for t in ts])
    numel = (ctypes.c_int64 * n)(*[t.numel() for t in per_rank[0]])
    return ptrs, numel, n


def _existing(lib, world, per_rank, variant):
    """Reference result: the existing entry point on a hosted-world context."""
    ctx = ctypes.c_void_p()
    assert lib.ga_create_local(world, 0, 8 << 20, ctypes.byref(ctx)) == 0, lib.ga_last_error()
    try:
        ptrs, numel, n = _table(per_rank)
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        assert lib.ga_average(ctx, ptrs, numel, n, st, variant) == 0, lib.ga_last_error()
        torch.cuda.synchronize()
        assert lib.ga_status(ctx) == 0
    finally:
        lib.ga_destroy(ctx)


def _static(lib, world, per_rank, variant):
    ptrs, numel, n = _table(per_rank)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    assert lib.ga_average_static(world, 0, ptrs, numel, n, st, variant) == 0, lib.ga_last_error()


def _oracle(host, world):
    return [oracle.fold_average_c([host[r][i].numpy() for r in range(world)]) for i in range(len(host[0]))]


def test_static_matches_context_call_twice_in_a_row(built):
    lib = built
    torch.cuda.set_device(0)
    world = 4
    shapes = oracle.MNIST_CNN_SHAPES                       # 4 ranks x 1.73 MB: inside the 8 MiB array
    for call, (variant, seed) in enumerate([(TWOSHOT, 1234), (ONESHOT, 4321), (TWOSHOT, 99)]):
        host = [oracle.seeded_grads(shapes, r, seed) for r in range(world)]
        want = [[t.cuda() for t in rank] for rank in host]
        got = [[t.cuda() for t in rank] for rank in host]
        _existing(lib, world, want, variant)
        _static(lib, world, got, variant)                  # back to back: no host sync between calls 1, 2, 3
        if call == 0:
            # second call straight after the first, different data, same array
            again_host = [oracle.seeded_grads(shapes, r, 777) for r in range(world)]
            again_want = [[t.cuda() for t in rank] for rank in again_host]
            again_got = [[t.cuda() for t in rank] for rank in again_host]
            _static(lib, world, again_got, ONESHOT)
            _existing(lib, world, again_want, ONESHOT)
            torch.cuda.synchronize()
            for r in range(world):
                for a, b in zip(again_got[r], again_want[r]):
                    assert torch.equal(a, b) and a.cpu().numpy().tobytes() == b.cpu().numpy().tobytes()
        torch.cuda.synchronize()
        assert lib.ga_static_status(0, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)) == 0, lib.ga_last_error()
        folded = _oracle(host, world)
        for r in range(world):
            for a, b, c in zip(got[r], want[r], folded):
                assert a.cpu().numpy().tobytes() == b.cpu().numpy().tobytes(), "call %d rank %d" % (call, r)
                assert a.cpu().numpy().tobytes() == c.tobytes(), "call %d rank %d vs oracle/fold.c" % (call, r)
    # a different world size afterwards reuses the same array from a different layout
    for world in (2, 8, 1):
        host = [oracle.seeded_grads(oracle.MINI_SHAPES, r) for r in range(world)]
        want = [[t.cuda() for t in rank] for rank in host]
        got = [[t.cuda() for t in rank] for rank in host]
        _existing(lib, world, want, TWOSHOT)
        _static(lib, world, got, TWOSHOT)
        torch.cuda.synchronize()
        for r in range(world):
            for a, b in zip(got[r], want[r]):
                assert a.cpu().numpy().tobytes() == b.cpu().numpy().tobytes()


def test_static_takes_no_device_memory_after_its_first_call(built):
    """The array is reserved when the library's device code is first used on the GPU; later calls
    must leave cudaMemGetInfo's free figure where it was."""
    lib = built
    torch.cuda.set_device(0)
    world = 2
    host = [oracle.seeded_grads(oracle.MINI_SHAPES, r) for r in range(world)]
    dev = [[t.cuda() for t in rank] for rank in host]
    _static(lib, world, dev, TWOSHOT)                       # first use on this device (if the other tests did not run)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(0)
    for variant in (ONESHOT, TWOSHOT, ONESHOT):
        dev = [[t.cuda() for t in rank] for rank in host]   # torch's cached blocks: no new cudaMalloc
        _static(lib, world, dev, variant)
        assert lib.ga_static_status(0, None) == 0
    free1, _ = torch.cuda.mem_get_info(0)
    assert free1 == free0
    want = _oracle(host, world)
    for r in range(world):
        for a, w in zip(dev[r], want):
            assert a.cpu().numpy().tobytes() == w.tobytes()


def test_static_rejects_what_does_not_fit(built):
    lib = built
    torch.cuda.set_device(0)
    world = 4
    big = [[torch.zeros(600_000, device="cuda")] for _ in range(world)]      # 4 x 2.4 MB > 8 MiB
    ptrs, numel, n = _table(big)
    assert lib.ga_average_static(world, 0, ptrs, numel, n, None, TWOSHOT) == -3
    assert b"8388608" in lib.ga_last_error()
    many = [[torch.zeros(4, device="cuda") for _ in range(40)] for _ in range(world)]   # 160 pointers > 128
    ptrs, numel, n = _table(many)
    assert lib.ga_average_static(world, 0, ptrs, numel, n, None, ONESHOT) == -3
    assert lib.ga_average_static(9, 0, ptrs, numel, n, None, ONESHOT) == -1
    assert all(float(t.sum()) == 0.0 for rank in big for t in rank)
