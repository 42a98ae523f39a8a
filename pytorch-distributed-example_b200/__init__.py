"""B200-native gradient averaging behind the reference's own call.

The one hot path of dblakely/pytorch-distributed-example -- `Trainer.average_gradients`
(`/root/reference/mnist/main.py:122-127`) and the toy all-reduce (`toy/main.py:19-20`) -- as a
fused sm_100a kernel (pack -> peer reduce in rank order -> divide -> unpack) in
`csrc/libgradavg.so`, reached through the C ABI of `include/gradavg.h`.

    from pytorch_distributed_example_b200 import distributed      # instead of torch's
    import pytorch_distributed_example_b200 as ga
    ga.install(Trainer)            # Trainer.average_gradients now runs the fused kernel
    ga.average_gradients(net)      # or call it directly

There is no CPU fallback: importing works without a GPU (so the host logic can be tested),
but every compute call raises if the CUDA library or a CUDA device is missing.
"""
from . import distributed  # noqa: F401
from ._lib import GradAvgError, lib_path, load_library  # noqa: F401
from .averager import (  # noqa: F401
    GradAverager, average_gradients, default_averager, reset_default_averager,
    ONESHOT, TWOSHOT, AUTO, ONESHOT_PUSH, TWOSHOT_PUSH, TWOSHOT_PIPE, HOSTED_DIRECT, VIEW, NVLS,
)
from .trainer import install  # noqa: F401
from .overlap import OverlappedAverager, plan_buckets  # noqa: F401

__all__ = ["distributed", "GradAverager", "GradAvgError", "average_gradients", "default_averager",
           "reset_default_averager", "install", "OverlappedAverager", "plan_buckets", "load_library", "lib_path", "ONESHOT", "TWOSHOT", "AUTO", "ONESHOT_PUSH", "TWOSHOT_PUSH", "TWOSHOT_PIPE", "HOSTED_DIRECT", "VIEW", "NVLS"]
