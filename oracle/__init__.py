"""oracle/ -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU restatements of the reference hot path (`/root/reference/mnist/main.py:122-127`,
`/root/reference/toy/main.py:9-25`).  Only `tests/`, `__graft_entry__.smoke()` and the
`cpu_baseline` / `--impl reference` legs of `bench.py` may import this package; the product
package `pytorch-distributed-example_b200` never does (a test enforces it).

Pinned how: the reference ships no tests, golden vectors or known answers for this path
(SURVEY.md section 4), so the pin is "outputs of the reference itself run here":
`tests/golden/make_golden.py` imports the real reference from `/root/reference` (line-211
typo patched in memory), runs its own `Trainer.average_gradients` / toy `foo`
on gloo and stores inputs + outputs under `tests/golden/`; `tests/test_oracle.py` checks
every function of this package against those files.
"""
from .host_fold import (  # noqa: F401
    fold_average, fold_average_c, fold_sum_i32, fold_sum_i32_c, build_c_oracle, fold_wide_average, fold_wide_average_c,
)
from .gradient_sets import (  # noqa: F401
    MNIST_CNN_SHAPES, MINI_SHAPES, resnet18_shapes, seeded_grads, seeded_positive_grads, numel_of,
)
from .metrics import backward_error, raw_relative_error  # noqa: F401
