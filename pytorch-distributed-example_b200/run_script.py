"""Run a reference-style training script on the fused path with nothing but its imports changed.

    python -m pytorch_distributed_example_b200.run_script path/to/mnist/main.py [--ga-synthetic N] [--ga-per-tensor] -- --rank 0 --world-size 2
    python -m pytorch_distributed_example_b200.run_script path/to/toy/main.py -- --rank 0 --world-size 2 --steps 3

The script is read as text and these substitutions are applied IN MEMORY (the file is not touched):
  * `from torch import distributed, nn`   -> torch's `nn`, this package's `distributed`   [mnist/main.py:7]
  * `import torch.distributed as dist`    -> `from pytorch_distributed_example_b200 import distributed as dist`  [toy/main.py:6]
  * `type=str,x`                          -> `type=str,`  (the syntax error at mnist/main.py:211)
It is then executed as a module (not as `__main__`), and before its `main()` is called:
  * a `Trainer` class, if the script defines one, gets the fused `average_gradients` (`install`),
  * with `--ga-synthetic N` a `get_dataloader(root, batch_size, world_size)` function, if defined, is
    replaced by N seeded MNIST-shaped batches per epoch (no dataset download on an air-gapped box, F7).
Everything else -- argument parsing, model, training loop, prints -- is the script's own code.
"""
import os
import sys
import types

SUBSTITUTIONS = (
    ("from torch import distributed, nn",
     "from torch import nn\nfrom pytorch_distributed_example_b200 import distributed"),
    ("import torch.distributed as dist",
     "from pytorch_distributed_example_b200 import distributed as dist"),
    ("type=str,x", "type=str,"),
)


def rewrite(source):
    applied = []
    for old, new in SUBSTITUTIONS:
        if old in source:
            source = source.replace(old, new)
            applied.append(old)
    return source, applied


def synthetic_loader_factory(n_batches):
    import torch

    def get_dataloader(root, batch_size, world_size):
        rank = 0
        try:
            import torch.distributed as td
            rank = td.get_rank() if td.is_initialized() else 0
        except Exception:
            pass
        g = torch.Generator().manual_seed(777 + rank)
        xs = torch.randn(n_batches, batch_size, 1, 28, 28, generator=g)
        ys = torch.randint(0, 10, (n_batches, batch_size), generator=g)
        train = list(zip(xs, ys))
        test = train[: max(1, n_batches // 10)]
        return train, test
    return get_dataloader


def load(path, synthetic=0, module_name="ga_user_script", install_trainer=True):
    """Returns the executed module (its `main` not yet called) and the substitutions applied."""
    source, applied = rewrite(open(path).read())
    mod = types.ModuleType(module_name)
    mod.__file__ = os.path.abspath(path)
    sys.modules[module_name] = mod
    exec(compile(source, mod.__file__, "exec"), mod.__dict__)
    if install_trainer and isinstance(getattr(mod, "Trainer", None), type) and hasattr(mod.Trainer, "average_gradients"):
        from .trainer import install
        install(mod.Trainer)
        applied.append("Trainer.average_gradients -> fused")
    if synthetic and callable(getattr(mod, "get_dataloader", None)):
        mod.get_dataloader = synthetic_loader_factory(synthetic)
        applied.append("get_dataloader -> %d synthetic batches" % synthetic)
    return mod, applied


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    if not argv or argv[0] in ("-h", "--help"):
        print(__doc__)
        return 0
    path = argv.pop(0)
    synthetic = 0
    if argv and argv[0] == "--ga-synthetic":
        synthetic = int(argv[1])
        argv = argv[2:]
    install_trainer = True
    if argv and argv[0] == "--ga-per-tensor":      # keep the script's own loop: one fused launch per all_reduce
        install_trainer = False
        argv = argv[1:]
    if argv and argv[0] == "--":
        argv = argv[1:]
    mod, applied = load(path, synthetic, install_trainer=install_trainer)
    print("[run_script] %s: %s" % (path, "; ".join(applied) or "no substitutions"), file=sys.stderr)
    old_argv = sys.argv
    sys.argv = [path] + argv
    try:
        mod.main()
    finally:
        sys.argv = old_argv
        from .averager import reset_default_averager
        try:
            reset_default_averager()
        except Exception:
            pass
    return 0


if __name__ == "__main__":
    sys.exit(main())
