"""Generates tests/golden/*.npz|json by running the REAL reference from /root/reference.

Run here (build container, CPU, gloo):   python tests/golden/make_golden.py
It cannot run on the GPU box (no /root/reference there); the files it writes are committed.

What is executed is the reference's own code, not a restatement:
  * mnist: `/root/reference/mnist/main.py` is loaded as text, the one-token syntax error on
    line 211 (`type=str,x` -> `type=str,`) is patched in memory, the module is
    exec'd, and its `Trainer.average_gradients` (`:122-127`) is called on an nn.Module whose
    `.grad`s were set from seeded tensors.  `Trainer.__init__` (`:56-63`) only stores its
    arguments, so any module works as `net`.
  * toy: `/root/reference/toy/main.py` is loaded unmodified; its `foo` (`:9-25`) is run with
    `sleep` replaced by a no-op and `random.seed(100 + rank)`; the printed lines are captured.

Outputs
  mini_ws{2,3,4,8}.npz   inputs of every rank + reference output (MINI_SHAPES, full arrays)
  mnist_cnn.json         per world size: sha256 of the reference output on the full 431 080-
                         element set (reference `Net`), element 0..7 of every tensor, and the
                         F3 statistics against the rank-order fold
  toy.json               per world size: the values drawn and the reduced sums printed
"""
import hashlib
import io
import json
import os
import sys
import types
import contextlib

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"


def load_reference_mnist():
    src = open(os.path.join(REF, "mnist", "main.py")).read()
    assert "type=str,x" in src
    src = src.replace("type=str,x", "type=str,")
    mod = types.ModuleType("reference_mnist_main")
    mod.__file__ = os.path.join(REF, "mnist", "main.py")
    exec(compile(src, mod.__file__, "exec"), mod.__dict__)
    return mod


def load_reference_toy():
    src = open(os.path.join(REF, "toy", "main.py")).read()
    mod = types.ModuleType("reference_toy_main")
    mod.__file__ = os.path.join(REF, "toy", "main.py")
    exec(compile(src, mod.__file__, "exec"), mod.__dict__)
    return mod


class Bag(torch.nn.Module):
    def __init__(self, shapes):
        super().__init__()
        self.ps = torch.nn.ParameterList([torch.nn.Parameter(torch.zeros(s)) for s in shapes])


def _mnist_job(rank, world, shapes, use_ref_net, positive):
    import warnings
    warnings.simplefilter("ignore")
    from oracle.gradient_sets import seeded_grads, seeded_positive_grads
    ref = load_reference_mnist()
    net = ref.Net() if use_ref_net else Bag(shapes)
    grads = (seeded_positive_grads if positive else seeded_grads)(shapes, rank)
    for p, g in zip(net.parameters(), grads):
        assert tuple(p.shape) == tuple(g.shape)
        p.grad = g.clone()
    trainer = ref.Trainer(net, None, None, None, torch.device("cpu"), True, False)
    trainer.average_gradients()                      # the reference's own method
    return [p.grad.numpy().copy() for p in net.parameters()]


def _toy_job(rank, world, steps):
    import random
    ref = load_reference_toy()
    ref.sleep = lambda s: None
    random.seed(100 + rank)
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        ref.foo(rank, world, steps)
    return buf.getvalue()


def main():
    from oracle import gloo_path, MINI_SHAPES, MNIST_CNN_SHAPES, seeded_grads, seeded_positive_grads
    from oracle import fold_average, raw_relative_error, backward_error
    meta = {"torch": torch.__version__, "worlds": {}}
    toy = {"torch": torch.__version__, "worlds": {}}
    for ws in (2, 3, 4, 8):
        res = gloo_path.run_world(ws, _mnist_job, (MINI_SHAPES, False, False))
        for r in range(1, ws):
            assert all(a.tobytes() == b.tobytes() for a, b in zip(res[0], res[r]))
        arrays = {}
        for r in range(ws):
            for i, g in enumerate(seeded_grads(MINI_SHAPES, r)):
                arrays["in_r%d_t%d" % (r, i)] = g.numpy()
        for i, o in enumerate(res[0]):
            arrays["out_t%d" % i] = o
        np.savez(os.path.join(HERE, "mini_ws%d.npz" % ws), **arrays)

        entry = {}
        for positive in (False, True):
            res = gloo_path.run_world(ws, _mnist_job, (MNIST_CNN_SHAPES, True, positive))
            for r in range(1, ws):
                assert all(a.tobytes() == b.tobytes() for a, b in zip(res[0], res[r]))
            flat = np.concatenate([o.ravel() for o in res[0]])
            gen = seeded_positive_grads if positive else seeded_grads
            ins = [np.concatenate([g.numpy().ravel() for g in gen(MNIST_CNN_SHAPES, r)]) for r in range(ws)]
            fold = fold_average(ins)
            entry["positive" if positive else "signed"] = {
                "sha256": hashlib.sha256(flat.tobytes()).hexdigest(),
                "fold_sha256": hashlib.sha256(fold.tobytes()).hexdigest(),
                "head": [[float(x) for x in o.ravel()[:8]] for o in res[0]],
                "n_differ_from_fold": int((flat != fold).sum()),
                "raw_rel_vs_fold": raw_relative_error(fold, flat),
                "backward_err_vs_fold": backward_error(fold, flat, ins),
            }
        meta["worlds"][str(ws)] = entry

        outs = gloo_path.run_world(ws, _toy_job, (3,))
        toy["worlds"][str(ws)] = outs
        print("ws", ws, "done", flush=True)
    json.dump(meta, open(os.path.join(HERE, "mnist_cnn.json"), "w"), indent=1)
    json.dump(toy, open(os.path.join(HERE, "toy.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
