"""CPU: the drop-in runner -- a reference-style script runs with only its imports substituted.

With /root/reference present (build container) the REAL scripts are loaded; on the GPU box a small
script written in the same style stands in.  world_size 1 is used here because that path needs no
collective (`mnist/main.py:193,236`), so it runs without a GPU; the multi-rank runs of the same
runner are in tests/test_gpu_multiproc.py."""
import os
import subprocess
import sys
import textwrap

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"

STYLE_SCRIPT = textwrap.dedent('''
    from __future__ import division, print_function
    import argparse
    import torch
    import torch.nn.functional as F
    from torch import distributed, nn

    class Trainer(object):
        def __init__(self, net, optimizer, loader, device, distributed):
            self.net, self.optimizer, self.loader, self.device, self.distributed = net, optimizer, loader, device, distributed
        def fit(self):
            for data, label in self.loader:
                data, label = data.to(self.device), label.to(self.device)
                loss = F.cross_entropy(self.net(data), label)
                self.optimizer.zero_grad()
                loss.backward()
                if self.distributed:
                    self.average_gradients()
                self.optimizer.step()
            print("final loss: {:.6f}".format(loss.item()))
        def average_gradients(self):
            world_size = distributed.get_world_size()
            for p in self.net.parameters():
                distributed.all_reduce(p.grad.data, op=distributed.reduce_op.SUM)
                p.grad.data /= float(world_size)

    def get_dataloader(root, batch_size, world_size):
        raise RuntimeError("would download a dataset")

    def main():
        parser = argparse.ArgumentParser()
        parser.add_argument('--backend', type=str, default='nccl')
        parser.add_argument('-i', '--init-method', type=str, default='tcp://127.0.0.1:23456')
        parser.add_argument('-r', '--rank', type=int)
        parser.add_argument('-s', '--world-size', type=int)
        parser.add_argument('--no-cuda', action='store_true')
        args = parser.parse_args()
        if args.world_size != 1:
            distributed.init_process_group(backend=args.backend, init_method=args.init_method,
                                           rank=args.rank, world_size=args.world_size)
        device = torch.device('cuda' if torch.cuda.is_available() and not args.no_cuda else 'cpu')
        torch.manual_seed(0)
        net = nn.Sequential(nn.Flatten(), nn.Linear(784, 32), nn.ReLU(), nn.Linear(32, 10)).to(device)
        train, _ = get_dataloader('data', 16, args.world_size)
        Trainer(net, torch.optim.Adam(net.parameters(), lr=1e-3), train, device, args.world_size != 1).fit()
        print("rank {} done".format(args.rank))
''')


def test_rewrite_rules():
    from pytorch_distributed_example_b200 import run_script
    src, applied = run_script.rewrite("from torch import distributed, nn\nx = dict(type=str,x\n)\n")
    assert "pytorch_distributed_example_b200 import distributed" in src and "type=str,\n" in src
    assert len(applied) == 2
    src, applied = run_script.rewrite("import torch.distributed as dist\n")
    assert src.startswith("from pytorch_distributed_example_b200 import distributed as dist")


def test_style_script_world1(tmp_path):
    script = tmp_path / "main.py"
    script.write_text(STYLE_SCRIPT)
    out = subprocess.run([sys.executable, "-m", "pytorch_distributed_example_b200.run_script", str(script),
                          "--ga-synthetic", "4", "--", "--rank", "0", "--world-size", "1", "--no-cuda"],
                         cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "rank 0 done" in out.stdout and "final loss" in out.stdout
    assert "Trainer.average_gradients -> fused" in out.stderr and "synthetic" in out.stderr


@pytest.mark.skipif(not os.path.exists(os.path.join(REF, "mnist", "main.py")), reason="reference tree not on this box")
def test_real_reference_scripts_load_and_run_world1():
    """The unmodified reference files: mnist/main.py (which does not even parse as shipped) loads,
    gets the fused method installed and trains one epoch of synthetic batches at world_size 1."""
    from pytorch_distributed_example_b200 import run_script
    mod, applied = run_script.load(os.path.join(REF, "mnist", "main.py"), synthetic=3, module_name="ref_mnist_probe")
    assert "type=str,x" in applied and "from torch import distributed, nn" in applied
    assert mod.Trainer.average_gradients.__name__ == "fused_average_gradients"
    assert mod.distributed.__name__.endswith("b200.distributed")
    toy, applied = run_script.load(os.path.join(REF, "toy", "main.py"), module_name="ref_toy_probe")
    assert applied == ["import torch.distributed as dist"] and toy.dist.__name__.endswith("b200.distributed")
    out = subprocess.run([sys.executable, "-m", "pytorch_distributed_example_b200.run_script",
                          os.path.join(REF, "mnist", "main.py"), "--ga-synthetic", "3", "--",
                          "--rank", "0", "--world-size", "1", "--no-cuda", "--epochs", "1", "--batch-size", "8"],
                         cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "Epoch: 1/1" in out.stdout
