#!/usr/bin/env python
"""Data-parallel MNIST-shaped training with fused B200 gradient averaging.

Same command line, model and training-step order as the reference script
(`/root/reference/mnist/main.py`: flags `:207-233`, `Net` `:130-147`, step order `:84-99`), written
from scratch around this repository's `distributed` stand-in and fused `average_gradients`:

    python examples/mnist_main.py --rank 0 --world-size 2 --synthetic 200 &
    python examples/mnist_main.py --rank 1 --world-size 2 --synthetic 200

Differences, all forced by the environment: each rank is put on its own GPU,
and `--synthetic N` replaces the FashionMNIST/MNIST downloads by N seeded MNIST-shaped batches.
"""
import argparse
import os
import sys
import time

import torch
import torch.nn.functional as F
from torch import nn

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pytorch_distributed_example_b200 as ga                      # noqa: E402
from pytorch_distributed_example_b200 import distributed           # noqa: E402  (instead of torch's)


class Net(nn.Module):
    """conv(1->20,5) pool conv(20->50,5) pool fc(800->500) fc(500->10) log_softmax: 431 080 weights."""

    def __init__(self):
        super().__init__()
        self.conv1, self.conv2 = nn.Conv2d(1, 20, 5), nn.Conv2d(20, 50, 5)
        self.fc1, self.fc2 = nn.Linear(800, 500), nn.Linear(500, 10)

    def forward(self, x):
        x = F.max_pool2d(F.relu(self.conv1(x)), 2)
        x = F.max_pool2d(F.relu(self.conv2(x)), 2)
        x = F.relu(self.fc1(x.flatten(1)))
        return F.log_softmax(self.fc2(x), dim=1)


class SyntheticBatches:
    """N batches of MNIST-shaped noise, seeded per rank (777 + rank), held in pinned
    host memory so that the `.to(device)` of the training loop is a real H2D copy."""

    def __init__(self, n_batches, batch_size, rank, pin=True):
        g = torch.Generator().manual_seed(777 + rank)
        self.x = torch.randn(n_batches, batch_size, 1, 28, 28, generator=g)
        self.y = torch.randint(0, 10, (n_batches, batch_size), generator=g)
        if pin and torch.cuda.is_available():
            self.x, self.y = self.x.pin_memory(), self.y.pin_memory()

    def __iter__(self):
        return iter(zip(self.x, self.y))

    def __len__(self):
        return self.x.shape[0]


class Meter:
    def __init__(self):
        self.total, self.n = 0.0, 0

    def add(self, value, count):
        self.total += value * count
        self.n += count

    @property
    def mean(self):
        return self.total / max(1, self.n)


class Trainer:
    """fit / train / evaluate / average_gradients, as the reference names them."""

    def __init__(self, net, optimizer, train_loader, test_loader, device, distributed, do_eval):
        self.net, self.optimizer = net, optimizer
        self.train_loader, self.test_loader = train_loader, test_loader
        self.device, self.distributed, self.do_eval = device, distributed, do_eval
        self.losses = []

    def average_gradients(self):
        ga.average_gradients(self.net.parameters())          # mnist/main.py:122-127, fused

    def train(self):
        loss_m, acc_m = Meter(), Meter()
        self.net.train()
        for data, label in self.train_loader:
            data, label = data.to(self.device, non_blocking=True), label.to(self.device, non_blocking=True)
            output = self.net(data)
            loss = F.cross_entropy(output, label)
            self.optimizer.zero_grad()
            loss.backward()
            if self.distributed:
                self.average_gradients()
            self.optimizer.step()
            value = loss.item()                                # the per-step host sync of the reference
            self.losses.append(value)
            loss_m.add(value, data.size(0))
            acc_m.add((output.argmax(1) == label).float().mean().item(), data.size(0))
        return loss_m.mean, acc_m.mean

    @torch.no_grad()
    def evaluate(self):
        loss_m, acc_m = Meter(), Meter()
        self.net.eval()
        for data, label in self.test_loader:
            data, label = data.to(self.device), label.to(self.device)
            output = self.net(data)
            loss_m.add(F.cross_entropy(output, label).item(), data.size(0))
            acc_m.add((output.argmax(1) == label).float().mean().item(), data.size(0))
        return loss_m.mean, acc_m.mean

    def fit(self, epochs):
        for epoch in range(1, epochs + 1):
            t0 = time.time()
            tl, ta = self.train()
            el, ea = self.evaluate() if (self.do_eval or epoch == epochs) else (0.0, 0.0)
            print("Epoch: %d/%d, train loss: %.6f, train acc: %.2f%%, test loss: %.6f, test acc: %.2f%%. (%.1f steps/s)"
                  % (epoch, epochs, tl, 100 * ta, el, 100 * ea, len(self.train_loader) / (time.time() - t0)))


def real_loaders(root, batch_size, world_size):
    from torch.utils import data
    from torchvision import datasets, transforms
    tf = transforms.Compose([transforms.ToTensor(), transforms.Normalize((0.1307,), (0.3081,))])
    train = datasets.FashionMNIST(root, train=True, transform=tf, download=True)
    sampler = data.distributed.DistributedSampler(train) if world_size > 1 else None
    return (data.DataLoader(train, batch_size=batch_size, shuffle=sampler is None, sampler=sampler),
            data.DataLoader(datasets.MNIST(root, train=False, transform=tf, download=True), batch_size=batch_size))


def parse_args(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("--backend", type=str, default="nccl", help="Name of the backend to use.")
    p.add_argument("-i", "--init-method", type=str, default="tcp://127.0.0.1:23456")
    p.add_argument("-r", "--rank", type=int, default=0)
    p.add_argument("-s", "--world-size", type=int, default=1)
    p.add_argument("--epochs", type=int, default=20)
    p.add_argument("--no-cuda", action="store_true")
    p.add_argument("--learning-rate", "-lr", type=float, default=1e-3)
    p.add_argument("--root", type=str, default="data")
    p.add_argument("--batch-size", type=int, default=128)
    p.add_argument("--eval", action="store_true", default=False)
    p.add_argument("--synthetic", type=int, default=0, metavar="N", help="N seeded synthetic batches per epoch")
    return p.parse_args(argv)


def main(argv=None):
    args = parse_args(argv)
    print(args)
    if args.no_cuda:
        raise SystemExit("--no-cuda: this build has no CPU path (the reference's gloo path is the CPU baseline)")
    if args.world_size != 1:
        distributed.init_process_group(backend=args.backend, init_method=args.init_method,
                                       rank=args.rank, world_size=args.world_size)
    device = torch.device("cuda", torch.cuda.current_device())
    # The reference builds Net() on every rank with no seed and no broadcast (mnist/main.py:185), so its
    # replicas start from DIFFERENT weights.  Here every rank builds its own and rank 0's are replicated
    # over the peer mappings (ga_broadcast_tensors) -- no reliance on a shared seed.
    net = Net().to(device)
    if args.world_size != 1:
        ga.default_averager().broadcast_parameters(net, root=0)
    opt = torch.optim.Adam(net.parameters(), lr=args.learning_rate)
    if args.synthetic:
        train = SyntheticBatches(args.synthetic, args.batch_size, args.rank)
        test = SyntheticBatches(max(1, args.synthetic // 10), args.batch_size, 10_000 + args.rank)
    else:
        train, test = real_loaders(args.root, args.batch_size, args.world_size)
    trainer = Trainer(net, opt, train, test, device, args.world_size != 1, args.eval)
    trainer.fit(args.epochs)
    if args.world_size != 1:
        ga.reset_default_averager()
        distributed.destroy_process_group()
    return trainer


if __name__ == "__main__":
    main()
