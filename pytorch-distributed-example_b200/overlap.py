"""Gradient averaging overlapped with backward (BASELINE configs[4]).

The reference averages after `loss.backward()` has returned (`/root/reference/mnist/main.py:92-95`):
for a many-tensor model the whole exchange is exposed.  Here the parameters are cut into buckets
in REVERSE `parameters()` order (the order autograd finishes them); a
post-accumulate-grad hook counts arrivals, and as soon as a bucket is complete its fused averaging
kernel is enqueued on a side stream with a capped grid, so it runs next to the rest of backward.
`average_gradients()` at the reference's call site then only waits for the side stream.

    ov = OverlappedAverager(net)            # after init_process_group, before training
    ...
    loss.backward()                         # buckets fire from the hooks
    ov.average_gradients()                  # replaces Trainer.average_gradients: join, nothing else
    optimizer.step()

Every rank builds the same buckets from the same shapes and fires them in bucket order, so the
calls stay collective.  Results are the same bits as the non-overlapped call: bucketing only
changes which tensors share a launch, never the per-element rank-order sum.
"""
import torch

from . import _lib
from .averager import GradAverager, bucket_bytes


def plan_buckets(numels, cap_bytes):
    """Indices of `parameters()` grouped into buckets, last parameter first; a bucket closes when
    adding the next tensor would exceed `cap_bytes` of bucket space (it always takes one tensor)."""
    buckets, cur, cur_bytes = [], [], 0
    for i in reversed(range(len(numels))):
        b = bucket_bytes([numels[i]])
        if cur and cur_bytes + b > cap_bytes:
            buckets.append(cur)
            cur, cur_bytes = [], 0
        cur.append(i)
        cur_bytes += b
    if cur:
        buckets.append(cur)
    return buckets


# What a hook-driven step costs on the host, measured on ResNet-18 (62 tensors) at N = 2/4/8 in
# round 1 (exposed time 396-541 us overlapped against <= 153 us for one call after backward) and
# again in round 2 with the leaner launch path: a few microseconds per hook plus one launch, one
# event record and one event wait per bucket.
HOOK_US_PER_TENSOR = 4.0
LAUNCH_US_PER_BUCKET = 40.0
LINK_GBPS = 600.0            # what the fused call reaches across GPUs at tens of MB (profiles/)


def overlap_pays(numels, world, n_buckets):
    """(pays, exchange_us, overhead_us): the time one fused call after backward would expose
    against what the hooks cost.  Overlap can hide at most the former and always pays the latter."""
    nbytes = sum(numels) * 4
    exchange_us = nbytes * 2.0 * (world - 1) / max(world, 1) / (LINK_GBPS * 1e3) if world > 1 else nbytes * 2.0 / 6.0e6
    overhead_us = HOOK_US_PER_TENSOR * len(numels) + LAUNCH_US_PER_BUCKET * n_buckets
    return exchange_us > overhead_us, exchange_us, overhead_us


class OverlappedAverager:
    def __init__(self, modules, averager=None, bucket_cap_bytes=25 << 20, max_blocks=32, variant=None, force=None):
        """modules: the replica (one rank per process) or the list of replicas of a hosted world.

        force=None: overlap is ENABLED only if the model above says it can pay (`self.enabled`,
        `self.reason`); otherwise no hooks are registered and `average_gradients()` is the plain
        fused call at the reference's call site.  force=True / False overrides the model."""
        self.modules = [modules] if isinstance(modules, torch.nn.Module) else list(modules)
        self.params = [[p for p in m.parameters() if p.requires_grad] for m in self.modules]
        n = len(self.params[0])
        if any(len(ps) != n for ps in self.params):
            raise _lib.GradAvgError(_lib.GA_EINVAL, "replicas have different parameter lists")
        numels = [p.numel() for p in self.params[0]]
        self.buckets = plan_buckets(numels, bucket_cap_bytes)
        self.avg = averager if averager is not None else (
            GradAverager(local_world=len(self.modules)) if len(self.modules) > 1 else GradAverager())
        if self.avg.n_local != len(self.modules):
            raise _lib.GradAvgError(_lib.GA_EINVAL, "averager hosts %d ranks, got %d replicas"
                                    % (self.avg.n_local, len(self.modules)))
        pays, ex_us, ov_us = overlap_pays(numels, self.avg.world_size, len(self.buckets))
        self.enabled = pays if force is None else bool(force)
        self.reason = ("one fused call after backward exposes about %.0f us; hooks and %d bucket launches cost about "
                       "%.0f us per step: overlap %s" % (ex_us, len(self.buckets), ov_us,
                                                         "can pay" if pays else "cannot pay, using the single call"))
        self.avg.reserve(max(bucket_bytes([numels[i] for i in b]) for b in self.buckets) if self.enabled
                         else bucket_bytes(numels))
        self.max_blocks = int(max_blocks)
        self.variant = variant
        self._bucket_of = {}
        for k, idx in enumerate(self.buckets):
            for i in idx:
                self._bucket_of[i] = k
        self._need = [len(idx) * len(self.modules) for idx in self.buckets]
        self._bucket_params = [[[ps[i] for i in idx] for ps in self.params] for idx in self.buckets]
        self._side = None
        self._ready = [torch.cuda.Event() for _ in self.buckets] if self.enabled else []   # created once, re-recorded every step
        self._done = torch.cuda.Event() if self.enabled else None
        self._handles = []
        self._reset()
        if self.enabled:
            for ps in self.params:
                for i, p in enumerate(ps):
                    self._handles.append(p.register_post_accumulate_grad_hook(self._make_hook(i)))

    def _reset(self):
        self._have = [0] * len(self.buckets)
        self._next = 0                       # buckets are launched strictly in order (collective semantics)
        self._launched = 0

    def _make_hook(self, i):
        k = self._bucket_of[i]
        need = self._need[k]

        def hook(_param):
            # autograd runs the hooks of one device's backward pass on one thread: no lock on this path
            self._have[k] += 1
            if k == self._next and self._have[k] == need:
                self._fire_ready()
        return hook

    def _fire_ready(self):
        while self._next < len(self.buckets) and self._have[self._next] == self._need[self._next]:
            self._launch(self._next)
            self._next += 1

    def _launch(self, k):
        dev = self.avg.device
        if self._side is None:
            self._side = torch.cuda.Stream(device=dev)
        ready = self._ready[k]
        ready.record(torch.cuda.current_stream(dev))   # gradients of this bucket are final at this point of backward
        self._side.wait_event(ready)
        # The gradients stay alive until the next zero_grad(), which comes after average_gradients()
        # has made the current stream wait for this kernel, so no record_stream is needed.
        with torch.cuda.stream(self._side):
            self.avg._run_params(self._bucket_params[k], self.variant, max_blocks=self.max_blocks)
        self._launched += 1

    def average_gradients(self):
        """Call where the reference calls `average_gradients()` (mnist/main.py:94-95): fires whatever
        did not fire from the hooks (parameters that got no gradient this step raise, as in the
        reference), then makes the current stream wait for the side stream."""
        if not self.enabled:
            self.avg._run_params(self.params, self.variant)
            return 1
        if self._next < len(self.buckets):
            for ps in self.params:
                for p in ps:
                    if p.grad is None:
                        raise AttributeError("parameter without .grad: 'NoneType' object has no attribute 'data'")
            while self._next < len(self.buckets):      # e.g. gradients set by hand, hooks never ran
                self._launch(self._next)
                self._next += 1
        if self._side is not None:
            self._done.record(self._side)
            torch.cuda.current_stream(self.avg.device).wait_event(self._done)
        launched = self._launched
        self._reset()
        return launched

    def remove(self):
        for h in self._handles:
            h.remove()
        self._handles = []
