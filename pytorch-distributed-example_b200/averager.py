"""Host side of the hot path: owns the C context, exchanges the IPC handles once, and turns a
parameter list into the pointer table the kernel reads (mirrors `Trainer.average_gradients`,
`/root/reference/mnist/main.py:122-127`)."""
import ctypes
import threading

import torch

from . import _lib
from ._lib import GradAvgError, check, load_library

AUTO, ONESHOT, TWOSHOT = _lib.VARIANT_AUTO, _lib.VARIANT_ONESHOT, _lib.VARIANT_TWOSHOT
ONESHOT_PUSH, TWOSHOT_PUSH, TWOSHOT_PIPE = _lib.VARIANT_ONESHOT_PUSH, _lib.VARIANT_TWOSHOT_PUSH, _lib.VARIANT_TWOSHOT_PIPE
HOSTED_DIRECT = _lib.VARIANT_HOSTED_DIRECT
VIEW, NVLS = _lib.VARIANT_VIEW, _lib.VARIANT_NVLS
_VIEW_VARIANTS = {"auto": AUTO, None: AUTO, AUTO: AUTO, "view": VIEW, VIEW: VIEW, "nvls": NVLS, NVLS: NVLS}
_VARIANTS = {"auto": AUTO, "oneshot": ONESHOT, "1shot": ONESHOT, "twoshot": TWOSHOT, "2shot": TWOSHOT,
             "oneshot_push": ONESHOT_PUSH, "twoshot_push": TWOSHOT_PUSH, "twoshot_pipe": TWOSHOT_PIPE, TWOSHOT_PIPE: TWOSHOT_PIPE,
             "hosted_direct": _lib.VARIANT_HOSTED_DIRECT, _lib.VARIANT_HOSTED_DIRECT: _lib.VARIANT_HOSTED_DIRECT,
             AUTO: AUTO, ONESHOT: ONESHOT, TWOSHOT: TWOSHOT, ONESHOT_PUSH: ONESHOT_PUSH, TWOSHOT_PUSH: TWOSHOT_PUSH,
             None: AUTO}


def _tiles(numel):
    return (numel * 4 + _lib.TILE_BYTES - 1) // _lib.TILE_BYTES


def bucket_bytes(numels):
    """Bytes a gradient set occupies in the bucket (each tensor padded to a 2 KiB tile)."""
    return sum(_tiles(n) for n in numels) * _lib.TILE_BYTES


def _dense(t):
    """True if the tensor's storage is one dense run of numel elements (default, channels_last or
    channels_last_3d layout).  The average is elementwise, so any such layout can be treated as a
    flat array as long as every rank uses the same one -- which data-parallel replicas do."""
    if t.is_contiguous():
        return True
    d = t.dim()
    if d == 4:
        return t.is_contiguous(memory_format=torch.channels_last)
    if d == 5:
        return t.is_contiguous(memory_format=torch.channels_last_3d)
    return False


def _require_cuda():
    if not torch.cuda.is_available():
        raise GradAvgError(_lib.GA_ESTATE, "no CUDA device: this path has no CPU fallback")


class GradAverager:
    """One per process (one rank per GPU), or one hosting all ranks on a single GPU.

    world_size/rank default to the initialised `torch.distributed` default group -- the one the
    reference script creates at `mnist/main.py:199-203` -- which is also what carries the 64-byte
    CUDA-IPC handles, once.  `local_world=W` hosts W ranks in this process on one device instead
    (no process group needed).
    """

    def __init__(self, world_size=None, rank=None, device=None, capacity_bytes=None, group=None,
                 local_world=None, variant=None):
        self._lib = load_library()
        self._ctx = None
        self._lock = threading.Lock()          # calls on one context are serialised (header contract)
        self._group = group
        self._variant = _VARIANTS[variant]
        self._capacity = capacity_bytes
        self._local = int(local_world) if local_world else 0
        if self._local:
            self.world_size, self.rank = self._local, 0
        else:
            import torch.distributed as dist
            self.world_size = int(world_size if world_size is not None else dist.get_world_size(group))
            self.rank = int(rank if rank is not None else dist.get_rank(group))
        if not 1 <= self.world_size <= _lib.MAX_WORLD:
            raise GradAvgError(_lib.GA_EINVAL, "world_size %d outside 1..%d" % (self.world_size, _lib.MAX_WORLD))
        self._device = device
        self._dev_obj = None
        self._launch_base = 0
        self._ptr_cache = {}
        self._plans = {}
        self._adopted = None                   # (param ids, gradient addresses, bytes) once adopt() has run
        if capacity_bytes:
            self._create(capacity_bytes)

    # -- context lifetime -------------------------------------------------------------------
    @property
    def n_local(self):
        return self._local or 1

    @property
    def device(self):
        d = self._dev_obj
        if d is None:
            if self._device is None:
                _require_cuda()
                self._device = torch.cuda.current_device()
            d = torch.device("cuda", self._device) if isinstance(self._device, int) else torch.device(self._device)
            self._dev_obj = d
        return d

    def _create(self, capacity_bytes):
        _require_cuda()
        dev = self.device.index if self.device.index is not None else torch.cuda.current_device()
        ctx = ctypes.c_void_p()
        if self._local:
            check(self._lib.ga_create_local(self.world_size, dev, int(capacity_bytes), ctypes.byref(ctx)))
        else:
            check(self._lib.ga_create(self.rank, self.world_size, dev, int(capacity_bytes), ctypes.byref(ctx)))
        self._ctx = ctx
        self._capacity = int(self._lib.ga_capacity(ctx))
        if not self._local:
            self._exchange_handles()

    def _exchange_handles(self):
        """all_gather of the 64-byte cudaIpcMemHandle over the script's own process group."""
        import torch.distributed as dist
        mine = ctypes.create_string_buffer(_lib.IPC_HANDLE_BYTES)
        check(self._lib.ga_export_handle(self._ctx, mine))
        if self.world_size == 1:
            check(self._lib.ga_open_peers(self._ctx, mine.raw))
            return
        backend = dist.get_backend(self._group)
        dev = self.device if "nccl" in str(backend) else torch.device("cpu")
        t = torch.frombuffer(bytearray(mine.raw), dtype=torch.uint8).to(dev)
        out = [torch.empty_like(t) for _ in range(self.world_size)]
        dist.all_gather(out, t, group=self._group)
        blob = b"".join(bytes(o.cpu().numpy().tobytes()) for o in out)
        check(self._lib.ga_open_peers(self._ctx, blob))
        dist.barrier(group=self._group)      # nobody signals into a pad that is not mapped yet

    def _ensure(self, need_bytes):
        if self._ctx is not None and need_bytes <= self._capacity:
            return
        if self._ctx is not None:
            if self._adopted is not None or int(self._lib.ga_view_bytes(self._ctx)):
                raise GradAvgError(_lib.GA_ECAPACITY, "this call needs a %d-byte bucket but the context (%d bytes) holds a "
                                   "gradient view region and cannot be re-created: size the averager with capacity_bytes "
                                   "up front" % (need_bytes, self._capacity))
            # Every rank sees the same sizes, so every rank takes this branch in the same call.
            self.close()
        self._create(max(need_bytes, 32 << 20))     # room for two one-shot staging halves of small sets

    def _release_views(self):
        """Adopted gradients point into the view region: before it is freed, every gradient that still lives
        there moves to memory of its own (same values), so no tensor is left pointing at freed memory."""
        ad, self._adopted = self._adopted, None
        if ad is None:
            return
        for p, a in zip(ad[3], ad[1]):
            g = p.grad
            if g is not None and g.data_ptr() == a:
                p.grad = g.clone()

    def close(self):
        if self._ctx is not None:
            self._release_views()
            if torch.cuda.is_available():
                torch.cuda.synchronize()
            if not self._local and self.world_size > 1:
                import torch.distributed as dist
                if dist.is_initialized():
                    dist.barrier(group=self._group)   # peers may still be reading my bucket
                    self._lib.ga_close_peers(self._ctx)
                    dist.barrier(group=self._group)   # nobody frees memory a peer still has mapped
            self._launch_base += int(self._lib.ga_launch_count(self._ctx))
            self._lib.ga_destroy(self._ctx)
            self._ctx = None
            self._ptr_cache.clear()
            self._adopted = None

    def __del__(self):
        try:
            if self._ctx is not None:
                self._release_views()
                self._lib.ga_destroy(self._ctx)
                self._ctx = None
        except Exception:
            pass

    # -- the hot path -----------------------------------------------------------------------
    @staticmethod
    def _grads_of(params):
        if isinstance(params, torch.nn.Module):
            params = params.parameters()
        grads = []
        for p in params:
            g = p.grad
            if g is None:
                # the reference would raise AttributeError on `None.data` (mnist/main.py:126)
                raise AttributeError("parameter without .grad: 'NoneType' object has no attribute 'data'")
            grads.append(g.data)
        return grads

    def _table(self, per_rank_tensors, dtype):
        n = len(per_rank_tensors[0])
        key = (len(per_rank_tensors), n)
        arrs = self._ptr_cache.get(key)
        if arrs is None:
            arrs = ((ctypes.c_void_p * (n * len(per_rank_tensors)))(), (ctypes.c_int64 * n)())
            self._ptr_cache[key] = arrs
        ptrs, numel = arrs
        dev = self.device
        k = 0
        for tensors in per_rank_tensors:
            if len(tensors) != n:
                raise GradAvgError(_lib.GA_EINVAL, "hosted ranks passed different tensor counts")
            for i, t in enumerate(tensors):
                if t.dtype != dtype or not t.is_cuda or t.device != dev:
                    raise GradAvgError(_lib.GA_EINVAL, "tensor %d: need %s on %s, got %s on %s (no CPU fallback)"
                                       % (i, dtype, dev, t.dtype, t.device))
                if not _dense(t):
                    raise GradAvgError(_lib.GA_EINVAL, "tensor %d is not dense (strides %s)" % (i, tuple(t.stride())))
                ptrs[k] = t.data_ptr()
                k += 1
        for i, t in enumerate(per_rank_tensors[0]):
            numel[i] = t.numel()
        return ptrs, numel, n

    def _run_tensors(self, per_rank_tensors, op, dtype, variant):
        if len(per_rank_tensors) != self.n_local:
            raise GradAvgError(_lib.GA_EINVAL, "got %d tensor lists for %d hosted rank(s)"
                               % (len(per_rank_tensors), self.n_local))
        with self._lock:
            numels = [t.numel() for t in per_rank_tensors[0]]
            self._ensure(bucket_bytes(numels))
            ptrs, numel, n = self._table(per_rank_tensors, dtype)
            stream = torch.cuda.current_stream(self.device).cuda_stream
            v = self._variant if variant is None else _VARIANTS[variant]
            check(self._lib.ga_allreduce_tensors(self._ctx, ptrs, numel, n, op, ctypes.c_void_p(stream), v))

    def _run_params(self, per_rank_params, variant, max_blocks=0):
        """Hot path of the training loop.  The element counts of a parameter list never change, so
        they are laid out once per list (keyed by the identity of its parameters); every call then
        only re-reads the gradient pointers and checks dtype / device / contiguity."""
        flat = [p for ps in per_rank_params for p in ps]
        if self._adopted is not None and self._try_view(flat, variant):
            return
        if variant in (VIEW, NVLS, "view", "nvls"):
            raise GradAvgError(_lib.GA_ESTATE, "variant %r needs adopted gradients that are still in their views "
                               "(adopt(); zero_grad(set_to_none=False))" % (variant,))
        key = tuple(map(id, flat))
        plan = self._plans.get(key)
        if plan is None:
            if len(self._plans) > 16:
                self._plans.clear()
            n = len(per_rank_params[0])
            if any(len(ps) != n for ps in per_rank_params):
                raise GradAvgError(_lib.GA_EINVAL, "hosted ranks passed different parameter counts")
            def _n(p):
                if p.grad is None:      # the reference would raise AttributeError on `None.data` (mnist/main.py:126)
                    raise AttributeError("parameter without .grad: 'NoneType' object has no attribute 'data'")
                return p.grad.numel()
            numels = [_n(p) for p in per_rank_params[0]]
            for ps in per_rank_params[1:]:
                if [_n(p) for p in ps] != numels:
                    raise GradAvgError(_lib.GA_EINVAL, "hosted ranks passed different parameter sizes")
            # `flat` is kept in the plan: while the parameters are alive their ids cannot be reused
            plan = (n, (ctypes.c_void_p * len(flat))(), (ctypes.c_int64 * n)(*numels), bucket_bytes(numels), flat, numels)
            self._plans[key] = plan
        n, ptrs, numel, need, _keepalive, numels = plan
        with self._lock:
            self._ensure(need)
            dev_index = self.device.index
            f32 = torch.float32
            k = 0
            for p in flat:
                g = p.grad
                if g is None:
                    # the reference would raise AttributeError on `None.data` (mnist/main.py:126)
                    raise AttributeError("parameter without .grad: 'NoneType' object has no attribute 'data'")
                if g.numel() != numels[k % n]:
                    del self._plans[key]
                    raise GradAvgError(_lib.GA_EINVAL, "gradient %d changed size since the first call with this "
                                       "parameter list (%d -> %d elements)" % (k % n, numels[k % n], g.numel()))
                if g.dtype is not f32 or g.get_device() != dev_index or not _dense(g):
                    raise GradAvgError(_lib.GA_EINVAL, "gradient %d: need dense float32 on cuda:%s, got %s on %s, strides %s "
                                       "(no CPU fallback)" % (k % n, dev_index, g.dtype, g.device, tuple(g.stride())))
                ptrs[k] = g.data_ptr()
                k += 1
            stream = torch.cuda.current_stream(self.device).cuda_stream
            v = self._variant if variant is None else _VARIANTS[variant]
            if max_blocks:
                check(self._lib.ga_allreduce_tensors_ex(self._ctx, ptrs, numel, n, _lib.OP_F32_AVG, ctypes.c_void_p(stream), v,
                                                        int(max_blocks)))
            else:
                check(self._lib.ga_average(self._ctx, ptrs, numel, n, ctypes.c_void_p(stream), v))

    def average_gradients(self, params, variant=None):
        """`Trainer.average_gradients` (mnist/main.py:122-127) for one rank: every `p.grad` is
        replaced IN PLACE by the rank-order sum over all ranks divided by float(world_size).
        Stream-ordered on torch's current stream; returns after the launch."""
        if isinstance(params, torch.nn.Module):
            params = params.parameters()
        if self.n_local != 1:
            raise GradAvgError(_lib.GA_EINVAL, "this averager hosts %d ranks: use average_gradients_local" % self.n_local)
        self._run_params([list(params)], variant)

    def average_gradients_local(self, per_rank_params, variant=None):
        """Same for a local world: one parameter list per hosted rank, rank order."""
        if len(per_rank_params) != self.n_local:
            raise GradAvgError(_lib.GA_EINVAL, "need %d parameter lists" % self.n_local)
        self._run_params([list(ps.parameters() if isinstance(ps, torch.nn.Module) else ps) for ps in per_rank_params],
                         variant)

    def all_reduce_(self, tensors, average=False, variant=None):
        """In-place SUM (or average) of a list of CUDA tensors (fp32 or int32), one list per hosted
        rank -- the op under `distributed.all_reduce(t, op=SUM)` (mnist/main.py:126, toy/main.py:20)."""
        if torch.is_tensor(tensors):
            tensors = [[tensors]]
        elif tensors and torch.is_tensor(tensors[0]):
            tensors = [list(tensors)]
        dtype = tensors[0][0].dtype
        if dtype == torch.float32:
            op = _lib.OP_F32_AVG if average else _lib.OP_F32_SUM
        elif dtype == torch.int32 and not average:
            op = _lib.OP_I32_SUM
        else:
            raise GradAvgError(_lib.GA_EINVAL, "unsupported dtype/op: %s average=%s" % (dtype, average))
        self._run_tensors(tensors, op, dtype, variant)

    def average_host(self, per_rank_host_tensors, variant=None):
        """Host-buffer form (`ga_average_host`): pinned CPU fp32 tensors, H2D -> reduce -> D2H,
        stream-ordered; the caller synchronises the stream before reading the tensors."""
        if per_rank_host_tensors and torch.is_tensor(per_rank_host_tensors[0]):
            per_rank_host_tensors = [list(per_rank_host_tensors)]
        if len(per_rank_host_tensors) != self.n_local:
            raise GradAvgError(_lib.GA_EINVAL, "got %d tensor lists for %d hosted rank(s)"
                               % (len(per_rank_host_tensors), self.n_local))
        with self._lock:
            numels = [t.numel() for t in per_rank_host_tensors[0]]
            self._ensure(bucket_bytes(numels))
            n = len(numels)
            ptrs = (ctypes.c_void_p * (n * len(per_rank_host_tensors)))()
            k = 0
            for ts in per_rank_host_tensors:
                if [t.numel() for t in ts] != numels:
                    raise GradAvgError(_lib.GA_EINVAL, "hosted ranks passed different tensor sizes")
                for t in ts:
                    if t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous():
                        raise GradAvgError(_lib.GA_EINVAL, "average_host needs contiguous CPU fp32 tensors")
                    ptrs[k] = t.data_ptr()
                    k += 1
            numel = (ctypes.c_int64 * n)(*numels)
            stream = torch.cuda.current_stream(self.device).cuda_stream
            v = self._variant if variant is None else _VARIANTS[variant]
            check(self._lib.ga_average_host(self._ctx, ptrs, numel, n, ctypes.c_void_p(stream), v))

    # -- gradients as views of the peer-mapped region ---------------------------
    def _create_view(self, nbytes, multicast):
        """Collective: allocates the view region and maps every peer's (mnist/main.py:199-203's group
        carries the handles, like the bucket's).  If the multicast kind cannot be set up on this box (the
        descriptors cannot be passed between the rank processes, the driver refuses the multicast object)
        every rank sees the same error, and the region is made again without it."""
        if multicast:
            try:
                return self._create_view_kind(nbytes, True)
            except GradAvgError:
                check(self._lib.ga_view_destroy(self._ctx))
        return self._create_view_kind(nbytes, False)

    def _create_view_kind(self, nbytes, multicast):
        flags = _lib.VIEW_MULTICAST if multicast else 0
        check(self._lib.ga_view_create(self._ctx, int(nbytes), flags))
        if self._local or self.world_size == 1:
            if not self._local:
                mine = ctypes.create_string_buffer(_lib.VIEW_HANDLE_BYTES)
                check(self._lib.ga_view_export(self._ctx, mine))
                check(self._lib.ga_view_open(self._ctx, mine.raw))
            return
        import torch.distributed as dist
        mine = ctypes.create_string_buffer(_lib.VIEW_HANDLE_BYTES)
        check(self._lib.ga_view_export(self._ctx, mine))
        backend = dist.get_backend(self._group)
        dev = self.device if "nccl" in str(backend) else torch.device("cpu")
        t = torch.frombuffer(bytearray(mine.raw), dtype=torch.uint8).to(dev)
        out = [torch.empty_like(t) for _ in range(self.world_size)]
        dist.all_gather(out, t, group=self._group)
        blob = b"".join(bytes(o.cpu().numpy().tobytes()) for o in out)
        check(self._lib.ga_view_open(self._ctx, blob))
        dist.barrier(group=self._group)      # every rank has fetched the descriptors and joined the multicast object
        check(self._lib.ga_view_bind(self._ctx))
        dist.barrier(group=self._group)      # every rank's memory is bound before anybody issues multimem.*

    def create_view(self, nbytes, multicast=True):
        """Flat form: a `nbytes` region per hosted rank for `view_tensor` / `allreduce_view`."""
        with self._lock:
            self._ensure(int(nbytes))
            if not int(self._lib.ga_view_bytes(self._ctx)):
                self._create_view(int(nbytes), multicast)
            elif int(self._lib.ga_view_bytes(self._ctx)) < int(nbytes):
                raise GradAvgError(_lib.GA_ECAPACITY, "the view region is smaller than %d bytes" % nbytes)

    @property
    def has_multicast(self):
        return bool(self._ctx is not None and self._lib.ga_view_has_multicast(self._ctx))

    def view_tensor(self, nbytes=None, local_rank=0, dtype=torch.float32, offset=0, shape=None):
        """Tensor aliasing hosted rank `local_rank`'s view region."""
        base = self._lib.ga_view_ptr(self._ctx, local_rank)
        if not base:
            raise GradAvgError(_lib.GA_ESTATE, "no view region (create_view / adopt)")
        nbytes = int(self._lib.ga_view_bytes(self._ctx)) - offset if nbytes is None else nbytes
        t = self._view(int(base) + int(offset), int(nbytes), dtype)
        return t if shape is None else t.view(shape)

    def allreduce_view(self, nbytes, op=_lib.OP_F32_AVG, variant=None):
        """`ga_average_view`: in place on the first nbytes of every rank's view region."""
        stream = torch.cuda.current_stream(self.device).cuda_stream
        with self._lock:
            check(self._lib.ga_average_view(self._ctx, int(nbytes), op, ctypes.c_void_p(stream), _VIEW_VARIANTS[variant]))

    def adopt(self, modules, multicast=True, optimizer=None):
        """Makes every `p.grad` a view into the peer-mapped region, so that `average_gradients` works
        in place: no pack, no unpack (the steps either side only touch p.grad: mnist/main.py:91-92,96).

        modules: the replica (or its parameters) of this rank, or one per hosted rank.  Existing
        gradients are copied into their views.  The views survive only while nothing replaces
        `p.grad`: use `optimizer.zero_grad(set_to_none=False)` -- pass `optimizer=` and its
        `zero_grad` gets that default.  If a gradient is found outside its view at averaging time
        (`zero_grad()` dropped it and autograd re-allocated it), that call silently takes the drop-in
        pack/unpack path instead, so correctness never depends on the views.  Collective."""
        if isinstance(modules, torch.nn.Module) or not isinstance(modules, (list, tuple)):
            modules = [modules]
        if len(modules) != self.n_local and self.n_local == 1:
            modules = [modules]                   # a plain list of parameters
        per_rank = [list(m.parameters()) if isinstance(m, torch.nn.Module) else list(m) for m in modules]
        if len(per_rank) != self.n_local:
            raise GradAvgError(_lib.GA_EINVAL, "need %d parameter lists" % self.n_local)
        numels = [p.numel() for p in per_rank[0]]
        if any([p.numel() for p in ps] != numels for ps in per_rank[1:]):
            raise GradAvgError(_lib.GA_EINVAL, "hosted ranks passed different parameter sizes")
        need = bucket_bytes(numels)
        with self._lock:
            if self._adopted is not None:
                raise GradAvgError(_lib.GA_ESTATE, "this averager has already adopted a parameter list")
            self._ensure(max(need, 32 << 20))
            self._create_view(need, multicast)
            addrs = []
            for l, ps in enumerate(per_rank):
                off = 0
                for p, n in zip(ps, numels):
                    if p.dtype != torch.float32 or not p.is_cuda or p.device != self.device:
                        raise GradAvgError(_lib.GA_EINVAL, "adopt needs float32 parameters on %s" % (self.device,))
                    g = self.view_tensor(n * 4, l, torch.float32, off, tuple(p.shape))
                    if p.grad is not None:
                        g.copy_(p.grad)
                    else:
                        g.zero_()
                    p.grad = g
                    addrs.append(g.data_ptr())
                    off += _tiles(n) * _lib.TILE_BYTES
            torch.cuda.current_stream(self.device).synchronize()
            flat = [p for ps in per_rank for p in ps]
            self._adopted = (tuple(map(id, flat)), addrs, need, flat)
        if optimizer is not None:
            import functools
            optimizer.zero_grad = functools.partial(optimizer.zero_grad, set_to_none=False)
        return self

    def _try_view(self, flat, variant):
        """In-place averaging if `flat` is the adopted list and every gradient still sits in its view."""
        ad = self._adopted
        if ad is None or variant not in _VIEW_VARIANTS:
            return False
        ids, addrs, need, _keep = ad
        if len(flat) != len(ids):
            return False
        for p, i, a in zip(flat, ids, addrs):
            g = p.grad
            if id(p) != i or g is None or g.data_ptr() != a:
                return False
        stream = torch.cuda.current_stream(self.device).cuda_stream
        with self._lock:
            check(self._lib.ga_average_view(self._ctx, need, _lib.OP_F32_AVG, ctypes.c_void_p(stream), _VIEW_VARIANTS[variant]))
        return True

    # -- parameter broadcast -----------------------------------------------------
    def broadcast_(self, tensors, root=0):
        """Replicates rank `root`'s tensors into every other rank's (in place).  The reference never
        makes its replicas start equal (no seed, no broadcast: mnist/main.py:185).  One list per
        hosted rank; any dense 4-byte-multiple dtype (moved as 32-bit words)."""
        if torch.is_tensor(tensors):
            tensors = [[tensors]]
        elif tensors and torch.is_tensor(tensors[0]):
            tensors = [list(tensors)]
        if len(tensors) != self.n_local:
            raise GradAvgError(_lib.GA_EINVAL, "got %d tensor lists for %d hosted rank(s)" % (len(tensors), self.n_local))
        words = []
        for ts in tensors:
            row = []
            for t in ts:
                if not t.is_cuda or t.device != self.device or not _dense(t) or (t.numel() * t.element_size()) % 4:
                    raise GradAvgError(_lib.GA_EINVAL, "broadcast needs dense CUDA tensors of a 4-byte-multiple size on %s" % (self.device,))
                row.append((t.data_ptr(), t.numel() * t.element_size() // 4))
            words.append(row)
        n = len(words[0])
        if any([w[1] for w in row] != [w[1] for w in words[0]] for row in words[1:]):
            raise GradAvgError(_lib.GA_EINVAL, "hosted ranks passed different tensor sizes")
        with self._lock:
            self._ensure(bucket_bytes([w[1] for w in words[0]]))
            ptrs = (ctypes.c_void_p * (n * len(words)))(*[w[0] for row in words for w in row])
            numel = (ctypes.c_int64 * n)(*[w[1] for w in words[0]])
            stream = torch.cuda.current_stream(self.device).cuda_stream
            check(self._lib.ga_broadcast_tensors(self._ctx, ptrs, numel, n, int(root), ctypes.c_void_p(stream)))

    def broadcast_parameters(self, modules, root=0):
        """Parameters and buffers of the replica(s) become rank `root`'s."""
        if isinstance(modules, torch.nn.Module):
            modules = [modules]
        lists = [[t.data for t in list(m.parameters()) + list(m.buffers()) if t.numel()] for m in modules]
        self.broadcast_(lists, root=root)

    # -- bucket-direct form (size sweep, NCCL comparison) -------------------------------------
    def reserve(self, nbytes):
        with self._lock:
            self._ensure(int(nbytes))

    def _view(self, ptr, nbytes, dtype):
        n = nbytes // 4

        class _Holder:
            pass
        h = _Holder()
        h.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4" if dtype == torch.float32 else "<i4",
                                      "data": (int(ptr), False), "version": 2}
        return torch.as_tensor(h, device=self.device)

    def send_view(self, nbytes=None, local_rank=0, dtype=torch.float32):
        """Tensor aliasing this rank's bucket (what the peers read)."""
        return self._view(self._lib.ga_send_ptr(self._ctx, local_rank), nbytes or self._capacity, dtype)

    def recv_view(self, nbytes=None, local_rank=0, dtype=torch.float32):
        """Tensor aliasing this rank's result region of the bucket-direct calls."""
        return self._view(self._lib.ga_recv_ptr(self._ctx, local_rank), nbytes or self._capacity, dtype)

    def bucket_acquire(self):
        stream = torch.cuda.current_stream(self.device).cuda_stream
        with self._lock:
            check(self._lib.ga_bucket_acquire(self._ctx, ctypes.c_void_p(stream)))

    def allreduce_bucket(self, nbytes, op=_lib.OP_F32_AVG, variant=None):
        stream = torch.cuda.current_stream(self.device).cuda_stream
        v = self._variant if variant is None else _VARIANTS[variant]
        with self._lock:
            check(self._lib.ga_allreduce_bucket(self._ctx, int(nbytes), op, ctypes.c_void_p(stream), v))

    # -- introspection ----------------------------------------------------------------------
    def set_option(self, option, value):
        check(self._lib.ga_set_option(self._ctx, option, int(value)))

    def read_trace(self):
        """Phase timestamps of the last launch as an int64 array [hosted ranks, blocks, slots]
        (needs set_option(OPT_TRACE, 1)); see ga_read_trace in include/gradavg.h."""
        import numpy as np
        buf = (ctypes.c_uint64 * (1024 * self.n_local * 16))()
        b, k = ctypes.c_int(), ctypes.c_int()
        check(self._lib.ga_read_trace(self._ctx, buf, len(buf), ctypes.byref(b), ctypes.byref(k)))
        a = np.frombuffer(buf, dtype=np.uint64)[: self.n_local * b.value * k.value].astype(np.int64)
        return a.reshape(self.n_local, b.value, k.value)

    def pick_variant(self, nbytes):
        return self._lib.ga_pick_variant(self._ctx, int(nbytes))

    @property
    def launch_count(self):
        """Kernels launched by this averager, across re-creations of its context."""
        return self._launch_base + (int(self._lib.ga_launch_count(self._ctx)) if self._ctx is not None else 0)

    @property
    def last_grid(self):
        b, t = ctypes.c_int(), ctypes.c_int()
        check(self._lib.ga_last_grid(self._ctx, ctypes.byref(b), ctypes.byref(t)))
        return b.value, t.value

    @property
    def capacity(self):
        return self._capacity or 0

    def check_status(self):
        """Raises GradAvgError(GA_ETIMEOUT) if a kernel gave up on a peer. Synchronises first."""
        torch.cuda.synchronize(self.device)
        check(self._lib.ga_status(self._ctx))


_default = None


def default_averager():
    """The process-wide averager bound to the default process group, created on first use and
    sized from the first parameter list it sees."""
    global _default
    if _default is None:
        _default = GradAverager()
    return _default


def reset_default_averager():
    global _default
    if _default is not None:
        _default.close()
    _default = None


def average_gradients(params, variant=None):
    """Drop-in body for `Trainer.average_gradients` (mnist/main.py:122-127): accepts the module
    or its `parameters()`."""
    default_averager().average_gradients(params, variant=variant)
