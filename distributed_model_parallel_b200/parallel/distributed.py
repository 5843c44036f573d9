"""DistributedDataParallel: one process per GPU, bucketed gradient Reducer, peer-memory all-reduce.

Same call shape as ``torch.nn.parallel.DistributedDataParallel`` (ctor signature at
torch/nn/parallel/distributed.py:655-675; semantics inventoried in SURVEY.md section 2.5), which the reference
only discusses in its README (/root/reference/Readme.md:10-15,145-157) and never calls.  Differences that
matter on B200:

* the per-bucket collective is our own NVLink peer-memory kernel (one-shot / two-shot by size) launched from
  the autograd hooks on a high-priority side stream -- ``comm_backend="peer"`` (default on CUDA);
  ``"torch"`` keeps ``torch.distributed.all_reduce`` (gloo on CPU, NCCL on GPU) as the baseline path;
* ``fused_sgd=dict(lr=..., momentum=..., weight_decay=...)`` moves the parameters into flat fp32 master
  arenas and fuses 1/W scaling, bf16 cast and the SGD-momentum update into that same kernel; ``self.optimizer``
  is then a facade whose ``step()`` only publishes learning-rate changes.

``state_dict()`` keys carry the ``module.`` prefix exactly like torch's wrappers, so checkpoints written by the
reference's DP script (data_parallel.py:147-151) load unchanged.
"""
from __future__ import annotations

import contextlib
import itertools
from typing import Any, Callable, List, Optional

import torch
import torch.distributed as dist
import torch.nn as nn

from .comm import broadcast_coalesced_group
from .reducer import DEFAULT_BUCKET_CAP_MB, DEFAULT_FIRST_BUCKET_BYTES, PeerBackend, Reducer, TorchDistBackend

BROADCAST_BUCKET_BYTES = 250 * 1024 * 1024


def _find_tensors(obj) -> List[torch.Tensor]:
    if isinstance(obj, torch.Tensor):
        return [obj]
    if isinstance(obj, (list, tuple)):
        return list(itertools.chain.from_iterable(map(_find_tensors, obj)))
    if isinstance(obj, dict):
        return list(itertools.chain.from_iterable(map(_find_tensors, obj.values())))
    return []


def _reachable_params(outputs) -> set:
    """ids of the leaf tensors whose AccumulateGrad nodes are reachable from ``outputs``."""
    seen, found, stack = set(), set(), []
    for t in _find_tensors(outputs):
        if t.grad_fn is not None:
            stack.append(t.grad_fn)
    while stack:
        fn = stack.pop()
        if fn in seen:
            continue
        seen.add(fn)
        var = getattr(fn, "variable", None)
        if var is not None:
            found.add(id(var))
        for nxt, _ in fn.next_functions:
            if nxt is not None:
                stack.append(nxt)
        # custom Functions that deliver gradients directly list their parameters explicitly
        for p in getattr(fn, "_dmp_params", ()):
            found.add(id(p))
    return found


class FusedSGD:
    """Optimizer facade for ``DistributedDataParallel(fused_sgd=...)``.

    The update itself happens inside the bucket all-reduce kernel during backward, so ``step()`` has nothing
    left to do except push hyper-parameter changes (LR schedulers mutate ``param_groups[0]['lr']``) to the
    device-resident hyper-parameter block the kernel reads.  ``state_dict`` exposes the flat momentum buffers
    per parameter in the usual ``torch.optim.SGD`` shape."""

    def __init__(self, reducer: Reducer, cfg: dict):
        self.reducer = reducer
        self.defaults = dict(lr=cfg.get("lr", 0.1), momentum=cfg.get("momentum", 0.0), weight_decay=cfg.get("weight_decay", 0.0),
                             nesterov=bool(cfg.get("nesterov", False)), dampening=0.0)
        self.param_groups = [dict(self.defaults, params=list(reducer.params), initial_lr=self.defaults["lr"])]
        self._pushed = None
        self._host = torch.empty(4, dtype=torch.float32).pin_memory() if torch.cuda.is_available() else torch.empty(4)
        self.push_hyperparameters()

    def push_hyperparameters(self):
        g = self.param_groups[0]
        cur = (float(g["lr"]), float(g["momentum"]), float(g["weight_decay"]))
        if cur != self._pushed:
            self._host[0], self._host[1], self._host[2], self._host[3] = cur[0], cur[1], cur[2], 1.0
            self.reducer.hp.copy_(self._host, non_blocking=True)
            self._pushed = cur

    def step(self, closure=None):
        loss = closure() if closure is not None else None
        self.push_hyperparameters()
        return loss

    def zero_grad(self, set_to_none: bool = False):
        pass  # the fused kernel leaves the gradient buckets zeroed

    def state_dict(self):
        state = {}
        for bk in self.reducer.buckets:
            for i, p, off in zip(bk.param_indices, bk.params, bk.offsets):
                state[i] = {"momentum_buffer": bk.fused.momentum[off: off + p.numel()].view_as(p).clone()}
        groups = [{k: v for k, v in self.param_groups[0].items() if k != "params"}]
        groups[0]["params"] = list(range(len(self.reducer.params)))
        return {"state": state, "param_groups": groups}

    def load_state_dict(self, sd):
        for bk in self.reducer.buckets:
            for i, p, off in zip(bk.param_indices, bk.params, bk.offsets):
                st = sd["state"].get(i) or sd["state"].get(str(i))
                if st is not None and st.get("momentum_buffer") is not None:
                    bk.fused.momentum[off: off + p.numel()].view_as(p).copy_(st["momentum_buffer"])
        for k, v in sd["param_groups"][0].items():
            if k != "params":
                self.param_groups[0][k] = v
        self.push_hyperparameters()


class DistributedDataParallel(nn.Module):
    def __init__(self, module: nn.Module, device_ids=None, output_device=None, dim: int = 0, broadcast_buffers: bool = True,
                 init_sync: bool = True, process_group=None, bucket_cap_mb: Optional[float] = None,
                 find_unused_parameters: bool = False, check_reduction: bool = False, gradient_as_bucket_view: bool = False,
                 static_graph: bool = False, *, comm_backend: Optional[str] = None, fused_sgd: Optional[dict] = None,
                 first_bucket_bytes: int = DEFAULT_FIRST_BUCKET_BYTES):
        super().__init__()
        params_all = [p for p in module.parameters()]
        if not any(p.requires_grad for p in params_all):
            raise RuntimeError("DistributedDataParallel is not needed when a module doesn't have any parameter that "
                               "requires a gradient.")
        if device_ids is not None and len(device_ids) > 1:
            raise ValueError("device_ids can only be None or contain a single element.")
        dev_types = {p.device.type for p in params_all}
        if len(dev_types) != 1:
            raise ValueError("DistributedDataParallel's input module must be on the same type of devices, but input "
                             "module parameters locate in %s." % dev_types)
        self.device_type = next(iter(dev_types))
        devices = {p.device for p in params_all}
        self.is_multi_device_module = len(devices) > 1
        if self.is_multi_device_module and device_ids is not None:
            raise ValueError("DistributedDataParallel device_ids and output_device arguments only work with "
                             "single-device modules.")
        self.device_ids = None if device_ids is None or self.device_type == "cpu" else [torch.device("cuda", d).index
                                                                                         if not isinstance(d, torch.device)
                                                                                         else d.index for d in device_ids]
        self.output_device = output_device if output_device is not None else (self.device_ids[0] if self.device_ids else None)
        self.module = module
        self.dim = dim
        self.broadcast_buffers = broadcast_buffers
        self.find_unused_parameters = find_unused_parameters
        self.static_graph = static_graph
        self.gradient_as_bucket_view = gradient_as_bucket_view
        self.require_backward_grad_sync = True
        self.require_forward_param_sync = True
        self.process_group = process_group if process_group is not None else (dist.group.WORLD if dist.is_initialized() else None)
        self.world_size = dist.get_world_size(self.process_group) if self.process_group is not None else 1
        self.rank = dist.get_rank(self.process_group) if self.process_group is not None else 0
        self.bucket_bytes_cap = int((bucket_cap_mb if bucket_cap_mb is not None else DEFAULT_BUCKET_CAP_MB) * 1024 * 1024)
        self.parameters_to_ignore = set(getattr(module, "_ddp_params_and_buffers_to_ignore", []))

        # ---- parameter / buffer lists (torch: _build_params_for_reducer)
        names, params, seen = [], [], set()
        for mname, m in module.named_modules():
            for pname, p in m.named_parameters(recurse=False):
                full = "%s.%s" % (mname, pname) if mname else pname
                if not p.requires_grad or full in self.parameters_to_ignore or id(p) in seen:
                    continue
                seen.add(id(p))
                names.append(full)
                params.append(p)
        self._param_names, self._params = names, params
        self.modules_buffers = [b for n, b in module.named_buffers() if n not in self.parameters_to_ignore]

        # ---- init sync: verify shapes, then rank 0's params + buffers everywhere
        if init_sync and self.world_size > 1:
            self._verify_param_shape_across_processes()
            self._sync_module_states()

        # ---- reducer
        if comm_backend is None:
            comm_backend = "peer" if self.device_type == "cuda" else "torch"
        if find_unused_parameters and fused_sgd is not None:
            raise ValueError("find_unused_parameters=True cannot be combined with fused_sgd: the update runs inside the bucket "
                             "kernels during backward, before it is known which parameters no rank used (their weight decay "
                             "and momentum would still be applied); use comm_backend='peer' without fused_sgd")
        self._ctor = dict(comm_backend=comm_backend, fused_sgd=dict(fused_sgd) if fused_sgd is not None else None,
                          first_bucket_bytes=int(first_bucket_bytes))
        self._build_reducer()
        self._rebuild_checked = False
        self._join_active, self._join_iters = False, 0
        self.logger_data = {"backend": self.backend.name, "world_size": self.world_size,
                            "bucket_cap_bytes": self.bucket_bytes_cap}

    def _build_reducer(self):
        """Back end + Reducer (+ fused optimizer facade, peer-visible buffer arena) from the stored constructor arguments; also
        what ``__setstate__`` runs after unpickling (torch distributed.py:1306-1336 rebuilds its reducer the same way)."""
        comm_backend, fused_sgd = self._ctor["comm_backend"], self._ctor["fused_sgd"]
        params, names = self._params, self._param_names
        if comm_backend == "peer":
            if self.device_type != "cuda":
                raise ValueError("comm_backend='peer' needs CUDA parameters")
            self.backend = PeerBackend(self.process_group, params[0].device)
        elif comm_backend == "torch":
            if self.process_group is None:
                raise RuntimeError("comm_backend='torch' needs an initialised process group")
            self.backend = TorchDistBackend(self.process_group)
        else:
            raise ValueError("unknown comm_backend %r" % comm_backend)
        if fused_sgd is not None and not self.backend.supports_fused:
            raise ValueError("fused_sgd needs comm_backend='peer'")
        self.reducer = Reducer(params, self.backend, self.bucket_bytes_cap, self._ctor["first_bucket_bytes"],
                               self.find_unused_parameters, self.gradient_as_bucket_view, names, fused_sgd, self.static_graph)
        self.optimizer = FusedSGD(self.reducer, fused_sgd) if fused_sgd is not None else None
        self._buffer_arena = None
        if isinstance(self.backend, PeerBackend) and self.backend.symm is not None and self.modules_buffers and self.broadcast_buffers:
            self._build_buffer_arena()
        self._static_unused = None

    # ------------------------------------------------------------------------------------------ init helpers
    def _verify_param_shape_across_processes(self):
        shapes = [tuple(p.shape) for p in self._params]
        gathered = [None] * self.world_size
        dist.all_gather_object(gathered, shapes, group=self.process_group)
        for r, s in enumerate(gathered):
            if s != shapes:
                raise RuntimeError("DDP parameter shapes differ between rank %d and rank %d" % (self.rank, r))

    def _sync_module_states(self, src: int = 0):
        tensors = [p.detach() for p in self._params] + list(self.modules_buffers)
        broadcast_coalesced_group(self.process_group, tensors, BROADCAST_BUCKET_BYTES, src=src)

    def _build_buffer_arena(self):
        """Moves every buffer (BN running statistics, counters) into one peer-visible byte arena so that the per-forward
        rank-0 buffer broadcast (torch distributed.py:2214-2257) is a single peer-store kernel instead of NCCL calls."""
        offs, total = [], 0
        for b in self.modules_buffers:
            offs.append(total)
            total += (b.numel() * b.element_size() + 15) // 16 * 16
        arena = self.backend.symm.alloc(total)
        raw = arena.local.view(torch.uint8)
        for b, o in zip(self.modules_buffers, offs):
            nb = b.numel() * b.element_size()
            v = raw[o: o + nb].view(b.dtype).view(b.shape)
            v.copy_(b)
            b.data = v
        self._buffer_arena = arena
        self._buffer_bytes = total
        self._buffer_state = self.backend.symm.new_state()
        from .symm import ptr_array
        self._buffer_dst = ptr_array([p for r, p in enumerate(arena.peer_ptrs) if r != 0])

    def _sync_buffers(self):
        if self._buffer_arena is not None:
            from .. import _native
            dev = self._params[0].device
            _native.check(self.backend.lib.dmp_broadcast(self._buffer_arena.peer_ptrs[0], self._buffer_dst, self.world_size - 1, None,
                                                         None, self._buffer_bytes, self._buffer_state.ptrs, self.rank, self.world_size,
                                                         1 if self.rank == 0 else 0, 8, torch.cuda.current_stream(dev).cuda_stream),
                          "buffer broadcast")
            return
        if self.modules_buffers and self.world_size > 1:
            with torch.no_grad():
                broadcast_coalesced_group(self.process_group, self.modules_buffers, BROADCAST_BUCKET_BYTES, src=0)

    # ------------------------------------------------------------------------------------------ forward
    def forward(self, *inputs, **kwargs):
        with torch.autograd.profiler.record_function("DistributedDataParallel.forward"):
            will_sync = torch.is_grad_enabled() and self.require_backward_grad_sync
            if self.optimizer is not None and will_sync and not (self.device_type == "cuda" and torch.cuda.is_current_stream_capturing()):
                # the fused kernel reads lr / momentum / weight decay from device memory DURING backward: publish what a stock
                # LR scheduler or warm-up wrote into param_groups since the last step (a tuple comparison when nothing changed)
                self.optimizer.push_hyperparameters()
            if self._join_active:
                self._join_count(1)  # tells the ranks that already ran out of data that one more iteration is coming
                self._join_iters += 1
            if will_sync and not self._rebuild_checked and self.reducer.stats["iterations"] >= 1:
                self._rebuild_checked = True
                self.reducer.maybe_rebuild_buckets()
            if self.require_forward_param_sync and self.broadcast_buffers and self.training:
                self._sync_buffers()
            if self.device_ids and self.device_type == "cuda":
                dev = torch.device("cuda", self.device_ids[0])
                inputs = _to_device(inputs, dev)
                kwargs = _to_device(kwargs, dev)
            out = self.module(*inputs, **kwargs)
            if will_sync:
                unused = ()
                if self.find_unused_parameters:
                    if self.static_graph and self._static_unused is not None:
                        unused = self._static_unused  # static_graph: the set found in the first iteration holds for all
                    else:
                        reach = _reachable_params(out)
                        unused = [i for i, p in enumerate(self._params) if id(p) not in reach]
                        if self.static_graph:
                            self._static_unused = unused
                self.reducer.require_sync = True
                self.reducer.prepare_for_backward(unused)
            else:
                self.reducer.require_sync = False
            self.require_forward_param_sync = self.require_backward_grad_sync
            return out

    @contextlib.contextmanager
    def no_sync(self):
        """Skip gradient synchronisation inside the context; gradients accumulate locally and the first
        forward/backward after it synchronises the sum (torch distributed.py:1474-1500)."""
        old = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        try:
            yield
        finally:
            self.require_backward_grad_sync = old

    # ------------------------------------------------------------------------------------------ uneven inputs
    def _join_count(self, mine: int) -> int:
        dev = self._params[0].device if dist.get_backend(self.process_group) == "nccl" else "cpu"
        t = torch.tensor([mine], dtype=torch.int32, device=dev)
        dist.all_reduce(t, group=self.process_group)
        return int(t.item())

    @contextlib.contextmanager
    def join(self, divide_by_initial_world_size: bool = True, enable: bool = True):
        """Training with a different number of batches per rank (torch distributed.py ``join``, the Join hook
        ``_DDPJoinHook``).  Inside the context every forward first all-reduces a one ("I am still training"); a rank that
        leaves the context keeps answering with a zero and, while the sum is positive, *shadows* the iteration of the others:
        the per-forward buffer broadcast and one all-reduce per gradient bucket with a zero contribution, in bucket order,
        so that nobody blocks in a collective.  Gradients are divided by the initial world size (the only mode offered).
        With the fused all-reduce + SGD kernel a shadowing rank applies the same update as everyone else and stays in
        sync; otherwise the parameters of the rank that trained longest are broadcast to all when the last rank leaves.
        The reference never has to deal with this: its loaders give every rank the same number of batches
        (model_parallel.py:92-97)."""
        if not divide_by_initial_world_size:
            raise ValueError("join(divide_by_initial_world_size=False) is not supported: the 1/W scale is compiled into the "
                             "bucket kernels' launch arguments")
        if not enable or self.world_size == 1:
            yield
            return
        self._join_active, self._join_iters = True, 0
        try:
            yield
        finally:
            self._join_active = False
            while self._join_count(0) > 0:
                if self.require_forward_param_sync and self.broadcast_buffers and self.training:
                    self._sync_buffers()
                self.reducer.shadow_iteration()
            # whoever trained longest owns the final model (ties: highest rank), unless the fused optimizer kept all in sync
            best = self._join_iters * self.world_size + self.rank
            dev = self._params[0].device if dist.get_backend(self.process_group) == "nccl" else "cpu"
            t = torch.tensor([best], dtype=torch.int64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.process_group)
            if self.optimizer is None:
                self._sync_module_states(src=int(t.item()) % self.world_size)

    # ------------------------------------------------------------------------------------------ hooks / misc API
    def register_comm_hook(self, state: Any, hook: Callable):
        """``hook(state, bucket) -> Future[Tensor]``; replaces the built-in bucket all-reduce."""
        if not callable(hook):
            raise TypeError("Communication hook must be callable.")
        if self.reducer.fused_cfg is not None:
            raise RuntimeError("communication hooks cannot be combined with fused_sgd")
        self.reducer.comm_hook = hook
        self.reducer.comm_hook_state = state

    def _set_static_graph(self):
        """The autograd graph is the same in every iteration: the set of unused parameters is searched once, in the first
        synchronised forward, and reused (torch distributed.py ``_set_static_graph``)."""
        self.static_graph = True
        self.reducer.static_graph = True

    def _get_ddp_logging_data(self) -> dict:
        """Construction-time and run-time facts about this wrapper (torch distributed.py ``_get_ddp_logging_data`` / the
        ``dist.Logger`` it feeds): back end, world size, bucket layout, iteration / launch / rebuild counters.  The trainers
        write it into their JSONL logs once per epoch."""
        red = self.reducer
        d = dict(self.logger_data)
        d.update(rank=self.rank, num_parameters=len(self._params), num_buffers=len(self.modules_buffers),
                 num_buckets=len(red.buckets), bucket_sizes=red.bucket_sizes(),
                 total_parameter_bytes=sum(p.numel() * p.element_size() for p in self._params),
                 find_unused_parameters=self.find_unused_parameters, gradient_as_bucket_view=self.gradient_as_bucket_view,
                 static_graph=self.static_graph, broadcast_buffers=self.broadcast_buffers,
                 fused_optimizer=self.optimizer is not None, iterations=red.stats.get("iterations", 0),
                 buckets_launched=red.stats.get("buckets_launched", 0), rebuilds=red.stats.get("rebuilds", 0),
                 globally_unused_last=red.stats.get("globally_unused", 0),
                 comm_kernel_launches=getattr(self.backend, "kernel_launches", None),
                 nvls_launches=getattr(self.backend, "nvls_launches", None))
        return d

    def check_health(self):
        if hasattr(self.backend, "check_health"):
            self.backend.check_health()

    def start_watchdog(self, interval_s: float = 5.0, on_fault=None):
        """Background thread that reads the communicator status word every ``interval_s`` seconds (on its own CUDA stream) and
        reports a timed-out peer through ``on_fault`` (default: print and exit 70).  Returns the ``CommWatchdog``; call
        ``.stop()`` on it when training ends.  A no-op probe with the torch/gloo back end, which has its own timeouts."""
        from .watchdog import CommWatchdog

        backend = self.backend
        stream = None

        def probe():
            nonlocal stream
            if not hasattr(backend, "check_health") or getattr(backend, "state", None) is None:
                return
            if stream is None:
                stream = torch.cuda.Stream(device=backend.device)
            with torch.cuda.device(backend.device), torch.cuda.stream(stream):
                backend.check_health()

        return CommWatchdog(probe, interval_s, on_fault).start()

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        r = super().load_state_dict(state_dict, strict=strict)
        self.reducer.sync_shadow()
        return r

    def __getstate__(self):
        """Process group, back end, Reducer and everything that holds peer mappings stay behind (torch distributed.py
        :1306-1320); the fused optimizer's momentum travels as a plain state_dict."""
        d = self.__dict__.copy()
        opt = d.get("optimizer")
        d["_pickled_optimizer_state"] = opt.state_dict() if opt is not None else None
        for k in ("reducer", "backend", "process_group", "optimizer", "_buffer_arena", "_buffer_state", "_buffer_dst"):
            d.pop(k, None)
        # parameters / buffers are views into peer-visible arenas here; pickle them as ordinary tensors
        return d

    def __setstate__(self, state):
        """Rebuilds the process-group handle, back end, Reducer (hooks included), fused optimizer facade and buffer arena in
        the unpickling process (torch distributed.py:1322-1336).  Needs ``torch.distributed`` initialised for world sizes > 1."""
        opt_sd = state.pop("_pickled_optimizer_state", None)
        self.__dict__.update(state)
        self.process_group = dist.group.WORLD if dist.is_initialized() else None
        self.world_size = dist.get_world_size(self.process_group) if self.process_group is not None else 1
        self.rank = dist.get_rank(self.process_group) if self.process_group is not None else 0
        # the unpickled parameters are fresh tensors: refresh the lists the Reducer is built from
        byname = dict(self.module.named_parameters())
        self._params = [byname[n] for n in self._param_names]
        self.modules_buffers = [b for n, b in self.module.named_buffers() if n not in self.parameters_to_ignore]
        self._build_reducer()
        if opt_sd is not None and self.optimizer is not None:
            self.optimizer.load_state_dict(opt_sd)
        self._rebuild_checked = False
        self._join_active, self._join_iters = False, 0


def _to_device(obj, dev):
    if isinstance(obj, torch.Tensor):
        return obj.to(dev, non_blocking=True) if obj.device != dev else obj
    if isinstance(obj, tuple):
        return tuple(_to_device(o, dev) for o in obj)
    if isinstance(obj, list):
        return [_to_device(o, dev) for o in obj]
    if isinstance(obj, dict):
        return {k: _to_device(v, dev) for k, v in obj.items()}
    return obj
