"""Persistent replicas and a CUDA-graph step for the single-process DataParallel path.

torch's DataParallel -- the path the reference's DP trainer runs (/root/reference/code/distributed_training/data_parallel.py:77,
99-120, dissected in its README :17-143) -- re-creates every replica on every forward: it flattens all parameters, broadcasts them,
clones ~150 module objects per device and wires the copies in, then reduces the replicas' gradients tensor by tensor.  Here the
replicas are built ONCE:

* device 0 keeps the user's module, whose parameters already live in the flat arenas of ``FlatFusedSGD`` (fp32 master, bf16
  shadow, fp32 gradient);
* every other device gets a deep copy whose parameters are views into arenas of the SAME layout, so
  - the per-forward parameter "broadcast_coalesced" is ONE kernel on device 0 that stores the master arena into every replica
    and writes their bf16 shadows in the same pass (``dmp_broadcast`` with its fp32 -> bf16 arm; an NVLS multicast store when the
    arenas sit in a multicast buffer),
  - "ReduceAddCoalesced" is ONE kernel on device 0 that peer-loads every replica's gradient arena, adds it to its own and
    zeroes the sources (``dmp_reduce_add_peer_z``), optionally applying the SGD update in the same pass;
* buffers (BatchNorm running statistics) are copied from device 0 as one flat arena as well; only device 0's survive a step,
  exactly like torch's replica 0 sharing storage with the base module.

``DataParallelStep`` goes one step further than thread-per-replica execution: each replica's forward + loss + backward is
captured in a CUDA graph on its own device, and a training step is scatter -> broadcast -> N graph launches from ONE thread ->
reduce(+SGD).  The thread-per-replica ``DataParallel.forward`` stays available (and is what ``bench.py`` reports as
``dp_value``); the graph step is reported next to it."""
from __future__ import annotations

import copy
from typing import List, Optional, Sequence

import torch
import torch.nn as nn

from .. import _native
from ..ops import functional as OF
from . import comm


def _is_flat_fused(opt) -> bool:
    return all(hasattr(opt, a) for a in ("master", "grad", "momentum_buf", "_offsets", "_params", "hp"))


class ReplicaSet:
    def __init__(self, module: nn.Module, devices: Sequence[torch.device], optimizer):
        if not _is_flat_fused(optimizer):
            raise TypeError("persistent replicas need a FlatFusedSGD optimizer over the module's parameters")
        self.module, self.devices, self.opt = module, list(devices), optimizer
        self.lib = _native.kernels()
        comm._ensure_peer_access(self.devices)
        base_params = list(optimizer._params)
        ids = {id(p): i for i, p in enumerate(base_params)}
        named = [(n, p) for n, p in module.named_parameters() if id(p) in ids]
        if len(named) != len(base_params):
            raise ValueError("the optimizer must own exactly the module's trainable parameters")
        self.numel = optimizer._numel
        self.replicas: List[nn.Module] = [module]
        self.master = [optimizer.master]
        self.shadow = [optimizer.shadow]
        self.grad = [optimizer.grad]
        for d in self.devices[1:]:
            with torch.cuda.device(d):
                rep = copy.deepcopy(module).to(d)
                master = torch.zeros(self.numel, dtype=torch.float32, device=d)
                shadow = torch.zeros(self.numel, dtype=torch.bfloat16, device=d) if optimizer.shadow is not None else None
                grad = torch.zeros(self.numel, dtype=torch.float32, device=d)
                rp = dict(rep.named_parameters())
                for name, p0 in named:
                    off = optimizer._offsets[ids[id(p0)]]
                    p = rp[name]
                    p.data = master[off: off + p.numel()].view_as(p)
                    p.grad = grad[off: off + p.numel()].view_as(p)
                    p._dmp_direct = True
                    if shadow is not None:
                        p._dmp_shadow = shadow[off: off + p.numel()].view_as(p)
            self.replicas.append(rep)
            self.master.append(master)
            self.shadow.append(shadow)
            self.grad.append(grad)
        # buffers: one flat byte arena per device in module.buffers() order
        self._buf_meta = []
        total = 0
        for b in module.buffers():
            nb = b.numel() * b.element_size()
            self._buf_meta.append((total, nb))
            total += (nb + 15) // 16 * 16
        self.buf_bytes = total
        self.buf_arena = []
        for k, d in enumerate(self.devices):
            arena = torch.zeros(max(total, 16), dtype=torch.uint8, device=d)
            for b, (o, nb) in zip(self.replicas[k].buffers(), self._buf_meta):
                v = arena[o: o + nb].view(b.dtype).view(b.shape)
                v.copy_(b)
                b.data = v
            self.buf_arena.append(arena)
        from .symm import ptr_array
        self._master_dst = ptr_array([m.data_ptr() for m in self.master[1:]])
        self._shadow_dst = ptr_array([s.data_ptr() for s in self.shadow[1:]]) if optimizer.shadow is not None else None
        self._buf_dst = ptr_array([a.data_ptr() for a in self.buf_arena[1:]])
        self._grad_src = ptr_array([g.data_ptr() for g in self.grad])
        self.kernel_launches = 0

    # --------------------------------------------------------------------------------------------------------------
    def broadcast(self, stream0: Optional[torch.cuda.Stream] = None):
        """Device 0's parameters (fp32 master -> every replica's master + bf16 shadow) and buffers, two kernels on device 0."""
        n = len(self.devices)
        if n < 2:
            return
        d0 = self.devices[0]
        s0 = stream0 if stream0 is not None else torch.cuda.current_stream(d0)
        for d in self.devices[1:]:  # the destinations must be idle before the source-side kernel overwrites them
            with torch.cuda.device(d):
                ev = torch.cuda.Event()
                ev.record(torch.cuda.current_stream(d))
            s0.wait_event(ev)
        with torch.cuda.device(d0):
            _native.check(self.lib.dmp_broadcast(self.master[0].data_ptr(), self._master_dst, n - 1, None, self._shadow_dst, self.numel * 4,
                                                 None, 0, n, 1, 0, s0.cuda_stream), "dp parameter broadcast")
            if self.buf_bytes:
                _native.check(self.lib.dmp_broadcast(self.buf_arena[0].data_ptr(), self._buf_dst, n - 1, None, None,
                                                     (self.buf_bytes + 15) // 16 * 16, None, 0, n, 1, 0, s0.cuda_stream),
                              "dp buffer broadcast")
            self.kernel_launches += 2
        with torch.cuda.device(d0):
            done = torch.cuda.Event()
            done.record(s0)
        for d in self.devices[1:]:
            torch.cuda.current_stream(d).wait_event(done)

    def reduce_grads(self, fused_sgd: bool = False, scale: float = 1.0, stream0: Optional[torch.cuda.Stream] = None):
        """Sum of every replica's gradient arena into device 0's (sources zeroed), or -- ``fused_sgd`` -- straight into the SGD
        update of device 0's master / momentum / shadow: one kernel on device 0 reading its peers over NVLink."""
        n = len(self.devices)
        d0 = self.devices[0]
        s0 = stream0 if stream0 is not None else torch.cuda.current_stream(d0)
        for m in OF.all_side_stream_managers():
            m.join_all()
        for d in self.devices[1:]:
            with torch.cuda.device(d):
                ev = torch.cuda.Event()
                ev.record(torch.cuda.current_stream(d))
            s0.wait_event(ev)
        opt = self.opt
        with torch.cuda.device(d0):
            if fused_sgd:
                opt.push_hyperparameters()
                _native.check(self.lib.dmp_reduce_add_peer_z(0, self._grad_src, n, self.numel, float(scale), None, opt.master.data_ptr(),
                                                             opt.momentum_buf.data_ptr(), _native.ptr(opt.shadow), opt.hp.data_ptr(),
                                                             1 if opt.param_groups[0]["nesterov"] else 0, 1, 0, s0.cuda_stream),
                              "dp reduce + sgd")
            else:
                _native.check(self.lib.dmp_reduce_add_peer_z(0, self._grad_src, n, self.numel, float(scale), self.grad[0].data_ptr(), None,
                                                             None, None, None, 0, 0, 0, s0.cuda_stream), "dp gradient reduce")
                # the sources other than device 0's own arena are zeroed by a second tiny pass (device 0's arena holds the result)
                for k in range(1, n):
                    self.grad[k].zero_()
            self.kernel_launches += 1
            done = torch.cuda.Event()
            done.record(s0)
        for d in self.devices[1:]:
            torch.cuda.current_stream(d).wait_event(done)


class DataParallelStep:
    """One training step of a ``DataParallel`` model with every replica's forward + loss + backward captured in a CUDA graph
    on its own device.  ``step(x, y)``: x, y on the host (pinned) or on device 0; returns device-0 stats {loss, top-1, top-5}."""

    def __init__(self, dp, optimizer, warmup_iters: int = 2):
        self.dp, self.opt = dp, optimizer
        self.devices = list(dp.devices)
        self.rs = dp.attach_optimizer(optimizer)
        self.warmup_iters = warmup_iters
        self.graphs = None
        self.stats = None
        self.launches_per_iteration = 0
        self.steps = 0

    def _iteration(self, k):
        rep = self.rs.replicas[k]
        out = rep(self.x[k])
        loss = OF.cross_entropy(out if out.dtype in (torch.float32, torch.bfloat16) else out.float(), self.y[k], self.stat[k]) * self.frac[k]
        loss.backward()

    def _setup(self, x, y):
        n = len(self.devices)
        B = x.shape[0]
        sizes = [c.shape[0] for c in x.chunk(n, 0)]
        if len(sizes) != n:
            raise ValueError("batch of %d rows cannot be split over %d devices" % (B, n))
        self.sizes = sizes
        self.x = [torch.empty((s,) + tuple(x.shape[1:]), dtype=x.dtype, device=d) for s, d in zip(sizes, self.devices)]
        self.y = [torch.empty((s,), dtype=y.dtype, device=d) for s, d in zip(sizes, self.devices)]
        self.stat = [torch.zeros(3, dtype=torch.float32, device=d) for d in self.devices]
        self.frac = [float(s) / float(B) for s in sizes]
        self.stats = torch.zeros(3, dtype=torch.float32, device=self.devices[0])
        self._stats_host = torch.zeros(3, dtype=torch.float32).pin_memory()
        self._stat0 = [torch.zeros(3, dtype=torch.float32, device=self.devices[0]) for _ in self.devices]
        self._copy_inputs(x, y)
        self.rs.broadcast()
        self.cstreams = [torch.cuda.Stream(device=d) for d in self.devices]
        self.graphs = []
        # device 0's BatchNorm buffers are advanced by the warm-up iterations (the replicas' are overwritten by the next
        # broadcast anyway): keep a copy and put it back after the captures
        saved = self.rs.buf_arena[0].clone()
        for k, d in enumerate(self.devices):
            with torch.cuda.device(d):
                s = self.cstreams[k]
                s.wait_stream(torch.cuda.current_stream(d))
                with torch.cuda.stream(s):
                    for _ in range(self.warmup_iters):  # lazy initialisation must not end up inside the capture
                        self._iteration(k)
                torch.cuda.current_stream(d).wait_stream(s)
                torch.cuda.synchronize(d)
                self.rs.grad[k].zero_()
                g = torch.cuda.CUDAGraph()
                n0 = _native.launch_count()
                with torch.cuda.graph(g, stream=s):
                    self._iteration(k)
                self.graphs.append(g)
                if k == 0:
                    self.launches_per_iteration = (_native.launch_count() - n0) * n + 3  # + broadcast x2 + reduce
        self.rs.buf_arena[0].copy_(saved)

    def _copy_inputs(self, x, y):
        off = 0
        for k, d in enumerate(self.devices):
            s = self.sizes[k]
            with torch.cuda.device(d):
                self.x[k].copy_(x[off: off + s], non_blocking=True)
                self.y[k].copy_(y[off: off + s], non_blocking=True)
            off += s

    def __call__(self, x, y):
        if self.graphs is None:
            self._setup(x, y)
        self._copy_inputs(x, y)
        self.rs.broadcast()
        for k, d in enumerate(self.devices):
            with torch.cuda.device(d):
                self.graphs[k].replay()
        self.rs.reduce_grads(fused_sgd=True)
        # statistics: sum of the replicas' {loss * n_k / B, hits} on device 0
        d0 = self.devices[0]
        with torch.cuda.device(d0):
            s0 = torch.cuda.current_stream(d0)
            self.stats.zero_()
            for k, d in enumerate(self.devices):
                if k:
                    self._stat0[k].copy_(self.stat[k], non_blocking=True)
                    src = self._stat0[k]
                else:
                    src = self.stat[0]
                self.stats[0] += src[0] * self.frac[k]
                self.stats[1:] += src[1:]
        self.steps += 1
        return self.stats

    def read_stats(self):
        self._stats_host.copy_(self.stats, non_blocking=True)
        torch.cuda.current_stream(self.devices[0]).synchronize()
        return self._stats_host.tolist()
