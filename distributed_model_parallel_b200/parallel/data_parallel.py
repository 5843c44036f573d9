"""DataParallel: single process, one thread per replica -- same call shape as ``torch.nn.DataParallel``.

``DataParallel(module, device_ids=None, output_device=None, dim=0)`` is what the reference's DP trainer wraps its
network in (/root/reference/code/distributed_training/data_parallel.py:77).  Forward = scatter inputs ->
replicate module (coalesced broadcast of parameters + buffers) -> parallel_apply (threads) -> gather outputs;
backward runs Gather.backward (scatter of the output gradient) and Broadcast.backward (reduce-add of replica
gradients onto device 0).  torch/nn/parallel/data_parallel.py:136-290 is the behavioural model; every
device-to-device transfer is one of our peer-memory kernels (``parallel/comm.py``).

``state_dict()`` keys carry the ``module.`` prefix, so the reference's best-accuracy checkpoints
(data_parallel.py:147-151) round-trip.  CPU "devices" are accepted so that the container logic is testable without
GPUs.
"""
from __future__ import annotations

import warnings
from itertools import chain
from typing import Any, Optional, Sequence

import torch
import torch.nn as nn

from ..utils.tracing import nvtx_range

from .comm import _dev
from .parallel_apply import parallel_apply
from .replicate import replicate
from .scatter_gather import gather, scatter_kwargs


def _check_balance(devices):
    props = [torch.cuda.get_device_properties(d) for d in devices if d.type == "cuda"]
    if len(props) < 2:
        return
    for name, values in (("memory", [p.total_memory for p in props]), ("cores", [p.multi_processor_count for p in props])):
        lo, hi = min(values), max(values)
        if lo / hi < 0.75:
            warnings.warn("There is an imbalance between your GPUs (%s): device %d has less than 75%% of device %d. "
                          "Exclude the weaker device via device_ids." % (name, values.index(lo), values.index(hi)))
            return


class DataParallel(nn.Module):
    def __init__(self, module: nn.Module, device_ids: Optional[Sequence] = None, output_device=None, dim: int = 0):
        super().__init__()
        if device_ids is None:
            if not torch.cuda.is_available():
                self.module, self.device_ids = module, []
                self.dim = dim
                return
            device_ids = list(range(torch.cuda.device_count()))
        self.devices = [_dev(d) for d in device_ids]
        self.device_ids = [d.index if d.type == "cuda" else d for d in self.devices]
        self.output_device = _dev(output_device) if output_device is not None else self.devices[0]
        self.src_device_obj = self.devices[0]
        self.dim = dim
        self.module = module
        if all(d.type == "cuda" for d in self.devices):
            _check_balance(self.devices)
        if len(self.devices) == 1:
            self.module.to(self.devices[0])
        self._mc_arena = None
        self._mc_tried = False
        self._replica_set = None

    def forward(self, *inputs: Any, **kwargs: Any) -> Any:
        with torch.autograd.profiler.record_function("DataParallel.forward"):
            if not self.device_ids:
                return self.module(*inputs, **kwargs)
            for t in chain(self.module.parameters(), self.module.buffers()):
                if t.device != self.src_device_obj:
                    raise RuntimeError("module must have its parameters and buffers on device %s (device_ids[0]) but found "
                                       "one of them on device: %s" % (self.src_device_obj, t.device))
            with nvtx_range("dp.scatter"):
                inputs, module_kwargs = self.scatter(inputs, kwargs, self.devices)
            if not inputs and not module_kwargs:  # forward without arguments
                inputs, module_kwargs = ((),), ({},)
            if len(self.devices) == 1:
                return self.module(*inputs[0], **module_kwargs[0])
            rs = self._replica_set
            if rs is not None and len(inputs) == len(self.devices):
                # persistent replicas: parameters and buffers travel as two flat kernels, nothing is cloned or concatenated
                with nvtx_range("dp.broadcast"):
                    rs.broadcast()
                with nvtx_range("dp.parallel_apply"):
                    outputs = self.parallel_apply(rs.replicas, inputs, module_kwargs)
                with nvtx_range("dp.gather"):
                    out = self.gather(outputs, self.output_device)
                if torch.is_grad_enabled():
                    out = _map_tensors(out, lambda t: _ReduceAfterBackward.apply(t, rs) if t.requires_grad else t)
                return out
            with nvtx_range("dp.replicate"):
                replicas = self.replicate(self.module, self.devices[: len(inputs)])
            with nvtx_range("dp.parallel_apply"):
                outputs = self.parallel_apply(replicas, inputs, module_kwargs)
            with nvtx_range("dp.gather"):
                return self.gather(outputs, self.output_device)

    def attach_optimizer(self, optimizer):
        """Switches this wrapper to persistent replicas laid out like ``optimizer``'s flat arenas (``FlatFusedSGD``): the
        per-forward replicate() / per-backward tensor-wise reduction become one broadcast kernel and one reduction kernel
        (see ``parallel/dp_engine.py``).  Returns the ``ReplicaSet``.  Call again after re-creating the optimizer."""
        from .dp_engine import ReplicaSet
        if len(self.devices) < 2 or not all(d.type == "cuda" for d in self.devices):
            return None
        self._replica_set = ReplicaSet(self.module, self.devices, optimizer)
        return self._replica_set

    def _multicast_arena(self):
        """NVLS multicast arena for the per-forward parameter/buffer broadcast (created once; None if unsupported or
        disabled with DMP_DP_MULTICAST=0)."""
        import os
        if self._mc_tried:
            return self._mc_arena
        self._mc_tried = True
        if os.environ.get("DMP_DP_MULTICAST", "1") == "0" or not all(d.type == "cuda" for d in self.devices) or len(self.devices) < 2:
            return None
        try:
            from .symm import MulticastArena
            total = sum(t.numel() * t.element_size() + 512 for t in chain(self.module.parameters(), self.module.buffers()))
            self._mc_arena = MulticastArena([d.index for d in self.devices], total + (1 << 20))
        except Exception as e:  # noqa: BLE001  (no NVSwitch multicast on this platform: keep peer stores)
            warnings.warn("DataParallel: NVLS multicast unavailable (%s); broadcasting with peer stores" % str(e)[:120])
            self._mc_arena = None
        return self._mc_arena

    def replicate(self, module, devices):
        from . import comm
        arena = self._multicast_arena() if len(devices) == len(self.devices) else None
        with comm.use_multicast_arena(arena):
            return replicate(module, devices, not torch.is_grad_enabled())

    def scatter(self, inputs, kwargs, devices):
        return scatter_kwargs(inputs, kwargs, devices, dim=self.dim)

    def parallel_apply(self, replicas, inputs, kwargs):
        return parallel_apply(replicas, inputs, kwargs, self.devices[: len(replicas)])

    def gather(self, outputs, output_device):
        return gather(outputs, output_device, dim=self.dim)


class _ReduceAfterBackward(torch.autograd.Function):
    """Identity on the gathered output whose backward queues, once per backward pass, the reduction of the persistent
    replicas' gradient arenas onto device 0 (it runs when the whole backward has been issued on every device)."""

    @staticmethod
    def forward(ctx, t, rs):
        ctx.rs = rs
        return t.view_as(t)

    @staticmethod
    def backward(ctx, g):
        rs = ctx.rs
        if not getattr(rs, "_reduce_queued", False):
            rs._reduce_queued = True

            def _cb():
                rs._reduce_queued = False
                rs.reduce_grads(fused_sgd=False)

            torch.autograd.Variable._execution_engine.queue_callback(_cb)
        return g, None


def _map_tensors(obj, fn):
    if isinstance(obj, torch.Tensor):
        return fn(obj)
    if isinstance(obj, tuple) and hasattr(obj, "_fields"):
        return type(obj)(*(_map_tensors(o, fn) for o in obj))
    if isinstance(obj, (list, tuple)):
        return type(obj)(_map_tensors(o, fn) for o in obj)
    if isinstance(obj, dict):
        return type(obj)((k, _map_tensors(v, fn)) for k, v in obj.items())
    return obj


def data_parallel(module: nn.Module, inputs: Any, device_ids: Optional[Sequence] = None, output_device=None, dim: int = 0,
                  module_kwargs: Optional[Any] = None) -> Any:
    """Functional form: evaluate ``module(inputs)`` in parallel across ``device_ids``."""
    if not isinstance(inputs, tuple):
        inputs = (inputs,) if inputs is not None else ()
    if device_ids is None:
        device_ids = list(range(torch.cuda.device_count()))
    devices = [_dev(d) for d in device_ids]
    out_dev = _dev(output_device) if output_device is not None else devices[0]
    src = devices[0]
    for t in chain(module.parameters(), module.buffers()):
        if t.device != src:
            raise RuntimeError("module must have its parameters and buffers on device %s (device_ids[0]) but found one of "
                               "them on device: %s" % (src, t.device))
    inputs, module_kwargs = scatter_kwargs(inputs, module_kwargs, devices, dim)
    if not inputs and not module_kwargs:
        inputs, module_kwargs = ((),), ({},)
    if len(devices) == 1:
        return module(*inputs[0], **module_kwargs[0])
    used = devices[: len(inputs)]
    replicas = replicate(module, used, not torch.is_grad_enabled())
    outputs = parallel_apply(replicas, inputs, module_kwargs, used)
    return gather(outputs, out_dev, dim)
