"""Module replication for the single-process DataParallel path.

Per forward, parameters and buffers of the base module (on ``devices[0]``) are broadcast in coalesced chunks to
every other device and shallow module clones are wired to the copies; replica 0 keeps sharing storage with the
base module, which is how BatchNorm running statistics persist.  The broadcast is differentiable: its backward
is a peer-load reduce-add of every replica's gradients onto device 0.

Behavioural model: torch/nn/parallel/replicate.py:83-214 and _functions.py:10-79 (``Broadcast`` /
``ReduceAddCoalesced``), i.e. README :30-68 of the reference.  The data movement itself is ours
(``comm.broadcast_coalesced`` = source-side peer-store kernel, ``comm.reduce_add_coalesced`` = destination-side
peer-load kernel) instead of ``torch._C._broadcast_coalesced`` / ``nccl.reduce``.
"""
from __future__ import annotations

from typing import Dict, List, Sequence

import torch
import torch.nn as nn

from . import comm


class Broadcast(torch.autograd.Function):
    @staticmethod
    def forward(ctx, target_devices, *inputs):
        ctx.target_devices = list(target_devices)
        if len(inputs) == 0:
            return tuple()
        ctx.num_inputs = len(inputs)
        ctx.input_device = inputs[0].device
        outputs = comm.broadcast_coalesced(inputs, ctx.target_devices)
        non_differentiables = []
        for idx, needs in enumerate(ctx.needs_input_grad[1:]):
            if not needs:
                non_differentiables.extend(output[idx] for output in outputs[1:])
        ctx.mark_non_differentiable(*non_differentiables)
        # replica 0 must be distinct autograd outputs too: hand out views of the inputs
        first = [t.view_as(t) for t in inputs]
        return tuple(first) + tuple(t for tensors in outputs[1:] for t in tensors)

    @staticmethod
    def backward(ctx, *grad_outputs):
        return (None,) + ReduceAddCoalesced.apply(ctx.input_device, ctx.num_inputs, *grad_outputs)


class ReduceAddCoalesced(torch.autograd.Function):
    @staticmethod
    def forward(ctx, destination, num_inputs, *grads):
        ctx.target_devices = [grads[i].device for i in range(0, len(grads), num_inputs)]
        grads_ = [grads[i: i + num_inputs] for i in range(0, len(grads), num_inputs)]
        return tuple(comm.reduce_add_coalesced(grads_, destination))

    @staticmethod
    def backward(ctx, *grad_outputs):
        return (None, None) + Broadcast.apply(ctx.target_devices, *grad_outputs)


def _broadcast_coalesced_reshape(tensors: Sequence[torch.Tensor], devices, detach: bool) -> List[List[torch.Tensor]]:
    if len(tensors) == 0:
        return [[] for _ in devices]
    if detach:
        return comm.broadcast_coalesced(tensors, devices)
    copies = Broadcast.apply(devices, *tensors)
    n = len(tensors)
    return [list(copies[i: i + n]) for i in range(0, len(copies), n)]


def replicate(network: nn.Module, devices: Sequence, detach: bool = False) -> List[nn.Module]:
    devices = [comm._dev(d) for d in devices]
    num_replicas = len(devices)
    params = list(network.parameters())
    param_indices = {p: i for i, p in enumerate(params)}
    param_copies = _broadcast_coalesced_reshape(params, devices, detach)

    buffers = list(network.buffers())
    buffers_rg = [b for b in buffers if b.requires_grad and not detach]
    buffers_not_rg = [b for b in buffers if not (b.requires_grad and not detach)]
    buffer_indices_rg = {b: i for i, b in enumerate(buffers_rg)}
    buffer_indices_not_rg = {b: i for i, b in enumerate(buffers_not_rg)}
    buffer_copies_rg = _broadcast_coalesced_reshape(buffers_rg, devices, detach)
    buffer_copies_not_rg = _broadcast_coalesced_reshape(buffers_not_rg, devices, True)

    modules = list(network.modules())
    module_copies: List[List[nn.Module]] = [[] for _ in devices]
    module_indices: Dict[nn.Module, int] = {}
    for i, module in enumerate(modules):
        module_indices[module] = i
        for j in range(num_replicas):
            replica = module._replicate_for_data_parallel()
            replica._former_parameters = {}
            module_copies[j].append(replica)

    for i, module in enumerate(modules):
        for key, child in module._modules.items():
            if child is None:
                for j in range(num_replicas):
                    module_copies[j][i]._modules[key] = None
            else:
                idx = module_indices[child]
                for j in range(num_replicas):
                    setattr(module_copies[j][i], key, module_copies[j][idx])
        for key, param in module._parameters.items():
            if param is None:
                for j in range(num_replicas):
                    module_copies[j][i]._parameters[key] = None
            else:
                idx = param_indices[param]
                for j in range(num_replicas):
                    replica = module_copies[j][i]
                    p = param_copies[j][idx]
                    # parameters of a replica are plain (non-leaf) tensors so gradients flow back through Broadcast
                    setattr(replica, key, p)
                    replica._former_parameters[key] = p
        for key, buf in module._buffers.items():
            if buf is None:
                for j in range(num_replicas):
                    module_copies[j][i]._buffers[key] = None
            else:
                if buf.requires_grad and not detach:
                    copies, idx = buffer_copies_rg, buffer_indices_rg[buf]
                else:
                    copies, idx = buffer_copies_not_rg, buffer_indices_not_rg[buf]
                for j in range(num_replicas):
                    setattr(module_copies[j][i], key, copies[j][idx])
    return [module_copies[j][0] for j in range(num_replicas)]
