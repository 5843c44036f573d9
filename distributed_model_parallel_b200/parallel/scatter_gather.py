"""Recursive scatter / gather over nested containers, with autograd Functions whose backward is the other.

Behavioural model: torch/nn/parallel/scatter_gather.py:49-150 and _functions.py:82-155 (the code
``nn.DataParallel`` runs per forward, executed by the reference at
/root/reference/code/distributed_training/data_parallel.py:77,108 and dissected in its README :17-143).
Tensors move with the peer-memory kernels in ``comm.py``; non-tensors are replicated by reference.
"""
from __future__ import annotations

import warnings
from typing import Any, List, Optional, Sequence

import torch

from . import comm


def _is_namedtuple(obj) -> bool:
    return isinstance(obj, tuple) and hasattr(obj, "_asdict") and hasattr(obj, "_fields")


class Scatter(torch.autograd.Function):
    @staticmethod
    def forward(ctx, target_devices, chunk_sizes, dim, input):
        ctx.dim = dim
        ctx.input_device = input.device
        outputs = comm.scatter(input, target_devices, chunk_sizes, dim)
        return tuple(outputs)

    @staticmethod
    def backward(ctx, *grad_output):
        return None, None, None, Gather.apply(ctx.input_device, ctx.dim, *grad_output)


class Gather(torch.autograd.Function):
    @staticmethod
    def forward(ctx, target_device, dim, *inputs):
        ctx.target_device = target_device
        ctx.dim = dim
        ctx.input_devices = tuple(i.device for i in inputs)
        if all(t.dim() == 0 for t in inputs) and dim == 0:
            inputs = tuple(t.view(1) for t in inputs)
            warnings.warn("Was asked to gather along dimension 0, but all input tensors were scalars; will instead "
                          "unsqueeze and return a vector.")
            ctx.unsqueezed_scalar = True
        else:
            ctx.unsqueezed_scalar = False
        ctx.input_sizes = tuple(i.size(ctx.dim) for i in inputs)
        return comm.gather(inputs, ctx.dim, ctx.target_device)

    @staticmethod
    def backward(ctx, grad_output):
        scattered = Scatter.apply(ctx.input_devices, ctx.input_sizes, ctx.dim, grad_output)
        if ctx.unsqueezed_scalar:
            scattered = tuple(g[0] for g in scattered)
        return (None, None) + scattered


def scatter(inputs, target_devices: Sequence, dim: int = 0):
    """Slice tensors into approximately equal chunks across devices; duplicate references to everything else."""

    def scatter_map(obj):
        if isinstance(obj, torch.Tensor):
            return Scatter.apply(list(target_devices), None, dim, obj)
        if _is_namedtuple(obj):
            return [type(obj)(*args) for args in zip(*map(scatter_map, obj))]
        if isinstance(obj, tuple) and len(obj) > 0:
            return list(zip(*map(scatter_map, obj)))
        if isinstance(obj, list) and len(obj) > 0:
            return [list(i) for i in zip(*map(scatter_map, obj))]
        if isinstance(obj, dict) and len(obj) > 0:
            return [type(obj)(i) for i in zip(*map(scatter_map, obj.items()))]
        return [obj for _ in target_devices]

    try:
        return scatter_map(inputs)
    finally:
        scatter_map = None  # break the closure's reference cycle


def scatter_kwargs(inputs, kwargs, target_devices: Sequence, dim: int = 0):
    """Scatter positional and keyword arguments, padding the shorter list so both have one entry per chunk."""
    scattered_inputs = scatter(inputs, target_devices, dim) if inputs else []
    scattered_kwargs = scatter(kwargs, target_devices, dim) if kwargs else []
    if len(scattered_inputs) < len(scattered_kwargs):
        scattered_inputs = list(scattered_inputs) + [() for _ in range(len(scattered_kwargs) - len(scattered_inputs))]
    elif len(scattered_kwargs) < len(scattered_inputs):
        scattered_kwargs = list(scattered_kwargs) + [{} for _ in range(len(scattered_inputs) - len(scattered_kwargs))]
    return tuple(scattered_inputs), tuple(scattered_kwargs)


def gather(outputs: Any, target_device, dim: int = 0) -> Any:
    """Inverse of scatter: concatenate per-device results (recursively for containers) on ``target_device``."""

    def gather_map(outs):
        out = outs[0]
        if isinstance(out, torch.Tensor):
            return Gather.apply(target_device, dim, *outs)
        if out is None:
            return None
        if isinstance(out, dict):
            if not all(len(out) == len(d) for d in outs):
                raise ValueError("All dicts must have the same number of keys")
            return type(out)((k, gather_map([d[k] for d in outs])) for k in out)
        if _is_namedtuple(out):
            return type(out)._make(map(gather_map, zip(*outs)))
        return type(out)(map(gather_map, zip(*outs)))

    try:
        return gather_map(outputs)
    finally:
        gather_map = None
