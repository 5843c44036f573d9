"""Layer-split partitioning of ``nn.Sequential``-izable models into pipeline stages.

The reference hard-codes ``Sequential(conv1, bn1, layers[0:3])`` / ``layers[6r-3:6r+3]`` /
``Sequential(layers[15:], conv2, bn2, Reshape1(), linear)`` and therefore only works for exactly 4 ranks
(/root/reference/code/distributed_training/model_parallel.py:102-103,129,143-144; SURVEY.md defect Q1).
Here the cut points come from a balanced contiguous partition (native DP in csrc/runtime/pipeline.cpp) of a
per-layer cost vector; ``reference_split`` reproduces the reference's 3/6/6/2 block split on request.
"""
from __future__ import annotations

import ctypes as C
import time
from typing import List, Optional, Sequence

import torch
import torch.nn as nn

from ... import _native

_native.register_runtime_sigs({
    "dmp_partition_balanced": (C.c_int, [C.c_int, C.POINTER(C.c_double), C.c_int, C.POINTER(C.c_int32)]),
    "dmp_schedule_generate": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int32), C.c_int]),
    "dmp_schedule_simulate": (C.c_double, [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.POINTER(C.c_int32)]),
})


def partition_balanced(costs: Sequence[float], num_stages: int) -> List[int]:
    """Boundaries ``b`` (len W+1) minimising the largest stage cost; stage s owns layers [b[s], b[s+1])."""
    n = len(costs)
    if num_stages < 1 or num_stages > n:
        raise ValueError("cannot split %d layers into %d stages" % (n, num_stages))
    arr = (C.c_double * n)(*[float(c) for c in costs])
    out = (C.c_int32 * (num_stages + 1))()
    rc = _native.runtime().dmp_partition_balanced(n, arr, num_stages, out)
    if rc != 0:
        raise RuntimeError("dmp_partition_balanced failed (%d)" % rc)
    return [int(x) for x in out]


def layer_costs(seq: nn.Sequential, mode: str = "params", sample: Optional[torch.Tensor] = None, repeats: int = 3) -> List[float]:
    """Per-layer cost: ``params`` (parameter count + 1), ``uniform``, or ``time`` (measured forward+backward seconds
    on ``sample``, which also accounts for activation sizes)."""
    layers = list(seq)
    if mode == "uniform":
        return [1.0] * len(layers)
    if mode == "params":
        return [1.0 + sum(p.numel() for p in l.parameters()) for l in layers]
    if mode == "time":
        if sample is None:
            raise ValueError("mode='time' needs a sample input")
        costs, x = [], sample
        for li, l in enumerate(layers):
            xin = x.detach().requires_grad_(x.is_floating_point() and li > 0)  # the network input itself needs no gradient
            best = float("inf")
            for _ in range(repeats):
                if x.is_cuda:
                    torch.cuda.synchronize()
                t0 = time.perf_counter()
                y = l(xin)
                if y.requires_grad:
                    y.float().sum().backward()
                if x.is_cuda:
                    torch.cuda.synchronize()
                best = min(best, time.perf_counter() - t0)
            costs.append(best)
            x = y.detach()
        for l in layers:
            l.zero_grad(set_to_none=True)
        return costs
    if mode == "graph_time":
        # What a layer costs INSIDE a captured stage: forward + backward of the layer replayed from a CUDA graph and timed with
        # events.  Eager timing (mode "time") mostly counts kernel launches -- the host cannot issue them as fast as a B200
        # retires them -- while the pipeline stages replay graphs, where only device time is left.
        if sample is None or not sample.is_cuda:
            raise ValueError("mode='graph_time' needs a CUDA sample input")
        costs, x = [], sample
        side = torch.cuda.Stream(device=sample.device)
        for li, l in enumerate(layers):
            xin = x.detach().clone().requires_grad_(x.is_floating_point() and li > 0)  # the network input needs no gradient
            side.wait_stream(torch.cuda.current_stream(sample.device))
            with torch.cuda.stream(side):
                for _ in range(2):  # eager warm-up: lazy initialisation must not land in the capture
                    y = l(xin)
                    if y.requires_grad:
                        torch.autograd.backward(y, torch.ones_like(y))
            torch.cuda.current_stream(sample.device).wait_stream(side)
            torch.cuda.synchronize(sample.device)
            try:
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=side):
                    y = l(xin)
                    if y.requires_grad:
                        torch.autograd.backward(y, torch.ones_like(y))
                g.replay()
                torch.cuda.synchronize(sample.device)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(repeats):
                    g.replay()
                e1.record()
                torch.cuda.synchronize(sample.device)
                costs.append(e0.elapsed_time(e1) / repeats * 1e-3)
                del g
            except Exception:  # a layer that cannot be captured: eager device time
                torch.cuda.synchronize(sample.device)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                y = l(xin)
                if y.requires_grad:
                    torch.autograd.backward(y, torch.ones_like(y))
                e1.record()
                torch.cuda.synchronize(sample.device)
                costs.append(e0.elapsed_time(e1) * 1e-3)
            x = y.detach()
        for l in layers:
            l.zero_grad(set_to_none=True)
        return costs
    raise ValueError("unknown cost mode %r" % mode)


def measured_boundaries(seq: nn.Sequential, num_stages: int, sample: torch.Tensor, device, rank: int, group=None,
                        mode: str = "graph_time") -> List[int]:
    """Stage boundaries from per-layer costs measured on ``device`` by rank 0 (the whole model visits that GPU once, layer
    by layer) and broadcast, so every stage adopts the same cut.  ``sample``: one MICRO-batch of stage-0 input."""
    import torch.distributed as dist

    b = torch.zeros(num_stages + 1, dtype=torch.int64, device=device)
    if rank == 0:
        seq.to(device)
        buffers = [t.detach().clone() for t in seq.buffers()]  # train-mode passes must not leak into BatchNorm statistics
        costs = layer_costs(seq, mode, sample=sample.to(device))
        with torch.no_grad():
            for t, v in zip(seq.buffers(), buffers):
                t.copy_(v)
        seq.to("cpu")
        b.copy_(torch.tensor(partition_balanced(costs, num_stages), dtype=torch.int64))
    if dist.is_available() and dist.is_initialized():
        dist.broadcast(b, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    return [int(v) for v in b.tolist()]


def split_sequential(seq: nn.Sequential, num_stages: int, costs: Optional[Sequence[float]] = None,
                     boundaries: Optional[Sequence[int]] = None) -> List[nn.Sequential]:
    layers = list(seq)
    if boundaries is None:
        boundaries = partition_balanced(costs if costs is not None else layer_costs(seq), num_stages)
    if len(boundaries) != num_stages + 1 or boundaries[0] != 0 or boundaries[-1] != len(layers):
        raise ValueError("boundaries must run from 0 to %d in %d steps" % (len(layers), num_stages))
    return [nn.Sequential(*layers[boundaries[s]: boundaries[s + 1]]) for s in range(num_stages)]


def reference_split(model, stem_relu: bool = True) -> List[nn.Sequential]:
    """The reference's 4-rank MobileNetV2 split: stem + blocks 0-2 | blocks 3-8 | blocks 9-14 | blocks 15-16 + head.
    ``stem_relu=False`` additionally reproduces its missing stem ReLU (defect Q2)."""
    seq = model.as_sequential(stem_relu=stem_relu)  # [stem, 17 blocks, head conv, Reshape1, linear] = 21 layers
    return split_sequential(seq, 4, boundaries=[0, 4, 10, 16, len(seq)])
