"""SyncBatchNorm: batch statistics combined across all ranks of a process group.

Behavioural model: ``torch.nn.SyncBatchNorm`` (torch/nn/modules/batchnorm.py:615-905, _functions.py:7-209):
forward = local statistics -> cross-rank combine -> normalise; backward = local {sum dy, sum dy*xhat} ->
cross-rank sum -> dx.  The reference only mentions it ("prepare for sync batch norm layers",
/root/reference/Readme.md:151); BASELINE.json config 3 (ResNet-50 DDP) needs it.

On CUDA the combine is NOT a library collective and not even a separate launch: the BatchNorm kernel accumulates its
per-channel double sums in a peer-visible workspace, and between its statistics pass and its normalisation pass block 0
barriers with the peers (flag pads in peer memory) and peer-loads every rank's partial sums (csrc/comm/syncbn.cu,
``dmp_syncbn_fwd`` / ``dmp_syncbn_bwd``: ONE cooperative launch per layer and direction).  Shapes that cannot be launched
cooperatively, and ragged batches (``equal_batches=False``), use the split path: statistics kernel -> small fp64 one-shot
all-reduce kernel -> finalize -> apply.  CPU tensors (gloo tests) use ``torch.distributed.all_reduce`` inside a small
autograd Function.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F

from .. import _native


class SyncContext:
    """Cross-rank reduction service shared by all SyncBatchNorm layers of a model (one ordered stream of tiny
    collectives -> one communicator state)."""

    def __init__(self, process_group=None):
        self.group = process_group if process_group is not None else (dist.group.WORLD if dist.is_initialized() else None)
        self.world_size = dist.get_world_size(self.group) if self.group is not None else 1
        self.rank = dist.get_rank(self.group) if self.group is not None else 0
        self._symm = None
        self._state = None
        self._slots = {}
        self.kernel_launches = 0
        self.equal_batches = True

    def _slot(self, key, n, device):
        """Peer-visible double[n] slot + local result buffer for one (layer, direction)."""
        ent = self._slots.get(key)
        if ent is None or ent[0].local.numel() < n * 8:
            if self._symm is None:
                from .symm import ProcessSymmGroup
                self._symm = ProcessSymmGroup(self.group, device)
                self._state = self._symm.new_state()
            buf = self._symm.alloc(n * 8)
            ent = (buf, torch.empty(n, dtype=torch.float64, device=device))
            self._slots[key] = ent
        return ent

    def fused_workspace(self, key, C, device):
        """(peer-visible workspace, local totals buffer) of one (layer, direction) for the single-launch kernels; allocated
        collectively on first use (every rank meets its layers in the same order)."""
        ent = self._slots.get(("fused",) + tuple(key))
        if ent is None:
            if self._symm is None:
                from .symm import ProcessSymmGroup
                self._symm = ProcessSymmGroup(self.group, device)
                self._state = self._symm.new_state()
            buf = self._symm.alloc(int(_native.kernels().dmp_syncbn_ws_bytes(C)))
            buf.local.view(torch.uint8).zero_()
            torch.cuda.synchronize(device)
            dist.barrier(group=self.group)  # nobody may start accumulating into a peer-read workspace before all are zeroed
            ent = (buf, torch.empty(2 * C, dtype=torch.float64, device=device))
            self._slots[("fused",) + tuple(key)] = ent
        return ent

    def all_reduce_sum(self, key, vec: torch.Tensor) -> torch.Tensor:
        """Sum of the double vector ``vec`` over all ranks (same ``key`` order on every rank)."""
        if self.world_size == 1:
            return vec
        if not vec.is_cuda:
            out = vec.clone()
            dist.all_reduce(out, group=self.group)
            return out
        n = (vec.numel() + 1) // 2 * 2
        buf, out = self._slot(key, n, vec.device)
        sym = buf.view(torch.float64, n)
        sym[: vec.numel()].copy_(vec)
        lib = _native.kernels()
        _native.check(lib.dmp_allreduce_fused(0, 2, buf.ptrs, self._state.ptrs, self.rank, self.world_size, n, 1.0, out.data_ptr(),
                                              None, None, None, None, 0, None, None, 0, 1, torch.cuda.current_stream().cuda_stream),
                      "syncbn allreduce")
        self.kernel_launches += 1
        return out[: vec.numel()]

    # used by ops.functional._ConvBNAct
    def reduce_stats(self, key, accum: torch.Tensor, count: int):
        """Returns (global sums, global element count).  With ``equal_batches`` (default) every rank is assumed to hold the
        same number of rows, so the count is ``count * world_size`` without a device->host read; set it to False for
        ragged batches (costs one host sync per layer)."""
        if self.equal_batches:
            return self.all_reduce_sum((key, "f"), accum).contiguous(), float(count) * self.world_size
        v = torch.cat([accum, torch.tensor([float(count), 0.0], dtype=torch.float64, device=accum.device)])
        r = self.all_reduce_sum((key, "f"), v)
        return r[:-2].contiguous(), float(r[-2])

    def reduce_grads(self, key, accum: torch.Tensor) -> torch.Tensor:
        return self.all_reduce_sum((key, "b"), accum).contiguous()


class _SyncBNReference(torch.autograd.Function):
    """Plain-PyTorch SyncBN (CPU / oracle path)."""

    @staticmethod
    def forward(ctx, x, weight, bias, running_mean, running_var, eps, momentum, sync: SyncContext, key):
        C = x.shape[1]
        red = [0] + list(range(2, x.dim()))
        xd = x.double()
        local = torch.cat([xd.sum(red), (xd * xd).sum(red), torch.tensor([x.numel() / C], dtype=torch.float64, device=x.device)])
        tot = sync.all_reduce_sum((key, "f"), local)
        n = tot[-1]
        mean = tot[:C] / n
        var = (tot[C: 2 * C] / n - mean * mean).clamp_min(0)
        invstd = torch.rsqrt(var + eps)
        if running_mean is not None:
            with torch.no_grad():
                running_mean.mul_(1 - momentum).add_(momentum * mean.to(running_mean.dtype))
                running_var.mul_(1 - momentum).add_(momentum * (var * n / (n - 1).clamp_min(1)).to(running_var.dtype))
        shape = [1, C] + [1] * (x.dim() - 2)
        xhat = (xd - mean.view(shape)) * invstd.view(shape)
        ctx.save_for_backward(xhat.to(x.dtype), weight, invstd)
        ctx.sync, ctx.key, ctx.n = sync, key, n
        w = weight.double().view(shape) if weight is not None else 1.0
        b = bias.double().view(shape) if bias is not None else 0.0
        return (xhat * w + b).to(x.dtype)

    @staticmethod
    def backward(ctx, dy):
        xhat, weight, invstd = ctx.saved_tensors
        C = dy.shape[1]
        red = [0] + list(range(2, dy.dim()))
        shape = [1, C] + [1] * (dy.dim() - 2)
        dyd, xh = dy.double(), xhat.double()
        local = torch.cat([dyd.sum(red), (dyd * xh).sum(red)])
        tot = ctx.sync.all_reduce_sum((ctx.key, "b"), local)
        s1, s2 = tot[:C] / ctx.n, tot[C:] / ctx.n
        w = weight.double() if weight is not None else torch.ones(C, dtype=torch.float64, device=dy.device)
        dx = (w * invstd).view(shape) * (dyd - s1.view(shape) - xh * s2.view(shape))
        dgamma = local[C:].to(dy.dtype) if weight is not None else None
        dbeta = local[:C].to(dy.dtype) if weight is not None else None
        return dx.to(dy.dtype), dgamma, dbeta, None, None, None, None, None, None


class SyncBatchNorm(nn.BatchNorm2d):
    """Drop-in for ``nn.BatchNorm2d`` whose training statistics span the process group.  Inside the fused
    ``conv_bn_act`` op the cross-GPU combine happens between our statistics and apply kernels; called as a plain
    module it normalises its input the same way."""

    _next_key = 0

    def __init__(self, num_features, eps=1e-5, momentum=0.1, affine=True, track_running_stats=True, process_group=None,
                 sync_context: Optional[SyncContext] = None):
        super().__init__(num_features, eps, momentum, affine, track_running_stats)
        self.process_group = process_group
        self._dmp_sync = sync_context
        self._dmp_key = SyncBatchNorm._next_key
        SyncBatchNorm._next_key += 1

    def _context(self) -> SyncContext:
        if self._dmp_sync is None:
            self._dmp_sync = SyncContext(self.process_group)
        return self._dmp_sync

    def forward(self, x):
        sync = self._context()
        if not self.training or sync.world_size == 1:
            return super().forward(x)
        if x.is_cuda:
            from ..ops import functional as OF
            if OF._use_kernels(x):
                return OF.batch_norm_act(x, self, relu=False)
        if self.track_running_stats and self.num_batches_tracked is not None:
            self.num_batches_tracked.add_(1)
        mom = self.momentum if self.momentum is not None else 1.0 / float(self.num_batches_tracked)
        return _SyncBNReference.apply(x, self.weight, self.bias, self.running_mean if self.track_running_stats else None,
                                      self.running_var if self.track_running_stats else None, self.eps, mom, sync, self._dmp_key)

    @classmethod
    def convert_sync_batchnorm(cls, module: nn.Module, process_group=None, _ctx: Optional[SyncContext] = None) -> nn.Module:
        """Recursively replace BatchNorm layers with SyncBatchNorm sharing parameters, buffers and one SyncContext
        (torch: ``nn.SyncBatchNorm.convert_sync_batchnorm``, batchnorm.py:845)."""
        ctx = _ctx if _ctx is not None else SyncContext(process_group)
        out = module
        if isinstance(module, nn.modules.batchnorm._BatchNorm) and not isinstance(module, SyncBatchNorm):
            out = cls(module.num_features, module.eps, module.momentum, module.affine, module.track_running_stats, process_group, ctx)
            if module.affine:
                with torch.no_grad():
                    out.weight = module.weight
                    out.bias = module.bias
            out.running_mean = module.running_mean
            out.running_var = module.running_var
            out.num_batches_tracked = module.num_batches_tracked
            out.training = module.training
        for name, child in module.named_children():
            new = cls.convert_sync_batchnorm(child, process_group, ctx)
            if new is not child:
                setattr(out, name, new)
        return out


convert_sync_batchnorm = SyncBatchNorm.convert_sync_batchnorm
