"""Single-device fused SGD over a flat parameter arena (``dmp_sgd_flat`` kernel).

Used where there is no cross-GPU gradient reduction to fuse with: single-GPU runs, each pipeline stage's private
optimizer (the reference builds one SGD per stage, /root/reference/code/distributed_training/
model_parallel.py:105,131,146) and DataParallel's device-0 update (data_parallel.py:90-91,111).

All parameters are moved into one fp32 master arena; ``param.grad`` becomes a view into a matching flat
gradient arena that the op layer accumulates into directly; a bf16 shadow arena feeds the bf16 kernels.
``step()`` is ONE kernel launch (update + shadow refresh + gradient zeroing) regardless of the number of
parameter tensors, and reads lr/momentum/weight-decay from device memory so it can live inside a CUDA graph
while LR schedulers keep mutating ``param_groups``.
"""
from __future__ import annotations

from typing import Iterable

import torch

from .. import _native

ALIGN = 8


class FlatFusedSGD(torch.optim.Optimizer):
    def __init__(self, params: Iterable[torch.nn.Parameter], lr: float = 0.1, momentum: float = 0.0, weight_decay: float = 0.0,
                 nesterov: bool = False, bf16_shadow: bool = True):
        params = [p for p in params if p.requires_grad]
        defaults = dict(lr=lr, momentum=momentum, weight_decay=weight_decay, nesterov=nesterov, dampening=0.0)
        super().__init__(params, defaults)
        if len(self.param_groups) != 1:
            raise ValueError("FlatFusedSGD supports a single parameter group")
        self._params = params
        dev = params[0].device
        self._cuda = dev.type == "cuda"
        offs, off = [], 0
        for p in params:
            if p.dtype != torch.float32:
                raise ValueError("FlatFusedSGD keeps fp32 master weights; got %s" % p.dtype)
            offs.append(off)
            off += (p.numel() + ALIGN - 1) // ALIGN * ALIGN
        self._offsets, self._numel = offs, off
        self.master = torch.zeros(off, dtype=torch.float32, device=dev)
        self.momentum_buf = torch.zeros(off, dtype=torch.float32, device=dev)
        self.grad = torch.zeros(off, dtype=torch.float32, device=dev)
        self.shadow = torch.zeros(off, dtype=torch.bfloat16, device=dev) if (bf16_shadow and self._cuda) else None
        for p, o in zip(params, offs):
            mv = self.master[o: o + p.numel()].view_as(p)
            mv.copy_(p.data)
            p.data = mv
            gv = self.grad[o: o + p.numel()].view_as(p)
            if p.grad is not None:
                gv.copy_(p.grad)
            p.grad = gv
            p._dmp_direct = True
            if self.shadow is not None:
                sv = self.shadow[o: o + p.numel()].view_as(p)
                sv.copy_(mv)
                p._dmp_shadow = sv
        self.hp = torch.tensor([lr, momentum, weight_decay, 1.0], dtype=torch.float32, device=dev)
        self._host = torch.empty(4, dtype=torch.float32)
        if self._cuda:
            self._host = self._host.pin_memory()
        self._pushed = (float(lr), float(momentum), float(weight_decay))
        self.kernel_launches = 0

    def push_hyperparameters(self):
        g = self.param_groups[0]
        cur = (float(g["lr"]), float(g["momentum"]), float(g["weight_decay"]))
        if cur != self._pushed:
            self._host[0], self._host[1], self._host[2], self._host[3] = cur[0], cur[1], cur[2], 1.0
            self.hp.copy_(self._host, non_blocking=True)
            self._pushed = cur

    def sync_shadow(self):
        if self.shadow is not None:
            self.shadow.copy_(self.master)

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if not torch.cuda.is_current_stream_capturing() if self._cuda else True:
            self.push_hyperparameters()
        # gradients that autograd re-pointed away from the arena (set_to_none, out-of-place accumulation) are folded back
        for p, o in zip(self._params, self._offsets):
            gv = self.grad[o: o + p.numel()].view_as(p)
            if p.grad is None:
                p.grad = gv
            elif p.grad.data_ptr() != gv.data_ptr():
                gv.add_(p.grad)
                p.grad = gv
        if self._cuda:
            lib = _native.kernels()
            _native.check(lib.dmp_sgd_flat(self.master.data_ptr(), _native.ptr(self.shadow), self.grad.data_ptr(), 0,
                                           self.momentum_buf.data_ptr(), self._numel, self.hp.data_ptr(),
                                           1 if self.param_groups[0]["nesterov"] else 0, 1,
                                           torch.cuda.current_stream().cuda_stream), "sgd_flat")
            self.kernel_launches += 1
        else:
            g = self.param_groups[0]
            d = self.grad + g["weight_decay"] * self.master
            self.momentum_buf.mul_(g["momentum"]).add_(d)
            upd = d + g["momentum"] * self.momentum_buf if g["nesterov"] else self.momentum_buf
            self.master.add_(upd, alpha=-g["lr"])
            self.grad.zero_()
        return loss

    def zero_grad(self, set_to_none: bool = False):
        pass  # step() leaves the gradient arena zeroed; gradients stay views into it

    def state_dict(self):
        state = {i: {"momentum_buffer": self.momentum_buf[o: o + p.numel()].view_as(p).clone()}
                 for i, (p, o) in enumerate(zip(self._params, self._offsets))}
        groups = [{k: v for k, v in self.param_groups[0].items() if k != "params"}]
        groups[0]["params"] = list(range(len(self._params)))
        return {"state": state, "param_groups": groups}

    def load_state_dict(self, sd):
        for i, (p, o) in enumerate(zip(self._params, self._offsets)):
            st = sd["state"].get(i) or sd["state"].get(str(i))
            if st is not None and st.get("momentum_buffer") is not None:
                self.momentum_buf[o: o + p.numel()].view_as(p).copy_(st["momentum_buffer"])
        for k, v in sd["param_groups"][0].items():
            if k != "params":
                self.param_groups[0][k] = v
        self.push_hyperparameters()
