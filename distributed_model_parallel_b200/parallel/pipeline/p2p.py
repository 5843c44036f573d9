"""Stage-boundary tensor transport over ``torch.distributed`` point-to-point (NCCL on GPU, gloo on CPU).

Replaces the reference's wire protocol (/root/reference/code/distributed_training/distributed_layers.py:7-62):
three blocking messages per tensor (ndim, size -- both as float32 -- then the payload), ``torch.rand`` receive
buffers, an ``input*1.0`` copy and two device->host syncs per tensor (SURVEY.md defects Q3/Q4).  Here:

* a tensor is preceded by at most ONE fixed-size int64 header {ndim, dtype, layout, d0..d5}; once a stage has seen a
  shape the header is skipped (static mode) and the receive can be posted before the data is needed;
* receive buffers are ``torch.empty``; channels-last activations travel as their physical NHWC bytes;
* ``SendForwardRecvBackward`` / ``RecvForwardSendBackward`` are the autograd-aware equivalents of the reference's
  ``ForwardSend_BackwardReceive`` / ``ForwardReceive_BackwardSend``; the micro-batch executor uses the same
  primitives batched with ``batch_isend_irecv`` so opposite-direction transfers between two stages cannot deadlock.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

HEADER_LEN = 9
_DTYPES = [torch.float32, torch.bfloat16, torch.float16, torch.int64, torch.int32, torch.uint8, torch.float64, torch.bool]
_DTYPE_CODE = {d: i for i, d in enumerate(_DTYPES)}


class TensorMeta:
    __slots__ = ("shape", "dtype", "channels_last")

    def __init__(self, shape, dtype, channels_last=False):
        self.shape, self.dtype, self.channels_last = tuple(int(s) for s in shape), dtype, bool(channels_last)

    @staticmethod
    def of(t: torch.Tensor) -> "TensorMeta":
        cl = t.dim() == 4 and not t.is_contiguous() and t.is_contiguous(memory_format=torch.channels_last)
        return TensorMeta(t.shape, t.dtype, cl)

    def with_batch(self, b: int) -> "TensorMeta":
        return TensorMeta((b,) + self.shape[1:], self.dtype, self.channels_last)

    def pack(self) -> torch.Tensor:
        if len(self.shape) > HEADER_LEN - 3:
            raise ValueError("tensors with more than %d dims are not supported on stage boundaries" % (HEADER_LEN - 3))
        h = [len(self.shape), _DTYPE_CODE[self.dtype], int(self.channels_last)] + list(self.shape)
        return torch.tensor(h + [0] * (HEADER_LEN - len(h)), dtype=torch.int64)

    @staticmethod
    def unpack(h: torch.Tensor) -> "TensorMeta":
        v = [int(x) for x in h.tolist()]
        return TensorMeta(v[3: 3 + v[0]], _DTYPES[v[1]], bool(v[2]))

    def empty(self, device) -> torch.Tensor:
        if self.channels_last:
            n, c, h, w = self.shape
            return torch.empty((n, h, w, c), dtype=self.dtype, device=device).permute(0, 3, 1, 2)
        return torch.empty(self.shape, dtype=self.dtype, device=device)

    def __eq__(self, o):
        return isinstance(o, TensorMeta) and (self.shape, self.dtype, self.channels_last) == (o.shape, o.dtype, o.channels_last)

    def __repr__(self):
        return "TensorMeta(%s, %s%s)" % (self.shape, self.dtype, ", channels_last" if self.channels_last else "")


def wire_view(t: torch.Tensor) -> torch.Tensor:
    """A contiguous view/copy of the bytes that go on the wire (NHWC for channels-last tensors)."""
    if t.dim() == 4 and not t.is_contiguous() and t.is_contiguous(memory_format=torch.channels_last):
        return t.permute(0, 2, 3, 1)
    return t.contiguous()


def _comm_device(group, like: torch.Tensor) -> torch.device:
    return like.device


def send_tensor(t: torch.Tensor, dst: int, group=None, header: bool = True) -> None:
    """Blocking send of one tensor (optionally preceded by its header)."""
    if header:
        dist.send(TensorMeta.of(t).pack().to(t.device), dst, group=group)
    dist.send(wire_view(t.detach()), dst, group=group)


def recv_tensor(src: int, device, group=None, meta: Optional[TensorMeta] = None) -> torch.Tensor:
    """Blocking receive; with ``meta=None`` the header is received first (one small message, one host sync)."""
    if meta is None:
        h = torch.empty(HEADER_LEN, dtype=torch.int64, device=device)
        dist.recv(h, src, group=group)
        meta = TensorMeta.unpack(h.cpu())
    buf = meta.empty(device)
    dist.recv(wire_view(buf), src, group=group)
    return buf


class SendForwardRecvBackward(torch.autograd.Function):
    """forward: send the activation to ``dst`` and pass it through; backward: the incoming grad_output is a
    placeholder -- the real gradient is received from ``dst``.  (reference: ForwardSend_BackwardReceive)"""

    @staticmethod
    def forward(ctx, x, dst, group, header):
        ctx.dst, ctx.group, ctx.header = dst, group, header
        ctx.meta = TensorMeta.of(x)
        send_tensor(x, dst, group, header)
        return x.view_as(x)

    @staticmethod
    def backward(ctx, grad_placeholder):
        g = recv_tensor(ctx.dst, grad_placeholder.device, ctx.group, None if ctx.header else ctx.meta)
        return g, None, None, None


class RecvForwardSendBackward(torch.autograd.Function):
    """forward: receive the activation from ``src`` into a leaf placeholder; backward: send grad_output back to
    ``src``.  (reference: generate_recv + ForwardReceive_BackwardSend, without the rand fill and the *1.0 copy)"""

    @staticmethod
    def forward(ctx, placeholder, src, group, meta, header):
        ctx.src, ctx.group, ctx.header = src, group, header
        return recv_tensor(src, placeholder.device, group, None if header else meta)

    @staticmethod
    def backward(ctx, grad_output):
        send_tensor(grad_output, ctx.src, ctx.group, ctx.header)
        return None, None, None, None, None


def recv_forward(src: int, device, group=None, meta: Optional[TensorMeta] = None, requires_grad: bool = True) -> torch.Tensor:
    """Autograd-connected receive: gradients flowing into the result are sent back to ``src`` during backward."""
    ph = torch.zeros((), device=device, requires_grad=requires_grad)
    return RecvForwardSendBackward.apply(ph, src, group, meta, meta is None)


def send_forward(x: torch.Tensor, dst: int, group=None, header: bool = True) -> torch.Tensor:
    return SendForwardRecvBackward.apply(x, dst, group, header)


def backward_placeholder(y: torch.Tensor) -> torch.Tensor:
    """Zero-storage stand-in for the reference's ``output.clone().detach()`` dummy gradient."""
    return torch.zeros((), dtype=y.dtype, device=y.device).expand_as(y)


# ------------------------------------------------------------------------------------------------ batched p2p
def batched(ops: Sequence[Tuple[str, torch.Tensor, int]], group=None) -> List:
    """Issue sends/receives as one group (``("send"|"recv", tensor, peer)``) and return the work handles.
    Grouping makes opposite-direction transfers between the same two stages deadlock-free."""
    if not ops:
        return []
    p2p = [dist.P2POp(dist.isend if kind == "send" else dist.irecv, t, peer, group) for kind, t, peer in ops]
    return dist.batch_isend_irecv(p2p)


def wait_all(works) -> None:
    for w in works:
        w.wait()
