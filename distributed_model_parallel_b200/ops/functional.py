"""Functional op layer: every hot op of the training step, with two implementations each.

* CUDA tensors run the hand-written sm_100a kernels of ``libdmp_b200.so`` (tcgen05 GEMM for 1x1
  convolutions / linears, depthwise stencil, fused BatchNorm statistics + apply(+residual,+ReLU), fused
  cross-entropy).  Activations are bf16 ``channels_last`` tensors (logical NCHW, physical NHWC), weights are
  fp32 masters with an optional bf16 shadow maintained by the fused optimizer (``param._dmp_shadow``).
* CPU tensors run a plain PyTorch restatement -- the oracle of the numerics tests and the path the gloo/CPU
  plumbing tests exercise.

There is no silent fallback on a GPU: if the native library cannot be loaded the CUDA path raises.
The ops mirror what the reference model calls (/root/reference/code/distributed_training/model/
mobilenetv2.py:31-36,68-76: conv -> bn -> relu chains, residual add, avg_pool2d, view, linear) and its
loss (utils.py:58, data_parallel.py:109).
"""
from __future__ import annotations

import os
import threading
from collections import Counter
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import _native

# which implementation served each op call (reported by bench.py / tests; keys like "conv_pw:tcgen05")
path_counts: Counter = Counter()
_tls = threading.local()

# Callbacks fired when an op has accumulated a gradient straight into ``param.grad`` (bypassing
# autograd's AccumulateGrad).  DistributedDataParallel registers its Reducer here.
grad_ready_callbacks: list = []


_ref_lock = threading.Lock()
_ref_depth = 0  # > 0 while some thread is inside use_reference_ops()


class use_reference_ops:
    """Context manager: run the plain PyTorch restatement even for CUDA tensors (numerics oracle / A-B timing).  The switch is
    process-wide on purpose -- DataParallel's replica threads, started inside the context, must see it -- so it is a counter
    under a lock rather than a saved-and-restored flag: contexts entered and left by different threads in any order cannot leave
    it stuck on or off."""

    def __enter__(self):
        global _ref_depth
        with _ref_lock:
            _ref_depth += 1
        return self

    def __exit__(self, *exc):
        global _ref_depth
        with _ref_lock:
            _ref_depth -= 1
        return False


def _use_kernels(x: torch.Tensor) -> bool:
    return x.is_cuda and _ref_depth == 0


def _note(key: str) -> None:
    path_counts[key] += 1


def _sp() -> int:
    return torch.cuda.current_stream().cuda_stream


def _k():
    return _native.kernels()


def _ck(rc, what):
    if rc != 0:
        _native.check(rc, what)


def _cl_empty(n, c, h, w, device, dtype=torch.bfloat16):
    return torch.empty((n, c, h, w), device=device, dtype=dtype, memory_format=torch.channels_last)


def _as_cl_bf16(x: torch.Tensor) -> torch.Tensor:
    if x.dtype != torch.bfloat16:
        x = x.to(torch.bfloat16)
    if x.dim() == 4 and not x.is_contiguous(memory_format=torch.channels_last):
        x = x.contiguous(memory_format=torch.channels_last)
    return x


def weight_bf16(w: torch.Tensor) -> torch.Tensor:
    """bf16 view of a weight: the optimizer-maintained shadow when present, else a cast."""
    sh = getattr(w, "_dmp_shadow", None)
    if sh is not None:
        return sh
    return w.detach().to(torch.bfloat16)


class _SideStreams:
    """One side stream per device for weight-gradient kernels, joined back into the stream that ran backward by a callback
    queued on the autograd engine (so ``.grad`` is complete when ``backward()`` returns, also inside a CUDA-graph capture)."""

    NSIDE = max(1, int(os.environ.get("DMP_WGRAD_STREAMS", "2")))  # side streams per device, used round-robin

    def __init__(self):
        self.streams = {}
        self.pending = {}      # device index -> [main stream, side streams...] with unjoined work
        self._queued = False
        self._rr = 0

    def side(self, dev):
        sts = self.streams.get(dev.index)
        if sts is None:
            sts = [torch.cuda.Stream(device=dev) for _ in range(self.NSIDE)]
            self.streams[dev.index] = sts
        self._rr = (self._rr + 1) % len(sts)
        return sts[self._rr]

    def fork(self, dev):
        """A side stream ordered after everything issued so far on the current stream of ``dev``."""
        main = torch.cuda.current_stream(dev)
        side = self.side(dev)
        if main.cuda_stream == side.cuda_stream:
            return None
        side.wait_stream(main)
        if dev.index not in self.pending or not self._queued:
            self.pending[dev.index] = [main] + list(self.streams[dev.index])
            if not self._queued:
                try:
                    torch.autograd.Variable._execution_engine.queue_callback(self.join_all)
                    self._queued = True
                except RuntimeError:  # not inside a backward pass (direct call): the caller joins
                    pass
        return side

    def join_all(self):
        self._queued = False
        for idx, sts in self.pending.items():
            main = sts[0]
            cur = torch.cuda.current_stream(torch.device("cuda", idx))
            for side in sts[1:]:
                main.wait_stream(side)
                if cur.cuda_stream != main.cuda_stream:  # backward() was called from another stream than the forward ran on
                    cur.wait_stream(side)
        self.pending.clear()

    def gradient_streams(self, dev):
        """Every stream that may still be writing gradients on ``dev`` in the running backward pass (the stream backward runs
        on and the side streams): a collective launched from a gradient-ready callback must order itself after all of them."""
        return list(self.pending.get(dev.index, ()))


_tls_side = threading.local()
_all_side = []          # every manager ever created (one per thread that ran a backward with side-stream work)
_all_side_lock = threading.Lock()


def side_streams() -> _SideStreams:
    # one manager per Python thread: DataParallel replicas run backward on one autograd thread per device
    m = getattr(_tls_side, "m", None)
    if m is None:
        m = _tls_side.m = _SideStreams()
        with _all_side_lock:
            _all_side.append(m)
    return m


def all_side_stream_managers():
    """Every thread's side-stream manager: a caller that consumes gradients written by several autograd threads (DataParallel's
    gradient reduction) joins them all first."""
    with _all_side_lock:
        return list(_all_side)


def _deliver_grad(param: Optional[torch.Tensor], writer, side_inputs=()):
    """Produce the gradient of ``param``.

    ``side_inputs``: the tensors the writer's kernels read.  When given (and the side stream is enabled) the kernels are launched
    on the per-device side stream, forked from the current stream, so they overlap with the data-gradient chain.

    ``writer(dst, accumulate)`` launches the kernel(s) writing the fp32 gradient into ``dst``.
    If ``param.grad`` is a pre-allocated fp32 buffer flagged ``_dmp_direct`` (a Reducer bucket view), the
    kernel accumulates straight into it, grad-ready callbacks fire and ``None`` is returned to autograd;
    otherwise a fresh tensor is returned and autograd accumulates as usual.
    """
    if param is None or not param.requires_grad:
        return None
    g = param.grad
    if g is not None and getattr(param, "_dmp_direct", False) and g.dtype == torch.float32 and g.is_contiguous():
        side = side_streams().fork(g.device) if (_WGRAD_SIDE and side_inputs and g.is_cuda) else None
        if side is None:
            writer(g, True)
            for cb in grad_ready_callbacks:
                cb(param)
            return None
        with torch.cuda.stream(side):
            for t in side_inputs:
                if t is not None:
                    t.record_stream(side)  # the caching allocator must not hand these out again before the side work is done
            writer(g, True)
            for cb in grad_ready_callbacks:  # a bucket launched from here orders itself after the side stream
                cb(param)
        return None
    dst = torch.empty_like(param, dtype=torch.float32, memory_format=torch.contiguous_format)
    writer(dst, False)
    return dst if param.dtype == torch.float32 else dst.to(param.dtype)


# ----------------------------------------------------------------------------------------------------
# convolution kinds
# ----------------------------------------------------------------------------------------------------
def conv_kind(conv: nn.Conv2d, x: torch.Tensor) -> str:
    kh, kw = conv.kernel_size
    cin, cout, g = conv.in_channels, conv.out_channels, conv.groups
    if conv.bias is not None or conv.dilation != (1, 1) or conv.padding_mode != "zeros":
        return "generic"
    if (kh, kw) == (1, 1) and g == 1 and conv.padding == (0, 0) and cin % 8 == 0 and cout % 8 == 0:
        return "pw"
    if (kh, kw) == (3, 3) and g == cin == cout and conv.padding == (1, 1) and conv.stride in ((1, 1), (2, 2)) \
            and cin % 8 == 0:
        return "dw"
    if g == 1 and kh == kw and kh % 2 == 1 and kh > 1 and conv.padding == (kh // 2, kh // 2) and cin % 64 == 0 and cout % 8 == 0 \
            and kh * kw <= 49:
        if conv.stride == (1, 1):
            return "igemm"
        # stride 2: TMA element strides forward / weight gradient, parity-class data gradient (needs even H, W and Cout % 64 == 0)
        if conv.stride == (2, 2) and x.shape[2] % 2 == 0 and x.shape[3] % 2 == 0 and cout % 64 == 0:
            return "igemm"
    if g == 1 and cin <= 4 and cout % 8 == 0 and conv.stride[0] == conv.stride[1] and conv.padding[0] == conv.padding[1] \
            and cin * kh * kw >= _STEM_GEMM_MIN_TAPS and _STEM_GEMM:
        return "stem_gemm"  # large stems (7x7): im2col rows + tcgen05 GEMMs
    if g == 1 and cin <= 4 and cout % 8 == 0 and conv.stride[0] == conv.stride[1] and conv.padding[0] == conv.padding[1] \
            and cin * kh * kw * cout * 4 <= 48 * 1024:
        return "stem"
    return "generic"


def _pack_filter(w, cout, cin, kh, kw, mode):
    """bf16 operand of the implicit-GEMM kernels from the [Cout][Cin][KH][KW] parameter (its bf16 shadow when the optimizer keeps
    one), in one launch: mode 0 [Cout][kh][kw][Cin]; 1 [Cin][kh][kw][Cout]; 2 the same with flipped taps."""
    src = getattr(w, "_dmp_shadow", None)
    if src is None:
        src = w.detach()
    if not src.is_contiguous():
        src = src.contiguous()
    rows = cout if mode == 0 else cin
    out = torch.empty((rows, kh * kw * (cin if mode == 0 else cout)), device=w.device, dtype=torch.bfloat16)
    _ck(_k().dmp_pack_filter(src.data_ptr(), 1 if src.dtype == torch.bfloat16 else 0, out.data_ptr(), cout, cin, kh, kw, mode, _sp()),
        "filter pack")
    return out


def _conv_fwd_cuda(x, w, conv, kind, stats=None):
    """``stats``: optional {double[2C]; ticket} buffer -- the producer then also accumulates the BatchNorm batch sums of its
    output (only the "pw" and "dw" kinds have such an epilogue)."""
    lib = _k()
    sp = _sp()
    N = x.shape[0]
    if kind == "pw":
        if conv.stride != (1, 1):
            x = x[:, :, ::conv.stride[0], ::conv.stride[1]].contiguous(memory_format=torch.channels_last)
        _, cin, H, W = x.shape
        cout = conv.out_channels
        y = _cl_empty(N, cout, H, W, x.device)
        wb = weight_bf16(w)
        if stats is not None:
            _ck(lib.dmp_gemm_bf16_colstats(x.data_ptr(), cin, wb.data_ptr(), cin, y.data_ptr(), cout, N * H * W, cout, cin,
                                           stats.data_ptr(), sp), "conv1x1 fwd gemm + stats")
            _note("conv_pw:tcgen05+bnstats")
        else:
            _ck(lib.dmp_gemm_bf16(0, x.data_ptr(), cin, wb.data_ptr(), cin, y.data_ptr(), cout, N * H * W, cout, cin, None, 0,
                                  1, sp), "conv1x1 fwd gemm")
            _note("conv_pw:tcgen05")
        return y, x
    if kind == "dw":
        _, C, H, W = x.shape
        s = conv.stride[0]
        Ho, Wo = (H + 2 - 3) // s + 1, (W + 2 - 3) // s + 1
        y = _cl_empty(N, C, Ho, Wo, x.device)
        wb = weight_bf16(w)
        if stats is not None:
            _ck(lib.dmp_dwconv3x3_fwd_colstats(x.data_ptr(), wb.data_ptr(), 1, y.data_ptr(), N, H, W, C, s, stats.data_ptr(), sp),
                "dwconv fwd + stats")
            _note("conv_dw:stencil+bnstats")
        else:
            _ck(lib.dmp_dwconv3x3_fwd(x.data_ptr(), wb.data_ptr(), 1, y.data_ptr(), N, H, W, C, s, sp), "dwconv fwd")
            _note("conv_dw:stencil")
        return y, x
    if kind == "stem":
        _, cin, H, W = x.shape
        kh, kw = conv.kernel_size
        s, p = conv.stride[0], conv.padding[0]
        Ho, Wo = (H + 2 * p - kh) // s + 1, (W + 2 * p - kw) // s + 1
        if x.dtype == torch.float32 and x.is_contiguous():
            layout = 0
        else:
            x = _as_cl_bf16(x)
            layout = 1
        y = _cl_empty(N, conv.out_channels, Ho, Wo, x.device)
        _ck(lib.dmp_stem_conv_fwd(x.data_ptr(), layout, w.data_ptr(), 0, y.data_ptr(), N, cin, H, W, conv.out_channels, kh, kw,
                                  s, p, sp), "stem conv fwd")
        _note("conv_stem:direct")
        return y, x
    if kind == "stem_gemm":
        _, cin, H, W = x.shape
        kh, kw = conv.kernel_size
        s, p = conv.stride[0], conv.padding[0]
        cout = conv.out_channels
        Ho, Wo = (H + 2 * p - kh) // s + 1, (W + 2 * p - kw) // s + 1
        if x.dtype == torch.float32 and x.is_contiguous():
            layout = 0
        else:
            x = _as_cl_bf16(x)
            layout = 1
        T = cin * kh * kw
        Kp = (T + 7) // 8 * 8
        npix = N * Ho * Wo
        cols = torch.empty((npix, Kp), device=x.device, dtype=torch.bfloat16)
        _ck(lib.dmp_stem_im2col(x.data_ptr(), layout, cols.data_ptr(), N, cin, H, W, kh, kw, s, p, Kp, sp), "stem im2col")
        wpad = torch.empty((cout, Kp), device=x.device, dtype=torch.bfloat16)
        _ck(lib.dmp_pack_rows_bf16(w.data_ptr(), 0 if w.dtype == torch.float32 else 1, wpad.data_ptr(), cout, T, Kp, sp), "stem filter pack")
        y = _cl_empty(N, cout, Ho, Wo, x.device)
        _ck(lib.dmp_gemm_bf16(0, cols.data_ptr(), Kp, wpad.data_ptr(), Kp, y.data_ptr(), cout, npix, cout, T, None, 0, 1, sp), "stem gemm fwd")
        _note("conv_stem:im2col+tcgen05")
        return y, cols  # the weight gradient re-uses the patch rows
    if kind == "igemm":
        _, cin, H, W = x.shape
        kh, kw = conv.kernel_size
        cout = conv.out_channels
        wp = _pack_filter(w, cout, cin, kh, kw, 0)  # [Cout][kh][kw][Cin]
        st, pd = conv.stride[0], conv.padding[0]
        y = _cl_empty(N, cout, (H + 2 * pd - kh) // st + 1, (W + 2 * pd - kw) // st + 1, x.device)
        _ck(lib.dmp_conv_igemm_fwd_strided(x.data_ptr(), wp.data_ptr(), None, y.data_ptr(), N, H, W, cin, cout, kh, kw, pd, st, sp),
            "conv igemm fwd")
        _note("conv_dense:tcgen05_igemm")
        return y, x
    # generic: library convolution (strided dense convolutions, 7x7 stem of ResNet-50)
    x = _as_cl_bf16(x)
    y = F.conv2d(x, weight_bf16(w), None, conv.stride, conv.padding, conv.dilation, conv.groups)
    _note("conv_generic:cudnn")
    return y.contiguous(memory_format=torch.channels_last), x


def _link_usable(link, x) -> bool:
    """Can this convolution's data-gradient kernel take over the BatchNorm-backward reductions of the op that produced x?"""
    return (link is not None and _FUSE_BN_BWD and link.consumers == 1 and link.y_raw is not None
            and link.y_raw.shape == x.shape and x.shape[1] % 8 == 0)


def _conv_bwd_cuda(dy, x, w, conv, kind, need_dx, link_in=None):
    """dy: [N,Cout,Ho,Wo] channels_last bf16. Returns (dx or None, dw_writer).  With a usable ``link_in`` the returned dx is
    the ReLU-masked gradient g and the producer's BatchNorm reductions have been accumulated (link_in.reduced)."""
    lib = _k()
    sp = _sp()
    N = x.shape[0]
    dx = None
    if kind == "pw":
        _, cin, H, W = x.shape  # x is the (possibly strided-subsampled) GEMM input saved by forward
        cout = conv.out_channels
        M = N * H * W
        if need_dx:
            wb = weight_bf16(w)
            dxs = _cl_empty(N, cin, H, W, x.device)
            if conv.stride == (1, 1) and _link_usable(link_in, x):
                L = link_in
                _ck(lib.dmp_gemm_bf16_dgrad_bnbwd(dy.data_ptr(), cout, wb.data_ptr(), cin, dxs.data_ptr(), cin, M, cin, cout,
                                                  L.y_raw.data_ptr(), cin, L.mean.data_ptr(), L.invstd.data_ptr(),
                                                  _native.ptr(L.gamma), _native.ptr(L.beta), 1 if L.relu else 0,
                                                  L.bsums.data_ptr(), sp), "conv1x1 dgrad gemm + bn-bwd reductions")
                L.reduced = True
                _note("conv_pw:dgrad+bnbwd")
            else:
                _ck(lib.dmp_gemm_bf16(1, dy.data_ptr(), cout, wb.data_ptr(), cin, dxs.data_ptr(), cin, M, cin, cout, None, 0, 1,
                                      sp), "conv1x1 dgrad gemm")
            if conv.stride != (1, 1):
                s0, s1 = conv.stride
                Hin, Win = conv._dmp_in_hw
                dx = torch.zeros((N, cin, Hin, Win), device=x.device, dtype=torch.bfloat16).contiguous(
                    memory_format=torch.channels_last)
                dx[:, :, ::s0, ::s1] = dxs
            else:
                dx = dxs

        def write_dw(dst, accumulate):
            if not accumulate:
                dst.zero_()
            _ck(lib.dmp_gemm_bf16(2, dy.data_ptr(), cout, x.data_ptr(), cin, dst.data_ptr(), cin, cout, cin, M, None, 2, 0,
                                  _sp()), "conv1x1 wgrad gemm")

        return dx, write_dw
    if kind == "dw":
        _, C, H, W = x.shape
        s = conv.stride[0]
        if need_dx:
            wb = weight_bf16(w)
            dx = _cl_empty(N, C, H, W, x.device)
            _ck(lib.dmp_dwconv3x3_dgrad(dy.data_ptr(), wb.data_ptr(), 1, dx.data_ptr(), N, H, W, C, s, sp), "dwconv dgrad")

        def write_dw(dst, accumulate):
            _ck(lib.dmp_dwconv3x3_wgrad(dy.data_ptr(), x.data_ptr(), dst.data_ptr(), N, H, W, C, s, 1 if accumulate else 0, _sp()),
                "dwconv wgrad")

        return dx, write_dw
    if kind == "stem":
        _, cin, H, W = x.shape
        kh, kw = conv.kernel_size
        s, p = conv.stride[0], conv.padding[0]
        layout = 0 if x.dtype == torch.float32 else 1
        if need_dx:
            raise RuntimeError("stem convolution does not provide an input gradient")

        def write_dw(dst, accumulate):
            _ck(lib.dmp_stem_conv_wgrad(dy.data_ptr(), x.data_ptr(), layout, dst.data_ptr(), N, cin, H, W, conv.out_channels,
                                        kh, kw, s, p, 1 if accumulate else 0, _sp()), "stem conv wgrad")

        return None, write_dw
    if kind == "stem_gemm":
        if need_dx:
            raise RuntimeError("stem convolution does not provide an input gradient")
        cols = x  # [pixels, Kp] patch rows saved by the forward
        cout = conv.out_channels
        T = conv.in_channels * conv.kernel_size[0] * conv.kernel_size[1]
        npix, Kp = cols.shape
        dy2 = dy if dy.is_contiguous(memory_format=torch.channels_last) else dy.contiguous(memory_format=torch.channels_last)

        def write_dw(dst, accumulate):
            if not accumulate:
                dst.zero_()
            _ck(lib.dmp_gemm_bf16(2, dy2.data_ptr(), cout, cols.data_ptr(), Kp, dst.data_ptr(), T, cout, T, npix, None, 2, 0, _sp()),
                "stem gemm wgrad")

        return None, write_dw
    if kind == "igemm":
        _, cin, H, W = x.shape
        kh, kw = conv.kernel_size
        cout = conv.out_channels
        st = conv.stride[0]
        if need_dx:
            if st == 2:
                # four output-parity implicit GEMMs over dy, each with the taps that reach that parity (no inserted zeros)
                wd = _pack_filter(w, cout, cin, kh, kw, 1)
                dx = _cl_empty(N, cin, H, W, x.device)
                _ck(lib.dmp_conv_igemm_dgrad_s2(dy.data_ptr(), wd.data_ptr(), dx.data_ptr(), N, H, W, cin, cout, kh, kw,
                                                conv.padding[0], sp), "conv igemm dgrad (stride 2)")
            elif cout % 64 == 0:
                # dx = conv(dy, flipped filter with in/out channels swapped): the same implicit-GEMM kernel
                wd = _pack_filter(w, cout, cin, kh, kw, 2)
                dx = _cl_empty(N, cin, H, W, x.device)
                _ck(lib.dmp_conv_igemm_fwd(dy.data_ptr(), wd.data_ptr(), None, dx.data_ptr(), N, H, W, cout, cin, kh, kw,
                                           conv.padding[0], sp), "conv igemm dgrad")
            else:
                dx = torch.ops.aten.convolution_backward(dy, x, weight_bf16(w), None, conv.stride, conv.padding, conv.dilation, False,
                                                         (0, 0), conv.groups, (True, False, False))[0]

        def write_dw(dst, accumulate):
            # the epilogue accumulates straight into the [Cout][Cin][KH][KW] gradient: no temporary, no permute, no add
            if not accumulate:
                dst.zero_()
            _ck(lib.dmp_conv_igemm_wgrad_strided(dy.data_ptr(), x.data_ptr(), dst.data_ptr(), N, H, W, cin, cout, kh, kw,
                                                 conv.padding[0], st, 1, _sp()), "conv igemm wgrad")

        return dx, write_dw
    # generic
    wb = weight_bf16(w)
    gi, gw, _ = torch.ops.aten.convolution_backward(dy, x, wb, None, conv.stride, conv.padding, conv.dilation, False, (0, 0),
                                                    conv.groups, (need_dx, True, False))

    def write_dw(dst, accumulate):
        if accumulate:
            dst.add_(gw.to(torch.float32))
        else:
            dst.copy_(gw)

    return gi, write_dw


# ----------------------------------------------------------------------------------------------------
# fused conv -> (BatchNorm) -> (+residual) -> (ReLU)
# ----------------------------------------------------------------------------------------------------
def _bn_workspace(bn, which: str, C: int, dev):
    """Per-layer, per-device zero-initialised workspace of the fused BatchNorm kernels (double accum[2][2C] + parity)."""
    store = bn.__dict__.setdefault("_dmp_ws", {})
    key = (dev.index, which)
    ws = store.get(key)
    need = int(_k().dmp_bn_fused_ws_bytes(C)) // 8
    if ws is None or ws.numel() < need:
        ws = torch.zeros(need, dtype=torch.float64, device=dev)
        store[key] = ws
    return ws


# BatchNorm batch statistics from the producing convolution's epilogue (DMP_BN_EPILOGUE_STATS=0 disables): the GEMM reads its
# staged bf16 output tile column-wise out of shared memory, the depthwise stencil reduces per block; the normalisation is then
# one elementwise pass (dmp_bn_apply_sums) instead of the cooperative statistics -> grid barrier -> apply kernel.
_EPILOGUE_STATS = os.environ.get("DMP_BN_EPILOGUE_STATS", "1") == "1"
# Opt-in fusions (numerically verified, tests/test_gpu_fused_chain.py; measured slower than the separate kernels at batch 128 on
# B200, profiles/README.md): DMP_FUSE_BN_CHAIN=1 runs expand -> bn1 -> relu -> depthwise -> bn2 -> relu as one chain in which
# relu(bn1(.)) is applied while the stencil loads its input (the register-bound stencil pays more for the extra arithmetic than
# the saved elementwise pass costs); DMP_FUSE_BN_BWD=1 lets the consumer's data-gradient kernel produce a BatchNorm's backward
# reductions (the GEMM epilogue warps, one per scheduler, become the kernel's bound).
_FUSE_CHAIN = os.environ.get("DMP_FUSE_BN_CHAIN", "0") == "1"
_FUSE_BN_BWD = os.environ.get("DMP_FUSE_BN_BWD", "0") == "1"
# Weight-gradient kernels run on a side stream (DMP_WGRAD_SIDE_STREAM=0 disables): the data-gradient chain is the critical
# path of backward, every kernel of a batch-128 step is latency- rather than throughput-bound, and the weight gradient of a
# layer depends on nothing that follows it -- so it overlaps with the next BatchNorm backward / data gradient.
_WGRAD_SIDE = os.environ.get("DMP_WGRAD_SIDE_STREAM", "1") == "1"
# Stems with more than 27 taps (ResNet's 7x7) run as im2col rows + tcgen05 GEMMs (DMP_STEM_GEMM=0: the per-tap kernels).
_STEM_GEMM = os.environ.get("DMP_STEM_GEMM", "1") == "1"
# Cin*KH*KW from which a stem takes the GEMM path: 27 = also the 3x3 RGB stems (MobileNetV2 step 3.33 -> 3.29 ms); the per-tap
# kernels (fp32 input, no bf16 rounding of the image) remain for smaller stems and behind DMP_STEM_GEMM_MIN_TAPS=28.
_STEM_GEMM_MIN_TAPS = int(os.environ.get("DMP_STEM_GEMM_MIN_TAPS", "27"))


class _BnLink:
    """Hand-over between an op that ends in act(BatchNorm(y_raw)) and the single convolution that consumes its output.

    The consumer's data-gradient kernel can apply the ReLU mask and accumulate the BatchNorm backward's two reductions
    (sum g, sum g*xhat) in its own epilogue; it then sets ``reduced`` and the producer's backward is a single elementwise pass.
    Only valid when the activation has exactly one consumer (``consumers`` is counted at forward time and checked at backward
    time); models declare that with ``conv_bn_act(..., single_consumer=True)``."""
    __slots__ = ("y_raw", "mean", "invstd", "gamma", "beta", "relu", "bsums", "reduced", "consumers", "count")

    def __init__(self):
        self.y_raw = self.mean = self.invstd = self.gamma = self.beta = self.bsums = None
        self.relu, self.reduced, self.consumers, self.count = False, False, 0, 0.0


def _bn_epilogue_stats_ok(bn, sync, kind) -> bool:
    """Can the convolution's epilogue produce this BatchNorm's batch statistics (train mode, local statistics, a producer
    kind that has the epilogue, plain momentum)?"""
    if not _EPILOGUE_STATS or bn is None or sync is not None or kind not in ("pw", "dw"):
        return False
    training = bn.training or (bn.running_mean is None)
    return training and (bn.momentum is not None or not bn.track_running_stats)


def _bn_sums_buffer(bn, C: int, dev, which: str = "sums"):
    """Per-layer, per-device {double sums[2C]; int ticket}: zero on creation, re-zeroed by its consumer kernel after each use
    (``sums``: forward batch sums, ``bsums``: backward reductions)."""
    store = bn.__dict__.setdefault("_dmp_ws", {})
    key = (dev.index, which)
    buf = store.get(key)
    need = int(_k().dmp_bn_sums_bytes(C)) // 8
    if buf is None or buf.numel() != need:
        buf = torch.zeros(need, dtype=torch.float64, device=dev)
        store[key] = buf
    return buf


def _bn_forward_cuda(y_raw, gamma, beta, residual, bn, relu, sync, sums=None):
    """Train/eval BatchNorm (+residual, +ReLU) of a channels-last bf16 tensor. Returns (out, mean, invstd, count, training).
    ``sums``: batch sums already accumulated by the producing convolution (see _bn_epilogue_stats_ok)."""
    lib = _k()
    sp = _sp()
    N, C, Ho, Wo = y_raw.shape
    M = N * Ho * Wo
    dev = y_raw.device
    training = bn.training or (bn.running_mean is None)
    if sums is not None:
        mean = torch.empty(C, device=dev, dtype=torch.float32)
        invstd = torch.empty(C, device=dev, dtype=torch.float32)
        out = torch.empty_like(y_raw)
        track = bn.track_running_stats and bn.running_mean is not None
        _ck(lib.dmp_bn_apply_sums(y_raw.data_ptr(), 1, _native.ptr(residual), out.data_ptr(), M, C, _native.ptr(gamma), _native.ptr(beta),
                                  _native.ptr(bn.running_mean if track else None), _native.ptr(bn.running_var if track else None),
                                  _native.ptr(bn.num_batches_tracked if track else None), float(bn.momentum or 0.1), float(bn.eps),
                                  mean.data_ptr(), invstd.data_ptr(), sums.data_ptr(), 1 if relu else 0, sp), "bn apply (epilogue sums)")
        _note("bn:apply_from_epilogue_sums")
        return out, mean, invstd, float(M), True
    if training and sync is None and (bn.momentum is not None or not bn.track_running_stats):
        # single cooperative launch: statistics -> grid barrier -> normalise (+residual, +ReLU), running stats, counter
        mean = torch.empty(C, device=dev, dtype=torch.float32)
        invstd = torch.empty(C, device=dev, dtype=torch.float32)
        out = torch.empty_like(y_raw)
        track = bn.track_running_stats
        rc = lib.dmp_bn_fwd_fused(y_raw.data_ptr(), 1, _native.ptr(residual), out.data_ptr(), M, C, _native.ptr(gamma), _native.ptr(beta),
                                  _native.ptr(bn.running_mean if track else None), _native.ptr(bn.running_var if track else None),
                                  _native.ptr(bn.num_batches_tracked if track else None), float(bn.momentum or 0.1), float(bn.eps),
                                  mean.data_ptr(), invstd.data_ptr(), _bn_workspace(bn, "f", C, dev).data_ptr(), 1 if relu else 0, sp)
        if rc == 0:
            _note("bn:fused1")
            return out, mean, invstd, float(M), True
        if rc != 1:
            _ck(rc, "bn fwd fused")
    if training and sync is not None and sync.equal_batches and (bn.momentum is not None or not bn.track_running_stats):
        # SyncBatchNorm in ONE launch: statistics -> grid barrier -> cross-GPU flag barrier + peer loads of all partial sums ->
        # grid barrier -> normalise (+residual, +ReLU)
        ws, total = sync.fused_workspace((bn._dmp_key, "f"), C, dev)
        mean = torch.empty(C, device=dev, dtype=torch.float32)
        invstd = torch.empty(C, device=dev, dtype=torch.float32)
        out = torch.empty_like(y_raw)
        track = bn.track_running_stats
        rc = lib.dmp_syncbn_fwd(y_raw.data_ptr(), _native.ptr(residual), out.data_ptr(), M, C, _native.ptr(gamma), _native.ptr(beta),
                                _native.ptr(bn.running_mean if track else None), _native.ptr(bn.running_var if track else None),
                                _native.ptr(bn.num_batches_tracked if track else None), float(bn.momentum or 0.1), float(bn.eps),
                                mean.data_ptr(), invstd.data_ptr(), ws.ptrs, sync._state.ptrs, sync.rank, sync.world_size,
                                total.data_ptr(), 1 if relu else 0, sp)
        if rc == 0:
            sync.kernel_launches += 1
            _note("bn:syncbn_fused1")
            return out, mean, invstd, float(M) * sync.world_size, True
        if rc != 1:
            _ck(rc, "syncbn fwd fused")
    if training:
        accum = torch.empty(2 * C, device=dev, dtype=torch.float64)
        mean = torch.empty(C, device=dev, dtype=torch.float32)
        invstd = torch.empty(C, device=dev, dtype=torch.float32)
        scale = torch.empty(C, device=dev, dtype=torch.float32)
        shift = torch.empty(C, device=dev, dtype=torch.float32)
        _ck(lib.dmp_bn_stats(y_raw.data_ptr(), 1, M, C, accum.data_ptr(), sp), "bn stats")
        count = float(M)
        if sync is not None:
            accum, count = sync.reduce_stats(bn._dmp_key, accum, M)
        mom = bn.momentum
        if bn.track_running_stats and bn.num_batches_tracked is not None:
            bn.num_batches_tracked.add_(1)
            if mom is None:
                mom = 1.0 / float(bn.num_batches_tracked)
        rm = bn.running_mean if bn.track_running_stats else None
        rv = bn.running_var if bn.track_running_stats else None
        _ck(lib.dmp_bn_finalize(accum.data_ptr(), count, _native.ptr(gamma), _native.ptr(beta), _native.ptr(rm), _native.ptr(rv),
                                float(mom if mom is not None else 0.1), float(bn.eps), mean.data_ptr(), invstd.data_ptr(),
                                scale.data_ptr(), shift.data_ptr(), C, sp), "bn finalize")
    else:
        invstd = torch.rsqrt(bn.running_var + bn.eps)
        mean = bn.running_mean
        g = gamma if gamma is not None else torch.ones_like(invstd)
        scale = (g * invstd).float()
        shift = ((beta if beta is not None else 0) - mean * scale).float()
        count = float(M)
    out = torch.empty_like(y_raw)
    _ck(lib.dmp_bn_apply(y_raw.data_ptr(), 1, scale.data_ptr(), shift.data_ptr(), _native.ptr(residual), out.data_ptr(), M, C,
                         1 if relu else 0, sp), "bn apply")
    _note("bn:fused")
    return out, mean, invstd, count, training


def _bn_backward_cuda(ctx, dout, out, y_raw, mean, invstd, gamma, bn, relu, needs_gamma, needs_beta):
    """Returns (dy wrt the BN input, dres or None, dgamma, dbeta)."""
    lib = _k()
    sp = _sp()
    N, C, Ho, Wo = y_raw.shape
    M = N * Ho * Wo
    dres_t = None
    link = getattr(ctx, "link_out", None)
    if link is not None and link.reduced:
        # the consumer's data-gradient kernel already masked the gradient and accumulated sum g / sum g*xhat: one elementwise pass
        link.reduced = False
        gw, gb = (bn.weight if needs_gamma else None), (bn.bias if needs_beta else None)
        direct = all(p is None or (p.grad is not None and getattr(p, "_dmp_direct", False) and p.grad.dtype == torch.float32
                                   and p.grad.is_contiguous()) for p in (gw, gb))
        if direct:
            dg_t, db_t, acc_flag = (gw.grad if gw is not None else None), (gb.grad if gb is not None else None), 1
        else:
            dg_t = torch.empty(C, device=y_raw.device, dtype=torch.float32) if gw is not None else None
            db_t = torch.empty(C, device=y_raw.device, dtype=torch.float32) if gb is not None else None
            acc_flag = 0
        dy = dout if dout.is_contiguous(memory_format=torch.channels_last) else dout.contiguous(memory_format=torch.channels_last)
        _ck(lib.dmp_bn_bwd_dx_sums(dy.data_ptr(), y_raw.data_ptr(), 1, mean.data_ptr(), invstd.data_ptr(), _native.ptr(gamma),
                                   dy.data_ptr(), _native.ptr(dg_t), _native.ptr(db_t), acc_flag, M, C, link.bsums.data_ptr(), sp),
            "bn bwd dx (producer-side sums)")
        _note("bn:bwd_dx_from_sums")
        if direct:
            for p in (gw, gb):
                if p is not None:
                    for cb in grad_ready_callbacks:
                        cb(p)
            return dy, None, None, None
        return dy, None, dg_t, db_t
    if ctx.bn_train and ctx.sync is None:
        # single cooperative launch: reductions -> grid barrier -> dx (+dres), dgamma/dbeta accumulated in place
        gw, gb = (bn.weight if needs_gamma else None), (bn.bias if needs_beta else None)
        direct = all(p is None or (p.grad is not None and getattr(p, "_dmp_direct", False) and p.grad.dtype == torch.float32
                                   and p.grad.is_contiguous()) for p in (gw, gb))
        dy = torch.empty_like(y_raw)
        dres_t = torch.empty_like(y_raw) if ctx.has_res else None
        if direct:
            dg_t, db_t, acc_flag = (gw.grad if gw is not None else None), (gb.grad if gb is not None else None), 1
        else:
            dg_t = torch.empty(C, device=y_raw.device, dtype=torch.float32) if gw is not None else None
            db_t = torch.empty(C, device=y_raw.device, dtype=torch.float32) if gb is not None else None
            acc_flag = 0
        # without a residual the ReLU mask is recomputed from the BN input: one tensor less to read in both passes
        y_mask = out if (relu and ctx.has_res) else None
        rc = lib.dmp_bn_bwd_fused(dout.data_ptr(), _native.ptr(y_mask), y_raw.data_ptr(), 1, mean.data_ptr(), invstd.data_ptr(),
                                  _native.ptr(gamma), _native.ptr(bn.bias), dy.data_ptr(), _native.ptr(dres_t), _native.ptr(dg_t),
                                  _native.ptr(db_t), acc_flag,
                                  M, C, 1 if relu else 0, _bn_workspace(bn, "b", C, y_raw.device).data_ptr(), sp)
        if rc == 0:
            if direct:
                for p in (gw, gb):
                    if p is not None:
                        for cb in grad_ready_callbacks:
                            cb(p)
                return dy, dres_t, None, None
            return dy, dres_t, dg_t, db_t
        if rc != 1:
            _ck(rc, "bn bwd fused")
    if ctx.bn_train and ctx.sync is not None and ctx.sync.equal_batches:
        sync = ctx.sync
        gw, gb = (bn.weight if needs_gamma else None), (bn.bias if needs_beta else None)
        (dg_t, db_t), acc_flag, direct = _param_grad_targets((gw, gb), C, y_raw.device)
        ws, total = sync.fused_workspace((bn._dmp_key, "b"), C, y_raw.device)
        dy = torch.empty_like(y_raw)
        dres_t = torch.empty_like(y_raw) if ctx.has_res else None
        rc = lib.dmp_syncbn_bwd(dout.data_ptr(), _native.ptr(out if relu else None), y_raw.data_ptr(), mean.data_ptr(), invstd.data_ptr(),
                                _native.ptr(gamma), dy.data_ptr(), _native.ptr(dres_t), _native.ptr(dg_t), _native.ptr(db_t), acc_flag,
                                M, C, 1 if relu else 0, ws.ptrs, sync._state.ptrs, sync.rank, sync.world_size, total.data_ptr(), sp)
        if rc == 0:
            sync.kernel_launches += 1
            if direct:
                for p in (gw, gb):
                    if p is not None:
                        for cb in grad_ready_callbacks:
                            cb(p)
                return dy, dres_t, None, None
            return dy, dres_t, dg_t, db_t
        if rc != 1:
            _ck(rc, "syncbn bwd fused")
    if ctx.bn_train:
        dy = torch.empty_like(y_raw)
        dres_t = torch.empty_like(y_raw) if ctx.has_res else None
        accum = torch.empty(2 * C, device=y_raw.device, dtype=torch.float64)
        _ck(lib.dmp_bn_bwd_reduce(dout.data_ptr(), _native.ptr(out), y_raw.data_ptr(), 1, mean.data_ptr(), invstd.data_ptr(), M, C,
                                  1 if relu else 0, accum.data_ptr(), sp), "bn bwd reduce")
        accum_g = accum
        if ctx.sync is not None:
            accum_g = ctx.sync.reduce_grads(bn._dmp_key, accum)
        _ck(lib.dmp_bn_bwd_dx(dout.data_ptr(), _native.ptr(out), y_raw.data_ptr(), 1, mean.data_ptr(), invstd.data_ptr(),
                              _native.ptr(gamma), accum_g.data_ptr(), ctx.count, dy.data_ptr(), _native.ptr(dres_t), M, C,
                              1 if relu else 0, sp), "bn bwd dx")

        def write_dgamma(dst, accumulate):
            _ck(lib.dmp_bn_bwd_params(accum.data_ptr(), dst.data_ptr(), None, C, 1 if accumulate else 0, _sp()), "bn dgamma")

        def write_dbeta(dst, accumulate):
            _ck(lib.dmp_bn_bwd_params(accum.data_ptr(), None, dst.data_ptr(), C, 1 if accumulate else 0, _sp()), "bn dbeta")

        dgamma = _deliver_grad(bn.weight, write_dgamma) if needs_gamma else None
        dbeta = _deliver_grad(bn.bias, write_dbeta) if needs_beta else None
        return dy, dres_t, dgamma, dbeta
    # eval-mode BN: y = x*scale + shift with constant statistics
    g = dout.float()
    if relu:
        g = torch.where(out > 0, g, torch.zeros_like(g))
    if ctx.has_res:
        dres_t = g.to(torch.bfloat16)
    sc = (gamma if gamma is not None else 1.0) * invstd
    dy = (g * sc.view(1, -1, 1, 1)).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    xh = (y_raw.float() - mean.view(1, -1, 1, 1)) * invstd.view(1, -1, 1, 1)
    dgamma = (g * xh).sum((0, 2, 3)) if needs_gamma else None
    dbeta = g.sum((0, 2, 3)) if needs_beta else None
    return dy, dres_t, dgamma, dbeta


class _ConvBNAct(torch.autograd.Function):
    """CUDA path. Inputs: x, conv weight, bn weight, bn bias, residual (the last three may be None)."""

    @staticmethod
    def forward(ctx, x, w, gamma, beta, residual, conv, bn, relu, sync, link_out=None, link_in=None):
        lib = _k()
        sp = _sp()
        kind = conv_kind(conv, x)
        ctx.link_out, ctx.link_in = None, link_in
        if link_in is not None:
            link_in.consumers += 1
        if kind not in ("stem", "stem_gemm"):
            x = _as_cl_bf16(x)
        if kind == "pw" and conv.stride != (1, 1):
            conv._dmp_in_hw = (x.shape[2], x.shape[3])
        sums = _bn_sums_buffer(bn, conv.out_channels, x.device) if _bn_epilogue_stats_ok(bn, sync, kind) else None
        y_raw, x_saved = _conv_fwd_cuda(x, w, conv, kind, sums)
        ctx.conv, ctx.bn, ctx.relu, ctx.kind, ctx.sync = conv, bn, relu, kind, sync
        ctx.has_res = residual is not None
        ctx.x_needs_grad = ctx.needs_input_grad[0]
        if residual is not None:
            residual = _as_cl_bf16(residual)
        if bn is None:
            if relu or residual is not None:
                out = torch.empty_like(y_raw)
                _ck(lib.dmp_add_act(y_raw.data_ptr(), _native.ptr(residual), out.data_ptr(), y_raw.numel(), 1 if relu else 0,
                                    sp), "add_act")
            else:
                out = y_raw
            ctx.save_for_backward(x_saved, w, out if relu else None)
            ctx.bn_train = False
            return out
        out, mean, invstd, ctx.count, ctx.bn_train = _bn_forward_cuda(y_raw, gamma, beta, residual, bn, relu, sync, sums)
        if link_out is not None and ctx.bn_train and sync is None and residual is None:
            # this op's output has a single consumer: let its data-gradient kernel produce the reductions of our backward
            link_out.y_raw, link_out.mean, link_out.invstd, link_out.gamma, link_out.beta = y_raw, mean, invstd, gamma, beta
            link_out.relu, link_out.count = relu, ctx.count
            link_out.bsums = _bn_sums_buffer(bn, y_raw.shape[1], y_raw.device, "bsums")
            ctx.link_out = link_out
        ctx.save_for_backward(x_saved, w, out if relu else None, y_raw, mean, invstd, gamma)
        return out

    @staticmethod
    def backward(ctx, dout):
        lib = _k()
        sp = _sp()
        conv, bn, relu, kind = ctx.conv, ctx.bn, ctx.relu, ctx.kind
        dout = _as_cl_bf16(dout)
        dres = None
        if bn is None:
            x, w, out = ctx.saved_tensors
            if relu:
                dy = torch.empty_like(dout)
                _ck(lib.dmp_relu_bwd(dout.data_ptr(), out.data_ptr(), dy.data_ptr(), dout.numel(), sp), "relu bwd")
            else:
                dy = dout
            if ctx.has_res:
                dres = dy
            dgamma = dbeta = None
        else:
            x, w, out, y_raw, mean, invstd, gamma = ctx.saved_tensors
            dy, dres, dgamma, dbeta = _bn_backward_cuda(ctx, dout, out, y_raw, mean, invstd, gamma, bn, relu,
                                                        ctx.needs_input_grad[2], ctx.needs_input_grad[3])
        dx, write_dw = _conv_bwd_cuda(dy, x, w, conv, kind, ctx.x_needs_grad, ctx.link_in)
        dw = _deliver_grad(conv.weight, write_dw, (dy, x)) if ctx.needs_input_grad[1] else None
        return dx, dw, dgamma, dbeta, dres, None, None, None, None, None, None


class _BNAct(torch.autograd.Function):
    """BatchNorm (+residual, +ReLU) without a convolution in front (bare BN / SyncBatchNorm modules)."""

    @staticmethod
    def forward(ctx, x, gamma, beta, residual, bn, relu, sync):
        x = _as_cl_bf16(x)
        ctx.bn, ctx.relu, ctx.sync = bn, relu, sync
        ctx.has_res = residual is not None
        if residual is not None:
            residual = _as_cl_bf16(residual)
        out, mean, invstd, ctx.count, ctx.bn_train = _bn_forward_cuda(x, gamma, beta, residual, bn, relu, sync)
        ctx.save_for_backward(out if relu else None, x, mean, invstd, gamma)
        return out

    @staticmethod
    def backward(ctx, dout):
        out, x, mean, invstd, gamma = ctx.saved_tensors
        dout = _as_cl_bf16(dout)
        dy, dres, dgamma, dbeta = _bn_backward_cuda(ctx, dout, out, x, mean, invstd, gamma, ctx.bn, ctx.relu,
                                                    ctx.needs_input_grad[1], ctx.needs_input_grad[2])
        return dy, dgamma, dbeta, dres, None, None, None


def _conv_bn_act_reference(x, conv, bn, relu, residual):
    y = F.conv2d(x, conv.weight, conv.bias, conv.stride, conv.padding, conv.dilation, conv.groups)
    if bn is not None:
        y = bn(y)
    if residual is not None:
        y = y + residual
    return F.relu(y) if relu else y


def conv_bn_act(x: torch.Tensor, conv: nn.Conv2d, bn: Optional[nn.Module] = None, relu: bool = False,
                residual: Optional[torch.Tensor] = None, single_consumer: bool = False) -> torch.Tensor:
    """``act(bn(conv(x)) + residual)`` with optional pieces.  ``bn`` may be BatchNorm2d or our SyncBatchNorm.

    ``single_consumer=True`` is the caller's promise that the returned activation feeds exactly one further ``conv_bn_act``
    (and nothing else): that convolution's data-gradient kernel then also produces this BatchNorm's backward reductions."""
    if not _use_kernels(x):
        _note("conv_bn_act:reference")
        return _conv_bn_act_reference(x, conv, bn, relu, residual)
    sync = getattr(bn, "_dmp_sync", None) if bn is not None else None
    if sync is not None and not (bn.training and sync.world_size > 1):
        sync = None
    gamma = bn.weight if bn is not None else None
    beta = bn.bias if bn is not None else None
    link_in = getattr(x, "_dmp_bn_link", None) if torch.is_grad_enabled() else None
    link_out = _BnLink() if (single_consumer and bn is not None and residual is None and _FUSE_BN_BWD and torch.is_grad_enabled()) else None
    out = _ConvBNAct.apply(x, conv.weight, gamma, beta, residual, conv, bn, relu, sync, link_out, link_in)
    if link_out is not None and link_out.y_raw is not None:
        out._dmp_bn_link = link_out
    return out


# ----------------------------------------------------------------------------------------------------
# expand 1x1 -> bn1 -> relu -> depthwise 3x3 -> bn2 -> relu  as one chain (the first two thirds of the reference's Block,
# /root/reference/code/distributed_training/model/mobilenetv2.py:32-33)
# ----------------------------------------------------------------------------------------------------
def _param_grad_targets(params, C, dev):
    """(tensors to hand to a kernel that writes per-channel parameter gradients, accumulate flag, direct?) for BatchNorm
    weight/bias: straight into ``param.grad`` when that is a flat-arena view, else fresh buffers."""
    direct = all(p is None or (p.grad is not None and getattr(p, "_dmp_direct", False) and p.grad.dtype == torch.float32
                               and p.grad.is_contiguous()) for p in params)
    if direct:
        return [p.grad if p is not None else None for p in params], 1, True
    return [torch.empty(C, device=dev, dtype=torch.float32) if p is not None else None for p in params], 0, False


class _PwRawStats(torch.autograd.Function):
    """y_raw = conv1x1(x), with the BatchNorm batch sums of y_raw accumulated by the GEMM epilogue into ``sums``.
    The normalisation itself happens inside the consumer (``_DwBnInBn``) while it loads y_raw."""

    @staticmethod
    def forward(ctx, x, w, conv, sums):
        x = _as_cl_bf16(x)
        ctx.conv = conv
        ctx.x_needs_grad = ctx.needs_input_grad[0]
        y_raw, x_saved = _conv_fwd_cuda(x, w, conv, "pw", sums)
        ctx.save_for_backward(x_saved, w)
        return y_raw

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dy = _as_cl_bf16(dy)
        dx, write_dw = _conv_bwd_cuda(dy, x, w, ctx.conv, "pw", ctx.x_needs_grad)
        dw = _deliver_grad(ctx.conv.weight, write_dw, (dy, x)) if ctx.needs_input_grad[1] else None
        return dx, dw, None, None


class _DwBnInBn(torch.autograd.Function):
    """a2 = relu(bn2(depthwise(relu(bn1(y1_raw))))) given the RAW expand output y1_raw and bn1's batch sums.

    Forward: ONE stencil kernel normalises y1_raw on load, convolves, accumulates bn2's batch sums and publishes bn1's
    statistics; one elementwise kernel applies bn2.  relu(bn1(y1_raw)) never exists in memory.
    Backward: bn2 backward (one elementwise pass when the consumer's data-gradient kernel produced its reductions, else the
    cooperative kernel) -> depthwise weight gradient with the activation re-created on load -> depthwise data gradient fused
    with bn1's ReLU mask and reductions -> bn1's elementwise dx pass."""

    @staticmethod
    def forward(ctx, y1, g1, b1, w2, g2, b2, bn1, bn2, conv2, sums1, link_out):
        lib = _k()
        sp = _sp()
        N, C, H, W = y1.shape
        dev = y1.device
        s = conv2.stride[0]
        Ho, Wo = (H + 2 - 3) // s + 1, (W + 2 - 3) // s + 1
        y2 = _cl_empty(N, C, Ho, Wo, dev)
        mean1 = torch.empty(C, device=dev, dtype=torch.float32)
        invstd1 = torch.empty(C, device=dev, dtype=torch.float32)
        sums2 = _bn_sums_buffer(bn2, C, dev)
        t1 = bn1.track_running_stats and bn1.running_mean is not None
        _ck(lib.dmp_dwconv3x3_fwd_bnin_colstats(y1.data_ptr(), weight_bf16(w2).data_ptr(), y2.data_ptr(), N, H, W, C, s,
                                                sums1.data_ptr(), _native.ptr(g1), _native.ptr(b1), float(bn1.eps),
                                                float(bn1.momentum or 0.1), _native.ptr(bn1.running_mean if t1 else None),
                                                _native.ptr(bn1.running_var if t1 else None),
                                                _native.ptr(bn1.num_batches_tracked if t1 else None), mean1.data_ptr(),
                                                invstd1.data_ptr(), sums2.data_ptr(), sp), "dwconv fwd (bn1 on load) + bn2 sums")
        _note("conv_dw:stencil+bn_in+bnstats")
        a2, mean2, invstd2, count2, _ = _bn_forward_cuda(y2, g2, b2, None, bn2, True, None, sums2)
        ctx.bn1, ctx.bn2, ctx.conv2 = bn1, bn2, conv2
        ctx.link_out = None
        if link_out is not None:
            link_out.y_raw, link_out.mean, link_out.invstd, link_out.gamma, link_out.beta = y2, mean2, invstd2, g2, b2
            link_out.relu, link_out.count = True, count2
            link_out.bsums = _bn_sums_buffer(bn2, C, dev, "bsums")
            ctx.link_out = link_out
        ctx.save_for_backward(y1, y2, mean1, invstd1, mean2, invstd2, g1, b1, w2, g2, b2)
        return a2

    @staticmethod
    def backward(ctx, da2):
        lib = _k()
        sp = _sp()
        y1, y2, mean1, invstd1, mean2, invstd2, g1, b1, w2, g2, b2 = ctx.saved_tensors
        bn1, bn2, conv2 = ctx.bn1, ctx.bn2, ctx.conv2
        N, C, H, W = y1.shape
        _, _, Ho, Wo = y2.shape
        s = conv2.stride[0]
        dev = y1.device
        da2 = _as_cl_bf16(da2)
        # ---- bn2 backward -> dy2 (gradient of the depthwise convolution's raw output)
        ctx.bn_train, ctx.sync, ctx.has_res, ctx.count = True, None, False, float(N * Ho * Wo)
        dy2, _, dg2, db2 = _bn_backward_cuda(ctx, da2, None, y2, mean2, invstd2, g2, bn2, True, ctx.needs_input_grad[4],
                                             ctx.needs_input_grad[5])
        # ---- depthwise weight gradient; its input relu(bn1(y1)) is re-created while loading y1
        def write_dw(dst, accumulate):
            _ck(lib.dmp_dwconv3x3_wgrad_bnin(dy2.data_ptr(), y1.data_ptr(), dst.data_ptr(), N, H, W, C, s, 1 if accumulate else 0,
                                             mean1.data_ptr(), invstd1.data_ptr(), _native.ptr(g1), _native.ptr(b1), _sp()),
                "dwconv wgrad (bn1 on load)")

        dw2 = _deliver_grad(conv2.weight, write_dw, (dy2, y1)) if ctx.needs_input_grad[3] else None
        # ---- depthwise data gradient fused with bn1's ReLU mask and backward reductions, then bn1's elementwise dx pass
        bs1 = _bn_sums_buffer(bn1, C, dev, "bsums")
        g = _cl_empty(N, C, H, W, dev)
        _ck(lib.dmp_dwconv3x3_dgrad_bnbwd(dy2.data_ptr(), weight_bf16(w2).data_ptr(), g.data_ptr(), N, H, W, C, s, y1.data_ptr(),
                                          mean1.data_ptr(), invstd1.data_ptr(), _native.ptr(g1), _native.ptr(b1), bs1.data_ptr(), sp),
            "dwconv dgrad + bn1 bwd reductions")
        p1 = (bn1.weight if ctx.needs_input_grad[1] else None, bn1.bias if ctx.needs_input_grad[2] else None)
        (dg1_t, db1_t), acc_flag, direct = _param_grad_targets(p1, C, dev)
        _ck(lib.dmp_bn_bwd_dx_sums(g.data_ptr(), y1.data_ptr(), 1, mean1.data_ptr(), invstd1.data_ptr(), _native.ptr(g1), g.data_ptr(),
                                   _native.ptr(dg1_t), _native.ptr(db1_t), acc_flag, N * H * W, C, bs1.data_ptr(), sp), "bn1 bwd dx")
        _note("bn:bwd_dx_from_sums")
        if direct:
            for p in p1:
                if p is not None:
                    for cb in grad_ready_callbacks:
                        cb(p)
            dg1_t = db1_t = None
        return g, dg1_t, db1_t, dw2, dg2, db2, None, None, None, None, None


def _chain_ok(x, conv1, bn1, conv2, bn2) -> bool:
    if not (_FUSE_CHAIN and _EPILOGUE_STATS and _use_kernels(x) and torch.is_grad_enabled()):
        return False
    if bn1 is None or bn2 is None or not (bn1.training and bn2.training):
        return False
    for bn in (bn1, bn2):
        if getattr(bn, "_dmp_sync", None) is not None or bn.momentum is None or not isinstance(bn, nn.BatchNorm2d):
            return False
    if conv_kind(conv1, x) != "pw" or conv1.stride != (1, 1):
        return False
    C = conv2.in_channels
    if not (conv2.kernel_size == (3, 3) and conv2.groups == C == conv2.out_channels and conv2.padding == (1, 1)
            and conv2.stride in ((1, 1), (2, 2)) and conv2.bias is None and C % 16 == 0 and conv1.out_channels == C):
        return False
    if conv2.stride == (2, 2) and (x.shape[2] % 2 or x.shape[3] % 2):
        return False
    return True


def expand_dw_bn_act(x: torch.Tensor, conv1: nn.Conv2d, bn1, conv2: nn.Conv2d, bn2, single_consumer: bool = False) -> torch.Tensor:
    """``relu(bn2(conv2(relu(bn1(conv1(x))))))`` for a 1x1 expand followed by a depthwise 3x3 (train mode).  On the kernel
    path the intermediate activation between bn1 and the depthwise convolution is never materialised; everywhere else (CPU,
    eval mode, SyncBatchNorm, unsupported shapes) this is two ``conv_bn_act`` calls."""
    if not _chain_ok(x, conv1, bn1, conv2, bn2):
        out = conv_bn_act(x, conv1, bn1, relu=True, single_consumer=True)
        return conv_bn_act(out, conv2, bn2, relu=True, single_consumer=single_consumer)
    dev = x.device
    sums1 = _bn_sums_buffer(bn1, conv1.out_channels, dev)
    y1 = _PwRawStats.apply(x, conv1.weight, conv1, sums1)
    link_out = _BnLink() if (single_consumer and _FUSE_BN_BWD) else None
    a2 = _DwBnInBn.apply(y1, bn1.weight, bn1.bias, conv2.weight, bn2.weight, bn2.bias, bn1, bn2, conv2, sums1, link_out)
    if link_out is not None:
        a2._dmp_bn_link = link_out
    return a2


class _IdentityConv:
    """Stand-in so a bare BatchNorm can reuse the fused op: y_raw = x."""
    kernel_size, stride, padding, dilation, groups, bias, padding_mode = (1, 1), (1, 1), (0, 0), (1, 1), 1, None, "zeros"


def batch_norm_act(x: torch.Tensor, bn: nn.Module, relu: bool = False, residual: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``act(bn(x) + residual)`` on the kernel path (x channels-last bf16), reference path otherwise."""
    if not _use_kernels(x) or x.dim() != 4 or x.shape[1] % 8 != 0:
        y = F.batch_norm(x, bn.running_mean, bn.running_var, bn.weight, bn.bias, bn.training or bn.running_mean is None,
                         bn.momentum if bn.momentum is not None else 0.1, bn.eps)
        if residual is not None:
            y = y + residual
        return F.relu(y) if relu else y
    sync = getattr(bn, "_dmp_sync", None)
    if sync is not None and not (bn.training and sync.world_size > 1):
        sync = None
    return _BNAct.apply(x, bn.weight, bn.bias, residual, bn, relu, sync)


# ----------------------------------------------------------------------------------------------------
# pooling / flatten
# ----------------------------------------------------------------------------------------------------
class _GlobalAvgPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, relu):
        x = _as_cl_bf16(x)
        N, C, H, W = x.shape
        y = torch.empty((N, C), device=x.device, dtype=torch.bfloat16)
        _ck(_k().dmp_avgpool_fwd(x.data_ptr(), y.data_ptr(), N, H * W, C, 1 if relu else 0, _sp()), "avgpool fwd")
        ctx.relu = relu
        ctx.shape = (N, C, H, W)
        ctx.save_for_backward(x if relu else None)
        _note("avgpool:kernel")
        return y

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        N, C, H, W = ctx.shape
        dy = dy.to(torch.bfloat16).contiguous()
        dx = _cl_empty(N, C, H, W, dy.device)
        _ck(_k().dmp_avgpool_bwd(dy.data_ptr(), _native.ptr(x), dx.data_ptr(), N, H * W, C, 1 if ctx.relu else 0, _sp()),
            "avgpool bwd")
        return dx, None


def global_avgpool_flatten(x: torch.Tensor, relu: bool = False) -> torch.Tensor:
    """``flatten(avg_pool2d(relu?(x), full spatial extent))`` -> [N, C] (the reference's Reshape1 / forward tail)."""
    if _use_kernels(x) and x.shape[1] % 8 == 0:
        return _GlobalAvgPool.apply(x, relu)
    _note("avgpool:reference")
    if relu:
        x = F.relu(x)
    return F.adaptive_avg_pool2d(x, 1).flatten(1)


class _MaxPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, s, p):
        x = _as_cl_bf16(x)
        N, C, H, W = x.shape
        Ho, Wo = (H + 2 * p - k) // s + 1, (W + 2 * p - k) // s + 1
        y = _cl_empty(N, C, Ho, Wo, x.device)
        idx = torch.empty((N, Ho, Wo, C), device=x.device, dtype=torch.uint8)
        _ck(_k().dmp_maxpool_fwd(x.data_ptr(), y.data_ptr(), idx.data_ptr(), N, H, W, C, k, s, p, _sp()), "maxpool fwd")
        ctx.geom = (N, C, H, W, k, s, p)
        ctx.save_for_backward(idx)
        _note("maxpool:kernel")
        return y

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        N, C, H, W, k, s, p = ctx.geom
        dy = _as_cl_bf16(dy)
        dx = _cl_empty(N, C, H, W, dy.device)
        _ck(_k().dmp_maxpool_bwd(dy.data_ptr(), idx.data_ptr(), dx.data_ptr(), N, H, W, C, k, s, p, _sp()), "maxpool bwd")
        return dx, None, None, None


def max_pool2d(x: torch.Tensor, kernel_size: int, stride: Optional[int] = None, padding: int = 0) -> torch.Tensor:
    """Max pooling on the kernel path (channels-last bf16, gather backward from the saved winning taps); the ResNet-50 stem
    pool and VGG's 2x2 pools (model zoo of /root/reference/code/distributed_training/data_parallel.py:58-73)."""
    stride = kernel_size if stride is None else stride
    if _use_kernels(x) and x.dim() == 4 and x.shape[1] % 8 == 0 and kernel_size <= 15:
        return _MaxPool.apply(x, int(kernel_size), int(stride), int(padding))
    _note("maxpool:reference")
    return F.max_pool2d(x, kernel_size, stride, padding)


# ----------------------------------------------------------------------------------------------------
# transformer pieces (BERT-base pipeline config; nothing of this exists in the reference)
# ----------------------------------------------------------------------------------------------------
class _LayerNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, res, gamma, beta, ln):
        shp = x.shape
        Hd = shp[-1]
        x2 = x.reshape(-1, Hd)
        x2 = x2 if (x2.dtype == torch.bfloat16 and x2.is_contiguous()) else x2.to(torch.bfloat16).contiguous()
        r2 = None
        if res is not None:
            r2 = res.reshape(-1, Hd)
            r2 = r2 if (r2.dtype == torch.bfloat16 and r2.is_contiguous()) else r2.to(torch.bfloat16).contiguous()
        rows = x2.shape[0]
        y = torch.empty_like(x2)
        z = torch.empty_like(x2) if r2 is not None else x2
        mean = torch.empty(rows, device=x.device, dtype=torch.float32)
        rstd = torch.empty(rows, device=x.device, dtype=torch.float32)
        _ck(_k().dmp_layernorm_fwd(x2.data_ptr(), _native.ptr(r2), gamma.data_ptr(), beta.data_ptr(), y.data_ptr(),
                                   z.data_ptr() if r2 is not None else None, mean.data_ptr(), rstd.data_ptr(), rows, Hd, float(ln.eps),
                                   _sp()), "layernorm fwd")
        ctx.ln, ctx.has_res, ctx.shape = ln, res is not None, shp
        ctx.save_for_backward(z, gamma, mean, rstd)
        _note("layernorm:kernel")
        return y.reshape(shp)

    @staticmethod
    def backward(ctx, dy):
        z, gamma, mean, rstd = ctx.saved_tensors
        ln = ctx.ln
        rows, Hd = z.shape
        dy2 = dy.reshape(rows, Hd)
        dy2 = dy2 if (dy2.dtype == torch.bfloat16 and dy2.is_contiguous()) else dy2.to(torch.bfloat16).contiguous()
        dz = torch.empty_like(z)
        (dg_t, db_t), acc, direct = _param_grad_targets((ln.weight, ln.bias), Hd, z.device)
        if not direct:
            dg_t.zero_(); db_t.zero_()
        _ck(_k().dmp_layernorm_bwd(dy2.data_ptr(), z.data_ptr(), gamma.data_ptr(), mean.data_ptr(), rstd.data_ptr(), dz.data_ptr(),
                                   dg_t.data_ptr(), db_t.data_ptr(), rows, Hd, _sp()), "layernorm bwd")
        if direct:
            for p in (ln.weight, ln.bias):
                for cb in grad_ready_callbacks:
                    cb(p)
            dg_t = db_t = None
        dzr = dz.reshape(ctx.shape)
        return dzr, (dzr if ctx.has_res else None), dg_t, db_t, None


def layer_norm(x: torch.Tensor, ln: nn.LayerNorm, residual: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``LayerNorm(x + residual)`` over the last dimension: one kernel forward (the sum is written once, for backward), one
    backward (dx = dresidual, dgamma, dbeta)."""
    Hd = x.shape[-1]
    if _use_kernels(x) and Hd % 8 == 0 and Hd <= 2048 and ln.elementwise_affine and ln.bias is not None:
        return _LayerNorm.apply(x, residual, ln.weight, ln.bias, ln)
    _note("layernorm:reference")
    z = x if residual is None else x + residual
    return F.layer_norm(z.float(), (Hd,), ln.weight, ln.bias, ln.eps).to(z.dtype)


class _Gelu(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        x = x if (x.dtype == torch.bfloat16 and x.is_contiguous()) else x.to(torch.bfloat16).contiguous()
        y = torch.empty_like(x)
        _ck(_k().dmp_gelu_fwd(x.data_ptr(), y.data_ptr(), x.numel(), _sp()), "gelu fwd")
        ctx.save_for_backward(x)
        _note("gelu:kernel")
        return y

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        dy = dy if (dy.dtype == torch.bfloat16 and dy.is_contiguous()) else dy.to(torch.bfloat16).contiguous()
        dx = torch.empty_like(x)
        _ck(_k().dmp_gelu_bwd(dy.data_ptr(), x.data_ptr(), dx.data_ptr(), x.numel(), _sp()), "gelu bwd")
        return dx


def gelu(x: torch.Tensor) -> torch.Tensor:
    if _use_kernels(x) and x.numel() % 8 == 0:
        return _Gelu.apply(x)
    _note("gelu:reference")
    return F.gelu(x)


def _bgemm(layout, A, lda, a_sh, a_sb, B, ldb, b_sh, b_sb, D, ldd, d_sh, d_sb, M, N, K, nh, nb, what):
    _ck(_k().dmp_gemm_bf16_batched(layout, A, lda, a_sh, a_sb, B, ldb, b_sh, b_sb, D, ldd, d_sh, d_sb, M, N, K, nh, nb, _sp()), what)


class _Attention(torch.autograd.Function):
    """Softmax attention on the fused projection qkv [B, S, 3, heads, d] (bf16): six batched tcgen05 GEMMs (every head of every
    sequence in one persistent launch each, operands addressed in place through 4-D tensor maps) and a row softmax."""

    @staticmethod
    def forward(ctx, qkv, heads):
        B, S, three, nh, d = qkv.shape
        qkv = qkv if (qkv.dtype == torch.bfloat16 and qkv.is_contiguous()) else qkv.to(torch.bfloat16).contiguous()
        dev = qkv.device
        ld = 3 * nh * d                       # row pitch of Q / K / V inside the fused projection
        sb = S * ld                           # batch stride
        esz = 2
        q, k, v = qkv.data_ptr(), qkv.data_ptr() + nh * d * esz, qkv.data_ptr() + 2 * nh * d * esz
        P = torch.empty((B, nh, S, S), device=dev, dtype=torch.bfloat16)
        out = torch.empty((B, S, nh, d), device=dev, dtype=torch.bfloat16)
        _bgemm(0, q, ld, d, sb, k, ld, d, sb, P.data_ptr(), S, S * S, nh * S * S, S, S, d, nh, B, "attention scores")
        _ck(_k().dmp_softmax_fwd(P.data_ptr(), P.data_ptr(), B * nh * S, S, 1.0 / (d ** 0.5), _sp()), "attention softmax")
        _bgemm(1, P.data_ptr(), S, S * S, nh * S * S, v, ld, d, sb, out.data_ptr(), nh * d, d, S * nh * d, S, d, S, nh, B, "attention PV")
        ctx.save_for_backward(qkv, P)
        _note("attention:tcgen05_batched")
        return out.view(B, S, nh * d)

    @staticmethod
    def backward(ctx, dout):
        qkv, P = ctx.saved_tensors
        B, S, _, nh, d = qkv.shape
        dev = qkv.device
        ld, sb, esz = 3 * nh * d, S * 3 * nh * d, 2
        q, k, v = qkv.data_ptr(), qkv.data_ptr() + nh * d * esz, qkv.data_ptr() + 2 * nh * d * esz
        do = dout.reshape(B, S, nh, d)
        do = do if (do.dtype == torch.bfloat16 and do.is_contiguous()) else do.to(torch.bfloat16).contiguous()
        dqkv = torch.empty_like(qkv)
        dq, dk, dv = dqkv.data_ptr(), dqkv.data_ptr() + nh * d * esz, dqkv.data_ptr() + 2 * nh * d * esz
        lo, so = nh * d, S * nh * d           # dO / context layout [B, S, heads, d]
        dP = torch.empty_like(P)
        # dV = P^T dO, dP = dO V^T, dS = softmax'(dP), dQ = dS K, dK = dS^T Q
        _bgemm(2, P.data_ptr(), S, S * S, nh * S * S, do.data_ptr(), lo, d, so, dv, ld, d, sb, S, d, S, nh, B, "attention dV")
        _bgemm(0, do.data_ptr(), lo, d, so, v, ld, d, sb, dP.data_ptr(), S, S * S, nh * S * S, S, S, d, nh, B, "attention dP")
        _ck(_k().dmp_softmax_bwd(dP.data_ptr(), P.data_ptr(), dP.data_ptr(), B * nh * S, S, 1.0 / (d ** 0.5), _sp()), "attention softmax bwd")
        _bgemm(1, dP.data_ptr(), S, S * S, nh * S * S, k, ld, d, sb, dq, ld, d, sb, S, d, S, nh, B, "attention dQ")
        _bgemm(2, dP.data_ptr(), S, S * S, nh * S * S, q, ld, d, sb, dk, ld, d, sb, S, d, S, nh, B, "attention dK")
        return dqkv, None


def attention(qkv: torch.Tensor, heads: int) -> torch.Tensor:
    """Multi-head softmax attention of a fused projection ``qkv`` [B, S, 3*H] -> context [B, S, H]."""
    B, S, H3 = qkv.shape
    Hd = H3 // 3
    d = Hd // heads
    if _use_kernels(qkv) and d % 8 == 0 and S % 8 == 0:
        return _Attention.apply(qkv.view(B, S, 3, heads, d), heads)
    _note("attention:reference")
    q, k, v = qkv.view(B, S, 3, heads, d).permute(2, 0, 3, 1, 4)
    a = F.scaled_dot_product_attention(q, k, v)
    return a.transpose(1, 2).reshape(B, S, Hd)


# ----------------------------------------------------------------------------------------------------
# linear
# ----------------------------------------------------------------------------------------------------
class _Linear(torch.autograd.Function):
    """y = x W^T + b.  Wide layers use the tcgen05 GEMM (bf16 out); skinny classifiers (out_features < 64
    or not a multiple of 8) use the warp-dot kernel with fp32 logits."""

    @staticmethod
    def forward(ctx, x, w, b, lin):
        lib = _k()
        sp = _sp()
        shp = x.shape
        x2 = x.reshape(-1, shp[-1])
        if x2.dtype != torch.bfloat16 or not x2.is_contiguous():
            x2 = x2.to(torch.bfloat16).contiguous()
        B, K = x2.shape
        Nf = w.shape[0]
        skinny = Nf < 64 or Nf % 8 != 0 or K % 8 != 0
        ctx.skinny, ctx.lin, ctx.shape_in = skinny, lin, shp
        if skinny:
            y = torch.empty((B, Nf), device=x.device, dtype=torch.float32)
            _ck(lib.dmp_linear_small_fwd(x2.data_ptr(), w.data_ptr(), 0, _native.ptr(b), y.data_ptr(), B, Nf, K, sp), "linear fwd")
            _note("linear:skinny")
        else:
            y = torch.empty((B, Nf), device=x.device, dtype=torch.bfloat16)
            wb = weight_bf16(w)
            bf = b if (b is None or b.dtype == torch.float32) else b.float()
            _ck(lib.dmp_gemm_bf16(0, x2.data_ptr(), K, wb.data_ptr(), K, y.data_ptr(), Nf, B, Nf, K, _native.ptr(bf), 0, 1, sp),
                "linear fwd gemm")
            _note("linear:tcgen05")
        ctx.save_for_backward(x2, w)
        return y.reshape(*shp[:-1], Nf)

    @staticmethod
    def backward(ctx, dy):
        lib = _k()
        sp = _sp()
        x2, w = ctx.saved_tensors
        lin = ctx.lin
        B, K = x2.shape
        Nf = w.shape[0]
        need_dx = ctx.needs_input_grad[0]
        dx = None
        if ctx.skinny:
            dy2 = dy.reshape(B, Nf).float().contiguous()
            if need_dx:
                dx = torch.empty((B, K), device=dy.device, dtype=torch.bfloat16)
                _ck(lib.dmp_linear_small_bwd(dy2.data_ptr(), x2.data_ptr(), w.data_ptr(), 0, dx.data_ptr(), None, None, B, Nf, K, 0,
                                             sp), "linear dx")

            def write_dw(dst, accumulate):
                _ck(lib.dmp_linear_small_bwd(dy2.data_ptr(), x2.data_ptr(), w.data_ptr(), 0, None, dst.data_ptr(), None, B, Nf, K,
                                             1 if accumulate else 0, _sp()), "linear dw")

            def write_db(dst, accumulate):
                if accumulate:
                    dst.add_(dy2.sum(0))
                else:
                    torch.sum(dy2, 0, out=dst)
        else:
            dy2 = dy.reshape(B, Nf)
            if dy2.dtype != torch.bfloat16 or not dy2.is_contiguous():
                dy2 = dy2.to(torch.bfloat16).contiguous()
            if need_dx:
                wb = weight_bf16(w)
                dx = torch.empty((B, K), device=dy.device, dtype=torch.bfloat16)
                _ck(lib.dmp_gemm_bf16(1, dy2.data_ptr(), Nf, wb.data_ptr(), K, dx.data_ptr(), K, B, K, Nf, None, 0, 1, sp),
                    "linear dgrad gemm")

            def write_dw(dst, accumulate):
                if not accumulate:
                    dst.zero_()
                _ck(lib.dmp_gemm_bf16(2, dy2.data_ptr(), Nf, x2.data_ptr(), K, dst.data_ptr(), K, Nf, K, B, None, 2, 0, _sp()),
                    "linear wgrad gemm")

            def write_db(dst, accumulate):
                s = dy2.float().sum(0)
                if accumulate:
                    dst.add_(s)
                else:
                    dst.copy_(s)
        dw = _deliver_grad(lin.weight, write_dw) if ctx.needs_input_grad[1] else None
        db = _deliver_grad(lin.bias, write_db) if (lin.bias is not None and ctx.needs_input_grad[2]) else None
        if dx is not None:
            dx = dx.reshape(ctx.shape_in)
        return dx, dw, db, None


def linear(x: torch.Tensor, lin: nn.Linear) -> torch.Tensor:
    if not _use_kernels(x):
        _note("linear:reference")
        return F.linear(x, lin.weight.to(x.dtype), lin.bias.to(x.dtype) if lin.bias is not None else None)
    return _Linear.apply(x, lin.weight, lin.bias, lin)


# ----------------------------------------------------------------------------------------------------
# loss
# ----------------------------------------------------------------------------------------------------
class _CrossEntropy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target, stats_out):
        B, C = logits.shape
        if logits.dtype not in (torch.float32, torch.bfloat16) or not logits.is_contiguous():
            logits = logits.float().contiguous()
        dl = torch.empty_like(logits)
        rows = torch.empty(3 * B, device=logits.device, dtype=torch.float32)
        out = stats_out if stats_out is not None else torch.empty(3, device=logits.device, dtype=torch.float32)
        _ck(_k().dmp_xent_fwd_bwd(logits.data_ptr(), target.data_ptr(), dl.data_ptr(), rows.data_ptr(), out.data_ptr(), B, C,
                                  0 if logits.dtype == torch.float32 else 1, 1.0 / B, _sp()), "cross entropy")
        ctx.save_for_backward(dl)
        _note("xent:fused")
        return out[0].clone()

    @staticmethod
    def backward(ctx, g):
        (dl,) = ctx.saved_tensors
        return dl * g.to(dl.dtype), None, None


def cross_entropy(logits: torch.Tensor, target: torch.Tensor, stats_out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Mean softmax cross-entropy.  On CUDA one fused kernel also produces dlogits and, if ``stats_out`` (3 floats)
    is given, {loss, top-1 hits, top-5 hits} for the metrics the reference logs."""
    if not _use_kernels(logits):
        _note("xent:reference")
        loss = F.cross_entropy(logits.float(), target)
        if stats_out is not None:
            with torch.no_grad():
                k = min(5, logits.shape[1])
                top = logits.topk(k, 1).indices
                stats_out[0] = loss.detach()
                stats_out[1] = (top[:, 0] == target).sum()
                stats_out[2] = (top == target[:, None]).any(1).sum()
        return loss
    if target.dtype != torch.int64:
        target = target.long()
    return _CrossEntropy.apply(logits, target.contiguous(), stats_out)
