"""Gradient Reducer: buckets, autograd hooks, in-order launch, finalize -- the Python face of
``csrc/runtime/reducer.cpp`` (bookkeeping) and ``csrc/comm/collectives.cu`` (the fused all-reduce).

Behaviour follows the c10d Reducer that torch's DistributedDataParallel constructs
(torch/nn/parallel/distributed.py:1183-1304; studied in /root/reference/Readme.md:145-157):

* bucket assignment by size with a small first bucket (1 MiB) and ``bucket_cap_mb`` (25 MiB) after it,
  computed on the parameter list and *reversed* so bucket 0 holds the last layers (whose gradients
  arrive first);
* one hook per parameter; a bucket launches when all its gradients are in **and** every lower-indexed
  bucket has launched (all ranks issue the same sequence of collectives);
* a callback queued on the autograd engine finalizes the iteration (waits for the collectives, writes
  gradients back, checks that every parameter was accounted for);
* after the first synchronised iteration the buckets are rebuilt once in observed gradient-ready order.

Two communication back ends share this logic: ``TorchDistBackend`` (``torch.distributed.all_reduce``;
gloo on CPU, NCCL on GPU -- the baseline) and ``PeerBackend`` (our NVLink peer-memory kernels, optionally
fused with the SGD update).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Callable, List, Optional, Sequence

import torch
import torch.distributed as dist

from .. import _native
from ..ops import functional as OF
from ..utils.tracing import nvtx_range

_native.register_runtime_sigs({
    "dmp_bucket_assign": (C.c_int, [C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.c_int,
                                    C.POINTER(C.c_int32)]),
    "dmp_reducer_create": (C.c_void_p, [C.c_int, C.c_int, C.POINTER(C.c_int32)]),
    "dmp_reducer_destroy": (None, [C.c_void_p]),
    "dmp_reducer_prepare": (None, [C.c_void_p]),
    "dmp_reducer_mark_ready": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_int32), C.c_int]),
    "dmp_reducer_finalize": (C.c_int, [C.c_void_p, C.POINTER(C.c_int32), C.c_int]),
    "dmp_reducer_num_launched": (C.c_int, [C.c_void_p]),
    "dmp_reducer_iterations": (C.c_longlong, [C.c_void_p]),
    "dmp_reducer_ready_order": (C.c_int, [C.c_void_p, C.POINTER(C.c_int32), C.c_int]),
    "dmp_reducer_unused_flags": (None, [C.c_void_p, C.POINTER(C.c_int32)]),
    "dmp_reducer_set_plans": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int]),
    "dmp_reducer_clear_plans": (None, [C.c_void_p]),
    "dmp_reducer_grad_ready": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_int32), C.c_int]),
    "dmp_reducer_finish": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_int32), C.c_int]),
    "dmp_reducer_native_launches": (C.c_longlong, [C.c_void_p]),
    "dmp_reducer_plan_bytes": (C.c_int, []),
    "dmp_reducer_fns_bytes": (C.c_int, []),
})

PLAN_CALLER, PLAN_ALLREDUCE_FUSED, PLAN_NVLS, PLAN_NVLS_SGD, PLAN_SGD_FLAT, PLAN_NONE, PLAN_NVLS_THEN_SGD = range(7)


class BucketPlan(C.Structure):
    """Mirror of ``struct DmpBucketPlan`` (csrc/runtime/reducer.cpp): how the native hook launches one bucket."""
    _fields_ = [("kind", C.c_int), ("algo", C.c_int), ("dtype", C.c_int), ("rank", C.c_int), ("world", C.c_int),
                ("nesterov", C.c_int), ("zero_after", C.c_int), ("max_blocks", C.c_int), ("numel", C.c_longlong),
                ("scale", C.c_float),
                ("bufs", C.c_void_p), ("state", C.c_void_p), ("out", C.c_void_p), ("master", C.c_void_p),
                ("momentum", C.c_void_p), ("shadow", C.c_void_p), ("hp", C.c_void_p), ("master_ptrs", C.c_void_p),
                ("shadow_ptrs", C.c_void_p), ("mc_grad", C.c_void_p), ("mc_master", C.c_void_p), ("mc_shadow", C.c_void_p),
                ("grad", C.c_void_p)]


class LaunchFns(C.Structure):
    """Mirror of ``struct DmpLaunchFns``: entry points of the kernel library (or test stubs) the native hook calls."""
    _fields_ = [(n, C.c_void_p) for n in ("fork", "join", "allreduce_fused", "allreduce_nvls", "allreduce_nvls_sgd", "sgd_flat")]


def kernel_launch_fns() -> LaunchFns:
    """The real entry points, taken from libdmp_b200.so."""
    lib = _native.kernels()
    f = LaunchFns()
    for field, sym in (("fork", "dmp_stream_fork"), ("join", "dmp_stream_join"), ("allreduce_fused", "dmp_allreduce_fused"),
                       ("allreduce_nvls", "dmp_allreduce_nvls"), ("allreduce_nvls_sgd", "dmp_allreduce_nvls_sgd"),
                       ("sgd_flat", "dmp_sgd_flat")):
        setattr(f, field, C.cast(getattr(lib, sym), C.c_void_p).value)
    return f

DEFAULT_FIRST_BUCKET_BYTES = 1024 * 1024
DEFAULT_BUCKET_CAP_MB = 25
ALIGN_ELEMS = 8  # every parameter starts on a 16-byte boundary even in the bf16 shadow


def compute_bucket_assignment_by_size(sizes_bytes: Sequence[int], limits: Sequence[int],
                                      keys: Optional[Sequence[int]] = None) -> List[List[int]]:
    """Greedy size-capped bucketing (native).  Returns buckets as lists of parameter indices, ordered by
    smallest index.  Mirrors ``dist._compute_bucket_assignment_by_size`` for dense parameters."""
    n = len(sizes_bytes)
    if n == 0:
        return []
    rt = _native.runtime()
    sz = (C.c_int64 * n)(*[int(s) for s in sizes_bytes])
    ky = (C.c_int32 * n)(*[int(k) for k in (keys if keys is not None else [0] * n)])
    lim = (C.c_int64 * len(limits))(*[int(min(l, 2 ** 62)) for l in limits])
    out = (C.c_int32 * n)()
    nb = rt.dmp_bucket_assign(n, sz, ky, lim, len(limits), out)
    buckets: List[List[int]] = [[] for _ in range(nb)]
    for i in range(n):
        buckets[out[i]].append(i)
    return buckets


_EMPTY: List[int] = []


class NativeReducerState:
    """Thin RAII wrapper over the C++ reducer state machine."""
    native_launch = False

    def __init__(self, bucket_of_param: Sequence[int], n_buckets: int):
        self.rt = _native.runtime()
        n = len(bucket_of_param)
        arr = (C.c_int32 * n)(*[int(b) for b in bucket_of_param])
        self.h = self.rt.dmp_reducer_create(n, n_buckets, arr)
        if not self.h:
            raise RuntimeError("dmp_reducer_create failed")
        self.n, self.nb = n, n_buckets
        self._out = (C.c_int32 * max(1, n_buckets))()

    def prepare(self):
        self.rt.dmp_reducer_prepare(self.h)

    def mark_ready(self, p: int, unused: bool = False) -> List[int]:
        k = self.rt.dmp_reducer_mark_ready(self.h, p, 1 if unused else 0, self._out, self.nb)
        if k < 0:
            if k == -1:
                raise RuntimeError(
                    "parameter %d was marked ready twice in one iteration (a parameter shared by several layers, or "
                    "several forward passes before one backward, is not supported without static_graph)" % p)
            raise RuntimeError("reducer mark_ready failed with code %d" % k)
        return [self._out[i] for i in range(k)]

    def finalize(self) -> List[int]:
        miss = (C.c_int32 * self.n)()
        m = self.rt.dmp_reducer_finalize(self.h, miss, self.n)
        return [miss[i] for i in range(min(m, self.n))]

    def ready_order(self) -> List[int]:
        out = (C.c_int32 * self.n)()
        k = self.rt.dmp_reducer_ready_order(self.h, out, self.n)
        return [out[i] for i in range(min(k, self.n))]

    def num_launched(self) -> int:
        return self.rt.dmp_reducer_num_launched(self.h)

    # ---- native launch path: the whole hook (mark -> launch decision -> stream fork -> collective) is one C++ call
    def set_plans(self, fns: LaunchFns, plans: Sequence[BucketPlan], comm_stream: int, side_streams: Sequence[int]) -> bool:
        """Installs one launch plan per bucket.  Returns True when every bucket launches natively."""
        assert self.rt.dmp_reducer_plan_bytes() == C.sizeof(BucketPlan) and self.rt.dmp_reducer_fns_bytes() == C.sizeof(LaunchFns)
        arr = (BucketPlan * len(plans))(*plans)
        side = (C.c_void_p * max(1, len(side_streams)))(*[C.c_void_p(s) for s in side_streams])
        self._keep = (fns, arr, side)
        rc = self.rt.dmp_reducer_set_plans(self.h, C.addressof(fns), C.addressof(arr), len(plans), C.c_void_p(comm_stream),
                                           C.addressof(side), len(side_streams))
        if rc < 0:
            raise RuntimeError("dmp_reducer_set_plans failed with code %d" % rc)
        self.native_launch = rc == 1
        return self.native_launch

    def clear_plans(self):
        self.rt.dmp_reducer_clear_plans(self.h)
        self.native_launch = False

    def grad_ready(self, p: int, unused: bool, cur_stream: int) -> List[int]:
        """The autograd hook's one native call.  Returns the buckets the CALLER must launch (empty on the native path)."""
        k = self.rt.dmp_reducer_grad_ready(self.h, p, 1 if unused else 0, C.c_void_p(cur_stream), self._out, self.nb)
        if k == 0:
            return _EMPTY
        if k < 0:
            if k == -1:
                raise RuntimeError(
                    "parameter %d was marked ready twice in one iteration (a parameter shared by several layers, or "
                    "several forward passes before one backward, is not supported without static_graph)" % p)
            if k <= -100:
                _native.check(-100 - k, "bucket collective (native reducer path)")
            raise RuntimeError("reducer grad_ready failed with code %d" % k)
        return [self._out[i] for i in range(k)]

    def finish(self, cur_stream: int) -> List[int]:
        miss = (C.c_int32 * self.n)()
        m = self.rt.dmp_reducer_finish(self.h, C.c_void_p(cur_stream), miss, self.n)
        if m <= -100:
            _native.check(-100 - m, "communication stream join (native reducer path)")
        return [miss[i] for i in range(min(m, self.n))]

    def native_launches(self) -> int:
        return int(self.rt.dmp_reducer_native_launches(self.h))

    def __del__(self):
        try:
            if self.h:
                self.rt.dmp_reducer_destroy(self.h)
                self.h = None
        except Exception:
            pass


# ------------------------------------------------------------------------------------------------------
# communication back ends
# ------------------------------------------------------------------------------------------------------
class TorchDistBackend:
    """Bucket all-reduce through ``torch.distributed`` (gloo on CPU, NCCL on GPU).  Default-hook semantics:
    pre-divide by the world size, then all-reduce-sum (torch default_hooks.py:26-32)."""

    name = "torch.distributed"
    supports_fused = False

    def __init__(self, group=None):
        self.group = group if group is not None else dist.group.WORLD
        self.world_size = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)

    def alloc_flat(self, numel, dtype, device, symmetric=False):
        return torch.zeros(numel, dtype=dtype, device=device), None

    def pad_multiple(self, dtype) -> int:
        return ALIGN_ELEMS

    def launch(self, bucket: "Bucket"):
        if bucket.flat.is_cuda:  # gradients may have been written on the op layer's side stream as well
            cur = torch.cuda.current_stream(bucket.flat.device)
            for st in OF.side_streams().gradient_streams(bucket.flat.device):
                if st.cuda_stream != cur.cuda_stream:
                    cur.wait_stream(st)
        bucket.flat.div_(self.world_size)
        bucket.work = dist.all_reduce(bucket.flat, group=self.group, async_op=True)

    def wait(self, bucket: "Bucket"):
        if bucket.work is not None:
            bucket.work.wait()
            bucket.work = None

    def begin_iteration(self):
        pass

    def end_iteration(self):
        pass


class PeerBackend:
    """Bucket all-reduce with our peer-memory kernels on a dedicated high-priority stream.

    One-shot for buckets up to ``oneshot_max_bytes`` (env ``DMP_ONESHOT_MAX_BYTES``; default by rank count: 8 MiB at 2 ranks,
    4 MiB up to 4, 2 MiB beyond, from the measured crossovers), two-shot above.  With a ``FusedSGDState`` attached to the bucket the same kernel applies the 1/W scale,
    the bf16->fp32 cast and the SGD-momentum update, writes the new parameters (fp32 master + bf16 shadow)
    and zeroes the gradient bucket for the next iteration.
    """

    name = "peer-kernels"
    supports_fused = True
    NVLS_FUSED_MIN_WORLD = 4

    def __init__(self, group=None, device=None, oneshot_max_bytes: Optional[int] = None, max_blocks: Optional[int] = None):
        from .symm import DTYPE_CODE, ProcessSymmGroup

        self.DTYPE_CODE = DTYPE_CODE
        self.lib = _native.kernels()
        self.group = group if group is not None else (dist.group.WORLD if dist.is_initialized() else None)
        self.world_size = dist.get_world_size(self.group) if self.group is not None else 1
        self.rank = dist.get_rank(self.group) if self.group is not None else 0
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.symm = ProcessSymmGroup(self.group, self.device) if self.world_size > 1 else None
        self.state = self.symm.new_state() if self.symm is not None else None
        # one-shot (every rank reads every peer's whole bucket) against two-shot (reduce-scatter + all-gather): the crossover
        # moves down with the rank count -- 4..16 MB at 2 ranks, 1..4 MB at 8 (profiles/allreduce_sweep_*gpu.json)
        table = 8 << 20 if self.world_size <= 2 else (4 << 20 if self.world_size <= 4 else 2 << 20)
        self.oneshot_max_bytes = int(os.environ.get("DMP_ONESHOT_MAX_BYTES", oneshot_max_bytes or table))
        self.max_blocks = int(os.environ.get("DMP_COMM_MAX_BLOCKS", max_blocks or 0))
        self.stream = torch.cuda.Stream(device=self.device, priority=-1)
        self._launched_any = False
        self.kernel_launches = 0
        # Opt-in (DMP_DDP_NVLS=1): fp32 buckets of at least DMP_DDP_NVLS_MIN_BYTES (default 8 MiB) without a fused optimizer
        # live in an NVLS multicast buffer and are reduced inside the switch (multimem.ld_reduce / multimem.st).  Off by
        # default: at W = 2 the peer-load two-shot is faster (profiles/README.md); meant for W = 8.
        self.nvls = os.environ.get("DMP_DDP_NVLS", "0") == "1" and self.world_size > 1
        self.nvls_min_bytes = int(os.environ.get("DMP_DDP_NVLS_MIN_BYTES", 8 << 20))
        self.nvls_launches = 0
        # NVLS all-reduce FUSED with SGD (gradient bucket, fp32 master and bf16 shadow in multicast buffers; see
        # allreduce_nvls_sgd_kernel).  DMP_DDP_NVLS_FUSED: "0" off, "1" on, "auto" (default) = on from NVLS_FUSED_MIN_WORLD ranks
        # for buckets of at least DMP_DDP_NVLS_FUSED_MIN_BYTES -- the switch-side reduction moves (4+10)/W bytes per element over
        # each GPU's link where the peer-load two-shot moves 10 (W-1)/W, which only pays once W is large enough
        mode = os.environ.get("DMP_DDP_NVLS_FUSED", "auto")
        self.nvls_fused = self.world_size > 1 and (mode == "1" or (mode == "auto" and self.world_size >= self.NVLS_FUSED_MIN_WORLD))
        self.nvls_fused_min_bytes = int(os.environ.get("DMP_DDP_NVLS_FUSED_MIN_BYTES", 1 << 20))
        self.nvls_fused_launches = 0
        # Buckets of at least DMP_DDP_NVLS_SPLIT_MIN_BYTES are reduced inside the switch and then updated by the flat SGD kernel
        # on this rank's copy, instead of one kernel that multicasts master + shadow + zeroed gradients (10 bytes per element
        # into every GPU).  Measured at 8 ranks (profiles/allreduce_sweep_8gpu_r2.json), single kernel vs all-reduce + SGD vs
        # NCCL + SGD: 1 MB 34 / 35 / 36 us, 4 MB 60 / 43 / 63, 16 MB 122 / 79 / 118, 64 MB 507 / 239 / 348, 256 MB 2009 / 890 / 1019.
        self.nvls_split_min_bytes = int(os.environ.get("DMP_DDP_NVLS_SPLIT_MIN_BYTES", 2 << 20))
        self.nvls_split_launches = 0

    def pad_multiple(self, dtype) -> int:
        esz = torch.empty((), dtype=dtype).element_size()
        m = max(ALIGN_ELEMS, (16 // esz) * max(1, self.world_size))
        if self.nvls_fused:
            m = max(m, 8 * self.world_size)  # the NVLS fused kernel works on groups of 8 elements per rank slice
        return m

    def wants_nvls_fused(self, numel, dtype) -> bool:
        return self.nvls_fused and dtype == torch.float32 and numel * 4 >= self.nvls_fused_min_bytes

    def alloc_flat(self, numel, dtype, device, symmetric=True):
        if self.symm is None or not symmetric:
            return torch.zeros(numel, dtype=dtype, device=device), None
        esz = torch.empty((), dtype=dtype).element_size()
        if symmetric == "nvls_fused":  # a multicast buffer or nothing: the caller checked wants_nvls_fused on every rank alike
            from .symm import ProcessMulticast
            try:
                mc = ProcessMulticast(numel * esz, self.group, self.device)
                mc.local.zero_()
                return mc.view(dtype, numel), mc
            except RuntimeError:  # no multicast on this platform (every rank takes this branch together)
                self.nvls_fused = False
                symmetric = True
        if self.nvls and symmetric == "nvls_ok" and dtype == torch.float32 and numel * esz >= self.nvls_min_bytes:
            from .symm import ProcessMulticast
            try:
                mc = ProcessMulticast(numel * esz, self.group, self.device)
                return mc.view(dtype, numel), mc
            except RuntimeError:  # no multicast on this platform (every rank takes this branch together)
                self.nvls = False
        buf = self.symm.alloc(numel * esz)
        return buf.view(dtype, numel), buf

    def begin_iteration(self):
        self._launched_any = False

    def launch(self, bucket: "Bucket"):
        cur = torch.cuda.current_stream(self.device)
        ev = torch.cuda.Event()
        ev.record(cur)
        self.stream.wait_event(ev)
        # the bucket's gradients were written on the stream backward runs on AND on the op layer's weight-gradient side stream
        for st in OF.side_streams().gradient_streams(self.device):
            if st.cuda_stream != cur.cuda_stream:
                self.stream.wait_stream(st)
        self._launched_any = True
        fused = bucket.fused
        W = self.world_size
        flat = bucket.flat
        nbytes = flat.numel() * flat.element_size()
        sp = self.stream.cuda_stream
        lib = self.lib
        if W == 1:
            if fused is not None:
                _native.check(lib.dmp_sgd_flat(fused.master.data_ptr(), _native.ptr(fused.shadow), flat.data_ptr(),
                                               self.DTYPE_CODE[flat.dtype], fused.momentum.data_ptr(), flat.numel(),
                                               fused.hp.data_ptr(), 1 if fused.nesterov else 0, 1, sp), "sgd_flat")
                self.kernel_launches += 1
            return
        algo = 0 if nbytes <= self.oneshot_max_bytes else 1
        scale = 1.0 / W
        if fused is not None and fused.nvls is not None and nbytes >= self.nvls_split_min_bytes:
            _native.check(lib.dmp_allreduce_nvls(fused.nvls[0], self.state.ptrs, self.rank, W, flat.numel(), scale, self.max_blocks, sp),
                          "allreduce nvls")
            _native.check(lib.dmp_sgd_flat(fused.master.data_ptr(), _native.ptr(fused.shadow), flat.data_ptr(),
                                           self.DTYPE_CODE[flat.dtype], fused.momentum.data_ptr(), flat.numel(),
                                           fused.hp.data_ptr(), 1 if fused.nesterov else 0, 1, sp), "sgd_flat after nvls")
            self.nvls_split_launches += 1
            self.kernel_launches += 2
            return
        if fused is not None and fused.nvls is not None:
            g_mc, m_mc, s_mc = fused.nvls
            _native.check(lib.dmp_allreduce_nvls_sgd(g_mc, m_mc, s_mc, fused.master.data_ptr(), fused.momentum.data_ptr(),
                                                     fused.hp.data_ptr(), 1 if fused.nesterov else 0, self.state.ptrs, self.rank, W,
                                                     flat.numel(), scale, self.max_blocks, sp), "allreduce nvls + sgd")
            self.nvls_fused_launches += 1
            self.kernel_launches += 1
            return
        if fused is None and hasattr(bucket.symm, "mc_ptr"):  # bucket lives in a multicast buffer: reduce inside the switch
            _native.check(lib.dmp_allreduce_nvls(bucket.symm.mc_ptr, self.state.ptrs, self.rank, W, flat.numel(), scale,
                                                 self.max_blocks, sp), "allreduce nvls")
            self.nvls_launches += 1
        elif fused is None:
            out = None
            if algo == 0:
                if bucket.scratch is None:
                    bucket.scratch = torch.empty_like(flat)
                out = bucket.scratch
            _native.check(lib.dmp_allreduce_fused(algo, self.DTYPE_CODE[flat.dtype], bucket.symm.ptrs, self.state.ptrs, self.rank,
                                                  W, flat.numel(), scale, _native.ptr(out), None, None, None, None, 0, None,
                                                  None, 0, self.max_blocks, sp), "allreduce")
        else:
            _native.check(lib.dmp_allreduce_fused(algo, self.DTYPE_CODE[flat.dtype], bucket.symm.ptrs, self.state.ptrs, self.rank,
                                                  W, flat.numel(), scale, None, fused.master.data_ptr(),
                                                  fused.momentum.data_ptr(), _native.ptr(fused.shadow), fused.hp.data_ptr(),
                                                  1 if fused.nesterov else 0,
                                                  fused.master_symm.ptrs if fused.master_symm is not None else None,
                                                  fused.shadow_symm.ptrs if fused.shadow_symm is not None else None, 1,
                                                  self.max_blocks, sp), "allreduce+sgd")
        self.kernel_launches += 1

    def native_plan(self, bucket: "Bucket") -> Optional[BucketPlan]:
        """What ``launch`` would do for this bucket, as data: the C++ reducer then launches it without coming back here."""
        def addr(x):
            return None if x is None else (x if isinstance(x, int) else C.cast(x, C.c_void_p).value)

        fused, W, flat = bucket.fused, self.world_size, bucket.flat
        pl = BucketPlan()
        pl.dtype, pl.rank, pl.world, pl.numel = self.DTYPE_CODE[flat.dtype], self.rank, W, flat.numel()
        pl.scale, pl.max_blocks = 1.0 / W, self.max_blocks
        pl.nesterov = 1 if (fused is not None and fused.nesterov) else 0
        if fused is not None:
            pl.master, pl.momentum, pl.hp = fused.master.data_ptr(), fused.momentum.data_ptr(), fused.hp.data_ptr()
            pl.shadow = _native.ptr(fused.shadow) or None
        if W == 1:
            pl.kind = PLAN_SGD_FLAT if fused is not None else PLAN_NONE
            pl.grad = flat.data_ptr()
            return pl
        pl.state = addr(self.state.ptrs)
        if fused is not None and fused.nvls is not None:
            pl.kind = PLAN_NVLS_THEN_SGD if flat.numel() * flat.element_size() >= self.nvls_split_min_bytes else PLAN_NVLS_SGD
            pl.mc_grad, pl.mc_master, pl.mc_shadow = fused.nvls
            pl.grad = flat.data_ptr()
            return pl
        if fused is None and hasattr(bucket.symm, "mc_ptr"):
            pl.kind, pl.mc_grad = PLAN_NVLS, bucket.symm.mc_ptr
            return pl
        pl.kind = PLAN_ALLREDUCE_FUSED
        pl.algo = 0 if flat.numel() * flat.element_size() <= self.oneshot_max_bytes else 1
        pl.bufs = addr(bucket.symm.ptrs)
        if fused is None:
            if pl.algo == 0:
                if bucket.scratch is None:
                    bucket.scratch = torch.empty_like(flat)
                pl.out = bucket.scratch.data_ptr()
            pl.zero_after = 0
        else:
            pl.master_ptrs = addr(fused.master_symm.ptrs) if fused.master_symm is not None else None
            pl.shadow_ptrs = addr(fused.shadow_symm.ptrs) if fused.shadow_symm is not None else None
            pl.zero_after = 1
        return pl

    def account_native(self, plans: Sequence[BucketPlan]):
        """Launch counters for one iteration the C++ reducer launched (every bucket exactly once)."""
        for pl in plans:
            if pl.kind != PLAN_NONE:
                self.kernel_launches += 1
            if pl.kind == PLAN_NVLS_THEN_SGD:
                self.nvls_split_launches += 1
                self.kernel_launches += 1
            if pl.kind == PLAN_NVLS_SGD:
                self.nvls_fused_launches += 1
            elif pl.kind == PLAN_NVLS:
                self.nvls_launches += 1

    def wait(self, bucket: "Bucket"):
        pass  # stream-ordered; joined once in end_iteration

    def end_iteration(self):
        if self._launched_any:
            torch.cuda.current_stream(self.device).wait_stream(self.stream)

    def check_health(self):
        if self.state is None:
            return
        code = self.lib.dmp_comm_status(self.state.local.data_ptr(), torch.cuda.current_stream(self.device).cuda_stream)
        if code != 0:
            raise RuntimeError("peer collective timed out waiting for rank %d (status %d): a peer died or the ranks "
                               "launched different collective sequences" % (code - 1, code))


class FusedSGDState:
    """Per-bucket optimizer state for the fused all-reduce + SGD kernel."""

    def __init__(self, master, momentum, shadow, hp, nesterov, master_symm=None, shadow_symm=None):
        self.master, self.momentum, self.shadow, self.hp, self.nesterov = master, momentum, shadow, hp, nesterov
        self.master_symm, self.shadow_symm = master_symm, shadow_symm
        self.nvls = None  # (gradient, master, shadow) multicast addresses when the bucket is reduced inside the switch


class Bucket:
    def __init__(self, index: int, param_indices: List[int], params: List[torch.Tensor], backend, symmetric=True):
        self.index = index
        self.param_indices = param_indices
        self.params = params
        dtype = params[0].dtype
        device = params[0].device
        self.offsets = []
        off = 0
        for p in params:
            self.offsets.append(off)
            off += (p.numel() + ALIGN_ELEMS - 1) // ALIGN_ELEMS * ALIGN_ELEMS
        mult = backend.pad_multiple(dtype)
        self.numel = (off + mult - 1) // mult * mult
        self.flat, self.symm = backend.alloc_flat(self.numel, dtype, device, symmetric)
        self.views = [self.flat[o: o + p.numel()].view_as(p) for o, p in zip(self.offsets, params)]
        self.work = None
        self.scratch = None
        self.fused: Optional[FusedSGDState] = None
        self.future = None

    def nbytes(self):
        return self.flat.numel() * self.flat.element_size()


class GradBucket:
    """What a communication hook receives (subset of torch.distributed.GradBucket)."""

    def __init__(self, bucket: Bucket, is_last: bool):
        self._b, self._last = bucket, is_last

    def index(self):
        return self._b.index

    def buffer(self):
        return self._b.flat

    def set_buffer(self, t):
        self._b.flat.copy_(t)

    def gradients(self):
        return list(self._b.views)

    def parameters(self):
        return list(self._b.params)

    def is_last(self):
        return self._last


class Reducer:
    def __init__(self, params: List[torch.nn.Parameter], backend, bucket_cap_bytes: int, first_bucket_bytes: int,
                 find_unused_parameters: bool = False, gradient_as_bucket_view: bool = False,
                 param_names: Optional[List[str]] = None, fused_sgd: Optional[dict] = None, static_graph: bool = False):
        self.params = params
        self.backend = backend
        self.bucket_cap_bytes = int(bucket_cap_bytes)
        self.first_bucket_bytes = int(first_bucket_bytes)
        self.find_unused = find_unused_parameters
        self.grad_as_view = gradient_as_bucket_view or fused_sgd is not None
        self.param_names = param_names or [str(i) for i in range(len(params))]
        self.fused_cfg = fused_sgd
        self.static_graph = static_graph
        self.require_sync = True
        self._comm_hook: Optional[Callable] = None
        self.comm_hook_state = None
        self._plans: List[BucketPlan] = []
        self._plans_dirty = True
        self._side_seen = None
        self._hooks_installed = False
        self._in_backward = False
        self._callback_queued = False
        self._rebuilt = fused_sgd is not None  # fused layout is fixed at construction
        self._unused_local: List[int] = []
        self._index_of = {id(p): i for i, p in enumerate(params)}
        self.stats = {"iterations": 0, "buckets_launched": 0, "rebuilds": 0}
        self.hp = None
        # initial assignment: model order, limits [first, cap], then reversed
        order = list(range(len(params)))
        self._build_buckets(self._assign(order, reverse=True))
        self._install_hooks()

    @property
    def comm_hook(self):
        return self._comm_hook

    @comm_hook.setter
    def comm_hook(self, hook):
        self._comm_hook = hook
        self._plans_dirty = True  # a communication hook takes the launches back to Python

    def _install_plans(self, side=None):
        """Hands the C++ reducer one launch plan per bucket (DMP_NATIVE_REDUCER=0 keeps every launch in Python).  Buckets a
        communication hook or a torch.distributed back end reduces cannot be launched natively: the hook then returns their
        ids instead."""
        self._plans_dirty = False
        self._plans = []
        native_ok = (os.environ.get("DMP_NATIVE_REDUCER", "1") != "0" and self._comm_hook is None
                     and hasattr(self.backend, "native_plan"))
        if not native_ok:
            self.native.clear_plans()
            return
        plans = [self.backend.native_plan(bk) for bk in self.buckets]
        if any(pl is None for pl in plans):
            self.native.clear_plans()
            return
        side = side or []
        self._plans = plans
        self.native.set_plans(kernel_launch_fns(), plans, self.backend.stream.cuda_stream, [s.cuda_stream for s in side])

    # -------------------------------------------------------------------------------- bucket layout
    def _assign(self, order: List[int], reverse: bool) -> List[List[int]]:
        sizes = [self.params[i].numel() * self.params[i].element_size() for i in order]
        keyset = {}
        keys = []
        for i in order:
            p = self.params[i]
            keys.append(keyset.setdefault((p.dtype, str(p.device)), len(keyset)))
        groups = compute_bucket_assignment_by_size(sizes, [self.first_bucket_bytes, self.bucket_cap_bytes], keys)
        buckets = [[order[j] for j in g] for g in groups]
        if reverse:
            buckets = list(reversed(buckets))
        return buckets

    def _build_buckets(self, assignment: List[List[int]]):
        self.buckets: List[Bucket] = []
        bucket_of = [0] * len(self.params)
        for b, idxs in enumerate(assignment):
            # plain gradient buckets (no fused optimizer state to ship to peers) may live in an NVLS multicast buffer
            mode = "nvls_ok" if self.fused_cfg is None else True
            if self.fused_cfg is not None and hasattr(self.backend, "wants_nvls_fused"):
                ps = [self.params[i] for i in idxs]
                numel = sum((p.numel() + ALIGN_ELEMS - 1) // ALIGN_ELEMS * ALIGN_ELEMS for p in ps)
                if bool(self.fused_cfg.get("bf16_shadow", True)) and self.backend.wants_nvls_fused(numel, ps[0].dtype):
                    mode = "nvls_fused"
            bk = Bucket(b, idxs, [self.params[i] for i in idxs], self.backend, mode)
            self.buckets.append(bk)
            for i in idxs:
                bucket_of[i] = b
        self.bucket_of = bucket_of
        self.slot_of = {}
        for bk in self.buckets:
            for k, i in enumerate(bk.param_indices):
                self.slot_of[i] = (bk, k)
        self.native = NativeReducerState(bucket_of, len(self.buckets))
        self._plans_dirty = True
        if self.fused_cfg is not None:
            self._attach_fused()
        if self.grad_as_view:
            for bk in self.buckets:
                for p, v in zip(bk.params, bk.views):
                    if self.fused_cfg is None and p.grad is not None:
                        v.copy_(p.grad)
                    p.grad = v
                    p._dmp_direct = True

    def _attach_fused(self):
        cfg = self.fused_cfg
        dev = self.params[0].device
        self.hp = torch.tensor([cfg.get("lr", 0.1), cfg.get("momentum", 0.0), cfg.get("weight_decay", 0.0), 1.0],
                               dtype=torch.float32, device=dev)
        use_shadow = bool(cfg.get("bf16_shadow", True))
        for bk in self.buckets:
            if bk.flat.dtype not in (torch.float32, torch.bfloat16):
                raise RuntimeError("fused SGD supports fp32/bf16 gradient buckets")
            via_switch = hasattr(bk.symm, "mc_ptr") and use_shadow  # the gradient bucket sits in a multicast buffer
            amode = "nvls_fused" if via_switch else True
            master, msymm = self.backend.alloc_flat(bk.numel, torch.float32, dev, amode)
            momentum = torch.zeros(bk.numel, dtype=torch.float32, device=dev)
            shadow = ssymm = None
            if use_shadow:
                shadow, ssymm = self.backend.alloc_flat(bk.numel, torch.bfloat16, dev, amode)
            for p, off in zip(bk.params, bk.offsets):
                if p.dtype != torch.float32:
                    raise RuntimeError("fused SGD expects fp32 master parameters (bf16 compute uses the shadow copy)")
                mv = master[off: off + p.numel()].view_as(p)
                mv.copy_(p.data)
                p.data = mv  # the parameter now lives inside the flat fp32 master arena
                if shadow is not None:
                    sv = shadow[off: off + p.numel()].view_as(p)
                    sv.copy_(mv)
                    p._dmp_shadow = sv
            bk.fused = FusedSGDState(master, momentum, shadow, self.hp, bool(cfg.get("nesterov", False)), msymm, ssymm)
            if via_switch and hasattr(msymm, "mc_ptr") and hasattr(ssymm, "mc_ptr"):
                bk.fused.nvls = (bk.symm.mc_ptr, msymm.mc_ptr, ssymm.mc_ptr)
            elif via_switch:
                raise RuntimeError("NVLS fused all-reduce: the master / shadow multicast buffers could not be created although the "
                                   "gradient bucket's was")

    def sync_shadow(self):
        """Re-derive the bf16 shadows from the fp32 masters (after load_state_dict or manual edits)."""
        for bk in self.buckets:
            if bk.fused is not None and bk.fused.shadow is not None:
                bk.fused.shadow.copy_(bk.fused.master)

    # -------------------------------------------------------------------------------- hooks
    def _install_hooks(self):
        if self._hooks_installed:
            return
        self._handles = []
        for i, p in enumerate(self.params):
            self._handles.append(p.register_post_accumulate_grad_hook(self._make_hook(i)))
        OF.grad_ready_callbacks.append(self._direct_ready)
        self._hooks_installed = True

    def remove_hooks(self):
        for h in getattr(self, "_handles", []):
            h.remove()
        if self._direct_ready in OF.grad_ready_callbacks:
            OF.grad_ready_callbacks.remove(self._direct_ready)
        self._hooks_installed = False

    def _make_hook(self, i):
        def hook(param):
            self._grad_ready(i, param)
        return hook

    def _direct_ready(self, param):
        i = self._index_of.get(id(param))
        if i is not None:
            self._grad_ready(i, param, direct=True)

    def prepare_for_backward(self, unused_indices: Sequence[int] = ()):
        """Called at the end of forward (when gradients will be synchronised by the coming backward)."""
        if self._plans_dirty:
            self._install_plans()
        self.native.prepare()
        self.backend.begin_iteration()
        self._in_backward = True
        self._callback_queued = False
        self._unused_local = list(unused_indices)
        self._direct_marked = set()

    def shadow_iteration(self):
        """One backward's worth of collectives with a zero contribution from this rank (DistributedDataParallel.join): every
        bucket is launched, in bucket order, exactly as the ranks that are still training launch theirs."""
        self.require_sync = True
        self.prepare_for_backward(())
        for bk in self.buckets:
            if bk.fused is None:
                bk.flat.zero_()
        self._unused_local = list(range(len(self.params)))
        for i in self._unused_local:
            self._mark(i, unused=True)
        self.finalize_backward()

    def _grad_ready(self, i: int, param, direct: bool = False):
        if not self.require_sync or not self._in_backward:
            return
        if direct:
            self._direct_marked.add(i)
        elif i in self._direct_marked:
            return  # autograd still visits AccumulateGrad (with an undefined gradient) after a direct delivery
        if not self._callback_queued:
            self._callback_queued = True
            torch.autograd.Variable._execution_engine.queue_callback(self.finalize_backward)
            for u in self._unused_local:  # locally unused parameters contribute zeros
                self._mark(u, unused=True)
        bk, k = self.slot_of[i]
        view = bk.views[k]
        g = param.grad
        if g is None:
            view.zero_()
        elif g.data_ptr() != view.data_ptr():  # gradient lives elsewhere (plain autograd tensor, another arena)
            view.copy_(g)
            if self.grad_as_view and not direct:
                param.grad = view
        self._mark(i, unused=False)

    def _mark(self, i: int, unused: bool):
        if unused:
            bk, k = self.slot_of[i]
            if bk.fused is None:
                bk.views[k].zero_()
        if self.native.native_launch:
            dev = self.backend.device
            # weight gradients may still be in flight on this thread's side streams: the C++ side orders the communication
            # stream after them (the list changes once, when the autograd thread creates its streams)
            side = OF.side_streams().streams.get(dev.index)
            if side is not self._side_seen:
                self._side_seen = side
                self.native.set_plans(kernel_launch_fns(), self._plans, self.backend.stream.cuda_stream,
                                      [s.cuda_stream for s in (side or [])])
            self.native.grad_ready(i, unused, torch.cuda.current_stream(dev).cuda_stream)  # launches what became launchable
            return
        for b in self.native.grad_ready(i, unused, 0):
            self._launch(self.buckets[b])

    def _launch(self, bk: Bucket):
        self.stats["buckets_launched"] += 1
        if self.comm_hook is not None:
            bk.future = self.comm_hook(self.comm_hook_state, GradBucket(bk, bk.index == len(self.buckets) - 1))
        else:
            with nvtx_range("ddp.bucket%d" % bk.index):
                self.backend.launch(bk)

    def finalize_backward(self):
        if not self._in_backward:
            return
        self._in_backward = False
        if self.native.native_launch:
            missing = self.native.finish(torch.cuda.current_stream(self.backend.device).cuda_stream)
            if not missing:
                self.backend.account_native(self._plans)
                self.stats["buckets_launched"] += len(self.buckets)
        else:
            missing = self.native.finalize()
        if missing:
            names = [self.param_names[i] for i in missing[:8]]
            raise RuntimeError(
                "DistributedDataParallel: %d parameter(s) received no gradient in this backward pass (e.g. %s). "
                "Pass find_unused_parameters=True if parts of the model are skipped." % (len(missing), ", ".join(names)))
        for bk in self.buckets:
            if bk.future is not None:
                res = bk.future.wait()
                if isinstance(res, (list, tuple)):
                    res = res[0]
                if res is not None and res.data_ptr() != bk.flat.data_ptr():
                    bk.flat.copy_(res.view_as(bk.flat))
                bk.future = None
            else:
                self.backend.wait(bk)
        self.backend.end_iteration()
        globally_unused = self._globally_unused()
        if not self.grad_as_view:
            for bk in self.buckets:
                for p, v in zip(bk.params, bk.views):
                    if id(p) in globally_unused:
                        continue
                    if p.grad is not None:
                        p.grad.copy_(v)
                    else:
                        p.grad = v.clone()
        elif globally_unused:
            for p in self.params:
                if id(p) in globally_unused:
                    p.grad = None
        self.stats["iterations"] += 1

    def _globally_unused(self) -> set:
        """Parameters no rank used in this iteration keep their ``.grad`` untouched (None with the usual set-to-none
        zero_grad), as in torch's reducer (its all-reduced ``local_used_map``): an optimizer with weight decay or momentum must
        not move a parameter that took no part.  One extra small all-reduce, only with find_unused_parameters=True; not
        offered with the fused optimizer, whose update has already run inside the bucket kernels."""
        W = getattr(self.backend, "world_size", 1)
        if not self.find_unused or self.fused_cfg is not None or W <= 1 or not (dist.is_available() and dist.is_initialized()):
            return set()
        group = self.backend.group
        dev = self.params[0].device if dist.get_backend(group) == "nccl" else "cpu"
        used = torch.ones(len(self.params), dtype=torch.int32)
        if self._unused_local:
            used[torch.tensor(self._unused_local, dtype=torch.long)] = 0
        used = used.to(dev)
        dist.all_reduce(used, group=group)
        idx = (used == 0).nonzero().flatten().tolist()
        self.stats["globally_unused"] = len(idx)
        return {id(self.params[i]) for i in idx}

    # -------------------------------------------------------------------------------- rebuild
    def maybe_rebuild_buckets(self) -> bool:
        """One-time re-bucketing in gradient-ready order (torch: ``reducer._rebuild_buckets()``)."""
        if self._rebuilt or self.find_unused or self.stats["iterations"] < 1:
            return False
        order = self.native.ready_order()
        if len(order) != len(self.params):
            return False
        self._rebuilt = True
        # every rank must agree: adopt rank 0's order
        if dist.is_available() and dist.is_initialized() and getattr(self.backend, "world_size", 1) > 1:
            t = torch.tensor(order, dtype=torch.int32, device=self.params[0].device if self.params[0].is_cuda else "cpu")
            dist.broadcast(t, src=dist.get_global_rank(self.backend.group, 0) if hasattr(dist, "get_global_rank") else 0,
                           group=self.backend.group)
            order = [int(x) for x in t.tolist()]
        assignment = self._assign(order, reverse=False)
        saved = [(p.grad.clone() if p.grad is not None else None) for p in self.params] if self.grad_as_view else None
        self._build_buckets(assignment)
        if saved is not None:
            for p, g in zip(self.params, saved):
                if g is not None:
                    p.grad.copy_(g)
        self.stats["rebuilds"] += 1
        return True

    def bucket_sizes(self) -> List[int]:
        return [bk.nbytes() for bk in self.buckets]
