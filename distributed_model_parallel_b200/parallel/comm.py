"""Coalesced device-to-device collectives for the single-process DataParallel path, and the coalesced
process-group broadcast DDP uses at construction / for buffers.

API shape follows ``torch.cuda.comm`` (torch/nn/parallel/comm.py:50-67 ``broadcast_coalesced``, :124-168
``reduce_add_coalesced``, :209 ``scatter``, :258 ``gather``) -- the functions the reference README singles out
as the C++ part of DataParallel (/root/reference/Readme.md:28-30,49-56,68,135).  On CUDA the data movement
is done by our peer-memory kernels:

* ``broadcast_coalesced``   one kernel on the source GPU storing each coalesced chunk into every replica
* ``reduce_add_coalesced``  one kernel on the destination GPU peer-loading and summing every replica's chunk
* ``gather``                one kernel on the destination GPU peer-loading every replica's slice
* ``scatter``               each target GPU peer-loads its own slice

CPU "devices" (used by the unit tests and the no-GPU plumbing path) take a plain-PyTorch route.
"""
from __future__ import annotations

from collections import defaultdict
from typing import Dict, Iterable, List, Optional, Sequence

import threading

import torch
import torch.distributed as dist

from .. import _native

DEFAULT_BUFFER_BYTES = 10 * 1024 * 1024  # torch default: 10 MiB coalescing buffer
stats: Dict[str, int] = {}               # which transport served broadcast chunks (multicast vs peer stores)


def _dev(d) -> torch.device:
    if isinstance(d, torch.device):
        return d
    if isinstance(d, int):
        return torch.device("cuda", d)
    return torch.device(d)


def take_tensors(tensors: Iterable[torch.Tensor], size_limit: int):
    """Group tensors into per-dtype chunks of at most ``size_limit`` bytes (order preserved within a dtype)."""
    buf: Dict[torch.dtype, list] = defaultdict(lambda: [[], 0])
    for t in tensors:
        entry = buf[t.dtype]
        nbytes = t.numel() * t.element_size()
        if entry[1] + nbytes > size_limit and entry[1] > 0:
            yield entry[0]
            entry[0], entry[1] = [], 0
            entry = buf[t.dtype]
        entry[0].append(t)
        entry[1] += nbytes
    for entry in buf.values():
        if entry[0]:
            yield entry[0]


def flatten_dense(tensors: Sequence[torch.Tensor], pad_to: int = 1) -> torch.Tensor:
    flat = torch.cat([t.reshape(-1) for t in tensors]) if len(tensors) > 1 else tensors[0].reshape(-1).clone()
    if pad_to > 1 and flat.numel() % pad_to:
        flat = torch.cat([flat, flat.new_zeros(pad_to - flat.numel() % pad_to)])
    return flat


def unflatten_dense(flat: torch.Tensor, like: Sequence[torch.Tensor]) -> List[torch.Tensor]:
    out, off = [], 0
    for t in like:
        n = t.numel()
        out.append(flat[off: off + n].view(t.shape))
        off += n
    return out


def _reorder(chunks_out: List[List[torch.Tensor]], chunks_in: List[List[torch.Tensor]], originals: Sequence[torch.Tensor]):
    pos = {id(t): i for i, t in enumerate(originals)}
    res: List[Optional[torch.Tensor]] = [None] * len(originals)
    for co, ci in zip(chunks_out, chunks_in):
        for o, i in zip(co, ci):
            res[pos[id(i)]] = o
    return res


def _pad_elems(dtype) -> int:
    return 16 // torch.empty((), dtype=dtype).element_size()


# ------------------------------------------------------------------------------------------------------
# process-group coalesced broadcast (DDP init sync and per-forward buffer sync)
# ------------------------------------------------------------------------------------------------------
def broadcast_coalesced_group(group, tensors: Sequence[torch.Tensor], bucket_bytes: int, src: int = 0) -> None:
    """In-place broadcast of ``tensors`` from group-rank ``src`` in coalesced chunks of <= bucket_bytes."""
    if not tensors:
        return
    gsrc = dist.get_global_rank(group, src) if group is not None and group is not dist.group.WORLD else src
    with torch.no_grad():
        for chunk in take_tensors(tensors, bucket_bytes):
            flat = flatten_dense(chunk)
            dist.broadcast(flat, src=gsrc, group=group)
            for t, f in zip(chunk, unflatten_dense(flat, chunk)):
                t.copy_(f)


# ------------------------------------------------------------------------------------------------------
# single-process, multi-device collectives
# ------------------------------------------------------------------------------------------------------
def _all_cuda(devs) -> bool:
    return all(d.type == "cuda" for d in devs)


_peer_enabled = set()
_arena_tls = threading.local()  # .arena: MulticastArena used by broadcast_coalesced while THIS thread's DataParallel forward replicates


class use_multicast_arena:
    """Route ``broadcast_coalesced`` destination buffers into an NVLS multicast arena (one multimem.st pass per chunk)."""

    def __init__(self, arena):
        self.arena = arena

    def __enter__(self):
        self._old = getattr(_arena_tls, "arena", None)
        _arena_tls.arena = self.arena
        if self.arena is not None:
            self.arena.reset()
        return self

    def __exit__(self, *exc):
        _arena_tls.arena = self._old
        return False



def _ensure_peer_access(devs: Sequence[torch.device]):
    key = tuple(sorted(d.index for d in devs))
    if key in _peer_enabled:
        return
    lib = _native.kernels()
    for d in devs:
        with torch.cuda.device(d):
            for e in devs:
                if e != d:
                    # peer access is switched on through torch (a tiny device-to-device copy in each direction) so that
                    # torch's allocator and our kernels share one consistent view of which pairs are enabled
                    torch.zeros(4, device=d).to(e)
                    torch.zeros(4, device=e).to(d)
    for d in devs:
        torch.cuda.synchronize(d)
    _peer_enabled.add(key)


def broadcast_coalesced(tensors: Sequence[torch.Tensor], devices: Sequence, buffer_size: int = DEFAULT_BUFFER_BYTES,
                        bf16_copy: bool = False) -> List[List[torch.Tensor]]:
    """Copies of ``tensors`` (all on ``devices[0]``) on every device; entry 0 aliases the inputs."""
    devs = [_dev(d) for d in devices]
    tensors = list(tensors)
    if not tensors:
        return [[] for _ in devs]
    outputs: List[List[torch.Tensor]] = [list(tensors)] + [[None] * len(tensors) for _ in devs[1:]]
    if len(devs) == 1:
        return outputs
    with torch.no_grad():  # a raw data movement op: differentiability is added by replicate.Broadcast
        return _broadcast_coalesced_impl(tensors, devs, outputs, buffer_size)


def _broadcast_coalesced_impl(tensors, devs, outputs, buffer_size):
    src_dev = tensors[0].device
    chunks = list(take_tensors(tensors, buffer_size))
    pos = {id(t): i for i, t in enumerate(tensors)}
    if src_dev.type == "cuda" and _all_cuda(devs):
        _ensure_peer_access(devs)
        lib = _native.kernels()
        from .symm import ptr_array

        src_stream = torch.cuda.current_stream(src_dev)
        # destinations must be idle before they are overwritten by the source-side kernel
        for d in devs[1:]:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(d))
            src_stream.wait_event(ev)
        with torch.cuda.device(src_dev):
            arena = getattr(_arena_tls, "arena", None)
            if arena is not None and [d.index for d in devs] != arena.devices[: len(devs)]:
                arena = None
            for chunk in chunks:
                flat = flatten_dense(chunk, _pad_elems(chunk[0].dtype))
                nbytes = flat.numel() * flat.element_size()
                off = arena.take(nbytes) if (arena is not None and len(devs) == len(arena.devices)) else None
                if off is not None:
                    # NVLS: one multimem.st pass lands in every replica's copy of the arena
                    dsts = [arena.views[k + 1][off: off + nbytes].view(flat.dtype) for k in range(len(devs) - 1)]
                    _native.check(lib.dmp_broadcast(flat.data_ptr(), None, 0, arena.mc_ptr + off, None, nbytes, None, 0, len(devs), 1,
                                                    0, src_stream.cuda_stream), "broadcast (multicast)")
                    stats["multicast_chunks"] = stats.get("multicast_chunks", 0) + 1
                else:
                    dsts = [torch.empty(flat.numel(), dtype=flat.dtype, device=d) for d in devs[1:]]
                    arr = ptr_array([t.data_ptr() for t in dsts])
                    _native.check(lib.dmp_broadcast(flat.data_ptr(), arr, len(dsts), None, None, nbytes, None, 0, len(devs), 1, 0,
                                                    src_stream.cuda_stream), "broadcast")
                    stats["peer_store_chunks"] = stats.get("peer_store_chunks", 0) + 1
                for k, dflat in enumerate(dsts):
                    for t, v in zip(chunk, unflatten_dense(dflat, chunk)):
                        outputs[k + 1][pos[id(t)]] = v
                flat.record_stream(src_stream)
        done = torch.cuda.Event()
        done.record(src_stream)
        for d in devs[1:]:
            torch.cuda.current_stream(d).wait_event(done)
        return outputs
    for k, d in enumerate(devs[1:]):
        for chunk in chunks:
            flat = flatten_dense(chunk).to(d, copy=True)
            for t, v in zip(chunk, unflatten_dense(flat, chunk)):
                outputs[k + 1][pos[id(t)]] = v
    return outputs


def reduce_add_coalesced(inputs: Sequence[Sequence[torch.Tensor]], destination=None, buffer_size: int = DEFAULT_BUFFER_BYTES,
                         scale: float = 1.0) -> List[torch.Tensor]:
    """Element-wise sum over devices of lists of tensors; result on ``destination`` (default: first input's device)."""
    n_dev = len(inputs)
    if n_dev == 0:
        return []
    ref = list(inputs[0])
    dest = _dev(destination) if destination is not None else ref[0].device
    if n_dev == 1 and scale == 1.0:
        return [t.to(dest) for t in ref]
    devs = [list(inp)[0].device for inp in inputs]
    pos = {id(t): i for i, t in enumerate(ref)}
    result: List[Optional[torch.Tensor]] = [None] * len(ref)
    ref_chunks = list(take_tensors(ref, buffer_size))
    index_chunks = [[pos[id(t)] for t in c] for c in ref_chunks]
    if dest.type == "cuda" and _all_cuda(devs) and all(c[0].dtype in (torch.float32, torch.bfloat16) for c in ref_chunks):
        _ensure_peer_access(list({d for d in devs} | {dest}))
        lib = _native.kernels()
        from .symm import DTYPE_CODE, ptr_array

        dst_stream = torch.cuda.current_stream(dest)
        flats_per_chunk = []
        for idxs in index_chunks:
            flats = []
            for k, inp in enumerate(inputs):
                inp = list(inp)
                with torch.cuda.device(devs[k]):
                    f = flatten_dense([inp[i] for i in idxs], _pad_elems(inp[idxs[0]].dtype))
                    ev = torch.cuda.Event()
                    ev.record(torch.cuda.current_stream(devs[k]))
                dst_stream.wait_event(ev)
                flats.append(f)
            flats_per_chunk.append(flats)
        with torch.cuda.device(dest):
            for idxs, flats in zip(index_chunks, flats_per_chunk):
                out = torch.empty(flats[0].numel(), dtype=torch.float32, device=dest)
                arr = ptr_array([f.data_ptr() for f in flats])
                _native.check(lib.dmp_reduce_add_peer(DTYPE_CODE[flats[0].dtype], arr, len(flats), flats[0].numel(), float(scale),
                                                      out.data_ptr(), None, None, None, None, 0, 0, dst_stream.cuda_stream),
                              "reduce_add_peer")
                like = [ref[i] for i in idxs]
                for i, v in zip(idxs, unflatten_dense(out, like)):
                    result[i] = v if ref[i].dtype == torch.float32 else v.to(ref[i].dtype)
        done = torch.cuda.Event()
        done.record(dst_stream)
        for d in set(devs):
            if d != dest:
                torch.cuda.current_stream(d).wait_event(done)  # sources may be reused only after the reduction read them
        return result
    for idxs in index_chunks:
        acc = None
        for inp in inputs:
            inp = list(inp)
            f = flatten_dense([inp[i] for i in idxs]).to(dest)
            acc = f.clone() if acc is None else acc.add_(f)
        if scale != 1.0:
            acc.mul_(scale)
        for i, v in zip(idxs, unflatten_dense(acc, [ref[i] for i in idxs])):
            result[i] = v
    return result


def reduce_add(inputs: Sequence[torch.Tensor], destination=None) -> torch.Tensor:
    return reduce_add_coalesced([[t] for t in inputs], destination)[0]


def scatter(tensor: torch.Tensor, devices: Sequence, chunk_sizes: Optional[Sequence[int]] = None, dim: int = 0,
            streams=None) -> List[torch.Tensor]:
    """Split ``tensor`` along ``dim`` into ``len(devices)`` chunks (``torch.chunk`` sizes unless given) placed on the devices."""
    devs = [_dev(d) for d in devices]
    if chunk_sizes is None:
        chunks = list(tensor.chunk(len(devs), dim))
    else:
        if sum(chunk_sizes) != tensor.size(dim):
            raise ValueError("chunk_sizes must sum to the scattered dimension")
        chunks = list(tensor.split(list(chunk_sizes), dim))
    outs = []
    src = tensor.device
    if src.type == "cuda":
        ready = torch.cuda.Event()
        ready.record(torch.cuda.current_stream(src))
    for c, d in zip(chunks, devs):
        if d == src:
            outs.append(c)
            continue
        if src.type == "cuda" and d.type == "cuda" and c.is_contiguous() and (c.numel() * c.element_size()) % 16 == 0 \
                and c.data_ptr() % 16 == 0:
            _ensure_peer_access([src, d])
            with torch.cuda.device(d):
                st = torch.cuda.current_stream(d)
                st.wait_event(ready)
                o = torch.empty_like(c, device=d)
                _native.check(_native.kernels().dmp_copy_peer(c.data_ptr(), o.data_ptr(), c.numel() * c.element_size(), 0,
                                                              st.cuda_stream), "scatter copy_peer")
                c.record_stream(st)
            outs.append(o)
        elif src.type == "cpu" and d.type == "cuda":
            outs.append(c.contiguous().pin_memory().to(d, non_blocking=True) if torch.cuda.is_available() else c.to(d))
        else:
            outs.append(c.to(d))
    return outs


def gather(tensors: Sequence[torch.Tensor], dim: int = 0, destination=None) -> torch.Tensor:
    """Concatenate per-device tensors along ``dim`` on ``destination`` (default: the first tensor's device)."""
    tensors = list(tensors)
    dest = _dev(destination) if destination is not None else tensors[0].device
    if dest.type == "cuda" and all(t.is_cuda for t in tensors) and dim == 0 and \
            all(t.is_contiguous() and (t.numel() * t.element_size()) % 16 == 0 and t.data_ptr() % 16 == 0 for t in tensors):
        devs = [t.device for t in tensors]
        _ensure_peer_access(list(set(devs) | {dest}))
        from .symm import ptr_array
        import ctypes as C

        dst_stream = torch.cuda.current_stream(dest)
        for t in tensors:
            if t.device != dest:
                ev = torch.cuda.Event()
                with torch.cuda.device(t.device):
                    ev.record(torch.cuda.current_stream(t.device))
                dst_stream.wait_event(ev)
        shape = list(tensors[0].shape)
        shape[0] = sum(t.shape[0] for t in tensors)
        with torch.cuda.device(dest):
            out = torch.empty(shape, dtype=tensors[0].dtype, device=dest)
            arr = ptr_array([t.data_ptr() for t in tensors])
            nb = (C.c_longlong * len(tensors))(*[t.numel() * t.element_size() for t in tensors])
            _native.check(_native.kernels().dmp_gather_peer(arr, nb, len(tensors), out.data_ptr(), 0, dst_stream.cuda_stream),
                          "gather_peer")
        done = torch.cuda.Event()
        done.record(dst_stream)
        for t in tensors:
            if t.device != dest:
                torch.cuda.current_stream(t.device).wait_event(done)
        return out
    return torch.cat([t.to(dest) for t in tensors], dim)
