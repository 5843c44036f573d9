"""Peer-addressable ("symmetric") memory for the communication kernels.

Two flavours, same interface (``SymmBuffer.local`` tensor + ``peer_ptrs`` for every rank):

* ``ProcessSymmGroup``  -- one process per GPU (DistributedDataParallel, SyncBatchNorm).  Buffers are raw
  ``cudaMalloc`` allocations exported with CUDA IPC; handles travel through the ``torch.distributed`` store
  (control plane only) and every rank maps every peer's buffer into its own address space.
* ``DeviceSymmGroup``   -- one process driving several GPUs (DataParallel).  Buffers are ordinary per-device
  tensors; ``cudaDeviceEnablePeerAccess`` makes them addressable from kernels on the other devices.

NVLink/NVSwitch does the rest: kernels in ``csrc/comm/collectives.cu`` dereference the peer pointers
directly.  Nothing here has a counterpart in the reference, whose DP/DDP paths delegate all communication
to torch/NCCL.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import torch
import torch.distributed as dist

from .. import _native

# (the C signatures of the communication entry points are declared centrally in _native.py)

DTYPE_CODE = {torch.float32: 0, torch.bfloat16: 1, torch.float64: 2}


def ptr_array(ptrs: Sequence[int]):
    """ctypes ``void*[n]`` from integer addresses (kept alive by the caller)."""
    arr = (C.c_void_p * len(ptrs))()
    for i, p in enumerate(ptrs):
        arr[i] = p
    return arr


class _RawCuda:
    """Exposes a raw device allocation through ``__cuda_array_interface__`` so torch can alias it."""

    def __init__(self, ptr: int, nbytes: int, owner):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}
        self._owner = owner


class SymmBuffer:
    """``nbytes`` of device memory on every rank; ``local`` aliases this rank's copy, ``peer_ptrs[r]`` is the
    address of rank r's copy as seen from this rank."""

    def __init__(self, local: torch.Tensor, peer_ptrs: List[int], rank: int, owner=None):
        self.local = local
        self.peer_ptrs = list(peer_ptrs)
        self.rank = rank
        self.nbytes = local.numel() * local.element_size()
        self._ptr_array = ptr_array(self.peer_ptrs)
        self._owner = owner

    @property
    def ptrs(self):
        return self._ptr_array

    def view(self, dtype: torch.dtype, numel: Optional[int] = None, byte_offset: int = 0) -> torch.Tensor:
        t = self.local.view(torch.uint8)[byte_offset:]
        if numel is not None:
            t = t[: numel * torch.empty((), dtype=dtype).element_size()]
        return t.view(dtype)

    def sub(self, byte_offset: int, nbytes: int) -> "SymmBuffer":
        loc = self.local.view(torch.uint8)[byte_offset: byte_offset + nbytes]
        return SymmBuffer(loc, [p + byte_offset for p in self.peer_ptrs], self.rank, owner=self)


class _IpcAllocation:
    def __init__(self, lib, ptr, imported):
        self.lib, self.ptr, self.imported = lib, ptr, imported

    def __del__(self):
        try:
            for p in self.imported:
                self.lib.dmp_ipc_close(p)
            self.lib.dmp_symm_free(self.ptr)
        except Exception:
            pass


class ProcessSymmGroup:
    """Symmetric memory across the ranks of a ``torch.distributed`` process group (one process per GPU)."""

    def __init__(self, group=None, device: Optional[torch.device] = None):
        self.group = group if group is not None else dist.group.WORLD
        self.rank = dist.get_rank(self.group)
        self.world_size = dist.get_world_size(self.group)
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.lib = _native.kernels()
        if self.world_size > self.lib.dmp_comm_max_ranks():
            raise RuntimeError("world size %d exceeds the %d ranks the peer kernels support" %
                               (self.world_size, self.lib.dmp_comm_max_ranks()))

    def alloc(self, nbytes: int) -> SymmBuffer:
        nbytes = max(256, (int(nbytes) + 255) // 256 * 256)
        with torch.cuda.device(self.device):
            p = C.c_void_p()
            _native.check(self.lib.dmp_symm_malloc(nbytes, C.byref(p)), "symm_malloc")
            hbytes = self.lib.dmp_ipc_handle_bytes()
            handle = C.create_string_buffer(hbytes)
            _native.check(self.lib.dmp_ipc_export(p, handle), "ipc_export")
            gathered = [None] * self.world_size
            dist.all_gather_object(gathered, (bytes(handle.raw), int(self.device.index or 0)), group=self.group)
            ptrs, imported = [], []
            for r, (h, _dev) in enumerate(gathered):
                if r == self.rank:
                    ptrs.append(p.value)
                    continue
                q = C.c_void_p()
                hb = C.create_string_buffer(h, hbytes)
                _native.check(self.lib.dmp_ipc_import(hb, C.byref(q)), "ipc_import")
                ptrs.append(q.value)
                imported.append(q.value)
            owner = _IpcAllocation(self.lib, p.value, imported)
            local = torch.as_tensor(_RawCuda(p.value, nbytes, owner), device=self.device)
        dist.barrier(group=self.group)
        return SymmBuffer(local, ptrs, self.rank, owner=owner)

    def new_state(self) -> SymmBuffer:
        """Fresh communicator state (flag pad + epochs + status) -- one per ordered stream of collectives."""
        return self.alloc(int(self.lib.dmp_comm_state_bytes()))


def comm_status_words(state: SymmBuffer) -> torch.Tensor:
    """The four trailing uint32 words of a communicator state: [0] status (0 healthy, 1 + peer = that peer never arrived),
    [1] spin timeout in milliseconds (0 = built-in default), [2..3] reserved."""
    words = int(_native.kernels().dmp_comm_state_bytes()) // 4
    return state.local.view(torch.int32)[words - 4: words]


def set_comm_timeout_ms(state: SymmBuffer, ms: int) -> None:
    """Bounds every cross-GPU spin-wait of the collectives that use ``state`` (fault detection: a dead or diverged peer
    turns into a status code within ``ms`` instead of a hung kernel).  The reference has no timeout anywhere
    (model_parallel.py:57-58 passes none to init_process_group)."""
    comm_status_words(state)[1] = int(ms)


def reset_comm_status(state: SymmBuffer) -> None:
    comm_status_words(state)[0] = 0


class ProcessMulticast:
    """NVLS multicast buffer across the ranks of a process group (one process per GPU).

    ``local`` is a uint8 tensor over this rank's copy; ``mc_ptr`` is the multicast address: ``multimem.ld_reduce`` on
    ``mc_ptr + off`` returns the switch-side sum of every rank's bytes at ``off``, ``multimem.st`` writes all copies at once.
    Rank 0 creates the driver object and passes its POSIX file descriptor to the peers over an abstract-namespace Unix
    socket (SCM_RIGHTS); the socket name travels through ``torch.distributed``.  Raises RuntimeError if the platform has no
    multicast support -- callers keep the peer-load kernels in that case."""

    def __init__(self, nbytes: int, group=None, device: Optional[torch.device] = None):
        import os
        import socket
        import uuid

        self.group = group if group is not None else dist.group.WORLD
        self.rank = dist.get_rank(self.group)
        self.world_size = dist.get_world_size(self.group)
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.lib = _native.kernels()
        dev = int(self.device.index or 0)
        ok = bool(self.lib.dmp_mc_supported((C.c_int * 1)(dev), 1))
        oks = [None] * self.world_size
        dist.all_gather_object(oks, ok, group=self.group)
        if not all(oks):
            raise RuntimeError("NVLS multicast is not supported on every rank's device")
        handle, size = C.c_void_p(), C.c_longlong()
        src = dist.get_global_rank(self.group, 0) if self.group is not dist.group.WORLD else 0
        if self.rank == 0:
            fd = C.c_int(-1)
            _native.check(self.lib.dmp_mc_xp_create(dev, self.world_size, int(nbytes), C.byref(handle), C.byref(fd), C.byref(size)),
                          "mc_xp_create")
            name = "\0dmp_mc_%d_%s" % (os.getpid(), uuid.uuid4().hex)
            srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            srv.bind(name)
            srv.listen(self.world_size)
            dist.broadcast_object_list([name], src=src, group=self.group)
            srv.settimeout(60.0)
            for _ in range(self.world_size - 1):
                conn, _addr = srv.accept()
                socket.send_fds(conn, [b"fd"], [fd.value])
                conn.recv(1)  # the peer has the descriptor: safe to close ours afterwards
                conn.close()
            srv.close()
            os.close(fd.value)
        else:
            box = [None]
            dist.broadcast_object_list(box, src=src, group=self.group)
            cli = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            cli.settimeout(60.0)
            cli.connect(box[0])
            _msg, fds, _flags, _addr = socket.recv_fds(cli, 16, 1)
            cli.send(b"k")
            cli.close()
            _native.check(self.lib.dmp_mc_xp_import(dev, int(fds[0]), self.world_size, int(nbytes), C.byref(handle), C.byref(size)),
                          "mc_xp_import")
            os.close(fds[0])
        self._handle = handle
        _native.check(self.lib.dmp_mc_xp_add_device(handle), "mc_xp_add_device")
        dist.barrier(group=self.group)  # every device must be part of the object before the first bind
        uc, mc = C.c_void_p(), C.c_void_p()
        _native.check(self.lib.dmp_mc_xp_bind_map(handle, C.byref(uc), C.byref(mc)), "mc_xp_bind_map")
        dist.barrier(group=self.group)
        self.size = int(size.value)
        self.mc_ptr = int(mc.value)
        self.local = torch.as_tensor(_RawCuda(uc.value, self.size, self), device=self.device)

    def view(self, dtype: torch.dtype, numel: int, byte_offset: int = 0) -> torch.Tensor:
        esz = torch.empty((), dtype=dtype).element_size()
        return self.local[byte_offset: byte_offset + numel * esz].view(dtype)

    def __del__(self):
        try:
            if getattr(self, "_handle", None):
                self.lib.dmp_mc_xp_destroy(self._handle)
                self._handle = None
        except Exception:
            pass


class DeviceSymmGroup:
    """Peer-addressable buffers across several devices of one process (DataParallel)."""

    def __init__(self, devices: Sequence[int]):
        self.devices = [int(d) for d in devices]
        self.world_size = len(self.devices)
        self.lib = _native.kernels()
        for d in self.devices:
            with torch.cuda.device(d):
                for e in self.devices:
                    if e != d:
                        torch.zeros(4, device=torch.device("cuda", d)).to(torch.device("cuda", e))  # torch enables the pair
                        torch.zeros(4, device=torch.device("cuda", e)).to(torch.device("cuda", d))

    def alloc(self, nbytes: int) -> List[torch.Tensor]:
        nbytes = max(256, (int(nbytes) + 255) // 256 * 256)
        return [torch.zeros(nbytes, dtype=torch.uint8, device=torch.device("cuda", d)) for d in self.devices]


class MulticastArena:
    """NVLS multicast-backed replica storage for one process driving several GPUs.

    ``views[d]`` is a uint8 tensor over device d's copy, ``mc_ptr`` the multicast address: a ``multimem.st`` issued by ONE GPU
    to ``mc_ptr + off`` lands at offset ``off`` of every device's copy.  ``take(nbytes)`` hands out offsets bump-pointer style;
    ``reset()`` rewinds (DataParallel rewinds at the start of every forward, when the previous replicas are dead)."""

    def __init__(self, devices: Sequence[int], nbytes: int):
        self.lib = _native.kernels()
        self.devices = [int(d) for d in devices]
        n = len(self.devices)
        devs = (C.c_int * n)(*self.devices)
        if not self.lib.dmp_mc_supported(devs, n):
            raise RuntimeError("NVLS multicast is not supported on these devices")
        handle, mc, size = C.c_void_p(), C.c_void_p(), C.c_longlong()
        uc = (C.c_void_p * n)()
        _native.check(self.lib.dmp_mc_create(devs, n, int(nbytes), C.byref(handle), uc, C.byref(mc), C.byref(size)), "mc_create")
        self._handle, self.mc_ptr, self.size = handle, mc.value, int(size.value)
        self.views = [torch.as_tensor(_RawCuda(uc[i], self.size, self), device=torch.device("cuda", d))
                      for i, d in enumerate(self.devices)]
        self._off = 0

    def reset(self):
        self._off = 0

    def take(self, nbytes: int) -> Optional[int]:
        nbytes = (int(nbytes) + 255) // 256 * 256
        if self._off + nbytes > self.size:
            return None
        off = self._off
        self._off += nbytes
        return off

    def __del__(self):
        try:
            if self._handle:
                self.lib.dmp_mc_destroy(self._handle)
                self._handle = None
        except Exception:
            pass
