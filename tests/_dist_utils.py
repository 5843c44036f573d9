"""Helpers for multi-process CPU tests: gloo, world_size > 1, rendezvous on 127.0.0.1."""
import os
import socket
import sys
import traceback

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _entry(rank, world, port, fn, args):
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    torch.set_num_threads(2)
    import faulthandler
    faulthandler.dump_traceback_later(240, exit=True)
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, world_size=world, rank=rank)
    try:
        fn(rank, world, *args)
        dist.barrier()
    finally:
        dist.destroy_process_group()


def run_gloo(fn, world=2, *args):
    """Runs ``fn(rank, world, *args)`` in ``world`` gloo processes; raises if any of them fails."""
    for attempt in range(3):  # the probed port can be taken again before rank 0 listens on it: pick another one and retry
        try:
            mp.spawn(_entry, args=(world, free_port(), fn, args), nprocs=world, join=True)
            return
        except Exception as e:  # noqa: BLE001
            if "EADDRINUSE" not in str(e) or attempt == 2:
                raise
