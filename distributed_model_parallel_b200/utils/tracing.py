"""NVTX ranges around the framework's phases: one per gradient bucket launch, per pipeline micro-batch forward/backward,
per DataParallel phase (scatter / replicate / parallel_apply / gather) and per captured train step.

The reference has no tracing at all beyond host ``time.time()`` deltas (utils.py:41-72); torch's wrappers mark their forward
with ``record_function`` (torch/nn/parallel/data_parallel.py:174, distributed.py:1694).  Ranges are emitted only when
``DMP_NVTX=1`` (or after ``enable(True)``): an NVTX push/pop is cheap but not free, and the steady state of this framework is
a replayed CUDA graph in which host-side ranges would mean nothing.  ``events`` additionally keeps the last ranges with
their host duration so that tests and the JSONL logger can see them without a profiler attached.
"""
from __future__ import annotations

import contextlib
import os
import time
from collections import deque

import torch

_enabled = os.environ.get("DMP_NVTX", "0") == "1"
events: deque = deque(maxlen=4096)  # (name, host milliseconds)


def enable(on: bool = True) -> None:
    global _enabled
    _enabled = bool(on)


def enabled() -> bool:
    return _enabled


@contextlib.contextmanager
def nvtx_range(name: str):
    if not _enabled:
        yield
        return
    cuda = torch.cuda.is_available()
    if cuda:
        torch.cuda.nvtx.range_push(name)
    t0 = time.perf_counter()
    try:
        yield
    finally:
        events.append((name, (time.perf_counter() - t0) * 1e3))
        if cuda:
            torch.cuda.nvtx.range_pop()


def mark(name: str) -> None:
    if _enabled:
        events.append((name, 0.0))
        if torch.cuda.is_available():
            torch.cuda.nvtx.mark(name)
