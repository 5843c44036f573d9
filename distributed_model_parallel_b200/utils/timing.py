"""Device-side timing (the reference times with an unsynchronised host clock that includes the data loader,
/root/reference/code/distributed_training/utils.py:41-72; SURVEY.md defect Q9)."""
from __future__ import annotations

import time

import torch


class StepTimer:
    """CUDA-event timer when a GPU is present, ``perf_counter`` otherwise.  ``max_over_ranks`` reduces with MAX."""

    def __init__(self, device=None):
        self.cuda = torch.cuda.is_available() and (device is None or torch.device(device).type == "cuda")
        self._t0 = None
        if self.cuda:
            self._e0, self._e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def start(self):
        if self.cuda:
            self._e0.record()
        else:
            self._t0 = time.perf_counter()

    def stop_ms(self) -> float:
        if self.cuda:
            self._e1.record()
            self._e1.synchronize()
            return self._e0.elapsed_time(self._e1)
        return (time.perf_counter() - self._t0) * 1e3

    @staticmethod
    def max_over_ranks(ms: float, group=None) -> float:
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()):
            return ms
        dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
        return float(t.item())
