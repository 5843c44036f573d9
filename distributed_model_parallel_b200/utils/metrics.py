"""Accuracy and running-average helpers (reference: accuracy() at
/root/reference/code/distributed_training/utils.py:215-229 and the running sums of data_parallel.py:112-120)."""
from __future__ import annotations

from typing import List, Sequence

import torch


@torch.no_grad()
def accuracy(output: torch.Tensor, target: torch.Tensor, topk: Sequence[int] = (1,)) -> List[torch.Tensor]:
    """Top-k accuracy in percent, one 1-element tensor per k (same contract as the reference helper)."""
    maxk = min(max(topk), output.shape[1])
    batch = target.size(0)
    _, pred = output.float().topk(maxk, 1, True, True)
    correct = pred.t().eq(target.view(1, -1).expand(maxk, batch))
    return [correct[: min(k, maxk)].reshape(-1).float().sum(0, keepdim=True).mul_(100.0 / batch) for k in topk]


class AverageMeter:
    """Running mean that keeps device tensors on the device (no .item() per iteration; defect K16 of the survey)."""

    def __init__(self):
        self.sum, self.count = 0.0, 0

    def update(self, value, n: int = 1):
        self.sum = self.sum + (value.detach() if isinstance(value, torch.Tensor) else value) * n
        self.count += n

    @property
    def avg(self) -> float:
        if self.count == 0:
            return 0.0
        s = self.sum
        return float(s.item() if isinstance(s, torch.Tensor) else s) / self.count
