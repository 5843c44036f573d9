"""Linear learning-rate warm-up that composes with any torch LR scheduler.

Stands in for ``pytorch_warmup.LinearWarmup(optimizer, warmup_period=10)`` + ``dampen()``, which the reference
calls once per epoch next to ``CosineAnnealingLR`` (/root/reference/code/distributed_training/
model_parallel.py:107-108,117-118; data_parallel.py:93-96,163-164) and which is not installable here.
The schedule: after the k-th ``dampen()`` the scheduler's learning rate is multiplied by ``min(1, (k+1)/period)``.
"""
from __future__ import annotations


class LinearWarmup:
    def __init__(self, optimizer, warmup_period: int = 10, last_step: int = -1):
        if warmup_period < 1:
            raise ValueError("warmup_period must be >= 1")
        self.optimizer = optimizer
        self.warmup_period = int(warmup_period)
        self.last_step = last_step
        self._undampened = None
        self.dampen()

    def factor(self, step: int) -> float:
        return min(1.0, (step + 1) / self.warmup_period)

    def dampen(self, step: int | None = None):
        """Apply the warm-up factor to the learning rates currently set by the wrapped scheduler."""
        self.last_step = self.last_step + 1 if step is None else step
        f = self.factor(self.last_step)
        for g in self.optimizer.param_groups:
            g["lr"] = g["lr"] * f
        return f

    def state_dict(self):
        return {"warmup_period": self.warmup_period, "last_step": self.last_step}

    def load_state_dict(self, sd):
        self.warmup_period, self.last_step = sd["warmup_period"], sd["last_step"]
