"""Pieces shared by the DP / DDP / MP trainers: LR schedule, train/val epoch loops, device placement."""
from __future__ import annotations

import math
import time
from typing import Optional

import torch
import torch.nn.functional as F

from ..ops import functional as OF
from ..optim import LinearWarmup
from ..utils import AverageMeter, accuracy


class CosineWithWarmup:
    """``CosineAnnealingLR(T_max=epochs)`` stepped once per epoch + ``LinearWarmup(warmup_period)`` dampening, i.e. the
    reference's schedule (/root/reference/code/distributed_training/model_parallel.py:107-108,117-118), without the
    deprecated ``scheduler.step(epoch)`` form (SURVEY.md defect Q12)."""

    def __init__(self, optimizer, base_lr: float, epochs: int, warmup_period: int = 10):
        self.optimizer, self.base_lr, self.epochs = optimizer, base_lr, max(1, epochs)
        self.warmup = LinearWarmup(optimizer, warmup_period, last_step=-1) if warmup_period > 0 else None
        self.epoch = 0
        self._apply()

    def _apply(self):
        lr = 0.5 * self.base_lr * (1 + math.cos(math.pi * min(self.epoch, self.epochs) / self.epochs))
        if self.warmup is not None:
            lr *= self.warmup.factor(self.epoch)
        for g in self.optimizer.param_groups:
            g["lr"] = lr
        if hasattr(self.optimizer, "push_hyperparameters"):
            self.optimizer.push_hyperparameters()

    def step(self):
        self.epoch += 1
        self._apply()

    def state_dict(self):
        return {"epoch": self.epoch}

    def load_state_dict(self, sd):
        self.epoch = sd["epoch"]
        self._apply()


def limit(loader, steps: int):
    for i, batch in enumerate(loader):
        if steps and i >= steps:
            return
        yield batch


def train_epoch(step, loader, device, cfg, rank: int = 0):
    """One epoch through a ``TrainStep`` (fused loss/accuracy on device, one host read per print interval)."""
    loss_m, acc_m, batch_t, data_t = AverageMeter(), AverageMeter(), AverageMeter(), AverageMeter()
    t0 = time.perf_counter()
    n = 0
    for i, (x, y) in enumerate(limit(loader, cfg.steps_per_epoch)):
        data_t.update(time.perf_counter() - t0)
        stats = step(x, y)
        loss_m.update(stats[0].clone(), y.shape[0])
        acc_m.update(stats[1].clone() * (100.0 / y.shape[0]), y.shape[0])
        n += 1
        if rank == 0 and cfg.print_freq and i % cfg.print_freq == 0:
            s = step.read_stats()
            print("train_loss: %.4f train_acc1 %.2f" % (s[0], 100.0 * s[1] / y.shape[0]))
        batch_t.update(time.perf_counter() - t0)
        t0 = time.perf_counter()
    return loss_m.avg, acc_m.avg, batch_t.avg, data_t.avg


@torch.no_grad()
def validate(model, loader, device, cfg):
    model.eval()
    loss_m, acc_m = AverageMeter(), AverageMeter()
    stats = torch.zeros(3, device=device)
    for x, y in limit(loader, cfg.steps_per_epoch):
        x, y = x.to(device, non_blocking=True), y.to(device, non_blocking=True)
        out = model(x)
        loss = OF.cross_entropy(out if out.dtype in (torch.float32, torch.bfloat16) else out.float(), y, stats)
        loss_m.update(loss.detach().clone(), y.shape[0])
        acc_m.update(stats[1].clone() * (100.0 / y.shape[0]), y.shape[0])
    model.train()
    return loss_m.avg, acc_m.avg
