"""Per-epoch logging: the reference's append-only text line plus a structured JSONL record.

Reference format (/root/reference/code/distributed_training/model_parallel.py:119-124, data_parallel.py:167-171):
``step:<e>  loss_train:<..>  acc1_train:<..>  loss_val:<..>  acc1_val:<..>  [time_per_batch:<..>  time_load_perbatch:<..>]``.
Unlike the reference the log directory is created on demand (its ``open(...,'a')`` needs ./log to pre-exist)."""
from __future__ import annotations

import json
import os
import time
from typing import Optional


class EpochLogger:
    def __init__(self, text_path: Optional[str] = None, jsonl_path: Optional[str] = None, rank: int = 0):
        self.text_path, self.jsonl_path, self.rank = text_path, jsonl_path, rank
        for p in (text_path, jsonl_path):
            if p and os.path.dirname(p):
                os.makedirs(os.path.dirname(p), exist_ok=True)

    def log_epoch(self, epoch: int, loss_train: float, acc1_train: float, loss_val: float, acc1_val: float,
                  time_per_batch: Optional[float] = None, time_load_perbatch: Optional[float] = None, **extra):
        line = "\n" + "step:%s  loss_train:%s  acc1_train:%s  loss_val:%s  acc1_val:%s" % (epoch, loss_train, acc1_train, loss_val, acc1_val)
        if time_per_batch is not None:
            line += "  time_per_batch:%s  time_load_perbatch:%s" % (time_per_batch, time_load_perbatch)
        if self.text_path:
            with open(self.text_path, "a") as f:
                f.write(line)
        if self.jsonl_path:
            rec = dict(ts=time.time(), rank=self.rank, epoch=epoch, loss_train=loss_train, acc1_train=acc1_train, loss_val=loss_val,
                       acc1_val=acc1_val, time_per_batch=time_per_batch, time_load_perbatch=time_load_perbatch, **extra)
            with open(self.jsonl_path, "a") as f:
                f.write(json.dumps(rec) + "\n")
        return line.strip()

    def log_step(self, **rec):
        if self.jsonl_path:
            with open(self.jsonl_path, "a") as f:
                f.write(json.dumps(dict(ts=time.time(), rank=self.rank, **rec)) + "\n")


def get_lr(optimizer) -> float:
    for g in optimizer.param_groups:
        return g["lr"]
    return 0.0
