"""Checkpoint / resume.

* ``save_best`` / ``load_checkpoint`` follow the reference DP script: on a new best validation accuracy write
  ``{'net': state_dict, 'acc': acc, 'epoch': epoch}`` to ``./checkpoint/ckpt.pth``; ``--resume`` restores weights,
  best accuracy and start epoch (/root/reference/code/distributed_training/data_parallel.py:80-87,143-155).  Keys keep
  the ``module.`` prefix of the DataParallel / DistributedDataParallel wrapper.  Optimizer / scheduler / warm-up
  state are saved too when given (the reference drops them, SURVEY.md defect Q11).
* Pipeline stages save one shard per stage; ``merge_pipeline_shards`` maps them back to the single-model key space
  (the reference's model-parallel script has no checkpointing at all).
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional

import torch


def save_best(net, acc: float, epoch: int, path: str = "./checkpoint/ckpt.pth", optimizer=None, scheduler=None, warmup=None) -> str:
    state = {"net": net.state_dict(), "acc": acc, "epoch": epoch}
    if optimizer is not None:
        state["optimizer"] = optimizer.state_dict()
    if scheduler is not None:
        state["scheduler"] = scheduler.state_dict()
    if warmup is not None:
        state["warmup"] = warmup.state_dict()
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)
    torch.save(state, path)
    return path


def load_checkpoint(net, path: str = "./checkpoint/ckpt.pth", optimizer=None, scheduler=None, warmup=None, map_location="cpu"):
    """Returns (best_acc, start_epoch)."""
    if not os.path.isfile(path):
        raise FileNotFoundError("Error: no checkpoint found at %s" % path)
    ckpt = torch.load(path, map_location=map_location, weights_only=False)
    sd = ckpt["net"]
    own = net.state_dict().keys()
    if all(k.startswith("module.") for k in sd) and not any(k.startswith("module.") for k in own):
        sd = {k[len("module."):]: v for k, v in sd.items()}          # wrapped checkpoint into a bare module
    elif not any(k.startswith("module.") for k in sd) and all(k.startswith("module.") for k in own):
        sd = {"module." + k: v for k, v in sd.items()}               # bare checkpoint into a wrapper
    net.load_state_dict(sd)
    if optimizer is not None and "optimizer" in ckpt:
        optimizer.load_state_dict(ckpt["optimizer"])
    if scheduler is not None and "scheduler" in ckpt:
        scheduler.load_state_dict(ckpt["scheduler"])
    if warmup is not None and "warmup" in ckpt:
        warmup.load_state_dict(ckpt["warmup"])
    return ckpt.get("acc", 0.0), ckpt.get("epoch", 0)


def save_pipeline_shard(stage_module, stage: int, key_map: Dict[str, str], directory: str, epoch: int, acc: float = 0.0) -> str:
    """``key_map``: stage-local state_dict key -> key in the full model (see ``stage_key_map``)."""
    os.makedirs(directory, exist_ok=True)
    sd = {key_map.get(k, k): v for k, v in stage_module.state_dict().items()}
    path = os.path.join(directory, "stage%d.pth" % stage)
    torch.save({"net": sd, "epoch": epoch, "acc": acc, "stage": stage}, path)
    return path


def merge_pipeline_shards(directory: str, num_stages: int) -> Dict[str, torch.Tensor]:
    merged: Dict[str, torch.Tensor] = {}
    for s in range(num_stages):
        merged.update(torch.load(os.path.join(directory, "stage%d.pth" % s), map_location="cpu", weights_only=False)["net"])
    return merged


def stage_key_map(full_model, stage_module) -> Dict[str, str]:
    """Maps a stage's state_dict keys to the full model's by tensor identity (stages share the model's tensors)."""
    by_id = {}
    for k, v in list(full_model.named_parameters()) + list(full_model.named_buffers()):
        by_id[id(v)] = k
    out = {}
    for k, v in list(stage_module.named_parameters()) + list(stage_module.named_buffers()):
        if id(v) in by_id:
            out[k] = by_id[id(v)]
    return out
