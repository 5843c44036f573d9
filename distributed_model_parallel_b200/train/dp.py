"""Single-process DataParallel trainer -- the framework's counterpart of the reference's ``data_parallel.py``
(/root/reference/code/distributed_training/data_parallel.py:1-172): model on device 0 wrapped in
``DataParallel``, CrossEntropy, SGD(momentum 0.9, wd 1e-4), cosine schedule with 10-epoch linear warm-up,
best-accuracy checkpoint with ``--resume``, one log line per epoch in ``<log_dir>/data_para_<batch>.txt``."""
from __future__ import annotations

import os

import torch

from ..data import prepare_dataloader
from ..models import build_model
from ..optim import FlatFusedSGD
from ..parallel.data_parallel import DataParallel
from ..utils import EpochLogger, TrainConfig, get_lr, load_checkpoint, save_best
from .common import CosineWithWarmup, train_epoch, validate
from .engine import TrainStep


def main(cfg: TrainConfig):
    torch.manual_seed(cfg.seed)
    device = torch.device("cuda:0" if torch.cuda.is_available() else "cpu")
    _, train_loader, val_loader = prepare_dataloader(cfg)
    # always wrapped, like the reference (data_parallel.py:74-76): checkpoints then carry the ``module.`` key prefix whatever the
    # device count; with fewer than two GPUs the wrapper simply calls the module
    net = DataParallel(build_model(cfg.model).to(device))
    multi = device.type == "cuda" and torch.cuda.device_count() > 1
    optimizer = FlatFusedSGD(net.parameters(), lr=cfg.lr, momentum=cfg.momentum, weight_decay=cfg.weight_decay)
    sched = CosineWithWarmup(optimizer, cfg.lr, cfg.epochs, cfg.warmup_period)
    best_acc, start_epoch = -1.0, 0  # the first validated epoch always writes a checkpoint
    if cfg.resume:
        # the checkpoint is written after its epoch's validation, with the optimizer and scheduler state of that moment: go
        # on with the next epoch (the reference restarts AT the stored epoch, data_parallel.py:86-87 -- it has no such state)
        best_acc, last_epoch = load_checkpoint(net, cfg.checkpoint, optimizer, sched)
        start_epoch = last_epoch + 1
        optimizer.sync_shadow()
        print("==> resumed from %s (epoch %d, acc %.2f)" % (cfg.checkpoint, start_epoch, best_acc))
    # DataParallel spawns replica threads per forward, which a CUDA graph cannot capture: graphs only on one device
    step = TrainStep(net if multi else net.module, optimizer, use_graph=cfg.cuda_graph and not multi)
    log = EpochLogger(os.path.join(cfg.log_dir, "data_para_%d.txt" % cfg.batch_size), os.path.join(cfg.log_dir, "dp.jsonl"))
    for epoch in range(start_epoch, cfg.epochs):
        tl, ta, bt, dt = train_epoch(step, train_loader, device, cfg)
        vl, va = validate(net, val_loader, device, cfg)
        sched.step()
        if va > best_acc:
            print("Saving..")
            save_best(net, va, epoch, cfg.checkpoint, optimizer, sched)
            best_acc = va
        print(log.log_epoch(epoch, tl, ta, vl, va, bt, dt, lr=get_lr(optimizer)))
    return best_acc
