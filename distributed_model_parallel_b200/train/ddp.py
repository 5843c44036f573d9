"""Multi-process DistributedDataParallel trainer (one process per GPU).  NOT IN REFERENCE as code -- the reference
only studies DDP in its README (:10-15,145-157); this is north-star scope with the same recipe as the DP trainer.
Launch with ``torchrun`` or let ``cli.py ddp --world-size N`` spawn the workers on 127.0.0.1."""
from __future__ import annotations

import os

import torch
import torch.distributed as dist

from ..data import prepare_dataloader
from ..models import build_model
from ..optim import FlatFusedSGD
from ..parallel import DistributedDataParallel
from ..parallel.sync_batchnorm import convert_sync_batchnorm
from ..utils import EpochLogger, TrainConfig, get_lr, load_checkpoint, save_best
from .common import CosineWithWarmup, train_epoch, validate
from .engine import TrainStep


def worker(rank: int, world_size: int, cfg: TrainConfig, init_pg: bool = True):
    use_cuda = torch.cuda.is_available() and cfg.dist_backend == "nccl"
    if use_cuda:
        torch.cuda.set_device(rank % torch.cuda.device_count())
    if init_pg:
        dist.init_process_group(cfg.dist_backend, init_method=cfg.dist_url, world_size=world_size, rank=rank)
    device = torch.device("cuda", torch.cuda.current_device()) if use_cuda else torch.device("cpu")
    torch.manual_seed(cfg.seed)
    sampler, train_loader, val_loader = prepare_dataloader(cfg, distributed=True, rank=rank, world_size=world_size)
    net = build_model(cfg.model).to(device)
    if cfg.sync_bn:
        net = convert_sync_batchnorm(net)
    sgd = dict(lr=cfg.lr, momentum=cfg.momentum, weight_decay=cfg.weight_decay)
    peer = use_cuda and cfg.comm_backend == "peer"
    net = DistributedDataParallel(net, device_ids=[device.index] if use_cuda else None, bucket_cap_mb=cfg.bucket_cap_mb,
                                  comm_backend="peer" if peer else "torch", fused_sgd=sgd if peer else None)
    optimizer = net.optimizer if peer else (FlatFusedSGD(net.parameters(), **sgd) if use_cuda else torch.optim.SGD(net.parameters(), **sgd))
    sched = CosineWithWarmup(optimizer, cfg.lr, cfg.epochs, cfg.warmup_period)
    best_acc, start_epoch = -1.0, 0  # the first validated epoch always writes a checkpoint
    if cfg.resume:
        best_acc, last_epoch = load_checkpoint(net, cfg.checkpoint, optimizer, sched)
        start_epoch = last_epoch + 1  # the stored epoch is complete (see train/dp.py)
    step = TrainStep(net, optimizer, use_graph=cfg.cuda_graph and use_cuda)
    log = EpochLogger(os.path.join(cfg.log_dir, "ddp_%d.txt" % cfg.batch_size) if rank == 0 else None,
                      os.path.join(cfg.log_dir, "ddp_rank%d.jsonl" % rank), rank)
    for epoch in range(start_epoch, cfg.epochs):
        if sampler is not None:
            sampler.set_epoch(epoch)
        tl, ta, bt, dt = train_epoch(step, train_loader, device, cfg, rank)
        vl, va = validate(net, val_loader, device, cfg)
        sched.step()
        if rank == 0 and va > best_acc:
            save_best(net, va, epoch, cfg.checkpoint, optimizer, sched)
        best_acc = max(best_acc, va)
        imgs_per_s = cfg.batch_size * world_size / bt if bt else None  # bt: seconds per (per-rank) batch on this rank
        line = log.log_epoch(epoch, tl, ta, vl, va, bt, dt, lr=get_lr(optimizer), images_per_s=imgs_per_s,
                             ddp=net._get_ddp_logging_data())
        if rank == 0:
            print(line)
        net.check_health()
    if init_pg:
        dist.barrier()
        dist.destroy_process_group()
    return best_acc


def main(cfg: TrainConfig):
    if "RANK" in os.environ:  # launched by torchrun
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        cfg.dist_url = "env://"
        return worker(int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), cfg)
    import torch.multiprocessing as mp
    mp.spawn(worker, args=(cfg.world_size, cfg), nprocs=cfg.world_size)
