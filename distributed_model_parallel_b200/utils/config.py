"""One config object for the three trainers; every flag is honoured (the reference parses and ignores ``data``,
``--batch-size``, ``--workers`` and ``--dataset-type``, /root/reference/code/distributed_training/
model_parallel.py:17,29-36,92,97 -- SURVEY.md defect Q5)."""
from __future__ import annotations

import argparse
from dataclasses import asdict, dataclass, fields


@dataclass
class TrainConfig:
    data: str = "./data"                 # dataset root (positional in the reference MP script)
    dataset_type: str = "synthetic"      # synthetic | CIFAR10 | Imagenet | CUB200 | Place365
    model: str = "mobilenetv2"
    dist_url: str = "tcp://127.0.0.1:1224"
    dist_backend: str = "nccl"
    world_size: int = 1
    lr: float = 0.4
    epochs: int = 90
    batch_size: int = 512
    val_batch_size: int = 1000
    workers: int = 12
    weight_decay: float = 1e-4
    momentum: float = 0.9
    warmup_period: int = 10
    resume: bool = False
    checkpoint: str = "./checkpoint/ckpt.pth"
    log_dir: str = "./log"
    microbatches: int = 4
    schedule: str = "1f1b"               # naive | gpipe | 1f1b
    loss_on: str = "last"                # last | first (reference ring)
    bucket_cap_mb: float = 25.0
    comm_backend: str = "peer"           # peer | torch
    sync_bn: bool = False
    steps_per_epoch: int = 0             # 0 = full dataset
    print_freq: int = 30
    seed: int = 0
    cuda_graph: bool = True
    reference_split: bool = False        # reproduce the reference's 3/6/6/2 MobileNetV2 split (needs world_size 4)
    partition: str = ""                  # pipeline cost model: time | params | uniform ("" = time on GPUs, params on CPU)

    def to_dict(self):
        return asdict(self)


def build_parser(description: str = "distributed_model_parallel_b200 trainer") -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description=description)
    p.add_argument("data", nargs="?", default=TrainConfig.data, metavar="DIR", help="path to dataset")
    p.add_argument("--dist-url", default=TrainConfig.dist_url)
    p.add_argument("--world-size", type=int, default=TrainConfig.world_size, help="number of single-GPU processes")
    p.add_argument("--dist-backend", default=TrainConfig.dist_backend)
    p.add_argument("--lr", "--learning-rate", type=float, default=TrainConfig.lr, dest="lr")
    p.add_argument("--epochs", type=int, default=TrainConfig.epochs)
    p.add_argument("-type", "--dataset-type", default=TrainConfig.dataset_type, dest="dataset_type")
    p.add_argument("-b", "--batch-size", type=int, default=TrainConfig.batch_size, dest="batch_size")
    p.add_argument("--val-batch-size", type=int, default=TrainConfig.val_batch_size)
    p.add_argument("-j", "--workers", type=int, default=TrainConfig.workers)
    p.add_argument("--wd", "--weight-decay", type=float, default=TrainConfig.weight_decay, dest="weight_decay")
    p.add_argument("--momentum", type=float, default=TrainConfig.momentum)
    p.add_argument("--warmup-period", type=int, default=TrainConfig.warmup_period)
    p.add_argument("--resume", "-r", action="store_true")
    p.add_argument("--checkpoint", default=TrainConfig.checkpoint)
    p.add_argument("--log-dir", default=TrainConfig.log_dir)
    p.add_argument("--model", default=TrainConfig.model)
    p.add_argument("--microbatches", type=int, default=TrainConfig.microbatches)
    p.add_argument("--schedule", default=TrainConfig.schedule, choices=["naive", "gpipe", "1f1b"])
    p.add_argument("--loss-on", default=TrainConfig.loss_on, choices=["first", "last"])
    p.add_argument("--bucket-cap-mb", type=float, default=TrainConfig.bucket_cap_mb)
    p.add_argument("--comm-backend", default=TrainConfig.comm_backend, choices=["peer", "torch"])
    p.add_argument("--sync-bn", action="store_true")
    p.add_argument("--steps-per-epoch", type=int, default=0)
    p.add_argument("--print-freq", type=int, default=TrainConfig.print_freq)
    p.add_argument("--seed", type=int, default=0)
    p.add_argument("--no-cuda-graph", action="store_false", dest="cuda_graph")
    p.add_argument("--reference-split", action="store_true")
    p.add_argument("--partition", default=TrainConfig.partition, choices=["", "time", "params", "uniform"])
    return p


def config_from_args(ns: argparse.Namespace) -> TrainConfig:
    names = {f.name for f in fields(TrainConfig)}
    return TrainConfig(**{k: v for k, v in vars(ns).items() if k in names})
