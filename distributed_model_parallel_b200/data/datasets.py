"""Dataset factory and loaders.

``DatasetCollection(type, path, compose_train, compose_val).init()`` covers the reference's four dataset types
(/root/reference/code/distributed_training/dataset/dataset_collection.py:28-69: Imagenet folders, CUB-200-2011 via
its three index files, CIFAR10, Places365) plus ``synthetic``.  ``prepare_dataloader`` is the reference's dead
helper (utils.py:12-31) made live: pinned memory, an honoured batch size / worker count and an optional
``DistributedSampler`` (the reference never shards data across ranks)."""
from __future__ import annotations

import os
from typing import Optional, Tuple

import torch
from torch.utils.data import DataLoader, Dataset
from torch.utils.data.distributed import DistributedSampler

from .synthetic import SyntheticImages


class CUBDataset(Dataset):
    """CUB-200-2011 images listed by a (filepath, target) table; targets are 1-based in the index files."""

    base_folder = "CUB_200_2011/images"

    def __init__(self, root, table, transform=None, loader=None):
        self.root = os.path.expanduser(root)
        self.table = table.reset_index(drop=True)
        self.transform = transform
        if loader is None:
            from torchvision.datasets.folder import default_loader as loader
        self.loader = loader

    def __len__(self):
        return len(self.table)

    def __getitem__(self, idx):
        row = self.table.iloc[idx]
        img = self.loader(os.path.join(self.root, self.base_folder, row.filepath))
        if self.transform is not None:
            img = self.transform(img)
        return img, int(row.target) - 1


class DatasetCollection:
    TYPES = ("Imagenet", "CUB200", "CIFAR10", "Place365", "synthetic")

    def __init__(self, type, path, compose_train=None, compose_val=None, download: bool = False):
        if type not in self.TYPES:
            raise ValueError("unknown dataset type %r (have %s)" % (type, ", ".join(self.TYPES)))
        self.type, self.path, self.download = type, path, download
        self.compose = {"train": compose_train, "val": compose_val}

    def init(self) -> Tuple[Dataset, Dataset]:
        if self.type == "synthetic":
            return SyntheticImages(50000, seed=0), SyntheticImages(10000, seed=1)
        import torchvision.datasets as tvd
        if self.type == "Imagenet":
            return (tvd.ImageFolder(os.path.join(self.path, "train"), self.compose["train"]),
                    tvd.ImageFolder(os.path.join(self.path, "val"), self.compose["val"]))
        if self.type == "CIFAR10":
            return (tvd.CIFAR10(root=self.path, train=True, transform=self.compose["train"], download=self.download),
                    tvd.CIFAR10(root=self.path, train=False, transform=self.compose["val"], download=self.download))
        if self.type == "Place365":
            return (tvd.Places365(root=self.path, split="train-standard", small=True, download=False, transform=self.compose["train"]),
                    tvd.Places365(root=self.path, split="val", small=True, download=False, transform=self.compose["val"]))
        import pandas as pd  # CUB200
        base = os.path.join(self.path, "CUB_200_2011")
        images = pd.read_csv(os.path.join(base, "images.txt"), sep=" ", names=["img_id", "filepath"])
        labels = pd.read_csv(os.path.join(base, "image_class_labels.txt"), sep=" ", names=["img_id", "target"])
        split = pd.read_csv(os.path.join(base, "train_test_split.txt"), sep=" ", names=["img_id", "is_training_img"])
        data = images.merge(labels, on="img_id").merge(split, on="img_id")
        return (CUBDataset(self.path, data[data.is_training_img == 1], self.compose["train"]),
                CUBDataset(self.path, data[data.is_training_img == 0], self.compose["val"]))


def cifar_transforms():
    import torchvision.transforms as T
    norm = T.Normalize((0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010))
    return (T.Compose([T.RandomCrop(32, padding=4), T.RandomHorizontalFlip(), T.ToTensor(), norm]), T.Compose([T.ToTensor(), norm]))


def prepare_dataloader(cfg, distributed: bool = False, rank: int = 0, world_size: int = 1):
    """Returns (train_sampler, train_loader, val_loader) for ``cfg`` (a TrainConfig)."""
    tr, va = (None, None) if cfg.dataset_type == "synthetic" else cifar_transforms()
    train_ds, val_ds = DatasetCollection(cfg.dataset_type, cfg.data, tr, va).init()
    sampler: Optional[DistributedSampler] = None
    if distributed and world_size > 1:
        sampler = DistributedSampler(train_ds, num_replicas=world_size, rank=rank, shuffle=True, seed=cfg.seed)
    pin = torch.cuda.is_available()
    train_loader = DataLoader(train_ds, batch_size=cfg.batch_size, shuffle=sampler is None, num_workers=cfg.workers, pin_memory=pin,
                              sampler=sampler, drop_last=False, persistent_workers=cfg.workers > 0)
    val_loader = DataLoader(val_ds, batch_size=cfg.val_batch_size, shuffle=False, num_workers=cfg.workers, pin_memory=pin,
                            persistent_workers=cfg.workers > 0)
    return sampler, train_loader, val_loader


def sample_batch(cfg, batch: int) -> torch.Tensor:
    """One random input batch of the shape ``cfg``'s model trains on (cost probing for the pipeline partitioner)."""
    if cfg.model.startswith("bert"):
        return torch.randint(0, 30522, (batch, 128))
    hw = 224 if cfg.model == "resnet50" or cfg.dataset_type in ("Imagenet", "CUB200", "Place365") else 32
    return torch.randn(batch, 3, hw, hw)
