"""Data pipeline: DistributedSampler, synthetic CIFAR-shaped data, CIFAR-10 file reader.

Reference: train_ddp.py:81-150 (CIFAR-10 via torchvision, RandomCrop(32, pad 4) + HFlip +
Normalize for train, Normalize for test, DistributedSampler(shuffle=True) for train only, the
validation loader is NOT sharded, batch_size is per GPU, drop_last=False).
"""
from __future__ import annotations

import math
import os
import pickle
from typing import Iterator, Optional

import numpy as np
import torch
from torch.utils.data import DataLoader, Dataset, Sampler

CIFAR_MEAN = (0.4914, 0.4822, 0.4465)
CIFAR_STD = (0.2470, 0.2435, 0.2616)


class DistributedSampler(Sampler[int]):
    """Index-stream-identical re-implementation of torch.utils.data.DistributedSampler
    (torch/utils/data/distributed.py:66-157): ceil(len/world) samples per rank, a
    ``randperm(len, Generator(seed + epoch))`` shared by all ranks, head-wrap padding,
    ``indices[rank::world]``."""

    def __init__(self, dataset, num_replicas: Optional[int] = None, rank: Optional[int] = None,
                 shuffle: bool = True, seed: int = 0, drop_last: bool = False) -> None:
        if num_replicas is None or rank is None:
            import torch.distributed as dist
            if not (dist.is_available() and dist.is_initialized()):
                raise RuntimeError("Requires distributed package to be available")
            num_replicas = dist.get_world_size() if num_replicas is None else num_replicas
            rank = dist.get_rank() if rank is None else rank
        if rank >= num_replicas or rank < 0:
            raise ValueError(f"Invalid rank {rank}, rank should be in [0, {num_replicas - 1}]")
        self.dataset = dataset
        self.num_replicas = num_replicas
        self.rank = rank
        self.epoch = 0
        self.drop_last = drop_last
        n = len(self.dataset)
        if self.drop_last and n % self.num_replicas != 0:
            self.num_samples = math.ceil((n - self.num_replicas) / self.num_replicas)
        else:
            self.num_samples = math.ceil(n / self.num_replicas)
        self.total_size = self.num_samples * self.num_replicas
        self.shuffle = shuffle
        self.seed = seed

    def __iter__(self) -> Iterator[int]:
        n = len(self.dataset)
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            indices = torch.randperm(n, generator=g).tolist()
        else:
            indices = list(range(n))
        if not self.drop_last:
            padding = self.total_size - len(indices)
            if padding <= len(indices):
                indices += indices[:padding]
            else:
                indices += (indices * math.ceil(padding / len(indices)))[:padding]
        else:
            indices = indices[: self.total_size]
        assert len(indices) == self.total_size
        indices = indices[self.rank: self.total_size: self.num_replicas]
        assert len(indices) == self.num_samples
        return iter(indices)

    def __len__(self) -> int:
        return self.num_samples

    def set_epoch(self, epoch: int) -> None:
        self.epoch = epoch


class SyntheticImages(Dataset):
    """Deterministic random images with the post-transform CIFAR shape/dtype: float32
    ``[3, S, S]`` (already normalised) and an int64 label."""

    def __init__(self, size: int = 50000, image_size: int = 32, num_classes: int = 10,
                 seed: int = 1234) -> None:
        g = torch.Generator().manual_seed(seed)
        # a modest pool reused cyclically keeps host memory small for ImageNet-sized images
        pool = min(size, 4096 if image_size <= 64 else 512)
        self.images = torch.randn(pool, 3, image_size, image_size, generator=g)
        self.labels = torch.randint(0, num_classes, (size,), generator=g)
        self.size = size
        self.pool = pool

    def __len__(self) -> int:
        return self.size

    def __getitem__(self, i: int):
        return self.images[i % self.pool], self.labels[i]


class CIFAR10Files(Dataset):
    """CIFAR-10 from the standard ``cifar-10-batches-py`` pickles (no torchvision, no download)."""

    def __init__(self, root: str, train: bool, augment: bool, seed: int = 0, raw: bool = False) -> None:
        # raw=True: items are the stored uint8 CHW images; crop / flip / normalise then run on
        # the GPU once per batch (GpuAugment, --gpu-augment) instead of here once per sample
        self.raw = raw
        base = os.path.join(root, "cifar-10-batches-py")
        names = [f"data_batch_{i}" for i in range(1, 6)] if train else ["test_batch"]
        data, labels = [], []
        for n in names:
            with open(os.path.join(base, n), "rb") as f:
                d = pickle.load(f, encoding="latin1")
            data.append(np.asarray(d["data"], dtype=np.uint8).reshape(-1, 3, 32, 32))
            labels.extend(d["labels"] if "labels" in d else d["fine_labels"])
        self.data = torch.from_numpy(np.concatenate(data))
        self.labels = torch.tensor(labels, dtype=torch.int64)
        self.augment = augment
        self.mean = torch.tensor(CIFAR_MEAN).view(3, 1, 1)
        self.std = torch.tensor(CIFAR_STD).view(3, 1, 1)

    @staticmethod
    def available(root: str) -> bool:
        return os.path.isfile(os.path.join(root, "cifar-10-batches-py", "data_batch_1"))

    def __len__(self) -> int:
        return self.data.shape[0]

    def __getitem__(self, i: int):
        img = self.data[i]
        if self.raw:
            return img, self.labels[i]
        if self.augment:
            # RandomCrop(32, padding=4) + RandomHorizontalFlip (train_ddp.py:91-96)
            padded = torch.zeros(3, 40, 40, dtype=torch.uint8)
            padded[:, 4:36, 4:36] = img
            dy, dx = torch.randint(0, 9, (2,)).tolist()
            img = padded[:, dy:dy + 32, dx:dx + 32]
            if torch.rand(()) < 0.5:
                img = img.flip(2)
        x = img.float().div_(255.0)
        x = (x - self.mean) / self.std
        return x, self.labels[i]


def build_datasets(args):
    if not getattr(args, "synthetic", False) and not CIFAR10Files.available(args.data_dir):
        # The reference downloads CIFAR-10 on rank 0 (train_ddp.py:103-112).  Without network a
        # silent switch to random data would fill metrics_rank0.csv with numbers that look real,
        # so a missing / mistyped --data-dir is an error; synthetic data has to be asked for.
        raise FileNotFoundError(
            f"no CIFAR-10 files under {os.path.join(args.data_dir, 'cifar-10-batches-py')!r} "
            "(this build does not download); point --data-dir at the extracted "
            "cifar-10-batches-py directory's parent or pass --synthetic")
    if getattr(args, "synthetic", False):
        size = getattr(args, "synthetic_size", 50000)
        img = getattr(args, "image_size", 32)
        ncls = getattr(args, "num_classes", 10)
        return (SyntheticImages(size, img, ncls, seed=1234),
                SyntheticImages(max(size // 5, 1), img, ncls, seed=4321))
    raw = bool(getattr(args, "gpu_augment", False)) and torch.cuda.is_available()
    return (CIFAR10Files(args.data_dir, train=True, augment=True, raw=raw),
            CIFAR10Files(args.data_dir, train=False, augment=False, raw=raw))


def get_dataloaders(args, rank: int, world_size: int):
    """(train_loader, val_loader, train_sampler) — reference train_ddp.py:81-150."""
    train_ds, test_ds = build_datasets(args)
    distributed = world_size > 1
    sampler = DistributedSampler(train_ds, num_replicas=world_size, rank=rank,
                                 shuffle=True) if distributed else None
    pin = torch.cuda.is_available()
    # the reference's loaders respawn their 4 worker processes at every epoch, twice (train and
    # validation; persistent_workers defaults to False, torch dataloader.py:264-265): same batches
    # here, but the workers stay alive between epochs
    keep = args.workers > 0
    train_loader = DataLoader(train_ds, batch_size=args.batch_size, shuffle=(sampler is None),
                              sampler=sampler, num_workers=args.workers, pin_memory=pin,
                              drop_last=False, persistent_workers=keep)
    val_loader = DataLoader(test_ds, batch_size=args.batch_size, shuffle=False,
                            num_workers=args.workers, pin_memory=pin, drop_last=False,
                            persistent_workers=keep)
    if getattr(train_ds, "raw", False):
        # uint8 batches: the loops apply these on the device (trainer.py, _device_batch)
        train_loader.gpu_transform = GpuAugment(seed=getattr(args, "seed", 0) + rank, train=True)
        val_loader.gpu_transform = GpuAugment(train=False)
    return train_loader, val_loader, sampler


class GpuAugment:
    """RandomCrop(32, padding=4) + RandomHorizontalFlip + Normalize on the device for a uint8
    batch (one kernel per batch instead of PIL per sample)."""

    def __init__(self, pad: int = 4, mean=CIFAR_MEAN, std=CIFAR_STD, seed: int = 0,
                 train: bool = True) -> None:
        self.pad, self.mean, self.std, self.train = pad, mean, std, train
        self.gen = torch.Generator().manual_seed(seed)

    def __call__(self, images_u8: torch.Tensor) -> torch.Tensor:
        from .ops import functional as Fn
        n = images_u8.shape[0]
        if not self.train:
            # test transform (ToTensor + Normalize, train_ddp.py:98-101): centred crop, no flip
            p = torch.tensor([[self.pad, self.pad, 0]], dtype=torch.int32).repeat(n, 1)
            return Fn.augment_crop_flip_norm(images_u8, p.to(images_u8.device, non_blocking=True),
                                             self.mean, self.std, self.pad)
        p = torch.stack([torch.randint(0, 2 * self.pad + 1, (n,), generator=self.gen),
                         torch.randint(0, 2 * self.pad + 1, (n,), generator=self.gen),
                         torch.randint(0, 2, (n,), generator=self.gen)], dim=1).to(torch.int32)
        return Fn.augment_crop_flip_norm(images_u8, p.to(images_u8.device, non_blocking=True),
                                         self.mean, self.std, self.pad)
