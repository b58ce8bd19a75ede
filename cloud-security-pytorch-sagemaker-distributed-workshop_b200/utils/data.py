"""Datasets, sharding and the batch loader.

Reference behaviour being reproduced (``notebooks/code/cifar10-distributed-native-cpu.py:42-84``,
``cifar10-distributed-smddp-gpu.py:55-99``): CIFAR-10 from a local directory, the augmenting
transform ``RandomCrop(32, padding=4) + RandomHorizontalFlip + ToTensor + Normalize`` applied to
both train and test sets, ``DistributedSampler`` (shuffle, seed 0, ``set_epoch`` never called) on
the train set, an unsharded shuffled test loader.

B200-first redesign: the dataset is held as one ``uint8 [N,32,32,3]`` array in pinned host memory;
a batch is an index gather into a pinned staging buffer, one async H2D copy of *uint8* pixels
(4x fewer bytes than the reference's fp32 tensors), and the crop/flip/normalise runs as a single
GPU kernel (``ops.gpu_augment``) that writes the NHWC bf16 activation the first
convolution consumes.  There is no network here, so a class-structured synthetic CIFAR-10
(``SyntheticCIFAR10``) stands in when no dataset files are present.
"""
from __future__ import annotations

import math
import os
from typing import Iterator, Optional, Tuple

import numpy as np
import torch

CIFAR10_MEAN = (0.4914, 0.4822, 0.4465)
CIFAR10_STD = (0.2023, 0.1994, 0.2010)


class SyntheticCIFAR10:
    """Deterministic, learnable stand-in for CIFAR-10 (uint8 HWC images, int64 labels).

    Each class owns a smooth low-frequency colour template; a sample is template + noise, so a
    CNN can reach high accuracy in a few epochs (used by the convergence tests) while shapes and
    dtypes match the real dataset (50,000 / 10,000 images of 32x32x3).
    """

    num_classes = 10

    def __init__(self, train: bool = True, size: Optional[int] = None, seed: int = 1234,
                 hw: int = 32, channels: int = 3, noise: float = 40.0):
        n = size if size is not None else (50000 if train else 10000)
        rng = np.random.RandomState(seed)           # templates shared by train and test
        yy, xx = np.meshgrid(np.linspace(0, 1, hw), np.linspace(0, 1, hw), indexing="ij")
        templates = np.zeros((self.num_classes, hw, hw, channels), dtype=np.float32)
        for c in range(self.num_classes):
            for ch in range(channels):
                fx, fy, ph = rng.uniform(0.5, 3.0), rng.uniform(0.5, 3.0), rng.uniform(0, 2 * math.pi)
                templates[c, :, :, ch] = 127.5 + 90.0 * np.sin(2 * math.pi * (fx * xx + fy * yy) + ph)
        srng = np.random.RandomState(seed + (1 if train else 2))
        self.targets = srng.randint(0, self.num_classes, size=n).astype(np.int64)
        data = templates[self.targets] + srng.normal(0.0, noise, size=(n, hw, hw, channels)).astype(np.float32)
        self.data = np.clip(data, 0, 255).astype(np.uint8)
        self.train = train

    def __len__(self) -> int:
        return len(self.targets)

    def __getitem__(self, i) -> Tuple[torch.Tensor, int]:
        x = torch.from_numpy(self.data[i]).permute(2, 0, 1).float().div_(255.0)
        return x, int(self.targets[i])


class SyntheticImageNet:
    """ImageNet-shaped synthetic set for the ResNet-50 config of BASELINE.json (uint8 224x224x3)."""

    num_classes = 1000

    def __init__(self, size: int = 4096, hw: int = 224, seed: int = 7):
        rng = np.random.RandomState(seed)
        self.targets = rng.randint(0, self.num_classes, size=size).astype(np.int64)
        self.data = rng.randint(0, 256, size=(size, hw, hw, 3), dtype=np.uint8)

    def __len__(self):
        return len(self.targets)


def load_cifar10(root: str, train: bool, synthetic: bool = False, size: Optional[int] = None):
    """Real CIFAR-10 when the torchvision files exist under ``root`` (``download=False``, as in the
    reference), else the synthetic stand-in.  Returns an object with ``.data`` (uint8 NHWC) and
    ``.targets``."""
    if not synthetic and root and os.path.isdir(os.path.join(root, "cifar-10-batches-py")):
        try:
            import torchvision
            ds = torchvision.datasets.CIFAR10(root=root, train=train, download=False)
            out = SyntheticCIFAR10.__new__(SyntheticCIFAR10)
            out.data = np.ascontiguousarray(ds.data)
            out.targets = np.asarray(ds.targets, dtype=np.int64)
            out.train = train
            return out
        except Exception:   # corrupted / partial copy: fall through to synthetic
            pass
    return SyntheticCIFAR10(train=train, size=size)


class ShardSampler:
    """Index sharding with ``torch.utils.data.DistributedSampler`` semantics: permutation from
    ``seed + epoch``, padded to a multiple of ``num_replicas``, strided ``rank::num_replicas``.
    ``set_epoch`` exists; the reference never calls it (SURVEY K-D2) and the parity trainer
    keeps that behaviour unless asked otherwise."""

    def __init__(self, n: int, num_replicas: int = 1, rank: int = 0, shuffle: bool = True,
                 seed: int = 0, drop_last: bool = False):
        self.n, self.num_replicas, self.rank = n, num_replicas, rank
        self.shuffle, self.seed, self.epoch, self.drop_last = shuffle, seed, 0, drop_last
        if drop_last and n % num_replicas:
            self.num_samples = math.ceil((n - num_replicas) / num_replicas)
        else:
            self.num_samples = math.ceil(n / num_replicas)
        self.total_size = self.num_samples * num_replicas

    def set_epoch(self, epoch: int) -> None:
        self.epoch = epoch

    def indices(self) -> np.ndarray:
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            idx = torch.randperm(self.n, generator=g).numpy()
        else:
            idx = np.arange(self.n)
        if not self.drop_last:
            pad = self.total_size - len(idx)
            if pad > 0:
                reps = math.ceil(pad / len(idx))
                idx = np.concatenate([idx, np.tile(idx, reps)[:pad]])
        else:
            idx = idx[: self.total_size]
        return idx[self.rank: self.total_size: self.num_replicas]

    def __len__(self) -> int:
        return self.num_samples

    def __iter__(self):
        return iter(self.indices().tolist())


class BatchLoader:
    """Pinned-memory uint8 batch loader with double-buffered async H2D.

    Yields ``(images_u8 [B,H,W,C] on device, labels int64 [B] on device)``; augmentation and
    normalisation happen downstream on the GPU (``ops.gpu_augment``).  On CPU it yields CPU
    tensors.  ``len(loader.sampler)`` / ``len(loader.dataset)`` mirror the attributes the
    reference's log lines read (``...smddp-gpu.py:129-135``).
    """

    def __init__(self, dataset, batch_size: int, sampler: Optional[ShardSampler] = None,
                 shuffle: bool = False, device: Optional[torch.device] = None, seed: int = 0,
                 drop_last: bool = False):
        self.dataset = dataset
        self.batch_size = int(batch_size)
        self.sampler = sampler if sampler is not None else ShardSampler(len(dataset), 1, 0, shuffle=shuffle, seed=seed)
        self.device = torch.device(device) if device is not None else torch.device("cpu")
        self.drop_last = drop_last
        self._data = torch.from_numpy(dataset.data)
        self._targets = torch.from_numpy(np.asarray(dataset.targets, dtype=np.int64))
        self._cuda = self.device.type == "cuda"
        if self._cuda:
            shape = (self.batch_size,) + tuple(self._data.shape[1:])
            self._stage_x = [torch.empty(shape, dtype=torch.uint8).pin_memory() for _ in range(2)]
            self._stage_y = [torch.empty((self.batch_size,), dtype=torch.int64).pin_memory() for _ in range(2)]
            self._dev_x = [torch.empty(shape, dtype=torch.uint8, device=self.device) for _ in range(2)]
            self._dev_y = [torch.empty((self.batch_size,), dtype=torch.int64, device=self.device) for _ in range(2)]
            self._copy_stream = torch.cuda.Stream(device=self.device)
            self._events = [torch.cuda.Event() for _ in range(2)]
            self._consumed = [torch.cuda.Event() for _ in range(2)]

    def __len__(self) -> int:
        n = len(self.sampler)
        return n // self.batch_size if self.drop_last else math.ceil(n / self.batch_size)

    def _batches(self):
        idx = self.sampler.indices()
        nb = len(self)
        for b in range(nb):
            yield torch.from_numpy(idx[b * self.batch_size: (b + 1) * self.batch_size])

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        if not self._cuda:
            for sel in self._batches():
                yield self._data[sel], self._targets[sel]
            return
        cur = torch.cuda.current_stream(self.device)
        pending = None
        slot = 0
        for sel in self._batches():
            n = len(sel)
            self._consumed[slot].synchronize()          # staging + device slot free again
            torch.index_select(self._data, 0, sel, out=self._stage_x[slot][:n])
            torch.index_select(self._targets, 0, sel, out=self._stage_y[slot][:n])
            with torch.cuda.stream(self._copy_stream):
                self._dev_x[slot][:n].copy_(self._stage_x[slot][:n], non_blocking=True)
                self._dev_y[slot][:n].copy_(self._stage_y[slot][:n], non_blocking=True)
                self._events[slot].record(self._copy_stream)
            if pending is not None:
                ps, pn = pending
                cur.wait_event(self._events[ps])
                yield self._dev_x[ps][:pn], self._dev_y[ps][:pn]
                self._consumed[ps].record(cur)
            pending = (slot, n)
            slot ^= 1
        if pending is not None:
            ps, pn = pending
            cur.wait_event(self._events[ps])
            yield self._dev_x[ps][:pn], self._dev_y[ps][:pn]
            self._consumed[ps].record(cur)


def augment_reference(images_u8: torch.Tensor, generator: Optional[torch.Generator] = None,
                      train_transform: bool = True, mean=CIFAR10_MEAN, std=CIFAR10_STD,
                      offsets: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Plain-PyTorch oracle of the reference transform on a uint8 NHWC batch -> fp32 NCHW.

    ``offsets`` ([B,3] int: dy, dx in 0..8, flip in {0,1}) may be supplied so the GPU kernel and
    this oracle use identical random draws.
    """
    B, H, W, C = images_u8.shape
    x = images_u8.permute(0, 3, 1, 2).float() / 255.0
    if train_transform:
        if offsets is None:
            offsets = torch.stack([
                torch.randint(0, 9, (B,), generator=generator),
                torch.randint(0, 9, (B,), generator=generator),
                torch.randint(0, 2, (B,), generator=generator)], dim=1)
        xp = torch.nn.functional.pad(x, (4, 4, 4, 4))
        offsets = offsets.to(torch.int64)
        ar_h, ar_w = torch.arange(H), torch.arange(W)
        rows = offsets[:, 0:1] + ar_h[None, :]                                   # [B,H]
        wsrc = torch.where(offsets[:, 2:3] > 0, (W - 1) - ar_w[None, :], ar_w[None, :])
        cols = offsets[:, 1:2] + wsrc                                            # [B,W]
        bidx = torch.arange(B)[:, None, None]
        x = xp[bidx, :, rows[:, :, None], cols[:, None, :]].permute(0, 3, 1, 2).contiguous()
    m = torch.tensor(mean, dtype=torch.float32).view(1, C, 1, 1)
    s = torch.tensor(std, dtype=torch.float32).view(1, C, 1, 1)
    return (x - m) / s
