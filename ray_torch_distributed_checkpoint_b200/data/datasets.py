"""Synthetic datasets with the shapes of the reference's data (no network on the box).

``FashionMNISTSynthetic`` stands in for ``torchvision.datasets.FashionMNIST`` as used at
``/root/reference/my_ray_module.py:38-67``: 60 000 train / 10 000 test uint8 images of 1x28x28 and
int64 labels in [0, 10), normalised with ``ToTensor() + Normalize((0.5,), (0.5,))``. The pixels are
class-dependent (a fixed random template per class plus noise) so that training actually learns and
accuracy is a meaningful signal in tests.
"""
from __future__ import annotations

import os
import struct
from typing import Dict, Optional, Tuple

import numpy as np
import torch
from torch.utils.data import Dataset as TorchDataset

_LABELS = {0: "T-Shirt", 1: "Trouser", 2: "Pullover", 3: "Dress", 4: "Coat", 5: "Sandal", 6: "Shirt", 7: "Sneaker",
           8: "Bag", 9: "Ankle Boot"}


def get_labels_map() -> Dict[int, str]:
    """Class id -> label (ref: ``get_labels_map``, my_ray_module.py:79-91)."""
    return dict(_LABELS)


def _make_images(n: int, seed: int, noise: float = 48.0) -> Tuple[np.ndarray, np.ndarray]:
    rng = np.random.default_rng(seed)
    templates = np.random.default_rng(1234).integers(0, 256, size=(10, 28, 28)).astype(np.float32)
    # smooth the templates a little so classes differ in structure, not only in pixel noise
    templates = (templates + np.roll(templates, 1, 1) + np.roll(templates, 1, 2) + np.roll(templates, -1, 1)) / 4.0
    labels = rng.integers(0, 10, size=n, dtype=np.int64)
    imgs = templates[labels] + rng.normal(0.0, noise, size=(n, 28, 28)).astype(np.float32)
    return np.clip(imgs, 0, 255).astype(np.uint8), labels


class FashionMNISTSynthetic(TorchDataset):
    """Map-style dataset. ``raw=False`` yields ``(float32[1,28,28] normalised, int label)`` like the
    torchvision dataset with the reference's transform; ``raw=True`` yields ``(uint8[28,28], label)``."""

    def __init__(self, train: bool = True, size: Optional[int] = None, seed: int = 0, raw: bool = False):
        n = size if size is not None else (60000 if train else 10000)
        self.data_np, self.targets_np = _make_images(n, seed + (0 if train else 1))
        self.data = torch.from_numpy(self.data_np)          # uint8 [N,28,28], like torchvision's .data
        self.targets = torch.from_numpy(self.targets_np)
        self.raw = raw
        self.train = train

    def __len__(self) -> int:
        return self.data.shape[0]

    def __getitem__(self, idx):
        img = self.data[idx]
        if self.raw:
            return img, int(self.targets[idx])
        x = (img.to(torch.float32) / 255.0 - 0.5) / 0.5
        return x.unsqueeze(0), int(self.targets[idx])

    def normalized(self) -> torch.Tensor:
        """All images as float32 [N,1,28,28], normalised."""
        return ((self.data.to(torch.float32) / 255.0 - 0.5) / 0.5).unsqueeze(1)


def write_fashion_mnist_idx(root: str, n_train: int = 60000, n_test: int = 10000, seed: int = 0) -> str:
    """Write synthetic data in the on-disk format ``torchvision.datasets.FashionMNIST`` reads
    (``<root>/FashionMNIST/raw/{train,t10k}-{images-idx3,labels-idx1}-ubyte``), so code that uses the
    torchvision dataset runs without a download."""
    raw = os.path.join(os.path.expanduser(root), "FashionMNIST", "raw")
    os.makedirs(raw, exist_ok=True)
    for prefix, n, s in (("train", n_train, seed), ("t10k", n_test, seed + 1)):
        imgs, labels = _make_images(n, s)
        with open(os.path.join(raw, f"{prefix}-images-idx3-ubyte"), "wb") as f:
            f.write(struct.pack(">IIII", 2051, n, 28, 28))
            f.write(imgs.tobytes())
        with open(os.path.join(raw, f"{prefix}-labels-idx1-ubyte"), "wb") as f:
            f.write(struct.pack(">II", 2049, n))
            f.write(labels.astype(np.uint8).tobytes())
    return raw


class ImageNetSynthetic(TorchDataset):
    """Random ``float32[3,224,224]`` images and labels in [0, 1000) (ResNet-50 configuration)."""

    def __init__(self, size: int = 1281167, seed: int = 0, classes: int = 1000, resolution: int = 224):
        self.size, self.seed, self.classes, self.res = size, seed, classes, resolution

    def __len__(self):
        return self.size

    def __getitem__(self, idx):
        g = torch.Generator().manual_seed(self.seed * 1000003 + idx)
        return torch.randn(3, self.res, self.res, generator=g), int(torch.randint(0, self.classes, (1,), generator=g))


class TokenSynthetic(TorchDataset):
    """Random token sequences ``int64[seq_len]`` with next-token targets (GPT-2 configuration)."""

    def __init__(self, size: int = 100000, seq_len: int = 1024, vocab: int = 50257, seed: int = 0):
        self.size, self.seq_len, self.vocab, self.seed = size, seq_len, vocab, seed

    def __len__(self):
        return self.size

    def __getitem__(self, idx):
        g = torch.Generator().manual_seed(self.seed * 1000003 + idx)
        toks = torch.randint(0, self.vocab, (self.seq_len + 1,), generator=g)
        return toks[:-1], toks[1:]
