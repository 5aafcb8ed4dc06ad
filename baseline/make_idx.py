"""Write synthetic FashionMNIST-shape idx files (random uint8 pixels, labels 0..9) where
torchvision.datasets.FashionMNIST looks for them. Standalone on purpose: the reference arm imports
nothing from the framework under test."""
import os
import struct

import numpy as np


def write(root: str, n_train: int, n_test: int, seed: int = 0) -> str:
    raw = os.path.join(os.path.expanduser(root), "FashionMNIST", "raw")
    os.makedirs(raw, exist_ok=True)
    rng = np.random.default_rng(seed)
    for prefix, n in (("train", n_train), ("t10k", n_test)):
        with open(os.path.join(raw, f"{prefix}-images-idx3-ubyte"), "wb") as f:
            f.write(struct.pack(">IIII", 2051, n, 28, 28))
            left = n
            while left > 0:
                m = min(left, 65536)
                f.write(rng.integers(0, 256, size=(m, 784), dtype=np.uint8).tobytes())
                left -= m
        with open(os.path.join(raw, f"{prefix}-labels-idx1-ubyte"), "wb") as f:
            f.write(struct.pack(">II", 2049, n))
            f.write(rng.integers(0, 10, size=n, dtype=np.uint8).tobytes())
    return raw
