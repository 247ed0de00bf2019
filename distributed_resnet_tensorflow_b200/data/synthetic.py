"""Synthetic datasets of the reference's shapes (there is no dataset on the build machines):
random uint8 images + random labels, optionally with a learnable signal so that convergence tests
have something to fit."""
from __future__ import annotations

from typing import Iterator, Tuple

import numpy as np


def synthetic_cifar(n: int, num_classes: int = 10, seed: int = 0, learnable: bool = True):
    """(images uint8 [n, 3072] CHW, labels int32 [n]). With ``learnable`` the label is encoded as
    a class-dependent brightness pattern, so a network can reach high precision quickly."""
    rng = np.random.default_rng(seed)
    labels = rng.integers(0, num_classes, size=n).astype(np.int32)
    imgs = rng.integers(0, 256, size=(n, 3, 32, 32)).astype(np.float32)
    if learnable:
        proto = np.random.default_rng(1234).uniform(-60, 60, size=(num_classes, 3, 4, 4)).astype(np.float32)
        imgs = 0.35 * imgs + 80 + np.kron(proto[labels], np.ones((1, 1, 8, 8), np.float32))
    return np.clip(imgs, 0, 255).astype(np.uint8).reshape(n, 3072), labels


def synthetic_imagenet_batches(batch: int, num_classes: int = 1001, hw: int = 224, seed: int = 0,
                               pool: int = 4) -> Iterator[Tuple[np.ndarray, np.ndarray]]:
    """Endless stream cycling over ``pool`` pre-generated uint8 HWC batches."""
    rng = np.random.default_rng(seed)
    batches = [(rng.integers(0, 256, size=(batch, hw * hw * 3), dtype=np.uint8),
                rng.integers(1, num_classes, size=(batch,)).astype(np.int32)) for _ in range(pool)]
    i = 0
    while True:
        yield batches[i % pool]
        i += 1
