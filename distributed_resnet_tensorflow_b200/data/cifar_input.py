"""Legacy CIFAR input with the call shape of the reference's ``cifar_input.build_input``
(cifar_input.py:21-115; used by its evaluators, resnet_cifar_eval.py:88-89).

``build_input(dataset, data_path, batch_size, mode)`` takes a file glob, supports cifar10 and
cifar100 (fine label at offset 1), and yields ``(images float32 [B,32,32,3], labels one-hot
float32 [B,classes])``.  Train mode: pad 2 pixels per side (36x36), random 32x32 crop, random flip,
per-image standardisation, shuffled (the reference's RandomShuffleQueue of 16*B records); eval
mode: standardisation only, file order (FIFO queue).  Unlike the tf.data pipeline of the trainers
this one pads by 2, not 4.
"""
from __future__ import annotations

import glob
from typing import Iterator, Tuple

import numpy as np

from . import cifar, preprocess

LEGACY_PAD = 2


def build_input(dataset: str, data_path: str, batch_size: int, mode: str, seed: int = 0,
                num_epochs=None) -> Iterator[Tuple[np.ndarray, np.ndarray]]:
    if dataset not in ("cifar10", "cifar100"):
        raise ValueError("Not supported dataset %s" % dataset)
    num_classes = 10 if dataset == "cifar10" else 100
    files = sorted(glob.glob(data_path))
    if not files:
        raise FileNotFoundError(f"no files match {data_path!r}")
    imgs, labs = cifar.load_records(files, dataset)
    training = mode == "train"
    step = 0
    for bi, bl in cifar.CifarBatches(imgs, labs, batch_size, training, num_epochs, seed):
        x = preprocess.cifar_preprocess_batch(bi, training, LEGACY_PAD, seed, step, step * batch_size)
        onehot = np.zeros((len(bl), num_classes), np.float32)
        onehot[np.arange(len(bl)), bl] = 1.0
        step += 1
        yield x, onehot
