"""CIFAR-10 / CIFAR-100 binary-record readers and batch iterators.

Record layouts (reference resnet_cifar_main.py:157-182, cifar_input.py:36-45):
  cifar10 : 1 label byte + 3072 image bytes (CHW, 32x32x3)
  cifar100: 1 coarse-label byte + 1 fine-label byte + 3072 image bytes (label_offset = 1)
File names (resnet_cifar_main.py:140-154): ``<dir>/cifar-10-batches-bin/data_batch_{1..5}.bin`` and
``test_batch.bin``; the evaluator takes a glob (``--eval_data_path=.../test_batch*``,
resnet_cifar_eval.py:88-89).

Images stay uint8 CHW on the host; padding / cropping / flipping / standardisation run on the
device (csrc/aug.cu) - or in ``data/preprocess.py`` for the CPU oracle / torch backend.  As in
the reference every replica reads ALL files and shuffles independently (no sharding).
"""
from __future__ import annotations

import glob
import os
from typing import Iterator, List, Optional, Tuple

import numpy as np

IMAGE_BYTES = 3 * 32 * 32
NUM_TRAIN, NUM_TEST = 50000, 10000


def record_layout(dataset: str) -> Tuple[int, int, int]:
    """(label_bytes, label_offset, record_bytes)"""
    if dataset == "cifar10":
        return 1, 0, 1 + IMAGE_BYTES
    if dataset == "cifar100":
        return 2, 1, 2 + IMAGE_BYTES
    raise ValueError(f"Not supported dataset {dataset}")


def train_filenames(data_dir: str, dataset: str = "cifar10") -> List[str]:
    if dataset == "cifar10":
        d = os.path.join(data_dir, "cifar-10-batches-bin")
        return [os.path.join(d, f"data_batch_{i}.bin") for i in range(1, 6)]
    return [os.path.join(data_dir, "cifar-100-binary", "train.bin")]


def eval_filenames(path_or_glob: str, dataset: str = "cifar10") -> List[str]:
    if os.path.isdir(path_or_glob):
        if dataset == "cifar10":
            return [os.path.join(path_or_glob, "cifar-10-batches-bin", "test_batch.bin")]
        return [os.path.join(path_or_glob, "cifar-100-binary", "test.bin")]
    return sorted(glob.glob(path_or_glob))


def load_records(files: List[str], dataset: str = "cifar10") -> Tuple[np.ndarray, np.ndarray]:
    """Returns (images uint8 [N, 3072] CHW, labels int32 [N])."""
    label_bytes, label_offset, rec = record_layout(dataset)
    imgs, labs = [], []
    for f in files:
        raw = np.fromfile(f, dtype=np.uint8)
        if raw.size % rec != 0:
            raise ValueError(f"{f}: size {raw.size} is not a multiple of the record size {rec}")
        raw = raw.reshape(-1, rec)
        labs.append(raw[:, label_offset].astype(np.int32))
        imgs.append(raw[:, label_bytes:])
    if not imgs:
        raise FileNotFoundError("no CIFAR record files given")
    return np.ascontiguousarray(np.concatenate(imgs)), np.concatenate(labs)


def write_records(path: str, images_chw_u8: np.ndarray, labels: np.ndarray, dataset: str = "cifar10"):
    """Inverse of load_records (used by tests and to materialise synthetic datasets)."""
    label_bytes, label_offset, rec = record_layout(dataset)
    n = images_chw_u8.shape[0]
    out = np.zeros((n, rec), np.uint8)
    out[:, label_offset] = labels.astype(np.uint8)
    out[:, label_bytes:] = images_chw_u8.reshape(n, IMAGE_BYTES)
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    out.tofile(path)


class CifarBatches:
    """Infinite (or ``num_epochs``-bounded) iterator of (images uint8 [B,3072], labels int32 [B]).

    Training: full-epoch shuffle then repeat - ``dataset.shuffle(50000) ... .repeat(num_epochs)
    .batch(B)`` (resnet_cifar_main.py:215-241); batches may straddle epoch boundaries exactly as
    tf.data's ``repeat().batch()`` does.  Eval: sequential."""

    def __init__(self, images: np.ndarray, labels: np.ndarray, batch_size: int, training: bool,
                 num_epochs: Optional[int] = None, seed: int = 0):
        self.images, self.labels = images, labels
        self.B, self.training = batch_size, training
        self.num_epochs = num_epochs
        self.rng = np.random.default_rng(seed)

    def __iter__(self) -> Iterator[Tuple[np.ndarray, np.ndarray]]:
        n = self.images.shape[0]
        epoch = 0
        carry_i = np.zeros((0, IMAGE_BYTES), np.uint8)
        carry_l = np.zeros((0,), np.int32)
        while self.num_epochs is None or epoch < self.num_epochs:
            order = self.rng.permutation(n) if self.training else np.arange(n)
            imgs = np.concatenate([carry_i, self.images[order]]) if carry_i.size else self.images[order]
            labs = np.concatenate([carry_l, self.labels[order]]) if carry_l.size else self.labels[order]
            full = (imgs.shape[0] // self.B) * self.B
            for s in range(0, full, self.B):
                yield imgs[s:s + self.B], labs[s:s + self.B]
            carry_i, carry_l = imgs[full:], labs[full:]
            epoch += 1
        if carry_i.shape[0]:
            yield carry_i, carry_l   # final partial batch, like tf.data's batch()
