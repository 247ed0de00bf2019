"""Variable initialisation with the reference's initialisers, keyed by TF variable name.

Everything is produced on the host from one seed, in TF layout (conv kernels HWIO, dense kernel
[in, out]) so that (a) every replica starts from identical weights - the reference gets this from
the chief initialising shared ps variables / ``hvd.BroadcastGlobalVariablesHook(0)``
(resnet_cifar_main_horovod.py:316) - and (b) both executors and the checkpoint code share a single
source of truth.

Initialisers: conv = variance_scaling(fan_in, truncated normal,
stddev sqrt(1/fan_in)), no bias (resnet_model_official.py:87-91); dense = glorot uniform + zero bias
(tf.layers.dense default); BN gamma=1, beta=0, moving_mean=0, moving_variance=1.
"""
from __future__ import annotations

import zlib
from typing import Dict

import numpy as np

from .spec import NetSpec


def _truncated_normal(rng: np.random.Generator, shape, stddev: float) -> np.ndarray:
    """Samples N(0, stddev) re-drawing values beyond 2 stddev (tf.truncated_normal)."""
    out = rng.standard_normal(size=shape)
    bad = np.abs(out) > 2.0
    while bad.any():
        out[bad] = rng.standard_normal(size=int(bad.sum()))
        bad = np.abs(out) > 2.0
    return (out * stddev).astype(np.float32)


def init_variables(spec: NetSpec, seed: int = 0, with_slots: bool = True) -> Dict[str, np.ndarray]:
    """All global variables of the training graph: trainables, BN moving statistics,
    ``<var>/Momentum`` slots and ``global_step`` (405 entries for ImageNet-50, 403 for CIFAR-50)."""
    rng = np.random.default_rng(seed)
    out: Dict[str, np.ndarray] = {}
    for name, shape in spec.trainable_variables():
        if name.endswith("/kernel") and len(shape) == 4:
            fan_in = shape[0] * shape[1] * shape[2]
            out[name] = _truncated_normal(rng, shape, float(np.sqrt(1.0 / fan_in)))
        elif name == "dense/kernel":
            limit = float(np.sqrt(6.0 / (shape[0] + shape[1])))
            out[name] = rng.uniform(-limit, limit, size=shape).astype(np.float32)
        elif name.endswith("/gamma"):
            out[name] = np.ones(shape, np.float32)
        else:  # beta, dense/bias
            out[name] = np.zeros(shape, np.float32)
    for name, shape in spec.moving_variables():
        out[name] = (np.ones if name.endswith("moving_variance") else np.zeros)(shape, np.float32)
    if with_slots:
        for name, shape in spec.trainable_variables():
            out[f"{name}/Momentum"] = np.zeros(shape, np.float32)
    out["global_step"] = np.zeros((), np.int64)
    return out


def variables_checksum(variables: Dict[str, np.ndarray]) -> float:
    """Cheap order-independent checksum used to assert that replicas agree."""
    total = 0.0
    for name in sorted(variables):
        v = variables[name]
        total += float(np.asarray(v, dtype=np.float64).sum()) * (1 + (zlib.crc32(name.encode()) % 7))
    return total
