"""Separate evaluation process (reference resnet_cifar_eval.py:85-141, resnet_imagenet_eval.py:154-210).

Polls ``--log_root`` for the ``checkpoint`` state file, restores the latest checkpoint by variable
name (only the non-Momentum entries - the eval graph has no optimizer), runs ``eval_batch_count``
batches in inference mode (BatchNorm uses the moving statistics), and writes ``Precision`` /
``Best Precision`` summaries at the checkpoint's ``global_step``; then sleeps 60 s, or stops with
``--eval_once``.  Differences to the reference, on purpose: it sleeps instead of busy-spinning while
no checkpoint exists yet and it skips re-evaluating an unchanged checkpoint.
"""
from __future__ import annotations

import logging
import os
import time
from typing import Dict, Iterator, Optional, Tuple

import numpy as np

from ..ckpt.saver import Saver, get_checkpoint_state
from ..config import Config
from ..models.init import init_variables
from ..models.spec import build_spec
from ..utils.events import EventFileWriter
from .backends import DistContext, NativeBackend, TorchBackend
from .loop import choose_backend, local_device

log = logging.getLogger("drn")


def eval_batches(cfg: Config, spec, batch: int, device: Optional[str] = None) -> Iterator[Tuple]:
    if spec.dataset == "imagenet":
        if cfg.paths.eval_data_path and not cfg.train.synthetic_data:
            from ..data.imagenet import ImageNetBatches, ImageNetBatchesGPU
            from .loop import use_gpu_decode
            gpu = use_gpu_decode(cfg, device)
            while True:
                if gpu:
                    yield from ImageNetBatchesGPU(cfg.paths.eval_data_path, batch, False, 1, device=device)
                else:
                    yield from ImageNetBatches(cfg.paths.eval_data_path, batch, False, 1)
        from ..data.synthetic import synthetic_imagenet_batches
        for imgs, labs in synthetic_imagenet_batches(batch, spec.num_classes, spec.image_size, 99):
            yield imgs, labs, np.zeros((batch, 3), np.int32)
    from ..data import cifar
    if cfg.paths.eval_data_path and not cfg.train.synthetic_data:
        files = cifar.eval_filenames(cfg.paths.eval_data_path, cfg.model.dataset)
        imgs, labs = cifar.load_records(files, cfg.model.dataset)
    else:
        from ..data.synthetic import synthetic_cifar
        imgs, labs = synthetic_cifar(2048, spec.num_classes, seed=cfg.train.seed + 7)
    while True:
        for i, l in cifar.CifarBatches(imgs, labs, batch, False, 1):
            if len(l) == batch:
                yield i, l, None


def evaluate(cfg: Config, max_rounds: Optional[int] = None) -> Dict[str, float]:
    spec = build_spec(cfg.model.dataset, cfg.model.resnet_size, cfg.model.resnet_version)
    backend_name = choose_backend(cfg)
    ctx = DistContext()
    device = local_device(cfg, ctx, backend_name)
    B = cfg.eval.eval_batch_size
    if backend_name == "native":
        be = NativeBackend(spec, B, ctx, device=device, weight_decay=cfg.train.weight_decay,
                           use_graph=False)
    else:
        be = TorchBackend(spec, B, ctx, device=device, dtype="fp32" if device == "cpu" else cfg.train.dtype,
                          weight_decay=cfg.train.weight_decay)
    names = [n for n, _ in spec.trainable_variables()] + [n for n, _ in spec.moving_variables()] + ["global_step"]
    writer = EventFileWriter(cfg.paths.eval_dir) if cfg.paths.eval_dir else None
    batches = eval_batches(cfg, spec, B, device if backend_name == "native" else None)
    best_precision, last_path, rounds = 0.0, None, 0
    result: Dict[str, float] = {}
    while True:
        st = get_checkpoint_state(cfg.paths.log_root)
        path = None
        if st is not None:
            p = st.model_checkpoint_path
            path = p if os.path.isabs(p) else os.path.join(cfg.paths.log_root, p)
        if path is None or not os.path.exists(path + ".index"):
            log.info("No model to eval yet at %s", cfg.paths.log_root)
        elif path != last_path:
            try:
                restored = Saver.restore(path, names)
            except (OSError, ValueError, KeyError) as e:
                log.error("Cannot restore checkpoint: %s", e)
                restored = None
            if restored is not None:
                variables = init_variables(spec, 0, with_slots=False)
                variables.update(restored)
                be.import_variables(variables)
                log.info("Loading checkpoint %s", path)
                total, correct, loss_sum = 0, 0, 0.0
                for _ in range(cfg.eval.eval_batch_count):
                    imgs, labs, crop = next(batches)
                    r = be.eval_batch(imgs, labs, crop)
                    correct += r["correct"]
                    loss_sum += r["loss_sum"]
                    total += len(labs)
                precision = correct / max(total, 1)
                best_precision = max(precision, best_precision)
                step = int(restored["global_step"])
                l2 = sum(float((np.asarray(v, np.float64) ** 2).sum()) for k, v in restored.items()
                         if k in dict(spec.trainable_variables()))
                cost = loss_sum / max(total, 1) + cfg.train.weight_decay * 0.5 * l2
                if writer is not None:
                    writer.add_scalars({"Precision": precision, "Best Precision": best_precision,
                                        "cost": cost}, step)
                    writer.flush()
                log.info("loss: %.3f, precision: %.3f, best precision: %.3f", cost, precision, best_precision)
                result = {"precision": precision, "best_precision": best_precision, "loss": cost,
                          "global_step": step}
                last_path = path
        rounds += 1
        if cfg.eval.eval_once or (max_rounds is not None and rounds >= max_rounds):
            break
        time.sleep(cfg.eval.eval_interval_secs)
    if writer is not None:
        writer.close()
    return result
