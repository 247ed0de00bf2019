"""Entry points with the reference's names and flags."""
from __future__ import annotations

import logging
import sys
from typing import Dict, Optional, Sequence

from .. import config as _config
from ..utils import flags as _flags


def _setup_logging() -> None:
    logging.basicConfig(level=logging.INFO, format="%(asctime)s %(levelname)s %(name)s: %(message)s",
                        stream=sys.stdout, force=True)


def _parse(argv: Optional[Sequence[str]], defaults: Dict, entry: str, **forced):
    fs = _flags.make_flagset(**defaults)
    values, _ = fs.parse(argv)
    if forced:
        values = values.replace(**{k: v for k, v in forced.items() if not values.is_set(k)})
    return _config.from_flags(values, entry=entry)


def cifar_main(argv: Optional[Sequence[str]] = None):
    """resnet_cifar_main.py: serial or parameter-server (SyncReplicas) CIFAR training."""
    _setup_logging()
    cfg = _parse(argv, _flags.CIFAR_MAIN, "train")
    if cfg.mode != "train":
        return {}
    from ..train.loop import train
    return train(cfg)


def cifar_main_horovod(argv: Optional[Sequence[str]] = None):
    """resnet_cifar_main_horovod.py: allreduce CIFAR training under a one-process-per-GPU launcher
    (torchrun / mpirun / srun: rank and world size come from the launcher's environment, the
    counterpart of hvd.init(), resnet_cifar_main_horovod.py:342).  The only difference from cifar_main
    is that this entry point IS the allreduce path: --use_horovod defaults to True here (the reference
    needs the flag passed explicitly, scripts/submit-horovod-train-mac.sh:5); everything else - flags,
    hooks, checkpoints - is the shared training loop."""
    _setup_logging()
    cfg = _parse(argv, _flags.CIFAR_MAIN, "train", use_horovod=True)
    if cfg.mode != "train":
        return {}
    from ..train.loop import train
    return train(cfg)


def imagenet_main(argv: Optional[Sequence[str]] = None):
    """resnet_imagenet_main.py: ImageNet training (PS mode as in the reference; --use_horovod=True
    selects the allreduce path; without --job_name it trains serially)."""
    _setup_logging()
    cfg = _parse(argv, _flags.IMAGENET_MAIN, "train")
    if cfg.mode != "train":
        return {}
    from ..train.loop import train
    return train(cfg)


def cifar_eval(argv: Optional[Sequence[str]] = None):
    """resnet_cifar_eval.py"""
    _setup_logging()
    cfg = _parse(argv, _flags.CIFAR_EVAL, "eval")
    if cfg.mode == "train":
        raise SystemExit("resnet_cifar_eval only evaluates (the reference calls an undefined train() "
                         "here, resnet_cifar_eval.py:169); use resnet_cifar_main for training")
    from ..train.evaluate import evaluate
    return evaluate(cfg)


def imagenet_eval(argv: Optional[Sequence[str]] = None):
    """resnet_imagenet_eval.py"""
    _setup_logging()
    cfg = _parse(argv, _flags.IMAGENET_EVAL, "eval")
    from ..train.evaluate import evaluate
    return evaluate(cfg)
