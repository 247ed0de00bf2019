"""Learning-rate schedules of the reference, as pure functions of the global step.

CIFAR (resnet_cifar_main.py:287-307): 0.1, x0.1 at 40k, 60k and 80k steps.
ImageNet (resnet_imagenet_main.py:223-247, "intel resnet_50_8_nodes version"): linear warm-up from
0.1 to 0.4 over 6240 steps, 0.4 until 37440, then 0.04, 0.004 (until 99840), 0.0004.

The reference applies the value computed from the global step fetched by the PREVIOUS run call
(one-step lag) and starts from 0.1 (CIFAR) / 0.4 (ImageNet) - ``LaggedSchedule`` reproduces that;
``schedule(step)`` is the lag-free table.  All rates scale with ``base_lr / 0.1``.
"""
from __future__ import annotations

from typing import Callable


def cifar_lr(step: int, base_lr: float = 0.1) -> float:
    k = base_lr / 0.1
    if step < 40000:
        return 0.1 * k
    if step < 60000:
        return 0.01 * k
    if step < 80000:
        return 0.001 * k
    return 0.0001 * k


def imagenet_lr(step: int, base_lr: float = 0.1) -> float:
    k = base_lr / 0.1
    if step < 6240:
        return (0.1 + 0.3 * step / 6240.0) * k
    if step < 37440:
        return 0.4 * k
    if step < 74880:
        return 0.1 * 0.4 * k
    if step < 99840:
        return 0.01 * 0.4 * k
    return 0.001 * 0.4 * k


def schedule_for(dataset: str) -> Callable[[int, float], float]:
    return imagenet_lr if dataset == "imagenet" else cifar_lr


class LaggedSchedule:
    """Reference hook semantics: ``before_step()`` returns the rate to use for the coming step,
    ``after_step(global_step)`` derives the next one from the step counter that run returned."""

    def __init__(self, dataset: str, base_lr: float = 0.1, constant: bool = False):
        self._fn = schedule_for(dataset)
        self._base = base_lr
        self._constant = constant
        k = base_lr / 0.1
        self._lr = (0.4 if dataset == "imagenet" else 0.1) * k   # `begin()` of the hooks
        self._const_value = 0.1 * k  # SyncReplicas compat: the apply op only ever saw the constant

    def before_step(self) -> float:
        return self._const_value if self._constant else self._lr

    def after_step(self, global_step_seen: int) -> None:
        self._lr = self._fn(global_step_seen, self._base)
