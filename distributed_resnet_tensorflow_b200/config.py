"""Typed configuration derived from parsed flags - replaces the reference's global ``FLAGS`` reads
inside model code (resnet_model.py:34,71-74,102) and its in-code constants:
weight decay 2e-4 / 1e-4 (resnet_cifar_main.py:99, resnet_imagenet_main.py:271), momentum 0.9,
base LR 0.1, log cadences 20 / 40 steps, checkpoint cadences 60 s / 600 s.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional

from .utils.flags import FlagValues


@dataclass(frozen=True)
class ClusterConfig:
    """Who am I and who else is there. #ps / #workers are implied by the host-list lengths, as in
    the reference (resnet_cifar_main.py:364-369)."""
    job_name: Optional[str]          # None = serial
    task_index: int
    ps_hosts: List[str]
    worker_hosts: List[str]
    use_horovod: bool
    sync_replicas: bool

    @property
    def num_ps(self) -> int:
        return len(self.ps_hosts)

    @property
    def num_workers(self) -> int:
        return len(self.worker_hosts)

    @property
    def is_chief(self) -> bool:
        return self.task_index == 0

    @property
    def distributed(self) -> bool:
        return self.job_name is not None or self.use_horovod


@dataclass(frozen=True)
class ModelConfig:
    dataset: str
    resnet_size: int
    resnet_version: int
    num_classes: int
    model: str                       # resnet | lrnet
    hidden_units: int
    image_size: int
    data_format: str


@dataclass(frozen=True)
class TrainConfig:
    batch_size: int
    train_steps: int
    num_epochs: int
    base_lr: float
    learning_rate: float             # LRNet only
    weight_decay: float
    momentum: float
    optimizer: str
    compat_ps_constant_lr: bool
    log_every: int
    summary_every: int
    save_checkpoint_secs: int
    save_checkpoint_steps: int
    keep_checkpoint_max: int
    seed: int
    backend: str
    dtype: str
    cuda_graph: bool
    allreduce_algo: str
    max_steps_per_run: int
    legacy_input: bool
    synthetic_data: bool
    profile: bool
    input_decode: str = "auto"
    ps_pull: str = "push"
    input_workers: int = 0
    deterministic: bool = False


@dataclass(frozen=True)
class PathConfig:
    train_data_path: str
    eval_data_path: str
    train_dir: str
    eval_dir: str
    log_root: str


@dataclass(frozen=True)
class EvalConfig:
    eval_batch_count: int
    eval_once: bool
    eval_interval_secs: int
    eval_batch_size: int


@dataclass(frozen=True)
class Config:
    cluster: ClusterConfig
    model: ModelConfig
    train: TrainConfig
    paths: PathConfig
    eval: EvalConfig
    mode: str
    num_gpus: int
    num_intra_threads: int
    num_inter_threads: int
    init_method: str


def _hosts(s: Optional[str]) -> List[str]:
    return [h for h in (s or "").split(",") if h]


def num_classes_for(dataset: str) -> int:
    """resnet_cifar_main.py:340-343, resnet_imagenet_main.py:266-267."""
    if dataset == "cifar10":
        return 10
    if dataset == "cifar100":
        return 100
    if dataset == "imagenet":
        return 1001
    raise ValueError(f"unknown dataset {dataset!r}")


def from_flags(f: FlagValues, entry: str = "train") -> Config:
    dataset = f.dataset
    imagenet = dataset == "imagenet"
    ps_hosts, worker_hosts = _hosts(f.ps_hosts), _hosts(f.worker_hosts)
    if f.num_ps:
        ps_hosts = [f"localhost:{2230 + i}" for i in range(f.num_ps)]
    if f.num_workers:
        worker_hosts = [f"localhost:{2220 + i}" for i in range(f.num_workers)]
    if f.job_name is not None and f.job_name not in ("ps", "worker"):
        raise ValueError(f"--job_name must be ps or worker, got {f.job_name!r}")
    if f.job_name is not None and f.task_index is None:
        raise ValueError("--task_index is required with --job_name")
    if f.job_name is not None and not f.sync_replicas and not f.use_horovod:
        # the reference crashes with AttributeError here (SURVEY.md 2.4 'asynchronous PS')
        raise ValueError("--job_name needs --sync_replicas=True: asynchronous parameter-server "
                         "training is not a working mode of the reference and is not provided")
    cluster = ClusterConfig(f.job_name, int(f.task_index or 0), ps_hosts, worker_hosts,
                            bool(f.use_horovod), bool(f.sync_replicas))
    wd = f.weight_decay if f.weight_decay >= 0 else (1e-4 if imagenet else 2e-4)
    if entry == "eval" and f.weight_decay < 0:
        wd = 2e-4  # both evaluators hard-code 2e-4 (resnet_imagenet_eval.py:232); only affects `cost`
    model = ModelConfig(dataset, f.resnet_size, f.resnet_version, num_classes_for(dataset), f.model,
                        f.hidden_units, 224 if imagenet else 32, f.data_format)
    train = TrainConfig(
        batch_size=f.batch_size, train_steps=f.train_steps, num_epochs=f.num_epochs,
        base_lr=f.base_lr, learning_rate=f.learning_rate, weight_decay=wd, momentum=f.momentum,
        optimizer=f.optimizer, compat_ps_constant_lr=f.compat_ps_constant_lr,
        log_every=f.log_every if f.log_every > 0 else (40 if imagenet else 20),
        summary_every=f.summary_every,
        save_checkpoint_secs=f.save_checkpoint_secs if f.save_checkpoint_secs >= 0 else (600 if imagenet else 60),
        save_checkpoint_steps=f.save_checkpoint_steps, keep_checkpoint_max=f.keep_checkpoint_max,
        seed=f.seed, backend=f.backend, dtype=f.dtype, cuda_graph=f.cuda_graph,
        allreduce_algo=f.allreduce_algo, max_steps_per_run=f.max_steps_per_run,
        legacy_input=f.legacy_input, synthetic_data=f.synthetic_data, profile=f.profile,
        input_decode=getattr(f, "input_decode", "auto"), ps_pull=getattr(f, "ps_pull", "push"),
        input_workers=getattr(f, "input_workers", 0), deterministic=bool(getattr(f, "deterministic", False)))
    paths = PathConfig(f.train_data_path, f.eval_data_path, f.train_dir, f.eval_dir, f.log_root)
    ev = EvalConfig(f.eval_batch_count, f.eval_once, f.eval_interval_secs,
                    f.batch_size if imagenet else 100)  # CIFAR evaluator: fixed 100 (resnet_cifar_eval.py:87)
    return Config(cluster, model, train, paths, ev, f.mode, f.num_gpus, f.num_intra_threads,
                  f.num_inter_threads, f.init_method)
