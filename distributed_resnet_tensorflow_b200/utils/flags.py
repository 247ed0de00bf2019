"""Command-line flags with the surface of the reference's ``tf.app.flags`` definitions.

The reference re-declares a copy of its flags in each entrypoint and lets model code read a
global ``FLAGS`` (resnet_cifar_main.py:30-88, resnet_model.py:30-34), which is how two of its
entrypoints end up reading a flag they never define. Here there is ONE
table; every entrypoint parses the same table (with its own defaults where the reference's differ)
into an immutable ``FlagValues`` that is turned into typed configs (``config.py``) - model code
never sees a global.

Syntax accepted (absl-compatible): ``--flag=value``, ``--flag value``, ``--boolflag``,
``--noboolflag``, ``--boolflag=True|False|true|false|1|0|yes|no``.
"""
from __future__ import annotations

import sys
from dataclasses import dataclass, field
from typing import Any, Dict, Iterable, List, Optional, Sequence


class FlagError(ValueError):
    pass


@dataclass(frozen=True)
class FlagDef:
    name: str
    kind: str  # 'string' | 'integer' | 'float' | 'bool'
    default: Any
    help: str


_TRUE = {"true", "t", "1", "yes", "y"}
_FALSE = {"false", "f", "0", "no", "n"}


def _convert(fd: FlagDef, raw: str) -> Any:
    if fd.kind == "string":
        return raw
    if fd.kind == "integer":
        try:
            return int(raw)
        except ValueError as e:
            raise FlagError(f"--{fd.name}: expected an integer, got {raw!r}") from e
    if fd.kind == "float":
        try:
            return float(raw)
        except ValueError as e:
            raise FlagError(f"--{fd.name}: expected a float, got {raw!r}") from e
    if fd.kind == "bool":
        low = raw.strip().lower()
        if low in _TRUE:
            return True
        if low in _FALSE:
            return False
        raise FlagError(f"--{fd.name}: expected a boolean, got {raw!r}")
    raise AssertionError(fd.kind)


class FlagValues:
    """Immutable result of a parse; attribute access like the reference's FLAGS."""

    def __init__(self, values: Dict[str, Any], explicitly_set: Iterable[str]):
        object.__setattr__(self, "_values", dict(values))
        object.__setattr__(self, "_set", frozenset(explicitly_set))

    def __getattr__(self, name: str) -> Any:
        try:
            return self._values[name]
        except KeyError:
            raise AttributeError(f"unknown flag {name!r}") from None

    def __setattr__(self, name, value):
        raise AttributeError("FlagValues is read-only; use .replace()")

    def is_set(self, name: str) -> bool:
        return name in self._set

    def as_dict(self) -> Dict[str, Any]:
        return dict(self._values)

    def replace(self, **kw) -> "FlagValues":
        vals = dict(self._values)
        for k, v in kw.items():
            if k not in vals:
                raise FlagError(f"unknown flag {k!r}")
            vals[k] = v
        return FlagValues(vals, self._set | set(kw))


class FlagSet:
    def __init__(self, defs: Sequence[FlagDef] = ()):
        self._defs: Dict[str, FlagDef] = {}
        for d in defs:
            self.define(d)

    def define(self, d: FlagDef) -> None:
        if d.name in self._defs:
            raise FlagError(f"duplicate flag {d.name}")
        self._defs[d.name] = d

    def with_defaults(self, **overrides) -> "FlagSet":
        out = FlagSet()
        for d in self._defs.values():
            if d.name in overrides:
                d = FlagDef(d.name, d.kind, overrides[d.name], d.help)
            out.define(d)
        unknown = set(overrides) - set(self._defs)
        if unknown:
            raise FlagError(f"unknown flags in overrides: {sorted(unknown)}")
        return out

    def names(self) -> List[str]:
        return list(self._defs)

    def __contains__(self, name: str) -> bool:
        return name in self._defs

    def usage(self) -> str:
        lines = []
        for d in self._defs.values():
            lines.append(f"  --{d.name} ({d.kind}, default {d.default!r})\n      {d.help}")
        return "\n".join(lines)

    def parse(self, argv: Optional[Sequence[str]] = None, allow_unknown: bool = False):
        """Returns (FlagValues, positional_args)."""
        if argv is None:
            argv = sys.argv[1:]
        values = {d.name: d.default for d in self._defs.values()}
        explicitly: List[str] = []
        rest: List[str] = []
        i = 0
        argv = list(argv)
        while i < len(argv):
            tok = argv[i]
            i += 1
            if tok == "--":
                rest.extend(argv[i:])
                break
            if not tok.startswith("-") or tok == "-":
                rest.append(tok)
                continue
            body = tok.lstrip("-")
            if body in ("help", "h", "helpfull"):
                print(self.usage())
                raise SystemExit(0)
            if "=" in body:
                name, raw = body.split("=", 1)
                has_val = True
            else:
                name, raw, has_val = body, "", False
            fd = self._defs.get(name)
            if fd is None and name.startswith("no") and name[2:] in self._defs \
                    and self._defs[name[2:]].kind == "bool" and not has_val:
                values[name[2:]] = False
                explicitly.append(name[2:])
                continue
            if fd is None:
                if allow_unknown:
                    rest.append(tok)
                    continue
                raise FlagError(f"unknown flag --{name}")
            if not has_val:
                if fd.kind == "bool":
                    # absl: a bare --flag means True; "--flag False" is NOT consumed as a value
                    values[name] = True
                    explicitly.append(name)
                    continue
                if i >= len(argv):
                    raise FlagError(f"--{name} needs a value")
                raw = argv[i]
                i += 1
            values[name] = _convert(fd, raw)
            explicitly.append(name)
        return FlagValues(values, explicitly), rest


def _s(name, default, help_):
    return FlagDef(name, "string", default, help_)


def _n(name, default, help_):
    return FlagDef(name, "integer", default, help_)


def _x(name, default, help_):
    return FlagDef(name, "float", default, help_)


def _b(name, default, help_):
    return FlagDef(name, "bool", default, help_)


# --------------------------------------------------------------------------------------------
# The reference's flag surface, followed by the flags this framework adds.
# --------------------------------------------------------------------------------------------
REFERENCE_FLAGS: List[FlagDef] = [
    _s("dataset", "cifar10", "cifar10, cifar100 or imagenet."),
    _s("mode", "train", "train or eval."),
    _s("train_data_path", "", "CIFAR: parent directory of cifar-10-batches-bin/. ImageNet: "
       "directory of train-*-of-01024 TFRecords. Empty = synthetic data."),
    _s("eval_data_path", "", "CIFAR: glob of test_batch*.bin files. ImageNet: directory of "
       "validation-*-of-00128 TFRecords. Empty = synthetic data."),
    _n("image_size", 32, "Image side length (used by the debug MLP only)."),
    _s("train_dir", "", "Directory for training summaries."),
    _s("eval_dir", "", "Directory for evaluation summaries."),
    _n("eval_batch_count", 50, "Number of batches per evaluation round."),
    _b("eval_once", False, "Evaluate the latest checkpoint once and exit."),
    _s("log_root", "", "Checkpoint directory shared by trainer and evaluator."),
    _n("num_gpus", 0, "GPUs per host; worker i uses GPU (task_index % num_gpus). 0 = CPU."),
    _n("task_index", None, "Index of this task inside its job; worker 0 is the chief."),
    _n("replicas_to_aggregate", None, "Accepted for compatibility; always overwritten with the "
       "number of workers, as in the reference."),
    _n("train_steps", 2000, "Stop after this many global steps (honoured by every trainer here; "
       "the reference's CIFAR trainers ignore it)."),
    _n("num_epochs", 90, "Dataset repeat count (ImageNet)."),
    _n("batch_size", 32, "Per-worker batch size."),
    _x("learning_rate", 0.01, "Learning rate of the debug MLP; ResNets use the built-in schedules."),
    _b("sync_replicas", False, "Synchronous parameter-server mode (required with --job_name)."),
    _b("existing_servers", False, "Accepted for compatibility; servers are always in-process."),
    _s("ps_hosts", "localhost:2222", "Comma-separated host:port list; its LENGTH is the number of "
       "parameter-server shards."),
    _s("worker_hosts", "localhost:2223,localhost:2224", "Comma-separated host:port list; its "
       "LENGTH is the number of workers (replicas)."),
    _s("job_name", None, "ps or worker; unset = serial (single-process) training."),
    _s("data_format", "channels_first", "External tensor layout convention (checkpoints / data "
       "files are layout independent; compute is always NHWC bf16 on device)."),
    _n("num_intra_threads", 0, "CPU intra-op threads (0 = library default)."),
    _n("num_inter_threads", 0, "CPU inter-op threads (0 = library default)."),
    _b("use_horovod", False, "Allreduce (Horovod-style) gradient averaging instead of the "
       "parameter-server path."),
    _n("hidden_units", 100, "Hidden units of the debug MLP (LRNet)."),
]

EXTENSION_FLAGS: List[FlagDef] = [
    _n("resnet_size", 50, "Network depth: CIFAR 6n+2 (20, 32, 44, 50, 56, 110, ...); ImageNet "
       "18, 34, 50, 101, 152, 200."),
    _n("resnet_version", 2, "2 = pre-activation (the reference's model), 1 = original post-activation."),
    _s("model", "resnet", "resnet, or lrnet for the reference's plumbing MLP (logist_model.py)."),
    _n("num_ps", 0, "Sugar: synthesise --ps_hosts with this many localhost entries."),
    _n("num_workers", 0, "Sugar: synthesise --worker_hosts with this many localhost entries."),
    _s("backend", "auto", "auto | native (sm_100a kernels) | torch (PyTorch oracle / CPU)."),
    _s("dtype", "bf16", "bf16 or fp32 compute (fp32 only with --backend=torch)."),
    _x("weight_decay", -1.0, "L2 coefficient; <0 = reference default (2e-4 CIFAR, 1e-4 ImageNet)."),
    _x("momentum", 0.9, "SGD momentum (reference: 0.9)."),
    _s("optimizer", "mom", "mom or sgd (reference HParams.optimizer)."),
    _x("base_lr", 0.1, "Base learning rate of the schedules (reference: 0.1)."),
    _b("compat_ps_constant_lr", False, "Reproduce the reference's PS-mode behaviour of applying a "
       "constant 0.1 learning rate (SURVEY.md 2.9 item 1)."),
    _n("save_checkpoint_secs", -1, "Checkpoint period in seconds (<0: 60 CIFAR, 600 ImageNet)."),
    _n("save_checkpoint_steps", 0, "Checkpoint every N steps instead of by time (0 = off)."),
    _n("keep_checkpoint_max", 5, "Number of recent checkpoints kept."),
    _n("log_every", -1, "Log period in steps (<0: 20 CIFAR, 40 ImageNet, as the reference)."),
    _n("summary_every", 100, "Summary period in steps."),
    _n("eval_interval_secs", 60, "Evaluator polling period."),
    _n("seed", 0, "Base RNG seed (initialisation is identical on all replicas)."),
    _b("synthetic_data", False, "Force synthetic inputs even if a data path is given."),
    _s("input_decode", "auto", "ImageNet JPEG decode + resize: cpu (PIL on the host, as the reference), "
       "gpu (nvJPEG batch decode - hardware JPEG engine when present - + one resize/crop kernel), "
       "auto = gpu when training on a CUDA device with a GPU decoder available."),
    _n("input_workers", 0, "ImageNet host input path: number of reader PROCESSES (0 = threads in the "
       "training process, as the reference's num_threads=5)."),
    _b("profile", False, "Print the parameter / FLOP table (tfprof-style) and per-kernel timings."),
    _b("cuda_graph", True, "Capture the training step in a CUDA graph (native backend)."),
    _s("allreduce_algo", "auto", "Gradient allreduce of the Horovod path: auto (one-shot for buckets <= 4 MB "
       "when NVLS multicast is available, else two-shot) | oneshot | twoshot | nccl (library baseline) | "
       "none; suffix _p2p (or plain p2p) forces peer loads/stores instead of multimem instructions."),
    _s("ps_pull", "push", "Parameter-server path, how workers get the new weights: push (the shard owner "
       "stores them into every replica) | pull (the first convolution that consumes a layer fetches it from "
       "the owners over NVLink)."),
    _b("deterministic", False, "Native backend: bit-reproducible steps - every floating-point accumulation "
       "across thread blocks happens in a fixed order (ordered split-K, ordered BatchNorm / loss reductions, "
       "peer loads in rank order instead of switch reductions). Slower; for parity runs."),
    _n("max_steps_per_run", 0, "Stop this process after N local steps (testing resume)."),
    _s("init_method", "", "torch.distributed init method; default env:// or file-based for "
       "--job_name launches."),
    _b("legacy_input", False, "CIFAR: use the legacy queue-reader augmentation (pad 4 instead of 8, "
       "cifar_input.py:66-75)."),
]


def make_flagset(**default_overrides) -> FlagSet:
    return FlagSet(REFERENCE_FLAGS + EXTENSION_FLAGS).with_defaults(**default_overrides)


CIFAR_MAIN = dict()
IMAGENET_MAIN = dict(dataset="imagenet", batch_size=128, train_steps=200)
CIFAR_EVAL = dict(mode="eval")
IMAGENET_EVAL = dict(dataset="imagenet", mode="eval", batch_size=128, num_epochs=3000)
