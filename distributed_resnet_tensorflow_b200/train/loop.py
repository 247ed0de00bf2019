"""Training driver: the framework's equivalent of the reference's ``train()`` + ``main()``
(resnet_cifar_main.py:250-399, resnet_cifar_main_horovod.py:249-403, resnet_imagenet_main.py:185-319)
and of what ``tf.train.MonitoredTrainingSession`` does around the hot loop [SURVEY.md 3.1-3.3]:

  * role / rank discovery from ``--job_name --task_index --ps_hosts --worker_hosts`` (PS mode) or the
    launcher's environment (allreduce mode, ``--use_horovod=True``)
  * chief initialises or restores from the latest checkpoint in ``--log_root``; the others receive
    the variables (BroadcastGlobalVariablesHook / "non-chief waits for chief" equivalent)
  * per step: learning-rate hook (one-step lag, as in the reference) -> train step -> logging
    every 20/40 steps on every worker -> summaries every 100 steps (chief) -> time-based
    checkpoints (chief; 60 s CIFAR, 600 s ImageNet) -> stop at ``--train_steps``
  * final checkpoint on exit.
"""
from __future__ import annotations

import logging
import os
import queue
import threading
import time
from typing import Dict, Iterator, Optional, Tuple

import numpy as np

from ..ckpt.saver import Saver, latest_checkpoint
from ..config import Config
from ..models.init import init_variables
from ..models.spec import NetSpec, build_spec
from ..utils.events import EventFileWriter
from .backends import DistContext, NativeBackend, TorchBackend
from .lr import LaggedSchedule

log = logging.getLogger("drn")


# ------------------------------------------------------------------------------------------------
# rendezvous
# ------------------------------------------------------------------------------------------------
def _env_int(*names: str) -> Optional[int]:
    for n in names:
        v = os.environ.get(n)
        if v not in (None, ""):
            return int(v)
    return None


def make_context(cfg: Config, native: bool) -> DistContext:
    c = cfg.cluster
    if c.use_horovod and c.job_name is None:
        rank = _env_int("RANK", "OMPI_COMM_WORLD_RANK", "PMI_RANK", "SLURM_PROCID") or 0
        world = _env_int("WORLD_SIZE", "OMPI_COMM_WORLD_SIZE", "PMI_SIZE", "SLURM_NTASKS") or 1
        return DistContext(rank, world, world, 0, "worker", rank, "allreduce" if world > 1 else "serial")
    if c.job_name is None:
        return DistContext()
    W, P = c.num_workers, c.num_ps
    mode = "allreduce" if c.use_horovod else "ps"
    if native or mode == "allreduce":
        # shards are hosted on the workers' GPUs: the process group contains workers only
        if c.job_name == "ps":
            return DistContext(-1, W, W, P, "ps", -1, mode)
        return DistContext(c.task_index, W, W, min(P, W), "worker", c.task_index, mode if W > 1 else "serial")
    if c.job_name == "ps":
        return DistContext(W + c.task_index, W + P, W, P, "ps", -1, "ps")
    return DistContext(c.task_index, W + P, W, P, "worker", c.task_index, "ps")


def init_process_group(cfg: Config, ctx: DistContext, device_type: str) -> None:
    import torch.distributed as dist
    if not ctx.distributed or ctx.rank < 0 or dist.is_initialized():
        return
    backend = "nccl" if device_type == "cuda" else "gloo"
    init = cfg.init_method
    if not init:
        if cfg.cluster.job_name is not None:
            host, port = cfg.cluster.worker_hosts[0].rsplit(":", 1)
            host = "127.0.0.1" if host == "localhost" else host
            init = f"tcp://{host}:{int(port) + 1000}"
        else:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29500")
            init = "env://"
    dist.init_process_group(backend, init_method=init, rank=ctx.rank, world_size=ctx.world)
    # collectives among the workers only (checkpoint agreement): with CPU parameter-server tasks in
    # the process group the ps ranks sit in their serve loop and must not be part of them
    ctx.worker_group = (dist.new_group(ranks=list(range(ctx.num_workers)))
                        if ctx.num_workers < ctx.world else None)


# ------------------------------------------------------------------------------------------------
# data
# ------------------------------------------------------------------------------------------------
def use_gpu_decode(cfg: Config, device: Optional[str]) -> bool:
    """--input_decode: `gpu` needs a CUDA device and a GPU JPEG decoder (the library's nvJPEG binding,
    csrc/jpeg.cu, or torchvision's); `auto` takes the device path whenever that is available - JPEG
    decode, VGG resize and crop then run on the GPU (drn_jpeg_decode_batch + drn_resize_crop_batch)
    and the host only parses records; `cpu` is the reference's host pipeline."""
    mode = cfg.train.input_decode
    if mode == "cpu" or device is None or not str(device).startswith("cuda"):
        return False
    from ..data.imagenet import gpu_decode_available
    ok = gpu_decode_available()
    if mode == "gpu" and not ok:
        raise RuntimeError("--input_decode=gpu needs the library's nvJPEG binding or torchvision with CUDA JPEG decoding")
    return ok


def training_batches(cfg: Config, spec: NetSpec, ctx: DistContext, device: Optional[str] = None) -> Iterator[Tuple]:
    B = cfg.train.batch_size
    seed = cfg.train.seed + 1000 * max(ctx.worker_index, 0)  # replicas shuffle independently
    if spec.dataset == "imagenet":
        if cfg.paths.train_data_path and not cfg.train.synthetic_data:
            from ..data.imagenet import ImageNetBatches, ImageNetBatchesGPU
            if use_gpu_decode(cfg, device):
                return iter(ImageNetBatchesGPU(cfg.paths.train_data_path, B, True, cfg.train.num_epochs, seed,
                                               device=device))
            if cfg.train.input_workers > 0:
                from ..data.imagenet import MultiProcessBatches
                return iter(MultiProcessBatches(cfg.paths.train_data_path, B, True, cfg.train.num_epochs, seed,
                                                workers=cfg.train.input_workers))
            return iter(ImageNetBatches(cfg.paths.train_data_path, B, True, cfg.train.num_epochs, seed))
        from ..data.synthetic import synthetic_imagenet_batches

        def gen():
            rng = np.random.default_rng(seed)
            for imgs, labs in synthetic_imagenet_batches(B, spec.num_classes, spec.image_size, seed):
                crop = np.stack([np.zeros(B), np.zeros(B), rng.integers(0, 2, B)], 1).astype(np.int32)
                yield imgs, labs, crop
        return gen()
    from ..data import cifar
    if cfg.paths.train_data_path and not cfg.train.synthetic_data:
        imgs, labs = cifar.load_records(cifar.train_filenames(cfg.paths.train_data_path, cfg.model.dataset),
                                        cfg.model.dataset)
    else:
        from ..data.synthetic import synthetic_cifar
        imgs, labs = synthetic_cifar(4096, spec.num_classes, seed=cfg.train.seed)
    it = cifar.CifarBatches(imgs, labs, B, True, None, seed)
    return ((i, l, None) for i, l in it)


# ------------------------------------------------------------------------------------------------
# backend construction
# ------------------------------------------------------------------------------------------------
def choose_backend(cfg: Config) -> str:
    b = cfg.train.backend
    if b != "auto":
        return b
    try:
        import torch
        has_gpu = torch.cuda.is_available() and cfg.num_gpus > 0
    except Exception:
        has_gpu = False
    if has_gpu and cfg.model.model == "resnet" and cfg.train.dtype == "bf16":
        return "native"
    return "torch"


def local_device(cfg: Config, ctx: DistContext, backend: str) -> str:
    """worker i uses GPU (task_index % num_gpus) (resnet_cifar_main.py:382-390)."""
    if cfg.num_gpus <= 0 or ctx.role == "ps":
        return "cpu"
    import torch
    if not torch.cuda.is_available():
        if backend == "native":
            raise RuntimeError("--backend=native needs a CUDA device")
        return "cpu"
    local = _env_int("LOCAL_RANK")
    idx = local if local is not None else max(ctx.worker_index, 0) % min(cfg.num_gpus, torch.cuda.device_count())
    return f"cuda:{idx}"


def build_backend(cfg: Config, spec: NetSpec, ctx: DistContext, backend: str, device: str, batch: int):
    t = cfg.train
    pad = 2 if t.legacy_input else 4
    if backend == "native":
        return NativeBackend(spec, batch, ctx, device=device, weight_decay=t.weight_decay,
                             momentum=t.momentum if t.optimizer == "mom" else 0.0, seed=t.seed,
                             cifar_pad=pad, use_graph=t.cuda_graph, algo=t.allreduce_algo, ps_pull=t.ps_pull,
                             deterministic=t.deterministic)
    dtype = "fp32" if device == "cpu" else t.dtype
    return TorchBackend(spec, batch, ctx, device=device, dtype=dtype, weight_decay=t.weight_decay,
                        momentum=t.momentum if t.optimizer == "mom" else 0.0, seed=t.seed, cifar_pad=pad)


def profile_report(spec: NetSpec, batch: int) -> str:
    """tfprof-style parameter / FLOP table (the one profiling capability of the reference,
    resnet_single.py:58-66)."""
    lines = [f"{'variable':40s} {'shape':>22s} {'params':>12s}"]
    for name, shp in spec.trainable_variables():
        lines.append(f"{name:40s} {str(tuple(shp)):>22s} {int(np.prod(shp)):12d}")
    lines.append(f"total trainable parameters: {spec.num_params()}")
    macs = spec.forward_macs_per_image()
    lines.append(f"forward: {2 * macs / 1e9:.3f} GFLOP/image; forward+backward ~ {6 * macs / 1e9:.3f} GFLOP/image; "
                 f"{6 * macs * batch / 1e12:.3f} TFLOP/step at batch {batch}")
    return "\n".join(lines)


# ------------------------------------------------------------------------------------------------
# main loop
# ------------------------------------------------------------------------------------------------
def train(cfg: Config) -> Dict[str, float]:
    import torch
    if cfg.num_intra_threads:
        torch.set_num_threads(cfg.num_intra_threads)
    if cfg.num_inter_threads:
        try:
            torch.set_num_interop_threads(cfg.num_inter_threads)
        except RuntimeError:
            pass
    if cfg.model.model == "lrnet":
        from .lrnet import train_lrnet
        return train_lrnet(cfg)
    spec = build_spec(cfg.model.dataset, cfg.model.resnet_size, cfg.model.resnet_version)
    backend_name = choose_backend(cfg)
    ctx = make_context(cfg, native=(backend_name == "native"))
    if cfg.train.profile and ctx.is_chief:
        log.info("\n%s", profile_report(spec, cfg.train.batch_size))
    if ctx.rank < 0:
        return _idle_ps(cfg)
    device = local_device(cfg, ctx, backend_name)
    if device.startswith("cuda"):
        torch.cuda.set_device(torch.device(device))
    init_process_group(cfg, ctx, "cuda" if device.startswith("cuda") else "cpu")
    be = build_backend(cfg, spec, ctx, backend_name, device, cfg.train.batch_size)
    log.info("role=%s rank=%d/%d mode=%s backend=%s device=%s", ctx.role, ctx.rank, ctx.world, ctx.mode,
             backend_name, device)

    # ---- initialise / restore (chief decides; identical on every replica) -------------------------
    variables = init_variables(spec, cfg.train.seed)
    ckpt = latest_checkpoint(cfg.paths.log_root) if cfg.paths.log_root else None
    if ckpt is not None:
        restored = Saver.restore(ckpt)
        variables.update(restored)
        log.info("restored %s (global_step %d)", ckpt, int(variables["global_step"]))
    be.import_variables(variables)
    if ctx.distributed:
        _assert_replicas_agree(variables, ctx, device)
        if isinstance(be, NativeBackend) and be.comm is not None:
            # BroadcastGlobalVariablesHook(0) (resnet_cifar_main_horovod.py:316): chief's master,
            # momentum and BN statistics overwrite every replica's through the NVLink broadcast kernel
            be.comm.broadcast_arena(be.trainer.engine, root=0)

    if ctx.role == "ps":
        while be.ps_serve_step():
            pass
        return {}
    if ctx.is_chief and cfg.paths.log_root:
        try:
            os.remove(os.path.join(cfg.paths.log_root, "train.done"))
        except OSError:
            pass

    # ---- hooks ------------------------------------------------------------------------------------
    t = cfg.train
    sched = LaggedSchedule(cfg.model.dataset, t.base_lr,
                           constant=(t.compat_ps_constant_lr and ctx.mode == "ps"))
    sched.after_step(be.global_step)  # resume: derive the rate from the restored step
    saver = Saver(cfg.paths.log_root, t.keep_checkpoint_max) if (cfg.paths.log_root and ctx.is_chief) else None
    summary_dir = cfg.paths.train_dir if cfg.model.dataset == "imagenet" else (cfg.paths.eval_dir or cfg.paths.train_dir)
    writer = EventFileWriter(summary_dir) if (ctx.is_chief and summary_dir) else None
    pipe = _StepPipe(be)
    feed = _Prefetcher(training_batches(cfg, spec, ctx, device if backend_name == "native" else None),
                       t.batch_size, pipe.prepare)
    start_step = be.global_step
    t_log, s_log = time.time(), start_step
    metrics: Dict[str, float] = {}
    handle = None
    local_steps = 0
    sharded = ctx.distributed and ctx.mode == "ps"     # saving is a collective over the workers
    try:
        while be.global_step < t.train_steps:
            if t.max_steps_per_run and local_steps >= t.max_steps_per_run:
                break
            lr = sched.before_step()
            staged = feed.next()
            if staged is None:
                log.info("input exhausted")
                break
            # the step is only ENQUEUED here (native backend): the host keeps running ahead, staging
            # the next batch while the device works, and looks at the metrics only when a hook needs them
            handle = pipe.submit(staged, lr)
            local_steps += 1
            if t.profile and local_steps == 3 and ctx.is_chief and not ctx.distributed and hasattr(be, "profile_step"):
                log.info("per-kernel device time of one training step:\n%s", be.profile_step())
            step = be.global_step
            sched.after_step(step)
            want_log = step % t.log_every == 0
            want_summary = writer is not None and step % t.summary_every == 0
            if want_log or want_summary:
                metrics = pipe.metrics(handle)
            if want_log:
                now = time.time()
                sps = (step - s_log) / max(now - t_log, 1e-9)
                t_log, s_log = now, step
                log.info("step = %d, loss = %.5f, precision = %.4f, lr = %.5f (%.2f global_step/sec, %.1f images/sec)",
                         step, metrics["cost"], metrics["precision"], lr, sps,
                         sps * t.batch_size * max(ctx.num_workers, 1))
                if writer is not None:
                    writer.add_scalars({"global_step/sec": sps}, step)
            if want_summary:
                writer.add_scalars({"cross_entropy": metrics["cross_entropy"], "cost": metrics["cost"],
                                    "learning_rate": lr, "Precision": metrics["precision"]}, step)
                writer.flush()
            if _checkpoint_due(t, ctx, saver, step, sharded, device):
                _save(be, saver, ctx)
        if handle is not None:
            metrics = pipe.metrics(handle)
    except Exception as e:
        from .native import CollectiveTimeout
        if isinstance(e, CollectiveTimeout):
            # the context is gone and the peers are in the same state: no checkpoint can be written;
            # leave with a distinct code so that the launcher restarts the job from the last one
            log.error("%s - exiting with code %d, resume from the latest checkpoint in %s", e,
                      EXIT_COLLECTIVE_TIMEOUT, cfg.paths.log_root or "(none)")
            feed.close()
            logging.shutdown()
            os._exit(EXIT_COLLECTIVE_TIMEOUT)
        raise
    finally:
        feed.close()
    # final checkpoint (MonitoredTrainingSession saves on exit)
    if cfg.paths.log_root and not (saver is not None and saver.last_save_step == be.global_step
                                   and not sharded):
        _save(be, saver, ctx)
    if ctx.distributed and ctx.mode == "ps" and isinstance(be, TorchBackend):
        be.ps_signal_stop()
    if writer is not None:
        writer.close()
    if ctx.is_chief and cfg.paths.log_root:
        open(os.path.join(cfg.paths.log_root, "train.done"), "w").write(str(be.global_step))
    metrics["global_step"] = be.global_step
    return metrics


EXIT_COLLECTIVE_TIMEOUT = 86


def _checkpoint_due(t, ctx: DistContext, saver: Optional[Saver], step: int, sharded: bool, device: str) -> bool:
    """CheckpointSaverHook: every --save_checkpoint_steps steps, else every --save_checkpoint_secs
    seconds (60 s CIFAR / 600 s ImageNet, resnet_cifar_main.py:313,329).  With sharded optimizer
    state (parameter-server mode) saving is a collective, so the chief's clock decides for everybody:
    at every logging step it broadcasts one flag to the other workers."""
    if t.save_checkpoint_steps:
        return (saver is not None or sharded) and step % t.save_checkpoint_steps == 0
    if t.save_checkpoint_secs <= 0:
        return False
    mine = saver is not None and time.time() - saver.last_save_time >= t.save_checkpoint_secs
    if not sharded:
        return mine
    if step % t.log_every != 0:
        return False
    import torch
    import torch.distributed as dist
    flag = torch.tensor([1 if mine else 0], dtype=torch.int32,
                        device=device if device.startswith("cuda") else "cpu")
    dist.broadcast(flag, src=0, group=getattr(ctx, "worker_group", None))
    return bool(int(flag.item()))


class _StepPipe:
    """One interface over both backends: the native one enqueues steps and hands out metrics on
    request (prepare_batch / submit_step / fetch_metrics), the PyTorch one runs a step synchronously."""

    def __init__(self, be):
        self.be = be
        self.pipelined = hasattr(be, "submit_step")

    def prepare(self, batch):
        return self.be.prepare_batch(*batch) if self.pipelined else batch

    def submit(self, staged, lr: float):
        if self.pipelined:
            return self.be.submit_step(staged, lr)
        imgs, labs, crop = staged
        return self.be.train_step(imgs, labs, lr, crop)

    def metrics(self, handle):
        return self.be.fetch_metrics(handle) if self.pipelined else handle


class _Prefetcher:
    """Input thread: pulls batches from the reader, drops short ones, copies them into the backend's
    pinned staging ring (``prepare``) and keeps two of them queued, so that batch assembly overlaps
    the device step (tf.data's prefetch in the reference, resnet_cifar_main.py:228-236)."""

    def __init__(self, batches: Iterator[Tuple], batch_size: int, prepare, depth: int = 2):
        self._q: "queue.Queue" = queue.Queue(maxsize=depth)
        self._stop = threading.Event()
        self._thread = threading.Thread(target=self._run, args=(batches, batch_size, prepare), daemon=True,
                                        name="drn-input")
        self._thread.start()

    def _run(self, batches, batch_size, prepare) -> None:
        try:
            for imgs, labs, crop in batches:
                if self._stop.is_set():
                    return
                if len(labs) != batch_size:
                    continue
                item = ("ok", prepare((imgs, labs, crop)))
                while not self._stop.is_set():
                    try:
                        self._q.put(item, timeout=0.2)
                        break
                    except queue.Full:
                        pass
            self._put_final(("end", None))
        except BaseException as e:          # surfaced on the training thread
            self._put_final(("error", e))

    def _put_final(self, item) -> None:
        while not self._stop.is_set():
            try:
                self._q.put(item, timeout=0.2)
                return
            except queue.Full:
                pass

    def next(self):
        kind, val = self._q.get()
        if kind == "ok":
            return val
        if kind == "error":
            raise val
        return None

    def close(self) -> None:
        """Stop the thread and wait for it: a daemon thread that is still inside numpy / a reader when the
        interpreter shuts down can take the process down with it."""
        self._stop.set()
        try:
            while True:
                self._q.get_nowait()
        except queue.Empty:
            pass
        if self._thread is not threading.current_thread():
            self._thread.join(timeout=10.0)


def _save(be, saver: Optional[Saver], ctx: DistContext) -> None:
    if ctx.distributed and ctx.mode == "ps":
        variables = be.export_variables()  # collective over all ranks
    elif saver is not None:
        variables = be.export_variables()
    else:
        return
    if saver is not None:
        path = saver.save(variables, be.global_step)
        log.info("saved checkpoint %s", path)


def _assert_replicas_agree(variables, ctx: DistContext, device: str) -> None:
    """Cheap stand-in for the reference's rank-0 broadcast: all replicas initialise from the same
    seed / checkpoint; verify it with a checksum all-reduce instead of moving 100+ MB."""
    import torch
    import torch.distributed as dist
    from ..models.init import variables_checksum
    c = torch.tensor([variables_checksum(variables)], dtype=torch.float64,
                     device=device if device.startswith("cuda") else "cpu")
    lo, hi = c.clone(), c.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    if float(lo) != float(hi):
        raise RuntimeError("replicas disagree on the initial variables (different seed or checkpoint?)")


def _idle_ps(cfg: Config) -> Dict[str, float]:
    """``--job_name=ps`` with the native backend: parameter shard i lives on worker i's GPU, so the
    ps process has nothing to serve; it stays up (like ``server.join()``) until training ends."""
    log.info("ps task %d: shards are GPU-resident on the workers; waiting for training to finish",
             cfg.cluster.task_index)
    done = os.path.join(cfg.paths.log_root or ".", "train.done")
    t0 = time.time()
    while not os.path.exists(done) and time.time() - t0 < 7 * 24 * 3600:
        time.sleep(1.0)
    return {}
