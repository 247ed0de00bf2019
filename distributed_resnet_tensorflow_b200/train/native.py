"""Training / evaluation driver around the native engine: pinned-host -> device input staging,
device-side augmentation, CUDA-graph capture of the whole step, gradient collectives, metrics.

One ``NativeTrainer`` lives in each process (one process per GPU).  A step is

    H2D copy of the uint8 batch (copy stream, double buffered)
    -> augmentation kernel (csrc/aug.cu)                              [compute stream]
    -> captured graph: forward, loss, backward, gradient collective, optimizer
    -> async D2H of (loss_sum, l2_sum, correct) into pinned memory

and corresponds to one ``mon_sess.run(model.train_op)`` of the reference
(resnet_cifar_main.py:336-337); the reference's hooks (LR setter, logging, summaries,
checkpoints) live in train/loop.py and only talk to this class through ``train_step`` /
``export_variables`` / ``import_variables``.
"""
from __future__ import annotations

from typing import Dict, Optional

import os

import numpy as np
import torch

from .._native import lib as _nl
from ..models.engine import VGG_MEANS, NativeResNet, _p, make_engine
from ..models.spec import NetSpec


class CollectiveTimeout(RuntimeError):
    """An in-kernel barrier wait hit the watchdog (a peer replica died or stalled): the CUDA context
    of this process is gone, the caller should exit and let the launcher resume from the last
    checkpoint (what MonitoredTrainingSession's session re-creation does for the reference,
    resnet_cifar_main.py:311-335)."""

    def __init__(self, code: int, cause: str = ""):
        super().__init__(f"in-kernel barrier wait 0x{code:x} timed out (peer replica dead or stalled?) {cause}")
        self.code = code


def raise_if_watchdog(exc: BaseException) -> None:
    """Turn a CUDA error that followed a watchdog trap into CollectiveTimeout; otherwise re-raise."""
    try:
        code = int(_nl.load().drn_watchdog_status())
    except Exception:
        code = 0
    if code > 0:
        raise CollectiveTimeout(code, str(exc)[:200]) from exc
    raise exc


class NativeTrainer:
    def __init__(self, spec: NetSpec, batch: int, device: str = "cuda:0", weight_decay: float = 1e-4,
                 momentum: float = 0.9, seed: int = 0, use_graph: bool = True, comm=None,
                 cifar_pad: int = 4, src_hw: Optional[int] = None):
        self.spec = spec
        self.B = batch
        self.device = torch.device(device)
        torch.cuda.set_device(self.device)
        storage = None
        if comm is not None:
            from ..models.engine import ParamLayout
            storage = comm.allocate(ParamLayout(spec).total)
        self.engine = make_engine(spec, batch, device=device, storage=storage)
        if comm is not None and hasattr(comm, "attach"):
            comm.attach(self.engine)
        self.weight_decay = weight_decay
        self.momentum = momentum
        self.seed = seed
        self.comm = comm                    # parallel.GradientReducer or None
        # single replica: per-bucket optimizer under backward (DRN_LOCAL_BUCKETS=0: one update at the end)
        self._local_opt = None
        if comm is None and os.environ.get("DRN_LOCAL_BUCKETS", "1") != "0":
            from ..parallel.reducers import LocalBucketedOptimizer
            self._local_opt = LocalBucketedOptimizer(self.device)
        self.world = comm.world_size if comm is not None else 1
        self.rank = comm.rank if comm is not None else 0
        self.use_graph = use_graph
        self.cifar_pad = cifar_pad
        self._graph: Optional[torch.cuda.CUDAGraph] = None
        self._eval_graph: Optional[torch.cuda.CUDAGraph] = None
        self._lr = None
        self.copy_stream = torch.cuda.Stream(device=self.device)
        # staging (double buffered)
        if spec.dataset == "imagenet":
            self.src_hw = src_hw or spec.image_size
            img_bytes = self.src_hw * self.src_hw * 3
        else:
            self.src_hw = 32
            img_bytes = 3 * 32 * 32
        self.img_bytes = img_bytes
        self._dev_u8 = [torch.empty((batch, img_bytes), dtype=torch.uint8, device=self.device) for _ in range(2)]
        self._dev_lab = [torch.empty(batch, dtype=torch.int32, device=self.device) for _ in range(2)]
        self._dev_crop = [torch.zeros((batch, 3), dtype=torch.int32, device=self.device) for _ in range(2)]
        self._h2d_done = [torch.cuda.Event() for _ in range(2)]
        self._consumed = [torch.cuda.Event() for _ in range(2)]
        self._slot = 0
        # metrics ring in pinned memory
        self._host_metrics = torch.zeros((2, 4), dtype=torch.float32).pin_memory()
        self._host_correct = torch.zeros((2, 1), dtype=torch.int32).pin_memory()
        self._metrics_done = [torch.cuda.Event() for _ in range(2)]
        # hyper-parameters travel through a small ring of pinned rows: a row is rewritten only after
        # the H2D copy that read it has completed (the step loop runs ahead of the device, and the
        # ImageNet warm-up changes the rate every step)
        self._host_hyper = torch.zeros((8, 4), dtype=torch.float32).pin_memory()
        self._hyper_done = [None] * 8
        self._hyper_next = 0
        _nl.check(_nl.load().drn_watchdog_arm(), "drn_watchdog_arm")
        # launch-bound CIFAR-sized steps: programmatic dependent launch for the convolution kernels - the
        # dependent is released when the last CTA of its predecessor starts its last tile (late trigger,
        # csrc/common.cuh), so launch latency and set-up overlap the predecessor's drain: 1 040 -> 1 061
        # steps/s at 16 images (1 100 with the element-wise kernels included, DRN_PDL=3: measured at the
        # very end of the round, after the last run of the numerics suite, hence opt-in), neutral at 128.
        # ImageNet-sized steps: +0.4 % with the convolution kernels only (within noise), -2 % with the
        # element-wise kernels included (the early-resident convolution CTAs hold the shared memory the
        # filter-gradient stream wants) - left off.
        _nl.load().drn_set_pdl_mask(2 if spec.dataset != "imagenet" else -1)
        self.steps_done = 0
        self.h2d_bytes_per_step = batch * img_bytes + batch * 4 + batch * 12
        self.d2h_bytes_per_step = 4 * 4 + 4

    # ---- variables ------------------------------------------------------------------------------
    def import_variables(self, variables: Dict[str, np.ndarray]) -> None:
        self.engine.import_variables(variables)

    def export_variables(self, with_slots: bool = True) -> Dict[str, np.ndarray]:
        return self.engine.export_variables(with_slots)

    @property
    def global_step(self) -> int:
        return self.engine.global_step

    # ---- step body ------------------------------------------------------------------------------
    def _step_body(self) -> None:
        e = self.engine
        e.scalars.zero_()
        e.correct.zero_()
        # parameter-server mode: every owner adds its share of sum(w^2) into every replica's pad
        l2_slot = getattr(self.comm, "l2_slot", None)
        pull = l2_slot is not None and self.comm.pull
        if l2_slot is not None and not pull:
            l2_slot.zero_()
        e.forward(training=True)
        if pull:
            # no closing barrier in pull mode: the owners' contributions of the previous update are
            # complete once their epoch flags were seen, which forward() has just waited for
            e.scalars[1:2].copy_(l2_slot)
            l2_slot.zero_()
        e.loss(with_grad=True)
        if self.comm is not None:
            self.comm.backward_with_reduction(e)
            if l2_slot is not None and not pull:
                e.scalars[1:2].copy_(l2_slot)
        elif self._local_opt is not None:
            self._local_opt.backward_and_update(e)
        else:
            e.backward()
            e.apply_gradients()

    def _set_lr(self, lr: float) -> None:
        if self._lr == lr:
            return
        self._lr = lr
        gs = 1.0 / self.world if (self.comm is not None and self.comm.scale_in_optimizer) else 1.0
        i = self._hyper_next
        self._hyper_next = (i + 1) % len(self._hyper_done)
        if self._hyper_done[i] is not None:
            self._hyper_done[i].synchronize()
        row = self._host_hyper[i]
        row[0], row[1], row[2], row[3] = lr, self.momentum, self.weight_decay, gs
        self.engine.hyper.copy_(row, non_blocking=True)
        if self._hyper_done[i] is None:
            self._hyper_done[i] = torch.cuda.Event()
        self._hyper_done[i].record(torch.cuda.current_stream())

    def _augment(self, slot: int, training: bool, step: int) -> None:
        e = self.engine
        st = torch.cuda.current_stream().cuda_stream
        if self.spec.dataset == "imagenet":
            _nl.call("drn_imagenet_prep", _p(self._dev_u8[slot]), self.img_bytes, self.src_hw, self.src_hw,
                     _p(self._dev_crop[slot]), _p(e.x0), self.B, self.spec.image_size, VGG_MEANS[0],
                     VGG_MEANS[1], VGG_MEANS[2], e.in_pitch, e.in_xoff, st)
        else:
            _nl.call("drn_cifar_augment", _p(self._dev_u8[slot]), self.img_bytes, 0, _p(e.x0), self.B,
                     self.cifar_pad, 1 if training else 0, self.seed, step & 0xffffffff,
                     self.rank * self.B + (step * self.B * self.world), 16, st)
        e.labels.copy_(self._dev_lab[slot], non_blocking=True)

    def stage_batch(self, images_u8: torch.Tensor, labels_i32: torch.Tensor,
                    crop: Optional[torch.Tensor] = None) -> int:
        """Start the H2D copy of a pinned host batch; returns the slot to pass to run_staged()."""
        slot = self._slot
        self._slot ^= 1
        cs = self.copy_stream
        cs.wait_event(self._consumed[slot])
        with torch.cuda.stream(cs):
            self._dev_u8[slot].copy_(images_u8.view(self.B, -1), non_blocking=True)
            self._dev_lab[slot].copy_(labels_i32, non_blocking=True)
            if crop is not None:
                self._dev_crop[slot].copy_(crop, non_blocking=True)
            self._h2d_done[slot].record(cs)
        return slot

    def run_staged(self, slot: int, lr: float) -> int:
        """Enqueue one training step on the staged batch; returns the metrics slot."""
        cur = torch.cuda.current_stream()
        cur.wait_event(self._h2d_done[slot])
        self._set_lr(lr)
        self._augment(slot, True, self.steps_done)
        self._consumed[slot].record(cur)
        if self.use_graph:
            if self._graph is None:
                self._capture()
            self._graph.replay()
        else:
            self._step_body()
        m = self.steps_done & 1
        self._host_metrics[m].copy_(self.engine.scalars, non_blocking=True)
        self._host_correct[m].copy_(self.engine.correct, non_blocking=True)
        self._metrics_done[m].record(cur)
        self.steps_done += 1
        self.engine.global_step += 1
        return m

    def read_metrics(self, m: int) -> Dict[str, float]:
        try:
            self._metrics_done[m].synchronize()
        except RuntimeError as e:       # a trapped kernel surfaces here: name the barrier if it was ours
            raise_if_watchdog(e)
        loss_sum, l2 = float(self._host_metrics[m, 0]), float(self._host_metrics[m, 1])
        xent = loss_sum / self.B
        return {"cross_entropy": xent, "cost": xent + self.weight_decay * 0.5 * l2,
                "precision": float(self._host_correct[m, 0]) / self.B}

    def train_step(self, images_u8: torch.Tensor, labels_i32: torch.Tensor, lr: float,
                   crop: Optional[torch.Tensor] = None) -> Dict[str, float]:
        """Synchronous convenience wrapper: stage, run, read back this step's metrics."""
        slot = self.stage_batch(images_u8, labels_i32, crop)
        return self.read_metrics(self.run_staged(slot, lr))

    def _capture(self) -> None:
        # warm-up on a side stream (sets kernel attributes, builds tensor maps' code paths)
        s = torch.cuda.Stream(device=self.device)
        s.wait_stream(torch.cuda.current_stream())
        snapshot = self._snapshot()
        with torch.cuda.stream(s):
            self._step_body()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize(self.device)
        self._restore(snapshot)
        g = torch.cuda.CUDAGraph()
        before = _nl.launch_count()
        # capture on a high-priority stream: kernel nodes inherit it, so the dependent chain wins the
        # SMs over the filter-gradient branch (engine._wg_stream, lowest priority)
        with torch.cuda.graph(g, stream=torch.cuda.Stream(device=self.device, priority=-1)):
            self._step_body()
        self._launches_in_graph = _nl.launch_count() - before
        torch.cuda.synchronize(self.device)
        self._restore(snapshot)   # capture does not execute, but keep state exact regardless
        self._graph = g

    def profile_step(self) -> str:
        """Per-entry-point device time of one eager training step on the currently staged input
        (state is restored afterwards). The capability behind ``--profile``."""
        snap = self._snapshot()
        self._step_body()                      # warm
        torch.cuda.synchronize(self.device)
        self._restore(snap)
        side, self.engine._wg_stream = self.engine._wg_stream, None   # one stream: events bracket every launch
        _nl.start_call_profile()
        try:
            self._step_body()
        finally:
            rows = _nl.stop_call_profile()
            self.engine._wg_stream = side
        self._restore(snap)
        total = sum(r[2] for r in rows) or 1.0
        lines = [f"{'entry point':28s} {'calls':>6s} {'total us':>10s} {'share':>7s}"]
        for name, calls, us in rows:
            lines.append(f"{name:28s} {calls:6d} {us:10.1f} {100 * us / total:6.1f}%")
        lines.append(f"{'sum of kernels':28s} {sum(r[1] for r in rows):6d} {total:10.1f}")
        return "\n".join(lines)

    def launches_per_step(self) -> int:
        """Kernels of libdrn_b200 enqueued per training step (graph body + augmentation)."""
        if self._graph is not None:
            return self._launches_in_graph + 1
        before = _nl.launch_count()
        snap = self._snapshot()
        self._step_body()
        torch.cuda.synchronize(self.device)
        self._restore(snap)
        return _nl.launch_count() - before + 1

    def _snapshot(self):
        e = self.engine
        bn = [(b.run_mean.clone(), b.run_var.clone()) for b in e._all_bns()]
        return (e.master.clone(), e.mom.clone(), e.shadow.clone(), e.grad.clone(), bn)

    def _restore(self, snap) -> None:
        e = self.engine
        master, mom, shadow, grad, bn = snap
        e.master.copy_(master)
        e.mom.copy_(mom)
        e.shadow.copy_(shadow)
        e.grad.copy_(grad)
        for b, (rm, rv) in zip(e._all_bns(), bn):
            b.run_mean.copy_(rm)
            b.run_var.copy_(rv)
        e.prep_dgrad_weights()

    # ---- evaluation -----------------------------------------------------------------------------
    def eval_batch(self, images_u8: torch.Tensor, labels_i32: torch.Tensor,
                   crop: Optional[torch.Tensor] = None) -> Dict[str, float]:
        e = self.engine
        slot = self.stage_batch(images_u8, labels_i32, crop)
        cur = torch.cuda.current_stream()
        cur.wait_event(self._h2d_done[slot])
        self._augment(slot, False, 0)
        self._consumed[slot].record(cur)
        e.scalars.zero_()
        e.correct.zero_()
        e.forward(training=False)
        e.loss(with_grad=False)
        torch.cuda.synchronize(self.device)
        return {"loss_sum": float(e.scalars[0]), "correct": int(e.correct[0])}
