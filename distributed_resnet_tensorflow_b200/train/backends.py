"""Executors behind the training / evaluation loops.

``TorchBackend``   plain PyTorch (models/torch_ref.py): CPU plumbing runs, v1 networks, fp32 oracle
                   mode, and the only backend that needs no GPU.  Data parallelism over
                   torch.distributed (gloo on CPU, nccl on GPU): allreduce (Horovod path) or a true
                   parameter-server exchange (reduce to the shard owner, owner applies momentum-SGD,
                   broadcast back) in which ``--job_name=ps`` processes really hold the shards.
``NativeBackend``  the sm_100a engine + in-kernel NVLink collectives (train/native.py, parallel/).

Both expose ``train_step(batch, lr) -> metrics``, ``eval_batch``, ``export_variables`` /
``import_variables`` so that hooks, checkpoints and the evaluator are backend agnostic.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import numpy as np
import torch

from ..data import preprocess
from ..models.spec import NetSpec
from ..models.torch_ref import TorchResNet


class DistContext:
    """Rank bookkeeping for one process."""

    def __init__(self, rank: int = 0, world: int = 1, num_workers: int = 1, num_ps: int = 0,
                 role: str = "worker", worker_index: int = 0, mode: str = "serial"):
        self.rank, self.world = rank, world
        self.num_workers, self.num_ps = num_workers, num_ps
        self.role, self.worker_index, self.mode = role, worker_index, mode

    @property
    def is_chief(self) -> bool:
        return self.role == "worker" and self.worker_index == 0

    @property
    def distributed(self) -> bool:
        return self.world > 1


def _shard_bounds(n: int, parts: int, i: int) -> Tuple[int, int]:
    per = (n + parts - 1) // parts
    return min(per * i, n), min(per * (i + 1), n)


class TorchBackend:
    def __init__(self, spec: NetSpec, batch: int, ctx: DistContext, device: str = "cpu",
                 dtype: str = "fp32", weight_decay: float = 2e-4, momentum: float = 0.9,
                 seed: int = 0, cifar_pad: int = 4):
        self.spec, self.B, self.ctx = spec, batch, ctx
        self.device = torch.device(device)
        self.wd, self.mu, self.seed, self.cifar_pad = weight_decay, momentum, seed, cifar_pad
        tdtype = torch.bfloat16 if dtype == "bf16" else torch.float32
        self.model = TorchResNet(spec, device=device, dtype=tdtype,
                                 channels_last=self.device.type == "cuda")
        self.steps_done = 0
        self._ps_master: Optional[torch.Tensor] = None   # ps role: owned slice of the flat arena
        self._ps_mom: Optional[torch.Tensor] = None

    # ---- variables ------------------------------------------------------------------------------
    def import_variables(self, variables: Dict[str, np.ndarray]) -> None:
        self.model.import_variables(variables)
        self._ps_master = None

    def export_variables(self, with_slots: bool = True) -> Dict[str, np.ndarray]:
        if self.ctx.mode == "ps" and self.ctx.distributed and self.ctx.role == "worker":
            self._ps_sync_momentum_to_workers()
        return self.model.export_variables(with_slots)

    @property
    def global_step(self) -> int:
        return self.model.global_step

    # ---- input ------------------------------------------------------------------------------------
    def _prep(self, images_u8: np.ndarray, training: bool, crop: Optional[np.ndarray]) -> torch.Tensor:
        if self.spec.dataset == "imagenet":
            s = self.spec.image_size
            imgs = images_u8.reshape(-1, s, s, 3)
            out = np.stack([preprocess.imagenet_prep_oracle(imgs[i], bool(crop[i, 2]) if crop is not None else False)
                            for i in range(imgs.shape[0])])
        else:
            base = self.ctx.worker_index * self.B + self.steps_done * self.B * max(self.ctx.num_workers, 1)
            out = preprocess.cifar_preprocess_batch(images_u8.reshape(-1, 3072), training, self.cifar_pad,
                                                    self.seed, self.steps_done, base)
        return torch.from_numpy(out).to(self.device)

    # ---- steps --------------------------------------------------------------------------------------
    def train_step(self, images_u8: np.ndarray, labels: np.ndarray, lr: float,
                   crop: Optional[np.ndarray] = None) -> Dict[str, float]:
        import torch.distributed as dist
        m, ctx = self.model, self.ctx
        images = self._prep(images_u8, True, crop)
        lab = torch.from_numpy(np.asarray(labels)).long().to(self.device)
        cost, xent, prec = m.loss_and_grads(images, lab, self.wd)
        if ctx.distributed and ctx.mode == "allreduce":
            flat = m.flat_grads()
            dist.all_reduce(flat)
            flat /= ctx.num_workers
            m.set_flat_grads(flat)
            m.apply_gradients(lr, self.mu, self.wd)
        elif ctx.distributed and ctx.mode == "ps":
            self._ps_worker_exchange(lr)
        else:
            m.apply_gradients(lr, self.mu, self.wd)
        self.steps_done += 1
        return {"cost": float(cost), "cross_entropy": float(xent), "precision": float(prec)}

    def eval_batch(self, images_u8: np.ndarray, labels: np.ndarray,
                   crop: Optional[np.ndarray] = None) -> Dict[str, float]:
        images = self._prep(images_u8, False, crop)
        lab = torch.from_numpy(np.asarray(labels)).long().to(self.device)
        loss, correct = self.model.evaluate(images, lab)
        return {"loss_sum": loss * images.shape[0], "correct": correct}

    # ---- parameter-server exchange (torch.distributed point-to-point collectives) -------------------
    # ranks: workers 0..W-1, ps W..W+P-1. Shard j of the flat parameter vector lives on ps j.
    def _flat_params(self) -> torch.Tensor:
        return torch.cat([self.model.params[n].detach().reshape(-1) for n in self.model._names])

    def _set_flat_params(self, flat: torch.Tensor) -> None:
        off = 0
        with torch.no_grad():
            for n in self.model._names:
                p = self.model.params[n]
                p.copy_(flat[off:off + p.numel()].view_as(p))
                off += p.numel()

    CMD_STEP, CMD_STOP, CMD_GATHER = 0.0, 1.0, 2.0

    def _ps_round(self, flat: Optional[torch.Tensor], lr: float, cmd: float) -> Optional[torch.Tensor]:
        """Worker side of one exchange with every ps shard: push (reduce to the owner), broadcast
        (lr, command) from the chief, pull (broadcast from the owner)."""
        import torch.distributed as dist
        ctx = self.ctx
        n = self._flat_params().numel() if flat is None else flat.numel()
        new_w = torch.empty(n, device=self.device)
        ctl = torch.tensor([lr, cmd], dtype=torch.float32, device=self.device)
        for j in range(ctx.num_ps):
            lo, hi = _shard_bounds(n, ctx.num_ps, j)
            owner = ctx.num_workers + j
            g = flat[lo:hi].contiguous() if flat is not None else torch.zeros(hi - lo, device=self.device)
            dist.reduce(g, dst=owner)
            dist.broadcast(ctl, src=0)
            buf = torch.empty(hi - lo, device=self.device)
            dist.broadcast(buf, src=owner)
            new_w[lo:hi] = buf
        return new_w

    def _ps_worker_exchange(self, lr: float) -> None:
        new_w = self._ps_round(self.model.flat_grads(), lr, self.CMD_STEP)
        self._set_flat_params(new_w)
        self.model.global_step += 1

    def ps_serve_step(self) -> bool:
        """One round of a parameter-server task: accumulate the workers' gradients, apply
        momentum-SGD on the owned shard, publish it. Returns False after the stop command."""
        import torch.distributed as dist
        ctx = self.ctx
        j = ctx.rank - ctx.num_workers
        if self._ps_master is None:
            flat = self._flat_params()
            self._n = flat.numel()
            lo, hi = _shard_bounds(self._n, ctx.num_ps, j)
            self._ps_master = flat[lo:hi].clone()
            mom = torch.cat([self.model.momentum[n].reshape(-1) for n in self.model._names])
            self._ps_mom = mom[lo:hi].clone()
        ctl = torch.zeros(2, dtype=torch.float32, device=self.device)
        cmd = self.CMD_STEP
        for k in range(ctx.num_ps):
            lo, hi = _shard_bounds(self._n, ctx.num_ps, k)
            owner = ctx.num_workers + k
            g = torch.zeros(hi - lo, device=self.device)
            dist.reduce(g, dst=owner)
            dist.broadcast(ctl, src=0)
            cmd = float(ctl[1])
            if k == j and cmd == self.CMD_STEP:
                g /= ctx.num_workers                       # SyncReplicas takes the MEAN of N gradients
                g += self.wd * self._ps_master
                self._ps_mom.mul_(self.mu).add_(g)
                self._ps_master.add_(self._ps_mom, alpha=-float(ctl[0]))
            buf = self._ps_master if k == j else torch.empty(hi - lo, device=self.device)
            dist.broadcast(buf, src=owner)
        if cmd == self.CMD_GATHER:
            self._ps_gather_momentum()
        return cmd != self.CMD_STOP

    def ps_signal_stop(self) -> None:
        """Workers call this once after their last step so that ps tasks leave their serve loop."""
        self._ps_round(None, 0.0, self.CMD_STOP)

    def _ps_sync_momentum_to_workers(self) -> None:
        """Momentum slots live on the ps tasks; fetch them for checkpointing. Collective over the
        workers (the ps tasks join from inside their serve loop), like the reference's Saver
        reading ps-resident variables."""
        self._ps_round(None, 0.0, self.CMD_GATHER)
        self._ps_gather_momentum()

    def _ps_gather_momentum(self) -> None:
        import torch.distributed as dist
        ctx = self.ctx
        n = self._flat_params().numel()
        full = torch.zeros(n, device=self.device)
        if ctx.role == "ps" and self._ps_mom is not None:
            j = ctx.rank - ctx.num_workers
            lo, hi = _shard_bounds(n, ctx.num_ps, j)
            full[lo:hi] = self._ps_mom
        dist.all_reduce(full)
        off = 0
        for name in self.model._names:
            mbuf = self.model.momentum[name]
            mbuf.copy_(full[off:off + mbuf.numel()].view_as(mbuf))
            off += mbuf.numel()


class NativeBackend:
    def __init__(self, spec: NetSpec, batch: int, ctx: DistContext, device: str = "cuda:0",
                 weight_decay: float = 1e-4, momentum: float = 0.9, seed: int = 0, cifar_pad: int = 4,
                 use_graph: bool = True, algo: str = "auto", ps_pull: str = "push", deterministic: bool = False):
        from ..train.native import NativeTrainer
        self.spec, self.B, self.ctx = spec, batch, ctx
        if deterministic:
            # process-wide library switch (include/drn_b200.h: drn_set_deterministic); the reducers
            # consult it too (peer loads in rank order instead of switch reductions)
            from .._native import lib as _nl
            _nl.call("drn_set_deterministic", 1)
        comm = None
        if ctx.distributed:
            from ..parallel import make_reducer
            comm = make_reducer("ps" if ctx.mode == "ps" else "allreduce", ctx.rank, ctx.world, device,
                                algo=algo, num_ps=ctx.num_ps or None, pull=ps_pull)
        self.comm = comm
        self.trainer = NativeTrainer(spec, batch, device=device, weight_decay=weight_decay,
                                     momentum=momentum, seed=seed, use_graph=use_graph, comm=comm,
                                     cifar_pad=cifar_pad)
        # pinned staging ring: slot s is refilled (possibly by the input thread) only after the H2D
        # copy that read it has completed; `_free` carries (slot, event of that copy)
        self.PIN_SLOTS = 3
        self._pin_i = [torch.empty((batch, self.trainer.img_bytes), dtype=torch.uint8).pin_memory()
                       for _ in range(self.PIN_SLOTS)]
        self._pin_l = [torch.empty(batch, dtype=torch.int32).pin_memory() for _ in range(self.PIN_SLOTS)]
        self._pin_c = [torch.zeros((batch, 3), dtype=torch.int32).pin_memory() for _ in range(self.PIN_SLOTS)]
        self._pin_ev = [torch.cuda.Event() for _ in range(self.PIN_SLOTS)]
        import queue
        self._free: "queue.Queue" = queue.Queue()
        for s_ in range(self.PIN_SLOTS):
            self._free.put((s_, None))

    def import_variables(self, variables):
        self.trainer.import_variables(variables)

    def export_variables(self, with_slots: bool = True):
        if self.comm is not None and hasattr(self.comm, "gather_master"):
            torch.cuda.synchronize()
            self.comm.gather_master(self.trainer.engine)
        return self.trainer.export_variables(with_slots)

    @property
    def global_step(self) -> int:
        return self.trainer.global_step

    # ---- pipelined interface (train/loop.py): prepare on the input thread, submit without a host
    # sync, fetch metrics only when somebody looks at them -------------------------------------------
    def prepare_batch(self, images_u8, labels, crop=None):
        """Copy one host batch into a free pinned slot (thread-safe; blocks while all slots are in
        flight). Returns an opaque handle for submit_step()."""
        slot, ev = self._free.get()
        if ev is not None:
            try:
                ev.synchronize()
            except RuntimeError as e:
                from .native import raise_if_watchdog
                raise_if_watchdog(e)
        on_device = isinstance(images_u8, torch.Tensor) and images_u8.is_cuda
        if not on_device:   # (GPU JPEG decode delivers device pixels: no pinned bounce for them)
            self._pin_i[slot].copy_(torch.from_numpy(np.ascontiguousarray(images_u8).reshape(self.B, -1)))
        self._pin_l[slot].copy_(torch.from_numpy(np.asarray(labels, np.int32)))
        if crop is not None:
            self._pin_c[slot].copy_(torch.from_numpy(np.asarray(crop, np.int32)))
        return slot, (images_u8.view(self.B, -1) if on_device else None)

    def _stage_prepared(self, staged) -> int:
        slot, dev_pixels = staged
        pix = dev_pixels if dev_pixels is not None else self._pin_i[slot]
        c = self._pin_c[slot] if self.spec.dataset == "imagenet" else None
        dslot = self.trainer.stage_batch(pix, self._pin_l[slot], c)
        self._pin_ev[slot].record(self.trainer.copy_stream)
        self._free.put((slot, self._pin_ev[slot]))
        return dslot

    def submit_step(self, staged, lr: float) -> int:
        """Enqueue one training step; returns the handle fetch_metrics() takes. No host sync."""
        return self.trainer.run_staged(self._stage_prepared(staged), lr)

    def fetch_metrics(self, handle: int):
        """Metrics of a submitted step (waits for it). Valid until two more steps were submitted."""
        return self.trainer.read_metrics(handle)

    def train_step(self, images_u8, labels, lr, crop=None):
        return self.fetch_metrics(self.submit_step(self.prepare_batch(images_u8, labels, crop), lr))

    def profile_step(self) -> str:
        return self.trainer.profile_step()

    def eval_batch(self, images_u8, labels, crop=None):
        slot, dev_pixels = self.prepare_batch(images_u8, labels, crop)
        pix = dev_pixels if dev_pixels is not None else self._pin_i[slot]
        c = self._pin_c[slot] if self.spec.dataset == "imagenet" else None
        out = self.trainer.eval_batch(pix, self._pin_l[slot], c)    # synchronises the device
        self._free.put((slot, None))
        return out
