"""Synchronous data-parallel strategies of the reference, re-expressed on one NVSwitch box.

``AllreduceReducer``  - the Horovod path (resnet_model.py:114-116): gradient buckets are averaged by
    the in-kernel allreduce (csrc/comm.cu) on a side stream as soon as backward has produced them,
    followed (same stream) by the momentum-SGD update of that bucket; only the last bucket is
    exposed.
``ParameterServerReducer`` - the SyncReplicasOptimizer path (resnet_model.py:102-112): the flat
    arena is partitioned over ``num_ps`` shard owners; one fused kernel per bucket does gradient
    push + accumulate + momentum apply + weight pull.  ``num_ps`` and ``num_workers`` stay
    independent quantities as in the reference's ``--ps_hosts`` / ``--worker_hosts`` lists.

Both keep the reference's semantics: mean gradient over replicas, learning rate not scaled by the
number of workers, per-replica BatchNorm statistics, one barrier ("token") per step.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional, Tuple

import torch
import torch.distributed as dist

from .._native import lib as _nl
from .symm import SymmetricBuffer

CH_ALLREDUCE, CH_PS, CH_BCAST, CH_BARRIER = 0, 1, 2, 3


def _p(t) -> int:
    return 0 if t is None else t.data_ptr()


class _ReducerBase:
    scale_in_optimizer = False

    def __init__(self, rank: int, world: int, device, bucket_elems: int = 4 << 20,
                 comm_blocks: int = 32, provider: Optional[str] = None, algo: str = "auto"):
        self.rank, self.world_size = rank, world
        self.device = torch.device(device)
        self.bucket_elems = bucket_elems
        self.comm_blocks = comm_blocks
        self.provider = provider
        self.algo = algo
        self.comm_stream = torch.cuda.Stream(device=self.device)
        self._ev_pool: List[torch.cuda.Event] = []
        self._ev_next = 0
        pad_bytes = _nl.load().drn_comm_pad_bytes()
        self.pads = SymmetricBuffer(pad_bytes, self.device, provider=provider)
        self.total = 0

    # ---- storage handed to the engine ---------------------------------------------------------
    def allocate(self, total_elems: int) -> dict:
        """Symmetric gradient (fp32) and working-weight (bf16) arenas."""
        self.total = total_elems
        self.grad_buf = SymmetricBuffer(total_elems * 4, self.device, provider=self.provider)
        self.shadow_buf = SymmetricBuffer(total_elems * 2, self.device, provider=self.provider)
        return {"grad": self.grad_buf.view(torch.float32), "shadow": self.shadow_buf.view(torch.bfloat16)}

    def _event(self) -> torch.cuda.Event:
        if self._ev_next == len(self._ev_pool):
            self._ev_pool.append(torch.cuda.Event())
        ev = self._ev_pool[self._ev_next]
        self._ev_next += 1
        return ev

    def use_multicast(self) -> bool:
        if self.algo in ("p2p", "oneshot_p2p"):
            return False
        return bool(self.grad_buf.multicast_ptr) and self.algo in ("auto", "multicast", "nvls", "twoshot", "oneshot")

    def barrier(self) -> None:
        _nl.call("drn_comm_barrier", self.pads.ptr_array(), self.rank, self.world_size, CH_BARRIER,
                 torch.cuda.current_stream().cuda_stream)

    def broadcast_arena(self, engine, root: int = 0) -> None:
        """Chief's variables -> every replica (BroadcastGlobalVariablesHook(0) equivalent): master,
        momentum are staged through the symmetric gradient arena."""
        st = torch.cuda.current_stream().cuda_stream
        for arena in (engine.master, engine.mom):
            engine.grad.copy_(arena)
            _nl.call("drn_bcast_f32", self.grad_buf.ptr_array(), self.pads.ptr_array(), self.rank,
                     self.world_size, root, self.total, self.comm_blocks, CH_BCAST, st)
            arena.copy_(engine.grad)
        engine.grad.zero_()
        for bn in engine._all_bns():
            dist.broadcast(bn.run_mean, root)
            dist.broadcast(bn.run_var, root)
        engine.refresh_working_weights()
        torch.cuda.synchronize(self.device)

    # ---- bucketing ------------------------------------------------------------------------------
    def _run_backward(self, engine, on_bucket) -> None:
        """engine.backward() with ranges coalesced into buckets of >= bucket_elems (last first)."""
        pending: List[Tuple[int, int]] = []
        self._ev_next = 0

        def hook(lo: int, hi: int, last: bool):
            pending.append((lo, hi))
            lo_all = min(p[0] for p in pending)
            hi_all = max(p[1] for p in pending)
            if last or hi_all - lo_all >= self.bucket_elems:
                pending.clear()
                on_bucket(lo_all, hi_all)

        engine.backward(range_hook=hook)


class AllreduceReducer(_ReducerBase):
    """Bucketed in-kernel allreduce + per-bucket optimizer on the communication stream."""

    def backward_with_reduction(self, engine) -> None:
        cur = torch.cuda.current_stream()
        cs = self.comm_stream
        mc = self.grad_buf.mc() if self.use_multicast() else 0
        skip = self.algo == "none"

        def on_bucket(lo: int, hi: int):
            ev = self._event()
            ev.record(cur)
            cs.wait_event(ev)
            # the filter gradients of this range run on the engine's second stream: the bucket waits
            # for them here, the dependent chain on `cur` never does
            wg = engine.last_wgrad_event()
            if wg is not None:
                cs.wait_event(wg)
            with torch.cuda.stream(cs):
                st = cs.cuda_stream
                if not skip:
                    _nl.call("drn_allreduce_f32", self.grad_buf.ptr_array(), mc, self.pads.ptr_array(),
                             self.rank, self.world_size, lo, hi - lo, 1.0 / self.world_size,
                             self.comm_blocks, CH_ALLREDUCE, st)
                _nl.call("drn_sgd_momentum", _p(engine.master[lo:]), _p(engine.grad[lo:]),
                         _p(engine.mom[lo:]), _p(engine.shadow[lo:]), hi - lo, _p(engine.hyper),
                         _p(engine.scalars[1:]), 1, st)

        self._run_backward(engine, on_bucket)
        cur.wait_stream(cs)


class ParameterServerReducer(_ReducerBase):
    """Sharded optimizer state on ``num_ps`` owners, fused push/apply/pull kernel per bucket."""

    def __init__(self, rank, world, device, num_ps: Optional[int] = None, **kw):
        super().__init__(rank, world, device, **kw)
        self.num_ps = num_ps or world
        if not (1 <= self.num_ps <= world):
            raise ValueError(f"num_ps={self.num_ps} must be in [1, {world}]")

    def backward_with_reduction(self, engine) -> None:
        cur = torch.cuda.current_stream()
        cs = self.comm_stream

        def on_bucket(lo: int, hi: int):
            ev = self._event()
            ev.record(cur)
            cs.wait_event(ev)
            # the filter gradients of this range run on the engine's second stream: the bucket waits
            # for them here, the dependent chain on `cur` never does
            wg = engine.last_wgrad_event()
            if wg is not None:
                cs.wait_event(wg)
            with torch.cuda.stream(cs):
                _nl.call("drn_ps_update", self.grad_buf.ptr_array(), self.shadow_buf.ptr_array(),
                         _p(engine.master), _p(engine.mom), _p(engine.hyper), self.rank,
                         self.world_size, self.num_ps, lo, hi - lo, self.pads.ptr_array(),
                         _p(engine.scalars[1:]), self.comm_blocks, CH_PS, cs.cuda_stream)

        self._run_backward(engine, on_bucket)
        cur.wait_stream(cs)

    def gather_master(self, engine) -> None:
        """Collect every shard's fp32 master / momentum slices on all ranks (checkpointing): each
        bucket was split evenly over the owners, so replay the same partition."""
        for arena in (engine.master, engine.mom):
            dist.all_reduce(self._owned_only(engine, arena), op=dist.ReduceOp.SUM)

    def _owned_only(self, engine, arena):
        # zero everything this rank does not own, then a SUM allreduce reassembles the arena
        mask = torch.zeros_like(arena, dtype=torch.bool)
        for lo, hi in engine.bucket_ranges(self.bucket_elems):
            n4 = (hi - lo) // 4
            per = (n4 + self.num_ps - 1) // self.num_ps
            b = min(per * self.rank, n4)
            e = min(b + per, n4)
            if self.rank < self.num_ps:
                mask[lo + 4 * b: lo + 4 * e] = True
        arena.mul_(mask)
        return arena


def make_reducer(mode: str, rank: int, world: int, device, algo: str = "auto",
                 num_ps: Optional[int] = None, **kw):
    if mode in ("allreduce", "horovod"):
        return AllreduceReducer(rank, world, device, algo=algo, **kw)
    if mode in ("ps", "parameter_server"):
        return ParameterServerReducer(rank, world, device, num_ps=num_ps, algo=algo, **kw)
    raise ValueError(f"unknown parallel mode {mode!r}")


class LocalBucketedOptimizer:
    """Single replica: nothing to reduce, but the momentum-SGD update still runs bucket by bucket on
    a side stream as soon as backward has finished an arena range (the way both data-parallel
    strategies do), so only the last bucket's update is exposed at the end of the step."""

    def __init__(self, device, bucket_elems: int = 4 << 20):
        self.device = torch.device(device)
        self.bucket_elems = bucket_elems
        self.stream = torch.cuda.Stream(device=self.device)
        self._ev_pool: List[torch.cuda.Event] = []
        self._ev_next = 0

    def _event(self) -> torch.cuda.Event:
        if self._ev_next == len(self._ev_pool):
            self._ev_pool.append(torch.cuda.Event())
        ev = self._ev_pool[self._ev_next]
        self._ev_next += 1
        return ev

    def backward_and_update(self, engine) -> None:
        cur = torch.cuda.current_stream()
        cs = self.stream
        self._ev_next = 0
        pending: List[Tuple[int, int]] = []

        def hook(lo: int, hi: int, last: bool):
            pending.append((lo, hi))
            lo_all = min(p[0] for p in pending)
            hi_all = max(p[1] for p in pending)
            if not (last or hi_all - lo_all >= self.bucket_elems):
                return
            pending.clear()
            ev = self._event()
            ev.record(cur)
            cs.wait_event(ev)
            wg = engine.last_wgrad_event()      # filter gradients of the range (second stream)
            if wg is not None:
                cs.wait_event(wg)
            _nl.call("drn_sgd_momentum", _p(engine.master[lo_all:]), _p(engine.grad[lo_all:]),
                     _p(engine.mom[lo_all:]), _p(engine.shadow[lo_all:]), hi_all - lo_all, _p(engine.hyper),
                     _p(engine.scalars[1:]), 1, cs.cuda_stream)

        engine.backward(range_hook=hook)
        cur.wait_stream(cs)
