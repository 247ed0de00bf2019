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

``--allreduce_algo`` (``algo``):
    twoshot   every rank reduces 1/N of the bucket and writes the result to all ranks (fused
              reduce-scatter + all-gather, csrc/comm.cu allreduce_f32_kernel), then momentum-SGD
    oneshot   every rank reduces the WHOLE bucket itself and applies momentum-SGD in the same kernel
              (allreduce_sgd_oneshot_kernel): no stores to peers, best for small buckets
    auto      oneshot for buckets of at most ONESHOT_MAX_BYTES when the fabric offers multicast
              (the switch does the additions), twoshot otherwise
    nccl      torch.distributed.all_reduce on the bucket (the library baseline; 1/N in the optimizer)
    none      no reduction (measures exposed communication by difference)
  a ``_p2p`` suffix (``twoshot_p2p``, ``oneshot_p2p``, or plain ``p2p``) forces peer loads/stores
  instead of NVLS multimem instructions; ``multicast`` / ``nvls`` are aliases of twoshot over NVLS.
``--ps_pull`` (``pull``): push (owner stores the new weights into every replica) or pull (the first
  convolution that consumes a layer fetches it from the owners, drn_conv_igemm_ex).
"""
from __future__ import annotations

import ctypes
from typing import List, Optional, Tuple

import torch
import torch.distributed as dist

from .._native import lib as _nl
from .symm import SymmetricBuffer

CH_ALLREDUCE, CH_PS, CH_BCAST, CH_BARRIER = 0, 1, 2, 3
ONESHOT_MAX_BYTES = 256 << 10   # measured on 8 B200 (profiles/bench_8gpu_r2.jsonl): at 3 MB the one-shot kernel loses to two-shot (1.145 vs 1.116 ms/step on CIFAR)
ALGOS = ("auto", "oneshot", "twoshot", "nccl", "none", "p2p", "oneshot_p2p", "twoshot_p2p", "multicast", "nvls")


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
        if algo not in ALGOS:
            raise ValueError(f"unknown --allreduce_algo {algo!r}; one of {', '.join(ALGOS)}")
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
        """NVLS multimem instructions unless the fabric has no multicast or a *_p2p algorithm was asked for."""
        if self.algo.endswith("p2p") or _nl.load().drn_deterministic():
            return False      # deterministic mode: peers are read in a fixed rank order
        return bool(self.grad_buf.multicast_ptr)

    def algorithm_for(self, nbytes: int) -> str:
        """oneshot | twoshot | nccl | none for a bucket of ``nbytes`` (see the module docstring)."""
        a = self.algo
        if a in ("nccl", "none"):
            return a
        if a.startswith("oneshot"):
            return "oneshot"
        if a == "auto" and nbytes <= ONESHOT_MAX_BYTES and self.use_multicast():
            return "oneshot"
        return "twoshot"

    def barrier(self) -> None:
        _nl.call("drn_comm_barrier", self.pads.ptr_array(), self.rank, self.world_size, CH_BARRIER,
                 torch.cuda.current_stream().cuda_stream)

    def broadcast_arena(self, engine, root: int = 0) -> None:
        """Chief's variables -> every replica (BroadcastGlobalVariablesHook(0) equivalent): master,
        momentum are staged through the symmetric gradient arena."""
        st = torch.cuda.current_stream().cuda_stream
        mc = self.grad_buf.mc() if self.use_multicast() else 0
        bns = engine._all_bns()
        stats = torch.cat([t for bn in bns for t in (bn.run_mean, bn.run_var)])   # moving statistics ride along
        n_stats = (stats.numel() + 3) // 4 * 4
        assert n_stats <= self.total
        for arena, n in ((engine.master, self.total), (engine.mom, self.total), (stats, n_stats)):
            engine.grad[:arena.numel()].copy_(arena)
            _nl.call("drn_bcast_f32", self.grad_buf.ptr_array(), mc, self.pads.ptr_array(), self.rank,
                     self.world_size, root, n, self.comm_blocks, CH_BCAST, st)
            arena.copy_(engine.grad[:arena.numel()])
        engine.grad.zero_()
        off = 0
        for bn in bns:
            for t in (bn.run_mean, bn.run_var):
                t.copy_(stats[off:off + t.numel()])
                off += t.numel()
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

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        # the NCCL arm leaves the sum in the arena; the optimizer kernel applies 1/N (hyper[3])
        self.scale_in_optimizer = self.algo == "nccl"
        self.last_algorithms: List[str] = []

    def backward_with_reduction(self, engine) -> None:
        cur = torch.cuda.current_stream()
        cs = self.comm_stream
        mc = self.grad_buf.mc() if self.use_multicast() else 0
        self.last_algorithms = []

        def on_bucket(lo: int, hi: int):
            ev = self._event()
            ev.record(cur)
            cs.wait_event(ev)
            # the filter gradients of this range run on the engine's second stream: the bucket waits
            # for them here, the dependent chain on `cur` never does
            wg = engine.last_wgrad_event()
            if wg is not None:
                cs.wait_event(wg)
            kind = self.algorithm_for((hi - lo) * 4)
            self.last_algorithms.append(kind)
            with torch.cuda.stream(cs):
                st = cs.cuda_stream
                if kind == "oneshot":
                    _nl.call("drn_allreduce_sgd_oneshot", self.grad_buf.ptr_array(), mc, _p(engine.master),
                             _p(engine.mom), _p(engine.shadow), _p(engine.hyper), self.rank, self.world_size,
                             lo, hi - lo, self.pads.ptr_array(), _p(engine.scalars[1:]), self.comm_blocks,
                             CH_ALLREDUCE, st)
                    return
                if kind == "twoshot":
                    _nl.call("drn_allreduce_f32", self.grad_buf.ptr_array(), mc, self.pads.ptr_array(),
                             self.rank, self.world_size, lo, hi - lo, 1.0 / self.world_size,
                             self.comm_blocks, CH_ALLREDUCE, st)
                elif kind == "nccl":
                    dist.all_reduce(engine.grad[lo:hi])
                _nl.call("drn_sgd_momentum", _p(engine.master[lo:]), _p(engine.grad[lo:]),
                         _p(engine.mom[lo:]), _p(engine.shadow[lo:]), hi - lo, _p(engine.hyper),
                         _p(engine.scalars[1:]), 1, st)

        self._run_backward(engine, on_bucket)
        cur.wait_stream(cs)


class ParameterServerReducer(_ReducerBase):
    """Sharded optimizer state on ``num_ps`` owners, fused push/apply/pull kernel per bucket."""

    def __init__(self, rank, world, device, num_ps: Optional[int] = None, pull: str = "push", **kw):
        super().__init__(rank, world, device, **kw)
        self.num_ps = num_ps or world
        if not (1 <= self.num_ps <= world):
            raise ValueError(f"num_ps={self.num_ps} must be in [1, {world}]")
        if pull not in ("push", "pull"):
            raise ValueError("pull must be 'push' or 'pull'")
        self.pull = pull == "pull"
        lib = _nl.load()
        self._l2_off = lib.drn_comm_pad_offset(2)
        self.l2_slot = self.pads.view(torch.float32, self._l2_off, 1)
        self._buckets: Optional[List[Tuple[int, int]]] = None

    # ---- shard geometry -----------------------------------------------------------------------
    def per_owner(self, lo: int, hi: int) -> int:
        return int(_nl.load().drn_shard_per(hi - lo, self.num_ps))

    def attach(self, engine) -> None:
        """Called once by the trainer: vector variables (BatchNorm gamma/beta, dense bias) are read
        from the bf16 working arena on EVERY replica - only the shard owner has the fp32 master up to
        date - and, in pull mode, every convolution gets the description of where its filter bank
        lives (drn_pull_desc)."""
        import numpy as np
        self._buckets = engine.bucket_ranges(self.bucket_elems)
        if len(self._buckets) > 64:
            raise ValueError("more than 64 buckets: raise bucket_elems")
        maps = np.zeros(len(self._buckets), dtype=[("lo", "<i8"), ("per", "<i8"), ("num_ps", "<i4"), ("bucket", "<i4")])
        for i, (lo, hi) in enumerate(self._buckets):
            maps[i] = (lo, self.per_owner(lo, hi), self.num_ps, i)
        self._maps_host = maps
        self._maps = torch.from_numpy(maps.view(np.uint8).copy()).to(self.device)
        rows = []
        for name in engine.layout.order:
            e = engine.layout.entries[name]
            if e.kind == "vector":
                rows.append((e.offset, e.numel, self.bucket_of(e.offset)))
        tab = np.zeros(len(rows), dtype=[("off", "<i8"), ("count", "<i4"), ("map", "<i4")])
        for i, r in enumerate(rows):
            tab[i] = r
        self._vec_table = torch.from_numpy(tab.view(np.uint8).copy()).to(self.device)
        self._vec_entries = len(rows)
        engine.pre_forward = self.sync_vectors
        if self.pull:
            engine.pull_plan = _PullPlan(self, engine)

    def bucket_of(self, offset: int) -> int:
        for i, (lo, hi) in enumerate(self._buckets):
            if lo <= offset < hi:
                return i
        raise KeyError(offset)

    def sync_vectors(self, engine) -> None:
        _nl.call("drn_ps_sync_vectors", self.shadow_buf.ptr_array(), _p(engine.shadow), _p(engine.vec_f32),
                 _p(self._vec_table), self._vec_entries, _p(self._maps), self.pads.local.data_ptr(),
                 self.rank, self.world_size, 1 if self.pull else 0, torch.cuda.current_stream().cuda_stream)

    def allocate(self, total_elems: int) -> dict:
        st = super().allocate(total_elems)
        st["vec_f32"] = torch.zeros(total_elems, device=self.device)
        return st

    def backward_with_reduction(self, engine) -> None:
        cur = torch.cuda.current_stream()
        cs = self.comm_stream
        mc = self.grad_buf.mc() if self.use_multicast() else 0
        index = {r: i for i, r in enumerate(self._buckets)} if self._buckets else {}

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
                _nl.call("drn_ps_update", self.grad_buf.ptr_array(), mc, self.shadow_buf.ptr_array(),
                         _p(engine.master), _p(engine.mom), _p(engine.hyper), self.rank,
                         self.world_size, self.num_ps, lo, hi - lo, self.pads.ptr_array(),
                         1, 1 if self.pull else 0, index.get((lo, hi), 0), self.comm_blocks, CH_PS,
                         cs.cuda_stream)

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
            n = hi - lo
            per = self.per_owner(lo, hi)
            b = min(per * self.rank, n)
            e = min(b + per, n)
            if self.rank < self.num_ps:
                mask[lo + b: lo + e] = True
        arena.mul_(mask)
        return arena


class _PullPlan:
    """Per-layer ``drn_pull_desc`` structures (pull mode): where the layer's filter bank lives in the
    arena, which bucket / owners hold it, and two counter words for the in-kernel hand-off."""

    def __init__(self, red: ParameterServerReducer, engine):
        names = [n for n in engine.layout.order if engine.layout.entries[n].kind != "vector"]
        self._counters = torch.zeros(2 * len(names), dtype=torch.int32, device=red.device)
        self._shadows = red.shadow_buf.ptr_array()
        self.desc = {}
        for i, name in enumerate(names):
            e = engine.layout.entries[name]
            b = red.bucket_of(e.offset)
            lo, hi = red._buckets[b]
            count = (e.numel + 7) // 8 * 8        # entries are 64-element aligned: the tail is padding
            d = _nl.PullDesc(ctypes.cast(self._shadows, ctypes.c_void_p), red.world_size, red.rank,
                             red.pads.local.data_ptr(), engine.shadow.data_ptr(),
                             self._counters[2 * i:].data_ptr(), lo, red.per_owner(lo, hi), red.num_ps, b,
                             e.offset, count)
            self.desc[name] = d

    def get(self, name: str):
        return self.desc.get(name)


def make_reducer(mode: str, rank: int, world: int, device, algo: str = "auto",
                 num_ps: Optional[int] = None, pull: str = "push", **kw):
    if mode in ("allreduce", "horovod"):
        return AllreduceReducer(rank, world, device, algo=algo, **kw)
    if mode in ("ps", "parameter_server"):
        return ParameterServerReducer(rank, world, device, num_ps=num_ps, algo=algo, pull=pull, **kw)
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
