"""Multi-GPU worker launched by tests/test_comm_gpu.py through torch.distributed.run.
Checks the in-kernel collectives against NCCL and a 2-replica training step against first
principles.  Prints 'COMM_WORKER_OK' on rank 0 when every check passed."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"]); world = int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    dist.init_process_group("nccl", device_id=torch.device(dev))
    from distributed_resnet_tensorflow_b200._native import lib as nl
    from distributed_resnet_tensorflow_b200.parallel.reducers import (CH_ALLREDUCE, AllreduceReducer,
                                                                      ParameterServerReducer)
    from distributed_resnet_tensorflow_b200.parallel.symm import SymmetricBuffer

    provider = os.environ.get("DRN_SYMM", "auto")
    # ---------------- raw allreduce vs NCCL, p2p and (if available) multicast -------------------
    n = 1 << 20
    red = AllreduceReducer(rank, world, dev, provider=provider)
    store = red.allocate(n)
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    for use_mc in (False, True):
        if use_mc and not red.grad_buf.multicast_ptr:
            continue
        x = torch.randn(n, device=dev, generator=g)
        store["grad"].copy_(x)
        want = x.clone()
        dist.all_reduce(want)
        want /= world
        torch.cuda.synchronize(); dist.barrier()
        for (lo, cnt) in ((0, n // 2), (n // 2, n // 2)):
            nl.call("drn_allreduce_f32", red.grad_buf.ptr_array(), red.grad_buf.mc() if use_mc else 0,
                    red.pads.ptr_array(), rank, world, lo, cnt, 1.0 / world, 16, CH_ALLREDUCE,
                    torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize(); dist.barrier()
        err = (store["grad"] - want).abs().max().item()
        assert err < 1e-5, ("allreduce", use_mc, err)
        # every rank must hold bit-identical results
        chk = store["grad"].double().sum().reshape(1)
        allc = [torch.zeros_like(chk) for _ in range(world)]
        dist.all_gather(allc, chk)
        assert all(float(c) == float(allc[0]) for c in allc), "ranks disagree bitwise"
        if rank == 0:
            print(f"allreduce multicast={use_mc} provider={red.grad_buf.provider} maxerr={err:.2e}", flush=True)

    # ---------------- race stress: many back-to-back reductions with random skew between ranks ----
    # ("race stress ... with random stream delays").  Inputs are small integers so the
    # expected mean is exact in fp32; any torn read, missed flag or early overwrite shows up as a
    # non-zero error.  Consecutive calls reuse the same buffer and the same signal pads, which is the
    # situation the 0<->1 flag protocol has to survive inside a replayed CUDA graph.
    iters = int(os.environ.get("DRN_STRESS_ITERS", "2000"))
    rng = np.random.default_rng(1234 + rank)
    worst = torch.zeros((), device=dev)
    ramp = (torch.arange(n, device=dev) % 13).float()
    stream = torch.cuda.current_stream().cuda_stream
    for it in range(iters):
        if rng.random() < 0.5:
            torch.cuda._sleep(int(rng.integers(1, 300000)))   # up to ~150 us of skew on this rank only
        store["grad"].copy_(ramp * (rank + 1) + float(it % 97))
        use_mc = bool(red.grad_buf.multicast_ptr) and (it % 2 == 1)
        nl.call("drn_allreduce_f32", red.grad_buf.ptr_array(), red.grad_buf.mc() if use_mc else 0,
                red.pads.ptr_array(), rank, world, 0, n, 1.0 / world, 8 + 8 * (it % 3), CH_ALLREDUCE, stream)
        expect = ramp * ((world + 1) / 2.0) + float(it % 97)
        worst = torch.maximum(worst, (store["grad"] - expect).abs().max())
    torch.cuda.synchronize(); dist.barrier()
    assert float(worst) == 0.0, ("allreduce stress", float(worst))
    if rank == 0:
        print(f"allreduce stress: {iters} skewed back-to-back reductions exact", flush=True)

    # ---------------- broadcast of the chief's variables (C3) ------------------------------------
    from distributed_resnet_tensorflow_b200.models.init import init_variables as _iv
    from distributed_resnet_tensorflow_b200.models.spec import build_spec as _bs
    from distributed_resnet_tensorflow_b200.train.native import NativeTrainer as _NT
    _spec = _bs("cifar10", 8)
    _comm = AllreduceReducer(rank, world, dev, provider=provider)
    _tr = _NT(_spec, 8, device=dev, comm=_comm, use_graph=False)
    _tr.import_variables(_iv(_spec, seed=100 + rank))         # every rank starts DIFFERENT
    _comm.broadcast_arena(_tr.engine, root=0)
    want_w = _tr.engine.master.clone(); dist.broadcast(want_w, 0)
    assert torch.equal(_tr.engine.master, want_w), "broadcast_arena: master differs from the chief's"
    chief = _iv(_spec, seed=100)
    got = _tr.export_variables()
    assert np.array_equal(got["conv2d_3/kernel"], chief["conv2d_3/kernel"])
    assert torch.equal(_tr.engine.shadow.float(), _tr.engine.master.bfloat16().float())
    if rank == 0:
        print("bcast: chief variables on all ranks ok", flush=True)
    del _tr, _comm

    # ---------------- end-to-end: 2-replica step == mean-gradient step --------------------------
    from distributed_resnet_tensorflow_b200.models.init import init_variables
    from distributed_resnet_tensorflow_b200.models.spec import build_spec
    from distributed_resnet_tensorflow_b200.train.native import NativeTrainer
    spec = build_spec("cifar10", 8)
    B = 16
    variables = init_variables(spec, seed=3)
    rng = np.random.default_rng(7 + rank)
    imgs = torch.from_numpy(rng.integers(0, 256, size=(B, 3072), dtype=np.uint8)).pin_memory()
    labs = torch.from_numpy(rng.integers(0, 10, size=(B,), dtype=np.int32)).pin_memory()
    results = {}
    has_mc = bool(red.grad_buf.multicast_ptr)
    modes = ["allreduce", "allreduce_oneshot", "allreduce_oneshot_p2p", "allreduce_nccl", "ps", "ps_p2p", "ps_pull"]
    for mode in modes:
        nps = max(1, world - 1) if world > 2 else world
        if mode.startswith("allreduce"):
            algo = {"allreduce": "twoshot", "allreduce_oneshot": "oneshot", "allreduce_oneshot_p2p": "oneshot_p2p",
                    "allreduce_nccl": "nccl"}[mode]
            comm = AllreduceReducer(rank, world, dev, provider=provider, algo=algo)
        else:
            comm = ParameterServerReducer(rank, world, dev, num_ps=nps, provider=provider,
                                          algo="p2p" if mode == "ps_p2p" else "auto",
                                          pull="pull" if mode == "ps_pull" else "push")
        is_ps = mode.startswith("ps")
        tr = NativeTrainer(spec, B, device=dev, weight_decay=2e-4, comm=comm,
                           use_graph=mode in ("allreduce", "allreduce_oneshot", "ps_pull"))
        tr.import_variables(variables)
        m = tr.train_step(imgs, labs, lr=0.1)
        torch.cuda.synchronize(); dist.barrier()
        if is_ps:
            comm.gather_master(tr.engine)
        w = tr.engine.master.clone()      # after ONE step (atomics make later steps diverge chaotically)
        for _ in range(2):
            m = tr.train_step(imgs, labs, lr=0.1)
        torch.cuda.synchronize(); dist.barrier()
        if is_ps:
            comm.gather_master(tr.engine)
        if mode == "ps_pull":
            # the working arena of a replica is completed by the forward pass that consumes it
            tr.engine.forward(training=False)
            torch.cuda.synchronize(); dist.barrier()
        sh = tr.engine.shadow.float().clone()
        # identical weights on every rank
        for t, name in ((sh, "shadow"), (tr.engine.master.clone(), "master")):
            ref = t.clone(); dist.broadcast(ref, 0)
            assert torch.equal(ref, t), f"{mode}: {name} differs between ranks"
        assert torch.equal(sh, tr.engine.master.bfloat16().float()), f"{mode}: working copy != bf16(master)"
        if is_ps:
            # every replica computes with the same gamma / beta / bias: the fp32 image of the working copy
            if mode != "ps_pull":
                comm.sync_vectors(tr.engine); torch.cuda.synchronize()
            e = tr.engine
            for name in ("batch_normalization/gamma", "batch_normalization_3/beta", "dense/bias"):
                ent = e.layout.entries[name]
                got = e.vec_f32[ent.offset:ent.offset + ent.numel]
                assert torch.equal(got, e.master[ent.offset:ent.offset + ent.numel].bfloat16().float()), (mode, name)
            assert abs(m["cost"] - m["cross_entropy"]) > 1e-3, f"{mode}: L2 term missing from the cost"
        results[mode] = w
        assert np.isfinite(m["cost"])
        if rank == 0:
            print(f"{mode}: 3 steps ok, cost={m['cost']:.4f} xent={m['cross_entropy']:.4f}"
                  f" multicast={comm.use_multicast()}", flush=True)
        del tr, comm
    # all transports implement the same mathematics
    for mode in modes[1:]:
        diff = (results["allreduce"] - results[mode]).abs().max().item()
        if rank == 0:
            print(f"allreduce-vs-{mode} max weight diff after one step: {diff:.3e}", flush=True)
        assert diff < 1e-3, ("allreduce vs " + mode + " weights after one step", diff)

    # deterministic mode across replicas: two runs of three steps end in bit-identical weights
    from distributed_resnet_tensorflow_b200._native import lib as _nl
    _nl.call("drn_set_deterministic", 1)
    for mode in ("allreduce", "ps"):
        finals = []
        for _ in range(2):
            if mode == "allreduce":
                comm = AllreduceReducer(rank, world, dev, provider=provider, algo="auto")
            else:
                comm = ParameterServerReducer(rank, world, dev, num_ps=world, provider=provider)
            assert not comm.use_multicast(), "deterministic mode reads peers in rank order"
            tr = NativeTrainer(spec, B, device=dev, weight_decay=2e-4, comm=comm, use_graph=True)
            tr.import_variables(variables)
            for _ in range(3):
                tr.train_step(imgs, labs, lr=0.1)
            torch.cuda.synchronize(); dist.barrier()
            if mode == "ps":
                comm.gather_master(tr.engine)
            finals.append(tr.engine.master.clone())
            del tr, comm
        assert torch.equal(finals[0], finals[1]), f"deterministic {mode}: {int((finals[0] != finals[1]).sum())} weights differ"
        if rank == 0:
            print(f"deterministic {mode}: two runs of 3 steps bit-identical", flush=True)
    _nl.call("drn_set_deterministic", -1)

    # single-replica reference: mean of the per-replica gradients applied once
    tr1 = NativeTrainer(spec, B, device=dev, weight_decay=2e-4, comm=None, use_graph=False)
    tr1.import_variables(variables)
    tr1.rank, tr1.world = rank, world      # same augmentation stream as the data-parallel run
    e = tr1.engine
    tr1.stage_batch(imgs, labs); torch.cuda.synchronize()
    tr1._augment(0, True, 0)
    e.scalars.zero_(); e.correct.zero_()
    e.forward(True); e.loss(True); e.backward()
    gmean = e.grad.clone(); dist.all_reduce(gmean); gmean /= world
    e.grad.copy_(gmean)
    tr1.world = 1
    tr1._set_lr(0.1)
    e.apply_gradients()
    comm = AllreduceReducer(rank, world, dev, provider=provider)
    tr2 = NativeTrainer(spec, B, device=dev, weight_decay=2e-4, comm=comm, use_graph=False)
    tr2.import_variables(variables)
    tr2.train_step(imgs, labs, lr=0.1)
    torch.cuda.synchronize()
    d = (tr2.engine.master - e.master).abs().max().item()
    if rank == 0:
        print(f"allreduce dp step vs mean-gradient step: max weight diff {d:.3e}", flush=True)
    assert d < 1e-3, ("dp step vs mean-gradient step", d)
    dist.barrier()
    if rank == 0:
        print("COMM_WORKER_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
