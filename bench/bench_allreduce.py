"""Device-timed microbenchmark of the in-kernel collectives (csrc/comm.cu) against NCCL, max over ranks.
Launch with torch.distributed.run on N GPUs.  Sizes: CIFAR-50 gradient (3 MB fp32), a 16 MB bucket,
the whole ImageNet-50 gradient (102 MB fp32).

For the fused kernels the row also gives the fraction of the NVLink roofline: the bytes that must
cross this GPU's links in the busier direction, divided by 770 GB/s (the measured per-direction peer
bandwidth of the profiling recipe), divided by the measured time.
  two-shot allreduce        in: (N-1)/N * S (my slice from the other ranks)   out: the same
  one-shot allreduce + SGD  in: (N-1) * S without NVLS; S with NVLS (the switch adds)
  ps_update, push           owner of S/N: in (N-1) * S/N of fp32 gradients, out (N-1) * S/N / 2 of bf16 weights
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
LINK_GBS = 770.0


def timed(fn, iters, dev):
    for _ in range(5):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters * 1e3], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return round(float(t), 2)


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    dist.init_process_group("nccl", device_id=torch.device(dev))
    from distributed_resnet_tensorflow_b200._native import lib as nl
    from distributed_resnet_tensorflow_b200.parallel.reducers import CH_ALLREDUCE, CH_PS, ParameterServerReducer
    n_max = 25_600_000
    red = ParameterServerReducer(rank, world, dev, num_ps=world)
    store = red.allocate(n_max)
    master = torch.zeros(n_max, device=dev)
    mom = torch.zeros(n_max, device=dev)
    hyper = torch.tensor([0.0, 0.9, 0.0, 1.0], device=dev)     # lr 0: the arenas stay finite over many calls
    ref = torch.zeros(n_max, device=dev)
    has_mc = bool(red.grad_buf.multicast_ptr)
    p = lambda t: t.data_ptr()
    for n, blocks in ((758_784, 16), (758_784, 32), (4 << 20, 32), (n_max, 32), (n_max, 64)):
        st = torch.cuda.current_stream().cuda_stream
        S = 4.0 * n
        mcp = lambda mc: red.grad_buf.mc() if mc else 0

        def twoshot(mc):
            nl.call("drn_allreduce_f32", red.grad_buf.ptr_array(), mcp(mc), red.pads.ptr_array(), rank, world, 0, n,
                    1.0 / world, blocks, CH_ALLREDUCE, st)

        def oneshot(mc):
            nl.call("drn_allreduce_sgd_oneshot", red.grad_buf.ptr_array(), mcp(mc), p(master), p(mom),
                    p(store["shadow"]), p(hyper), rank, world, 0, n, red.pads.ptr_array(), 0, blocks, CH_ALLREDUCE, st)

        def ps(mc, pull):
            nl.call("drn_ps_update", red.grad_buf.ptr_array(), mcp(mc), red.shadow_buf.ptr_array(), p(master), p(mom),
                    p(hyper), rank, world, world, 0, n, red.pads.ptr_array(), 0, 1 if pull else 0, 0, blocks, CH_PS, st)

        def nccl():
            dist.all_reduce(ref[:n])
            ref[:n].mul_(1.0 / world)

        row = {"elems": n, "bytes": int(S), "blocks": blocks, "ranks": world,
               "twoshot_p2p_us": timed(lambda: twoshot(False), 30, dev),
               "oneshot_sgd_p2p_us": timed(lambda: oneshot(False), 30, dev),
               "ps_push_p2p_us": timed(lambda: ps(False, False), 30, dev),
               "ps_publish_for_pull_p2p_us": timed(lambda: ps(False, True), 30, dev)}
        if has_mc:
            row["twoshot_multimem_us"] = timed(lambda: twoshot(True), 30, dev)
            row["oneshot_sgd_multimem_us"] = timed(lambda: oneshot(True), 30, dev)
            row["ps_push_multimem_us"] = timed(lambda: ps(True, False), 30, dev)
        row["nccl_allreduce_plus_scale_us"] = timed(nccl, 30, dev)
        f = (world - 1) / world
        us = lambda bytes_: bytes_ / (LINK_GBS * 1e3)          # microseconds at the link bandwidth
        best2 = min(row["twoshot_p2p_us"], row.get("twoshot_multimem_us", 1e30))
        row["twoshot_busbw_GBs"] = round(2 * f * S / best2 * 1e-3, 1)
        row["twoshot_roofline_frac"] = round(us(f * S) / best2, 3)
        best1 = min(row["oneshot_sgd_p2p_us"], row.get("oneshot_sgd_multimem_us", 1e30))
        row["oneshot_roofline_frac"] = round(us(S if has_mc else (world - 1) * S) / best1, 3)
        bestp = min(row["ps_push_p2p_us"], row.get("ps_push_multimem_us", 1e30))
        row["ps_push_roofline_frac"] = round(us(f * S) / bestp, 3)    # ingress (N-1)*S/N of fp32 > egress of bf16
        if rank == 0:
            print(json.dumps(row), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
