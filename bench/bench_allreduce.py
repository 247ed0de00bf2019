"""In-kernel allreduce (csrc/comm.cu) vs NCCL all_reduce (+ separate 1/N scale), device-timed, max over
ranks. Launch with torch.distributed.run on N GPUs. Sizes: CIFAR-50 gradient (3 MB fp32), a 16 MB
bucket, the whole ImageNet-50 gradient (102 MB fp32)."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


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
    return float(t)


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    dist.init_process_group("nccl", device_id=torch.device(dev))
    from distributed_resnet_tensorflow_b200._native import lib as nl
    from distributed_resnet_tensorflow_b200.parallel.reducers import CH_ALLREDUCE, AllreduceReducer
    n_max = 25_600_000
    red = AllreduceReducer(rank, world, dev)
    buf = red.allocate(n_max)["grad"]
    ref = torch.zeros(n_max, device=dev)
    rows = []
    for n, blocks in ((758_784, 16), (758_784, 32), (4 << 20, 32), (n_max, 32), (n_max, 64)):
        st = torch.cuda.current_stream().cuda_stream

        def ours(mc):
            nl.call("drn_allreduce_f32", red.grad_buf.ptr_array(), red.grad_buf.mc() if mc else 0,
                    red.pads.ptr_array(), rank, world, 0, n, 1.0 / world, blocks, CH_ALLREDUCE, st)

        def nccl():
            dist.all_reduce(ref[:n])
            ref[:n].mul_(1.0 / world)

        row = {"elems": n, "bytes": 4 * n, "blocks": blocks, "ranks": world,
               "p2p_us": timed(lambda: ours(False), 30, dev)}
        if red.grad_buf.multicast_ptr:
            row["multimem_us"] = timed(lambda: ours(True), 30, dev)
        row["nccl_plus_scale_us"] = timed(nccl, 30, dev)
        best = min(row["p2p_us"], row.get("multimem_us", 1e30))
        row["busbw_GBs_best"] = round(2 * (world - 1) / world * 4 * n / best * 1e-3, 1)
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
