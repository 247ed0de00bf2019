"""Single-rank launches of the collective kernels (world size 1: the peer pointers are this GPU's own
buffers) - for `ncu`, which must never wrap a multi-rank command.  Shows the instruction mix and the
local HBM behaviour of allreduce_f32_kernel / ps_update_kernel; link-level numbers come from
bench/bench_allreduce.py on 8 GPUs."""
import os, sys, torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29777")
torch.cuda.set_device(0)
dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device("cuda:0"))
from distributed_resnet_tensorflow_b200._native import lib as nl
from distributed_resnet_tensorflow_b200.parallel.reducers import CH_ALLREDUCE, CH_PS, ParameterServerReducer
n = 25_600_000
red = ParameterServerReducer(0, 1, "cuda:0", num_ps=1)
store = red.allocate(n)
master = torch.zeros(n, device="cuda"); mom = torch.zeros(n, device="cuda")
hyper = torch.tensor([0.0, 0.9, 0.0, 1.0], device="cuda")
st = torch.cuda.current_stream().cuda_stream
mc = red.grad_buf.mc()
print("multicast address available at world size 1:", bool(mc))
for _ in range(4):
    nl.call("drn_allreduce_f32", red.grad_buf.ptr_array(), mc, red.pads.ptr_array(), 0, 1, 0, n, 1.0, 64, CH_ALLREDUCE, st)
    nl.call("drn_ps_update", red.grad_buf.ptr_array(), mc, red.shadow_buf.ptr_array(), master.data_ptr(), mom.data_ptr(),
            hyper.data_ptr(), 0, 1, 1, 0, n, red.pads.ptr_array(), 1, 0, 0, 64, CH_PS, st)
torch.cuda.synchronize()
e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
e0.record()
nl.call("drn_allreduce_f32", red.grad_buf.ptr_array(), mc, red.pads.ptr_array(), 0, 1, 0, n, 1.0, 64, CH_ALLREDUCE, st)
e1.record()
nl.call("drn_ps_update", red.grad_buf.ptr_array(), mc, red.shadow_buf.ptr_array(), master.data_ptr(), mom.data_ptr(),
        hyper.data_ptr(), 0, 1, 1, 0, n, red.pads.ptr_array(), 1, 0, 0, 64, CH_PS, st)
e2.record(); torch.cuda.synchronize()
print(f"allreduce (1 rank, 102 MB): {e0.elapsed_time(e1) * 1e3:.1f} us; ps_update: {e1.elapsed_time(e2) * 1e3:.1f} us")
dist.destroy_process_group()
