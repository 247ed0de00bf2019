"""Times the BN elementwise kernels on ResNet-50 sized tensors (L2 flushed)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_resnet_tensorflow_b200._native import lib as nl
nl.load()
st = lambda: torch.cuda.current_stream().cuda_stream
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for M, C in ((128 * 56 * 56, 256), (128 * 56 * 56, 64), (128 * 14 * 14, 1024), (128 * 7 * 7, 2048)):
    y = torch.randn((M, C), device="cuda").bfloat16(); dm = torch.randn_like(y); ad = torch.randn_like(y); out = torch.empty_like(y)
    ss = torch.rand((C, 2), device="cuda"); coef = torch.rand((C, 4), device="cuda")
    def t(fn):
        for _ in range(2): fn()
        best = 1e9
        for _ in range(5):
            flush.zero_(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); torch.cuda.synchronize(); best = min(best, e0.elapsed_time(e1) * 1e3)
        return best
    a = t(lambda: nl.call("drn_bn_apply", y.data_ptr(), ss.data_ptr(), 0, out.data_ptr(), M, C, 1, st()))
    b = t(lambda: nl.call("drn_bn_bwd_apply", dm.data_ptr(), y.data_ptr(), coef.data_ptr(), 0, out.data_ptr(), M, C, st()))
    c = t(lambda: nl.call("drn_bn_bwd_apply", dm.data_ptr(), y.data_ptr(), coef.data_ptr(), ad.data_ptr(), out.data_ptr(), M, C, st()))
    by = y.numel() * 2
    print(f"M={M} C={C}: apply {a:.1f} us {2*by/a*1e-3:.0f} GB/s | bwd_apply {b:.1f} us {3*by/b*1e-3:.0f} GB/s | +addend {c:.1f} us {4*by/c*1e-3:.0f} GB/s", flush=True)
