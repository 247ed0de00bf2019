"""Times drn_bn_bwd_finalize_apply / drn_bn_finalize_apply on ResNet-50 sized tensors (L2 flushed)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_resnet_tensorflow_b200._native import lib as nl
nl.load()
st = lambda: torch.cuda.current_stream().cuda_stream
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
res = []
for M, C in ((128 * 56 * 56, 256), (128 * 56 * 56, 64), (128 * 28 * 28, 512), (128 * 28 * 28, 128), (128 * 14 * 14, 1024), (128 * 14 * 14, 256), (128 * 7 * 7, 2048)):
    y = torch.randn((M, C), device="cuda").bfloat16(); dm = torch.randn_like(y); ad = torch.randn_like(y); out = torch.empty_like(y)
    stats = torch.rand((2, C), device="cuda"); gamma = torch.rand(C, device="cuda"); beta = torch.rand(C, device="cuda")
    mi = torch.rand((C, 2), device="cuda"); ss = torch.rand((C, 2), device="cuda"); dg = torch.zeros(C, device="cuda"); db = torch.zeros(C, device="cuda")
    rm = torch.zeros(C, device="cuda"); rv = torch.ones(C, device="cuda")
    def t(fn):
        for _ in range(2): fn()
        best = 1e9
        for _ in range(5):
            flush.zero_(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); torch.cuda.synchronize(); best = min(best, e0.elapsed_time(e1) * 1e3)
        return best
    p = lambda t_: t_.data_ptr()
    a = t(lambda: nl.call("drn_bn_finalize_apply", p(y), p(stats), C, M, p(gamma), p(beta), p(rm), p(rv), 0.997, 1e-5, p(ss), p(mi), p(out), C, 1, st()))
    b = t(lambda: nl.call("drn_bn_bwd_finalize_apply", p(dm), p(y), p(stats), C, M, p(gamma), p(mi), p(dg), p(db), 0, p(out), C, st()))
    c = t(lambda: nl.call("drn_bn_bwd_finalize_apply", p(dm), p(y), p(stats), C, M, p(gamma), p(mi), p(dg), p(db), p(ad), p(out), C, st()))
    by = y.numel() * 2
    print(f"M={M:7d} C={C:4d}: fwd {a:6.1f} us {2*by/a*1e-3:5.0f} GB/s | bwd {b:6.1f} us {3*by/b*1e-3:5.0f} GB/s | bwd+add {c:6.1f} us {4*by/c*1e-3:5.0f} GB/s", flush=True)
