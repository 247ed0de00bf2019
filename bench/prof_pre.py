"""conv1-role 1x1 convolutions: plain conv on a materialised activation vs the BatchNorm-prologue
variant on the raw tensor (plus the stand-alone BN apply it replaces). L2 flushed."""
import ctypes, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_resnet_tensorflow_b200._native import lib as nl
from distributed_resnet_tensorflow_b200.ops.conv import ConvGeom
nl.load()
st = lambda: torch.cuda.current_stream().cuda_stream
p = lambda t: t.data_ptr()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def t(fn, iters=5):
    for _ in range(2): fn()
    best = 1e9
    for _ in range(iters):
        flush.zero_(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); best = min(best, e0.elapsed_time(e1) * 1e3)
    return best
for H, C, K, s in ((56, 256, 64, 1), (28, 512, 128, 1), (14, 1024, 256, 1), (7, 2048, 512, 1), (56, 256, 512, 2), (28, 512, 1024, 2), (56, 256, 128, 1)):
    N = 128
    g = ConvGeom.tf_fixed_padding(N, H, H, C, K, 1, s); d = g.desc()
    x = torch.randn((N, H, H, C), device="cuda").bfloat16(); a = torch.empty_like(x)
    w = (torch.randn((K, 1, 1, C), device="cuda") / C ** 0.5).bfloat16()
    ss = torch.rand((C, 2), device="cuda"); y = torch.empty((N, g.P, g.Q, K), dtype=torch.bfloat16, device="cuda")
    sb = torch.zeros((2, K), device="cuda")
    ta = t(lambda: nl.call("drn_bn_apply", p(x), p(ss), 0, p(a), N * H * H, C, 1, st()))
    tc = t(lambda: nl.call("drn_conv_igemm", ctypes.byref(d), p(a), p(w), p(y), 0, p(sb), K, 0, 0, 0, st()))
    tp = t(lambda: nl.call("drn_conv_igemm_pre", ctypes.byref(d), p(x), p(ss), p(w), p(y), 0, p(sb), K, 0, 0, 0, st()))
    print(f"H={H:2d} C={C:4d} K={K:4d} s{s}: bn_apply {ta:6.1f} us + conv {tc:6.1f} us = {ta+tc:6.1f} | conv_pre {tp:6.1f} us", flush=True)
    dy = torch.randn((N, g.P, g.Q, K), device="cuda").bfloat16(); dw = torch.zeros((K, 1, 1, C), device="cuda")
    tw = t(lambda: nl.call("drn_conv_wgrad", ctypes.byref(d), p(a), p(dy), p(dw), 0, st()))
    twp = t(lambda: nl.call("drn_conv_wgrad_pre", ctypes.byref(d), p(x), p(ss), p(dy), p(dw), 0, st()))
    print(f"                          wgrad {tw:6.1f} us | wgrad_pre {twp:6.1f} us", flush=True)
