"""Runs ONE convolution configuration a few times (for `ncu --set full` captures and quick timing).
usage: prof_one.py fprop|wgrad N H C K ksize stride mode(0|1|2) stats(0|1) [iters]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_resnet_tensorflow_b200.ops.conv import ConvGeom, conv_fprop, conv_wgrad

kind = sys.argv[1]
N, H, C, K, ks, st, mode, stats = map(int, sys.argv[2:10])
iters = int(sys.argv[10]) if len(sys.argv) > 10 else 5
force_bn = int(sys.argv[11]) if len(sys.argv) > 11 else 0
lo = (ks - 1) // 2
g = ConvGeom(N, H, H, C, K, ks, ks, st, lo, ks - 1 - lo)
gen = torch.Generator(device="cuda").manual_seed(0)
x = torch.randn((N, H, H, C), device="cuda", generator=gen).bfloat16()
w = (torch.randn((K, ks, ks, C), device="cuda", generator=gen) / (ks * ks * C) ** 0.5).bfloat16()
side = torch.randn((N, g.P, g.Q, K), device="cuda", generator=gen).bfloat16() if mode else None
ss = torch.rand((K, 2), device="cuda", generator=gen) if mode == 2 else None
sb = torch.zeros((2, K), device="cuda") if stats else None
y = torch.empty((N, g.P, g.Q, K), dtype=torch.bfloat16, device="cuda")
dy = torch.randn((N, g.P, g.Q, K), device="cuda", generator=gen).bfloat16()
dw = torch.zeros((K, ks, ks, C), device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

def run():
    if kind == "fprop":
        conv_fprop(x, w, g, out=y, side=side, stats=sb, bn_scale_shift=ss, mode=mode, force_bn=force_bn)
    else:
        conv_wgrad(x, dy, g, dw=dw)

for _ in range(3):
    run()
torch.cuda.synchronize()
ts = []
for _ in range(iters):
    flush.zero_()  # evict L2 between timed iterations
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); run(); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1) * 1e3)
flops = 2.0 * N * g.P * g.Q * K * C * ks * ks
byts = 2.0 * (x.numel() + w.numel() + y.numel() * (2 if side is not None else 1)) if kind == "fprop" else 2.0 * (x.numel() + dy.numel()) + 4.0 * dw.numel()
best = min(ts)
print(f"{kind} N{N} H{H} C{C} K{K} k{ks} s{st} mode{mode} stats{stats} bn{force_bn}: best {best:.1f} us  median {sorted(ts)[len(ts)//2]:.1f} us  "
      f"{flops / best * 1e-6:.1f} TFLOP/s  {byts / best * 1e-3:.0f} GB/s (algorithmic bytes, L2 flushed)")
