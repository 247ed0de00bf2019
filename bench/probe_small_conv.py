"""Fixed cost of the implicit-GEMM kernel on CIFAR-sized problems: N launches of one configuration inside
a CUDA graph (no launch overhead between them), microseconds per launch."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_resnet_tensorflow_b200.ops.conv import ConvGeom, conv_fprop, conv_wgrad

def run(N, H, C, K, ks, stats, reps=200):
    lo = (ks - 1) // 2
    g = ConvGeom(N, H, H, C, K, ks, ks, 1, lo, ks - 1 - lo)
    x = torch.randn((N, H, H, C), device="cuda").bfloat16()
    w = (torch.randn((K, ks, ks, C), device="cuda") / (ks * ks * C) ** 0.5).bfloat16()
    y = torch.empty((N, g.P, g.Q, K), dtype=torch.bfloat16, device="cuda")
    sb = torch.zeros((2, K), device="cuda") if stats else None
    dy = torch.randn_like(y); dw = torch.zeros((K, ks, ks, C), device="cuda")
    out = {}
    for name, fn in (("fprop", lambda: conv_fprop(x, w, g, out=y, stats=sb)), ("wgrad", lambda: conv_wgrad(x, dy, g, dw=dw))):
        fn(); torch.cuda.synchronize()
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=s):
                for _ in range(reps): fn()
        torch.cuda.synchronize()
        gr.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); gr.replay(); e1.record(); torch.cuda.synchronize()
        out[name] = e0.elapsed_time(e1) * 1e3 / reps
    print(f"N={N:3d} H={H:2d} C={C:3d} K={K:3d} k={ks} stats={stats}: fprop {out['fprop']:6.2f} us  wgrad {out['wgrad']:6.2f} us", flush=True)

for N in (16, 128):
    run(N, 32, 16, 16, 1, 0); run(N, 32, 16, 16, 3, 0); run(N, 32, 16, 16, 3, 1); run(N, 32, 64, 64, 3, 1)
    run(N, 16, 32, 32, 3, 1); run(N, 8, 64, 64, 3, 1); run(N, 8, 64, 64, 1, 1)
