"""Per-shape sweep of the implicit-GEMM N tile (force_bn) and side-ring depth (DRN_NRES) over the
convolution roles of ResNet-50 (fprop and dgrad-as-fprop), L2 flushed between timed iterations.
Prints one line per (shape, role) with every configuration and the best one."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_resnet_tensorflow_b200.ops.conv import ConvGeom, conv_fprop

N = 128
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def time_cfg(g, mode, stats, bn, nres, iters=5):
    if nres is None:
        os.environ.pop("DRN_NRES", None)
    else:
        os.environ["DRN_NRES"] = str(nres)
    gen = torch.Generator(device="cuda").manual_seed(0)
    x = torch.randn((g.N, g.H, g.W, g.C), device="cuda", generator=gen).bfloat16()
    w = (torch.randn((g.K, g.R, g.S, g.C), device="cuda", generator=gen) / (g.R * g.S * g.C) ** 0.5).bfloat16()
    side = torch.randn((g.N, g.P, g.Q, g.K), device="cuda", generator=gen).bfloat16() if mode else None
    ss = torch.rand((g.K, 2), device="cuda", generator=gen) if mode == 2 else None
    sb = torch.zeros((2, g.K), device="cuda") if stats else None
    y = torch.empty((g.N, g.P, g.Q, g.K), dtype=torch.bfloat16, device="cuda")
    run = lambda: conv_fprop(x, w, g, out=y, side=side, stats=sb, bn_scale_shift=ss, mode=mode, force_bn=bn)
    try:
        for _ in range(2):
            run()
        torch.cuda.synchronize()
    except Exception as e:  # configuration does not fit
        return None
    best = 1e9
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); run(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e3)
    return best


# (name, H, C, K, ksize, mode, stats)
ROLES = []
for H, cn in ((56, 64), (28, 128), (14, 256), (7, 512)):
    cw = cn * 4
    ROLES += [(f"conv3 fwd  {H}", H, cn, cw, 1, 1, 1), (f"conv1 dgrad {H}", H, cn, cw, 1, 2, 1),
              (f"conv3 dgrad {H}", H, cw, cn, 1, 2, 1), (f"conv1 fwd  {H}", H, cw, cn, 1, 0, 1),
              (f"conv2 fwd  {H}", H, cn, cn, 3, 0, 1), (f"conv2 dgrad {H}", H, cn, cn, 3, 2, 1)]

if __name__ == "__main__":
    for name, H, C, K, ks, mode, stats in ROLES:
        lo = (ks - 1) // 2
        g = ConvGeom(N, H, H, C, K, ks, ks, 1, lo, ks - 1 - lo)
        res = {}
        for bn in (64, 128, 256):
            if K % bn:
                continue
            for nres in ((2, 4) if mode else (2,)):
                t = time_cfg(g, mode, stats, bn, nres)
                if t is not None:
                    res[(bn, nres)] = t
        best = min(res, key=res.get)
        auto = time_cfg(g, mode, stats, 0, None)
        cells = " ".join(f"bn{b}/r{r}={t:6.1f}" for (b, r), t in sorted(res.items()))
        print(f"{name:16s} C={C:4d} K={K:4d} k{ks} mode{mode}: {cells}  -> best bn{best[0]}/r{best[1]} {res[best]:.1f} us | policy {auto:.1f} us", flush=True)
