"""CTA-pair (cta_group::2) variant of the implicit-GEMM kernel vs the single-CTA kernel: outputs,
fused statistics, timing.  DRN_TWO_CTA is read at every launch."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_resnet_tensorflow_b200.ops.conv import ConvGeom, conv_fprop
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def run(g, mode, stats, two, bn, time_it=False):
    os.environ["DRN_TWO_CTA"] = "2" if two else "0"
    gen = torch.Generator(device="cuda").manual_seed(0)
    x = torch.randn((g.N, g.H, g.W, g.C), device="cuda", generator=gen).bfloat16()
    w = (torch.randn((g.K, g.R, g.S, g.C), device="cuda", generator=gen) / (g.R * g.S * g.C) ** 0.5).bfloat16()
    side = torch.randn((g.N, g.P, g.Q, g.K), device="cuda", generator=gen).bfloat16() if mode else None
    ss = torch.rand((g.K, 2), device="cuda", generator=gen) - 0.3 if mode == 2 else None
    sb = torch.zeros((2, g.K), device="cuda") if stats else None
    y = torch.full((g.N, g.P, g.Q, g.K), float("nan"), dtype=torch.bfloat16, device="cuda")
    f = lambda: conv_fprop(x, w, g, out=y, side=side, stats=sb, bn_scale_shift=ss, mode=mode, force_bn=bn)
    f(); torch.cuda.synchronize()
    best = None
    if time_it:
        best = 1e9
        for _ in range(5):
            flush.zero_(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); f(); e1.record(); torch.cuda.synchronize(); best = min(best, e0.elapsed_time(e1) * 1e3)
        sb = None
    return y.float(), (sb.clone() if sb is not None else None), best
ok = True
for (N, H, C, K, ks, mode, stats, bn) in [(8, 14, 256, 256, 1, 0, 0, 128), (8, 14, 256, 256, 3, 0, 1, 128), (5, 7, 512, 512, 3, 1, 1, 256),
                                          (3, 14, 1024, 256, 1, 2, 1, 128), (7, 7, 512, 2048, 1, 1, 1, 128), (8, 28, 128, 128, 3, 2, 1, 128)]:
    lo = (ks - 1) // 2
    g = ConvGeom(N, H, H, C, K, ks, ks, 1, lo, ks - 1 - lo)
    y1, s1, _ = run(g, mode, stats, False, bn)
    y2, s2, _ = run(g, mode, stats, True, bn)
    err = (y1 - y2).abs().max().item(); nan = bool(torch.isnan(y2).any())
    serr = (s1 - s2).abs().max().item() / (s1.abs().max().item() + 1e-6) if s1 is not None else 0.0
    good = (not nan) and err <= 0.02 * y1.abs().max().item() + 1e-3 and serr < 2e-3
    ok &= good
    print(f"N{N} H{H} C{C} K{K} k{ks} mode{mode} stats{stats} bn{bn}: max|diff| {err:.4f} nan={nan} stats rel {serr:.2e} {'OK' if good else 'FAIL'}", flush=True)
print("CHECK_TWO", "OK" if ok else "FAIL", flush=True)
if ok and len(sys.argv) > 1:
    for (H, C, K, ks, mode, stats, bn) in [(14, 256, 256, 3, 0, 1, 128), (14, 256, 256, 3, 2, 1, 128), (7, 512, 512, 3, 0, 1, 256), (7, 512, 512, 3, 2, 1, 256),
                                           (28, 128, 128, 3, 0, 1, 128), (14, 1024, 256, 1, 0, 1, 128), (14, 256, 1024, 1, 1, 1, 128), (7, 2048, 512, 1, 0, 1, 128),
                                           (7, 512, 2048, 1, 1, 1, 128), (28, 512, 128, 1, 0, 1, 128), (14, 1024, 256, 1, 2, 1, 128), (7, 2048, 512, 1, 2, 1, 256)]:
        lo = (ks - 1) // 2
        g = ConvGeom(128, H, H, C, K, ks, ks, 1, lo, ks - 1 - lo)
        t1 = run(g, mode, stats, False, bn, True)[2]; t2 = run(g, mode, stats, True, bn, True)[2]
        print(f"H{H} C{C} K{K} k{ks} mode{mode} bn{bn}: 1-CTA {t1:6.1f} us  CTA-pair {t2:6.1f} us", flush=True)
