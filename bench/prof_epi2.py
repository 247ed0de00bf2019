"""Kernel-level A/B of the BatchNorm-backward tail of a bottleneck block (L2 flushed):
  old: dgrad(conv1) with mask+reductions epilogue -> element-wise bn_bwd_finalize_apply (+ shortcut gradient)
  new: dgrad statistics-only -> bn_bwd_finalize -> dgrad with the BNBWD_APPLY epilogue."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_resnet_tensorflow_b200._native import lib as nl
from distributed_resnet_tensorflow_b200._native.lib import EPI_BNBWD, EPI_BNBWD_APPLY
from distributed_resnet_tensorflow_b200.ops.conv import ConvGeom, conv_fprop, conv_fprop_epi2
nl.load()
st = lambda: torch.cuda.current_stream().cuda_stream
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def t(fn, n=5):
    for _ in range(2): fn()
    best = 1e9
    for _ in range(n):
        flush.zero_(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); best = min(best, e0.elapsed_time(e1) * 1e3)
    return best
p = lambda x: x.data_ptr()
for (H, Cred, Cout) in ((56, 64, 256), (28, 128, 512), (14, 256, 1024), (7, 512, 2048)):
    N = 128
    g = ConvGeom(N, H, H, Cred, Cout, 1, 1, 1, 0, 0)
    dy = torch.randn((N, H, H, Cred), device="cuda").bfloat16(); w = (torch.randn((Cout, 1, 1, Cred), device="cuda") / Cred ** 0.5).bfloat16()
    x = torch.randn((N, H, H, Cout), device="cuda").bfloat16(); gsc = torch.randn_like(x); out = torch.empty_like(x); dm = torch.empty_like(x)
    ss = torch.rand((Cout, 2), device="cuda"); coef = torch.rand((Cout, 4), device="cuda"); stats = torch.zeros((2, Cout), device="cuda")
    gamma = torch.rand(Cout, device="cuda"); mi = torch.rand((Cout, 2), device="cuda"); dg = torch.zeros(Cout, device="cuda"); db = torch.zeros(Cout, device="cuda")
    M = N * H * H
    a1 = t(lambda: conv_fprop(dy, w, g, out=dm, side=x, stats=stats, bn_scale_shift=ss, mode=EPI_BNBWD))
    a2 = t(lambda: nl.call("drn_bn_bwd_finalize_apply", p(dm), p(x), p(stats), Cout, M, p(gamma), p(mi), p(dg), p(db), p(gsc), p(out), Cout, st()))
    res = [f"H={H:3d} {Cred}->{Cout}: old {a1:6.1f} + {a2:6.1f} = {a1 + a2:6.1f} us |"]
    for bn in (0, 128, 256):
        for nres in ("2", "3", "4"):
            os.environ["DRN_NRES2"] = nres
            try:
                b1 = t(lambda: conv_fprop_epi2(dy, w, g, x, None, stats=stats, bn_scale_shift=ss, mode=EPI_BNBWD, store=False, force_bn=bn))
                b2 = t(lambda: conv_fprop_epi2(dy, w, g, x, gsc, out=out, bn_scale_shift=ss, bn_coef=coef, mode=EPI_BNBWD_APPLY, force_bn=bn))
                res.append(f" bn{bn}/r{nres}: {b1:5.1f}+{b2:5.1f}={b1 + b2:6.1f}")
            except RuntimeError as e:
                res.append(f" bn{bn}/r{nres}: n/a")
    print("".join(res), flush=True)
