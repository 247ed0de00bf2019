"""Runs ONE of the remaining hot kernels a few times on a ResNet-50 stage-1 sized tensor (for ncu captures
and quick timing).  usage: prof_misc.py pre_conv | bn_fwd | mask_reduce | maxpool | wgrad_pre"""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_resnet_tensorflow_b200._native import lib as nl
from distributed_resnet_tensorflow_b200.ops.conv import ConvGeom

nl.load()
what = sys.argv[1]
st = lambda: torch.cuda.current_stream().cuda_stream
p = lambda t: t.data_ptr() if t is not None else 0
N, H = 128, 56
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

if what in ("pre_conv", "wgrad_pre"):
    C, K = 256, 64
    g = ConvGeom.tf_fixed_padding(N, H, H, C, K, 1, 1); d = g.desc()
    x = torch.randn((N, H, H, C), device="cuda").bfloat16()
    w = (torch.randn((K, 1, 1, C), device="cuda") / C ** 0.5).bfloat16()
    ss = torch.rand((C, 2), device="cuda"); y = torch.empty((N, H, H, K), dtype=torch.bfloat16, device="cuda")
    sb = torch.zeros((2, K), device="cuda"); dy = torch.randn((N, H, H, K), device="cuda").bfloat16()
    dw = torch.zeros((K, 1, 1, C), device="cuda")
    if what == "pre_conv":
        run = lambda: nl.call("drn_conv_igemm_pre", ctypes.byref(d), p(x), p(ss), p(w), p(y), 0, p(sb), K, 0, 0, 0, st())
        nbytes = 2.0 * (x.numel() + y.numel())
    else:
        run = lambda: nl.call("drn_conv_wgrad_pre", ctypes.byref(d), p(x), p(ss), p(dy), p(dw), 0, st())
        nbytes = 2.0 * (x.numel() + dy.numel())
elif what == "bn_fwd":
    M, C = N * H * H, 64
    y = torch.randn((M, C), device="cuda").bfloat16(); out = torch.empty_like(y)
    stats = torch.rand((2, C), device="cuda") * M; gamma = torch.rand(C, device="cuda"); beta = torch.rand(C, device="cuda")
    mi = torch.rand((C, 2), device="cuda"); ss = torch.rand((C, 2), device="cuda")
    rm = torch.zeros(C, device="cuda"); rv = torch.ones(C, device="cuda")
    run = lambda: nl.call("drn_bn_finalize_apply", p(y), p(stats), C, M, p(gamma), p(beta), p(rm), p(rv), 0.997, 1e-5,
                          p(ss), p(mi), p(out), C, 1, st())
    nbytes = 4.0 * y.numel()
elif what == "mask_reduce":
    C = 256
    da = torch.randn((N, H, H, C), device="cuda").bfloat16(); t = torch.randn_like(da); y = torch.randn_like(da)
    dm = torch.empty_like(da); ss = torch.rand((C, 2), device="cuda") - 0.5; stats = torch.zeros((2, C), device="cuda")
    run = lambda: nl.call("drn_bn_bwd_mask_reduce", p(da), p(t), 1, N, H, H, 0, 0, 0.0, p(y), p(ss), p(dm), p(stats), C, C, st())
    nbytes = 8.0 * da.numel()
elif what == "maxpool":
    C, Hs = 64, 112
    x = torch.randn((N, Hs, Hs, C), device="cuda").bfloat16()
    out = torch.empty((N, 56, 56, C), dtype=torch.bfloat16, device="cuda"); arg = torch.empty((N, 56, 56, C), dtype=torch.uint8, device="cuda")
    run = lambda: nl.call("drn_maxpool_fwd", p(x), p(out), p(arg), N, Hs, Hs, C, 3, 2, 0, 56, 56, st())
    nbytes = 2.0 * (x.numel() + out.numel()) + arg.numel()
else:
    raise SystemExit(__doc__)

for _ in range(3):
    run()
torch.cuda.synchronize()
ts = []
for _ in range(5):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); run(); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1) * 1e3)
print(f"{what}: best {min(ts):.1f} us  {nbytes / min(ts) * 1e-3:.0f} GB/s (algorithmic bytes, L2 flushed)")
