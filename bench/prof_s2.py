"""Strided 3x3 data gradient: zero-insertion + fused BN-backward epilogue vs four parity phases +
interleaving mask/reduce kernel. L2 flushed."""
import ctypes, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_resnet_tensorflow_b200._native import lib as nl
from distributed_resnet_tensorflow_b200.ops.conv import ConvGeom
nl.load()
st = lambda: torch.cuda.current_stream().cuda_stream
p = lambda t: 0 if t is None else t.data_ptr()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def t(fn, iters=5):
    for _ in range(2): fn()
    best = 1e9
    for _ in range(iters):
        flush.zero_(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); best = min(best, e0.elapsed_time(e1) * 1e3)
    return best
N = 128
for H, C in ((56, 128), (28, 256), (14, 512)):
    K = C
    g = ConvGeom.tf_fixed_padding(N, H, H, C, K, 3, 2); d = g.desc()
    w = (torch.randn((K, 3, 3, C), device="cuda") / (9 * C) ** 0.5).bfloat16()
    dy = torch.randn((N, g.P, g.Q, K), device="cuda").bfloat16()
    y = torch.randn((N, H, H, C), device="cuda").bfloat16(); ss = torch.rand((C, 2), device="cuda")
    dm = torch.empty_like(y); stats = torch.zeros((2, C), device="cuda")
    wd = torch.empty(K * 9 * C, dtype=torch.bfloat16, device="cuda"); nl.call("drn_weight_prep_dgrad", p(w), p(wd), K, 3, 3, C, st())
    need = nl.load().drn_conv_dgrad_scratch_bytes(ctypes.byref(d)); scratch = torch.zeros(need // 2, dtype=torch.bfloat16, device="cuda")
    banks = torch.empty(9 * K * C, dtype=torch.bfloat16, device="cuda"); nl.call("drn_weight_prep_dgrad_s2", p(w), p(banks), K, C, st())
    ph = torch.empty((4, N, H // 2, H // 2, C), dtype=torch.bfloat16, device="cuda")
    t_old = t(lambda: nl.call("drn_conv_dgrad", ctypes.byref(d), p(dy), p(wd), p(dm), p(y), p(stats), C, p(ss), 2, p(scratch), need, st()))
    t_ph = t(lambda: nl.call("drn_conv_dgrad_s2", ctypes.byref(d), p(dy), p(banks), p(ph), st()))
    t_mr = t(lambda: nl.call("drn_bn_bwd_mask_reduce", p(ph), 0, 4, N, H, H, 0, 0, 0.0, p(y), p(ss), p(dm), p(stats), C, C, st()))
    print(f"H={H} C={C}: zero-insertion dgrad+BNBWD {t_old:6.1f} us | phases {t_ph:6.1f} + interleave/mask {t_mr:6.1f} = {t_ph + t_mr:6.1f} us", flush=True)
