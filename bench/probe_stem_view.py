"""Space-to-depth stem as a 4x1 convolution over an overlapping 64-channel view (see
drn_conv_igemm_view): correctness vs torch and timing vs the plain 4x4 / C=16 formulation."""
import ctypes, os, sys
import torch
import torch.nn.functional as F
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_resnet_tensorflow_b200._native import lib as nl
from distributed_resnet_tensorflow_b200._native.lib import ConvView
from distributed_resnet_tensorflow_b200.ops.conv import ConvGeom, conv_fprop, conv_wgrad


def run(B, H, K=64, time_it=False):
    WP = H + 4
    gen = torch.Generator(device="cuda").manual_seed(0)
    xpad = torch.zeros((B, H, WP, 16), dtype=torch.bfloat16, device="cuda")
    xpad[:, :, 2:2 + H, :] = torch.randn((B, H, H, 16), device="cuda", generator=gen).bfloat16()
    w = (torch.randn((K, 4, 4, 16), device="cuda", generator=gen) / 16.0).bfloat16()
    dy = torch.randn((B, H, H, K), device="cuda", generator=gen).bfloat16()
    g_old = ConvGeom(B, H, H, 16, K, 4, 4, 1, 2, 1)
    x_dense = xpad[:, :, 2:2 + H, :].contiguous()
    # view: [B, H, W'=H, C'=64], W stride 16 elements, rows WP*16, images H*WP*16; 4x1 filter
    d = ConvGeom(B, H, H, 64, K, 4, 1, 1, 2, 1).desc()
    d.Q = H
    v = ConvView(16, WP * 16, H * WP * 16, 0, 0)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    y = torch.empty((B, H, H, K), dtype=torch.bfloat16, device="cuda")
    stats = torch.zeros((2, K), device="cuda")
    nl.call("drn_conv_igemm_view", ctypes.byref(d), ctypes.byref(v), xpad.data_ptr(), w.data_ptr(), y.data_ptr(),
            None, stats.data_ptr(), K, None, 0, 0, st)
    dw = torch.zeros((K, 4, 4, 16), device="cuda")
    nl.call("drn_conv_wgrad_view", ctypes.byref(d), ctypes.byref(v), xpad.data_ptr(), dy.data_ptr(), dw.data_ptr(), 0, st)
    torch.cuda.synchronize()
    xf = F.pad(x_dense.float().permute(0, 3, 1, 2), (2, 1, 2, 1))
    ref = F.conv2d(xf, w.float().permute(0, 3, 1, 2)).permute(0, 2, 3, 1)
    refw = torch.nn.grad.conv2d_weight(xf, (K, 16, 4, 4), dy.float().permute(0, 3, 1, 2)).permute(0, 2, 3, 1)
    e1 = (y.float() - ref).abs().max().item() / ref.abs().max().item()
    e2 = (dw - refw).abs().max().item() / refw.abs().max().item()
    print(f"B={B} H={H}: fprop relerr {e1:.3e}  wgrad relerr {e2:.3e}", "OK" if e1 < 0.02 and e2 < 3e-3 else "FAIL", flush=True)
    if time_it:
        def t(fn):
            for _ in range(3): fn()
            torch.cuda.synchronize()
            e0, e1_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10): fn()
            e1_.record(); torch.cuda.synchronize()
            return e0.elapsed_time(e1_) * 100
        print("  view  fprop %.1f us  wgrad %.1f us" % (
            t(lambda: nl.call("drn_conv_igemm_view", ctypes.byref(d), ctypes.byref(v), xpad.data_ptr(), w.data_ptr(), y.data_ptr(), None, stats.data_ptr(), K, None, 0, 0, st)),
            t(lambda: nl.call("drn_conv_wgrad_view", ctypes.byref(d), ctypes.byref(v), xpad.data_ptr(), dy.data_ptr(), dw.data_ptr(), 0, st))))
        print("  plain fprop %.1f us  wgrad %.1f us" % (
            t(lambda: conv_fprop(x_dense, w, g_old, out=y, stats=stats)), t(lambda: conv_wgrad(x_dense, dy, g_old, dw=dw))), flush=True)


run(2, 16)
run(3, 12)
run(128, 112, time_it=True)
