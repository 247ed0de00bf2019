"""Randomised (fixed seed) sweep of the implicit-GEMM convolution over shapes, strides, epilogue modes
and launch variants, against PyTorch fp32 references - the shape-sweep of SURVEY.md section 4.  Every
case draws a ragged batch/size so that partial row tiles, single-tile grids and multi-round grids
are all exercised; outputs sit between guard bands."""
import ctypes

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _cases(n, seed):
    rng = np.random.default_rng(seed)
    out = []
    chans = [16, 32, 64, 128, 256, 512]
    while len(out) < n:
        C, K = int(rng.choice(chans)), int(rng.choice(chans))
        ks = int(rng.choice([1, 3]))
        st = int(rng.choice([1, 2]))
        H = int(rng.integers(4, 29))
        if st == 2 and H % 2:
            H += 1
        N = int(rng.integers(1, 10))
        mode = int(rng.choice([0, 1, 2]))
        two = int(rng.choice([0, 2])) if C >= 64 else 0
        out.append((N, H, C, K, ks, st, mode, two))
    return out


@pytest.mark.parametrize("N,H,C,K,ks,st,mode,two", _cases(28, 2024))
def test_fprop_random_shapes(N, H, C, K, ks, st, mode, two, monkeypatch):
    from distributed_resnet_tensorflow_b200.ops.conv import ConvGeom, conv_fprop
    monkeypatch.setenv("DRN_TWO_CTA", str(two))
    g = ConvGeom.tf_fixed_padding(N, H, H, C, K, ks, st)
    gen = torch.Generator(device="cuda").manual_seed(N * 1000 + H)
    x = torch.randn((N, H, H, C), device="cuda", generator=gen).bfloat16()
    w = (torch.randn((K, ks, ks, C), device="cuda", generator=gen) / (ks * ks * C) ** 0.5).bfloat16()
    side = torch.randn((N, g.P, g.Q, K), device="cuda", generator=gen).bfloat16() if mode else None
    ss = (torch.rand((K, 2), device="cuda", generator=gen) - 0.3) if mode == 2 else None
    n = N * g.P * g.Q * K
    raw = torch.full((n + 8192,), -777.0, dtype=torch.bfloat16, device="cuda")
    y = raw[4096:4096 + n].view(N, g.P, g.Q, K)
    stats = torch.zeros((2, K), device="cuda")
    conv_fprop(x, w, g, out=y, side=side, stats=stats, bn_scale_shift=ss, mode=mode)
    torch.cuda.synchronize()
    assert bool((raw[:4096] == -777.0).all()) and bool((raw[-4096:] == -777.0).all())
    xf = F.pad(x.float().permute(0, 3, 1, 2), (g.pad_lo, g.pad_hi, g.pad_lo, g.pad_hi))
    ref = F.conv2d(xf, w.float().permute(0, 3, 1, 2), stride=st).permute(0, 2, 3, 1)
    if mode == 1:
        ref = ref + side.float()
    elif mode == 2:
        ref = ref * ((side.float() * ss[:, 0] + ss[:, 1]) > 0)
    assert (y.float() - ref).abs().max() <= 0.02 * ref.abs().max() + 2e-3
    yq = y.float().reshape(-1, K)
    second = yq * side.float().reshape(-1, K) if mode == 2 else yq * yq
    assert torch.allclose(stats[0], yq.sum(0), rtol=2e-3, atol=0.3)
    assert torch.allclose(stats[1], second.sum(0), rtol=2e-3, atol=0.3)


@pytest.mark.parametrize("N,H,C,K,ks,st,mode,two", _cases(14, 7))
def test_wgrad_and_dgrad_random_shapes(N, H, C, K, ks, st, mode, two):
    from distributed_resnet_tensorflow_b200._native import lib as nl
    from distributed_resnet_tensorflow_b200.ops.conv import ConvGeom, conv_wgrad
    nl.load()
    g = ConvGeom.tf_fixed_padding(N, H, H, C, K, ks, st)
    gen = torch.Generator(device="cuda").manual_seed(N * 77 + H)
    x = torch.randn((N, H, H, C), device="cuda", generator=gen).bfloat16()
    w = (torch.randn((K, ks, ks, C), device="cuda", generator=gen) / (ks * ks * C) ** 0.5).bfloat16()
    dy = torch.randn((N, g.P, g.Q, K), device="cuda", generator=gen).bfloat16()
    xf = F.pad(x.float().permute(0, 3, 1, 2), (g.pad_lo, g.pad_hi, g.pad_lo, g.pad_hi))
    dw = conv_wgrad(x, dy, g)
    refw = torch.nn.grad.conv2d_weight(xf, (K, C, ks, ks), dy.float().permute(0, 3, 1, 2), stride=st).permute(0, 2, 3, 1)
    assert (dw - refw).abs().max() <= 3e-3 * refw.abs().max() + 2e-3
    st_ = torch.cuda.current_stream().cuda_stream
    d = g.desc()
    wd = torch.empty(K * ks * ks * C, dtype=torch.bfloat16, device="cuda")
    nl.call("drn_weight_prep_dgrad", w.data_ptr(), wd.data_ptr(), K, ks, ks, C, st_)
    need = nl.load().drn_conv_dgrad_scratch_bytes(ctypes.byref(d))
    scratch = torch.zeros(max(need // 2, 8), dtype=torch.bfloat16, device="cuda")
    dx = torch.empty((N, H, H, C), dtype=torch.bfloat16, device="cuda")
    nl.call("drn_conv_dgrad", ctypes.byref(d), dy.data_ptr(), wd.data_ptr(), dx.data_ptr(), 0, 0, 0, 0, 0,
            scratch.data_ptr(), need, st_)
    torch.cuda.synchronize()
    refx = torch.nn.grad.conv2d_input(xf.shape, w.float().permute(0, 3, 1, 2), dy.float().permute(0, 3, 1, 2), stride=st)
    refx = refx[:, :, g.pad_lo:g.pad_lo + H, g.pad_lo:g.pad_lo + H].permute(0, 2, 3, 1)
    assert (dx.float() - refx).abs().max() <= 0.02 * refx.abs().max() + 2e-3
