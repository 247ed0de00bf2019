"""Fused / batched / view entry points against the plain entry points they replace, and guard-band
checks for out-of-bounds writes (compute-sanitizer is not available on the GPU pool, so the tests
place every output between two sentinel regions instead)."""
import ctypes

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _nl():
    from distributed_resnet_tensorflow_b200._native import lib
    lib.load()
    return lib


def _st():
    return torch.cuda.current_stream().cuda_stream


def _p(t):
    return 0 if t is None else t.data_ptr()


class Guarded:
    """A tensor carved out of the middle of a larger allocation whose borders hold a sentinel."""

    def __init__(self, shape, dtype, guard_elems=4096):
        n = int(np.prod(shape))
        self.raw = torch.empty(n + 2 * guard_elems, dtype=dtype, device="cuda")
        self.sentinel = -1234.0 if dtype.is_floating_point else 0x5A
        self.raw.fill_(self.sentinel)
        self.g = guard_elems
        self.t = self.raw[guard_elems:guard_elems + n].view(shape)

    def intact(self):
        lo, hi = self.raw[:self.g], self.raw[-self.g:]
        return bool((lo == self.sentinel).all()) and bool((hi == self.sentinel).all())


@pytest.mark.parametrize("N,H,C,K,ks,st", [(5, 7, 512, 512, 3, 1), (3, 14, 256, 64, 1, 1),
                                           (3, 28, 64, 64, 3, 2), (7, 8, 64, 64, 3, 1)])
def test_conv_outputs_stay_inside_their_buffers(N, H, C, K, ks, st):
    """Row counts that are not multiples of the 128-row tile: TMA stores must clip, statistics and
    split-K reductions must not touch memory outside their arrays."""
    from distributed_resnet_tensorflow_b200._native.lib import ConvDesc
    from distributed_resnet_tensorflow_b200.ops.conv import ConvGeom
    nl = _nl()
    g = ConvGeom.tf_fixed_padding(N, H, H, C, K, ks, st)
    gen = torch.Generator(device="cuda").manual_seed(0)
    x = torch.randn((N, H, H, C), device="cuda", generator=gen).bfloat16()
    w = (torch.randn((K, ks, ks, C), device="cuda", generator=gen) / (ks * ks * C) ** 0.5).bfloat16()
    y = Guarded((N, g.P, g.Q, K), torch.bfloat16)
    stats = Guarded((2, K), torch.float32)
    stats.t.zero_()
    d = g.desc()
    nl.call("drn_conv_igemm", ctypes.byref(d), _p(x), _p(w), _p(y.t), 0, _p(stats.t), K, 0, 0, 0, _st())
    torch.cuda.synchronize()
    assert y.intact() and stats.intact()
    xf = F.pad(x.float().permute(0, 3, 1, 2), (g.pad_lo, g.pad_hi, g.pad_lo, g.pad_hi))
    ref = F.conv2d(xf, w.float().permute(0, 3, 1, 2), stride=st).permute(0, 2, 3, 1)
    assert (y.t.float() - ref).abs().max() <= 0.02 * ref.abs().max() + 1e-3
    s_ref = y.t.float().reshape(-1, K).sum(0)
    assert (stats.t[0] - s_ref).abs().max() <= 2e-2 * s_ref.abs().max() + 0.5
    dy = torch.randn((N, g.P, g.Q, K), device="cuda", generator=gen).bfloat16()
    dw = Guarded((K, ks, ks, C), torch.float32)
    dw.t.zero_()
    nl.call("drn_conv_wgrad", ctypes.byref(d), _p(x), _p(dy), _p(dw.t), 0, _st())
    torch.cuda.synchronize()
    assert dw.intact()
    refw = torch.nn.grad.conv2d_weight(xf, (K, C, ks, ks), dy.float().permute(0, 3, 1, 2), stride=st)
    refw = refw.permute(0, 2, 3, 1)
    assert (dw.t - refw).abs().max() <= 2e-3 * refw.abs().max() + 1e-3
    # data gradient through the first-class entry point, strided case included
    wd = torch.empty(K * ks * ks * C, dtype=torch.bfloat16, device="cuda")
    nl.call("drn_weight_prep_dgrad", _p(w), _p(wd), K, ks, ks, C, _st())
    need = nl.load().drn_conv_dgrad_scratch_bytes(ctypes.byref(d))
    scratch = Guarded((max(need // 2, 8),), torch.bfloat16)
    dx = Guarded((N, H, H, C), torch.bfloat16)
    nl.call("drn_conv_dgrad", ctypes.byref(d), _p(dy), _p(wd), _p(dx.t), 0, 0, 0, 0, 0,
            _p(scratch.t), need, _st())
    torch.cuda.synchronize()
    assert dx.intact() and scratch.intact()
    refx = torch.nn.grad.conv2d_input(xf.shape, w.float().permute(0, 3, 1, 2),
                                      dy.float().permute(0, 3, 1, 2), stride=st)
    refx = refx[:, :, g.pad_lo:g.pad_lo + H, g.pad_lo:g.pad_lo + H].permute(0, 2, 3, 1)
    assert (dx.t.float() - refx).abs().max() <= 0.02 * refx.abs().max() + 1e-3


@pytest.mark.parametrize("M,C,relu", [(1000, 64, 1), (25088, 256, 1), (333, 2048, 0)])
def test_bn_finalize_apply_equals_the_two_step_path(M, C, relu):
    nl = _nl()
    gen = torch.Generator(device="cuda").manual_seed(2)
    y = (torch.randn((M, C), device="cuda", generator=gen) * 2 + 0.5).bfloat16()
    gamma = torch.rand(C, device="cuda", generator=gen) + 0.5
    beta = torch.randn(C, device="cuda", generator=gen) * 0.1
    stats = torch.zeros((2, C), device="cuda")
    nl.call("drn_bn_stats", _p(y), _p(stats), C, M, C, _st())
    outs = []
    for fused in (False, True):
        rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
        ss = torch.empty((C, 2), device="cuda")
        mi = torch.empty((C, 2), device="cuda")
        out = Guarded((M, C), torch.bfloat16)
        if fused:
            nl.call("drn_bn_finalize_apply", _p(y), _p(stats), C, M, _p(gamma), _p(beta), _p(rm), _p(rv),
                    0.997, 1e-5, _p(ss), _p(mi), _p(out.t), C, relu, _st())
        else:
            nl.call("drn_bn_finalize", _p(stats), C, M, _p(gamma), _p(beta), _p(rm), _p(rv), 0.997, 1e-5,
                    _p(ss), _p(mi), C, _st())
            nl.call("drn_bn_apply", _p(y), _p(ss), 0, _p(out.t), M, C, relu, _st())
        torch.cuda.synchronize()
        assert out.intact()
        outs.append((out.t.clone(), ss, mi, rm, rv))
    a, b = outs
    assert torch.equal(a[0], b[0])
    for i in range(1, 5):
        assert torch.allclose(a[i], b[i], rtol=1e-6, atol=1e-7)
    # and against first principles
    yf = y.float()
    mean, var = yf.mean(0), yf.var(0, unbiased=False)
    ref = (yf - mean) * torch.rsqrt(var + 1e-5) * gamma + beta
    if relu:
        ref = ref.relu()
    assert (b[0].float() - ref).abs().max() < 0.05
    assert torch.allclose(b[3], mean * 0.003, atol=1e-4)


@pytest.mark.parametrize("M,C,with_addend", [(1000, 64, False), (6272, 512, True)])
def test_bn_bwd_finalize_apply_equals_the_two_step_path(M, C, with_addend):
    nl = _nl()
    gen = torch.Generator(device="cuda").manual_seed(3)
    y = torch.randn((M, C), device="cuda", generator=gen).bfloat16()
    dm = torch.randn((M, C), device="cuda", generator=gen).bfloat16()
    addend = torch.randn((M, C), device="cuda", generator=gen).bfloat16() if with_addend else None
    gamma = torch.rand(C, device="cuda", generator=gen) + 0.5
    yf, df = y.float(), dm.float()
    mean, var = yf.mean(0), yf.var(0, unbiased=False)
    invstd = torch.rsqrt(var + 1e-5)
    mi = torch.stack([mean, invstd], 1).contiguous()
    stats = torch.stack([df.sum(0), (df * yf).sum(0)]).contiguous()
    res = []
    for fused in (False, True):
        dgamma, dbeta = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
        dy = Guarded((M, C), torch.bfloat16)
        if fused:
            nl.call("drn_bn_bwd_finalize_apply", _p(dm), _p(y), _p(stats), C, M, _p(gamma), _p(mi),
                    _p(dgamma), _p(dbeta), _p(addend), _p(dy.t), C, _st())
        else:
            coef = torch.empty((C, 4), device="cuda")
            nl.call("drn_bn_bwd_finalize", _p(stats), C, M, _p(gamma), _p(mi), _p(dgamma), _p(dbeta),
                    _p(coef), C, _st())
            nl.call("drn_bn_bwd_apply", _p(dm), _p(y), _p(coef), _p(addend), _p(dy.t), M, C, _st())
        torch.cuda.synchronize()
        assert dy.intact()
        res.append((dy.t.clone(), dgamma, dbeta))
    assert torch.equal(res[0][0], res[1][0])
    assert torch.allclose(res[0][1], res[1][1], rtol=1e-5, atol=1e-5)
    assert torch.allclose(res[0][2], res[1][2], rtol=1e-5, atol=1e-5)
    # first principles: gradient of BN(y)*gamma w.r.t. y given upstream dm
    xhat = (yf - mean) * invstd
    ref = gamma * invstd * (df - df.mean(0) - xhat * (df * xhat).mean(0))
    if with_addend:
        ref = ref + addend.float()
    assert (res[1][0].float() - ref).abs().max() < 0.06
    assert torch.allclose(res[1][1], (df * xhat).sum(0), rtol=2e-3, atol=0.05)


def test_batched_dgrad_bank_prep_equals_per_layer_prep():
    nl = _nl()
    gen = torch.Generator(device="cuda").manual_seed(4)
    layers = [(64, 1, 64), (64, 9, 64), (256, 1, 64), (16, 9, 16), (128, 9, 128), (1024, 1, 512), (80, 49, 16)]
    total = sum(k * rs * c for k, rs, c in layers)
    arena = torch.randn(total, device="cuda", generator=gen).bfloat16()
    want = torch.empty_like(arena)
    got = Guarded((total,), torch.bfloat16)
    rows, off, tiles = [], 0, 0
    for k, rs, c in layers:
        r = int(round(rs ** 0.5))
        nl.call("drn_weight_prep_dgrad", arena.data_ptr() + 2 * off, want.data_ptr() + 2 * off, k, r, r, c, _st())
        rows.append((off, off, k, rs, c, tiles))
        tiles += rs * ((k + 63) // 64) * ((c + 63) // 64)
        off += k * rs * c
    table = np.zeros(len(rows), dtype=np.dtype([("s", "<i8"), ("d", "<i8"), ("k", "<i4"), ("rs", "<i4"),
                                                ("c", "<i4"), ("t", "<i4")]))
    for i, r in enumerate(rows):
        table[i] = r
    assert table.dtype.itemsize == 32
    dev_table = torch.from_numpy(table.view(np.uint8).copy()).cuda()
    nl.call("drn_weight_prep_dgrad_batched", _p(arena), _p(got.t), _p(dev_table), len(rows), tiles, _st())
    torch.cuda.synchronize()
    assert got.intact()
    assert torch.equal(got.t, want)


def test_view_entry_points_equal_the_dense_4x4_convolution():
    """The ImageNet stem trick: a [B,H,W+4,16] tensor with two zero pixels on each side of every row,
    viewed with a W stride of 16 elements as [B,H,W,64], turns the 4x4 filter over 16 channels into a
    4x1 filter over 64 'channels'.  fprop and wgrad over the view must equal the dense 4x4 problem."""
    from distributed_resnet_tensorflow_b200._native import lib as L
    from distributed_resnet_tensorflow_b200.ops.conv import ConvGeom
    nl = _nl()
    B, H, K = 3, 24, 64
    pitch = H + 4
    gen = torch.Generator(device="cuda").manual_seed(5)
    x0 = torch.zeros((B, H, pitch, 16), dtype=torch.bfloat16, device="cuda")
    x0[:, :, 2:2 + H, :] = torch.randn((B, H, H, 16), device="cuda", generator=gen).bfloat16()
    w = (torch.randn((K, 4, 4, 16), device="cuda", generator=gen) / 16.0).bfloat16()
    # dense oracle: 4x4 stride 1, pad (2 before, 1 after) in both directions
    xf = F.pad(x0[:, :, 2:2 + H, :].float().permute(0, 3, 1, 2), (2, 1, 2, 1))
    ref = F.conv2d(xf, w.float().permute(0, 3, 1, 2)).permute(0, 2, 3, 1)
    g = ConvGeom(B, H, H, 64, K, 4, 1, 1, 2, 1)
    d = g.desc()
    d.Q = H
    view = L.ConvView(16, pitch * 16, H * pitch * 16, 0, 0)
    y = Guarded((B, H, H, K), torch.bfloat16)
    stats = torch.zeros((2, K), device="cuda")
    # the [K][4][1][64] filter bank of the view is the same memory as [K][4][4][16]
    nl.call("drn_conv_igemm_view", ctypes.byref(d), ctypes.byref(view), _p(x0), _p(w), _p(y.t), 0,
            _p(stats), K, 0, 0, 0, _st())
    torch.cuda.synchronize()
    assert y.intact()
    assert (y.t.float() - ref).abs().max() <= 0.02 * ref.abs().max() + 1e-3
    dy = torch.randn((B, H, H, K), device="cuda", generator=gen).bfloat16()
    dw = Guarded((K, 4, 4, 16), torch.float32)
    dw.t.zero_()
    nl.call("drn_conv_wgrad_view", ctypes.byref(d), ctypes.byref(view), _p(x0), _p(dy), _p(dw.t), 0, _st())
    torch.cuda.synchronize()
    assert dw.intact()
    refw = torch.nn.grad.conv2d_weight(xf, (K, 16, 4, 4), dy.float().permute(0, 3, 1, 2)).permute(0, 2, 3, 1)
    assert (dw.t - refw).abs().max() <= 2e-3 * refw.abs().max() + 1e-3


PRE_CASES = [(8, 14, 256, 128, 1, 1), (5, 7, 512, 64, 1, 1), (4, 28, 256, 512, 1, 2), (3, 56, 64, 256, 1, 1),
         # filters larger than 1x1 (taps in the zero padding must stay zero) and the 16 / 32-channel tiles
         (4, 32, 16, 16, 3, 1), (4, 32, 16, 32, 3, 2), (4, 32, 16, 32, 1, 2), (4, 16, 32, 32, 3, 1),
         (4, 16, 32, 64, 3, 2), (5, 8, 64, 64, 3, 1), (3, 14, 256, 256, 3, 1), (3, 28, 128, 128, 3, 2),
         (2, 7, 512, 512, 3, 1)]


@pytest.mark.parametrize("N,H,C,K,ks,st", PRE_CASES)
def test_conv_with_batchnorm_prologue_equals_apply_then_conv(N, H, C, K, ks, st):
    """drn_conv_igemm_pre (BN+ReLU applied to the A tile in shared memory) against drn_bn_apply
    followed by the plain convolution: identical bf16 inputs to the tensor core, so the outputs and
    the fused statistics must agree to rounding of the accumulation order."""
    from distributed_resnet_tensorflow_b200.ops.conv import ConvGeom
    nl = _nl()
    g = ConvGeom.tf_fixed_padding(N, H, H, C, K, ks, st)
    gen = torch.Generator(device="cuda").manual_seed(6)
    x = (torch.randn((N, H, H, C), device="cuda", generator=gen) * 1.5).bfloat16()
    w = (torch.randn((K, ks, ks, C), device="cuda", generator=gen) / (ks * ks * C) ** 0.5).bfloat16()
    ss = torch.stack([torch.rand(C, device="cuda", generator=gen) + 0.5,
                      torch.randn(C, device="cuda", generator=gen) * 0.3], 1).contiguous()
    a = torch.empty_like(x)
    nl.call("drn_bn_apply", _p(x), _p(ss), 0, _p(a), N * H * H, C, 1, _st())
    d = g.desc()
    want = torch.empty((N, g.P, g.Q, K), dtype=torch.bfloat16, device="cuda")
    st_want = torch.zeros((2, K), device="cuda")
    nl.call("drn_conv_igemm", ctypes.byref(d), _p(a), _p(w), _p(want), 0, _p(st_want), K, 0, 0, 0, _st())
    got = Guarded((N, g.P, g.Q, K), torch.bfloat16)
    st_got = torch.zeros((2, K), device="cuda")
    nl.call("drn_conv_igemm_pre", ctypes.byref(d), _p(x), _p(ss), _p(w), _p(got.t), 0, _p(st_got), K, 0, 0, 0, _st())
    torch.cuda.synchronize()
    assert got.intact()
    assert (got.t.float() - want.float()).abs().max() <= 0.01 * want.float().abs().max() + 1e-3
    assert torch.allclose(st_got, st_want, rtol=2e-3, atol=0.5)
    # first principles
    act = torch.relu(x.float() * ss[:, 0] + ss[:, 1]).permute(0, 3, 1, 2)
    act = F.pad(act, (g.pad_lo, g.pad_hi, g.pad_lo, g.pad_hi))     # zero padding AFTER the activation
    ref = F.conv2d(act, w.float().permute(0, 3, 1, 2), stride=st).permute(0, 2, 3, 1)
    assert (got.t.float() - ref).abs().max() <= 0.03 * ref.abs().max() + 1e-3


@pytest.mark.parametrize("N,H,C,K,ks,st", PRE_CASES)
def test_wgrad_with_batchnorm_prologue_equals_apply_then_wgrad(N, H, C, K, ks, st):
    from distributed_resnet_tensorflow_b200.ops.conv import ConvGeom
    nl = _nl()
    g = ConvGeom.tf_fixed_padding(N, H, H, C, K, ks, st)
    gen = torch.Generator(device="cuda").manual_seed(7)
    x = (torch.randn((N, H, H, C), device="cuda", generator=gen) * 1.5).bfloat16()
    dy = torch.randn((N, g.P, g.Q, K), device="cuda", generator=gen).bfloat16()
    ss = torch.stack([torch.rand(C, device="cuda", generator=gen) + 0.5,
                      torch.randn(C, device="cuda", generator=gen) * 0.3], 1).contiguous()
    a = torch.empty_like(x)
    nl.call("drn_bn_apply", _p(x), _p(ss), 0, _p(a), N * H * H, C, 1, _st())
    d = g.desc()
    want = torch.zeros((K, ks, ks, C), device="cuda")
    nl.call("drn_conv_wgrad", ctypes.byref(d), _p(a), _p(dy), _p(want), 0, _st())
    got = Guarded((K, ks, ks, C), torch.float32)
    got.t.zero_()
    nl.call("drn_conv_wgrad_pre", ctypes.byref(d), _p(x), _p(ss), _p(dy), _p(got.t), 0, _st())
    torch.cuda.synchronize()
    assert got.intact()
    assert (got.t - want).abs().max() <= 1e-3 * want.abs().max() + 1e-3


@pytest.mark.parametrize("N,H,C,K", [(4, 16, 64, 128), (2, 28, 128, 128), (3, 14, 256, 64)])
def test_stride2_dgrad_by_parity_phases(N, H, C, K):
    """drn_conv_dgrad_s2 (four stride-1 phase convolutions over dy) + the interleaving mask/reduce
    kernel against torch's conv2d_input of the 3x3/stride-2 convolution."""
    from distributed_resnet_tensorflow_b200.ops.conv import ConvGeom
    nl = _nl()
    g = ConvGeom.tf_fixed_padding(N, H, H, C, K, 3, 2)
    assert (g.pad_lo, g.pad_hi, g.P) == (1, 1, H // 2)
    gen = torch.Generator(device="cuda").manual_seed(8)
    w = (torch.randn((K, 3, 3, C), device="cuda", generator=gen) / (9 * C) ** 0.5).bfloat16()
    dy = torch.randn((N, g.P, g.Q, K), device="cuda", generator=gen).bfloat16()
    banks = Guarded((9 * K * C,), torch.bfloat16)
    nl.call("drn_weight_prep_dgrad_s2", _p(w), _p(banks.t), K, C, _st())
    phases = Guarded((4, N, H // 2, H // 2, C), torch.bfloat16)
    d = g.desc()
    nl.call("drn_conv_dgrad_s2", ctypes.byref(d), _p(dy), _p(banks.t), _p(phases.t), _st())
    torch.cuda.synchronize()
    assert banks.intact() and phases.intact()
    ref = torch.nn.grad.conv2d_input((N, C, H + 2, H + 2), w.float().permute(0, 3, 1, 2),
                                     dy.float().permute(0, 3, 1, 2), stride=2)
    ref = ref[:, :, 1:1 + H, 1:1 + H].permute(0, 2, 3, 1)
    got = torch.empty((N, H, H, C), device="cuda")
    for a in range(2):
        for b in range(2):
            got[:, a::2, b::2, :] = phases.t[a * 2 + b].float()
    assert (got - ref).abs().max() <= 0.02 * ref.abs().max() + 1e-3
    # interleave + mask + reductions in one pass
    y = torch.randn((N, H, H, C), device="cuda", generator=gen).bfloat16()
    ss = torch.stack([torch.ones(C, device="cuda"), torch.zeros(C, device="cuda")], 1).contiguous()
    dm = Guarded((N, H, H, C), torch.bfloat16)
    stats = torch.zeros((2, C), device="cuda")
    nl.call("drn_bn_bwd_mask_reduce", _p(phases.t), 0, 4, N, H, H, 0, 0, 0.0, _p(y), _p(ss), _p(dm.t), _p(stats),
            C, C, _st())
    torch.cuda.synchronize()
    assert dm.intact()
    want = got.bfloat16().float() * (y.float() > 0)
    assert (dm.t.float() - want).abs().max() <= 1e-2 * want.abs().max() + 1e-3
    assert torch.allclose(stats[0], dm.t.float().reshape(-1, C).sum(0), rtol=1e-3, atol=0.05)
    assert torch.allclose(stats[1], (dm.t.float() * y.float()).reshape(-1, C).sum(0), rtol=1e-3, atol=0.05)


@pytest.mark.parametrize("N,H,C,K,ks,mode,bn", [(8, 14, 256, 256, 3, 0, 128), (5, 7, 512, 512, 3, 1, 256),
                                                (3, 14, 1024, 256, 1, 2, 128), (7, 7, 512, 2048, 1, 1, 128)])
def test_cta_pair_kernel_equals_single_cta_kernel(N, H, C, K, ks, mode, bn, monkeypatch):
    """tcgen05 cta_group::2 variant (two CTAs of a cluster share one 256-row MMA, each stages half
    of the weight tile) against the single-CTA kernel: same products in the same order, so outputs
    are identical; includes an odd number of row tiles (the last pair has an empty partner)."""
    from distributed_resnet_tensorflow_b200.ops.conv import ConvGeom, conv_fprop
    lo = (ks - 1) // 2
    g = ConvGeom(N, H, H, C, K, ks, ks, 1, lo, ks - 1 - lo)
    gen = torch.Generator(device="cuda").manual_seed(9)
    x = torch.randn((N, H, H, C), device="cuda", generator=gen).bfloat16()
    w = (torch.randn((K, ks, ks, C), device="cuda", generator=gen) / (ks * ks * C) ** 0.5).bfloat16()
    side = torch.randn((N, g.P, g.Q, K), device="cuda", generator=gen).bfloat16() if mode else None
    ss = (torch.rand((K, 2), device="cuda", generator=gen) - 0.3) if mode == 2 else None
    outs = []
    for two in ("0", "2"):
        monkeypatch.setenv("DRN_TWO_CTA", two)
        y = Guarded((N, g.P, g.Q, K), torch.bfloat16)
        st = torch.zeros((2, K), device="cuda")
        conv_fprop(x, w, g, out=y.t, side=side, stats=st, bn_scale_shift=ss, mode=mode, force_bn=bn)
        torch.cuda.synchronize()
        assert y.intact()
        outs.append((y.t.clone(), st))
    assert torch.equal(outs[0][0], outs[1][0])
    assert torch.allclose(outs[0][1], outs[1][1], rtol=1e-4, atol=1e-3)
    xf = F.pad(x.float().permute(0, 3, 1, 2), (lo, ks - 1 - lo, lo, ks - 1 - lo))
    ref = F.conv2d(xf, w.float().permute(0, 3, 1, 2)).permute(0, 2, 3, 1)
    if mode == 1:
        ref = ref + side.float()
    elif mode == 2:
        ref = ref * ((side.float() * ss[:, 0] + ss[:, 1]) > 0)
    assert (outs[1][0].float() - ref).abs().max() <= 0.02 * ref.abs().max() + 1e-3


@pytest.mark.parametrize("N,H,C,K,ks,mode", [(16, 32, 16, 16, 3, 0), (8, 16, 32, 64, 3, 1), (6, 14, 256, 1024, 1, 1),
                                             (128, 8, 64, 64, 3, 0), (2, 56, 64, 256, 1, 1)])
def test_finalize_by_the_last_cta_equals_bn_finalize(N, H, C, K, ks, mode):
    """drn_conv_igemm_fin: the last CTA of the statistics-producing convolution computes (scale, shift),
    (mean, 1/std) and the moving averages; against drn_bn_finalize on the same sums.  Twice in a row:
    the counter must come back to zero."""
    from distributed_resnet_tensorflow_b200._native.lib import BnFin
    from distributed_resnet_tensorflow_b200.ops.conv import ConvGeom
    nl = _nl()
    g = ConvGeom.tf_fixed_padding(N, H, H, C, K, ks, 1)
    gen = torch.Generator(device="cuda").manual_seed(11)
    x = torch.randn((N, H, H, C), device="cuda", generator=gen).bfloat16()
    w = (torch.randn((K, ks, ks, C), device="cuda", generator=gen) / (ks * ks * C) ** 0.5).bfloat16()
    side = torch.randn((N, g.P, g.Q, K), device="cuda", generator=gen).bfloat16() if mode else None
    gamma = torch.rand(K, device="cuda", generator=gen) + 0.5
    beta = torch.randn(K, device="cuda", generator=gen)
    d = g.desc()
    count = N * g.P * g.Q
    y = torch.empty((N, g.P, g.Q, K), dtype=torch.bfloat16, device="cuda")
    counter = torch.zeros(1, dtype=torch.int32, device="cuda")
    rm, rv = torch.zeros(K, device="cuda"), torch.ones(K, device="cuda")
    rm2, rv2 = rm.clone(), rv.clone()
    for _ in range(2):
        stats = torch.zeros((2, K), device="cuda")
        ss, mi = torch.zeros((K, 2), device="cuda"), torch.zeros((K, 2), device="cuda")
        fin = BnFin(_p(counter), count, _p(gamma), _p(beta), _p(rm), _p(rv), 0.9, 1e-5, _p(ss), _p(mi))
        nl.call("drn_conv_igemm_fin", ctypes.byref(d), None, 0, None, ctypes.byref(fin), _p(x), _p(w), _p(y),
                _p(side) if side is not None else 0, _p(stats), K, 0, mode, 0, _st())
        ss2, mi2 = torch.zeros_like(ss), torch.zeros_like(mi)
        nl.call("drn_bn_finalize", _p(stats), K, count, _p(gamma), _p(beta), _p(rm2), _p(rv2), 0.9, 1e-5, _p(ss2),
                _p(mi2), K, _st())
        torch.cuda.synchronize()
        assert int(counter) == 0
        assert torch.equal(ss, ss2) and torch.equal(mi, mi2)
        assert torch.equal(rm, rm2) and torch.equal(rv, rv2)
        assert float(ss.abs().sum()) > 0
