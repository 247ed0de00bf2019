"""Two-side epilogues of the implicit-GEMM kernel (csrc/conv_igemm.cu: DRN_EPI_BNBWD_ADD,
DRN_EPI_BNBWD_APPLY, statistics-only passes) against a plain PyTorch fp32 restatement.  They finish
the BatchNorm+ReLU backward of the layer in front inside the data-gradient kernel (reference:
the autodiff of batch_norm_relu, resnet_model_official.py:41-50,154)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(N, H, C, K, seed):
    from distributed_resnet_tensorflow_b200.ops.conv import ConvGeom
    g = ConvGeom(N, H, H, C, K, 1, 1, 1, 0, 0)
    gen = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn((N, H, H, C), device="cuda", generator=gen).bfloat16()
    w = (torch.randn((K, 1, 1, C), device="cuda", generator=gen) / C ** 0.5).bfloat16()
    side = torch.randn((N, H, H, K), device="cuda", generator=gen).bfloat16()
    side2 = torch.randn((N, H, H, K), device="cuda", generator=gen).bfloat16()
    ss = torch.stack([torch.rand(K, device="cuda", generator=gen) + 0.5,
                      torch.randn(K, device="cuda", generator=gen) * 0.3], 1).contiguous()
    coef = torch.zeros((K, 4), device="cuda")
    coef[:, 0] = torch.rand(K, device="cuda", generator=gen) + 0.5
    coef[:, 1] = torch.randn(K, device="cuda", generator=gen) * 0.1
    coef[:, 2] = torch.randn(K, device="cuda", generator=gen) * 0.1
    acc = x.float().reshape(-1, C) @ w.float().reshape(K, C).t()
    return g, x, w, side, side2, ss, coef, acc.reshape(N, H, H, K)


def _close(got, want, tol=0.02):
    got, want = got.float(), want.float()
    assert torch.isfinite(got).all()
    assert (got - want).abs().max() <= tol * want.abs().max() + 2e-2, float((got - want).abs().max())


@pytest.mark.parametrize("shape", [(4, 8, 64, 256), (3, 7, 256, 512), (2, 14, 128, 128), (4, 8, 64, 64)])
def test_bnbwd_add_and_stats_only(shape):
    from distributed_resnet_tensorflow_b200._native.lib import EPI_BNBWD, EPI_BNBWD_ADD
    from distributed_resnet_tensorflow_b200.ops.conv import conv_fprop, conv_fprop_epi2
    g, x, w, side, side2, ss, coef, acc = _setup(*shape, seed=1)
    K = shape[3]
    mask = (side.float() * ss[:, 0] + ss[:, 1]) > 0
    want = ((acc + side2.float()) * mask)
    stats = torch.zeros((2, K), device="cuda")
    got = conv_fprop_epi2(x, w, g, side, side2, stats=stats, bn_scale_shift=ss, mode=EPI_BNBWD_ADD)
    torch.cuda.synchronize()
    _close(got, want)
    gq = got.float()
    s1, s2 = gq.sum((0, 1, 2)), (gq * side.float()).sum((0, 1, 2))
    assert torch.allclose(stats[0], s1, rtol=2e-3, atol=0.5) and torch.allclose(stats[1], s2, rtol=2e-3, atol=0.5)
    # statistics only: same sums, nothing written
    stats2 = torch.zeros((2, K), device="cuda")
    assert conv_fprop_epi2(x, w, g, side, side2, stats=stats2, bn_scale_shift=ss, mode=EPI_BNBWD_ADD, store=False) is None
    torch.cuda.synchronize()
    assert torch.allclose(stats2, stats, rtol=1e-4, atol=0.2)
    # ... and the one-side mode without a store
    ref_stats = torch.zeros((2, K), device="cuda")
    conv_fprop(x, w, g, side=side, stats=ref_stats, bn_scale_shift=ss, mode=EPI_BNBWD)
    only = torch.zeros((2, K), device="cuda")
    conv_fprop_epi2(x, w, g, side, None, stats=only, bn_scale_shift=ss, mode=EPI_BNBWD, store=False)
    torch.cuda.synchronize()
    assert torch.allclose(only, ref_stats, rtol=1e-4, atol=0.2)


@pytest.mark.parametrize("shape", [(4, 8, 64, 256), (3, 7, 256, 512), (2, 14, 128, 128)])
@pytest.mark.parametrize("first", [False, True])
def test_bnbwd_apply(shape, first):
    from distributed_resnet_tensorflow_b200._native.lib import EPI_BNBWD_APPLY
    from distributed_resnet_tensorflow_b200.ops.conv import conv_fprop_epi2
    g, x, w, side, side2, ss, coef, acc = _setup(*shape, seed=2)
    mask = (side.float() * ss[:, 0] + ss[:, 1]) > 0
    v = acc + side2.float() if first else acc
    want = coef[:, 0] * (v * mask) + coef[:, 1] * side.float() + coef[:, 2]
    if not first:
        want = want + side2.float()
    guard = torch.full((shape[0], shape[1], shape[1], shape[3]), 7.0, device="cuda").bfloat16()
    got = conv_fprop_epi2(x, w, g, side, side2, out=guard, bn_scale_shift=ss, bn_coef=coef, side2_first=first,
                          mode=EPI_BNBWD_APPLY)
    torch.cuda.synchronize()
    _close(got, want)
