"""Each hand-written sm_100a kernel against a plain PyTorch fp32 reference of the same op."""
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


@pytest.mark.parametrize("name,N,H,C,K,ks,st", [
    ("1x1", 8, 14, 256, 128, 1, 1), ("3x3", 4, 28, 64, 64, 3, 1), ("3x3s2", 4, 28, 128, 128, 3, 2),
    ("1x1s2", 4, 28, 256, 512, 1, 2), ("cifar16", 8, 32, 16, 16, 3, 1), ("cifar32s2", 8, 16, 16, 32, 3, 2),
    ("tail", 5, 7, 512, 512, 3, 1)])
def test_conv_fprop_and_wgrad(name, N, H, C, K, ks, st):
    from distributed_resnet_tensorflow_b200.ops.conv import ConvGeom, conv_fprop, conv_wgrad
    g = ConvGeom.tf_fixed_padding(N, H, H, C, K, ks, st)
    gen = torch.Generator(device="cuda").manual_seed(0)
    x = torch.randn((N, H, H, C), device="cuda", generator=gen).bfloat16()
    w = (torch.randn((K, ks, ks, C), device="cuda", generator=gen) / (ks * ks * C) ** 0.5).bfloat16()
    y = conv_fprop(x, w, g)
    xf = F.pad(x.float().permute(0, 3, 1, 2), (g.pad_lo, g.pad_hi, g.pad_lo, g.pad_hi))
    ref = F.conv2d(xf, w.float().permute(0, 3, 1, 2), stride=st).permute(0, 2, 3, 1)
    assert (y.float() - ref).abs().max() <= 0.02 * ref.abs().max() + 1e-3
    dy = torch.randn((N, g.P, g.Q, K), device="cuda", generator=gen).bfloat16()
    dw = conv_wgrad(x, dy, g)
    refw = torch.nn.grad.conv2d_weight(xf, (K, C, ks, ks), dy.float().permute(0, 3, 1, 2), stride=st)
    refw = refw.permute(0, 2, 3, 1)
    assert (dw - refw).abs().max() <= 2e-3 * refw.abs().max() + 1e-3


@pytest.mark.parametrize("ks,st", [(3, 1), (3, 2), (1, 1)])
def test_dgrad_via_flipped_filters(ks, st):
    """dx from the forward kernel on (zero-upsampled) dy and the flipped/transposed filter bank."""
    from distributed_resnet_tensorflow_b200.ops.conv import ConvGeom, conv_fprop
    nl = _nl()
    N, H, C, K = 4, 16, 64, 128
    g = ConvGeom.tf_fixed_padding(N, H, H, C, K, ks, st)
    gen = torch.Generator(device="cuda").manual_seed(1)
    w = (torch.randn((K, ks, ks, C), device="cuda", generator=gen) / (ks * ks * C) ** 0.5).bfloat16()
    dy = torch.randn((N, g.P, g.Q, K), device="cuda", generator=gen).bfloat16()
    wd = torch.empty(K * ks * ks * C, dtype=torch.bfloat16, device="cuda")
    nl.call("drn_weight_prep_dgrad", _p(w), _p(wd), K, ks, ks, C, _st())
    dg = g.dgrad_geom()
    src = dy
    if st > 1:
        src = torch.empty((N, dg.H, dg.W, K), dtype=torch.bfloat16, device="cuda")
        nl.call("drn_upsample_zero", _p(dy), _p(src), N, g.P, g.Q, dg.H, dg.W, K, st, _st())
    dx = conv_fprop(src, wd, dg)
    assert tuple(dx.shape) == (N, H, H, C)
    from distributed_resnet_tensorflow_b200.ops.conv import conv_dgrad
    dx2 = conv_dgrad(dy, wd, g)            # the C entry point that derives the geometry itself
    assert torch.equal(dx, dx2)
    xz = torch.zeros((N, C, H + g.pad_lo + g.pad_hi, H + g.pad_lo + g.pad_hi), device="cuda")
    ref = torch.nn.grad.conv2d_input(xz.shape, w.float().permute(0, 3, 1, 2), dy.float().permute(0, 3, 1, 2), stride=st)
    ref = ref[:, :, g.pad_lo:g.pad_lo + H, g.pad_lo:g.pad_lo + H].permute(0, 2, 3, 1)
    assert (dx.float() - ref).abs().max() <= 0.02 * ref.abs().max() + 1e-3


def test_bn_forward_and_backward_chain():
    nl = _nl()
    M, C = 4 * 14 * 14, 256
    gen = torch.Generator(device="cuda").manual_seed(2)
    y = (torch.randn((M, C), device="cuda", generator=gen) * 2 + 0.5).bfloat16()
    gamma = torch.rand(C, device="cuda", generator=gen) + 0.5
    beta = torch.randn(C, device="cuda", generator=gen) * 0.2
    rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    stats = torch.zeros((2, C), device="cuda")
    ss, mi, coef = (torch.zeros((C, k), device="cuda") for k in (2, 2, 4))
    nl.call("drn_bn_stats", _p(y), _p(stats), C, M, C, _st())
    yf = y.float()
    assert torch.allclose(stats[0], yf.sum(0), rtol=1e-4, atol=1e-2)
    assert torch.allclose(stats[1], (yf * yf).sum(0), rtol=1e-4, atol=1e-2)
    nl.call("drn_bn_finalize", _p(stats), C, M, _p(gamma), _p(beta), _p(rm), _p(rv), 0.997, 1e-5, _p(ss), _p(mi), C, _st())
    a = torch.empty_like(y)
    nl.call("drn_bn_apply", _p(y), _p(ss), 0, _p(a), M, C, 1, _st())
    trm, trv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    yr = yf.clone().requires_grad_(True)
    gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    ref = F.relu(F.batch_norm(yr, trm, trv, gr, br, training=True, momentum=0.003, eps=1e-5))
    assert (a.float() - ref).abs().max() < 0.03
    assert torch.allclose(rm, trm, atol=1e-5) and torch.allclose(rv, trv, atol=1e-5)
    # backward: mask + reductions, finalize, apply (+ addend)
    da = torch.randn((M, C), device="cuda", generator=gen).bfloat16()
    add = torch.randn((M, C), device="cuda", generator=gen).bfloat16()
    dm = torch.empty_like(y)
    bst = torch.zeros((2, C), device="cuda")
    nl.call("drn_bn_bwd_mask_reduce", _p(da), 0, 0, 1, M, 1, 0, 0, 0.0, _p(y), _p(ss), _p(dm), _p(bst), C, C, _st())
    dg, db = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
    nl.call("drn_bn_bwd_finalize", _p(bst), C, M, _p(gamma), _p(mi), _p(dg), _p(db), _p(coef), C, _st())
    dy = torch.empty_like(y)
    nl.call("drn_bn_bwd_apply", _p(dm), _p(y), _p(coef), _p(add), _p(dy), M, C, _st())
    ref.backward(da.float())
    assert torch.allclose(dg, gr.grad, rtol=2e-2, atol=0.3) and torch.allclose(db, br.grad, rtol=2e-2, atol=0.3)
    want = yr.grad + add.float()
    assert (dy.float() - want).abs().max() < 0.05 * want.abs().max()


def test_bn_inference_params():
    nl = _nl()
    C = 64
    g, b, rm, rv = (torch.rand(C, device="cuda") + 0.5 for _ in range(4))
    ss = torch.zeros((C, 2), device="cuda")
    nl.call("drn_bn_inference_params", _p(g), _p(b), _p(rm), _p(rv), 1e-5, _p(ss), C, _st())
    sc = g / torch.sqrt(rv + 1e-5)
    assert torch.allclose(ss[:, 0], sc, rtol=1e-5) and torch.allclose(ss[:, 1], b - rm * sc, rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("H", [16, 15, 7])
def test_maxpool_same_padding_fwd_bwd(H):
    nl = _nl()
    N, C = 3, 64
    P = (H + 1) // 2
    total_pad = max((P - 1) * 2 + 3 - H, 0)
    lo, hi = total_pad // 2, total_pad - total_pad // 2          # TF "SAME": the extra pixel goes after
    gen = torch.Generator(device="cuda").manual_seed(3)
    x = torch.randn((N, H, H, C), device="cuda", generator=gen).bfloat16()
    out = torch.empty((N, P, P, C), dtype=torch.bfloat16, device="cuda")
    arg = torch.empty((N, P, P, C), dtype=torch.uint8, device="cuda")
    nl.call("drn_maxpool_fwd", _p(x), _p(out), _p(arg), N, H, H, C, 3, 2, lo, P, P, _st())
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    ref = F.max_pool2d(F.pad(xr, (lo, hi, lo, hi), value=float("-inf")), 3, 2)
    assert torch.equal(out.float(), ref.permute(0, 2, 3, 1))
    dy = torch.randn((N, P, P, C), device="cuda", generator=gen).bfloat16()
    dx = torch.full_like(x, float("nan"))                        # every pixel must be written
    nl.call("drn_maxpool_bwd", _p(dy), _p(arg), _p(dx), N, H, H, C, 3, 2, lo, P, P, _st())
    ref.backward(dy.float().permute(0, 3, 1, 2))
    assert (dx.float() - xr.grad.permute(0, 2, 3, 1)).abs().max() < 0.02


def test_bnrelu_avgpool_and_its_gradient():
    nl = _nl()
    N, HW, C = 6, 49, 512
    gen = torch.Generator(device="cuda").manual_seed(4)
    y = torch.randn((N, 7, 7, C), device="cuda", generator=gen).bfloat16()
    ss = torch.stack([torch.rand(C, device="cuda", generator=gen) + 0.5,
                      torch.randn(C, device="cuda", generator=gen) * 0.2], 1).contiguous()
    out = torch.empty((N, C), dtype=torch.bfloat16, device="cuda")
    nl.call("drn_bnrelu_avgpool", _p(y), _p(ss), _p(out), N, HW, C, _st())
    a = F.relu(y.float() * ss[:, 0] + ss[:, 1])
    assert (out.float() - a.mean(dim=(1, 2))).abs().max() < 0.01
    dp = torch.randn((N, C), device="cuda", generator=gen).bfloat16()
    dm = torch.empty_like(y)
    st = torch.zeros((2, C), device="cuda")
    nl.call("drn_bn_bwd_mask_reduce", 0, _p(dp), 3, N, 7, 7, 0, 0, 1.0 / HW, _p(y), _p(ss), _p(dm), _p(st), C, C, _st())
    want = (a > 0) * dp.float()[:, None, None, :] / HW
    assert (dm.float() - want).abs().max() < 1e-3 + 0.01 * want.abs().max()
    assert torch.allclose(st[0], dm.float().sum(dim=(0, 1, 2)), rtol=1e-3, atol=1e-3)


def test_softmax_xent_forward_backward_precision():
    nl = _nl()
    B, classes, ld = 64, 1001, 1024
    gen = torch.Generator(device="cuda").manual_seed(5)
    logits = torch.randn((B, ld), device="cuda", generator=gen).bfloat16()
    bias = torch.randn(classes, device="cuda", generator=gen) * 0.1
    labels = torch.randint(0, classes, (B,), device="cuda", generator=gen, dtype=torch.int32)
    dlog = torch.empty_like(logits)
    dbias, probs = torch.zeros(classes, device="cuda"), torch.zeros((B, classes), device="cuda")
    loss, correct = torch.zeros(1, device="cuda"), torch.zeros(1, dtype=torch.int32, device="cuda")
    nl.call("drn_softmax_xent", _p(logits), ld, _p(bias), _p(labels), B, classes, 1.0 / B, _p(dlog), _p(dbias),
            _p(probs), _p(loss), _p(correct), _st())
    z = (logits.float()[:, :classes] + bias).requires_grad_(True)
    ref = F.cross_entropy(z, labels.long(), reduction="sum")
    ref.backward()
    assert abs(float(loss) - float(ref.detach())) < 1e-3 * float(ref.detach())
    assert int(correct) == int((z.argmax(1) == labels).sum())
    assert torch.allclose(probs, F.softmax(z.detach(), 1), atol=1e-5)
    assert (dlog.float()[:, :classes] - z.grad / B).abs().max() < 1e-4 and torch.all(dlog[:, classes:] == 0)
    assert torch.allclose(dbias, z.grad.sum(0) / B, atol=1e-4)


def test_momentum_sgd_with_weight_decay():
    nl = _nl()
    n = 1 << 16
    gen = torch.Generator(device="cuda").manual_seed(6)
    w, g, m = (torch.randn(n, device="cuda", generator=gen) for _ in range(3))
    w0, g0, m0 = w.clone(), g.clone(), m.clone()
    shadow = torch.empty(n, dtype=torch.bfloat16, device="cuda")
    hyper = torch.tensor([0.1, 0.9, 2e-4, 0.5], device="cuda")
    l2 = torch.zeros(1, device="cuda")
    nl.call("drn_sgd_momentum", _p(w), _p(g), _p(m), _p(shadow), n, _p(hyper), _p(l2), 1, _st())
    m_ref = 0.9 * m0 + (0.5 * g0 + 2e-4 * w0)
    w_ref = w0 - 0.1 * m_ref
    assert torch.allclose(m, m_ref, atol=1e-6) and torch.allclose(w, w_ref, atol=1e-6)
    assert torch.all(g == 0) and torch.equal(shadow, w.bfloat16())
    assert abs(float(l2) - float((w0 * w0).sum())) < 1e-3 * float((w0 * w0).sum())


def test_cifar_augment_matches_cpu_oracle():
    from distributed_resnet_tensorflow_b200.data import preprocess
    nl = _nl()
    B = 16
    rng = np.random.default_rng(7)
    recs = rng.integers(0, 256, (B, 3072), dtype=np.uint8)
    d = torch.from_numpy(recs).cuda()
    out = torch.empty((B, 32, 32, 16), dtype=torch.bfloat16, device="cuda")
    for training, pad in ((1, 4), (1, 2), (0, 4)):
        nl.call("drn_cifar_augment", _p(d), 3072, 0, _p(out), B, pad, training, 12345, 9, 1000, 16, _st())
        ref = preprocess.cifar_preprocess_batch(recs, bool(training), pad, 12345, 9, 1000)
        got = out.float().cpu().numpy()
        assert np.abs(got[..., :3] - ref).max() < 0.03 and np.all(got[..., 3:] == 0)


def test_imagenet_prep_matches_cpu_oracle():
    from distributed_resnet_tensorflow_b200.data import preprocess
    nl = _nl()
    B, S = 4, 224
    rng = np.random.default_rng(8)
    imgs = rng.integers(0, 256, (B, S, S, 3), dtype=np.uint8)
    crop = np.array([[0, 0, 0], [0, 0, 1], [0, 0, 1], [0, 0, 0]], np.int32)
    d, c = torch.from_numpy(imgs).cuda(), torch.from_numpy(crop).cuda()
    out = torch.empty((B, S // 2, S // 2, 16), dtype=torch.bfloat16, device="cuda")
    m = preprocess.VGG_MEANS_01
    nl.call("drn_imagenet_prep", _p(d), S * S * 3, S, S, _p(c), _p(out), B, S, float(m[0]), float(m[1]), float(m[2]), 0, 0, _st())
    got = out.float().cpu().numpy()
    # the zero-bordered layout the stem reads: pitch 116 pixels, image at column 2
    outp = torch.zeros((B, S // 2, S // 2 + 4, 16), dtype=torch.bfloat16, device="cuda")
    nl.call("drn_imagenet_prep", _p(d), S * S * 3, S, S, _p(c), _p(outp), B, S, float(m[0]), float(m[1]), float(m[2]),
            S // 2 + 4, 2, _st())
    gotp = outp.float().cpu().numpy()
    assert np.all(gotp[:, :, :2] == 0) and np.all(gotp[:, :, -2:] == 0) and np.array_equal(gotp[:, :, 2:-2], got)
    for b in range(B):
        ref = preprocess.space_to_depth16(preprocess.imagenet_prep_oracle(imgs[b], bool(crop[b, 2])))
        assert np.abs(got[b] - ref).max() < 0.005


def test_resize_bilinear_matches_cpu_oracle():
    from distributed_resnet_tensorflow_b200.data import preprocess
    nl = _nl()
    img = np.random.default_rng(9).integers(0, 256, (37, 53, 3), dtype=np.uint8)
    d = torch.from_numpy(img).cuda()
    out = torch.empty((64, 91, 3), device="cuda")
    nl.call("drn_resize_bilinear", _p(d), 37, 53, _p(out), 64, 91, _st())
    assert np.abs(out.cpu().numpy() - preprocess.resize_bilinear(img, 64, 91)).max() < 1e-2
