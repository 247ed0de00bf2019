"""Native sm_100a engine vs the PyTorch fp32 oracle: logits, loss, gradients, BN statistics and
the weights after one momentum-SGD step, for small CIFAR / ImageNet v2 networks."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a @ b) / (a.norm() * b.norm() + 1e-30))


def _run(dataset, size, batch, seed=0, custom=None, version=2):
    from distributed_resnet_tensorflow_b200.models.engine import make_engine
    from distributed_resnet_tensorflow_b200.models.init import init_variables
    from distributed_resnet_tensorflow_b200.models.spec import build_spec
    from distributed_resnet_tensorflow_b200.models.torch_ref import TorchResNet

    if custom is not None:
        from distributed_resnet_tensorflow_b200.models.spec import imagenet_resnet
        spec = imagenet_resnet(size, version=version, custom=custom)
    else:
        spec = build_spec(dataset, size, version)
    variables = init_variables(spec, seed)
    # make BN affine parameters non-trivial so their gradients are exercised
    rng = np.random.default_rng(seed + 1)
    for k in variables:
        if k.endswith("/gamma"):
            variables[k] = (1.0 + 0.1 * rng.standard_normal(variables[k].shape)).astype(np.float32)
        if k.endswith("/beta"):
            variables[k] = (0.1 * rng.standard_normal(variables[k].shape)).astype(np.float32)
    dev = "cuda:0"
    ref = TorchResNet(spec, device=dev)
    ref.import_variables(variables)
    r16 = TorchResNet(spec, device=dev, dtype=torch.bfloat16)   # bf16 noise floor
    r16.import_variables(variables)
    eng = make_engine(spec, batch, device=dev)
    eng.import_variables(variables)

    g = torch.Generator(device="cpu").manual_seed(seed)
    images = torch.randn((batch, spec.image_size, spec.image_size, 3), generator=g)
    images = images.bfloat16().float()  # both sides see identical (bf16-representable) inputs
    labels = torch.randint(0, spec.num_classes, (batch,), generator=g)
    wd, lr, mu = 1e-4, 0.1, 0.9

    cost, xent, prec = ref.loss_and_grads(images.to(dev), labels.to(dev), wd)
    r16.loss_and_grads(images.to(dev), labels.to(dev), wd)

    eng.set_input_float(images, labels)
    eng.set_hyper(lr, mu, wd, 1.0)
    eng.scalars.zero_(); eng.correct.zero_()
    eng.forward(training=True)
    eng.loss(with_grad=True)
    eng.backward()
    torch.cuda.synchronize()
    loss_native = float(eng.scalars[0]) / batch
    assert abs(loss_native - float(xent)) < 0.05 * max(1.0, abs(float(xent))), (loss_native, float(xent))

    # gradients, tensor by tensor, in device layout
    worst = 1.0
    host_g = eng.grad.cpu().numpy()
    for name in eng.layout.order:
        e = eng.layout.entries[name]
        got = torch.from_numpy(host_g[e.offset:e.offset + e.numel])
        want_tf = ref._to_tf_layout(name, ref.params[name].grad.detach().cpu().numpy())
        want = torch.from_numpy(eng.layout.to_dev(name, want_tf).reshape(-1))
        if float(want.norm()) < 1e-12:
            continue
        w16_tf = r16._to_tf_layout(name, r16.params[name].grad.detach().float().cpu().numpy())
        w16 = torch.from_numpy(eng.layout.to_dev(name, w16_tf).reshape(-1))
        c, c16 = _cos(got, want), _cos(w16, want)
        ratio = float(got.norm() / want.norm())
        worst = min(worst, c)
        # the native bf16 pipeline must track the fp32 oracle as well as PyTorch's own bf16 path does
        assert c > c16 - 0.04 and c > 0.85 and 0.8 < ratio < 1.25, (name, c, c16, ratio)
    # BN moving statistics after one step
    for bn in eng._all_bns():
        rm = ref.buffers[f"{bn.spec.name}/moving_mean"]
        rv = ref.buffers[f"{bn.spec.name}/moving_variance"]
        assert torch.allclose(bn.run_mean, rm, atol=3e-3, rtol=5e-2), bn.spec.name
        assert torch.allclose(bn.run_var, rv, atol=3e-3, rtol=5e-2), bn.spec.name
    # one optimizer step on both sides
    ref.apply_gradients(lr, mu, wd)
    eng.apply_gradients()
    torch.cuda.synchronize()
    out = eng.export_variables()
    refv = ref.export_variables()
    for name, _ in spec.trainable_variables():
        a, b = torch.from_numpy(out[name]), torch.from_numpy(refv[name])
        assert _cos(a, b) > (0.999 if version == 2 else 0.998), name
    return worst


def test_cifar_resnet8_matches_oracle():
    _run("cifar10", 8, 32)


def test_cifar_resnet14_tail_batch():
    _run("cifar10", 14, 10)  # M not a multiple of 128 anywhere


def test_imagenet_resnet18_matches_oracle():
    _run("imagenet", 18, 4)


def test_imagenet_bottleneck_matches_oracle():
    # every code path of ResNet-50 (bottlenecks, stride-1 and stride-2 projections, strided 3x3
    # dgrad) at a depth where bf16 round-off does not yet dominate the comparison
    _run("imagenet", 14, 8, custom=("bottleneck", [1, 1, 1, 1]))


# ---- post-activation (v1) networks on the native engine (models/engine_v1.py) --------------------
def test_v1_cifar_resnet8_matches_oracle():
    _run("cifar10", 8, 32, version=1)


def test_v1_imagenet_resnet18_matches_oracle():
    _run("imagenet", 18, 4, version=1)


def test_v1_imagenet_bottleneck_matches_oracle():
    _run("imagenet", 14, 8, custom=("bottleneck", [1, 1, 1, 1]), version=1)


def test_v1_imagenet_resnet50_trains():
    from distributed_resnet_tensorflow_b200.models.init import init_variables
    from distributed_resnet_tensorflow_b200.models.spec import build_spec
    from distributed_resnet_tensorflow_b200.train.native import NativeTrainer
    spec = build_spec("imagenet", 50, 1)
    B = 16
    tr = NativeTrainer(spec, B, device="cuda:0", weight_decay=1e-4, use_graph=True)
    tr.import_variables(init_variables(spec, seed=0))
    rng = np.random.default_rng(0)
    imgs = torch.from_numpy(rng.integers(0, 256, size=(B, 224 * 224 * 3), dtype=np.uint8)).pin_memory()
    labs = torch.from_numpy(rng.integers(0, 1001, size=(B,), dtype=np.int32)).pin_memory()
    losses = [tr.train_step(imgs, labs, lr=0.05)["cross_entropy"] for _ in range(12)]
    assert all(np.isfinite(losses)), losses
    assert losses[-1] < losses[0] - 0.5, losses


def test_imagenet_resnet50_trains():
    """Full ResNet-50: loss is finite and falls when the same batch is repeated."""
    from distributed_resnet_tensorflow_b200.models.init import init_variables
    from distributed_resnet_tensorflow_b200.models.spec import build_spec
    from distributed_resnet_tensorflow_b200.train.native import NativeTrainer
    spec = build_spec("imagenet", 50)
    B = 16
    tr = NativeTrainer(spec, B, device="cuda:0", weight_decay=1e-4, use_graph=True)
    tr.import_variables(init_variables(spec, seed=0))
    rng = np.random.default_rng(0)
    imgs = torch.from_numpy(rng.integers(0, 256, size=(B, 224 * 224 * 3), dtype=np.uint8)).pin_memory()
    labs = torch.from_numpy(rng.integers(0, 1001, size=(B,), dtype=np.int32)).pin_memory()
    losses = [tr.train_step(imgs, labs, lr=0.05)["cross_entropy"] for _ in range(12)]
    assert all(np.isfinite(losses)), losses
    assert losses[-1] < losses[0] - 0.5, losses
