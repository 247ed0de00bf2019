"""Training-level parity of the native engine with the PyTorch fp32 oracle ("loss curve
within tolerance of the oracle for 200 steps from identical init"), and the per-tensor gradient check
on the FULL ResNet-50 at the benchmark batch size (at random init with noise inputs the gradient DIRECTION
of most tensors of a 50-layer network is lost to bf16 round-off by any bf16 implementation - PyTorch's own
bf16 path agrees with fp32 to a cosine of ~0.35 - so the full-size check asserts magnitudes everywhere and
directions where bf16 can hold them; the direction checks proper run on the shallower networks of
test_engine_gpu.py)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_loss_curve_tracks_the_fp32_oracle_for_200_steps():
    from distributed_resnet_tensorflow_b200.data.synthetic import synthetic_cifar
    from distributed_resnet_tensorflow_b200.models.init import init_variables
    from distributed_resnet_tensorflow_b200.models.spec import build_spec
    from distributed_resnet_tensorflow_b200.train.backends import DistContext, NativeBackend, TorchBackend
    spec = build_spec("cifar10", 8)
    B, steps, win = 32, 200, 20
    variables = init_variables(spec, seed=0)
    nb = NativeBackend(spec, B, DistContext(), device="cuda:0", weight_decay=2e-4, seed=0)
    tb = TorchBackend(spec, B, DistContext(), device="cuda:0", dtype="fp32", weight_decay=2e-4, seed=0)
    nb.import_variables(variables)
    tb.import_variables(variables)
    imgs, labs = synthetic_cifar(512, 10, seed=3, learnable=True)
    ln, lt = [], []
    for s in range(steps):
        lo = (s * B) % 512
        bi, bl = imgs[lo:lo + B], labs[lo:lo + B]
        ln.append(nb.train_step(bi, bl, 0.05)["cross_entropy"])
        lt.append(tb.train_step(bi, bl, 0.05)["cross_entropy"])
    ln, lt = np.asarray(ln).reshape(-1, win).mean(1), np.asarray(lt).reshape(-1, win).mean(1)
    print("native", np.round(ln, 3), "\noracle", np.round(lt, 3))
    assert np.all(np.isfinite(ln))
    assert lt[-1] < 0.5 * lt[0] and ln[-1] < 0.5 * ln[0], (ln, lt)          # both learn the task
    assert np.all(np.abs(ln - lt) <= 0.2 * lt + 0.08), (ln, lt)             # and along the same curve


def test_full_resnet50_gradients_at_batch_128():
    from distributed_resnet_tensorflow_b200.models.engine import make_engine
    from distributed_resnet_tensorflow_b200.models.init import init_variables
    from distributed_resnet_tensorflow_b200.models.spec import build_spec
    from distributed_resnet_tensorflow_b200.models.torch_ref import TorchResNet
    spec = build_spec("imagenet", 50)
    B = 128
    variables = init_variables(spec, 0)
    dev = "cuda:0"
    g = torch.Generator(device="cpu").manual_seed(0)
    images = torch.randn((B, 224, 224, 3), generator=g).bfloat16().float()
    labels = torch.randint(0, 1001, (B,), generator=g)
    ref = TorchResNet(spec, device=dev)
    ref.import_variables(variables)
    ref.loss_and_grads(images.to(dev), labels.to(dev), 1e-4)
    want = {n: ref._to_tf_layout(n, ref.params[n].grad.detach().cpu().numpy()) for n in ref._names}
    del ref
    r16 = TorchResNet(spec, device=dev, dtype=torch.bfloat16)
    r16.import_variables(variables)
    r16.loss_and_grads(images.to(dev), labels.to(dev), 1e-4)
    w16 = {n: r16._to_tf_layout(n, r16.params[n].grad.detach().float().cpu().numpy()) for n in r16._names}
    del r16
    torch.cuda.empty_cache()
    eng = make_engine(spec, B, device=dev)
    eng.import_variables(variables)
    eng.set_input_float(images, labels)
    eng.set_hyper(0.1, 0.9, 1e-4, 1.0)
    eng.scalars.zero_(); eng.correct.zero_()
    eng.forward(True); eng.loss(True); eng.backward()
    torch.cuda.synchronize()
    host_g = eng.grad.cpu().numpy()

    def cos(a, b):
        a, b = a.astype(np.float64).ravel(), b.astype(np.float64).ravel()
        return float(a @ b / (np.linalg.norm(a) * np.linalg.norm(b) + 1e-30))

    worst, worst16, below, checked, ratios = 1.0, 1.0, [], 0, []
    for name in eng.layout.order:
        e = eng.layout.entries[name]
        got = host_g[e.offset:e.offset + e.numel]
        w = eng.layout.to_dev(name, want[name]).reshape(-1)
        if np.linalg.norm(w) < 1e-12:
            continue
        w16d = eng.layout.to_dev(name, w16[name]).reshape(-1)
        c, c16 = cos(got, w), cos(w16d, w)
        ratio, ratio16 = np.linalg.norm(got) / np.linalg.norm(w), np.linalg.norm(w16d) / np.linalg.norm(w)
        ratios.append((ratio, ratio16, name))
        # the gradient MAGNITUDE of every tensor survives bf16 even where its direction does not
        assert 0.5 < ratio < 2.0 and abs(np.log(ratio / ratio16)) < 0.5, (name, ratio, ratio16)
        worst, worst16 = min(worst, c), min(worst16, c16)
        if c < 0.98:
            below.append((name, round(c, 4), round(c16, 4)))
        # At random init with noise inputs the early layers' gradients of a 50-layer network are dominated
        # by bf16 round-off for ANY bf16 implementation (PyTorch's own bf16 path agrees with fp32 to a
        # cosine of only ~0.35 on the stem), so the yardstick is that path: never worse than it by more
        # than a hair, and within 0.97 of fp32 wherever bf16 can be (cosine of the torch bf16 path > 0.99).
        # (where even that path has lost the fp32 direction - cosine below 0.9 - two noise vectors are
        # being compared and nothing can be asserted)
        if c16 >= 0.9:
            checked += 1
            assert c > c16 - 0.05, (name, c, c16)
        assert c16 < 0.99 or c > 0.97, (name, c, c16)
    print(f"worst cosine vs fp32: native {worst:.4f}, torch bf16 {worst16:.4f}; {checked} tensors checked against "
          f"the torch bf16 path; {len(below)} tensors below 0.98; norm ratio native/fp32 in "
          f"[{min(r[0] for r in ratios):.3f}, {max(r[0] for r in ratios):.3f}], torch bf16/fp32 in "
          f"[{min(r[1] for r in ratios):.3f}, {max(r[1] for r in ratios):.3f}]")
    assert checked >= 2
