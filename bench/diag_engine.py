"""Per-tensor gradient agreement: native engine vs fp32 oracle, next to bf16-autocast oracle vs fp32
oracle (the noise floor of bf16 training)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_resnet_tensorflow_b200.models.engine import NativeResNet
from distributed_resnet_tensorflow_b200.models.init import init_variables
from distributed_resnet_tensorflow_b200.models.spec import build_spec
from distributed_resnet_tensorflow_b200.models.torch_ref import TorchResNet

def cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a @ b) / (a.norm() * b.norm() + 1e-30))

def main(dataset, size, batch, seed=0):
    spec = build_spec(dataset, size)
    variables = init_variables(spec, seed)
    rng = np.random.default_rng(seed + 1)
    for k in variables:
        if k.endswith("/gamma"): variables[k] = (1.0 + 0.1 * rng.standard_normal(variables[k].shape)).astype(np.float32)
        if k.endswith("/beta"): variables[k] = (0.1 * rng.standard_normal(variables[k].shape)).astype(np.float32)
    dev = "cuda:0"
    ref = TorchResNet(spec, device=dev); ref.import_variables(variables)
    r16 = TorchResNet(spec, device=dev, dtype=torch.bfloat16); r16.import_variables(variables)
    eng = NativeResNet(spec, batch, device=dev); eng.import_variables(variables)
    g = torch.Generator(device="cpu").manual_seed(seed)
    images = torch.randn((batch, spec.image_size, spec.image_size, 3), generator=g).bfloat16().float()
    labels = torch.randint(0, spec.num_classes, (batch,), generator=g)
    _, xent, _ = ref.loss_and_grads(images.to(dev), labels.to(dev), 1e-4)
    _, xent16, _ = r16.loss_and_grads(images.to(dev), labels.to(dev), 1e-4)
    eng.set_input_float(images, labels); eng.scalars.zero_(); eng.correct.zero_()
    eng.forward(True); eng.loss(True); eng.backward(); torch.cuda.synchronize()
    print(f"{dataset}-{size} B={batch} loss fp32 {float(xent):.5f} bf16-oracle {float(xent16):.5f} native {float(eng.scalars[0])/batch:.5f}")
    hg = eng.grad.cpu().numpy()
    for name in eng.layout.order:
        e = eng.layout.entries[name]
        got = torch.from_numpy(hg[e.offset:e.offset + e.numel])
        want = torch.from_numpy(eng.layout.to_dev(name, ref._to_tf_layout(name, ref.params[name].grad.cpu().numpy())).reshape(-1))
        w16 = torch.from_numpy(eng.layout.to_dev(name, r16._to_tf_layout(name, r16.params[name].grad.float().cpu().numpy())).reshape(-1))
        if float(want.norm()) < 1e-12: continue
        print(f"  {name:34s} native cos {cos(got, want):.4f} ratio {float(got.norm()/want.norm()):.3f} | bf16-oracle cos {cos(w16, want):.4f} ratio {float(w16.norm()/want.norm()):.3f}")

if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]), int(sys.argv[3]))
