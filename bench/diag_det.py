"""Diagnostic: per-tensor comparison of one training step's gradients between the ordinary and the
deterministic mode (and between two ordinary runs)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from distributed_resnet_tensorflow_b200._native import lib as nl
from distributed_resnet_tensorflow_b200.models.engine import make_engine
import test_deterministic_gpu as T

spec, variables, images, labels = T._problem("cifar10", 20, 64, 2)


def one(det):
    nl.call("drn_set_deterministic", 1 if det else 0)
    eng = make_engine(spec, 64, device="cuda:0")
    eng.import_variables(variables)
    eng.set_hyper(0.05, 0.9, 1e-4, 1.0)
    eng.set_input_float(images, labels)
    eng.scalars.zero_(); eng.correct.zero_(); eng.grad.zero_()
    eng.forward(training=True)
    eng.loss(with_grad=True)
    eng.backward()
    torch.cuda.synchronize()
    stats = {bn.spec.name: (bn.fstats.clone(), bn.bstats.clone()) for bn in eng._all_bns()}
    return eng, eng.grad.clone(), stats, eng.logits.clone()


def cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float(a @ b / (a.norm() * b.norm() + 1e-30))


e0, g0, s0, l0 = one(False)
e1, g1, s1, l1 = one(False)
e2, g2, s2, l2 = one(True)
print("ordinary vs ordinary", cos(g0, g1), "ordinary vs det", cos(g0, g2))
print("logits max diff ord/ord", float((l0.float() - l1.float()).abs().max()), "ord/det", float((l0.float() - l2.float()).abs().max()))
for name in list(s0)[:40]:
    f0, b0 = s0[name]; f2, b2 = s2[name]
    print(f"{name:40s} fstats rel {float((f0 - f2).abs().max() / (f0.abs().max() + 1e-30)):.2e}  bstats rel {float((b0 - b2).abs().max() / (b0.abs().max() + 1e-30)):.2e}")
for name in e0.layout.order[:60]:
    e = e0.layout.entries[name]
    a, b = g0[e.offset:e.offset + e.numel], g2[e.offset:e.offset + e.numel]
    print(f"{name:50s} cos {cos(a, b):.6f}  norm ratio {float(b.norm() / (a.norm() + 1e-30)):.4f}")
