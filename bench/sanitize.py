"""Small programs for compute-sanitizer (memcheck / racecheck / synccheck / initcheck).

    python bench/sanitize.py                       # must exit 0 on its own first
    compute-sanitizer --tool memcheck python bench/sanitize.py

Runs every kernel family of the training and evaluation steps once on tiny problems: CIFAR ResNet-8
(3x3 stem, basic blocks, strided projections) and ImageNet ResNet-50 at batch 2 (space-to-depth stem
view, max pool, bottlenecks, strided dgrad through zero insertion, 1001-way classifier), plus the
no-allocation dgrad entry point.  SURVEY.md section 5.2: the reference has no sanitizer story."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_resnet_tensorflow_b200.models.init import init_variables  # noqa: E402
from distributed_resnet_tensorflow_b200.models.spec import build_spec  # noqa: E402
from distributed_resnet_tensorflow_b200.train.native import NativeTrainer  # noqa: E402


def run(dataset, size, batch, steps):
    spec = build_spec(dataset, size)
    tr = NativeTrainer(spec, batch, device="cuda:0", weight_decay=1e-4, use_graph=False)
    tr.import_variables(init_variables(spec, seed=0))
    rng = np.random.default_rng(0)
    npix = spec.image_size * spec.image_size * 3
    imgs = torch.from_numpy(rng.integers(0, 256, size=(batch, npix), dtype=np.uint8)).pin_memory()
    labs = torch.from_numpy(rng.integers(0, spec.num_classes, size=(batch,), dtype=np.int32)).pin_memory()
    for _ in range(steps):
        m = tr.train_step(imgs, labs, lr=0.05)
    e = tr.eval_batch(imgs, labs)
    torch.cuda.synchronize()
    assert np.isfinite(m["cost"]), m
    print(f"{dataset}-{size} B={batch}: train cost {m['cost']:.4f}, eval {e}", flush=True)


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which in ("all", "cifar"):
        run("cifar10", 8, 8, 2)
    if which in ("all", "imagenet"):
        run("imagenet", 50, 2, 1)
    print("SANITIZE_PROGRAM_OK", flush=True)
