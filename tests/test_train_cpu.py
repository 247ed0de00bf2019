"""Serial training, checkpoint / resume and the separate evaluator on the CPU (torch backend)."""
import glob
import os

import numpy as np
import torch

from distributed_resnet_tensorflow_b200 import config as C
from distributed_resnet_tensorflow_b200.ckpt.saver import Saver, latest_checkpoint
from distributed_resnet_tensorflow_b200.utils import flags as F
from distributed_resnet_tensorflow_b200.utils.events import read_events


def _cfg(argv, defaults=F.CIFAR_MAIN, entry="train"):
    return C.from_flags(F.make_flagset(**defaults).parse(argv)[0], entry=entry)


def test_serial_training_learns_and_checkpoints(tmp_path):
    from distributed_resnet_tensorflow_b200.train.loop import train
    root = str(tmp_path / "ckpt")
    cfg = _cfg(["--resnet_size=8", "--batch_size=32", "--train_steps=30", f"--log_root={root}",
                f"--eval_dir={tmp_path / 'train'}", "--log_every=10", "--summary_every=10",
                "--save_checkpoint_steps=15", "--base_lr=0.05"])
    m = train(cfg)
    assert m["global_step"] == 30 and np.isfinite(m["cost"])
    assert m["precision"] > 0.3          # the synthetic set carries a learnable class signal
    assert latest_checkpoint(root).endswith("model.ckpt-30")
    v = Saver.restore(latest_checkpoint(root))
    assert int(v["global_step"]) == 30 and len(v) == 2 * 26 + 2 * 7 + 1   # trainables + Momentum, BN stats, step
    assert np.abs(v["conv2d/kernel/Momentum"]).max() > 0
    ev = [e for f in glob.glob(str(tmp_path / "train" / "events.*")) for e in read_events(f)]
    tags = {t for _, _, s in ev for t in s}
    assert {"cross_entropy", "cost", "learning_rate", "Precision"} <= tags


def test_resume_continues_step_counter_and_momentum(tmp_path):
    from distributed_resnet_tensorflow_b200.train.loop import train
    root = str(tmp_path / "ckpt")
    common = ["--resnet_size=8", "--batch_size=16", f"--log_root={root}", "--log_every=1000"]
    train(_cfg(common + ["--train_steps=4"]))
    v4 = Saver.restore(latest_checkpoint(root))
    m = train(_cfg(common + ["--train_steps=7"]))          # resume-by-resubmit
    assert m["global_step"] == 7
    v7 = Saver.restore(latest_checkpoint(root))
    assert int(v4["global_step"]) == 4 and int(v7["global_step"]) == 7
    assert not np.array_equal(v4["dense/kernel"], v7["dense/kernel"])
    # an uninterrupted 7-step run from the same seed ends in the same place as 4 + 3 resumed steps?
    # (input order differs after a restart, as in the reference, so only sanity-check magnitudes)
    assert np.abs(v7["dense/kernel/Momentum"]).max() > 0


def test_evaluator_restores_and_reports(tmp_path):
    from distributed_resnet_tensorflow_b200.train.evaluate import evaluate
    from distributed_resnet_tensorflow_b200.train.loop import train
    root = str(tmp_path / "ckpt")
    train(_cfg(["--resnet_size=8", "--batch_size=32", "--train_steps=25", f"--log_root={root}",
                "--log_every=1000", "--base_lr=0.05"]))
    ecfg = _cfg(["--resnet_size=8", f"--log_root={root}", f"--eval_dir={tmp_path / 'eval'}", "--eval_once=True",
                 "--eval_batch_count=3"], F.CIFAR_EVAL, "eval")
    r = evaluate(ecfg)
    assert r["global_step"] == 25 and 0.0 <= r["precision"] <= 1.0 and r["best_precision"] >= r["precision"]
    ev = [e for f in glob.glob(str(tmp_path / "eval" / "events.*")) for e in read_events(f)]
    assert any("Best Precision" in s and step == 25 for _, step, s in ev)


def test_evaluator_waits_without_checkpoint(tmp_path):
    from distributed_resnet_tensorflow_b200.train.evaluate import evaluate
    ecfg = _cfg(["--resnet_size=8", f"--log_root={tmp_path / 'none'}", "--eval_once=True"], F.CIFAR_EVAL, "eval")
    assert evaluate(ecfg) == {}


def test_v1_and_imagenet_specs_run_on_torch_backend():
    from distributed_resnet_tensorflow_b200.models.init import init_variables
    from distributed_resnet_tensorflow_b200.models.spec import build_spec, imagenet_resnet
    from distributed_resnet_tensorflow_b200.models.torch_ref import TorchResNet
    for spec, hw in ((build_spec("cifar10", 8, version=1), 32),
                     (imagenet_resnet(18, custom=("bottleneck", [1, 1, 1, 1])), 64)):
        m = TorchResNet(spec)
        m.import_variables(init_variables(spec, 0))
        x = torch.randn(2, hw, hw, 3)
        y = torch.randint(0, spec.num_classes, (2,))
        cost, xent, prec = m.loss_and_grads(x, y, 1e-4)
        assert torch.isfinite(cost)
        m.apply_gradients(0.1, 0.9, 1e-4)
        out = m.export_variables()
        assert set(out) == set(init_variables(spec, 0))


def test_lrnet_plumbing_model(tmp_path):
    from distributed_resnet_tensorflow_b200.train.loop import train
    m = train(_cfg(["--model=lrnet", "--batch_size=32", "--train_steps=40", "--learning_rate=0.01", "--log_every=1000"]))
    assert m["global_step"] == 40 and np.isfinite(m["cost"])


def test_profile_report_lists_params_and_flops():
    from distributed_resnet_tensorflow_b200.models.spec import imagenet_resnet
    from distributed_resnet_tensorflow_b200.train.loop import profile_report
    rep = profile_report(imagenet_resnet(50), 128)
    assert "25551401" in rep and "8.178 GFLOP/image" in rep


def test_resnet_single_train_eval_mode(tmp_path):
    import subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    res = subprocess.run([sys.executable, os.path.join(root, "resnet_single.py"), "--mode=train_eval", "--resnet_size=8",
                          "--batch_size=8", "--train_steps=6", "--eval_every=3", "--eval_batch_count=1",
                          f"--log_root={tmp_path / 'ckpt'}", "--log_every=3"], capture_output=True, text=True, timeout=600)
    out = res.stdout + res.stderr
    assert res.returncode == 0, out[-3000:]
    assert "total trainable parameters" in out and out.count("best precision") == 2 and "step = 6" in out


def test_prefetcher_keeps_order_drops_short_batches_and_forwards_errors():
    """The input thread of the training loop (train/loop.py:_Prefetcher)."""
    import numpy as np
    from distributed_resnet_tensorflow_b200.train.loop import _Prefetcher

    def batches(n, bad_at=None):
        for i in range(n):
            if bad_at is not None and i == bad_at:
                raise ValueError("reader failed")
            size = 4 if i != 2 else 3          # one short batch (end of an epoch)
            yield np.full((size, 2), i), np.arange(size), None

    seen = []
    pf = _Prefetcher(batches(6), 4, lambda b: ("prepared", int(b[0][0, 0])))
    while True:
        item = pf.next()
        if item is None:
            break
        seen.append(item)
    pf.close()
    assert seen == [("prepared", i) for i in (0, 1, 3, 4, 5)]
    pf = _Prefetcher(batches(6, bad_at=1), 4, lambda b: b)
    assert pf.next()[0][0, 0] == 0
    import pytest
    with pytest.raises(ValueError):
        pf.next()
    pf.close()


def test_new_flags_reach_the_config():
    from distributed_resnet_tensorflow_b200 import config as C
    from distributed_resnet_tensorflow_b200.utils import flags as F
    v, _ = F.make_flagset(**F.IMAGENET_MAIN).parse(["--ps_pull=pull", "--input_workers=6", "--allreduce_algo=oneshot_p2p",
                                                    "--resnet_version=1", "--deterministic"])
    cfg = C.from_flags(v)
    assert cfg.train.ps_pull == "pull" and cfg.train.input_workers == 6 and cfg.train.deterministic is True
    assert cfg.train.allreduce_algo == "oneshot_p2p" and cfg.model.resnet_version == 1
    cfg = C.from_flags(F.make_flagset().parse([])[0])
    assert cfg.train.ps_pull == "push" and cfg.train.input_workers == 0 and cfg.train.allreduce_algo == "auto"
    assert cfg.train.deterministic is False
