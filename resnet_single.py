#!/usr/bin/env python
"""resnet_single - single-process train + eval, the capability of the reference's (stale)
resnet_single.py:51-213: ``--mode=train`` trains serially and prints the tfprof-style
parameter/FLOP table first (resnet_single.py:58-66); ``--mode=eval`` runs the checkpoint-polling
evaluator; ``--mode=train_eval`` alternates: train ``--eval_every`` steps, evaluate once, repeat."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from distributed_resnet_tensorflow_b200 import config as C  # noqa: E402
from distributed_resnet_tensorflow_b200.cli import _setup_logging  # noqa: E402
from distributed_resnet_tensorflow_b200.utils import flags as F  # noqa: E402


def main(argv):
    _setup_logging()
    fs = F.make_flagset(profile=True)
    fs.define(F.FlagDef("eval_every", "integer", 1000, "train_eval mode: steps between evaluations."))
    values, _ = fs.parse(argv)
    if values.job_name is not None:
        raise SystemExit("resnet_single is the single-process entry point; use resnet_cifar_main for --job_name runs")
    from distributed_resnet_tensorflow_b200.train.evaluate import evaluate
    from distributed_resnet_tensorflow_b200.train.loop import train
    if values.mode == "train":
        return train(C.from_flags(values))
    if values.mode == "eval":
        return evaluate(C.from_flags(values, entry="eval"))
    if values.mode != "train_eval":
        raise SystemExit("--mode must be train, eval or train_eval")
    total, done, result = values.train_steps, 0, {}
    while done < total:
        done = min(done + values.eval_every, total)
        train(C.from_flags(values.replace(train_steps=done, profile=False if done > values.eval_every else values.profile)))
        result = evaluate(C.from_flags(values.replace(eval_once=True), entry="eval"))
    return result


if __name__ == "__main__":
    main(sys.argv[1:])
