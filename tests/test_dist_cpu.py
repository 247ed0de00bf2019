"""Multi-process plumbing on the CPU (gloo): the reference's 1-ps / 2-worker local smoke run
(scripts/submit_mac_dist.sh) and the allreduce path, checked against first principles."""
import os
import subprocess
import sys

import numpy as np
import pytest

from distributed_resnet_tensorflow_b200 import config as C
from distributed_resnet_tensorflow_b200.ckpt.saver import Saver, latest_checkpoint
from distributed_resnet_tensorflow_b200.utils import flags as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STEPS, BATCH, SIZE = 3, 8, 8


def _expected_two_replica_weights(seed=0):
    """Two replicas, each on its own batch stream; mean gradient; one momentum-SGD apply per step."""
    from distributed_resnet_tensorflow_b200.models.init import init_variables
    from distributed_resnet_tensorflow_b200.models.spec import build_spec
    from distributed_resnet_tensorflow_b200.train.backends import DistContext, TorchBackend
    from distributed_resnet_tensorflow_b200.train.loop import training_batches
    from distributed_resnet_tensorflow_b200.train.lr import LaggedSchedule
    import torch
    torch.set_num_threads(1)      # bit-for-bit comparable with the single-threaded subprocesses
    cfg = C.from_flags(F.make_flagset().parse([f"--resnet_size={SIZE}", f"--batch_size={BATCH}"])[0])
    spec = build_spec("cifar10", SIZE)
    reps, its = [], []
    for i in range(2):
        ctx = DistContext(i, 1, 2, 0, "worker", i, "serial")
        be = TorchBackend(spec, BATCH, ctx, weight_decay=2e-4, seed=seed)
        be.import_variables(init_variables(spec, seed))
        reps.append(be)
        its.append(training_batches(cfg, spec, ctx))
    sched = LaggedSchedule("cifar10")
    for step in range(STEPS):
        lr = sched.before_step()
        grads = []
        for be, it in zip(reps, its):
            imgs, labs, _ = next(it)
            x = be._prep(imgs, True, None)
            be.model.loss_and_grads(x, torch.from_numpy(labs).long(), 2e-4)
            grads.append(be.model.flat_grads())
            be.steps_done += 1
        mean = (grads[0] + grads[1]) / 2
        for be in reps:
            be.model.set_flat_grads(mean)
            be.model.apply_gradients(lr, 0.9, 2e-4)
        sched.after_step(step + 1)
    return reps[0].model.export_variables()


def _run(cmds, timeout=600):
    procs = [subprocess.Popen(c, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for c in cmds]
    outs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=timeout)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append(out)
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o[-3000:]
    return outs


def _compare(got, want, with_momentum=True):
    for k, v in want.items():
        if k == "global_step" or (k.endswith("/Momentum") and not with_momentum):
            continue
        if "moving_" in k:
            continue        # BN moving statistics are per replica
        assert np.allclose(got[k], v, rtol=2e-4, atol=2e-5), k


@pytest.mark.slow
def test_one_ps_two_workers_sync_replicas(tmp_path):
    root = str(tmp_path / "ckpt")
    port = 23000 + os.getpid() % 2000
    common = [f"--ps_hosts=localhost:{port + 500}", f"--worker_hosts=localhost:{port},localhost:{port + 1}",
              "--sync_replicas=True", f"--resnet_size={SIZE}", f"--batch_size={BATCH}", f"--train_steps={STEPS}",
              "--num_gpus=0", f"--log_root={root}", "--log_every=1", "--backend=torch", "--num_intra_threads=1"]
    py = [sys.executable, os.path.join(ROOT, "resnet_cifar_main.py")]
    outs = _run([py + ["--job_name=ps", "--task_index=0"] + common,
                 py + ["--job_name=worker", "--task_index=0"] + common,
                 py + ["--job_name=worker", "--task_index=1"] + common])
    assert "role=ps" in outs[0] and "step = 3" in outs[1] and "step = 3" in outs[2]
    got = Saver.restore(latest_checkpoint(root))
    assert int(got["global_step"]) == STEPS
    _compare(got, _expected_two_replica_weights())


@pytest.mark.slow
def test_parameter_server_mode_saves_time_based_checkpoints(tmp_path):
    """MonitoredTrainingSession saves every save_checkpoint_secs in PS mode too (resnet_cifar_main.py:329);
    with sharded state the save is a collective, so the chief's clock is broadcast to the workers."""
    root = str(tmp_path / "ckpt")
    port = 27000 + os.getpid() % 2000
    common = [f"--ps_hosts=localhost:{port + 500}", f"--worker_hosts=localhost:{port},localhost:{port + 1}",
              "--sync_replicas=True", f"--resnet_size={SIZE}", f"--batch_size={BATCH}", "--train_steps=90",
              "--num_gpus=0", f"--log_root={root}", "--log_every=2", "--backend=torch", "--num_intra_threads=1",
              "--save_checkpoint_secs=1"]
    py = [sys.executable, os.path.join(ROOT, "resnet_cifar_main.py")]
    outs = _run([py + ["--job_name=ps", "--task_index=0"] + common,
                 py + ["--job_name=worker", "--task_index=0"] + common,
                 py + ["--job_name=worker", "--task_index=1"] + common])
    saved = [l for l in outs[1].splitlines() if "saved checkpoint" in l]
    assert len(saved) >= 2, outs[1][-2000:]          # at least one timed save before the final one
    assert "saved checkpoint" not in outs[2]         # only the chief writes
    assert int(Saver.restore(latest_checkpoint(root))["global_step"]) == 90


@pytest.mark.slow
def test_two_rank_allreduce_matches_parameter_server(tmp_path):
    root = str(tmp_path / "ckpt")
    port = 25000 + os.getpid() % 2000
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "resnet_cifar_main_horovod.py"),
           "--use_horovod=True", f"--resnet_size={SIZE}", f"--batch_size={BATCH}", f"--train_steps={STEPS}",
           "--num_gpus=0", f"--log_root={root}", "--log_every=1", "--backend=torch", "--num_intra_threads=1"]
    _run([cmd])
    got = Saver.restore(latest_checkpoint(root))
    assert int(got["global_step"]) == STEPS
    _compare(got, _expected_two_replica_weights())


def test_context_mapping_for_ps_and_workers():
    from distributed_resnet_tensorflow_b200.train.loop import make_context
    base = ["--sync_replicas=True", "--ps_hosts=a:1,b:2", "--worker_hosts=c:1,d:2,e:3"]
    wk = C.from_flags(F.make_flagset().parse(base + ["--job_name=worker", "--task_index=2"])[0])
    ps = C.from_flags(F.make_flagset().parse(base + ["--job_name=ps", "--task_index=1"])[0])
    c = make_context(wk, native=False)
    assert (c.rank, c.world, c.num_workers, c.num_ps, c.role, c.mode) == (2, 5, 3, 2, "worker", "ps")
    c = make_context(ps, native=False)
    assert (c.rank, c.world, c.role) == (4, 5, "ps")
    c = make_context(wk, native=True)      # GPU mode: shards live on the workers' GPUs
    assert (c.rank, c.world, c.num_ps, c.role) == (2, 3, 2, "worker")
    assert make_context(ps, native=True).rank == -1
