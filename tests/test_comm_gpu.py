"""In-kernel NVLink collectives and the two data-parallel transports on >= 2 GPUs."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _launch(nproc, env_extra=None, port=29611):
    env = dict(os.environ)
    env.update(env_extra or {})
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(ROOT, "tests", "_comm_worker.py")]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    out = res.stdout + res.stderr
    sys.stdout.write("\n".join(l for l in out.splitlines() if "allreduce" in l or "steps ok" in l or "deterministic" in l) + "\n")
    assert res.returncode == 0 and "COMM_WORKER_OK" in out, out[-4000:]
    return out


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_collectives_and_transports_all_gpus():
    _launch(min(torch.cuda.device_count(), 8))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_collectives_ipc_provider_two_gpus():
    _launch(2, {"DRN_SYMM": "ipc"}, port=29612)
