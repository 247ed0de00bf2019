"""Command-line entry points on one GPU with the native backend: train -> checkpoint -> separate
eval process -> resume; and cross-backend checkpoint compatibility."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, timeout=900):
    res = subprocess.run([sys.executable] + args, cwd=ROOT, capture_output=True, text=True, timeout=timeout)
    assert res.returncode == 0, (res.stdout + res.stderr)[-4000:]
    return res.stdout + res.stderr


def test_cifar_train_eval_resume_native(tmp_path):
    from distributed_resnet_tensorflow_b200.ckpt.saver import Saver, latest_checkpoint
    root = str(tmp_path / "ckpt")
    common = ["--resnet_size=14", "--batch_size=64", "--num_gpus=1", f"--log_root={root}",
              f"--eval_dir={tmp_path / 'ev'}", "--log_every=20", "--base_lr=0.05"]
    out = _run([os.path.join(ROOT, "resnet_cifar_main.py"), "--train_steps=80"] + common)
    assert "backend=native" in out and "step = 80" in out
    train_prec = float(out.split("step = 80,")[1].split("precision = ")[1].split(",")[0])
    assert train_prec > 0.5, out[-800:]     # learnable synthetic data: the native path really trains
    v = Saver.restore(latest_checkpoint(root))
    assert int(v["global_step"]) == 80 and len(v) == 2 * 44 + 2 * 13 + 1
    ev = _run([os.path.join(ROOT, "resnet_cifar_eval.py"), "--resnet_size=14", "--num_gpus=1", f"--log_root={root}",
               f"--eval_dir={tmp_path / 'ev2'}", "--eval_once=True", "--eval_batch_count=5"])
    assert "best precision" in ev
    prec = float(ev.split("precision:")[1].split(",")[0])
    loss = float(ev.split("loss:")[1].split(",")[0])
    # (after 80 steps the 0.997-decay moving statistics are far from converged, so inference-mode
    #  precision is low by construction - exactly as it would be in the reference)
    # the same checkpoint evaluated by the PyTorch backend agrees (TF-name / layout map is consistent)
    ev_t = _run([os.path.join(ROOT, "resnet_cifar_eval.py"), "--resnet_size=14", "--num_gpus=1", "--backend=torch",
                 "--dtype=fp32", f"--log_root={root}", "--eval_once=True", "--eval_batch_count=5"])
    prec_t = float(ev_t.split("precision:")[1].split(",")[0])
    loss_t = float(ev_t.split("loss:")[1].split(",")[0])
    assert abs(prec - prec_t) < 0.05 and abs(loss - loss_t) < 0.05 * max(1.0, loss_t), (prec, prec_t, loss, loss_t)
    out2 = _run([os.path.join(ROOT, "resnet_cifar_main.py"), "--train_steps=100"] + common)
    assert "restored" in out2 and "step = 100" in out2


def test_v1_networks_train_and_evaluate_through_the_cli(tmp_path):
    """--resnet_version=1 (post-activation) on the native backend: train, checkpoint, evaluate."""
    root = str(tmp_path / "ckpt")
    out = _run([os.path.join(ROOT, "resnet_cifar_main.py"), "--resnet_version=1", "--resnet_size=14", "--batch_size=64",
                "--num_gpus=1", f"--log_root={root}", "--train_steps=80", "--log_every=20", "--base_lr=0.05"])
    assert "backend=native" in out and "step = 80" in out
    prec = float(out.split("step = 80,")[1].split("precision = ")[1].split(",")[0])
    assert prec > 0.5, out[-800:]
    ev = _run([os.path.join(ROOT, "resnet_cifar_eval.py"), "--resnet_version=1", "--resnet_size=14", "--num_gpus=1",
               f"--log_root={root}", "--eval_once=True", "--eval_batch_count=5"])
    assert "best precision" in ev


def test_imagenet_main_serial_native(tmp_path):
    out = _run([os.path.join(ROOT, "resnet_imagenet_main.py"), "--resnet_size=18", "--batch_size=16", "--num_gpus=1",
                "--train_steps=12", "--log_every=4", f"--log_root={tmp_path / 'ckpt'}", f"--train_dir={tmp_path / 'tr'}"])
    assert "backend=native" in out and "step = 12" in out


@pytest.mark.parametrize("decoder", ["native", "torchvision"])
def test_gpu_jpeg_input_path_matches_the_host_path(tmp_path, decoder):
    """TFRecord shards of synthetic JPEGs through both ImageNet input paths: PIL decode + numpy
    resize/crop on the host vs nvJPEG batch decode + drn_resize_crop_batch on the device.  Both
    draw the same crop decisions; the decoders may differ by a grey level or two."""
    import io
    import numpy as np
    import torch
    from PIL import Image
    from distributed_resnet_tensorflow_b200.data import imagenet as im
    from distributed_resnet_tensorflow_b200.data.tfrecord import make_example, write_tfrecords
    if decoder == "torchvision":
        pytest.importorskip("torchvision")
    else:
        from distributed_resnet_tensorflow_b200.data import jpeg_native
        assert jpeg_native.available(), "libdrn_b200.so was built without the nvJPEG binding"
    rng = np.random.default_rng(0)
    records = []
    for i in range(12):
        h, w = int(rng.integers(260, 420)), int(rng.integers(260, 520))
        yy, xx = np.mgrid[0:h, 0:w]
        img = np.stack([(yy * 255 // h), (xx * 255 // w), ((yy + xx) * 255 // (h + w))], 2).astype(np.uint8)
        img = (img.astype(np.int32) + rng.integers(-6, 7, img.shape)).clip(0, 255).astype(np.uint8)
        if i == 5:
            img = img[:, :, 0]            # one greyscale JPEG, as ImageNet has
        buf = io.BytesIO()
        Image.fromarray(img).save(buf, format="JPEG", quality=92)
        records.append(make_example({"image/encoded": [buf.getvalue()], "image/class/label": [i + 1]}))
    shard = str(tmp_path / "validation-00000-of-00128")
    write_tfrecords(shard, records)
    for training in (False, True):
        host = list(im.ImageNetBatches(str(tmp_path), 4, training, 1, seed=3, files=[shard]))
        it = im.ImageNetBatchesGPU(str(tmp_path), 4, training, 1, seed=3, files=[shard], device="cuda:0",
                                   decoder=decoder)
        assert (it.native is not None) == (decoder == "native"), it.decoder_name
        dev = list(it)
        assert len(host) == len(dev) == 3
        for (hi, hl, hc), (di, dl, dc) in zip(host, dev):
            assert di.is_cuda and di.dtype == torch.uint8 and tuple(di.shape) == hi.shape
            assert np.array_equal(hl, dl) and np.array_equal(hc, dc)
            diff = np.abs(di.cpu().numpy().astype(np.int32) - hi.astype(np.int32))
            assert diff.mean() < 1.5 and np.percentile(diff, 99) <= 8, (diff.mean(), diff.max())


@pytest.mark.parametrize("pull", ["push", "pull"])
def test_two_worker_parameter_server_cli_saves_timed_checkpoints(tmp_path, pull):
    """resnet_cifar_main.py as two --job_name=worker processes on two GPUs (native backend: parameter
    shard i lives on worker i's GPU), both weight paths, with the reference's time-based checkpoint
    hook (every worker joins the collective save the chief's clock asks for) and an evaluation of
    the result by a separate process."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from distributed_resnet_tensorflow_b200.ckpt.saver import Saver, latest_checkpoint
    root = str(tmp_path / "ckpt")
    port = 24000 + os.getpid() % 2000
    common = [os.path.join(ROOT, "resnet_cifar_main.py"), "--job_name=worker", "--sync_replicas=True",
              f"--ps_hosts=localhost:{port + 500},localhost:{port + 501}",
              f"--worker_hosts=localhost:{port},localhost:{port + 1}", "--resnet_size=8", "--batch_size=32",
              "--num_gpus=2", f"--log_root={root}", "--log_every=500", "--train_steps=9000", "--base_lr=0.05",
              "--save_checkpoint_secs=1", f"--ps_pull={pull}", f"--eval_dir={tmp_path / 'ev'}"]
    procs = [subprocess.Popen([sys.executable] + common + [f"--task_index={i}"], cwd=ROOT, stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for i in range(2)]
    outs = [p.communicate(timeout=600)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o[-3000:]
    assert "mode=ps backend=native" in outs[0] and "step = 9000" in outs[0] and "step = 9000" in outs[1]
    assert outs[0].count("saved checkpoint") >= 2, outs[0][-2000:]
    v = Saver.restore(latest_checkpoint(root))
    assert int(v["global_step"]) == 9000
    prec = float(outs[0].split("step = 9000,")[1].split("precision = ")[1].split(",")[0])
    assert prec > 0.5, outs[0][-800:]


def test_deterministic_flag_reproduces_a_run_bit_for_bit(tmp_path):
    """--deterministic: two training runs from the same seed write identical checkpoints (weights,
    momentum, moving statistics), through the whole product path - input thread, CUDA graph, side
    stream for the filter gradients, bucketed optimizer."""
    from distributed_resnet_tensorflow_b200.ckpt.saver import Saver, latest_checkpoint
    got = []
    for run in ("a", "b"):
        root = str(tmp_path / run)
        out = _run([os.path.join(ROOT, "resnet_cifar_main.py"), "--resnet_size=14", "--batch_size=64", "--num_gpus=1",
                    f"--log_root={root}", "--train_steps=40", "--log_every=20", "--base_lr=0.05", "--deterministic"])
        assert "backend=native" in out and "step = 40" in out
        got.append(Saver.restore(latest_checkpoint(root)))
    a, b = got
    assert a.keys() == b.keys()
    diff = [k for k in a if not np.array_equal(np.asarray(a[k]), np.asarray(b[k]))]
    assert not diff, f"{len(diff)} of {len(a)} variables differ, e.g. {diff[:5]}"
