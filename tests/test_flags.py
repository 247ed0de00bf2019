import pytest

from distributed_resnet_tensorflow_b200 import config as C
from distributed_resnet_tensorflow_b200.utils import flags as F


def test_reference_flag_surface_present():
    fs = F.make_flagset()
    for name in ["dataset", "mode", "train_data_path", "eval_data_path", "image_size", "train_dir", "eval_dir",
                 "eval_batch_count", "eval_once", "log_root", "num_gpus", "task_index", "replicas_to_aggregate",
                 "train_steps", "num_epochs", "batch_size", "learning_rate", "sync_replicas", "existing_servers",
                 "ps_hosts", "worker_hosts", "job_name", "data_format", "num_intra_threads", "num_inter_threads",
                 "use_horovod", "hidden_units", "resnet_size"]:
        assert name in fs, name


def test_defaults_match_reference():
    v, _ = F.make_flagset(**F.CIFAR_MAIN).parse([])
    assert (v.dataset, v.batch_size, v.train_steps, v.data_format) == ("cifar10", 32, 2000, "channels_first")
    assert v.ps_hosts == "localhost:2222" and v.worker_hosts == "localhost:2223,localhost:2224"
    assert v.job_name is None and v.task_index is None and v.eval_batch_count == 50
    v, _ = F.make_flagset(**F.IMAGENET_MAIN).parse([])
    assert (v.dataset, v.batch_size, v.train_steps, v.num_epochs) == ("imagenet", 128, 200, 90)
    v, _ = F.make_flagset(**F.IMAGENET_EVAL).parse([])
    assert (v.mode, v.num_epochs) == ("eval", 3000)


def test_boolean_syntaxes():
    fs = F.make_flagset()
    for argv, want in [(["--sync_replicas=True"], True), (["--sync_replicas=False"], False),
                       (["--sync_replicas"], True), (["--nosync_replicas"], False),
                       (["--sync_replicas=true"], True), (["--sync_replicas=0"], False)]:
        assert fs.parse(argv)[0].sync_replicas is want, argv
    with pytest.raises(F.FlagError):
        fs.parse(["--sync_replicas=maybe"])


def test_value_forms_and_errors():
    fs = F.make_flagset()
    v, rest = fs.parse(["--batch_size", "64", "--job_name=worker", "--task_index=1", "pos"])
    assert v.batch_size == 64 and v.job_name == "worker" and v.task_index == 1 and rest == ["pos"]
    assert v.is_set("batch_size") and not v.is_set("train_steps")
    with pytest.raises(F.FlagError):
        fs.parse(["--no_such_flag=1"])
    with pytest.raises(F.FlagError):
        fs.parse(["--batch_size=abc"])


def test_world_size_from_host_lists():
    v, _ = F.make_flagset().parse(["--job_name=worker", "--task_index=0", "--sync_replicas=True",
                                   "--ps_hosts=a:1,b:2", "--worker_hosts=c:1,d:2,e:3,f:4"])
    cfg = C.from_flags(v)
    assert cfg.cluster.num_ps == 2 and cfg.cluster.num_workers == 4 and cfg.cluster.is_chief
    v, _ = F.make_flagset().parse(["--num_ps=8", "--num_workers=8"])
    cfg = C.from_flags(v)
    assert cfg.cluster.num_ps == 8 and cfg.cluster.num_workers == 8


def test_config_defaults_and_validation():
    cfg = C.from_flags(F.make_flagset(**F.CIFAR_MAIN).parse([])[0])
    assert cfg.train.weight_decay == 2e-4 and cfg.train.log_every == 20 and cfg.train.save_checkpoint_secs == 60
    assert cfg.model.num_classes == 10 and cfg.eval.eval_batch_size == 100
    cfg = C.from_flags(F.make_flagset(**F.IMAGENET_MAIN).parse([])[0])
    assert cfg.train.weight_decay == 1e-4 and cfg.train.log_every == 40 and cfg.train.save_checkpoint_secs == 600
    assert cfg.model.num_classes == 1001
    assert C.from_flags(F.make_flagset().parse(["--dataset=cifar100"])[0]).model.num_classes == 100
    with pytest.raises(ValueError):  # async PS is not a working mode of the reference
        C.from_flags(F.make_flagset().parse(["--job_name=worker", "--task_index=0"])[0])
    with pytest.raises(ValueError):
        C.from_flags(F.make_flagset().parse(["--job_name=chief", "--task_index=0", "--sync_replicas=True"])[0])


def test_input_decode_flag_reaches_the_config_and_defaults_to_host_path():
    from distributed_resnet_tensorflow_b200.train.loop import use_gpu_decode
    v, _ = F.make_flagset(**F.IMAGENET_MAIN).parse([])
    assert v.input_decode == "auto"
    cfg = C.from_flags(v)
    assert cfg.train.input_decode == "auto"
    assert use_gpu_decode(cfg, "cuda:0") is False          # auto stays on the host path (measured equal)
    assert use_gpu_decode(cfg, None) is False
    v, _ = F.make_flagset(**F.IMAGENET_MAIN).parse(["--input_decode=cpu"])
    assert use_gpu_decode(C.from_flags(v), "cuda:0") is False
