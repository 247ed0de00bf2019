"""Which BatchNorm+ReLU applications become prologues of their consuming convolution (models/engine.py).
Pure policy, no GPU: the default is the measured rule (first 1x1 convolution of the 256/512-channel
bottlenecks), DRN_FUSE_PRE=all fuses whatever a prologue variant exists for."""
import importlib


def _engine(monkeypatch, **env):
    for k in ("DRN_FUSE_PRE", "DRN_FUSE_BN0", "DRN_FUSE_BN2", "DRN_FUSE_BN0_MAXC"):
        monkeypatch.delenv(k, raising=False)
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    import distributed_resnet_tensorflow_b200.models.engine as eng
    return importlib.reload(eng)


def test_prologue_variants_exist_for_the_tiles_the_networks_use(monkeypatch):
    eng = _engine(monkeypatch)
    ok = eng._pre_supported
    assert ok(16, 16) and ok(16, 32) and ok(32, 32) and ok(32, 64) and ok(64, 64) and ok(256, 64) and ok(512, 2048)
    assert not ok(16, 64) and not ok(32, 16) and not ok(48, 64) and not ok(64, 16)


def test_default_policy_is_the_measured_rule(monkeypatch):
    from distributed_resnet_tensorflow_b200.models.spec import build_spec
    eng = _engine(monkeypatch)
    spec = build_spec("imagenet", 50)
    fused = [eng._fuse_bn0(b) for b in spec.blocks]
    # stage 1: blocks 2-3 (256 channels in, no projection); stage 2: block 1 (256 in, projection) and 2-4 (512 in)
    assert fused == [False, True, True, True, True, True, True] + [False] * 9
    assert not any(eng._fuse_inner(b, i) for b in spec.blocks for i in range(1, len(b.convs)))
    cifar = build_spec("cifar10", 50)
    assert not any(eng._fuse_bn0(b, True) for b in cifar.blocks)      # measured slower on CIFAR-sized steps
    assert eng.FIN_IN_CONV is False


def test_fuse_all_and_none(monkeypatch):
    from distributed_resnet_tensorflow_b200.models.spec import build_spec
    eng = _engine(monkeypatch, DRN_FUSE_PRE="all")
    for spec in (build_spec("imagenet", 50), build_spec("cifar10", 20), build_spec("imagenet", 18)):
        assert all(eng._fuse_bn0(b) for b in spec.blocks)
        assert all(eng._fuse_inner(b, i) for b in spec.blocks for i in range(1, len(b.convs)))
    eng = _engine(monkeypatch, DRN_FUSE_PRE="none")
    assert not any(eng._fuse_bn0(b) for b in build_spec("imagenet", 50).blocks)
    eng = _engine(monkeypatch, DRN_FUSE_BN2="128")
    spec = build_spec("imagenet", 50)
    got = [eng._fuse_inner(b, 2) for b in spec.blocks]
    assert got == [True] * 7 + [False] * 9          # last 1x1 convolution of the 64- and 128-channel bottlenecks
    _engine(monkeypatch)                               # leave the module in its default state
