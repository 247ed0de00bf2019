from distributed_resnet_tensorflow_b200.train.lr import LaggedSchedule, cifar_lr, imagenet_lr


def test_cifar_table():
    assert cifar_lr(0) == 0.1 and cifar_lr(39999) == 0.1
    assert abs(cifar_lr(40000) - 0.01) < 1e-12 and abs(cifar_lr(59999) - 0.01) < 1e-12
    assert abs(cifar_lr(60000) - 0.001) < 1e-12 and abs(cifar_lr(80000) - 0.0001) < 1e-12


def test_imagenet_table():
    assert abs(imagenet_lr(0) - 0.1) < 1e-12
    assert abs(imagenet_lr(3120) - 0.25) < 1e-9
    assert abs(imagenet_lr(6240) - 0.4) < 1e-12 and abs(imagenet_lr(37439) - 0.4) < 1e-12
    assert abs(imagenet_lr(37440) - 0.04) < 1e-12 and abs(imagenet_lr(74880) - 0.004) < 1e-12
    assert abs(imagenet_lr(99840) - 0.0004) < 1e-12


def test_one_step_lag_and_start_values():
    s = LaggedSchedule("cifar10")
    assert s.before_step() == 0.1
    s.after_step(40000)
    assert abs(s.before_step() - 0.01) < 1e-12
    s = LaggedSchedule("imagenet")
    assert s.before_step() == 0.4          # `begin()` of the reference hook
    s.after_step(1)
    assert abs(s.before_step() - (0.1 + 0.3 / 6240.0)) < 1e-9


def test_ps_compat_constant():
    s = LaggedSchedule("cifar10", constant=True)
    s.after_step(90000)
    assert s.before_step() == 0.1
