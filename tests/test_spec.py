import numpy as np
import pytest

from distributed_resnet_tensorflow_b200.models.init import init_variables
from distributed_resnet_tensorflow_b200.models.spec import build_spec, cifar_resnet, imagenet_resnet


def test_variable_counts_match_survey_appendix_a():
    s = imagenet_resnet(50)
    assert len(s.trainable_variables()) == 153 and len(s.moving_variables()) == 98
    assert len(init_variables(s)) == 405 and s.num_params() == 25551401
    c = cifar_resnet(50)
    assert len(c.trainable_variables()) == 152 and len(init_variables(c)) == 403
    assert c.num_params() == 758618


def test_creation_order_names():
    names = [n for n, _ in imagenet_resnet(50).trainable_variables()]
    assert names[:8] == ["conv2d/kernel", "batch_normalization/gamma", "batch_normalization/beta",
                         "conv2d_1/kernel", "conv2d_2/kernel", "batch_normalization_1/gamma",
                         "batch_normalization_1/beta", "conv2d_3/kernel"]
    assert names[-4:] == ["batch_normalization_48/gamma", "batch_normalization_48/beta", "dense/kernel", "dense/bias"]
    shapes = dict(imagenet_resnet(50).trainable_variables())
    assert shapes["conv2d/kernel"] == (7, 7, 3, 64) and shapes["conv2d_1/kernel"] == (1, 1, 64, 256)
    assert shapes["conv2d_52/kernel"] == (1, 1, 512, 2048) and shapes["dense/kernel"] == (2048, 1001)
    cshapes = dict(cifar_resnet(50).trainable_variables())
    assert cshapes["conv2d/kernel"] == (3, 3, 3, 16) and cshapes["conv2d_1/kernel"] == (1, 1, 16, 16)
    assert cshapes["conv2d_51/kernel"] == (3, 3, 64, 64) and cshapes["dense/kernel"] == (64, 10)


def test_conv_shape_table():
    convs = imagenet_resnet(50).convs()
    assert len(convs) == 53
    strided = [(c.cin, c.cout, c.ksize) for c in convs if c.stride == 2]
    assert strided == [(3, 64, 7), (256, 512, 1), (128, 128, 3), (512, 1024, 1), (256, 256, 3), (1024, 2048, 1), (512, 512, 3)]
    assert abs(2 * imagenet_resnet(50).forward_macs_per_image() / 1e9 - 8.18) < 0.01
    assert len(cifar_resnet(50).convs()) == 52


@pytest.mark.parametrize("size", [18, 34, 101, 152, 200])
def test_other_sizes_build(size):
    from distributed_resnet_tensorflow_b200.models.spec import IMAGENET_BLOCKS
    s = imagenet_resnet(size)
    kind, layers = IMAGENET_BLOCKS[size]
    per = 2 if kind == "basic" else 3
    nb = sum(layers)
    assert s.dense_in == (512 if kind == "basic" else 2048)
    assert len(s.convs()) == 1 + per * nb + 4
    assert len(s.bns()) == per * nb + 1
    assert len(s.trainable_variables()) == len(s.convs()) + 2 * len(s.bns()) + 2


def test_invalid_sizes():
    with pytest.raises(ValueError):
        cifar_resnet(51)
    with pytest.raises(ValueError):
        imagenet_resnet(49)


def test_v1_and_cifar100():
    s = build_spec("cifar100", 20, version=1)
    assert s.num_classes == 100 and s.final_bn is None and s.stem_bn is not None
    assert len(s.bns()) == 1 + 3 * 3 * 2 + 3   # stem + 2 per block + 3 projection BNs


def test_init_statistics():
    v = init_variables(cifar_resnet(8), seed=1)
    k = v["conv2d_2/kernel"]
    assert abs(k.std() - np.sqrt(1.0 / (3 * 3 * 16))) < 0.02 and np.abs(k).max() <= 2.0 * np.sqrt(1.0 / 144) + 1e-6
    assert np.all(v["batch_normalization/gamma"] == 1) and np.all(v["batch_normalization/moving_variance"] == 1)
    lim = np.sqrt(6.0 / (64 + 10))
    assert np.abs(v["dense/kernel"]).max() <= lim and v["global_step"] == 0
    assert np.all(v["conv2d/kernel/Momentum"] == 0)
