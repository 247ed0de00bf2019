"""Device parameter layouts (no GPU needed): TF <-> device transforms and the space-to-depth
re-expression of the 7x7/2 ImageNet stem."""
import numpy as np
import torch
import torch.nn.functional as F

from distributed_resnet_tensorflow_b200.data.preprocess import space_to_depth16
from distributed_resnet_tensorflow_b200.models.init import init_variables
from distributed_resnet_tensorflow_b200.models.spec import cifar_resnet, imagenet_resnet


def _layout(spec):
    from distributed_resnet_tensorflow_b200.models.engine import ParamLayout
    return ParamLayout(spec)


def test_roundtrip_all_variables():
    for spec in (imagenet_resnet(18), cifar_resnet(8)):
        lay = _layout(spec)
        v = init_variables(spec, seed=2)
        for name, shp in spec.trainable_variables():
            dev = lay.to_dev(name, v[name])
            assert dev.shape == lay.entries[name].dev_shape
            assert np.array_equal(lay.to_tf(name, dev), v[name]), name
        offs = sorted((e.offset, e.numel) for e in lay.entries.values())
        for (o1, n1), (o2, _) in zip(offs, offs[1:]):
            assert o1 + n1 <= o2 and o2 % 64 == 0
        assert lay.total % 1024 == 0


def test_s2d_stem_equals_7x7_stride2_conv():
    spec = imagenet_resnet(18)
    lay = _layout(spec)
    rng = np.random.default_rng(0)
    w = rng.standard_normal((7, 7, 3, 64)).astype(np.float32)
    img = rng.standard_normal((2, 32, 32, 3)).astype(np.float32)
    # reference: explicit pad 3/3 then VALID stride 2 (conv2d_fixed_padding)
    x = torch.from_numpy(img).permute(0, 3, 1, 2)
    ref = F.conv2d(F.pad(x, (3, 3, 3, 3)), torch.from_numpy(w).permute(3, 2, 0, 1), stride=2)
    # s2d formulation: 4x4 stride-1 conv, pad 2 before / 1 after, over [H/2, W/2, 16]
    wd = lay.to_dev("conv2d/kernel", w)                                   # [64, 4, 4, 16]
    xs = torch.from_numpy(np.stack([space_to_depth16(i) for i in img])).permute(0, 3, 1, 2)
    out = F.conv2d(F.pad(xs, (2, 1, 2, 1)), torch.from_numpy(wd).permute(0, 3, 1, 2))
    assert out.shape == ref.shape and torch.allclose(out, ref, atol=1e-4)
    mask = lay.stem_mask()
    assert mask.sum() == 64 * 7 * 7 * 3 and np.all(wd[mask == 0] == 0)


def test_dense_padding_and_classes():
    from distributed_resnet_tensorflow_b200.models.engine import padded_classes
    assert [padded_classes(n) for n in (10, 100, 1000, 1001)] == [16, 128, 1024, 1024]
    lay = _layout(imagenet_resnet(50))
    e = lay.entries["dense/kernel"]
    assert e.dev_shape == (1024, 2048)
    k = np.random.default_rng(0).standard_normal((2048, 1001)).astype(np.float32)
    d = lay.to_dev("dense/kernel", k)
    assert np.all(d[1001:] == 0) and np.array_equal(d[:1001], k.T)
