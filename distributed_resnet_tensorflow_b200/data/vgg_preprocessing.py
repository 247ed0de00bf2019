"""The reference's ``vgg_preprocessing`` module surface (vgg_preprocessing.py:336-363) on numpy
arrays: ``preprocess_image(image, output_height, output_width, is_training)`` and the two
branches.  Images are float HWC in [0,1] as produced by ``convert_image_dtype`` in the reference's
record parser (resnet_imagenet_main.py:143); the channel means are therefore applied in [0,1] units.
The implementation lives in data/preprocess.py (which also serves as the oracle of the device
kernels)."""
from __future__ import annotations

import numpy as np

from . import preprocess as _p

_R_MEAN, _G_MEAN, _B_MEAN = _p.R_MEAN / 255.0, _p.G_MEAN / 255.0, _p.B_MEAN / 255.0
_RESIZE_SIDE_MIN, _RESIZE_SIDE_MAX = _p.RESIZE_SIDE_MIN, _p.RESIZE_SIDE_MAX


def preprocess_for_train(image, output_height, output_width, resize_side_min=_RESIZE_SIDE_MIN,
                         resize_side_max=_RESIZE_SIDE_MAX, rng=None):
    assert output_height == output_width
    rng = rng or np.random.default_rng()
    h, w = image.shape[:2]
    side = int(rng.integers(resize_side_min, resize_side_max + 1))
    nh, nw = _p.smallest_size_at_least(h, w, side)
    img = _p.resize_bilinear(image, nh, nw)
    oy = int(rng.integers(0, nh - output_height + 1))
    ox = int(rng.integers(0, nw - output_width + 1))
    img = img[oy:oy + output_height, ox:ox + output_width]
    if rng.integers(0, 2):
        img = img[:, ::-1]
    return (img - _p.VGG_MEANS_01).astype(np.float32)


def preprocess_for_eval(image, output_height, output_width, resize_side=_RESIZE_SIDE_MIN):
    h, w = image.shape[:2]
    nh, nw = _p.smallest_size_at_least(h, w, resize_side)
    img = _p.central_crop(_p.resize_bilinear(image, nh, nw), output_height, output_width)
    return (img - _p.VGG_MEANS_01).astype(np.float32)


def preprocess_image(image, output_height, output_width, is_training=False,
                     resize_side_min=_RESIZE_SIDE_MIN, resize_side_max=_RESIZE_SIDE_MAX, rng=None):
    if is_training:
        return preprocess_for_train(image, output_height, output_width, resize_side_min, resize_side_max, rng)
    return preprocess_for_eval(image, output_height, output_width, resize_side_min)
