"""CPU (numpy) restatements of the reference's input preprocessing - the oracles for the device
kernels in csrc/aug.cu and the input path of the torch backend.

CIFAR (resnet_cifar_main.py:185-200 / legacy cifar_input.py:66-75): zero-pad 4 (legacy: 2) pixels per
side, random 32x32 crop, random horizontal flip, per-image standardisation
``(x - mean) / max(std, 1/sqrt(3072))``.
ImageNet (vgg_preprocessing.py:196-363): image in [0,1]; train: aspect-preserving bilinear resize
(align_corners=False) so the short side is uniform in [256, 512], random 224 crop, random flip;
eval: short side 256, central crop; then subtract the channel means 123.68/116.78/103.94 - which the
reference applies to the [0,1]-scaled image (resnet_imagenet_main.py:143), reproduced here as
means/255.
Randomness: Philox4x32-10 keyed by (seed, step), counter = global image index - bit-identical to
the device kernels.
"""
from __future__ import annotations

from typing import Tuple

import numpy as np

R_MEAN, G_MEAN, B_MEAN = 123.68, 116.78, 103.94
VGG_MEANS_01 = np.array([R_MEAN, G_MEAN, B_MEAN], np.float32) / 255.0
RESIZE_SIDE_MIN, RESIZE_SIDE_MAX = 256, 512

_M0, _M1, _W0, _W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85


def philox4x32_10(ctr: Tuple[int, int, int, int], key: Tuple[int, int]) -> Tuple[int, int, int, int]:
    c0, c1, c2, c3 = ctr
    k0, k1 = key
    for _ in range(10):
        p0, p1 = _M0 * c0, _M1 * c2
        hi0, lo0 = p0 >> 32, p0 & 0xFFFFFFFF
        hi1, lo1 = p1 >> 32, p1 & 0xFFFFFFFF
        c0, c1, c2, c3 = (hi1 ^ c1 ^ k0) & 0xFFFFFFFF, lo1, (hi0 ^ c3 ^ k1) & 0xFFFFFFFF, lo0
        k0, k1 = (k0 + _W0) & 0xFFFFFFFF, (k1 + _W1) & 0xFFFFFFFF
    return c0, c1, c2, c3


def cifar_crop_params(seed: int, step: int, index: int, pad: int) -> Tuple[int, int, int]:
    """(oy, ox, flip) exactly as drn_cifar_augment draws them."""
    r = philox4x32_10((index & 0xFFFFFFFF, (index >> 32) & 0xFFFFFFFF, step & 0xFFFFFFFF, 0),
                      (seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF))
    return r[0] % (2 * pad + 1), r[1] % (2 * pad + 1), r[2] & 1


def per_image_standardization(img: np.ndarray) -> np.ndarray:
    x = img.astype(np.float32)
    n = x.size
    mean = x.mean(dtype=np.float64)
    var = max((x.astype(np.float64) ** 2).mean() - mean * mean, 0.0)
    adj = max(np.sqrt(var), 1.0 / np.sqrt(n))
    return ((x - mean) / adj).astype(np.float32)


def cifar_preprocess(record_chw_u8: np.ndarray, training: bool, pad: int = 4, seed: int = 0,
                     step: int = 0, index: int = 0) -> np.ndarray:
    """One CIFAR record (3072 uint8, CHW) -> float32 HWC [32, 32, 3]."""
    img = record_chw_u8.reshape(3, 32, 32).transpose(1, 2, 0).astype(np.float32)
    if training:
        oy, ox, flip = cifar_crop_params(seed, step, index, pad)
        padded = np.zeros((32 + 2 * pad, 32 + 2 * pad, 3), np.float32)
        padded[pad:pad + 32, pad:pad + 32] = img
        img = padded[oy:oy + 32, ox:ox + 32]
        if flip:
            img = img[:, ::-1]
    return per_image_standardization(img)


def cifar_preprocess_batch(records: np.ndarray, training: bool, pad: int = 4, seed: int = 0,
                           step: int = 0, index_base: int = 0) -> np.ndarray:
    return np.stack([cifar_preprocess(records[i], training, pad, seed, step, index_base + i)
                     for i in range(records.shape[0])])


# ------------------------------------------------------------------------------------------------
# VGG-style ImageNet preprocessing
# ------------------------------------------------------------------------------------------------
def smallest_size_at_least(height: int, width: int, smallest_side: int) -> Tuple[int, int]:
    """vgg_preprocessing.py:229-256 (float32 arithmetic, truncation)."""
    h, w, s = np.float32(height), np.float32(width), np.float32(smallest_side)
    scale = s / w if h > w else s / h
    return int(h * scale), int(w * scale)


def resize_bilinear(img: np.ndarray, new_h: int, new_w: int) -> np.ndarray:
    """TF1 ``resize_bilinear(align_corners=False)``: src = dst * (in / out), no half-pixel shift."""
    h, w = img.shape[:2]
    x = img.astype(np.float32)
    ys = np.arange(new_h, dtype=np.float32) * np.float32(h / new_h)
    xs = np.arange(new_w, dtype=np.float32) * np.float32(w / new_w)
    y0 = np.floor(ys).astype(np.int64)
    x0 = np.floor(xs).astype(np.int64)
    y1 = np.minimum(y0 + 1, h - 1)
    x1 = np.minimum(x0 + 1, w - 1)
    ly = (ys - y0)[:, None, None]
    lx = (xs - x0)[None, :, None]
    top = x[y0][:, x0] + (x[y0][:, x1] - x[y0][:, x0]) * lx
    bot = x[y1][:, x0] + (x[y1][:, x1] - x[y1][:, x0]) * lx
    return top + (bot - top) * ly


def central_crop(img: np.ndarray, ch: int, cw: int) -> np.ndarray:
    h, w = img.shape[:2]
    oy, ox = (h - ch) // 2, (w - cw) // 2
    return img[oy:oy + ch, ox:ox + cw]


def vgg_preprocess(img01: np.ndarray, training: bool, out_hw: int = 224,
                   rng: np.random.Generator = None) -> np.ndarray:
    """float [0,1] HWC image of any size -> float32 [out_hw, out_hw, 3], mean-subtracted."""
    h, w = img01.shape[:2]
    if training:
        rng = rng or np.random.default_rng()
        side = int(rng.integers(RESIZE_SIDE_MIN, RESIZE_SIDE_MAX + 1))
        nh, nw = smallest_size_at_least(h, w, side)
        img = resize_bilinear(img01, nh, nw)
        oy = int(rng.integers(0, nh - out_hw + 1))
        ox = int(rng.integers(0, nw - out_hw + 1))
        img = img[oy:oy + out_hw, ox:ox + out_hw]
        if rng.integers(0, 2):
            img = img[:, ::-1]
    else:
        nh, nw = smallest_size_at_least(h, w, RESIZE_SIDE_MIN)
        img = central_crop(resize_bilinear(img01, nh, nw), out_hw, out_hw)
    return (img - VGG_MEANS_01).astype(np.float32)


def vgg_crop_params(h: int, w: int, training: bool, out_hw: int = 224,
                    rng: np.random.Generator = None) -> Tuple[int, int, int, int, int]:
    """(new_h, new_w, off_y, off_x, flip) of the reference's VGG preprocessing for an h x w image
    (vgg_preprocessing.py:284-333): training = random short side in [256, 512], random crop, random
    flip; evaluation = short side 256, central crop.  Shared by the host and the GPU input paths so
    that both make identical decisions from the same generator."""
    if training:
        rng = rng or np.random.default_rng()
        side = int(rng.integers(RESIZE_SIDE_MIN, RESIZE_SIDE_MAX + 1))
        nh, nw = smallest_size_at_least(h, w, side)
        oy = int(rng.integers(0, nh - out_hw + 1))
        ox = int(rng.integers(0, nw - out_hw + 1))
        return nh, nw, oy, ox, int(rng.integers(0, 2))
    nh, nw = smallest_size_at_least(h, w, RESIZE_SIDE_MIN)
    return nh, nw, (nh - out_hw) // 2, (nw - out_hw) // 2, 0


def vgg_resize_crop_u8(img_u8: np.ndarray, training: bool, out_hw: int = 224,
                       rng: np.random.Generator = None) -> Tuple[np.ndarray, int]:
    """Host half of the device pipeline: decoded uint8 HWC -> resized + cropped uint8 [hw, hw, 3]
    plus the flip decision; scaling to [0,1], flip, mean subtraction and the space-to-depth
    layout happen in drn_imagenet_prep.  Quantising the resized image back to uint8 loses at most
    half a grey level (1/510) relative to the reference's float pipeline."""
    h, w = img_u8.shape[:2]
    nh, nw, oy, ox, flip = vgg_crop_params(h, w, training, out_hw, rng)
    img = resize_bilinear(img_u8, nh, nw)[oy:oy + out_hw, ox:ox + out_hw]
    return np.clip(np.rint(img), 0, 255).astype(np.uint8), flip


def imagenet_prep_oracle(u8_hwc: np.ndarray, flip: bool) -> np.ndarray:
    """What drn_imagenet_prep produces for one cropped image, before space-to-depth: float32 HWC."""
    img = u8_hwc.astype(np.float32) * np.float32(1.0 / 255.0)
    if flip:
        img = img[:, ::-1]
    return img - VGG_MEANS_01


def space_to_depth16(img_hwc: np.ndarray) -> np.ndarray:
    """[H, W, 3] -> [H/2, W/2, 16]: channel (dy*2+dx)*3 + c, channels 12..15 zero (the layout the
    4x4/1 re-expression of the 7x7/2 stem convolution consumes)."""
    h, w, _ = img_hwc.shape
    x = img_hwc.reshape(h // 2, 2, w // 2, 2, 3).transpose(0, 2, 1, 3, 4).reshape(h // 2, w // 2, 12)
    return np.concatenate([x, np.zeros((h // 2, w // 2, 4), x.dtype)], axis=2)
