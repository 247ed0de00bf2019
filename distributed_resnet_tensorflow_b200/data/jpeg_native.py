"""Batched GPU JPEG decoding through the library's own nvJPEG binding (csrc/jpeg.cu).

The reference decodes JPEGs on the host inside tf.data (``tf.image.decode_jpeg`` on five map
threads, resnet_imagenet_main.py:138-168).  Here a whole batch of encoded byte strings goes to the
GPU: the hardware JPEG engine when the device has one, nvJPEG's GPU-hybrid backend otherwise; the
output is planar RGB in device memory, which ``drn_resize_crop_batch`` turns into the 224x224 crops.
"""
from __future__ import annotations

import ctypes
from typing import List, Sequence, Tuple

import numpy as np
import torch

from .._native import lib as _nl

BACKEND_NAMES = {2: "hardware JPEG engine", 1: "nvJPEG GPU hybrid", 0: "nvJPEG default"}


class NativeJpegDecoder:
    def __init__(self, device="cuda:0", prefer_hardware: bool = True):
        self.device = torch.device(device)
        lib = _nl.load()
        h = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            _nl.check(lib.drn_jpeg_create(1 if prefer_hardware else 0, ctypes.byref(h)), "drn_jpeg_create")
        self._h = h
        self.backend = BACKEND_NAMES.get(int(lib.drn_jpeg_backend(h)), "?")
        self._arena = None      # one growing device buffer for the decoded planes of a batch

    def close(self) -> None:
        if self._h:
            _nl.load().drn_jpeg_destroy(self._h)
            self._h = None

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    def decode(self, blobs: Sequence[bytes]) -> Tuple[torch.Tensor, np.ndarray, np.ndarray, np.ndarray]:
        """Returns (arena uint8 CUDA, offsets int64 [B], wh int32 [B,2], status int32 [B]): image i is
        ``arena[offsets[i] : offsets[i] + 3*h*w]`` viewed as [3, h, w]; status != 0 = decode it on the host.
        The arena is reused by the next call."""
        lib = _nl.load()
        B = len(blobs)
        keep = [np.frombuffer(b, dtype=np.uint8) for b in blobs]
        ptrs = (ctypes.c_void_p * B)(*[k.ctypes.data for k in keep])
        lens = (ctypes.c_size_t * B)(*[len(b) for b in blobs])
        wh = np.zeros((B, 2), np.int32)
        _nl.check(lib.drn_jpeg_info(self._h, ptrs, lens, B, wh.ctypes.data), "drn_jpeg_info")
        sizes = 3 * wh[:, 0].astype(np.int64) * wh[:, 1].astype(np.int64)
        sizes = (sizes + 255) // 256 * 256
        offsets = np.concatenate([[0], np.cumsum(sizes)[:-1]]).astype(np.int64)
        total = int(sizes.sum()) + 256
        if self._arena is None or self._arena.numel() < total:
            self._arena = torch.empty(int(total * 1.25), dtype=torch.uint8, device=self.device)
        base = self._arena.data_ptr()
        outs = (ctypes.c_void_p * B)(*[base + int(o) for o in offsets])
        status = np.zeros(B, np.int32)
        with torch.cuda.device(self.device):
            _nl.check(lib.drn_jpeg_decode_batch(self._h, ptrs, lens, B, outs, wh.ctypes.data, status.ctypes.data,
                                                ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)),
                      "drn_jpeg_decode_batch")
        return self._arena, offsets, wh, status


def available() -> bool:
    try:
        return torch.cuda.is_available() and hasattr(_nl.load(), "drn_jpeg_create")
    except Exception:
        return False
