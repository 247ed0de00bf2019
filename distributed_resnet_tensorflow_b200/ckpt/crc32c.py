"""CRC-32C (Castagnoli) and TensorFlow's masking of it.

Uses the native ``drn_crc32c`` (csrc/hostutil.cu) when libdrn_b200.so is present, otherwise a
table-driven pure-Python fallback (slow, but checkpoints stay writable without the native library).
"""
from __future__ import annotations

import ctypes

_MASK_DELTA = 0xA282EAD8
_native = None
_table = None


def _get_native():
    global _native
    if _native is None:
        try:
            from .._native import lib as nl
            lib = nl.load(build_if_missing=False)
            _native = getattr(lib, "drn_crc32c", False) or False
        except Exception:
            _native = False
    return _native


def _py_table():
    global _table
    if _table is None:
        poly = 0x82F63B78
        t = []
        for i in range(256):
            c = i
            for _ in range(8):
                c = (c >> 1) ^ poly if c & 1 else c >> 1
            t.append(c)
        _table = t
    return _table


def crc32c(data, crc: int = 0) -> int:
    """CRC-32C of a bytes-like object; pass the previous value to continue a running CRC."""
    mv = memoryview(data).cast("B")
    fn = _get_native()
    if fn:
        n = len(mv)
        if n == 0:
            return crc & 0xFFFFFFFF
        if mv.readonly:
            buf = bytes(mv)
            return int(fn(ctypes.c_char_p(buf), n, crc))
        arr = (ctypes.c_ubyte * n).from_buffer(mv)
        return int(fn(ctypes.cast(arr, ctypes.c_void_p), n, crc))
    t = _py_table()
    c = crc ^ 0xFFFFFFFF
    for b in mv:
        c = t[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def mask(crc: int) -> int:
    """tensorflow/core/lib/hash/crc32c.h: rotate right by 15 bits and add a constant."""
    crc &= 0xFFFFFFFF
    return (((crc >> 15) | (crc << 17)) + _MASK_DELTA) & 0xFFFFFFFF


def unmask(masked: int) -> int:
    rot = (masked - _MASK_DELTA) & 0xFFFFFFFF
    return ((rot >> 17) | (rot << 15)) & 0xFFFFFFFF


def masked_crc32c(data) -> int:
    return mask(crc32c(data))
