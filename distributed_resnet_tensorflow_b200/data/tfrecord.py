"""TFRecord files and tf.Example protos without TensorFlow (ImageNet shards of the reference:
``train-%05d-of-01024`` / ``validation-%05d-of-00128``, resnet_imagenet_main.py:103-155).

Record framing: uint64 length, uint32 masked-crc32c(length), payload, uint32 masked-crc32c(payload).
Indexing/verification of a whole file is done natively (csrc/hostutil.cu: drn_tfrecord_index) when
the library is available.
"""
from __future__ import annotations

import ctypes
import struct
from typing import Dict, Iterator, List, Union

import numpy as np

from ..ckpt.crc32c import masked_crc32c
from ..ckpt.tensorbundle import _field_bytes, _field_varint, _parse_fields, _varint


def write_tfrecords(path: str, payloads: List[bytes]) -> None:
    with open(path, "wb") as f:
        for p in payloads:
            header = struct.pack("<Q", len(p))
            f.write(header + struct.pack("<I", masked_crc32c(header)) + p +
                    struct.pack("<I", masked_crc32c(p)))


def read_tfrecords(path: str, verify: bool = True) -> Iterator[bytes]:
    with open(path, "rb") as f:
        data = f.read()
    buf = np.frombuffer(data, dtype=np.uint8)
    try:
        from .._native import lib as nl
        lib = nl.load(build_if_missing=False)
        fn = lib.drn_tfrecord_index
    except Exception:
        fn = None
    if fn is not None and len(data):
        cap = max(16, len(data) // 16)
        offs = np.zeros(cap, np.uint64)
        lens = np.zeros(cap, np.uint64)
        n = fn(buf.ctypes.data_as(ctypes.c_void_p), len(data), offs.ctypes.data_as(ctypes.c_void_p),
               lens.ctypes.data_as(ctypes.c_void_p), cap, 1 if verify else 0)
        if n < 0:
            raise ValueError(f"{path}: corrupt TFRecord at index {-n - 1}")
        for i in range(int(n)):
            o, ln = int(offs[i]), int(lens[i])
            yield data[o:o + ln]
        return
    pos = 0
    while pos + 12 <= len(data):
        (ln,) = struct.unpack_from("<Q", data, pos)
        if verify and struct.unpack_from("<I", data, pos + 8)[0] != masked_crc32c(data[pos:pos + 8]):
            raise ValueError(f"{path}: corrupt length checksum at {pos}")
        payload = data[pos + 12:pos + 12 + ln]
        if verify and struct.unpack_from("<I", data, pos + 12 + ln)[0] != masked_crc32c(payload):
            raise ValueError(f"{path}: corrupt payload checksum at {pos}")
        yield payload
        pos += 12 + ln + 4


Feature = Union[List[bytes], List[int], List[float]]


def parse_example(buf: bytes) -> Dict[str, Feature]:
    """tf.Example{features=1: Features{feature=1: map<string, Feature{bytes_list=1|float_list=2|
    int64_list=3}>}}"""
    out: Dict[str, Feature] = {}
    for num, _, features in _parse_fields(buf):
        if num != 1:
            continue
        for n2, _, entry in _parse_fields(features):
            if n2 != 1:
                continue
            key, val = "", None
            for n3, _, x in _parse_fields(entry):
                if n3 == 1:
                    key = x.decode()
                elif n3 == 2:
                    val = x
            if val is None:
                out[key] = []
                continue
            for n4, _, lst in _parse_fields(val):
                if n4 == 1:      # BytesList
                    out[key] = [v for k, _, v in _parse_fields(lst) if k == 1]
                elif n4 == 2:    # FloatList (packed or not)
                    vals: List[float] = []
                    for k, wt, v in _parse_fields(lst):
                        if k != 1:
                            continue
                        if wt == 2:
                            vals.extend(struct.unpack(f"<{len(v) // 4}f", v))
                        else:
                            vals.append(struct.unpack("<f", struct.pack("<I", v))[0])
                    out[key] = vals
                elif n4 == 3:    # Int64List (packed or not)
                    ints: List[int] = []
                    for k, wt, v in _parse_fields(lst):
                        if k != 1:
                            continue
                        if wt == 2:
                            pos = 0
                            while pos < len(v):
                                x, pos = _read_varint_signed(v, pos)
                                ints.append(x)
                        else:
                            ints.append(v if v < (1 << 63) else v - (1 << 64))
                    out[key] = ints
    return out


def _read_varint_signed(buf: bytes, pos: int):
    shift = result = 0
    while True:
        b = buf[pos]
        pos += 1
        result |= (b & 0x7F) << shift
        if not b & 0x80:
            break
        shift += 7
    if result >= 1 << 63:
        result -= 1 << 64
    return result, pos


def make_example(features: Dict[str, Feature]) -> bytes:
    """Encoder used by tests / synthetic shard generation."""
    entries = b""
    for key, vals in features.items():
        if vals and isinstance(vals[0], bytes):
            feat = _field_bytes(1, b"".join(_field_bytes(1, v) for v in vals))
        elif vals and isinstance(vals[0], float):
            feat = _field_bytes(2, _field_bytes(1, struct.pack(f"<{len(vals)}f", *vals)))
        else:
            feat = _field_bytes(3, _field_bytes(1, b"".join(_varint(int(v)) for v in vals)))
        entries += _field_bytes(1, _field_bytes(1, key.encode()) + _field_bytes(2, feat))
    return _field_bytes(1, entries)
