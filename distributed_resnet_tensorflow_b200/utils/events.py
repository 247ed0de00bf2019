"""TensorBoard event files (``events.out.tfevents.*``) without TensorFlow.

The reference writes scalar summaries ``cross_entropy``, ``cost``, ``learning_rate``
(resnet_model.py:82,88,93), ``Precision`` (resnet_cifar_main.py:274-278) through SummarySaverHook,
and ``Precision`` / ``Best Precision`` from the evaluators via tf.summary.FileWriter
(resnet_cifar_eval.py:93,122-136).  Here the Event/Summary protos are hand-encoded and framed as
TFRecords (length, masked crc32c(length), payload, masked crc32c(payload)).
"""
from __future__ import annotations

import os
import socket
import struct
import time
from typing import Dict, Iterator, List, Optional, Tuple

from ..ckpt.crc32c import masked_crc32c
from ..ckpt.tensorbundle import _field_bytes, _field_varint, _parse_fields, _varint


def _field_double(num: int, v: float) -> bytes:
    return _varint((num << 3) | 1) + struct.pack("<d", v)


def _field_float(num: int, v: float) -> bytes:
    return _varint((num << 3) | 5) + struct.pack("<f", v)


def encode_scalar_summary(values: Dict[str, float]) -> bytes:
    """Summary{ repeated Value value = 1 { string tag = 1; float simple_value = 2; } }"""
    out = b""
    for tag, v in values.items():
        out += _field_bytes(1, _field_bytes(1, tag.encode()) + _field_float(2, float(v)))
    return out


def encode_event(wall_time: float, step: int, summary: Optional[bytes] = None,
                 file_version: Optional[str] = None) -> bytes:
    """Event{ double wall_time = 1; int64 step = 2; string file_version = 3; Summary summary = 5; }"""
    out = _field_double(1, wall_time) + _field_varint(2, int(step))
    if file_version is not None:
        out += _field_bytes(3, file_version.encode())
    if summary is not None:
        out += _field_bytes(5, summary)
    return out


def frame_record(payload: bytes) -> bytes:
    header = struct.pack("<Q", len(payload))
    return (header + struct.pack("<I", masked_crc32c(header)) + payload +
            struct.pack("<I", masked_crc32c(payload)))


class EventFileWriter:
    def __init__(self, logdir: str, filename_suffix: str = ""):
        os.makedirs(logdir or ".", exist_ok=True)
        name = f"events.out.tfevents.{int(time.time())}.{socket.gethostname()}.{os.getpid()}{filename_suffix}"
        self.path = os.path.join(logdir or ".", name)
        self._f = open(self.path, "ab")
        self._write(encode_event(time.time(), 0, file_version="brain.Event:2"))

    def _write(self, event: bytes) -> None:
        self._f.write(frame_record(event))

    def add_scalars(self, values: Dict[str, float], step: int, wall_time: Optional[float] = None) -> None:
        self._write(encode_event(wall_time or time.time(), step, encode_scalar_summary(values)))

    def flush(self) -> None:
        self._f.flush()

    def close(self) -> None:
        self._f.flush()
        self._f.close()


def read_events(path: str, verify: bool = True) -> Iterator[Tuple[float, int, Dict[str, float]]]:
    """Yields (wall_time, step, {tag: value}) - used by tests and by tools that post-process runs."""
    with open(path, "rb") as f:
        data = f.read()
    pos = 0
    while pos + 12 <= len(data):
        (ln,) = struct.unpack_from("<Q", data, pos)
        if verify and struct.unpack_from("<I", data, pos + 8)[0] != masked_crc32c(data[pos:pos + 8]):
            raise ValueError("event file: corrupt length checksum")
        payload = data[pos + 12:pos + 12 + ln]
        if verify and struct.unpack_from("<I", data, pos + 12 + ln)[0] != masked_crc32c(payload):
            raise ValueError("event file: corrupt payload checksum")
        pos += 12 + ln + 4
        wall, step, scalars = 0.0, 0, {}
        for num, wt, v in _parse_fields(payload):
            if num == 1 and wt == 1:
                wall = struct.unpack("<d", struct.pack("<Q", v))[0]
            elif num == 2:
                step = v
            elif num == 5:
                for n2, _, val in _parse_fields(v):
                    if n2 != 1:
                        continue
                    tag, sv = "", None
                    for n3, w3, x in _parse_fields(val):
                        if n3 == 1:
                            tag = x.decode()
                        elif n3 == 2 and w3 == 5:
                            sv = struct.unpack("<f", struct.pack("<I", x))[0]
                    if sv is not None:
                        scalars[tag] = sv
        yield wall, step, scalars
