"""TensorFlow checkpoint ("TensorBundle V2") reader and writer without TensorFlow.

A checkpoint ``prefix`` consists of
  ``prefix.index``                 - LevelDB-format SSTable: key "" -> BundleHeaderProto,
                                     key <variable name> -> BundleEntryProto
  ``prefix.data-0000i-of-0000N``   - raw little-endian tensor bytes
and the directory holds a text ``checkpoint`` state file naming the latest prefix.

This is what ``tf.train.Saver`` (used by the reference's evaluators, resnet_cifar_eval.py:92-109) and
``MonitoredTrainingSession``'s CheckpointSaverHook (resnet_cifar_main.py:311-335) read and write;
layout per SURVEY.md Appendix B.  The writer emits one shard, no compression; the reader also
accepts multi-shard bundles (written by PS jobs) and snappy-compressed index blocks.
"""
from __future__ import annotations

import os
import struct
from typing import Dict, Iterable, List, Optional, Tuple

import numpy as np

from .crc32c import crc32c, mask, masked_crc32c, unmask

TABLE_MAGIC = 0xDB4775248B80FB57
BLOCK_RESTART_INTERVAL = 16
BLOCK_SIZE = 256 * 1024

# tensorflow/core/framework/types.proto
DT_FLOAT, DT_DOUBLE, DT_INT32, DT_UINT8, DT_INT64, DT_BOOL, DT_BFLOAT16, DT_HALF = 1, 2, 3, 4, 9, 10, 14, 19
_NP_TO_DT = {np.dtype(np.float32): DT_FLOAT, np.dtype(np.float64): DT_DOUBLE,
             np.dtype(np.int32): DT_INT32, np.dtype(np.uint8): DT_UINT8,
             np.dtype(np.int64): DT_INT64, np.dtype(np.bool_): DT_BOOL,
             np.dtype(np.float16): DT_HALF}
_DT_TO_NP = {v: k for k, v in _NP_TO_DT.items()}
_DT_TO_NP[DT_BFLOAT16] = np.dtype(np.uint16)  # raw bits


# ---------------------------------------------------------------------------------------------
# protobuf wire helpers (only what the two bundle messages need)
# ---------------------------------------------------------------------------------------------
def _varint(n: int) -> bytes:
    if n < 0:
        n += 1 << 64
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _read_varint(buf, pos: int) -> Tuple[int, int]:
    shift = result = 0
    while True:
        b = buf[pos]
        pos += 1
        result |= (b & 0x7F) << shift
        if not b & 0x80:
            return result, pos
        shift += 7


def _field_varint(num: int, value: int) -> bytes:
    return _varint(num << 3) + _varint(value)


def _field_bytes(num: int, payload: bytes) -> bytes:
    return _varint((num << 3) | 2) + _varint(len(payload)) + payload


def _field_fixed32(num: int, value: int) -> bytes:
    return _varint((num << 3) | 5) + struct.pack("<I", value & 0xFFFFFFFF)


def _parse_fields(buf: bytes) -> List[Tuple[int, int, object]]:
    out, pos = [], 0
    while pos < len(buf):
        tag, pos = _read_varint(buf, pos)
        num, wt = tag >> 3, tag & 7
        if wt == 0:
            v, pos = _read_varint(buf, pos)
        elif wt == 2:
            ln, pos = _read_varint(buf, pos)
            v = bytes(buf[pos:pos + ln])
            pos += ln
        elif wt == 5:
            v = struct.unpack_from("<I", buf, pos)[0]
            pos += 4
        elif wt == 1:
            v = struct.unpack_from("<Q", buf, pos)[0]
            pos += 8
        else:
            raise ValueError(f"unsupported wire type {wt}")
        out.append((num, wt, v))
    return out


def encode_header(num_shards: int = 1) -> bytes:
    """BundleHeaderProto{num_shards=1, endianness=2 (LITTLE=0, omitted), version=3{producer=1}}"""
    version = _field_varint(1, 1)
    return _field_varint(1, num_shards) + _field_bytes(3, version)


def encode_entry(dtype: int, shape: Iterable[int], shard_id: int, offset: int, size: int,
                 crc: int) -> bytes:
    """BundleEntryProto{dtype=1, shape=2, shard_id=3, offset=4, size=5, crc32c=6 (fixed32)}"""
    dims = b"".join(_field_bytes(2, _field_varint(1, int(d))) for d in shape)
    out = _field_varint(1, dtype) + _field_bytes(2, dims)
    if shard_id:
        out += _field_varint(3, shard_id)
    if offset:
        out += _field_varint(4, offset)
    out += _field_varint(5, size) + _field_fixed32(6, crc)
    return out


def decode_entry(buf: bytes) -> dict:
    e = {"dtype": 0, "shape": [], "shard_id": 0, "offset": 0, "size": 0, "crc32c": 0, "slices": 0}
    for num, _, v in _parse_fields(buf):
        if num == 1:
            e["dtype"] = v
        elif num == 2:
            for n2, _, dim in _parse_fields(v):
                if n2 == 2:
                    size = 0
                    for n3, _, x in _parse_fields(dim):
                        if n3 == 1:
                            size = x if x < (1 << 63) else x - (1 << 64)
                    e["shape"].append(size)
        elif num == 3:
            e["shard_id"] = v
        elif num == 4:
            e["offset"] = v
        elif num == 5:
            e["size"] = v
        elif num == 6:
            e["crc32c"] = v
        elif num == 7:
            e["slices"] += 1
    return e


# ---------------------------------------------------------------------------------------------
# SSTable (LevelDB table format)
# ---------------------------------------------------------------------------------------------
class _BlockBuilder:
    def __init__(self):
        self.buf = bytearray()
        self.restarts = [0]
        self.counter = 0
        self.last_key = b""

    def add(self, key: bytes, value: bytes) -> None:
        shared = 0
        if self.counter < BLOCK_RESTART_INTERVAL:
            m = min(len(self.last_key), len(key))
            while shared < m and self.last_key[shared] == key[shared]:
                shared += 1
        else:
            self.restarts.append(len(self.buf))
            self.counter = 0
        self.buf += _varint(shared) + _varint(len(key) - shared) + _varint(len(value))
        self.buf += key[shared:] + value
        self.last_key = key
        self.counter += 1

    def finish(self) -> bytes:
        out = bytes(self.buf) + b"".join(struct.pack("<I", r) for r in self.restarts)
        return out + struct.pack("<I", len(self.restarts))

    def size_estimate(self) -> int:
        return len(self.buf) + 4 * len(self.restarts) + 4

    def empty(self) -> bool:
        return not self.buf


def _block_handle(offset: int, size: int) -> bytes:
    return _varint(offset) + _varint(size)


def _shortest_separator(a: bytes, b: bytes) -> bytes:
    """LevelDB's BytewiseComparator::FindShortestSeparator."""
    m = min(len(a), len(b))
    i = 0
    while i < m and a[i] == b[i]:
        i += 1
    if i >= m:
        return a
    if a[i] < 0xFF and a[i] + 1 < b[i]:
        return a[:i] + bytes([a[i] + 1])
    return a


def _short_successor(a: bytes) -> bytes:
    for i, ch in enumerate(a):
        if ch != 0xFF:
            return a[:i] + bytes([ch + 1])
    return a


def write_sstable(path: str, items: List[Tuple[bytes, bytes]]) -> None:
    """items must be sorted bytewise by key."""
    out = bytearray()
    index = _BlockBuilder()
    data = _BlockBuilder()
    pending: Optional[Tuple[bytes, bytes]] = None  # (last key of finished block, handle)

    def flush_block(b: _BlockBuilder) -> bytes:
        raw = b.finish()
        handle = _block_handle(len(out), len(raw))
        trailer_type = b"\x00"  # kNoCompression
        crc = mask(crc32c(trailer_type, crc32c(raw)))
        out.extend(raw)
        out.extend(trailer_type + struct.pack("<I", crc))
        return handle

    for key, value in items:
        if pending is not None:
            index.add(_shortest_separator(pending[0], key), pending[1])
            pending = None
        data.add(key, value)
        if data.size_estimate() >= BLOCK_SIZE:
            h = flush_block(data)
            pending = (data.last_key, h)
            data = _BlockBuilder()
    if not data.empty():
        h = flush_block(data)
        pending = (data.last_key, h)
    if pending is not None:
        index.add(_short_successor(pending[0]), pending[1])
    meta_handle = flush_block(_BlockBuilder())
    index_handle = flush_block(index)
    footer = meta_handle + index_handle
    footer += b"\x00" * (40 - len(footer))
    footer += struct.pack("<Q", TABLE_MAGIC)
    out.extend(footer)
    with open(path, "wb") as f:
        f.write(out)


def _snappy_decompress(src: bytes) -> bytes:
    n, pos = _read_varint(src, 0)
    out = bytearray()
    while pos < len(src):
        tag = src[pos]
        pos += 1
        kind = tag & 3
        if kind == 0:
            ln = tag >> 2
            if ln >= 60:
                nb = ln - 59
                ln = int.from_bytes(src[pos:pos + nb], "little")
                pos += nb
            ln += 1
            out += src[pos:pos + ln]
            pos += ln
            continue
        if kind == 1:
            ln = ((tag >> 2) & 7) + 4
            off = ((tag >> 5) << 8) | src[pos]
            pos += 1
        elif kind == 2:
            ln = (tag >> 2) + 1
            off = src[pos] | (src[pos + 1] << 8)
            pos += 2
        else:
            ln = (tag >> 2) + 1
            off = int.from_bytes(src[pos:pos + 4], "little")
            pos += 4
        for _ in range(ln):
            out.append(out[-off])
    if len(out) != n:
        raise ValueError("snappy: length mismatch")
    return bytes(out)


def _read_block(buf: bytes, offset: int, size: int, verify: bool = True) -> bytes:
    raw = buf[offset:offset + size]
    ctype = buf[offset + size]
    stored = struct.unpack_from("<I", buf, offset + size + 1)[0]
    if verify:
        actual = crc32c(bytes([ctype]), crc32c(raw))
        if unmask(stored) != actual:
            raise ValueError("sstable: block checksum mismatch")
    if ctype == 0:
        return raw
    if ctype == 1:
        return _snappy_decompress(raw)
    raise ValueError(f"sstable: unknown compression {ctype}")


def _iter_block(block: bytes):
    num_restarts = struct.unpack_from("<I", block, len(block) - 4)[0]
    end = len(block) - 4 - 4 * num_restarts
    pos, key = 0, b""
    while pos < end:
        shared, pos = _read_varint(block, pos)
        non_shared, pos = _read_varint(block, pos)
        vlen, pos = _read_varint(block, pos)
        key = key[:shared] + block[pos:pos + non_shared]
        pos += non_shared
        yield key, block[pos:pos + vlen]
        pos += vlen


def read_sstable(path: str) -> List[Tuple[bytes, bytes]]:
    with open(path, "rb") as f:
        buf = f.read()
    if len(buf) < 48 or struct.unpack_from("<Q", buf, len(buf) - 8)[0] != TABLE_MAGIC:
        raise ValueError(f"{path}: not an SSTable (bad magic)")
    footer = buf[-48:]
    pos = 0
    _, pos = _read_varint(footer, pos)
    _, pos = _read_varint(footer, pos)
    ioff, pos = _read_varint(footer, pos)
    isize, pos = _read_varint(footer, pos)
    out = []
    for _, handle in _iter_block(_read_block(buf, ioff, isize)):
        boff, p2 = _read_varint(handle, 0)
        bsize, _ = _read_varint(handle, p2)
        out.extend(_iter_block(_read_block(buf, boff, bsize)))
    return out


# ---------------------------------------------------------------------------------------------
# bundle API
# ---------------------------------------------------------------------------------------------
def _data_name(prefix: str, shard: int, num_shards: int) -> str:
    return f"{prefix}.data-{shard:05d}-of-{num_shards:05d}"


def write_bundle(prefix: str, tensors: Dict[str, np.ndarray]) -> None:
    """Writes ``prefix.index`` + ``prefix.data-00000-of-00001`` atomically (temp files + rename),
    so a polling evaluator never sees a torn checkpoint."""
    os.makedirs(os.path.dirname(os.path.abspath(prefix)), exist_ok=True)
    names = sorted(tensors, key=lambda s: s.encode())
    items: List[Tuple[bytes, bytes]] = [(b"", encode_header(1))]
    data_tmp = _data_name(prefix, 0, 1) + ".tmp"
    offset = 0
    with open(data_tmp, "wb") as f:
        for name in names:
            arr = np.asarray(tensors[name])
            if arr.dtype not in _NP_TO_DT:
                raise TypeError(f"{name}: unsupported dtype {arr.dtype}")
            shape = arr.shape  # np.ascontiguousarray would promote 0-d arrays to 1-d
            raw = arr.astype(arr.dtype.newbyteorder("<"), copy=False).tobytes(order="C")
            f.write(raw)
            items.append((name.encode(), encode_entry(_NP_TO_DT[arr.dtype], shape, 0, offset,
                                                      len(raw), masked_crc32c(raw))))
            offset += len(raw)
    index_tmp = prefix + ".index.tmp"
    write_sstable(index_tmp, items)
    os.replace(data_tmp, _data_name(prefix, 0, 1))
    os.replace(index_tmp, prefix + ".index")


class BundleReader:
    def __init__(self, prefix: str):
        self.prefix = prefix
        self.entries: Dict[str, dict] = {}
        self.num_shards = 1
        for key, value in read_sstable(prefix + ".index"):
            if key == b"":
                for num, _, v in _parse_fields(value):
                    if num == 1:
                        self.num_shards = v
                    if num == 2 and v != 0:
                        raise ValueError("big-endian bundles are not supported")
                continue
            self.entries[key.decode()] = decode_entry(value)

    def keys(self) -> List[str]:
        return sorted(self.entries)

    def has_tensor(self, name: str) -> bool:
        return name in self.entries

    def shape_and_dtype(self, name: str):
        e = self.entries[name]
        return tuple(e["shape"]), _DT_TO_NP[e["dtype"]]

    def get_tensor(self, name: str, verify: bool = True) -> np.ndarray:
        e = self.entries[name]
        if e["slices"]:
            raise NotImplementedError(f"{name}: partitioned (sliced) variables are not supported")
        with open(_data_name(self.prefix, e["shard_id"], self.num_shards), "rb") as f:
            f.seek(e["offset"])
            raw = f.read(e["size"])
        if len(raw) != e["size"]:
            raise ValueError(f"{name}: truncated data file")
        if verify and unmask(e["crc32c"]) != crc32c(raw):
            raise ValueError(f"{name}: tensor checksum mismatch")
        dt = _DT_TO_NP[e["dtype"]]
        return np.frombuffer(raw, dtype=dt).reshape(e["shape"]).copy()

    def read_all(self) -> Dict[str, np.ndarray]:
        return {k: self.get_tensor(k) for k in self.keys()}


def read_bundle(prefix: str) -> Dict[str, np.ndarray]:
    return BundleReader(prefix).read_all()
