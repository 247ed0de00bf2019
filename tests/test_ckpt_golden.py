"""Byte-level golden for the TensorBundle V2 writer: the expected .index file
of a two-tensor checkpoint is assembled HERE, independently of ckpt/tensorbundle.py, from the
documented formats - LevelDB table (prefix-compressed entries, restart array, 5-byte block trailer
with a masked CRC-32C, metaindex + index blocks, 48-byte footer with the table magic) holding
BundleHeaderProto / BundleEntryProto messages - and compared byte for byte with what the writer
produces.  The CRC used by this assembly is a bit-by-bit implementation pinned by the standard
check value, so a bug shared with ckpt/crc32c.py cannot hide.  (The reference gets all of this from
tf.train.Saver, resnet_cifar_eval.py:92-109.)"""
import struct

import numpy as np

from distributed_resnet_tensorflow_b200.ckpt.tensorbundle import write_bundle


def _crc32c_bitwise(data: bytes) -> int:
    crc = 0xFFFFFFFF
    for byte in data:
        crc ^= byte
        for _ in range(8):
            crc = (crc >> 1) ^ (0x82F63B78 if crc & 1 else 0)
    return crc ^ 0xFFFFFFFF


def _masked(data: bytes) -> int:
    c = _crc32c_bitwise(data)
    return (((c >> 15) | (c << 17)) + 0xA282EAD8) & 0xFFFFFFFF


def _varint(n: int) -> bytes:
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        out.append(b | (0x80 if n else 0))
        if not n:
            return bytes(out)


def _block(entries) -> bytes:
    """LevelDB block: (shared, non_shared, value_len, key delta, value)*, restarts, num_restarts.
    Fewer than 16 keys here, so there is exactly one restart point at offset 0."""
    body, prev = bytearray(), b""
    for key, value in entries:
        shared = 0
        while shared < min(len(prev), len(key)) and prev[shared] == key[shared]:
            shared += 1
        body += _varint(shared) + _varint(len(key) - shared) + _varint(len(value)) + key[shared:] + value
        prev = key
    body += struct.pack("<I", 0) + struct.pack("<I", 1)
    return bytes(body)


def _with_trailer(block: bytes) -> bytes:
    return block + b"\x00" + struct.pack("<I", _masked(block + b"\x00"))   # type 0 = uncompressed


def test_crc32c_check_value():
    assert _crc32c_bitwise(b"123456789") == 0xE3069283
    from distributed_resnet_tensorflow_b200.ckpt.crc32c import crc32c
    assert crc32c(b"123456789") == 0xE3069283


def test_index_file_matches_hand_assembled_bytes(tmp_path):
    a = np.asarray([1.0, 2.0], np.float32)
    step = np.asarray(7, np.int64)
    prefix = str(tmp_path / "model.ckpt-7")
    write_bundle(prefix, {"global_step": step, "a": a})

    # data file: tensors in key order, raw little-endian
    data = a.tobytes() + step.tobytes()
    assert open(prefix + ".data-00000-of-00001", "rb").read() == data

    # BundleHeaderProto{num_shards=1 (field 1), endianness=LITTLE=0 (default, omitted), version{producer=1}}
    header = bytes([0x08, 0x01]) + bytes([0x1A, 0x02, 0x08, 0x01])
    # BundleEntryProto{dtype=1, shape=2, shard_id=3 (0, omitted), offset=4, size=5, crc32c=6 fixed32}
    entry_a = (bytes([0x08, 0x01])                                  # DT_FLOAT
               + bytes([0x12, 0x04, 0x12, 0x02, 0x08, 0x02])         # shape { dim { size: 2 } }
               + bytes([0x28, 0x08])                                 # size 8 (offset 0 omitted)
               + bytes([0x35]) + struct.pack("<I", _masked(a.tobytes())))
    entry_s = (bytes([0x08, 0x09])                                  # DT_INT64
               + bytes([0x12, 0x00])                                 # scalar: empty TensorShapeProto
               + bytes([0x20, 0x08])                                 # offset 8
               + bytes([0x28, 0x08])                                 # size 8
               + bytes([0x35]) + struct.pack("<I", _masked(step.tobytes())))
    data_block = _block([(b"", header), (b"a", entry_a), (b"global_step", entry_s)])
    metaindex = _block([])
    off_meta = len(data_block) + 5
    off_index = off_meta + len(metaindex) + 5
    handle_data = _varint(0) + _varint(len(data_block))
    handle_meta = _varint(off_meta) + _varint(len(metaindex))
    # index block: one entry, key = a short successor of the last data key ("global_step" -> "h")
    index = _block([(b"h", handle_data)])
    handle_index = _varint(off_index) + _varint(len(index))
    footer = (handle_meta + handle_index).ljust(40, b"\x00") + struct.pack("<Q", 0xDB4775248B80FB57)
    expected = _with_trailer(data_block) + _with_trailer(metaindex) + _with_trailer(index) + footer

    got = open(prefix + ".index", "rb").read()
    assert got[-8:] == expected[-8:], "table magic"
    assert got == expected, (got.hex(), expected.hex())
