import os
import struct

import numpy as np
import pytest

from distributed_resnet_tensorflow_b200.ckpt import crc32c as C
from distributed_resnet_tensorflow_b200.ckpt.saver import CheckpointState, Saver, get_checkpoint_state, latest_checkpoint
from distributed_resnet_tensorflow_b200.ckpt.tensorbundle import (BundleReader, _snappy_decompress, read_bundle,
                                                                   read_sstable, write_bundle, write_sstable)
from distributed_resnet_tensorflow_b200.models.init import init_variables
from distributed_resnet_tensorflow_b200.models.spec import cifar_resnet


def test_crc32c_known_vectors():
    assert C.crc32c(b"") == 0
    assert C.crc32c(b"123456789") == 0xE3069283
    assert C.crc32c(bytes(32)) == 0x8A9136AA
    assert C.crc32c(b"6789", C.crc32c(b"12345")) == 0xE3069283
    assert C.unmask(C.mask(0x12345678)) == 0x12345678


def test_crc32c_python_fallback_matches_native(monkeypatch):
    data = np.random.default_rng(0).integers(0, 256, 10007, dtype=np.uint8).tobytes()
    fast = C.crc32c(data)
    monkeypatch.setattr(C, "_native", False)
    assert C.crc32c(data) == fast


def test_sstable_roundtrip_many_keys(tmp_path):
    items = sorted((f"key/{i:05d}".encode(), os.urandom(i % 97)) for i in range(3000))
    p = str(tmp_path / "t.sst")
    write_sstable(p, items)
    assert read_sstable(p) == items
    with open(p, "rb") as f:
        blob = f.read()
    assert struct.unpack("<Q", blob[-8:])[0] == 0xDB4775248B80FB57


def test_sstable_detects_corruption(tmp_path):
    p = str(tmp_path / "t.sst")
    write_sstable(p, [(b"a", b"1" * 100), (b"b", b"2" * 100)])
    blob = bytearray(open(p, "rb").read())
    blob[10] ^= 0xFF
    open(p, "wb").write(bytes(blob))
    with pytest.raises(ValueError):
        read_sstable(p)


def test_snappy_decoder():
    # literal "abcd" then copy (offset 4, len 8) -> "abcdabcdabcd"
    comp = bytes([12]) + bytes([(4 - 1) << 2]) + b"abcd" + bytes([((8 - 4) << 2) | 1, 4])
    assert _snappy_decompress(comp) == b"abcdabcdabcd"


def test_bundle_roundtrip_full_model(tmp_path):
    variables = init_variables(cifar_resnet(50), seed=3)
    prefix = str(tmp_path / "model.ckpt-7")
    write_bundle(prefix, variables)
    assert os.path.exists(prefix + ".index") and os.path.exists(prefix + ".data-00000-of-00001")
    r = BundleReader(prefix)
    assert len(r.keys()) == 403 and r.keys() == sorted(variables, key=lambda s: s.encode())
    assert r.shape_and_dtype("conv2d/kernel") == ((3, 3, 3, 16), np.dtype(np.float32))
    assert r.shape_and_dtype("global_step") == ((), np.dtype(np.int64))
    back = r.read_all()
    for k, v in variables.items():
        assert back[k].dtype == np.asarray(v).dtype and np.array_equal(back[k], v), k


def test_bundle_detects_corrupt_tensor(tmp_path):
    prefix = str(tmp_path / "m")
    write_bundle(prefix, {"a": np.arange(100, dtype=np.float32)})
    f = prefix + ".data-00000-of-00001"
    blob = bytearray(open(f, "rb").read())
    blob[17] ^= 1
    open(f, "wb").write(bytes(blob))
    with pytest.raises(ValueError):
        read_bundle(prefix)


def test_saver_state_file_and_rotation(tmp_path):
    root = str(tmp_path / "log_root")
    s = Saver(root, keep_max=5)
    for step in range(1, 8):
        s.save({"w": np.full(3, step, np.float32), "global_step": np.asarray(step, np.int64)}, step)
    st = get_checkpoint_state(root)
    assert st.model_checkpoint_path == "model.ckpt-7"
    assert st.all_model_checkpoint_paths == [f"model.ckpt-{i}" for i in range(3, 8)]
    assert not os.path.exists(os.path.join(root, "model.ckpt-1.index"))
    assert os.path.exists(os.path.join(root, "model.ckpt-3.index"))
    assert latest_checkpoint(root).endswith("model.ckpt-7")
    assert int(Saver.restore(latest_checkpoint(root))["global_step"]) == 7
    text = open(os.path.join(root, "checkpoint")).read()
    assert text.splitlines()[0] == 'model_checkpoint_path: "model.ckpt-7"'
    assert CheckpointState.from_text(text).all_model_checkpoint_paths[-1] == "model.ckpt-7"


def test_restore_subset_like_evaluator(tmp_path):
    variables = init_variables(cifar_resnet(8))
    prefix = str(tmp_path / "model.ckpt-0")
    write_bundle(prefix, variables)
    names = [k for k in variables if not k.endswith("/Momentum")]
    got = Saver.restore(prefix, names)
    assert set(got) == set(names)
    with pytest.raises(KeyError):
        Saver.restore(prefix, ["does/not/exist"])
