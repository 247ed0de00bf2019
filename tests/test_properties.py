"""Property-based tests (hypothesis) of the host-side formats: whatever is written must be read back
bit for bit, for arbitrary names, shapes and payloads (hypothesis-driven sweeps)."""
import os

import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

from distributed_resnet_tensorflow_b200.ckpt import crc32c as crcmod
from distributed_resnet_tensorflow_b200.ckpt import tensorbundle as tb
from distributed_resnet_tensorflow_b200.data import preprocess, tfrecord

_names = st.text(alphabet="abcdefghijklmnopqrstuvwxyz0123456789_/", min_size=1, max_size=40)
_dtypes = st.sampled_from([np.float32, np.int32, np.int64])
_shapes = st.lists(st.integers(0, 5), min_size=0, max_size=4)
_slow = settings(max_examples=25, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture])


@_slow
@given(st.dictionaries(_names, st.tuples(_dtypes, _shapes), min_size=1, max_size=12), st.integers(0, 2 ** 31 - 1))
def test_bundle_roundtrip_arbitrary_tensors(tmp_path_factory, spec, seed):
    rng = np.random.default_rng(seed)
    tensors = {}
    for name, (dt, shape) in spec.items():
        a = rng.standard_normal(shape) * 100 if dt is np.float32 else rng.integers(-2 ** 31, 2 ** 31 - 1, shape)
        tensors[name] = np.asarray(a).astype(dt).reshape(shape)
    prefix = str(tmp_path_factory.mktemp("b") / "model.ckpt-7")
    tb.write_bundle(prefix, tensors)
    back = tb.read_bundle(prefix)
    assert sorted(back) == sorted(tensors)
    for k, v in tensors.items():
        assert back[k].dtype == v.dtype and back[k].shape == v.shape
        assert back[k].tobytes() == v.tobytes()


@_slow
@given(st.lists(st.binary(min_size=0, max_size=3000), min_size=0, max_size=20))
def test_tfrecord_roundtrip_arbitrary_payloads(tmp_path_factory, payloads):
    path = str(tmp_path_factory.mktemp("r") / "shard")
    tfrecord.write_tfrecords(path, payloads)
    assert list(tfrecord.read_tfrecords(path)) == payloads


@settings(max_examples=60, deadline=None)
@given(st.binary(min_size=0, max_size=5000), st.integers(0, 4999))
def test_crc32c_is_incremental_and_masks_invertibly(data, cut):
    cut = min(cut, len(data))
    whole = crcmod.crc32c(data)
    assert crcmod.crc32c(data[cut:], crcmod.crc32c(data[:cut])) == whole
    assert crcmod.unmask(crcmod.mask(whole)) == whole


@settings(max_examples=60, deadline=None)
@given(st.integers(32, 900), st.integers(32, 900), st.booleans(), st.integers(0, 2 ** 31 - 1))
def test_vgg_crop_parameters_always_fit(h, w, training, seed):
    nh, nw, oy, ox, flip = preprocess.vgg_crop_params(h, w, training, 224, np.random.default_rng(seed))
    # float32 truncation as in the reference (tf.to_int32(height * scale)) may land one pixel short
    assert min(nh, nw) >= 255 and (min(nh, nw) <= 512 if training else min(nh, nw) in (255, 256))
    assert 0 <= oy <= nh - 224 and 0 <= ox <= nw - 224 and flip in (0, 1)
    # aspect ratio preserved to rounding
    assert abs(nh / nw - h / w) <= max(1.0 / nw, 1.0 / nh) * (h / w + 1) + 1e-6
