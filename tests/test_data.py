import io
import os

import numpy as np
import pytest

from distributed_resnet_tensorflow_b200.data import cifar, preprocess, tfrecord
from distributed_resnet_tensorflow_b200.data.synthetic import synthetic_cifar


def test_cifar_record_roundtrip_10_and_100(tmp_path):
    imgs, labs = synthetic_cifar(37, 10, seed=1)
    p = str(tmp_path / "cifar-10-batches-bin" / "data_batch_1.bin")
    cifar.write_records(p, imgs, labs, "cifar10")
    assert os.path.getsize(p) == 37 * 3073
    i2, l2 = cifar.load_records([p], "cifar10")
    assert np.array_equal(i2, imgs) and np.array_equal(l2, labs)
    labs100 = (labs * 7) % 100
    p100 = str(tmp_path / "train.bin")
    cifar.write_records(p100, imgs, labs100, "cifar100")
    assert os.path.getsize(p100) == 37 * 3074      # coarse + fine label bytes (cifar_input.py:41-45)
    i3, l3 = cifar.load_records([p100], "cifar100")
    assert np.array_equal(i3, imgs) and np.array_equal(l3, labs100)
    assert cifar.train_filenames("/d")[0].endswith("cifar-10-batches-bin/data_batch_1.bin")
    assert len(cifar.train_filenames("/d")) == 5


def test_cifar_batches_shuffle_and_epochs():
    imgs, labs = synthetic_cifar(50, 10, seed=2)
    it = cifar.CifarBatches(imgs, labs, 16, True, num_epochs=2, seed=5)
    sizes = [len(l) for _, l in it]
    assert sum(sizes) == 100 and sizes[:-1] == [16] * 6 and sizes[-1] == 4
    ev = list(cifar.CifarBatches(imgs, labs, 25, False, num_epochs=1))
    assert np.array_equal(np.concatenate([l for _, l in ev]), labs)     # eval keeps file order


def test_philox_known_answer():
    # Random123 known-answer test for philox4x32-10
    assert preprocess.philox4x32_10((0, 0, 0, 0), (0, 0)) == (0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8)
    assert preprocess.philox4x32_10((0xFFFFFFFF,) * 4, (0xFFFFFFFF, 0xFFFFFFFF)) == \
        (0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD)


def test_cifar_preprocess_semantics():
    rng = np.random.default_rng(0)
    rec = rng.integers(0, 256, 3072, dtype=np.uint8)
    ev = preprocess.cifar_preprocess(rec, training=False)
    assert ev.shape == (32, 32, 3) and abs(ev.mean()) < 1e-4 and abs(ev.std() - 1) < 1e-3
    hwc = rec.reshape(3, 32, 32).transpose(1, 2, 0).astype(np.float32)
    assert np.allclose(ev, (hwc - hwc.mean()) / hwc.std(), atol=1e-4)
    flat = np.full(3072, 7, np.uint8)        # zero variance: divisor is 1/sqrt(3072), result 0
    assert np.allclose(preprocess.cifar_preprocess(flat, False), 0)
    seen = set()
    for idx in range(40):
        oy, ox, flip = preprocess.cifar_crop_params(3, 11, idx, 4)
        assert 0 <= oy <= 8 and 0 <= ox <= 8 and flip in (0, 1)
        seen.add((oy, ox, flip))
    assert len(seen) > 20
    for pad in (4, 2):                        # tf.data pipeline pads 4 per side, the legacy reader 2
        oy, ox, flip = preprocess.cifar_crop_params(3, 11, 5, pad)
        tr = preprocess.cifar_preprocess(rec, True, pad=pad, seed=3, step=11, index=5)
        padded = np.zeros((32 + 2 * pad, 32 + 2 * pad, 3), np.float32)
        padded[pad:pad + 32, pad:pad + 32] = hwc
        crop = padded[oy:oy + 32, ox:ox + 32]
        crop = crop[:, ::-1] if flip else crop
        assert np.allclose(tr, preprocess.per_image_standardization(crop), atol=1e-5)


def test_vgg_resize_semantics():
    assert preprocess.smallest_size_at_least(500, 375, 256) == (341, 256)
    assert preprocess.smallest_size_at_least(375, 500, 256) == (256, 341)
    img = np.arange(4 * 6 * 3, dtype=np.float32).reshape(4, 6, 3)
    same = preprocess.resize_bilinear(img, 4, 6)
    assert np.allclose(same, img)
    up = preprocess.resize_bilinear(img, 8, 12)       # align_corners=False: src = dst * 0.5
    assert np.allclose(up[0, 0], img[0, 0]) and np.allclose(up[2, 2], img[1, 1])
    assert np.allclose(up[1, 1], (img[0, 0] + img[0, 1] + img[1, 0] + img[1, 1]) / 4)
    assert np.allclose(up[7, 11], img[3, 5])          # clamped at the border
    c = preprocess.central_crop(np.zeros((300, 260, 3)), 224, 224)
    assert c.shape == (224, 224, 3)


def test_vgg_preprocess_shapes_and_means():
    rng = np.random.default_rng(0)
    img = rng.random((300, 400, 3)).astype(np.float32)
    ev = preprocess.vgg_preprocess(img, training=False)
    assert ev.shape == (224, 224, 3)
    tr = preprocess.vgg_preprocess(img, training=True, rng=np.random.default_rng(1))
    assert tr.shape == (224, 224, 3)
    assert np.allclose(preprocess.VGG_MEANS_01 * 255, [123.68, 116.78, 103.94])
    u8 = (img * 255).astype(np.uint8)
    crop, flip = preprocess.vgg_resize_crop_u8(u8, False)
    dev = preprocess.imagenet_prep_oracle(crop, flip)
    assert np.abs(dev - ev).max() < 2.5 / 255        # uint8 quantisation of the two pipelines


def test_space_to_depth_layout():
    img = np.arange(8 * 8 * 3, dtype=np.float32).reshape(8, 8, 3)
    s = preprocess.space_to_depth16(img)
    assert s.shape == (4, 4, 16) and np.all(s[..., 12:] == 0)
    for dy in range(2):
        for dx in range(2):
            for c in range(3):
                assert s[1, 2, (dy * 2 + dx) * 3 + c] == img[2 + dy, 4 + dx, c]


def test_tfrecord_and_example_roundtrip(tmp_path):
    from PIL import Image
    buf = io.BytesIO()
    Image.fromarray(np.random.default_rng(0).integers(0, 256, (40, 50, 3), dtype=np.uint8)).save(buf, format="JPEG")
    ex = tfrecord.make_example({"image/encoded": [buf.getvalue()], "image/format": [b"jpeg"],
                                "image/class/label": [417], "image/object/bbox/xmin": [0.1, 0.25]})
    p = str(tmp_path / "train-00000-of-01024")
    tfrecord.write_tfrecords(p, [ex, ex])
    recs = list(tfrecord.read_tfrecords(p))
    assert len(recs) == 2 and recs[0] == ex
    parsed = tfrecord.parse_example(recs[1])
    assert parsed["image/class/label"] == [417] and parsed["image/format"] == [b"jpeg"]
    assert np.allclose(parsed["image/object/bbox/xmin"], [0.1, 0.25])
    blob = bytearray(open(p, "rb").read())
    blob[30] ^= 0xFF
    open(p, "wb").write(bytes(blob))
    with pytest.raises(ValueError):
        list(tfrecord.read_tfrecords(p))


def test_imagenet_batches_from_shards(tmp_path):
    from PIL import Image
    from distributed_resnet_tensorflow_b200.data.imagenet import ImageNetBatches, filenames
    assert filenames(True, "/d")[1023].endswith("train-01023-of-01024")
    assert filenames(False, "/d")[127].endswith("validation-00127-of-00128")
    rng = np.random.default_rng(0)
    exs = []
    for i in range(6):
        buf = io.BytesIO()
        Image.fromarray(rng.integers(0, 256, (260 + 10 * i, 300, 3), dtype=np.uint8)).save(buf, format="JPEG")
        exs.append(tfrecord.make_example({"image/encoded": [buf.getvalue()], "image/class/label": [i + 1]}))
    f = str(tmp_path / "validation-00000-of-00128")
    tfrecord.write_tfrecords(f, exs)
    batches = list(ImageNetBatches(str(tmp_path), 4, False, num_epochs=1, files=[f]))
    assert [b[0].shape for b in batches] == [(4, 224 * 224 * 3), (2, 224 * 224 * 3)]
    assert list(batches[0][1]) == [1, 2, 3, 4] and batches[0][2].shape == (4, 3)


def test_legacy_build_input_and_vgg_module(tmp_path):
    from distributed_resnet_tensorflow_b200.data import cifar_input, vgg_preprocessing
    imgs, labs = synthetic_cifar(40, 100, seed=3)
    p = str(tmp_path / "test_batch.bin")
    cifar.write_records(p, imgs, labs % 100, "cifar100")
    batches = list(cifar_input.build_input("cifar100", str(tmp_path / "test_batch*"), 10, "eval", num_epochs=1))
    assert len(batches) == 4 and batches[0][0].shape == (10, 32, 32, 3) and batches[0][1].shape == (10, 100)
    assert np.array_equal(batches[0][1].argmax(1), (labs % 100)[:10])
    tr = next(iter(cifar_input.build_input("cifar100", p, 8, "train")))
    assert tr[0].shape == (8, 32, 32, 3) and abs(tr[0][0].mean()) < 1e-3
    with pytest.raises(ValueError):
        next(iter(cifar_input.build_input("mnist", p, 8, "train")))
    img = np.random.default_rng(0).random((300, 280, 3)).astype(np.float32)
    out = vgg_preprocessing.preprocess_image(img, 224, 224, is_training=False)
    assert out.shape == (224, 224, 3) and np.allclose(out, preprocess.vgg_preprocess(img, False))
    assert vgg_preprocessing.preprocess_image(img, 224, 224, True, rng=np.random.default_rng(1)).shape == (224, 224, 3)


def test_multi_process_readers_deliver_every_example_once(tmp_path):
    """--input_workers: reader processes over disjoint shard subsets (host JPEG path)."""
    import io
    from PIL import Image
    from distributed_resnet_tensorflow_b200.data import imagenet as im
    from distributed_resnet_tensorflow_b200.data.tfrecord import make_example, write_tfrecords
    rng = np.random.default_rng(0)
    files, label = [], 1
    for s_ in range(2):
        recs = []
        for _ in range(6):
            img = rng.integers(0, 255, (260, 300, 3), dtype=np.uint8)
            buf = io.BytesIO()
            Image.fromarray(img).save(buf, format="JPEG", quality=80)
            recs.append(make_example({"image/encoded": [buf.getvalue()], "image/class/label": [label]}))
            label += 1
        f = str(tmp_path / ("validation-%05d-of-00128" % s_))
        write_tfrecords(f, recs)
        files.append(f)
    mp = im.MultiProcessBatches(str(tmp_path), 3, False, 1, seed=0, workers=2, num_threads=2, files=files)
    labels = []
    for imgs, labs, crop in mp:
        assert imgs.shape == (3, 224 * 224 * 3) and imgs.dtype == np.uint8 and crop.shape == (3, 3)
        labels.extend(int(x) for x in labs)
    mp.close()
    assert sorted(labels) == list(range(1, 13))
