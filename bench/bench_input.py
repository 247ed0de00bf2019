"""Throughput of the two ImageNet input paths on synthetic JPEGs of ImageNet-like size
(500x375, quality 90): host (PIL decode + numpy resize/crop, 5 threads as the reference) vs device
(nvJPEG batch decode + drn_resize_crop_batch)."""
import io, os, sys, tempfile, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from PIL import Image
from distributed_resnet_tensorflow_b200.data import imagenet as im
from distributed_resnet_tensorflow_b200.data.tfrecord import make_example, write_tfrecords

def main(n=1024, batch=128):
    rng = np.random.default_rng(0)
    recs = []
    base = rng.integers(0, 255, (48, 64, 3), dtype=np.uint8)
    for i in range(n):
        img = np.asarray(Image.fromarray(base).resize((500, 375), Image.BILINEAR))
        img = (img.astype(np.int32) + rng.integers(-20, 21, img.shape)).clip(0, 255).astype(np.uint8)
        buf = io.BytesIO(); Image.fromarray(img).save(buf, format="JPEG", quality=90)
        recs.append(make_example({"image/encoded": [buf.getvalue()], "image/class/label": [i % 1000 + 1]}))
    d = tempfile.mkdtemp(); shard = os.path.join(d, "train-00000-of-01024"); write_tfrecords(shard, recs)
    print(f"{n} JPEGs, {sum(len(r) for r in recs) / n / 1024:.0f} KiB each")
    for name, cls, kw in (("host", im.ImageNetBatches, {}), ("device", im.ImageNetBatchesGPU, {"device": "cuda:0"})):
        it = iter(cls(d, batch, True, 2, seed=1, files=[shard], **kw))
        next(it)                      # warm-up (thread pool, nvJPEG handles)
        t0 = time.time(); k = 0
        for b in it:
            k += len(b[1])
        dt = time.time() - t0
        print(f"{name:6s}: {k / dt:8.0f} images/s ({k} images in {dt:.2f} s)", flush=True)

if __name__ == "__main__":
    main()
