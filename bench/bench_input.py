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
    shards = []
    for j in range(8):                 # the same records spread over 8 shards for the reader processes
        f = os.path.join(d, "train-%05d-of-01024" % (j + 1)); write_tfrecords(f, recs[j::8] * 4); shards.append(f)
    print(f"{n} JPEGs, {sum(len(r) for r in recs) / n / 1024:.0f} KiB each")
    import torch
    ncpu = os.cpu_count() or 1
    print(f"host cores: {ncpu}")
    arms = [("host, 5 threads (reference's num_threads=5)", im.ImageNetBatches, {}),
            (f"host, {min(8, ncpu)} reader processes x 5 threads", im.MultiProcessBatches, {"workers": min(8, ncpu)}),
            ("device, torchvision nvJPEG binding", im.ImageNetBatchesGPU, {"device": "cuda:0", "decoder": "torchvision"}),
            ("device, libdrn nvJPEG binding (csrc/jpeg.cu), 1 decoder thread", im.ImageNetBatchesGPU,
             {"device": "cuda:0", "decoder": "native", "decode_threads": 1}),
            ("device, libdrn nvJPEG binding, 3 decoder threads", im.ImageNetBatchesGPU,
             {"device": "cuda:0", "decoder": "native", "decode_threads": 3}),
            ("device, libdrn nvJPEG binding, 6 decoder threads", im.ImageNetBatchesGPU,
             {"device": "cuda:0", "decoder": "native", "decode_threads": 6, "num_threads": 8}),
            ("device, libdrn nvJPEG binding, 10 decoder threads", im.ImageNetBatchesGPU,
             {"device": "cuda:0", "decoder": "native", "decode_threads": 10, "num_threads": 8})]
    for name, cls, kw in arms:
        try:
            obj = cls(d, batch, True, 6, seed=1, files=[shard] if cls is not im.MultiProcessBatches else shards, **kw)
        except Exception as e:
            print(f"{name}: unavailable ({e!r})"); continue
        it = iter(obj)
        next(it)                      # warm-up (thread pool, nvJPEG handles, process start)
        torch.cuda.synchronize()
        t0 = time.time(); k = 0
        for b in it:
            k += len(b[1])
        torch.cuda.synchronize()
        dt = time.time() - t0
        extra = f" [{obj.decoder_name}]" if hasattr(obj, "decoder_name") else ""
        print(f"{name}{extra}: {k / dt:8.0f} images/s ({k} images in {dt:.2f} s)", flush=True)
        if hasattr(obj, "close"):
            obj.close()

if __name__ == "__main__":
    main()
