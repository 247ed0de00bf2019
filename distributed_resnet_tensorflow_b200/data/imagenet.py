"""ImageNet TFRecord input (reference resnet_imagenet_main.py:103-183).

Pipeline per replica (unsharded, as in the reference): shuffle the shard list (buffer 1024) ->
read records -> parse tf.Example (``image/encoded``, ``image/class/label``) -> decode JPEG (PIL,
worker threads standing in for tf.data's ``num_threads=5``) -> VGG resize + crop on the host ->
shuffle buffer of 1500 examples -> batches of uint8 224x224x3 crops + labels + flip flags.
Scaling, flipping, mean subtraction and the space-to-depth layout run on the device.
"""
from __future__ import annotations

import io
import os
from concurrent.futures import ThreadPoolExecutor
from typing import Iterator, List, Optional, Tuple

import numpy as np

from . import preprocess
from .tfrecord import parse_example, read_tfrecords

FILE_SHUFFLE_BUFFER = 1024
SHUFFLE_BUFFER = 1500
NUM_IMAGES = {"train": 1281167, "validation": 50000}


def filenames(is_training: bool, data_dir: str) -> List[str]:
    if is_training:
        return [os.path.join(data_dir, "train-%05d-of-01024" % i) for i in range(1024)]
    return [os.path.join(data_dir, "validation-%05d-of-00128" % i) for i in range(128)]


def decode_image(encoded: bytes) -> np.ndarray:
    from PIL import Image
    img = Image.open(io.BytesIO(encoded)).convert("RGB")
    return np.asarray(img, dtype=np.uint8)


def parse_record(payload: bytes) -> Tuple[bytes, int]:
    ex = parse_example(payload)
    enc = ex.get("image/encoded", [b""])
    lab = ex.get("image/class/label", [-1])
    return (enc[0] if enc else b""), int(lab[0] if lab else -1)


class ImageNetBatches:
    def __init__(self, data_dir: str, batch_size: int, training: bool, num_epochs: Optional[int] = None,
                 seed: int = 0, num_threads: int = 5, files: Optional[List[str]] = None):
        self.files = files if files is not None else [f for f in filenames(training, data_dir) if os.path.exists(f)]
        if not self.files:
            raise FileNotFoundError(f"no ImageNet TFRecord shards under {data_dir!r}")
        self.B, self.training, self.num_epochs = batch_size, training, num_epochs
        self.rng = np.random.default_rng(seed)
        self.pool = ThreadPoolExecutor(max_workers=max(1, num_threads))

    def _examples(self) -> Iterator[Tuple[np.ndarray, int, int]]:
        epoch = 0
        while self.num_epochs is None or epoch < self.num_epochs:
            files = list(self.files)
            if self.training:
                self.rng.shuffle(files)

            def work(payload, seed):
                enc, lab = parse_record(payload)
                img = decode_image(enc)
                crop, flip = preprocess.vgg_resize_crop_u8(img, self.training, 224,
                                                           np.random.default_rng(seed))
                return crop, lab, flip

            pending = []
            for f in files:
                for payload in read_tfrecords(f):
                    pending.append(self.pool.submit(work, payload, int(self.rng.integers(0, 2 ** 31))))
                    if len(pending) >= 4 * self.B:
                        yield pending.pop(0).result()
            for p in pending:
                yield p.result()
            epoch += 1

    def __iter__(self) -> Iterator[Tuple[np.ndarray, np.ndarray, np.ndarray]]:
        """Yields (images uint8 [B, 224*224*3], labels int32 [B], crop int32 [B,3] = (0,0,flip))."""
        buf: List[Tuple[np.ndarray, int, int]] = []
        out: List[Tuple[np.ndarray, int, int]] = []
        cap = SHUFFLE_BUFFER if self.training else 1

        def emit():
            imgs = np.stack([o[0].reshape(-1) for o in out])
            labs = np.asarray([o[1] for o in out], np.int32)
            crop = np.asarray([[0, 0, o[2]] for o in out], np.int32)
            out.clear()
            return imgs, labs, crop

        for ex in self._examples():
            buf.append(ex)
            if len(buf) >= cap:
                i = int(self.rng.integers(0, len(buf))) if self.training else 0
                buf[i], buf[-1] = buf[-1], buf[i]
                out.append(buf.pop())
                if len(out) == self.B:
                    yield emit()
        if self.training:
            self.rng.shuffle(buf)
        for ex in buf:
            out.append(ex)
            if len(out) == self.B:
                yield emit()
        if out:
            yield emit()


def gpu_decode_available() -> bool:
    """True when JPEG batches can be decoded on the GPU: the library's own nvJPEG binding
    (csrc/jpeg.cu) or, failing that, torchvision's."""
    try:
        from . import jpeg_native
        if jpeg_native.available():
            return True
    except Exception:  # pragma: no cover
        pass
    try:
        import torch
        import torchvision  # noqa: F401
        from torchvision.io import decode_jpeg  # noqa: F401
        return torch.cuda.is_available()
    except Exception:  # pragma: no cover - depends on the image
        return False


class ImageNetBatchesGPU(ImageNetBatches):
    """Same stream of examples, but the pixels never touch the host: records are parsed on worker
    threads (tf.Example only), a whole batch of JPEG byte strings is decoded by nvJPEG on the GPU
    (``torchvision.io.decode_jpeg(..., device=...)``), and ONE launch of ``drn_resize_crop_batch``
    produces the uint8 224x224 crops that ``drn_imagenet_prep`` consumes.  The reference decodes and
    resizes on five CPU threads (resnet_imagenet_main.py:138-168).  Measured (bench/bench_input.py):
    both paths deliver ~1 000 images/s per process - torchvision's binding still Huffman-decodes on
    the host - so this path buys CPU threads, not throughput; a B200 trains at >13 000 img/s, so real
    data needs either many ranks' worth of cores or the nvJPEG hardware backend (not wired up).

    The shuffle buffer holds encoded examples (smaller than decoded crops).  Files nvJPEG rejects
    (CMYK, the PNG hiding in the ImageNet train set) are decoded with PIL and uploaded.
    Yields (images uint8 CUDA [B, 224*224*3], labels int32 [B], crop int32 [B,3] = (0,0,flip))."""

    def __init__(self, data_dir: str, batch_size: int, training: bool, num_epochs: Optional[int] = None,
                 seed: int = 0, num_threads: int = 5, files: Optional[List[str]] = None,
                 device: str = "cuda:0", out_hw: int = 224, decoder: str = "auto",
                 decode_threads: Optional[int] = None):
        super().__init__(data_dir, batch_size, training, num_epochs, seed, num_threads, files)
        import threading
        import torch
        self.device = torch.device(device)
        self.out_hw = out_hw
        self.native = None
        self._tls = threading.local()
        # whole batches are decoded concurrently by `decode_threads` workers, each with its own nvJPEG
        # decoder and CUDA stream (the nvJPEG calls release the GIL); results are yielded in order
        if decode_threads is None:
            # measured (profiles/bench_input_threads_r2.log, 16 host cores): 3 887 / 6 604 / 7 166 / 9 667
            # images/s with 1 / 3 / 6 / 10 decoder threads; keep a few cores for the record parsers
            decode_threads = int(os.environ.get("DRN_DECODE_THREADS", min(8, max(3, (os.cpu_count() or 8) // 4))))
        self.decode_threads = max(1, decode_threads)
        if decoder in ("auto", "native"):
            try:
                from .jpeg_native import NativeJpegDecoder, available
                if available():
                    self.native = NativeJpegDecoder(self.device)
            except Exception:
                if decoder == "native":
                    raise
        if self.native is None:
            self.decode_threads = 1
        self.decoder_name = self.native.backend if self.native is not None else "torchvision nvJPEG binding"

    def _thread_decoder(self):
        d = getattr(self._tls, "dec", None)
        if d is None:
            import torch
            from .jpeg_native import NativeJpegDecoder
            d = self._tls.dec = NativeJpegDecoder(self.device)
            self._tls.stream = torch.cuda.Stream(device=self.device)
        return d

    def _decode_native(self, items):
        """csrc/jpeg.cu: one nvJPEG batch -> planar RGB planes in one device arena."""
        import torch
        arena, offsets, wh, status = self._thread_decoder().decode([enc for enc, _, _ in items])
        base = arena.data_ptr()
        ptr_list, dims, keep = [], [], []
        for i, (enc, _, _) in enumerate(items):
            if status[i] == 0:
                ptr_list.append(base + int(offsets[i]))
                dims.append((int(wh[i, 1]), int(wh[i, 0])))
            else:       # not decodable on the GPU (CMYK, PNG in disguise, ...): host decoder, then upload
                t = torch.from_numpy(decode_image(enc)).permute(2, 0, 1).contiguous().to(self.device)
                keep.append(t)
                ptr_list.append(t.data_ptr())
                dims.append((int(t.shape[1]), int(t.shape[2])))
        return ptr_list, dims, keep

    def _encoded(self) -> Iterator[Tuple[bytes, int, int]]:
        """(jpeg bytes, label, per-example seed) in file order (shuffled shard list when training)."""
        epoch = 0
        while self.num_epochs is None or epoch < self.num_epochs:
            files = list(self.files)
            if self.training:
                self.rng.shuffle(files)
            for f in files:
                for enc, lab in self.pool.map(parse_record, read_tfrecords(f)):
                    yield enc, lab, int(self.rng.integers(0, 2 ** 31))
            epoch += 1

    def _decode_batch(self, items):
        import torch
        if self.native is not None:
            self._thread_decoder()
            with torch.cuda.stream(self._tls.stream):
                return self._decode_batch_on_stream(items)
        return self._decode_batch_on_stream(items)

    def _decode_batch_on_stream(self, items):
        import torch
        from .._native import lib as _nl
        if self.native is None:
            from torchvision.io import ImageReadMode, decode_jpeg
        B = len(items)
        if self.native is not None:
            ptr_list, dims, imgs = self._decode_native(items)
        else:
            bufs = [torch.frombuffer(bytearray(enc), dtype=torch.uint8) for enc, _, _ in items]
            try:
                imgs = decode_jpeg(bufs, device=self.device, mode=ImageReadMode.RGB)
            except Exception:
                imgs = []
                for (enc, _, _), b in zip(items, bufs):
                    try:
                        imgs.append(decode_jpeg(b, device=self.device, mode=ImageReadMode.RGB))
                    except Exception:   # not a baseline/progressive RGB or grey JPEG: host decoder
                        imgs.append(torch.from_numpy(decode_image(enc)).permute(2, 0, 1).contiguous().to(self.device))
            imgs = [im.contiguous() for im in imgs]
            ptr_list = [im.data_ptr() for im in imgs]
            dims = [(int(im.shape[1]), int(im.shape[2])) for im in imgs]
        meta = np.zeros((B, 6), np.int32)
        flips = np.zeros(B, np.int32)
        for i, ((_, _, seed), (h, w)) in enumerate(zip(items, dims)):
            nh, nw, oy, ox, flip = preprocess.vgg_crop_params(h, w, self.training, self.out_hw,
                                                              np.random.default_rng(seed))
            meta[i] = (h, w, nh, nw, oy, ox)
            flips[i] = flip
        ptrs = torch.tensor(ptr_list, dtype=torch.int64).to(self.device)
        dmeta = torch.from_numpy(meta).to(self.device)
        out = torch.empty((B, self.out_hw * self.out_hw * 3), dtype=torch.uint8, device=self.device)
        with torch.cuda.device(self.device):
            _nl.call("drn_resize_crop_batch", ptrs.data_ptr(), dmeta.data_ptr(), out.data_ptr(), B, self.out_hw,
                     torch.cuda.current_stream().cuda_stream)
            torch.cuda.current_stream().synchronize()   # `imgs` may be freed; consumers use other streams
        labs = np.asarray([lab for _, lab, _ in items], np.int32)
        crop = np.stack([np.zeros(B, np.int32), np.zeros(B, np.int32), flips], 1).astype(np.int32)
        return out, labs, crop

    def _encoded_batches(self):
        buf: List[Tuple[bytes, int, int]] = []
        out: List[Tuple[bytes, int, int]] = []
        cap = SHUFFLE_BUFFER if self.training else 1
        for ex in self._encoded():
            buf.append(ex)
            if len(buf) >= cap:
                i = int(self.rng.integers(0, len(buf))) if self.training else 0
                buf[i], buf[-1] = buf[-1], buf[i]
                out.append(buf.pop())
                if len(out) == self.B:
                    yield out
                    out = []
        if self.training:
            self.rng.shuffle(buf)
        for ex in buf:
            out.append(ex)
            if len(out) == self.B:
                yield out
                out = []
        if out:
            yield out

    def __iter__(self):
        if self.decode_threads <= 1:
            for items in self._encoded_batches():
                yield self._decode_batch(items)
            return
        from collections import deque
        pool = ThreadPoolExecutor(max_workers=self.decode_threads)
        pending = deque()
        try:
            for items in self._encoded_batches():
                pending.append(pool.submit(self._decode_batch, items))
                if len(pending) > self.decode_threads:
                    yield pending.popleft().result()
            while pending:
                yield pending.popleft().result()
        finally:
            pool.shutdown(wait=False)


# ------------------------------------------------------------------------------------------------
# reader processes
# ------------------------------------------------------------------------------------------------
def _reader_process(q, data_dir, batch_size, training, num_epochs, seed, num_threads, files):
    try:
        it = ImageNetBatches(data_dir, batch_size, training, num_epochs, seed, num_threads, files)
        for batch in it:
            q.put(batch)
        q.put(None)
    except BaseException as e:   # surfaced in the consumer
        q.put(e)


class MultiProcessBatches:
    """``workers`` reader processes, each running the host pipeline (TFRecord -> tf.Example -> JPEG
    decode -> VGG resize/crop) over its own share of the shard files with its own shuffle stream;
    batches are taken from the processes in turn.  One Python process decodes ~1 000 images/s with
    five threads (the reference's ``num_threads=5``, resnet_imagenet_main.py:166-168) and a B200
    consumes fifteen times that, so the host path needs processes, not threads, to scale with cores.
    ``--input_workers`` selects the count."""

    def __init__(self, data_dir: str, batch_size: int, training: bool, num_epochs: Optional[int] = None,
                 seed: int = 0, workers: int = 4, num_threads: int = 5, files: Optional[List[str]] = None,
                 depth: int = 3):
        import multiprocessing as mp
        files = files if files is not None else [f for f in filenames(training, data_dir) if os.path.exists(f)]
        if not files:
            raise FileNotFoundError(f"no ImageNet TFRecord shards under {data_dir!r}")
        workers = max(1, min(workers, len(files)))
        ctx = mp.get_context("spawn")
        self._procs, self._queues = [], []
        for w in range(workers):
            q = ctx.Queue(maxsize=depth)
            p = ctx.Process(target=_reader_process, daemon=True,
                            args=(q, data_dir, batch_size, training, num_epochs, seed + 7919 * w, num_threads,
                                  files[w::workers]))
            p.start()
            self._procs.append(p)
            self._queues.append(q)

    def __iter__(self):
        live = list(range(len(self._queues)))
        while live:
            for w in list(live):
                item = self._queues[w].get()
                if item is None:
                    live.remove(w)
                    continue
                if isinstance(item, BaseException):
                    self.close()
                    raise item
                yield item

    def close(self) -> None:
        for p in self._procs:
            if p.is_alive():
                p.terminate()

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass
