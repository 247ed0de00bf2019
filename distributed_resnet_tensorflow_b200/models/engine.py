"""Native sm_100a executor of a pre-activation ``NetSpec``: static buffers, one flat parameter
arena, and a fixed schedule of hand-written kernels (csrc/*.cu) that is captured in a CUDA graph.

Data flow of one v2 block (reference resnet_model_official.py:94-175), forward:

    x (raw residual stream, bf16 NHWC)  --stats from producer epilogue-->  bn_finalize
    a0 = bn_apply(x)                               shortcut = x  or  conv1x1(a0)  [projection]
    y1 = conv(a0) [+stats] -> a1 = bn_apply(y1) -> y2 = conv(a1) [+stats] -> a2 -> ...
    out = conv(a_last) + shortcut [+stats for the next block]        (residual add in the epilogue)

backward: every dgrad is the same implicit-GEMM kernel on flipped filters; its epilogue applies
the ReLU/BN mask and accumulates the two BN-backward reductions, ``bn_bwd_apply`` turns the masked
gradient into the gradient of the raw tensor (adding the identity-shortcut gradient on the way),
wgrads accumulate straight into the fp32 gradient arena that the optimizer / collectives consume.

Gradient averaging across replicas, the optimizer step and the PS-style sharded update live in
``parallel/``; this class only exposes the arenas (``master``, ``grad``, ``mom``, ``shadow``).
"""
from __future__ import annotations

import ctypes
import os
from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from .._native import lib as _nl
from .._native.lib import EPI_BNBWD, EPI_BNBWD_ADD, EPI_BNBWD_APPLY, EPI_PLAIN, EPI_RESADD
from ..ops.conv import ConvGeom
from .spec import BlockSpec, BNSpec, ConvSpec, NetSpec

BN_DECAY = 0.997
# filter gradients on a second stream (DRN_WGRAD_STREAM=0 keeps everything on one stream)
WGRAD_SIDE_STREAM = os.environ.get("DRN_WGRAD_STREAM", "1") != "0"
# first BatchNorm+ReLU of a bottleneck block applied inside its consumers (drn_conv_igemm_pre /
# drn_conv_wgrad_pre) instead of being written to memory; DRN_FUSE_BN0=0 restores the separate pass
FUSE_BN0 = os.environ.get("DRN_FUSE_BN0", "1") != "0"
# data gradient of the strided 3x3 convolutions as four parity-phase convolutions (no zero insertion)
DGRAD_S2_PHASES = os.environ.get("DRN_DGRAD_S2", "1") != "0"
# timing experiments only (results are wrong): leave out a class of kernels to see what the step would
# gain if it were free - "wgrad", "bnbwd", "bnfwd", comma separated
_DEBUG_SKIP = frozenset(filter(None, os.environ.get("DRN_DEBUG_SKIP", "").split(",")))


# BatchNorm+ReLU as a prologue of the consuming convolution (and of its filter gradient):
#   DRN_FUSE_PRE=default  the measured policy below (first 1x1 convolution of the 256/512-channel bottlenecks)
#   DRN_FUSE_PRE=all      every convolution for which a prologue variant exists (any filter size, 16/32/64k
#                         channels): no standalone BatchNorm pass is left in the forward pass
#   DRN_FUSE_PRE=none     never (same as DRN_FUSE_BN0=0)
# Measured on one B200 (bench/run_r2_pre.sh, profiles/fuse_pre_all_r2.txt): "all" is correct (oracle tests
# pass with it) and SLOWER everywhere - CIFAR-50 966 -> 710-726 steps/s at 16 images, 534 -> 396 at 128,
# ImageNet-50 14 539 -> 10 660 img/s: a 3x3 layer re-activates its tile once per filter tap, and the
# transform hop (TMA -> four warps -> MMA) adds latency to every k-step of the short 16/32-channel reductions,
# which costs more than the BatchNorm launches it removes.
FUSE_PRE = os.environ.get("DRN_FUSE_PRE", "default")
# the BatchNorm finalize (sums -> scale/shift, moving averages) done by the last CTA of the convolution that
# accumulates the sums, when the consumer of that BatchNorm is prologue-fused (no launch in between).
# Measured: 726 -> 710 steps/s (CIFAR, all fused), noise on ImageNet - the serial tail of the last CTA costs
# what the tiny launch did.  Off by default.
FIN_IN_CONV = os.environ.get("DRN_FIN_IN_CONV", "0") != "0"


def _pre_supported(cin: int, cout: int) -> bool:
    """A prologue variant of the implicit-GEMM kernel exists (csrc/conv_igemm.cu: launch_conv)."""
    if cin == 16:
        return cout in (16, 32)
    if cin == 32:
        return cout in (32, 64)
    return cin % 64 == 0 and cout % 64 == 0


def _fuse_all(blk: BlockSpec, small_images: bool) -> bool:
    return FUSE_PRE == "all"


def _fuse_bn0(blk: BlockSpec, small_images: bool = False) -> bool:
    """Where the prologue variant is a measured win (bench/prof_pre.py and bench.py A/B on one B200):
    1x1 first convolution over 256 or 512 input channels (at 1024 the A tile is re-transformed once
    per 128-column N tile and the gain vanishes); with a projection shortcut only at 256 channels."""
    c0 = blk.convs[0]
    if not FUSE_BN0 or FUSE_PRE == "none":
        return False
    if _fuse_all(blk, small_images):
        return _pre_supported(c0.cin, c0.cout) and (blk.proj is None or _pre_supported(blk.proj.cin, blk.proj.cout))
    if c0.ksize != 1 or c0.stride != 1 or len(blk.convs) != 3:
        return False
    if blk.cin not in (256, 512, 1024) or blk.cin > int(os.environ.get("DRN_FUSE_BN0_MAXC", "512")):
        return False
    return blk.proj is None or blk.cin == 256


def _fuse_inner(blk: BlockSpec, i: int, small_images: bool = False) -> bool:
    """The BatchNorm+ReLU in front of convolution i >= 1 of the block applied by that convolution's
    prologue (and by its filter gradient's) to the raw output of convolution i-1, so the activation is
    never stored.  ImageNet: off (DRN_FUSE_BN2 = largest channel count to fuse in front of a bottleneck's
    last 1x1 convolution; measured no gain, profiles/ceilings_and_fuse_bn2_r2.txt)."""
    c = blk.convs[i]
    if FUSE_PRE == "none":
        return False
    if _fuse_all(blk, small_images):
        return _pre_supported(c.cin, c.cout)
    maxc = int(os.environ.get("DRN_FUSE_BN2", "0"))
    return (len(blk.convs) == 3 and i == 2 and c.ksize == 1 and c.stride == 1 and c.cin % 64 == 0 and c.cin <= maxc)


def _fuse_bnbwd(blk: BlockSpec) -> bool:
    """BatchNorm backward of the block's FIRST BatchNorm finished inside the data-gradient kernel of
    its 1x1 consumer (two passes of the cheap convolution: statistics, then apply) instead of writing
    the masked gradient and running an element-wise pass over the wide tensor.  It moves 2W - N fewer
    bytes per block (W/N = wide/narrow activation), but measured on B200 it does NOT pay: the
    mask+statistics epilogue is bound by its instruction count, not by HBM (a statistics-only pass
    over 256 MB takes 97 us where the store-ing one takes 107 us), so the two passes cost what the
    fused-epilogue + element-wise pair costs (bench/prof_epi2.py: 233-264 us vs 243 us at 56x56,
    140 vs 142 us at 28x28, worse below) and the step time is unchanged (8.80 ms either way).
    Therefore OFF by default; DRN_FUSE_BNBWD=1 enables it for feature maps >= DRN_FUSE_BNBWD_MINHW."""
    c0 = blk.convs[0]
    if os.environ.get("DRN_FUSE_BNBWD", "0") != "1" or c0.ksize != 1 or c0.stride != 1 or len(blk.convs) != 3:
        return False
    if blk.hin < int(os.environ.get("DRN_FUSE_BNBWD_MINHW", "14")):
        return False
    return blk.proj is None or blk.proj.stride == 1


BN_EPS = 1e-5
VGG_MEANS = (123.68 / 255.0, 116.78 / 255.0, 103.94 / 255.0)  # vgg_preprocessing.py:37-39 (in [0,1] units)


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


def padded_classes(n: int) -> int:
    if n <= 16:
        return 16
    if n <= 32:
        return 32
    if n <= 256:
        return _round_up(n, 64)
    return _round_up(n, 256)


# ------------------------------------------------------------------------------------------------
# parameter arena layout
# ------------------------------------------------------------------------------------------------
@dataclass
class ParamEntry:
    name: str
    tf_shape: Tuple[int, ...]
    dev_shape: Tuple[int, ...]
    offset: int
    kind: str  # conv | stem_cifar | stem_s2d | dense_kernel | vector

    @property
    def numel(self) -> int:
        return int(np.prod(self.dev_shape))


def _s2d_index_maps():
    """For the 7x7/2 ImageNet stem rewritten as a 4x4/1 convolution over the space-to-depth input:
    (r, s, c) of the 7x7x3 filter -> (a_y, a_x, ch16) of the 4x4x16 filter."""
    idx = []
    for r in range(7):
        ay, by = divmod(r + 1, 2)
        for s in range(7):
            ax, bx = divmod(s + 1, 2)
            for c in range(3):
                idx.append((r, s, c, ay, ax, (by * 2 + bx) * 3 + c))
    return idx


class ParamLayout:
    ALIGN = 64

    def __init__(self, spec: NetSpec):
        self.spec = spec
        self.entries: Dict[str, ParamEntry] = {}
        self.order: List[str] = []
        off = 0
        stem_name = f"{spec.stem.name}/kernel"
        for name, shp in spec.trainable_variables():
            if name == stem_name:
                if spec.dataset == "imagenet":
                    kind, dshape = "stem_s2d", (shp[3], 4, 4, 16)
                else:
                    kind, dshape = "stem_cifar", (shp[3], shp[0], shp[1], 16)
            elif name.endswith("/kernel") and len(shp) == 4:
                kind, dshape = "conv", (shp[3], shp[0], shp[1], shp[2])
            elif name == "dense/kernel":
                kind, dshape = "dense_kernel", (padded_classes(shp[1]), shp[0])
            else:
                kind, dshape = "vector", tuple(shp)
            e = ParamEntry(name, tuple(shp), dshape, off, kind)
            self.entries[name] = e
            self.order.append(name)
            off = _round_up(off + e.numel, self.ALIGN)
        self.total = _round_up(off, 1024)

    def to_dev(self, name: str, arr: np.ndarray) -> np.ndarray:
        e = self.entries[name]
        arr = np.asarray(arr, np.float32)
        assert tuple(arr.shape) == e.tf_shape, (name, arr.shape, e.tf_shape)
        if e.kind == "conv":
            return np.ascontiguousarray(arr.transpose(3, 0, 1, 2))
        if e.kind == "stem_cifar":
            out = np.zeros(e.dev_shape, np.float32)
            out[:, :, :, :3] = arr.transpose(3, 0, 1, 2)
            return out
        if e.kind == "stem_s2d":
            out = np.zeros(e.dev_shape, np.float32)
            for r, s, c, ay, ax, ch in _s2d_index_maps():
                out[:, ay, ax, ch] = arr[r, s, c, :]
            return out
        if e.kind == "dense_kernel":
            out = np.zeros(e.dev_shape, np.float32)
            out[: arr.shape[1], :] = arr.T
            return out
        return arr

    def to_tf(self, name: str, dev: np.ndarray) -> np.ndarray:
        e = self.entries[name]
        dev = np.asarray(dev, np.float32).reshape(e.dev_shape)
        if e.kind == "conv":
            return np.ascontiguousarray(dev.transpose(1, 2, 3, 0))
        if e.kind == "stem_cifar":
            return np.ascontiguousarray(dev[:, :, :, :3].transpose(1, 2, 3, 0))
        if e.kind == "stem_s2d":
            out = np.zeros(e.tf_shape, np.float32)
            for r, s, c, ay, ax, ch in _s2d_index_maps():
                out[r, s, c, :] = dev[:, ay, ax, ch]
            return out
        if e.kind == "dense_kernel":
            return np.ascontiguousarray(dev[: e.tf_shape[1], :].T)
        return dev.copy()

    def stem_mask(self) -> Optional[np.ndarray]:
        """1 where the s2d stem filter has a real tap, 0 at structural zeros."""
        e = self.entries[f"{self.spec.stem.name}/kernel"]
        if e.kind != "stem_s2d":
            return None
        m = np.zeros(e.dev_shape, np.float32)
        for r, s, c, ay, ax, ch in _s2d_index_maps():
            m[:, ay, ax, ch] = 1.0
        return m


# ------------------------------------------------------------------------------------------------
# engine
# ------------------------------------------------------------------------------------------------
def _p(t) -> int:
    return 0 if t is None else t.data_ptr()


class _BN:
    """Per-BatchNorm device state."""

    def __init__(self, spec: BNSpec, eng: "NativeResNet"):
        self.spec = spec
        C = spec.channels
        self.C = C
        self.gamma = eng.param_view(f"{spec.name}/gamma")
        self.beta = eng.param_view(f"{spec.name}/beta")
        self.dgamma = eng.grad_view(f"{spec.name}/gamma")
        self.dbeta = eng.grad_view(f"{spec.name}/beta")
        self.run_mean = torch.zeros(C, device=eng.device)
        self.run_var = torch.ones(C, device=eng.device)
        self.fstats = eng.alloc_stats(C)
        self.bstats = eng.alloc_stats(C)
        self.ss = torch.zeros((C, 2), device=eng.device)
        self.mi = torch.zeros((C, 2), device=eng.device)
        self.coef = torch.zeros((C, 4), device=eng.device)
        self.fin_counter = eng.alloc_fin_counter()
        self._fin = None

    def fin(self, count: int):
        """drn_bn_fin for a convolution that accumulates this BatchNorm's forward sums over ``count``
        elements per channel (training mode: moving averages are updated)."""
        if self._fin is None or self._fin.count != count:
            self._fin = _nl.BnFin(_p(self.fin_counter), count, _p(self.gamma), _p(self.beta), _p(self.run_mean),
                                  _p(self.run_var), BN_DECAY, BN_EPS, _p(self.ss), _p(self.mi))
        return self._fin


class _Conv:
    def __init__(self, spec: ConvSpec, eng: "NativeResNet", geom: ConvGeom, need_dgrad: bool,
                 s2_phases: bool = False):
        self.spec = spec
        self.g = geom
        self.desc = geom.desc()
        name = f"{spec.name}/kernel"
        self.w = eng.shadow_view(name)
        self.dw = eng.grad_view(name)
        self.wd = None
        self.dg = None
        self.up = None
        self.wd_s2 = None
        if (need_dgrad and s2_phases and DGRAD_S2_PHASES and geom.R == 3 and geom.S == 3 and geom.stride == 2
                and (geom.pad_lo, geom.pad_hi) == (1, 1) and geom.H % 2 == 0 and geom.W % 2 == 0):
            # four phase banks, see drn_conv_dgrad_s2; refreshed by prep_dgrad_weights()
            self.wd_s2 = torch.empty(9 * geom.K * geom.C, dtype=torch.bfloat16, device=eng.device)
            eng._s2_convs.append(self)
        if need_dgrad:
            self.wd = eng.alloc_dgrad_bank(name, geom.K * geom.R * geom.S * geom.C)
            self.dg = geom.dgrad_geom()
            self.dg_desc = self.dg.desc()
            if geom.stride > 1 and geom.R > 1 and self.wd_s2 is None:
                self.up = torch.zeros((geom.N, self.dg.H, self.dg.W, geom.K), dtype=torch.bfloat16,
                                      device=eng.device)


class NativeResNet:
    def __init__(self, spec: NetSpec, batch: int, device="cuda:0", storage: Optional[dict] = None):
        """``storage`` may supply pre-allocated ``grad`` (fp32) / ``shadow`` (bf16) arenas of at
        least ``ParamLayout(spec).total`` elements - the collectives need them in symmetric memory."""
        if spec.version != 2 and type(self) is NativeResNet:
            raise ValueError("post-activation (v1) networks run on models.engine_v1.NativeResNetV1; "
                             "use models.engine.make_engine(spec, ...)")
        _nl.load()
        self.spec = spec
        self.B = batch
        self.device = torch.device(device)
        self.layout = ParamLayout(spec)
        n = self.layout.total
        dev = self.device
        self.master = torch.zeros(n, device=dev)
        storage = storage or {}
        self.grad = storage["grad"][:n] if "grad" in storage else torch.zeros(n, device=dev)
        self.mom = torch.zeros(n, device=dev)
        self.shadow = storage["shadow"][:n] if "shadow" in storage else \
            torch.zeros(n, dtype=torch.bfloat16, device=dev)
        assert self.grad.dtype == torch.float32 and self.shadow.dtype == torch.bfloat16
        self.grad.zero_()
        # parameter-server mode (parallel/reducers.py): gamma / beta / bias are read from an fp32 image
        # of the bf16 working arena instead of the master (only the shard owner's master is current);
        # `pre_forward` refreshes it, `pull_plan` tells each convolution where its filter bank lives
        self.vec_f32 = storage["vec_f32"][:n] if "vec_f32" in storage else None
        self.pre_forward = None
        self.pull_plan = None
        self.hyper = torch.tensor([0.1, 0.9, 1e-4, 1.0], device=dev)   # lr, momentum, wd, grad_scale
        self.scalars = torch.zeros(4, device=dev)       # [0] loss_sum, [1] l2_sum
        self.correct = torch.zeros(1, dtype=torch.int32, device=dev)
        self.global_step = 0
        self._stats_chunks: List[torch.Tensor] = []
        self._stats_arena = torch.zeros(2 * 2 * sum(b.channels for b in spec.bns()) * 2 + 4096, device=dev)
        self._stats_off = 0
        self._fin_counters = torch.zeros(len(list(spec.bns())) + 8, dtype=torch.int32, device=dev)
        self._fin_next = 0
        self._prefinalized = set()      # BatchNorms whose finalize ran inside the producing convolution (this forward)
        self.small_images = spec.dataset != "imagenet"
        self._wd_arena = torch.zeros(self.layout.total + 64 * len(self.layout.order), dtype=torch.bfloat16, device=dev)
        self._wd_off = 0
        self._wd_entries: List[Tuple[str, int]] = []
        self._s2_convs: List["_Conv"] = []
        m = self.layout.stem_mask()
        self.stem_mask = torch.from_numpy(m).to(dev).reshape(-1) if m is not None else None
        self._build()

    # ---- arena views ------------------------------------------------------------------------
    def _view(self, arena, name):
        e = self.layout.entries[name]
        return arena[e.offset:e.offset + e.numel]

    def param_view(self, name):
        return self._view(self.vec_f32 if self.vec_f32 is not None else self.master, name)

    def grad_view(self, name):
        return self._view(self.grad, name)

    def shadow_view(self, name):
        return self._view(self.shadow, name)

    def alloc_dgrad_bank(self, name: str, numel: int) -> torch.Tensor:
        """Slice of the arena holding the flipped/transposed filter banks of all layers (one
        batched kernel refreshes them after every update)."""
        off = self._wd_off
        self._wd_off = _round_up(off + numel, 64)
        assert self._wd_off <= self._wd_arena.numel()
        self._wd_entries.append((name, off))
        return self._wd_arena[off:off + numel]

    def alloc_stats(self, C: int) -> torch.Tensor:
        off = self._stats_off
        self._stats_off += 2 * C
        assert self._stats_off <= self._stats_arena.numel()
        return self._stats_arena[off:off + 2 * C].view(2, C)

    def alloc_fin_counter(self) -> torch.Tensor:
        i = self._fin_next
        self._fin_next += 1
        assert self._fin_next <= self._fin_counters.numel()
        return self._fin_counters[i:i + 1]

    def _act(self, n, h, w, c, dtype=torch.bfloat16):
        return torch.empty((n, h, w, c), dtype=dtype, device=self.device)

    # ---- plan ---------------------------------------------------------------------------------
    def _geom(self, c: ConvSpec, cin_override: Optional[int] = None) -> ConvGeom:
        lo, hi = c.pad
        return ConvGeom(self.B, c.hin, c.hin, cin_override or c.cin, c.cout, c.ksize, c.ksize,
                        c.stride, lo, hi)

    def _build(self):
        s, B = self.spec, self.B
        # stem
        self.stem_view = None
        if s.dataset == "imagenet":
            # space-to-depth stem: x0 is [B,112,116,16] with two zero pixels left and right of each
            # row; viewed as [B,112,112,64] with a W stride of 16 elements, the four horizontal
            # taps of the 4x4 filter become ONE 128-byte TMA row -> a 4x1 convolution, K = 4 x 64.
            self.in_h = s.image_size // 2
            self.in_pitch, self.in_xoff = self.in_h + 4, 2
            stem_geom = ConvGeom(B, self.in_h, self.in_h, 64, s.stem.cout, 4, 1, 1, 2, 1)
            self.stem_view = _nl.ConvView(16, self.in_pitch * 16, self.in_h * self.in_pitch * 16, 0, 0)
            self.x0 = torch.zeros((B, self.in_h, self.in_pitch, 16), dtype=torch.bfloat16, device=self.device)
        else:
            self.in_h = s.image_size
            self.in_pitch, self.in_xoff = self.in_h, 0
            stem_geom = self._geom(s.stem, cin_override=16)
            self.x0 = self._act(B, self.in_h, self.in_h, 16)
        self.labels = torch.zeros(B, dtype=torch.int32, device=self.device)
        self.stem = _Conv(s.stem, self, stem_geom, need_dgrad=False)
        if self.stem_view is not None:
            self.stem.desc.Q = self.in_h       # no padding along W in the view (the borders are data)
        self.stem_P = self.stem_Q = stem_geom.P
        self.y_stem = self._act(B, self.stem_P, self.stem_Q, s.stem.cout)
        if s.maxpool:
            hp = (stem_geom.P + 1) // 2
            self.pool_out = self._act(B, hp, hp, s.stem.cout)
            self.pool_arg = torch.empty((B, hp, hp, s.stem.cout), dtype=torch.uint8, device=self.device)
        # blocks
        self.blocks = []
        for blk in s.blocks:
            self.blocks.append(self._plan_block(blk))
        self.final_bn = _BN(s.final_bn, self)
        self.pooled = torch.empty((B, s.dense_in), dtype=torch.bfloat16, device=self.device)
        kp = padded_classes(s.num_classes)
        self.kp = kp
        self.dense_geom = ConvGeom(B, 1, 1, s.dense_in, kp, 1, 1, 1, 0, 0)
        self.dense_desc = self.dense_geom.desc()
        self.dense_w = self.shadow_view("dense/kernel")
        self.dense_dw = self.grad_view("dense/kernel")
        self.dense_wd = self.alloc_dgrad_bank("dense/kernel", kp * s.dense_in)
        self.dense_dgeom = ConvGeom(B, 1, 1, kp, s.dense_in, 1, 1, 1, 0, 0)
        self.dense_ddesc = self.dense_dgeom.desc()
        self.dense_bias = self.param_view("dense/bias")
        self.dense_dbias = self.grad_view("dense/bias")
        self.logits = torch.empty((B, kp), dtype=torch.bfloat16, device=self.device)
        self.dlogits = torch.empty((B, kp), dtype=torch.bfloat16, device=self.device)
        self.d_pooled = torch.empty((B, s.dense_in), dtype=torch.bfloat16, device=self.device)
        # backward scratch: a few buffers of the largest activation size
        max_elems = max([self.y_stem.numel()] + [b["max_elems"] for b in self.blocks])
        n_scratch = 6 + (4 if WGRAD_SIDE_STREAM else 0)   # longer reuse distance under overlap
        self._scratch = [torch.empty(max_elems, dtype=torch.bfloat16, device=self.device) for _ in range(n_scratch)]
        # filter gradients have no consumer before the optimizer, so they run on a second,
        # lower-priority stream and fill the SMs the persistent data-gradient kernels leave idle in
        # their last round of tiles (1 CTA/SM kernels, 196 tiles on 148 SMs in stage 3, ...)
        self._wg_stream = torch.cuda.Stream(device=self.device, priority=0) if WGRAD_SIDE_STREAM else None
        self._wg_last = None
        self._buf_event = {}
        self._ev_pool, self._ev_next = [], 0

    def _plan_block(self, blk: BlockSpec) -> dict:
        B = self.B
        d = {"spec": blk}
        d["bn0"] = _BN(blk.bns[0], self)
        d["fuse"] = _fuse_bn0(blk, self.small_images)
        d["fuse_bwd"] = _fuse_bnbwd(blk)
        d["fuse_in"] = [d["fuse"]] + [_fuse_inner(blk, i, self.small_images) for i in range(1, len(blk.convs))]
        d["a0"] = None if d["fuse"] else self._act(B, blk.hin, blk.hin, blk.cin)
        sizes = [B * blk.hin * blk.hin * blk.cin]
        if blk.proj is not None:
            d["proj"] = _Conv(blk.proj, self, self._geom(blk.proj), need_dgrad=True)
            pg = d["proj"].g
            d["proj_cdesc"] = ConvGeom(B, pg.P, pg.Q, pg.K, pg.C, 1, 1, 1, 0, 0).desc()
            d["shortcut"] = self._act(B, blk.proj.hout, blk.proj.hout, blk.cout)
            sizes.append(d["shortcut"].numel())
        convs, bns, ys, acts = [], [], [], []
        for i, c in enumerate(blk.convs):
            convs.append(_Conv(c, self, self._geom(c), need_dgrad=True, s2_phases=i >= 1))
            last = i + 1 == len(blk.convs)
            y = self._act(B, c.hout, c.hout, c.cout)
            ys.append(y)
            sizes.append(y.numel())
            if not last:
                bns.append(_BN(blk.bns[i + 1], self))
                fused = d["fuse_in"][i + 1]    # consumed raw by the next convolution's prologue
                acts.append(None if fused else self._act(B, c.hout, c.hout, c.cout))
        d.update(convs=convs, bns=bns, ys=ys, acts=acts, max_elems=max(sizes))
        return d

    # ---- variables in / out -----------------------------------------------------------------
    def import_variables(self, variables: Dict[str, np.ndarray]) -> None:
        host_w = np.zeros(self.layout.total, np.float32)
        host_m = np.zeros(self.layout.total, np.float32)
        for name in self.layout.order:
            e = self.layout.entries[name]
            host_w[e.offset:e.offset + e.numel] = self.layout.to_dev(name, variables[name]).reshape(-1)
            slot = f"{name}/Momentum"
            if slot in variables:
                host_m[e.offset:e.offset + e.numel] = self.layout.to_dev(name, variables[slot]).reshape(-1)
        self.master.copy_(torch.from_numpy(host_w))
        self.mom.copy_(torch.from_numpy(host_m))
        self.grad.zero_()
        for bn in self._all_bns():
            bn.run_mean.copy_(torch.from_numpy(np.asarray(variables[f"{bn.spec.name}/moving_mean"], np.float32)))
            bn.run_var.copy_(torch.from_numpy(np.asarray(variables[f"{bn.spec.name}/moving_variance"], np.float32)))
        self.global_step = int(variables.get("global_step", 0))
        self.refresh_working_weights()

    def export_variables(self, with_slots: bool = True) -> Dict[str, np.ndarray]:
        torch.cuda.synchronize(self.device)
        w = self.master.cpu().numpy()
        m = self.mom.cpu().numpy()
        out: Dict[str, np.ndarray] = {}
        for name in self.layout.order:
            e = self.layout.entries[name]
            out[name] = self.layout.to_tf(name, w[e.offset:e.offset + e.numel])
            if with_slots:
                out[f"{name}/Momentum"] = self.layout.to_tf(name, m[e.offset:e.offset + e.numel])
        for bn in self._all_bns():
            out[f"{bn.spec.name}/moving_mean"] = bn.run_mean.cpu().numpy()
            out[f"{bn.spec.name}/moving_variance"] = bn.run_var.cpu().numpy()
        out["global_step"] = np.asarray(self.global_step, np.int64)
        return out

    def _all_bns(self) -> List[_BN]:
        out = []
        for b in self.blocks:
            out.append(b["bn0"])
            out.extend(b["bns"])
        out.append(self.final_bn)
        return out

    def _all_convs(self) -> List[_Conv]:
        out = [self.stem]
        for b in self.blocks:
            if "proj" in b:
                out.append(b["proj"])
            out.extend(b["convs"])
        return out

    # ---- kernels ------------------------------------------------------------------------------
    @staticmethod
    def _st() -> int:
        return torch.cuda.current_stream().cuda_stream

    def refresh_working_weights(self) -> None:
        """bf16 working copy + flipped dgrad banks from the fp32 master (after import / update)."""
        st = self._st()
        _nl.call("drn_f32_to_bf16", _p(self.master), _p(self.shadow), self.layout.total, st)
        self.prep_dgrad_weights()

    def _build_prep_table(self) -> None:
        rows, tiles = [], 0
        for name, dst_off in self._wd_entries:
            e = self.layout.entries[name]
            if e.kind == "dense_kernel":
                K, RS, C = e.dev_shape[0], 1, e.dev_shape[1]
            else:
                K, RS, C = e.dev_shape[0], e.dev_shape[1] * e.dev_shape[2], e.dev_shape[3]
            rows.append((e.offset, dst_off, K, RS, C, tiles))
            tiles += RS * ((K + 63) // 64) * ((C + 63) // 64)
        tab = np.zeros(len(rows), dtype=[("src", "<i8"), ("dst", "<i8"), ("K", "<i4"), ("RS", "<i4"),
                                         ("C", "<i4"), ("tb", "<i4")])
        for i, r in enumerate(rows):
            tab[i] = r
        self._prep_table = torch.from_numpy(tab.view(np.uint8).copy()).to(self.device)
        self._prep_layers, self._prep_tiles = len(rows), tiles

    def prep_dgrad_weights(self, stream_handle: Optional[int] = None) -> None:
        """Flipped + transposed bf16 banks of every layer that has a data gradient (one launch).
        forward(training=True) issues it on the filter-gradient stream, under the forward pass;
        backward() waits for it."""
        if not hasattr(self, "_prep_table"):
            self._build_prep_table()
        st = self._st() if stream_handle is None else stream_handle
        _nl.call("drn_weight_prep_dgrad_batched", _p(self.shadow), _p(self._wd_arena), _p(self._prep_table),
                 self._prep_layers, self._prep_tiles, st)
        for c in self._s2_convs:
            _nl.call("drn_weight_prep_dgrad_s2", _p(c.w), _p(c.wd_s2), c.g.K, c.g.C, st)

    def _pull(self, conv_name: Optional[str]):
        return self.pull_plan.get(f"{conv_name}/kernel") if (self.pull_plan is not None and conv_name) else None

    def _conv(self, desc, x, w, y, side=None, stats=None, ss=None, mode=EPI_PLAIN, pre=None, pull=None,
              view=None, fin=None):
        """``pre``: (scale, shift) table of a BatchNorm+ReLU to apply to ``x`` on the fly; ``pull``:
        drn_pull_desc of the layer when its weights are still on the parameter-server shards; ``fin``:
        the _BN whose sums go to ``stats`` and whose finalize the kernel's last CTA should do."""
        if fin is not None and stats is not None:
            count = y.numel() // fin.C
            _nl.call("drn_conv_igemm_fin", ctypes.byref(desc), ctypes.byref(view) if view is not None else None,
                     _p(pre), ctypes.byref(pull) if pull is not None else None, ctypes.byref(fin.fin(count)),
                     _p(x), _p(w), _p(y), _p(side), _p(stats), stats.stride(0), _p(ss), mode, 0, self._st())
            self._prefinalized.add(id(fin))
            return
        if pull is not None or view is not None:
            _nl.call("drn_conv_igemm_ex", ctypes.byref(desc), ctypes.byref(view) if view is not None else None,
                     _p(pre), ctypes.byref(pull) if pull is not None else None, _p(x), _p(w), _p(y), _p(side),
                     _p(stats), stats.stride(0) if stats is not None else 0, _p(ss), mode, 0, self._st())
            return
        if pre is not None:
            _nl.call("drn_conv_igemm_pre", ctypes.byref(desc), _p(x), _p(pre), _p(w), _p(y), _p(side), _p(stats),
                     stats.stride(0) if stats is not None else 0, _p(ss), mode, 0, self._st())
            return
        _nl.call("drn_conv_igemm", ctypes.byref(desc), _p(x), _p(w), _p(y), _p(side), _p(stats),
                 stats.stride(0) if stats is not None else 0, _p(ss), mode, 0, self._st())

    def _stem_conv(self, stats, fin=None) -> None:
        self._conv(self.stem.desc, self.x0, self.stem.w, self.y_stem, stats=stats,
                   pull=self._pull(self.spec.stem.name), view=self.stem_view, fin=fin)

    def _fin_for(self, bn: _BN, consumer_fused: bool, training: bool):
        """The BatchNorm to finalize inside the convolution that produces its sums: only when nothing but
        prologue-fused consumers read its tables next (a standalone bn_finalize_apply finalizes itself)."""
        return bn if (FIN_IN_CONV and training and consumer_fused) else None

    def _finalize_once(self, bn: _BN, count: int, training: bool) -> None:
        if id(bn) in self._prefinalized:
            return
        self._bn_finalize(bn, count, training)

    def _event(self):
        if self._ev_next == len(self._ev_pool):
            self._ev_pool.append(torch.cuda.Event())
        ev = self._ev_pool[self._ev_next]
        self._ev_next += 1
        return ev

    def _on_wgrad_stream(self, launch, dy_idx: Optional[int]) -> None:
        """Run ``launch(stream_handle)`` after everything issued so far on the current stream, on the
        filter-gradient stream; remember which scratch buffer it reads so that _get() can order the
        next writer of that buffer behind it."""
        side = self._wg_stream
        if side is None:
            launch(self._st())
            return
        ready = self._event()
        ready.record(torch.cuda.current_stream())
        side.wait_event(ready)
        launch(side.cuda_stream)
        done = self._event()
        done.record(side)
        self._wg_last = done
        if dy_idx is not None:
            self._buf_event[dy_idx] = done

    def last_wgrad_event(self):
        """Event recorded behind the most recent filter-gradient launch (None when everything runs on
        one stream); a consumer of the gradient arena on another stream waits on it."""
        return self._wg_last

    def _join_wgrads(self) -> None:
        if self._wg_last is not None:
            torch.cuda.current_stream().wait_event(self._wg_last)
            self._wg_last = None

    def _wgrad(self, conv: _Conv, x, dy, dy_idx: Optional[int] = None, pre=None):
        if "wgrad" in _DEBUG_SKIP:
            return
        if pre is not None:
            self._on_wgrad_stream(lambda st: _nl.call("drn_conv_wgrad_pre", ctypes.byref(conv.desc), _p(x), _p(pre),
                                                      _p(dy), _p(conv.dw), 0, st), dy_idx)
            return
        self._on_wgrad_stream(lambda st: _nl.call("drn_conv_wgrad", ctypes.byref(conv.desc), _p(x), _p(dy),
                                                  _p(conv.dw), 0, st), dy_idx)

    def _bn_finalize(self, bn: _BN, count: int, training: bool):
        st = self._st()
        if training:
            _nl.call("drn_bn_finalize", _p(bn.fstats), bn.fstats.stride(0), count, _p(bn.gamma),
                     _p(bn.beta), _p(bn.run_mean), _p(bn.run_var), BN_DECAY, BN_EPS, _p(bn.ss),
                     _p(bn.mi), bn.C, st)
        else:
            _nl.call("drn_bn_inference_params", _p(bn.gamma), _p(bn.beta), _p(bn.run_mean),
                     _p(bn.run_var), BN_EPS, _p(bn.ss), bn.C, st)

    def _bn_fwd(self, bn: _BN, y, out, training: bool, relu: bool = True):
        """a = relu(bn(y)): statistics -> scale/shift -> apply, one launch in training mode."""
        count = y.numel() // bn.C
        if "bnfwd" in _DEBUG_SKIP:
            return
        if training:
            _nl.call("drn_bn_finalize_apply", _p(y), _p(bn.fstats), bn.fstats.stride(0), count, _p(bn.gamma),
                     _p(bn.beta), _p(bn.run_mean), _p(bn.run_var), BN_DECAY, BN_EPS, _p(bn.ss), _p(bn.mi),
                     _p(out), bn.C, 1 if relu else 0, self._st())
        else:
            self._bn_finalize(bn, count, False)
            self._bn_apply(bn, y, out, relu=relu)

    def _bn_apply(self, bn: _BN, y, out, addend=None, relu: bool = True):
        M = y.numel() // bn.C
        _nl.call("drn_bn_apply", _p(y), _p(bn.ss), _p(addend), _p(out), M, bn.C, 1 if relu else 0, self._st())

    def _dgrad(self, conv: _Conv, dy, out, side=None, stats=None, ss=None, mode=EPI_PLAIN):
        """Data gradient of ``conv`` given dy; writes [B, H, W, Cin] into ``out``. Strided
        convolutions use the per-layer zero-insertion buffer ``conv.up`` as scratch."""
        _nl.call("drn_conv_dgrad", ctypes.byref(conv.desc), _p(dy), _p(conv.wd), _p(out), _p(side),
                 _p(stats), stats.stride(0) if stats is not None else 0, _p(ss), mode, _p(conv.up),
                 conv.up.numel() * 2 if conv.up is not None else 0, self._st())
        if conv.up is not None:
            _nl.add_launches(1)   # zero-insertion kernel + implicit GEMM

    # ---- forward ------------------------------------------------------------------------------
    def forward(self, training: bool = True) -> None:
        """Consumes self.x0 / self.labels; produces self.logits (and saves what backward needs)."""
        s = self.spec
        st = self._st()
        if self.pre_forward is not None:
            self.pre_forward(self)
        self._prefinalized.clear()
        if training:
            self._stats_arena.zero_()
            self._ev_next = 0
            if self.pull_plan is None:
                self._on_wgrad_stream(self.prep_dgrad_weights, None)
        first_bn = self.blocks[0]["bn0"]
        if s.maxpool:
            self._stem_conv(None)
            g = self.stem.g
            hp = self.pool_out.shape[1]
            _nl.call("drn_maxpool_fwd", _p(self.y_stem), _p(self.pool_out), _p(self.pool_arg), self.B,
                     self.stem_P, self.stem_Q, g.K, 3, 2, 0, hp, hp, st)
            if training:
                _nl.call("drn_bn_stats", _p(self.pool_out), _p(first_bn.fstats), first_bn.fstats.stride(0),
                         self.pool_out.numel() // g.K, g.K, st)
            x = self.pool_out
        else:
            self._stem_conv(first_bn.fstats if training else None,
                            fin=self._fin_for(first_bn, self.blocks[0]["fuse"], training))
            x = self.y_stem
        for i, b in enumerate(self.blocks):
            last_block = i + 1 == len(self.blocks)
            nxt = self.final_bn if last_block else self.blocks[i + 1]["bn0"]
            # (the final BatchNorm is applied inside the pooling kernel: only its tables are needed)
            nxt_fused = True if last_block else self.blocks[i + 1]["fuse"]
            x = self._block_fwd(b, x, nxt, training, nxt_fused)
        self.x_final = x
        fb = self.final_bn
        hw = x.shape[1] * x.shape[2]
        self._finalize_once(fb, x.numel() // fb.C, training)
        _nl.call("drn_bnrelu_avgpool", _p(x), _p(fb.ss), _p(self.pooled), self.B, hw, fb.C, st)
        self._conv(self.dense_desc, self.pooled, self.dense_w, self.logits, pull=self._pull("dense"))
        if training and self.pull_plan is not None:
            # pull mode: the flipped dgrad banks can only be derived once every layer's first consumer
            # has fetched its weights, i.e. after the forward pass
            self._on_wgrad_stream(self.prep_dgrad_weights, None)

    def _block_fwd(self, b: dict, x, next_bn: _BN, training: bool, next_fused: bool = False):
        bn0 = b["bn0"]
        b["x"] = x
        fuse_in = b["fuse_in"]
        if b["fuse"]:
            # only the per-channel scale/shift (and the moving averages) are computed - by the convolution
            # that produced x or here; the consumers of relu(bn0(x)) apply it to their tiles of the raw x
            # in shared memory
            self._finalize_once(bn0, x.numel() // bn0.C, training)
            h, pre = x, bn0.ss
        else:
            self._bn_fwd(bn0, x, b["a0"], training)
            h, pre = b["a0"], None
        shortcut = x
        if "proj" in b:
            pj = b["proj"]
            self._conv(pj.desc, h, pj.w, b["shortcut"], pre=pre, pull=self._pull(pj.spec.name))
            shortcut = b["shortcut"]
        convs = b["convs"]
        for i, c in enumerate(convs):
            last = i + 1 == len(convs)
            y = b["ys"][i]
            if last:
                self._conv(c.desc, h, c.w, y, side=shortcut, mode=EPI_RESADD, pre=pre,
                           stats=next_bn.fstats if training else None, pull=self._pull(c.spec.name),
                           fin=self._fin_for(next_bn, next_fused, training))
            else:
                bn = b["bns"][i]
                fused = fuse_in[i + 1]
                self._conv(c.desc, h, c.w, y, stats=bn.fstats if training else None, pre=pre,
                           pull=self._pull(c.spec.name), fin=self._fin_for(bn, fused, training))
                if fused:
                    self._finalize_once(bn, y.numel() // bn.C, training)
                    h, pre = y, bn.ss
                else:
                    self._bn_fwd(bn, y, b["acts"][i], training)
                    h, pre = b["acts"][i], None
        return b["ys"][-1]

    # ---- loss ---------------------------------------------------------------------------------
    def loss(self, with_grad: bool = True) -> None:
        """scalars[0] += sum of per-example cross-entropies; correct += #hits; dlogits, dbias."""
        _nl.call("drn_softmax_xent", _p(self.logits), self.kp, _p(self.dense_bias), _p(self.labels),
                 self.B, self.spec.num_classes, 1.0 / self.B,
                 _p(self.dlogits) if with_grad else 0, _p(self.dense_dbias) if with_grad else 0, 0,
                 _p(self.scalars), _p(self.correct), self._st())

    # ---- backward -----------------------------------------------------------------------------
    def _get(self, like):
        """Borrow a scratch buffer shaped like ``like`` (stream order makes reuse safe)."""
        i = self._pop_free()
        return i, self._scratch[i][: like.numel()].view(like.shape)

    def _pop_free(self) -> int:
        i = self._free.pop(0)            # FIFO: the buffer released longest ago
        ev = self._buf_event.pop(i, None)
        if ev is not None:               # a filter-gradient launch may still be reading it
            torch.cuda.current_stream().wait_event(ev)
        return i

    def _put(self, i: int) -> None:
        self._free.append(i)

    def range_marks(self) -> List[int]:
        """Arena offsets at which the stem, each block, and the head (final BN + dense) start."""
        ent = self.layout.entries
        marks = [0]
        for b in self.spec.blocks:
            marks.append(ent[f"{b.bns[0].name}/gamma"].offset)
        marks.append(ent[f"{self.spec.final_bn.name}/gamma"].offset)
        marks.append(self.layout.total)
        return marks

    def range_sequence(self) -> List[Tuple[int, int, bool]]:
        """(lo, hi, last) arena ranges in the order backward() completes them."""
        m = self.range_marks()
        seq = [(m[-2], m[-1], False)]
        for i in range(len(self.spec.blocks) - 1, -1, -1):
            seq.append((m[i + 1], m[i + 2], False))
        seq.append((m[0], m[1], True))
        return seq

    def bucket_ranges(self, bucket_elems: int) -> List[Tuple[int, int]]:
        out, pend = [], []
        for lo, hi, last in self.range_sequence():
            pend.append((lo, hi))
            lo_all, hi_all = min(p[0] for p in pend), max(p[1] for p in pend)
            if last or hi_all - lo_all >= bucket_elems:
                pend = []
                out.append((lo_all, hi_all))
        return out

    def backward(self, range_hook=None) -> None:
        """``range_hook(lo, hi, last)`` is called as soon as the gradient arena range [lo, hi) is
        final (head first, stem last) so that collectives can overlap with the rest of backward."""
        s = self.spec
        st = self._st()
        marks = self.range_marks()
        self._free = list(range(len(self._scratch)))
        self._buf_event.clear()
        self._join_wgrads()          # the dgrad filter banks refreshed under forward()
        # dense
        self._conv(self.dense_ddesc, self.dlogits, self.dense_wd, self.d_pooled)
        self._on_wgrad_stream(lambda h: _nl.call("drn_conv_wgrad", ctypes.byref(self.dense_desc), _p(self.pooled),
                                                 _p(self.dlogits), _p(self.dense_dw), 0, h), None)
        # final BN + ReLU + average pool
        fb = self.final_bn
        x = self.x_final
        Bn, H, W, C = x.shape
        dm_i, dm = self._get(x)
        _nl.call("drn_bn_bwd_mask_reduce", 0, _p(self.d_pooled), 3, Bn, H, W, 0, 0, 1.0 / (H * W),
                 _p(x), _p(fb.ss), _p(dm), _p(fb.bstats), fb.bstats.stride(0), C, st)
        g_i, g = self._get(x)
        self._bn_bwd(fb, dm, x, None, g)
        self._put(dm_i)
        # the hooks order their consumer behind last_wgrad_event() themselves: no join here
        if range_hook is not None:
            range_hook(marks[-2], marks[-1], False)
        for bi in range(len(self.blocks) - 1, -1, -1):
            g_i, g = self._block_bwd(self.blocks[bi], g_i, g)
            if range_hook is not None:
                range_hook(marks[bi + 1], marks[bi + 2], False)
        # stem
        if s.maxpool:
            sg = self.stem.g
            hp = self.pool_out.shape[1]
            d_i, d_stem = self._get(self.y_stem)
            _nl.call("drn_maxpool_bwd", _p(g), _p(self.pool_arg), _p(d_stem), self.B, self.stem_P, self.stem_Q,
                     sg.K, 3, 2, 0, hp, hp, st)
            self._put(g_i)
            g_i, g = d_i, d_stem
        if self.stem_view is not None:
            self._on_wgrad_stream(lambda h: _nl.call("drn_conv_wgrad_view", ctypes.byref(self.stem.desc),
                                                     ctypes.byref(self.stem_view), _p(self.x0), _p(g),
                                                     _p(self.stem.dw), 0, h), g_i)
        else:
            self._wgrad(self.stem, self.x0, g, g_i)
        self._put(g_i)
        self._join_wgrads()
        if self.stem_mask is not None:
            _nl.call("drn_mul_inplace_f32", _p(self.stem.dw), _p(self.stem_mask),
                     self.stem_mask.numel(), st)
        if range_hook is not None:
            range_hook(marks[0], marks[1], True)

    def _dgrad_epi2(self, conv: _Conv, dy, out, side, side2, bn: _BN, mode, stats: bool, first: bool = False):
        """Stride-1 data gradient of ``conv`` with one of the two-side epilogues (out=None: statistics only)."""
        _nl.call("drn_conv_igemm_epi2", ctypes.byref(conv.dg_desc), _p(dy), _p(conv.wd), _p(out), _p(side), _p(side2),
                 _p(bn.bstats) if stats else 0, bn.bstats.stride(0) if stats else 0, _p(bn.ss),
                 _p(bn.coef) if mode == EPI_BNBWD_APPLY else 0, 1 if first else 0, mode, 0, self._st())

    def _bn_bwd_coef(self, bn: _BN, count: int):
        _nl.call("drn_bn_bwd_finalize", _p(bn.bstats), bn.bstats.stride(0), count, _p(bn.gamma), _p(bn.mi),
                 _p(bn.dgamma), _p(bn.dbeta), _p(bn.coef), bn.C, self._st())

    def _bn_bwd(self, bn: _BN, dm, y, addend, out):
        if "bnbwd" in _DEBUG_SKIP:
            return
        st = self._st()
        count = y.numel() // bn.C
        _nl.call("drn_bn_bwd_finalize_apply", _p(dm), _p(y), _p(bn.bstats), bn.bstats.stride(0), count,
                 _p(bn.gamma), _p(bn.mi), _p(bn.dgamma), _p(bn.dbeta), _p(addend), _p(out), bn.C, st)

    def _block_bwd(self, b: dict, g_i: int, g):
        """g: gradient w.r.t. the block output (scratch[g_i]); returns (index, tensor) of the
        gradient w.r.t. the block input."""
        st = self._st()
        convs, bns, ys, acts = b["convs"], b["bns"], b["ys"], b["acts"]
        x, a0, bn0 = b["x"], b["a0"], b["bn0"]
        dy_i, dy = g_i, g  # the residual add passes g unchanged to the last conv of the main path
        for i in range(len(convs) - 1, 0, -1):
            c, bn = convs[i], bns[i - 1]
            y_prev, a_prev = ys[i - 1], acts[i - 1]
            if a_prev is None:      # fuse_last: the filter gradient rebuilds relu(bn(y_prev)) itself
                self._wgrad(c, y_prev, dy, dy_i, pre=bn.ss)
            else:
                self._wgrad(c, a_prev, dy, dy_i)
            dm_i, dm = self._get(y_prev)
            if c.wd_s2 is not None:
                # strided 3x3: four parity-phase convolutions over dy, interleaved by the mask kernel
                ph_i, ph = self._get(y_prev)
                _nl.call("drn_conv_dgrad_s2", ctypes.byref(c.desc), _p(dy), _p(c.wd_s2), _p(ph), st)
                _nl.add_launches(3)
                Bn_, Hn_, Wn_, Cn_ = y_prev.shape
                _nl.call("drn_bn_bwd_mask_reduce", _p(ph), 0, 4, Bn_, Hn_, Wn_, 0, 0, 0.0, _p(y_prev), _p(bn.ss),
                         _p(dm), _p(bn.bstats), bn.bstats.stride(0), Cn_, st)
                self._put(ph_i)
            else:
                self._dgrad(c, dy, dm, side=y_prev, stats=bn.bstats, ss=bn.ss, mode=EPI_BNBWD)
            if dy_i != g_i:
                self._put(dy_i)
            n_i, ndy = self._get(y_prev)
            self._bn_bwd(bn, dm, y_prev, None, ndy)
            self._put(dm_i)
            dy_i, dy = n_i, ndy
        c0 = convs[0]
        pre0 = bn0.ss if b["fuse"] else None
        if b["fuse"]:
            a0 = x          # the filter gradients rebuild relu(bn0(x)) from the raw tensor
        self._wgrad(c0, a0, dy, dy_i, pre=pre0)
        Bn, H, W, C = x.shape
        if "proj" in b:
            pj = b["proj"]
            self._wgrad(pj, a0, g, g_i, pre=pre0)
            if b["fuse_bwd"]:
                # t = data gradient through the projection (computed once); the cheap 1x1 main-path
                # gradient is recomputed: pass 1 reduces (t + dgrad) under the mask, pass 2 applies
                t_i, t = self._get(x)
                self._conv(pj.dg_desc, g, pj.wd, t)
                self._put(g_i)
                self._dgrad_epi2(c0, dy, None, x, t, bn0, EPI_BNBWD_ADD, stats=True)
                self._bn_bwd_coef(bn0, x.numel() // C)
                gin_i, gin = self._get(x)
                self._dgrad_epi2(c0, dy, gin, x, t, bn0, EPI_BNBWD_APPLY, stats=False, first=True)
                if dy_i != g_i:
                    self._put(dy_i)
                self._put(t_i)
                return gin_i, gin
            da_i, da = self._get(x)
            self._dgrad(c0, dy, da)
            if dy_i != g_i:
                self._put(dy_i)
            dm_i, dm = self._get(x)
            if pj.g.stride == 1:
                t_i, t = self._get(x)
                self._conv(pj.dg_desc, g, pj.wd, t)
                _nl.call("drn_bn_bwd_mask_reduce", _p(da), _p(t), 1, Bn, H, W, 0, 0, 0.0, _p(x),
                         _p(bn0.ss), _p(dm), _p(bn0.bstats), bn0.bstats.stride(0), C, st)
            else:
                # compact data gradient of the 1x1/stride-s projection: a 1x1/1 conv over g
                t_i = self._pop_free()
                t = self._scratch[t_i][: Bn * pj.g.P * pj.g.Q * C].view(Bn, pj.g.P, pj.g.Q, C)
                self._conv(b["proj_cdesc"], g, pj.wd, t)
                _nl.call("drn_bn_bwd_mask_reduce", _p(da), _p(t), 2, Bn, H, W, pj.g.P, pj.g.Q, 0.0,
                         _p(x), _p(bn0.ss), _p(dm), _p(bn0.bstats), bn0.bstats.stride(0), C, st)
            self._put(da_i)
            self._put(t_i)
            self._put(g_i)
            gin_i, gin = self._get(x)
            self._bn_bwd(bn0, dm, x, None, gin)
            self._put(dm_i)
            return gin_i, gin
        if b["fuse_bwd"]:
            self._dgrad_epi2(c0, dy, None, x, None, bn0, EPI_BNBWD, stats=True)
            self._bn_bwd_coef(bn0, x.numel() // C)
            gin_i, gin = self._get(x)
            self._dgrad_epi2(c0, dy, gin, x, g, bn0, EPI_BNBWD_APPLY, stats=False)
            if dy_i != g_i:
                self._put(dy_i)
            self._put(g_i)
            return gin_i, gin
        dm_i, dm = self._get(x)
        self._dgrad(c0, dy, dm, side=x, stats=bn0.bstats, ss=bn0.ss, mode=EPI_BNBWD)
        if dy_i != g_i:
            self._put(dy_i)
        gin_i, gin = self._get(x)
        self._bn_bwd(bn0, dm, x, g, gin)
        self._put(dm_i)
        self._put(g_i)
        return gin_i, gin

    # ---- optimizer (single replica) -----------------------------------------------------------
    def apply_gradients(self, want_l2: bool = True) -> None:
        st = self._st()
        _nl.call("drn_sgd_momentum", _p(self.master), _p(self.grad), _p(self.mom), _p(self.shadow),
                 self.layout.total, _p(self.hyper), _p(self.scalars[1:]) if want_l2 else 0, 1, st)

    def set_hyper(self, lr: float, momentum: float, weight_decay: float, grad_scale: float = 1.0):
        self.hyper.copy_(torch.tensor([lr, momentum, weight_decay, grad_scale], dtype=torch.float32),
                         non_blocking=True)

    # ---- inputs ---------------------------------------------------------------------------------
    def set_input_float(self, images_nhwc: torch.Tensor, labels: torch.Tensor) -> None:
        """Load already-preprocessed float images [B,H,W,3] (test / oracle-comparison path; the
        production path fills x0 with the device augmentation kernels in data/device_pipeline.py)."""
        x = images_nhwc.to(self.device, torch.float32)
        B, H, W, _ = x.shape
        if self.spec.dataset == "imagenet":
            # space-to-depth: channel = (dy*2+dx)*3 + c
            x = x.view(B, H // 2, 2, W // 2, 2, 3).permute(0, 1, 3, 2, 4, 5).reshape(B, H // 2, W // 2, 12)
            pad = 4
        else:
            pad = 13
        x = torch.nn.functional.pad(x, (0, pad))
        if self.in_pitch != self.in_h:
            self.x0[:, :, self.in_xoff:self.in_xoff + self.in_h, :].copy_(x.to(torch.bfloat16))
        else:
            self.x0.copy_(x.to(torch.bfloat16))
        self.labels.copy_(labels.to(self.device, torch.int32))

    # ---- convenience: one full step (no graph) ----------------------------------------------------
    def train_step_eager(self) -> None:
        self.scalars.zero_()
        self.correct.zero_()
        self.forward(training=True)
        self.loss(with_grad=True)
        self.backward()
        self.apply_gradients()


def make_engine(spec: NetSpec, batch: int, device="cuda:0", storage: Optional[dict] = None):
    """Native executor for ``spec``: pre-activation (v2, the reference's networks) or
    post-activation (v1, the north-star's extra family)."""
    if spec.version == 1:
        from .engine_v1 import NativeResNetV1
        return NativeResNetV1(spec, batch, device=device, storage=storage)
    return NativeResNet(spec, batch, device=device, storage=storage)
