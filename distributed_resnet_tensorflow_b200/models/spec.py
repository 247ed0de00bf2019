"""Architecture descriptions of the ResNets the reference builds, as plain data.

A ``NetSpec`` is a list of layers in *TensorFlow creation order*, which is also what fixes the
checkpoint variable names (``conv2d``, ``conv2d_1`` ... / ``batch_normalization_k`` / ``dense``;
SURVEY.md Appendix A).  Both executors consume it: the PyTorch oracle (models/torch_ref.py) and the
native sm_100a engine (models/engine.py).

Reference: resnet_model_official.py:94-366 (v2 pre-activation blocks, CIFAR 6n+2 and ImageNet
18/34/50/101/152/200).  ``version=1`` adds the original post-activation ResNet the north-star asks
for (conv-BN-ReLU, ReLU after the shortcut add); it has no counterpart in the reference.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

IMAGENET_BLOCKS = {
    18: ("basic", [2, 2, 2, 2]),
    34: ("basic", [3, 4, 6, 3]),
    50: ("bottleneck", [3, 4, 6, 3]),
    101: ("bottleneck", [3, 4, 23, 3]),
    152: ("bottleneck", [3, 8, 36, 3]),
    200: ("bottleneck", [3, 24, 36, 3]),
}


@dataclass(frozen=True)
class ConvSpec:
    name: str       # TF layer name: conv2d, conv2d_1, ...
    cin: int
    cout: int
    ksize: int
    stride: int
    hin: int        # input spatial size (square)

    @property
    def pad(self) -> Tuple[int, int]:
        """(before, after) zero padding - conv2d_fixed_padding, resnet_model_official.py:53-91."""
        total = self.ksize - 1
        return total // 2, total - total // 2

    @property
    def hout(self) -> int:
        lo, hi = self.pad
        return (self.hin + lo + hi - self.ksize) // self.stride + 1

    @property
    def params(self) -> int:
        return self.ksize * self.ksize * self.cin * self.cout

    def macs_per_image(self) -> int:
        return self.hout * self.hout * self.params


@dataclass(frozen=True)
class BNSpec:
    name: str       # batch_normalization, batch_normalization_1, ...
    channels: int


@dataclass(frozen=True)
class BlockSpec:
    kind: str                    # 'basic' | 'bottleneck'
    version: int                 # 1 | 2
    cin: int
    cout: int
    stride: int
    hin: int
    convs: Tuple[ConvSpec, ...]  # main path, in order
    bns: Tuple[BNSpec, ...]      # v2: bn0 (pre-activation of the input) then one per conv except the last
                                 # v1: one per conv (after it)
    proj: Optional[ConvSpec]     # 1x1 projection shortcut (first block of every stage)
    proj_bn: Optional[BNSpec]    # v1 only

    @property
    def hout(self) -> int:
        return self.convs[-1].hout


@dataclass(frozen=True)
class NetSpec:
    dataset: str                 # 'cifar' | 'imagenet'
    resnet_size: int
    version: int
    num_classes: int
    image_size: int
    stem: ConvSpec
    stem_bn: Optional[BNSpec]    # v1 only
    maxpool: bool                # 3x3/2 SAME after the stem (ImageNet)
    blocks: Tuple[BlockSpec, ...]
    final_bn: Optional[BNSpec]   # v2 only
    dense_in: int
    pool_size: int               # final (global) average pool window

    # ---- derived ------------------------------------------------------------------------
    def convs(self) -> List[ConvSpec]:
        """All convolutions in TF creation order."""
        out = [self.stem]
        for b in self.blocks:
            if self.version == 2:
                if b.proj is not None:
                    out.append(b.proj)
                out.extend(b.convs)
            else:
                out.extend(b.convs)
                if b.proj is not None:
                    out.append(b.proj)
        return out

    def bns(self) -> List[BNSpec]:
        out = []
        if self.stem_bn is not None:
            out.append(self.stem_bn)
        for b in self.blocks:
            out.extend(b.bns)
            if b.proj_bn is not None:
                out.append(b.proj_bn)
        if self.final_bn is not None:
            out.append(self.final_bn)
        return out

    def trainable_variables(self) -> List[Tuple[str, Tuple[int, ...]]]:
        """(TF name, TF shape) of every trainable variable in creation order: conv kernels are
        HWIO, dense kernel is [in, out]."""
        order: List[Tuple[str, Tuple[int, ...]]] = []

        def conv(c: ConvSpec):
            order.append((f"{c.name}/kernel", (c.ksize, c.ksize, c.cin, c.cout)))

        def bn(b: BNSpec):
            order.append((f"{b.name}/gamma", (b.channels,)))
            order.append((f"{b.name}/beta", (b.channels,)))

        conv(self.stem)
        if self.stem_bn is not None:
            bn(self.stem_bn)
        for blk in self.blocks:
            if self.version == 2:
                bn(blk.bns[0])
                if blk.proj is not None:
                    conv(blk.proj)
                for i, c in enumerate(blk.convs):
                    conv(c)
                    if i + 1 < len(blk.convs):
                        bn(blk.bns[i + 1])
            else:
                for c, b in zip(blk.convs, blk.bns):
                    conv(c)
                    bn(b)
                if blk.proj is not None:
                    conv(blk.proj)
                    bn(blk.proj_bn)
        if self.final_bn is not None:
            bn(self.final_bn)
        order.append(("dense/kernel", (self.dense_in, self.num_classes)))
        order.append(("dense/bias", (self.num_classes,)))
        return order

    def moving_variables(self) -> List[Tuple[str, Tuple[int, ...]]]:
        out = []
        for b in self.bns():
            out.append((f"{b.name}/moving_mean", (b.channels,)))
            out.append((f"{b.name}/moving_variance", (b.channels,)))
        return out

    def num_params(self) -> int:
        n = 0
        for _, shp in self.trainable_variables():
            k = 1
            for s in shp:
                k *= s
            n += k
        return n

    def forward_macs_per_image(self) -> int:
        return sum(c.macs_per_image() for c in self.convs()) + self.dense_in * self.num_classes


class _Namer:
    def __init__(self):
        self.conv = 0
        self.bn = 0

    def conv_name(self) -> str:
        n = "conv2d" if self.conv == 0 else f"conv2d_{self.conv}"
        self.conv += 1
        return n

    def bn_name(self) -> str:
        n = "batch_normalization" if self.bn == 0 else f"batch_normalization_{self.bn}"
        self.bn += 1
        return n


def _make_block(nm: _Namer, kind: str, version: int, cin: int, filters: int, stride: int, hin: int,
                project: bool) -> BlockSpec:
    cout = 4 * filters if kind == "bottleneck" else filters
    if version == 2:
        bn0 = BNSpec(nm.bn_name(), cin)
        proj = ConvSpec(nm.conv_name(), cin, cout, 1, stride, hin) if project else None
        convs, bns = [], [bn0]
        if kind == "bottleneck":
            c1 = ConvSpec(nm.conv_name(), cin, filters, 1, 1, hin)
            b1 = BNSpec(nm.bn_name(), filters)
            c2 = ConvSpec(nm.conv_name(), filters, filters, 3, stride, hin)
            b2 = BNSpec(nm.bn_name(), filters)
            c3 = ConvSpec(nm.conv_name(), filters, cout, 1, 1, c2.hout)
            convs, bns = [c1, c2, c3], [bn0, b1, b2]
        else:
            c1 = ConvSpec(nm.conv_name(), cin, filters, 3, stride, hin)
            b1 = BNSpec(nm.bn_name(), filters)
            c2 = ConvSpec(nm.conv_name(), filters, filters, 3, 1, c1.hout)
            convs, bns = [c1, c2], [bn0, b1]
        return BlockSpec(kind, 2, cin, cout, stride, hin, tuple(convs), tuple(bns), proj, None)
    # version 1: conv-bn-relu ... ; shortcut projection conv-bn; relu after the add
    convs, bns = [], []
    if kind == "bottleneck":
        c1 = ConvSpec(nm.conv_name(), cin, filters, 1, 1, hin)
        convs.append(c1); bns.append(BNSpec(nm.bn_name(), filters))
        c2 = ConvSpec(nm.conv_name(), filters, filters, 3, stride, hin)
        convs.append(c2); bns.append(BNSpec(nm.bn_name(), filters))
        c3 = ConvSpec(nm.conv_name(), filters, cout, 1, 1, c2.hout)
        convs.append(c3); bns.append(BNSpec(nm.bn_name(), cout))
    else:
        c1 = ConvSpec(nm.conv_name(), cin, filters, 3, stride, hin)
        convs.append(c1); bns.append(BNSpec(nm.bn_name(), filters))
        c2 = ConvSpec(nm.conv_name(), filters, filters, 3, 1, c1.hout)
        convs.append(c2); bns.append(BNSpec(nm.bn_name(), filters))
    proj = proj_bn = None
    if project:
        proj = ConvSpec(nm.conv_name(), cin, cout, 1, stride, hin)
        proj_bn = BNSpec(nm.bn_name(), cout)
    return BlockSpec(kind, 1, cin, cout, stride, hin, tuple(convs), tuple(bns), proj, proj_bn)


def _stages(nm, kind, version, cin, hin, filters_list, blocks_list, strides_list):
    blocks = []
    for filters, nblocks, stride in zip(filters_list, blocks_list, strides_list):
        for i in range(nblocks):
            b = _make_block(nm, kind, version, cin, filters, stride if i == 0 else 1, hin,
                            project=(i == 0))
            blocks.append(b)
            cin, hin = b.cout, b.hout
    return blocks, cin, hin


def cifar_resnet(resnet_size: int = 50, num_classes: int = 10, version: int = 2) -> NetSpec:
    """resnet_model_official.py:217-278 - 6n+2 layers, basic blocks, 16/32/64 filters."""
    if resnet_size % 6 != 2:
        raise ValueError(f"resnet_size must be 6n + 2: {resnet_size}")
    n = (resnet_size - 2) // 6
    nm = _Namer()
    stem = ConvSpec(nm.conv_name(), 3, 16, 3, 1, 32)
    stem_bn = BNSpec(nm.bn_name(), 16) if version == 1 else None
    blocks, cout, hout = _stages(nm, "basic", version, 16, 32, [16, 32, 64], [n, n, n], [1, 2, 2])
    final_bn = BNSpec(nm.bn_name(), cout) if version == 2 else None
    return NetSpec("cifar", resnet_size, version, num_classes, 32, stem, stem_bn, False,
                   tuple(blocks), final_bn, cout, hout)


def imagenet_resnet(resnet_size: int = 50, num_classes: int = 1001, version: int = 2,
                    custom: Optional[Tuple[str, List[int]]] = None) -> NetSpec:
    """resnet_model_official.py:281-366 - 7x7/2 stem, 3x3/2 max pool, 4 stages.
    ``custom=(block kind, blocks per stage)`` builds the same generator with another depth
    (imagenet_resnet_v2_generator accepts any ``layers``)."""
    if custom is not None:
        kind, layers = custom
    else:
        if resnet_size not in IMAGENET_BLOCKS:
            raise ValueError(f"Not a valid resnet_size: {resnet_size}")
        kind, layers = IMAGENET_BLOCKS[resnet_size]
    nm = _Namer()
    stem = ConvSpec(nm.conv_name(), 3, 64, 7, 2, 224)
    stem_bn = BNSpec(nm.bn_name(), 64) if version == 1 else None
    hin = (stem.hout + 1) // 2  # max pool 3x3/2 SAME
    blocks, cout, hout = _stages(nm, kind, version, 64, hin, [64, 128, 256, 512], layers,
                                 [1, 2, 2, 2])
    final_bn = BNSpec(nm.bn_name(), cout) if version == 2 else None
    return NetSpec("imagenet", resnet_size, version, num_classes, 224, stem, stem_bn, True,
                   tuple(blocks), final_bn, cout, hout)


def build_spec(dataset: str, resnet_size: int = 50, version: int = 2,
               num_classes: Optional[int] = None) -> NetSpec:
    if dataset in ("cifar10", "cifar100", "cifar"):
        nc = num_classes if num_classes is not None else (100 if dataset == "cifar100" else 10)
        return cifar_resnet(resnet_size, nc, version)
    if dataset == "imagenet":
        return imagenet_resnet(resnet_size, num_classes if num_classes is not None else 1001, version)
    raise ValueError(f"unknown dataset {dataset!r}")
