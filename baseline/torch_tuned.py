"""Tuned stock-PyTorch arm: what a careful PyTorch user gets WITHOUT any of this repository's
kernels - torch.nn modules in channels_last, bf16 autocast (cuDNN / cuBLAS through ATen), fused
momentum-SGD, DistributedDataParallel (NCCL) for N > 1, and the whole step (forward, backward,
optimizer) captured in ONE CUDA graph.  It replaces the eager restatement in torch_baseline.py as the
bar the hand-written path is compared against (the eager stand-in overstated the margin).

Same network (the reference's ResNet v2, resnet_model_official.py:94-366, built from models/spec.py),
same loss (mean cross-entropy + L2 on every trainable variable via the optimizer's weight_decay),
same batch, same synthetic uint8 input with the preprocessing done on the device.
"""
from __future__ import annotations

import os
import sys

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

VGG_MEANS = (123.68 / 255.0, 116.78 / 255.0, 103.94 / 255.0)


def _conv(c):
    lo, hi = c.pad
    if lo == hi:
        return nn.Conv2d(c.cin, c.cout, c.ksize, c.stride, lo, bias=False)
    # fixed_padding of the reference (resnet_model_official.py:53-77): explicit asymmetric pad, VALID conv
    return nn.Sequential(nn.ZeroPad2d((lo, hi, lo, hi)), nn.Conv2d(c.cin, c.cout, c.ksize, c.stride, 0, bias=False))


def _bn(b):
    return nn.BatchNorm2d(b.channels, eps=1e-5, momentum=1.0 - 0.997)


class _BlockV2(nn.Module):
    def __init__(self, blk):
        super().__init__()
        self.bn0 = _bn(blk.bns[0])
        self.proj = _conv(blk.proj) if blk.proj is not None else None
        self.convs = nn.ModuleList([_conv(c) for c in blk.convs])
        self.bns = nn.ModuleList([_bn(b) for b in blk.bns[1:]])

    def forward(self, x):
        a = F.relu(self.bn0(x))
        shortcut = self.proj(a) if self.proj is not None else x
        h = a
        for i, conv in enumerate(self.convs):
            h = conv(h)
            if i + 1 < len(self.convs):
                h = F.relu(self.bns[i](h))
        return h + shortcut


class ResNetV2(nn.Module):
    def __init__(self, spec):
        super().__init__()
        assert spec.version == 2
        self.spec = spec
        self.stem = _conv(spec.stem)
        self.blocks = nn.Sequential(*[_BlockV2(b) for b in spec.blocks])
        self.final_bn = _bn(spec.final_bn)
        self.dense = nn.Linear(spec.dense_in, spec.num_classes)

    def forward(self, x):
        x = self.stem(x)
        if self.spec.maxpool:
            x = F.max_pool2d(F.pad(x, (0, 1, 0, 1), value=float("-inf")), 3, 2)
        x = self.blocks(x)
        x = F.relu(self.final_bn(x))
        return self.dense(torch.flatten(F.adaptive_avg_pool2d(x, 1), 1))


class TorchTuned:
    def __init__(self, spec, batch, device, weight_decay, world, momentum=0.9, lr=0.1, graph=True):
        self.spec, self.B, self.device, self.world = spec, batch, torch.device(device), world
        torch.backends.cudnn.benchmark = True
        torch.manual_seed(0)
        model = ResNetV2(spec).to(self.device).to(memory_format=torch.channels_last)
        self.model = model
        self.ddp = None
        if world > 1:
            from torch.nn.parallel import DistributedDataParallel as DDP
            self.ddp = DDP(model, device_ids=[self.device.index], gradient_as_bucket_view=True, static_graph=True)
        self.opt = torch.optim.SGD(model.parameters(), lr=lr, momentum=momentum, weight_decay=weight_decay,
                                   fused=True)
        self.mean = torch.tensor(VGG_MEANS, device=self.device).view(1, 3, 1, 1)
        s = spec.image_size
        self.static_u8 = torch.zeros((batch, s * s * 3), dtype=torch.uint8, device=self.device)
        self.static_lab = torch.zeros(batch, dtype=torch.int64, device=self.device)
        self.static_loss = torch.zeros((), device=self.device)
        self.graph = None
        self.want_graph = graph
        self.graphed = False

    def _preprocess(self, u8):
        s = self.spec.image_size
        if self.spec.dataset == "imagenet":
            x = u8.view(self.B, s, s, 3).permute(0, 3, 1, 2).float()
            x = x * (1.0 / 255.0) - self.mean
        else:
            x = u8.view(self.B, 3, s, s).float()
            m = x.mean(dim=(1, 2, 3), keepdim=True)
            sd = x.std(dim=(1, 2, 3), keepdim=True, unbiased=False).clamp_min(1.0 / (3072 ** 0.5))
            x = (x - m) / sd
        return x.contiguous(memory_format=torch.channels_last)

    def _body(self):
        net = self.ddp if self.ddp is not None else self.model
        with torch.autocast("cuda", dtype=torch.bfloat16):
            logits = net(self._preprocess(self.static_u8))
        loss = F.cross_entropy(logits.float(), self.static_lab)
        loss.backward()
        self.opt.step()
        self.static_loss.copy_(loss.detach())

    def _eager_step(self):
        self.opt.zero_grad(set_to_none=True)
        self._body()

    def prepare(self):
        """Warm up (cuDNN autotune, DDP bucket rebuild) and capture the step in a CUDA graph."""
        s = torch.cuda.Stream(device=self.device)
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(12 if self.ddp is not None else 4):
                self._eager_step()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        if not self.want_graph:
            return
        try:
            g = torch.cuda.CUDAGraph()
            self.opt.zero_grad(set_to_none=True)
            with torch.cuda.graph(g):
                self._body()
            self.graph = g
            self.graphed = True
        except Exception as e:           # e.g. NCCL capture refused: stay eager, say so
            self.graph = None
            self.graph_error = repr(e)[:200]
            torch.cuda.synchronize()

    def load(self, u8, labels):
        """Feed the next batch: device tensors are copied into the graph's static inputs; pinned host
        tensors go through a double-buffered H2D copy on a side stream first (overlaps the running step)."""
        if u8.is_cuda:
            self.static_u8.copy_(u8.view(self.B, -1), non_blocking=True)
            self.static_lab.copy_(labels, non_blocking=True)
            return
        if not hasattr(self, "_stage"):
            self._stage = [(torch.empty_like(self.static_u8), torch.empty(self.B, dtype=labels.dtype, device=self.device))
                           for _ in range(2)]
            self._copy_stream = torch.cuda.Stream(device=self.device)
            self._stage_ev = [torch.cuda.Event(), torch.cuda.Event()]
            self._used_ev = [None, None]
            self._slot = 0
        i = self._slot
        self._slot ^= 1
        cs = self._copy_stream
        if self._used_ev[i] is not None:
            cs.wait_event(self._used_ev[i])
        with torch.cuda.stream(cs):
            self._stage[i][0].copy_(u8.view(self.B, -1), non_blocking=True)
            self._stage[i][1].copy_(labels, non_blocking=True)
            self._stage_ev[i].record(cs)
        cur = torch.cuda.current_stream()
        cur.wait_event(self._stage_ev[i])
        self.static_u8.copy_(self._stage[i][0], non_blocking=True)
        self.static_lab.copy_(self._stage[i][1], non_blocking=True)
        ev = torch.cuda.Event()
        ev.record(cur)
        self._used_ev[i] = ev

    def step(self):
        if self.graph is not None:
            self.graph.replay()
        else:
            self._eager_step()
        return self.static_loss
