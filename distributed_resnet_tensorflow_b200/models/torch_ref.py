"""Plain-PyTorch executor of a ``NetSpec``: the numerical oracle for the sm_100a engine, the CPU
path for plumbing tests, and (on GPU, with cuDNN + NCCL) the stand-in "reference build" that
``baseline/`` times.

Semantics follow the reference exactly: explicit padding for strided
convolutions, BN decay 0.997 / eps 1e-5 with per-replica batch statistics, mean softmax
cross-entropy + wd * sum(0.5*||v||^2) over ALL trainable variables, momentum-SGD without Nesterov
(resnet_model.py:69-121, resnet_model_official.py:41-214).
"""
from __future__ import annotations

from typing import Dict, Iterable, List, Optional, Tuple

import numpy as np
import torch
import torch.nn.functional as F

from .spec import BlockSpec, BNSpec, ConvSpec, NetSpec

BN_DECAY = 0.997
BN_EPS = 1e-5


class TorchResNet:
    """Functional ResNet over a dict of tensors keyed by TF variable name (torch layouts:
    conv OIHW, dense [out, in])."""

    def __init__(self, spec: NetSpec, device="cpu", dtype=torch.float32, channels_last=False):
        self.spec = spec
        self.device = torch.device(device)
        self.dtype = dtype            # compute dtype (bf16 => autocast-like casting of conv/dense)
        self.channels_last = channels_last
        self.params: Dict[str, torch.Tensor] = {}    # trainable, fp32 master
        self.buffers: Dict[str, torch.Tensor] = {}   # BN moving statistics
        self.momentum: Dict[str, torch.Tensor] = {}
        self.global_step = 0
        self._names = [n for n, _ in spec.trainable_variables()]

    # ---- variables in / out (TF layout numpy) -----------------------------------------------
    @staticmethod
    def _to_torch_layout(name: str, arr: np.ndarray) -> np.ndarray:
        if name.endswith("/kernel") or name.endswith("/kernel/Momentum"):
            if arr.ndim == 4:
                return np.ascontiguousarray(arr.transpose(3, 2, 0, 1))   # HWIO -> OIHW
            return np.ascontiguousarray(arr.T)                           # [in,out] -> [out,in]
        return arr

    @staticmethod
    def _to_tf_layout(name: str, arr: np.ndarray) -> np.ndarray:
        if name.endswith("/kernel") or name.endswith("/kernel/Momentum"):
            if arr.ndim == 4:
                return np.ascontiguousarray(arr.transpose(2, 3, 1, 0))   # OIHW -> HWIO
            return np.ascontiguousarray(arr.T)
        return arr

    def import_variables(self, variables: Dict[str, np.ndarray]) -> None:
        for name, _ in self.spec.trainable_variables():
            t = torch.from_numpy(self._to_torch_layout(name, np.asarray(variables[name], np.float32)))
            self.params[name] = t.to(self.device).clone().requires_grad_(True)
            slot = f"{name}/Momentum"
            if slot in variables:
                m = torch.from_numpy(self._to_torch_layout(slot, np.asarray(variables[slot], np.float32)))
                self.momentum[name] = m.to(self.device).clone()
            else:
                self.momentum[name] = torch.zeros_like(self.params[name])
        for name, _ in self.spec.moving_variables():
            self.buffers[name] = torch.from_numpy(np.asarray(variables[name], np.float32)).to(self.device).clone()
        if "global_step" in variables:
            self.global_step = int(variables["global_step"])

    def export_variables(self, with_slots: bool = True) -> Dict[str, np.ndarray]:
        out: Dict[str, np.ndarray] = {}
        for name in self._names:
            out[name] = self._to_tf_layout(name, self.params[name].detach().float().cpu().numpy())
            if with_slots:
                out[f"{name}/Momentum"] = self._to_tf_layout(
                    f"{name}/Momentum", self.momentum[name].detach().float().cpu().numpy())
        for name, t in self.buffers.items():
            out[name] = t.detach().float().cpu().numpy()
        out["global_step"] = np.asarray(self.global_step, np.int64)
        return out

    # ---- layers ---------------------------------------------------------------------------
    def _conv(self, x, c: ConvSpec):
        w = self.params[f"{c.name}/kernel"]
        if self.dtype != torch.float32:
            w = w.to(self.dtype)
        lo, hi = c.pad
        if c.stride > 1 and (lo or hi):
            x = F.pad(x, (lo, hi, lo, hi))       # fixed_padding, then VALID
            return F.conv2d(x, w, stride=c.stride)
        if lo == hi:
            return F.conv2d(x, w, stride=c.stride, padding=lo)
        x = F.pad(x, (lo, hi, lo, hi))
        return F.conv2d(x, w, stride=c.stride)

    def _bn(self, x, b: BNSpec, training: bool, relu: bool = True):
        g, be = self.params[f"{b.name}/gamma"], self.params[f"{b.name}/beta"]
        rm, rv = self.buffers[f"{b.name}/moving_mean"], self.buffers[f"{b.name}/moving_variance"]
        y = F.batch_norm(x.float(), rm, rv, g, be, training=training, momentum=1.0 - BN_DECAY,
                         eps=BN_EPS)
        if relu:
            y = F.relu(y)
        return y.to(self.dtype) if self.dtype != torch.float32 else y

    def _block(self, x, blk: BlockSpec, training: bool):
        if blk.version == 2:
            shortcut = x
            a = self._bn(x, blk.bns[0], training)
            if blk.proj is not None:
                shortcut = self._conv(a, blk.proj)
            h = a
            for i, c in enumerate(blk.convs):
                h = self._conv(h, c)
                if i + 1 < len(blk.convs):
                    h = self._bn(h, blk.bns[i + 1], training)
            return h + shortcut
        shortcut = x
        if blk.proj is not None:
            shortcut = self._bn(self._conv(x, blk.proj), blk.proj_bn, training, relu=False)
        h = x
        for i, (c, b) in enumerate(zip(blk.convs, blk.bns)):
            h = self._bn(self._conv(h, c), b, training, relu=(i + 1 < len(blk.convs)))
        return F.relu(h + shortcut)

    def forward(self, images_nhwc: torch.Tensor, training: bool) -> torch.Tensor:
        """images: [B, H, W, 3] float (already preprocessed). Returns fp32 logits [B, classes]."""
        s = self.spec
        x = images_nhwc.permute(0, 3, 1, 2)
        if self.channels_last:
            x = x.contiguous(memory_format=torch.channels_last)
        x = x.to(self.dtype)
        x = self._conv(x, s.stem)
        if s.stem_bn is not None:
            x = self._bn(x, s.stem_bn, training)
        if s.maxpool:
            x = F.max_pool2d(F.pad(x, (0, 1, 0, 1), value=float("-inf")), 3, 2)  # SAME: 0 before, 1 after
        for blk in s.blocks:
            x = self._block(x, blk, training)
        if s.final_bn is not None:
            x = self._bn(x, s.final_bn, training)
        x = x.float().mean(dim=(2, 3))
        w, b = self.params["dense/kernel"], self.params["dense/bias"]
        if self.dtype != torch.float32:
            return (x.to(self.dtype) @ w.to(self.dtype).t()).float() + b
        return x @ w.t() + b

    # ---- loss / step ------------------------------------------------------------------------
    def l2_loss(self) -> torch.Tensor:
        return sum(0.5 * (p.float() ** 2).sum() for p in self.params.values())

    def loss_and_grads(self, images, labels, weight_decay: float):
        """Returns (cost, xent, precision) and leaves d(cost)/d(param) in ``.grad``; the weight
        decay gradient is NOT included (added in apply_gradients, like the fused native update -
        mathematically identical to differentiating wd * l2_loss)."""
        for p in self.params.values():
            p.grad = None
        logits = self.forward(images, training=True)
        xent = F.cross_entropy(logits, labels)
        xent.backward()
        with torch.no_grad():
            precision = (logits.argmax(dim=1) == labels).float().mean()
            cost = xent + weight_decay * self.l2_loss()
        return cost.detach(), xent.detach(), precision

    def flat_grads(self) -> torch.Tensor:
        return torch.cat([self.params[n].grad.reshape(-1) for n in self._names])

    def set_flat_grads(self, flat: torch.Tensor) -> None:
        off = 0
        for n in self._names:
            p = self.params[n]
            k = p.numel()
            p.grad = flat[off:off + k].view_as(p).clone()
            off += k

    @torch.no_grad()
    def apply_gradients(self, lr: float, momentum: float, weight_decay: float) -> None:
        for n in self._names:
            p, m = self.params[n], self.momentum[n]
            g = p.grad + weight_decay * p
            m.mul_(momentum).add_(g)
            p.add_(m, alpha=-lr)
        self.global_step += 1

    @torch.no_grad()
    def evaluate(self, images, labels) -> Tuple[float, int]:
        logits = self.forward(images, training=False)
        loss = F.cross_entropy(logits, labels)
        return float(loss), int((logits.argmax(dim=1) == labels).sum())
