"""Stand-in "reference build": the reference's algorithm restated in stock PyTorch - cuDNN/cuBLAS
kernels through ATen, NCCL collectives through torch.distributed - for the same model, batch,
precision (bf16 autocast-style casting, fp32 master weights and BN) and synthetic data as bench.py.

This is NOT the product path and contains none of this repository's kernels; it is the measured
bar the hand-written path is compared against because the real reference (TensorFlow 1.3/1.4 +
Horovod) cannot run on this image.

  allreduce mode: NCCL all_reduce of the flat gradient, replicated momentum-SGD (Horovod path).
  ps mode:        NCCL reduce_scatter of the flat gradient, sharded momentum-SGD, NCCL all_gather
                  of the updated weights (SyncReplicas / parameter-server path).
"""
from __future__ import annotations

import os
import sys

import torch
import torch.distributed as dist
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from distributed_resnet_tensorflow_b200.models.torch_ref import TorchResNet  # noqa: E402

VGG_MEANS = (123.68 / 255.0, 116.78 / 255.0, 103.94 / 255.0)


class TorchBaseline:
    def __init__(self, spec, batch, device, variables, weight_decay, world, mode="allreduce",
                 momentum=0.9):
        self.spec, self.B, self.device, self.world, self.mode = spec, batch, torch.device(device), world, mode
        self.wd, self.mu = weight_decay, momentum
        torch.backends.cudnn.benchmark = True
        self.model = TorchResNet(spec, device=device, dtype=torch.bfloat16, channels_last=True)
        self.model.import_variables(variables)
        self.names = [n for n, _ in spec.trainable_variables()]
        self.mean = torch.tensor(VGG_MEANS, device=self.device).view(1, 1, 1, 3)

    def to_device(self, imgs_u8, labels):
        return imgs_u8.to(self.device, non_blocking=True), labels.to(self.device, non_blocking=True)

    def _preprocess(self, u8):
        s = self.spec.image_size
        x = u8.view(self.B, s, s, 3).float() if self.spec.dataset == "imagenet" else \
            u8.view(self.B, 3, s, s).permute(0, 2, 3, 1).float()
        if self.spec.dataset == "imagenet":
            return x * (1.0 / 255.0) - self.mean
        m = x.mean(dim=(1, 2, 3), keepdim=True)
        sd = x.std(dim=(1, 2, 3), keepdim=True, unbiased=False).clamp_min(1.0 / (3072 ** 0.5))
        return (x - m) / sd

    def step_device(self, batch, lr):
        u8, labels = batch
        m = self.model
        images = self._preprocess(u8)
        cost, xent, prec = m.loss_and_grads(images, labels.long(), self.wd)
        if self.world > 1:
            flat = m.flat_grads()
            if self.mode == "allreduce":
                dist.all_reduce(flat)
                flat /= self.world
                m.set_flat_grads(flat)
                m.apply_gradients(lr, self.mu, self.wd)
            else:
                self._ps_update(flat, lr)
        else:
            m.apply_gradients(lr, self.mu, self.wd)
        return xent

    def _ps_update(self, flat, lr):
        m = self.model
        n = flat.numel()
        per = (n + self.world - 1) // self.world
        pad = per * self.world - n
        if pad:
            flat = F.pad(flat, (0, pad))
        shard = torch.empty(per, device=self.device)
        dist.reduce_scatter_tensor(shard, flat)
        shard /= self.world
        if not hasattr(self, "_w_shard"):
            w = torch.cat([m.params[k].detach().reshape(-1) for k in self.names])
            w = F.pad(w, (0, pad))
            r = dist.get_rank()
            self._w_shard = w[r * per:(r + 1) * per].clone()
            self._m_shard = torch.zeros_like(self._w_shard)
        self._m_shard.mul_(self.mu).add_(shard + self.wd * self._w_shard)
        self._w_shard.add_(self._m_shard, alpha=-lr)
        full = torch.empty(per * self.world, device=self.device)
        dist.all_gather_into_tensor(full, self._w_shard)
        off = 0
        with torch.no_grad():
            for k in self.names:
                p = m.params[k]
                p.copy_(full[off:off + p.numel()].view_as(p))
                off += p.numel()

    def step_host(self, imgs_u8, labels, lr):
        batch = self.to_device(imgs_u8, labels)
        return float(self.step_device(batch, lr))
