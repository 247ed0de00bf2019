"""The reference's plumbing model "LRNet" (logist_model.py): a 3072 -> hidden_units -> 10 MLP with a
clipped-log cross-entropy and Adam(--learning_rate), used to debug the distributed wiring without
convolution cost (swapped in at resnet_cifar_main.py:257).  Here: ``--model=lrnet`` on any trainer;
runs on the torch backend, serial or allreduce/PS-style gradient averaging over torch.distributed."""
from __future__ import annotations

import logging
from typing import Dict

import numpy as np
import torch

from ..config import Config

log = logging.getLogger("drn")


class LRNet(torch.nn.Module):
    def __init__(self, image_size: int = 32, hidden_units: int = 100, num_classes: int = 10, seed: int = 0):
        super().__init__()
        g = torch.Generator().manual_seed(seed)
        d = image_size * image_size * 3
        # truncated_normal(stddev=1/image_size) for hid_w, 1/sqrt(hidden) for sm_w (logist_model.py:36-52)
        self.hid_w = torch.nn.Parameter(torch.nn.init.trunc_normal_(torch.empty(d, hidden_units), std=1.0 / image_size, generator=g))
        self.hid_b = torch.nn.Parameter(torch.zeros(hidden_units))
        self.sm_w = torch.nn.Parameter(torch.nn.init.trunc_normal_(torch.empty(hidden_units, num_classes), std=1.0 / hidden_units ** 0.5, generator=g))
        self.sm_b = torch.nn.Parameter(torch.zeros(num_classes))

    def forward(self, x):
        hid = torch.relu(x.reshape(x.shape[0], -1) @ self.hid_w + self.hid_b)
        return torch.softmax(hid @ self.sm_w + self.sm_b, dim=1)

    @staticmethod
    def loss(probs, onehot):
        # cross_entropy = -sum(y * log(clip(p, 1e-10, 1)))   (logist_model.py:57-58)
        return -(onehot * torch.log(probs.clamp(1e-10, 1.0))).sum()


def train_lrnet(cfg: Config) -> Dict[str, float]:
    import torch.distributed as dist
    from ..data import cifar, preprocess
    from ..data.synthetic import synthetic_cifar
    from .loop import init_process_group, make_context
    ctx = make_context(cfg, native=True)   # workers-only process group, gradients averaged
    if ctx.rank < 0:
        return {}
    init_process_group(cfg, ctx, "cpu")
    model = LRNet(cfg.model.image_size, cfg.model.hidden_units, cfg.model.num_classes, cfg.train.seed)
    opt = torch.optim.Adam(model.parameters(), lr=cfg.train.learning_rate)
    if cfg.paths.train_data_path and not cfg.train.synthetic_data:
        imgs, labs = cifar.load_records(cifar.train_filenames(cfg.paths.train_data_path, cfg.model.dataset), cfg.model.dataset)
    else:
        imgs, labs = synthetic_cifar(2048, cfg.model.num_classes, cfg.train.seed)
    it = iter(cifar.CifarBatches(imgs, labs, cfg.train.batch_size, True, None, cfg.train.seed + ctx.rank))
    step, metrics = 0, {}
    while step < cfg.train.train_steps:
        bi, bl = next(it)
        x = torch.from_numpy(np.stack([preprocess.per_image_standardization(
            r.reshape(3, 32, 32).transpose(1, 2, 0)) for r in bi]))
        y = torch.nn.functional.one_hot(torch.from_numpy(bl).long(), cfg.model.num_classes).float()
        probs = model(x)
        loss = LRNet.loss(probs, y)
        opt.zero_grad()
        loss.backward()
        if ctx.distributed:
            for p in model.parameters():
                dist.all_reduce(p.grad)
                p.grad /= ctx.num_workers
        opt.step()
        step += 1
        prec = float((probs.argmax(1) == torch.from_numpy(bl).long()).float().mean())
        metrics = {"cost": float(loss.detach()), "precision": prec, "global_step": step}
        if step % cfg.train.log_every == 0:
            log.info("step = %d, loss = %.4f, precision = %.4f", step, float(loss.detach()), prec)
    return metrics
