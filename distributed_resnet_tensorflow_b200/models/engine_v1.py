"""Native sm_100a executor for the post-activation ("v1", He et al. 2015) ResNets.

The reference ships only the pre-activation v2 generators (resnet_model_official.py:15-29,94-214);
v1 is the additional model family the north-star asks for (``--resnet_version=1``).
Same kernels, different wiring:

    forward   y_i = conv(h) [+ batch statistics in the epilogue] -> a_i = relu(bn_i(y_i)) -> ...
              out = relu(bn_last(y_last) + shortcut),  shortcut = x  or  bn_p(conv1x1(x))
    backward  dz = g * [out > 0] with the BatchNorm-backward reductions of bn_last (and bn_p) in the
              same pass (drn_relu_bwd_reduce); every other BatchNorm+ReLU backward is the mask +
              reduction epilogue of the data-gradient kernel, exactly as in v2; the gradient of the
              block input is dgrad(conv_0) with the shortcut gradient added in the epilogue.

Checkpoint names follow the same tf.layers auto-naming in creation order (conv, bn, conv, bn, ...,
projection conv, projection bn), see models/spec.py.
"""
from __future__ import annotations

import ctypes
from typing import List

import torch

from .._native import lib as _nl
from .._native.lib import EPI_BNBWD, EPI_RESADD
from ..ops.conv import ConvGeom
from .engine import (NativeResNet, WGRAD_SIDE_STREAM, _BN, _Conv, _p, padded_classes)
from .spec import BlockSpec


class NativeResNetV1(NativeResNet):
    # ---- plan ---------------------------------------------------------------------------------
    def _build(self):
        s, B = self.spec, self.B
        self.stem_view = None
        if s.dataset == "imagenet":
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
            self.stem.desc.Q = self.in_h
        self.stem_P = self.stem_Q = stem_geom.P
        self.y_stem = self._act(B, self.stem_P, self.stem_Q, s.stem.cout)
        self.a_stem = self._act(B, self.stem_P, self.stem_Q, s.stem.cout)
        self.stem_bn = _BN(s.stem_bn, self)
        if s.maxpool:
            hp = (stem_geom.P + 1) // 2
            self.pool_out = self._act(B, hp, hp, s.stem.cout)
            self.pool_arg = torch.empty((B, hp, hp, s.stem.cout), dtype=torch.uint8, device=self.device)
        self.blocks = [self._plan_block(blk) for blk in s.blocks]
        self.final_bn = None
        # global average pool through the fused BN+ReLU+pool kernel with an identity (scale, shift):
        # the last block's output is already activated
        ident = torch.zeros((s.dense_in, 2), device=self.device)
        ident[:, 0] = 1.0
        self.ident_ss = ident
        self.dummy_stats = torch.zeros((2, s.dense_in), device=self.device)
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
        max_elems = max([self.y_stem.numel()] + [b["max_elems"] for b in self.blocks])
        n_scratch = 8 + (4 if WGRAD_SIDE_STREAM else 0)
        self._scratch = [torch.empty(max_elems, dtype=torch.bfloat16, device=self.device) for _ in range(n_scratch)]
        self._wg_stream = torch.cuda.Stream(device=self.device, priority=0) if WGRAD_SIDE_STREAM else None
        self._wg_last = None
        self._buf_event = {}
        self._ev_pool, self._ev_next = [], 0

    def _plan_block(self, blk: BlockSpec) -> dict:
        B = self.B
        d = {"spec": blk}
        sizes = [B * blk.hin * blk.hin * blk.cin]
        convs, bns, ys, acts = [], [], [], []
        for i, c in enumerate(blk.convs):
            convs.append(_Conv(c, self, self._geom(c), need_dgrad=True, s2_phases=i >= 1))
            bns.append(_BN(blk.bns[i], self))
            y = self._act(B, c.hout, c.hout, c.cout)
            ys.append(y)
            sizes.append(y.numel())
            if i + 1 < len(blk.convs):
                acts.append(self._act(B, c.hout, c.hout, c.cout))
        d["out"] = self._act(B, blk.hout, blk.hout, blk.cout)
        if blk.proj is not None:
            pj = _Conv(blk.proj, self, self._geom(blk.proj), need_dgrad=True)
            if pj.g.stride > 1 and pj.up is None:
                # 1x1 / stride-2 projection: its data gradient goes through the generic zero-insertion path
                pj.up = torch.zeros((B, pj.dg.H, pj.dg.W, pj.g.K), dtype=torch.bfloat16, device=self.device)
            d["proj"] = pj
            d["proj_bn"] = _BN(blk.proj_bn, self)
            d["yp"] = self._act(B, blk.hout, blk.hout, blk.cout)
            d["shortcut"] = self._act(B, blk.hout, blk.hout, blk.cout)
        d.update(convs=convs, bns=bns, ys=ys, acts=acts, max_elems=max(sizes + [d["out"].numel()]))
        return d

    def _all_bns(self) -> List[_BN]:
        out = [self.stem_bn]
        for b in self.blocks:
            out.extend(b["bns"])
            if "proj_bn" in b:
                out.append(b["proj_bn"])
        return out

    def range_marks(self) -> List[int]:
        """Arena offsets at which the stem, each block and the head (dense) start."""
        ent = self.layout.entries
        marks = [0]
        for b in self.spec.blocks:
            marks.append(ent[f"{b.convs[0].name}/kernel"].offset)
        marks.append(ent["dense/kernel"].offset)
        marks.append(self.layout.total)
        return marks

    # ---- forward ------------------------------------------------------------------------------
    def forward(self, training: bool = True) -> None:
        s = self.spec
        st = self._st()
        if self.pre_forward is not None:
            self.pre_forward(self)
        if training:
            self._stats_arena.zero_()
            self._ev_next = 0
            if self.pull_plan is None:
                self._on_wgrad_stream(self.prep_dgrad_weights, None)
        self._stem_conv(self.stem_bn.fstats if training else None)
        self._bn_fwd(self.stem_bn, self.y_stem, self.a_stem, training)
        x = self.a_stem
        if s.maxpool:
            g = self.stem.g
            hp = self.pool_out.shape[1]
            _nl.call("drn_maxpool_fwd", _p(self.a_stem), _p(self.pool_out), _p(self.pool_arg), self.B,
                     self.stem_P, self.stem_Q, g.K, 3, 2, 0, hp, hp, st)
            x = self.pool_out
        for b in self.blocks:
            x = self._block_fwd_v1(b, x, training)
        self.x_final = x
        hw = x.shape[1] * x.shape[2]
        _nl.call("drn_bnrelu_avgpool", _p(x), _p(self.ident_ss), _p(self.pooled), self.B, hw, x.shape[3], st)
        self._conv(self.dense_desc, self.pooled, self.dense_w, self.logits, pull=self._pull("dense"))
        if training and self.pull_plan is not None:
            self._on_wgrad_stream(self.prep_dgrad_weights, None)

    def _block_fwd_v1(self, b: dict, x, training: bool):
        b["x"] = x
        shortcut = x
        if "proj" in b:
            pj, bnp = b["proj"], b["proj_bn"]
            self._conv(pj.desc, x, pj.w, b["yp"], stats=bnp.fstats if training else None,
                       pull=self._pull(pj.spec.name))
            self._bn_fwd(bnp, b["yp"], b["shortcut"], training, relu=False)
            shortcut = b["shortcut"]
        h = x
        convs = b["convs"]
        for i, c in enumerate(convs):
            bn, y = b["bns"][i], b["ys"][i]
            self._conv(c.desc, h, c.w, y, stats=bn.fstats if training else None, pull=self._pull(c.spec.name))
            if i + 1 < len(convs):
                self._bn_fwd(bn, y, b["acts"][i], training)
                h = b["acts"][i]
            else:
                self._bn_finalize(bn, y.numel() // bn.C, training)
                self._bn_apply(bn, y, b["out"], addend=shortcut, relu=True)
        return b["out"]

    # ---- backward -----------------------------------------------------------------------------
    def backward(self, range_hook=None) -> None:
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
        if range_hook is not None:
            range_hook(marks[-2], marks[-1], False)
        # average pool: broadcast d_pooled / (H*W) over the plane (the [x > 0] factor of the kernel is
        # the last block's own ReLU mask, which _block_bwd_v1 applies again - idempotent)
        x = self.x_final
        Bn, H, W, C = x.shape
        g_i, g = self._get(x)
        self.dummy_stats.zero_()
        _nl.call("drn_bn_bwd_mask_reduce", 0, _p(self.d_pooled), 3, Bn, H, W, 0, 0, 1.0 / (H * W),
                 _p(x), _p(self.ident_ss), _p(g), _p(self.dummy_stats), self.dummy_stats.stride(0), C, st)
        for bi in range(len(self.blocks) - 1, -1, -1):
            g_i, g = self._block_bwd_v1(self.blocks[bi], g_i, g)
            if range_hook is not None:
                range_hook(marks[bi + 1], marks[bi + 2], False)
        # stem: max pool, BatchNorm + ReLU, convolution
        sg = self.stem.g
        if s.maxpool:
            hp = self.pool_out.shape[1]
            d_i, d_act = self._get(self.y_stem)
            _nl.call("drn_maxpool_bwd", _p(g), _p(self.pool_arg), _p(d_act), self.B, self.stem_P, self.stem_Q,
                     sg.K, 3, 2, 0, hp, hp, st)
            self._put(g_i)
            g_i, g = d_i, d_act
        bn = self.stem_bn
        dm_i, dm = self._get(self.y_stem)
        _nl.call("drn_bn_bwd_mask_reduce", _p(g), 0, 0, self.B, self.stem_P, self.stem_Q, 0, 0, 0.0,
                 _p(self.y_stem), _p(bn.ss), _p(dm), _p(bn.bstats), bn.bstats.stride(0), sg.K, st)
        self._put(g_i)
        g_i, g = self._get(self.y_stem)
        self._bn_bwd(bn, dm, self.y_stem, None, g)
        self._put(dm_i)
        if self.stem_view is not None:
            self._on_wgrad_stream(lambda h: _nl.call("drn_conv_wgrad_view", ctypes.byref(self.stem.desc),
                                                     ctypes.byref(self.stem_view), _p(self.x0), _p(g),
                                                     _p(self.stem.dw), 0, h), g_i)
        else:
            self._wgrad(self.stem, self.x0, g, g_i)
        self._put(g_i)
        self._join_wgrads()
        if self.stem_mask is not None:
            _nl.call("drn_mul_inplace_f32", _p(self.stem.dw), _p(self.stem_mask), self.stem_mask.numel(), st)
        if range_hook is not None:
            range_hook(marks[0], marks[1], True)

    def _block_bwd_v1(self, b: dict, g_i: int, g):
        """g = gradient w.r.t. the block output (after its ReLU); returns (index, tensor) of the
        gradient w.r.t. the block input."""
        st = self._st()
        convs, bns, ys, acts = b["convs"], b["bns"], b["ys"], b["acts"]
        x, out = b["x"], b["out"]
        last = len(convs) - 1
        has_proj = "proj" in b
        M = out.numel() // out.shape[3]
        C = out.shape[3]
        dz_i, dz = self._get(out)
        bnl = bns[last]
        _nl.call("drn_relu_bwd_reduce", _p(g), _p(out), _p(ys[last]), _p(b["yp"]) if has_proj else 0, _p(dz),
                 _p(bnl.bstats), _p(b["proj_bn"].bstats) if has_proj else 0, bnl.bstats.stride(0), M, C, st)
        self._put(g_i)
        dy_i, dy = self._get(out)
        self._bn_bwd(bnl, dz, ys[last], None, dy)
        if has_proj:
            dyp_i, dyp = self._get(out)
            self._bn_bwd(b["proj_bn"], dz, b["yp"], None, dyp)
        for i in range(last, 0, -1):
            c, bn = convs[i], bns[i - 1]
            y_prev, a_prev = ys[i - 1], acts[i - 1]
            self._wgrad(c, a_prev, dy, dy_i)
            dm_i, dm = self._get(y_prev)
            if c.wd_s2 is not None:
                ph_i, ph = self._get(y_prev)
                _nl.call("drn_conv_dgrad_s2", ctypes.byref(c.desc), _p(dy), _p(c.wd_s2), _p(ph), st)
                _nl.add_launches(3)
                Bn_, Hn_, Wn_, Cn_ = y_prev.shape
                _nl.call("drn_bn_bwd_mask_reduce", _p(ph), 0, 4, Bn_, Hn_, Wn_, 0, 0, 0.0, _p(y_prev), _p(bn.ss),
                         _p(dm), _p(bn.bstats), bn.bstats.stride(0), Cn_, st)
                self._put(ph_i)
            else:
                self._dgrad(c, dy, dm, side=y_prev, stats=bn.bstats, ss=bn.ss, mode=EPI_BNBWD)
            self._put(dy_i)
            n_i, ndy = self._get(y_prev)
            self._bn_bwd(bn, dm, y_prev, None, ndy)
            self._put(dm_i)
            dy_i, dy = n_i, ndy
        c0 = convs[0]
        self._wgrad(c0, x, dy, dy_i)
        gin_i, gin = self._get(x)
        if has_proj:
            pj = b["proj"]
            self._wgrad(pj, x, dyp, dyp_i)
            t_i, t = self._get(x)
            self._dgrad(c0, dy, t)
            self._dgrad(pj, dyp, gin, side=t, mode=EPI_RESADD)
            self._put(t_i)
            self._put(dyp_i)
        else:
            self._dgrad(c0, dy, gin, side=dz, mode=EPI_RESADD)   # identity shortcut: + dz in the epilogue
        self._put(dz_i)
        self._put(dy_i)
        return gin_i, gin
