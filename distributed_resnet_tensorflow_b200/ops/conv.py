"""Convolution ops backed by the tcgen05 implicit-GEMM kernels (csrc/conv_igemm.cu, conv_wgrad.cu).

Layouts: activations NHWC bf16, filters KRSC bf16 (``[Cout, R, S, Cin]``), filter gradients
KRSC fp32. The reference builds these with ``tf.layers.conv2d`` + explicit ``tf.pad`` for strided
convolutions (resnet_model_official.py:53-91); here padding is part of the TMA descriptor.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass

import torch

from .._native import lib as _nl
from .._native.lib import ConvDesc, EPI_BNBWD, EPI_PLAIN, EPI_RESADD  # noqa: F401


@dataclass(frozen=True)
class ConvGeom:
    """Static geometry of one convolution (per-GPU batch included)."""
    N: int
    H: int
    W: int
    C: int
    K: int
    R: int
    S: int
    stride: int
    pad_lo: int
    pad_hi: int

    @property
    def P(self) -> int:
        return (self.H + self.pad_lo + self.pad_hi - self.R) // self.stride + 1

    @property
    def Q(self) -> int:
        return (self.W + self.pad_lo + self.pad_hi - self.S) // self.stride + 1

    def desc(self) -> ConvDesc:
        return ConvDesc(self.N, self.H, self.W, self.C, self.K, self.R, self.S, self.stride,
                        self.pad_lo, self.pad_hi, self.P, self.Q)

    @staticmethod
    def tf_fixed_padding(N, H, W, C, K, ksize, stride) -> "ConvGeom":
        """Padding rule of the reference's conv2d_fixed_padding (resnet_model_official.py:80-91):
        stride 1 -> SAME; stride > 1 -> explicit (k-1)//2 before, rest after, then VALID."""
        total = ksize - 1
        lo = total // 2
        hi = total - lo
        return ConvGeom(N, H, W, C, K, ksize, ksize, stride, lo, hi)

    def dgrad_geom(self) -> "ConvGeom":
        """Geometry of the stride-1 convolution over dy (zero-upsampled when stride > 1) that
        yields dx; its filters are the flipped/transposed bank from ``weight_prep``."""
        # dx[h] = sum_r' dy_up[h + r' - (R-1-pad_lo)] * Wflip[r'],  dy_up[u] = dy[u/stride] (u % stride == 0)
        uh = (self.P - 1) * self.stride + 1
        uw = (self.Q - 1) * self.stride + 1
        pad_hi_h = self.H - uh + self.pad_lo
        pad_hi_w = self.W - uw + self.pad_lo
        assert pad_hi_h == pad_hi_w and self.R == self.S, "square geometry expected"
        return ConvGeom(self.N, uh, uw, self.K, self.C, self.R, self.S, 1,
                        self.R - 1 - self.pad_lo, pad_hi_h)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def conv_fprop(x, w, g: ConvGeom, out=None, side=None, stats=None, bn_scale_shift=None,
               mode: int = EPI_PLAIN, force_bn: int = 0):
    """y = conv(x, w) [+ side | masked by BN/ReLU backward]; optional per-channel sum / sumsq
    accumulated into ``stats`` (float32 ``[2, ld]``, must be zeroed by the caller)."""
    assert x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16
    assert x.is_contiguous() and w.is_contiguous()
    assert tuple(x.shape) == (g.N, g.H, g.W, g.C), (tuple(x.shape), g)
    assert w.numel() == g.K * g.R * g.S * g.C
    if out is None:
        out = torch.empty((g.N, g.P, g.Q, g.K), dtype=torch.bfloat16, device=x.device)
    stats_ld = 0
    if stats is not None:
        assert stats.dtype == torch.float32 and stats.dim() == 2 and stats.shape[0] == 2
        stats_ld = stats.stride(0)
    d = g.desc()
    _nl.call("drn_conv_igemm", ctypes.byref(d), _ptr(x), _ptr(w), _ptr(out), _ptr(side),
             _ptr(stats), stats_ld, _ptr(bn_scale_shift), mode, force_bn, _stream())
    return out


def conv_wgrad(x, dy, g: ConvGeom, dw=None, split_k: int = 0):
    """dw[K,R,S,C] (fp32) += x (*) dy. Allocates a zeroed dw when none is given."""
    assert x.dtype == torch.bfloat16 and dy.dtype == torch.bfloat16
    assert x.is_contiguous() and dy.is_contiguous()
    if dw is None:
        dw = torch.zeros((g.K, g.R, g.S, g.C), dtype=torch.float32, device=x.device)
    d = g.desc()
    _nl.call("drn_conv_wgrad", ctypes.byref(d), _ptr(x), _ptr(dy), _ptr(dw), split_k, _stream())
    return dw


def conv_dgrad(dy, w_dgrad, g: ConvGeom, out=None, side=None, stats=None, bn_scale_shift=None,
               mode: int = EPI_PLAIN, scratch=None):
    """dx[N,H,W,C] for the forward geometry ``g`` (filters prepared by drn_weight_prep_dgrad).
    ``scratch`` (bf16) is required for stride > 1; allocated here when missing."""
    d = g.desc()
    lib = _nl.load()
    need = lib.drn_conv_dgrad_scratch_bytes(ctypes.byref(d))
    if need and scratch is None:
        scratch = torch.empty(need // 2, dtype=torch.bfloat16, device=dy.device)
    if out is None:
        out = torch.empty((g.N, g.H, g.W, g.C), dtype=torch.bfloat16, device=dy.device)
    stats_ld = stats.stride(0) if stats is not None else 0
    _nl.call("drn_conv_dgrad", ctypes.byref(d), _ptr(dy), _ptr(w_dgrad), _ptr(out), _ptr(side), _ptr(stats),
             stats_ld, _ptr(bn_scale_shift), mode, _ptr(scratch), scratch.numel() * 2 if scratch is not None else 0,
             _stream())
    return out


def conv_fprop_epi2(x, w, g: ConvGeom, side, side2, out=None, stats=None, bn_scale_shift=None, bn_coef=None,
                    side2_first: bool = False, mode: int = _nl.EPI_BNBWD_ADD, store: bool = True, force_bn: int = 0):
    """Two-side epilogues of the implicit-GEMM kernel (drn_conv_igemm_epi2): BNBWD_ADD / BNBWD_APPLY, or a
    statistics-only BNBWD / BNBWD_ADD pass with ``store=False`` (returns None)."""
    if store and out is None:
        out = torch.empty((g.N, g.P, g.Q, g.K), dtype=torch.bfloat16, device=x.device)
    stats_ld = stats.stride(0) if stats is not None else 0
    d = g.desc()
    _nl.call("drn_conv_igemm_epi2", ctypes.byref(d), _ptr(x), _ptr(w), _ptr(out) if store else _ptr(None), _ptr(side),
             _ptr(side2), _ptr(stats), stats_ld, _ptr(bn_scale_shift), _ptr(bn_coef), 1 if side2_first else 0, mode,
             force_bn, _stream())
    return out if store else None
