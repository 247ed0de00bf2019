"""Numerics + timing probe of the tcgen05 conv kernels against torch (cuDNN fp32) references.

Run on a B200: ``python bench/probe_conv.py [--start I] [--only fprop|wgrad] [--time]``.
Each case prints one line; a CUDA fault ends the process with exit code 3 after printing the
index so that ``bench/run_probe.sh`` can resume at the next case in a fresh process.
"""
from __future__ import annotations

import argparse
import os
import sys
import time

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from distributed_resnet_tensorflow_b200.ops.conv import (EPI_BNBWD, EPI_PLAIN, EPI_RESADD, ConvGeom,
                                                         conv_fprop, conv_wgrad)
from distributed_resnet_tensorflow_b200._native import lib as nl

# (name, N, H, C, K, ksize, stride, pad_lo, pad_hi, mode, stats, force_bn)
FPROP_CASES = [
    ("1x1 c64 k64 M128", 2, 8, 64, 64, 1, 1, 0, 0, EPI_PLAIN, False, 0),
    ("1x1 c64 k64 M512", 8, 8, 64, 64, 1, 1, 0, 0, EPI_PLAIN, False, 0),
    ("1x1 c256 k256", 8, 8, 256, 256, 1, 1, 0, 0, EPI_PLAIN, False, 0),
    ("1x1 c256 k128", 8, 8, 256, 128, 1, 1, 0, 0, EPI_PLAIN, False, 0),
    ("1x1 c128 k512 bn256", 8, 8, 128, 512, 1, 1, 0, 0, EPI_PLAIN, False, 0),
    ("3x3 c64 k64", 2, 16, 64, 64, 3, 1, 1, 1, EPI_PLAIN, False, 0),
    ("3x3 c128 k128 h14", 4, 14, 128, 128, 3, 1, 1, 1, EPI_PLAIN, False, 0),
    ("3x3 s2 c64 k128", 2, 16, 64, 128, 3, 2, 1, 1, EPI_PLAIN, False, 0),
    ("1x1 s2 c64 k256", 2, 16, 64, 256, 1, 2, 0, 0, EPI_PLAIN, False, 0),
    ("3x3 c16 k16", 4, 16, 16, 16, 3, 1, 1, 1, EPI_PLAIN, False, 0),
    ("3x3 c32 k32", 4, 16, 32, 32, 3, 1, 1, 1, EPI_PLAIN, False, 0),
    ("3x3 c16 k64", 4, 16, 16, 64, 3, 1, 1, 1, EPI_PLAIN, False, 0),
    ("4x4 pad2/1 c16 k64 (stem s2d)", 2, 16, 16, 64, 4, 1, 2, 1, EPI_PLAIN, False, 0),
    ("tail M=147 1x1 c64 k64", 3, 7, 64, 64, 1, 1, 0, 0, EPI_PLAIN, False, 0),
    ("tail M=147 3x3 c64 k64", 3, 7, 64, 64, 3, 1, 1, 1, EPI_PLAIN, False, 0),
    ("resadd+stats 1x1 c64 k256", 4, 8, 64, 256, 1, 1, 0, 0, EPI_RESADD, True, 0),
    ("stats 3x3 c64 k64", 4, 8, 64, 64, 3, 1, 1, 1, EPI_PLAIN, True, 0),
    ("stats 1x1 c64 k512 (2 n-tiles)", 4, 8, 64, 512, 1, 1, 0, 0, EPI_PLAIN, True, 0),
    ("bnbwd+stats 1x1 c256 k64", 4, 8, 256, 64, 1, 1, 0, 0, EPI_BNBWD, True, 0),
    ("bnbwd+stats 3x3 c64 k128", 4, 8, 64, 128, 3, 1, 1, 1, EPI_BNBWD, True, 0),
    ("dense 2048->1024", 128, 1, 2048, 1024, 1, 1, 0, 0, EPI_PLAIN, False, 0),
    ("rn50 1x1 64->256 @56 N32", 32, 56, 64, 256, 1, 1, 0, 0, EPI_PLAIN, True, 0),
    ("rn50 3x3 64->64 @56 N32", 32, 56, 64, 64, 3, 1, 1, 1, EPI_PLAIN, True, 0),
    ("rn50 1x1 256->64 @56 N32", 32, 56, 256, 64, 1, 1, 0, 0, EPI_PLAIN, True, 0),
    ("rn50 3x3 512->512 @7 N32", 32, 7, 512, 512, 3, 1, 1, 1, EPI_PLAIN, True, 0),
    ("rn50 1x1 1024->2048 s2 @14 N32", 32, 14, 1024, 2048, 1, 2, 0, 0, EPI_PLAIN, True, 0),
]

WGRAD_CASES = [
    ("1x1 c64 k64 M128", 2, 8, 64, 64, 1, 1, 0, 0),
    ("1x1 c64 k128", 8, 8, 64, 128, 1, 1, 0, 0),
    ("1x1 c256 k256", 8, 8, 256, 256, 1, 1, 0, 0),
    ("1x1 c128 k64", 8, 8, 128, 64, 1, 1, 0, 0),
    ("3x3 c64 k64", 2, 16, 64, 64, 3, 1, 1, 1),
    ("3x3 s2 c64 k128", 2, 16, 64, 128, 3, 2, 1, 1),
    ("1x1 s2 c64 k256", 2, 16, 64, 256, 1, 2, 0, 0),
    ("3x3 c16 k16", 4, 16, 16, 16, 3, 1, 1, 1),
    ("3x3 c32 k32", 4, 16, 32, 32, 3, 1, 1, 1),
    ("3x3 c16 k64", 4, 16, 16, 64, 3, 1, 1, 1),
    ("4x4 pad2/1 c16 k64", 2, 16, 16, 64, 4, 1, 2, 1),
    ("tail M=147 3x3 c64 k64", 3, 7, 64, 64, 3, 1, 1, 1),
    ("dense 2048->1024", 128, 1, 2048, 1024, 1, 1, 0, 0),
    ("rn50 3x3 64->64 @56 N32", 32, 56, 64, 64, 3, 1, 1, 1),
    ("rn50 1x1 256->64 @56 N32", 32, 56, 256, 64, 1, 1, 0, 0),
    ("rn50 3x3 512->512 @7 N32", 32, 7, 512, 512, 3, 1, 1, 1),
    ("3x3 c128 k128 h14 (1152 cols)", 4, 14, 128, 128, 3, 1, 1, 1),
    ("3x3 s2 c128 k128 h28", 4, 28, 128, 128, 3, 2, 1, 1),
    ("3x3 c32 k64 (288 cols)", 4, 16, 32, 64, 3, 1, 1, 1),
    ("rn50 stem 4x4 c16 k64 @112 N32", 32, 112, 16, 64, 4, 1, 2, 1),
    ("rn50 1x1 64->256 @56 N32", 32, 56, 64, 256, 1, 1, 0, 0),
    ("rn50 1x1 2048->512 @7 N32", 32, 7, 2048, 512, 1, 1, 0, 0),
]


def ref_conv(x, w, g):
    xf = x.float().permute(0, 3, 1, 2)
    wf = w.float().permute(0, 3, 1, 2)
    xf = F.pad(xf, (g.pad_lo, g.pad_hi, g.pad_lo, g.pad_hi))
    return F.conv2d(xf, wf, stride=g.stride).permute(0, 2, 3, 1).contiguous()


def time_it(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3  # us


def run_fprop(i, case, do_time):
    name, N, H, C, K, ks, st, plo, phi, mode, want_stats, force_bn = case
    g = ConvGeom(N, H, H, C, K, ks, ks, st, plo, phi)
    gen = torch.Generator(device="cuda").manual_seed(1234 + i)
    x = torch.randn((N, H, H, C), device="cuda", generator=gen).bfloat16()
    w = (torch.randn((K, ks, ks, C), device="cuda", generator=gen) / (ks * ks * C) ** 0.5).bfloat16()
    side = None
    ss = None
    if mode != EPI_PLAIN:
        side = torch.randn((N, g.P, g.Q, K), device="cuda", generator=gen).bfloat16()
    if mode == EPI_BNBWD:
        ss = torch.stack([torch.rand(K, device="cuda", generator=gen) + 0.5,
                          torch.randn(K, device="cuda", generator=gen) * 0.3], dim=1).contiguous()
    stats = torch.zeros((2, K), device="cuda") if want_stats else None
    y = conv_fprop(x, w, g, side=side, stats=stats, bn_scale_shift=ss, mode=mode, force_bn=force_bn)
    torch.cuda.synchronize()
    ref = ref_conv(x, w, g)
    if mode == EPI_RESADD:
        ref = ref + side.float()
    elif mode == EPI_BNBWD:
        mask = (side.float() * ss[:, 0] + ss[:, 1]) > 0
        ref = ref * mask
    err = (y.float() - ref).abs().max().item()
    scale = ref.abs().max().item()
    msg = f"FPROP {i:2d} {name:36s} maxerr {err:.4e} (ref max {scale:.3f})"
    ok = err <= 0.02 * scale + 1e-3
    if want_stats:
        yq = y.float()
        s1 = yq.sum(dim=(0, 1, 2))
        s2 = (yq * side.float()).sum(dim=(0, 1, 2)) if mode == EPI_BNBWD else (yq * yq).sum(dim=(0, 1, 2))
        e1 = ((stats[0] - s1).abs().max() / (s1.abs().max() + 1e-6)).item()
        e2 = ((stats[1] - s2).abs().max() / (s2.abs().max() + 1e-6)).item()
        msg += f" stats relerr {e1:.2e} {e2:.2e}"
        ok = ok and e1 < 2e-3 and e2 < 2e-3
    if not ok:
        bad = ((y.float() - ref).abs() > 0.02 * scale + 1e-3)
        idx = bad.nonzero()
        msg += f" BAD={bad.sum().item()}/{bad.numel()} first={idx[0].tolist() if len(idx) else None}"
        # which rows / columns are wrong (helps to tell descriptor bugs from pipeline bugs)
        rows = bad.reshape(-1, K).any(dim=1).nonzero().flatten()
        cols = bad.reshape(-1, K).any(dim=0).nonzero().flatten()
        msg += f" badrows[{len(rows)}]={rows[:8].tolist()} badcols[{len(cols)}]={cols[:8].tolist()}"
    if do_time:
        us = time_it(lambda: conv_fprop(x, w, g, out=y, side=side, stats=stats, bn_scale_shift=ss,
                                        mode=mode, force_bn=force_bn))
        flops = 2.0 * N * g.P * g.Q * K * C * ks * ks
        bytes_ = 2.0 * (x.numel() + w.numel() + y.numel() * (2 if side is not None else 1))
        xf = x.float().permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last).bfloat16()
        wf = w.permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last)
        pad = plo if plo == phi else 0
        if plo == phi:
            us_ref = time_it(lambda: F.conv2d(xf, wf, stride=st, padding=pad))
        else:
            us_ref = float("nan")
        msg += (f" | {us:8.1f} us {flops / us * 1e-6:7.1f} TFLOP/s {bytes_ / us * 1e-3:7.1f} GB/s"
                f" (cudnn {us_ref:8.1f} us)")
    print(("ok   " if ok else "FAIL ") + msg, flush=True)
    return ok


def run_wgrad(i, case, do_time):
    name, N, H, C, K, ks, st, plo, phi = case
    g = ConvGeom(N, H, H, C, K, ks, ks, st, plo, phi)
    gen = torch.Generator(device="cuda").manual_seed(4321 + i)
    x = torch.randn((N, H, H, C), device="cuda", generator=gen).bfloat16()
    dy = torch.randn((N, g.P, g.Q, K), device="cuda", generator=gen).bfloat16()
    dw = conv_wgrad(x, dy, g)
    torch.cuda.synchronize()
    xf = x.float().permute(0, 3, 1, 2)
    xf = F.pad(xf, (plo, phi, plo, phi)).requires_grad_(False)
    wshape = (K, C, ks, ks)
    ref = torch.nn.grad.conv2d_weight(xf, wshape, dy.float().permute(0, 3, 1, 2), stride=st)
    ref = ref.permute(0, 2, 3, 1).contiguous()
    err = (dw - ref).abs().max().item()
    scale = ref.abs().max().item()
    ok = err <= 2e-3 * scale + 1e-3
    msg = f"WGRAD {i:2d} {name:36s} maxerr {err:.4e} (ref max {scale:.3f})"
    if not ok:
        bad = (dw - ref).abs() > 2e-3 * scale + 1e-3
        idx = bad.nonzero()
        msg += f" BAD={bad.sum().item()}/{bad.numel()} first={idx[0].tolist() if len(idx) else None}"
        kk = bad.reshape(K, -1).any(dim=1).nonzero().flatten()
        taps = bad.reshape(K, ks * ks, C).any(dim=2).any(dim=0).nonzero().flatten()
        cc = bad.reshape(K * ks * ks, C).any(dim=0).nonzero().flatten()
        msg += f" badK[{len(kk)}]={kk[:6].tolist()} badtaps={taps.tolist()} badC[{len(cc)}]={cc[:6].tolist()}"
    if do_time:
        def f():
            dw.zero_()
            conv_wgrad(x, dy, g, dw=dw)
        us = time_it(f)
        flops = 2.0 * N * g.P * g.Q * K * C * ks * ks
        msg += f" | {us:8.1f} us {flops / us * 1e-6:7.1f} TFLOP/s (incl. memset)"
    print(("ok   " if ok else "FAIL ") + msg, flush=True)
    return ok


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--start", type=int, default=0)
    ap.add_argument("--only", default="")
    ap.add_argument("--time", action="store_true")
    ap.add_argument("--progress", default="gpurun_out/probe_progress.txt")
    args = ap.parse_args()
    nl.load()
    cases = []
    if args.only in ("", "fprop"):
        cases += [("fprop", c) for c in FPROP_CASES]
    if args.only in ("", "wgrad"):
        cases += [("wgrad", c) for c in WGRAD_CASES]
    os.makedirs(os.path.dirname(args.progress) or ".", exist_ok=True)
    nfail = 0
    for i, (kind, case) in enumerate(cases):
        if i < args.start:
            continue
        with open(args.progress, "w") as f:
            f.write(str(i))
        try:
            ok = run_fprop(i, case, args.time) if kind == "fprop" else run_wgrad(i, case, args.time)
            nfail += 0 if ok else 1
        except Exception as e:  # CUDA fault: the context is gone, restart in a new process
            print(f"CRASH {kind} {i} {case[0]}: {type(e).__name__}: {str(e)[:300]}", flush=True)
            try:
                print("watchdog:", hex(nl.load().drn_watchdog_status()), flush=True)
            except Exception:
                pass
            sys.stdout.flush()
            os._exit(3)
    with open(args.progress, "w") as f:
        f.write(str(len(cases)))
    print(f"DONE failures={nfail}", flush=True)
    os._exit(0 if nfail == 0 else 1)


if __name__ == "__main__":
    main()
