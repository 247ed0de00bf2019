"""Filter-gradient kernel on the small-spatial ResNet-50 layers: time vs split-K factor, and with the
epilogue / the MMAs switched off (DRN_WGRAD_DEBUG, set by the caller) to see what bounds it."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_resnet_tensorflow_b200.ops.conv import ConvGeom, conv_wgrad

SHAPES = [(14, 256, 256, 3), (14, 1024, 256, 1), (14, 256, 1024, 1), (7, 512, 512, 3), (7, 2048, 512, 1), (7, 512, 2048, 1),
          (28, 128, 128, 3), (28, 512, 128, 1)]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for H, C, K, ks in SHAPES:
    lo = (ks - 1) // 2
    g = ConvGeom(128, H, H, C, K, ks, ks, 1, lo, ks - 1 - lo)
    x = torch.randn((128, H, H, C), device="cuda").bfloat16()
    dy = torch.randn((128, H, H, K), device="cuda").bfloat16()
    dw = torch.zeros((K, ks, ks, C), device="cuda")
    row = []
    for split in ([0] if os.environ.get("DRN_WGRAD_DEBUG") else [0, 1, 2, 4, 8, 16]):
        for _ in range(2):
            conv_wgrad(x, dy, g, dw=dw, split_k=split)
        ts = []
        for _ in range(5):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); conv_wgrad(x, dy, g, dw=dw, split_k=split); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        row.append(f"split {split}: {min(ts):.1f}")
    fl = 2.0 * 128 * H * H * K * C * ks * ks
    print(f"H{H} C{C} K{K} k{ks} ({fl*1e-9:.1f} GFLOP): " + "  ".join(row), flush=True)
