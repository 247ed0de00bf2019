"""TMA box-row-length probe (csrc/probe_tma.cu): copy bandwidth with 64-byte vs 128-byte box rows."""
import ctypes, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_resnet_tensorflow_b200._native import lib as nl
lib = nl.load()
lib.drn_probe_tma_copy.restype = ctypes.c_int
lib.drn_probe_tma_copy.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for rows, cols in ((401408, 256), (401408, 64), (100352, 512)):
    src = torch.randn((rows, cols), device="cuda").bfloat16(); dst = torch.zeros_like(src)
    for bc in (32, 64):
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        nl.check(lib.drn_probe_tma_copy(src.data_ptr(), dst.data_ptr(), rows, cols, bc, st), "probe")
        torch.cuda.synchronize(); assert torch.equal(src, dst)
        best = 1e9
        for _ in range(5):
            flush.zero_(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); lib.drn_probe_tma_copy(src.data_ptr(), dst.data_ptr(), rows, cols, bc, st); e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) * 1e3)
        print(f"[{rows} x {cols}] box 128 x {bc} cols ({bc * 2}-byte rows): {best:7.1f} us  {2 * src.numel() * 2 / best * 1e-3:6.0f} GB/s", flush=True)
