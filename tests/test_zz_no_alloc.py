"""drn_conv_dgrad_static: the strided data gradient for callers that cannot allocate.

Some users of the C library cannot supply the scratch buffer drn_conv_dgrad asks for - their
process has already given all GPU memory to its own pool, or they sit in a latency-critical loop
where cudaMalloc/cudaFree (which synchronise the device) are not allowed.  The static variant keeps
a fixed 16 MiB scratch inside the library instead.  This test drives it straight through the C ABI
(ctypes, raw pointers) and checks, on two consecutive calls with different problems, that it
produces what the ordinary entry point produces with a caller-allocated buffer - the second call
also proves that contents left behind by the first do not leak into the result - and that an
oversized problem is refused with a readable message.

The file name sorts last so that a `pytest -x` run has exercised everything else before this.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from distributed_resnet_tensorflow_b200._native import lib as nl
    return nl, nl.load()


def _problem(N, H, C, K, ks, stride, seed):
    from distributed_resnet_tensorflow_b200.ops.conv import ConvGeom
    g = ConvGeom.tf_fixed_padding(N, H, H, C, K, ks, stride)
    gen = torch.Generator(device="cuda").manual_seed(seed)
    w = (torch.randn((K, ks, ks, C), device="cuda", generator=gen) / (ks * ks * C) ** 0.5).bfloat16()
    dy = torch.randn((N, g.P, g.Q, K), device="cuda", generator=gen).bfloat16()
    return g, w, dy


def _both_ways(nl, lib, g, w, dy):
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    d = g.desc()
    wd = torch.empty(w.numel(), dtype=torch.bfloat16, device="cuda")
    nl.check(lib.drn_weight_prep_dgrad(w.data_ptr(), wd.data_ptr(), g.K, g.R, g.S, g.C, st), "weight_prep")
    need = lib.drn_conv_dgrad_scratch_bytes(ctypes.byref(d))
    assert 0 < need <= (16 << 20)
    scratch = torch.empty(need, dtype=torch.uint8, device="cuda")
    ref = torch.empty((g.N, g.H, g.W, g.C), dtype=torch.bfloat16, device="cuda")
    got = torch.full_like(ref, float("nan"))
    nl.check(lib.drn_conv_dgrad(ctypes.byref(d), dy.data_ptr(), wd.data_ptr(), ref.data_ptr(), None, None, 0,
                                None, 0, scratch.data_ptr(), need, st), "drn_conv_dgrad")
    nl.check(lib.drn_conv_dgrad_static(ctypes.byref(d), dy.data_ptr(), wd.data_ptr(), got.data_ptr(), None, None,
                                       0, None, 0, st), "drn_conv_dgrad_static")
    torch.cuda.synchronize()
    return ref.float(), got.float()


def test_static_scratch_variant_matches_on_two_consecutive_calls():
    nl, lib = _lib()
    # call 1: a larger problem fills a good part of the built-in scratch ...
    g1, w1, dy1 = _problem(8, 28, 128, 128, 3, 2, seed=11)
    ref1, got1 = _both_ways(nl, lib, g1, w1, dy1)
    assert torch.isfinite(got1).all()
    assert (got1 - ref1).abs().max() <= 0.02 * ref1.abs().max() + 1e-3
    # ... call 2: a different, smaller problem reuses the same bytes and must not see the leftovers
    g2, w2, dy2 = _problem(3, 14, 64, 256, 3, 2, seed=12)
    ref2, got2 = _both_ways(nl, lib, g2, w2, dy2)
    assert torch.isfinite(got2).all()
    assert (got2 - ref2).abs().max() <= 0.02 * ref2.abs().max() + 1e-3


def test_static_scratch_variant_back_to_back_without_sync_and_without_allocating():
    """Two different problems enqueued on ONE stream with no synchronisation in between (the second
    overwrites the scratch the first has just read: stream order must be enough), and the amount of
    free device memory is the same before and after warmed calls - the entry point allocates nothing."""
    nl, lib = _lib()
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    probs = [_problem(8, 28, 128, 128, 3, 2, seed=21), _problem(4, 14, 256, 64, 3, 2, seed=22)]
    prepared = []
    for g, w, dy in probs:
        wd = torch.empty(w.numel(), dtype=torch.bfloat16, device="cuda")
        nl.check(lib.drn_weight_prep_dgrad(w.data_ptr(), wd.data_ptr(), g.K, g.R, g.S, g.C, st), "weight_prep")
        d = g.desc()
        need = lib.drn_conv_dgrad_scratch_bytes(ctypes.byref(d))
        scratch = torch.empty(need, dtype=torch.uint8, device="cuda")
        ref = torch.empty((g.N, g.H, g.W, g.C), dtype=torch.bfloat16, device="cuda")
        nl.check(lib.drn_conv_dgrad(ctypes.byref(d), dy.data_ptr(), wd.data_ptr(), ref.data_ptr(), None, None, 0,
                                    None, 0, scratch.data_ptr(), need, st), "drn_conv_dgrad")
        got = torch.full_like(ref, float("nan"))
        prepared.append((d, dy, wd, ref, got))
    # warm call (resolves the array's address on this device), then measure free memory
    d, dy, wd, ref, got = prepared[0]
    nl.check(lib.drn_conv_dgrad_static(ctypes.byref(d), dy.data_ptr(), wd.data_ptr(), got.data_ptr(), None, None,
                                       0, None, 0, st), "drn_conv_dgrad_static")
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    for d, dy, wd, ref, got in prepared:          # no synchronisation between the two calls
        nl.check(lib.drn_conv_dgrad_static(ctypes.byref(d), dy.data_ptr(), wd.data_ptr(), got.data_ptr(), None,
                                           None, 0, None, 0, st), "drn_conv_dgrad_static")
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after == free_before, (free_before, free_after)
    for d, dy, wd, ref, got in prepared:
        r, g_ = ref.float(), got.float()
        assert torch.isfinite(g_).all()
        assert (g_ - r).abs().max() <= 0.02 * r.abs().max() + 1e-3


def test_static_scratch_variant_refuses_oversized_problems():
    nl, lib = _lib()
    g, w, dy = _problem(1, 8, 64, 64, 3, 2, seed=13)
    d = g.desc()
    d.N = 4096                      # pretend batch: 2*4096*7*7*64 bytes would still fit ...
    d.H = d.W = 224                 # ... but at 224x224 the zero-inserted dy is far beyond 16 MiB
    d.P = d.Q = 112
    assert lib.drn_conv_dgrad_scratch_bytes(ctypes.byref(d)) > (16 << 20)
    rc = lib.drn_conv_dgrad_static(ctypes.byref(d), dy.data_ptr(), w.data_ptr(), dy.data_ptr(), None, None, 0, None,
                                   0, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    msg = lib.drn_last_error().decode()
    assert "scratch" in msg and "drn_conv_dgrad" in msg
