"""ctypes binding of libdrn_b200.so (C ABI in include/drn_b200.h).

The library is built in-tree by ``_native/build.py``. On a machine with a GPU the ops in
``distributed_resnet_tensorflow_b200.ops`` REQUIRE it: there is no silent PyTorch fallback on
CUDA tensors - a missing library raises ``NativeLibraryError``.
"""
from __future__ import annotations

import ctypes
import os
from pathlib import Path

HERE = Path(__file__).resolve().parent
LIB_PATH = HERE / "libdrn_b200.so"


class NativeLibraryError(RuntimeError):
    pass


class ConvDesc(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int) for n in
                ("N", "H", "W", "C", "K", "R", "S", "stride", "pad_lo", "pad_hi", "P", "Q")]


class PullDesc(ctypes.Structure):
    """drn_pull_desc (include/drn_b200.h)."""
    _fields_ = [("shadows", ctypes.c_void_p), ("nranks", ctypes.c_int), ("rank", ctypes.c_int),
                ("my_pad", ctypes.c_void_p), ("local_arena", ctypes.c_void_p), ("counters", ctypes.c_void_p),
                ("bucket_lo", ctypes.c_longlong), ("per", ctypes.c_longlong), ("num_ps", ctypes.c_int),
                ("bucket", ctypes.c_int), ("off", ctypes.c_longlong), ("count", ctypes.c_longlong)]


class BnFin(ctypes.Structure):
    """drn_bn_fin (include/drn_b200.h): BatchNorm finalize done by the last CTA of the producing convolution."""
    _fields_ = [("counter", ctypes.c_void_p), ("count", ctypes.c_longlong), ("gamma", ctypes.c_void_p),
                ("beta", ctypes.c_void_p), ("run_mean", ctypes.c_void_p), ("run_var", ctypes.c_void_p),
                ("decay", ctypes.c_float), ("eps", ctypes.c_float), ("scale_shift", ctypes.c_void_p),
                ("mean_invstd", ctypes.c_void_p)]


class ConvView(ctypes.Structure):
    _fields_ = [("stride_w", ctypes.c_longlong), ("stride_h", ctypes.c_longlong),
                ("stride_n", ctypes.c_longlong), ("pad_w_lo", ctypes.c_int), ("pad_w_hi", ctypes.c_int)]


EPI_PLAIN, EPI_RESADD, EPI_BNBWD, EPI_BNBWD_ADD, EPI_BNBWD_APPLY = 0, 1, 2, 3, 4

_lib = None

_vp = ctypes.c_void_p
_i = ctypes.c_int
_f = ctypes.c_float
_ll = ctypes.c_longlong

# name -> (restype, argtypes)
_SIGNATURES = {
    "drn_last_error": (ctypes.c_char_p, []),
    "drn_watchdog_status": (_i, []),
    "drn_watchdog_arm": (_i, []),
    "drn_num_sms": (_i, []),
    "drn_set_pdl_mask": (_i, [_i]),
    "drn_set_deterministic": (_i, [_i]),
    "drn_deterministic": (_i, []),
    "drn_conv_igemm": (_i, [ctypes.POINTER(ConvDesc), _vp, _vp, _vp, _vp, _vp, _i, _vp, _i, _i, _vp]),
    "drn_conv_wgrad": (_i, [ctypes.POINTER(ConvDesc), _vp, _vp, _vp, _i, _vp]),
    "drn_conv_igemm_ex": (_i, [ctypes.POINTER(ConvDesc), ctypes.POINTER(ConvView), _vp, ctypes.POINTER(PullDesc), _vp, _vp, _vp, _vp, _vp, _i, _vp, _i, _i, _vp]),
    "drn_conv_igemm_fin": (_i, [ctypes.POINTER(ConvDesc), ctypes.POINTER(ConvView), _vp, ctypes.POINTER(PullDesc), ctypes.POINTER(BnFin), _vp, _vp, _vp, _vp, _vp, _i, _vp, _i, _i, _vp]),
    "drn_conv_igemm_epi2": (_i, [ctypes.POINTER(ConvDesc), _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _i, _i, _vp]),
    "drn_conv_igemm_view": (_i, [ctypes.POINTER(ConvDesc), ctypes.POINTER(ConvView), _vp, _vp, _vp, _vp, _vp, _i, _vp, _i, _i, _vp]),
    "drn_resize_crop_batch": (_i, [_vp, _vp, _vp, _i, _i, _vp]),
    "drn_weight_prep_dgrad_s2": (_i, [_vp, _vp, _i, _i, _vp]),
    "drn_conv_dgrad_s2": (_i, [ctypes.POINTER(ConvDesc), _vp, _vp, _vp, _vp]),
    "drn_conv_wgrad_pre": (_i, [ctypes.POINTER(ConvDesc), _vp, _vp, _vp, _vp, _i, _vp]),
    "drn_conv_igemm_pre": (_i, [ctypes.POINTER(ConvDesc), _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _i, _i, _vp]),
    "drn_conv_wgrad_view": (_i, [ctypes.POINTER(ConvDesc), ctypes.POINTER(ConvView), _vp, _vp, _vp, _i, _vp]),
    "drn_conv_dgrad_scratch_bytes": (_ll, [ctypes.POINTER(ConvDesc)]),
    "drn_conv_dgrad": (_i, [ctypes.POINTER(ConvDesc), _vp, _vp, _vp, _vp, _vp, _i, _vp, _i, _vp, _ll, _vp]),
    "drn_conv_dgrad_static": (_i, [ctypes.POINTER(ConvDesc), _vp, _vp, _vp, _vp, _vp, _i, _vp, _i, _vp]),
    "drn_bn_finalize": (_i, [_vp, _i, _ll, _vp, _vp, _vp, _vp, _f, _f, _vp, _vp, _i, _vp]),
    "drn_bn_finalize_apply": (_i, [_vp, _vp, _i, _ll, _vp, _vp, _vp, _vp, _f, _f, _vp, _vp, _vp, _i, _i, _vp]),
    "drn_bn_bwd_finalize_apply": (_i, [_vp, _vp, _vp, _i, _ll, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp]),
    "drn_bn_inference_params": (_i, [_vp, _vp, _vp, _vp, _f, _vp, _i, _vp]),
    "drn_bn_apply": (_i, [_vp, _vp, _vp, _vp, _ll, _i, _i, _vp]),
    "drn_bn_stats": (_i, [_vp, _vp, _i, _ll, _i, _vp]),
    "drn_bn_pair_reduce": (_i, [_vp, _vp, _vp, _i, _ll, _i, _vp]),
    "drn_bn_bwd_mask_reduce": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _f, _vp, _vp, _vp, _vp, _i, _i, _vp]),
    "drn_relu_bwd_reduce": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _ll, _i, _vp]),
    "drn_bn_bwd_finalize": (_i, [_vp, _i, _ll, _vp, _vp, _vp, _vp, _vp, _i, _vp]),
    "drn_bn_bwd_apply": (_i, [_vp, _vp, _vp, _vp, _vp, _ll, _i, _vp]),
    "drn_maxpool_fwd": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "drn_maxpool_bwd": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "drn_bnrelu_avgpool": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp]),
    "drn_softmax_xent": (_i, [_vp, _i, _vp, _vp, _i, _i, _f, _vp, _vp, _vp, _vp, _vp, _vp]),
    "drn_sgd_momentum": (_i, [_vp, _vp, _vp, _vp, _ll, _vp, _vp, _i, _vp]),
    "drn_weight_prep_dgrad": (_i, [_vp, _vp, _i, _i, _i, _i, _vp]),
    "drn_weight_prep_dgrad_batched": (_i, [_vp, _vp, _vp, _i, _i, _vp]),
    "drn_upsample_zero": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "drn_f32_to_bf16": (_i, [_vp, _vp, _ll, _vp]),
    "drn_mul_inplace_f32": (_i, [_vp, _vp, _ll, _vp]),
    "drn_cifar_augment": (_i, [_vp, _i, _i, _vp, _i, _i, _i, ctypes.c_ulonglong, ctypes.c_uint, _ll, _i, _vp]),
    "drn_imagenet_prep": (_i, [_vp, _ll, _i, _i, _vp, _vp, _i, _i, _f, _f, _f, _i, _i, _vp]),
    "drn_resize_bilinear": (_i, [_vp, _i, _i, _vp, _i, _i, _vp]),
    "drn_jpeg_create": (_i, [_i, ctypes.POINTER(_vp)]),
    "drn_jpeg_backend": (_i, [_vp]),
    "drn_jpeg_destroy": (_i, [_vp]),
    "drn_jpeg_info": (_i, [_vp, _vp, _vp, _i, _vp]),
    "drn_jpeg_decode_batch": (_i, [_vp, _vp, _vp, _i, _vp, _vp, _vp, _vp]),
    "drn_crc32c": (ctypes.c_uint32, [_vp, ctypes.c_size_t, ctypes.c_uint32]),
    "drn_tfrecord_index": (_ll, [_vp, ctypes.c_size_t, _vp, _vp, _ll, _i]),
    "drn_comm_pad_bytes": (_i, []),
    "drn_allreduce_f32": (_i, [_vp, _vp, _vp, _i, _i, _ll, _ll, _f, _i, _i, _vp]),
    "drn_ps_update": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _ll, _ll, _vp, _i, _i, _i, _i, _i, _vp]),
    "drn_shard_per": (_ll, [_ll, _i]),
    "drn_allreduce_sgd_oneshot": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _ll, _ll, _vp, _vp, _i, _i, _vp]),
    "drn_bcast_f32": (_i, [_vp, _vp, _vp, _i, _i, _i, _ll, _i, _i, _vp]),
    "drn_ps_sync_vectors": (_i, [_vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _i, _i, _vp]),
    "drn_comm_pad_offset": (_i, [_i]),
    "drn_comm_barrier": (_i, [_vp, _i, _i, _i, _vp]),
    "drn_ipc_alloc": (_i, [ctypes.c_size_t, ctypes.POINTER(_vp), _vp]),
    "drn_ipc_open": (_i, [_vp, ctypes.POINTER(_vp)]),
}


def register(name: str, restype, argtypes) -> None:
    """Declare one more C function (used by the op modules next to their wrappers)."""
    _SIGNATURES[name] = (restype, argtypes)
    if _lib is not None and hasattr(_lib, name):
        fn = getattr(_lib, name)
        fn.restype, fn.argtypes = restype, argtypes


def load(build_if_missing: bool = True):
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        if not build_if_missing or os.environ.get("DRN_NO_BUILD"):
            raise NativeLibraryError(f"{LIB_PATH} not found; run python -m "
                                     "distributed_resnet_tensorflow_b200._native.build")
        from . import build as _build
        _build.build()
    try:
        lib = ctypes.CDLL(str(LIB_PATH))
    except OSError as e:  # pragma: no cover
        raise NativeLibraryError(f"cannot load {LIB_PATH}: {e}") from e
    for name, (restype, argtypes) in _SIGNATURES.items():
        if hasattr(lib, name):
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = restype, argtypes
    _lib = lib
    return lib


def available() -> bool:
    try:
        load(build_if_missing=False)
        return True
    except NativeLibraryError:
        return False


def check(status: int, what: str = "") -> None:
    if status != 0:
        msg = load().drn_last_error().decode(errors="replace")
        raise RuntimeError(f"libdrn_b200 {what} failed: {msg}")


_launch_count = 0


def reset_launch_count() -> None:
    global _launch_count
    _launch_count = 0


def launch_count() -> int:
    """Number of kernel-launching C calls made through ``call`` since the last reset (every
    ``drn_*`` launcher enqueues exactly one kernel of this library)."""
    return _launch_count


def add_launches(n: int) -> None:
    """For entry points that enqueue more than one kernel (e.g. strided drn_conv_dgrad)."""
    global _launch_count
    _launch_count += n


_profile = None   # list of (name, start_event, end_event) while per-call timing is on


def start_call_profile() -> None:
    """Time every subsequent ``call`` with a pair of CUDA events on the current stream
    (``--profile``: the per-kernel device-time table; do not use while capturing a graph)."""
    global _profile
    _profile = []


def stop_call_profile():
    """Returns [(entry point, calls, total_us)] sorted by time and switches timing off."""
    global _profile
    import torch
    torch.cuda.synchronize()
    agg = {}
    for name, e0, e1 in _profile or []:
        c, t = agg.get(name, (0, 0.0))
        agg[name] = (c + 1, t + e0.elapsed_time(e1) * 1e3)
    _profile = None
    return sorted(((n, c, t) for n, (c, t) in agg.items()), key=lambda r: -r[2])


_NVTX = os.environ.get("DRN_NVTX", "0") == "1"


def call(name: str, *args) -> None:
    global _launch_count
    _launch_count += 1
    if _profile is not None:
        import torch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn = getattr(load(), name)
        check(fn(*args), name)
        e1.record()
        _profile.append((name, e0, e1))
        return
    lib = load()
    fn = getattr(lib, name, None)
    if fn is None:
        raise NativeLibraryError(f"symbol {name} missing from {LIB_PATH} (stale build?)")
    if _NVTX:
        # DRN_NVTX=1: one NVTX range per native call, so timeline tools show the engine's structure
        import torch
        torch.cuda.nvtx.range_push(name)
        try:
            check(fn(*args), name)
        finally:
            torch.cuda.nvtx.range_pop()
        return
    check(fn(*args), name)
