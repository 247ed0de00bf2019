"""Build libdrn_b200.so (the sm_100a kernels) in-tree with nvcc.

Usage: ``python -m distributed_resnet_tensorflow_b200._native.build [--force] [--verbose]``

The library has no PyTorch/pybind dependency: it is plain CUDA C++ with a C ABI
(``include/drn_b200.h``) and is loaded with ctypes, so a full rebuild takes seconds per file
and the resulting ``.so`` travels with the source tree to the GPU box.
"""
from __future__ import annotations

import concurrent.futures
import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
REPO = HERE.parent.parent
CSRC = REPO / "csrc"
OBJ = HERE / "obj"
LIB = HERE / "libdrn_b200.so"

ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-rdc=true", "-Xcompiler", "-fPIC",
          "--expt-relaxed-constexpr", "-I", str(REPO / "include"), "-I", str(CSRC)]


def _nvcc() -> str:
    cand = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(cand):
        raise RuntimeError("nvcc not found")
    return cand


def _sources() -> list[Path]:
    return sorted(CSRC.glob("*.cu"))


def _stamp(src: Path) -> str:
    h = hashlib.sha1()
    h.update(src.read_bytes())
    for hdr in sorted(list(CSRC.glob("*.cuh")) + list((REPO / "include").glob("*.h"))):
        h.update(hdr.read_bytes())
    h.update(" ".join(COMMON + ARCH_FLAGS).encode())
    return h.hexdigest()


def _compile(src: Path, verbose: bool) -> Path:
    obj = OBJ / (src.stem + ".o")
    stamp_file = OBJ / (src.stem + ".stamp")
    stamp = _stamp(src)
    if obj.exists() and stamp_file.exists() and stamp_file.read_text() == stamp:
        return obj
    cmd = [_nvcc(), *ARCH_FLAGS, *COMMON, "-c", str(src), "-o", str(obj)]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
        print(" ".join(cmd), flush=True)
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{res.stdout}\n{res.stderr}")
    if verbose and (res.stdout or res.stderr):
        print(res.stdout, res.stderr, flush=True)
    stamp_file.write_text(stamp)
    return obj


def build(force: bool = False, verbose: bool = False) -> Path:
    OBJ.mkdir(parents=True, exist_ok=True)
    if force:
        for f in OBJ.glob("*"):
            f.unlink()
    srcs = _sources()
    if not srcs:
        raise RuntimeError(f"no .cu sources under {CSRC}")
    with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile(s, verbose), srcs))
    newest = max(o.stat().st_mtime for o in objs)
    if LIB.exists() and LIB.stat().st_mtime >= newest and not force:
        return LIB
    cmd = [_nvcc(), *ARCH_FLAGS, "-shared", "-Xcompiler", "-fPIC", "-o", str(LIB),
           *[str(o) for o in objs], "-lcudart", "-lnvjpeg", "-Xlinker", "-rpath", "-Xlinker", "/usr/local/cuda/lib64"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    return LIB


if __name__ == "__main__":
    lib = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(lib)
