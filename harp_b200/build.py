"""In-tree build of the native libraries.

``libharp_b200.so``  — sm_100a CUDA kernels + C ABI (``include/harp_b200.h``), built with nvcc.
``libharp_host.so``  — C++17 host runtime (pooled allocator, parsers, schedulers), built with g++.

The shared objects live next to this file so that they travel with a repo
snapshot to a GPU box (they are git-ignored, not gpurun-ignored).  Nothing is
JIT-compiled at import time: ``python -m harp_b200.build`` (or
``__graft_entry__.build()``) is the only thing that invokes a compiler.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
PKG = Path(__file__).resolve().parent
OBJ = ROOT / "build" / "obj"
CUDA_LIB = PKG / "libharp_b200.so"
HOST_LIB = PKG / "libharp_host.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
    "-I", str(ROOT / "include"),
] + (["-DHB_GRAM_DEBUG"] if os.environ.get("HB_GRAM_DEBUG") else []) + [f"-D{k}={v}" for k, v in os.environ.items() if k in ("HB_SGD_THREADS", "HB_SGD_PF")]
GXX_FLAGS = ["-O3", "-std=c++17", "-fPIC", "-pthread", "-Wall", "-I", str(ROOT / "include")]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _stamp(src: Path, flags: list[str]) -> str:
    h = hashlib.sha256()
    h.update(src.read_bytes())
    for inc in sorted(list(CSRC.glob("*.cuh")) + list((ROOT / "include").glob("*.h")) + list(CSRC.glob("host/*.h"))):
        h.update(inc.read_bytes())
    h.update(" ".join(flags).encode())
    return h.hexdigest()


def _compile(cmd_prefix: list[str], src: Path, flags: list[str], verbose: bool) -> Path:
    OBJ.mkdir(parents=True, exist_ok=True)
    obj = OBJ / (src.stem + ".o")
    stamp_file = OBJ / (src.stem + ".stamp")
    stamp = _stamp(src, flags)
    if obj.exists() and stamp_file.exists() and stamp_file.read_text() == stamp:
        return obj
    cmd = cmd_prefix + flags + ["-c", str(src), "-o", str(obj)]
    res = subprocess.run(cmd, capture_output=True, text=True)
    log = OBJ / (src.stem + ".log")
    log.write_text(res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError(f"compile failed: {' '.join(cmd)}\n{res.stdout}\n{res.stderr}")
    if verbose:
        sys.stderr.write(res.stderr)
    stamp_file.write_text(stamp)
    return obj


def build_cuda(verbose: bool = False) -> Path:
    nvcc = _nvcc()
    srcs = sorted(CSRC.glob("*.cu"))
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile([nvcc], s, NVCC_FLAGS, verbose), srcs))
    newest = max(o.stat().st_mtime for o in objs)
    if CUDA_LIB.exists() and CUDA_LIB.stat().st_mtime >= newest:
        return CUDA_LIB
    cmd = [nvcc, "-shared", "-o", str(CUDA_LIB)] + [str(o) for o in objs] + ["-cudart", "static"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"link failed: {' '.join(cmd)}\n{res.stderr}")
    return CUDA_LIB


def build_host(verbose: bool = False) -> Path | None:
    srcs = sorted((CSRC / "host").glob("*.cpp"))
    if not srcs:
        return None
    gxx = shutil.which("g++") or "g++"
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile([gxx], s, GXX_FLAGS, verbose), srcs))
    newest = max(o.stat().st_mtime for o in objs)
    if HOST_LIB.exists() and HOST_LIB.stat().st_mtime >= newest:
        return HOST_LIB
    cmd = [gxx, "-shared", "-pthread", "-o", str(HOST_LIB)] + [str(o) for o in objs]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"link failed: {' '.join(cmd)}\n{res.stderr}")
    return HOST_LIB


def build_all(verbose: bool = False) -> None:
    build_cuda(verbose)
    build_host(verbose)


if __name__ == "__main__":
    build_all(verbose="-v" in sys.argv)
    print("built:", CUDA_LIB, HOST_LIB if HOST_LIB.exists() else "(no host lib)")
