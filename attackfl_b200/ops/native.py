"""Build + load the native libraries through their C ABI (``include/attackfl_b200.h``).

Two in-tree shared objects (git-ignored, shipped to the GPU box by ``gpurun``):

* ``attackfl_b200/lib/libattackfl_b200.so`` — ``nvcc -gencode arch=compute_100a,code=sm_100a`` build of
  ``csrc/*.cu``; operates on device pointers and CUDA streams.
* ``attackfl_b200/lib/libattackfl_cpu.so`` — g++ build of ``csrc/cpu_backend.cpp`` (the same kernel cores);
  operates on host pointers.

``lib_for(tensor)`` returns the right handle for a tensor's device and **raises** if the CUDA library
is missing on a machine with a GPU: there is no silent PyTorch fall-back on the GPU path.
"""
from __future__ import annotations

import ctypes
import glob
import os
import shutil
import subprocess
import sys
from typing import List, Optional

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
CSRC = os.path.join(ROOT, "csrc")
LIBDIR = os.path.join(ROOT, "attackfl_b200", "lib")
CUDA_LIB = os.path.join(LIBDIR, "libattackfl_b200.so")
CPU_LIB = os.path.join(LIBDIR, "libattackfl_cpu.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC", "-Xptxas", "-v"]
GXX_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-shared", "-fno-fast-math"]


def _digest(sources: List[str], flags: List[str]) -> str:
    """Content hash of the sources and flags: the staleness test does not depend on file times (a snapshot copied to
    another machine keeps contents, not necessarily mtimes)."""
    import hashlib
    h = hashlib.sha1(" ".join(flags).encode())
    for s in sorted(sources):
        h.update(os.path.basename(s).encode())
        with open(s, "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()


def _stamp_ok(target: str, digest: str) -> bool:
    try:
        with open(target + ".sha1") as fh:
            return os.path.exists(target) and fh.read().strip() == digest
    except OSError:
        return False


def _write_stamp(target: str, digest: str) -> None:
    with open(target + ".sha1", "w") as fh:
        fh.write(digest)


def _sources():
    cu = sorted(glob.glob(os.path.join(CSRC, "*.cu")))
    hdr = sorted(glob.glob(os.path.join(CSRC, "*.h")) + glob.glob(os.path.join(CSRC, "*.cuh"))
                 + glob.glob(os.path.join(ROOT, "include", "*.h")))
    return cu, hdr


def nvcc_path() -> Optional[str]:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    return None


def build_cuda(force: bool = False, verbose: bool = False) -> str:
    """Compiles every ``csrc/*.cu`` for sm_100a into one shared object (cross-compiles without a GPU)."""
    cu, hdr = _sources()
    digest = _digest(cu + hdr, NVCC_FLAGS)
    if not force and _stamp_ok(CUDA_LIB, digest):
        return CUDA_LIB
    nvcc = nvcc_path()
    if nvcc is None:
        raise RuntimeError("nvcc not found; cannot build the sm_100a library")
    os.makedirs(LIBDIR, exist_ok=True)
    objs = [os.path.join(LIBDIR, os.path.basename(src) + ".o") for src in cu]
    logs = []

    def compile_one(src: str, obj: str) -> str:
        obj_digest = _digest([src] + hdr, NVCC_FLAGS)
        if not force and _stamp_ok(obj, obj_digest):
            return ""
        cmd = [nvcc] + NVCC_FLAGS + ["-I", os.path.join(ROOT, "include"), "-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        _write_stamp(obj, obj_digest)
        return f"$ {' '.join(cmd)}\n{r.stdout}{r.stderr}"

    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 2)) as pool:      # one nvcc process per translation unit
        logs = [l for l in pool.map(compile_one, cu, objs) if l]
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC", "-o", CUDA_LIB] + objs
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    _write_stamp(CUDA_LIB, digest)
    with open(os.path.join(LIBDIR, "build_cuda.log"), "w") as fh:
        fh.write("\n".join(logs))
    if verbose:
        print("\n".join(logs))
    return CUDA_LIB


def build_cpu(force: bool = False) -> str:
    src = os.path.join(CSRC, "cpu_backend.cpp")
    _, hdr = _sources()
    digest = _digest([src] + hdr, GXX_FLAGS)
    if not force and _stamp_ok(CPU_LIB, digest):
        return CPU_LIB
    os.makedirs(LIBDIR, exist_ok=True)
    cmd = ["g++"] + GXX_FLAGS + ["-I", os.path.join(ROOT, "include"), src, "-o", CPU_LIB]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"g++ failed:\n{r.stdout}\n{r.stderr}")
    _write_stamp(CPU_LIB, digest)
    return CPU_LIB


_cuda: Optional[ctypes.CDLL] = None
_cpu: Optional[ctypes.CDLL] = None


def _declare(lib: ctypes.CDLL) -> ctypes.CDLL:
    lib.afl_last_error.restype = ctypes.c_char_p
    lib.afl_launch_count.restype = ctypes.c_longlong
    return lib


def cpu_lib() -> ctypes.CDLL:
    global _cpu
    if _cpu is None:
        _cpu = _declare(ctypes.CDLL(build_cpu()))
    return _cpu


def cuda_lib() -> ctypes.CDLL:
    global _cuda
    if _cuda is None:
        # always ask the builder: it is a no-op when the content stamp matches, and a stale binary can never be loaded
        build_cuda(force=bool(os.environ.get("ATTACKFL_REBUILD")))
        _cuda = _declare(ctypes.CDLL(CUDA_LIB))
    return _cuda


def lib_for(t: torch.Tensor) -> ctypes.CDLL:
    return cuda_lib() if t.is_cuda else cpu_lib()


def check(lib: ctypes.CDLL, rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib.afl_last_error().decode(errors="replace")
        raise RuntimeError(f"native call {what} failed (rc={rc}): {msg}")


def stream_ptr(t: torch.Tensor) -> ctypes.c_void_p:
    if t.is_cuda:
        return ctypes.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)
    return ctypes.c_void_p(0)


def ptr(t: Optional[torch.Tensor]) -> ctypes.c_void_p:
    if t is None:
        return ctypes.c_void_p(0)
    assert t.is_contiguous(), "native ops need contiguous tensors"
    return ctypes.c_void_p(t.data_ptr())


def launch_count() -> int:
    return int(cuda_lib().afl_launch_count()) if _cuda is not None else 0


def reset_launch_count() -> None:
    if _cuda is not None:
        _cuda.afl_reset_launch_count()


if __name__ == "__main__":
    print(build_cpu(force="--force" in sys.argv))
    print(build_cuda(force="--force" in sys.argv, verbose="-v" in sys.argv))
