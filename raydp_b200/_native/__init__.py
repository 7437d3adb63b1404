"""In-tree native libraries: build helpers and loaders.

Two shared objects live next to this file (git-ignored, shipped to the GPU box by gpurun):

* ``librdp_store.so``   -- C++ shared-memory object store (``csrc/store``), CPU only, built
  with g++ on first use if missing.
* ``librdp_kernels.so`` -- every sm_100a CUDA kernel plus the communicator bootstrap
  (``csrc/kernels``, ``csrc/comm``), built by ``__graft_entry__.build()`` with
  ``nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo``.

Both expose a plain C ABI (``include/*.h``) and are loaded with ctypes, so no torch headers
are compiled and a rebuild is seconds.
"""
from __future__ import annotations

import ctypes
import glob
import os
import shutil
import subprocess
import threading
from typing import List, Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(os.path.dirname(_HERE))
CSRC = os.path.join(REPO_ROOT, "csrc")
INCLUDE = os.path.join(REPO_ROOT, "include")
STORE_LIB = os.path.join(_HERE, "librdp_store.so")
KERNELS_LIB = os.path.join(_HERE, "librdp_kernels.so")

_lock = threading.Lock()
_store = None
_kernels = None

NVCC_ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def _newer(target: str, sources: List[str]) -> bool:
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources if os.path.exists(s))


def build_store(force: bool = False, verbose: bool = False) -> str:
    srcs = sorted(glob.glob(os.path.join(CSRC, "store", "*.cc")))
    deps = srcs + sorted(glob.glob(os.path.join(INCLUDE, "*.h")))
    if not force and _newer(STORE_LIB, deps):
        return STORE_LIB
    if not srcs:
        if os.path.exists(STORE_LIB):
            return STORE_LIB
        raise RuntimeError("csrc/store sources are missing and librdp_store.so is not built")
    cxx = shutil.which("g++") or shutil.which("c++")
    if cxx is None:
        raise RuntimeError("g++ not found; cannot build the object store")
    tmp = STORE_LIB + f".tmp{os.getpid()}"
    cmd = [cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-o", tmp, *srcs,
           "-lrt", "-lpthread", "-ldl"]
    if verbose:
        print(" ".join(cmd))
    subprocess.check_call(cmd)
    os.replace(tmp, STORE_LIB)
    return STORE_LIB


def kernel_sources() -> List[str]:
    return sorted(glob.glob(os.path.join(CSRC, "kernels", "*.cu")) +
                  glob.glob(os.path.join(CSRC, "comm", "*.cu")))


def nvcc_path() -> Optional[str]:
    cand = shutil.which("nvcc")
    if cand:
        return cand
    cand = "/usr/local/cuda/bin/nvcc"
    return cand if os.path.exists(cand) else None


def build_kernels(force: bool = False, verbose: bool = False, jobs: int = 0) -> str:
    """Compile every .cu to an object (in parallel) and link ``librdp_kernels.so``."""
    srcs = kernel_sources()
    hdrs = sorted(glob.glob(os.path.join(INCLUDE, "*.h")) +
                  glob.glob(os.path.join(CSRC, "kernels", "*.cuh")) +
                  glob.glob(os.path.join(CSRC, "comm", "*.cuh")))
    if not force and _newer(KERNELS_LIB, srcs + hdrs):
        return KERNELS_LIB
    nvcc = nvcc_path()
    if nvcc is None:
        raise RuntimeError("nvcc not found; cannot build CUDA kernels")
    objdir = os.path.join(REPO_ROOT, "build", "obj")
    os.makedirs(objdir, exist_ok=True)
    common = [nvcc, *NVCC_ARCH, "-lineinfo", "-O3", "-std=c++17", "--expt-relaxed-constexpr",
              "-Xcompiler", "-fPIC", "-I", INCLUDE, "-I", os.path.join(CSRC, "kernels"),
              "-I", os.path.join(CSRC, "comm")]
    procs = []
    objs = []
    for src in srcs:
        obj = os.path.join(objdir, os.path.basename(src) + ".o")
        objs.append(obj)
        if not force and _newer(obj, [src] + hdrs):
            continue
        cmd = common + ["-Xptxas", "-v", "-c", src, "-o", obj]
        if verbose:
            print(" ".join(cmd))
        log = open(obj + ".log", "w")
        procs.append((src, subprocess.Popen(cmd, stdout=log, stderr=subprocess.STDOUT), log))
        if jobs and len(procs) >= jobs:
            _drain(procs)
    _drain(procs)
    tmp = KERNELS_LIB + f".tmp{os.getpid()}"
    link = [nvcc, *NVCC_ARCH, "-shared", "-Xcompiler", "-fPIC", "-o", tmp, *objs, "-ldl"]
    if verbose:
        print(" ".join(link))
    subprocess.check_call(link)
    os.replace(tmp, KERNELS_LIB)
    return KERNELS_LIB


def _drain(procs) -> None:
    failed = []
    for src, proc, log in procs:
        rc = proc.wait()
        log.close()
        if rc != 0:
            failed.append(src)
    procs.clear()
    if failed:
        msgs = []
        for src in failed:
            path = os.path.join(REPO_ROOT, "build", "obj", os.path.basename(src) + ".o.log")
            with open(path) as fh:
                msgs.append(f"--- {src}\n" + fh.read()[-4000:])
        raise RuntimeError("nvcc failed:\n" + "\n".join(msgs))


def store_lib() -> ctypes.CDLL:
    """The object-store library, built on first use (needs only g++)."""
    global _store
    if _store is not None:
        return _store
    with _lock:
        if _store is None:
            path = build_store()
            lib = ctypes.CDLL(path)
            _declare_store(lib)
            _store = lib
    return _store


def kernels_available() -> bool:
    return os.path.exists(KERNELS_LIB)


def kernels_lib() -> ctypes.CDLL:
    """The CUDA kernel library.  Never falls back silently: on a GPU box a missing
    extension is an error (``__graft_entry__.build()`` creates it)."""
    global _kernels
    if _kernels is not None:
        return _kernels
    with _lock:
        if _kernels is None:
            if not os.path.exists(KERNELS_LIB):
                raise RuntimeError(
                    f"{KERNELS_LIB} is missing: run `python -c 'import __graft_entry__ as g; "
                    "g.build()'` (nvcc, sm_100a) before using raydp_b200 GPU ops")
            lib = ctypes.CDLL(KERNELS_LIB)
            from raydp_b200.ops import _abi
            _abi.declare(lib)
            _kernels = lib
    return _kernels


def _declare_store(lib: ctypes.CDLL) -> None:
    c = ctypes
    vp, u64, i64 = c.c_void_p, c.c_uint64, c.c_int64
    lib.rdp_store_create.argtypes = [c.c_char_p, u64, u64, c.POINTER(vp)]
    lib.rdp_store_create.restype = c.c_int
    lib.rdp_store_open.argtypes = [c.c_char_p, c.c_int, c.POINTER(vp)]
    lib.rdp_store_open.restype = c.c_int
    lib.rdp_store_seal.argtypes = [vp]
    lib.rdp_store_seal.restype = c.c_int
    for fn in ("rdp_store_data", "rdp_store_meta"):
        getattr(lib, fn).argtypes = [vp]
        getattr(lib, fn).restype = vp
    for fn in ("rdp_store_data_size", "rdp_store_meta_size", "rdp_store_created_ns"):
        getattr(lib, fn).argtypes = [vp]
        getattr(lib, fn).restype = u64
    lib.rdp_store_shrink.argtypes = [vp, u64]
    lib.rdp_store_shrink.restype = c.c_int
    lib.rdp_store_incref.argtypes = [vp, i64]
    lib.rdp_store_incref.restype = i64
    lib.rdp_store_pin.argtypes = [vp]
    lib.rdp_store_pin.restype = c.c_int
    lib.rdp_store_close.argtypes = [vp]
    lib.rdp_store_close.restype = c.c_int
    lib.rdp_store_unlink.argtypes = [c.c_char_p]
    lib.rdp_store_unlink.restype = c.c_int
    lib.rdp_store_exists.argtypes = [c.c_char_p]
    lib.rdp_store_exists.restype = c.c_int
    lib.rdp_store_unlink_prefix.argtypes = [c.c_char_p]
    lib.rdp_store_unlink_prefix.restype = i64
    lib.rdp_store_usage.argtypes = [c.c_char_p, c.POINTER(u64)]
    lib.rdp_store_usage.restype = i64
    lib.rdp_store_parallel_copy.argtypes = [vp, vp, u64, c.c_int]
    lib.rdp_store_parallel_copy.restype = None
