"""Loader/builder for the in-tree sm_100a kernel library (``lib/libveles_b200.so``).

The library is plain ``nvcc -shared`` output with a C ABI (see ``include/veles_b200.h``);
Python talks to it through :mod:`ctypes`, the C++ inference runtime links it directly.
It replaces the reference's run-time ``nvcc`` JIT + kernel cache
(``veles/accelerated_units.py:298-419, 605-673``): kernels are compiled once, ahead of
time, for exactly one target (sm_100a).
"""
from __future__ import annotations

import ctypes
import os
import shutil
import subprocess
import threading

PKG_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC_DIR = os.path.join(PKG_DIR, "csrc")
LIB_DIR = os.path.join(PKG_DIR, "lib")
LIB_PATH = os.path.join(LIB_DIR, "libveles_b200.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared",
]

_lock = threading.Lock()
_lib = None


PDL_DEFAULT = False


def kernel_sources():
    return sorted(os.path.join(CSRC_DIR, f) for f in os.listdir(CSRC_DIR) if f.endswith(".cu"))


def _stale():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = kernel_sources() + [os.path.join(CSRC_DIR, f) for f in os.listdir(CSRC_DIR)
                               if f.endswith((".cuh", ".h"))]
    return any(os.path.getmtime(d) > t for d in deps)


def find_nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    return None


def build(force=False, verbose=False):
    """Compile every ``csrc/*.cu`` into one shared object for sm_100a."""
    with _lock:
        if not force and not _stale():
            return LIB_PATH
        nvcc = find_nvcc()
        if nvcc is None:
            raise RuntimeError("nvcc not found: cannot build %s" % LIB_PATH)
        os.makedirs(LIB_DIR, exist_ok=True)
        objs = []
        procs = []
        build_dir = os.path.join(LIB_DIR, "obj")
        os.makedirs(build_dir, exist_ok=True)
        for src in kernel_sources():
            obj = os.path.join(build_dir, os.path.basename(src)[:-3] + ".o")
            objs.append(obj)
            cmd = [nvcc] + [f for f in NVCC_FLAGS if f != "-shared"] + ["-I", CSRC_DIR, "-c", src, "-o", obj]
            procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
        for cmd, p in procs:
            out, _ = p.communicate()
            if p.returncode != 0:
                raise RuntimeError("nvcc failed: %s\n%s" % (" ".join(cmd), out.decode(errors="replace")))
            if verbose and out:
                print(out.decode(errors="replace"))
        tmp = LIB_PATH + ".tmp"
        cmd = [nvcc, "-shared", "-o", tmp] + objs + ["-lcudart"]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
        if r.returncode != 0:
            raise RuntimeError("link failed: %s\n%s" % (" ".join(cmd), r.stdout.decode(errors="replace")))
        os.replace(tmp, LIB_PATH)
        return LIB_PATH


class KernelError(RuntimeError):
    pass


def lib():
    """Return the loaded library; builds it when missing.  Raises loudly if it cannot
    be made available — there is no silent eager fallback on a GPU box."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
    path = LIB_PATH
    if _stale():
        if find_nvcc() is not None:
            path = build()
        elif not os.path.exists(path):
            raise KernelError("kernel library %s is missing and nvcc is unavailable" % path)
    handle = ctypes.CDLL(path, mode=ctypes.RTLD_GLOBAL)
    handle.vb_last_error.restype = ctypes.c_char_p
    handle.vb_col_sum_scratch_elems.restype = ctypes.c_longlong
    # programmatic dependent launch between consecutive kernels (root.common.engine.pdl / VELES_B200_PDL)
    from ..config import root, get
    pdl = os.environ.get("VELES_B200_PDL")
    handle.vb_set_pdl(ctypes.c_int(int(pdl) if pdl is not None else int(bool(get(root.common.engine.pdl, PDL_DEFAULT)))))
    _lib = handle
    return _lib


def check(rc, what):
    if rc != 0:
        msg = lib().vb_last_error()
        raise KernelError("%s failed (%d): %s" % (what, rc, msg.decode(errors="replace") if msg else "?"))
