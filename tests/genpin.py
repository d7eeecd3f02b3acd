"""ctypes loader of tools/genpin.c (deterministic synthetic collections for the full-size parity
pins; test tooling).  Builds tools/_build/libgenpin.so + genpin with gcc when they are missing."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tools", "genpin.c")
BUILD = os.path.join(ROOT, "tools", "_build")
LIB = os.path.join(BUILD, "libgenpin.so")
EXE = os.path.join(BUILD, "genpin")
_lib = None


def build():
    os.makedirs(BUILD, exist_ok=True)
    if not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(SRC):
        subprocess.check_call(["gcc", "-O2", "-march=x86-64-v3", "-shared", "-fPIC", "-DGENPIN_NO_MAIN", "-o", LIB,
                               SRC, "-lpthread"])
    if not os.path.exists(EXE) or os.path.getmtime(EXE) < os.path.getmtime(SRC):
        subprocess.check_call(["gcc", "-O2", "-march=x86-64-v3", "-Wno-unused-function", "-o", EXE, SRC, "-lpthread"])


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB)
        u64, vp = C.c_uint64, C.c_void_p
        L.genpin_reference.argtypes = [u64, u64, vp]
        L.genpin_reference.restype = None
        L.genpin_text.argtypes = [u64, u64, u64, C.c_double, u64, C.c_int, C.c_int, vp, vp, u64, C.POINTER(u64)]
        L.genpin_text.restype = u64
        L.genpin_text_cap.argtypes = [u64, u64, C.c_double, C.c_int]
        L.genpin_text_cap.restype = u64
        _lib = L
    return _lib


def make_text(ref_bp, copies, div, seed, svs=0, first=0, count=None, threads=None, out=None):
    """-> (ref uint8[ref_bp], text uint8[N + 64] (64 bytes of 0x02 padding), N, ndocs); `out`: a
    caller-provided uint8 buffer (e.g. pinned memory) of at least text_cap() + 64 bytes."""
    L = lib()
    if count is None:
        count = copies - first
    if threads is None:
        threads = min(32, os.cpu_count() or 1)
    cap = int(L.genpin_text_cap(ref_bp, count, div, svs))
    ref = np.empty(ref_bp, np.uint8)
    if out is None:
        out = np.empty(cap + 64, np.uint8)
    assert out.nbytes >= cap + 64
    nd = C.c_uint64(0)
    N = int(L.genpin_text(ref_bp, first, count, div, seed, svs, threads, ref.ctypes.data_as(C.c_void_p),
                          out.ctypes.data_as(C.c_void_p), cap, C.byref(nd)))
    if N == 0:
        raise RuntimeError("genpin_text: buffer overflow")
    out[N:N + 64] = 2
    return ref, out, N, int(nd.value)


def text_cap(ref_bp, copies, div, svs=0):
    return int(lib().genpin_text_cap(ref_bp, copies, div, svs)) + 64


def write_files(ref_bp, copies, div, seed, svs, prefix, threads=8):
    build()
    subprocess.check_call([EXE, str(ref_bp), str(copies), repr(div), str(seed), str(svs), prefix, str(threads)])
