"""ctypes view of oracle/liboracle.so — the CPU checker (test infrastructure only).

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may use this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
SO = os.path.join(ORACLE_DIR, "liboracle.so")

GEN_UNIT, GEN_BITS, GEN_DENORM = 0, 1, 2

_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    src = os.path.join(ORACLE_DIR, "oracle.c")
    if not os.path.exists(SO) or os.path.getmtime(SO) < os.path.getmtime(src):
        subprocess.run(["make", "-C", ORACLE_DIR], check=True, capture_output=True)
    L = C.CDLL(SO)
    u32, u64, sz, i32, f32 = C.c_uint32, C.c_uint64, C.c_size_t, C.c_int, C.c_float
    vp = C.c_void_p
    L.oracle_stream_key.restype = u32
    L.oracle_stream_key.argtypes = [u32, u32]
    L.oracle_hash32.restype = u32
    L.oracle_hash32.argtypes = [u32, u64]
    L.oracle_gen_f32_bits.restype = u32
    L.oracle_gen_f32_bits.argtypes = [u32, u64, i32]
    L.oracle_gen_bf16_bits.restype = C.c_uint16
    L.oracle_gen_bf16_bits.argtypes = [u32, u64, i32]
    L.oracle_fill_f32.restype = None
    L.oracle_fill_f32.argtypes = [vp, u64, sz, u32, u32, i32]
    L.oracle_fill_bf16.restype = None
    L.oracle_fill_bf16.argtypes = [vp, u64, sz, u32, u32, i32]
    L.oracle_bf16_to_f32.restype = f32
    L.oracle_bf16_to_f32.argtypes = [C.c_uint16]
    L.oracle_f32_to_bf16_rne.restype = C.c_uint16
    L.oracle_f32_to_bf16_rne.argtypes = [f32]
    for sfx in ("f32", "bf16"):
        getattr(L, f"oracle_add_{sfx}").restype = None
        getattr(L, f"oracle_add_{sfx}").argtypes = [vp, vp, vp, sz, i32]
        getattr(L, f"oracle_copy_{sfx}").restype = None
        getattr(L, f"oracle_copy_{sfx}").argtypes = [vp, vp, sz, i32]
        getattr(L, f"oracle_scale_{sfx}").restype = None
        getattr(L, f"oracle_scale_{sfx}").argtypes = [vp, f32, vp, sz, i32]
        getattr(L, f"oracle_triad_{sfx}").restype = None
        getattr(L, f"oracle_triad_{sfx}").argtypes = [vp, f32, vp, vp, sz, i32]
    L.oracle_checksum.restype = None
    L.oracle_checksum.argtypes = [vp, sz, u64, C.POINTER(u64), i32]
    L.oracle_vector_add_selftest.restype = i32
    L.oracle_vector_add_selftest.argtypes = [sz, u32, C.c_longlong]
    L.oracle_max_threads.restype = i32
    L.oracle_max_threads.argtypes = []
    _lib = L
    return L


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


NP = {"f32": np.uint32, "bf16": np.uint16}


def fill(dtype, start, n, seed, stream, mode=GEN_UNIT):
    """Return the generator's bit patterns (uint32 for f32, uint16 for bf16)."""
    out = np.empty(n, dtype=NP[dtype])
    getattr(lib(), f"oracle_fill_{dtype}")(_p(out), start, n, seed, stream, mode)
    return out


def op(name, dtype, n, s=None, x=None, y=None, threads=1):
    """Run one oracle loop on bit-pattern arrays; returns the output bit patterns.

    add(x, y) -> x+y ; copy(x) ; scale(s, x) -> s*x ; triad(s, x, y) -> fma(s, x, y)
    """
    out = np.empty(n, dtype=NP[dtype])
    L = lib()
    if name == "add":
        getattr(L, f"oracle_add_{dtype}")(_p(out), _p(x), _p(y), n, threads)
    elif name == "copy":
        getattr(L, f"oracle_copy_{dtype}")(_p(out), _p(x), n, threads)
    elif name == "scale":
        getattr(L, f"oracle_scale_{dtype}")(_p(out), s, _p(x), n, threads)
    elif name == "triad":
        getattr(L, f"oracle_triad_{dtype}")(_p(out), s, _p(x), _p(y), n, threads)
    else:
        raise ValueError(name)
    return out


def checksum(arr, word_offset=0, threads=1):
    arr = np.ascontiguousarray(arr)
    out = (C.c_uint64 * 2)()
    lib().oracle_checksum(_p(arr), arr.nbytes, word_offset, out, threads)
    return int(out[0]), int(out[1])


def f32_from_bits(bits):
    return np.array([bits], dtype=np.uint32).view(np.float32)[0].item()
