"""kubernetes-with-nvidia-gpu_b200 — B200-native GPU validation workload (harness-side binding).

The product is native: ``libb200v.so`` (sm_100a kernels behind the C ABI in ``include/b200v.h``)
and the ``gpu-validate`` pod entrypoint that replaces the ``nvidia-smi`` payload of the reference's
Step 9 pod (/root/reference/README.md:302-319).  This module is only the ctypes view of that ABI
that ``bench.py``, ``__graft_entry__.py`` and ``tests/`` use; it adds no compute path of its own
and there is no CPU fallback: if the library is missing or no GPU is usable, calls raise.

The directory name has a hyphen, so import it with
``importlib.import_module("kubernetes-with-nvidia-gpu_b200")``.
"""
import ctypes as C
import os
import subprocess

import numpy as np

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG_DIR)
LIB_PATH = os.path.join(PKG_DIR, "libb200v.so")
LAUNCHER_PATH = os.path.join(PKG_DIR, "gpu-validate")

OK, ERR_INVALID, ERR_NO_DEVICE, ERR_CUDA, ERR_ARCH, ERR_ALIGN, ERR_PEER, ERR_MISMATCH, ERR_NOMEM = range(9)
F32, BF16 = 0, 1
OP_ADD, OP_COPY, OP_SCALE, OP_TRIAD = 0, 1, 2, 3
OPS = {"add": OP_ADD, "copy": OP_COPY, "scale": OP_SCALE, "triad": OP_TRIAD}
DTYPES = {"f32": F32, "bf16": BF16}
GEN_UNIT, GEN_BITS, GEN_DENORM = 0, 1, 2
NP_BITS = {"f32": np.uint32, "bf16": np.uint16}


class B200VError(RuntimeError):
    def __init__(self, status, name, text):
        super().__init__(f"{name} ({status}): {text}")
        self.status = status
        self.name = name


class DeviceInfo(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("device", C.c_int32),
        ("name", C.c_char * 128), ("pci_bus_id", C.c_char * 32), ("uuid", C.c_char * 48),
        ("cc_major", C.c_int32), ("cc_minor", C.c_int32), ("sm_count", C.c_int32),
        ("clock_khz", C.c_int32), ("mem_clock_khz", C.c_int32), ("mem_bus_bits", C.c_int32),
        ("hbm_bytes", C.c_uint64), ("hbm_free_bytes", C.c_uint64), ("l2_bytes", C.c_uint64),
        ("driver_version", C.c_int32), ("runtime_version", C.c_int32),
        ("ecc_enabled", C.c_int32), ("reserved", C.c_int32),
    ]


class Timing(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("iters", C.c_int32),
        ("best_ms", C.c_float), ("median_ms", C.c_float), ("mean_ms", C.c_float),
        ("bytes_per_launch", C.c_uint64),
    ]


def build(verbose=False):
    """Compile libb200v.so and gpu-validate in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    r = subprocess.run(["make", "-C", PKG_DIR, "all"], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError("building libb200v failed:\n" + r.stdout + r.stderr)


_SIGS = None


def _signatures():
    vp, sz, u64, u32, i32, f32 = C.c_void_p, C.c_size_t, C.c_uint64, C.c_uint32, C.c_int, C.c_float
    pi = C.POINTER(C.c_int)
    pvp = C.POINTER(C.c_void_p)
    s = {
        "b200v_abi_version": (i32, []),
        "b200v_last_error": (C.c_char_p, []),
        "b200v_status_name": (C.c_char_p, [i32]),
        "b200v_device_count": (i32, [pi]),
        "b200v_device_info_get": (i32, [i32, C.POINTER(DeviceInfo)]),
        "b200v_set_device": (i32, [i32]),
        "b200v_get_device": (i32, [pi]),
        "b200v_peer_matrix": (i32, [i32, pi]),
        "b200v_require_sm100": (i32, [i32]),
        "b200v_malloc": (i32, [pvp, sz]),
        "b200v_free": (i32, [vp]),
        "b200v_host_alloc": (i32, [pvp, sz]),
        "b200v_host_free": (i32, [vp]),
        "b200v_host_register": (i32, [vp, sz]),
        "b200v_host_unregister": (i32, [vp]),
        "b200v_memcpy_h2d": (i32, [vp, vp, sz, vp]),
        "b200v_memcpy_d2h": (i32, [vp, vp, sz, vp]),
        "b200v_memcpy_d2d": (i32, [vp, vp, sz, vp]),
        "b200v_memset": (i32, [vp, i32, sz, vp]),
        "b200v_stream_create": (i32, [pvp]),
        "b200v_stream_destroy": (i32, [vp]),
        "b200v_stream_sync": (i32, [vp]),
        "b200v_device_sync": (i32, []),
        "b200v_fill": (i32, [i32, vp, u64, sz, u32, u32, i32, vp]),
        "b200v_add_f32": (i32, [vp, vp, vp, sz, vp]),
        "b200v_copy_f32": (i32, [vp, vp, sz, vp]),
        "b200v_scale_f32": (i32, [vp, f32, vp, sz, vp]),
        "b200v_triad_f32": (i32, [vp, f32, vp, vp, sz, vp]),
        "b200v_add_bf16": (i32, [vp, vp, vp, sz, vp]),
        "b200v_copy_bf16": (i32, [vp, vp, sz, vp]),
        "b200v_scale_bf16": (i32, [vp, f32, vp, sz, vp]),
        "b200v_triad_bf16": (i32, [vp, f32, vp, vp, sz, vp]),
        "b200v_stream_op": (i32, [i32, i32, vp, vp, vp, f32, sz, vp]),
        "b200v_op_bytes": (u64, [i32, i32, sz]),
        "b200v_vector_add_host": (i32, [vp, vp, vp, sz]),
        "b200v_vector_add_host_release": (i32, []),
        "b200v_vector_add_host_config": (i32, [sz, i32]),
        "b200v_checksum_workspace_bytes": (sz, []),
        "b200v_checksum": (i32, [vp, sz, u64, vp, sz, vp, vp]),
        "b200v_checksum_static": (i32, [vp, sz, u64, vp, vp]),
        "b200v_event_create": (i32, [pvp]),
        "b200v_event_destroy": (i32, [vp]),
        "b200v_event_record": (i32, [vp, vp]),
        "b200v_event_sync": (i32, [vp]),
        "b200v_event_elapsed_ms": (i32, [vp, vp, C.POINTER(C.c_float)]),
        "b200v_time_op": (i32, [i32, i32, vp, vp, vp, f32, sz, i32, i32, vp, C.POINTER(Timing)]),
        "b200v_shard_range": (i32, [u64, i32, i32, C.POINTER(u64), C.POINTER(u64)]),
        "b200v_peer_enable": (i32, [i32, i32]),
        "b200v_peer_copy": (i32, [vp, i32, vp, i32, sz, vp]),
        "b200v_peer_pull": (i32, [vp, vp, sz, vp]),
        "b200v_peer_push": (i32, [vp, vp, sz, vp]),
    }
    return s


_lib = None


def lib():
    """Load libb200v.so (never a substitute).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C kubernetes-with-nvidia-gpu_b200`). There is no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    for name, (res, args) in _signatures().items():
        fn = getattr(L, name)  # AttributeError here = header and library disagree
        fn.restype = res
        fn.argtypes = args
    if L.b200v_abi_version() != 1:
        raise RuntimeError("libb200v ABI version mismatch")
    _lib = L
    return L


def check(status):
    if status != OK:
        L = lib()
        raise B200VError(status, L.b200v_status_name(status).decode(), L.b200v_last_error().decode())


# ---------------------------------------------------------------------------------------------
# small conveniences over the ABI (no logic of their own)

def device_count():
    n = C.c_int(0)
    check(lib().b200v_device_count(C.byref(n)))
    return n.value


def device_info(device=0):
    info = DeviceInfo()
    info.struct_size = C.sizeof(DeviceInfo)
    check(lib().b200v_device_info_get(device, C.byref(info)))
    return info


def shard_range(n, rank, world):
    b, c = C.c_uint64(0), C.c_uint64(0)
    check(lib().b200v_shard_range(n, rank, world, C.byref(b), C.byref(c)))
    return b.value, c.value


class DeviceBuffer:
    """cudaMalloc'd bytes owned through the ABI (b200v_malloc / b200v_free)."""

    def __init__(self, nbytes):
        self.nbytes = int(nbytes)
        p = C.c_void_p(None)
        check(lib().b200v_malloc(C.byref(p), max(self.nbytes, 1)))
        self.ptr = p.value

    def at(self, byte_offset):
        return C.c_void_p(self.ptr + int(byte_offset))

    def upload(self, arr, byte_offset=0):
        arr = np.ascontiguousarray(arr)
        assert byte_offset + arr.nbytes <= self.nbytes
        check(lib().b200v_memcpy_h2d(self.at(byte_offset), arr.ctypes.data_as(C.c_void_p), arr.nbytes, None))
        check(lib().b200v_device_sync())
        return self

    def download(self, dtype, count, byte_offset=0):
        out = np.empty(int(count), dtype=dtype)
        assert byte_offset + out.nbytes <= self.nbytes
        check(lib().b200v_memcpy_d2h(out.ctypes.data_as(C.c_void_p), self.at(byte_offset), out.nbytes, None))
        check(lib().b200v_device_sync())
        return out

    def free(self):
        if self.ptr:
            check(lib().b200v_free(C.c_void_p(self.ptr)))
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class PinnedBuffer:
    """Pinned host bytes (b200v_host_alloc) viewed as a numpy array."""

    def __init__(self, nbytes):
        self.nbytes = int(nbytes)
        p = C.c_void_p(None)
        check(lib().b200v_host_alloc(C.byref(p), max(self.nbytes, 1)))
        self.ptr = p.value

    def view(self, dtype):
        n = self.nbytes // np.dtype(dtype).itemsize
        return np.ctypeslib.as_array(C.cast(self.ptr, C.POINTER(C.c_uint8)), shape=(self.nbytes,)).view(dtype)[:n]

    def free(self):
        if self.ptr:
            check(lib().b200v_host_free(C.c_void_p(self.ptr)))
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def stream_op(op, dtype, dst, x, y, s, n, stream=None):
    """dst = op(x, y, s) on device pointers (ints / c_void_p); enqueue only."""
    check(lib().b200v_stream_op(OPS[op], DTYPES[dtype], dst, x, y if y is not None else None, s, n, stream))


def checksum(ptr, nbytes, word_offset=0, workspace=None):
    """(sum, weighted sum) of a device buffer — see b200v_checksum."""
    L = lib()
    own = workspace is None
    ws = workspace or DeviceBuffer(L.b200v_checksum_workspace_bytes())
    out = DeviceBuffer(16)
    try:
        check(L.b200v_checksum(ptr, nbytes, word_offset, C.c_void_p(ws.ptr), ws.nbytes, C.c_void_p(out.ptr), None))
        r = out.download(np.uint64, 2)
    finally:
        out.free()
        if own:
            ws.free()
    return int(r[0]), int(r[1])
