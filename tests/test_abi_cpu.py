"""CPU tier: the C-ABI library loads, exports what include/b200v.h declares, and its host-only
entry points behave (no compute call is made here — there is no GPU in this tier)."""
import ctypes as C
import os
import re
import subprocess

import pytest

from b200v_pkg import pkg, ROOT


@pytest.fixture(scope="module")
def L():
    if not os.path.exists(pkg.LIB_PATH):
        pkg.build()
    return pkg.lib()


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200v.h")).read()
    return sorted(set(re.findall(r"B200V_API[^;(]*?\b(b200v_\w+)\s*\(", text)))


def test_header_symbols_are_exported(L):
    names = declared_symbols()
    assert len(names) >= 40
    out = subprocess.run(["nm", "-D", "--defined-only", pkg.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r" T (b200v_\w+)", out))
    missing = [n for n in names if n not in exported]
    assert not missing, f"declared in b200v.h but not exported: {missing}"
    # and the binding knows every one of them
    assert sorted(pkg._signatures()) == names


def test_library_is_sm100a_only(L):
    out = subprocess.run(["cuobjdump", "-lelf", pkg.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_no_oracle_in_product(L):
    """The product never links or names the oracle (oracle/ is test infrastructure)."""
    out = subprocess.run(["nm", "-D", pkg.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle_" not in out
    ldd = subprocess.run(["ldd", pkg.LIB_PATH], capture_output=True, text=True).stdout
    assert "liboracle" not in ldd and "libtorch" not in ldd and "libcudart" not in ldd  # cudart is static
    for dirpath, _, files in os.walk(os.path.join(ROOT, "kubernetes-with-nvidia-gpu_b200")):
        for f in files:
            if f.endswith((".cu", ".cuh", ".cpp", ".h", ".py", "Makefile")):
                src = open(os.path.join(dirpath, f)).read()
                assert "liboracle" not in src and "oracle.h" not in src and "oracle_lib" not in src, f
    # the developer tools stay clear of it too: only tests/, __graft_entry__.smoke() and bench.py's CPU legs use the oracle
    for f in os.listdir(os.path.join(ROOT, "tools")):
        if f.endswith((".py", ".cu", ".c", ".sh", "Makefile")):
            src = open(os.path.join(ROOT, "tools", f)).read()
            assert "liboracle" not in src and "oracle.h" not in src and "oracle_lib" not in src, f


def test_status_names_and_version(L):
    assert L.b200v_abi_version() == 1
    assert L.b200v_status_name(0) == b"B200V_OK"
    assert L.b200v_status_name(2) == b"B200V_ERR_NO_DEVICE"
    assert L.b200v_status_name(99) == b"B200V_ERR_UNKNOWN"


def test_shard_range(L):
    for n, world in ((1 << 32, 8), (1 << 32, 4), (1 << 32, 2), (50_000, 8), (7, 8), (0, 3), (10, 3)):
        spans = [pkg.shard_range(n, r, world) for r in range(world)]
        assert spans[0][0] == 0
        for (b0, c0), (b1, _) in zip(spans, spans[1:]):
            assert b0 + c0 == b1
        assert spans[-1][0] + spans[-1][1] == n
        counts = [c for _, c in spans]
        assert max(counts) - min(counts) <= 1
    assert pkg.shard_range(1 << 32, 3, 8) == (3 << 29, 1 << 29)
    b, c = C.c_uint64(), C.c_uint64()
    assert L.b200v_shard_range(10, 3, 3, C.byref(b), C.byref(c)) == pkg.ERR_INVALID
    assert b"rank 3" in L.b200v_last_error()
    assert L.b200v_shard_range(10, 0, 0, C.byref(b), C.byref(c)) == pkg.ERR_INVALID
    assert L.b200v_shard_range(10, 0, 1, None, None) == pkg.ERR_INVALID


def test_op_bytes(L):
    n = 1 << 28
    assert L.b200v_op_bytes(pkg.OP_ADD, pkg.F32, n) == 12 * n == 3_221_225_472
    assert L.b200v_op_bytes(pkg.OP_TRIAD, pkg.F32, n) == 12 * n
    assert L.b200v_op_bytes(pkg.OP_COPY, pkg.F32, n) == 8 * n
    assert L.b200v_op_bytes(pkg.OP_SCALE, pkg.BF16, n) == 4 * n
    assert L.b200v_op_bytes(pkg.OP_TRIAD, pkg.BF16, 1 << 32) == 6 << 32


def _gpu_visible():
    try:
        return pkg.device_count() > 0
    except Exception:
        return False


@pytest.mark.skipif(_gpu_visible(), reason="checks the no-GPU failure mode")
def test_fails_loudly_without_a_gpu(L):
    """No CPU fallback: with no device every device entry point reports an error."""
    n = C.c_int(-1)
    rc = L.b200v_device_count(C.byref(n))
    assert rc == pkg.ERR_NO_DEVICE and n.value == 0
    assert L.b200v_last_error()
    p = C.c_void_p()
    assert L.b200v_malloc(C.byref(p), 1024) != pkg.OK
    with pytest.raises(pkg.B200VError):
        pkg.device_info(0)
    buf = (C.c_float * 8)()
    assert L.b200v_vector_add_host(buf, buf, buf, 8) != pkg.OK


def test_argument_checks_need_no_gpu(L):
    assert L.b200v_stream_op(9, pkg.F32, None, None, None, 0.0, 0, None) == pkg.ERR_INVALID
    assert L.b200v_stream_op(pkg.OP_ADD, 7, None, None, None, 0.0, 0, None) == pkg.ERR_INVALID
    assert L.b200v_add_f32(None, None, None, 16, None) == pkg.ERR_INVALID
    assert L.b200v_fill(pkg.F32, None, 0, 4, 1, 0, pkg.GEN_UNIT, None) == pkg.ERR_INVALID
    assert L.b200v_fill(pkg.F32, C.c_void_p(256), 0, 4, 1, 0, 9, None) == pkg.ERR_INVALID
    assert L.b200v_vector_add_host(None, None, None, 4) == pkg.ERR_INVALID
    assert L.b200v_checksum(None, 16, 0, None, 0, None, None) == pkg.ERR_INVALID
