"""CPU tier: the oracle against the exact-arithmetic known-answer vectors (tests/golden/)."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import oracle_lib as O

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def load(name):
    with open(os.path.join(GOLD, name)) as f:
        return json.load(f)


def cols(rows, dtype):
    a = np.array(rows, dtype=np.uint64)
    return [np.ascontiguousarray(a[:, i].astype(dtype)) for i in range(a.shape[1])]


def test_kat_f32_add_scale_triad():
    k = load("kat_f32.json")
    a, b, want = cols(k["add"], np.uint32)
    np.testing.assert_array_equal(O.op("add", "f32", len(a), x=a, y=b), want)

    s, c, want = cols(k["scale"], np.uint32)
    for sv in np.unique(s):
        m = s == sv
        got = O.op("scale", "f32", int(m.sum()), s=O.f32_from_bits(int(sv)), x=np.ascontiguousarray(c[m]))
        np.testing.assert_array_equal(got, want[m])

    s, c, b, fused, unfused = cols(k["triad"], np.uint32)
    assert (fused != unfused).sum() >= 250  # the vectors really separate fma from mul+add
    for sv in np.unique(s):
        m = s == sv
        got = O.op("triad", "f32", int(m.sum()), s=O.f32_from_bits(int(sv)),
                   x=np.ascontiguousarray(c[m]), y=np.ascontiguousarray(b[m]))
        np.testing.assert_array_equal(got, fused[m])


def test_kat_bf16():
    k = load("kat_bf16.json")
    L = O.lib()
    for b32, want in k["rne"]:
        f = np.array([b32], dtype=np.uint32).view(np.float32)[0]
        assert L.oracle_f32_to_bf16_rne(C.c_float(f)) == want, hex(b32)
    a, b, want = cols(k["add"], np.uint16)
    np.testing.assert_array_equal(O.op("add", "bf16", len(a), x=a, y=b), want)
    rows = np.array(k["scale"], dtype=np.uint64)
    for sv in np.unique(rows[:, 0]):
        m = rows[:, 0] == sv
        c = np.ascontiguousarray(rows[m, 1].astype(np.uint16))
        got = O.op("scale", "bf16", len(c), s=O.f32_from_bits(int(sv)), x=c)
        np.testing.assert_array_equal(got, rows[m, 2].astype(np.uint16))
    rows = np.array(k["triad"], dtype=np.uint64)
    for sv in np.unique(rows[:, 0]):
        m = rows[:, 0] == sv
        c = np.ascontiguousarray(rows[m, 1].astype(np.uint16))
        b = np.ascontiguousarray(rows[m, 2].astype(np.uint16))
        got = O.op("triad", "bf16", len(c), s=O.f32_from_bits(int(sv)), x=c, y=b)
        np.testing.assert_array_equal(got, rows[m, 3].astype(np.uint16))


def test_kat_generator_and_checksum():
    k = load("kat_gen.json")
    L = O.lib()
    for seed, stream, key in k["keys"]:
        assert L.oracle_stream_key(seed, stream) == key
    for key, i, h in k["hash"]:
        assert L.oracle_hash32(key, i) == h
    for key, i, mode, bits in k["f32"]:
        assert L.oracle_gen_f32_bits(key, i, mode) == bits
    for key, i, mode, bits in k["bf16"]:
        assert L.oracle_gen_bf16_bits(key, i, mode) == bits
    for hexbuf, off, s0, s1 in k["checksum"]:
        buf = np.frombuffer(bytes.fromhex(hexbuf), dtype=np.uint8)
        assert O.checksum(buf, off) == (s0, s1)


@pytest.mark.parametrize("dtype", ["f32", "bf16"])
@pytest.mark.parametrize("mode", [O.GEN_UNIT, O.GEN_BITS, O.GEN_DENORM])
def test_serial_equals_openmp(dtype, mode):
    n = 100_003
    x = O.fill(dtype, 7, n, 11, 0, mode)
    y = O.fill(dtype, 7, n, 11, 1, mode)
    for name in ("add", "copy", "scale", "triad"):
        one = O.op(name, dtype, n, s=3.0, x=x, y=y, threads=1)
        many = O.op(name, dtype, n, s=3.0, x=x, y=y, threads=0)
        np.testing.assert_array_equal(one, many)
    assert O.checksum(x, 5, threads=1) == O.checksum(x, 5, threads=0)


def test_fill_is_index_derived_and_finite():
    whole = O.fill("f32", 0, 5000, 3, 2, O.GEN_BITS)
    part = O.fill("f32", 1234, 1000, 3, 2, O.GEN_BITS)
    np.testing.assert_array_equal(whole[1234:2234], part)
    assert ((whole >> 23) & 0xFF).max() < 0xFF
    unit = O.fill("f32", 0, 5000, 3, 2, O.GEN_UNIT).view(np.float32)
    assert unit.min() >= 0.0 and unit.max() < 1.0
    hi = O.fill("bf16", (1 << 32) - 10, 20, 3, 2, O.GEN_BITS)
    lo = O.fill("bf16", 0, 20, 3, 2, O.GEN_BITS)
    assert not np.array_equal(hi[10:], lo[:10])  # index bits above 2^32 matter


def test_checksum_shards_add_up():
    x = O.fill("f32", 0, 40_001, 5, 0, O.GEN_BITS)
    whole = O.checksum(x)
    cut = 12_345
    a = O.checksum(x[:cut], 0)
    b = O.checksum(x[cut:], cut)
    M = (1 << 64) - 1
    assert ((a[0] + b[0]) & M, (a[1] + b[1]) & M) == whole
    y = x.copy()
    y[[100, 200]] = y[[200, 100]]  # a swap keeps the plain sum, breaks the weighted one
    cy = O.checksum(y)
    assert cy[0] == whole[0] and cy[1] != whole[1]


def test_config0_vector_add_contract_on_host():
    """BASELINE.json configs[0]: N=50,000 fp32, PASSED/FAILED decided by the host loop."""
    L = O.lib()
    assert L.oracle_vector_add_selftest(50_000, 20261106, -1) == 0
    assert L.oracle_vector_add_selftest(50_000, 20261106, 49_999) == 1
    assert L.oracle_vector_add_selftest(1, 1, -1) == 0
    assert L.oracle_vector_add_selftest(0, 1, -1) == 0


def test_oracle_under_sanitizers():
    """The oracle's C sources under AddressSanitizer + UBSan (oracle/selftest.c): ragged sizes on
    exact-size heap blocks, serial and threaded.  The sanitized build must run clean and print what the
    plain build prints; spot lines are re-derived through the ctypes view."""
    import subprocess
    odir = O.ORACLE_DIR
    r = subprocess.run(["make", "-C", odir, "sanitize"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=1:abort_on_error=0", UBSAN_OPTIONS="print_stacktrace=1", OMP_NUM_THREADS="3")
    plain = subprocess.run([os.path.join(odir, "selftest")], capture_output=True, text=True, env=env, timeout=300)
    san = subprocess.run([os.path.join(odir, "selftest_san")], capture_output=True, text=True, env=env, timeout=600)
    assert plain.returncode == 0 and san.returncode == 0, san.stderr[-3000:]
    assert "ERROR" not in san.stderr and "runtime error" not in san.stderr, san.stderr[-3000:]
    assert plain.stdout == san.stdout and plain.stdout.strip().endswith("selftest 0 1 0")
    lines = plain.stdout.splitlines()
    assert len(lines) == 3 * 6 * 2 * 8 + 1
    n, s = 100003, -0.333251953125
    x = O.fill("f32", (1 << 32) - 5, n, 7, 0, O.GEN_BITS)
    y = O.fill("f32", (1 << 32) - 5, n, 7, 1, O.GEN_BITS)
    c = O.checksum(O.op("triad", "f32", n, s=s, x=x, y=y), 3)
    assert f"triad f32 mode=1 n={n} {c[0]:016x}:{c[1]:016x}" in lines
    hx = O.fill("bf16", 11, n, 7, 0, O.GEN_DENORM)
    hy = O.fill("bf16", 11, n, 7, 1, O.GEN_DENORM)
    c = O.checksum(O.op("add", "bf16", n, x=hx, y=hy), 3)
    assert f"add bf16 mode=2 n={n} {c[0]:016x}:{c[1]:016x}" in lines
