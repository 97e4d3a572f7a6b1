"""GPU tier: the CUDA path, called through the C ABI, against the CPU oracle — bit-exact.

Everything here goes through libb200v.so (ctypes); the oracle (oracle/liboracle.so) is only the
checker.  Tolerance: none.  Integer/byte/index work and IEEE fp32/bf16 results must match bit for
bit (BASELINE.json: "bit-exact against a host C reference loop, triad using explicit fmaf").
"""
import ctypes as C
import json
import os

import numpy as np
import pytest

import oracle_lib as O
from b200v_pkg import pkg

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden")
SEED = 20261106
M64 = (1 << 64) - 1


@pytest.fixture(scope="module")
def L():
    L = pkg.lib()
    assert pkg.device_count() >= 1
    pkg.check(L.b200v_set_device(0))
    pkg.check(L.b200v_require_sm100(0))
    return L


def esz(dtype):
    return 4 if dtype == "f32" else 2


def run_op(L, op, dtype, x, y, s=3.0, offset_elems=0):
    """Upload bit patterns, run one kernel through the ABI, return the output bit patterns.
    offset_elems shifts the arrays inside their allocations (alignment cases): one int for all
    three, or a (dst, x, y) tuple."""
    n = len(x)
    e = esz(dtype)
    od, ox, oy = (offset_elems,) * 3 if isinstance(offset_elems, int) else offset_elems
    dx = pkg.DeviceBuffer(n * e + ox * e + 64).upload(x, ox * e)
    dy = pkg.DeviceBuffer(n * e + oy * e + 64)
    if y is not None:
        dy.upload(y, oy * e)
    pad = od * e
    dd = pkg.DeviceBuffer(n * e + pad + 64)
    pkg.check(L.b200v_memset(C.c_void_p(dd.ptr), 0xA5, dd.nbytes, None))
    pkg.stream_op(op, dtype, dd.at(pad), dx.at(ox * e), dy.at(oy * e) if y is not None else None, s, n)
    pkg.check(L.b200v_device_sync())
    out = dd.download(pkg.NP_BITS[dtype], n, pad)
    guard = dd.download(np.uint8, 64, pad + n * e)
    assert (guard == 0xA5).all(), "kernel wrote past the end of its output"
    if pad:
        assert (dd.download(np.uint8, pad) == 0xA5).all(), "kernel wrote before the start of its output"
    for b in (dx, dy, dd):
        b.free()
    return out


# ------------------------------------------------------------------------------------------------
def test_discovery(L):
    info = pkg.device_info(0)
    assert info.cc_major == 10
    assert info.sm_count >= 100
    assert info.hbm_bytes > 100 << 30
    assert info.l2_bytes > 32 << 20
    assert info.name and info.uuid.startswith(b"GPU-") and info.pci_bus_id
    assert info.runtime_version >= 12080 and info.driver_version >= info.runtime_version - 1000
    n = pkg.device_count()
    m = (C.c_int * (n * n))()
    pkg.check(L.b200v_peer_matrix(n, m))
    assert all(m[i * n + i] == 1 for i in range(n))


@pytest.mark.parametrize("dtype", ["f32", "bf16"])
@pytest.mark.parametrize("mode", [O.GEN_UNIT, O.GEN_BITS, O.GEN_DENORM])
def test_fill_matches_oracle(L, dtype, mode):
    e = esz(dtype)
    # (start, n, element offset of the destination): both vector paths of the kernel — first index of every vector a
    # multiple of the vector length (start + head aligned: the hoisted-xor path) or not — across the 2^32 index boundary,
    # with ragged heads and tails
    for start, n, off in ((0, 1, 0), (0, 50_000, 0), (5, 100_003, 0), ((1 << 32) - 777, 4099, 0),
                          ((1 << 34) + 1, 1 << 20, 0), (9, 1237, 1), (9, 1237, 3), (0, 7, 5),
                          ((1 << 32) - 4096, 8192 + 3, 0), ((1 << 32) - 8, 64, 0), ((1 << 33) + 16, 100_001, 0),
                          (7, 40_001, 1), (3, 40_001, 5), ((1 << 32) - 1029, 4099, 5), (8, (1 << 20) + 9, 0),
                          (1, 40_001, 1), ((1 << 32) - 1027, 4099, 5), ((1 << 32) - 1031, 4099, 1)):
        buf = pkg.DeviceBuffer(n * e + off * e + 64)
        pkg.check(L.b200v_memset(C.c_void_p(buf.ptr), 0x5A, buf.nbytes, None))
        pkg.check(L.b200v_fill(pkg.DTYPES[dtype], buf.at(off * e), start, n, SEED, 3, mode, None))
        got = buf.download(pkg.NP_BITS[dtype], n, off * e)
        np.testing.assert_array_equal(got, O.fill(dtype, start, n, SEED, 3, mode))
        assert (buf.download(np.uint8, 64 - 0, off * e + n * e)[:32] == 0x5A).all()
        buf.free()


SIZES = [1, 2, 7, 31, 50_000, 8191, 8192, 8193, 16384 + 5, (1 << 20) + 3, 3 * (1 << 20) + 12345]


@pytest.mark.parametrize("dtype", ["f32", "bf16"])
@pytest.mark.parametrize("op", ["add", "copy", "scale", "triad"])
def test_ops_bit_exact(L, op, dtype):
    for mode in (O.GEN_UNIT, O.GEN_BITS, O.GEN_DENORM):
        for n in SIZES:
            x = O.fill(dtype, 11, n, SEED, 0, mode)
            y = O.fill(dtype, 11, n, SEED, 1, mode)
            for s in (3.0, -0.333251953125):
                want = O.op(op, dtype, n, s=s, x=x, y=y)
                got = run_op(L, op, dtype, x, y if op in ("add", "triad") else None, s)
                np.testing.assert_array_equal(got, want, err_msg=f"{op} {dtype} mode={mode} n={n} s={s}")
                if op in ("copy", "add"):
                    break


@pytest.mark.parametrize("dtype", ["f32", "bf16"])
@pytest.mark.parametrize("offset", [1, 2, 4, 8, 12])
def test_ops_on_unaligned_views(L, dtype, offset):
    """Element offsets walk through the scalar (not 16-B aligned), 128-bit and 256-bit paths."""
    n = 70_001
    x = O.fill(dtype, 0, n, SEED, 0, O.GEN_BITS)
    y = O.fill(dtype, 0, n, SEED, 1, O.GEN_BITS)
    for op in ("add", "copy", "scale", "triad"):
        want = O.op(op, dtype, n, s=3.0, x=x, y=y)
        got = run_op(L, op, dtype, x, y if op in ("add", "triad") else None, 3.0, offset_elems=offset)
        np.testing.assert_array_equal(got, want, err_msg=f"{op} {dtype} offset={offset}")


def test_random_shapes_and_alignments(L):
    """Seeded fuzz: random op, dtype, generator mode, scalar, length and an independent element
    offset for each of the three arrays (so the 256-bit, 128-bit and element-wise dispatch paths are
    all hit with mixed alignments)."""
    rng = np.random.default_rng(20261106)
    ops = ["add", "copy", "scale", "triad"]
    scalars = [3.0, -0.75, 1.0009765625, 0.333251953125, -2.5e-20, 6.0e18]
    for case in range(120):
        op = ops[rng.integers(4)]
        dtype = "f32" if rng.integers(2) else "bf16"
        mode = int(rng.integers(3))
        n = int(rng.integers(1, 400_000)) if case % 4 else int(rng.integers(1, 70))
        offs = tuple(int(v) for v in rng.integers(0, 17, size=3))
        s = scalars[rng.integers(len(scalars))]
        x = O.fill(dtype, int(rng.integers(0, 1 << 40)), n, SEED + case, 0, mode)
        y = O.fill(dtype, int(rng.integers(0, 1 << 40)), n, SEED + case, 1, mode)
        want = O.op(op, dtype, n, s=s, x=x, y=y)
        got = run_op(L, op, dtype, x, y if op in ("add", "triad") else None, s, offset_elems=offs)
        np.testing.assert_array_equal(got, want, err_msg=f"case {case}: {op} {dtype} mode={mode} n={n} offs={offs} s={s}")


def test_golden_vectors_through_the_gpu(L):
    """The exact-arithmetic fixtures (tests/golden) straight through the kernels."""
    with open(os.path.join(GOLD, "kat_f32.json")) as f:
        k = json.load(f)
    a = np.array(k["add"], dtype=np.uint64)
    got = run_op(L, "add", "f32", a[:, 0].astype(np.uint32), a[:, 1].astype(np.uint32))
    np.testing.assert_array_equal(got, a[:, 2].astype(np.uint32))
    t = np.array(k["triad"], dtype=np.uint64)
    for sv in np.unique(t[:, 0]):
        m = t[:, 0] == sv
        got = run_op(L, "triad", "f32", t[m, 1].astype(np.uint32), t[m, 2].astype(np.uint32), O.f32_from_bits(int(sv)))
        np.testing.assert_array_equal(got, t[m, 3].astype(np.uint32))      # fused result ...
    assert (t[:, 3] != t[:, 4]).sum() >= 250                                  # ... which mul+add would miss
    sc = np.array(k["scale"], dtype=np.uint64)
    for sv in np.unique(sc[:, 0]):
        m = sc[:, 0] == sv
        got = run_op(L, "scale", "f32", sc[m, 1].astype(np.uint32), None, O.f32_from_bits(int(sv)))
        np.testing.assert_array_equal(got, sc[m, 2].astype(np.uint32))
    with open(os.path.join(GOLD, "kat_bf16.json")) as f:
        kb = json.load(f)
    a = np.array(kb["add"], dtype=np.uint64)
    got = run_op(L, "add", "bf16", a[:, 0].astype(np.uint16), a[:, 1].astype(np.uint16))
    np.testing.assert_array_equal(got, a[:, 2].astype(np.uint16))
    t = np.array(kb["triad"], dtype=np.uint64)
    for sv in np.unique(t[:, 0]):
        m = t[:, 0] == sv
        got = run_op(L, "triad", "bf16", t[m, 1].astype(np.uint16), t[m, 2].astype(np.uint16), O.f32_from_bits(int(sv)))
        np.testing.assert_array_equal(got, t[m, 3].astype(np.uint16))
    sc = np.array(kb["scale"], dtype=np.uint64)
    for sv in np.unique(sc[:, 0]):
        m = sc[:, 0] == sv
        got = run_op(L, "scale", "bf16", sc[m, 1].astype(np.uint16), None, O.f32_from_bits(int(sv)))
        np.testing.assert_array_equal(got, sc[m, 2].astype(np.uint16))
    # bf16 RNE alone: scale by 1.0 cannot be used (inputs are already bf16), so round through
    # add with +0: widen(hi) + widen(lo-part) is not expressible either -> covered by add/triad above.


def test_infinities_and_bf16_nan(L):
    """Edge values the generator never emits.  fp32: +-Inf operands wherever the result is not a NaN
    (NaN *payloads* are the one thing IEEE leaves open, and x86 and the GPU pick different ones, which
    is why the workload keeps NaN-producing inputs out).  bf16: NaN results too, because both sides
    canonicalise every NaN to 0x7FFF when rounding to bf16."""
    inf, ninf, big, one, zero, nzero, den = 0x7F800000, 0xFF800000, 0x7F7FFFFF, 0x3F800000, 0, 0x80000000, 0x00000001
    xs = np.array([inf, ninf, inf, big, big, one, den, zero, nzero, nzero, inf, ninf], dtype=np.uint32)
    ys = np.array([one, big, inf, big, one, ninf, den, nzero, nzero, zero, big, ninf], dtype=np.uint32)
    for op, s in (("add", 0.0), ("triad", 3.0), ("triad", -2.0), ("scale", 3.0), ("scale", -0.0), ("copy", 0.0)):
        if op == "triad" and s < 0:
            m = np.array([not (x in (inf, ninf) and y in (inf, ninf)) for x, y in zip(xs, ys)])  # drop inf-inf
        elif op == "scale" and s == 0:
            m = np.array([x not in (inf, ninf) for x in xs])                                      # drop 0*inf
        else:
            m = np.ones(len(xs), dtype=bool)
        x, y = np.ascontiguousarray(xs[m]), np.ascontiguousarray(ys[m])
        want = O.op(op, "f32", len(x), s=s, x=x, y=y)
        assert not (((want & 0x7FFFFFFF) > 0x7F800000).any()), "test vector produces a NaN; fix the mask"
        got = run_op(L, op, "f32", x, y if op in ("add", "triad") else None, s)
        np.testing.assert_array_equal(got, want, err_msg=f"{op} s={s}")
    hx = np.array([0x7F80, 0xFF80, 0x7F80, 0x7FC0, 0x7F7F, 0x0001, 0x8000, 0xFFFF, 0x7F80, 0x3F80], dtype=np.uint16)
    hy = np.array([0xFF80, 0x7F80, 0x7F80, 0x3F80, 0x7F7F, 0x8001, 0x8000, 0x0000, 0x0000, 0x7FC1], dtype=np.uint16)
    for op, s in (("add", 0.0), ("triad", 3.0), ("triad", 0.0), ("scale", 0.0), ("scale", -1.5)):
        want = O.op(op, "bf16", len(hx), s=s, x=hx, y=hy)
        got = run_op(L, op, "bf16", hx, hy if op in ("add", "triad") else None, s)
        np.testing.assert_array_equal(got, want, err_msg=f"bf16 {op} s={s}")
    assert (O.op("add", "bf16", len(hx), x=hx, y=hy)[:2] == 0x7FFF).all()          # inf-inf -> canonical NaN


def test_argument_errors_on_device(L):
    buf = pkg.DeviceBuffer(1 << 20)
    p = buf.ptr
    vp = C.c_void_p
    assert L.b200v_add_f32(vp(p), vp(p), vp(p + 4096), 256, None) == pkg.ERR_INVALID          # dst == a
    assert b"overlap" in L.b200v_last_error()
    assert L.b200v_triad_f32(vp(p + 512), 3.0, vp(p), vp(p + 8192), 256, None) == pkg.ERR_INVALID  # partial overlap
    assert L.b200v_add_f32(vp(p + 2), vp(p + 4096), vp(p + 8192), 16, None) == pkg.ERR_ALIGN
    assert L.b200v_add_bf16(vp(p + 1), vp(p + 4096), vp(p + 8192), 16, None) == pkg.ERR_ALIGN
    assert L.b200v_add_f32(vp(p), vp(p + 4096), None, 16, None) == pkg.ERR_INVALID
    assert L.b200v_copy_f32(vp(p), vp(p + 4096), 0, None) == pkg.OK                              # empty is fine
    ws = pkg.DeviceBuffer(64)
    assert L.b200v_checksum(vp(p), 1024, 0, vp(ws.ptr), 64, vp(p + 4096), None) == pkg.ERR_INVALID
    assert b"workspace" in L.b200v_last_error()
    pkg.check(L.b200v_device_sync())
    buf.free()
    ws.free()


@pytest.mark.parametrize("nbytes,off", [(0, 0), (1, 0), (3, 0), (4, 0), (17, 4), (4096, 0), (4099, 8), (1_000_003, 12),
                                        ((64 << 20) + 6, 4)])
def test_checksum_matches_oracle(L, nbytes, off):
    rng = np.random.default_rng(nbytes + off)
    host = rng.integers(0, 256, size=nbytes, dtype=np.uint8)
    buf = pkg.DeviceBuffer(nbytes + off + 16).upload(host, off) if nbytes else pkg.DeviceBuffer(64)
    for word_offset in (0, (1 << 32) - 3, 123456789012):
        assert pkg.checksum(buf.at(off), nbytes, word_offset) == O.checksum(host, word_offset)
    buf.free()


def test_checksum_workspace_reuse_and_shards(L):
    n = 3_000_001
    x = O.fill("f32", 0, n, SEED, 4, O.GEN_BITS)
    buf = pkg.DeviceBuffer(n * 4).upload(x)
    ws = pkg.DeviceBuffer(L.b200v_checksum_workspace_bytes())
    pkg.check(L.b200v_memset(C.c_void_p(ws.ptr), 0xFF, ws.nbytes, None))      # dirty scratch must not matter
    whole = pkg.checksum(buf.at(0), n * 4, 0, ws)
    assert whole == O.checksum(x)
    assert pkg.checksum(buf.at(0), n * 4, 0, ws) == whole                       # second use of the same scratch
    total = [0, 0]
    for r in range(3):
        b, c = pkg.shard_range(n, r, 3)
        s = pkg.checksum(buf.at(4 * b), 4 * c, b, ws)
        total = [(total[0] + s[0]) & M64, (total[1] + s[1]) & M64]
    assert tuple(total) == whole
    buf.free()
    ws.free()


@pytest.mark.parametrize("pinned", [True, False])
def test_vector_add_host_buffers(L, pinned):
    """The cuda-samples call shape: host arrays in, host array out, H2D/D2H inside the call."""
    for n in (50_000, 1, (8 << 20), (20 << 20) + 77):
        a = O.fill("f32", 0, n, SEED, 0, O.GEN_UNIT)
        b = O.fill("f32", 0, n, SEED, 1, O.GEN_UNIT)
        want = O.op("add", "f32", n, x=a, y=b)
        if pinned:
            bufs = [pkg.PinnedBuffer(n * 4) for _ in range(3)]
            ha, hb, hc = (x.view(np.uint32) for x in bufs)
            ha[:] = a
            hb[:] = b
        else:
            ha, hb, hc = a.copy(), b.copy(), np.zeros(n, dtype=np.uint32)
        hc[:] = 0
        vp = lambda arr: arr.ctypes.data_as(C.c_void_p)
        pkg.check(L.b200v_vector_add_host(vp(ha), vp(hb), vp(hc), n))
        np.testing.assert_array_equal(np.asarray(hc), want)
        pkg.check(L.b200v_vector_add_host(vp(ha), vp(hb), vp(hc), n))          # staging buffers are reused
        np.testing.assert_array_equal(np.asarray(hc), want)
        del ha, hb, hc
        if pinned:
            for x in bufs:
                x.free()
    pkg.check(L.b200v_vector_add_host_release())


def test_vector_add_host_with_registered_arrays_and_pipeline_shapes(L):
    """Host arrays the caller already owns, pinned in place with b200v_host_register; and every chunking the pipeline
    can be given (one chunk, a tail that is ramped down, chunk sizes that do not divide n) gives the same bits."""
    n = (5 << 20) + 12345
    a = O.fill("f32", 3, n, SEED, 0, O.GEN_BITS)
    b = O.fill("f32", 3, n, SEED, 1, O.GEN_BITS)
    c = np.zeros(n, dtype=np.uint32)
    want = O.op("add", "f32", n, x=a, y=b)
    vp = lambda arr: arr.ctypes.data_as(C.c_void_p)
    for arr in (a, b, c):
        pkg.check(L.b200v_host_register(vp(arr), arr.nbytes))
    try:
        for chunk_elems, slots in ((8 << 20, 3), (1 << 20, 2), (1 << 18, 4), ((1 << 20) + 24, 3), (1024, 8), (1 << 23, 1)):
            pkg.check(L.b200v_vector_add_host_config(chunk_elems, slots))
            c[:] = 0
            pkg.check(L.b200v_vector_add_host(vp(a), vp(b), vp(c), n))
            np.testing.assert_array_equal(c, want, err_msg=f"chunk {chunk_elems} slots {slots}")
        # arguments are checked; the call is synchronous also when it fails
        assert L.b200v_vector_add_host(None, vp(b), vp(c), n) == pkg.ERR_INVALID
        assert L.b200v_vector_add_host_config(8, 3) == pkg.ERR_INVALID and L.b200v_vector_add_host_config(1 << 20, 0) == pkg.ERR_INVALID
        assert L.b200v_host_register(None, 16) == pkg.ERR_INVALID
    finally:
        pkg.check(L.b200v_vector_add_host_config(8 << 20, 3))      # back to the defaults
        pkg.check(L.b200v_vector_add_host_release())
        for arr in (a, b, c):
            pkg.check(L.b200v_host_unregister(vp(arr)))


# ------------------------------------------------------------------------------------------------
# BASELINE.json sizes: verified through the checksum property (shards of the host oracle's output
# add up to the device's checksum) plus bit-exact windows brought back to the host.

def oracle_checksum_chunked(op, dtype, n, start, s, mode=O.GEN_UNIT, chunk=1 << 24):
    tot = [0, 0]
    for off in range(0, n, chunk):
        m = min(chunk, n - off)
        x = O.fill(dtype, start + off, m, SEED, 0, mode)
        y = O.fill(dtype, start + off, m, SEED, 1, mode)
        out = O.op(op, dtype, m, s=s, x=x, y=y, threads=0)
        per_word = 1 if dtype == "f32" else 2
        assert off % per_word == 0
        c = O.checksum(out, off // per_word, threads=0)
        tot = [(tot[0] + c[0]) & M64, (tot[1] + c[1]) & M64]
    return tuple(tot)


@pytest.mark.parametrize("op,dtype,log2n", [("add", "f32", 28), ("triad", "f32", 28), ("triad", "bf16", 28),
                                            ("copy", "f32", 28), ("scale", "bf16", 27)])
def test_baseline_size_checksum_and_windows(L, op, dtype, log2n):
    n = (1 << log2n)
    e = esz(dtype)
    start = (1 << 32) - (1 << 20)            # global indices cross the 2^32 boundary
    dx, dy, dd = (pkg.DeviceBuffer(n * e) for _ in range(3))
    pkg.check(L.b200v_fill(pkg.DTYPES[dtype], C.c_void_p(dx.ptr), start, n, SEED, 0, O.GEN_UNIT, None))
    pkg.check(L.b200v_fill(pkg.DTYPES[dtype], C.c_void_p(dy.ptr), start, n, SEED, 1, O.GEN_UNIT, None))
    pkg.stream_op(op, dtype, C.c_void_p(dd.ptr), C.c_void_p(dx.ptr), C.c_void_p(dy.ptr), 3.0, n)
    got = pkg.checksum(C.c_void_p(dd.ptr), n * e, 0)
    assert got == oracle_checksum_chunked(op, dtype, n, start, 3.0)
    for woff in (0, n // 3 + 1, n - 65_536):
        m = 65_536
        x = O.fill(dtype, start + woff, m, SEED, 0, O.GEN_UNIT)
        y = O.fill(dtype, start + woff, m, SEED, 1, O.GEN_UNIT)
        np.testing.assert_array_equal(dd.download(pkg.NP_BITS[dtype], m, woff * e), O.op(op, dtype, m, s=3.0, x=x, y=y))
    for b in (dx, dy, dd):
        b.free()


def test_full_2p32_triad_one_gpu(L):
    """BASELINE.json configs[3] at its full size on one GPU: triad over 2^32 fp32 elements (3 x 16 GiB),
    whole-output checksum against the host oracle computed in chunks on all host cores, plus windows
    on both sides of element 2^31 and at the very end (64-bit indexing)."""
    n = 1 << 32
    info = pkg.device_info(0)
    if info.hbm_free_bytes < 50 << 30:
        pytest.skip("needs 48 GiB of free HBM")
    vp = C.c_void_p
    dx, dy, dd = (pkg.DeviceBuffer(4 * n) for _ in range(3))
    pkg.check(L.b200v_fill(pkg.F32, vp(dx.ptr), 0, n, SEED, 0, O.GEN_UNIT, None))
    pkg.check(L.b200v_fill(pkg.F32, vp(dy.ptr), 0, n, SEED, 1, O.GEN_UNIT, None))
    pkg.check(L.b200v_triad_f32(vp(dd.ptr), 3.0, vp(dx.ptr), vp(dy.ptr), n, None))
    got = [0, 0]
    piece = 1 << 30                      # device checksum in 4 GiB pieces (the no-workspace entry's limit), summed
    ws = pkg.DeviceBuffer(L.b200v_checksum_workspace_bytes())
    for off in range(0, n, piece):
        c = pkg.checksum(dd.at(4 * off), 4 * piece, off, ws)
        got = [(got[0] + c[0]) & M64, (got[1] + c[1]) & M64]
    for woff in (0, (1 << 31) - 32_768, n - 65_536):
        m = 65_536
        x = O.fill("f32", woff, m, SEED, 0)
        y = O.fill("f32", woff, m, SEED, 1)
        np.testing.assert_array_equal(dd.download(np.uint32, m, 4 * woff), O.op("triad", "f32", m, s=3.0, x=x, y=y))
    for b in (dx, dy, dd, ws):
        b.free()
    assert tuple(got) == oracle_checksum_chunked("triad", "f32", n, 0, 3.0, chunk=1 << 26)


def test_sharding_and_repetition_do_not_change_the_result(L):
    """Size-independent properties at a BASELINE size (2^28 fp32 triad): running the vector as 3, 5 or 8
    contiguous shards (ragged, so shard starts land on every alignment path) gives the same bits as one
    launch (checksum of checksums), and running it again gives the same bits again (idempotence)."""
    n = 1 << 28
    vp = C.c_void_p
    dx, dy, whole, parts = (pkg.DeviceBuffer(4 * n) for _ in range(4))
    pkg.check(L.b200v_fill(pkg.F32, vp(dx.ptr), 0, n, SEED, 0, O.GEN_BITS, None))
    pkg.check(L.b200v_fill(pkg.F32, vp(dy.ptr), 0, n, SEED, 1, O.GEN_BITS, None))
    pkg.check(L.b200v_triad_f32(vp(whole.ptr), -0.75, vp(dx.ptr), vp(dy.ptr), n, None))
    ref = pkg.checksum(vp(whole.ptr), 4 * n)
    for world in (3, 5, 8):
        pkg.check(L.b200v_memset(vp(parts.ptr), 0, 4 * n, None))
        total = [0, 0]
        for r in range(world):
            b, c = pkg.shard_range(n - 3, r, world)          # n-3: ragged shard sizes
            pkg.check(L.b200v_triad_f32(parts.at(4 * b), -0.75, dx.at(4 * b), dy.at(4 * b), c, None))
            cs = pkg.checksum(parts.at(4 * b), 4 * c, b)
            total = [(total[0] + cs[0]) & M64, (total[1] + cs[1]) & M64]
        b = n - 3                                             # the 3 elements left over
        pkg.check(L.b200v_triad_f32(parts.at(4 * b), -0.75, dx.at(4 * b), dy.at(4 * b), 3, None))
        cs = pkg.checksum(parts.at(4 * b), 12, b)
        total = [(total[0] + cs[0]) & M64, (total[1] + cs[1]) & M64]
        assert tuple(total) == ref, world
        assert pkg.checksum(vp(parts.ptr), 4 * n) == ref
    pkg.check(L.b200v_triad_f32(vp(whole.ptr), -0.75, vp(dx.ptr), vp(dy.ptr), n, None))
    assert pkg.checksum(vp(whole.ptr), 4 * n) == ref
    for b in (dx, dy, whole, parts):
        b.free()


def test_time_op_reports_device_times(L):
    n = 1 << 26
    dx, dy, dd = (pkg.DeviceBuffer(n * 4) for _ in range(3))
    pkg.check(L.b200v_fill(pkg.F32, C.c_void_p(dx.ptr), 0, n, SEED, 0, O.GEN_UNIT, None))
    pkg.check(L.b200v_fill(pkg.F32, C.c_void_p(dy.ptr), 0, n, SEED, 1, O.GEN_UNIT, None))
    t = pkg.Timing()
    t.struct_size = C.sizeof(pkg.Timing)
    pkg.check(L.b200v_time_op(pkg.OP_TRIAD, pkg.F32, C.c_void_p(dd.ptr), C.c_void_p(dx.ptr), C.c_void_p(dy.ptr),
                              3.0, n, 3, 10, None, C.byref(t)))
    assert t.iters == 10 and t.bytes_per_launch == 12 * n
    assert 0 < t.best_ms <= t.median_ms
    gbs = t.bytes_per_launch / t.best_ms * 1e-6
    assert 1000 < gbs < 12000, gbs           # an HBM3e part, and not a no-op
    for b in (dx, dy, dd):
        b.free()


# ------------------------------------------------------------------------------------------------
# Size-independent identities of the four operations, at BASELINE.json's full size (2^28 elements).  None of them
# needs the oracle at that size: the two sides of each identity are computed by DIFFERENT kernels (or the generator's
# own checksum, which the host loop supplies) and must agree bit for bit over the whole vector.

@pytest.mark.parametrize("dtype", ["f32", "bf16"])
def test_algebraic_identities_at_full_size(L, dtype):
    n = 1 << 28
    e = esz(dtype)
    vp = C.c_void_p
    start = (1 << 33) + 12                                      # global indices beyond 2^32
    dx, dy, d1, d2 = (pkg.DeviceBuffer(n * e) for _ in range(4))
    ws = pkg.DeviceBuffer(L.b200v_checksum_workspace_bytes())

    def ck(buf):
        return pkg.checksum(vp(buf.ptr), n * e, 0, ws)

    def op(name, dst, x, y, s):
        pkg.stream_op(name, dtype, vp(dst.ptr), vp(x.ptr), vp(y.ptr) if y is not None else None, s, n)

    for mode in (O.GEN_BITS, O.GEN_UNIT):
        pkg.check(L.b200v_fill(pkg.DTYPES[dtype], vp(dx.ptr), start, n, SEED, 0, mode, None))
        pkg.check(L.b200v_fill(pkg.DTYPES[dtype], vp(dy.ptr), start, n, SEED, 1, mode, None))
        cx, cy = ck(dx), ck(dy)
        # the generator on the device == the generator on the host, over the whole vector
        words = 4 // e
        want = [0, 0]
        for off in range(0, n, 1 << 24):
            c = O.checksum(O.fill(dtype, start + off, 1 << 24, SEED, 0, mode), off // words, threads=0)
            want = [(want[0] + c[0]) & M64, (want[1] + c[1]) & M64]
        assert cx == tuple(want)
        # copy is the identity, and idempotent
        op("copy", d1, dx, None, 0.0)
        assert ck(d1) == cx
        op("copy", d2, d1, None, 0.0)
        assert ck(d2) == cx
        # scale by 1 is the identity (every finite value, denormals and signed zeros included)
        op("scale", d1, dx, None, 1.0)
        assert ck(d1) == cx
        # addition commutes
        op("add", d1, dx, dy, 0.0)
        op("add", d2, dy, dx, 0.0)
        c_add = ck(d1)
        assert ck(d2) == c_add
        # triad with s = 1 is the addition (one rounding either way)
        op("triad", d2, dx, dy, 1.0)
        assert ck(d2) == c_add
        # doubling: 2 * x == x + x (exact, or the same overflow to infinity)
        op("scale", d1, dx, None, 2.0)
        op("add", d2, dx, dx, 0.0)
        assert ck(d1) == ck(d2)
        if mode == O.GEN_UNIT:
            # triad with s = 0 returns the addend: fma(0, x, y) == y for x, y >= 0
            op("triad", d1, dx, dy, 0.0)
            assert ck(d1) == cy
    for b in (dx, dy, d1, d2, ws):
        b.free()
