"""GPU tier, collected last on purpose (see `pytest --collect-only`): b200v_checksum_static.

Some callers of libb200v cannot allocate device memory when they need a checksum — a burn-in that
has filled all free HBM with its own arrays, a probe inside a process whose allocator already owns
the GPU, a hook that must not call cudaMalloc/cudaFree because they synchronise the device.  For
them the library carries the reduction's scratch inside its own image.  This test calls that entry
point through the C ABI and requires exactly the result of the ordinary b200v_checksum (which gets
a caller-allocated workspace) — bit for bit, the tolerance every test here uses — on two calls in
a row, because the second call finds the first call's leftovers in the shared array.  The small
buffer is also brought back to the host and summed by the CPU oracle, so the two CUDA entry points
cannot agree on a shared mistake.  A second case repeats this on GPU 1 after GPU 0 where the box has
two GPUs: every device holds its own copy of the array, found through its own address look-up.
It stands alone (fresh process) and also runs after every other GPU test.
"""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as O
from b200v_pkg import pkg

pytestmark = pytest.mark.gpu

SEED = 20261106


def test_static_scratch_checksum_matches_workspace_checksum_twice_in_a_row():
    L = pkg.lib()
    pkg.check(L.b200v_set_device(0))
    pkg.check(L.b200v_require_sm100(0))
    vp = C.c_void_p

    n_big, n_small = (96 << 20) + 3, 70_001            # fp32 elements: many CTAs, then few
    big = pkg.DeviceBuffer(4 * n_big)
    small = pkg.DeviceBuffer(4 * n_small + 8)
    pkg.check(L.b200v_fill(pkg.F32, vp(big.ptr), 5, n_big, SEED, 6, pkg.GEN_BITS, None))
    pkg.check(L.b200v_fill(pkg.F32, small.at(4), 9, n_small, SEED, 7, pkg.GEN_BITS, None))

    out = pkg.DeviceBuffer(32)                           # inputs and outputs still come from the caller

    def static_sum(ptr, nbytes, word_offset, slot):
        pkg.check(L.b200v_checksum_static(ptr, nbytes, word_offset, out.at(16 * slot), None))

    # two calls back to back on the same stream, different sizes and word offsets, no sync between
    static_sum(vp(big.ptr), 4 * n_big, (1 << 32) - 11, 0)
    static_sum(small.at(4), 4 * n_small + 3, 77, 1)      # ragged byte count, 4-byte-aligned start
    got = out.download(np.uint64, 4)

    want_big = pkg.checksum(vp(big.ptr), 4 * n_big, (1 << 32) - 11)       # existing entry point, own workspace
    want_small = pkg.checksum(small.at(4), 4 * n_small + 3, 77)
    assert (int(got[0]), int(got[1])) == want_big
    assert (int(got[2]), int(got[3])) == want_small
    # ... and what the CPU oracle says about the same bytes, copied back to the host
    host_bytes = small.download(np.uint8, 4 * n_small + 3, byte_offset=4)
    assert want_small == O.checksum(host_bytes, 77)

    # and again in the other order: the array now holds the small call's partials
    static_sum(small.at(4), 4 * n_small + 3, 77, 0)
    static_sum(vp(big.ptr), 4 * n_big, (1 << 32) - 11, 1)
    got = out.download(np.uint64, 4)
    assert (int(got[0]), int(got[1])) == want_small
    assert (int(got[2]), int(got[3])) == want_big

    # empty input is the empty sum; sizes over the documented limit are refused, not truncated
    static_sum(vp(big.ptr), 0, 0, 0)
    assert tuple(out.download(np.uint64, 2)) == (0, 0)
    rc = L.b200v_checksum_static(vp(big.ptr), (1 << 32) + 4, 0, vp(out.ptr), None)
    assert rc == pkg.ERR_INVALID
    assert b"at most" in L.b200v_last_error()
    assert L.b200v_checksum_static(vp(big.ptr), 16, 0, None, None) == pkg.ERR_INVALID
    pkg.check(L.b200v_device_sync())
    for b in (big, small, out):
        b.free()


@pytest.mark.multigpu
def test_static_scratch_is_per_device():
    """GPU 0 first, then GPU 1, then GPU 0 again: each device resolves and keeps its own array address."""
    if pkg.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    L = pkg.lib()
    vp = C.c_void_p
    n = (3 << 20) + 5
    want = None
    for dev in (0, 1, 0):
        pkg.check(L.b200v_set_device(dev))
        buf = pkg.DeviceBuffer(4 * n)
        out = pkg.DeviceBuffer(16)
        pkg.check(L.b200v_fill(pkg.F32, vp(buf.ptr), 123, n, SEED, 11, pkg.GEN_BITS, None))
        for _ in range(2):                                   # twice in a row on this device
            pkg.check(L.b200v_checksum_static(vp(buf.ptr), 4 * n, 9, vp(out.ptr), None))
            got = tuple(int(x) for x in out.download(np.uint64, 2))
            assert got == pkg.checksum(vp(buf.ptr), 4 * n, 9)
            if want is None:
                want = O.checksum(buf.download(np.uint32, n), 9)
            assert got == want                               # same data on every device: same sum, oracle-checked
        buf.free()
        out.free()
    pkg.check(L.b200v_set_device(0))
