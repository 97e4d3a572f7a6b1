"""GPU tier, collected last on purpose (see `pytest --collect-only`): b200v_checksum_static.

Some callers of libb200v cannot allocate device memory when they need a checksum — a burn-in that
has filled all free HBM with its own arrays, a probe inside a process whose allocator already owns
the GPU, a hook that must not call cudaMalloc/cudaFree because they synchronise the device.  For
them the library carries the reduction's scratch inside its own image.  This test calls that entry
point through the C ABI and requires exactly the result of the ordinary b200v_checksum (which gets
a caller-allocated workspace) — bit for bit, the tolerance every test here uses — on two calls in
a row, because the second call finds the first call's leftovers in the shared array.
It stands alone (fresh process) and also runs after every other GPU test.
"""
import ctypes as C

import numpy as np
import pytest

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
