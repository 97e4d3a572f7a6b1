#!/usr/bin/env python3
"""Launch every kernel of the library once at N = 2^28 words (after a warm-up launch each), for
`ncu --set full -k regex:'stream_kernel|checksum_kernel|fill_kernel'`.  Developer tool."""
import ctypes as C
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("kubernetes-with-nvidia-gpu_b200")
L = pkg.lib()
pkg.check(L.b200v_set_device(0))
vp = C.c_void_p
NB = 1 << 30                                   # bytes per array
x, y, d = (pkg.DeviceBuffer(NB) for _ in range(3))
ws = pkg.DeviceBuffer(L.b200v_checksum_workspace_bytes())
out = pkg.DeviceBuffer(16)
for rep in range(2):                           # rep 0 = warm-up, rep 1 = the launches to look at
    for dtype, esz in (("f32", 4), ("bf16", 2)):
        n = NB // esz
        pkg.check(L.b200v_fill(pkg.DTYPES[dtype], vp(x.ptr), 0, n, 1, 0, pkg.GEN_UNIT, None))
        pkg.check(L.b200v_fill(pkg.DTYPES[dtype], vp(y.ptr), 0, n, 1, 1, pkg.GEN_UNIT, None))
        for op in ("copy", "scale", "add", "triad"):
            pkg.stream_op(op, dtype, vp(d.ptr), vp(x.ptr), vp(y.ptr), 3.0, n)
    pkg.check(L.b200v_checksum(vp(d.ptr), NB, 0, vp(ws.ptr), ws.nbytes, vp(out.ptr), None))
    pkg.check(L.b200v_device_sync())
print("ok")
