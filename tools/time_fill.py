#!/usr/bin/env python3
"""Time b200v_fill (1 GiB, every dtype and mode) with device events.  Developer tool."""
import ctypes as C
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("kubernetes-with-nvidia-gpu_b200")
if os.environ.get("B200V_LIB"):
    pkg.LIB_PATH = os.environ["B200V_LIB"]
L = pkg.lib()
pkg.check(L.b200v_set_device(0))
vp = C.c_void_p
NB = 1 << 30
d = pkg.DeviceBuffer(NB)
e0, e1 = vp(None), vp(None)
pkg.check(L.b200v_event_create(C.byref(e0)))
pkg.check(L.b200v_event_create(C.byref(e1)))
for dtype, esz in (("f32", 4), ("bf16", 2)):
    for mode in (0, 1, 2):
        n = NB // esz
        for start in (0, 12345):      # first index a multiple of the vector length (the usual case), or not
            best = 1e9
            for it in range(8):
                pkg.check(L.b200v_event_record(e0, None))
                pkg.check(L.b200v_fill(pkg.DTYPES[dtype], vp(d.ptr), start, n, 1, 0, mode, None))
                pkg.check(L.b200v_event_record(e1, None))
                pkg.check(L.b200v_event_sync(e1))
                ms = C.c_float(0)
                pkg.check(L.b200v_event_elapsed_ms(e0, e1, C.byref(ms)))
                if it >= 2:
                    best = min(best, ms.value)
            print(f"fill {dtype} mode {mode} start {start:>5}: {best:.4f} ms = {NB / best * 1e-6:.0f} GB/s written")
