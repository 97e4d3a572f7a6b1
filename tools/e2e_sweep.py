#!/usr/bin/env python3
"""Sweep the host-buffer vectorAdd pipeline (b200v_vector_add_host_config) and print the raw PCIe
copy rates next to it, so the end-to-end number can be read against its own bound.
Developer tool (round-1 tables under profiles/); not part of the product."""
import ctypes as C
import importlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("kubernetes-with-nvidia-gpu_b200")
L = pkg.lib()
pkg.check(L.b200v_set_device(0))
vp = C.c_void_p
N = 1 << 28
nbytes = 4 * N
hA, hB, hC = (pkg.PinnedBuffer(nbytes) for _ in range(3))
import numpy as np
hA.view(np.uint32)[:] = 0x3F000000
hB.view(np.uint32)[:] = 0x3E800000
d = pkg.DeviceBuffer(nbytes)
d2 = pkg.DeviceBuffer(nbytes)
s1, s2 = vp(None), vp(None)
pkg.check(L.b200v_stream_create(C.byref(s1)))
pkg.check(L.b200v_stream_create(C.byref(s2)))
out = {"raw": {}, "pipeline": []}


def wall(fn, reps=3):
    fn()
    best = 1e9
    for _ in range(reps):
        pkg.check(L.b200v_device_sync())
        t0 = time.perf_counter()
        fn()
        pkg.check(L.b200v_device_sync())
        best = min(best, time.perf_counter() - t0)
    return best


h2d = wall(lambda: pkg.check(L.b200v_memcpy_h2d(vp(d.ptr), vp(hA.ptr), nbytes, s1)))
d2h = wall(lambda: pkg.check(L.b200v_memcpy_d2h(vp(hC.ptr), vp(d.ptr), nbytes, s1)))


def both():
    pkg.check(L.b200v_memcpy_h2d(vp(d.ptr), vp(hA.ptr), nbytes, s1))
    pkg.check(L.b200v_memcpy_d2h(vp(hC.ptr), vp(d2.ptr), nbytes, s2))


bi = wall(both)
out["raw"] = {"h2d_gbs": nbytes / h2d * 1e-9, "d2h_gbs": nbytes / d2h * 1e-9,
              "h2d_while_d2h_each_gbs": nbytes / bi * 1e-9, "bidir_total_gbs": 2 * nbytes / bi * 1e-9}
print("raw", json.dumps(out["raw"]), flush=True)
bound_ms = 2 * nbytes / (out["raw"]["h2d_gbs"] * 1e9) * 1e3
print(f"bound: 2 GiB H2D alone = {bound_ms:.2f} ms -> {12 * N / bound_ms * 1e-6:.1f} GB/s algorithmic", flush=True)

for chunk_mib in (4, 8, 16, 32, 64, 128, 256):
    for slots in (2, 3, 4, 6):
        pkg.check(L.b200v_vector_add_host_config((chunk_mib << 20) // 4, slots))
        t = wall(lambda: pkg.check(L.b200v_vector_add_host(vp(hA.ptr), vp(hB.ptr), vp(hC.ptr), N)), reps=3)
        row = {"chunk_mib": chunk_mib, "slots": slots, "ms": t * 1e3, "gbs": 12 * N / t * 1e-9}
        out["pipeline"].append(row)
        print(json.dumps(row), flush=True)
pkg.check(L.b200v_vector_add_host_release())
best = min(out["pipeline"], key=lambda r: r["ms"])
print("best", json.dumps(best))
if len(sys.argv) > 1:
    json.dump(out, open(sys.argv[1], "w"), indent=1)
