#!/usr/bin/env python3
"""What bounds the end-to-end (host-buffer) vectorAdd at N GPUs of one box?

    python tools/e2e_scaling.py out.json                                  # N = 1
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P tools/e2e_scaling.py out.json                     # N ranks, one per GPU

Every rank drives its own GPU; all ranks start each measurement together (gloo barrier on the host, so
no kernel is parked on any GPU) and the job's time is the max over ranks of the wall time.  Measured:

  raw      pinned H2D alone, D2H alone, both directions at once, and the step's own mix (2 GiB in while
           1 GiB goes out) — per GPU and summed over the N GPUs: the PCIe / root-complex ceiling;
  host     what the host's DRAM gives the same cores: every rank copying 1 GiB with its share of the
           host threads (the oracle's OpenMP copy loop) at the same time: the host-DRAM ceiling;
  pipeline b200v_vector_add_host over chunk size x slots, N ranks at once: the e2e number itself;
  bound    per N: the step's bytes through the lower of the two ceilings.

Developer tool (evidence under profiles/r02_e2e_scaling*.json); not part of the product.  The oracle's copy
loop is used here only as a host memory-bandwidth probe.
"""
import ctypes as C
import importlib
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

world_env = int(os.environ.get("WORLD_SIZE", "1"))
cores = len(os.sched_getaffinity(0))
threads = max(1, cores // world_env)
os.environ["OMP_NUM_THREADS"] = str(threads)

import numpy as np  # noqa: E402

pkg = importlib.import_module("kubernetes-with-nvidia-gpu_b200")
ranks_mod = importlib.import_module("kubernetes-with-nvidia-gpu_b200.ranks")
L = pkg.lib()
R = ranks_mod.Ranks(backend="gloo" if world_env > 1 else None)
rank, world = R.rank, R.world
pkg.check(L.b200v_set_device(R.local_rank))
vp = C.c_void_p
N = 1 << 28
GIB = 4 * N
quick = "--quick" in sys.argv


def sync():
    pkg.check(L.b200v_device_sync())


def job_seconds(fn, reps=3):
    """All ranks run fn at the same time; returns the best-of-reps of (max over ranks of the wall time)."""
    fn()
    sync()
    best = 1e9
    for _ in range(reps):
        R.barrier()
        t0 = time.perf_counter()
        fn()
        sync()
        dt = R.max(time.perf_counter() - t0)
        best = min(best, dt)
    return best


hA, hB, hC = (pkg.PinnedBuffer(GIB) for _ in range(3))
hA.view(np.uint32)[:] = 0x3F000000
hB.view(np.uint32)[:] = 0x3E800000
hC.view(np.uint32)[:] = 0
dX, dY, dZ = (pkg.DeviceBuffer(GIB) for _ in range(3))
s1, s2 = vp(None), vp(None)
pkg.check(L.b200v_stream_create(C.byref(s1)))
pkg.check(L.b200v_stream_create(C.byref(s2)))

out = {"n_gpus": world, "host_cores": cores, "host_threads_per_rank": threads, "bytes_per_array": GIB}

# ---- topology, as far as the box shows it ---------------------------------------------------------
if rank == 0:
    topo = {}
    try:
        nodes = sorted(d for d in os.listdir("/sys/devices/system/node") if d.startswith("node"))
        topo["numa_nodes"] = nodes
    except Exception as e:
        topo["numa_nodes"] = repr(e)
    gpus = []
    for d in range(pkg.device_count()):
        info = pkg.device_info(d)
        bus = info.pci_bus_id.decode().lower()
        entry = {"gpu": d, "pci": bus}
        for key in ("numa_node", "current_link_speed", "current_link_width", "max_link_speed", "max_link_width"):
            try:
                entry[key] = open(f"/sys/bus/pci/devices/{bus[-12:]}/{key}").read().strip()
            except Exception:
                entry[key] = None
        gpus.append(entry)
    topo["gpus"] = gpus
    try:
        topo["lscpu"] = [ln for ln in subprocess.run(["lscpu"], capture_output=True, text=True).stdout.splitlines()
                         if any(k in ln for k in ("Model name", "Socket", "NUMA", "CPU(s):", "Thread", "Hypervisor", "L3"))]
    except Exception:
        pass
    try:
        mem = {ln.split(":")[0]: ln.split(":")[1].strip() for ln in open("/proc/meminfo") if ln.startswith(("MemTotal", "MemAvailable", "Hugepagesize"))}
        topo["meminfo"] = mem
    except Exception:
        pass
    out["topology"] = topo

# ---- raw copies --------------------------------------------------------------------------------------
def h2d():
    pkg.check(L.b200v_memcpy_h2d(vp(dX.ptr), vp(hA.ptr), GIB, s1))


def d2h():
    pkg.check(L.b200v_memcpy_d2h(vp(hC.ptr), vp(dZ.ptr), GIB, s2))


def both():
    h2d()
    d2h()


def step_mix():          # the step's own traffic: 2 GiB in (two arrays) while 1 GiB goes out
    pkg.check(L.b200v_memcpy_h2d(vp(dX.ptr), vp(hA.ptr), GIB, s1))
    pkg.check(L.b200v_memcpy_h2d(vp(dY.ptr), vp(hB.ptr), GIB, s1))
    pkg.check(L.b200v_memcpy_d2h(vp(hC.ptr), vp(dZ.ptr), GIB, s2))


raw = {}
t = job_seconds(h2d)
raw["h2d_gbs_per_gpu"], raw["h2d_gbs_sum"] = GIB / t * 1e-9, world * GIB / t * 1e-9
t = job_seconds(d2h)
raw["d2h_gbs_per_gpu"], raw["d2h_gbs_sum"] = GIB / t * 1e-9, world * GIB / t * 1e-9
t = job_seconds(both)
raw["bidir_each_gbs_per_gpu"], raw["bidir_total_gbs_sum"] = GIB / t * 1e-9, 2 * world * GIB / t * 1e-9
t = job_seconds(step_mix)
raw["step_mix_ms"] = t * 1e3
raw["step_mix_e2e_bound_gbs_sum"] = world * 12 * N / t * 1e-9      # 12 B/element through the copy engines alone
out["raw"] = raw
if rank == 0:
    print("raw", json.dumps(raw), flush=True)

# ---- host DRAM: every rank copies 1 GiB with its share of the host threads, all at once ----------------
import oracle_lib as O  # noqa: E402  (memory-bandwidth probe only)
OL = O.lib()
src = np.full(N, 0x3F000000, dtype=np.uint32)
dst = np.zeros(N, dtype=np.uint32)


def host_copy():
    OL.oracle_copy_f32(dst.ctypes.data_as(vp), src.ctypes.data_as(vp), N, threads)


t = job_seconds(host_copy)
host = {"copy_gbs_sum_algorithmic": world * 2 * GIB / t * 1e-9,          # read + write
        "copy_gbs_sum_with_write_allocate": world * 3 * GIB / t * 1e-9,  # + the read-for-ownership of the destination
        "threads_per_rank": threads, "ranks": world}


def host_add():
    OL.oracle_add_f32(dst.ctypes.data_as(vp), src.ctypes.data_as(vp), hA.view(np.uint32).ctypes.data_as(vp), N, threads)


t = job_seconds(host_add)
host["add_gbs_sum_algorithmic"] = world * 3 * GIB / t * 1e-9
host["add_gbs_sum_with_write_allocate"] = world * 4 * GIB / t * 1e-9
out["host"] = host
if rank == 0:
    print("host", json.dumps(host), flush=True)
del src, dst

# ---- the pipeline itself ---------------------------------------------------------------------------------
pipeline = []
chunks = (32,) if quick else (8, 16, 32, 64, 128, 256)
slot_set = (3,) if quick else (2, 3, 4)
for chunk_mib in chunks:
    for slots in slot_set:
        pkg.check(L.b200v_vector_add_host_config((chunk_mib << 20) // 4, slots))
        t = job_seconds(lambda: pkg.check(L.b200v_vector_add_host(vp(hA.ptr), vp(hB.ptr), vp(hC.ptr), N)), reps=3)
        row = {"chunk_mib": chunk_mib, "slots": slots, "ms": t * 1e3, "e2e_gbs_sum": world * 12 * N / t * 1e-9}
        pipeline.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
pkg.check(L.b200v_vector_add_host_release())
out["pipeline"] = pipeline
best = min(pipeline, key=lambda r: r["ms"])
default = next(r for r in pipeline if r["chunk_mib"] == 32 and r["slots"] == 3)
# the step needs 3 GiB of PCIe traffic per GPU and 3 GiB of host-DRAM traffic per GPU (2 read by the H2D DMA,
# 1 written by the D2H DMA); its ceiling is whichever of the two the box runs out of first
pcie_bound = raw["step_mix_e2e_bound_gbs_sum"]
dram_bound = host["copy_gbs_sum_with_write_allocate"]      # bytes/s the host DRAM has been seen to move
out["summary"] = {"n_gpus": world, "e2e_default_gbs": default["e2e_gbs_sum"], "e2e_best_gbs": best["e2e_gbs_sum"],
                  "best_shape": {"chunk_mib": best["chunk_mib"], "slots": best["slots"]},
                  "copy_engine_bound_gbs": pcie_bound, "host_dram_seen_gbs": dram_bound,
                  "e2e_over_copy_engine_bound": best["e2e_gbs_sum"] / pcie_bound}
if rank == 0:
    print("summary", json.dumps(out["summary"]), flush=True)
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    if args:
        json.dump(out, open(args[0], "w"), indent=1)
R.close()
