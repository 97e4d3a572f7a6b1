#!/usr/bin/env python3
"""What bounds the end-to-end (host-buffer) vectorAdd at N GPUs of one box?

    python tools/e2e_scaling.py out.json                                  # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node W --master-addr 127.0.0.1 \
        --master-port P tools/e2e_scaling.py out.json                     # W ranks, one per GPU

One launch with W ranks measures every N in {1, 2, 4, 8} <= W: for a given N only ranks 0..N-1 are
active, the others idle at the barriers (gloo, on the host: no kernel is parked on any GPU).  All active
ranks start each measurement together; the job's time is the max over ranks of the wall time.

  raw      pinned H2D alone, D2H alone, both directions at once, and the step's own mix (2 GiB in while
           1 GiB goes out) — per GPU and summed over the N GPUs: what the copy engines / PCIe / root
           complex deliver with no kernel and no pipeline in the way;
  solo     each GPU alone, in turn (is every GPU's link the same?);
  pairs    GPU 0 together with each other GPU (which GPUs share an upstream path?);
  host     what the host's DRAM gives the CPU: all W ranks copying / adding 1 GiB arrays at once with their
           share of the host threads (plain OpenMP loops, tools/hostbw.c);
  pipeline b200v_vector_add_host over chunk size x slots with N ranks at once: the e2e number itself,
           with cudaHostAlloc buffers and, once, with transparent-huge-page-backed registered buffers.

Developer tool (evidence: profiles/r02_e2e_scaling*.json); not part of the product.
"""
import ctypes as C
import importlib
import json
import mmap
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

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


def job_seconds(fn, active=True, reps=3):
    """Active ranks run fn at the same time; returns best-of-reps of (max over ranks of the wall time)."""
    if active:
        fn()
        sync()
    best = 1e9
    for _ in range(reps):
        R.barrier()
        t0 = time.perf_counter()
        if active:
            fn()
            sync()
        dt = R.max(time.perf_counter() - t0 if active else 0.0)
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

out = {"ranks_launched": world, "host_cores": cores, "host_threads_per_rank": threads, "bytes_per_array": GIB}

# ---- topology, as far as the box shows it ---------------------------------------------------------
if rank == 0:
    topo = {}
    try:
        topo["numa_nodes"] = sorted(d for d in os.listdir("/sys/devices/system/node") if d.startswith("node"))
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
        topo["lscpu"] = [ln.strip() for ln in subprocess.run(["lscpu"], capture_output=True, text=True).stdout.splitlines()
                         if any(k in ln for k in ("Model name", "Socket", "NUMA", "CPU(s):", "Thread", "Hypervisor", "L3"))]
    except Exception:
        pass
    try:
        topo["meminfo"] = {ln.split(":")[0]: ln.split(":")[1].strip() for ln in open("/proc/meminfo")
                           if ln.startswith(("MemTotal", "MemAvailable", "Hugepagesize", "AnonHugePages"))}
        topo["thp"] = open("/sys/kernel/mm/transparent_hugepage/enabled").read().strip()
    except Exception:
        pass
    out["topology"] = topo
    print("topology", json.dumps(topo), flush=True)


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


def pipeline_call(a=hA, b=hB, c=hC):
    pkg.check(L.b200v_vector_add_host(vp(a.ptr), vp(b.ptr), vp(c.ptr), N))


def raw_block(active, m):
    r = {}
    t = job_seconds(h2d, active)
    r["h2d_gbs_per_gpu"], r["h2d_gbs_sum"] = GIB / t * 1e-9, m * GIB / t * 1e-9
    t = job_seconds(d2h, active)
    r["d2h_gbs_per_gpu"], r["d2h_gbs_sum"] = GIB / t * 1e-9, m * GIB / t * 1e-9
    t = job_seconds(both, active)
    r["bidir_each_gbs_per_gpu"], r["bidir_total_gbs_sum"] = GIB / t * 1e-9, 2 * m * GIB / t * 1e-9
    t = job_seconds(step_mix, active)
    r["step_mix_ms"] = t * 1e3
    r["step_mix_e2e_bound_gbs_sum"] = m * 12 * N / t * 1e-9      # 12 B/element through the copy engines alone
    return r


sizes = [m for m in (1, 2, 4, 8) if m <= world]
by_n = {}
for m in sizes:
    active = rank < m
    entry = {"raw": raw_block(active, m)}
    rows = []
    shapes = [(32, 3)] if quick else [(16, 3), (32, 2), (32, 3), (32, 4), (64, 3), (128, 3)]
    for chunk_mib, slots in shapes:
        if active:
            pkg.check(L.b200v_vector_add_host_config((chunk_mib << 20) // 4, slots))
        t = job_seconds(pipeline_call, active)
        rows.append({"chunk_mib": chunk_mib, "slots": slots, "ms": t * 1e3, "e2e_gbs_sum": m * 12 * N / t * 1e-9})
    entry["pipeline"] = rows
    best = min(rows, key=lambda r: r["ms"])
    default = next(r for r in rows if r["chunk_mib"] == 32 and r["slots"] == 3)
    entry["summary"] = {"n_gpus": m, "e2e_default_gbs": default["e2e_gbs_sum"], "e2e_best_gbs": best["e2e_gbs_sum"],
                        "best_shape": [best["chunk_mib"], best["slots"]],
                        "copy_engine_bound_gbs": entry["raw"]["step_mix_e2e_bound_gbs_sum"],
                        "e2e_default_over_copy_engine_bound": default["e2e_gbs_sum"] / entry["raw"]["step_mix_e2e_bound_gbs_sum"]}
    by_n[str(m)] = entry
    if rank == 0:
        print(f"N={m}", json.dumps(entry["raw"]), flush=True)
        print(f"N={m}", json.dumps(entry["summary"]), flush=True)
if rank < world:
    pkg.check(L.b200v_vector_add_host_config((32 << 20) // 4, 3))
out["by_n"] = by_n

# ---- every GPU alone, then GPU 0 with each other GPU ---------------------------------------------------
if world > 1 and not quick:
    solo = []
    for k in range(world):
        t1 = job_seconds(h2d, rank == k, reps=2)
        t2 = job_seconds(d2h, rank == k, reps=2)
        solo.append({"gpu": k, "h2d_gbs": GIB / t1 * 1e-9, "d2h_gbs": GIB / t2 * 1e-9})
    out["solo"] = solo
    pairs = []
    for k in range(1, world):
        t = job_seconds(step_mix, rank in (0, k), reps=2)
        pairs.append({"gpus": [0, k], "step_mix_ms": t * 1e3, "step_mix_e2e_bound_gbs_sum": 2 * 12 * N / t * 1e-9})
    out["pairs_with_gpu0"] = pairs
    if rank == 0:
        print("solo", json.dumps(solo), flush=True)
        print("pairs", json.dumps(pairs), flush=True)

# ---- host DRAM as the CPU sees it: all ranks copy / add 1 GiB arrays at once ----------------------------
HB = C.CDLL(os.path.join(ROOT, "tools", "libhostbw.so"))      # make -C tools
HB.hostbw_copy.argtypes = [vp, vp, C.c_size_t, C.c_int]
HB.hostbw_add.argtypes = [vp, vp, vp, C.c_size_t, C.c_int]
src = np.full(N, 0x3F000000, dtype=np.uint32)
dst = np.zeros(N, dtype=np.uint32)


def host_copy():
    HB.hostbw_copy(dst.ctypes.data_as(vp), src.ctypes.data_as(vp), N, threads)


def host_add():
    HB.hostbw_add(dst.ctypes.data_as(vp), src.ctypes.data_as(vp), hA.view(np.uint32).ctypes.data_as(vp), N, threads)


t = job_seconds(host_copy)
host = {"ranks": world, "threads_per_rank": threads,
        "copy_gbs_sum_algorithmic": world * 2 * GIB / t * 1e-9,          # read + write
        "copy_gbs_sum_with_write_allocate": world * 3 * GIB / t * 1e-9}  # + the read-for-ownership of the destination
t = job_seconds(host_add)
host["add_gbs_sum_algorithmic"] = world * 3 * GIB / t * 1e-9
host["add_gbs_sum_with_write_allocate"] = world * 4 * GIB / t * 1e-9
out["host"] = host
if rank == 0:
    print("host", json.dumps(host), flush=True)
del src, dst

# ---- the pipeline once more with huge-page-backed, registered host buffers (all ranks) ------------------
try:
    maps = []
    bufs = []
    for _ in range(3):
        mm = mmap.mmap(-1, GIB + (2 << 20), flags=mmap.MAP_PRIVATE | mmap.MAP_ANONYMOUS)
        try:
            mm.madvise(mmap.MADV_HUGEPAGE)
        except Exception:
            pass
        arr = np.frombuffer(mm, dtype=np.uint8)
        base = arr.ctypes.data
        off = (-base) % (2 << 20)
        view = arr[off:off + GIB].view(np.uint32)
        view[:] = 0x3F000000                       # touch: pages are faulted in (as huge pages where the kernel agrees)
        pkg.check(L.b200v_host_register(vp(view.ctypes.data), GIB))
        maps.append(mm)
        bufs.append(view)

    class _P:                                      # duck-typed stand-in for PinnedBuffer
        def __init__(self, v):
            self.ptr = v.ctypes.data

    pa, pb, pc = (_P(v) for v in bufs)
    t = job_seconds(lambda: pipeline_call(pa, pb, pc))
    thp = {"e2e_gbs_sum": world * 12 * N / t * 1e-9, "ms": t * 1e3, "n_gpus": world,
           "what": "mmap + MADV_HUGEPAGE + b200v_host_register instead of b200v_host_alloc, default pipeline shape"}
    try:
        thp["anon_huge_pages_kb"] = int(next(ln for ln in open("/proc/self/smaps_rollup") if ln.startswith("AnonHugePages")).split()[1])
    except Exception:
        pass
    for v in bufs:
        pkg.check(L.b200v_host_unregister(vp(v.ctypes.data)))
    out["pipeline_thp_registered"] = thp
    if rank == 0:
        print("thp", json.dumps(thp), flush=True)
except Exception as e:                              # the probe is optional: record why it did not run
    out["pipeline_thp_registered"] = {"error": repr(e)}
    if rank == 0:
        print("thp", repr(e), flush=True)

pkg.check(L.b200v_vector_add_host_release())
if rank == 0:
    table = []
    for m in sizes:
        s = by_n[str(m)]["summary"]
        table.append({"n_gpus": m, "e2e_gbs": round(s["e2e_default_gbs"], 1), "e2e_best_shape_gbs": round(s["e2e_best_gbs"], 1),
                      "copy_engine_bound_gbs": round(s["copy_engine_bound_gbs"], 1),
                      "host_dram_seen_by_cpu_gbs": round(host["add_gbs_sum_with_write_allocate"], 1),
                      "e2e_over_bound": round(s["e2e_default_gbs"] / s["copy_engine_bound_gbs"], 3)})
    out["table"] = table
    print("table", json.dumps(table), flush=True)
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    if args:
        json.dump(out, open(args[0], "w"), indent=1)
R.close()
