#!/usr/bin/env python3
"""bench.py — headline benchmark of the B200 validation workload (BASELINE.json configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A step is one pass of the hot path over one batch: ONE launch of the fp32 vectorAdd kernel
(C = A + B, N = 2^28 elements, 3 x 1 GiB arrays) per GPU, through the C ABI (b200v_add_f32).
Inputs are synthetic (index-hash generator, filled on the device before the timed region) and
the three arrays are 3 GiB >> the 126 MB L2, so no L2 flush is needed between launches.
With N GPUs every rank runs the same per-GPU workload on its own device (weak scaling, no
data-path collective: the vectors shard with no communication); torch.distributed/NCCL is used
only for the barrier and the max-over-ranks of the device-timed durations.

One JSON line on stdout (rank 0).  `value` = whole-job GB/s of algorithmic bytes (12 B/element);
`e2e` = the same metric through b200v_vector_add_host with pinned HOST buffers (H2D + kernel +
D2H inside the timed region); `roofline` = the add kernel against the measured HBM copy peak;
`cpu_baseline` = the oracle's OpenMP loop on this box's host cores (a reported baseline).
The oracle is only ever executed inside the cpu_baseline / --impl reference legs; at N = 1 that leg
also yields the parity check of the whole timed output (its checksum must equal the device's).
`--impl reference` times that CPU loop as the reference arm: the reference repo itself has no
compute code for this path (its Step 9 pod runs `nvidia-smi`), so the CPU restatement in oracle/
is the reference implementation.
"""
import argparse
import ctypes as C
import importlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "vectorAdd/triad HBM GB/s per GPU & node at 1/2/4/8 B200, % of ~8 TB/s peak"
LOG2_N = 28
N_ELEMS = 1 << LOG2_N
BYTES_PER_STEP = 12 * N_ELEMS          # 12 B per fp32 element for add/triad
NOMINAL_GBS = 8000.0
FALLBACK_GBS = 6650.0                  # B200_PROFILING.md fallback when MEASURED_PEAKS.json is absent
SEED = 20261106
def _cpu_sample_log2():
    """Elements per CPU step in the reference arm / cpu_baseline: the full 2^28 workload (3 GiB of
    arrays, larger than any host L3, like the GPU's) when host RAM allows, else 2^26."""
    try:
        import psutil
        return 28 if psutil.virtual_memory().available > (16 << 30) else 26
    except Exception:
        return 26


CPU_SAMPLE_LOG2 = _cpu_sample_log2()


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    except Exception:
        return FALLBACK_GBS, "B200_PROFILING.md fallback 6.65 TB/s (of fallback)"


def ncu_traffic_per_launch():
    """dram__bytes_read.sum + dram__bytes_write.sum of one add launch, from the committed ncu capture."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f)["add_f32_2p28"]["dram_bytes_per_launch"]
    except Exception:
        return None


# ------------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU through NVML while the timed region runs."""

    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown",
               0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake"}

    def __init__(self, uuid, period=0.005):
        super().__init__(daemon=True)
        self.period = period
        self.samples = []          # (t, sm_mhz, reasons_mask, power_w)
        self.marks = {}
        self.stop_flag = False
        self.ok = False
        self.sm_max = None
        try:
            import pynvml
            self.nv = pynvml
            pynvml.nvmlInit()
            try:
                self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode() if isinstance(uuid, str) else uuid)
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(0)
            self.sm_max = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # NVML not loadable: report that, never invent clocks
            self.err = repr(e)

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        while not self.stop_flag:
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                try:
                    pw = nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0
                except Exception:
                    pw = None
                self.samples.append((time.perf_counter(), mhz, mask, pw))
            except Exception:
                pass
            time.sleep(self.period)

    def mark(self, name):
        self.marks[name] = time.perf_counter()

    def summary(self):
        if not self.ok:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "note": "NVML unavailable: " + getattr(self, "err", "")}
        t0, t1 = self.marks.get("timed_start", 0), self.marks.get("timed_end", float("inf"))
        inside = [s for s in self.samples if t0 <= s[0] <= t1]
        window = "timed region"
        if len(inside) < 3:   # very short timed region: widen to warm-up .. end of the device-timed legs
            inside = [s for s in self.samples if self.marks.get("load_start", 0) <= s[0] <= self.marks.get("load_end", t1)]
            window = "warm-up + timed + triad legs (timed region shorter than 3 samples)"
        if not inside:
            inside = self.samples
        mhz = sorted(s[1] for s in inside)
        mask = 0
        for s in inside:
            mask |= s[2]
        pw = [s[3] for s in inside if s[3] is not None]
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None, "sm_max_mhz": self.sm_max,
                "reasons": [n for b, n in self.REASONS.items() if mask & b],
                "power_w_max": max(pw) if pw else None, "samples": len(inside), "window": window}


# ------------------------------------------------------------------------------------------------
def cpu_step_runner(log2n):
    """Returns (run_one_step, cores, n): the oracle's OpenMP vectorAdd on host memory."""
    import numpy as np
    import oracle_lib as O
    n = 1 << log2n
    a = O.fill("f32", 0, n, SEED, 0)
    b = O.fill("f32", 0, n, SEED, 1)
    c = np.zeros(n, dtype=np.uint32)
    L = O.lib()
    pa, pb, pc = (x.ctypes.data_as(C.c_void_p) for x in (a, b, c))

    # every host core this process may run on; explicit, because torchrun exports OMP_NUM_THREADS=1
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)

    def step():
        L.oracle_add_f32(pc, pa, pb, n, cores)

    step.keep = (a, b, c)
    step.checksum = lambda: O.checksum(c, 0, threads=cores)   # of the last pass's output (oracle, used as the checker)
    return step, cores, n


def cpu_baseline(budget_s=10.0):
    step, cores, n = cpu_step_runner(CPU_SAMPLE_LOG2)
    step()
    t0 = time.perf_counter()
    reps = 0
    while True:
        step()
        reps += 1
        dt = time.perf_counter() - t0
        if dt >= budget_s or reps >= 20000:
            break
    gbs = 12.0 * n * reps / dt * 1e-9
    out = {"value": round(gbs, 2), "unit": "GB/s", "cores": cores, "kind": "port",
           "sample": f"oracle_add_f32 (OpenMP, -O3 -ffp-contract=off), N=2^{CPU_SAMPLE_LOG2} fp32 per pass, "
                     f"{reps} passes in {dt:.1f} s; the reference repo has no code for this path, so the port is the baseline"}
    # The CPU leg computes C = A + B over the same seeded inputs as rank 0's GPU shard (global indices
    # [0, 2^28)) when the sample is the full workload: its checksum is the parity check of the whole output.
    if n == N_ELEMS:
        cs = step.checksum()
        out["output_checksum"] = ["%016x" % cs[0], "%016x" % cs[1]]
    return out


def run_reference(args, rank):
    """Reference arm: the CPU implementation of the path on this box's host cores."""
    if rank != 0:
        return
    step, cores, n = cpu_step_runner(CPU_SAMPLE_LOG2)
    for _ in range(max(args.warmup, 1)):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    gbs = 12.0 * n * args.steps / dt * 1e-9
    sample = (f"each step = oracle_add_f32 (OpenMP, {cores} threads) over N=2^{CPU_SAMPLE_LOG2} fp32 elements resident in "
              f"host RAM, {'the full' if CPU_SAMPLE_LOG2 == LOG2_N else 'a bounded sample of the'} 2^{LOG2_N} workload")
    line = {
        "impl": "reference", "metric": METRIC, "value": round(gbs, 2), "unit": "GB/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"vectorAdd fp32 C=A+B, N=2^{LOG2_N} per GPU (BASELINE.json configs[1]); CPU arm runs a "
                               f"2^{CPU_SAMPLE_LOG2}-element sample per step", "bytes_per_element": 12},
        "cpu_baseline": {"value": round(gbs, 2), "unit": "GB/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": round(gbs, 2), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="vectoradd", choices=["vectoradd", "triad-sharded"],
                    help="vectoradd: configs[1], 2^28 fp32 per GPU (weak, the default and the graded line); "
                         "triad-sharded: configs[3], 2^32 fp32 in total split over the ranks (strong)")
    ap.add_argument("--e2e-steps", type=int, default=8)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    if args.impl == "reference":
        run_reference(args, int(os.environ.get("RANK", "0")))
        return 0

    import numpy as np
    pkg = importlib.import_module("kubernetes-with-nvidia-gpu_b200")
    ranks_mod = importlib.import_module("kubernetes-with-nvidia-gpu_b200.ranks")
    L = pkg.lib()          # raises if libb200v.so is missing: no fallback
    R = ranks_mod.Ranks(backend="nccl" if int(os.environ.get("WORLD_SIZE", "1")) > 1 else None,
                        device_sync=lambda: pkg.check(L.b200v_device_sync()))
    rank, local_rank, world = R.rank, R.local_rank, R.world
    pkg.check(L.b200v_set_device(local_rank))
    pkg.check(L.b200v_require_sm100(local_rank))
    info = pkg.device_info(local_rank)

    sharded = args.workload == "triad-sharded"
    if sharded:
        total = 1 << 32
        start, n_local = pkg.shard_range(total, rank, world)
        main_op, main_name = pkg.OP_TRIAD, "triad"
    else:
        n_local = N_ELEMS
        start = rank * N_ELEMS          # rank r owns [r*N, (r+1)*N) of the job's global index space
        total = world * N_ELEMS
        main_op, main_name = pkg.OP_ADD, "add"
    bytes_local = 12 * n_local
    bytes_job = 12 * total

    vp = C.c_void_p
    nbytes = 4 * n_local
    dA, dB, dC = (pkg.DeviceBuffer(nbytes) for _ in range(3))
    pkg.check(L.b200v_fill(pkg.F32, vp(dA.ptr), start, n_local, SEED, 0, pkg.GEN_UNIT, None))
    pkg.check(L.b200v_fill(pkg.F32, vp(dB.ptr), start, n_local, SEED, 1, pkg.GEN_UNIT, None))
    stream = vp(None)
    pkg.check(L.b200v_stream_create(C.byref(stream)))
    ev0, ev1 = vp(None), vp(None)
    pkg.check(L.b200v_event_create(C.byref(ev0)))
    pkg.check(L.b200v_event_create(C.byref(ev1)))

    sampler = ClockSampler(info.uuid.decode())
    sampler.start()

    def timed(fn, steps):
        R.barrier()
        pkg.check(L.b200v_event_record(ev0, stream))
        for _ in range(steps):
            fn()
        pkg.check(L.b200v_event_record(ev1, stream))
        pkg.check(L.b200v_event_sync(ev1))
        R.barrier()
        ms = C.c_float(0)
        pkg.check(L.b200v_event_elapsed_ms(ev0, ev1, C.byref(ms)))
        return ms.value

    add = lambda: pkg.check(L.b200v_add_f32(vp(dC.ptr), vp(dA.ptr), vp(dB.ptr), n_local, stream))
    triad = lambda: pkg.check(L.b200v_triad_f32(vp(dC.ptr), 3.0, vp(dA.ptr), vp(dB.ptr), n_local, stream))
    main_fn, other_fn = (triad, add) if sharded else (add, triad)

    # ---- device-resident leg (value): W warm-up steps, then exactly K timed steps ----
    sampler.mark("load_start")
    for _ in range(args.warmup):
        main_fn()
    sampler.mark("timed_start")
    ms_local = timed(main_fn, args.steps)
    sampler.mark("timed_end")
    ms_total = R.max(ms_local)
    ms_per_step = ms_total / args.steps
    value = bytes_job / ms_per_step * 1e-6

    # checksum of what was just timed (whole output of every rank, summed); at N = 1 the cpu_baseline leg
    # below recomputes the same vector with the oracle and the two must agree
    local_checksum = pkg.checksum(vp(dC.ptr), nbytes, start)
    job_checksum = R.sum_checksums(local_checksum)

    # the other 12 B/element kernel as a second, separately timed line (triad is the north-star kernel)
    for _ in range(args.warmup):
        other_fn()
    other_steps = max(min(args.steps, 200), 3)
    other_ms = R.max(timed(other_fn, other_steps)) / other_steps
    sampler.mark("load_end")
    other_gbs = bytes_job / other_ms * 1e-6

    # ---- end-to-end leg: host buffers through the public call, copies inside the timed region ----
    for b in (dA, dB, dC):
        b.free()
    n_e2e = N_ELEMS
    e2e_bytes = 4 * n_e2e
    hA, hB, hC = (pkg.PinnedBuffer(e2e_bytes) for _ in range(3))
    va, vb, vc = (h.view(np.uint32) for h in (hA, hB, hC))
    # host inputs = the same generator stream, synthesised on the device and copied out before the timed region
    tmp = pkg.DeviceBuffer(e2e_bytes)
    for h, sid in ((hA, 0), (hB, 1)):
        pkg.check(L.b200v_fill(pkg.F32, vp(tmp.ptr), start, n_e2e, SEED, sid, pkg.GEN_UNIT, None))
        pkg.check(L.b200v_memcpy_d2h(vp(h.ptr), vp(tmp.ptr), e2e_bytes, None))
        pkg.check(L.b200v_device_sync())
    tmp.free()
    e2e_call = lambda: pkg.check(L.b200v_vector_add_host(vp(hA.ptr), vp(hB.ptr), vp(hC.ptr), n_e2e))
    e2e_call()   # warm-up: allocates the staging buffers
    e2e_call()
    R.barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        e2e_call()     # synchronous: returns when hC is complete in host memory
    e2e_ms_local = (time.perf_counter() - t0) * 1e3
    R.barrier()
    e2e_ms = R.max(e2e_ms_local) / args.e2e_steps
    e2e_value = world * 12 * n_e2e / e2e_ms * 1e-6
    # the host result must be the vector the device-resident leg produced (same inputs, same kernel, plus the
    # H2D/D2H pipeline): compare checksums; at N = 1 and the vectoradd workload that is the timed output itself
    def host_checksum(words, word_offset):
        idx = (np.arange(words.size, dtype=np.uint64) + np.uint64(word_offset)) & np.uint64(0xFFFFFFFF) | np.uint64(1)
        w64 = words.astype(np.uint64)
        return int(w64.sum(dtype=np.uint64)), int((w64 * idx).sum(dtype=np.uint64))
    if sharded:
        e2e_ok = True          # different op and shard: covered by tests/test_gpu_parity.py::test_vector_add_host_buffers
        e2e_note = "not compared in the sharded workload"
    else:
        e2e_ok = host_checksum(vc, start) == local_checksum
        e2e_note = "host result checksum == device-resident result checksum"
    e2e_ok = R.all_ok(e2e_ok)
    pkg.check(L.b200v_vector_add_host_release())
    del va, vb, vc
    for h in (hA, hB, hC):
        h.free()

    sampler.stop_flag = True
    sampler.join(timeout=2)
    R.barrier()
    if rank != 0:
        R.close()
        return 0

    peak, peak_src = measured_peak()
    achieved = bytes_local / (ms_local / args.steps) * 1e-6     # rank 0's kernel, GB/s
    other_name = "add" if sharded else "triad"
    line = {
        "metric": METRIC, "value": round(value, 1), "unit": "GB/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms_per_step, 5), "higher_is_better": True,
        "scaling": "strong" if sharded else "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {
            "workload": ("triad fp32 a=fma(s,c,b), 2^32 elements in total sharded over the ranks (BASELINE.json configs[3])" if sharded else
                         f"vectorAdd fp32 C=A+B, N=2^{LOG2_N} elements per GPU (BASELINE.json configs[1])") + ", one kernel launch per step",
            "elements_per_gpu": n_local, "bytes_per_step_per_gpu": bytes_local, "bytes_per_element": 12,
            "parallelism": f"shard{world} (one process per GPU, no collective on the data path)",
            "l2": f"the three arrays are {3 * nbytes >> 20} MiB per GPU, far larger than the 126 MB L2: no flush between steps",
            "gpu": info.name.decode(), "sm_count": info.sm_count,
            "inputs": f"index-hash generator, UNIT mode (24-bit k * 2^-24), seed {SEED}, streams 0/1, filled on the device before the timed region",
        },
        "roofline": {
            "bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
            "frac": round(achieved / peak, 4), "traffic": None if sharded else ncu_traffic_per_launch(),
            "peak_source": peak_src, "kernel": f"b200v::stream_kernel<{main_name}, f32, 256-bit> via b200v_{main_name}_f32",
            "algorithmic_bytes_per_launch": bytes_local,
            "frac_of_nominal_8000": round(achieved / NOMINAL_GBS, 4),
        },
        other_name: {"node_gbs": round(other_gbs, 1), "gbs_per_gpu": round(other_gbs / world, 1), "ms_per_step": round(other_ms, 5),
                     "steps": other_steps, "frac_of_measured": round(other_gbs / world / peak, 4),
                     "frac_of_nominal_8000": round(other_gbs / world / NOMINAL_GBS, 4)},
        "e2e": {"value": round(e2e_value, 2), "unit": "GB/s", "h2d_bytes_per_step": 8 * n_e2e * world,
                "d2h_bytes_per_step": 4 * n_e2e * world, "ms_per_step": round(e2e_ms, 3), "steps": args.e2e_steps,
                "api": f"b200v_vector_add_host(hA, hB, hC, 2^{LOG2_N}) per rank with pinned host buffers; host wall clock around "
                       "the synchronous call, max over ranks; bound by PCIe (2 GiB in + 1 GiB out per step per GPU)"},
        "gpu_launches": args.steps * world, "gpu_launches_per_rank": args.steps,
        "parity": {"job_checksum": ["%016x" % job_checksum[0], "%016x" % job_checksum[1]],
                   "e2e_matches_device": e2e_ok, "e2e_check": e2e_note,
                   "device_matches_cpu_baseline": None},
        "clocks": sampler.summary(),
    }
    parity_ok = True
    if not args.no_cpu and world == 1:      # reported on rank 0 at N=1 only
        line["cpu_baseline"] = cpu_baseline()
        want = line["cpu_baseline"].get("output_checksum")
        if want is not None and not sharded:   # whole 2^28 output, GPU vs the oracle's loop: bit-exact or the run fails
            parity_ok = want == line["parity"]["job_checksum"]
            line["parity"]["device_matches_cpu_baseline"] = parity_ok
    print(json.dumps(line), flush=True)
    R.close()
    return 0 if (parity_ok and e2e_ok) else 1


if __name__ == "__main__":
    sys.exit(main())
