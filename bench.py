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
`cpu_baseline` = the oracle's OpenMP loop on this box's host cores (a reported baseline, N = 1).
Further keys, all measured in the same run:
  `triad`               the north-star kernel at the same size, separately timed;
  `sustained`           >= 3 s of back-to-back add (and triad) launches under one event pair, with the
                        NVML clock/power window over exactly that stretch (the loaded operating point);
  `triad_sharded_2p32`  BASELINE.json configs[3]: 2^32 fp32 elements in total, contiguous shards over
                        the ranks (strong scaling), device-timed, max over ranks, every shard's checksum
                        compared with the oracle's;
  `parity`              at every N: each rank's whole 2^28-element output against the oracle's loop over
                        the same global indices (checksum), AND over ranks; the run fails if false;
  `native_node`         rank 0, after every rank has freed its buffers: the pod entrypoint's own per-GPU shard
                        launcher (`gpu-validate --mode node`, fork/exec + pipe barrier) on the same 2^32-element
                        triad;
  `burnin`, `p2p`       rank 0: BASELINE.json configs[4] through the Job manifest's own command (`gpu-validate
                        --mode full --hbm-fraction F --fail-on-throttle`, with a shorter burn-in): per-GPU triad
                        burn-in over a share of every GPU's HBM with every pass verified, then all ordered GPU
                        pairs over NVLink by copy engine, peer loads and peer stores (`p2p` is null at N = 1).
                        These legs run the product binaries as subprocesses.
The oracle is only ever executed as the checker (parity legs) or as the timed CPU baseline
(`cpu_baseline`, `--impl reference`), never on the GPU arm's product path.
`--impl reference` times the CPU loop as the reference arm: the reference repo itself has no
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
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "vectorAdd/triad HBM GB/s per GPU & node at 1/2/4/8 B200, % of ~8 TB/s peak"
LOG2_N = 28
N_ELEMS = 1 << LOG2_N
BYTES_PER_STEP = 12 * N_ELEMS          # 12 B per fp32 element for add/triad
SHARDED_LOG2 = 32                      # configs[3]: 2^32 fp32 elements in total
NOMINAL_GBS = 8000.0
NVLINK_GBS = 900.0
FALLBACK_GBS = 6650.0                  # B200_PROFILING.md fallback when MEASURED_PEAKS.json is absent
SEED = 20261106
M64 = (1 << 64) - 1


def host_cores():
    return len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)


def _cpu_sample_log2():
    """Elements per CPU step in the reference arm / cpu_baseline: the full 2^28 workload (3 GiB of
    arrays, larger than any host L3, like the GPU's) when host RAM allows, else 2^26."""
    try:
        import psutil
        return 28 if psutil.virtual_memory().available > (16 << 30) else 26
    except Exception:
        return 26


CPU_SAMPLE_LOG2 = _cpu_sample_log2()


def workload_config(world):
    """The `config` object of BOTH arms: it names the workload (BASELINE.json configs[1]) and nothing
    that differs between the GPU arm and the CPU reference arm."""
    return {
        "workload": f"vectorAdd fp32 C=A+B, N=2^{LOG2_N} elements per GPU (BASELINE.json configs[1]), one pass over the arrays per step",
        "elements_per_gpu": N_ELEMS, "bytes_per_step_per_gpu": BYTES_PER_STEP, "bytes_per_element": 12,
        "parallelism": f"shard{world} (one process per GPU, no collective on the data path)",
        "l2": f"the three arrays are {3 * 4 * N_ELEMS >> 20} MiB per GPU, far larger than the 126 MB L2 (and any host L3): no flush between steps",
        "inputs": f"index-hash generator, UNIT mode (24-bit k * 2^-24), seed {SEED}, streams 0/1, resident before the timed region",
    }


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
    """Samples SM clock, power and throttle reasons of one GPU through NVML while the bench runs;
    `summary(a, b)` reports the samples between two named marks."""

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

    def _window(self, a, b):
        t0, t1 = self.marks.get(a, 0), self.marks.get(b, float("inf"))
        return [s for s in self.samples if t0 <= s[0] <= t1]

    def _describe(self, inside, window):
        mhz = sorted(s[1] for s in inside)
        mask = 0
        for s in inside:
            mask |= s[2]
        pw = [s[3] for s in inside if s[3] is not None]
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None, "sm_mhz_min": mhz[0] if mhz else None, "sm_max_mhz": self.sm_max,
                "reasons": [n for b, n in self.REASONS.items() if mask & b],
                "power_w_max": max(pw) if pw else None, "power_w_median": sorted(pw)[len(pw) // 2] if pw else None,
                "samples": len(inside), "window": window}

    def summary(self, a="timed_start", b="timed_end", widen=("load_start", "load_end")):
        if not self.ok:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "note": "NVML unavailable: " + getattr(self, "err", "")}
        inside, window = self._window(a, b), f"{a} .. {b}"
        if len(inside) < 3 and widen:   # very short timed region: widen to warm-up .. end of the device-timed legs
            inside = self._window(*widen)
            window = f"{widen[0]} .. {widen[1]} (the {a} .. {b} region is shorter than 3 samples)"
        if not inside:
            inside, window = self.samples, "whole run"
        return self._describe(inside, window)


# ------------------------------------------------------------------------------------------------
# CPU side: the oracle as the timed baseline and as the checker

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
    cores = host_cores()

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


def oracle_shard_checksum(op, start, count, word_offset, threads, s=3.0, chunk=1 << 24):
    """The oracle as the checker: checksum of op(gen(stream 0), gen(stream 1), s) over global indices
    [start, start+count), word indices counted from word_offset, computed in chunks so host RAM never
    holds the arrays (the same walk as tests/test_gpu_parity.py::oracle_checksum_chunked)."""
    import oracle_lib as O
    tot = [0, 0]
    for off in range(0, count, chunk):
        m = min(chunk, count - off)
        x = O.fill("f32", start + off, m, SEED, 0)
        y = O.fill("f32", start + off, m, SEED, 1)
        out = O.op(op, "f32", m, s=s, x=x, y=y, threads=threads)
        c = O.checksum(out, word_offset + off, threads=threads)
        tot = [(tot[0] + c[0]) & M64, (tot[1] + c[1]) & M64]
    return tot[0], tot[1]


def hex_pair(p):
    return ["%016x" % p[0], "%016x" % p[1]]


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
              f"host RAM, {'the full' if CPU_SAMPLE_LOG2 == LOG2_N else 'a bounded sample of the'} 2^{LOG2_N}-element per-GPU workload")
    line = {
        "impl": "reference", "metric": METRIC, "value": round(gbs, 2), "unit": "GB/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.gpus),
        "cpu_baseline": {"value": round(gbs, 2), "unit": "GB/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": round(gbs, 2), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# rank 0: the product's own binaries as subprocesses (native shard launcher, NVLink check)

def run_native(pkg, mode_args, js, timeout):
    """Runs gpu-validate with --json js; returns a dict: rc, verdict line, parsed json (or None), summary lines."""
    cmd = [pkg.LAUNCHER_PATH, *mode_args, "--json", js]
    env = dict(os.environ, OMP_NUM_THREADS=str(host_cores()))     # torchrun exports OMP_NUM_THREADS=1
    t0 = time.perf_counter()
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env)
        rc, out = r.returncode, r.stdout
    except subprocess.TimeoutExpired:
        rc, out = -1, f"timed out after {timeout} s"
    wall = time.perf_counter() - t0
    lines = out.strip().splitlines()
    data = None
    try:
        with open(js) as f:
            data = json.load(f)
    except Exception:
        pass
    return {"rc": rc, "verdict": lines[-1] if lines else "", "json": data, "wall_s": round(wall, 1),
            "summary_lines": [ln for ln in lines if ln.startswith(("[node]", "[p2p]", "reason:"))],
            "cmd": "gpu-validate " + " ".join(mode_args)}


def native_node_leg(pkg, world, iters):
    with tempfile.TemporaryDirectory() as td:
        r = run_native(pkg, ["--mode", "node", "--gpus", str(world), "--log2n", str(SHARDED_LOG2), "--iters", str(iters)],
                       os.path.join(td, "node.json"), 600)
    d = r.pop("json")
    out = dict(r, what="the pod entrypoint's per-GPU shard launcher (fork/exec one child per GPU, pipe start barrier, max over ranks), "
                       f"triad over 2^{SHARDED_LOG2} fp32 elements in total (BASELINE.json configs[3]); every shard verified against the host loop's checksum")
    if d:
        out.update({"gpus": d["gpus"], "elements": d["elements"], "node_gbs": d["node_gbs"], "node_gbs_from": d.get("node_gbs_from"),
                    "max_mean_ms": d["max_mean_ms"], "max_best_ms": d["max_best_ms"], "node_best_gbs": d.get("node_best_gbs"),
                    "gbs_per_gpu": round(d["node_gbs"] / d["gpus"], 1),
                    "frac_of_nominal_8000_per_gpu": round(d["node_gbs"] / d["gpus"] / NOMINAL_GBS, 4),
                    "verified": bool(d["ok"]) and all(k["ok"] == 1 for k in d["ranks"]),
                    "ranks": [{"rank": k["rank"], "ok": k["ok"], "elements": k["elements"], "mean_ms": k["mean_ms"], "best_ms": k["best_ms"]}
                              for k in d["ranks"]]})
    out["ok"] = out["rc"] == 0 and out["verdict"] == "Test PASSED" and bool(d) and out.get("verified", False)
    return out


def native_full_leg(pkg, world, p2p_bytes, seconds, hbm_fraction):
    """The Job manifest's command (deploy/job-node-validate.yaml: burn-in over a share of every GPU's HBM with every
    pass verified and thermal / hardware slowdown failing the run, then the all-pairs NVLink check) with a shorter
    burn-in; returns (burnin, p2p) — the two halves of BASELINE.json configs[4]."""
    args = ["--mode", "full", "--gpus", str(world), "--seconds", f"{seconds:g}", "--hbm-fraction", f"{hbm_fraction:g}",
            "--fail-on-throttle", "--p2p-bytes", str(p2p_bytes), "--iters", "12"]
    with tempfile.TemporaryDirectory() as td:
        js = os.path.join(td, "full.json")
        r = run_native(pkg, args, js, 800)
        try:
            with open(js + ".p2p") as f:
                pj = json.load(f)
        except Exception:
            pj = None
    d = r.pop("json")
    common = {k: r[k] for k in ("cmd", "rc", "verdict", "wall_s")}
    burn = dict(common, summary_lines=[ln for ln in r["summary_lines"] if not ln.startswith("[p2p]")],
                what="per-GPU triad burn-in, one process per GPU, arrays sized to a share of each GPU's free HBM, EVERY pass verified against "
                     "the host loop's checksum, NVML clocks / temperature / power / ECC read under load (BASELINE.json configs[4], first half)")
    if d:
        burn.update({"gpus": d["gpus"], "elements_per_gpu": d["ranks"][0]["elements"], "hbm_fraction": hbm_fraction,
                     "passes_per_gpu": [k["passes"] for k in d["ranks"]], "node_gbs": d["node_gbs"], "gbs_per_gpu": round(d["node_gbs"] / d["gpus"], 1),
                     "frac_of_nominal_8000_per_gpu": round(d["node_gbs"] / d["gpus"] / NOMINAL_GBS, 4),
                     "verified": bool(d["ok"]) and all(k["ok"] == 1 for k in d["ranks"]),
                     "under_load": [{"rank": k["rank"], "sm_min_mhz": k["sm_min_mhz"], "temp_max_c": k["temp_max_c"], "power_max_w": k["power_max_w"],
                                     "clock_limit_reasons": k["clock_limit_reasons"]} for k in d["ranks"]]})
    burn["ok"] = r["rc"] == 0 and r["verdict"] == "Test PASSED" and bool(d) and burn.get("verified", False)
    if world < 2:
        return burn, None
    p2p = dict(common, summary_lines=[ln for ln in r["summary_lines"] if ln.startswith("[p2p]")],
               what="all ordered GPU pairs over NVLink (BASELINE.json configs[4], second half): cudaMemcpyPeer, a kernel loading from the peer, a kernel "
                    "storing into the peer; then every GPU sending and receiving at once; every transfer verified by checksum on the destination")
    if pj:
        p2p.update({"gpus": pj["gpus"], "bytes_per_transfer": pj["bytes"], "pairs": len(pj["pairs"]), "verified": bool(pj["ok"]),
                    "nvlink_gbs_per_direction": pj.get("nvlink_gbs_per_direction", NVLINK_GBS), "gbs": pj.get("summary")})
    p2p["ok"] = r["rc"] == 0 and r["verdict"] == "Test PASSED" and bool(pj) and p2p.get("verified", False)
    return burn, p2p


# ------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--e2e-steps", type=int, default=8)
    ap.add_argument("--sustained-seconds", type=float, default=3.0, help="length of each sustained leg (add, triad)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (N = 1) and every oracle parity leg")
    ap.add_argument("--no-sharded", action="store_true", help="skip the 2^32-element sharded triad leg")
    ap.add_argument("--no-native", action="store_true", help="skip the gpu-validate subprocess legs (native_node, p2p)")
    ap.add_argument("--p2p-bytes", type=int, default=1 << 30)
    ap.add_argument("--burnin-seconds", type=float, default=4.0, help="length of the burn-in inside the gpu-validate --mode full leg")
    ap.add_argument("--hbm-fraction", type=float, default=0.5, help="share of each GPU's free HBM the burn-in arrays cover")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    world_env = int(os.environ.get("WORLD_SIZE", "1"))
    # torchrun exports OMP_NUM_THREADS=1; the oracle legs (the checker) get this rank's share of the host cores
    threads = max(1, host_cores() // world_env)
    os.environ["OMP_NUM_THREADS"] = str(threads)

    if args.impl == "reference":
        os.environ["OMP_NUM_THREADS"] = str(host_cores())
        run_reference(args, int(os.environ.get("RANK", "0")))
        return 0

    import numpy as np
    pkg = importlib.import_module("kubernetes-with-nvidia-gpu_b200")
    ranks_mod = importlib.import_module("kubernetes-with-nvidia-gpu_b200.ranks")
    L = pkg.lib()          # raises if libb200v.so is missing: no fallback
    R = ranks_mod.Ranks(backend="nccl" if world_env > 1 else None,
                        device_sync=lambda: pkg.check(L.b200v_device_sync()))
    rank, local_rank, world = R.rank, R.local_rank, R.world
    pkg.check(L.b200v_set_device(local_rank))
    pkg.check(L.b200v_require_sm100(local_rank))
    info = pkg.device_info(local_rank)

    vp = C.c_void_p
    stream = vp(None)
    pkg.check(L.b200v_stream_create(C.byref(stream)))
    ev0, ev1 = vp(None), vp(None)
    pkg.check(L.b200v_event_create(C.byref(ev0)))
    pkg.check(L.b200v_event_create(C.byref(ev1)))

    sampler = ClockSampler(info.uuid.decode())
    sampler.start()

    def timed(fn, steps):
        """K launches between two events on the launching stream, barrier + sync on both sides; local ms."""
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

    def timed_for(fn, seconds, est_ms):
        """At least `seconds` of back-to-back launches under ONE event pair; returns (local ms, launches)."""
        launches = max(int(seconds * 1e3 / est_ms) + 1, 3)
        return timed(fn, launches), launches

    # ---- configs[1]: vectorAdd, 2^28 fp32 per GPU; rank r owns [r*N, (r+1)*N) of the job's index space ----
    n_local = N_ELEMS
    start = rank * N_ELEMS
    total = world * N_ELEMS
    bytes_local = 12 * n_local
    bytes_job = 12 * total
    nbytes = 4 * n_local
    dA, dB, dC = (pkg.DeviceBuffer(nbytes) for _ in range(3))
    # inputs are filled on the launching stream: the warm-up cannot start before they are complete
    pkg.check(L.b200v_fill(pkg.F32, vp(dA.ptr), start, n_local, SEED, 0, pkg.GEN_UNIT, stream))
    pkg.check(L.b200v_fill(pkg.F32, vp(dB.ptr), start, n_local, SEED, 1, pkg.GEN_UNIT, stream))
    pkg.check(L.b200v_stream_sync(stream))

    add = lambda: pkg.check(L.b200v_add_f32(vp(dC.ptr), vp(dA.ptr), vp(dB.ptr), n_local, stream))
    triad = lambda: pkg.check(L.b200v_triad_f32(vp(dC.ptr), 3.0, vp(dA.ptr), vp(dB.ptr), n_local, stream))

    # ---- device-resident leg (value): W warm-up steps, then exactly K timed steps ----
    sampler.mark("load_start")
    for _ in range(args.warmup):
        add()
    sampler.mark("timed_start")
    ms_local = timed(add, args.steps)
    sampler.mark("timed_end")
    ms_total = R.max(ms_local)
    ms_per_step = ms_total / args.steps
    value = bytes_job / ms_per_step * 1e-6

    # checksum of what was just timed: this rank's whole output, word indices counted from its first global
    # index, so the ranks' pairs add up to the checksum of the job's whole vector
    local_checksum = pkg.checksum(vp(dC.ptr), nbytes, start)
    job_checksum = R.sum_checksums(local_checksum)

    # the other 12 B/element kernel as a second, separately timed line (triad is the north-star kernel)
    for _ in range(args.warmup):
        triad()
    triad_steps = max(min(args.steps, 200), 3)
    triad_ms = R.max(timed(triad, triad_steps)) / triad_steps
    sampler.mark("load_end")
    triad_gbs = bytes_job / triad_ms * 1e-6

    # ---- sustained legs: >= S seconds of back-to-back launches, one event pair, NVML window over exactly that ----
    sustained = {}
    if args.sustained_seconds > 0:
        for name, fn in (("add", add), ("triad", triad)):
            sampler.mark(f"sustained_{name}_start")
            ms_s, launches = timed_for(fn, args.sustained_seconds, ms_local / args.steps)
            sampler.mark(f"sustained_{name}_end")
            ms_s_max = R.max(ms_s)
            gbs = bytes_job * launches / ms_s_max * 1e-6
            sustained[name] = {"gbs": round(gbs, 1), "gbs_per_gpu": round(gbs / world, 1), "launches_per_gpu": launches,
                               "seconds": round(ms_s_max * 1e-3, 3), "ms_per_launch": round(ms_s_max / launches, 5),
                               "frac_of_nominal_8000_per_gpu": round(gbs / world / NOMINAL_GBS, 4),
                               "clocks": sampler.summary(f"sustained_{name}_start", f"sustained_{name}_end", widen=None)}
        sustained["gbs"] = sustained["add"]["gbs"]
        sustained["vs_value"] = round(sustained["add"]["gbs"] / value, 4)
        sustained["what"] = (f"back-to-back b200v_add_f32 / b200v_triad_f32 launches at N=2^{LOG2_N} per GPU for >= {args.sustained_seconds:g} s each, "
                             "one event pair around all of them, max over ranks; clocks = rank 0's GPU over exactly that stretch")

    for b in (dA, dB, dC):
        b.free()

    # ---- configs[3]: triad over 2^32 fp32 elements in total, contiguous shards (strong scaling) ----
    sharded = None
    sharded_ok = True
    if not args.no_sharded:
        tot32 = 1 << SHARDED_LOG2
        s_begin, s_count = pkg.shard_range(tot32, rank, world)
        sb = 4 * s_count
        sOut, sX, sY = (pkg.DeviceBuffer(sb) for _ in range(3))     # out = fma(3, x, y)
        pkg.check(L.b200v_fill(pkg.F32, vp(sX.ptr), s_begin, s_count, SEED, 0, pkg.GEN_UNIT, stream))
        pkg.check(L.b200v_fill(pkg.F32, vp(sY.ptr), s_begin, s_count, SEED, 1, pkg.GEN_UNIT, stream))
        pkg.check(L.b200v_stream_sync(stream))
        striad = lambda: pkg.check(L.b200v_triad_f32(vp(sOut.ptr), 3.0, vp(sX.ptr), vp(sY.ptr), s_count, stream))
        for _ in range(args.warmup):
            striad()
        pkg.check(L.b200v_memset(vp(sOut.ptr), 0x5A, sb, stream))   # the timed launches, not the warm-up, must produce the result
        s_steps = max(min(args.steps, 50), 3)
        s_ms_local = timed(striad, s_steps)
        s_ms = R.max(s_ms_local) / s_steps
        s_local_ck = pkg.checksum(vp(sOut.ptr), sb, s_begin)
        s_job_ck = R.sum_checksums(s_local_ck)
        for b in (sOut, sX, sY):
            b.free()
        sharded = {"elements_total": tot32, "elements_this_rank": s_count, "scaling": "strong", "steps": s_steps,
                   "ms_per_step": round(s_ms, 5), "node_gbs": round(12 * tot32 / s_ms * 1e-6, 1),
                   "gbs_per_gpu": round(12 * tot32 / s_ms * 1e-6 / world, 1),
                   "frac_of_nominal_8000_per_gpu": round(12 * tot32 / s_ms * 1e-6 / world / NOMINAL_GBS, 4),
                   "job_checksum": hex_pair(s_job_ck), "matches_oracle": None,
                   "what": f"b200v_triad_f32 over this rank's b200v_shard_range share of 2^{SHARDED_LOG2} fp32 elements (BASELINE.json configs[3]); "
                           "device-timed, max over ranks; the output is scrubbed after the warm-up and every rank's checksum is compared with the oracle's"}
        if not args.no_cpu:      # the oracle, as the checker, over this rank's shard
            t0 = time.perf_counter()
            want = oracle_shard_checksum("triad", s_begin, s_count, s_begin, threads)
            sharded_ok = R.all_ok(want == s_local_ck)
            sharded["matches_oracle"] = sharded_ok
            sharded["oracle_seconds"] = round(time.perf_counter() - t0, 1)

    # ---- end-to-end leg: host buffers through the public call, copies inside the timed region ----
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
    # H2D/D2H pipeline): compare checksums
    def host_checksum(words, word_offset):
        idx = (np.arange(words.size, dtype=np.uint64) + np.uint64(word_offset)) & np.uint64(0xFFFFFFFF) | np.uint64(1)
        w64 = words.astype(np.uint64)
        return int(w64.sum(dtype=np.uint64)), int((w64 * idx).sum(dtype=np.uint64))
    e2e_ok = R.all_ok(host_checksum(vc, start) == local_checksum)
    pkg.check(L.b200v_vector_add_host_release())
    del va, vb, vc
    for h in (hA, hB, hC):
        h.free()

    # ---- parity of the timed output against the oracle, at every N ----
    # N = 1: the cpu_baseline leg below recomputes the same vector (and is the timed CPU baseline as well).
    # N > 1: every rank runs the oracle, as the checker, over ITS OWN global index range; AND over ranks.
    parity_ok = None
    parity_how = None
    if not args.no_cpu and world > 1:
        t0 = time.perf_counter()
        want = oracle_shard_checksum("add", start, n_local, start, threads)
        parity_ok = R.all_ok(want == local_checksum)
        parity_how = (f"every rank: oracle loop over its own global indices [r*2^{LOG2_N}, (r+1)*2^{LOG2_N}) on {threads} host threads, "
                      f"checksum == the device's ({time.perf_counter() - t0:.1f} s on rank 0); AND over ranks")

    sampler.stop_flag = True
    sampler.join(timeout=2)
    R.barrier()

    # ---- rank 0 alone: the product's own launcher binaries; the other ranks wait on the host (no GPU work) ----
    native_node = burnin = p2p = None
    if rank == 0 and not args.no_native and os.path.exists(pkg.LAUNCHER_PATH):
        native_node = native_node_leg(pkg, world, iters=20)
        burnin, p2p = native_full_leg(pkg, world, args.p2p_bytes, args.burnin_seconds, args.hbm_fraction)
    R.host_barrier()
    if rank != 0:
        R.close()
        return 0

    peak, peak_src = measured_peak()
    achieved = bytes_local / (ms_local / args.steps) * 1e-6     # rank 0's kernel, GB/s
    cfg = workload_config(world)
    line = {
        "metric": METRIC, "value": round(value, 1), "unit": "GB/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms_per_step, 5), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": cfg,
        "device": {"gpu": info.name.decode(), "sm_count": info.sm_count, "host_cores": host_cores()},
        "roofline": {
            "bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
            "frac": round(achieved / peak, 4), "traffic": ncu_traffic_per_launch(),
            "peak_source": peak_src, "kernel": "b200v::stream_kernel<add, f32, 256-bit> via b200v_add_f32",
            "algorithmic_bytes_per_launch": bytes_local,
            "frac_of_nominal_8000": round(achieved / NOMINAL_GBS, 4),
        },
        "triad": {"node_gbs": round(triad_gbs, 1), "gbs_per_gpu": round(triad_gbs / world, 1), "ms_per_step": round(triad_ms, 5),
                  "steps": triad_steps, "frac_of_measured": round(triad_gbs / world / peak, 4),
                  "frac_of_nominal_8000": round(triad_gbs / world / NOMINAL_GBS, 4)},
        "sustained": sustained or None,
        "triad_sharded_2p32": sharded,
        "e2e": {"value": round(e2e_value, 2), "unit": "GB/s", "h2d_bytes_per_step": 8 * n_e2e * world,
                "d2h_bytes_per_step": 4 * n_e2e * world, "ms_per_step": round(e2e_ms, 3), "steps": args.e2e_steps,
                "api": f"b200v_vector_add_host(hA, hB, hC, 2^{LOG2_N}) per rank with pinned host buffers; host wall clock around "
                       "the synchronous call, max over ranks; bound by PCIe (2 GiB in + 1 GiB out per step per GPU)"},
        "gpu_launches": args.steps * world, "gpu_launches_per_rank": args.steps,
        "parity": {"job_checksum": hex_pair(job_checksum),
                   "e2e_matches_device": e2e_ok, "e2e_check": "host result checksum == device-resident result checksum, every rank",
                   "device_matches_cpu_baseline": parity_ok, "how": parity_how},
        "native_node": native_node, "burnin": burnin, "p2p": p2p,
        "clocks": sampler.summary(),
    }
    if not args.no_cpu and world == 1:      # the timed CPU baseline is reported on rank 0 at N = 1 only
        line["cpu_baseline"] = cpu_baseline()
        want = line["cpu_baseline"].get("output_checksum")
        if want is not None:   # whole 2^28 output, GPU vs the oracle's loop: bit-exact or the run fails
            parity_ok = want == line["parity"]["job_checksum"]
            line["parity"]["device_matches_cpu_baseline"] = parity_ok
            line["parity"]["how"] = "the cpu_baseline leg's output (oracle loop over the same 2^28 global indices): checksum == the device's"
    print(json.dumps(line), flush=True)
    R.close()
    ok = (parity_ok is not False) and e2e_ok and sharded_ok
    ok = ok and all(leg is None or leg["ok"] for leg in (native_node, burnin, p2p))
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
