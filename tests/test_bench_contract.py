"""bench.py's JSON contract.  CPU tier: the reference arm (the CPU loop) prints the agreed keys.
GPU tier: our arm prints one line with value/e2e/roofline/cpu_baseline/clocks and passes its own
in-run parity checks."""
import json
import os
import subprocess
import sys

import pytest

from b200v_pkg import ROOT

BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
             "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches"}


def run_bench(*args, timeout=600):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, r.stderr[-3000:]
    lines = [ln for ln in r.stdout.strip().splitlines() if ln.startswith("{")]
    assert len(lines) == 1, r.stdout
    return json.loads(lines[0])


def test_reference_arm_line():
    d = run_bench("--impl", "reference", "--steps", "2", "--warmup", "1")
    assert BASE_KEYS <= set(d) and d["impl"] == "reference"
    assert d["unit"] == "GB/s" and d["higher_is_better"] is True and d["vs_baseline"] is None and d["dtype"] == "f32"
    assert d["e2e"] == {"value": d["value"], "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and "2^" in cb["sample"]
    assert "workload" in d["config"] and "model" not in d["config"]
    assert d["value"] > 0.5 and d["gpu_launches"] == 0
    # both arms print the SAME config object for the same N (it names the workload, nothing arm-specific)
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    assert d["config"] == bench.workload_config(1)
    assert bench.workload_config(8)["parallelism"].startswith("shard8")


def test_oracle_shard_checksums_add_up():
    """The N > 1 parity leg: every rank checks its own global index range; the shards' checksums add up to the
    whole vector's, and a chunked walk equals a one-shot one."""
    import importlib.util
    import oracle_lib as O
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    n, world = 300_007, 3
    x = O.fill("f32", 0, n, bench.SEED, 0)
    y = O.fill("f32", 0, n, bench.SEED, 1)
    for op in ("add", "triad"):
        whole = O.checksum(O.op(op, "f32", n, s=3.0, x=x, y=y))
        parts = [0, 0]
        for r in range(world):
            b, c = r * n // world, (r + 1) * n // world - r * n // world
            ck = bench.oracle_shard_checksum(op, b, c, b, threads=2, chunk=1 << 15)
            parts = [(parts[0] + ck[0]) & bench.M64, (parts[1] + ck[1]) & bench.M64]
        assert tuple(parts) == whole


@pytest.mark.gpu
def test_our_arm_line():
    d = run_bench("--steps", "30", "--warmup", "3", "--e2e-steps", "2", "--sustained-seconds", "0.5", "--burnin-seconds", "1.5", timeout=900)
    assert BASE_KEYS <= set(d) and "impl" not in d
    assert d["metric"].startswith("vectorAdd/triad HBM GB/s") and d["scaling"] == "weak" and d["data"] == "synthetic"
    assert d["n_gpus"] == 1 and d["steps"] == 30 and d["gpu_launches"] == 30
    assert 3000 < d["value"] < 12000
    rf = d["roofline"]
    assert rf["bound"] == "hbm" and rf["unit"] == "GB/s" and abs(rf["frac"] - rf["achieved"] / rf["peak"]) < 1e-3
    assert rf["algorithmic_bytes_per_launch"] == 12 << 28
    e = d["e2e"]
    assert e["h2d_bytes_per_step"] == 8 << 28 and e["d2h_bytes_per_step"] == 4 << 28 and 5 < e["value"] < 200
    assert e["value"] < d["value"] / 10          # a host-buffer number, not a copy of the device-resident one
    # whole 2^28-element output: device checksum == the oracle's (computed in the cpu_baseline leg), and the
    # host-buffer result == the device-resident result
    assert d["parity"]["device_matches_cpu_baseline"] is True and d["parity"]["e2e_matches_device"] is True
    assert d["parity"]["job_checksum"] == d["cpu_baseline"]["output_checksum"]
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    assert "sm_mhz" in d["clocks"] and "reasons" in d["clocks"]
    assert "workload" in d["config"] and "model" not in d["config"] and "gpu" not in d["config"]
    assert d["device"]["gpu"] == "NVIDIA B200" and d["device"]["sm_count"] == 148
    # sustained: back-to-back launches for the asked time, within a few % of the K-step value, clocks over that window
    su = d["sustained"]
    for k in ("add", "triad"):
        assert su[k]["seconds"] >= 0.45 and su[k]["launches_per_gpu"] > 500 and 0.9 < su[k]["gbs"] / d["value"] < 1.1
        assert su[k]["clocks"]["samples"] >= 10 and su[k]["clocks"]["power_w_max"] > 200
    assert su["gbs"] == su["add"]["gbs"]
    # configs[3]: 2^32-element triad (one shard at N = 1), device checksum == the oracle's
    sh = d["triad_sharded_2p32"]
    assert sh["elements_total"] == 1 << 32 and sh["matches_oracle"] is True and sh["scaling"] == "strong"
    assert 3000 < sh["node_gbs"] < 12000
    # the product's own shard launcher, run by the bench as a subprocess
    nn = d["native_node"]
    assert nn["ok"] and nn["verified"] and nn["verdict"] == "Test PASSED" and nn["elements"] == 1 << 32 and nn["gpus"] == 1
    assert nn["node_gbs_from"] == "max over ranks of mean ms" and 3000 < nn["node_gbs"] < 12000
    # the Job manifest's command (burn-in over a share of the HBM, every pass verified; then the NVLink check)
    bi = d["burnin"]
    assert bi["ok"] and bi["verified"] and bi["verdict"] == "Test PASSED" and "--mode full" in bi["cmd"] and "--fail-on-throttle" in bi["cmd"]
    assert bi["elements_per_gpu"] > (1 << 32) and min(bi["passes_per_gpu"]) >= 2 and 3000 < bi["gbs_per_gpu"] < 12000
    assert bi["under_load"][0]["power_max_w"] > 200
    assert d["p2p"] is None          # one GPU: no pair to check


def _bench_module():
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    return bench


def test_clock_sampler_windows():
    """The NVML window logic with synthetic samples: a leg's summary covers exactly the samples between its marks;
    the headline window widens to the loaded stretch when the timed region is shorter than three samples."""
    bench = _bench_module()
    s = bench.ClockSampler.__new__(bench.ClockSampler)       # no NVML in this tier: fill the fields by hand
    s.ok, s.sm_max, s.marks = True, 1965, {}
    s.samples = [(t * 0.01, 1965 if t < 50 else 1700, 0 if t < 50 else 0x4, 200.0 if t < 50 else 990.0) for t in range(100)]
    s.marks.update(load_start=0.10, timed_start=0.20, timed_end=0.215, load_end=0.40, sustained_add_start=0.60, sustained_add_end=0.90)
    head = s.summary()
    assert head["samples"] == 31 and "shorter than 3 samples" in head["window"] and head["sm_mhz"] == 1965 and head["reasons"] == []
    leg = s.summary("sustained_add_start", "sustained_add_end", widen=None)
    assert leg["samples"] == 31 and leg["sm_mhz"] == 1700 and leg["sm_mhz_min"] == 1700 and leg["reasons"] == ["sw_power_cap"]
    assert leg["power_w_max"] == 990.0 and leg["window"] == "sustained_add_start .. sustained_add_end"
    s.marks.update(timed_end=0.30)
    assert s.summary()["samples"] == 11 and s.summary()["window"] == "timed_start .. timed_end"
    s.ok = False
    s.err = "no NVML"
    assert s.summary()["sm_mhz"] is None and "NVML unavailable" in s.summary()["note"]


def test_committed_launch_summary_matches_its_raw_list(tmp_path):
    """profiles/r02_ncu_launches_summary.txt is what tools/ncu_summarise.py makes of the committed raw ncu launch list."""
    raw = os.path.join(ROOT, "profiles", "r02_ncu_launches_raw.csv")
    out = tmp_path / "summary.txt"
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summarise.py"), "launches", raw, str(out)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    fresh = [ln for ln in out.read_text().splitlines() if not ln.startswith("#")]
    kept = [ln for ln in open(os.path.join(ROOT, "profiles", "r02_ncu_launches_summary.txt")).read().splitlines() if not ln.startswith("#")]
    assert fresh == kept and len(fresh) >= 5
    big_add = next(ln for ln in fresh if "stream_kernel<0, 0, 32, 4, 512, 0, 0, 1>" in ln)
    assert 440 < float(big_add.split("x")[1].split("us")[0]) < 460           # the headline kernel: ~0.45 ms per launch under ncu too
