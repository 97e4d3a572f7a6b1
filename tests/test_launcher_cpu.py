"""CPU tier: the pod entrypoint's contract where no GPU exists — verdict line, exit codes, the
host reference loop (BASELINE.json configs[0]) cross-checked against the oracle, and the pod
manifest's envelope against the reference's Step 9 manifest."""
import os
import re
import subprocess

import pytest
import yaml

import oracle_lib as O
from b200v_pkg import pkg, ROOT

SEED = 20261106


@pytest.fixture(scope="module")
def exe():
    if not os.path.exists(pkg.LAUNCHER_PATH):
        pkg.build()
    return pkg.LAUNCHER_PATH


def run(exe, *args, env=None):
    e = dict(os.environ)
    if env:
        e.update(env)
    return subprocess.run([exe, *args], capture_output=True, text=True, timeout=300, env=e)


def last_line(r):
    return r.stdout.strip().splitlines()[-1]


def _gpu_visible():
    try:
        return pkg.device_count() > 0
    except Exception:
        return False


no_gpu = pytest.mark.skipif(_gpu_visible(), reason="checks the no-GPU verdict")


@no_gpu
def test_default_mode_fails_loudly_without_gpu(exe):
    """No device => Test FAILED and a non-zero exit (restartPolicy: OnFailure acts on it); never a CPU pass."""
    r = run(exe)
    assert r.returncode == 2
    assert last_line(r) == "Test FAILED"
    assert "Test PASSED" not in r.stdout
    assert "no usable CUDA device" in r.stdout


@no_gpu
@pytest.mark.parametrize("mode", ["bench", "node", "p2p", "burnin", "full", "list"])
def test_every_gpu_mode_fails_without_gpu(exe, mode):
    r = run(exe, "--mode", mode, "--seconds", "1")
    assert r.returncode != 0
    assert "Test PASSED" not in r.stdout
    assert last_line(r) == "Test FAILED"


def test_usage_errors(exe):
    for bad in (["--bogus"], ["--mode"], ["--dtype", "fp64"], ["--mode", "nonsense"], ["--iters", "0"]):
        r = run(exe, *bad)
        assert r.returncode == 64, bad
        assert last_line(r) == "Test FAILED"


@pytest.mark.parametrize("n", [50_000, 1, 65_537])
def test_host_reference_matches_oracle(exe, n):
    """configs[0]: the launcher's own host loop (hostcheck.cpp) and the oracle are written
    separately from the same spec; their checksums over the same inputs must agree."""
    r = run(exe, "--mode", "host-reference", "--n", str(n))
    assert r.returncode == 0 and last_line(r) == "Test PASSED"
    m = re.search(r"add_checksum=([0-9a-f]{16}):([0-9a-f]{16}) triad_checksum=([0-9a-f]{16}):([0-9a-f]{16})", r.stdout)
    assert m
    a = O.fill("f32", 0, n, SEED, 0)
    b = O.fill("f32", 0, n, SEED, 1)
    add = O.checksum(O.op("add", "f32", n, x=a, y=b))
    triad = O.checksum(O.op("triad", "f32", n, s=3.0, x=a, y=b))
    assert (int(m.group(1), 16), int(m.group(2), 16)) == add
    assert (int(m.group(3), 16), int(m.group(4), 16)) == triad
    assert "no GPU is used" in r.stdout


def test_host_reference_fault_injection_turns_the_verdict(exe):
    r = run(exe, "--mode", "host-reference", "--inject-fault")
    assert r.returncode == 1 and last_line(r) == "Test FAILED"


def test_launcher_is_self_contained(exe):
    """The image is a base plus this ONE file (reference README.md:312-314's slot): beyond glibc the
    binary may need nothing the image would have to install — no libcudart, libstdc++, libgomp,
    libb200v or oracle — because only driver libraries arrive from the container runtime."""
    ldd = subprocess.run(["ldd", exe], capture_output=True, text=True).stdout
    needed = {ln.split()[0] for ln in ldd.strip().splitlines()}
    allowed = re.compile(r"^(linux-vdso\.so\.1|libc\.so\.6|libm\.so\.6|libdl\.so\.2|libpthread\.so\.0|librt\.so\.1|/lib64/ld-linux-x86-64\.so\.2)$")
    extra = sorted(n for n in needed if not allowed.match(n))
    assert not extra, f"gpu-validate needs libraries a bare glibc image does not have: {extra}"
    archs = set(re.findall(r"sm_\d+a?", subprocess.run(["cuobjdump", "-lelf", exe], capture_output=True, text=True).stdout))
    assert archs == {"sm_100a"}


def test_dockerfile_is_base_plus_one_file(exe):
    """The image recipe installs nothing: FROM, one COPY of the built binary, ENTRYPOINT == the pod's command."""
    deploy = os.path.join(ROOT, "kubernetes-with-nvidia-gpu_b200", "deploy")
    lines = [ln.strip() for ln in open(os.path.join(deploy, "Dockerfile")) if ln.strip() and not ln.startswith("#")]
    verbs = [ln.split()[0] for ln in lines]
    assert verbs.count("FROM") == 1 and verbs.count("COPY") == 1 and "RUN" not in verbs and "ADD" not in verbs, lines
    copy = next(ln for ln in lines if ln.startswith("COPY")).split()
    src, dst = copy[1], copy[2]
    assert os.path.samefile(os.path.join(ROOT, src), exe), "the Dockerfile copies something other than the built launcher"
    entry = next(ln for ln in lines if ln.startswith("ENTRYPOINT"))
    entrypoint = yaml.safe_load(entry[len("ENTRYPOINT"):])
    assert entrypoint == [dst]
    pod = yaml.safe_load(open(os.path.join(deploy, "pod-cuda-vector-add.yaml")))
    job = yaml.safe_load(open(os.path.join(deploy, "job-node-validate.yaml")))
    assert pod["spec"]["containers"][0]["command"] == entrypoint
    assert job["spec"]["template"]["spec"]["containers"][0]["command"] == entrypoint
    tag = re.search(r"-t (\S+)", open(os.path.join(deploy, "Dockerfile")).read()).group(1)
    assert pod["spec"]["containers"][0]["image"] == tag == job["spec"]["template"]["spec"]["containers"][0]["image"]


def test_exit_code_table(exe):
    """Every library status maps to a distinct, documented pod exit code — including the two that need
    hardware this pool does not have (not sm_100 -> 3; no peer access -> 5): reference README.md:353-357's
    'Pod Cannot Access GPU' checklist is what exit 2 points at."""
    r = run(exe, "--mode", "exit-codes")
    assert r.returncode == 0
    table = {m.group(1): int(m.group(2)) for m in re.finditer(r"status \d+ (B200V_\w+)\s+-> exit (\d+)", r.stdout)}
    assert table == {"B200V_OK": 0, "B200V_ERR_INVALID": 4, "B200V_ERR_NO_DEVICE": 2, "B200V_ERR_CUDA": 4, "B200V_ERR_ARCH": 3,
                     "B200V_ERR_ALIGN": 4, "B200V_ERR_PEER": 5, "B200V_ERR_MISMATCH": 1, "B200V_ERR_NOMEM": 4}
    assert "usage error             -> exit 64" in r.stdout
    assert "Test PASSED" not in r.stdout and "Test FAILED" not in r.stdout      # a table, not a verdict
    integration = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    for name, code in table.items():
        if name in ("B200V_ERR_ARCH", "B200V_ERR_PEER", "B200V_ERR_NO_DEVICE", "B200V_ERR_MISMATCH"):
            assert re.search(rf"\b{code}\b", integration)


def test_pod_manifest_keeps_the_reference_envelope():
    """Same envelope as the reference's Step 9 pod (reference README.md:304-318); only image/command/args differ."""
    deploy = os.path.join(ROOT, "kubernetes-with-nvidia-gpu_b200", "deploy")
    pod = yaml.safe_load(open(os.path.join(deploy, "pod-cuda-vector-add.yaml")))
    ref_readme = "/root/reference/README.md"
    if os.path.exists(ref_readme):
        text = open(ref_readme).read()
        block = text[text.index("apiVersion: v1\nkind: Pod"):]
        ref = yaml.safe_load(block[:block.index("EOF")])
    else:  # the GPU box has no /root/reference: the same structure, written out
        ref = {"apiVersion": "v1", "kind": "Pod", "metadata": {"name": "cuda-vector-add"},
               "spec": {"restartPolicy": "OnFailure", "containers": [{
                   "name": "cuda-vector-add", "image": "nvidia/cuda:12.1.0-base-ubuntu22.04",
                   "command": ["/bin/sh", "-c"], "args": ["nvidia-smi"],
                   "resources": {"limits": {"nvidia.com/gpu": 1}}}]}}

    def envelope(p):
        c = dict(p["spec"]["containers"][0])
        for k in ("image", "command", "args"):
            c.pop(k)
        return {"apiVersion": p["apiVersion"], "kind": p["kind"], "metadata": p["metadata"],
                "restartPolicy": p["spec"]["restartPolicy"], "n_containers": len(p["spec"]["containers"]), "container": c}

    assert envelope(pod) == envelope(ref)
    c = pod["spec"]["containers"][0]
    assert c["command"] == ["/usr/local/bin/gpu-validate"] and c["args"] == []
    job = yaml.safe_load(open(os.path.join(deploy, "job-node-validate.yaml")))
    jc = job["spec"]["template"]["spec"]["containers"][0]
    assert jc["resources"]["limits"]["nvidia.com/gpu"] == 8 and "--mode" in jc["args"]


def test_host_loops_under_sanitizers_and_against_the_oracle():
    """hostcheck.cpp (the launcher's host reference) under AddressSanitizer + UBSan, and line by line
    against oracle/selftest: both print `op dtype mode n checksum` for the same seeded cases, and the two
    independent implementations must print the same lines."""
    pkg_dir = os.path.join(ROOT, "kubernetes-with-nvidia-gpu_b200")
    r = subprocess.run(["make", "-C", pkg_dir, "hostcheck-selftest"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    r = subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "selftest"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    env = dict(os.environ, OMP_NUM_THREADS="3", ASAN_OPTIONS="detect_leaks=1")
    plain = subprocess.run([os.path.join(pkg_dir, "hostcheck_selftest")], capture_output=True, text=True, env=env, timeout=300)
    san = subprocess.run([os.path.join(pkg_dir, "hostcheck_selftest_san")], capture_output=True, text=True, env=env, timeout=600)
    assert plain.returncode == 0 and san.returncode == 0, san.stderr[-3000:]
    assert "ERROR" not in san.stderr and "runtime error" not in san.stderr, san.stderr[-3000:]
    assert plain.stdout == san.stdout and plain.stdout.strip().endswith("hostcheck selftest ok")
    orc = subprocess.run([os.path.join(ROOT, "oracle", "selftest")], capture_output=True, text=True, env=env, timeout=300)
    assert orc.returncode == 0
    ours = {ln for ln in plain.stdout.splitlines() if " mode=" in ln}
    theirs = {ln for ln in orc.stdout.splitlines() if " mode=" in ln}
    assert len(ours) == 2 * 3 * 6 * 4 - 0 or len(ours) > 100
    assert ours == theirs
