"""CPU tier: what nvcc/ptxas actually emitted for sm_100a ("nvcc -arch=sm_100a compile + SASS
grep for LDG.E...256").  A compiler upgrade that quietly drops the Blackwell-only 256-bit global accesses,
loses the programmatic-dependent-launch handshake or starts spilling would still pass every parity test and
only show up as a slower bench; these checks read the shipped library's SASS and resource usage instead."""
import json
import os
import re
import subprocess

import pytest

from b200v_pkg import pkg, ROOT

CSRC = os.path.join(ROOT, "kubernetes-with-nvidia-gpu_b200", "csrc")


def demangle(names):
    r = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True, check=True)
    return r.stdout.strip().splitlines()


@pytest.fixture(scope="module")
def sass():
    """{demangled kernel name: SASS text} of libb200v.so"""
    assert os.path.exists(pkg.LIB_PATH), "libb200v.so has not been built"
    out = subprocess.run(["cuobjdump", "-sass", pkg.LIB_PATH], capture_output=True, text=True, check=True).stdout
    parts = re.split(r"^\s*Function : (\S+)\s*$", out, flags=re.M)
    mangled, bodies = parts[1::2], parts[2::2]
    return dict(zip(demangle(mangled), bodies))


@pytest.fixture(scope="module")
def resources():
    """{demangled kernel name: {REG, STACK, LOCAL, SHARED}}"""
    out = subprocess.run(["cuobjdump", "-res-usage", pkg.LIB_PATH], capture_output=True, text=True, check=True).stdout
    found = re.findall(r"Function (\S+):\s*\n\s*REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", out)
    names = demangle([f[0] for f in found])
    return {n: {"REG": int(f[1]), "STACK": int(f[2]), "SHARED": int(f[3]), "LOCAL": int(f[4])} for n, f in zip(names, found)}


def stream_kernels(d, vecb=None):
    """stream_kernel<OP, BF16, VECB, UNROLL, BLOCK, LD, ST, PDL> instantiations (optionally one vector width)"""
    out = {}
    for name, body in d.items():
        m = re.match(r"void b200v::stream_kernel<(\d), (true|false), (\d+), (\d+), (\d+), (\d), (\d), (true|false)>", name)
        if m and (vecb is None or int(m.group(3)) == vecb):
            out[name] = (tuple(m.groups()), body)
    return out


def test_only_sm100a_code_is_shipped():
    for path in (pkg.LIB_PATH, pkg.LAUNCHER_PATH):
        elf = subprocess.run(["cuobjdump", "-lelf", path], capture_output=True, text=True).stdout
        ptx = subprocess.run(["cuobjdump", "-lptx", path], capture_output=True, text=True).stdout
        assert set(re.findall(r"sm_\d+a?", elf)) == {"sm_100a"}
        assert set(re.findall(r"sm_\d+a?", ptx)) <= {"sm_100a"}      # any embedded PTX is arch-specific too: no JIT to other GPUs


def test_every_dispatched_shape_is_in_the_library(sass):
    """csrc/tuning.h names the launch shapes; each one must exist for all four ops and both dtypes."""
    tuning = open(os.path.join(CSRC, "tuning.h")).read()
    shapes = {m.group(1): (int(m.group(2)), int(m.group(3)))
              for m in re.finditer(r"struct (Shape\w+) \{ static constexpr int unroll = (\d+), block = (\d+); \};", tuning)}
    assert {"ShapeLarge2", "ShapeMedium", "ShapeSmall", "ShapeLarge1024"} <= set(shapes)
    have = {(g[0], g[1], int(g[2]), int(g[3]), int(g[4])) for g, _ in stream_kernels(sass).values()}
    for op in "0123":
        for bf in ("true", "false"):
            for shape in ("ShapeMedium", "ShapeSmall"):
                u, b = shapes[shape]
                assert (op, bf, 32, u, b) in have, (op, bf, shape)
    u, b = shapes["ShapeLarge2"]
    assert ("0", "false", 32, u, b) in have and ("3", "false", 32, u, b) in have       # fp32 add / triad, >= 512 MiB per array


def test_256_bit_global_access_in_every_vector_kernel(sass):
    ks = stream_kernels(sass, vecb=32)
    assert len(ks) == 24          # 4 ops x 2 dtypes x 3 size classes
    for name, ((op, bf, vecb, unroll, block, ld, st, pdl), body) in ks.items():
        loads = len(re.findall(r"LDG\.E\.ENL2\.256", body))
        stores = len(re.findall(r"STG\.E\.ENL2\.256", body))
        arrays_in = 2 if op in ("0", "3") else 1
        # the tile body: UNROLL loads per input array and UNROLL stores, all 256-bit (ptxas may duplicate the loop)
        assert loads >= int(unroll) * arrays_in and stores >= int(unroll), (name, loads, stores)
        assert loads % (int(unroll) * arrays_in) == 0 and stores % int(unroll) == 0, (name, loads, stores)
    for name, (_, body) in stream_kernels(sass, vecb=16).items():
        assert "LDG.E.128" in body and "STG.E.128" in body, name
    for k in ("b200v::peer_pull_kernel", "b200v::peer_push_kernel"):
        body = next(b for n, b in sass.items() if n.startswith(k))
        assert len(re.findall(r"LDG\.E\.ENL2\.256", body)) >= 4 and len(re.findall(r"STG\.E\.ENL2\.256", body)) >= 4, k


def test_programmatic_dependent_launch_handshake(sass):
    """griddepcontrol.launch_dependents -> PREEXIT, griddepcontrol.wait -> ACQBULK, before any global access."""
    for name, (_, body) in stream_kernels(sass).items():
        code = [ln for ln in body.splitlines() if re.search(r"/\*[0-9a-f]{4}\*/", ln)]
        first = {key: next((i for i, ln in enumerate(code) if re.search(pat, ln)), None)
                 for key, pat in (("PREEXIT", r"\bPREEXIT\b"), ("ACQBULK", r"\bACQBULK\b"), ("mem", r"\b(LDG|STG)\."))}
        assert first["PREEXIT"] is not None and first["ACQBULK"] is not None, name
        assert first["PREEXIT"] < first["ACQBULK"] < first["mem"], (name, first)


def test_fp_contract_discipline(sass):
    """Bit-exactness: only triad may contain an FMA (the explicit __fmaf_rn); add must be FADD, scale FMUL;
    nothing may use the flush-to-zero forms."""
    for name, ((op, bf, *_), body) in stream_kernels(sass).items():
        # (F2I.FTZ / I2F appear in ptxas's integer-division preamble; the data path is FADD / FMUL / FFMA)
        assert not re.search(r"\b(FADD|FMUL|FFMA)(\.[A-Z0-9]+)*\.FTZ", body), name
        has_ffma = bool(re.search(r"\bFFMA\b", body))
        if op == "3":
            assert has_ffma, name
        else:
            assert not has_ffma, name
        if op == "0":
            assert re.search(r"\bFADD\b", body), name
        if op == "2":
            assert re.search(r"\bFMUL\b", body), name


def test_no_spills_and_expected_registers(resources):
    ks = {n: r for n, r in resources.items() if "b200v::" in n}
    assert len(ks) >= 50
    for name, r in ks.items():
        assert r["STACK"] == 0 and r["LOCAL"] == 0, f"{name} spills: {r}"
    big = next(r for n, r in ks.items() if n.startswith("void b200v::stream_kernel<0, false, 32, 4, 512,"))
    # 4 x 2 x 256-bit loads in flight = 64 data registers + addressing; 128 regs x 512 threads would still fit one CTA per SM
    assert 72 <= big["REG"] <= 128, big


def test_static_scratch_array_is_in_the_image():
    """b200v_checksum_static's scratch: 256 + 1024 x 16 bytes of __device__ memory in the library image."""
    out = subprocess.run(["cuobjdump", "-res-usage", pkg.LIB_PATH], capture_output=True, text=True, check=True).stdout
    assert re.search(r"GLOBAL:16640\b", out)


def test_committed_evidence_names_the_shipped_kernel(sass):
    """profiles/traffic.json (the roofline.traffic source) and the committed SASS listing must be of the shape
    csrc/tuning.h ships for fp32 add at 2^28 elements."""
    tuning = open(os.path.join(CSRC, "tuning.h")).read()
    m = re.search(r"struct ShapeLarge2 \{ static constexpr int unroll = (\d+), block = (\d+); \};", tuning)
    unroll, block = m.group(1), m.group(2)
    src = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["add_f32_2p28"]["source"]
    assert f"stream_kernel<add,f32,32,{unroll},{block},PDL>" in src.replace(" ", ""), src
    listing = os.path.join(ROOT, "profiles", "r02_sass_stream_add_f32.txt")
    assert os.path.exists(listing), "run tools/dump_sass.py"
    text = open(listing).read()
    assert f"stream_kernel<0, false, 32, {unroll}, {block}, 0, 0, true>" in text
    name = next(n for n in sass if n.startswith(f"void b200v::stream_kernel<0, false, 32, {unroll}, {block}, 0, 0, true>"))
    live = [re.sub(r"/\*[0-9a-f]+\*/", "", ln).strip() for ln in sass[name].splitlines() if re.search(r"/\*[0-9a-f]{4}\*/", ln)]
    kept = [ln.strip() for ln in text.splitlines() if ln.strip() and not ln.startswith("#")]
    ops = lambda lines: [re.split(r"\s+", ln.split(";")[0].strip())[0] for ln in lines if ln]
    # same instruction stream (opcodes in order) as the committed listing: regenerate the listing when the kernel changes
    assert ops(live) == ops(kept), "profiles/r02_sass_stream_add_f32.txt is stale: run tools/dump_sass.py"
