#!/usr/bin/env python3
"""Writes profiles/r02_sass_stream_add_f32.txt: the SASS of the shipped fp32 vectorAdd kernel at the
BASELINE size (csrc/tuning.h ShapeLarge2), from the built libb200v.so.  tests/test_sass_cpu.py compares
the live library against it, so regenerate it whenever that kernel changes."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "kubernetes-with-nvidia-gpu_b200", "libb200v.so")
tuning = open(os.path.join(ROOT, "kubernetes-with-nvidia-gpu_b200", "csrc", "tuning.h")).read()
m = re.search(r"struct ShapeLarge2 \{ static constexpr int unroll = (\d+), block = (\d+); \};", tuning)
want = f"void b200v::stream_kernel<0, false, 32, {m.group(1)}, {m.group(2)}, 0, 0, true>"
out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
parts = re.split(r"^\s*Function : (\S+)\s*$", out, flags=re.M)
names = subprocess.run(["c++filt"], input="\n".join(parts[1::2]), capture_output=True, text=True, check=True).stdout.strip().splitlines()
body = next(b for n, b in zip(names, parts[2::2]) if n.startswith(want))
lines = [re.sub(r"/\*[0-9a-f]+\*/", "", ln).rstrip() for ln in body.splitlines() if re.search(r"/\*[0-9a-f]{4}\*/", ln)]
res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True, check=True).stdout
mangled = next(p for p, n in zip(parts[1::2], names) if n.startswith(want))
usage = re.search(re.escape(mangled) + r":\s*\n\s*(REG:\S+ STACK:\S+ SHARED:\S+ LOCAL:\S+)", res).group(1)
count = lambda pat: sum(1 for ln in lines if re.search(pat, ln))
path = os.path.join(ROOT, "profiles", "r02_sass_stream_add_f32.txt")
with open(path, "w") as f:
    f.write(f"# {want}(void*, void const*, void const*, float, unsigned long)\n")
    f.write("# cuobjdump -sass kubernetes-with-nvidia-gpu_b200/libb200v.so  (nvcc 12.9, -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false)\n")
    f.write(f"# {usage}\n")
    f.write(f"# LDG.E.ENL2.256 x {count(r'LDG\.E\.ENL2\.256')}, STG.E.ENL2.256 x {count(r'STG\.E\.ENL2\.256')}, "
            f"PREEXIT x {count(r'PREEXIT')}, ACQBULK x {count(r'ACQBULK')}, FADD x {count(r'FADD')}, FFMA x {count(r'FFMA')}\n")
    for ln in lines:
        f.write(ln.strip() + "\n")
print(path, len(lines), "instructions;", usage)
