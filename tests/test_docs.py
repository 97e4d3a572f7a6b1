"""Docs stay honest: every evidence file the documents cite exists, and every ABI symbol the
integration guide names is declared in the header."""
import glob
import os
import re

from b200v_pkg import ROOT


def read(name):
    with open(os.path.join(ROOT, name)) as f:
        return f.read()


def test_cited_profile_files_exist():
    cited = set()
    for doc in ("DESIGN.md", "BASELINE.md", "INTEGRATION.md", os.path.join("profiles", "README.md"),
                os.path.join("kubernetes-with-nvidia-gpu_b200", "csrc", "tuning.h"),
                os.path.join("kubernetes-with-nvidia-gpu_b200", "csrc", "kernels.cu")):
        text = read(doc)
        cited |= set(re.findall(r"(?<![A-Za-z0-9_])(?:profiles/)?(r\d\d_[A-Za-z0-9_.{},*-]+|traffic\.json)", text))
    assert len(cited) > 15
    missing = []
    for c in sorted(cited):
        c = c.rstrip(".,")
        names = [c]
        while any("{" in n for n in names):          # expand every {a,b,c} group
            out = []
            for n in names:
                m = re.search(r"\{([^{}]*)\}", n)
                out += [n[:m.start()] + alt + n[m.end():] for alt in m.group(1).split(",")] if m else [n.replace("{", "")]
            names = out
        for n in names:
            pattern = os.path.join(ROOT, "profiles", n if "." in os.path.basename(n) or "*" in n else n + "*")
            if not glob.glob(pattern):
                missing.append(n)
    assert not missing, f"cited but not in profiles/: {missing}"


def test_integration_guide_names_real_symbols():
    header = read(os.path.join("include", "b200v.h"))
    declared = set(re.findall(r"\b(b200v_\w+)\s*\(", header))
    for doc in ("INTEGRATION.md", "DESIGN.md"):
        for sym in set(re.findall(r"\b(b200v_[a-z0-9_]+)\b", read(doc))):
            if sym in ("b200v_pkg",):
                continue
            assert sym in declared or sym + "_get" in declared or any(d.startswith(sym) for d in declared), (doc, sym)
