import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run through gpurun)")
    config.addinivalue_line("markers", "multigpu: needs at least 2 visible GPUs (skips itself otherwise)")


def pytest_sessionstart(session):
    """A fresh checkout has no built artefacts (they are git-ignored): build the product and the
    oracle once, in-tree, before any test needs them (nvcc cross-compiles without a GPU)."""
    import importlib
    pkg = importlib.import_module("kubernetes-with-nvidia-gpu_b200")
    need = [pkg.LIB_PATH, pkg.LAUNCHER_PATH, os.path.join(ROOT, "oracle", "liboracle.so")]
    if not all(os.path.exists(p) for p in need):
        import __graft_entry__
        __graft_entry__.build()
