"""Import helper: the package directory has a hyphen in its name."""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

pkg = importlib.import_module("kubernetes-with-nvidia-gpu_b200")
