"""Import helper: the product's ctypes binding lives in orb-ygz-slam_b200/ (not an importable name)."""
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "orb-ygz-slam_b200")
if PKG not in sys.path:
    sys.path.insert(0, PKG)
import ygzorb_ctypes as Y  # noqa: E402,F401
