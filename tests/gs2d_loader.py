"""Import helper: the package name starts with a digit, so it cannot be written in an import statement."""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
gs2d = importlib.import_module("2dgaussiansplatting_b200")
