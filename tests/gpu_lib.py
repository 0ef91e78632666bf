"""Test-side access to the product package (hyphenated directory name -> importlib)."""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

rs = importlib.import_module("gpu-rscode_b200")
