"""pytest configuration: registers the `gpu` marker and puts the repo root on sys.path."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
if os.path.dirname(os.path.abspath(__file__)) not in sys.path:
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (run with -m gpu on the B200 box)")


def pytest_sessionstart(session):
    """Build the in-tree libraries once if they are missing (fresh checkout): the product library with
    nvcc for sm_100a, the oracle with gcc.  Building is not a fallback -- without librsgpu.so the GPU
    tests fail loudly."""
    import subprocess
    if not os.path.exists(os.path.join(ROOT, "gpu-rscode_b200", "librsgpu.so")) or \
            not os.path.exists(os.path.join(ROOT, "gpu-rscode_b200", "RS")):
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "gpu-rscode_b200"), "-j8"], check=False)
    if not os.path.exists(os.path.join(ROOT, "oracle", "liboracle.so")):
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle")], check=False)
