"""Every kernel family on buffers that end exactly at unmapped memory (tests/tools/guard_probe.py): one byte read or
written out of bounds faults instead of going unnoticed inside an allocator's slack.  Runs in a child process -- an
illegal address would take the CUDA context of the test session with it."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


def test_no_kernel_touches_memory_outside_its_buffers():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    r = subprocess.run([sys.executable, os.path.join(HERE, "tools", "guard_probe.py")], capture_output=True, text=True, timeout=900)
    if r.returncode == 77:
        pytest.skip("cuda-python (cuda.bindings) is not importable here")
    assert r.returncode == 0 and "guard ok:" in r.stdout, (r.stdout[-1500:] + "\n" + r.stderr[-3000:])


def test_the_guard_catches_an_overrun():
    """Self-test of the probe: told a length 64 bytes beyond the real end of the buffers, the same call must fault."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    r = subprocess.run([sys.executable, os.path.join(HERE, "tools", "guard_probe.py"), "--selftest"], capture_output=True, text=True,
                       timeout=300)
    if r.returncode == 77:
        pytest.skip("cuda-python (cuda.bindings) is not importable here")
    assert r.returncode != 0 and "UNNOTICED" not in r.stdout, r.stdout[-1000:] + r.stderr[-1000:]
