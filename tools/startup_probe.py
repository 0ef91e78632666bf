#!/usr/bin/env python3
"""Where a cold process spends its start-up (GPU box): cuInit, primary context, library load, first context of
the engine, first launch.  One line per phase; run it in a fresh process each time.
    python tools/startup_probe.py [--k 4 --p 2]"""
import argparse
import ctypes as C
import os
import sys
import time

ap = argparse.ArgumentParser()
ap.add_argument("--k", type=int, default=4)
ap.add_argument("--p", type=int, default=2)
a = ap.parse_args()
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
t = [time.perf_counter()]
names = []


def lap(name):
    t.append(time.perf_counter())
    names.append(name)


cu = C.CDLL("libcuda.so.1")
lap("dlopen libcuda")
rc = cu.cuInit(0)
lap("cuInit rc=%d" % rc)
rt = C.CDLL(os.path.join(ROOT, "gpu-rscode_b200", "librsgpu.so"))
lap("dlopen librsgpu.so (%.1f MB)" % (os.path.getsize(os.path.join(ROOT, "gpu-rscode_b200", "librsgpu.so")) / 1e6))
n = C.c_int(0)
rt.rsgpu_device_count(C.byref(n))
lap("rsgpu_device_count -> %d" % n.value)
ctx = C.c_void_p()
rt.rsgpu_create.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int]
rc = rt.rsgpu_create(C.byref(ctx), 0, a.k, a.p)
lap("rsgpu_create rc=%d (primary context + module load + set-up kernels)" % rc)
ctx2 = C.c_void_p()
rc = rt.rsgpu_create(C.byref(ctx2), 0, a.k, a.p)
lap("second rsgpu_create rc=%d" % rc)
for nm, a0, a1 in zip(names, t, t[1:]):
    print("%8.1f ms  %s" % ((a1 - a0) * 1e3, nm))
print("%8.1f ms  total" % ((t[-1] - t[0]) * 1e3))
