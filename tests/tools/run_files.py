#!/usr/bin/env python3
"""File-level runs on the GPU box (BASELINE configs C1 and C5-style): the RS binary against the
reference's own CPU-RS binary on the same file.  Prints one JSON line per case.

    python tests/tools/run_files.py c1                       # RS(4,6), 64 MiB file, erase natives {2,3}
    python tests/tools/run_files.py big --k 32 --p 8 --gib 2 # host-file encode/decode, all visible GPUs
"""
import argparse
import hashlib
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from rsdata import splitmix64_bytes  # noqa: E402

RS = os.path.join(ROOT, "gpu-rscode_b200", "RS")
CPU_RS = os.path.join(ROOT, "oracle", "_ref", "CPU-RS")


def run(cmd, cwd):
    t0 = time.perf_counter()
    r = subprocess.run(cmd, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    dt = time.perf_counter() - t0
    if r.returncode != 0:
        raise SystemExit(f"{cmd} failed:\n{r.stdout}")
    return dt, r.stdout


def printed_ms(out, what):
    m = re.search(r"^Total (?:GPU|CPU) %s time(?: using multiple devices)?: ([0-9.]+)ms" % what, out, re.M)
    return float(m.group(1)) if m else None


def sha_file(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for blk in iter(lambda: f.read(1 << 24), b""):
            h.update(blk)
    return h.hexdigest()


def case(k, p, size, seed, erase, with_cpu, workdir):
    n = k + p
    d_gpu = os.path.join(workdir, "gpu")
    os.makedirs(d_gpu)
    data = splitmix64_bytes(seed, size)
    with open(os.path.join(d_gpu, "f.bin"), "wb") as f:
        f.write(data.tobytes())
    del data
    order = list(range(k))
    for slot, par in zip(erase, range(k, n)):
        order[slot] = par
    conf = "".join(f"_{i}_f.bin\n" for i in order)
    res = {"k": k, "n": n, "file_bytes": size, "erased_natives": list(erase), "conf_order": order}

    t_enc, out = run([RS, "-k", str(k), "-n", str(n), "-e", "f.bin"], d_gpu)
    res["gpu_encode_wall_s"] = t_enc
    res["gpu_encode_pipeline_ms"] = printed_ms(out, "encoding")
    open(os.path.join(d_gpu, "conf"), "w").write(conf)
    t_dec, out = run([RS, "-d", "-i", "f.bin", "-c", "conf", "-o", "out.bin"], d_gpu)
    res["gpu_decode_wall_s"] = t_dec
    res["gpu_decode_pipeline_ms"] = printed_ms(out, "decoding")
    res["gpu_roundtrip_ok"] = sha_file(os.path.join(d_gpu, "out.bin")) == sha_file(os.path.join(d_gpu, "f.bin"))
    res["gpu_encode_GBps_input_wall"] = size / t_enc / 1e9
    res["gpu_decode_GBps_input_wall"] = size / t_dec / 1e9

    if with_cpu and os.path.exists(CPU_RS):
        d_cpu = os.path.join(workdir, "cpu")
        os.makedirs(d_cpu)
        shutil.copy(os.path.join(d_gpu, "f.bin"), os.path.join(d_cpu, "f.bin"))
        t, out = run([CPU_RS, "-k", str(k), "-n", str(n), "-e", "f.bin"], d_cpu)
        res["cpu_encode_wall_s"] = t
        res["cpu_encode_kernel_ms"] = printed_ms(out, "encoding")
        open(os.path.join(d_cpu, "conf"), "w").write(conf)
        t, out = run([CPU_RS, "-d", "-i", "f.bin", "-c", "conf", "-o", "out.bin"], d_cpu)
        res["cpu_decode_wall_s"] = t
        res["cpu_decode_kernel_ms"] = printed_ms(out, "decoding")
        res["fragments_bit_exact_vs_CPU_RS"] = all(
            sha_file(os.path.join(d_gpu, f"_{i}_f.bin")) == sha_file(os.path.join(d_cpu, f"_{i}_f.bin")) for i in range(n))
        res["decoded_bit_exact_vs_CPU_RS"] = sha_file(os.path.join(d_gpu, "out.bin")) == sha_file(os.path.join(d_cpu, "out.bin"))
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("what", choices=["c1", "big"])
    ap.add_argument("--k", type=int, default=32)
    ap.add_argument("--p", type=int, default=8)
    ap.add_argument("--gib", type=float, default=2.0)
    ap.add_argument("--dir", default=None, help="work directory (default: a temp dir under /dev/shm if present)")
    args = ap.parse_args()
    base = args.dir or ("/dev/shm" if os.path.isdir("/dev/shm") else None)
    work = tempfile.mkdtemp(prefix="rsfiles_", dir=base)
    try:
        if args.what == "c1":
            res = case(4, 2, 64 << 20, 0xC1, (2, 3), True, work)
            res["config"] = "C1: RS(4,6) encode+decode of a 64 MiB file, natives 2,3 erased, conf _0,_1,_4,_5"
        else:
            size = int(args.gib * (1 << 30))
            res = case(args.k, args.p, size, 0xC5, tuple(range(min(args.p, args.k))), False, work)
            res["config"] = f"RS({args.k},{args.k + args.p}) host-file encode + decode of {args.gib} GiB in {work}"
        print(json.dumps(res))
    finally:
        shutil.rmtree(work, ignore_errors=True)


if __name__ == "__main__":
    main()
