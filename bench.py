#!/usr/bin/env python3
"""bench.py -- headline benchmark of the RS GF(2^8) coding path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c3|c4]

One "step" = one pass of the hot path (the GF(2^8) generator-matrix multiply) over one stripe batch
of synthetic random fragments.  Default workload = BASELINE.json configs[1]: RS(10,14) encode of
4,194,304,000 input bytes (10 fragments x 400 MiB) resident in HBM, per GPU (weak scaling: stripes
shard across GPUs with no collective on the data path).  Inputs (4.2 GB) are far larger than L2
(126 MB), so every step streams from HBM without an explicit flush.

Prints ONE JSON line (rank 0).  `value` = input GB/s (1 GB = 1e9 B) over all GPUs with inputs already
in HBM; `e2e` = the same through the host-buffer C-ABI call (rsgpu_encode_host: pinned host memory ->
H2D -> kernel -> D2H, all inside the timed region); `roofline` = algorithmic bytes (k+p)*L per launch
over the measured kernel time, against the measured HBM copy peak; `cpu_baseline` = the reference's own
cpu-rs.c matrix_mul (oracle/_ref, else our C port) on this box's host cores.  Beside the headline the same
line carries `decode` (BASELINE configs[2]: GPU Gauss-Jordan inverse + reconstruct of 4 erased natives per
step) and `c4` (configs[3] stripe shape, RS(16,20)), each with its own kernel time and roofline fraction.

Self-check (after the timed regions, never inside them): every output row of what was just timed is compared
with the oracle's matrix_mul on the head, the tail and 64 random 64 KiB windows, plus a full-buffer
property; with --gpus N >= 2 rank 0 also stores parity rows into a peer GPU (rsgpu_encode_rows) and runs the
RS binary with two GPU workers on a tmpfs file, both against the oracle.  The oracle is only the checker here.
"""
import argparse
import ctypes
import importlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

MIB = 1 << 20
WORKLOADS = {
    # name: (k, p, fragment bytes L, mode, description)
    "c2": (10, 4, 400 * MIB, "encode", "RS(10,14) encode, 10 x 400 MiB device-resident fragments per GPU (BASELINE configs[1])"),
    "c3": (10, 4, 400 * MIB, "decode", "RS(10,14) decode, 4 erased natives, GPU Gauss-Jordan inverse + reconstruct (BASELINE configs[2])"),
    "c4": (16, 4, 256 * MIB, "encode", "RS(16,20) encode, 16 x 256 MiB fragments per GPU per step (BASELINE configs[3] stripe batch, sharded)"),
    "c1": (4, 2, 16 * MIB, "encode", "RS(4,6) encode of a 64 MiB file's fragments (BASELINE configs[0])"),
}
FALLBACK_HBM_GBS = 6650.0   # /opt/skills/guides/B200_PROFILING.md, used only if MEASURED_PEAKS.json is absent


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--frag-mib", type=float, default=0, help="override fragment size (MiB) for quick checks")
    ap.add_argument("--variant", type=int, default=0)
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--e2e-slots", type=int, default=0)
    ap.add_argument("--no-legs", action="store_true", help="skip the decode (c3) / c4 legs and the multi-GPU checks")
    ap.add_argument("--sustained-s", type=float, default=2.0,
                    help="after the timed steps, keep stepping for this many seconds and report the power-capped rate (0 = skip)")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference's own matrix_mul (oracle/_ref) or our C port of it (oracle/), all host cores
# ------------------------------------------------------------------------------------------------

def _cpu_matmul_fn():
    """returns (kind, callable(A, B, C, n, p, m)) -- ctypes releases the GIL during the call."""
    from oracle_lib import ORACLE_SO, REF_SO, Oracle, Reference
    u8p = ctypes.POINTER(ctypes.c_uint8)
    if os.path.exists(REF_SO):
        ref = Reference()
        return "reference", ref.L.matrix_mul, ref.gen_encoding_matrix
    orc = Oracle()
    return "port", orc.L.orc_matrix_mul, orc.gen_encoding_matrix


class CpuArm:
    """C = V * B on `threads` host threads, each on its own contiguous stripe slice (stripes are
    independent, which is also how the GPU path shards).  Buffers are made once; run(nbytes) times one
    pass over the first nbytes of input."""

    def __init__(self, k, p, threads, max_input_bytes):
        self.kind, self.fn, gen = _cpu_matmul_fn()
        self.k, self.p, self.threads = k, p, threads
        self.V = np.ascontiguousarray(gen(p, k))
        self.m_max = max(4096, int(max_input_bytes // (k * threads)))
        rng = np.random.default_rng(12345)
        self.datas = [rng.integers(0, 256, k * self.m_max, dtype=np.uint8) for _ in range(threads)]
        self.codes = [np.zeros(p * self.m_max, np.uint8) for _ in range(threads)]

    def run(self, input_bytes):
        """returns (seconds, input bytes processed)"""
        u8p = ctypes.POINTER(ctypes.c_uint8)
        m = min(self.m_max, max(4096, int(input_bytes // (self.k * self.threads))))

        def work(t):
            self.codes[t][: self.p * m] = 0      # matrix_mul accumulates into C (cpu-rs.c:194)
            self.fn(self.V.ctypes.data_as(u8p), self.datas[t].ctypes.data_as(u8p), self.codes[t].ctypes.data_as(u8p),
                    self.p, self.k, m)

        ths = [threading.Thread(target=work, args=(t,)) for t in range(self.threads)]
        t0 = time.perf_counter()
        for th in ths:
            th.start()
        for th in ths:
            th.join()
        return time.perf_counter() - t0, self.k * m * self.threads


def cpu_throughput(k, p, total_input_bytes, threads):
    arm = CpuArm(k, p, threads, total_input_bytes)
    dt, nb = arm.run(total_input_bytes)
    return nb / dt / 1e9, arm.kind, dt, nb


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference_arm(args, rank):
    """`--impl reference`: the reference's CPU implementation of the path, all host threads, rank 0 only."""
    if rank != 0:
        return
    k, p, L, mode, desc = WORKLOADS[args.workload]
    cores = host_cores()
    arm = CpuArm(k, p, cores, 512 * MIB)
    dt, nb = arm.run(64 * MIB)                       # calibration pass (untimed in the result)
    rate = nb / dt
    # bounded sample per step: the whole --steps K --warmup W run should take about two minutes at most
    budget_s = 100.0
    sample = int(min(512 * MIB, max(16 * MIB, rate * budget_s / max(1, args.steps + args.warmup))))
    for _ in range(args.warmup):
        arm.run(sample)
    t_total, b_total, kind = 0.0, 0, arm.kind
    for _ in range(args.steps):
        dt, nb = arm.run(sample)
        t_total += dt
        b_total += nb
    val = b_total / t_total / 1e9
    line = {
        "impl": "reference",
        "metric": "%s GB/s (input bytes) RS(%d,%d)" % (mode, k, k + p),
        "value": val, "unit": "GB/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * t_total / max(1, args.steps), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": desc, "k": k, "p": p, "note": "reference CPU path (cpu-rs.c matrix_mul) on host cores; "
                   "each step is a bounded %d MiB-input sample of the workload" % (sample // MIB)},
        "cpu_baseline": {"value": val, "unit": "GB/s", "cores": cores, "kind": kind,
                         "sample": "%d MiB input per step, %d threads, each thread its own stripe slice" % (sample // MIB, cores)},
        "e2e": {"value": val, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# clocks sampler
# ------------------------------------------------------------------------------------------------

class ClockSampler:
    """nvidia-smi in loop mode for the whole run; stop(t0, t1) keeps the samples taken between the wall
    times t0 (warm-up start) and t1 (end of the timed region)."""
    Q = "timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index):
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "50"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None

    def stop(self, t0, t1):
        import datetime
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
            out, _ = self.proc.communicate()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        rows = []
        for ln in out.splitlines():
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                ts = datetime.datetime.strptime(f[0], "%Y/%m/%d %H:%M:%S.%f").timestamp()
                rows.append((ts, float(f[1]), float(f[2]), float(f[3]),
                             [nm for nm, v in zip(names, f[4:8]) if v.lower().startswith("active")]))
            except ValueError:
                continue
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sel = [r for r in rows if t0 - 0.03 <= r[0] <= t1 + 0.03]
        window = "warm-up + timed region"
        if not sel:     # region shorter than the sampling period: take the sample nearest to it
            sel = [min(rows, key=lambda r: abs(r[0] - (t0 + t1) / 2))]
            window = "nearest sample (timed region shorter than the 50 ms sampling period)"
        reasons = sorted({nm for r in sel for nm in r[4]})
        return {"sm_mhz": statistics.median(r[1] for r in sel), "sm_max_mhz": max(r[2] for r in sel), "reasons": reasons,
                "power_w_max": max(r[3] for r in sel), "samples": len(sel), "window": window}


# ------------------------------------------------------------------------------------------------
# checker: the oracle on sampled column windows (outside every timed region)
# ------------------------------------------------------------------------------------------------

def _oracle():
    """The CPU checker: the reference's own cpu-rs.c (oracle/_ref) when it was built, else our C port.
    Used AFTER the timed regions only (head, tail and random 64 KiB windows per fragment)."""
    from oracle_lib import REF_SO, Oracle, Reference
    orc = Oracle()
    mm = Reference().matrix_mul if os.path.exists(REF_SO) else orc.matrix_mul
    return orc, mm


def oracle_window_check(mm, M, src, out, L, seed, n_random=64, win=64 << 10):
    """out (rows x L, device) must equal M (rows x k) * src (k x L, device) on the head window, the tail
    window and n_random random windows, every output row, against the oracle's matrix_mul."""
    import torch
    M = np.ascontiguousarray(M, np.uint8)
    win = min(win, L)
    rng = np.random.default_rng(seed)
    starts = [0, L - win] + [int(x) for x in rng.integers(0, max(1, L - win + 1), n_random)]
    bad = None
    for c0 in starts:
        a = src[:, c0:c0 + win].contiguous().cpu().numpy()
        got = out[:, c0:c0 + win].contiguous().cpu().numpy()
        want = mm(M, a)
        if not np.array_equal(got, want):
            r, c = np.argwhere(got != want)[0]
            bad = "MISMATCH vs oracle: row %d, byte %d" % (r, c0 + c)
            break
    return {"status": bad or "ok", "windows": len(starts), "window_bytes": win, "rows_checked": int(M.shape[0]),
            "checker": "oracle matrix_mul (cpu-rs.c), head + tail + %d random windows of every output row" % n_random}


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------

def load_peak():
    try:
        mp = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(mp["hbm_gbs"]), "measured"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback"


def time_steps(torch, step_pre, step_kernel, steps):
    """Runs `steps` steps; step_pre (may be None) is the host-blocking part of a step (decode: inverse +
    tables), step_kernel the multiply.  Returns (total ms over all steps, per-launch kernel ms list)."""
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    ev0.record()
    for i in range(steps):
        if step_pre:
            step_pre()
        kev[i][0].record()
        step_kernel()
        kev[i][1].record()
    ev1.record()
    torch.cuda.synchronize()
    return ev0.elapsed_time(ev1), [a.elapsed_time(b) for a, b in kev]


def side_leg(torch, rs, name, local_rank, stream, peak, steps, seed, sustained_s=0.0, want_check=True):
    """A further BASELINE config on its own buffers, reported beside the headline: kernel time, roofline
    fraction and an oracle window check of what was just timed."""
    k, p, L, mode, desc = WORKLOADS[name]
    dev = torch.device("cuda", local_rank)
    g = torch.Generator(device=dev).manual_seed(seed)
    data = torch.randint(0, 256, (k, L), dtype=torch.uint8, device=dev, generator=g)
    leg = {"workload": desc, "k": k, "p": p, "fragment_bytes": L, "mode": mode}
    with rs.RSContext(k, p, device=local_rank) as ctx:
        parity = torch.empty((p, L), dtype=torch.uint8, device=dev)
        ctx.encode(data, L, parity, L, L, stream)
        torch.cuda.synchronize()
        if mode == "encode":
            src, out, pre = data, parity, None

            def kern():
                ctx.encode(src, L, out, L, L, stream)
            algo = (k + p) * L
        else:
            e = p
            order = list(range(k, k + e)) + list(range(e, k))
            src = torch.cat([parity, data[e:]], dim=0).contiguous()
            out = torch.empty((e, L), dtype=torch.uint8, device=dev)

            def pre():
                ctx.set_survivors(order)

            def kern():
                ctx.decode_erased(src, L, out, L, L, stream)
            algo = (k + e) * L
        for _ in range(3):
            if pre:
                pre()
            kern()
        torch.cuda.synchronize()
        l0 = ctx.launch_count()
        total_ms, kms = time_steps(torch, pre, kern, steps)
        k_ms = statistics.mean(kms)
        ach = algo / (k_ms / 1e3) / 1e9
        leg.update({"steps": steps, "ms_per_step": total_ms / steps, "kernel_ms_mean": k_ms, "kernel_ms_min": min(kms),
                    "input_gbs": k * L * steps / (total_ms / 1e3) / 1e9, "achieved": ach, "peak": peak, "unit": "GB/s",
                    "frac": ach / peak, "algorithmic_bytes_per_launch": algo,
                    "launches_per_step": (ctx.launch_count() - l0) / steps})
        try:      # DRAM bytes per launch of this leg's kernel from its ncu --set full capture (profiles/traffic.json)
            tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(name)
            if tr:
                leg["traffic"] = tr["dram_bytes_per_launch"]
                leg["traffic_source"] = tr.get("source")
        except Exception:
            pass
        if mode == "decode":
            leg["step"] = "rsgpu_set_survivors (GPU Gauss-Jordan inverse + tables, blocks) + rsgpu_decode_erased, every step"
        if sustained_s > 0:
            n_sus = max(10, int(sustained_s * 1e3 / max(1e-3, total_ms / steps)))
            sm, _ = time_steps(torch, pre, kern, n_sus)
            leg["sustained"] = {"steps": n_sus, "ms_per_step": sm / n_sus, "achieved": algo / (sm / n_sus / 1e3) / 1e9,
                                "frac": algo / (sm / n_sus / 1e3) / 1e9 / peak}
        if want_check:
            orc, mm = _oracle()
            if mode == "encode":
                chk = oracle_window_check(mm, orc.gen_encoding_matrix(p, k), src, out, L, seed)
            else:
                inv, st = orc.invert_matrix(orc.gen_total_encoding_matrix(k, p)[order])
                erased = ctx.erased()
                chk = oracle_window_check(mm, inv[erased], src, out, L, seed)
                if st != 0:
                    chk["status"] = "oracle inverse outside the reference-correct domain (status %d)" % st
                elif not np.array_equal(ctx.decode_matrix(), inv):
                    chk["status"] = "MISMATCH: GPU Gauss-Jordan inverse differs from the oracle's"
                elif not bool((out == data[:e]).all()):
                    chk["status"] = "MISMATCH: decode did not return the erased natives"
                else:
                    chk["also"] = "inverse == oracle invert_matrix; full buffers == the erased natives"
            leg["self_check"] = chk
    del data, parity, src, out
    torch.cuda.empty_cache()
    return leg


def multi_gpu_checks(torch, rs, world):
    """Rank 0, world >= 2 (SCALE runs): hardware evidence for the two paths one GPU cannot show --
    parity rows stored straight into a PEER GPU's memory (rsgpu_encode_rows over NVLink), and the file
    front end with TWO GPU workers (RS binary, RS_GPUS=2, tmpfs).  Both checked against the oracle."""
    import hashlib
    import shutil
    import tempfile
    res = {}
    orc, mm = _oracle()
    # ---- peer scatter: parity rows 0,2 local, 1,3 on GPU 1 ----
    try:
        k, p, L = 10, 4, (32 << 20) + 16 * 5 + 3
        d0, d1 = torch.device("cuda", 0), torch.device("cuda", 1)
        data = torch.randint(0, 256, (k, L), dtype=torch.uint8, device=d0)
        rows = [torch.zeros(L, dtype=torch.uint8, device=(d1 if i % 2 else d0)) for i in range(p)]
        with rs.RSContext(k, p, device=0) as ctx:
            ctx.enable_peer(1)
            ctx.encode_rows(data, L, rows, L, torch.cuda.current_stream().cuda_stream)
            torch.cuda.synchronize()
        got = torch.stack([r.to(d0) for r in rows])
        chk = oracle_window_check(mm, orc.gen_encoding_matrix(p, k), data, got, L, 77, n_random=16)
        res["peer_scatter"] = {"status": chk["status"], "api": "rsgpu_enable_peer + rsgpu_encode_rows, rows 1 and 3 in GPU 1's memory",
                               "fragment_bytes": L, "windows": chk["windows"]}
        del data, rows, got
    except Exception as ex:      # reported, not fatal: the bench line must still be printed
        res["peer_scatter"] = {"status": "ERROR: %s" % ex}
    # ---- file front end with two GPU workers ----
    tmp = tempfile.mkdtemp(prefix="rsbench_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    try:
        k, p, size = 10, 4, (320 << 20) + 12345
        rng = np.random.default_rng(2024)
        blob = rng.integers(0, 256, size, dtype=np.uint8)
        name = "bench2gpu.bin"
        blob.tofile(os.path.join(tmp, name))
        env = dict(os.environ, RS_GPUS="2", RS_TIMING="1")
        for var in ("RANK", "LOCAL_RANK", "WORLD_SIZE"):
            env.pop(var, None)
        r1 = subprocess.run([rs.RS_BIN, "-k", str(k), "-n", str(k + p), "-e", name], cwd=tmp, env=env, capture_output=True, text=True, timeout=300)
        chunk = (size + k - 1) // k
        workers = [ln for ln in r1.stdout.splitlines() if ln.startswith("Device")]
        # parity fragments against the oracle on windows (natives are plain slices of the file)
        frag_ok = r1.returncode == 0
        if frag_ok:
            padded = np.zeros(k * chunk, np.uint8)
            padded[:size] = blob
            nat = padded.reshape(k, chunk)
            V = orc.gen_encoding_matrix(p, k)
            for c0 in (0, chunk - 65536, chunk // 3):
                want = mm(V, np.ascontiguousarray(nat[:, c0:c0 + 65536]))
                for i in range(p):
                    got = np.fromfile(os.path.join(tmp, "_%d_%s" % (k + i, name)), dtype=np.uint8, count=65536, offset=c0)
                    frag_ok = frag_ok and np.array_equal(got, want[i])
        conf = os.path.join(tmp, "conf")
        order = [10, 11, 12, 13, 4, 5, 6, 7, 8, 9]
        open(conf, "w").write("".join("_%d_%s\n" % (i, name) for i in order))
        r2 = subprocess.run([rs.RS_BIN, "-d", "-i", name, "-c", "conf", "-o", "back.bin"], cwd=tmp, env=env, capture_output=True, text=True, timeout=300)
        same = r2.returncode == 0 and hashlib.sha256(open(os.path.join(tmp, "back.bin"), "rb").read()).digest() == hashlib.sha256(blob.tobytes()).digest()
        dworkers = [ln for ln in r2.stdout.splitlines() if ln.startswith("Device")]
        both = len(workers) >= 2 and len(dworkers) >= 2
        status = "ok" if (frag_ok and same and both) else "FAILED: encode rc %d frag_ok %s, decode rc %d same %s, workers %d/%d; %s %s" % (
            r1.returncode, frag_ok, r2.returncode, same, len(workers), len(dworkers), r1.stderr[-200:], r2.stderr[-200:])
        res["file_two_gpu_workers"] = {"status": status, "cli": "RS -k 10 -n 14 -e / RS -d, RS_GPUS=2, %d-byte file in tmpfs, 4 natives erased" % size,
                                       "encode_workers": workers, "decode_workers": dworkers,
                                       "slabs_per_gpu": [ln.split("]", 1)[1].strip() for ln in (r1.stderr + r2.stderr).splitlines() if "coded" in ln]}
    except Exception as ex:
        res["file_two_gpu_workers"] = {"status": "ERROR: %s" % ex}
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return res


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference_arm(args, rank)
        return

    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (the product path has no CPU fallback)")
    rs = importlib.import_module("gpu-rscode_b200")
    sharding = importlib.import_module("gpu-rscode_b200.sharding")
    if not os.path.exists(rs.LIB_PATH) and local_rank == 0:
        rs.build()          # fresh checkout: compile in-tree (nvcc, sm_100a); there is no other path
    rs.lib()

    torch.cuda.set_device(local_rank)
    sampler = ClockSampler(local_rank) if (rank == 0 and not os.environ.get('BENCH_NO_SMI')) else None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    n_gpus = world
    peak, peak_src = load_peak()

    k, p, L, mode, desc = WORKLOADS[args.workload]
    if args.frag_mib:
        L = int(args.frag_mib * MIB) // 256 * 256
        desc += " [fragment size overridden to %d bytes]" % L
    dev = torch.device("cuda", local_rank)
    stream = torch.cuda.current_stream().cuda_stream

    ctx = rs.RSContext(k, p, device=local_rank)
    if args.variant or args.threads:
        ctx.set_tuning(variant=args.variant, threads=args.threads)

    # synthetic random fragments, device resident
    g = torch.Generator(device=dev).manual_seed(0xB200 + rank)
    data = torch.randint(0, 256, (k, L), dtype=torch.uint8, device=dev, generator=g)
    order = None
    if mode == "encode":
        rows_out = p
        out = torch.empty((p, L), dtype=torch.uint8, device=dev)
        src = data
        pre = None

        def kern():
            ctx.encode(src, L, out, L, L, stream)
        algo_bytes = (k + p) * L
    else:
        # decode: survivors = 4 parities + natives 4..9 in natural conf order; timed = GJ inverse + reconstruct
        e = p
        parity = torch.empty((p, L), dtype=torch.uint8, device=dev)
        ctx.encode(data, L, parity, L, L, stream)
        torch.cuda.synchronize()
        order = list(range(k, k + e)) + list(range(e, k))
        src = torch.cat([parity, data[e:]], dim=0).contiguous()
        del parity
        out = torch.empty((e, L), dtype=torch.uint8, device=dev)
        rows_out = e

        def pre():
            ctx.set_survivors(order)          # inverse + table build (blocks briefly; part of the step)

        def kern():
            ctx.decode_erased(src, L, out, L, L, stream)
        algo_bytes = (k + e) * L

    def step():
        if pre:
            pre()
        kern()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    t_wall0 = time.time()
    for _ in range(max(3, args.warmup)):
        step()
    barrier()

    launches0 = ctx.launch_count()
    barrier()
    elapsed_ms, kernel_ms = time_steps(torch, pre, kern, args.steps)
    barrier()
    launches = ctx.launch_count() - launches0
    clocks = sampler.stop(t_wall0, time.time()) if sampler else None

    # ---- sustained leg: the kernel draws enough power that a multi-second run sits on the 1 kW cap
    # (sw_power_cap, SM clock ~1.5 GHz); report that rate next to the burst one.  Not part of `value`. ----
    sustained = None
    if args.sustained_s > 0 and mode == "encode":
        s2 = ClockSampler(local_rank) if rank == 0 else None
        n_sus = max(10, int(args.sustained_s * 1e3 / max(1e-3, elapsed_ms / args.steps)))
        tw0 = time.time()
        sus_total, _ = time_steps(torch, pre, kern, n_sus)
        sus_ms = sus_total / n_sus
        sustained = {"steps": n_sus, "ms_per_step": sus_ms, "achieved": algo_bytes / (sus_ms / 1e3) / 1e9,
                     "input_gbs": k * L / (sus_ms / 1e3) / 1e9}
        if s2:
            c2 = s2.stop(tw0 + 0.5 * args.sustained_s, time.time())     # second half: settled clocks
            sustained["clocks"] = c2

    # ---- small-stripe leg: the same bytes as a batch of stripes with 64 KiB fragments, ONE rsgpu_encode_batch
    # call per step (stripe-major [stripe][fragment][64 KiB] inside the same buffers).  Not part of `value`. ----
    batch_leg = None
    if mode == "encode" and rank == 0 and L >= (1 << 20):
        Ls = 64 << 10
        n_st = (k * L) // (k * Ls)
        flat_in, flat_out = data.view(-1), out.view(-1)

        def batch_step():
            ctx.encode_batch(flat_in, Ls, k * Ls, flat_out, Ls, p * Ls, Ls, n_st, stream)
        for _ in range(3):
            batch_step()
        nb = max(5, min(args.steps, 20))
        l0 = ctx.launch_count()
        b_total, _ = time_steps(torch, None, batch_step, nb)
        b_ms = b_total / nb
        # oracle on a sample of the stripes: stripe s is [fragment][64 KiB] at flat offset s*k*Ls
        orc, mm = _oracle()
        V = orc.gen_encoding_matrix(p, k)
        rng = np.random.default_rng(99)
        bstat = "ok"
        for s_i in [0, n_st - 1] + [int(x) for x in rng.integers(0, n_st, 14)]:
            a = flat_in[s_i * k * Ls:(s_i + 1) * k * Ls].view(k, Ls).cpu().numpy()
            got = flat_out[s_i * p * Ls:(s_i + 1) * p * Ls].view(p, Ls).cpu().numpy()
            if not np.array_equal(got, mm(V, a)):
                bstat = "MISMATCH vs oracle in stripe %d" % s_i
                break
        batch_leg = {"fragment_bytes": Ls, "stripes_per_call": n_st, "ms_per_call": b_ms,
                     "input_gbs": n_st * k * Ls / (b_ms / 1e3) / 1e9, "achieved": n_st * (k + p) * Ls / (b_ms / 1e3) / 1e9,
                     "launches_per_call": (ctx.launch_count() - l0) // nb,
                     "self_check": bstat, "self_check_detail": "16 whole stripes, all parity rows, oracle matrix_mul",
                     "api": "rsgpu_encode_batch, stripe-major batch in the same device buffers"}
        step()                      # put the single-stripe parity back for the checks below
        torch.cuda.synchronize()

    t = torch.tensor([elapsed_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    elapsed_ms = float(t.item())
    value = sharding.job_throughput_gbs(k * L, n_gpus, args.steps, elapsed_ms)

    # ---- check of what was just timed (after the timed region; the oracle is only the checker here):
    # every output row on head, tail and 64 random 64 KiB windows against the oracle's matrix_mul, then a
    # full-buffer property on top (parity row 0 = XOR of the fragments / decode returns the natives).
    # EVERY rank checks its own buffers; rank 0 reports the worst. ----
    orc, mm = _oracle()
    if mode == "encode":
        chk = oracle_window_check(mm, orc.gen_encoding_matrix(p, k), src, out, L, 1000 + rank)
        x = data[0].clone()
        for j in range(1, k):
            x ^= data[j]
        if chk["status"] == "ok" and not bool((out[0] == x).all()):
            chk["status"] = "MISMATCH: parity row 0 is not the XOR of the fragments (full buffer)"
        chk["full_buffer"] = "parity row 0 == XOR of all fragments"
        del x
    else:
        inv, st = orc.invert_matrix(orc.gen_total_encoding_matrix(k, p)[order])
        chk = oracle_window_check(mm, inv[ctx.erased()], src, out, L, 1000 + rank)
        if chk["status"] == "ok" and (st != 0 or not np.array_equal(ctx.decode_matrix(), inv)):
            chk["status"] = "MISMATCH: GPU inverse differs from the oracle's (oracle status %d)" % st
        if chk["status"] == "ok" and not bool((out == data[:rows_out]).all()):
            chk["status"] = "MISMATCH: decode did not return the erased natives (full buffer)"
        chk["full_buffer"] = "decoded rows == the erased natives; inverse == oracle invert_matrix"
    check = chk["status"]
    if world > 1:
        flag = torch.tensor([0 if check == "ok" else 1], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MAX)
        if int(flag.item()) and check == "ok":
            check = "MISMATCH on another rank"
        chk["ranks_checked"] = world

    # ---- end to end through the host-buffer C-ABI call ----
    e2e = None
    if not args.no_e2e:
        e2e_L = L
        try:
            import psutil
            avail = psutil.virtual_memory().available
        except Exception:
            avail = 64 << 30
        need = (k + rows_out) * e2e_L * n_gpus
        note = ""
        while need > 0.5 * avail and e2e_L > 16 * MIB:
            e2e_L //= 2
            need = (k + rows_out) * e2e_L * n_gpus
            note = " (fragment size halved to fit pinned host memory)"
        h_in = torch.empty((k, e2e_L), dtype=torch.uint8).pin_memory()
        h_out = torch.empty((rows_out, e2e_L), dtype=torch.uint8).pin_memory()
        h_in.copy_(src[:, :e2e_L])
        torch.cuda.synchronize()
        if mode == "decode":
            ctx.set_survivors(order)

        def e2e_step():
            if mode == "encode":
                ctx.encode_host(h_in.data_ptr(), e2e_L, h_out.data_ptr(), e2e_L, e2e_L, args.e2e_slots)
            else:
                ctx.set_survivors(order)
                ctx.decode_erased_host(h_in.data_ptr(), e2e_L, h_out.data_ptr(), e2e_L, e2e_L, args.e2e_slots)
        e2e_steps = max(1, min(args.steps, 10))
        for _ in range(2):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()          # blocks until the parity bytes are in host memory
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
        e2e = {"value": n_gpus * k * e2e_L * e2e_steps / dt / 1e9, "unit": "GB/s",
               "h2d_bytes_per_step": k * e2e_L, "d2h_bytes_per_step": rows_out * e2e_L,
               "steps": e2e_steps, "fragment_bytes": e2e_L,
               "api": "%s (pinned host buffers, H2D+kernel+D2H pipelined inside the call)%s" % ("rsgpu_encode_host" if mode == "encode" else "rsgpu_set_survivors + rsgpu_decode_erased_host", note)}
        pcie = os.path.join(ROOT, "profiles", "pcie_peaks.json")
        if os.path.exists(pcie):     # measured host-link ceiling for this many concurrent GPUs (tools/pcie_probe.py)
            try:
                pk = json.load(open(pcie)).get(str(n_gpus))
                if pk:
                    e2e["peak"] = pk["encode_input_ceiling_gbs"]
                    e2e["frac"] = e2e["value"] / pk["encode_input_ceiling_gbs"]
                    e2e["peak_source"] = pk.get("source")
                    e2e["peak_note"] = ("the probe (two monolithic 2 GiB streams per GPU) is a floor of what the box can deliver: "
                                        "the pipeline's interleaved slices measure above it at N >= 2")
            except Exception:
                pass
        if rank == 0:   # the host-path result: the oracle on windows, and the device-path result of the same input
            hchk = oracle_window_check(mm, orc.gen_encoding_matrix(p, k) if mode == "encode" else inv[ctx.erased()],
                                       h_in, h_out, e2e_L, 5, n_random=16)
            e2e["self_check"] = hchk["status"]
            if hchk["status"] != "ok" or not bool((h_out.cuda(dev) == out[:, :e2e_L]).all()):
                check = "MISMATCH(e2e)"
        del h_in, h_out

    # ---- the other BASELINE configs, on their own buffers (every rank runs them so the GPUs stay in step;
    # rank 0 reports its own figures) ----
    del data, src, out
    ctx.close()
    torch.cuda.empty_cache()
    legs = {}
    if not args.no_legs and not args.frag_mib:
        n_leg = max(5, min(args.steps, 20))
        for name in ("c3", "c4"):
            if name == args.workload:
                continue
            legs["decode" if name == "c3" else name] = side_leg(torch, rs, name, local_rank, stream, peak, n_leg, 0xC0DE + rank,
                                                                 sustained_s=(1.0 if name == "c4" and args.sustained_s > 0 else 0.0),
                                                                 want_check=(rank == 0))
    barrier()
    multi = None
    if world >= 2 and rank == 0 and not args.no_legs:
        multi = multi_gpu_checks(torch, rs, world)

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel ----
    k_ms = statistics.mean(kernel_ms)
    achieved = algo_bytes / (k_ms / 1e3) / 1e9
    if sustained:
        sustained["frac"] = sustained["achieved"] / peak
        sustained["unit"] = "GB/s"
    if batch_leg:
        batch_leg["frac"] = batch_leg["achieved"] / peak
        batch_leg["unit"] = "GB/s"
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": None, "peak_source": peak_src, "kernel": "k_gf_matmul_fat", "kernel_ms_mean": k_ms,
                "kernel_ms_min": min(kernel_ms), "kernel_ms_median": statistics.median(kernel_ms),
                "algorithmic_bytes_per_launch": algo_bytes,
                "input_gbs_kernel_only": k * L / (k_ms / 1e3) / 1e9}
    prof = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(prof):
        try:
            tr = json.load(open(prof)).get(args.workload)
            if tr:
                roofline["traffic"] = tr["dram_bytes_per_launch"]
                roofline["traffic_source"] = tr.get("source")
        except Exception:
            pass

    cpu = None
    if n_gpus == 1 and not args.no_cpu:
        cores = host_cores()
        v1, kind, dt1, nb1 = cpu_throughput(k, p, 128 * MIB, 1)
        vN, kind, dtN, nbN = cpu_throughput(k, p, 2048 * MIB if cores >= 16 else 128 * MIB * cores, cores)
        cpu = {"value": vN, "unit": "GB/s", "cores": cores, "kind": kind,
               "sample": "%d MiB of input (same k,p; encode matrix), %d threads each on its own stripe slice, %.1f s" % (nbN // MIB, cores, dtN),
               "value_1core": v1, "sample_1core": "%d MiB input, %.1f s" % (nb1 // MIB, dt1)}

    self_check = {"status": check, "detail": chk}
    if multi:
        self_check.update(multi)
    for nm, lg in legs.items():
        if lg.get("self_check", {}).get("status", "ok") != "ok":
            self_check["status"] = "MISMATCH(%s)" % nm
    line = {
        "metric": "%s GB/s (input bytes) RS(%d,%d)" % (mode, k, k + p),
        "value": value, "unit": "GB/s", "n_gpus": n_gpus, "steps": args.steps, "warmup": max(3, args.warmup),
        "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": desc, "k": k, "p": p, "fragment_bytes": L, "input_bytes_per_gpu_per_step": k * L,
                   "l2_policy": "inputs (%.1f GB) larger than L2, no flush" % (k * L / 1e9),
                   "parallelism": "stripes sharded across %d GPU(s), no collective on the data path" % n_gpus,
                   "variant": args.variant, "self_check": self_check},
        "roofline": roofline,
        "cpu_baseline": cpu,
        "e2e": e2e,
        "gpu_launches": launches * n_gpus,
        "clocks": clocks,
        "sustained": sustained,
        "batch_small_stripes": batch_leg,
    }
    line.update(legs)      # "decode": BASELINE configs[2], "c4": configs[3] stripe shape -- own roofline each
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
