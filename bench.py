#!/usr/bin/env python
"""bench.py -- headline benchmark: HMC gradient evaluations per second on Bayesian logistic
regression (BASELINE.json configs[1]: N=1M observations, D=100, 4096 chains per GPU).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A step is one HMC transition (L=16 leapfrog steps) of all 4096 chains of this rank: 16 batched
value+gradient passes over the 800 MB observation matrix, plus momentum refresh, leapfrog updates,
Metropolis accept and dual-averaging. Chains shard across ranks with no data-path collective
(weak scaling: 4096 chains per GPU). One JSON line on stdout (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_OBS, D, CHAINS_PER_GPU, L = 1_000_000, 100, 4096, 16
ALG_BYTES_PER_GRAD_EVAL = 8 * N_OBS * D + N_OBS + 16 * D  # X once + y + state
ALG_FLOPS_PER_GRAD_EVAL = 4 * N_OBS * D                   # eta and X^T r contractions (FP64)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured"
    return 6650.0, "fallback"


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of one k_dense_glm launch, from the committed
    `ncu --set full` capture of this same command (profiles/ncu_dense_r1_v11.txt); None if absent."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "ncu_dense_r1_v11.txt")
    try:
        tot = 0.0
        for line in open(path):
            if line.startswith("dram__bytes_read.sum") or line.startswith("dram__bytes_write.sum"):
                unit = line[line.index("[") + 1:line.index("]")]
                val = float(line.split("=")[1])
                tot += val * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}[unit]
        return tot or None
    except (OSError, ValueError, KeyError):
        return None


def fp64_peak():
    p = os.path.join(ROOT, "profiles", "fp64_peaks_r1.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return max(j.get("dfma_tflops", 0.0), j.get("dmma_m8n8k4_tflops", 0.0)), "measured (tools/fp64_peaks.cu)"
    return 37.0, "fallback"


class ClockSampler(threading.Thread):
    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._stop_evt = index, [], threading.Event()

    def run(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        while not self._stop_evt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=3)
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in self.rows if len(r) > 2 + i)]
        mx = [int(r[1]) for r in self.rows if r[1].isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons}


def run_reference(args, rank, world):
    """The reference's own CPU path for this workload: the finite-difference force of
    hmc.rs:304-329 (2*d full model scorings per gradient), restated in oracle/ (the Rust crate
    cannot be compiled in this image). One chain per host thread."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    import oracle
    from fugue_b200 import models
    threads = os.cpu_count() or 1
    m, _ = models.logistic_regression(N=N_OBS, D=D)
    O = oracle.OracleProgram(m)
    rng = np.random.default_rng(0)
    n_coord = 2  # coordinates probed per step: 2*n_coord of the 2*D scorings of one FD force
    q = rng.normal(0, 0.1, (D, threads))

    def step():
        for k in range(n_coord):
            for sgn in (1.0, -1.0):
                qq = q.copy()
                qq[k] += sgn * 1e-5
                O.logp(qq, nthreads=threads)

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = (time.perf_counter() - t0) / max(1, args.steps)
    value = threads * (n_coord / D) / dt  # gradient evaluations per second, all threads
    sample = f"{threads} chains (1/thread) x {2 * n_coord} of the {2 * D} full-model scorings of one central-difference force per step; N={N_OBS}, D={D}"
    line = {"impl": "reference", "metric": "grad-evals/sec", "value": value, "unit": "grad-evals/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": {"workload": f"bayesian logistic regression N={N_OBS} D={D}, HMC force by central differences (reference hmc.rs:304-329)"},
            "cpu_baseline": {"value": value, "unit": "grad-evals/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "grad-evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def cpu_baseline():
    """Oracle (CPU port, analytic gradient -- cheaper than the reference's own FD force) on the
    host cores, bounded sample of the same workload."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    import oracle
    from fugue_b200 import models
    threads = os.cpu_count() or 1
    m, _ = models.logistic_regression(N=N_OBS, D=D)
    O = oracle.OracleProgram(m)
    rng = np.random.default_rng(0)
    q = rng.normal(0, 0.1, (D, threads))
    O.logp(q[:, :threads], grad=True, nthreads=threads)
    reps, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < 10.0:
        O.logp(q, grad=True, nthreads=threads)
        reps += 1
    dt = time.perf_counter() - t0
    return {"value": reps * threads / dt, "unit": "grad-evals/s", "cores": threads, "kind": "port",
            "sample": f"{reps} passes x {threads} chains (1/thread) of the analytic value+gradient oracle at N={N_OBS}, D={D} (the reference's own finite-difference force costs {2 * D}x a scoring per gradient: see --impl reference)"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--chains", type=int, default=CHAINS_PER_GPU)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch
    import torch.distributed as dist
    from fugue_b200 import api, models
    torch.cuda.set_device(local)
    api.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    Cn = args.chains
    m, _ = models.logistic_regression(N=N_OBS, D=D)          # same synthetic data on every rank (seed 42)
    P = api.Program(m)                                        # X, y resident in HBM from here on
    ext = torch.cuda.ExternalStream(P.stream)
    S = api.HmcSession(P, n_chains=Cn, n_warmup=10**9, config=api.HMCConfig(n_leapfrog=L), seed=1234, first_chain=rank * Cn)
    S.sync()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput: K transitions, CUDA events on the launching stream ----
    for _ in range(args.warmup):
        S.step(1)
    S.sync()
    P.kernel_timing(True)
    launches0, ge0 = P.launch_count, S.grad_evals()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    with torch.cuda.stream(ext):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        S.step(args.steps)
        e1.record()
        e1.synchronize()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop()
    barrier()
    kern_ms, kern_launches = P.kernel_time()
    P.kernel_timing(False)
    launches = P.launch_count - launches0
    grad_evals = S.grad_evals() - ge0
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = world * Cn * grad_evals / (ms_max * 1e-3)

    # ---- end to end through the C ABI with HOST buffers: per step the caller supplies the
    # step's randomness (the reference's `rng: &mut R` argument) from pinned host memory and reads
    # the new positions and log-joints back (what hmc_chain returns per draw) ----
    d = P.d_cont
    z_pin = torch.randn(d, Cn, dtype=torch.float64).pin_memory()
    u_pin = torch.rand(Cn, dtype=torch.float64).pin_memory()
    q_pin = torch.zeros(d, Cn, dtype=torch.float64).pin_memory()
    lp_pin = torch.zeros(Cn, dtype=torch.float64).pin_memory()
    zn, un, qn, lpn = z_pin.numpy(), u_pin.numpy(), q_pin.numpy(), lp_pin.numpy()
    import ctypes as C

    def e2e_step():
        S.set_explicit_randomness(zn, un)
        S.step(1)
        api.check(api.lib.fc_hmc_get(S.h, C.c_void_p(qn.ctypes.data), C.c_void_p(lpn.ctypes.data), None))

    e2e_step()
    barrier()
    ge1 = S.grad_evals()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = world * Cn * (S.grad_evals() - ge1) / float(te.item())

    if rank == 0:
        hbm, how = peaks()
        f64pk, f64how = fp64_peak()
        per_launch_ms = kern_ms / max(1, kern_launches)
        alg_bytes = ALG_BYTES_PER_GRAD_EVAL * Cn  # one launch = one gradient evaluation of Cn chains
        achieved = alg_bytes / (per_launch_ms * 1e-3) / 1e9
        tfl = ALG_FLOPS_PER_GRAD_EVAL * Cn / (per_launch_ms * 1e-3) / 1e12
        line = {
            "metric": "grad-evals/sec", "value": value, "unit": "grad-evals/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": f"BASELINE configs[1]: bayesian logistic regression N={N_OBS} D={D}, {Cn} HMC chains per GPU, L={L} leapfrog steps per transition",
                       "chains_per_gpu": Cn, "parallelism": f"chains sharded x{world}, no data-path collective",
                       "l2": "inputs larger than L2 (X = 800 MB streamed every gradient pass)"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm, "traffic": ncu_traffic(), "traffic_unit": "bytes per launch (ncu dram read+write, profiles/ncu_dense_r1_v11.txt)",
                         "kernel": "k_dense_glm", "launch_ms": per_launch_ms, "launches_timed": int(kern_launches), "peak_source": how,
                         "note": "algorithmic bytes = 801,001,600 B per chain-gradient (SURVEY 8d) x chains per launch; X is shared by the 128 chains of a CTA, so achieved exceeds the streaming roof and the FP64 pipe binds: see roofline_fp64"},
            "roofline_fp64": {"bound": "fp64 tensor/FMA pipe (DMMA)", "achieved": tfl, "peak": f64pk, "unit": "TFLOP/s", "frac": tfl / f64pk, "peak_source": f64how},
            "e2e": {"value": e2e_value, "unit": "grad-evals/s", "h2d_bytes_per_step": int((d + 1) * Cn * 8), "d2h_bytes_per_step": int((d + 1) * Cn * 8)},
            "gpu_launches": int(launches), "clocks": clocks,
            "kernel_share_of_step": kern_ms / ms,
        }
        if not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline()
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
