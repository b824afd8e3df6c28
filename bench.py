#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric on its configs, one JSON line on stdout (rank 0):

  grad-evals/sec   HMC on Bayesian logistic regression (configs[1]: N=1M observations, D=100, 4096
                   chains, L=16): the headline `value` (weak scaling: 4096 chains per GPU; at N>1 the
                   config as written -- 4096 chains over N GPUs -- is reported as `strong`)
  smc              SMC particle-updates/sec on the Gaussian mixture (configs[3]: adaptive_smc, 2^24
                   particles, rejuvenation_steps=1, global systematic resampling); at --gpus N the SAME
                   2^24 population is sharded over the N ranks (strong scaling)

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

An HMC step is one transition (L=16 leapfrog steps) of all chains of this rank: 16 batched
value+gradient passes over the 800 MB observation matrix, plus momentum refresh, leapfrog updates,
Metropolis accept and dual averaging. Chains shard across ranks with no data-path collective. An SMC
"run" is the whole tempering ladder; its exchange steps run on the device over peer memory.
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
WORKLOAD = f"BASELINE configs[1]: bayesian logistic regression N={N_OBS} D={D}, {CHAINS_PER_GPU} HMC chains per GPU, L={L} leapfrog steps per transition"
SMC_PARTICLES, SMC_R, SMC_OBS = 1 << 24, 1, 80
SMC_WORKLOAD = f"BASELINE configs[3]: gaussian mixture (n_obs={SMC_OBS}) via adaptive_smc, {SMC_PARTICLES} particles, rejuvenation_steps={SMC_R}, systematic resampling, ess_threshold=0.5"
SMC_ALG_BYTES_PER_UPDATE = 488   # SURVEY 8(d), C4 at R=1: reweight 24 + resample 288 + move 176
SMC_ALG_BYTES_PER_MOVE = 176     # ... of which one rejuvenation move: 120 read + 40 write + 16


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(path, kernel=None):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from a committed ncu capture; None if absent."""
    try:
        if path.endswith(".txt") and kernel is None:  # `ncu --set full` summary of one k_dense_glm launch
            tot = 0.0
            for line in open(os.path.join(ROOT, "profiles", path)):
                if line.startswith("dram__bytes_read.sum") or line.startswith("dram__bytes_write.sum"):
                    unit = line[line.index("[") + 1:line.index("]")]
                    tot += float(line.split("=")[1]) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}[unit]
            return tot or None
        for line in open(os.path.join(ROOT, "profiles", path)):  # launch-list summary: "<kernel> ... dram_bytes/launch=X"
            if line.startswith(kernel) and "dram_bytes/launch=" in line:
                return float(line.split("dram_bytes/launch=")[1].split()[0])
    except (OSError, ValueError, KeyError, IndexError):
        pass
    return None


def fp64_peak():
    p = os.path.join(ROOT, "profiles", "fp64_peaks_r1.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return max(j.get("dfma_tflops", 0.0), j.get("dmma_m8n8k4_tflops", 0.0)), "measured (tools/fp64_peaks.cu, profiles/fp64_peaks_r1.json)"
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


# =============================================================================================
# CPU arms (the oracle: test infrastructure, executed here only as the timed baseline)
# =============================================================================================
def _oracle():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    import oracle
    return oracle


def smc_cpu_baseline(threads):
    """Oracle adaptive_smc (smc.rs:455-581 restated) on a bounded population of the same model: 1 thread
    (one generator and one adaptation object threaded through all particles, as the reference) and all cores."""
    oracle = _oracle()
    from fugue_b200 import models
    m = models.gaussian_mixture(n_obs=SMC_OBS)[0]
    O = oracle.OracleProgram(m)
    out = {}
    for label, nt, n in (("single_thread", 1, 65536), ("all_cores", threads, 32768 * max(1, min(threads, 32)))):
        t0 = time.perf_counter()
        r = O.adaptive_smc(n, rejuvenation_steps=SMC_R, seed=1, nthreads=nt, values=False)
        dt = time.perf_counter() - t0
        out[label] = {"value": n * r["steps"] / dt, "cores": nt, "particles": n, "temper_steps": int(r["steps"]), "seconds": dt}
    return {"value": out["all_cores"]["value"], "unit": "particle-updates/s", "cores": threads, "kind": "port",
            "sample": f"oracle adaptive_smc on {out['all_cores']['particles']} particles of the same model ({out['all_cores']['temper_steps']} temper steps, {out['all_cores']['seconds']:.1f} s); next_beta and the resampler run on one thread, prior draws / scoring / moves on all",
            "single_thread": out["single_thread"]}


def run_reference(args, rank, world):
    """The reference's own CPU path for this workload: the finite-difference force of
    hmc.rs:304-329 (2*d full model scorings per gradient), restated in oracle/ (the Rust crate
    cannot be compiled in this image). One chain per host thread. Plus adaptive_smc on all cores."""
    if rank != 0:
        return
    oracle = _oracle()
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
    sample = f"{threads} chains (1/thread) x {2 * n_coord} of the {2 * D} full-model scorings of one central-difference force per step (the reference's force, hmc.rs:304-329); N={N_OBS}, D={D}"
    smc = smc_cpu_baseline(threads)
    line = {"impl": "reference", "metric": "grad-evals/sec", "value": value, "unit": "grad-evals/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": {"workload": WORKLOAD, "smc_workload": SMC_WORKLOAD},
            "cpu_baseline": {"value": value, "unit": "grad-evals/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "grad-evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "smc": {"metric": "SMC particle-updates/sec", "value": smc["value"], "unit": "particle-updates/s", "cpu_baseline": smc,
                    "e2e": {"value": smc["value"], "unit": "particle-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}}
    print(json.dumps(line), flush=True)


def cpu_baseline():
    """Oracle (CPU port, analytic gradient -- cheaper than the reference's own FD force) on the
    host cores, bounded sample of the same workload; all cores and one thread."""
    oracle = _oracle()
    from fugue_b200 import models
    threads = os.cpu_count() or 1
    m, _ = models.logistic_regression(N=N_OBS, D=D)
    O = oracle.OracleProgram(m)
    rng = np.random.default_rng(0)
    q = rng.normal(0, 0.1, (D, threads))
    O.logp(q[:, :threads], grad=True, nthreads=threads)
    reps, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < 8.0:
        O.logp(q, grad=True, nthreads=threads)
        reps += 1
    dt = time.perf_counter() - t0
    r1, t1 = 0, time.perf_counter()
    while time.perf_counter() - t1 < 4.0:
        O.logp(q[:, :1], grad=True, nthreads=1)
        r1 += 1
    d1 = time.perf_counter() - t1
    return {"value": reps * threads / dt, "unit": "grad-evals/s", "cores": threads, "kind": "port",
            "sample": f"{reps} passes x {threads} chains (1/thread) of the analytic value+gradient oracle at N={N_OBS}, D={D} (the reference's own finite-difference force costs {2 * D}x a scoring per gradient: see --impl reference)",
            "single_thread": {"value": r1 / d1, "cores": 1, "sample": f"{r1} analytic value+gradient passes of one chain"}}


# =============================================================================================
# SMC block (BASELINE configs[3])
# =============================================================================================
def smc_bench(args, rank, world, local, barrier, allmax):
    import torch
    from fugue_b200 import api, models, dist as D
    n_total = args.smc_particles
    cfg = api.SMCConfig(ess_threshold=0.5, rejuvenation_steps=SMC_R)
    m = models.gaussian_mixture(n_obs=SMC_OBS)[0]
    P = api.Program(m)
    ext = torch.cuda.ExternalStream(P.stream)
    lo, hi = D.shard_range(n_total, rank, world)
    n_local = hi - lo

    def session(seed):
        if world > 1:
            return D.ShardedSmc(P, n_total, cfg, seed).S
        return api.SmcSession(P, n_total, cfg, seed)

    def ladder(S):
        steps = 0
        while True:
            steps += 1
            if S.step() >= 1.0 or steps > 500:
                return steps

    # ---- device-resident: the population is in HBM (prior draws scored) when the clock starts ----
    reps = max(3, args.steps)
    times, steps_seen, launches = [], [], 0
    for r in range(max(1, min(args.warmup, 2)) + reps):
        S = session(1000 + r)
        barrier()
        l0 = P.launch_count
        with torch.cuda.stream(ext):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            k = ladder(S)
            e1.record()
            e1.synchronize()
        ms = allmax(e0.elapsed_time(e1))
        if r >= max(1, min(args.warmup, 2)):
            times.append(ms)
            steps_seen.append(k)
            launches += P.launch_count - l0
        counters = S.counters()
        S.close()
    updates = sum(n_total * k for k in steps_seen)
    value = updates / (sum(times) * 1e-3)

    # ---- per-phase CUDA-event timing (a separate run: the phase timer synchronises after every step) ----
    api.debug_set("smc_timing", 1)
    S = session(77)
    k_t = ladder(S)
    phase, nphase = S.phase_ms()
    S.close()
    api.debug_set("smc_timing", 0)
    move_ms = phase["move"] / max(1, nphase[4])

    # ---- end to end through the reference-facing call with HOST buffers: program creation (the model's data
    # host -> device), adaptive_smc, and the SMCResult read back into pinned host memory ----
    d = P.d_cont
    v_pin = torch.zeros(d, n_local, dtype=torch.float64).pin_memory()
    w_pin = torch.zeros(n_local, dtype=torch.float64).pin_memory()
    lw_pin = torch.zeros(n_local, dtype=torch.float64).pin_memory()
    import ctypes as C
    e2e_t, e2e_updates = [], 0
    for r in range(1 + min(reps, 3)):
        barrier()
        t0 = time.perf_counter()
        P2 = api.Program(models.gaussian_mixture(n_obs=SMC_OBS)[0])
        if world > 1:
            S = D.ShardedSmc(P2, n_total, cfg, 2000 + r).S
        else:
            S = api.SmcSession(P2, n_total, cfg, 2000 + r)
        k = ladder(S)
        le, ns = C.c_double(), C.c_uint64()
        api.check(api.lib.fc_smc_get(S.h, C.c_void_p(v_pin.data_ptr()), None, C.c_void_p(w_pin.data_ptr()), C.c_void_p(lw_pin.data_ptr()), C.byref(le), C.byref(ns)))
        dt = allmax(time.perf_counter() - t0)
        S.close()
        P2.close()
        if r > 0:
            e2e_t.append(dt)
            e2e_updates += n_total * k
    e2e_value = e2e_updates / sum(e2e_t)
    hbm, how = peaks()
    move_gbs = SMC_ALG_BYTES_PER_MOVE * n_local / (move_ms * 1e-3) / 1e9
    out = {
        "metric": "SMC particle-updates/sec", "value": value, "unit": "particle-updates/s", "n_gpus": world, "scaling": "strong",
        "config": {"workload": SMC_WORKLOAD, "particles": n_total, "particles_per_gpu": n_local,
                   "parallelism": f"population sharded rank-major x{world}; global log-sum-exp, exact carry and migration on the device over peer memory" if world > 1 else "one GPU",
                   "l2": f"population larger than L2 ({n_local * 144 / 1e6:.0f} MB of particle records per GPU)"},
        "temper_steps": steps_seen, "runs": reps, "ms_per_run": sum(times) / len(times), "ms_per_temper_step": sum(times) / sum(steps_seen),
        "roofline": {"bound": "hbm", "achieved": move_gbs, "peak": hbm, "unit": "GB/s", "frac": move_gbs / hbm,
                     "traffic": ncu_traffic("launches_c4_r2_v8_summary.txt", "k_smcx_move"), "traffic_unit": "bytes per launch at 2^24 particles (ncu dram read+write, profiles/launches_c4_r2_v8_summary.txt)",
                     "kernel": "k_smcx_move", "launch_ms": move_ms, "launches_timed": int(nphase[4]), "peak_source": how,
                     "note": f"algorithmic bytes = {SMC_ALG_BYTES_PER_MOVE} B per particle and move (SURVEY 8d) x particles per launch; the kernel scores the proposal in O(1) from the per-particle mixture statistics and is issue-bound (~1300 instructions per particle: Philox, Box-Muller, 9 log): see DESIGN.md 3.2"},
        "roofline_step": {"bound": "hbm", "achieved": value / world * SMC_ALG_BYTES_PER_UPDATE / 1e9, "peak": hbm, "unit": "GB/s", "frac": value / world * SMC_ALG_BYTES_PER_UPDATE / 1e9 / hbm,
                          "note": f"whole temper step per GPU: {SMC_ALG_BYTES_PER_UPDATE} B per particle-update (SURVEY 8d, R=1)"},
        "phase_ms_per_run": phase, "temper_steps_timed_run": k_t,
        "e2e": {"value": e2e_value, "unit": "particle-updates/s", "h2d_bytes_per_step": int(SMC_OBS * 8), "d2h_bytes_per_step": int((d + 2) * n_local * 8),
                "seconds_per_run": sum(e2e_t) / len(e2e_t), "includes": "fc_program_create, prior draw + scoring of the population, the ladder, final normalisation, values + weight + log_weight to pinned host memory (per rank: its shard)"},
        "gpu_launches": int(launches), "ess_evals_per_run": counters["ess_evals"], "fused_kernels": counters["fused"], "migrated_rank0_per_run": counters["migrated"],
    }
    P.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--chains", type=int, default=CHAINS_PER_GPU)
    ap.add_argument("--smc-particles", type=int, default=SMC_PARTICLES)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-smc", action="store_true")
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

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(x):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    Cn = args.chains
    m, _ = models.logistic_regression(N=N_OBS, D=D)          # same synthetic data on every rank (seed 42)
    t0 = time.perf_counter()
    P = api.Program(m)                                        # X, y resident in HBM from here on
    torch.cuda.synchronize()
    program_create_s = time.perf_counter() - t0
    ext = torch.cuda.ExternalStream(P.stream)

    def hmc_timed(S, n_chains_rank):
        """K transitions with CUDA events on the launching stream; returns (ms max over ranks, grad evals per chain, launches, clocks, kernel ms, kernel launches)"""
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
        return allmax(ms), ms, S.grad_evals() - ge0, P.launch_count - launches0, clocks, kern_ms, kern_launches

    # ---- weak scaling (the headline): 4096 chains per GPU ----
    S = api.HmcSession(P, n_chains=Cn, n_warmup=10**9, config=api.HMCConfig(n_leapfrog=L), seed=1234, first_chain=rank * Cn)
    S.sync()
    ms_max, ms, grad_evals, launches, clocks, kern_ms, kern_launches = hmc_timed(S, Cn)
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
    e2e_value = world * Cn * (S.grad_evals() - ge1) / allmax(time.perf_counter() - t0)
    S.close()

    # ---- the config as written at N > 1: 4096 chains over the N GPUs (strong scaling) ----
    strong = None
    if world > 1:
        Cs = max(128, Cn // world)
        S2 = api.HmcSession(P, n_chains=Cs, n_warmup=10**9, config=api.HMCConfig(n_leapfrog=L), seed=1234, first_chain=rank * Cs)
        S2.sync()
        ms2_max, _, ge2, _, _, _, _ = hmc_timed(S2, Cs)
        strong = {"value": world * Cs * ge2 / (ms2_max * 1e-3), "unit": "grad-evals/s", "chains_total": world * Cs, "chains_per_gpu": Cs, "ms_per_step": ms2_max / args.steps,
                  "scaling": "strong", "note": "BASELINE configs[1] as written: 4096 chains on N GPUs; X is shared by the 128 chains of a CTA, so fewer chains per GPU means fewer CTAs per row range, not less data per CTA"}
        S2.close()

    smc = None
    if not args.no_smc:
        smc = smc_bench(args, rank, world, local, barrier, allmax)

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
            "config": {"workload": WORKLOAD, "chains_per_gpu": Cn, "parallelism": f"chains sharded x{world}, no data-path collective",
                       "l2": "inputs larger than L2 (X = 800 MB streamed every gradient pass)", "smc_workload": SMC_WORKLOAD},
            # the roof that binds: X is shared by the 128 chains of a CTA, so the two contractions run on the FP64
            # tensor path (DMMA.8x8x4) and the FP64 pipe, not HBM, limits the kernel
            "roofline": {"bound": "tensor", "achieved": tfl, "peak": f64pk, "unit": "TFLOP/s", "frac": tfl / f64pk,
                         "traffic": ncu_traffic("ncu_dense_r1_v11.txt"), "traffic_unit": "bytes per launch (ncu dram read+write, profiles/ncu_dense_r1_v11.txt)",
                         "kernel": "k_dense_glm", "launch_ms": per_launch_ms, "launches_timed": int(kern_launches), "peak_source": f64how,
                         "note": f"FP64 tensor (DMMA m8n8k4) pipe: algorithmic flops = 4 N D = {ALG_FLOPS_PER_GRAD_EVAL:.3g} per chain-gradient (SURVEY 8d) x {Cn} chains per launch; 15 of 16 passes per transition are force-only"},
            "roofline_hbm_streaming": {"bound": "hbm", "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm, "peak_source": how,
                                       "note": "SURVEY 8d's per-chain streaming definition (801,001,600 B per chain-gradient x chains per launch): above 1 because X is read once per 128-chain CTA, not once per chain; it does not bind"},
            "e2e": {"value": e2e_value, "unit": "grad-evals/s", "h2d_bytes_per_step": int((d + 1) * Cn * 8), "d2h_bytes_per_step": int((d + 1) * Cn * 8),
                    "program_create_s": program_create_s, "program_create_note": "one-off: 800 MB of X + y host -> device and lowering, outside the per-step region"},
            "gpu_launches": int(launches), "clocks": clocks,
            "kernel_share_of_step": kern_ms / ms,
        }
        if strong:
            line["strong"] = strong
        if smc:
            line["smc"] = smc
        if not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline()
            if smc:
                line["smc"]["cpu_baseline"] = smc_cpu_baseline(os.cpu_count() or 1)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
