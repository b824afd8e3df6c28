"""Multi-GPU run of the sharded tempered SMC (in-library peer-memory exchange), one rank per GPU:
  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/smc_sharded_check.py [particles] [parity_particles]
(1) parity: one reweight + global systematic resample on the C4 mixture model; every rank records the
    global ancestors of the slots it produced, rank 0 gathers them and the weights and compares the
    whole index array, bit for bit, with the single-GPU resampler run on the gathered weights (itself
    bit-exact vs the sequential loop: tests/test_gpu_smc.py); migrated state == state[ancestors];
(2) throughput: adaptive_smc at BASELINE configs[3] (2^24 particles over all ranks,
    rejuvenation_steps=1), particle-updates/s with the time taken as the max over ranks.
Rank 0 prints one JSON line."""
import json, os, sys, time
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

rank, ws, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
if ws > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
from fugue_b200 import api, models, dist as D

api.set_device(lr)
n_total = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 24
n_par = int(sys.argv[2]) if len(sys.argv) > 2 else n_total
m = models.gaussian_mixture()[0]
out = {"world_size": ws, "particles": n_total, "exchange": "device peer memory (cudaIpc windows)" if ws > 1 else "single"}


def gather_cols(a, dtype=np.float64):
    """rank-major concatenation along the last axis, on every rank (host channel; test harness only)"""
    a = np.ascontiguousarray(a, dtype=np.float64)
    if ws == 1:
        return a.astype(dtype)
    sizes = [int(x[0]) for x in D.all_gather_np(np.array([float(a.shape[-1])]))]
    mx = max(sizes)
    pad = np.zeros(a.shape[:-1] + (mx,))
    pad[..., :a.shape[-1]] = a
    parts = D.all_gather_np(pad)
    return np.concatenate([p[..., :s] for p, s in zip(parts, sizes)], axis=-1).astype(dtype)


# ---- (1) parity of the global resample on n_par particles (ragged shards when n_par is odd) ----
sm = D.ShardedSmc(m, n_par, api.SMCConfig(ess_threshold=0.5, rejuvenation_steps=1), seed=3)
sm.S.record_offspring(True)
sm.S.reweight(0.02)
w = sm.S.weights()[1]
rb = sm.S.result()
vb, zb = rb.values.copy(), rb.dvalues.copy()
sm.S.resample_particles(1, [0.61])
jlo, jhi, idx = sm.S.last_offspring()
ra = sm.S.result()
W = gather_cols(w)
G = gather_cols(idx.astype(np.float64), np.int64)
Vb, Va = gather_cols(vb), gather_cols(ra.values)
Zb, Za = gather_cols(zb.astype(np.float64), np.int64), gather_cols(ra.dvalues.astype(np.float64), np.int64)
mig = D.all_reduce_np(np.array([float(sm.S.counters()["migrated"])]))[0]
if rank == 0:
    ref = api.resample_indices(1, W, [0.61]).astype(np.int64)
    out["resample_bit_exact"] = bool(G.shape == ref.shape and np.array_equal(G, ref))
    out["migration_exact"] = bool(np.array_equal(Va, Vb[:, ref]) and np.array_equal(Za, Zb[:, ref]))
    out["parity_particles"] = int(n_par)
    out["parity_migrated"] = int(mig)
sm.close()
del W, G, Vb, Va, Zb, Za

# ---- (2) throughput ----
for rep in range(2):  # first pass warms up allocations / peer mappings
    sm = D.ShardedSmc(m, n_total, api.SMCConfig(ess_threshold=0.5, rejuvenation_steps=1), seed=5 + rep)
    torch.cuda.synchronize()
    if ws > 1:
        dist.barrier()
    t0 = time.perf_counter()
    steps = 0
    while sm.step() < 1.0:
        steps += 1
    steps += 1
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    tt = torch.tensor([dt], dtype=torch.float64, device="cuda")
    if ws > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    dt = float(tt.item())
    r = sm.result(values=False)
    c = sm.S.counters()
    if rank == 0:
        out.update(temper_steps=steps, seconds=dt, particle_updates_per_s=n_total * steps / dt, log_evidence=r.log_evidence,
                   ess_evals=c["ess_evals"], fused=c["fused"], migrated_rank0=c["migrated"])
    sm.close()
if rank == 0:
    print(json.dumps(out), flush=True)
if ws > 1:
    dist.barrier()
    dist.destroy_process_group()
