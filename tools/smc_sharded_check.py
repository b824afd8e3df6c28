"""Multi-GPU run of the sharded tempered SMC (fugue_b200/dist.py) over NCCL, one rank per GPU:
  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/smc_sharded_check.py [particles]
(1) parity: one reweight + global systematic resample on the C4 mixture model; the ranks' ancestors
    are gathered on rank 0 and compared, bit for bit, with the single-GPU resampler run on the
    gathered weights (itself bit-exact vs the sequential loop: tests/test_gpu_smc.py);
(2) throughput: adaptive_smc_sharded at BASELINE configs[3] (2^24 particles over all ranks,
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
m = models.gaussian_mixture()[0]
out = {"world_size": ws, "particles": n_total, "backend": "nccl" if ws > 1 else "single"}

# ---- (1) parity of the global resample ----
n_par = min(n_total, 1 << 20) + 1  # ragged shards on purpose
sm = D.ShardedSmc(m, n_par, api.SMCConfig(ess_threshold=0.5, rejuvenation_steps=1), seed=3)
lse1, _ = D._lse_pair(sm._tri(0.02))
api.check(api.lib.fc_smc_shard_reweight(sm.S.h, 0.02, lse1))
sm.beta = 0.02
w = np.zeros(sm.n)
api.check(api.lib.fc_smc_shard_weights(sm.S.h, None, api._ptr(w)))
vb = sm.result().values.copy()
zb = sm.result().dvalues.copy()
(jlo, jhi), idx = sm.resample_global(u_raw=0.61, return_idx=True)
ra = sm.result()


def gather_rows(a):  # rank-major concatenation along the last axis, on every rank
    a = np.ascontiguousarray(a, dtype=np.float64)
    if ws == 1:
        return a
    sizes = [int(x[0]) for x in D.all_gather_np(np.array([float(a.shape[-1])]))]
    mx = max(sizes)
    pad = np.zeros(a.shape[:-1] + (mx,))
    pad[..., :a.shape[-1]] = a
    parts = D.all_gather_np(pad)
    return np.concatenate([p[..., :s] for p, s in zip(parts, sizes)], axis=-1)


W = gather_rows(w)
G = gather_rows((idx.astype(np.int64) + sm.lo).astype(np.float64)).astype(np.int64)
Vb, Va = gather_rows(vb), gather_rows(ra.values)
Zb, Za = gather_rows(zb.astype(np.float64)), gather_rows(ra.dvalues.astype(np.float64))
if rank == 0:
    ref = api.resample_indices(1, W, [0.61]).astype(np.int64)
    out["resample_bit_exact"] = bool(np.array_equal(G, ref))
    out["migration_exact"] = bool(np.array_equal(Va, Vb[:, ref]) and np.array_equal(Za, Zb[:, ref]))
    out["parity_particles"] = int(n_par)
mig = D.all_reduce_np(np.array([float(sm.migrated)]))[0]
out["parity_migrated"] = int(mig)
sm.close()

# ---- (2) throughput at C4 ----
cfg = api.SMCConfig(ess_threshold=0.5, rejuvenation_steps=1)
D.adaptive_smc_sharded(m, min(n_total, 1 << 18), cfg, seed=5, values=False)  # warm-up (module load, NCCL channels)
torch.cuda.synchronize()
if ws > 1:
    dist.barrier()
t0 = time.perf_counter()
big = D.ShardedSmc(m, n_total, cfg, seed=7)  # prior draw + first scoring of the shard
torch.cuda.synchronize()
t1 = time.perf_counter()
while big.step() < 1.0:
    pass
torch.cuda.synchronize()
dt = time.perf_counter() - t1
init_s = t1 - t0
res = big.result(values=False)
phase = dict(big.timers)
res.migrated = big.migrated
big.close()
dt = D.all_reduce_np(np.array([dt]), op="max")[0]
mig = D.all_reduce_np(np.array([float(res.migrated)]))[0]
wsum = D.all_reduce_np(np.array([res.weight.sum()]))[0]
if rank == 0:
    out.update(seconds=dt, init_seconds=init_s, phase_seconds=phase, temper_steps=int(res.n_temper_steps), particle_updates_per_s=n_total * res.n_temper_steps / dt,
               log_evidence=res.log_evidence, migrated_particles=int(mig), weight_sum=wsum)
    print(json.dumps(out))
if ws > 1:
    dist.barrier()
    dist.destroy_process_group()
