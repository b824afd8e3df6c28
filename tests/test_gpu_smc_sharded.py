"""Sharded tempered SMC (fugue_b200/dist.py ShardedSmc over the fc_smc_shard_* entry points):
world size 2, both ranks on cuda:0, gloo for the scalars (NCCL refuses two ranks on one device; the
2-GPU NCCL run of the same driver is tools/smc_sharded_check.py, recorded under profiles/).
Checks SURVEY 8e: the global systematic resample is bit-exact w.r.t. the sequential reference loop
over the rank-major population (oracle) on the weights the device holds, migration moves the right
particles, and the sharded ladder's evidence / posterior agree with the single-session run."""
import math
import os
import sys
import numpy as np
import pytest
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
Y = [1.5, 2.0, 2.5, 1.8, 2.2]


def _model():
    from fugue_b200.model import ModelBuilder, Normal, Bernoulli
    m = ModelBuilder()
    mu = m.sample("mu", Normal(0.0, 2.0))
    m.sample("flag", Bernoulli(0.3))  # a discrete plane rides along in the migration
    m.plate_observe("y", Normal(mu, 1.0), m.data(np.array(Y)))
    return m


def _worker(rank, ws, port, n_total, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    torch.cuda.set_device(0)
    from fugue_b200 import api, dist as D
    out = {}
    # ---- one reweight + global resample, inspected ----
    sm = D.ShardedSmc(_model(), n_total, api.SMCConfig(ess_threshold=0.5, rejuvenation_steps=1), seed=3)
    out["shard"] = (sm.lo, sm.hi)
    tri = sm._tri(0.05)
    lse1, _ = D._lse_pair(tri)
    api.check(api.lib.fc_smc_shard_reweight(sm.S.h, 0.05, lse1))
    sm.beta = 0.05
    w = np.zeros(sm.n)
    api.check(api.lib.fc_smc_shard_weights(sm.S.h, None, api._ptr(w)))
    v_before = sm.result().values[0].copy()
    f_before = sm.result().dvalues[0].copy()
    (jlo, jhi), idx = sm.resample_global(u_raw=0.37, return_idx=True)
    r = sm.result()
    out.update(w=w, v_before=v_before, f_before=f_before, slots=(jlo, jhi), idx=idx.copy(), v_after=r.values[0].copy(), f_after=r.dvalues[0].copy(),
               w_after=r.weight.copy(), lse1=lse1, migrated=sm.migrated)
    sm.close()
    # ---- the whole ladder ----
    res = D.adaptive_smc_sharded(_model(), n_total, api.SMCConfig(ess_threshold=0.5, rejuvenation_steps=2), seed=11)
    out.update(le=res.log_evidence, steps=res.n_temper_steps, mu=res.values[0].copy(), wt=res.weight.copy(), mig=res.migrated)
    q.put((rank, out))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_smc_world_size_2_on_one_gpu():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle
    from fugue_b200 import api
    ws, n_total = 2, 40001
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, ws, port, n_total, q)) for r in range(ws)]
    for p in procs:
        p.start()
    got = {}
    for _ in range(ws):
        r, o = q.get(timeout=600)
        got[r] = o
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    a, b = got[0], got[1]
    assert a["shard"] == (0, 20001) and b["shard"] == (20001, 40001)
    # (1) the global resample is the sequential loop over the rank-major population, bit for bit
    W = np.concatenate([a["w"], b["w"]])
    assert abs(W.sum() - 1.0) < 1e-9
    anc = oracle.resample(1, W, [0.37]).astype(np.int64)
    assert a["slots"][0] == 0 and a["slots"][1] == b["slots"][0] and b["slots"][1] == n_total
    glob = np.concatenate([a["idx"].astype(np.int64), b["idx"].astype(np.int64) + 20001])
    assert np.array_equal(glob, anc)
    # and it equals the single-GPU resampler on the same weights
    assert np.array_equal(api.resample_indices(1, W, [0.37]).astype(np.int64), anc)
    # (2) migration delivered exactly those ancestors' state (continuous and discrete planes)
    Vb, Fb = np.concatenate([a["v_before"], b["v_before"]]), np.concatenate([a["f_before"], b["f_before"]])
    Va, Fa = np.concatenate([a["v_after"], b["v_after"]]), np.concatenate([a["f_after"], b["f_after"]])
    assert np.array_equal(Va, Vb[anc]) and np.array_equal(Fa, Fb[anc])
    assert a["migrated"] + b["migrated"] == int((anc[:20001] >= 20001).sum() + (anc[20001:] < 20001).sum())
    wa = np.concatenate([a["w_after"], b["w_after"]])
    assert np.allclose(wa, 1.0 / n_total, rtol=1e-12)
    # (3) the ladder: global evidence identical on both ranks, close to the closed form, posterior right
    assert a["le"] == b["le"] and a["steps"] == b["steps"] and a["steps"] >= 2
    y = np.array(Y)
    n = len(y)
    # log marginal of y under mu ~ N(0, 2^2), y_i ~ N(mu, 1): multivariate normal with cov I + 4 11^T
    cov = np.eye(n) + 4.0 * np.ones((n, n))
    sign, logdet = np.linalg.slogdet(cov)
    log_z = -0.5 * (y @ np.linalg.solve(cov, y)) - 0.5 * logdet - 0.5 * n * math.log(2 * math.pi)
    assert abs(a["le"] - log_z) < 0.05
    mu = np.concatenate([a["mu"], b["mu"]])
    wt = np.concatenate([a["wt"], b["wt"]])
    assert abs(wt.sum() - 1.0) < 1e-9
    mean = (wt * mu).sum()
    var = (wt * (mu - mean) ** 2).sum()
    assert abs(mean - 1.90476) < 0.03 and abs(var - 0.190476) < 0.03
    single = api.adaptive_smc(_model(), n_total, api.SMCConfig(ess_threshold=0.5, rejuvenation_steps=2), seed=11)
    assert abs(single.log_evidence - a["le"]) < 0.05
    assert single.n_temper_steps == a["steps"]
