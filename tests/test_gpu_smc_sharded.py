"""Sharded tempered SMC inside the CUDA library (fc_smc_create_sharded / export / connect): world size 2,
ONE RANK PER GPU as separate processes (needs >= 2 GPUs: run with `gpurun --gpus 2`; skipped on a
one-GPU box, because ranks that spin on each other's flags must never share a GPU -- time-sliced
contexts are not guaranteed to run at the same time, B200_PROFILING.md). gloo only for the one-off swap
of the 128-byte slab descriptors; every exchange step of the ladder runs on the device through cudaIpc
peer windows over NVLink. The last 2-GPU run of this file is recorded under profiles/. On one GPU the
same kernels run with world size 1 (every other SMC test), the slot routing is tested on the CPU
(tests/test_dist_gloo.py) and the carry / routing / offset logic of the resample by
tests/test_gpu_smc.py::test_virtual_shards_resample_matches_single_population.
Checks SURVEY 8e: the global systematic / stratified resample is bit-exact w.r.t. the sequential
reference loop over the rank-major population (oracle) on the weights the device holds, migration
moves exactly those particles, and the sharded ladder's evidence / posterior agree with the
single-session run -- for the fused per-particle kernels and for the generic scoring engine."""
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
    m.sample("flag", Bernoulli(0.3))  # a discrete site rides along in the migration
    m.plate_observe("y", Normal(mu, 1.0), m.data(np.array(Y)))
    return m


def _worker(rank, ws, port, n_total, generic, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    torch.cuda.set_device(rank)
    from fugue_b200 import api, models, dist as D
    api.set_device(rank)
    api.debug_set("smc_generic", generic)
    out = {}
    try:
        for name, model in (("nn", _model()), ("mix", models.gaussian_mixture(n_obs=12)[0])):
            for method, u in ((1, [0.37]), (2, None)):
                # ---- one reweight + global resample, inspected ----
                sm = D.ShardedSmc(model, n_total, api.SMCConfig(resampling_method=method, ess_threshold=0.5, rejuvenation_steps=1), seed=3)
                sm.S.record_offspring(True)
                o = {"shard": (sm.lo, sm.hi), "fused": sm.S.counters()["fused"]}
                o["log_norm"] = sm.S.reweight(0.05)
                o["w"] = sm.S.weights()[1]
                rb = sm.S.result()
                o["v_before"], o["z_before"] = rb.values.copy(), rb.dvalues.copy()
                sm.S.resample_particles(method, u)
                jlo, jhi, idx = sm.S.last_offspring()
                ra = sm.S.result()
                o.update(slots=(jlo, jhi), idx=idx.copy(), v_after=ra.values.copy(), z_after=ra.dvalues.copy(), w_after=ra.weight.copy(),
                         migrated=sm.S.counters()["migrated"])
                if method == 2:  # the uniforms the device used for the stratified thresholds (global slot j)
                    o["resamples"] = 1
                sm.close()
                out[(name, method)] = o
        # ---- the whole ladder ----
        res = D.adaptive_smc_sharded(_model(), n_total, api.SMCConfig(ess_threshold=0.5, rejuvenation_steps=2), seed=11)
        out["ladder"] = dict(le=res.log_evidence, steps=res.n_temper_steps, mu=res.values[0].copy(), wt=res.weight.copy(), mig=res.migrated)
        # ---- a particle outside the likelihood support must not poison the final normalisation (-inf loglik) ----
        from fugue_b200.model import ModelBuilder, Normal, Uniform
        m = ModelBuilder()
        s = m.sample("s", Normal(0.0, 1.0))
        m.observe("y", Normal(0.5, s), 0.3)  # sigma <= 0 for half of the prior draws: log-likelihood -inf
        r0 = D.adaptive_smc_sharded(m, 4000, api.SMCConfig(rejuvenation_steps=0), seed=5)
        out["neg_inf"] = dict(w=r0.weight.copy(), s=r0.values[0].copy(), le=r0.log_evidence)
        # ---- unsupported on a sharded population: typed error, not a silent fallback ----
        try:
            sm = D.ShardedSmc(_model(), 1000, api.SMCConfig(resampling_method=0, rejuvenation_steps=1), seed=1)
            sm.S.reweight(0.1)
            sm.S.resample_particles(0, None)
            out["multinomial"] = "no error"
        except api.FugueError as e:
            out["multinomial"] = e.code
    except Exception as e:  # surface the failure in the parent instead of a queue timeout
        import traceback
        out["error"] = traceback.format_exc()
    q.put((rank, out))
    dist.barrier()
    dist.destroy_process_group()


def _stratified_uniforms(seed, tag, n):
    """the per-slot uniforms of the session's stratified stream (k_uniform_one keyed by the resample count)"""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle
    return oracle.philox_uniforms(seed, 0x8000000000000000, tag, n)


@pytest.mark.parametrize("generic", [0, 1])
def test_sharded_smc_world_size_2_one_rank_per_gpu(generic):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs: ranks that wait on each other must not share a GPU (run under gpurun --gpus 2)")
    import oracle
    from fugue_b200 import api
    ws, n_total = 2, 40001
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000) + generic
    procs = [ctx.Process(target=_worker, args=(r, ws, port, n_total, generic, q)) for r in range(ws)]
    for p in procs:
        p.start()
    got = {}
    for _ in range(ws):
        r, o = q.get(timeout=900)
        got[r] = o
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert "error" not in got[0], got[0]["error"]
    assert "error" not in got[1], got[1]["error"]
    for key in (("nn", 1), ("nn", 2), ("mix", 1), ("mix", 2)):
        a, b = got[0][key], got[1][key]
        method = key[1]
        assert a["shard"] == (0, 20001) and b["shard"] == (20001, 40001)
        assert a["fused"] == (not generic)
        assert a["log_norm"] == b["log_norm"]  # the global normaliser is replicated bit for bit
        # (1) the global resample is the sequential loop over the rank-major population, bit for bit
        W = np.concatenate([a["w"], b["w"]])
        assert abs(W.sum() - 1.0) < 1e-9
        us = [0.37] if method == 1 else _stratified_uniforms(3, 0x1000000000000000 | 1, n_total)
        anc = oracle.resample(method, W, us).astype(np.int64)
        assert a["slots"][0] == 0 and a["slots"][1] == b["slots"][0] and b["slots"][1] == n_total
        glob = np.concatenate([a["idx"].astype(np.int64), b["idx"].astype(np.int64)])
        assert np.array_equal(glob, anc)
        # (2) migration delivered exactly those ancestors' state (continuous and discrete sites)
        Vb, Zb = np.concatenate([a["v_before"], b["v_before"]], axis=1), np.concatenate([a["z_before"], b["z_before"]], axis=1)
        Va, Za = np.concatenate([a["v_after"], b["v_after"]], axis=1), np.concatenate([a["z_after"], b["z_after"]], axis=1)
        assert np.array_equal(Va, Vb[:, anc]) and np.array_equal(Za, Zb[:, anc])
        assert a["migrated"] + b["migrated"] == int((anc[:20001] >= 20001).sum() + (anc[20001:] < 20001).sum())
        wa = np.concatenate([a["w_after"], b["w_after"]])
        assert np.allclose(wa, 1.0 / n_total, rtol=1e-12)
    # (3) the ladder: global evidence identical on both ranks, close to the closed form, posterior right
    a, b = got[0]["ladder"], got[1]["ladder"]
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
    # (4) -inf log-likelihoods: zero weight for those particles, the others normalised (smc.rs:565-575)
    a, b = got[0]["neg_inf"], got[1]["neg_inf"]
    w, s = np.concatenate([a["w"], b["w"]]), np.concatenate([a["s"], b["s"]])
    assert np.isfinite(a["le"]) and a["le"] == b["le"]
    assert (s <= 0).any() and np.all(w[s <= 0] == 0.0) and abs(w.sum() - 1.0) < 1e-9 and np.all(w[s > 0.05] > 0.0)
    # (5) multinomial on a sharded population is a typed error (FC_ERR_UNSUPPORTED_OP = 1000)
    assert got[0]["multinomial"] == 1000 and got[1]["multinomial"] == 1000
