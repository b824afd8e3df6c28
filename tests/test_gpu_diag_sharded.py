"""Cross-rank R-hat / ESS: two ranks (gloo, both on cuda:0) each run their shard of the
chains (`first_chain` stream offsets) and combine per-rank device partials (fc_diag_partials) by
all-reduce; the result must equal the single-process device diagnostics of all chains, and the
sharded chains must be the slices of the single-process run."""
import os
import sys
import numpy as np
import pytest
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
C_TOTAL, N_WARM, N_DRAW = 10, 60, 240


def _run(first, n):
    from fugue_b200 import api, models
    m = models.hierarchical(G=6, per_group=5)[0]
    S = api.HmcSession(m, n_chains=n, n_warmup=N_WARM, config=api.HMCConfig(n_leapfrog=8), seed=21, first_chain=first)
    S.step(N_WARM)
    out = np.zeros((N_DRAW, S.program.d_cont, n))
    for t in range(N_DRAW):
        S.step(1)
        out[t] = S.position()
    S.close()
    return out


def _worker(rank, ws, port, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    torch.cuda.set_device(0)
    from fugue_b200 import dist as D
    lo, hi = D.shard_range(C_TOTAL, rank, ws)
    draws = _run(lo, hi - lo)
    dg = D.diagnostics_sharded(draws)
    dg_dev = D.diagnostics_sharded(torch.from_numpy(draws).cuda())  # the same draws resident in HBM (fc_diag_partials_dev)
    assert np.array_equal(dg["r_hat"], dg_dev["r_hat"]) and np.array_equal(dg["ess"], dg_dev["ess"])
    q.put((rank, dict(lo=lo, hi=hi, draws=draws, r_hat=dg["r_hat"], ess=dg["ess"])))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_rhat_ess_equal_single_process():
    sys.path.insert(0, ROOT)
    from fugue_b200 import api
    ws = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, ws, port, q)) for r in range(ws)]
    for p in procs:
        p.start()
    got = {}
    for _ in range(ws):
        r, o = q.get(timeout=600)
        got[r] = o
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    full = _run(0, C_TOTAL)
    # the shards ARE the slices of the single-process run (random streams keyed by the global chain index)
    for r in range(ws):
        assert np.array_equal(got[r]["draws"], full[:, :, got[r]["lo"]:got[r]["hi"]])
    ref = api.diagnostics(full)
    for r in range(ws):
        assert np.allclose(got[r]["r_hat"], ref["r_hat"], rtol=1e-10, atol=0)
        assert np.allclose(got[r]["ess"], ref["ess"], rtol=1e-8, atol=0)
    assert np.array_equal(got[0]["r_hat"], got[1]["r_hat"])
