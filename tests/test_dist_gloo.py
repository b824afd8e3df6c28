"""world_size-2 gloo tests (CPU) of the multi-GPU host logic in fugue_b200/dist.py: shard ranges,
cross-rank R-hat / ESS combination, the slab-descriptor exchange of a sharded SMC session, and the
rank-major exact-carry slot routing of a global systematic resample (the library's host arithmetic,
fc_smc_slot_range). Per-rank device results are stood in for by the CPU oracle /
numpy here; on the box they come from the CUDA library."""
import os
import sys
import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _seq_prefix(w, carry):
    S = np.zeros(len(w) + 1)
    c = carry
    S[0] = c
    for i, x in enumerate(w):
        c = c + x
        S[i + 1] = c
    return S


def _worker(rank, ws, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    from fugue_b200 import dist as D
    import oracle
    out = {}
    # ---- shard ranges ----
    out["shards"] = [D.shard_range(4099, r, ws) for r in range(ws)]
    # ---- R-hat / ESS across ranks ----
    rng = np.random.default_rng(7)
    S, Cn = 400, 6
    x = np.zeros((S, Cn))
    e = rng.standard_normal((S, Cn))
    for s in range(1, S):
        x[s] = 0.7 * x[s - 1] + e[s]
    x += rng.normal(size=Cn) * 0.3
    lo, hi = D.shard_range(Cn, rank, ws)
    xl = x[:, lo:hi]
    h = S // 2
    seg = np.stack([xl[:h].T, xl[h:2 * h].T], axis=1).reshape(-1, h)            # c0h0, c0h1, c1h0, ...
    seg_mean = seg.mean(axis=1)[None, :]
    seg_ss = ((seg - seg.mean(axis=1, keepdims=True)) ** 2).sum(axis=1)[None, :]
    out["rhat"] = float(D.combine_rhat(seg_mean, seg_ss, h)[0])
    out["rhat_ref"] = oracle.split_rhat(x)
    cm = xl.mean(axis=0)
    cen = xl - cm
    max_lag = min(S - 1, 2048)
    lagsum = np.array([[sum((cen[: S - t, c] * cen[t:, c]).sum() / S for c in range(xl.shape[1])) for t in range(max_lag + 1)]])
    out["ess"] = float(D.combine_ess(cm[None, :], lagsum, S)[0])
    out["ess_ref"] = oracle.ess_multichain(x)
    # ---- the one host exchange of a sharded SMC session: slab descriptors, any bytes ----
    got = D.exchange_bytes(bytes([rank]) * 128)
    out["exch"] = [g[0] for g in got] + [len(g) for g in got]
    # ---- global systematic resampling: exact carries + the library's host-side slot routing ----
    import torch
    from fugue_b200 import api
    n = 1000
    w = np.random.default_rng(11).random(n) ** 6
    w /= w.sum()
    results = {}
    for u in (0.37, 0.0, 0.999999):
        a, b = D.shard_range(n, rank, ws)
        # the exact running sum continues from the previous rank's end value (on the box: the device
        # mailbox chain of fc_smc.cu; here the same dependency through gloo)
        ends = np.zeros(ws)
        carry = 0.0
        for r in range(ws):
            if r == rank:
                Sl = _seq_prefix(w[a:b], carry)
                ends[r] = Sl[-1]
            t = torch.tensor([ends[r]], dtype=torch.float64)
            dist.broadcast(t, src=r)
            ends[r] = float(t.item())
            if r < rank:
                carry = ends[r]
        jlo, jhi = api.smc_slot_range(1, n, u, carry, ends[rank], rank == ws - 1)
        anc = []
        for j in range(jlo, jhi):
            thr = (u / float(n)) + (float(j) / float(n))            # smc.rs:258,264: two roundings
            i = int(np.searchsorted(Sl, thr, side="left"))          # first local i with S_i >= thr
            i = min(i, len(Sl) - 1)
            g = a + i                                               # global i
            anc.append(n - 1 if g == 0 else min(g - 1, n - 1))
        results[u] = (jlo, jhi, anc)
    out["resample"] = results
    out["resample_ref"] = {u: oracle.resample(1, w, [u]).tolist() for u in (0.37, 0.0, 0.999999)}
    q.put((rank, out))
    dist.barrier()
    dist.destroy_process_group()


def test_world_size_2_host_logic():
    ws = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, ws, port, q)) for r in range(ws)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in range(ws))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    r0, r1 = res[0], res[1]
    assert r0["shards"] == [(0, 2050), (2050, 4099)]
    for r in (r0, r1):
        assert abs(r["rhat"] - r["rhat_ref"]) < 1e-10 * r["rhat_ref"]
        assert abs(r["ess"] - r["ess_ref"]) < 1e-8 * r["ess_ref"]
        assert r["exch"] == [0, 1, 128, 128]
    for u, ref in r0["resample_ref"].items():
        j0, j1, a0 = r0["resample"][u]
        k0, k1, a1 = r1["resample"][u]
        # u == 0: slot 0 wraps to n-1 and is produced by the last rank
        if u == 0.0:
            assert ref[0] == 999
            got = {j: v for j, v in zip(range(j0, j1), a0)}
            got.update({j: v for j, v in zip(range(k0, k1), a1)})
            got.setdefault(0, 999)
            assert [got[j] for j in range(1000)] == ref
        else:
            assert j0 == 0 and j1 == k0 and k1 == 1000
            assert a0 + a1 == ref

