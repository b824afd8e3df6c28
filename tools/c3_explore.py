"""Scratch: which run shape makes C3 converge (R-hat) under the reference's adaptation rules."""
import json, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fugue_b200 import api, models

G = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
Cn = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
m, truth = models.hierarchical(G=G, per_group=15)
P = api.Program(m)
d = P.d_cont
names = P.sites


def run(label, n_w, n_s, q0=None, L=16):
    t0 = time.time()
    draws, lp, st = api.hmc_chain_dev(P, n_samples=n_s, n_warmup=n_w, config=api.HMCConfig(n_leapfrog=L, adapt_mass=True), n_chains=Cn, seed=2, q0=q0)
    torch.cuda.synchronize()
    t = time.time() - t0
    dg = api.diagnostics_dev(draws)
    rh, ess = dg["r_hat"], dg["ess"]
    print(names[:3], names[-4:]) if label.startswith("data-init jitter 0.5, 500") else None
    print(json.dumps(dict(label=label, n_w=n_w, n_s=n_s, L=L, sec=t, rhat_max=float(np.nanmax(rh)), rhat_med=float(np.nanmedian(rh)), n_bad=int(np.sum(~(rh < 1.05))),
                          ess_min=float(ess.min()), ess_med=float(np.median(ess)), acc=float(st["n_accept"].mean() / (n_w + n_s)), div=float(st["n_divergent"].mean() / (n_w + n_s)),
                          eps=float(np.median(st["step_size"])), glob={k: round(float(v), 4) for k, v in zip(names, dg["mean"]) if not k.startswith("alpha")})), flush=True)
    del draws, lp


def inits(jit, seed=0):
    return models.hierarchical_inits(truth, names, Cn, jit, seed)


def pilot(n, jit=0.5, L=16):
    S = api.HmcSession(P, n_chains=Cn, n_warmup=n, config=api.HMCConfig(n_leapfrog=L), seed=1, q0=inits(jit))
    S.step(n)
    q = S.position()
    S.close()
    return q


which = sys.argv[3] if len(sys.argv) > 3 else "all"
if which == "pilot":
    for n in (200, 500):
        run(f"pilot {n} from data-init jitter 0.5, then 500+500", 500, 500, pilot(n))
    run("pilot 300, then 300+300", 300, 300, pilot(300))
if which in ("all", "data"):
    run("data-init jitter 0.5, 500+500", 500, 500, inits(0.5))
    run("data-init jitter 0.5, 1000+500", 1000, 500, inits(0.5))
if which in ("all", "prior"):
    run("prior-init 3000+500", 3000, 500)
