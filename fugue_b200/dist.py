"""Multi-GPU plumbing: one process per GPU, `torch.distributed` (NCCL on the box, gloo in the CPU
tests). Chains and particles shard contiguously by rank; sampling itself needs no collective
(SURVEY 8e). The three exchange points north_star names live here:

  1. cross-chain R-hat / ESS moments        -> all-reduce of O(d * (lags + 3)) doubles
  2. global SMC weight log-sum-exp           -> on the device, inside the CUDA library: the tail of
  3. global systematic resampling               every weight kernel exchanges the rank's sums through
                                                peer-memory mailboxes; offspring are written straight
                                                into the destination rank's buffers over NVLink

The R-hat / ESS part of this file is HOST logic on small arrays plus collectives; the per-rank
numbers it combines come from the CUDA library (fc_diag_partials).
It never imports the oracle and has no CPU compute fallback for the device kernels.
"""
import numpy as np
import torch
import torch.distributed as dist


def world():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_range(n, rank, world_size):
    """Contiguous partition of n units: rank r owns [lo, hi). Earlier ranks take the remainder."""
    base, rem = divmod(int(n), int(world_size))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def _dev():
    if dist.is_initialized() and dist.get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def all_reduce_np(a, op="sum"):
    """All-reduce a numpy float64 array across ranks (identity when not initialised)."""
    a = np.ascontiguousarray(a, dtype=np.float64)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return a.copy()
    t = torch.from_numpy(a.copy()).to(_dev())
    dist.all_reduce(t, op={"sum": dist.ReduceOp.SUM, "max": dist.ReduceOp.MAX, "min": dist.ReduceOp.MIN}[op])
    return t.cpu().numpy()


def all_gather_np(a):
    """Gather equal-shaped float64 arrays from every rank -> list in rank order."""
    a = np.ascontiguousarray(a, dtype=np.float64)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [a.copy()]
    t = torch.from_numpy(a.copy()).to(_dev())
    out = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return [o.cpu().numpy() for o in out]


# ---------------------------------------------------------------------------------------------
# 1. R-hat / ESS from per-rank partials (diagnostics.rs:262-304, mcmc_utils.rs:285-339)
# ---------------------------------------------------------------------------------------------
def combine_rhat(seg_mean, seg_ss, n_per_seg):
    """split-R-hat of every parameter from THIS rank's per-segment means / sums of squares
    ([d][m_local] each; a segment is one half-chain of n_per_seg draws) and the other ranks'.
    Two all-reduces of d doubles: sum of means, then sum of squared deviations + sum of ss."""
    seg_mean = np.asarray(seg_mean, dtype=np.float64)
    seg_ss = np.asarray(seg_ss, dtype=np.float64)
    d = seg_mean.shape[0]
    m = all_reduce_np(np.array([seg_mean.shape[1]], dtype=np.float64))[0]
    if m < 2:
        return np.ones(d)
    n = float(n_per_seg)
    overall = all_reduce_np(seg_mean.sum(axis=1)) / m
    red = all_reduce_np(np.stack([((seg_mean - overall[:, None]) ** 2).sum(axis=1), (seg_ss / (n - 1.0)).sum(axis=1)]))
    b = n / (m - 1.0) * red[0]
    w = red[1] / m
    var_plus = ((n - 1.0) / n) * w + (1.0 / n) * b
    with np.errstate(invalid="ignore", divide="ignore"):
        return np.sqrt(var_plus / w)


def combine_ess(cmean, lagsum, n_draws):
    """Multi-chain ESS of every parameter from per-rank chain means [d][C_local] and chain-summed
    autocovariances [d][max_lag+1] (Geyer initial positive + monotone sequences,
    mcmc_utils.rs:295-339)."""
    cmean = np.asarray(cmean, dtype=np.float64)
    lagsum = all_reduce_np(lagsum)
    d, L1 = lagsum.shape
    max_lag = L1 - 1
    nf = float(n_draws)
    m = all_reduce_np(np.array([cmean.shape[1]], dtype=np.float64))[0]
    overall = all_reduce_np(cmean.sum(axis=1)) / m
    bsum = all_reduce_np(((cmean - overall[:, None]) ** 2).sum(axis=1))
    out = np.zeros(d)
    for k in range(d):
        mean_var = lagsum[k, 0] * nf / (nf - 1.0) / m
        if mean_var <= 0.0:
            out[k] = m * nf
            continue
        var_plus = mean_var * (nf - 1.0) / nf
        if m > 1:
            var_plus += bsum[k] / (m - 1.0)
        rho = lambda t: 1.0 - (mean_var - lagsum[k, t] / m) / var_plus
        rh = np.zeros(max_lag + 1)
        rh[0] = 1.0
        if max_lag >= 1:
            rh[1] = rho(1)
        t, max_t = 1, min(1, max_lag)
        while t + 2 <= max_lag:
            re, ro = rho(t + 1), rho(t + 2)
            if re + ro < 0.0:
                break
            rh[t + 1], rh[t + 2] = re, ro
            max_t = t + 2
            t += 2
        kk = 1
        while kk + 2 <= max_t:
            prev, cur = rh[kk - 1] + rh[kk], rh[kk + 1] + rh[kk + 2]
            if cur > prev:
                rh[kk + 1] = rh[kk + 2] = prev / 2.0
            kk += 2
        tau = max(-1.0 + 2.0 * rh[: max_t + 1].sum(), 1.0)
        out[k] = m * nf / tau
    return out


def diagnostics_sharded(draws_local):
    """R-hat and ESS over ALL ranks' chains. draws_local [S][d][C_local]: this rank's draws, a host array or a
    CUDA float64 tensor (api.hmc_chain_dev / HmcSession.sample_dev: read in place in HBM). Per-rank partials come
    from the CUDA library (fc_diag_partials / fc_diag_partials_dev)."""
    from . import api
    import ctypes as C
    on_device = hasattr(draws_local, "is_cuda") and draws_local.is_cuda
    if on_device:
        if draws_local.dim() != 3 or str(draws_local.dtype) != "torch.float64" or not draws_local.is_contiguous():
            raise ValueError("diagnostics_sharded: a device tensor must be contiguous float64 [S][d][C]")
        S, d, Cn = (int(v) for v in draws_local.shape)
        ptr = draws_local.data_ptr()
        fn = api.lib.fc_diag_partials_dev
    else:
        x = np.ascontiguousarray(draws_local, dtype=np.float64)
        S, d, Cn = x.shape
        ptr = x.ctypes.data
        fn = api.lib.fc_diag_partials
    nseg = 2 if S // 2 > 0 else 1
    max_lag = min(S - 1, 2048)
    seg_mean, seg_ss = np.zeros((d, nseg * Cn)), np.zeros((d, nseg * Cn))
    cmean, lagsum = np.zeros((d, Cn)), np.zeros((d, max_lag + 1))
    fn.restype = C.c_int32
    fn.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    api.check(fn(ptr, S, d, Cn, seg_mean.ctypes.data, seg_ss.ctypes.data, cmean.ctypes.data,
                 lagsum.ctypes.data if S >= 4 else None))
    rhat = combine_rhat(seg_mean, seg_ss, S // 2 if nseg == 2 else S)
    if S >= 4:
        ess = combine_ess(cmean, lagsum, S)
    else:
        total = all_reduce_np(np.array([float(S * Cn)]))[0]
        ess = np.full(d, max(total, 1.0))
    return dict(r_hat=rhat, ess=ess)


# ---------------------------------------------------------------------------------------------
# 2./3. adaptive_smc over all ranks (smc.rs:455-581): particles sharded rank-major. The exchange steps
#    (global weight log-sum-exp / ESS, the exact running-sum carry of the global systematic
#    resample, particle migration) run INSIDE the CUDA library on the device over peer memory
#    (include/fugue_cuda.h: fc_smc_create_sharded / fc_smc_export / fc_smc_connect); this file only
#    swaps the 128-byte slab descriptors once, through whatever process group is initialised.
# ---------------------------------------------------------------------------------------------
def exchange_bytes(b):
    """All-gather one equal-length bytes object per rank -> list in rank order (host channel; used
    once per sharded session to swap cudaIpc slab descriptors)."""
    rank, ws = world()
    if ws == 1:
        return [b]
    t = torch.frombuffer(bytearray(b), dtype=torch.uint8).to(_dev())
    out = [torch.empty_like(t) for _ in range(ws)]
    dist.all_gather(out, t)
    return [bytes(o.cpu().numpy().tobytes()) for o in out]


class ShardedSmc:
    """One rank's shard of a population of `num_particles` (rank-major, contiguous). Particle j of
    the global population uses the same random stream on any world size (first_particle offsets).
    Every rank must make the same sequence of calls."""

    def __init__(self, model, num_particles, config=None, seed=0):
        from . import api
        self.api = api
        self.rank, self.ws = world()
        self.n_total = int(num_particles)
        self.lo, self.hi = shard_range(self.n_total, self.rank, self.ws)
        self.n = self.hi - self.lo
        self.config = config or api.SMCConfig()
        self.S = api.SmcSession(model, self.n, self.config, seed, first_particle=self.lo,
                                shard=(self.rank, self.ws, self.n_total, exchange_bytes))
        self.P = self.S.program

    def close(self):
        self.S.close()

    def step(self):
        """One pass of the tempering loop (smc.rs:499-560) on the global population."""
        return self.S.step()

    def result(self, values=True):
        """This rank's shard of SMCResult: values, globally normalised weights, log-evidence."""
        r = self.S.result(values)
        r.migrated = self.S.counters()["migrated"]
        return r


def adaptive_smc_sharded(model, num_particles, config=None, seed=0, values=True):
    """adaptive_smc (smc.rs:455-581) with the population sharded over all ranks; every rank returns
    its shard of the result (the log-evidence and the weight normalisation are global)."""
    sm = ShardedSmc(model, num_particles, config, seed)
    while sm.step() < 1.0:
        pass
    r = sm.result(values)
    sm.close()
    return r
