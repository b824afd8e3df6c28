"""Multi-GPU plumbing: one process per GPU, `torch.distributed` (NCCL on the box, gloo in the CPU
tests). Chains and particles shard contiguously by rank; sampling itself needs no collective
(SURVEY 8e). The three exchange points north_star names live here:

  1. cross-chain R-hat / ESS moments        -> all-reduce of O(d * (lags + 3)) doubles
  2. global SMC weight log-sum-exp           -> all-reduce(MAX) then all-reduce(SUM)
  3. global systematic resampling            -> rank-major running sum (exact carries), ancestor
                                                routing / particle migration

Everything in this file is HOST logic on small arrays plus collectives; the per-rank numbers it
combines come from the CUDA library (fc_diag_partials, fc_exact_prefix_sums, the SMC session).
It never imports the oracle and has no CPU compute fallback for the device kernels.
"""
import math
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
    """R-hat and ESS over ALL ranks' chains. draws_local [S][d][C_local] (host array of this
    rank's draws). Per-rank partials come from the CUDA library (fc_diag_partials)."""
    from . import api
    import ctypes as C
    x = np.ascontiguousarray(draws_local, dtype=np.float64)
    S, d, Cn = x.shape
    nseg = 2 if S // 2 > 0 else 1
    max_lag = min(S - 1, 2048)
    seg_mean, seg_ss = np.zeros((d, nseg * Cn)), np.zeros((d, nseg * Cn))
    cmean, lagsum = np.zeros((d, Cn)), np.zeros((d, max_lag + 1))
    fn = api.lib.fc_diag_partials
    fn.restype = C.c_int32
    fn.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    api.check(fn(x.ctypes.data, S, d, Cn, seg_mean.ctypes.data, seg_ss.ctypes.data, cmean.ctypes.data,
                 lagsum.ctypes.data if S >= 4 else None))
    rhat = combine_rhat(seg_mean, seg_ss, S // 2 if nseg == 2 else S)
    if S >= 4:
        ess = combine_ess(cmean, lagsum, S)
    else:
        total = all_reduce_np(np.array([float(S * Cn)]))[0]
        ess = np.full(d, max(total, 1.0))
    return dict(r_hat=rhat, ess=ess)


# ---------------------------------------------------------------------------------------------
# 2. global weight log-sum-exp (numerical.rs:15-38 over all ranks' particles)
# ---------------------------------------------------------------------------------------------
def global_log_sum_exp(local_max, local_sumexp_fn):
    """Two collectives on scalars: all-reduce(MAX) of the local maxima, then all-reduce(SUM) of
    sum(exp(x - global_max)) which `local_sumexp_fn(global_max)` evaluates on this rank's
    particles (a device reduction). Edge rules as the reference: all -inf or empty -> -inf."""
    gmax = all_reduce_np(np.array([local_max], dtype=np.float64), op="max")[0]
    if gmax == -math.inf:
        return -math.inf
    if math.isnan(gmax) or gmax == math.inf:
        return gmax
    s = all_reduce_np(np.array([local_sumexp_fn(gmax)], dtype=np.float64))[0]
    return -math.inf if s == 0.0 else gmax + math.log(s)


# ---------------------------------------------------------------------------------------------
# 3. global systematic / stratified resampling across ranks (smc.rs:255-292)
#    Global particle order is rank-major, so the reference's sequential running sum over the
#    whole population is: rank 0's weights, then rank 1's starting from rank 0's end value, ...
# ---------------------------------------------------------------------------------------------
def chain_carries(local_total_fn):
    """Exact carry-in of every rank: carry[r] = the sequentially rounded running sum after the
    last particle of rank r-1. `local_total_fn(carry_in) -> end value` runs this rank's exact
    scan (fc_exact_prefix_sums with a carry). The chain is sequential over ranks (world_size
    small steps of one scalar each); returns (carry_in of this rank, list of all end values)."""
    rank, ws = world()
    ends = np.zeros(ws)
    carry = 0.0
    for r in range(ws):
        if r == rank:
            ends[r] = local_total_fn(carry)
        if ws > 1:
            t = torch.tensor([ends[r]], dtype=torch.float64, device=_dev())
            dist.broadcast(t, src=r)
            ends[r] = float(t.item())
        if r < rank:
            carry = ends[r]
    my_carry = 0.0 if rank == 0 else ends[rank - 1]
    return my_carry, ends


def systematic_threshold(j, n, u_raw):
    """threshold of output slot j exactly as coded: u = gen()/n; u + j/n (two roundings, smc.rs:258,264)."""
    nf = float(n)
    return (u_raw / nf) + (float(j) / nf)


def count_thresholds_le(x, n, u_raw):
    """#{ j in [0, n) : threshold_j <= x } by binary search (thresholds are monotone in j)."""
    lo, hi = 0, n
    while lo < hi:
        mid = (lo + hi) // 2
        if systematic_threshold(mid, n, u_raw) <= x:
            lo = mid + 1
        else:
            hi = mid
    return lo


def slot_range_of_rank(carry_in, end_value, n_total, u_raw, is_last):
    """Output slots [jlo, jhi) whose ancestor lives on this rank: the walk of smc.rs:263-268 stops
    inside this rank's particles iff carry_in < threshold_j <= end_value. Slots whose threshold is
    never reached (i == n) take particle n-1, i.e. the last rank; the u == 0 wrap (i == 0 -> n-1)
    also lands on the last rank."""
    jlo = count_thresholds_le(carry_in, n_total, u_raw)
    jhi = n_total if is_last else count_thresholds_le(end_value, n_total, u_raw)
    return jlo, max(jlo, jhi)


def migration_plan(slot_ranges, shard_bounds):
    """send_counts[src][dst] = number of offspring produced on rank src (slots slot_ranges[src])
    that belong to rank dst's output shard [shard_bounds[dst], shard_bounds[dst+1])."""
    ws = len(slot_ranges)
    m = np.zeros((ws, ws), dtype=np.int64)
    for s, (a, b) in enumerate(slot_ranges):
        for t in range(ws):
            lo, hi = max(a, shard_bounds[t]), min(b, shard_bounds[t + 1])
            m[s, t] = max(0, hi - lo)
    return m


def plan_global_systematic(carry_in, end_value, n_local, u_raw):
    """Collective: every rank contributes its exact (carry_in, end) pair; returns
    (my slot range, send-count matrix, shard bounds). The u == 0 wrap slot (j = 0 -> n-1) is
    assigned to the last rank as in the sequential code."""
    rank, ws = world()
    sizes = [int(v[0]) for v in all_gather_np(np.array([float(n_local)]))]
    bounds = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    n_total = int(bounds[-1])
    pairs = all_gather_np(np.array([carry_in, end_value]))
    ranges = []
    for r, (c, e) in enumerate(pairs):
        ranges.append(slot_range_of_rank(float(c), float(e), n_total, u_raw, r == ws - 1))
    if u_raw == 0.0 and n_total > 0:  # threshold_0 == 0: `cum < thr` fails at once, i = 0, (i-1) wraps to n-1
        ranges = [(max(a, 1), max(b, 1)) if r == 0 else (a, b) for r, (a, b) in enumerate(ranges)]
    return ranges[rank], migration_plan(ranges, bounds), bounds
