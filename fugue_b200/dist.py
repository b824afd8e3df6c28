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


# ---------------------------------------------------------------------------------------------
# 4. adaptive_smc over all ranks (smc.rs:455-581): particles sharded rank-major, every exchange
#    step of the ladder as a collective. Device work per rank: the fc_smc_shard_* entry points.
# ---------------------------------------------------------------------------------------------
def _ir_unsupported():
    from . import _ir
    return _ir.FC_ERR_UNSUPPORTED_OP


def _merge_tri(tris):
    """Fold per-rank (max, sum e, sum e^2) triples in rank order -- the same merge the single-GPU
    kernel applies to its per-CTA partials (fc_smc.cu tri_merge)."""
    m, s1, s2 = -math.inf, 0.0, 0.0
    for bm, b1, b2 in tris:
        if bm == -math.inf and b1 == 0.0:
            continue
        if m == -math.inf and s1 == 0.0:
            m, s1, s2 = bm, b1, b2
            continue
        mm = max(m, bm)
        fa = 1.0 if m == mm else math.exp(m - mm)
        fb = 1.0 if bm == mm else math.exp(bm - mm)
        s1, s2, m = s1 * fa + b1 * fb, s2 * fa * fa + b2 * fb * fb, mm
    return m, s1, s2


def _lse_pair(tri):
    m, s1, s2 = tri
    ninf = -math.inf
    if m != m or s1 != s1:
        return math.nan, math.nan
    lse1 = ninf if s1 == 0.0 else m + math.log(s1)
    lse2 = ninf if s2 == 0.0 else 2.0 * m + math.log(s2)
    return lse1, lse2


class ShardedSmc:
    """One rank's shard of a population of `num_particles` (rank-major, contiguous). Particle j of
    the global population uses the same random stream on any world size (first_particle offsets)."""

    def __init__(self, model, num_particles, config=None, seed=0):
        from . import api
        self.api = api
        self.rank, self.ws = world()
        self.n_total = int(num_particles)
        self.lo, self.hi = shard_range(self.n_total, self.rank, self.ws)
        self.n = self.hi - self.lo
        self.config = config or api.SMCConfig()
        self.seed = int(seed)
        self.S = api.SmcSession(model, self.n, self.config, seed, first_particle=self.lo)
        api.check(api.lib.fc_smc_shard_begin(self.S.h, self.n_total))
        self.P = self.S.program
        self.beta, self.log_evidence, self.steps = 0.0, 0.0, 0
        self.bounds = [shard_range(self.n_total, r, self.ws)[0] for r in range(self.ws)] + [self.n_total]
        self.migrated = 0  # particles that crossed ranks (all resamples)
        self.timers = {"next_beta": 0.0, "reweight": 0.0, "resample": 0.0, "rejuvenate": 0.0}  # host wall-clock per phase

    def close(self):
        self.S.close()

    # -- global (max, s1, s2) of lv = log_w + d_beta * ll: one all-gather of 3 doubles --
    def _tri(self, d_beta):
        import ctypes as C
        out = (C.c_double * 3)()
        self.api.check(self.api.lib.fc_smc_shard_tri(self.S.h, float(d_beta), out))
        return _merge_tri([tuple(t) for t in all_gather_np(np.array(list(out)))])

    def _ess_at(self, b):
        lse1, lse2 = _lse_pair(self._tri(b - self.beta))
        if not (math.isfinite(lse1) and math.isfinite(lse2)):
            return float(self.n_total)
        return math.exp(2.0 * lse1 - lse2)

    def _ess_multi(self, bs):
        """ESS at several candidate betas from ONE pass over the shard and ONE all-gather."""
        import ctypes as C
        K = len(bs)
        d = (C.c_double * K)(*[b - self.beta for b in bs])
        out = (C.c_double * (3 * K))()
        rc = self.api.lib.fc_smc_shard_tri_multi(self.S.h, d, K, out)
        if rc == _ir_unsupported():
            return [self._ess_at(b) for b in bs]
        self.api.check(rc)
        per_rank = all_gather_np(np.array(list(out)))
        res = []
        for k in range(K):
            lse1, lse2 = _lse_pair(_merge_tri([tuple(t[3 * k:3 * k + 3]) for t in per_rank]))
            res.append(float(self.n_total) if not (math.isfinite(lse1) and math.isfinite(lse2)) else math.exp(2.0 * lse1 - lse2))
        return res

    def next_beta(self, target_ess):
        """next_beta (smc.rs:588-622) on the global population. The bisection is walked three levels
        per round: the 7 midpoints the next three iterations can visit are computed with the reference's
        own expression 0.5 * (lo + hi), evaluated together, and the path through them is the one the
        sequential loop takes -- same floats, same decisions, a third of the collectives."""
        if self._ess_at(1.0) >= target_ess:
            return 1.0
        lo, hi = self.beta, 1.0
        it = 0
        if self.ws == 1:  # no collective latency to amortise: one candidate per pass (7x fewer exps)
            while it < 64:
                mid = 0.5 * (lo + hi)
                if mid == lo or mid == hi:
                    break
                if self._ess_at(mid) < target_ess:
                    hi = mid
                else:
                    lo = mid
                it += 1
            return min(max(hi, self.beta + 1e-9), 1.0)
        while it < 64:
            # candidate tree: level 1 (1 midpoint), level 2 (2), level 3 (4)
            c1 = 0.5 * (lo + hi)
            c2 = (0.5 * (lo + c1), 0.5 * (c1 + hi))
            c3 = (0.5 * (lo + c2[0]), 0.5 * (c2[0] + c1), 0.5 * (c1 + c2[1]), 0.5 * (c2[1] + hi))
            cands = [c1, c2[0], c2[1], c3[0], c3[1], c3[2], c3[3]]
            ess = dict(zip(cands, self._ess_multi(cands)))
            done = False
            for _ in range(3):
                if it >= 64:
                    break
                mid = 0.5 * (lo + hi)
                if mid == lo or mid == hi:  # fixed point of the remaining iterations
                    done = True
                    break
                if ess[mid] < target_ess:
                    hi = mid
                else:
                    lo = mid
                it += 1
            if done:
                break
        return min(max(hi, self.beta + 1e-9), 1.0)

    def _uniform(self):
        """The one uniform of a systematic resample, identical on every rank (keyed by seed, step)."""
        return float(np.random.Generator(np.random.Philox(key=[self.seed & (2**64 - 1), 0x5EED0000 + self.steps])).random())

    def resample_global(self, u_raw=None, return_idx=False):
        """Global systematic resample (smc.rs:255-272 over the rank-major population): exact carry
        chain, slot ranges, ancestor search on each rank, all-to-all migration. Returns this rank's
        global slot range and, with return_idx (tests), the local ancestor indices."""
        import ctypes as C
        api, lib, S = self.api, self.api.lib, self.S
        if u_raw is None:
            u_raw = self._uniform()

        def scan(carry):
            e = C.c_double()
            api.check(lib.fc_smc_shard_scan(S.h, float(carry), C.byref(e)))
            return e.value

        carry, ends = chain_carries(scan)
        (jlo, jhi), counts, bounds = plan_global_systematic(carry, ends[self.rank], self.n, u_raw)
        cnt = jhi - jlo
        dc, dd = self.P.d_cont, self.P.d_disc
        dev = torch.device("cuda", torch.cuda.current_device())
        idx = np.zeros(max(cnt, 1) if return_idx else 1, dtype=np.uint64)
        # my offspring per destination rank (slots are rank-major, so each destination's are contiguous).
        # Those that stay here go straight to their final local position; only the others are packed.
        send_f, send_i, col = [], [], 0
        for t in range(self.ws):
            c = int(counts[self.rank][t])
            a, b = jlo + col, jlo + col + c
            hp = idx[col:col + c].ctypes.data_as(C.c_void_p) if (c and return_idx) else None
            if t == self.rank or c == 0:
                if c:
                    api.check(lib.fc_smc_shard_offspring(S.h, self.n_total, float(u_raw), a, b, None, None, hp, a - int(bounds[self.rank])))
                send_f.append(torch.empty(0, dtype=torch.float64, device=dev))
                send_i.append(torch.empty(0, dtype=torch.int32, device=dev))
            else:
                f64 = torch.empty((dc + 3) * c, dtype=torch.float64, device=dev)
                i32 = torch.empty(max(dd, 1) * c, dtype=torch.int32, device=dev)
                api.check(lib.fc_smc_shard_offspring(S.h, self.n_total, float(u_raw), a, b, C.c_void_p(f64.data_ptr()), C.c_void_p(i32.data_ptr()), hp, 0))
                send_f.append(f64)
                send_i.append(i32[:dd * c])
            col += c
        recv_counts = [int(counts[s][self.rank]) for s in range(self.ws)]
        wire = [0 if s == self.rank else c for s, c in enumerate(recv_counts)]
        recv_f = [torch.empty((dc + 3) * c, dtype=torch.float64, device=dev) for c in wire]
        recv_i = [torch.empty(dd * c, dtype=torch.int32, device=dev) for c in wire]
        if self.ws > 1:
            self._all_to_all(recv_f, send_f)
            if dd:
                self._all_to_all(recv_i, send_i)
        # NCCL enqueues on torch's streams; the library installs on the program's own stream
        torch.cuda.synchronize()
        off = 0
        for s_rank, c in enumerate(recv_counts):
            if c and s_rank != self.rank:
                api.check(lib.fc_smc_shard_install(S.h, off, c, C.c_void_p(recv_f[s_rank].data_ptr()),
                                                   C.c_void_p(recv_i[s_rank].data_ptr()) if dd else None))
            off += c
        assert off == self.n, (off, self.n)
        torch.cuda.synchronize()
        api.check(lib.fc_smc_shard_resampled(S.h))
        self.migrated += sum(c for s_rank, c in enumerate(recv_counts) if s_rank != self.rank)
        return (jlo, jhi), (idx[:cnt] if return_idx else None)

    def _all_to_all(self, recv, send):
        if self.ws == 1:
            recv[0].copy_(send[0])
            return
        if dist.get_backend() == "nccl":
            dist.all_to_all(recv, send)
            return
        # gloo (CPU tests on one GPU): stage through host tensors with point-to-point exchanges
        for src in range(self.ws):
            for dst in range(self.ws):
                if src == dst == self.rank:
                    recv[src].copy_(send[dst])
                elif src == self.rank:
                    if send[dst].numel():
                        dist.send(send[dst].cpu(), dst)
                elif dst == self.rank and recv[src].numel():
                    buf = torch.empty(recv[src].shape, dtype=recv[src].dtype)
                    dist.recv(buf, src)
                    recv[src].copy_(buf)

    def step(self):
        """One pass of the tempering loop (smc.rs:499-560) on the global population."""
        api, lib, S = self.api, self.api.lib, self.S
        n = self.n_total
        target = min(max(self.config.ess_threshold * n, 1.0), float(n))
        import time
        self.steps += 1
        t0 = time.perf_counter()
        beta_new = self.next_beta(target)
        if self.steps >= 10000:
            beta_new = 1.0
        t1 = time.perf_counter()
        lse1, _ = _lse_pair(self._tri(beta_new - self.beta))
        log_norm = lse1
        self.log_evidence += log_norm
        api.check(lib.fc_smc_shard_reweight(S.h, float(beta_new), float(log_norm)))
        self.beta = beta_new
        t2 = time.perf_counter()
        t3 = t4 = t2
        if self.beta < 1.0:
            self.resample_global()
            t3 = time.perf_counter()
            if self.config.rejuvenation_steps:
                S.rejuvenate_particles(self.beta, self.config.rejuvenation_steps)
            t4 = time.perf_counter()
        for k, v in (("next_beta", t1 - t0), ("reweight", t2 - t1), ("resample", t3 - t2), ("rejuvenate", t4 - t3)):
            self.timers[k] += v
        return self.beta

    def result(self, values=True):
        """This rank's shard of SMCResult: values, normalised global weights, log-evidence."""
        import ctypes as C
        api, lib, S, P = self.api, self.api.lib, self.S, self.P
        v = np.zeros((P.d_cont, self.n)) if values else None
        dv = np.zeros((P.d_disc, self.n), dtype=np.int64) if values else None
        if values:
            api.check(lib.fc_smc_get(S.h, api._ptr(v), api._ptr(dv), None, None, None, None))
        lw = np.zeros(self.n)
        api.check(lib.fc_smc_shard_weights(S.h, api._ptr(lw), None))
        log_norm, _ = _lse_pair(self._tri(0.0))  # final normalisation (smc.rs:565-575): device reduction + one all-gather
        if math.isfinite(log_norm):
            lwn = lw - log_norm
            w = np.exp(lwn)
        else:
            lwn = np.full(self.n, -math.log(self.n_total))
            w = np.full(self.n, 1.0 / self.n_total)
        return self.api.SMCResult(v, dv, w, lwn, self.log_evidence, self.steps, list(P.sites))


def adaptive_smc_sharded(model, num_particles, config=None, seed=0, values=True):
    """adaptive_smc (smc.rs:455-581) with the population sharded over all ranks; every rank returns
    its shard of the result (the log-evidence is global). rejuvenation_steps == 0 degenerates to the
    single importance-sampling reweight (smc.rs:484-493), which needs only the global log-sum-exp."""
    from . import api
    cfg = config or api.SMCConfig()
    sm = ShardedSmc(model, num_particles, cfg, seed)
    if cfg.rejuvenation_steps == 0:
        lse1, _ = _lse_pair(sm._tri(1.0))
        sm.log_evidence = lse1
        api.check(api.lib.fc_smc_shard_reweight(sm.S.h, 1.0, 0.0))  # log_w = -ln N + ll (smc.rs:490-493), normalised in result()
        sm.beta, sm.steps = 1.0, 1
    else:
        while sm.step() < 1.0:
            pass
    r = sm.result(values)
    r.migrated = sm.migrated
    sm.close()
    return r
