"""Host-side mirror of the reference's inference entry points over the C ABI of
include/fugue_cuda.h (fugue_b200/libfugue_cuda.so: hand-written sm_100a kernels).

Names and argument meaning follow the reference:
  hmc_chain            src/inference/hmc.rs:566
  adaptive_mcmc_chain  src/inference/mh.rs:921
  adaptive_smc         src/inference/smc.rs:455
  HMCConfig / SMCConfig / ResamplingMethod   hmc.rs:105-135, smc.rs:132-189
  r_hat_f64 / effective_sample_size_multichain / summarize_f64_parameter
                       diagnostics.rs:230,331; mcmc_utils.rs:214
The difference from the reference is batching: every entry point runs `n_chains` independent
chains (or `num_particles` particles) resident on the GPU and returns arrays, not Vec<Trace>.

There is NO CPU fallback: if the CUDA library is missing, import of this module raises.
"""
import ctypes as C
import math
import os
import numpy as np
from . import _ir
from ._ir import FugueError

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libfugue_cuda.so")
if not os.path.exists(_SO):
    raise ImportError(f"{_SO} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                      "(fugue_b200 has no CPU fallback)")
lib = C.CDLL(_SO)

dp = C.POINTER(C.c_double)
i64p = C.POINTER(C.c_int64)
u64p = C.POINTER(C.c_uint64)
vp = C.c_void_p

lib.fc_last_error.restype = C.c_char_p


def _sig(name, *argtypes):
    f = getattr(lib, name)
    f.argtypes = list(argtypes)
    f.restype = C.c_int32
    return f


_sig("fc_device_count", C.POINTER(C.c_int32))
_sig("fc_set_device", C.c_int32)
_sig("fc_program_create", vp, C.c_size_t, vp, C.c_size_t, C.POINTER(vp))
_sig("fc_program_create_ex", vp, C.c_size_t, vp, C.c_size_t, vp, C.c_size_t, C.POINTER(vp))
_sig("fc_program_destroy", vp)
_sig("fc_program_check_structure", vp, vp, C.c_size_t)
_sig("fc_program_dims", vp, i64p, i64p)
_sig("fc_program_sites", vp, C.POINTER(C.POINTER(C.c_char_p)), C.POINTER(i64p), i64p)
_sig("fc_program_stream", vp, C.POINTER(vp))
_sig("fc_program_device", vp, C.POINTER(C.c_int32))
_sig("fc_program_launch_count", vp, u64p)
_sig("fc_program_kernel_timing", vp, C.c_int32)
_sig("fc_program_kernel_time", vp, dp, u64p)
_sig("fc_logp", vp, vp, vp, C.c_size_t, vp, vp, vp)
_sig("fc_logp_grad", vp, vp, vp, C.c_size_t, vp, vp)
_sig("fc_logp_grad_dev", vp, vp, vp, C.c_size_t, vp, vp)


def check(rc):
    if rc != 0:
        raise FugueError(rc, lib.fc_last_error().decode(errors="replace"))


def set_device(i):
    check(lib.fc_set_device(int(i)))


def device_count():
    n = C.c_int32()
    check(lib.fc_device_count(C.byref(n)))
    return n.value


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _ptr(a):
    return None if a is None else a.ctypes.data_as(vp)


class Program:
    """A model lowered to the flat op list, resident on the current CUDA device."""

    def __init__(self, builder):
        builder = getattr(builder, "builder", builder)  # a dsl.CompiledModel carries its op list
        self._builder = builder
        self._low = builder.lower()  # keeps ctypes arrays and numpy columns alive during create
        ops, n, cols, ncol = self._low
        h = vp()
        nodes, n_nodes = builder.nodes_lowered
        check(lib.fc_program_create_ex(C.cast(ops, vp), n, C.cast(cols, vp), ncol, C.cast(nodes, vp), n_nodes, C.byref(h)))
        self.h = h
        dc, dd = C.c_int64(), C.c_int64()
        check(lib.fc_program_dims(h, C.byref(dc), C.byref(dd)))
        self.d_cont, self.d_disc = dc.value, dd.value
        names = C.POINTER(C.c_char_p)()
        perm = i64p()
        nn = C.c_int64()
        check(lib.fc_program_sites(h, C.byref(names), C.byref(perm), C.byref(nn)))
        self.sites = [names[i].decode() for i in range(nn.value)]
        self.perm = [perm[i] for i in range(nn.value)]
        self._low = None  # columns were copied to the device

    def close(self):
        if getattr(self, "h", None):
            lib.fc_program_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def stream(self):
        s = vp()
        check(lib.fc_program_stream(self.h, C.byref(s)))
        return s.value or 0

    @property
    def device(self):
        dv = C.c_int32()
        check(lib.fc_program_device(self.h, C.byref(dv)))
        return dv.value

    @property
    def launch_count(self):
        n = C.c_uint64()
        check(lib.fc_program_launch_count(self.h, C.byref(n)))
        return n.value

    def kernel_timing(self, enable=True):
        check(lib.fc_program_kernel_timing(self.h, int(enable)))

    def kernel_time(self):
        """(summed ms, launches) of the dominant observation-streaming kernel since kernel_timing(True)."""
        t, n = C.c_double(), C.c_uint64()
        check(lib.fc_program_kernel_time(self.h, C.byref(t), C.byref(n)))
        return t.value, n.value

    def check_structure(self, builder):
        """Second trace of the same model must match (ErrorCode 302 otherwise)."""
        low = builder.lower()
        check(lib.fc_program_check_structure(self.h, C.cast(low[0], vp), low[1]))

    def _q(self, q, qd):
        q = _f64(q).reshape(self.d_cont, -1) if self.d_cont else np.zeros((0, np.asarray(qd).reshape(self.d_disc, -1).shape[1]))
        n = q.shape[1]
        qdd = None
        if self.d_disc:
            qdd = np.ascontiguousarray(qd, dtype=np.int64).reshape(self.d_disc, n)
        return q, qdd, n

    def logp(self, q, qd=None):
        """(log_prior, log_likelihood, log_factors) per chain: Trace accumulators, trace.rs:168-177."""
        q, qdd, n = self._q(q, qd)
        lp, ll, lf = np.zeros(n), np.zeros(n), np.zeros(n)
        check(lib.fc_logp(self.h, _ptr(q), _ptr(qdd), n, _ptr(lp), _ptr(ll), _ptr(lf)))
        return lp, ll, lf

    def logp_grad(self, q, qd=None):
        """(total log-joint [C], analytic gradient [d][C])."""
        q, qdd, n = self._q(q, qd)
        lp, g = np.zeros(n), np.zeros_like(q)
        check(lib.fc_logp_grad(self.h, _ptr(q), _ptr(qdd), n, _ptr(lp), _ptr(g)))
        return lp, g

    def logp_grad_dev(self, d_q, d_qd, n, d_lp, d_grad):
        """Device pointers (ints); asynchronous on self.stream."""
        check(lib.fc_logp_grad_dev(self.h, vp(d_q), vp(d_qd) if d_qd else None, n, vp(d_lp), vp(d_grad) if d_grad else None))


# ---------------------------------------------------------------------------------------------
# HMC (src/inference/hmc.rs)
# ---------------------------------------------------------------------------------------------
i32p = C.POINTER(C.c_int32)
_sig("fc_hmc_create", vp, vp, C.c_uint64, C.c_size_t, C.c_size_t, vp, vp, C.POINTER(vp))
_sig("fc_hmc_create_sharded", vp, vp, C.c_uint64, C.c_size_t, C.c_size_t, vp, vp, C.c_uint64, C.POINTER(vp))
_sig("fc_hmc_destroy", vp)
_sig("fc_hmc_step", vp, C.c_size_t)
_sig("fc_hmc_step_recorded", vp, vp, vp, u64p)
_sig("fc_hmc_sync", vp)
_sig("fc_hmc_get", vp, vp, vp, vp)
_sig("fc_hmc_trace", vp, vp, vp, vp)
_sig("fc_hmc_device_state", vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp))
_sig("fc_hmc_set_position", vp, vp)
_sig("fc_hmc_set_step_size", vp, C.c_double)
_sig("fc_hmc_set_n_leapfrog", vp, C.c_uint64)
_sig("fc_hmc_grad_evals", vp, u64p)
_sig("fc_hmc_iterations", vp, u64p, i32p)
_sig("fc_hmc_last_accept_prob", vp, vp, vp, vp)
_sig("fc_hmc_set_explicit_randomness", vp, vp, vp)
_sig("fc_hmc_find_reasonable_epsilon", vp, vp)
_sig("fc_hmc_run", vp, vp, C.c_uint64, C.c_size_t, C.c_size_t, C.c_size_t, vp, vp, vp)
_sig("fc_hmc_run_dev", vp, vp, C.c_uint64, C.c_size_t, C.c_size_t, C.c_size_t, vp, vp, vp)
_sig("fc_hmc_sample_dev", vp, C.c_size_t, vp, vp)
lib.fc_hmc_config_default.argtypes = [C.POINTER(_ir.fc_hmc_config)]
lib.fc_hmc_config_default.restype = None


class HMCConfig:
    """HMCConfig (hmc.rs:105-135), field for field; `grad_mode` 0 = analytic force on device,
    1 = the reference's central finite differences."""

    def __init__(self, n_leapfrog=16, target_accept=0.8, init_step_size=None, finite_diff_eps=1e-5, adapt_mass=False,
                 grad_mode=0):
        self.n_leapfrog, self.target_accept, self.init_step_size = n_leapfrog, target_accept, init_step_size
        self.finite_diff_eps, self.adapt_mass, self.grad_mode = finite_diff_eps, adapt_mass, grad_mode

    def _c(self):
        c = _ir.fc_hmc_config()
        c.n_leapfrog, c.target_accept = self.n_leapfrog, self.target_accept
        c.init_step_size = float("nan") if self.init_step_size is None else float(self.init_step_size)
        c.finite_diff_eps, c.adapt_mass, c.grad_mode = self.finite_diff_eps, int(self.adapt_mass), self.grad_mode
        return c


def _as_program(model):
    return model if isinstance(model, Program) else Program(model)


class HmcSession:
    """HmcSession (hmc.rs:643-919) for n_chains independent chains resident on the GPU."""

    def __init__(self, model, n_chains=1, n_warmup=0, config=None, seed=0, q0=None, qd0=None, first_chain=0):
        self.program = _as_program(model)
        self.config = config or HMCConfig()
        self.n_chains = int(n_chains)
        d, dd = self.program.d_cont, self.program.d_disc
        q0a = None if q0 is None else _f64(q0).reshape(d, self.n_chains)
        qd0a = None if qd0 is None else np.ascontiguousarray(qd0, dtype=np.int64).reshape(dd, self.n_chains)
        cfg = self.config._c()
        h = vp()
        check(lib.fc_hmc_create_sharded(self.program.h, C.byref(cfg), seed, self.n_chains, n_warmup, _ptr(q0a), _ptr(qd0a),
                                        first_chain, C.byref(h)))
        self.h = h
        self._n_leapfrog = max(1, int(self.config.n_leapfrog))

    def close(self):
        if getattr(self, "h", None):
            lib.fc_hmc_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def sites(self):
        return list(self.program.sites)

    def step(self, n=1):
        check(lib.fc_hmc_step(self.h, n))

    def sync(self):
        check(lib.fc_hmc_sync(self.h))

    def sample_dev(self, n):
        """n further transitions, every post-transition state kept in HBM (fc_hmc_sample_dev): CUDA tensors
        draws [n][d][C] and log_joint [n][C]."""
        draws = _device_f64((n, self.program.d_cont, self.n_chains), self.program.device)
        lp = _device_f64((n, self.n_chains), self.program.device)
        check(lib.fc_hmc_sample_dev(self.h, n, vp(draws.data_ptr()), vp(lp.data_ptr())))
        self.sync()
        return draws, lp

    def step_recorded(self):
        """HmcSession::step_recorded (hmc.rs:811-817): one transition; returns the leapfrog points
        (q [L+1][d][C], h [L+1][C]) -- LeapfrogPoint{q, h} (hmc.rs:331-343) for every chain."""
        L = self._n_leapfrog
        q = np.zeros((L + 1, self.program.d_cont, self.n_chains))
        h = np.zeros((L + 1, self.n_chains))
        n = C.c_uint64()
        check(lib.fc_hmc_step_recorded(self.h, _ptr(q), _ptr(h), C.byref(n)))
        return q[:n.value], h[:n.value]

    def position(self):
        q = np.zeros((self.program.d_cont, self.n_chains))
        check(lib.fc_hmc_get(self.h, _ptr(q), None, None))
        return q

    def log_joint(self):
        lp = np.zeros(self.n_chains)
        check(lib.fc_hmc_get(self.h, None, _ptr(lp), None))
        return lp

    def trace(self):
        """(log_prior, log_likelihood, log_factors) of the current state of every chain."""
        a, b, c = np.zeros(self.n_chains), np.zeros(self.n_chains), np.zeros(self.n_chains)
        check(lib.fc_hmc_trace(self.h, _ptr(a), _ptr(b), _ptr(c)))
        return a, b, c

    def stats(self):
        st = (_ir.fc_hmc_stats * self.n_chains)()
        check(lib.fc_hmc_get(self.h, None, None, C.cast(st, vp)))
        return dict(n_accept=np.array([s.n_accept for s in st]), n_divergent=np.array([s.n_divergent for s in st]),
                    step_size=np.array([s.step_size for s in st]))

    def step_size(self):
        return self.stats()["step_size"]

    def last_step_info(self):
        ap = np.zeros(self.n_chains)
        acc = np.zeros(self.n_chains, dtype=np.int32)
        div = np.zeros(self.n_chains, dtype=np.int32)
        check(lib.fc_hmc_last_accept_prob(self.h, _ptr(ap), _ptr(acc), _ptr(div)))
        return dict(accept_prob=ap, accepted=acc.astype(bool), divergent=div.astype(bool))

    def iterations(self):
        n, w = C.c_uint64(), C.c_int32()
        check(lib.fc_hmc_iterations(self.h, C.byref(n), C.byref(w)))
        return n.value

    def is_warming_up(self):
        n, w = C.c_uint64(), C.c_int32()
        check(lib.fc_hmc_iterations(self.h, C.byref(n), C.byref(w)))
        return bool(w.value)

    def grad_evals(self):
        n = C.c_uint64()
        check(lib.fc_hmc_grad_evals(self.h, C.byref(n)))
        return n.value

    def set_position(self, q):
        q = _f64(q).reshape(self.program.d_cont, self.n_chains)
        check(lib.fc_hmc_set_position(self.h, _ptr(q)))

    def set_step_size(self, eps):
        check(lib.fc_hmc_set_step_size(self.h, float(eps)))

    def set_n_leapfrog(self, l):
        check(lib.fc_hmc_set_n_leapfrog(self.h, int(l)))
        self._n_leapfrog = max(1, int(l))

    def set_explicit_randomness(self, z, u):
        z = _f64(z).reshape(self.program.d_cont, self.n_chains)
        u = _f64(u).reshape(self.n_chains)
        check(lib.fc_hmc_set_explicit_randomness(self.h, _ptr(z), _ptr(u)))

    def find_reasonable_epsilon(self):
        e = np.zeros(self.n_chains)
        check(lib.fc_hmc_find_reasonable_epsilon(self.h, _ptr(e)))
        return e

    def device_state(self):
        a, b, c = vp(), vp(), vp()
        check(lib.fc_hmc_device_state(self.h, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value


def hmc_chain(model, n_samples, n_warmup, config=None, n_chains=1, seed=0):
    """hmc_chain (hmc.rs:566-583) for n_chains independent chains.
    Returns (draws [n_samples][d][n_chains], log_joint [n_samples][n_chains], stats)."""
    P = _as_program(model)
    cfg = (config or HMCConfig())._c()
    d = P.d_cont
    draws = np.zeros((n_samples, d, n_chains))
    lp = np.zeros((n_samples, n_chains))
    st = (_ir.fc_hmc_stats * n_chains)()
    check(lib.fc_hmc_run(P.h, C.byref(cfg), seed, n_chains, n_warmup, n_samples, _ptr(draws), _ptr(lp), C.cast(st, vp)))
    stats = dict(n_accept=np.array([s.n_accept for s in st]), n_divergent=np.array([s.n_divergent for s in st]),
                 step_size=np.array([s.step_size for s in st]))
    return draws, lp, stats


def _device_f64(shape, device):
    """HBM for draws that stay on the device. torch is the allocator only (plumbing); the library reads
    and writes the buffer through its raw pointer."""
    import torch
    return torch.empty(shape, dtype=torch.float64, device=f"cuda:{device}")


def hmc_chain_dev(model, n_samples, n_warmup, config=None, n_chains=1, seed=0, q0=None):
    """hmc_chain with the draws left in HBM (fc_hmc_run_dev): returns (draws, log_joint, stats) where draws
    [n_samples][d][n_chains] and log_joint [n_samples][n_chains] are CUDA tensors; pass draws to
    diagnostics_dev. At the C3 shape (1024 chains x 500 draws x 10 004 parameters) that is 41 GB that never
    crosses PCIe. q0 [d][n_chains]: starting points (default: a prior draw per chain, hmc.rs:673)."""
    P = _as_program(model)
    if q0 is not None:
        S = HmcSession(P, n_chains=n_chains, n_warmup=n_warmup, config=config, seed=seed, q0=q0)
        S.step(n_warmup)
        draws, lp = S.sample_dev(n_samples)
        stats = S.stats()
        S.close()
        return draws, lp, stats
    cfg = (config or HMCConfig())._c()
    d = P.d_cont
    draws = _device_f64((n_samples, d, n_chains), P.device)
    lp = _device_f64((n_samples, n_chains), P.device)
    st = (_ir.fc_hmc_stats * n_chains)()
    check(lib.fc_hmc_run_dev(P.h, C.byref(cfg), seed, n_chains, n_warmup, n_samples, vp(draws.data_ptr()), vp(lp.data_ptr()), C.cast(st, vp)))
    stats = dict(n_accept=np.array([s.n_accept for s in st]), n_divergent=np.array([s.n_divergent for s in st]),
                 step_size=np.array([s.step_size for s in st]))
    return draws, lp, stats


# ---------------------------------------------------------------------------------------------
# single-site MH (src/inference/mh.rs)
# ---------------------------------------------------------------------------------------------
class fc_site_proposal(C.Structure):
    _fields_ = [("address", C.c_char_p), ("kind", C.c_int32), ("_pad", C.c_int32), ("lower", C.c_double), ("upper", C.c_double)]


_sig("fc_mh_create", vp, vp, C.c_uint64, C.c_size_t, C.c_size_t, vp, C.c_size_t, C.c_uint64, C.POINTER(vp))
_sig("fc_mh_destroy", vp)
_sig("fc_mh_step", vp, C.c_size_t)
_sig("fc_mh_get", vp, vp, vp, vp, vp, vp)
_sig("fc_mh_n_sites", vp, i64p)
_sig("fc_mh_run", vp, vp, C.c_uint64, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, vp, vp, vp, vp)


class SiteProposal:
    """SiteProposal (mh.rs:145-161)."""
    Gaussian = ("gaussian", 0, 0.0, 0.0)
    LogSpace = ("logspace", 1, 0.0, 0.0)
    PriorResample = ("prior", 3, 0.0, 0.0)

    @staticmethod
    def Reflect(lower, upper):
        return ("reflect", 2, float(lower), float(upper))


class MhSession:
    """C independent adaptive single-site MH chains on device (adaptive_mcmc_chain, mh.rs:921-1014)."""

    def __init__(self, model, n_chains=1, n_warmup=0, seed=0, overrides=None, target_accept=0.44, gamma=0.7, first_chain=0):
        self.program = _as_program(model)
        self.n_chains = int(n_chains)
        cfg = _ir.fc_mh_config(target_accept, gamma)
        ov, n_ov = None, 0
        if overrides:
            n_ov = len(overrides)
            ov = (fc_site_proposal * n_ov)()
            for i, (a, k) in enumerate(overrides.items()):
                ov[i].address, ov[i].kind, ov[i].lower, ov[i].upper = a.encode(), k[1], k[2], k[3]
        h = vp()
        check(lib.fc_mh_create(self.program.h, C.byref(cfg), seed, self.n_chains, n_warmup, C.cast(ov, vp) if ov else None, n_ov,
                               first_chain, C.byref(h)))
        self.h = h
        n = C.c_int64()
        check(lib.fc_mh_n_sites(h, C.byref(n)))
        self.n_sites = n.value

    def close(self):
        if getattr(self, "h", None):
            lib.fc_mh_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def step(self, n=1):
        check(lib.fc_mh_step(self.h, n))

    def get(self, scales=False):
        P, n = self.program, self.n_chains
        q = np.zeros((P.d_cont, n))
        qd = np.zeros((P.d_disc, n), dtype=np.int64)
        lp = np.zeros(n)
        st = (_ir.fc_mh_stats * n)()
        sc = np.zeros((self.n_sites, n)) if scales else None
        check(lib.fc_mh_get(self.h, _ptr(q), _ptr(qd), _ptr(lp), C.cast(st, vp), _ptr(sc)))
        out = dict(q=q, qd=qd, log_joint=lp, n_accept=np.array([s.n_accept for s in st]), n_steps=st[0].n_steps)
        if scales:
            out["scales"] = sc
        return out


def adaptive_mcmc_chain(model, n_samples, n_warmup, n_chains=1, seed=0, thin=1, target_accept=0.44, gamma=0.7):
    """adaptive_mcmc_chain (mh.rs:921-936) for n_chains chains.
    Returns (draws [n_kept][d_cont][C], ddraws [n_kept][d_disc][C], log_joint [n_kept][C], stats)."""
    P = _as_program(model)
    cfg = _ir.fc_mh_config(target_accept, gamma)
    kept = n_samples // thin
    draws = np.zeros((kept, P.d_cont, n_chains))
    ddraws = np.zeros((kept, P.d_disc, n_chains), dtype=np.int64)
    lp = np.zeros((kept, n_chains))
    st = (_ir.fc_mh_stats * n_chains)()
    check(lib.fc_mh_run(P.h, C.byref(cfg), seed, n_chains, n_warmup, n_samples, thin, _ptr(draws), _ptr(ddraws), _ptr(lp), C.cast(st, vp)))
    return draws, ddraws, lp, dict(n_accept=np.array([s.n_accept for s in st]), n_steps=st[0].n_steps)


# ---------------------------------------------------------------------------------------------
# SMC (src/inference/smc.rs) and the weight numerics (src/core/numerical.rs)
# ---------------------------------------------------------------------------------------------
_sig("fc_smc_run", vp, vp, C.c_uint64, C.c_size_t, vp, vp, vp, vp, dp, u64p)
_sig("fc_smc_create", vp, vp, C.c_uint64, C.c_size_t, C.c_uint64, C.POINTER(vp))
_sig("fc_smc_step", vp, dp)
_sig("fc_smc_get", vp, vp, vp, vp, vp, dp, u64p)
_sig("fc_smc_device_state", vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp))
_sig("fc_smc_destroy", vp)
_sig("fc_smc_prior_weights", vp)
_sig("fc_smc_normalize", vp, dp, dp)
_sig("fc_smc_resample_particles", vp, C.c_int32, vp, vp)
_sig("fc_smc_rejuvenate_particles", vp, C.c_double, C.c_uint64)
_sig("fc_smc_reweight", vp, C.c_double, dp)
_sig("fc_smc_weights", vp, vp, vp)
_sig("fc_smc_counters", vp, u64p, u64p, C.POINTER(C.c_int32))
_sig("fc_smc_create_sharded", vp, vp, C.c_uint64, C.c_size_t, C.c_uint64, C.c_uint64, C.c_int32, C.c_int32, C.POINTER(vp))
_sig("fc_smc_export", vp, vp)
_sig("fc_smc_connect", vp, vp)
_sig("fc_smc_record_offspring", vp, C.c_int32)
_sig("fc_smc_last_offspring", vp, u64p, u64p, vp)
_sig("fc_smc_resample_virtual_shards", vp, C.c_int32, C.c_int32, C.c_double, vp)
_sig("fc_smc_slot_range", C.c_int32, C.c_uint64, C.c_double, C.c_uint64, C.c_uint64, C.c_double, C.c_double, C.c_int32, u64p, u64p)
_sig("fc_debug_set", C.c_char_p, C.c_int32)
_sig("fc_smc_phase_ms", vp, vp, vp)
FC_SMC_HANDLE_BYTES = 128


def debug_set(key, value):
    check(lib.fc_debug_set(key.encode(), int(value)))


def smc_slot_range(method, n_total, u_raw, carry_in, end, is_last, seed=0, tag=0):
    """Output slots [jlo, jhi) of a global systematic / stratified resample whose ancestor lives on the rank
    whose exact running sum runs from carry_in to end (host arithmetic of the sharded routing)."""
    a, b = C.c_uint64(), C.c_uint64()
    check(lib.fc_smc_slot_range(method, n_total, float(u_raw), seed, tag, float(carry_in), float(end), int(is_last), C.byref(a), C.byref(b)))
    return a.value, b.value
_sig("fc_resample", C.c_int32, vp, C.c_size_t, vp, vp)
_sig("fc_resample_dev", C.c_int32, vp, C.c_size_t, vp, vp, vp)
_sig("fc_exact_prefix_sums", vp, C.c_size_t, vp)
_sig("fc_log_sum_exp", vp, C.c_size_t, dp)
_sig("fc_ess", vp, C.c_size_t, dp)
_sig("fc_normalize_log_probs", vp, C.c_size_t, vp)
_sig("fc_next_beta", vp, vp, C.c_size_t, C.c_double, C.c_double, dp)


class ResamplingMethod:
    """ResamplingMethod (smc.rs:132-160)."""
    Multinomial, Systematic, Stratified = 0, 1, 2


class SMCConfig:
    """SMCConfig (smc.rs:172-189)."""

    def __init__(self, resampling_method=ResamplingMethod.Systematic, ess_threshold=0.5, rejuvenation_steps=0):
        self.resampling_method, self.ess_threshold, self.rejuvenation_steps = resampling_method, ess_threshold, rejuvenation_steps

    def _c(self):
        return _ir.fc_smc_config(self.resampling_method, 0, self.ess_threshold, self.rejuvenation_steps)


class SMCResult:
    """SMCResult (smc.rs:360-373) as arrays: values [d_cont][P], dvalues [d_disc][P], weight, log_weight."""

    def __init__(self, values, dvalues, weight, log_weight, log_evidence, n_temper_steps, sites):
        self.values, self.dvalues, self.weight, self.log_weight = values, dvalues, weight, log_weight
        self.log_evidence, self.n_temper_steps, self.sites = log_evidence, n_temper_steps, sites


class SmcSession:
    """The particles of adaptive_smc resident on the device. With `shard=(rank, world, n_total,
    exchange)` this process holds one rank-major shard of a population spread over the GPUs of a box:
    `exchange(bytes) -> [bytes per rank]` is any host all-gather (torch.distributed in dist.py), used
    ONCE to swap the 128-byte slab descriptors; every later exchange step runs on the device over peer
    memory (include/fugue_cuda.h, fc_smc_create_sharded)."""

    def __init__(self, model, num_particles, config=None, seed=0, first_particle=0, shard=None):
        self.program = _as_program(model)
        self.n = int(num_particles)
        cfg = (config or SMCConfig())._c()
        h = vp()
        if shard is None:
            self.rank, self.world, self.n_total = 0, 1, self.n
            check(lib.fc_smc_create(self.program.h, C.byref(cfg), seed, self.n, first_particle, C.byref(h)))
            self.h = h
        else:
            self.rank, self.world, self.n_total, exchange = shard
            check(lib.fc_smc_create_sharded(self.program.h, C.byref(cfg), seed, self.n, first_particle, self.n_total, self.rank, self.world, C.byref(h)))
            self.h = h
            mine = (C.c_char * FC_SMC_HANDLE_BYTES)()
            check(lib.fc_smc_export(h, mine))
            parts = exchange(bytes(mine))
            assert len(parts) == self.world and all(len(b) == FC_SMC_HANDLE_BYTES for b in parts)
            allh = (C.c_char * (FC_SMC_HANDLE_BYTES * self.world)).from_buffer_copy(b"".join(parts))
            check(lib.fc_smc_connect(h, allh))

    def close(self):
        if getattr(self, "h", None):
            lib.fc_smc_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def step(self):
        b = C.c_double()
        check(lib.fc_smc_step(self.h, C.byref(b)))
        return b.value

    # ---- standalone pieces of the SMC surface (smc.rs:230-349, 698-790) on the resident particles ----
    def prior_weights(self):
        """smc_prior_particles' weighting: log_weight = log-likelihood (smc.rs:779-787)."""
        check(lib.fc_smc_prior_weights(self.h))
        return self

    def normalize(self):
        """normalize_particles (smc.rs:719-754); returns (log normaliser, ESS of the weights)."""
        ln, ess = C.c_double(), C.c_double()
        check(lib.fc_smc_normalize(self.h, C.byref(ln), C.byref(ess)))
        return ln.value, ess.value

    def effective_sample_size(self):
        """effective_sample_size(particles) (smc.rs:230-233)."""
        return self.normalize()[1]

    def reweight(self, beta_new):
        """One reweight of the tempering loop (smc.rs:507-531); returns the log normaliser."""
        ln = C.c_double()
        check(lib.fc_smc_reweight(self.h, float(beta_new), C.byref(ln)))
        return ln.value

    def weights(self):
        """(log weights as they are, globally normalised weights) of this rank's particles."""
        lw, w = np.zeros(self.n), np.zeros(self.n)
        check(lib.fc_smc_weights(self.h, _ptr(lw), _ptr(w)))
        return lw, w

    def resample_particles(self, method=ResamplingMethod.Systematic, uniforms=None):
        """resample_particles (smc.rs:326-349). Returns the ancestor indices (single-rank sessions)."""
        idx = np.zeros(self.n, dtype=np.uint64) if self.world == 1 else None
        us = None if uniforms is None else _f64(np.atleast_1d(uniforms))
        if us is not None and us.size != (1 if method == ResamplingMethod.Systematic else self.n):
            raise FugueError(_ir.FC_ERR_INVALID_ARGUMENT, "uniforms: 1 value for systematic, n otherwise")
        check(lib.fc_smc_resample_particles(self.h, method, _ptr(us), _ptr(idx)))
        return idx

    def resample_virtual_shards(self, shards, method=ResamplingMethod.Systematic, u_raw=0.5):
        """resample_particles with the population processed as `shards` virtual ranks (parity hook)."""
        idx = np.zeros(self.n, dtype=np.uint64)
        check(lib.fc_smc_resample_virtual_shards(self.h, int(shards), method, float(u_raw), _ptr(idx)))
        return idx

    def record_offspring(self, enable=True):
        check(lib.fc_smc_record_offspring(self.h, int(enable)))

    def last_offspring(self):
        """(jlo, jhi, global ancestor of every slot in [jlo, jhi)) of the last resample: the slots whose
        ancestor this rank holds."""
        a, b = C.c_uint64(), C.c_uint64()
        check(lib.fc_smc_last_offspring(self.h, C.byref(a), C.byref(b), None))
        idx = np.zeros(b.value - a.value, dtype=np.uint64)
        check(lib.fc_smc_last_offspring(self.h, C.byref(a), C.byref(b), _ptr(idx)))
        return a.value, b.value, idx

    def counters(self):
        m, e, f = C.c_uint64(), C.c_uint64(), C.c_int32()
        check(lib.fc_smc_counters(self.h, C.byref(m), C.byref(e), C.byref(f)))
        return dict(migrated=m.value, ess_evals=e.value, fused=bool(f.value))

    def phase_ms(self):
        """CUDA-event ms per phase of the tempering steps (needs debug_set("smc_timing", 1))."""
        ms, n = np.zeros(5), np.zeros(5, dtype=np.uint64)
        check(lib.fc_smc_phase_ms(self.h, _ptr(ms), _ptr(n)))
        return dict(zip(("next_beta", "reweight", "scan", "gather", "move"), ms.tolist())), n.tolist()

    def rejuvenate_particles(self, beta, rejuvenation_steps):
        """rejuvenate_particles (smc.rs:698-713): weights are left unchanged."""
        check(lib.fc_smc_rejuvenate_particles(self.h, float(beta), int(rejuvenation_steps)))
        return self

    def result(self, values=True):
        P, n = self.program, self.n
        v = np.zeros((P.d_cont, n)) if values else None
        dv = np.zeros((P.d_disc, n), dtype=np.int64) if values else None
        w, lw = np.zeros(n), np.zeros(n)
        le, ns = C.c_double(), C.c_uint64()
        check(lib.fc_smc_get(self.h, _ptr(v), _ptr(dv), _ptr(w), _ptr(lw), C.byref(le), C.byref(ns)))
        return SMCResult(v, dv, w, lw, le.value, ns.value, list(P.sites))


def adaptive_smc(model, num_particles, config=None, seed=0):
    """adaptive_smc (smc.rs:455-581)."""
    P = _as_program(model)
    n = int(num_particles)
    if n == 0:
        return SMCResult(np.zeros((P.d_cont, 0)), np.zeros((P.d_disc, 0), dtype=np.int64), np.zeros(0), np.zeros(0), 0.0, 0, list(P.sites))
    S = SmcSession(P, n, config, seed)
    while S.step() < 1.0:
        pass
    r = S.result()
    S.close()
    return r


def smc_prior_particles(model, num_particles, seed=0):
    """smc_prior_particles (smc.rs:764-790): prior draws weighted by their likelihood, resident on
    the device; returns the session holding them (see its prior_weights / normalize /
    resample_particles / rejuvenate_particles / result)."""
    return SmcSession(model, num_particles, SMCConfig(), seed).prior_weights()


def resample_particles(particles, method=ResamplingMethod.Systematic, uniforms=None):
    """resample_particles (smc.rs:326-349) on a session's particles (in place); returns it."""
    particles.resample_particles(method, uniforms)
    return particles


def rejuvenate_particles(particles, beta, rejuvenation_steps):
    """rejuvenate_particles (smc.rs:698-713) on a session's particles (in place)."""
    return particles.rejuvenate_particles(beta, rejuvenation_steps)


def normalize_particles(particles):
    """normalize_particles (smc.rs:719-754)."""
    particles.normalize()
    return particles


def resample_indices(method, weights, uniforms):
    """systematic / stratified / multinomial_indices (smc.rs:255-314) with explicit uniforms."""
    w = _f64(weights)
    us = _f64(np.atleast_1d(uniforms))
    idx = np.zeros(w.size, dtype=np.uint64)
    check(lib.fc_resample(method, _ptr(w), w.size, _ptr(us), _ptr(idx)))
    return idx


def exact_prefix_sums(weights):
    w = _f64(weights)
    S = np.zeros(w.size + 1)
    check(lib.fc_exact_prefix_sums(_ptr(w), w.size, _ptr(S)))
    return S


def log_sum_exp(x):
    x = _f64(x)
    o = C.c_double()
    check(lib.fc_log_sum_exp(_ptr(x), x.size, C.byref(o)))
    return o.value


def next_beta(beta, log_w, loglik, target_ess):
    """next_beta (smc.rs:588-622): the next tempering exponent keeping ESS >= target_ess."""
    lw, ll = _f64(log_w), _f64(loglik)
    assert lw.size == ll.size
    o = C.c_double()
    check(lib.fc_next_beta(_ptr(lw), _ptr(ll), ll.size, float(beta), float(target_ess), C.byref(o)))
    return o.value


def effective_sample_size(weights):
    w = _f64(weights)
    o = C.c_double()
    check(lib.fc_ess(_ptr(w), w.size, C.byref(o)))
    return o.value


def normalize_log_probs(x):
    x = _f64(x)
    out = np.zeros_like(x)
    check(lib.fc_normalize_log_probs(_ptr(x), x.size, _ptr(out)))
    return out


# ---------------------------------------------------------------------------------------------
# diagnostics (src/inference/diagnostics.rs, src/inference/mcmc_utils.rs)
# ---------------------------------------------------------------------------------------------
_sig("fc_split_rhat", vp, C.c_size_t, C.c_size_t, dp)
_sig("fc_classic_rhat", vp, C.c_size_t, C.c_size_t, dp)
_sig("fc_ess_multichain", vp, C.c_size_t, C.c_size_t, dp)
_sig("fc_diagnostics", vp, C.c_size_t, C.c_size_t, C.c_size_t, vp, vp, vp, vp)
_sig("fc_diagnostics_dev", vp, C.c_size_t, C.c_size_t, C.c_size_t, vp, vp, vp, vp)


def _one(fn, draws):
    d = _f64(draws)
    o = C.c_double()
    check(fn(_ptr(d), d.shape[0], d.shape[1], C.byref(o)))
    return o.value


def r_hat_f64(draws):
    """split-R-hat of one parameter; draws [S][C] (diagnostics.rs:230-260)."""
    return _one(lib.fc_split_rhat, draws)


def classic_r_hat_f64(draws):
    return _one(lib.fc_classic_rhat, draws)


def effective_sample_size_multichain(draws):
    """mcmc_utils.rs:214-224; draws [S][C]."""
    return _one(lib.fc_ess_multichain, draws)


def diagnostics(draws):
    """All parameters at once; draws [S][d][C] -> dict of arrays of length d."""
    x = _f64(draws)
    S, d, Cn = x.shape
    mean, std, rh, ess = np.zeros(d), np.zeros(d), np.zeros(d), np.zeros(d)
    check(lib.fc_diagnostics(_ptr(x), S, d, Cn, _ptr(mean), _ptr(std), _ptr(rh), _ptr(ess)))
    return dict(mean=mean, std=std, r_hat=rh, ess=ess)


def diagnostics_dev(draws):
    """diagnostics() on a CUDA tensor [S][d][C] (hmc_chain_dev's draws), read in place in HBM."""
    if draws.dim() != 3 or str(draws.dtype) != "torch.float64" or not draws.is_cuda or not draws.is_contiguous():
        raise ValueError("diagnostics_dev: draws must be a contiguous CUDA float64 tensor [S][d][C]")
    S, d, Cn = draws.shape
    mean, std, rh, ess = np.zeros(d), np.zeros(d), np.zeros(d), np.zeros(d)
    check(lib.fc_diagnostics_dev(vp(draws.data_ptr()), S, d, Cn, _ptr(mean), _ptr(std), _ptr(rh), _ptr(ess)))
    return dict(mean=mean, std=std, r_hat=rh, ess=ess)


def summarize_f64_parameter(draws):
    """ParameterSummary (diagnostics.rs:321-391) of one parameter; draws [S][C]. Mean/std/R-hat/ESS
    come from the device; the five quantiles use the reference's rule round((len-1)*p) on the
    sorted pooled draws (host sort: not on the hot path)."""
    x = _f64(draws)
    if x.size == 0:
        return dict(mean=math.nan, std=math.nan, quantiles={}, r_hat=math.nan, ess=0.0)
    dg = diagnostics(x[:, None, :])
    srt = np.sort(x.reshape(-1))
    qs = {}
    for name, p in (("2.5%", 0.025), ("25%", 0.25), ("50%", 0.5), ("75%", 0.75), ("97.5%", 0.975)):
        v = (srt.size - 1) * p
        idx = int(math.floor(v + 0.5))  # f64::round: half away from zero
        qs[name] = float(srt[idx])
    return dict(mean=float(dg["mean"][0]), std=float(dg["std"][0]), quantiles=qs, r_hat=float(dg["r_hat"][0]), ess=float(dg["ess"][0]))
