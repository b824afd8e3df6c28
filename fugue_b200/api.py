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
_sig("fc_program_destroy", vp)
_sig("fc_program_check_structure", vp, vp, C.c_size_t)
_sig("fc_program_dims", vp, i64p, i64p)
_sig("fc_program_sites", vp, C.POINTER(C.POINTER(C.c_char_p)), C.POINTER(i64p), i64p)
_sig("fc_program_stream", vp, C.POINTER(vp))
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
        self._builder = builder
        self._low = builder.lower()  # keeps ctypes arrays and numpy columns alive during create
        ops, n, cols, ncol = self._low
        h = vp()
        check(lib.fc_program_create(C.cast(ops, vp), n, C.cast(cols, vp), ncol, C.byref(h)))
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
