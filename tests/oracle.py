"""ctypes driver for oracle/libfugue_oracle.so (TEST INFRASTRUCTURE; the checker, never the product)."""
import ctypes as C
import os
import numpy as np
from fugue_b200 import _ir

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_lib = None
dp = C.POINTER(C.c_double)
ip = C.POINTER(C.c_int64)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(os.path.join(ROOT, "oracle", "libfugue_oracle.so"))
        for n in ("fo_dist_logp", "fo_categorical_lp", "fo_binomial_lp", "fo_discrete_uniform_lp", "fo_digamma",
                  "fo_log_sum_exp", "fo_log1p_exp", "fo_safe_ln", "fo_ess", "fo_next_beta", "fo_split_rhat",
                  "fo_classic_rhat", "fo_ess_multichain", "fo_dual_averaging", "fo_find_reasonable_epsilon"):
            getattr(L, n).restype = C.c_double
        L.fo_dist_logp.argtypes = [C.c_int, dp, C.c_double, C.c_int64]
        L.fo_dist_eval.argtypes = [C.c_int, dp, C.c_double, C.c_int64, dp]
        L.fo_categorical_lp.argtypes = [dp, C.c_size_t, C.c_uint64]
        L.fo_binomial_lp.argtypes = [C.c_uint64, C.c_double, C.c_uint64]
        L.fo_discrete_uniform_lp.argtypes = [C.c_int64, C.c_int64, C.c_int64]
        L.fo_digamma.argtypes = [C.c_double]
        L.fo_log_sum_exp.argtypes = [dp, C.c_size_t]
        L.fo_normalize_log_probs.argtypes = [dp, C.c_size_t, dp]
        L.fo_log1p_exp.argtypes = [C.c_double]
        L.fo_safe_ln.argtypes = [C.c_double]
        L.fo_program_create.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]
        L.fo_program_create_ex.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]
        L.fo_program_destroy.argtypes = [C.c_void_p]
        L.fo_program_dims.argtypes = [C.c_void_p, ip, ip]
        L.fo_program_site.argtypes = [C.c_void_p, C.c_int64, C.c_int]
        L.fo_program_site.restype = C.c_char_p
        L.fo_program_perm.argtypes = [C.c_void_p, C.c_int64, C.c_int]
        L.fo_program_perm.restype = C.c_int64
        L.fo_logp.argtypes = [C.c_void_p, dp, ip, C.c_size_t, dp, dp, dp, dp, C.c_int]
        L.fo_grad_fd.argtypes = [C.c_void_p, dp, ip, C.c_size_t, C.c_double, dp, C.POINTER(C.c_int32), C.c_int]
        L.fo_hmc_transition.argtypes = [C.c_void_p, ip, dp, dp, dp, C.c_double, C.c_double, C.c_uint64, dp, dp,
                                        C.c_int, C.c_double, dp]
        L.fo_find_reasonable_epsilon.argtypes = [C.c_void_p, ip, dp, C.c_double, dp, dp, dp, C.c_int, C.c_double]
        L.fo_dual_averaging.argtypes = [C.c_double, C.c_double, dp, C.c_size_t, dp]
        L.fo_welford.argtypes = [dp, C.c_size_t, C.c_size_t, dp]
        L.fo_dim_adapt.argtypes = [C.POINTER(C.c_uint8), C.c_size_t, C.c_double, C.c_double, dp]
        L.fo_resample.argtypes = [C.c_int, dp, C.c_size_t, dp, C.POINTER(C.c_uint64)]
        L.fo_ess.argtypes = [dp, C.c_size_t]
        L.fo_next_beta.argtypes = [C.c_double, dp, dp, C.c_size_t, C.c_double]
        for n in ("fo_split_rhat", "fo_classic_rhat", "fo_ess_multichain"):
            getattr(L, n).argtypes = [dp, C.c_size_t, C.c_size_t]
        L.fo_n_sites.argtypes = [C.c_void_p]
        L.fo_n_sites.restype = C.c_int64
        L.fo_mh_step.argtypes = [C.c_void_p, dp, ip, C.c_double, C.c_int64, C.c_int, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double,
                                 C.c_double, C.c_int64, dp]
        L.fo_mh_kind.argtypes = [C.c_void_p, dp, ip, C.c_int64]
        L.fo_tempered_step.argtypes = [C.c_void_p, dp, ip, C.c_int64, C.c_double, C.c_double, C.c_double, C.c_double, dp]
        L.fo_normalize_particles.argtypes = [dp, C.c_size_t, dp]
        L.fo_adaptive_smc.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_double, C.c_uint64, C.c_uint64, C.c_int, dp, ip, dp, dp, dp, C.POINTER(C.c_uint64)]
        _lib = L
    return _lib


def _d(a):
    return a.ctypes.data_as(dp)


def f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def dist_logp(dist, params, x=0.0, xi=0):
    P = f64(list(params) + [0.0] * (3 - len(params)))
    return lib().fo_dist_logp(dist, _d(P), float(x), int(xi))


def dist_eval(dist, params, x=0.0, xi=0):
    P = f64(list(params) + [0.0] * (3 - len(params)))
    out = np.zeros(5)
    lib().fo_dist_eval(dist, _d(P), float(x), int(xi), _d(out))
    return out


def log_sum_exp(x):
    x = f64(x)
    return lib().fo_log_sum_exp(_d(x), x.size)


def normalize_log_probs(x):
    x = f64(x)
    out = np.zeros_like(x)
    lib().fo_normalize_log_probs(_d(x), x.size, _d(out))
    return out


class OracleProgram:
    """Sequential CPU scorer of the same op list the device gets."""

    def __init__(self, builder):
        self._b = builder
        self._low = builder.lower()  # keep ctypes arrays (and numpy data) alive
        ops, n, cols, ncol = self._low
        h = C.c_void_p()
        nodes, n_nodes = builder.nodes_lowered
        rc = lib().fo_program_create_ex(C.cast(ops, C.c_void_p), n, C.cast(cols, C.c_void_p), ncol, C.cast(nodes, C.c_void_p), n_nodes, C.byref(h))
        if rc != 0:
            raise _ir.FugueError(rc, "oracle program create")
        self.h = h
        dc, dd = C.c_int64(), C.c_int64()
        lib().fo_program_dims(h, C.byref(dc), C.byref(dd))
        self.d_cont, self.d_disc = dc.value, dd.value

    def __del__(self):
        if getattr(self, "h", None):
            lib().fo_program_destroy(self.h)
            self.h = None

    def sites(self, discrete=False):
        n = self.d_disc if discrete else self.d_cont
        return [lib().fo_program_site(self.h, i, int(discrete)).decode() for i in range(n)]

    def _qd(self, qd, Cn):
        if self.d_disc == 0:
            return None, None
        a = np.ascontiguousarray(qd, dtype=np.int64).reshape(self.d_disc, Cn)
        return a, a.ctypes.data_as(ip)

    def logp(self, q, qd=None, grad=False, nthreads=1):
        if self.d_cont == 0:  # discrete-only program: the chain count comes from qd
            Cn = np.asarray(qd).reshape(self.d_disc, -1).shape[1]
            q = np.zeros((0, Cn))
        else:
            q = f64(q).reshape(self.d_cont, -1)
            Cn = q.shape[1]
        a, qdp = self._qd(qd, Cn)
        lp, ll, lf = np.zeros(Cn), np.zeros(Cn), np.zeros(Cn)
        g = np.zeros_like(q) if grad else None
        lib().fo_logp(self.h, _d(q), qdp, Cn, _d(lp), _d(ll), _d(lf), _d(g) if grad else None, nthreads)
        return (lp, ll, lf, g) if grad else (lp, ll, lf)

    # ---- single-site MH / tempered SMC with explicit randomness (oracle/fo_smc.hpp) ----
    def _state(self, q, qd):
        q = f64(q).reshape(self.d_cont).copy()
        d = np.ascontiguousarray(qd if qd is not None else np.zeros(self.d_disc), dtype=np.int64).reshape(self.d_disc).copy()
        return q, d

    def n_sites(self):
        return lib().fo_n_sites(self.h)

    def mh_step(self, q, qd, cur_lw, g, z, u, scale=1.0, kind=-1, lower=0.0, upper=0.0, prior_x=0.0, prior_xi=0):
        """single_site_mh_step (mh.rs:698-744) on global site g -> (q', qd', accept, log_alpha, lw')."""
        q, d = self._state(q, qd)
        out = np.zeros(3)
        lib().fo_mh_step(self.h, _d(q), d.ctypes.data_as(ip), cur_lw, g, kind, scale, z, u, lower, upper, prior_x, prior_xi, _d(out))
        return q, d, bool(out[0]), out[1], out[2]

    def mh_kind(self, q, qd, g):
        q, d = self._state(q, qd)
        return lib().fo_mh_kind(self.h, _d(q), d.ctypes.data_as(ip), g)

    def tempered_step(self, q, qd, k, beta, scale, z, u):
        """tempered_single_site_mh (smc.rs:631-688) on the k-th f64 site -> (q', accept, log_alpha, (lp, ll, lf))."""
        q, d = self._state(q, qd)
        out = np.zeros(5)
        lib().fo_tempered_step(self.h, _d(q), d.ctypes.data_as(ip), k, beta, scale, z, u, _d(out))
        return q, bool(out[0]), out[1], out[2:5].copy()

    def adaptive_smc(self, n, rejuvenation_steps=0, ess_threshold=0.5, method=1, seed=0, nthreads=1, values=True):
        """adaptive_smc (smc.rs:455-581) -> dict(values, dvalues, weight, log_weight, log_evidence, steps)."""
        v = np.zeros((self.d_cont, n)) if values else None
        dv = np.zeros((self.d_disc, n), dtype=np.int64) if values else None
        w, lw = np.zeros(n), np.zeros(n)
        le, st = C.c_double(), C.c_uint64()
        rc = lib().fo_adaptive_smc(self.h, n, method, ess_threshold, rejuvenation_steps, seed, nthreads, _d(v) if values else None,
                                   dv.ctypes.data_as(ip) if values else None, _d(w), _d(lw), C.byref(le), C.byref(st))
        if rc != 0:
            raise _ir.FugueError(rc, "oracle adaptive_smc: a prior family that is not restated")
        return dict(values=v, dvalues=dv, weight=w, log_weight=lw, log_evidence=le.value, steps=st.value)

    def grad_fd(self, q, qd=None, h=1e-5, nthreads=1):
        q = f64(q).reshape(self.d_cont, -1)
        Cn = q.shape[1]
        a, qdp = self._qd(qd, Cn)
        g = np.zeros_like(q)
        ok = np.zeros(Cn, dtype=np.int32)
        lib().fo_grad_fd(self.h, _d(q), qdp, Cn, h, _d(g), ok.ctypes.data_as(C.POINTER(C.c_int32)), nthreads)
        return g, ok

    def hmc_transition(self, q, lj, z, u, eps, l, m_inv=None, mass_sqrt=None, qd=None, grad_mode=0, h=1e-5):
        q = f64(q).copy()
        d = q.size
        m_inv = f64(np.ones(d) if m_inv is None else m_inv)
        mass_sqrt = f64(np.ones(d) if mass_sqrt is None else mass_sqrt)
        z = f64(z)
        ljc = C.c_double(lj)
        out = np.zeros(4)
        a, qdp = self._qd(qd, 1)
        lib().fo_hmc_transition(self.h, qdp, _d(q), C.byref(ljc), _d(z), u, eps, l, _d(m_inv), _d(mass_sqrt),
                                grad_mode, h, _d(out))
        return q, ljc.value, dict(accepted=bool(out[0]), divergent=bool(out[1]), accept_prob=out[2], lj_new=out[3])

    def find_reasonable_epsilon(self, q, lj, z, m_inv=None, mass_sqrt=None, qd=None, grad_mode=0, h=1e-5):
        q = f64(q)
        d = q.size
        m_inv = f64(np.ones(d) if m_inv is None else m_inv)
        mass_sqrt = f64(np.ones(d) if mass_sqrt is None else mass_sqrt)
        z = f64(z)
        a, qdp = self._qd(qd, 1)
        return lib().fo_find_reasonable_epsilon(self.h, qdp, _d(q), lj, _d(z), _d(m_inv), _d(mass_sqrt), grad_mode, h)


def normalize_particles(log_w):
    lw = f64(log_w)
    w = np.zeros_like(lw)
    lib().fo_normalize_particles(_d(lw), lw.size, _d(w))
    return w


def resample(method, w, us):
    w, us = f64(w), f64(np.atleast_1d(us))
    idx = np.zeros(w.size, dtype=np.uint64)
    lib().fo_resample(method, _d(w), w.size, _d(us), idx.ctypes.data_as(C.POINTER(C.c_uint64)))
    return idx


def ess(w):
    w = f64(w)
    return lib().fo_ess(_d(w), w.size)


def next_beta(beta, log_w, ll, target):
    log_w, ll = f64(log_w), f64(ll)
    return lib().fo_next_beta(beta, _d(log_w), _d(ll), ll.size, target)


def split_rhat(draws):
    d = f64(draws)
    return lib().fo_split_rhat(_d(d), d.shape[0], d.shape[1])


def classic_rhat(draws):
    d = f64(draws)
    return lib().fo_classic_rhat(_d(d), d.shape[0], d.shape[1])


def ess_multichain(draws):
    d = f64(draws)
    return lib().fo_ess_multichain(_d(d), d.shape[0], d.shape[1])


def dual_averaging(eps0, target, alphas):
    a = f64(alphas)
    out = np.zeros_like(a)
    frozen = lib().fo_dual_averaging(eps0, target, _d(a), a.size, _d(out))
    return out, frozen


def welford(x):
    x = f64(x)
    out = np.zeros(x.shape[1])
    lib().fo_welford(_d(x), x.shape[0], x.shape[1], _d(out))
    return out


def dim_adapt(accepts, target=0.44, gamma=0.7):
    a = np.ascontiguousarray(accepts, dtype=np.uint8)
    out = np.zeros(a.size)
    lib().fo_dim_adapt(a.ctypes.data_as(C.POINTER(C.c_uint8)), a.size, target, gamma, _d(out))
    return out


def philox_uniforms(seed, stream, tag, n):
    """The n uniforms of the device's Philox4x32-10 stream (fc_common.cuh): key from (seed, stream) by
    splitmix64, block i = counter (i, tag) gives uniforms 2i and 2i+1 (53 bits each). Test helper for
    resamplers that draw their own uniforms (stratified thresholds of an SMC session)."""
    M64 = (1 << 64) - 1
    z = (seed + 0x9E3779B97F4A7C15 * (stream + 1)) & M64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M64
    z ^= z >> 31
    m = (n + 1) // 2
    i = np.arange(m, dtype=np.uint64)
    c0 = (i & np.uint64(0xFFFFFFFF)).astype(np.uint64)
    c1 = (i >> np.uint64(32)).astype(np.uint64)
    c2 = np.full(m, tag & 0xFFFFFFFF, dtype=np.uint64)
    c3 = np.full(m, (tag >> 32) & 0xFFFFFFFF, dtype=np.uint64)
    a, b = z & 0xFFFFFFFF, (z >> 32) & 0xFFFFFFFF
    mask = np.uint64(0xFFFFFFFF)
    for _ in range(10):
        m0 = np.uint64(0xD2511F53) * c0
        m1 = np.uint64(0xCD9E8D57) * c2
        hi0, lo0, hi1, lo1 = m0 >> np.uint64(32), m0 & mask, m1 >> np.uint64(32), m1 & mask
        n0, n1, n2, n3 = hi1 ^ c1 ^ np.uint64(a), lo1, hi0 ^ c3 ^ np.uint64(b), lo0
        c0, c1, c2, c3 = n0, n1, n2, n3
        a, b = (a + 0x9E3779B9) & 0xFFFFFFFF, (b + 0xBB67AE85) & 0xFFFFFFFF
    x = (c0 << np.uint64(32)) | c1
    y = (c2 << np.uint64(32)) | c3
    u = np.empty(2 * m)
    u[0::2] = (x >> np.uint64(11)).astype(np.float64) * 2.0 ** -53
    u[1::2] = (y >> np.uint64(11)).astype(np.float64) * 2.0 ** -53
    return u[:n]
