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
