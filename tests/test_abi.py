"""CPU-side checks of the boundary: the C-ABI library loads and exports every symbol
include/fugue_cuda.h declares (no compute calls without a GPU), the ctypes mirrors match the
header's struct layouts, and host-side validation returns the reference's error codes."""
import ctypes
import os
import re
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HDR = open(os.path.join(ROOT, "include", "fugue_cuda.h")).read()
SO = os.path.join(ROOT, "fugue_b200", "libfugue_cuda.so")


def test_every_declared_symbol_is_exported():
    lib = ctypes.CDLL(SO)
    names = sorted(set(re.findall(r"\b(fc_[a-z0-9_]+)\s*\(", HDR)))
    assert len(names) >= 55
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_header_cites_reference_interfaces():
    for anchor in ("hmc.rs:566", "mh.rs:921", "smc.rs:455", "interpreters.rs:138-163", "trace.rs:168-177", "error.rs:40-58"):
        assert anchor in HDR, anchor


def test_ctypes_struct_layouts_match_header():
    from fugue_b200 import _ir
    assert ctypes.sizeof(_ir.fc_term) == 24
    assert ctypes.sizeof(_ir.fc_param) == 3 * 24 + 8 + 16
    assert ctypes.sizeof(_ir.fc_op) == 16 + 8 + 8 + 8 + 8 + 3 * ctypes.sizeof(_ir.fc_param) + 8  # + expr_first, expr_len
    assert ctypes.sizeof(_ir.fc_expr_node) == 24
    assert ctypes.sizeof(_ir.fc_column) == 24
    assert ctypes.sizeof(_ir.fc_hmc_config) == 40
    assert ctypes.sizeof(_ir.fc_smc_config) == 24


def test_defaults_match_reference_configs():
    lib = ctypes.CDLL(SO)
    from fugue_b200 import _ir
    h = _ir.fc_hmc_config()
    lib.fc_hmc_config_default(ctypes.byref(h))
    assert (h.n_leapfrog, h.target_accept, h.finite_diff_eps, h.adapt_mass) == (16, 0.8, 1e-5, 0)  # hmc.rs:125-135
    assert h.init_step_size != h.init_step_size                                                   # None
    s = _ir.fc_smc_config()
    lib.fc_smc_config_default(ctypes.byref(s))
    assert (s.resampling_method, s.ess_threshold, s.rejuvenation_steps) == (1, 0.5, 0)             # smc.rs:181-189
    m = _ir.fc_mh_config()
    lib.fc_mh_config_default(ctypes.byref(m))
    assert (m.target_accept, m.gamma) == (0.44, 0.7)                                               # mh.rs:946


def test_constructor_validation_codes():
    """Distribution::new error codes (src/core/distribution.rs `new` fns; src/error.rs:40-49)."""
    from fugue_b200 import model as M
    from fugue_b200._ir import FugueError
    cases = [(lambda: M.Normal(float("nan"), 1.0), 100), (lambda: M.Normal(0.0, -1.0), 101), (lambda: M.Bernoulli(1.5), 102),
             (lambda: M.Uniform(2.0, 1.0), 103), (lambda: M.Beta(0.0, 1.0), 104), (lambda: M.Gamma(1.0, 0.0), 105),
             (lambda: M.Poisson(-1.0), 105), (lambda: M.StudentT(0.0, 0.0, 1.0), 104), (lambda: M.Weibull(1.0, 0.0), 101)]
    for f, code in cases:
        with pytest.raises(FugueError) as e:
            f()
        assert e.value.code == code


def test_addr_macro_escaping():
    from fugue_b200.model import addr
    assert addr("y", 3) == "y#3" and addr("a#b") == "a\\#b" and addr("a\\b", 1) == "a\\\\b#1"  # address.rs:175-223


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "fugue_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "libfugue_oracle" not in src and "import oracle" not in src and "oracle/" not in src.replace("the oracle", ""), f
